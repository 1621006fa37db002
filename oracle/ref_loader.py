"""Imports the unmodified reference `segment_anything` from oracle/_ref (placed there by oracle/install_ref.py) -- TEST
INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use it."""
from __future__ import annotations

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "segment_anything", "predictor.py"))


def load():
    """Returns the reference's `segment_anything` module, or None when oracle/_ref is absent."""
    if not available():
        return None
    sys.dont_write_bytecode = True
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import segment_anything  # noqa: E402  (the copy under oracle/_ref)

    if not os.path.abspath(segment_anything.__file__).startswith(REF_ROOT):
        raise RuntimeError(f"another segment_anything shadows oracle/_ref: {segment_anything.__file__}")
    return segment_anything


def reference_predictor(cfg, state_dict):
    """SamPredictor(Sam) of the reference (SA/build_sam.py:55-107, SA/predictor.py:17-32) on CPU with `state_dict`
    loaded strictly.  `cfg` needs embed_dim / depth / num_heads / global_attn_indexes."""
    sa = load()
    if sa is None:
        raise RuntimeError("oracle/_ref is absent: run `python oracle/install_ref.py` where /root/reference exists")
    from segment_anything.build_sam import _build_sam

    sam = _build_sam(cfg.embed_dim, cfg.depth, cfg.num_heads, list(cfg.global_attn_indexes))
    sam.load_state_dict(state_dict, strict=True)
    return sa.SamPredictor(sam)
