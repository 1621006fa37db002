"""Error structure of two builds' ViT-H embeddings against the fp32 oracle (files written by tools/emb_dump.py)."""
import sys

import numpy as np


def rms(a):
    return float(np.sqrt((a.astype(np.float64) ** 2).mean()))


ref_e, ref_l = np.load(sys.argv[1] + "_emb.npy"), np.load(sys.argv[1] + "_low.npy")
for p in sys.argv[2:]:
    e, l = np.load(p + "_emb.npy"), np.load(p + "_low.npy")
    d = (e - ref_e)[0]                                   # [256][64][64]
    ch_mean = d.mean(axis=(1, 2))                        # error common to all tokens of a channel
    tok_mean = d.mean(axis=0)                            # error common to all channels of a token
    print(f"{p}: emb rel-RMS {rms(d) / rms(ref_e):.3e}  per-channel-mean part {rms(ch_mean) / rms(ref_e):.3e}  "
          f"per-token-mean part {rms(tok_mean) / rms(ref_e):.3e}  low-res rel-RMS {rms(l - ref_l) / rms(ref_l):.3e}  "
          f"on the ::4 grid {rms((l - ref_l)[:, :, ::4, ::4]) / rms(ref_l[:, :, ::4, ::4]):.3e}  "
          f"low-res mean err {float((l - ref_l).mean()):+.3e} (rms of ref {rms(ref_l):.3e})")
