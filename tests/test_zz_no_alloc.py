"""inksam_resize_u8_static: the allocation-free variant of inksam_resize_u8 (scratch in a static __device__ array of
inksam_resize_static_max_bytes() bytes; include/inksam_ops.h).  No counterpart in the reference.
Called through the C ABI and compared, byte for byte (the resize is integer work: the bar is bit-exact),
with the existing entry point on two calls in a row that share the scratch.  Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

from inklayer_b200 import _lib

pytestmark = pytest.mark.gpu


def ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _scaled(hw, long_side=1024):
    s = long_side / max(hw)
    return int(hw[0] * s + 0.5), int(hw[1] * s + 0.5)


def _reference(lib, src, hw, out_hw):
    """The existing entry point with a caller-owned workspace."""
    dst = torch.full((1024, 1024, 3), 9, dtype=torch.uint8, device="cuda")
    need = int(lib.inksam_resize_workspace_bytes(hw[0], hw[1], out_hw[0], out_hw[1]))
    ws = torch.empty(max(need, 1), dtype=torch.uint8, device="cuda")
    _lib.check(lib.inksam_resize_u8(ptr(src), hw[0], hw[1], hw[1] * 3, ptr(dst), out_hw[0], out_hw[1], 1024 * 3, ptr(ws), stream()))
    torch.cuda.synchronize()
    return dst


def test_static_scratch_resize_matches_the_workspace_entry_point_twice_in_a_row():
    lib = _lib.load()
    assert int(lib.inksam_resize_static_max_bytes()) == 16 << 20
    # the second call is smaller than the first: it runs on scratch that still holds the first call's tables and pixels
    cases = [(2000, 1500), (600, 750), (1333, 751)]
    srcs, refs, outs = [], [], []
    for hw in cases:
        img = np.random.default_rng(hw[0] * 7 + hw[1]).integers(0, 256, hw + (3,), dtype=np.uint8)
        src = torch.from_numpy(img).cuda()
        srcs.append(src)
        refs.append(_reference(lib, src, hw, _scaled(hw)))
        outs.append(torch.full((1024, 1024, 3), 9, dtype=torch.uint8, device="cuda"))
    torch.cuda.synchronize()
    torch_bytes = torch.cuda.memory_allocated()
    for i, hw in enumerate(cases):
        h, w = _scaled(hw)
        free_before = torch.cuda.mem_get_info()[0] if i > 0 else None  # (the first call may load the module lazily)
        _lib.check(lib.inksam_resize_u8_static(ptr(srcs[i]), hw[0], hw[1], hw[1] * 3, ptr(outs[i]), h, w, 1024 * 3, stream()))
        torch.cuda.synchronize()
        if free_before is not None:
            assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the device's free memory"
    assert torch.cuda.memory_allocated() == torch_bytes
    for i, hw in enumerate(cases):
        assert torch.equal(outs[i], refs[i]), f"call {i} ({hw}) differs from inksam_resize_u8"


def test_static_scratch_resize_rejects_sizes_beyond_the_array():
    lib = _lib.load()
    hw, out_hw = (9000, 1200), (7680, 1024)  # intermediate 9000 x 1024 x 3 = 27.6 MB > 16 MiB
    assert int(lib.inksam_resize_workspace_bytes(hw[0], hw[1], out_hw[0], out_hw[1])) > int(lib.inksam_resize_static_max_bytes())
    src = torch.zeros(16, dtype=torch.uint8, device="cuda")  # never read: the size check comes first
    rc = lib.inksam_resize_u8_static(ptr(src), hw[0], hw[1], hw[1] * 3, ptr(src), out_hw[0], out_hw[1], out_hw[1] * 3, stream())
    assert rc != 0
    msg = lib.inksam_last_error().decode()
    assert "static array holds" in msg and "inksam_resize_u8_static" in msg
    # identity size (no scratch needed) still works and copies
    img = torch.arange(5 * 7 * 3, dtype=torch.uint8, device="cuda").view(5, 7, 3)
    out = torch.zeros_like(img)
    _lib.check(lib.inksam_resize_u8_static(ptr(img), 5, 7, 21, ptr(out), 5, 7, 21, stream()))
    torch.cuda.synchronize()
    assert torch.equal(out, img)
