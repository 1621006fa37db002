"""Helpers for the -m gpu tests: typed access to libinksam's op-level C ABI with torch tensors as device memory."""
import ctypes

import torch

from inklayer_b200 import _lib

EPI = dict(BIAS_BF16=0, GELU=1, RESID=2, ROWADD=3, LN_PAD=4, LN_NCHW=5, RESID_LN=6, BIAS_F32=7, UPSCALE1=8, UPSCALE2=9)


def ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def gemm(epi, A, W, out, M=None, N=None, K=None, lda=None, ldo=None, cg=1, tf32=0, bn=256, bias=None, resid=None,
         rowadd=None, rowadd_mod=1, ln_w=None, ln_b=None, eps=1e-6, hyper=None, conv3=0, conv_c=0, win_remap=0, resid_mod=0):
    lib = _lib.load()
    M = A.shape[0] if M is None else M
    N = W.shape[0] if N is None else N
    K = W.shape[1] if K is None else K
    lda = (A.stride(0) if A.dim() == 2 else K) if lda is None else lda
    ldo = N if ldo is None else ldo
    _lib.check(lib.inksam_op_gemm(cg, tf32, bn, epi, ptr(A), lda, ptr(W), W.stride(0), M, N, K, ptr(out), ldo, ptr(bias),
                                  ptr(resid), ptr(rowadd), rowadd_mod, ptr(ln_w), ptr(ln_b), ctypes.c_float(eps),
                                  ptr(hyper), conv3, conv_c, win_remap, resid_mod, stream()))


def rel_err(got, ref):
    got, ref = got.float(), ref.float()
    return ((got - ref).abs().max() / (ref.abs().max() + 1e-12)).item()


def rel_rms(got, ref):
    got, ref = got.double(), ref.double()
    return ((got - ref).pow(2).mean().sqrt() / (ref.pow(2).mean().sqrt() + 1e-30)).item()
