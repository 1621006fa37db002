"""GPU bring-up check of the tcgen05 GEMM (run on the B200 box): each group runs in its own process
under a timeout so a deadlocked kernel cannot hang the call.  Usage: python tools/gemm_check.py [group]"""
import ctypes, os, subprocess, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EPI = dict(BIAS_BF16=0, GELU=1, RESID=2, ROWADD=3, LN_PAD=4, LN_NCHW=5, RESID_LN=6, BIAS_F32=7)


def load():
    lib = ctypes.CDLL(os.environ.get("INKSAM_LIB", os.path.join(ROOT, "inklayer_b200", "libinksam.so")))
    lib.inksam_last_error.restype = ctypes.c_char_p
    return lib


def ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def gemm(lib, cg, tf32, bn, epi, A, W, out, bias=None, resid=None, rowadd=None, rowadd_mod=1, ln_w=None, ln_b=None,
         eps=1e-6, M=None):
    import torch
    M = A.shape[0] if M is None else M
    N, K = W.shape
    rc = lib.inksam_op_gemm(cg, tf32, bn, epi, ptr(A), A.stride(0), ptr(W), W.stride(0), M, N, K, ptr(out),
                            out.stride(0) if out.dim() == 2 else N, ptr(bias), ptr(resid), ptr(rowadd), rowadd_mod,
                            ptr(ln_w), ptr(ln_b), ctypes.c_float(eps), ptr(None), 0, 0, 0, 0,
                            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError(lib.inksam_last_error().decode())


def run_group(group):
    import torch
    torch.backends.cuda.matmul.allow_tf32 = False
    lib = load()
    dev = "cuda"
    g = torch.Generator(device="cpu").manual_seed(0)
    cg = 2 if "cg2" in group else 1
    tf32 = 1 if "tf32" in group else 0
    dt = torch.float32 if tf32 else torch.bfloat16

    def rnd(*s, scale=1.0):
        return (torch.randn(*s, generator=g) * scale).to(dev)

    def check(name, got, ref, tol):
        err = (got.float() - ref).abs().max().item()
        rel = err / (ref.abs().max().item() + 1e-9)
        ok = rel < tol
        print(f"[{group}] {name}: max_abs_err={err:.4e} rel={rel:.3e} {'OK' if ok else 'FAIL'}", flush=True)
        return ok

    ok = True
    if "perf" in group:
        shapes = [(32768, 3840, 1280, 0), (32768, 1280, 1280, 2), (32768, 5120, 1280, 1), (32768, 1280, 5120, 2),
                  (4096, 5120, 1280, 1)]
        if "proj" in group:  # which part of the attn.proj epilogue costs: bf16 store / fp32 store / fp32 store + residual
            shapes = [(32768, 1280, 1280, 0), (32768, 1280, 1280, 7), (32768, 1280, 1280, 2)]
        for (M, N, K, epi) in shapes:
            A = rnd(M, K).to(dt); W = rnd(N, K, scale=0.03).to(dt); bias = rnd(N)
            out = torch.empty(M, N, device=dev, dtype=torch.float32 if epi in (2, 7) else torch.bfloat16)
            resid = torch.zeros(M, N, device=dev) if epi == 2 else None
            for _ in range(3):
                gemm(lib, cg, tf32, 256, epi, A, W, out, bias=bias, resid=resid)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                gemm(lib, cg, tf32, 256, epi, A, W, out, bias=bias, resid=resid)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            print(f"[{group}] M={M} N={N} K={K} epi={epi}: {ms:.3f} ms  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
            t0 = time.time()
            for _ in range(10):
                torch.matmul(A, W.t())
            torch.cuda.synchronize()
            e0.record()
            for _ in range(10):
                torch.matmul(A, W.t())
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            print(f"[{group}]    torch.matmul: {ms:.3f} ms  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
        return True

    shapes = [(256, 256, 64), (256, 256, 256), (4096, 3840, 1280), (112, 256, 256), (4096 + 77, 1280, 5120)]
    for (M, N, K) in shapes:
        A = rnd(M, K).to(dt); W = rnd(N, K, scale=0.05).to(dt); bias = rnd(N)
        ref = A.float() @ W.float().t() + bias
        tol = 2e-3 if tf32 else 1e-2
        if not tf32:
            out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
            gemm(lib, cg, tf32, 256, EPI["BIAS_BF16"], A, W, out, bias=bias)
            torch.cuda.synchronize()
            ok &= check(f"bias_bf16 {M}x{N}x{K}", out, ref, tol)
            out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
            gemm(lib, cg, tf32, 256, EPI["GELU"], A, W, out, bias=bias)
            torch.cuda.synchronize()
            ok &= check(f"gelu {M}x{N}x{K}", out, torch.nn.functional.gelu(ref), tol)
            resid = rnd(M, N)
            out = resid.clone()
            gemm(lib, cg, tf32, 256, EPI["RESID"], A, W, out, bias=bias, resid=out)
            torch.cuda.synchronize()
            ok &= check(f"resid(inplace) {M}x{N}x{K}", out, ref + resid, tol)
        out = torch.zeros(M, N, device=dev)
        gemm(lib, cg, tf32, 256, EPI["BIAS_F32"], A, W, out, bias=bias)
        torch.cuda.synchronize()
        ok &= check(f"bias_f32 {M}x{N}x{K}", out, ref, tol)
        radd = rnd(64, N)
        out = torch.zeros(M, N, device=dev)
        gemm(lib, cg, tf32, 256, EPI["ROWADD"], A, W, out, bias=bias, rowadd=radd, rowadd_mod=64)
        torch.cuda.synchronize()
        ok &= check(f"rowadd {M}x{N}x{K}", out, ref + radd[torch.arange(M, device=dev) % 64], tol)
        if N == 256 and tf32:
            resid = rnd(M, N); lw = rnd(N) * 0.1 + 1; lb = rnd(N) * 0.1
            out = torch.zeros(M, N, device=dev)
            gemm(lib, cg, tf32, 256, EPI["RESID_LN"], A, W, out, bias=bias, resid=resid, ln_w=lw, ln_b=lb, eps=1e-5)
            torch.cuda.synchronize()
            ok &= check(f"resid_ln {M}x{N}x{K}", out, torch.nn.functional.layer_norm(ref + resid, (N,), lw, lb, 1e-5), tol)
    return ok


if __name__ == "__main__":
    if len(sys.argv) > 1:
        sys.exit(0 if run_group(sys.argv[1]) else 1)
    rcs = {}
    for grp in ["bf16_cg1", "bf16_cg2", "tf32_cg1", "tf32_cg2", "perf_bf16_cg1", "perf_bf16_cg2"]:
        try:
            r = subprocess.run([sys.executable, __file__, grp], timeout=180)
            rcs[grp] = r.returncode
        except subprocess.TimeoutExpired:
            rcs[grp] = "TIMEOUT"
        print("GROUP", grp, "->", rcs[grp], flush=True)
    print(rcs)
