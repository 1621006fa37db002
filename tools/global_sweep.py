"""Times the global attention kernel alone over INKSAM_GLOBAL_KPOLY settings (read per launch), ViT-H and ViT-B shapes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
src = open(os.path.join(ROOT, "tools", "attn_bench.py")).read().split("for (B, heads, hd) in")[0]
exec(src)
shapes = [(8, 16, 80), (1, 16, 80), (8, 12, 64)]
for kp in sys.argv[1:] or ("0", "8", "16"):
    os.environ["INKSAM_GLOBAL_KPOLY"] = kp
    line = []
    for (B, heads, hd) in shapes:
        ms, tf, _ = run(0, B, heads, hd, n=10)
        line.append(f"B={B} hd={hd}: {ms*1e3:7.1f} us")
    print(f"global kpoly {kp:>2}  " + "   ".join(line), flush=True)
