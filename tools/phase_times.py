"""Development: per-phase cycle counts of the step kernel (needs the -DLGK_PROFILE build in build/liblgk_prof.so)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["LGK_LIB_PATH"] = os.path.join(ROOT, "build", os.environ.get("LGK_PROF_LIB", "liblgk_prof.so"))
import torch, helpers
from legged_gym_b200 import _lgk
from legged_gym_b200.utils import synthetic
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
per = int(sys.argv[2]) if len(sys.argv) > 2 else 0
def extra(cfg):
    cfg.terrain.materialize_heights = False
    cfg.env.envs_per_cta = per
env = helpers.make_env("go1_rough_full", N, device="cuda:0", seed=1, extra=extra)
fr = synthetic.make_frames(env, 2, seed=3)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0") if os.environ.get("LGK_FLUSH_L2") else None
for i in range(6):
    synthetic.load_frame(env, fr[i % 2])
    if flush is not None: flush.zero_()   # cold L2, like bench.py's kernel timing
    env.step(fr[i % 2]["actions"])
torch.cuda.synchronize()
lib = _lgk.load()
n = 4096 * 4 * 16
buf = (C.c_longlong * n)()
lib.lgk_debug_read_prof.argtypes = [C.c_void_p, C.c_int]
assert lib.lgk_debug_read_prof(buf, n) == 0
t = torch.tensor(list(buf), dtype=torch.float64).view(4096, 4, 16)
ncta = min(4096, (N + per - 1) // per) if per > 0 else int(sys.argv[3])   # per = 0: the library's own grid (first 4096 CTAs)
t = t[:ncta]
t0 = t[:, 0, 0].min()
names = ["load", "stage1", "barrier1", "stage2", "s3_A+barrier", "s3_BC"]
for w in (0, 1, 2, 3):
    d = t[:, w, 1:7] - t[:, w, 0:6]
    print(f"warp {w}: " + "  ".join(f"{names[k]} {d[:, k].mean():8.0f}" for k in range(6)), " total", float((t[:, w, 6] - t[:, w, 0]).mean()))
for w in (0, 3):
    print(f"  warp {w} stage 3 after barrier: base-height {float((t[:, w, 8] - t[:, w, 5]).mean()):.0f}  pass B {float((t[:, w, 9] - t[:, w, 8]).mean()):.0f}  ticket/arrive {float((t[:, w, 7] - t[:, w, 9]).mean()):.0f}  pass C {float((t[:, w, 6] - t[:, w, 7]).mean()):.0f}")
print(f"  warp 3 stage 1: reset decision {float((t[:, 3, 12] - t[:, 3, 1]).mean()):.0f}  penalised contacts {float((t[:, 3, 10] - t[:, 3, 12]).mean()):.0f}  feet {float((t[:, 3, 11] - t[:, 3, 10]).mean()):.0f}  rest {float((t[:, 3, 2] - t[:, 3, 11]).mean()):.0f};  warp 0 reset decision {float((t[:, 0, 12] - t[:, 0, 1]).mean()):.0f}")
d7 = t[:, 0, 7] - t[:, 0, 5]
print("stage3 pass B (+ticket wait for W3)", float(d7.mean()), " pass C", float((t[:, 0, 6] - t[:, 0, 7]).mean()))
print("kernel span cycles", float(t[:, 0, 6].max() - t0), " first-wave CTAs start spread", float(t[:, 0, 0].max() - t0))

# CTA timeline (globaltimer, ns): waves per SM, tail
m = 4096 * 4
buf2 = (C.c_longlong * m)()
lib.lgk_debug_read_prof_cta.argtypes = [C.c_void_p, C.c_int]
assert lib.lgk_debug_read_prof_cta(buf2, m) == 0
g = torch.tensor(list(buf2), dtype=torch.float64).view(4096, 4)[:ncta]
g0 = g[:, 0].min()
st, en, smid = g[:, 0] - g0, g[:, 1] - g0, g[:, 2].long()
print(f"timeline (us): last CTA start {st.max() / 1e3:.1f}  last W0 end {en.max() / 1e3:.1f}  mean CTA life {(en - st).mean() / 1e3:.1f}")
print("  SMs used", int(smid.unique().numel()), " CTAs per SM min/max", int(torch.bincount(smid).clamp(min=0)[smid.unique()].min()), int(torch.bincount(smid).max()))
qs = torch.tensor([0.1, 0.25, 0.5, 0.75, 0.9, 1.0], dtype=torch.float64)
print("  start quantiles us", [round(float(x) / 1e3, 1) for x in torch.quantile(st, qs)])
print("  end   quantiles us", [round(float(x) / 1e3, 1) for x in torch.quantile(en, qs)])
en3 = g[:, 3] - g0
print(f"  last W3 end {en3.max() / 1e3:.1f} us (CTA {int(en3.argmax())}); W3 - W0 end: mean {(en3 - en).mean() / 1e3:.2f} max {(en3 - en).max() / 1e3:.2f} us; second-latest W3 end {en3.sort().values[-2] / 1e3:.1f}")
life = (en - st)
order = torch.argsort(st)
k = max(ncta // 4, 1)
print("  CTA life by start order (us): first quarter %.1f  second %.1f  third %.1f  last %.1f" % tuple(float(life[order[i * k:(i + 1) * k]].mean()) / 1e3 for i in range(4)))

if per == 0:
    nb = int(sys.argv[4])
    for name, sl in (("big", slice(0, nb)), ("small", slice(nb, ncta))):
        print(f"  {name}: n {life[sl].numel()} life mean {life[sl].mean() / 1e3:.1f} us  start min {st[sl].min() / 1e3:.1f} max {st[sl].max() / 1e3:.1f}  end max {en[sl].max() / 1e3:.1f}")
        for w in (0, 1):
            d = t[sl, w, 1:7] - t[sl, w, 0:6]
            print(f"    warp {w}: " + "  ".join(f"{names[k]} {d[:, k].mean():8.0f}" for k in range(6 if w == 0 else 4)))
print("  smid of CTAs 0..23:", smid[:24].tolist())
print("  smid of CTAs 148..159:", smid[148:160].tolist(), " 296..303:", smid[296:304].tolist())
same = {}
for b_ in range(min(ncta, 888)):
    same.setdefault(int(smid[b_]), []).append(b_)
print("  first-wave CTAs on SM of CTA 0:", same[int(smid[0])], " on SM of CTA 1:", same[int(smid[1])])

