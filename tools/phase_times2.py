"""Development: per-phase cycle counts of the round-2 step kernel (lanes = envs) from the -DLGK_PROFILE build
(build/liblgk_prof.so).  Stamps (clock64, first sub-tile of each CTA, per logical warp):
  0 kernel top   1 loads landed   2 stage 1 done   3 past the CTA barrier   4 (W0..2) heads stored
  6 (W3) term values stored
usage: python tools/phase_times2.py [envs]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["LGK_LIB_PATH"] = os.path.join(ROOT, "build", "liblgk_prof.so")
import torch, helpers
from legged_gym_b200 import _lgk
from legged_gym_b200.utils import synthetic
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
def extra(cfg):
    cfg.terrain.materialize_heights = False
env = helpers.make_env("go1_rough_full", N, device="cuda:0", seed=1, extra=extra)
fr = synthetic.make_frames(env, 2, seed=3)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
for i in range(6):
    synthetic.load_frame(env, fr[i % 2])
    flush.zero_()
    env.step(fr[i % 2]["actions"])
torch.cuda.synchronize()
lib = _lgk.load()
n = 4096 * 4 * 16
buf = (C.c_longlong * n)()
lib.lgk_debug_read_prof.argtypes = [C.c_void_p, C.c_int]
assert lib.lgk_debug_read_prof(buf, n) == 0
ncta = min(4096, (N + 31) // 32)
t = torch.tensor(list(buf), dtype=torch.float64).view(4096, 4, 16)[:ncta]
names = {1: "load", 2: "stage1", 3: "barrier", 4: "head stores", 6: "term table"}
for w in range(4):
    seq = [0, 1, 2, 3, 4] if w < 3 else [0, 1, 2, 3, 6]
    parts = [f"{names[b]} {float((t[:, w, b] - t[:, w, a]).mean()):8.0f}" for a, b in zip(seq[:-1], seq[1:])]
    print(f"warp {w}: " + "  ".join(parts) + f"   total {float((t[:, w, seq[-1]] - t[:, w, 0]).mean()):8.0f}")
m = 4096 * 4
buf2 = (C.c_longlong * m)()
lib.lgk_debug_read_prof_cta.argtypes = [C.c_void_p, C.c_int]
assert lib.lgk_debug_read_prof_cta(buf2, m) == 0
g = torch.tensor(list(buf2), dtype=torch.float64).view(4096, 4)[:ncta]
g0 = g[:, 0].min()
st, en, en3 = g[:, 0] - g0, g[:, 1] - g0, g[:, 3] - g0
qs = torch.tensor([0.1, 0.25, 0.5, 0.75, 0.9, 1.0], dtype=torch.float64)
print("CTA start quantiles us", [round(float(x) / 1e3, 1) for x in torch.quantile(st, qs)])
print("W0 end quantiles us   ", [round(float(x) / 1e3, 1) for x in torch.quantile(en, qs)])
print("W3 end quantiles us   ", [round(float(x) / 1e3, 1) for x in torch.quantile(en3, qs)])
print(f"mean CTA life (start -> W3 end) {float((en3 - st).mean()) / 1e3:.1f} us; W0 life {float((en - st).mean()) / 1e3:.1f} us")
life = en3 - st
order = torch.argsort(st)
k = max(ncta // 4, 1)
print("CTA life by start order (us): " + "  ".join(f"{float(life[order[i * k:(i + 1) * k]].mean()) / 1e3:.1f}" for i in range(4)))
