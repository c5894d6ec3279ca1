"""normalize([0,0,qz,qw]) on torch-CPU / torch-CUDA vs the four kernel norm modes (GPU box)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from legged_gym_b200 import _lgk
from oracle.legged_oracle import normalize

lib = _lgk.load()
n = 20_000_000
g = torch.Generator().manual_seed(3)
q = torch.zeros(n, 4)
q[:, 2:] = torch.randn(n, 2, generator=g)
qc = q.cuda()
out = torch.empty(n, 2, device="cuda")
for H in (1, 187):
    # the reference normalises base_quat.repeat(1, H).view(-1, 4): check the reduction shape does not matter
    m = 200_000 if H > 1 else n
    want_cpu = normalize(q[:m].repeat(1, H).view(-1, 4))[:, 2:].view(m, H, 2)[:, 0]
    want_cuda = normalize(qc[:m].repeat(1, H).view(-1, 4))[:, 2:].view(m, H, 2)[:, 0].cpu()
    for mode in range(4):
        _lgk.check("probe", lib.lgk_debug_yaw_quat(_lgk.ptr(qc), m, mode, _lgk.ptr(out), _lgk.current_stream()))
        torch.cuda.synchronize()
        got = out[:m].cpu()
        print(f"H={H:3d} mode {mode}: vs torch-cpu {int((got != want_cpu).any(1).sum())} / {m}   vs torch-cuda {int((got != want_cuda).any(1).sum())} / {m}")
