"""Experiment driver (GPU box): per-role mbarrier wait shares of the pipeline kernel.
Build first with  MIBMI_EXTRA_NVCC_FLAGS=-DMIBMI_PROF python mibminet_b200/_build.py --force"""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth, _lib
from mibminet_b200.engine import Int8Net
C, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (64, 480)
B = 65536
PATH = os.environ.get('MIBMI_PATH', 'auto')
net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
net.set_path(PATH)
g = torch.Generator(device='cuda'); g.manual_seed(5)
x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda', generator=g)
for _ in range(3):
    y = net.forward(x)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    net.forward(x)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
w = torch.arange(1, 4 * B + 1, device='cuda', dtype=torch.int64)
print(f'{os.environ.get("MIBMI_LIB_PATH", "default")}: {ms:.4f} ms  {B / ms / 1e3:.1f} M trials/s  checksum {int((y.flatten().long() * w).sum())}')
buf = np.zeros(148 * 32 * 4, np.int64)
L = _lib.lib()
if not hasattr(L, 'mibmi_debug_prof'):
    sys.exit(0)
L.mibmi_debug_prof.argtypes = [ctypes.c_void_p]
assert L.mibmi_debug_prof(buf.ctypes.data) == 0
p = buf.reshape(148, 32, 4).astype(np.float64)
ntr = B / 148
roles = {'A (L1)': range(0, 8), 'B (L2/3)': range(8, 24), 'C (L4/5)': range(24, 26), 'P (TMA)': range(26, 27)}
names = {'A (L1)': ['full_x', 'empty_y1'], 'B (L2/3)': ['full_y1', 'empty_y3'], 'C (L4/5)': ['full_y3'], 'P (TMA)': ['empty_x']}
print(f'{C}x{T}: cycles per trial per CTA = {p[:, 0, 0].mean() / ntr:.0f}')
for r, ws in roles.items():
    q = p[:, list(ws), :]
    tot = q[..., 0].mean()
    s = '  '.join(f'wait {n} {100 * q[..., 1 + i].mean() / tot:.1f}%' for i, n in enumerate(names[r]))
    busy = tot - q[..., 1:].sum(-1).mean()
    print(f'{r:9s} busy {busy / ntr:7.0f} cyc/trial ({100 * busy / tot:.1f}%)  {s}')
    if r.startswith('A'):
        print('   per A warp busy cyc/trial:', ' '.join(f'{(q[:, i, 0] - q[:, i, 1:].sum(-1)).mean() / ntr:.0f}' for i in range(8)))
