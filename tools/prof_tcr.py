"""Experiment driver (GPU box): per-role mbarrier wait shares of tcr_forward_kernel (22 x 1125) (build with -DMIBMI_PROF)."""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth, _lib
from mibminet_b200.engine import Int8Net
C, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (22, 1125)
layout = sys.argv[3] if len(sys.argv) > 3 else 'bct'
B = 65536
net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
g = torch.Generator(device='cuda'); g.manual_seed(5)
x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda', generator=g)
if layout == 'btc':
    xw = torch.zeros((B, T, (C + 3) // 4 * 4), dtype=torch.int8, device='cuda'); xw[:, :, :C] = x.transpose(1, 2); x = xw
for _ in range(3):
    y = net.forward(x, layout=layout)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    net.forward(x, layout=layout)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
w = torch.arange(1, 4 * B + 1, device='cuda', dtype=torch.int64)
print(f'{os.environ.get("MIBMI_LIB_PATH", "default")} {layout}: {ms:.4f} ms  {B / ms / 1e3:.1f} M trials/s  checksum {int((y.flatten().long() * w).sum())}')
L = _lib.lib()
if not hasattr(L, 'mibmi_debug_prof_tcr'):
    sys.exit(0)
buf = np.zeros(148 * 32 * 4, np.int64)
L.mibmi_debug_prof_tcr.argtypes = [ctypes.c_void_p]
assert L.mibmi_debug_prof_tcr(buf.ctypes.data) == 0
p = buf.reshape(148, 32, 4).astype(np.float64)
ntr = B / 148
roles = {'A  L1 epilogue': (range(0, 12), ['t1_full', 'tile_done']), 'B  L2 epilogue': (range(12, 20), ['t2_full', 'y2_empty']),
         'C  L3/4/5': (range(20, 24), ['y2_full']), 'R  realign': (range(24, 28), ['raw_full', 'x_empty']),
         'P  bulk copy': (range(28, 29), ['raw_empty']),
         'M1 L1 mma': (range(29, 30), ['x_full', 't1_empty']), 'M2 L2 mma': (range(30, 31), ['y1_full', 't2_empty'])}
print(f'{C}x{T}: cycles per trial per CTA = {p[:, 0, 0].mean() / ntr:.0f}')
for r, (ws, names) in roles.items():
    q = p[:, list(ws), :]
    tot = q[..., 0].mean()
    s = '  '.join(f'wait {n} {100 * q[..., 1 + i].mean() / tot:.1f}%' for i, n in enumerate(names))
    busy = tot - q[..., 1:].sum(-1).mean()
    print(f'{r:15s} busy {busy / ntr:7.0f} cyc/trial ({100 * busy / tot:.1f}%)  {s}')
