"""Experiment driver (GPU box): per-role mbarrier wait shares of eegnet_tc_kernel (build with -DMIBMI_PROF)."""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth, _lib
from mibminet_b200.engine import EEGNetInt8
C, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (22, 1125)
B = 16384
p = convert.eegnet_params_from_state_dict(synth.make_eegnet_state_dict(C=C, T=T, F1=8, D=2, Nf=128, seed=0), C=C, T=T, F1=8, D=2)
net = EEGNetInt8(p, device='cuda:0')
g = torch.Generator(device='cuda'); g.manual_seed(5)
x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda', generator=g)
for _ in range(3):
    y = net.forward(x)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    net.forward(x)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f'{os.environ.get("MIBMI_LIB_PATH", "default")} {net.last_kernel}: {ms:.4f} ms  {B / ms / 1e3:.2f} M trials/s')
L = _lib.lib()
if not hasattr(L, 'mibmi_debug_prof_et'):
    sys.exit(0)
buf = np.zeros(148 * 32 * 4, np.int64)
L.mibmi_debug_prof_et.argtypes = [ctypes.c_void_p]
assert L.mibmi_debug_prof_et(buf.ctypes.data) == 0
q = buf.reshape(148, 32, 4).astype(np.float64)
ntr = B / 148
roles = {'L  layers 3-5': (range(0, 4), ['y2_full', '-', '-']), 'E2 pool': (range(4, 8), ['t2_full', 'y2_empty', '-']),
         'E1 requant': (range(8, 20), ['t1_full', 'a2_empty', '-']), 'S  staging': (range(20, 23), ['xs_empty', '-', '-']),
         'M  mma': (range(23, 24), ['xs_full', 'a2_full', 't1/t2_empty'])}
print(f'{C}x{T}: cycles per trial per CTA = {q[:, 0, 0].mean() / ntr:.0f}')
for r, (ws, names) in roles.items():
    w = q[:, list(ws), :]
    tot = w[..., 0].mean()
    s = '  '.join(f'wait {n} {100 * w[..., 1 + i].mean() / tot:.1f}%' for i, n in enumerate(names) if n != '-')
    busy = tot - w[..., 1:].sum(-1).mean()
    print(f'{r:15s} busy {busy / ntr:7.0f} cyc/trial ({100 * busy / tot:.1f}%)  {s}')
