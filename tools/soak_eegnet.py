"""Soak run (GPU box): eegnet_tc_kernel against eegnet_mma_kernel (and the oracle on a few trials) over random shapes and nets."""
import os, sys, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import eegnet_np
from mibminet_b200 import convert, synth
from mibminet_b200.engine import EEGNetInt8
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
bad = 0
for it in range(n):
    C = int(rng.integers(1, 65)); T = int(rng.integers(64, 1500)); D = int(rng.integers(1, 3)); B = int(rng.integers(1, 700))
    sd = synth.make_eegnet_state_dict(C=C, T=T, F1=8, D=D, Nf=128, seed=1000 + it, negative_gamma=bool(it & 1))
    p = convert.eegnet_params_from_state_dict(sd, C=C, T=T, F1=8, D=D)
    net = EEGNetInt8(p, device='cuda:0')
    x = synth.make_trials(B, C, T, seed=it, dist='gauss' if it % 3 else 'uniform')
    xd = torch.from_numpy(x).cuda()
    os.environ.pop('MIBMI_EEGNET_TC', None)
    y, taps = net.forward_taps(xd)
    k1 = net.last_kernel
    os.environ['MIBMI_EEGNET_TC'] = '0'
    yr, tr = net.forward_taps(xd)
    k2 = net.last_kernel
    ok = torch.equal(y, yr) and all(torch.equal(taps[k], tr[k]) for k in taps)
    nb = min(B, 6)
    ok_o = np.array_equal(y[:nb].cpu().numpy(), eegnet_np.forward(p, x[:nb]))
    print(f'{it:3d} C={C:2d} T={T:4d} D={D} B={B:3d} {k1} vs {k2}: {"ok" if ok else "MISMATCH"} oracle {"ok" if ok_o else "MISMATCH"}', flush=True)
    bad += (not ok) or (not ok_o)
print('mismatches:', bad)
sys.exit(1 if bad else 0)
