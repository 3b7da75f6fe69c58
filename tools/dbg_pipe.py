"""GPU debug driver: compare pipe kernel taps/logits with the oracle per layer."""
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import golden_np
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
C, T = int(sys.argv[1]), int(sys.argv[2])
B = int(sys.argv[3]) if len(sys.argv) > 3 else 300
p = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=1), C=C, T=T)
net = Int8Net(p, device='cuda:0')
print('fused_supported', net.fused_supported())
x = synth.make_trials(B, C, T, seed=3)
xg = torch.from_numpy(x).cuda()
net.set_path('generic')
yg, tg = net.forward_taps(xg)
net.set_path('fused')
yf, tf = net.forward_taps(xg)
torch.cuda.synchronize()
for k in ('y1', 'y2', 'y3', 'y4'):
    a, b = tg[k].cpu().numpy(), tf[k].cpu().numpy()
    bad = np.argwhere(a != b)
    print(k, 'mismatch', len(bad), 'of', a.size, bad[:6].tolist())
    if len(bad):
        i = tuple(bad[0]); print('  ref', a[i], 'got', b[i])
print('logits(taps) equal', bool((yg == yf).all().item()))
y2 = net.forward(xg); torch.cuda.synchronize()
print('logits(fast) equal', bool((yg == y2).all().item()), net.last_path)
