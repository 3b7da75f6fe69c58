"""GPU box, under compute-sanitizer: one small forward of every kernel family against the oracle."""
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import wolf_port, eegnet_np
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net, EEGNetInt8
for C, T, B in ((64, 480, 601), (22, 1125, 449), (16, 480, 77)):
    p = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=1), C=C, T=T)
    net = Int8Net(p, device='cuda:0')
    x = synth.make_trials(B, C, T, seed=2)
    ref = wolf_port.forward(p.to_wolf_arrays(), C, T, x)
    y = net.forward(torch.from_numpy(x).cuda()).cpu().numpy()
    print(C, T, net.last_kernel, np.array_equal(y, ref))
    if C in (64, 16):
        y = net.forward(torch.from_numpy(convert.to_wolf_input(x)).cuda(), layout='btc').cpu().numpy()
        print(C, T, 'btc', net.last_kernel, np.array_equal(y, ref))
    bufs = [torch.zeros((B + 8, 4), dtype=torch.int8, device='cuda') for _ in range(3)]
    net.forward_gather(torch.from_numpy(x).cuda(), [b.data_ptr() for b in bufs], 4)
    print(C, T, 'gather', all(np.array_equal(b.cpu().numpy()[4:4 + B], ref) for b in bufs))
C, T = 22, 1125
sd = synth.make_eegnet_state_dict(C=C, T=T, seed=1)
pe = convert.eegnet_params_from_state_dict(sd, C=C, T=T)
x = synth.make_trials(40, C, T, seed=3)
print('eegnet order', np.array_equal(EEGNetInt8(pe, device='cuda:0').forward(torch.from_numpy(x).cuda()).cpu().numpy(), eegnet_np.forward(pe, x)))
