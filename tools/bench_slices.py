"""GPU box: is the 64 x 480 rate a function of the batch size or of where the trials live?  65 536-trial slices of a 2^20-trial
array one by one, then growing batches from the start of the array."""
import sys, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
C, T = 64, 480
net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
xall = torch.randint(-127, 128, (1 << 20, C, T), dtype=torch.int8, device='cuda')
out = torch.empty((1 << 20, 4), dtype=torch.int8, device='cuda')
def t(x, reps=10):
    for _ in range(2):
        net.forward(x, out=out[:x.shape[0]])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        net.forward(x, out=out[:x.shape[0]])
    e1.record(); torch.cuda.synchronize()
    return x.shape[0] / (e0.elapsed_time(e1) / reps) / 1e3
print('slices of 65536:', ' '.join(f'{t(xall[k << 16:(k + 1) << 16]):.0f}' for k in range(16)))
for B in (65536, 98304, 131072, 196608, 262144, 524288, 1048576):
    print(f'B={B}: {t(xall[:B], 5):.1f} M trials/s', flush=True)
