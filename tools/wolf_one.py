"""GPU box: one Wolf-order forward (for an ncu launch list).  usage: wolf_one.py C T B [chunk]"""
import os, sys, torch
sys.path.insert(0, '/root/repo')
C, T, B = (int(v) for v in sys.argv[1:4])
if len(sys.argv) > 4:
    os.environ['MIBMI_WOLF_CHUNK'] = sys.argv[4]
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
xw = torch.randint(-127, 128, (B, T, (C + 3) // 4 * 4), dtype=torch.int8, device='cuda')
for _ in range(2):
    y = net.forward(xw, layout='btc')
torch.cuda.synchronize()
print(net.last_kernel, y.shape)
