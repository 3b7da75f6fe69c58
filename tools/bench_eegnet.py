"""GPU box: throughput of the EEGNet-order kernel (mibmi_eegnet_forward_int8), CUDA events."""
import sys, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth
from mibminet_b200.engine import EEGNetInt8
for C, T in ((22, 1125), (64, 480)):
    sd = synth.make_eegnet_state_dict(C=C, T=T, seed=0)
    net = EEGNetInt8(convert.eegnet_params_from_state_dict(sd, C=C, T=T), device='cuda:0')
    B = 8192
    x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda')
    for _ in range(2):
        net.forward(x)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        net.forward(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    macs = 8 * C * T * 128 + 16 * C * T
    print(f'EEGNet order {C}x{T}: {ms:.2f} ms per {B} trials = {B / ms / 1e3:.3f} M trials/s, {B * macs / ms / 1e9:.1f} TMAC/s')
