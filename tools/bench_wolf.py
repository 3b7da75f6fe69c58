"""GPU box: throughput of the Wolf-order [B][T][C_ALIGN] input through the tcgen05 kernel vs module order."""
import sys, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
for C in (64, 16):
    T, B = 480, 65536
    net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
    x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda')
    xw = x.transpose(1, 2).contiguous()
    ref = net.forward(x)
    assert torch.equal(net.forward(xw, layout='btc'), ref)
    for name, inp, lay in (('bct', x, 'bct'), ('btc', xw, 'btc')):
        for _ in range(3):
            net.forward(inp, layout=lay)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            net.forward(inp, layout=lay)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f'{C}x{T} {name}: {ms:.4f} ms  {B / ms / 1e3:.1f} M trials/s  {net.last_kernel}')
