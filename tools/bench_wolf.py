"""GPU box: throughput of the Wolf-order [B][T][C_ALIGN] input vs module order (tcgen05 K-major operand for 64 / 16 byte rows,
L2-resident reorder scratch for the others; MIBMI_WOLF_CHUNK sweeps the chunk)."""
import os, sys, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
shapes = [(64, 480), (16, 480), (22, 1125), (38, 480), (24, 480), (8, 480)]
chunks = [None] if len(sys.argv) < 2 else [int(v) for v in sys.argv[1].split(',')]
for C, T in shapes:
    B = 65536 if T == 480 else 32768
    x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda')
    xw = torch.zeros((B, T, (C + 3) // 4 * 4), dtype=torch.int8, device='cuda')
    xw[:, :, :C] = x.transpose(1, 2)
    for chunk in (chunks if C not in (64, 16) else [None]):
        if chunk:
            os.environ['MIBMI_WOLF_CHUNK'] = str(chunk)
        else:
            os.environ.pop('MIBMI_WOLF_CHUNK', None)
        net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
        ref = net.forward(x)
        assert torch.equal(net.forward(xw, layout='btc'), ref)
        for name, inp, lay in (('bct', x, 'bct'), ('btc', xw, 'btc')):
            for _ in range(3):
                net.forward(inp, layout=lay)
            n0 = net.launch_count
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                net.forward(inp, layout=lay)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            print(f'{C}x{T} {name} chunk={chunk}: {ms:.4f} ms  {B / ms / 1e3:.1f} M trials/s  {net.last_kernel} '
                  f'launches/call={(net.launch_count - n0) // 10}', flush=True)
        net.close()
    del x, xw
