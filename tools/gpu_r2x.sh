# round 2, call x: plain (REORDER_BN off) variant on the tcgen05 kernels: tests + throughput
set -x
O=gpurun_out/r2x; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "plain or fixtures or torch_op" > $O/pytest_plain.log 2>&1; tail -8 $O/pytest_plain.log
timeout 300 python - > $O/plain_speed.txt 2>&1 <<'PY'
import sys, torch, numpy as np
sys.path.insert(0, '.')
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
for C, T in ((64, 480), (22, 1125)):
    arrays = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T).to_wolf_arrays()
    B = 65536
    x = torch.randint(-127, 128, (B, C, T), dtype=torch.int8, device='cuda')
    for rb in (True, False):
        net = Int8Net(arrays, C=C, T=T, device='cuda:0', reorder_bn=rb)
        for _ in range(3): net.forward(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): net.forward(x)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f'{C}x{T} reorder_bn={rb}: {net.last_kernel} {ms:.4f} ms  {B / ms / 1e3:.1f} M trials/s', flush=True)
        if not rb:
            net.set_path('generic'); xs = x[:4096]
            for _ in range(2): net.forward(xs)
            e0.record(); net.forward(xs); e1.record(); torch.cuda.synchronize()
            print(f'   generic kernel: {4096 / e0.elapsed_time(e1) / 1e3:.2f} M trials/s', flush=True)
PY
cat $O/plain_speed.txt
