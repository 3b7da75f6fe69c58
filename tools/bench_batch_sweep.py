"""GPU box: config 3 of SURVEY 8(d) -- 22 x 1125, B = 1 .. 2^20 in powers of 4 (and 64 x 480 beside it): CUDA-event time per call."""
import json, sys, torch
sys.path.insert(0, '/root/repo')
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net
out = {}
for C, T in ((22, 1125), (64, 480)):
    net = Int8Net(convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=0), C=C, T=T), device='cuda:0')
    xall = torch.randint(-127, 128, (1 << 20, C, T), dtype=torch.int8, device='cuda')
    ref = net.forward(xall[:4096])
    for p in range(0, 21, 2):
        B = 1 << p
        x = xall[:B]
        reps = max(5, min(200, (1 << 22) // B))
        for _ in range(3):
            y = net.forward(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            y = net.forward(x)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / reps * 1e3
        assert torch.equal(y[:min(B, 4096)], ref[:min(B, 4096)])
        out[f'{C}x{T}:{B}'] = {'us_per_call': us, 'trials_per_s': B / us * 1e6, 'kernel': net.last_kernel}
        print(f'{C}x{T} B={B:8d}: {us:10.1f} us/call  {B / us:9.3f} M trials/s  {net.last_kernel}', flush=True)
    del xall
import os; os.makedirs('gpurun_out/r2bs', exist_ok=True); json.dump(out, open('gpurun_out/r2bs/batch_sweep.json', 'w'), indent=1)
