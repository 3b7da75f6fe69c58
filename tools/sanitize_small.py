"""GPU box, under compute-sanitizer (one tool per run): one small forward of every kernel family against the oracle.
PYTORCH_NO_CUDA_MEMCACHING=1 makes every tensor its own cudaMalloc block, so that a read past the end of a batch is an
out-of-bounds access for memcheck (the caching allocator would hide it inside a larger block)."""
import os, sys
os.environ.setdefault("PYTORCH_NO_CUDA_MEMCACHING", "1")
import numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import wolf_port, eegnet_np
from mibminet_b200 import convert, synth
from mibminet_b200.engine import Int8Net, EEGNetInt8
ok = True
def report(*a):
    global ok
    ok = ok and bool(a[-1])
    print(*a, flush=True)
# B chosen so that the batch ends at an odd byte for the shapes whose trials are not multiples of 16 bytes
for C, T, B in ((64, 480, 301), (22, 1125, 449), (16, 480, 77), (38, 480, 150), (19, 480, 75)):
    p = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=1), C=C, T=T)
    net = Int8Net(p, device='cuda:0')
    x = synth.make_trials(B, C, T, seed=2)
    ref = wolf_port.forward(p.to_wolf_arrays(), C, T, x)
    xd = torch.from_numpy(x).cuda()
    report(C, T, 'bct', net.forward(xd).cpu().numpy().shape, net.last_kernel, np.array_equal(net.forward(xd).cpu().numpy(), ref))
    y = net.forward(torch.from_numpy(convert.to_wolf_input(x)).cuda(), layout='btc').cpu().numpy()
    report(C, T, 'btc', net.last_kernel, np.array_equal(y, ref))
    bufs = [torch.zeros((B + 8, 4), dtype=torch.int8, device='cuda') for _ in range(3)]
    net.forward_gather(xd, [b.data_ptr() for b in bufs], 4)
    report(C, T, 'gather', all(np.array_equal(b.cpu().numpy()[4:4 + B], ref) for b in bufs))
    if (C, T) in ((22, 1125), (64, 480)):
        net.set_path('fused_pipe')
        report(C, T, 'pipe', np.array_equal(net.forward(xd).cpu().numpy(), ref), net.last_kernel)
        net.set_path('fused')
        yt, taps = net.forward_taps(xd[:40])
        report(C, T, 'fused taps', np.array_equal(yt.cpu().numpy(), ref[:40]))
        net.set_path('generic')
        report(C, T, 'generic', np.array_equal(net.forward(xd[:40]).cpu().numpy(), ref[:40]))
        net.set_path('auto')
        report(C, T, 'host', np.array_equal(net.forward_host(x[:200]).numpy(), ref[:200]))
C, T = 64, 480
p = convert.params_from_state_dict(synth.make_state_dict(C=C, T=T, seed=1), C=C, T=T)
net = Int8Net(p, device='cuda:0')
xs = synth.make_trials(3, C, 480 + 16 * 40, seed=4)
report('stream', np.array_equal(net.forward_stream(torch.from_numpy(xs).cuda()).cpu().numpy(), wolf_port.forward_stream(p.to_wolf_arrays(), C, T, xs)))
xf = (np.random.default_rng(0).standard_normal((5, C, T)) * 0.4 * p.input_scale).astype(np.float32)
q = net.quantize_input(torch.from_numpy(xf).cuda())
report('quantize', np.array_equal(q.cpu().numpy(), convert.quantize_input(xf, p.input_scale)))
C, T = 22, 1125
sd = synth.make_eegnet_state_dict(C=C, T=T, seed=1)
pe = convert.eegnet_params_from_state_dict(sd, C=C, T=T)
x = synth.make_trials(20, C, T, seed=3)
report('eegnet order', np.array_equal(EEGNetInt8(pe, device='cuda:0').forward(torch.from_numpy(x).cuda()).cpu().numpy(), eegnet_np.forward(pe, x)))
torch.cuda.synchronize()
print('ALL OK' if ok else 'MISMATCH')
sys.exit(0 if ok else 1)
