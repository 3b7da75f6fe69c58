# round 2: two-GPU check after the kernel changes of this session: PeerGather test, scaling lines at N = 2 (64x480, 22x1125, eegnet)
set -x
O=gpurun_out/r2l2; mkdir -p $O
timeout 600 python -m pytest tests/test_peer_gather_gpu.py -m gpu -x -q > $O/pytest_peer.log 2>&1; tail -3 $O/pytest_peer.log
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@"; }
run --steps 20 --warmup 5 > $O/scale_n2_default.json 2>$O/scale_n2_default.err; python -c "
import json; d=json.loads(open('$O/scale_n2_default.json').read().strip().splitlines()[-1]); print('64x480 n2', round(d['value']/1e6,1), (d.get('e2e') or {}).get('value'), d['config'].get('gather_checked'))"
run --steps 20 --warmup 5 --shape 22x1125 --no-e2e --no-cpu-baseline > $O/scale_n2_22.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/scale_n2_22.json').read().strip().splitlines()[-1]); print('22x1125 n2', round(d['value']/1e6,1), d['config'].get('gather_checked'))"
run --steps 20 --warmup 5 --workload eegnet --no-cpu-baseline > $O/scale_n2_eegnet.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/scale_n2_eegnet.json').read().strip().splitlines()[-1]); print('eegnet n2', round(d['value']/1e6,2))"
