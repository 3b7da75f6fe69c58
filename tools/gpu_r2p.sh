# round 2, call p (2 GPUs): PeerGather rendezvous test on two ranks, gather-mode A/B
set -x
O=gpurun_out/r2p; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 600 python -m pytest tests/test_peer_gather_gpu.py -x -q -s > $O/pytest_peer_gather.log 2>&1; tail -6 $O/pytest_peer_gather.log
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 "$@"; }
for g in peer root none nccl; do run --gather $g --no-e2e --no-cpu-baseline --sustained-seconds 0 2> $O/n2_$g.err | tail -1 > $O/n2_$g.json; done
run > $O/n2_default.json 2> $O/n2_default.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2p/n2_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d['value']/1e6,1), 'M/s', [round(r['kernel_ms_median'],4) for r in d.get('per_rank',[])], d['run']['gather_checked'], (d.get('e2e') or {}).get('value'), (d.get('e2e') or {}).get('h2d_gbs_per_rank'), (d.get('e2e') or {}).get('cpu_affinity'))
    except Exception as e: print(f, 'ERR', e)
PY
