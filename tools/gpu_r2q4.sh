# round 2 (4 GPUs): the N = 4 point of the scaling table, gather to root
set -x
O=gpurun_out/r2q4; mkdir -p $O
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 20 --warmup 5 "$@"; }
run > $O/n4_default.json 2> $O/n4_default.err
run --shape 22x1125 --no-e2e --no-cpu-baseline --sustained-seconds 0 2> $O/n4_22.err | tail -1 > $O/n4_22.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2q4/n4_*.json')):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f.split('/')[-1], round(d['value']/1e6,1), 'M/s', [round(r['kernel_ms_median'],4) for r in d.get('per_rank',[])], d['run']['gather_checked'], (d.get('e2e') or {}).get('value'), (d.get('e2e') or {}).get('h2d_gbs_per_rank'), (d.get('sustained') or {}).get('value'))
PY
