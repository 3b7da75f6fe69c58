# round 2, call q (8 GPUs): gather-mode A/B (all-gather / gather-to-root / none) with per-rank kernel times, default line with e2e
set -x
O=gpurun_out/r2q; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
numactl -H > $O/numa.txt 2>&1; lscpu | head -30 >> $O/numa.txt
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 "$@"; }
for g in peer root none; do run --gather $g --no-e2e --no-cpu-baseline --sustained-seconds 0 2> $O/n8_$g.err | tail -1 > $O/n8_$g.json; done
run --gather root > $O/n8_root_full.json 2> $O/n8_root_full.err
run --shape 22x1125 --gather root --no-e2e --no-cpu-baseline --sustained-seconds 0 2> $O/n8_22_root.err | tail -1 > $O/n8_22_root.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2q/n8_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d['value']/1e6,1), 'M/s', [round(r['kernel_ms_median'],4) for r in d.get('per_rank',[])], d['run']['gather_checked'], (d.get('e2e') or {}).get('value'), (d.get('e2e') or {}).get('h2d_gbs_per_rank'), (d.get('e2e') or {}).get('cpu_affinity'), (d.get('sustained') or {}).get('value'))
    except Exception as e: print(f, 'ERR', e)
PY
