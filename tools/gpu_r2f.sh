# round 2, closing run: all GPU tests (incl. the CUDA-graph capture tests and the 19x480 streaming instantiation), smoke, the bench lines,
# the ncu launch list of the default command, one ncu --set full capture of the streaming kernel
set -x
O=gpurun_out/r2f; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -4 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 300 python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -c 300 $O/bench_default.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>/dev/null
timeout 300 python bench.py --workload config1 > $O/bench_config1.json 2>/dev/null; cat $O/bench_config1.json | cut -c1-1500
timeout 300 python bench.py --shape 22x1125 --no-cpu-baseline --no-e2e > $O/bench_22.json 2>/dev/null
timeout 300 python bench.py --workload stream --no-cpu-baseline > $O/bench_stream.json 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2f/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d.get('roofline') or {}
        print(f.split('/')[-1], d.get('value'), d.get('unit'), 'frac', r.get('frac'), 'kernel', r.get('kernel'), 'sustained', (d.get('sustained') or {}).get('value'), 'e2e', (d.get('e2e') or {}).get('value'))
    except Exception as e: print(f,'ERR',e)
PY
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --sustained-seconds 0 > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --sustained-seconds 0 > $O/ncu_launches.log 2>&1
tail -3 $O/bench_launches.csv | cut -c1-200
timeout 300 python bench.py --workload stream --subjects 1024 --steps 1 --warmup 3 --no-cpu-baseline > $O/stream_small.json 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_forward_kernel -s 3 -c 1 -f -o $O/stream_full python bench.py --workload stream --subjects 1024 --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_stream.log 2>&1
ls -la $O
