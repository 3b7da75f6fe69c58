# round 2, closing run of this session: all GPU tests, smoke, every bench line, ncu launch list of the default command
set -x
O=gpurun_out/r2z3; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 300 python bench.py > $O/bench_default.json 2> $O/bench_default.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>/dev/null
timeout 300 python bench.py --shape 22x1125 --no-cpu-baseline > $O/bench_22.json 2>/dev/null
timeout 300 python bench.py --shape 22x1125 --layout btc --no-cpu-baseline --no-e2e > $O/bench_22_btc.json 2>/dev/null
timeout 300 python bench.py --shape 38x480 --layout btc --no-cpu-baseline --no-e2e > $O/bench_38_btc.json 2>/dev/null
timeout 300 python bench.py --shape 16x480 --no-cpu-baseline --no-e2e > $O/bench_16.json 2>/dev/null
timeout 300 python bench.py --workload stream --no-cpu-baseline > $O/bench_stream.json 2>/dev/null
timeout 300 python bench.py --workload eegnet --no-cpu-baseline > $O/bench_eegnet.json 2>/dev/null
timeout 300 python bench.py --workload eegnet --shape 064x480 --no-cpu-baseline > $O/bench_eegnet_64x480.json 2>/dev/null
timeout 300 python bench.py --workload config1 > $O/bench_config1.json 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2z3/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d.get('roofline') or {}
        print(f.split('/')[-1], d.get('value'), d.get('unit'), 'frac', r.get('frac'), 'kernel', r.get('kernel'), 'sustained', (d.get('sustained') or {}).get('value'), 'e2e', (d.get('e2e') or {}).get('value'))
    except Exception as e: print(f,'ERR',e)
PY
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --sustained-seconds 0 > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --sustained-seconds 0 > $O/ncu_launches.log 2>&1
tail -2 $O/bench_launches.csv | cut -c1-160
