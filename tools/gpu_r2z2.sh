# round 2, final check after the plain-variant change: all GPU tests, smoke, default bench line
set -x
O=gpurun_out/r2z2; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -4 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 300 python bench.py > $O/bench_default.json 2> $O/bench_default.err; python -c "
import json; d=json.loads(open('$O/bench_default.json').read().strip().splitlines()[-1]); print(d['value'], d['roofline']['frac'], d['sustained']['value'], d['e2e']['value'], d['cpu_baseline'])"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>/dev/null; cat $O/bench_ref.json | cut -c1-900
