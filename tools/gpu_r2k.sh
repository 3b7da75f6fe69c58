# round 2: layer 3 in the no-shift magic form (all kernels get m3 with sh3 = 0; the streaming kernel a compile-time FAST3 instantiation)
set -x
O=gpurun_out/r2k; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log
timeout 300 python bench.py --workload stream --no-cpu-baseline > $O/bench_stream.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/bench_stream.json').read().strip().splitlines()[-1]); print('stream', round(d['value']/1e6,1), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
for a in "" "--shape 22x1125" "--shape 22x1125 --layout btc" "--shape 16x480"; do timeout 300 python bench.py $a --no-cpu-baseline --no-e2e --sustained-seconds 0 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$a', round(d['value']/1e6,2), round(d['roofline']['frac'],4))"; done
