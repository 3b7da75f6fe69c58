# round 2: layer 3 without the shift inside the tcgen05 kernels' FAST instantiation
set -x
O=gpurun_out/r2n; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log
for a in "" "--shape 22x1125" "--shape 22x1125 --layout btc" "--shape 16x480" "--shape 38x480"; do timeout 300 python bench.py $a --no-cpu-baseline --no-e2e | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$a', round(d['value']/1e6,2), round(d['roofline']['frac'],4), 'sustained', round(d['sustained']['value']/1e6,2), d['sustained']['clocks']['sm_mhz'])"; done
