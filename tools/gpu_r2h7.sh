set -x
O=gpurun_out/r2h7; mkdir -p $O
timeout 300 python -m pytest tests/test_eegnet_order.py -m gpu -x -q > $O/pytest_eegnet.log 2>&1; tail -3 $O/pytest_eegnet.log
b() { timeout 200 python bench.py --workload eegnet --no-cpu-baseline "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,3), d['roofline']['kernel'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
b; b --shape 064x480; b; b --shape 064x480
