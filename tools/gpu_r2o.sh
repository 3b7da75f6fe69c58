# round 2: 22x1125 kernel with setmaxnreg (32 warps; realign warpgroup 88 registers, producer / MMA warpgroup 40)
set -x
O=gpurun_out/r2o2; mkdir -p $O
timeout 400 python -m pytest tests -m gpu -x -q -k "1125 or 22 or gather or graph" > $O/pytest_22.log 2>&1; tail -3 $O/pytest_22.log
b() { timeout 200 python bench.py --shape 22x1125 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
b; b --layout btc; b; b --layout btc
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 200 python tools/prof_tcr.py 22 1125 bct 2>&1 | tail -9
