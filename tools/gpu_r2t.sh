# round 2, call t: Wolf-order rows of 40 / 24 / 20 / 8 bytes as super-rows inside tc_forward_kernel (WOLF = 3)
set -x
O=gpurun_out/r2t; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "wolf or superrow" > $O/pytest_wolf.log 2>&1; tail -15 $O/pytest_wolf.log
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['gpu_launches'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for sh in 38x480 24x480 19x480 8x480; do
  echo "== $sh bct"; bench1 --shape $sh
  echo "== $sh btc"; bench1 --shape $sh --layout btc
done > $O/bench.txt 2>&1
cat $O/bench.txt
