# round 2, call c: state check of HEAD on a fresh box: GPU tests, bench lines for configs 2/3 (both layouts)
set -x
O=gpurun_out/r2c; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; tail -15 $O/pytest.log
bench1() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['gpu_launches'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for i in 1 2; do
  echo "== 64x480 bct"; bench1
  echo "== 22x1125 bct"; bench1 --shape 22x1125
  echo "== 22x1125 btc"; bench1 --shape 22x1125 --layout btc
  echo "== 16x480 bct"; bench1 --shape 16x480
  echo "== 38x480 bct"; bench1 --shape 38x480
done > $O/bench.txt 2>&1
cat $O/bench.txt
python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -1 $O/bench_default.json
