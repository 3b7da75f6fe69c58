# round 2: 22x1125 role balance A/B: layer-1 epilogue warps 12 -> 8, realign warps 4 -> 6 / 8
set -x
O=gpurun_out/r2aa; mkdir -p $O
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in base a8r8 a8r4 a8r6 base; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v bct"; bench1 --shape 22x1125
  echo "== $v btc"; bench1 --shape 22x1125 --layout btc
done > $O/bench.txt 2>&1
grep -E "^== |^[0-9]" $O/bench.txt | paste - -
export MIBMI_LIB_PATH=tools/variants/a8r8/libmibmi.so
timeout 600 python -m pytest tests -m gpu -x -q -k "22x1125 and not beyond" > $O/pytest_a8r8.log 2>&1; tail -2 $O/pytest_a8r8.log
