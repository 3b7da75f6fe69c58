set -x
O=gpurun_out/r2o3; mkdir -p $O
b() { timeout 200 python bench.py --shape 22x1125 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in nosetreg setreg nosetreg setreg; do
  if [ $v = setreg ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v bct"; b
  echo "== $v btc"; b --layout btc
done > $O/ab.txt 2>&1
grep -E "^== |^[0-9]" $O/ab.txt | paste - -
timeout 400 python -m pytest tests -m gpu -x -q -k "1125 or 22 or gather or graph" > $O/pytest_22.log 2>&1; tail -2 $O/pytest_22.log
