set -x
O=gpurun_out/r2h9; mkdir -p $O
b() { timeout 200 python bench.py --workload eegnet --no-cpu-baseline "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,3), d['roofline']['kernel'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in base sl100 sl400 base; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v 22x1125"; b
  echo "== $v 64x480"; b --shape 064x480
done > $O/ab.txt 2>&1
grep -E "^== |^[0-9]" $O/ab.txt | paste - -
