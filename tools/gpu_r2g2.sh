# round 2: tcr kernel with the first layer-1 epilogue set as a third layer-2 consumer (nohelp = previous form)
set -x
O=gpurun_out/r2g2; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log
b() { timeout 200 python bench.py --shape 22x1125 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in nohelp help nohelp help; do
  if [ $v = help ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v bct"; b
  echo "== $v btc"; b --layout btc
done > $O/ab.txt 2>&1
grep -E "^== |^[0-9]" $O/ab.txt | paste - -
export MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so
timeout 200 python tools/prof_tcr.py 22 1125 bct > $O/role_profile_bct.txt 2>&1; cat $O/role_profile_bct.txt
