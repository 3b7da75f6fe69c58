# round 2: tcr kernel, helper consumer set (default) with 4 / 5 realign warps
set -x
O=gpurun_out/r2g3; mkdir -p $O
b() { timeout 200 python bench.py --shape 22x1125 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in help rw5 help rw5; do
  if [ $v = help ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v bct"; b
  echo "== $v btc"; b --layout btc
done > $O/ab.txt 2>&1
grep -E "^== |^[0-9]" $O/ab.txt | paste - -
export MIBMI_LIB_PATH=tools/variants/rw5/libmibmi.so
timeout 600 python -m pytest tests -m gpu -x -q -k "1125 or 22" > $O/pytest_rw5.log 2>&1; tail -3 $O/pytest_rw5.log
export MIBMI_LIB_PATH=tools/variants/rw5prof/libmibmi.so
timeout 200 python tools/prof_tcr.py 22 1125 bct > $O/role_profile_rw5_bct.txt 2>&1; cat $O/role_profile_rw5_bct.txt
export MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so
timeout 200 python tools/prof_tcr.py 22 1125 btc > $O/role_profile_btc.txt 2>&1; cat $O/role_profile_btc.txt
