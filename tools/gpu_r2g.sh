# round 2, call g: realign-warp variants of the double-buffered kernel + ncu of the base build
set -x
O=gpurun_out/r2g; mkdir -p $O
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['gpu_launches'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in base rb6 rb6w5 rb2; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v 22x1125 bct"; bench1 --shape 22x1125
  echo "== $v 22x1125 btc"; bench1 --shape 22x1125 --layout btc
done > $O/bench.txt 2>&1
cat $O/bench.txt
unset MIBMI_LIB_PATH
timeout 300 python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tcr_forward -c 1 -s 4 -o $O/tcr_full -f python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/ncu_tcr.log 2>&1
tail -3 $O/ncu_tcr.log
