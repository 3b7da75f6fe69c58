# round 2, call v: tc_forward_kernel with 12 layer-1 epilogue warps (27 warps, 72 registers) vs 8
set -x
O=gpurun_out/r2v; mkdir -p $O
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 2 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['sustained']; print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['clocks']['sm_mhz'], d['clocks']['reasons'], '| sustained', round(s['value']/1e6,1), round(s['roofline_frac'],4), s['clocks']['sm_mhz'])"; }
for v in base aw12 base aw12; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v 64x480"; bench1
  echo "== $v 16x480"; bench1 --shape 16x480
  echo "== $v 38x480 btc"; bench1 --shape 38x480 --layout btc
done > $O/bench.txt 2>&1
cat $O/bench.txt
export MIBMI_LIB_PATH=tools/variants/aw12/libmibmi.so
timeout 600 python -m pytest tests -m gpu -x -q -k "not 22 and not tcr and not stream and not eegnet" > $O/pytest_aw12.log 2>&1; tail -3 $O/pytest_aw12.log
