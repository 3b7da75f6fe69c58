# round 2, call s: layer-1 sign correction on the FMA pipe (IMAD.HI.U32 x, 2, x) vs LEA.HI (ALU pipe), 22x1125
set -x
O=gpurun_out/r2s; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q -k "22 or tcr or wolf or unaligned or gather or taps" > $O/pytest_tcr.log 2>&1; tail -3 $O/pytest_tcr.log
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['gpu_launches'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in base nosignfma base nosignfma; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v 22x1125 bct"; bench1 --shape 22x1125
  echo "== $v 22x1125 btc"; bench1 --shape 22x1125 --layout btc
done > $O/bench.txt 2>&1
cat $O/bench.txt
