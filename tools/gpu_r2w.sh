# round 2, call w: 22x1125 layer-1 epilogue in half passes of 8 filters (two y1 commits per group)
set -x
O=gpurun_out/r2w; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q -k "22 or tcr or wolf or unaligned or gather or taps" > $O/pytest_tcr.log 2>&1; tail -3 $O/pytest_tcr.log
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['gpu_launches'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for i in 1 2; do
  echo "== 22x1125 bct"; bench1 --shape 22x1125
  echo "== 22x1125 btc"; bench1 --shape 22x1125 --layout btc
done > $O/bench.txt 2>&1
cat $O/bench.txt
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 180 python tools/prof_tcr.py 22 1125 bct > $O/prof_bct.txt 2>&1; cat $O/prof_bct.txt
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 180 python tools/prof_tcr.py 22 1125 btc > $O/prof_btc.txt 2>&1; cat $O/prof_btc.txt
