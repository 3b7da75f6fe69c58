# round 2, call b: the N = 32 / Wolf-order tcr kernel: tests, bench, role profile
set -x
O=gpurun_out/r2b; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; tail -15 $O/pytest.log
bench1() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['gpu_launches'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for i in 1 2; do
  echo "== 22x1125 bct"; bench1 --shape 22x1125
  echo "== 22x1125 btc"; bench1 --shape 22x1125 --layout btc
done > $O/bench.txt 2>&1
cat $O/bench.txt
export MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so
python tools/prof_tcr.py 22 1125 bct > $O/prof_bct.txt 2>&1; cat $O/prof_bct.txt
python tools/prof_tcr.py 22 1125 btc > $O/prof_btc.txt 2>&1; cat $O/prof_btc.txt
