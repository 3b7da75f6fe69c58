# round 2, call a: tests on the pruned build, N = 32 layer-2 probe, mbarrier wait-mode A/B, memcheck
set -x
O=gpurun_out/r2a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; tail -3 $O/pytest.log
./tools/probe_umma3 > $O/probe_umma3.txt 2>&1; cat $O/probe_umma3.txt
bench1() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in base w1 w2 base w1 w2; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v 64x480"; bench1
  echo "== $v 22x1125"; bench1 --shape 22x1125
  echo "== $v 16x480"; bench1 --shape 16x480
done > $O/ab_wait.txt 2>&1
cat $O/ab_wait.txt
for v in base w2; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v sustained 64x480"; python bench.py --steps 6000 --warmup 5 --no-cpu-baseline --no-e2e | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['clocks']; print(round(d['value']/1e6,1), 'M trials/s', c['sm_mhz'], 'MHz', c['reasons'], c['power_w_max'], 'W')"
done > $O/ab_sustained.txt 2>&1
cat $O/ab_sustained.txt
unset MIBMI_LIB_PATH
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_small.py > $O/sanitizer_memcheck.log 2>&1; tail -8 $O/sanitizer_memcheck.log
