# round 2, call y: layer-1 offset fold (MIBMI_TC_FOLD) where the epilogue warps bind (C < 64)
set -x
O=gpurun_out/r2y; mkdir -p $O
bench1() { timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 0 "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['roofline']['frac'],4), d['roofline']['kernel'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for v in base tcfold base tcfold; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  for sh in 16x480 38x480 8x480 64x480; do echo "== $v $sh"; bench1 --shape $sh; done
done > $O/bench.txt 2>&1
grep -E "^== |^[0-9]" $O/bench.txt | paste - -
