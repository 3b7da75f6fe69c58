# round 2, call r: 64x480 burst + sustained A/B of the mbarrier wait mode / poll interval / layer-1 offset fold
set -x
O=gpurun_out/r2r; mkdir -p $O
for v in base tcw1 tcw2 tcfold tcw1s256 base tcw1; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v"; timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --sustained-seconds 2 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['sustained']; c=s['clocks']
print('burst', round(d['value']/1e6,1), round(d['roofline']['frac'],4), '| sustained', round(s['value']/1e6,1), round(s['roofline_frac'],4), c['sm_mhz'], 'MHz', c['power_w_max'], 'W', c['reasons'])"
done > $O/ab_tc_wait.txt 2>&1
cat $O/ab_tc_wait.txt
