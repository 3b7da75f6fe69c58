# sustained (power-capped) rate of the 64 x 480 kernel: variants named on the command line against the default build
for v in base "$@" base "$@"; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  python bench.py --steps 6000 --warmup 5 --no-cpu-baseline --no-e2e | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); c=d['clocks']; print('$v', round(d['value']/1e6,1), 'M trials/s', c['sm_mhz'], 'MHz', c['reasons'], c['power_w_max'], 'W')"
done
