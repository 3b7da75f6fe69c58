for v in base nofold base nofold; do
  if [ $v = base ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo $v; python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['roofline']['frac'])"
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --shape 22x1125 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['roofline']['frac'])"
done
