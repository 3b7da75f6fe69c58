# round 2: streaming kernel A/B (stbase = previous build)
set -x
O=gpurun_out/r2st; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q -k "stream" > $O/pytest_stream.log 2>&1; tail -3 $O/pytest_stream.log
for v in stbase new stbase new; do
  if [ $v = new ]; then unset MIBMI_LIB_PATH; else export MIBMI_LIB_PATH=tools/variants/$v/libmibmi.so; fi
  echo "== $v"; timeout 300 python bench.py --workload stream --no-cpu-baseline | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']/1e6,1), 'M windows/s', d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done > $O/bench.txt 2>&1
grep -E "^== |^[0-9]" $O/bench.txt | paste - -
