set -x
O=gpurun_out/r2h3; mkdir -p $O
timeout 300 python -m pytest tests/test_eegnet_order.py -m gpu -x -q > $O/pytest_eegnet.log 2>&1; tail -15 $O/pytest_eegnet.log
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 200 python tools/prof_et.py 22 1125 2>&1 | tail -7
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 200 python tools/prof_et.py 64 480 2>&1 | tail -7
timeout 200 python bench.py --workload eegnet --no-cpu-baseline > $O/bench_eegnet.json 2>/dev/null; cut -c1-120 $O/bench_eegnet.json
timeout 200 python bench.py --workload eegnet --shape 64x480x --no-cpu-baseline > /dev/null 2>&1
