# allocation-free entry point: its test alone in a fresh process, then after all other GPU tests
set -x
O=gpurun_out/r2na; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/pytest_alone.log 2>&1; tail -15 $O/pytest_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_all.log 2>&1; tail -4 $O/pytest_all.log
