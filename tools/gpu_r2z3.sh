set -x
O=gpurun_out/r2z3; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "beyond_one_launch" > $O/pytest.log 2>&1; tail -5 $O/pytest.log
