# round 2: first runs of eegnet_tc_kernel (tcgen05 form of the EEGNet-order layers 1 + 2)
set -x
O=gpurun_out/r2h1; mkdir -p $O
timeout 300 python -m pytest tests/test_eegnet_order.py -m gpu -x -q > $O/pytest_eegnet.log 2>&1; tail -30 $O/pytest_eegnet.log
timeout 200 python bench.py --workload eegnet --no-cpu-baseline > $O/bench_eegnet.json 2>$O/bench_eegnet.err; cut -c1-400 $O/bench_eegnet.json; tail -3 $O/bench_eegnet.err
MIBMI_EEGNET_TC=0 timeout 200 python bench.py --workload eegnet --no-cpu-baseline > $O/bench_eegnet_mma.json 2>/dev/null; cut -c1-200 $O/bench_eegnet_mma.json
nvidia-smi --query-gpu=name,clocks.sm --format=csv,noheader
