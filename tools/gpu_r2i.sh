# round 2: evidence after the tcgen05 EEGNet-order kernel: all GPU tests, smoke, default + eegnet bench lines, ncu --set full of eegnet_tc_kernel
set -x
O=gpurun_out/r2i; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; tail -4 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 300 python bench.py > $O/bench_default.json 2> $O/bench_default.err; cut -c1-200 $O/bench_default.json
timeout 300 python bench.py --workload eegnet --no-cpu-baseline > $O/bench_eegnet.json 2>/dev/null; cut -c1-160 $O/bench_eegnet.json
timeout 300 python bench.py --workload eegnet --shape 064x480 --no-cpu-baseline > $O/bench_eegnet_64x480.json 2>/dev/null; cut -c1-160 $O/bench_eegnet_64x480.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:eegnet_tc_kernel -s 3 -c 1 -f -o $O/eegnet_tc_full python bench.py --workload eegnet --batch 4096 --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_eegnet.log 2>&1
ls -la $O
