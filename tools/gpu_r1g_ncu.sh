set -x
mkdir -p gpurun_out/r1g
ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 1 -s 4 -o gpurun_out/r1g/tc_full -f python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1g/ncu_tc.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tcr_forward -c 1 -s 4 -o gpurun_out/r1g/tcr_full -f python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1g/ncu_tcr.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:eegnet_mma -c 1 -s 2 -o gpurun_out/r1g/eegnet_full -f python bench.py --workload eegnet --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1g/ncu_eegnet.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g/tcr_launches.csv python bench.py --shape 22x1125 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r1g/ncu_tcr1.log 2>&1
ls -la gpurun_out/r1g | grep ncu-rep
