# round 2: ncu --set full of the final eegnet_tc_kernel (after the register-constant / counter changes)
set -x
O=gpurun_out/r2z5; mkdir -p $O
timeout 200 python bench.py --workload eegnet --batch 4096 --steps 1 --warmup 3 --no-cpu-baseline > $O/plain.json 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:eegnet_tc_kernel -s 3 -c 1 -f -o $O/eegnet_tc_final python bench.py --workload eegnet --batch 4096 --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu.log 2>&1
ls -la $O | tail -3
