set -x
mkdir -p gpurun_out/r1e
python bench.py --shape 22x1125 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1e/bench_22.json 2> gpurun_out/r1e/bench.err; cat gpurun_out/r1e/bench_22.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1e/launches.csv python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1e/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tcr_forward -c 1 -s 4 -o gpurun_out/r1e/tcr_full -f python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1e/ncu2.log 2>&1
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so python tools/prof_tcr.py 22 1125 | tee gpurun_out/r1e/prof.txt
ls -la gpurun_out/r1e
