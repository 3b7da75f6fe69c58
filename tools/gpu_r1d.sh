set -x
mkdir -p gpurun_out/r1d
python -m pytest tests -m gpu -x -q > gpurun_out/r1d/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1d/pytest.log
tail -3 gpurun_out/r1d/pytest.log
python bench.py > gpurun_out/r1d/bench.json 2> gpurun_out/r1d/bench.err; echo "bench exit $?"
cat gpurun_out/r1d/bench.json
python bench.py --shape 22x1125 --no-cpu-baseline --no-e2e > gpurun_out/r1d/bench_22.json 2>> gpurun_out/r1d/bench.err; cat gpurun_out/r1d/bench_22.json
python bench.py --workload stream --no-cpu-baseline > gpurun_out/r1d/bench_stream.json 2>> gpurun_out/r1d/bench.err; cat gpurun_out/r1d/bench_stream.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1d/bench_ref.json 2>> gpurun_out/r1d/bench.err; cat gpurun_out/r1d/bench_ref.json
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1d/b_pre.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1d/launches.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1d/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 1 -s 4 -o gpurun_out/r1d/tc_full -f python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1d/ncu2.log 2>&1
ls -la gpurun_out/r1d
