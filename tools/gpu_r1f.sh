set -x
mkdir -p gpurun_out/r1f
python -m pytest tests -m gpu -x -q > gpurun_out/r1f/pytest.log 2>&1; tail -2 gpurun_out/r1f/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1f/smoke.log 2>&1; tail -1 gpurun_out/r1f/smoke.log
python bench.py > gpurun_out/r1f/bench.json 2> gpurun_out/r1f/bench.err; cat gpurun_out/r1f/bench.json
python bench.py --shape 22x1125 --no-cpu-baseline > gpurun_out/r1f/bench_22.json 2>> gpurun_out/r1f/bench.err; cat gpurun_out/r1f/bench_22.json
python bench.py --workload stream --no-cpu-baseline > gpurun_out/r1f/bench_stream.json 2>> gpurun_out/r1f/bench.err; cat gpurun_out/r1f/bench_stream.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1f/bench_ref.json 2>> gpurun_out/r1f/bench.err; cat gpurun_out/r1f/bench_ref.json
ncu --set full --clock-control none --import-source on -k regex:stream_forward -c 1 -s 4 -o gpurun_out/r1f/stream_full -f python bench.py --workload stream --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1f/ncu_stream.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f/stream_launches.csv python bench.py --workload stream --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1f/ncu_stream1.log 2>&1
ls gpurun_out/r1f
