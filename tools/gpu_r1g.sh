set -x
mkdir -p gpurun_out/r1g
python -m pytest tests -m gpu -x -q > gpurun_out/r1g/pytest.log 2>&1; tail -2 gpurun_out/r1g/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1g/smoke.log 2>&1; tail -1 gpurun_out/r1g/smoke.log
python bench.py > gpurun_out/r1g/bench.json 2> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench.json
python bench.py --shape 22x1125 --no-cpu-baseline > gpurun_out/r1g/bench_22.json 2>> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench_22.json
python bench.py --shape 22x1125 --layout btc --batch 56832 > gpurun_out/r1g/bench_22_btc.json 2>> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench_22_btc.json
python bench.py --layout btc --no-cpu-baseline > gpurun_out/r1g/bench_64_btc.json 2>> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench_64_btc.json
python bench.py --workload stream --no-cpu-baseline > gpurun_out/r1g/bench_stream.json 2>> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench_stream.json
python bench.py --workload eegnet > gpurun_out/r1g/bench_eegnet.json 2>> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench_eegnet.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1g/bench_ref.json 2>> gpurun_out/r1g/bench.err; cat gpurun_out/r1g/bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r1g/ncu_bench.log 2>&1
tail -3 gpurun_out/r1g/bench.err
ls gpurun_out/r1g
