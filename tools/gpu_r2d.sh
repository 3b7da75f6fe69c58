# round 2, call d: role profile of the 22x1125 kernel, int8 tcgen05 peak, the new bench blocks (sustained, config1), memcheck
set -x
O=gpurun_out/r2d; mkdir -p $O
./tools/probe_i8_peak 1.0 256 > $O/i8_peak.json 2>&1; cat $O/i8_peak.json
./tools/probe_i8_peak 0.5 64 > $O/i8_peak_n64.json 2>&1; cat $O/i8_peak_n64.json
./tools/probe_i8_peak 0.5 32 > $O/i8_peak_n32.json 2>&1; cat $O/i8_peak_n32.json
./tools/probe_i8_peak 0.5 16 > $O/i8_peak_n16.json 2>&1; cat $O/i8_peak_n16.json
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so python tools/prof_tcr.py 22 1125 bct > $O/prof_bct.txt 2>&1; cat $O/prof_bct.txt
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so python tools/prof_tcr.py 22 1125 btc > $O/prof_btc.txt 2>&1; cat $O/prof_btc.txt
python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -c 1500 $O/bench_default.json
python bench.py --shape 22x1125 --no-cpu-baseline > $O/bench_22.json 2> $O/bench_22.err; tail -c 900 $O/bench_22.json
python bench.py --workload config1 > $O/bench_config1.json 2> $O/bench_config1.err; cat $O/bench_config1.json; tail -3 $O/bench_config1.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cat $O/bench_ref.json
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_small.py > $O/sanitizer_memcheck.log 2>&1; tail -8 $O/sanitizer_memcheck.log
