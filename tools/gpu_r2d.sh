# round 2, call d (second attempt: the first one hung in probe_i8_peak N = 64, a phase-aliasing bug of the probe's commit
# barrier, fixed): role profile of the 22x1125 kernel, int8 tcgen05 peak, the new bench blocks (sustained, config1), memcheck
set -x
O=gpurun_out/r2d; mkdir -p $O
for n in 256 128 64 32 16; do timeout 60 ./tools/probe_i8_peak 0.5 $n > $O/i8_peak_n$n.json 2>&1; cat $O/i8_peak_n$n.json; done
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 180 python tools/prof_tcr.py 22 1125 bct > $O/prof_bct.txt 2>&1; cat $O/prof_bct.txt
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 180 python tools/prof_tcr.py 22 1125 btc > $O/prof_btc.txt 2>&1; cat $O/prof_btc.txt
timeout 300 python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -c 1500 $O/bench_default.json
timeout 300 python bench.py --shape 22x1125 --no-cpu-baseline > $O/bench_22.json 2> $O/bench_22.err; tail -c 900 $O/bench_22.json
timeout 300 python bench.py --workload config1 > $O/bench_config1.json 2> $O/bench_config1.err; cat $O/bench_config1.json; tail -3 $O/bench_config1.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cat $O/bench_ref.json
timeout 420 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_small.py > $O/sanitizer_memcheck.log 2>&1; tail -8 $O/sanitizer_memcheck.log
