# round 2, call u: what bounds tc_forward_kernel when HBM does not (16x480): role profile + ncu; 64x480 role profile for comparison
set -x
O=gpurun_out/r2u; mkdir -p $O
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 180 python tools/prof_tc.py 16 480 > $O/prof_tc_16.txt 2>&1; cat $O/prof_tc_16.txt
MIBMI_LIB_PATH=tools/variants/prof/libmibmi.so timeout 180 python tools/prof_tc.py 64 480 > $O/prof_tc_64.txt 2>&1; cat $O/prof_tc_64.txt
timeout 300 python bench.py --shape 16x480 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 1 -s 4 -o $O/tc16_full -f python bench.py --shape 16x480 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/ncu_tc16.log 2>&1
tail -2 $O/ncu_tc16.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/plain64.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 1 -s 4 -o $O/tc64_full -f python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/ncu_tc64.log 2>&1
tail -2 $O/ncu_tc64.log
