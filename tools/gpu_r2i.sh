set -x
O=gpurun_out/r2i; mkdir -p $O
timeout 300 python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tcr_forward -c 1 -s 4 -o $O/tcr_full -f python bench.py --shape 22x1125 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --sustained-seconds 0 > $O/ncu_tcr.log 2>&1
tail -3 $O/ncu_tcr.log
