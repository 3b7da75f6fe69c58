"""Metric extract of an .ncu-rep for profiles/: `ncu -i rep --page raw --csv` filtered to the metrics the roofline lines quote.
usage: ncu_extract.py rep.ncu-rep out.csv"""
import csv, re, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
KEEP = re.compile(r"^(Kernel Name|dram__bytes|dram__cycles_active|gpu__dram_throughput|gpu__time_duration|sm__cycles_elapsed\.avg|"
                  r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared|l1tex__data_pipe_(lsu|tc)_wavefronts_mem_shared|"
                  r"launch__(registers_per_thread|grid_size|block_size|shared_mem_per_block_dynamic|waves_per_multiprocessor)|"
                  r"lts__t_sector_hit_rate|sm__inst_executed\.sum|smsp__inst_executed\.sum|smsp__issue_active|sm__warps_active|"
                  r"sm__pipe_tensor_cycles_active|sm__inst_executed_pipe_(alu|fma|lsu|tmem|tc|uniform|adu|cbu)\.|"
                  r"sm__icc_request_hit_rate|gcc__cache_requests_type_instruction|smsp__average_warps_issue_stalled_.*_per_issue_active|"
                  r"sm__throughput|l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld\.sum)")
with open(out, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["metric", "unit", "value"])
    for h, u, v in zip(hdr, units, vals):
        if KEEP.match(h) and ".max" not in h and ".min" not in h:
            w.writerow([h, u, v[:140]])
print(out)
