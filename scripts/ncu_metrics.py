"""Print the interesting raw metrics of an .ncu-rep (one line per kernel instance)."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed_pipe_xu.sum"]
stall = [k for k in h if "warps_issue_stalled" in k and "per_issue_active" in k]
for r in rows[2:]:
    print("==", r[h.index("Kernel Name")][:60])
    for k in keys:
        m = [x for x in h if x.endswith(k) or x == k]
        for x in m:
            print("  %-70s %s %s" % (x[-70:], r[h.index(x)], rows[1][h.index(x)]))
    st = sorted(((float(r[h.index(k)] or 0), k) for k in stall), reverse=True)[:7]
    for v, k in st:
        print("  stall %-50s %.3f" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))
