"""`ncu -i X.ncu-rep --page raw --csv` -> one compact block per kernel for profiles/ (time, grid, registers, executed
instructions, issue-active share, DRAM bytes, the stall reasons above 0.3 warps per issue)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
for line in sys.argv[2:]:
    print("# " + line)
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__inst_executed_pipe_tensor.sum"]
for r in rows[2:]:
    print("== " + r[hdr.index("Kernel Name")][:110])
    for k in KEYS:
        if k in hdr:
            print("   %-55s %s %s" % (k, r[hdr.index(k)], rows[1][hdr.index(k)]))
    st = [(h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), float(v.replace(",", "") or 0))
          for h, v in zip(hdr, r) if "average_warps_issue_stalled" in h and "not_issued" not in h]
    print("   stalls (warps per issue): " + ", ".join("%s %.2f" % s for s in sorted(st, key=lambda x: -x[1]) if s[1] > 0.3))
