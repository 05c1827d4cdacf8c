"""Reads an .ncu-rep (here, no GPU needed): key raw metrics per kernel and, from the source page, the
instructions with the most stall samples.  usage: python tools/ncu_regions.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:100])
    for h, v in zip(hdr, r):
        if h in KEYS or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and float(v or 0) > 0.3):
            print("   %-88s %s" % (h, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
cur, blocks = None, {}
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "Kernel Name":
        cur = r[1]; blocks[cur] = []
    elif cur is not None and r:
        blocks[cur].append(r)
for name, rs in blocks.items():
    h, data = rs[0], rs[1:]
    ia, isrc, isamp, iex = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    base = int(data[0][ia], 16)
    tot = sum(int(r[isamp]) for r in data); totex = sum(int(r[iex]) for r in data)
    print("==", name[:100], "samples", tot, "warp-instr", totex)
    # regions delimited by barriers / mbarrier waits
    acc_s = acc_e = 0; start = 0
    for r in data:
        acc_s += int(r[isamp]); acc_e += int(r[iex])
        if any(k in r[isrc] for k in ("BAR.SYNC", "SYNCS.PHASECHK", "SYNCS.ARRIVE", "UBLKCP", "EXIT")):
            off = int(r[ia], 16) - base
            if acc_e:
                print("   [%5x..%5x] samples %5.1f%% inst %5.1f%%  ends at %s" % (start, off, 100 * acc_s / tot, 100 * acc_e / totex, r[isrc].strip()[:50]))
            acc_s = acc_e = 0; start = off + 16
    for r in sorted(data, key=lambda r: -int(r[isamp]))[:top]:
        print("   %5x %5s %9s %s" % (int(r[ia], 16) - base, r[isamp], r[iex], r[isrc].strip()[:90]))
