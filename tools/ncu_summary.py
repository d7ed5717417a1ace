"""Writes a compact text summary of an ncu --set full report (first launch of each kernel)."""
import csv
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], stderr=subprocess.DEVNULL).decode()
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_bytes.sum", "smsp__cycles_active.avg"]
idx = [(w, hdr.index(w)) for w in want if w in hdr]
ki = hdr.index("Kernel Name")
seen = set()
with open(out, "w") as f:
    f.write("# from %s (ncu --set full --clock-control none); first captured launch of each kernel\n" % rep)
    for r in rows[2:]:
        name = r[ki].split("(")[0]
        if name in seen:
            continue
        seen.add(name)
        f.write("\n[%s]\n" % name)
        for w, i in idx:
            f.write("  %-66s %18s %s\n" % (w, r[i], units[i]))
print(open(out).read()[:400])
