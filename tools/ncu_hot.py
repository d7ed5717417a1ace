"""Summarises an `ncu --page source --csv` dump: total stall samples per reason and the hottest SASS lines."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
data = []
for r in rows[hi + 1:]:
    if r and r[0] in ("Kernel Name", "Address"):
        break  # first kernel instance only
    if len(r) == len(hdr):
        data.append(r)
col = {n: i for i, n in enumerate(hdr)}
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
tot = {n: sum(int(r[col[n]] or 0) for r in data) for n in stall_cols}
alls = sum(int(r[col["# Samples"]] or 0) for r in data)
print("total samples", alls, "instructions executed", sum(int(r[col["Instructions Executed"]] or 0) for r in data))
for n, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
    print("  %-24s %8d %5.1f%%" % (n, v, 100.0 * v / max(alls, 1)))
print("hottest lines:")
for r in sorted(data, key=lambda r: -int(r[col["# Samples"]] or 0))[:top]:
    reasons = sorted(((int(r[col[n]] or 0), n) for n in stall_cols), reverse=True)[:2]
    print("  %6s %7s  %-70s %s" % (r[col["# Samples"]], r[col["Instructions Executed"]], r[col["Source"]].strip()[:70],
                                   " ".join("%s=%d" % (n[6:], v) for v, n in reasons if v)))
