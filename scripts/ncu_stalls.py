"""Per-SASS-instruction stall summary of one ncu report (reads `ncu --page source --print-source sass`).
usage: python scripts/ncu_stalls.py report.ncu-rep [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(out.splitlines()))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[start], rows[start + 1:]
ix = {h: i for i, h in enumerate(hdr)}


def f(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0


tot = sum(f(r, "# Samples") for r in data)
print("kernel:", rows[0][1] if rows and len(rows[0]) > 1 else "?", " samples:", int(tot), " sass rows:", len(data))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = sorted(((sum(f(r, k) for r in data), k) for k in stalls), reverse=True)
print("stall totals:", ", ".join("%s %.1f%%" % (k[6:], 100 * v / tot) for v, k in agg if v > 0.005 * tot))
top = sorted(range(len(data)), key=lambda i: -f(data[i], "# Samples"))[:top_n]
for i in sorted(top):
    r = data[i]
    dom = max(stalls, key=lambda k: f(r, k))
    print("%5d %-72s %6d (%4.1f%%) %s" % (i, r[ix["Source"]].strip()[:72], f(r, "# Samples"),
                                          100 * f(r, "# Samples") / tot, dom[6:]))
