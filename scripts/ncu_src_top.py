"""Print the SASS lines of an `ncu --page source --csv` export that collect the most stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
print("kernel:", rows[0][1][:120], " lines:", len(body), " samples:", tot)
idx = sorted(range(len(body)), key=lambda i: -int(body[i][col["# Samples"]] or 0))[:n]
for i in sorted(idx):
    r = body[i]
    print("%5d %6s %5.1f%% exec=%-9s lsb=%-6s %s" % (i, r[col["# Samples"]], 100.0 * int(r[col["# Samples"]] or 0) / max(tot, 1),
          r[col["Instructions Executed"]], r[col["stall_long_sb"]], r[col["Source"]][:110]))
