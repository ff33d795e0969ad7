"""per-region LSU wavefront accounting from an ncu report's source page (regions split at BAR.SYNC)"""
import csv, sys, subprocess, re
rep = sys.argv[1]
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
if len(sys.argv) > 2 and sys.argv[2] == "cols":
    print("\n".join(hdr)); sys.exit()
data = rows[hi + 1:]
ia = hdr.index("Source")
cols = [c for c in hdr if "avefront" in c or c in ("# Samples", "Instructions Executed")]
idx = {c: hdr.index(c) for c in cols}
bars = [k for k, r in enumerate(data) if "BAR.SYNC" in r[ia]] + [len(data)]
def num(x):
    try: return float(x)
    except: return 0.0
prev = 0
print("columns:", cols)
for b in bars:
    tot = {c: 0.0 for c in cols}
    byop = {}
    for r in data[prev:b]:
        op = re.sub(r"@!?U?P\d+\s+", "", r[ia].strip()).split()[0].split(".")[0] if r[ia].strip() else "?"
        for c in cols: tot[c] += num(r[idx[c]])
        w = sum(num(r[idx[c]]) for c in cols if "avefront" in c and "Ideal" not in c and "Excessive" not in c)
        if w: byop[op] = byop.get(op, 0) + w
    print("region [%d,%d): " % (prev, b) + "  ".join("%s=%.1fM" % (c.replace("L1 Wavefronts", "W").replace("Wavefronts", "W"), tot[c] / 1e6) for c in cols))
    print("      by op: " + "  ".join("%s=%.1fM" % (k, v / 1e6) for k, v in sorted(byop.items(), key=lambda kv: -kv[1])))
    prev = b
