"""Top stalled SASS instructions of an ncu report's source page (CSV): python tools/ncu_hot.py file.ncu-rep [frac] [ctx]"""
import csv, subprocess, sys
rep = sys.argv[1]; frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01; ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
h = next(i for i, row in enumerate(r) if "Source" in row and "# Samples" in row)
H = r[h]; si = H.index("# Samples"); so = H.index("Source"); ie = H.index("Instructions Executed")
stall = [(i, n) for i, n in enumerate(H) if n.startswith("stall_") and "Not Issued" not in n]
rows = [x for x in r[h + 1:] if len(x) > si and x[si].isdigit()]
tot = sum(int(x[si]) for x in rows)
print("total samples", tot)
for k, x in enumerate(rows):
    s = int(x[si])
    if s >= tot * frac:
        top = sorted(((int(x[i] or 0), n) for i, n in stall), reverse=True)[:2]
        for c in range(max(0, k - ctx), k):
            print(f"        {'':14}      #{c:5d} {rows[c][so][:100]}")
        print(f"{s:7d} {100*s/tot:5.1f}% exec={x[ie]:>8} #{k:5d} {x[so][:90]:90} {top[0][1]}={top[0][0]} {top[1][1]}={top[1][0]}")
