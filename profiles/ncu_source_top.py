"""Top stalled SASS instructions of an .ncu-rep: python profiles/ncu_source_top.py REP [stall_col] [n]"""
import csv, subprocess, sys
rep = sys.argv[1]; col = sys.argv[2] if len(sys.argv) > 2 else "Warp Stall Sampling (All Samples)"; n = int(sys.argv[3]) if len(sys.argv) > 3 else 15
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; body = [r for r in rows[hi + 1:] if len(r) == len(h)]
ci, si, ai = h.index(col), h.index("Source"), h.index("Warp Stall Sampling (All Samples)")
f = lambda x: float(x) if x.replace('.', '', 1).isdigit() else 0.0
tot = sum(f(r[ci]) for r in body)
print(f"total {col}: {tot:.0f} over {len(body)} instructions")
for idx, r in sorted(enumerate(body), key=lambda t: -f(t[1][ci]))[:n]:
    print(f"{f(r[ci]):8.0f} all={f(r[ai]):7.0f} #{idx:5d} {r[si].strip()[:100]}")
