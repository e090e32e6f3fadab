"""Hot spots of an `ncu --page source --csv` export: python tools/src_hot.py file.source.csv [N]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))[2:]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
tot = sum(int(r[2]) for r in rows)
print("samples", tot, "instructions", len(rows))
for a in range(0, len(rows), 300):
    s = sum(int(r[2]) for r in rows[a:a + 300]); ex = max(int(r[5]) for r in rows[a:a + 300])
    n = sum(1 for r in rows[a:a + 300] if int(r[5]) == ex)
    if s: print("  [%4d..] samples %6d  max exec %8d (%d instrs)" % (a, s, ex, n))
c = collections.Counter(int(r[5]) for r in rows if int(r[5]) > 0)
print("exec-count groups:", sorted(c.items(), key=lambda x: -x[1])[:6])
top = sorted(range(len(rows)), key=lambda i: -int(rows[i][2]))[:N]
for i in sorted(top):
    r = rows[i]; print(i, r[1].strip()[:84], r[2], r[5])
