"""Table of per-launch times from tools/ablate.sh outputs: python tools/ablate_table.py 0 1 2 4"""
import csv, os, sys
def load(path):
    lines = [l for l in open(path) if l.startswith('"')]
    return [(x["Kernel Name"].split("(")[0].split("::")[-1], float(x["Metric Value"]) / 1e3) for x in csv.DictReader(lines)]
tabs = {}
for D in sys.argv[1:]:
    p = "gpurun_out/abl_%s.csv" % D
    if os.path.exists(p):
        t = load(p)
        if t:
            tabs[D] = t
n = min(len(t) for t in tabs.values())
print("idx kernel              " + " ".join("%8s" % D for D in tabs))
for i in range(n):
    print("%3d %-18s " % (i, list(tabs.values())[0][i][0][:18]) + " ".join("%8.1f" % tabs[D][i][1] for D in tabs))
print("sum                    " + " ".join("%8.0f" % sum(x[1] for x in tabs[D][:n]) for D in tabs))
