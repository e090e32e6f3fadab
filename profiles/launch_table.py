"""Per-launch table from an `ncu --metrics gpu__time_duration.sum --csv` log (one forward of our kernels).
usage: python profiles/launch_table.py A.csv [B.csv]"""
import csv
import sys


def load(path):
    lines = [l for l in open(path) if l.startswith('"')]
    out = []
    for x in csv.DictReader(lines):
        name = x["Kernel Name"]
        if "_kernel" not in name or "at::" in name:
            continue
        short = name.split("(")[0].split("::")[-1].replace("void ", "")
        out.append((short, x["Grid Size"], float(x["Metric Value"]) / 1e3))
    return out


def main():
    tabs = [load(p) for p in sys.argv[1:]]
    tabs = [[x for x in t if not x[0].startswith(("nms", "rescale"))] for t in tabs]
    n = min(len(t) for t in tabs)
    n = min(n, 90)
    for i in range(n):
        print("%3d " % i + " | ".join("%-18s %-14s %9.1f" % (t[i][0][:18], t[i][1], t[i][2]) for t in tabs))
    print("sum us: " + " | ".join("%.1f" % sum(x[2] for x in t[:n]) for t in tabs))


if __name__ == "__main__":
    main()
