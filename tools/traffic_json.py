"""DRAM traffic of the conv stack from an ncu metrics log:
   ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:conv_gemm_kernel --launch-skip 75 -c 75 --csv
usage: python tools/traffic_json.py conv_traffic.csv BATCH > profiles/rN_conv_traffic.json"""
import csv
import json
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    path, batch = sys.argv[1], int(sys.argv[2])
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    per = {}
    for r in rows:
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Name"].startswith("dram__bytes"):
            v *= UNIT[r["Metric Unit"]]
        per.setdefault(r["ID"], {})[r["Metric Name"]] = v
    rd = sum(m["dram__bytes_read.sum"] for m in per.values())
    wr = sum(m["dram__bytes_write.sum"] for m in per.values())
    print(json.dumps({"launches": len(per), "batch": batch, "dram_read_bytes": rd, "dram_write_bytes": wr,
                      "bytes_per_tile": (rd + wr) / batch,
                      "source": "ncu dram__bytes_read.sum + dram__bytes_write.sum over the %d conv_gemm_kernel launches of one "
                                "ay_forward (second forward of `bench.py --steps 2 --warmup 3`)" % len(per)}, indent=1))


if __name__ == "__main__":
    main()
