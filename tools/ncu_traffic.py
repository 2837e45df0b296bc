#!/usr/bin/env python3
"""profiles/ncu_traffic.json from an `ncu --page raw --csv` export of one step: per kernel (base name, largest launch) the DRAM
bytes (dram__bytes_read.sum + dram__bytes_write.sum) and the warp instructions (smsp__inst_executed.sum) per launch.
usage: ncu_traffic.py raw.csv bench_line.json "source text" > profiles/ncu_traffic.json"""
import csv
import json
import re
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    ri, wi, ii = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("smsp__inst_executed.sum")
    best = {}
    for r in rows[2:]:
        name = re.sub(r"<.*", "", r[ki].split("(")[0]).replace("void ", "").split("::")[-1].strip()
        b = float(r[ri].replace(",", "")) * UNIT[units[ri]] + float(r[wi].replace(",", "")) * UNIT[units[wi]]
        if name not in best or b > best[name][0]:
            best[name] = (b, float(r[ii].replace(",", "")))
    line = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    out = {"source": sys.argv[3], "P": line["config"]["P"], "num_rendered": line["config"]["num_rendered"],
           "kernels": {k: int(v[0]) for k, v in best.items()}, "warp_instructions": {k: int(v[1]) for k, v in best.items()}}
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
