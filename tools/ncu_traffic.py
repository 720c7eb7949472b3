"""Collects DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) from the committed `ncu --set full`
summaries under profiles/ into profiles/traffic.json, which bench.py reads for `roofline.traffic`.
    python tools/ncu_traffic.py kitti:k_hard_fwd=profiles/r01_kitti_k_hard_fwd_full_v2.txt ..."""
import json
import os
import re
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def parse(path):
    out = {}
    for line in open(os.path.join(ROOT, path)):
        m = re.match(r"^dram__bytes_(read|write)\.sum\s+([0-9.]+)\s+(\w+)", line)
        if m:
            out[m.group(1)] = int(round(float(m.group(2)) * UNIT[m.group(3)]))
    return out


def main():
    dst = os.path.join(ROOT, "profiles", "traffic.json")
    table = json.load(open(dst)) if os.path.exists(dst) else {}
    for arg in sys.argv[1:]:
        key, path = arg.split("=")
        rw = parse(path)
        table[key] = {"read": rw["read"], "write": rw["write"], "bytes": rw["read"] + rw["write"], "source": path}
    json.dump(table, open(dst, "w"), indent=1, sort_keys=True)
    print(json.dumps(table, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
