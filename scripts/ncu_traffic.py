"""Extract per-launch DRAM traffic and shared-memory wavefronts of the search kernel from committed
`ncu --set full ... --page raw --csv` exports and write profiles/ncu_traffic.json, which bench.py reads for
`roofline.traffic` (so the number in the bench line is traceable to a committed capture, per config).

    python scripts/ncu_traffic.py C2=profiles/r02_ncu_c2_raw.csv C3u16=profiles/r02_ncu_c3u16_raw.csv ...
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "": 1.0, "inst": 1.0,
        "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def parse(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h, units, body = rows[hi], rows[hi + 1], rows[hi + 2:]

    def col(name, r):
        i = h.index(name)
        return float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)
    out = []
    for r in body:
        if len(r) != len(h):
            continue
        out.append({"kernel": r[h.index("Kernel Name")].split("(")[0],
                    "dram_bytes": col("dram__bytes_read.sum", r) + col("dram__bytes_write.sum", r),
                    "dram_bytes_read": col("dram__bytes_read.sum", r),
                    "smem_wavefronts": col("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", r),
                    "inst_executed": col("smsp__inst_executed.sum", r),
                    "time_ms": col("gpu__time_duration.sum", r)})
    return out


def main():
    dst = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    table = json.load(open(dst)) if os.path.exists(dst) else {}
    for arg in sys.argv[1:]:
        cfg, path = arg.split("=", 1)
        launches = parse(path)
        top = max(launches, key=lambda d: d["time_ms"])
        top["source"] = os.path.relpath(os.path.abspath(path), ROOT)
        table[cfg] = top
    json.dump(table, open(dst, "w"), indent=1, sort_keys=True)
    print(json.dumps(table, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
