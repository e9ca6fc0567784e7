#!/usr/bin/env bash
# Usage (on the GPU box): scripts/ncu_export.sh <tag> <kernel-regex> <python args...>
# Runs the command plainly first, then under ncu --set full; exports text pages into gpurun_out/ (the binary
# report stays in /tmp because it embeds the whole cubin and exceeds the gpurun_out size limit).
set -u
tag=$1; shift; regex=$1; shift
mkdir -p gpurun_out
timeout -s KILL 300 python "$@" > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:$regex -s 1 -c 1 -f -o /tmp/${tag} python "$@" > gpurun_out/${tag}_ncu.log 2>&1
ncu -i /tmp/${tag}.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>&1
ncu -i /tmp/${tag}.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>&1
ncu -i /tmp/${tag}.ncu-rep --page source --csv > /tmp/${tag}_source.csv 2>&1
python - "$tag" <<'PY'
import csv, sys
tag = sys.argv[1]
rows = list(csv.reader(open(f"/tmp/{tag}_source.csv", errors="replace")))
hdr = None
for i, r in enumerate(rows):
    if any("Sampling" in c for c in r) or any(c.strip() == "Source" for c in r):
        hdr = i; break
if hdr is None:
    open(f"gpurun_out/{tag}_source_top.csv", "w").write("no header found\n" + "\n".join(",".join(r) for r in rows[:5]))
    sys.exit(0)
h = rows[hdr]
col = next((j for j, c in enumerate(h) if "Warp Stall Sampling (All" in c), None)
body = [r for r in rows[hdr + 1:] if len(r) == len(h)]
def val(r):
    try: return float(r[col])
    except Exception: return 0.0
if col is not None:
    body.sort(key=val, reverse=True)
# instruction mix over ALL SASS lines: executed warp-instructions and stall samples per opcode
import collections
icol = next((j for j, c in enumerate(h) if c.strip() == "Instructions Executed"), None)
scol = next((j for j, c in enumerate(h) if c.strip() == "Source"), 1)
mix, smp = collections.Counter(), collections.Counter()
def num(x):
    try: return float(x)
    except Exception: return 0.0
for r in body:
    t = r[scol].split()
    op = (t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")).rstrip(";")
    if icol is not None: mix[op] += num(r[icol])
    if col is not None: smp[op] += num(r[col])
ti, ts = sum(mix.values()) or 1.0, sum(smp.values()) or 1.0
with open(f"gpurun_out/{tag}_opcode_mix.txt", "w") as f:
    f.write(f"# executed warp-instructions {ti:.4e}, stall samples {ts:.0f}\n")
    for op, v in mix.most_common(60):
        f.write(f"{op:32s} {v:14.4e} {100 * v / ti:6.2f}%   samples {100 * smp[op] / ts:6.2f}%\n")
w = csv.writer(open(f"gpurun_out/{tag}_source_top.csv", "w"))
w.writerow(h)
for r in body[:400]:
    w.writerow(r)
PY
tail -1 gpurun_out/${tag}_plain.log; ls -la gpurun_out | tail -8
