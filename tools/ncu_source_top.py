"""Top stall sites of an ncu report with source correlation.  usage: ncu_source_top.py report.ncu-rep [N]
(reads `ncu -i report --page source --csv`; needs -lineinfo at compile time and --import-source on at capture)"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], []
for line in txt.splitlines():
    if line.startswith('"Kernel Name"'):
        if cur:
            blocks.append(cur)
        cur = [line]
    elif cur:
        cur.append(line)
if cur:
    blocks.append(cur)
for b in blocks[:1]:
    print(b[0][:150])
    rows = list(csv.reader(io.StringIO("\n".join(b[1:]))))
    hdr = rows[0]
    idx = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[1:] if len(r) == len(hdr)]
    tot = sum(float(r[idx["# Samples"]] or 0) for r in body)
    inst = sum(float(r[idx["Instructions Executed"]] or 0) for r in body)
    print("total samples %d, warp instructions %d" % (tot, inst))
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {h: sum(float(r[idx[h]] or 0) for r in body) for h in stalls}
    print("stall mix:", ", ".join("%s %.1f%%" % (h[6:], 100 * v / max(tot, 1)) for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    # opcode histogram
    ops = {}
    for r in body:
        op = r[idx["Source"]].split()[0] if r[idx["Source"]].split() else "?"
        if op.startswith("@"):
            op = r[idx["Source"]].split()[1]
        op = op.split(".")[0]
        ops[op] = ops.get(op, 0) + float(r[idx["Instructions Executed"]] or 0)
    print("opcode mix:", ", ".join("%s %.1f%%" % (k, 100 * v / max(inst, 1)) for k, v in sorted(ops.items(), key=lambda x: -x[1])[:16]))
    body.sort(key=lambda r: -float(r[idx["# Samples"]] or 0))
    for r in body[:top]:
        st = sorted(((float(r[idx[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
        print("%6.2f%%  %-70s %s" % (100 * float(r[idx["# Samples"]] or 0) / max(tot, 1), r[idx["Source"]][:70],
                                      " ".join("%s:%d" % (n, v) for v, n in st if v)))
