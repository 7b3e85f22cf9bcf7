"""Hot SASS instructions of one kernel from an .ncu-rep source page.
usage: ncu_hot.py rep kernel-substring [topN]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(raw)):
    if len(row) >= 2 and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": [], "hdr": None}
        blocks.append(cur)
    elif cur is not None:
        if cur["hdr"] is None:
            cur["hdr"] = row
        else:
            cur["rows"].append(row)
seen = set()
for b in blocks:
    if pat not in b["name"] or b["name"] in seen:
        continue
    seen.add(b["name"])
    h = b["hdr"]
    iS, iSamp, iEx = h.index("Source"), h.index("Warp Stall Sampling (All Samples)"), h.index("Instructions Executed")
    rows = b["rows"]
    tot = sum(int(r[iSamp] or 0) for r in rows)
    totex = sum(int(r[iEx] or 0) for r in rows)
    print("==", b["name"], "samples", tot, "warp-instr", totex, "sass lines", len(rows))
    # opcode histogram by executed count
    hist = {}
    for r in rows:
        op = r[iS].split()[0] if r[iS].split() else "?"
        if op.startswith("@"):
            op = r[iS].split()[1]
        op = op.split(".")[0]
        hist[op] = hist.get(op, 0) + int(r[iEx] or 0)
    print("  executed by opcode:", ", ".join("%s %.1f%%" % (k, 100.0 * v / max(totex, 1)) for k, v in sorted(hist.items(), key=lambda x: -x[1])[:18]))
    order = sorted(range(len(rows)), key=lambda i: -int(rows[i][iSamp] or 0))[:top]
    for i in sorted(order):
        r = rows[i]
        print("  %5d %5.1f%%  ex=%8s  %s" % (i, 100.0 * int(r[iSamp] or 0) / max(tot, 1), r[iEx], r[iS].strip()[:100]))
