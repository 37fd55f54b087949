"""Summarise an ncu source-page CSV (ncu -i X.ncu-rep --page source --csv --print-source sass):
per kernel, the stall-reason totals and the hottest SASS instructions."""
import csv, io, sys

path = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
blocks = open(path).read().split('"Kernel Name",')
seen = set()
for blk in blocks[1:]:
    lines = blk.splitlines()
    name = lines[0].strip('",')
    if pat not in name or name in seen:
        continue
    seen.add(name)
    rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = {h: 0 for h in stalls}
    data = []
    for k, r in enumerate(rows[1:]):
        try:
            n = int(r[col["# Samples"]])
        except Exception:
            continue
        for h in stalls:
            tot[h] += int(r[col[h]] or 0)
        best = max(stalls, key=lambda h: int(r[col[h]] or 0))
        data.append((n, k, r[col["Source"]].strip(), best, r[col["Instructions Executed"]]))
    total = sum(d[0] for d in data)
    print("=" * 100)
    print(name, "| samples", total, "| instructions", len(data))
    print("stall totals:", ", ".join(f"{h[6:]}={100 * v / max(total, 1):.1f}%" for h, v in
                                      sorted(tot.items(), key=lambda kv: -kv[1])[:8]))
    for n, k, src, best, ex in sorted(data, reverse=True)[:top]:
        print(f"{n:7d} {100 * n / max(total, 1):5.1f}%  #{k:5d} x{ex:>8s} {best[6:]:14s} {src[:90]}")
