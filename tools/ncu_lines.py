"""Top source lines by warp-stall samples from an ncu report (needs -lineinfo and --import-source on).
usage: ncu_lines.py <report.ncu-rep> [n]"""
import collections
import csv
import subprocess
import sys

rep, n = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file, hdr, agg = None, None, collections.OrderedDict()
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split('/')[-1]
        continue
    if len(r) > 3 and r[0] == "Line No":
        hdr = r
        iS, iE = hdr.index("# Samples"), hdr.index("Instructions Executed")
        stallcols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr and len(r) > iE and r[0].strip().isdigit():
        try:
            st = {h: int(r[i]) for i, h in stallcols if r[i] not in ("", "0")}
            agg[(cur_file, int(r[0]))] = (int(r[iS]), int(r[iE]), r[1][:80], st)
        except ValueError:
            pass
tot = sum(v[0] for v in agg.values())
kinds = collections.Counter()
for v in agg.values():
    kinds.update(v[3])
print("total samples", tot, {k: f"{100 * c / tot:.1f}%" for k, c in kinds.most_common(8)})
for (f, l), (s, e, src, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    top = ", ".join(f"{k[6:]} {100 * c / max(s, 1):.0f}%" for k, c in sorted(st.items(), key=lambda x: -x[1])[:2])
    print(f"{f[:22]}:{l:4d} {100 * s / tot:5.1f}%  {src.strip()[:72]:72s} [{top}]")
