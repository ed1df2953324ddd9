"""Summarise an .ncu-rep: per kernel, the source lines with the most warp-stall samples (needs -lineinfo + --import-source).
    python tools/ncu_lines.py gpurun_out/x.ncu-rep [top_n] [file_filter]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
flt = sys.argv[3] if len(sys.argv) > 3 else ""
col = sys.argv[4] if len(sys.argv) > 4 else "# Samples"     # e.g. "Instructions Executed"
kflt = sys.argv[5] if len(sys.argv) > 5 else ""           # substring of the kernel name
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
secs, cur = [], None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = {"file": r[1], "rows": []}; secs.append(cur)
    elif r[0] == "Function Name" and cur is not None:
        cur["fn"] = r[1]
    elif r[0] == "Line No" and cur is not None:
        cur["hdr"] = r
    elif cur is not None:
        cur["rows"].append(r)
for s in secs:
    if "hdr" not in s or flt not in s["file"] or kflt not in s.get("fn", ""):
        continue
    H = s["hdr"]; si = H.index(col)
    agg = collections.OrderedDict()
    for r in s["rows"]:
        if len(r) <= si or not r[si].isdigit() or not r[0]:
            continue
        k = (r[0], r[1].strip()[:110]); agg[k] = agg.get(k, 0) + int(r[si])
    tot = sum(agg.values())
    if not tot:
        continue
    print("==", s["file"].split("/")[-1], s.get("fn", "")[:80], "samples", tot)
    for (ln, src), v in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
        print("  %7d %5.1f%%  L%-4s %s" % (v, 100 * v / tot, ln, src))
