"""Top source lines of one profiled launch by stall samples / instructions executed.
usage: ncu_lines.py rep [launch-skip] [topN]    (needs -lineinfo + --import-source on)"""
import csv, io, subprocess, sys
rep = sys.argv[1]; skip = sys.argv[2] if len(sys.argv) > 2 else "0"; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
agg = {}; fname = ""; hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Name": fname = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Kernel Name": print(r[1][:120]); continue
    if r[0] == "Line No": hdr = r; ix = {h: i for i, h in enumerate(r)}; continue
    if hdr and len(r) == len(hdr) and "# Samples" in ix:
        try:
            ln = int(r[0])
        except ValueError:
            continue
        key = (fname, ln)
        e = agg.setdefault(key, {"src": r[1].strip()[:100], "smp": 0, "ins": 0, "st": {}})
        e["smp"] += int(r[ix["# Samples"]] or 0); e["ins"] += int(r[ix["Instructions Executed"]] or 0)
        for h, i in ix.items():
            if h.startswith("stall_") and "Not Issued" not in h:
                e["st"][h] = e["st"].get(h, 0) + int(r[i] or 0)
recs = [(e["smp"], e["ins"], k[0], k[1], e["src"], e["st"]) for k, e in agg.items()]
ts = sum(x[0] for x in recs) or 1; ti = sum(x[1] for x in recs) or 1
print(f"total samples {ts}, total warp instructions {ti}")
keys = ["stall_barrier", "stall_long_sb", "stall_short_sb", "stall_mio", "stall_lg", "stall_math", "stall_wait", "stall_not_selected", "stall_branch_resolving", "stall_dispatch"]
for smp, ins, f, ln, src, d in sorted(recs, reverse=True)[:top]:
    st = sorted(((int(d.get(k, "0") or 0), k[6:]) for k in keys), reverse=True)[:3]
    print(f"{100*smp/ts:5.1f}% smp {100*ins/ti:5.1f}% ins  {f}:{ln:<4d} {src}   [{' '.join(f'{n}:{v}' for v, n in st if v)}]")
