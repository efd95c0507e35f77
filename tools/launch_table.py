"""Mean device time per kernel from an ncu launch-list csv (tools/profile_round.sh)."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
d = collections.OrderedDict()
for r in rows[hdr + 1:]:
    name = r[4].split("(")[0][-52:]
    d.setdefault((name, r[8], r[7]), []).append(float(r[-1]) / 1e3)
tot = 0.0
for k, v in d.items():
    print(f"{k[0]:54s} grid {k[1]:16s} block {k[2]:14s} n={len(v):3d} mean={sum(v) / len(v):8.1f} us")
