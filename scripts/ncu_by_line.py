"""Aggregates an `ncu --page source --csv` export (SASS level) by CUDA source line, using nvdisasm -g line info of the same build.
Usage: ncu_by_line.py <source.csv> <nvdisasm -g output> <mangled kernel name substring> [top]"""
import csv, re, sys, collections
src_csv, dis, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
line_of = {}
cur = None
inside = False
for ln in open(dis):
    if ln.startswith(".text."):
        inside = kname in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ia, isamp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
base = None
agg = collections.defaultdict(lambda: [0, 0])
tot = [0, 0]
for r in rows[2:]:
    if len(r) <= iex or not r[ia].startswith("0x"):
        continue
    a = int(r[ia], 16)
    if base is None:
        base = a
    key = line_of.get(a - base, ("?", 0))
    s, e = int(r[isamp] or 0), int(r[iex] or 0)
    agg[key][0] += s; agg[key][1] += e
    tot[0] += s; tot[1] += e
print("total samples %d, warp instructions %d" % tuple(tot))
byfile = collections.defaultdict(lambda: [0, 0])
for (f, l), (s, e) in agg.items():
    byfile[f][0] += s; byfile[f][1] += e
for f, (s, e) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("%-40s samples %5.1f %%  instructions %5.1f %%" % (f, 100.0 * s / tot[0], 100.0 * e / tot[1]))
print()
for (f, l), (s, e) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-28s %5d  samples %5.2f %%  instructions %5.2f %%" % (f, l, 100.0 * s / tot[0], 100.0 * e / tot[1]))
if "--ranges" in sys.argv:
    pass
# optional: --ranges file:lo-hi=name,... buckets
for arg in sys.argv:
    if arg.startswith("--ranges="):
        buckets = []
        for spec in arg[len("--ranges="):].split(","):
            loc, name = spec.split("=")
            f, rng = loc.split(":")
            lo, hi = rng.split("-")
            buckets.append((f, int(lo), int(hi), name))
        out = collections.defaultdict(lambda: [0, 0])
        for (f, l), (s, e) in agg.items():
            name = "other"
            for bf, lo, hi, bn in buckets:
                if f == bf and lo <= l <= hi:
                    name = bn
                    break
            out[name][0] += s; out[name][1] += e
        print()
        for n_, (s, e) in sorted(out.items(), key=lambda kv: -kv[1][0]):
            print("%-30s samples %5.1f %%  instructions %5.1f %%" % (n_, 100.0 * s / tot[0], 100.0 * e / tot[1]))
