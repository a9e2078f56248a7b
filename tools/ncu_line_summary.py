"""Executed warp instructions per SOURCE LINE of one kernel: joins `ncu --page source --csv` (SASS rows with executed
counts and stall samples) with `nvdisasm --print-line-info` of the same cubin (instruction order is identical).

    cuobjdump -xelf all build/.../trunc_float_1_8.o; nvdisasm --print-line-info trunc_inst.sm_100a.cubin > dis.txt
    ncu -i rep.ncu-rep --page source --csv > src.csv
    python tools/ncu_line_summary.py src.csv dis.txt '<mangled kernel name>' <warp-elements>
"""
import collections
import csv
import re
import sys

src_csv, dis, name, elems = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
rows = list(csv.reader(open(src_csv)))
hi = [i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r][0]
ix = {k: i for i, k in enumerate(rows[hi])}
data = rows[hi + 1:]
seq, cur, on = [], None, False
for l in open(dis).read().splitlines():
    if l.startswith('//---') and '.text.' in l:
        on = ('.text.' + name + ' ') in l
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        seq.append((cur, m.group(2).strip()))
assert len(seq) == len(data), (len(seq), len(data))
per, perop, samp, tot = collections.Counter(), collections.defaultdict(collections.Counter), collections.Counter(), 0
for (cur, ins), r in zip(seq, data):
    n = int(r[ix['Instructions Executed']] or 0)
    toks = ins.split()
    if toks[0].startswith('@'):
        toks = toks[1:]
    per[cur] += n
    perop[cur][toks[0].split('.')[0]] += n
    samp[cur] += int(r[ix['# Samples']] or 0)
    tot += n
st = max(1, sum(samp.values()))
print("kernel %s: %d warp instructions, %.2f per warp-element" % (name, tot, tot / elems))
print("%-34s %9s %9s  top opcodes (per element)" % ("source line", "inst/elem", "samples%"))
for k, v in sorted(per.items(), key=lambda kv: -kv[1])[:45]:
    print("%-34s %9.2f %9.1f  %s" % ("%s:%d" % k if k else "?", v / elems, 100.0 * samp[k] / st,
                                     ", ".join("%s %.2f" % (o, c / elems) for o, c in perop[k].most_common(4))))
