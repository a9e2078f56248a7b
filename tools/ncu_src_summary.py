"""Summarise `ncu --page source --csv`: instruction mix (executed warp instructions by opcode,
weighted) and stall-reason totals.  Usage: python tools/ncu_src_summary.py file.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: k for k, h in enumerate(hdr)}
ops = collections.Counter()
samples = collections.Counter()
stalls = collections.Counter()
tot_inst = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    src = r[ix['Source']].strip()
    toks = src.split()
    if toks and toks[0].startswith('@'):
        toks = toks[1:]
    op = toks[0].split('.')[0] if toks else '?'
    n = int(r[ix['Instructions Executed']] or 0)
    ops[op] += n
    tot_inst += n
    samples[op] += int(r[ix['# Samples']] or 0)
    for h in hdr:
        if h.startswith('stall_') and '(Not Issued)' not in h:
            stalls[h] += int(r[ix[h]] or 0)
print("total warp instructions executed: %d" % tot_inst)
print("%-10s %8s %8s" % ("opcode", "inst %", "samples %"))
ts = sum(samples.values())
for op, n in ops.most_common(28):
    print("%-10s %7.2f%% %7.2f%%" % (op, 100.0 * n / tot_inst, 100.0 * samples[op] / max(ts, 1)))
print("stall reasons (all samples):")
t = sum(stalls.values())
for k, v in stalls.most_common(10):
    print("  %-24s %6.2f%%" % (k, 100.0 * v / max(t, 1)))
