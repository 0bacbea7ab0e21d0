#!/usr/bin/env python
"""Per-source-line instruction shares of one kernel: joins `nvdisasm -g` line info of the cubin
with the per-SASS-instruction execution counts of an ncu report (`--page source --csv`).
usage: line_profile.py <nvdisasm.txt> <mangled kernel name> <ncu_source.csv> <file.cu>"""
import collections, csv, re, sys
dis, kern, srccsv, cu = sys.argv[1:5]
lines = open(dis).read().split('\n')
start = [i for i, l in enumerate(lines) if l.startswith('//--------------------- .text.' + kern)][0]
cur = None; m = {}; ins = {}
for l in lines[start + 1:]:
    if l.startswith('//--------------------- .text.'): break
    t = re.search(r'//## File ".*?([^/"]+)", line (\d+)', l)
    if t: cur = int(t.group(2)); continue
    a = re.search(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', l)
    if a: m[int(a.group(1), 16)] = cur; ins[int(a.group(1), 16)] = a.group(2)
rows = list(csv.reader(open(srccsv)))
hdr = rows[1]; ia = hdr.index('Address'); ie = hdr.index('Instructions Executed'); iss = hdr.index('Warp Stall Sampling (All Samples)')
data = []
for r in rows[2:]:
    if len(r) < 10 or r[0] == 'Kernel Name': break
    data.append((int(r[ia], 16), int(r[ie]), int(r[iss])))
base = data[0][0]
ALU = ('LOP3', 'SHF', 'PRMT', 'IADD3', 'ISETP', 'SEL', 'VIMNMX', 'VIADD', 'LEA', 'PLOP3', 'POPC', 'FLO', 'BREV', 'IABS', 'VIMNMX3', 'MOV')
byline = collections.Counter(); alu = collections.Counter(); st = collections.Counter()
for a, e, s in data:
    ln = m.get(a - base); byline[ln] += e; st[ln] += s
    op = [o for o in ins.get(a - base, '').split() if not o.startswith('@')]
    if op and op[0].split('.')[0] in ALU: alu[ln] += e
tot = sum(byline.values()); tots = max(1, sum(st.values()))
src = open(cu).read().split('\n')
print("total warp instructions %d" % tot)
for ln, e in sorted(byline.items(), key=lambda x: (x[0] or 0)):
    if e > tot * 0.004:
        print("%4s %5.1f%% alu=%5.1f%% stall=%5.1f%%  %s" % (ln, 100 * e / tot, 100 * alu[ln] / tot, 100 * st[ln] / tots,
                                                        (src[ln - 1].strip()[:88] if ln and ln <= len(src) else '')))
print('ALU total %.1f%%' % (100 * sum(alu.values()) / tot))
