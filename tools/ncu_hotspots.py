"""Aggregate an `ncu --page source --csv` SASS dump by CUDA source line using nvdisasm line info.
usage: ncu_hotspots.py src.csv dis.txt kernel_mangled_name"""
import collections, csv, re, sys
src, dis, kern = sys.argv[1], sys.argv[2], sys.argv[3]
lines = open(dis).read().split('\n')
start = None
for i, l in enumerate(lines):
    if l.strip() == '.text.%s:' % kern or l.strip() == '%s:' % kern:
        start = i; break
cur = None; off2loc = {}
for l in lines[start:]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S+)', l)
    if m and cur: off2loc[int(m.group(1), 16)] = cur
    if l.strip().startswith('.section') and len(off2loc) > 100: break
rows = list(csv.reader(open(src)))
hdr = rows[1]; ai = hdr.index('Address'); ii = hdr.index('Instructions Executed'); si = hdr.index('# Samples'); so = hdr.index('Source')
base = None; agg = collections.Counter(); sagg = collections.Counter(); op = collections.Counter(); tot = 0
for r in rows[2:]:
    if len(r) <= ii or not r[ai].startswith('0x'): continue
    a = int(r[ai], 16)
    if base is None: base = a
    n = int(r[ii]); s = int(r[si]); loc = off2loc.get(a - base, ('?', 0))
    agg[loc] += n; sagg[loc] += s; tot += n
    t = r[so].split(); o = t[1] if t[0].startswith('@') else t[0]; op[o.split('.')[0]] += n
stot = sum(sagg.values())
print('total warp instructions', tot, 'static', len(off2loc))
byf = collections.Counter(); sbyf = collections.Counter()
for (f, l), n in agg.items(): byf[f] += n
for (f, l), n in sagg.items(): sbyf[f] += n
for f, n in byf.most_common(): print('%-28s inst %5.1f%%  samples %5.1f%%' % (f, 100 * n / tot, 100 * sbyf[f] / stot))
print('--- top lines (by samples)')
for (f, l), n in sagg.most_common(30): print('%-22s %5d inst %5.2f%% samples %5.2f%%' % (f, l, 100 * agg[(f, l)] / tot, 100 * n / stot))
print('--- opcodes')
print(', '.join('%s %.1f%%' % (k, 100 * v / tot) for k, v in op.most_common(18)))
st = [i for i, c in enumerate(hdr) if c.startswith('stall_') and 'Not Issued' not in c]
tt = collections.Counter()
for r in rows[2:]:
    if len(r) < len(hdr) - 2: continue
    for i in st:
        try: tt[hdr[i]] += int(r[i])
        except ValueError: pass
s = sum(tt.values())
print('--- stalls: ' + ', '.join('%s %.1f%%' % (k, 100 * v / s) for k, v in tt.most_common(8)))
