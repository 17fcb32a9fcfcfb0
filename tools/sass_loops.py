"""List the loops (backward branches) of one kernel in an `nvdisasm -c` listing with their static size.
usage: sass_loops.py dis.txt kernel_mangled_name [min_instructions]"""
import re, sys
dis, kern = sys.argv[1], sys.argv[2]
lo = int(sys.argv[3]) if len(sys.argv) > 3 else 200
lines = open(dis).read().split('\n')
start = next(i for i, l in enumerate(lines) if l.strip() in ('.text.%s:' % kern, '%s:' % kern))
labels = {}; insts = []
for l in lines[start + 1:]:
    if l.strip().startswith('.section') and len(insts) > 100: break
    m = re.match(r'\s*(\.L_x_\d+):', l)
    if m: labels[m.group(1)] = len(insts); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: insts.append(m.group(2))
print('static instructions', len(insts), '= %.1f KB' % (len(insts) * 16 / 1024))
loops = []
for i, t in enumerate(insts):
    m = re.search(r'\bBRA\S*\s+.*?(\.L_x_\d+)', t)
    if m and m.group(1) in labels and labels[m.group(1)] <= i: loops.append((i - labels[m.group(1)] + 1, labels[m.group(1)], i))
for n, a, b in sorted(loops, reverse=True):
    if n >= lo: print('loop %6d..%6d  %5d instructions  %.1f KB' % (a, b, n, n * 16 / 1024))
