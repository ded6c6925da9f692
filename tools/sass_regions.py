"""Static SASS instruction counts of one kernel by source file and by the '// ----' regions of a source file.
usage: sass_regions.py disasm.txt kernel_symbol_prefix [source.cuh]   (disasm = nvdisasm -g -c of the cubin)"""
import re, sys
lines = open(sys.argv[1]).read().split('\n')
sym = sys.argv[2]
start = end = None
for i, l in enumerate(lines):
    if start is None and l.startswith('.text.' + sym): start = i
    elif start is not None and l.startswith('.text.'): end = i; break
end = end or len(lines)
cur = None; cnt = {}
for l in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,6}\*/', l): cnt[cur] = cnt.get(cur, 0) + 1
tot = sum(cnt.values())
print('%s: %d instructions, %d bytes' % (lines[start][6:70], tot, tot * 16))
byfile = {}
for (f, l), n in cnt.items(): byfile[f] = byfile.get(f, 0) + n
for f, n in sorted(byfile.items(), key=lambda x: -x[1]): print('  %-32s %5d' % (f, n))
if len(sys.argv) > 3:
    name = sys.argv[3].split('/')[-1]
    src = open(sys.argv[3]).read().split('\n')
    marks = [(i + 1, l.strip()[:64]) for i, l in enumerate(src) if l.strip().startswith('// ----') or l.startswith('template') or l.startswith('static __global__') or l.startswith('__device__')]
    for i, (l0, nm) in enumerate(marks):
        l1 = marks[i + 1][0] if i + 1 < len(marks) else 10 ** 9
        n = sum(v for (f, l), v in cnt.items() if f == name and l0 <= l < l1)
        if n: print('    L%4d %5d  %s' % (l0, n, nm))
