"""Static size of photo_strip_kernel<3,2>'s row loop: instructions between the backward branch and its
target, opcode histogram, registers / spills.  Usage: python tools/sass_loop.py [build/photo_fused.o] [-v]"""
import re, subprocess, sys, collections
obj = next((a for a in sys.argv[1:] if not a.startswith('-')), 'monodepth2.jl_b200/build/photo_fused.o')
sass = subprocess.run(['cuobjdump', '-sass', obj], capture_output=True, text=True).stdout
fn = None; rows = []
for line in sass.splitlines():
    if 'Function :' in line:
        fn = line.split(':', 1)[1].strip(); continue
    if fn and 'photo_strip_kernelILi3ELi2ELb0' in fn:
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', line)
        if m: rows.append((int(m.group(1), 16), m.group(2).strip()))
back = [(a, int(re.search(r'0x([0-9a-f]+)', t).group(1), 16)) for a, t in rows if t.split()[0].startswith(('BRA', '@')) and 'BRA' in t and re.search(r'0x([0-9a-f]+)', t) and int(re.search(r'0x([0-9a-f]+)', t).group(1), 16) < a]
# the row loop is the longest backward branch
a, tgt = max(back, key=lambda p: p[0] - p[1])
body = [t for ad, t in rows if tgt <= ad <= a]
print(f'total {len(rows)} instr; row loop {len(body)} instr (0x{tgt:x}..0x{a:x})')
h = collections.Counter(re.sub(r'^@!?U?P\d+\s+', '', t).split()[0].split('.')[0] for t in body)
print(' '.join(f'{k}:{v}' for k, v in h.most_common(28)))
res = subprocess.run(['cuobjdump', '-res-usage', obj], capture_output=True, text=True).stdout.splitlines()
for i, l in enumerate(res):
    if 'photo_strip_kernelILi3ELi2ELb0' in l: print(res[i + 1].strip())
