"""Static size of photo_strip_kernel<3,2>'s row loop: instructions between the backward branch and its
target, opcode histogram, registers / spills.  Usage: python tools/sass_loop.py [build/photo_fused.o] [-v] [--json out.json]
--json records the histogram, the fp32 flops per lane and row (2 FFMA + FADD + FMUL + 4 FFMA2 + 2 FADD2 + 2 FMUL2), the
-res-usage line and the git blob hash of csrc/photo_strip.inl it was built from: bench.py's roofline reads it and
prints null when the kernel source has changed since."""
import re, subprocess, sys, collections, json, hashlib, os
obj = next((a for a in sys.argv[1:] if a.endswith('.o')), 'monodepth2.jl_b200/build/photo_fused.o')
sass = subprocess.run(['cuobjdump', '-sass', obj], capture_output=True, text=True).stdout
fn = None; rows = []
for line in sass.splitlines():
    if 'Function :' in line:
        fn = line.split(':', 1)[1].strip(); continue
    if fn and 'photo_strip_kernelILi3ELi2ELb0ELb1ELb0' in fn:
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', line)
        if m: rows.append((int(m.group(1), 16), m.group(2).strip()))
back = [(a, int(re.search(r'0x([0-9a-f]+)', t).group(1), 16)) for a, t in rows if t.split()[0].startswith(('BRA', '@')) and 'BRA' in t and re.search(r'0x([0-9a-f]+)', t) and int(re.search(r'0x([0-9a-f]+)', t).group(1), 16) < a]
# the steady-state row loop: the innermost of the long backward branches (the outer one also spans the generic copy of the
# body that the <= 6 edge rows of a segment run, STRIP_PEEL)
long_ = [p for p in back if (p[0] - p[1]) // 16 > 500]
a, tgt = min(long_, key=lambda p: p[0] - p[1])
body = [t for ad, t in rows if tgt <= ad <= a]
print(f'total {len(rows)} instr; row loop {len(body)} instr (0x{tgt:x}..0x{a:x})')
h = collections.Counter(re.sub(r'^@!?U?P\d+\s+', '', t).split()[0].split('.')[0] for t in body)
print(' '.join(f'{k}:{v}' for k, v in h.most_common(28)))
res = subprocess.run(['cuobjdump', '-res-usage', obj], capture_output=True, text=True).stdout.splitlines()
for i, l in enumerate(res):
    if 'photo_strip_kernelILi3ELi2ELb0ELb1ELb0' in l: print(res[i + 1].strip())

if '--json' in sys.argv:
    out = sys.argv[sys.argv.index('--json') + 1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    def blob(path):
        d = open(os.path.join(root, path), 'rb').read()
        return hashlib.sha1(b'blob %d\0' % len(d) + d).hexdigest()
    flops = 2 * h['FFMA'] + h['FADD'] + h['FMUL'] + 4 * h['FFMA2'] + 2 * h['FADD2'] + 2 * h['FMUL2']
    usage = [res[i + 1].strip() for i, l in enumerate(res) if 'photo_strip_kernelILi3ELi2ELb0ELb1ELb0' in l]
    json.dump({'kernel': 'photo_strip_kernel<3,2,false>', 'row_loop_instructions': len(body), 'histogram': dict(h.most_common()),
               'fp32_flops_per_lane_row': flops, 'fp32_flop_formula': '2 FFMA + FADD + FMUL + 4 FFMA2 + 2 FADD2 + 2 FMUL2 (static count of the row loop)',
               'res_usage': usage[0] if usage else None,
               'source_hash': {'monodepth2.jl_b200/csrc/photo_strip.inl': blob('monodepth2.jl_b200/csrc/photo_strip.inl')},
               'how': 'python tools/sass_loop.py --json ' + out}, open(out, 'w'), indent=1)
    print('wrote', out, 'flops per lane and row:', flops)
