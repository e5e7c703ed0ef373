import json, sys, glob
for f in sorted(glob.glob('gpurun_out/sweep_*.log')):
    lines = open(f).read().strip().splitlines()
    for l in lines:
        if l.startswith('{'):
            d = json.loads(l); print(f, lines[-1], '| step', round(d['ms_per_step'], 3), 'fused', round(d['roofline']['segments_ms']['fused'], 3), 'loss', d['config']['loss'])
