import sys, json
tag = sys.argv[1]
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print(tag, 'fused_ms', round(r['kernel_ms'], 3), 'partial_ms', round(r.get('partial_stage_ms', 0), 3), 'frac', round(r['frac'], 3),
              'step_ms', round(d['ms_per_step'], 3), d['config']['agg_path'])
    elif 'rror' in l:
        print(l.strip()[:200])
