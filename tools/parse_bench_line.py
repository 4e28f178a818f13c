import sys, json
tag = sys.argv[1]
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        if d.get('orderkey'):
            o = d['orderkey']
            print(tag, 'orderkey partial/exchange/final ms', round(o['partial_ms'], 2), round(o['exchange_ms'], 2), round(o['final_ms'], 2),
                  'G rows/s', round(o['rows_per_s'] / 1e9, 2), o['agg_path'], 'verified', d.get('verified'))
        print(tag, 'fused_ms', round(r['kernel_ms'], 3), 'partial_ms', round(r.get('partial_stage_ms', 0), 3), 'frac', round(r['frac'], 3),
              'step_ms', round(d['ms_per_step'], 3), d['config']['agg_path'])
    elif 'rror' in l:
        print(l.strip()[:200])
