import json,sys
for f in sys.argv[1:]:
    for l in open(f):
        l=l.strip()
        if not l.startswith('{'): continue
        d=json.loads(l); b=d.get('batched',{})
        print(f, 'C2 p50', d.get('p50_latency_ms'), 'e2e', d['e2e'].get('p50_latency_ms'), 'C5 ms', b.get('ms_per_step'), 'frac', b.get('roofline',{}).get('frac'), 'val', b.get('value'))
