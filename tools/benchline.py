import json,sys
for l in sys.stdin:
    l=l.strip()
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']
    print(sys.argv[1], 'prof_sum_ms_per_step', round(sum(0 for _ in [0]),1) if 'profile_ms' not in r else round(sum(r['profile_ms'].values())/d['steps'],2), 'value',d['value'],'ms/step',d['ms_per_step'],'e2e',d['e2e']['value'],'dom',r['kernel'],r['kernel_ms_per_launch'],'frac',r['frac'],'step_frac',r['whole_step_frac'], {k:v for k,v in r['kernel_share_of_step'].items() if v>0.02})
