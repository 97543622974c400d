"""One-line digest of a bench.py JSON line (stdin); argv[1] = label."""
import json
import sys
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'):
        continue
    d = json.loads(l)
    r = d['roofline']
    sim = d.get('sim', {})
    print(sys.argv[1], 'value', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'], 'sim', sim.get('value'),
          'sim_ms', sim.get('ms_per_step'), 'dom', r['kernel'][:12], r['kernel_ms_per_launch'], 'frac', r['frac'],
          'step_frac', r.get('whole_step_frac'), 'iters', d.get('step', {}).get('mean_iterations'),
          {k: v for k, v in r['kernel_share_of_step'].items() if v > 0.02},
          'sim_prof', {k: v for k, v in sim.get('profile_ms_one_step', {}).items() if v > 1.0})
