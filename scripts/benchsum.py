import sys, json
src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin      # never block on a terminal-less stdin by accident
for l in src:
    if l.startswith('{'):
        d = json.loads(l)
        print("ms/step", round(d['ms_per_step'], 2), "top", d['roofline']['kernel'], d['roofline']['frac'], "step frac", d['step_roofline']['frac'],
              "e2e", d['e2e'].get('ms_per_step'))
        tot = 0
        for k, v in d['kernels'].items():
            tot += v['calls_per_step'] * v['ms_per_call']
            print('  %-22s %4.1f x %7.3f ms  %s GB/s' % (k, v['calls_per_step'], v['ms_per_call'], v['achieved_gbs']))
        print('  our kernels total %.1f ms; other %.1f ms' % (tot, d['ms_per_step'] - tot))
    else:
        print(l.rstrip()[:200])
