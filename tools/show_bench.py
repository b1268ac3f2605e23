"""Pretty-prints a bench.py JSON line: python tools/show_bench.py gpurun_out/bench.log [--ops]"""
import json, sys
for path in [a for a in sys.argv[1:] if not a.startswith('--')]:
    l = [x for x in open(path) if x.startswith('{')]
    if not l:
        print(path, open(path).read()[-1500:]); continue
    d = json.loads(l[-1])
    r = d.get('roofline') or {}
    print('%s: value %.1f img/s  e2e %.1f  step %.1f TF (%.3f of sustained)  dom-op %.1f TF  all-ops %.1f TF  launches %s clocks %s' % (
        path, d['value'], d['e2e']['value'], d['step_tensor']['algorithmic_tflops_per_gpu'], d['step_tensor']['frac_of_sustained_peak'],
        r.get('achieved', 0), r.get('all_ops_tflops', 0), d.get('gpu_launches'), d.get('clocks')))
    if '--ops' in sys.argv and d.get('ops'):
        for o in d['ops']:
            if o['share'] is not None:
                print('   %-34s n=%d ms=%.4f tf=%.1f share=%.3f' % (o['op'], o['units'], o['ms'], o['tflops'], o['share']))
