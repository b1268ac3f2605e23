"""Top stalled SASS instructions of an .ncu-rep source page (read here).  python tools/ncu_hot.py rep [N] [launch-index]"""
import csv, io, subprocess, sys
sel = ['--launch-skip', sys.argv[3], '--launch-count', '1'] if len(sys.argv) > 3 else []
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv'] + sel, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) >= len(hdr) - 1 and r[ix['# Samples']].isdigit()]
tot = sum(int(r[ix['# Samples']] or 0) for r in body)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
print('total samples', tot)
for n, r in sorted(enumerate(body), key=lambda t: -int(t[1][ix['# Samples']] or 0))[:N]:
    s = int(r[ix['# Samples']] or 0)
    top = sorted(((int(r[ix[h]] or 0), h) for h in stalls), reverse=True)[:2]
    print('%5.1f%% line %4d exec %8s  %-70s %s' % (100.0 * s / tot, n, r[ix['Instructions Executed']], r[ix['Source']].strip()[:70],
                                              ' '.join('%s=%d' % (h[6:], v) for v, h in top if v)))
