"""Stall samples of an .ncu-rep source page in ADDRESS order, one line per instruction above a threshold
(and every sync / barrier / tensor / TMA instruction as a landmark):
  python tools/ncu_regions.py rep launch-index [min-samples]"""
import csv, io, subprocess, sys
sel = ['--launch-skip', sys.argv[2], '--launch-count', '1']
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv'] + sel, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) >= len(hdr) - 1 and r[ix['# Samples']].isdigit()]
half = len(body) // 2 if len(body) > 2 and body[0][ix['Source']] == body[len(body) // 2][ix['Source']] else len(body)
body = body[:half]
thr = int(sys.argv[3]) if len(sys.argv) > 3 else 100
tot = sum(int(r[ix['# Samples']]) for r in body)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
marks = ('SYNCS', 'UTCHMMA', 'UTMALDG', 'UTCBAR', 'LDTM', 'BAR.', 'WARPSYNC', 'LDG', 'ELECT')
print('total samples', tot, 'instructions', len(body))
acc = 0
for n, r in enumerate(body):
    s = int(r[ix['# Samples']])
    acc += s
    src = r[ix['Source']].strip()
    if s >= thr or any(m in src for m in marks):
        top = sorted(((int(r[ix[h]] or 0), h) for h in stalls), reverse=True)[:2]
        print('%5d cum %5.1f%% %6d exec %8s  %-64s %s' % (n, 100.0 * acc / tot, s, r[ix['Instructions Executed']], src[:64],
                                                        ' '.join('%s=%d' % (h[6:], v) for v, h in top if v)))
