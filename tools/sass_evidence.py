"""SASS evidence: per kernel of libapvae_b200.so, how many tcgen05 / TMEM / TMA instructions the sm_100a code holds.
  python tools/sass_evidence.py > profiles/r2_sass_evidence.txt
UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld (TMEM -> registers), UTMALDG / UTMASTG = bulk tensor load / store (TMA),
UTCBAR = tcgen05.commit, SYNCS = mbarrier operations (B200_PROFILING.md's mnemonics)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, 'adversarial-prior-vae_b200', 'libapvae_b200.so')
OPS = ('UTCHMMA', 'LDTM', 'UTMALDG', 'UTMASTG', 'UTCBAR', 'UTMAPF', 'SYNCS')


def main():
    sass = subprocess.run(['cuobjdump', '-sass', SO], capture_output=True, text=True, check=True).stdout
    counts, order, fn = collections.defaultdict(collections.Counter), [], None
    for line in sass.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            fn = m.group(1)
            order.append(fn)
            continue
        m = re.search(r'^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
        if m and fn:
            op = m.group(1)
            if op in OPS:
                counts[fn][op] += 1
            counts[fn]['_instructions'] += 1
    names = subprocess.run(['c++filt'], input='\n'.join(order), capture_output=True, text=True).stdout.splitlines()
    print('%s: %d kernels (cuobjdump -sass, sm_100a)' % (os.path.basename(SO), len(order)))
    print('%-110s %8s %s' % ('kernel', 'instr', ' '.join('%8s' % o for o in OPS)))
    tot = collections.Counter()
    for fn, name in zip(order, names):
        c = counts[fn]
        name = name.rsplit('(', 1)[0].replace('void ', '', 1) if name.endswith(')') else name
        print('%-110s %8d %s' % (name[:110], c['_instructions'], ' '.join('%8d' % c[o] for o in OPS)))
        tot.update({o: c[o] for o in OPS})
    print('%-110s %8s %s' % ('TOTAL', '', ' '.join('%8d' % tot[o] for o in OPS)))


if __name__ == '__main__':
    main()
