"""Runs a fixed list of layer cases (the hot layers of arch A at the bench chunk size) in ONE process, so
a single ncu invocation can capture all of them:  python tools/layer_suite.py [--reps R] [--only substr]
Each case = (label, environment overrides, bench_layer arguments)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_layer  # noqa: E402

N = '1000'
L192 = ['--k', '3', '--h', '16', '--w', '16', '--cin', '192', '--cout', '192', '--n', N]
L96_5 = ['--transpose', '--k', '5', '--h', '32', '--w', '32', '--cin', '96', '--cout', '96', '--n', N]
L96_3 = ['--transpose', '--k', '3', '--h', '32', '--w', '32', '--cin', '96', '--cout', '96', '--n', N]
L192_S2 = ['--transpose', '--k', '3', '--stride', '2', '--h', '16', '--w', '16', '--cin', '192', '--cout', '96', '--n', N]
D00 = ['--k', '3', '--h', '32', '--w', '32', '--cin', '3', '--cout', '3', '--n', N, '--outs', 'act']
D01 = ['--k', '3', '--h', '32', '--w', '32', '--cin', '3', '--cout', '16', '--n', N]
D10 = ['--k', '3', '--h', '32', '--w', '32', '--cin', '16', '--cout', '16', '--n', N, '--outs', 'act']
D20 = ['--k', '3', '--h', '32', '--w', '32', '--cin', '32', '--cout', '32', '--n', N, '--outs', 'act']
D30 = ['--k', '3', '--h', '32', '--w', '32', '--cin', '64', '--cout', '64', '--n', N, '--outs', 'act']
L128 = ['--k', '3', '--h', '16', '--w', '16', '--cin', '128', '--cout', '128', '--n', N, '--outs', 'act']
CASES = [
    ('pair_D30_1cta', {'APVAE_PATCH_2CTA': '0'}, D30),
    ('pair_D30_2cta', {'APVAE_PATCH_2CTA': '1'}, D30),
    ('pair_D40_1cta', {'APVAE_PATCH_2CTA': '0'}, L128),
    ('pair_D40_2cta', {'APVAE_PATCH_2CTA': '1'}, L128),
    ('pair_96x3_1cta', {'APVAE_PATCH_2CTA': '0'}, L96_3 + ['--res', 'even']),
    ('pair_96x3_2cta', {'APVAE_PATCH_2CTA': '1'}, L96_3 + ['--res', 'even']),
    ('tap192', {'APVAE_PATCH': '0'}, L192),
    ('tap192_res', {'APVAE_PATCH': '0'}, L192 + ['--res', 'same']),
    ('patch192', {}, L192),
    ('patch192_res', {}, L192 + ['--res', 'same']),
    ('patch96x5', {}, L96_5),
    ('patch96x5_res', {}, L96_5 + ['--res', 'same']),
    ('patch96x5_tpi1', {'APVAE_PATCH_TPI': '1'}, L96_5),
    ('patch96x5_macc1', {'APVAE_PATCH_MACC': '1'}, L96_5),
    ('patch96x3_even', {}, L96_3 + ['--res', 'even']),
    ('patch96x3', {}, L96_3),
    ('tap96x3_even', {'APVAE_PATCH': '0'}, L96_3 + ['--res', 'even']),
    ('tap192s2', {'APVAE_PATCH_PHASES': '0'}, L192_S2 + ['--outs', 'act']),
    ('phase192s2', {}, L192_S2 + ['--outs', 'act']),
    ('patch192_2cta', {'APVAE_PATCH_2CTA': '1'}, L192),
    ('patch192_res_2cta', {'APVAE_PATCH_2CTA': '1'}, L192 + ['--res', 'same']),
    ('patch96x5_2cta', {'APVAE_PATCH_2CTA': '1'}, L96_5),
    ('patch96x5_res_2cta', {'APVAE_PATCH_2CTA': '1'}, L96_5 + ['--res', 'same']),
    ('patch96x3_even_2cta', {'APVAE_PATCH_2CTA': '1'}, L96_3 + ['--res', 'even']),
    ('patch192_c1', {'APVAE_PATCH_CLUSTER': '1'}, L192),
    ('nar_D00_a4', {'APVAE_PATCH_ACAP': '4'}, D00),
    ('nar_D00_a8', {'APVAE_PATCH_ACAP': '8'}, D00),
    ('nar_D01_a4', {'APVAE_PATCH_ACAP': '4'}, D01),
    ('nar_D01_a8', {'APVAE_PATCH_ACAP': '8'}, D01),
    ('nar_D10_a4', {'APVAE_PATCH_ACAP': '4'}, D10),
    ('nar_D10_a8', {'APVAE_PATCH_ACAP': '8'}, D10),
    ('nar_D20_a4', {'APVAE_PATCH_ACAP': '4'}, D20),
    ('nar_D20_a8', {'APVAE_PATCH_ACAP': '8'}, D20),
    ('nar_D30_a4', {'APVAE_PATCH_ACAP': '4'}, D30),
    ('nar_D30_a8', {'APVAE_PATCH_ACAP': '8'}, D30),
    ('nar_96x3_a8', {'APVAE_PATCH_ACAP': '8'}, L96_3 + ['--res', 'even']),
    ('nar_96x3_a4', {'APVAE_PATCH_ACAP': '4'}, L96_3 + ['--res', 'even']),

    ('patch192_c2', {'APVAE_PATCH_CLUSTER': '2'}, L192),
    ('patch96x5_c1', {'APVAE_PATCH_CLUSTER': '1'}, L96_5),
    ('patch96x5_c2', {'APVAE_PATCH_CLUSTER': '2'}, L96_5),
    ('patch96x5_res_c1', {'APVAE_PATCH_CLUSTER': '1'}, L96_5 + ['--res', 'same']),
    ('patch96x3_even_c1', {'APVAE_PATCH_CLUSTER': '1'}, L96_3 + ['--res', 'even']),
]
# tuning probes: APVAE_PATCH_DEBUG skips loads (results are garbage, timing only)
for _d, _nm in ((1, 'noload'), (2, 'noB'), (3, 'noA')):
    for _c in (1, 2):
        CASES.append(('dbg192_%s_c%d' % (_nm, _c), {'APVAE_PATCH_DEBUG': str(_d), 'APVAE_PATCH_CLUSTER': str(_c)}, L192))
    CASES.append(('dbg96x5_%s' % _nm, {'APVAE_PATCH_DEBUG': str(_d)}, L96_5))
for _b in (2, 3):
    CASES.append(('dbg192_bst%d' % _b, {'APVAE_PATCH_BSTAGES': str(_b)}, L192))
for _d, _nm in ((4, 'noepi'), (5, 'nostore'), (14, 'noload_noepi'), (15, 'noload_nostore')):
    CASES.append(('dbx192_%s' % _nm, {'APVAE_PATCH_DEBUG': str(_d)}, L192))
    CASES.append(('dbx96x5_%s' % _nm, {'APVAE_PATCH_DEBUG': str(_d)}, L96_5))
CASES.append(('dbg192_ast2', {'APVAE_PATCH_ASTAGES': '2'}, L192))
for _nm, _a in (('D00', D00), ('D01', D01), ('D10', D10), ('D20', D20)):
    CASES.append(('dbn_%s_base' % _nm, {}, _a))
    for _d, _dn in ((4, 'noepi'), (5, 'nostore'), (1, 'noload'), (15, 'noload_nostore')):
        CASES.append(('dbn_%s_%s' % (_nm, _dn), {'APVAE_PATCH_DEBUG': str(_d)}, _a))
for _n in (125, 250, 500, 1000, 2000, 4000):
    _a = ['--k', '3', '--h', '32', '--w', '32', '--cin', '16', '--cout', '16', '--n', str(_n), '--outs', 'act']
    CASES.append(('dbs_n%d_base' % _n, {}, _a))
    CASES.append(('dbs_n%d_noepi' % _n, {'APVAE_PATCH_DEBUG': '14'}, _a))
    CASES.append(('dbs_n%d_nomma' % _n, {'APVAE_PATCH_DEBUG': '16'}, _a))
for _m in (1, 2):
    _a = ['--k', '3', '--h', '32', '--w', '32', '--cin', '16', '--cout', '16', '--n', '1000', '--outs', 'act']
    CASES.append(('dbs_macc%d_base' % _m, {'APVAE_PATCH_MACC': str(_m)}, _a))
    CASES.append(('dbs_macc%d_noepi' % _m, {'APVAE_PATCH_MACC': str(_m), 'APVAE_PATCH_DEBUG': '14'}, _a))


def main():
    reps, only = '5', None
    av = sys.argv[1:]
    if '--reps' in av:
        reps = av[av.index('--reps') + 1]
    if '--only' in av:
        only = av[av.index('--only') + 1].split(',')
    for label, env, args in CASES:
        if (only and not any(label.startswith(o) for o in only)) or (not only and label.startswith('db')):
            continue
        saved = {k: os.environ.get(k) for k in env}
        os.environ.update(env)
        sys.stdout.write('%-16s ' % label)
        sys.stdout.flush()
        bench_layer.main(args + ['--reps', reps])
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


if __name__ == '__main__':
    main()
