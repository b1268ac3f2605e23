"""Tuning options for the tools: `--tuning patch_2cta=1,patch_macc=2` -> dict for ExpConfig(tuning=...) / Plan(tuning=...).
The library itself reads no environment variables (apvae_tuning, include/apvae.h); for the A/B scripts' convenience the
TOOLS also accept the old APVAE_* variable names from the environment and translate them here."""
import os

_ENV = {  # variable -> (field, kind)
    'APVAE_PATCH': ('patch', 'switch'), 'APVAE_PATCH_MACC': ('patch_macc', 'int'), 'APVAE_PATCH_CLUSTER': ('patch_cluster', 'int'),
    'APVAE_PATCH_MAXN': ('patch_max_n', 'int'), 'APVAE_PATCH_PHASES': ('patch_phases', 'switch'),
    'APVAE_PATCH_MERGE': ('patch_merge', 'switch'), 'APVAE_PATCH_S2': ('patch_s2', 'switch'), 'APVAE_PATCH_FOLD': ('patch_fold', 'switch'),
    'APVAE_PATCH_TPI': ('patch_tpi', 'int'), 'APVAE_PATCH_2CTA': ('patch_2cta', 'switch'), 'APVAE_PATCH_2CTA_MINK': ('patch_2cta_min_k', 'int'),
    'APVAE_PATCH_ACAP': ('patch_a_cap', 'int'), 'APVAE_PATCH_ASTAGES': ('patch_a_stages', 'int'), 'APVAE_PATCH_BSTAGES': ('patch_b_stages', 'int'),
    'APVAE_TAP_2CTA': ('tap_2cta', 'switch'), 'APVAE_FAST_EPI': ('fast_epi', 'switch'), 'APVAE_FAST_MASK': ('fast_mask', 'switch'),
    'APVAE_TINY': ('tiny_conv', 'switch'), 'APVAE_SMALL_BLOCK': ('small_block', 'switch'), 'APVAE_HEADS_GEMM': ('heads_gemm', 'switch'),
    'APVAE_PATCH_EPI2': ('patch_epi2', 'switch'), 'APVAE_PATCH_DEBUG': ('debug', 'int'), 'APVAE_PATCH_TRACE': ('trace', 'switch'),
}


def parse(text=None, env=True):
    out = {}
    if env:
        for var, (field, kind) in _ENV.items():
            v = os.environ.get(var)
            if v is None or v == '':
                continue
            out[field] = (1 if int(v) else -1) if kind == 'switch' else int(v)
    for item in (text or '').split(','):
        if item.strip():
            k, v = item.split('=')
            out[k.strip()] = int(v)
    return out
