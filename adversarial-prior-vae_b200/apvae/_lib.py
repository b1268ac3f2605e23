"""ctypes binding of libapvae_b200.so (the C ABI in include/apvae.h).

There is no CPU fallback: if the shared library is missing, importing this
module raises, and every compute entry point fails with APVAE_ERR_NO_DEVICE
when no sm_100 GPU is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'libapvae_b200.so')

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_WEIGHTS, ERR_UNSUPPORTED, ERR_STATE = range(7)
ARCH_A, ARCH_B, ARCH_M = 0, 1, 2
PER_IMAGE_ROWS = 10
(ROW_LL, ROW_LB, ROW_ADV_LL, ROW_ADV_LB, ROW_RECON, ROW_DG, ROW_LZ_MAX, ROW_LZ_SUM,
 ROW_REAL_E, ROW_MSE) = range(10)
PER_SAMPLE_ROWS = 5
SROW_LOG_PX, SROW_LOG_Q, SROW_LOG_PN, SROW_DG, SROW_SQ_ERR = range(5)
RES_NONE, RES_SAME, RES_EVEN = 0, 1, 2
ACCUM_WEIGHT, ACCUM_BATCHES, ACCUM_MSE = 9, 10, 11
ACCUM_PN_ENERGY, ACCUM_PD_ENERGY, ACCUM_KL_ADV, ACCUM_LOG_Z_OP, ACCUM_PRIOR_WEIGHT = 12, 13, 14, 15, 16
ACCUM_SLOTS = 20
PRIOR_PN_ENERGY, PRIOR_PD_ENERGY, PRIOR_KL_ADV, PRIOR_LOG_Z_OP = range(4)
AUX_COUNT = 11
NOALLOC_MAX_SAMPLES = 65536


class ApvaeError(RuntimeError):
    def __init__(self, status, message):
        super(ApvaeError, self).__init__('apvae status %d: %s' % (status, message))
        self.status = status


TUNING_FIELDS = ('patch', 'patch_macc', 'patch_cluster', 'patch_max_n', 'patch_phases', 'patch_merge', 'patch_s2',
                 'patch_fold', 'patch_tpi', 'patch_2cta', 'patch_2cta_min_k', 'patch_a_cap', 'patch_a_stages', 'patch_b_stages',
                 'tap_2cta', 'fast_epi', 'fast_mask', 'tiny_conv', 'small_block', 'heads_gemm', 'patch_epi2', 'patch_split', 'debug', 'trace', 'patch_virtual')


class Tuning(C.Structure):
    """apvae_tuning: kernel-selection overrides; all-zero = defaults, switch fields -1 = off / +1 = on."""
    _fields_ = [(n, C.c_int) for n in TUNING_FIELDS]

    @classmethod
    def make(cls, options=None):
        t = cls()
        for k, v in (options or {}).items():
            if k not in TUNING_FIELDS:
                raise AttributeError('apvae_tuning has no field %r' % (k,))
            setattr(t, k, int(v))
        return t


class Config(C.Structure):
    _fields_ = [
        ('arch', C.c_int), ('x_h', C.c_int), ('x_w', C.c_int), ('x_c', C.c_int),
        ('z_dim', C.c_int), ('celeba', C.c_int),
        ('pull_back_energy_weight', C.c_float), ('min_logstd_of_q', C.c_float),
        ('epsilon', C.c_float),
        ('uniform_scale', C.c_int), ('use_truncated', C.c_int), ('bernoulli', C.c_int),
        ('max_batch', C.c_int), ('max_chunk', C.c_int), ('enable_grad', C.c_int),
        ('smallest_step', C.c_float), ('len_train', C.c_int), ('precise', C.c_int),
        ('nf_layers', C.c_int), ('latent_gan', C.c_int), ('operand_bf16', C.c_int),
        ('tuning', Tuning),
    ]


class ConvTerm(C.Structure):
    _fields_ = [
        ('src', C.c_int), ('transpose', C.c_int), ('k', C.c_int), ('stride', C.c_int),
        ('cin', C.c_int), ('cout', C.c_int), ('in_h', C.c_int), ('in_w', C.c_int),
        ('in_cp', C.c_int), ('kernel', C.POINTER(C.c_float)), ('bias', C.POINTER(C.c_float)),
    ]


class PlanInfo(C.Structure):
    _fields_ = [
        ('kc', C.c_int), ('npad', C.c_int), ('nt', C.c_int), ('n_tiles', C.c_int),
        ('cout', C.c_int), ('num_phases', C.c_int), ('phase_begin', C.c_int * 5),
        ('total_steps', C.c_int),
        ('bw', C.c_int), ('bh', C.c_int), ('bn', C.c_int), ('rows', C.c_int),
        ('wg', C.c_int), ('hg', C.c_int), ('tiles_x', C.c_int), ('tiles_y', C.c_int),
        ('ho', C.c_int), ('wo', C.c_int), ('os', C.c_int), ('nsrc', C.c_int),
        ('capacity_n', C.c_int),
        ('src_dims', (C.c_uint64 * 5) * 3), ('src_strides', (C.c_uint64 * 4) * 3),
        ('src_box', (C.c_uint32 * 5) * 3),
        ('macs_per_image', C.c_double),
        ('patch', C.c_int), ('patch_macc', C.c_int), ('patch_cluster', C.c_int),
        ('patch_rows_tile', C.c_int), ('patch_groups', C.c_int), ('patch_items', C.c_int),
        ('patch_a_stages', C.c_int), ('patch_b_stages', C.c_int), ('patch_prow', C.c_int * 2),
        ('patch_s2d', C.c_int), ('patch_nmaps', C.c_int), ('patch_map_src', C.c_int * 4), ('patch_map_ph', C.c_int * 4),
        ('patch_map_prow', C.c_int * 4), ('patch_phase', C.c_int), ('patch_out_scale', C.c_int), ('patch_npad', C.c_int),
        ('patch_fold', C.c_int), ('patch_epi2', C.c_int), ('patch_wv', C.c_int),
    ]


class Epilogue(C.Structure):
    _fields_ = [
        ('out_raw_dev', C.c_void_p), ('out_act_dev', C.c_void_p), ('ldo', C.c_int),
        ('bf16_zero_from', C.c_int),
        ('out_f32_dev', C.c_void_p), ('ld_f32', C.c_int), ('f32_cols', C.c_int),
        ('res_dev', C.c_void_p), ('res_mode', C.c_int), ('ld_res', C.c_int),
        ('mask_dev', C.c_void_p), ('ld_mask', C.c_int),
        ('tanh_cols', C.c_int),
        ('out_raw_lo_dev', C.c_void_p), ('out_act_lo_dev', C.c_void_p), ('res_lo_dev', C.c_void_p),
    ]


# every symbol include/apvae.h declares: (name, restype, argtypes)
_P = C.c_void_p
_SYMBOLS = [
    ('apvae_version', C.c_char_p, []),
    ('apvae_last_error', C.c_char_p, [_P]),
    ('apvae_default_config', C.c_int, [C.c_int, C.c_int, C.POINTER(Config)]),
    ('apvae_weight_count', C.c_int, [C.POINTER(Config), C.POINTER(C.c_int)]),
    ('apvae_weight_info', C.c_int, [C.POINTER(Config), C.c_int, C.c_char_p, C.c_int,
                                    C.POINTER(C.c_int64), C.POINTER(C.c_int)]),
    ('apvae_create', C.c_int, [C.POINTER(Config), C.c_int, C.POINTER(_P)]),
    ('apvae_destroy', C.c_int, [_P]),
    ('apvae_load_weights', C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(_P),
                                     C.POINTER(_P), C.POINTER(C.c_int)]),
    ('apvae_set_scalars', C.c_int, [_P, C.c_float, C.c_float, C.c_float]),
    ('apvae_encode', C.c_int, [_P, _P, C.c_int, _P, _P, _P]),
    ('apvae_decode', C.c_int, [_P, _P, C.c_int, _P, _P, _P]),
    ('apvae_energy', C.c_int, [_P, _P, _P, C.c_int, _P, _P]),
    ('apvae_is_nll', C.c_int, [_P, _P, _P, C.c_int, C.c_int, _P, C.c_uint64, _P, _P, _P]),
    ('apvae_sample_z', C.c_int, [_P, _P, C.c_int, C.c_int, _P, C.c_uint64, _P, _P, _P, _P]),
    ('apvae_accumulate_metrics', C.c_int, [_P, _P, _P, C.c_int, C.c_int, _P, _P]),
    ('apvae_allreduce_metrics', C.c_int, [_P, C.c_int, _P, _P, _P]),
    ('apvae_merge_metrics', C.c_int, [_P, C.c_int, _P, _P]),
    ('apvae_prior_stats', C.c_int, [_P, C.c_int, _P, C.c_uint64, _P, _P, _P, _P]),
    ('apvae_mala', C.c_int, [_P, _P, C.c_int, C.c_int, C.c_float, _P, _P, C.c_uint64, _P, _P, _P, _P, _P]),
    ('apvae_energy_grad', C.c_int, [_P, _P, C.c_int, _P, _P, _P, _P]),
    ('apvae_image_energy_grad', C.c_int, [_P, _P, C.c_int, _P, _P, _P, _P]),
    ('apvae_sample_images', C.c_int, [_P, _P, C.c_int, _P, _P]),
    ('apvae_latent_gan', C.c_int, [_P, _P, C.c_int, C.c_uint64, _P, _P]),
    ('apvae_discretized_logistic_log_prob', C.c_int, [_P, _P, _P, C.c_int64, C.c_int64, C.c_float, C.c_int, C.c_float,
                                                      C.c_int, C.c_float, C.c_int, C.c_float, _P, _P]),
    ('apvae_bernoulli_log_prob', C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P]),
    ('apvae_is_reduce', C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, _P, _P]),
    ('apvae_is_ll_noalloc', C.c_int, [_P, _P, _P, C.c_int64, C.c_float, C.c_int, C.c_float, C.c_int, C.c_float, C.c_int,
                                      C.c_float, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, _P, _P]),
    ('apvae_fill_normal', C.c_int, [_P, C.c_int64, C.c_uint64, C.c_uint64, C.c_int, _P]),
    ('apvae_fill_uniform', C.c_int, [_P, C.c_int64, C.c_uint64, C.c_uint64, _P]),
    ('apvae_launch_count', C.c_int, [_P, C.POINTER(C.c_int64)]),
    ('apvae_macs', C.c_int, [_P, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    ('apvae_op_count', C.c_int, [_P, C.c_int, C.POINTER(C.c_int)]),
    ('apvae_op_time', C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int,
                                C.POINTER(C.c_double), C.POINTER(C.c_float), _P]),
    ('apvae_aux_time', C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double),
                                 C.POINTER(C.c_float), _P]),
    ('apvae_plan_create', C.c_int, [C.POINTER(ConvTerm), C.c_int, C.c_int, C.c_int, C.POINTER(Tuning), C.POINTER(_P)]),
    ('apvae_plan_get_info', C.c_int, [_P, C.POINTER(PlanInfo)]),
    ('apvae_plan_copy_steps', C.c_int, [_P, C.POINTER(C.c_int32)]),
    ('apvae_plan_copy_weights', C.c_int, [_P, C.POINTER(C.c_uint16)]),
    ('apvae_plan_copy_bias', C.c_int, [_P, C.POINTER(C.c_float)]),
    ('apvae_plan_copy_patch_groups', C.c_int, [_P, C.POINTER(C.c_int32)]),
    ('apvae_plan_copy_patch_weights', C.c_int, [_P, C.POINTER(C.c_uint16)]),
    ('apvae_plan_copy_patch_bias', C.c_int, [_P, C.POINTER(C.c_float)]),
    ('apvae_plan_select_patch', C.c_int, [_P, C.c_int]),
    ('apvae_plan_destroy', C.c_int, [_P]),
    ('apvae_plan_run', C.c_int, [_P, C.c_int, C.POINTER(_P), C.c_int, C.POINTER(Epilogue)]),
    ('apvae_plan_bench', C.c_int, [_P, C.c_int, C.POINTER(_P), C.c_int, C.POINTER(Epilogue), C.c_int,
                                   C.POINTER(C.c_float)]),
]
SYMBOL_NAMES = [s[0] for s in _SYMBOLS]

if not os.path.exists(LIB_PATH):
    raise ImportError(
        'libapvae_b200.so not found at %s: build it with `python -c "import '
        '__graft_entry__ as g; g.build()"` (or make -C adversarial-prior-vae_b200/csrc). '
        'There is no CPU fallback.' % LIB_PATH)

lib = C.CDLL(LIB_PATH)
for _name, _res, _args in _SYMBOLS:
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


def last_error(handle=None):
    msg = lib.apvae_last_error(handle)
    return msg.decode('utf-8', 'replace') if msg else ''


def check(status, handle=None):
    if status != OK:
        raise ApvaeError(status, last_error(handle) or last_error(None))
