"""Generates tests/golden/reference_kats.npz from the REFERENCE's own known-answer tests.

TensorFlow 1.x cannot be imported here, so the reference's test modules cannot be run as
they are.  What CAN be executed is their pure-numpy answer code: this script pulls the
helper functions `safe_sigmoid` and `naive_discretized_logistic_pdf` out of
/root/reference/code/tfsnippet/tests/distributions/test_discretized.py with `ast`, executes
THAT source (nothing is copied into the repo), rebuilds the seeded payloads exactly as the
reference tests do (same np.random.seed and draw order) and stores inputs + the reference's
expected answers.  Run once in the authoring container (the reference tree does not exist on
the GPU box); the .npz is committed.

  python tests/golden/make_golden.py
"""
import ast
import os

import numpy as np

REF = '/root/reference/code/tfsnippet/tests'
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'reference_kats.npz')


def load_reference_functions(path, names):
    with open(path) as f:
        tree = ast.parse(f.read())
    ns = {'np': np}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, 'exec'), ns)
    return [ns[n] for n in names]


def main():
    safe_sigmoid, naive_pdf = load_reference_functions(
        os.path.join(REF, 'distributions', 'test_discretized.py'),
        ['safe_sigmoid', 'naive_discretized_logistic_pdf'])
    out = {}

    # test_discretized.py:232-310  test_log_prob payload (seed 1234, same draw order)
    np.random.seed(1234)
    x = np.random.normal(size=[7, 3, 2, 5, 4]).astype(np.float64)
    mean = 3 * np.random.uniform(size=[2, 1, 4]).astype(np.float64) - 1
    log_scale = np.random.normal(size=[3, 1, 5, 1]).astype(np.float64)
    bin_size, min_val, max_val = 1 / 256., -1., 2.
    out.update(dl_x=x, dl_mean=mean, dl_log_scale=log_scale,
               dl_params=np.array([bin_size, min_val, max_val]))
    out['dl_unbiased'] = naive_pdf(x, mean, log_scale, bin_size, min_val, max_val, biased_edges=False, group_ndims=0)
    out['dl_nomin'] = naive_pdf(x, mean, log_scale, bin_size, None, max_val, biased_edges=True, group_ndims=0)
    out['dl_nomax'] = naive_pdf(x, mean, log_scale, bin_size, min_val, None, biased_edges=True, group_ndims=0)
    out['dl_biased'] = naive_pdf(x, mean, log_scale, bin_size, min_val, max_val, biased_edges=True, group_ndims=0)
    out['dl_biased_g2'] = naive_pdf(x, mean, log_scale, bin_size, min_val, max_val, biased_edges=True, group_ndims=2)

    # test_discretized.py:312-334  test_prob_sum payload (probabilities over the 11 bins sum to 1)
    np.random.seed(1234)
    out['ps_x'] = np.reshape(np.linspace(0, 1, 11), [-1, 1, 1, 1, 1])
    out['ps_mean'] = np.random.uniform(size=[2, 1, 4]).astype(np.float64)
    out['ps_log_scale'] = np.random.normal(size=[3, 1, 5, 1]).astype(np.float64)

    # test_discretized.py:336-363  test_log_prob_extreme (the PixelCNN++ fallback branch, epsilon=1e-7)
    s = np.linspace(0, 20, 101)
    t = bin_size / (2 * np.exp(s))
    idx = np.where(safe_sigmoid(t) - safe_sigmoid(-t) < 1e-8)[0]
    out['ex_log_scale'] = s[idx]
    with np.errstate(divide='ignore'):
        out['ex_naive'] = naive_pdf(np.asarray(0.), np.asarray(0.), s[idx], bin_size, None, None,
                                    biased_edges=True, group_ndims=0)

    # ops/test_misc.py:39-87  log_sum_exp / log_mean_exp on linspace(0,10,1000).reshape(5,10,20)
    lx = np.linspace(0, 10, 1000).reshape([5, 10, 20])
    out['lse_x'] = lx
    out['lse_last'] = np.log(np.sum(np.exp(lx), axis=-1))
    out['lse_02'] = np.log(np.sum(np.exp(lx), axis=(0, 2)))
    out['lse_all'] = np.log(np.sum(np.exp(lx)))
    out['lme_last'] = np.log(np.mean(np.exp(lx), axis=-1))
    out['lme_02'] = np.log(np.mean(np.exp(lx), axis=(0, 2)))
    out['lme_all'] = np.log(np.mean(np.exp(lx)))

    # variational/test_evaluation.py:9-47 and test_objectives.py  payload (seed 1234)
    np.random.seed(1234)
    log_p = np.random.normal(size=[13]).astype(np.float32)
    log_q = np.random.normal(size=[7, 13]).astype(np.float32)
    out['vi_log_p'], out['vi_log_q'] = log_p, log_q
    d = (log_p - log_q).astype(np.float64)
    out['vi_is_ll'] = np.log(np.mean(np.exp(d), axis=0))
    out['vi_elbo'] = np.mean(d, axis=0)

    # trainer/test_evaluator.py:53-103  arange(6), batch_size 4: weighted 2.5, unweighted 3.0
    out['ev_batches'] = np.array([np.mean(np.arange(6.)[:4]), np.mean(np.arange(6.)[4:])])
    out['ev_weights'] = np.array([4., 2.])
    out['ev_expected'] = np.array([2.5, 3.0])

    # layers/convolutional/test_conv2d.py:117-137  1x1 conv == strided slice + matmul (seed 1234, reduced batch)
    np.random.seed(1234)
    cx = np.random.normal(size=[3, 2, 8, 9, 5]).astype(np.float32)
    ck = np.random.random(size=[1, 1, 5, 7]).astype(np.float32)
    cb = np.random.random(size=[7]).astype(np.float32)
    out.update(c11_x=cx, c11_k=ck, c11_b=cb)
    out['c11_s1'] = np.dot(cx, ck.reshape(5, 7)) + cb
    out['c11_s2'] = np.dot(cx[..., ::2, ::2, :], ck.reshape(5, 7)) + cb

    np.savez_compressed(OUT, **out)
    print('wrote', OUT, {k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    main()
