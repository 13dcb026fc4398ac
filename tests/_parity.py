"""Shared parity helpers of the GPU tests: the stated fp16 tolerance and the comparison against a reference run.

Tolerance (calibrated on the golden vectors): eps0 = max |log p_fp16 - log p_fp32| between the fp16
and fp32 reference runs over tokens with p >= 1e-4 is 0.015-0.016; the CUDA path must stay within
LOGP_TOL = max(2*eps0, 5e-2) = 5e-2 of the fp32 reference on those tokens, and give the same ranking of options /
same top-1 token wherever the reference's own margin exceeds 2*LOGP_TOL."""
import types

import numpy as np

LOGP_TOL = 5e-2
P_MIN = 1e-4
ANSWER_TOKEN = 5


def make_args(model_path, **kw):
    d = dict(model_path=model_path, num_batch=1, layer_num_per_shard=1, storage_location="gpu",
             max_activation_in_cpu=100, data_parallel=False, disk_folder="./temp", num_gen_token=1,
             resident_weights="never")
    d.update(kw)
    return types.SimpleNamespace(**d)


def check_against_reference(out, want32):
    got = np.stack([o[:, 0] for o in out])
    assert got.dtype == np.float16 and got.shape == want32.shape
    g, w = got.astype(np.float64), want32.astype(np.float64)
    m = w >= P_MIN
    dlog = np.abs(np.log(np.maximum(g[m], 1e-30)) - np.log(w[m]))
    assert dlog.max() <= LOGP_TOL, dlog.max()
    assert np.abs(g.sum(-1) - 1).max() <= 5e-3
    # top-1 token: identical wherever the reference's top-2 margin is rankable
    srt = np.sort(np.log(w), axis=-1)
    rankable = (srt[..., -1] - srt[..., -2]) > 2 * LOGP_TOL
    assert (g.argmax(-1)[rankable] == w.argmax(-1)[rankable]).all()
    # option ranking on the answer token
    gl, wl = np.log(np.maximum(g[..., ANSWER_TOKEN], 1e-30)), np.log(w[..., ANSWER_TOKEN])
    for p in range(w.shape[0]):
        order = np.argsort(-wl[p])
        gaps = -np.diff(wl[p][order])
        if (gaps > 2 * LOGP_TOL).all():
            assert (np.argsort(-gl[p]) == order).all()
    return dlog.max()
