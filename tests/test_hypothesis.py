"""Property tests with generated edge inputs (consequence (iii)):
clips assembled from the segments that stress the fixed-point path -- full-scale
constants, DC, digital silence, alternating +-32768 (FFT 16-bit wraps), loud noise (the
16-bit mel wrap), quiet noise (log2 of small values), steps and impulses.

CPU: the two independent restatements of the RTL (numpy, C) must agree bit for bit on
every stage.  GPU: the CUDA path must equal the C oracle bit for bit on batches of them."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import ref_c as RC
from oracle import ref_np as R

KINDS = ["max", "min", "dc", "zero", "alt", "loud", "quiet", "step", "impulse", "tone"]


def segment(kind, n, seed):
    rng = np.random.default_rng(seed)
    i = np.arange(n)
    if kind == "max":
        return np.full(n, 32767)
    if kind == "min":
        return np.full(n, -32768)
    if kind == "dc":
        return np.full(n, int(rng.integers(-3, 4)))
    if kind == "zero":
        return np.zeros(n, dtype=np.int64)
    if kind == "alt":
        return np.where((i + seed) & 1, 32767, -32768)
    if kind == "loud":
        return rng.integers(-32768, 32768, n)
    if kind == "quiet":
        return rng.integers(-6, 7, n)
    if kind == "step":
        return np.where(i < n // 2, -32768, 32767)
    if kind == "impulse":
        x = np.zeros(n, dtype=np.int64)
        x[int(rng.integers(0, n))] = int(rng.choice([-32768, 32767, 1, -1]))
        return x
    f = float(rng.uniform(20, 7900))
    return np.round(32767 * np.sin(2 * np.pi * f / 16000 * i + rng.uniform(0, 6.28)))


clip_plan = st.lists(st.tuples(st.sampled_from(KINDS), st.integers(1, 700), st.integers(0, 2 ** 31 - 1)),
                     min_size=1, max_size=6)


def build(plan):
    return np.concatenate([segment(k, n, s) for k, n, s in plan]).astype(np.int16)


@settings(max_examples=60, deadline=None, derandomize=True, database=None, suppress_health_check=[HealthCheck.too_slow])
@given(plan=clip_plan, nf=st.sampled_from([(32, 32), (32, 16), (16, 16)]))
def test_numpy_and_c_oracles_agree_on_edge_clips(plan, nf):
    x = build(plan)
    nfilt, ncep = nf
    K = max(1, min(4, R.num_frames(len(x)) + 1))              # include a zero-extended frame
    want, t = R.mfcc(x, nfilters=nfilt, nceptrums=ncep, n_frames=K, taps=True)
    cfg = RC.RefConfig(nfilters=nfilt, nceptrums=ncep)
    got, ct = RC.run(cfg, x, n_frames=K, taps=True)
    assert np.array_equal(got, want)
    assert np.array_equal(ct["fft"][..., 0], t["fft_r"]) and np.array_equal(ct["fft"][..., 1], t["fft_i"])
    for cname, name in (("window", "window"), ("power", "power"), ("mel", "filter"), ("log2", "log"), ("dct", "dct")):
        assert np.array_equal(np.asarray(ct[cname]).astype(np.int64), np.asarray(t[name]).astype(np.int64)), name


@pytest.mark.gpu
@settings(max_examples=12, deadline=None, derandomize=True, database=None, suppress_health_check=[HealthCheck.too_slow])
@given(plans=st.lists(clip_plan, min_size=1, max_size=5), nf=st.sampled_from([(32, 32), (16, 16)]))
def test_cuda_equals_oracle_on_edge_clips(plans, nf):
    import torch
    import mfcc_b200
    nfilt, ncep = nf
    clips = [build(p) for p in plans]
    n = max(len(c) for c in clips)
    x = np.zeros((len(clips), n), dtype=np.int16)
    for i, c in enumerate(clips):
        x[i, :len(c)] = c
    cfg = RC.RefConfig(nfilters=nfilt, nceptrums=ncep)
    K = cfg.num_frames(n) + 2
    op = _op(nfilt, ncep)
    got = op(torch.from_numpy(x).to("cuda:0"), n_frames=K).cpu().numpy()
    want = RC.run(cfg, x, n_frames=K)
    assert np.array_equal(got, want)


_ops = {}


def _op(nfilt, ncep):
    import mfcc_b200
    if (nfilt, ncep) not in _ops:
        _ops[(nfilt, ncep)] = mfcc_b200.MFCC(nfilters=nfilt, nceptrums=ncep, device="cuda:0")
    return _ops[(nfilt, ncep)]
