"""GPU parity tests (run on the B200 box): the CUDA path through the C ABI against the
CPU oracle, bit for bit (integer work: no tolerance)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def RC():
    from oracle import ref_c
    return ref_c


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import mfcc_b200
    return {(nf, nc): mfcc_b200.MFCC(nfilters=nf, nceptrums=nc) for nf, nc in
            [(32, 32), (32, 16), (16, 16)]}


def dev(a, dtype=np.int16):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=dtype)).cuda()


def u16(t):
    return t.cpu().numpy().view(np.uint16)


def edge_clips(n=4000, seed=0):
    rng = np.random.default_rng(seed)
    t = np.arange(n)
    return {
        "noise_full": rng.integers(-32768, 32768, n),
        "gauss2000": np.clip(rng.normal(0, 2000, n), -32768, 32767),
        "tones": 8000 * sum(np.sin(2 * np.pi * f * t / 16000) for f in (440.0, 1234.5, 5100.0)) / 3
                 + rng.normal(0, 50, n),
        "gated": np.clip(rng.normal(0, 2000, n), -32768, 32767) * ((t // 800) % 2),
        "zero": np.zeros(n), "max": np.full(n, 32767), "min": np.full(n, -32768),
        "alt": np.tile([-32768, 32767], n // 2), "dc1": np.ones(n),
        "impulse": (t == 777) * 32767, "alt4": np.tile([32767, 32767, -32768, -32768], n // 4),
    }


def test_wav_whole_chain_and_taps(ops, RC, wav):
    m = ops[(32, 32)]
    cfg = RC.RefConfig(nfilters=32, nceptrums=32)
    want, tw = RC.run(cfg, wav, taps=True)
    got, tg = m(dev(wav), taps=True)
    torch.cuda.synchronize()
    assert np.array_equal(tg["window"].cpu().numpy(), tw["window"])
    assert np.array_equal(tg["fft"].cpu().numpy(), tw["fft"])
    assert np.array_equal(tg["power"].cpu().numpy().view(np.uint32), tw["power"])
    assert np.array_equal(u16(tg["filter"]), tw["mel"])
    assert np.array_equal(u16(tg["log"]), tw["log2"])
    assert np.array_equal(tg["dct"].cpu().numpy(), tw["dct"])
    assert got.shape == (1046, 32) and np.array_equal(got.cpu().numpy(), want)
    # and without taps (the production instantiation)
    assert np.array_equal(m(dev(wav)).cpu().numpy(), want)


@pytest.mark.parametrize("key", [(32, 32), (32, 16), (16, 16)])
def test_edge_clips_batched(ops, RC, key):
    m = ops[key]
    cfg = RC.RefConfig(nfilters=key[0], nceptrums=key[1])
    clips = edge_clips()
    x = np.stack([np.asarray(v, dtype=np.float64).astype(np.int16) for v in clips.values()])
    want = RC.run(cfg, x, threads=4)
    got = m(dev(x)).cpu().numpy()
    for i, name in enumerate(clips):
        assert np.array_equal(got[i], want[i]), name


@pytest.mark.parametrize("n", [0, 1, 511, 512, 513, 681, 682, 683, 1000, 4421, 4422, 4423, 4591, 4592, 4593,
                               5951, 5952, 5953, 6121, 6122, 8671, 8672, 8673, 11393, 20001])
def test_ragged_lengths_and_tile_boundaries(ops, RC, n):
    """n around nfft, the hop and the tile boundaries: a tile is 24 frames (4422 = 170*23+512 fills
    one, 4592 starts the second, 8672 the third); 5952 = 170*32+512 is the fp32 kernel's tile."""
    m = ops[(32, 32)]
    cfg = RC.RefConfig()
    rng = np.random.default_rng(n)
    x = rng.integers(-32768, 32768, n).astype(np.int16)
    want = RC.run(cfg, x)
    got = m(dev(x)).cpu().numpy() if n else np.zeros((0, 32), np.int16)
    assert got.shape == want.shape and np.array_equal(got, want)
    # host-protocol frame count: zero-extended tail (software/main.c:95, :135-144)
    k = m.cfg.num_frames_host(n)
    if n:
        want = RC.run(cfg, x, n_frames=k)
        assert np.array_equal(m(dev(x), n_frames=k).cpu().numpy(), want)


def test_unaligned_rows_take_the_plain_load_path(ops, RC):
    m = ops[(32, 32)]
    cfg = RC.RefConfig()
    rng = np.random.default_rng(5)
    big = rng.integers(-32768, 32768, (3, 7003)).astype(np.int16)   # odd pitch: rows not 16-B aligned
    want = RC.run(cfg, big)
    assert np.array_equal(m(dev(big)).cpu().numpy(), want)
    off = dev(np.concatenate([[0, 0, 0], big[0]]))[3:]              # base pointer off by 6 bytes
    assert np.array_equal(m(off).cpu().numpy(), want[0])


def test_many_clips_full_length(ops, RC):
    """C2-shaped: 60 s clips, the four synthetic classes of the benchmark."""
    m = ops[(32, 32)]
    cfg = RC.RefConfig()
    import bench
    x = bench.synth_clips(8, 960000, torch.device("cuda"), first_clip=0)
    got = m(x).cpu().numpy()
    want = RC.run(cfg, x.cpu().numpy(), threads=RC.max_threads())
    assert got.shape == (8, 5645, 32) and np.array_equal(got, want)
    # the loud class drives the 16-bit mel wrap (filterbank.py:122)
    _, tp = RC.run(cfg, x[0].cpu().numpy()[:40000], taps=True)
    assert (tp["mel"] > 0).any()


def test_large_batch_takes_the_72_frame_tiles(ops, RC):
    """Batches that fill the machine several times over run the 2 x 12-warp shape with 72-frame
    tiles (24 clips x 79 tiles here); ragged length, so the last tile of every clip is partial
    and the tail is read by plain loads."""
    m = ops[(32, 32)]
    cfg = RC.RefConfig()
    import bench
    x = bench.synth_clips(24, 960000, torch.device("cuda"), first_clip=0)[:, :960000 - 37].contiguous()
    got = m(x).cpu().numpy()
    want = RC.run(cfg, x.cpu().numpy(), threads=RC.max_threads())
    assert got.shape == want.shape and np.array_equal(got, want)
    # zero-extended frame count of the host protocol on the same batch
    k = m.cfg.num_frames_host(x.shape[1])
    assert np.array_equal(m(x, n_frames=k).cpu().numpy(), RC.run(cfg, x.cpu().numpy(), n_frames=k, threads=RC.max_threads()))


def test_run_host_equals_device_path(ops, RC):
    m = ops[(32, 16)]
    cfg = RC.RefConfig(nfilters=32, nceptrums=16)
    rng = np.random.default_rng(11)
    x = rng.integers(-20000, 20000, (37, 30001)).astype(np.int16)
    want = RC.run(cfg, x, threads=4)
    xh = torch.from_numpy(x).pin_memory()
    assert np.array_equal(m.run_host(xh).numpy(), want)
    assert np.array_equal(m.run_host(torch.from_numpy(x)).numpy(), want)     # pageable too


def test_run_host_pipeline_shapes(RC):
    """mfcc_b200_run_host: chunks smaller than the batch (every staging slot reused several
    times), 2..8 chunks in flight, rows packed (one plain copy per chunk) and pitched (2-D
    copies), sample counts that are not multiples of 8, library-allocated pinned buffers."""
    import ctypes as C
    import mfcc_b200
    from mfcc_b200 import _lib
    from mfcc_b200.mfcc import pinned_empty
    m = mfcc_b200.MFCC(nfilters=32, nceptrums=32)
    cfg = RC.RefConfig(nfilters=32, nceptrums=32)
    rng = np.random.default_rng(12)
    for n, B in ((40000, 61), (39997, 23)):
        x = rng.integers(-32768, 32768, (B, n)).astype(np.int16)
        want = RC.run(cfg, x, threads=4)
        K = want.shape[1]
        xh = pinned_empty((B, n), torch.int16, write_combined=(n % 8 == 0))
        xh.copy_(torch.from_numpy(x))
        oh = pinned_empty((B, K, 32), torch.int16)
        for mb, slots in ((0.25, 2), (0.25, 3), (0.5, 8), (32, 4)):
            m.host_config(mb, slots)
            oh.zero_()
            assert np.array_equal(m.run_host(xh, out=oh).numpy(), want), (n, mb, slots)
        # pitched rows on both sides, straight through the C ABI
        pitch, opitch = n + 24, K * 32 + 40
        xp = np.zeros((B, pitch), np.int16)
        xp[:, :n] = x
        op_ = np.full((B, opitch), -1, np.int16)
        m.host_config(0.25, 3)
        _lib.check(_lib.lib().mfcc_b200_run_host(m._h.ptr, C.c_void_p(xp.ctypes.data), pitch, n, B, K,
                                                 C.c_void_p(op_.ctypes.data), opitch))
        assert np.array_equal(op_[:, :K * 32].reshape(B, K, 32), want) and (op_[:, K * 32:] == -1).all()
    with pytest.raises(_lib.MfccError):
        m.host_config(1, 1)
    with pytest.raises(_lib.MfccError):
        m.host_config(1, 9)


def test_run_host_error_leaves_nothing_in_flight(RC):
    """A failure in the middle of the staging pipeline (injected after chunk 2's host->device copy
    has been queued) is reported only after all three streams have drained: the caller may free
    its buffers at once, and the handle keeps working."""
    import mfcc_b200
    from mfcc_b200 import _lib
    m = mfcc_b200.MFCC(nfilters=32, nceptrums=16)
    cfg = RC.RefConfig(nfilters=32, nceptrums=16)
    rng = np.random.default_rng(13)
    x = rng.integers(-32768, 32768, (40, 64000)).astype(np.int16)
    want = RC.run(cfg, x, threads=4)
    m.host_config(0.5, 3)                                   # 4 clips per chunk, 10 chunks
    xh = torch.from_numpy(x).pin_memory()
    oh = torch.full(want.shape, -7, dtype=torch.int16).pin_memory()
    _lib.check(_lib.lib().mfcc_b200_debug_fail_chunk(m._h.ptr, 2))
    with pytest.raises(_lib.MfccError, match="injected failure"):
        m.run_host(xh, out=oh)
    assert torch.cuda.current_stream().query()
    got = oh.numpy().copy()
    # chunks 0 and 1 are complete, nothing of chunk 2.. was written, and nothing arrives later
    assert np.array_equal(got[:8], want[:8]) and (got[8:] == -7).all()
    del xh
    torch.cuda.synchronize()
    assert np.array_equal(oh.numpy(), got)
    _lib.check(_lib.lib().mfcc_b200_debug_fail_chunk(m._h.ptr, -1))
    assert np.array_equal(m.run_host(torch.from_numpy(x)).numpy(), want)


@pytest.mark.parametrize("frames", [100, 200])
def test_many_short_clips(ops, RC, frames):
    """2000 short clips in one launch: 200 frames each take the 72-frame tiles (three per clip,
    the last one partial), 100 frames each stay on the 24-frame tiles (the large shape would
    leave its second tile mostly empty)."""
    m = ops[(32, 32)]
    cfg = RC.RefConfig()
    rng = np.random.default_rng(frames)
    n = 512 + 170 * (frames - 1) + 5
    x = rng.integers(-32768, 32768, (2000, n)).astype(np.int16)
    got = m(dev(x)).cpu().numpy()
    want = RC.run(cfg, x, threads=RC.max_threads())
    assert got.shape == (2000, frames, 32) and np.array_equal(got, want)


# ---- size-independent properties at full scale ---------------------------------------
def test_properties_at_scale(ops):
    m = ops[(32, 32)]
    import bench
    x = bench.synth_clips(64, 960000, torch.device("cuda"), first_clip=0)
    a = m(x)
    # determinism / idempotence
    assert torch.equal(a, m(x))
    # clips are independent: a permutation of the batch permutes the result
    perm = torch.randperm(64, device="cuda")
    assert torch.equal(m(x[perm].contiguous()), a[perm])
    # frame k depends only on samples 170k-1 .. 170k+511: a window cut out of a clip
    # (with its predecessor sample zeroed by the cut) agrees from its second frame on
    cut = x[3, 170 * 100:170 * 100 + 170 * 40 + 512].contiguous()
    b = m(cut)
    assert torch.equal(b[1:], a[3, 101:141])
    # silence -> log2(0) := 0 everywhere -> all-zero cepstra
    z = torch.zeros(5000, dtype=torch.int16, device="cuda")
    assert int(m(z).abs().max()) == 0


# ---- per-stage operators -------------------------------------------------------------
def test_stage_operators(ops, RC, wav):
    from oracle import ref_np as R
    m = ops[(32, 32)]
    x = wav[:30000]
    _, tp = R.mfcc(x, nfilters=32, nceptrums=32, taps=True)
    y = m.preemph(dev(x))
    assert np.array_equal(y.cpu().numpy(), tp["preemph"])
    fr = m.frame(y)
    assert np.array_equal(fr.cpu().numpy(), tp["frame"])
    assert np.array_equal(m.frame(y, n_frames=fr.shape[0] + 3).cpu().numpy()[-1, -170:], np.zeros(170))
    w = m.window(fr)
    assert np.array_equal(w.cpu().numpy(), tp["window"])
    f = m.fft_stream(w)
    assert np.array_equal(f.cpu().numpy()[..., 0], tp["fft_r"]) and np.array_equal(f.cpu().numpy()[..., 1], tp["fft_i"])
    p = m.powspec(f)
    assert np.array_equal(p.cpu().numpy().view(np.uint32), tp["power"])
    mel = m.filterbank(p)
    assert np.array_equal(u16(mel), tp["filter"])
    lg = m.log2(mel)
    assert np.array_equal(u16(lg), tp["log"])
    d = m.dct_stream(lg)
    assert np.array_equal(d.cpu().numpy(), tp["dct"])
    assert np.array_equal(m.discard(d).cpu().numpy(), tp["dct"][:, :32])
    # arbitrary (not chain-produced) inputs, including full-scale wrap cases
    rng = np.random.default_rng(2)
    fr2 = rng.integers(-32768, 32768, (50, 512)).astype(np.int16)
    fr2[0] = np.tile([-32768, 32767], 256)
    rr, ri = R.fft_stream(fr2)
    f2 = m.fft_stream(dev(fr2)).cpu().numpy()
    assert np.array_equal(f2[..., 0], rr) and np.array_equal(f2[..., 1], ri)
    pw = rng.integers(0, 1 << 30, (40, 256)).astype(np.uint32)
    assert np.array_equal(u16(m.filterbank(dev(pw.view(np.int32), np.int32))), R.filterbank(pw, ntap=32))
    allm = np.arange(65536, dtype=np.uint16)
    assert np.array_equal(u16(m.log2(dev(allm.view(np.int16)))), R.log2fix(allm))
    L = rng.integers(0, 32768, (64, 32)).astype(np.int16)
    assert np.array_equal(m.dct_stream(dev(L)).cpu().numpy(), R.dct_stream(L, 32))


def test_stage_operators_16_filters(ops):
    from oracle import ref_np as R
    m = ops[(16, 16)]
    rng = np.random.default_rng(4)
    pw = rng.integers(0, 1 << 30, (40, 256)).astype(np.uint32)
    assert np.array_equal(u16(m.filterbank(dev(pw.view(np.int32), np.int32))), R.filterbank(pw, ntap=16))
    L = rng.integers(0, 32768, (64, 16)).astype(np.int16)
    assert np.array_equal(m.dct_stream(dev(L)).cpu().numpy(), R.dct_stream(L, 16))


# ---- streaming -----------------------------------------------------------------------
def test_streaming_matches_batch(ops, RC):
    m = ops[(32, 16)]
    cfg = RC.RefConfig(nfilters=32, nceptrums=16)
    rng = np.random.default_rng(9)
    S, steps = 70, 12
    hops = rng.integers(-30000, 30000, (S, steps, 170)).astype(np.int16)
    m.reset(S)
    outs = []
    for t in range(steps):
        outs.append(m.step(dev(hops[:, t])).cpu().numpy())
    got = np.stack(outs, axis=1)
    # zero state = 342 zeros already pushed after reset
    x = np.concatenate([np.zeros((S, 342), np.int16), hops.reshape(S, -1)], axis=1)
    want = RC.run(cfg, x, threads=4)
    assert want.shape == got.shape and np.array_equal(got, want)


def test_errors(ops):
    import mfcc_b200
    m = ops[(32, 32)]
    with pytest.raises(TypeError):
        m(torch.zeros(1000, device="cuda"))
    with pytest.raises(ValueError):
        m(torch.zeros(1000, dtype=torch.int16))
    with pytest.raises(ValueError):
        mfcc_b200.MFCC(nfft=384)                         # not a power of two, fft.py:351-353
    with pytest.raises(ValueError):
        mfcc_b200.MFCC(nfilters=64, nceptrums=65)
    # other sizes construct (chain of general stage kernels, tests/test_rtl_golden.py)
    assert mfcc_b200.MFCC(nfft=256)._composed and not m._composed


@pytest.mark.gpu
def test_all_kernels_driver():
    """profiles/all_kernels.py launches every kernel of the library (both filter sets, both
    modes, streaming, stages, post-processing, the librosa set) with its own cross-checks."""
    import os
    import runpy
    runpy.run_path(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                "profiles", "all_kernels.py"), run_name="__main__")


@pytest.mark.gpu
def test_streaming_groups_and_pinned_host_buffers(ops):
    """step(first_stream=...) on groups of streams and step() on pinned host tensors
    (the kernel reads / writes them over PCIe itself) give the columns of a plain step."""
    import torch
    op = ops[(32, 32)]
    S = 300
    g = torch.Generator().manual_seed(4)
    hops = [torch.randint(-32768, 32768, (S, 170), dtype=torch.int16, generator=g) for _ in range(4)]
    op.reset(S)
    want = [op.step(h.to("cuda:0")).cpu() for h in hops]
    op.reset(S)
    for h, w in zip(hops, want):
        hd = h.to("cuda:0")
        out = torch.empty((S, 32), dtype=torch.int16, device="cuda:0")
        for a, b in ((0, 7), (7, 160), (160, S)):
            op.step(hd[a:b], out=out[a:b], first_stream=a)
        assert torch.equal(out.cpu(), w)
    op.reset(S)
    oh = torch.empty((S, 32), dtype=torch.int16).pin_memory()
    for h, w in zip(hops, want):
        op.step(h.pin_memory(), out=oh)
        torch.cuda.synchronize()
        assert torch.equal(oh, w)
    with pytest.raises(ValueError):
        op.step(hops[0][:10].to("cuda:0"), first_stream=S - 5)
    with pytest.raises(ValueError):
        op.step(hops[0])                                        # pageable host memory
    op.reset(0)
