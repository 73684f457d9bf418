#!/usr/bin/env python
"""Golden vectors made by EXECUTING the reference's own RTL source.

    python tests/golden/make_rtl_golden.py          (build container only)

The reference hot path is nMigen RTL (mfcc/core/mfcc.py:19-117).  nmigen is not
installed and cannot be, so the reference's unmodified source files are imported from
/root/reference against oracle/hdl_shim (a small interpreter of the nMigen constructs
they use) and driven by a bench with the structure of the reference's own
`mfcc-sim` bench (mfcc/core/mfcc.py:120-183: feed samples then zero padding, keep
source.ready high, collect every stage's source stream on valid & ready, close a row
on `last`).  Every stage output of every frame is stored, so the fixtures pin each
stage of the oracle and of the CUDA path separately.

Writes tests/golden/rtl_<case>.npz with arrays
    samples, window, fft_r, fft_i, power, filter, log, dct, out   (+ nfilters, nceptrums)
and tests/golden/rtl_stage_vectors.npz (single-stage runs of the reference's own
bench stimuli: fft.py:490, dct_stream.py:81, log.py:149, filterbank.py:151).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, os.path.join(ROOT, "oracle", "hdl_shim"))
sys.path.insert(0, REF)

import contextlib  # noqa: E402
import io  # noqa: E402

from nmigen import *  # noqa: E402,F401,F403
from nmigen.sim import Simulator, Passive  # noqa: E402

with contextlib.redirect_stdout(io.StringIO()):
    from mfcc.core.mfcc import MFCC  # noqa: E402


def run_mfcc(samples, n_frames, nfilters, nceptrums, stall=None, seed=0, nfft=512, samplerate=16e3):
    """Drive the reference MFCC top level.  `stall` = probability that the bench
    withholds sink.valid / source.ready in a cycle (exercises the handshakes)."""
    with contextlib.redirect_stdout(io.StringIO()):       # calc_coeffs()/calc_filters() print
        dut = MFCC(nfft=nfft, samplerate=samplerate, nfilters=nfilters, nceptrums=nceptrums)
        sim = Simulator(dut)
    rng = np.random.default_rng(seed)
    taps = [("window", dut.window.source, "data"), ("fft_r", dut.fft_stream.source, "data_r"),
            ("fft_i", dut.fft_stream.source, "data_i"), ("power", dut.powspec.source, "data"),
            ("filter", dut.filterbank.source, "data"), ("log", dut.log2.source, "data"),
            ("dct", dut.dct_stream.source, "data"), ("out", dut.discard.source, "data"),
            ("frame", dut.frame.source, "data")]
    rows = {n: [] for n, _s, _f in taps}
    cur = {n: [] for n, _s, _f in taps}
    sig = [int(v) for v in samples]

    def bench():
        idx = 0
        yield dut.source.ready.eq(1)
        while min(len(r) for r in rows.values()) < n_frames:
            # drive the inputs for the coming cycle
            v = 1 if stall is None else int(rng.random() >= stall)
            yield dut.sink.valid.eq(v)
            yield dut.sink.data.eq(sig[idx] if idx < len(sig) else 0)      # padding, mfcc.py:150-153
            if stall is not None:
                yield dut.source.ready.eq(int(rng.random() >= stall))
            # sample every stream as it stands for this cycle (settled)
            took = (yield dut.sink.valid) and (yield dut.sink.ready)
            for name, src, field in taps:
                if (yield src.valid) and (yield src.ready):
                    cur[name].append((yield getattr(src, field)))
                    if (yield src.last):
                        rows[name].append(cur[name])
                        cur[name] = []
            yield
            if took:
                idx += 1
        yield dut.sink.valid.eq(0)

    sim.add_clock(1e-6)
    sim.add_sync_process(bench)
    sim.run()
    out = {}
    for n in rows:
        r = rows[n][:n_frames]
        assert len(r) == n_frames, (n, len(r))
        out[n] = np.array(r, dtype=np.int64)
    return out, sim.cycles


def case_inputs():
    wav = np.load(os.path.join(HERE, "f2bjrop1.npz"))["samples"].astype(np.int64)
    rng = np.random.default_rng(20260001)
    N = 512 + 170 * 7
    t = np.arange(N)
    cases = {
        # name: (samples, n_frames, nfilters, nceptrums, stall)
        "wav_head": (wav[:512 + 170 * 11], 12, 32, 32, None),
        "wav_mid_16c": (wav[60000:60000 + 512 + 170 * 5], 6, 32, 16, None),
        "wav_mid_16f": (wav[60000:60000 + 512 + 170 * 5], 6, 16, 16, None),
        "noise_full": (rng.integers(-32768, 32768, N), 8, 32, 32, None),
        "noise_full_16f": (rng.integers(-32768, 32768, N), 8, 16, 16, None),
        "gauss": (np.clip(rng.normal(0, 2000, N), -32768, 32767).astype(np.int64), 8, 32, 32, None),
        "tones": ((8000 * np.sin(2 * np.pi * 440 / 16000 * t) + 6000 * np.sin(2 * np.pi * 3100 / 16000 * t)
                   + 3000 * np.sin(2 * np.pi * 7000 / 16000 * t)).astype(np.int64), 8, 32, 32, None),
        "alt_fullscale": (np.where(t % 2 == 0, 32767, -32768), 4, 32, 32, None),
        "square_fullscale": (np.where((t // 7) % 2 == 0, 32767, -32768), 4, 32, 32, None),
        "const_max": (np.full(N, 32767), 3, 32, 32, None),
        "const_min": (np.full(N, -32768), 3, 32, 32, None),
        "zeros": (np.zeros(N, dtype=np.int64), 3, 32, 32, None),
        "impulse": (np.where(t == 300, 32767, 0), 4, 32, 32, None),
        "dc_one": (np.ones(N, dtype=np.int64), 3, 32, 32, None),
        # short clip: the bench pads with zeros like the hosts do (software/main.c:135-144)
        "short_padded": (wav[30000:30700], 4, 32, 32, None),
        # random back-pressure on both ends: results must not depend on timing
        "wav_stalled": (wav[60000:60000 + 512 + 170 * 3], 4, 32, 32, 0.3),
    }
    return cases


def _stream_stage(dut, rows_in, field_in="data", fields_out=("data",), n_rows_out=None):
    """Feed rows (lists of ints, `last` on the final element of each row) into a stage
    with sink/source stream endpoints; collect the source rows."""
    sim = Simulator(dut)
    flat = [(int(v), i == len(r) - 1) for r in rows_in for i, v in enumerate(r)]
    n_rows_out = len(rows_in) if n_rows_out is None else n_rows_out
    rows, cur = [], []

    def bench():
        idx = 0
        yield dut.source.ready.eq(1)
        while len(rows) < n_rows_out:
            if idx < len(flat):
                yield dut.sink.valid.eq(1)
                yield getattr(dut.sink, field_in).eq(flat[idx][0])
                yield dut.sink.last.eq(int(flat[idx][1]))
            else:
                yield dut.sink.valid.eq(0)
            took = (yield dut.sink.valid) and (yield dut.sink.ready)
            if (yield dut.source.valid) and (yield dut.source.ready):
                item = []
                for f in fields_out:
                    item.append((yield getattr(dut.source, f)))
                cur.append(item)
                if (yield dut.source.last):
                    rows.append(list(cur))
                    del cur[:]
            yield
            if took:
                idx += 1

    sim.add_clock(1e-6)
    sim.add_sync_process(bench)
    sim.run()
    return np.array(rows, dtype=np.int64)


def stage_vectors():
    """Single stages of the reference, instantiated with the arguments MFCC.elaborate
    uses (mfcc/core/mfcc.py:38-88), on the reference's own bench stimuli
    (tests/golden/stimuli.json) and on seeded random rows."""
    import json
    stim = json.load(open(os.path.join(HERE, "stimuli.json")))
    with contextlib.redirect_stdout(io.StringIO()):
        from mfcc.misc.fft import FFT
        from mfcc.core.fft_stream import FftStream
        from mfcc.core.dct_stream import DCTStream
        from mfcc.core.log import Log2Fix
        from mfcc.core.filterbank import FilterBank
        from mfcc.core.pow2 import PowerSpectrum
        from mfcc.core.window import WindowHamming
        from mfcc.misc.mul import Multiplier
    rng = np.random.default_rng(7)
    s16 = lambda v: ((np.asarray(v, dtype=np.int64) + 32768) % 65536) - 32768
    out = {}

    # FFT core with the bench of mfcc/misc/fft.py:497-525
    data = s16(stim["fft512_hex"])
    dut = FFT(size=512, i_width=16, o_width=16, m_width=16)
    sim = Simulator(dut)
    res = []

    def process():
        for i, n in enumerate(data):
            yield dut.i.addr.eq(i)
            yield dut.i.data.real.eq(int(n))
            yield dut.i.en.eq(1)
            yield
        yield dut.i.en.eq(0)
        yield dut.start.eq(1)
        yield
        yield dut.start.eq(0)
        yield
        while not (yield dut.ready):
            yield
        yield
        for i in range(dut.size // 2):
            yield dut.o.addr.eq(i)
            yield
            res.append(((yield dut.o.data.real), (yield dut.o.data.imag)))
    sim.add_clock(1e-6)
    sim.add_sync_process(process)
    sim.run()
    out["fft512_in"] = data
    out["fft512_out"] = np.array(res, dtype=np.int64)

    # FftStream on random full-scale rows (two frames back to back)
    rows = rng.integers(-32768, 32768, (3, 512))
    out["fftstream_in"] = rows
    out["fftstream_out"] = _stream_stage(FftStream(16, 512), rows, fields_out=("data_r", "data_i"))

    # DCTStream(16, 16) on the stimulus of dct_stream.py:81, DCTStream(16, 32) on log-range rows
    d16 = s16(stim["dct16_hex"])
    out["dct16_in"] = d16[None, :]
    out["dct16_out"] = _stream_stage(DCTStream(16, 16), d16[None, :])[:, :, 0]
    rows = rng.integers(0, 32737, (6, 32))
    out["dct32_in"] = rows
    out["dct32_out"] = _stream_stage(DCTStream(16, 32), rows)[:, :, 0]
    rows = rng.integers(-32768, 32768, (4, 32))            # full signed range (wraps)
    out["dct32s_in"] = rows
    out["dct32s_out"] = _stream_stage(DCTStream(16, 32), rows)[:, :, 0]

    # Log2Fix(16, 15): boundary values and a sweep
    vals = np.array([0, 1, 2, 3, 4, 5, 7, 8, 255, 256, 257, 1023, 1024, 4095, 4096, 32767, 32768, 65535,
                     stim["log2_value"] & 0xffff] + list(rng.integers(0, 65536, 45)))
    out["log2_in"] = vals[None, :]
    out["log2_out"] = _stream_stage(Log2Fix(16, 15, multiplier_cls=Multiplier), vals[None, :])[:, :, 0]

    # PowerSpectrum(16, 30)
    pr = rng.integers(-32768, 32768, (1, 64, 2))
    pr[0, :4] = [[-32768, -32768], [32767, 32767], [-32768, 32767], [0, -1]]
    dutp = PowerSpectrum(width=16, width_output=30, multiplier_cls=Multiplier)
    simp = Simulator(dutp)
    got = []

    def pbench():
        idx = 0
        yield dutp.source.ready.eq(1)
        while len(got) < pr.shape[1]:
            if idx < pr.shape[1]:
                yield dutp.sink.valid.eq(1)
                yield dutp.sink.data_r.eq(int(pr[0, idx, 0]))
                yield dutp.sink.data_i.eq(int(pr[0, idx, 1]))
            else:
                yield dutp.sink.valid.eq(0)
            took = (yield dutp.sink.valid) and (yield dutp.sink.ready)
            if (yield dutp.source.valid) and (yield dutp.source.ready):
                got.append((yield dutp.source.data))
            yield
            if took:
                idx += 1
    simp.add_clock(1e-6)
    simp.add_sync_process(pbench)
    simp.run()
    out["power_in"] = pr[0]
    out["power_out"] = np.array(got, dtype=np.int64)

    # FilterBank as MFCC builds it, 32 and 16 taps: constant input (filterbank.py:151),
    # random 30-bit rows, and all-ones (largest sums: drives the 16-bit output wrap)
    for ntap in (32, 16):
        rows = np.stack([np.full(256, stim["filterbank_const"]), rng.integers(0, 1 << 30, 256),
                         rng.integers(0, 1 << 20, 256), np.full(256, (1 << 30) - 1),
                         rng.integers(0, 1 << 30, 256)])
        mk = lambda: FilterBank(width=30, width_output=16, gain=18, sample_rate=16e3, nfft=512,
                                ntap=ntap, multiplier_cls=Multiplier)
        out["mel%d_in" % ntap] = rows
        out["mel%d_out" % ntap] = _stream_stage(mk(), rows)[:, :, 0]

    # WindowHamming(16, 512, 8): three frames so the frame-to-frame state is covered
    rows = rng.integers(-32768, 32768, (3, 512))
    rows[1] = 32767
    out["window_in"] = rows
    out["window_out"] = s16(_stream_stage(WindowHamming(width=16, nfft=512, precision=8), rows)[:, :, 0])
    return out


def frame_vectors():
    """Frame and Preemph on their own, for the parameter sets the reference itself uses:
    the constructor defaults and bench case of frame.py:50, :233-249 (400/160/512), the
    MFCC instance (512/170/512, mfcc.py:41-44), and two small ones; random gaps on the
    sink and random stalls on the source (the five back-pressure scenarios of
    frame.py:160-231 in one)."""
    with contextlib.redirect_stdout(io.StringIO()):
        from mfcc.core.frame import Frame
        from mfcc.core.preemph import Preemph
    rng = np.random.default_rng(11)
    out = {}
    for name, (wl, st, nfft, nfr) in {"400_160_512": (400, 160, 512, 6), "512_170_512": (512, 170, 512, 4),
                                      "96_96_128": (96, 96, 128, 5), "100_30_128": (100, 30, 128, 9),
                                      "25_8_32": (25, 8, 32, 12)}.items():      # frame.py:158, the reference's own bench
        n = wl + st * (nfr - 1) + 7
        x = rng.integers(-32768, 32768, n)
        dut = Frame(16, wl, st, nfft)
        sim = Simulator(dut)
        rows, cur = [], []
        gaps = np.random.default_rng(5)

        def bench():
            idx = 0
            while len(rows) < nfr:
                v = idx < n and gaps.random() < 0.7
                yield dut.sink.valid.eq(int(v))
                if v:
                    yield dut.sink.data.eq(int(x[idx]))
                yield dut.source.ready.eq(int(gaps.random() < 0.8))
                took = (yield dut.sink.valid) and (yield dut.sink.ready)
                if (yield dut.source.valid) and (yield dut.source.ready):
                    cur.append((yield dut.source.data))
                    if (yield dut.source.last):
                        rows.append(list(cur))
                        del cur[:]
                yield
                if took:
                    idx += 1
        sim.add_clock(1e-6)
        sim.add_sync_process(bench)
        sim.run()
        out["frame_%s_in" % name] = x
        out["frame_%s_out" % name] = np.array(rows, dtype=np.int64)
    # Preemph: full-scale corners then random
    x = np.concatenate([[32767, -32768, -32768, 32767, 32767, 0, -1, 1, 31, 32, 33, -31, -32, -33],
                        rng.integers(-32768, 32768, 200)])
    dut = Preemph(16)
    sim = Simulator(dut)
    got = []

    def pbench():
        idx = 0
        yield dut.source.ready.eq(1)
        while len(got) < len(x):
            yield dut.sink.valid.eq(int(idx < len(x)))
            if idx < len(x):
                yield dut.sink.data.eq(int(x[idx]))
            took = (yield dut.sink.valid) and (yield dut.sink.ready)
            if (yield dut.source.valid) and (yield dut.source.ready):
                got.append((yield dut.source.data))
            yield
            if took:
                idx += 1
    sim.add_clock(1e-6)
    sim.add_sync_process(pbench)
    sim.run()
    out["preemph_in"] = x
    out["preemph_out"] = np.array(got, dtype=np.int64)

    # PowerSpectrum(width=16) with the class default width_output = 24 (pow2.py:44, the bench at :74)
    with contextlib.redirect_stdout(io.StringIO()):
        from mfcc.core.pow2 import PowerSpectrum
        from mfcc.misc.mul import Multiplier
    pr = rng.integers(-32768, 32768, (48, 2))
    pr[:4] = [[-32768, -32768], [32767, 32767], [-32768, 32767], [0, -1]]
    dutp = PowerSpectrum(width=16, width_output=24, multiplier_cls=Multiplier)
    simp = Simulator(dutp)
    gotp = []

    def powbench():
        idx = 0
        yield dutp.source.ready.eq(1)
        while len(gotp) < len(pr):
            yield dutp.sink.valid.eq(int(idx < len(pr)))
            if idx < len(pr):
                yield dutp.sink.data_r.eq(int(pr[idx, 0]))
                yield dutp.sink.data_i.eq(int(pr[idx, 1]))
            took = (yield dutp.sink.valid) and (yield dutp.sink.ready)
            if (yield dutp.source.valid) and (yield dutp.source.ready):
                gotp.append((yield dutp.source.data))
            yield
            if took:
                idx += 1
    simp.add_clock(1e-6)
    simp.add_sync_process(powbench)
    simp.run()
    out["power24_in"] = pr
    out["power24_out"] = np.array(gotp, dtype=np.int64)

    # Log2Fix(37, 20), the reference's own bench (log.py:143-149: the value 2207315)
    with contextlib.redirect_stdout(io.StringIO()):
        from mfcc.core.log import Log2Fix
    vals = np.array([2207315, 0, 1, 2, 3, 255, 256, (1 << 37) - 1, 1 << 36, 1 << 20, 12345678901]
                    + [int(v) for v in rng.integers(1, 1 << 37, 21)], dtype=np.int64)
    out["log2_37_20_in"] = vals
    out["log2_37_20_out"] = _stream_stage(Log2Fix(37, 20, multiplier_cls=Multiplier), vals[None, :])[0, :, 0]

    # FilterBank with the widths of the reference's own bench (filterbank.py:148: width 31, output 32,
    # constant input 1234) and with the class defaults (24 / 24 / 16 taps, filterbank.py:38-48)
    with contextlib.redirect_stdout(io.StringIO()):
        from mfcc.core.filterbank import FilterBank
    rows = np.stack([np.full(256, 1234), rng.integers(0, 1 << 31, 256), np.full(256, (1 << 31) - 1),
                     np.zeros(256, dtype=np.int64), rng.integers(0, 1 << 12, 256)])
    out["mel_31_32_in"] = rows
    out["mel_31_32_out"] = _stream_stage(FilterBank(width=31, width_output=32, multiplier_cls=Multiplier), rows)[:, :, 0]
    rows = np.stack([np.full(256, 1234), rng.integers(0, 1 << 24, 256), np.full(256, (1 << 24) - 1)])
    out["mel_default_in"] = rows
    out["mel_default_out"] = _stream_stage(FilterBank(multiplier_cls=Multiplier), rows)[:, :, 0]
    return out


def size_vectors():
    """WindowHamming, FftStream and DCTStream at sizes other than the ones MFCC builds: the
    stage classes are generic in nfft (window.py:11, fft_stream.py:7, dct_stream.py:7).
    Random full-range rows, several rows back to back."""
    with contextlib.redirect_stdout(io.StringIO()):
        from mfcc.core.fft_stream import FftStream
        from mfcc.core.dct_stream import DCTStream
        from mfcc.core.window import WindowHamming
    rng = np.random.default_rng(23)
    s16 = lambda v: ((np.asarray(v, dtype=np.int64) + 32768) % 65536) - 32768
    out = {}
    for nfft, rows_n in ((8, 4), (16, 4), (64, 3), (256, 2), (1024, 2)):
        rows = rng.integers(-32768, 32768, (rows_n, nfft))
        rows[0, : nfft // 2] = 32767
        rows[0, nfft // 2:] = -32768
        out["fft%d_in" % nfft] = rows
        out["fft%d_out" % nfft] = s16(_stream_stage(FftStream(16, nfft), rows, fields_out=("data_r", "data_i")))
    for nfft, rows_n in ((2, 4), (4, 4), (8, 4), (64, 3), (128, 2)):
        rows = rng.integers(-32768, 32768, (rows_n, nfft))
        out["dct%d_in" % nfft] = rows
        out["dct%d_out" % nfft] = s16(_stream_stage(DCTStream(16, nfft), rows)[:, :, 0])
    for nfft, prec, rows_n in ((8, 8, 3), (64, 8, 3), (256, 6, 2), (1024, 8, 2), (2048, 10, 2)):
        rows = rng.integers(-32768, 32768, (rows_n, nfft))
        rows[1] = 32767
        with contextlib.redirect_stdout(io.StringIO()):
            dut = WindowHamming(width=16, nfft=nfft, precision=prec)
        out["window%d_%d_in" % (nfft, prec)] = rows
        out["window%d_%d_out" % (nfft, prec)] = s16(_stream_stage(dut, rows)[:, :, 0])
    return out


PARAM_SETS = {      # name: (nfft, samplerate, nfilters, nceptrums, frames)
    "256_8k_8_8": (256, 8e3, 8, 8, 4),
    "1024_16k_64_24": (1024, 16e3, 64, 24, 2),
    "512_44k1_32_13": (512, 44100, 32, 13, 3),
    "128_16k_16_16": (128, 16e3, 16, 16, 5),
    "512_16k_8_8": (512, 16e3, 8, 8, 2),
}


def param_vectors():
    """The reference MFCC top level built with other constructor arguments than its tops
    use (mfcc/core/mfcc.py:20-21: nfft, samplerate, nfilters and nceptrums are parameters),
    every stage tapped, on full-scale noise followed by a quiet stretch."""
    rng = np.random.default_rng(31)
    out = {}
    for name, (nfft, sr, nf, nc, nfr) in PARAM_SETS.items():
        n = nfft + (nfft // 3) * (nfr - 1)
        x = rng.integers(-32768, 32768, n)
        x[n // 2:] = rng.integers(-300, 300, n - n // 2)
        t0 = time.time()
        taps, cycles = run_mfcc(x, nfr, nf, nc, nfft=nfft, samplerate=sr)
        out[name + "_samples"] = x.astype(np.int16)
        for k, v in taps.items():
            out[name + "_" + k] = v.astype(np.int32)
        print("%-16s %d frames %7d cycles %.1f s" % (name, nfr, cycles, time.time() - t0))
    return out


def main():
    only = sys.argv[1:]
    if only == ["params"]:
        pv = param_vectors()
        np.savez_compressed(os.path.join(HERE, "rtl_param_vectors.npz"), **pv)
        return
    if only == ["sizes"]:
        zv = size_vectors()
        np.savez_compressed(os.path.join(HERE, "rtl_size_vectors.npz"), **zv)
        print("size vectors:", {k: v.shape for k, v in zv.items()})
        return
    if only == ["frames"]:
        fv = frame_vectors()
        np.savez_compressed(os.path.join(HERE, "rtl_frame_vectors.npz"), **fv)
        print("frame vectors:", {k: v.shape for k, v in fv.items()})
        return
    for name, (x, nfr, nf, nc, stall) in case_inputs().items():
        if only and name not in only:
            continue
        t0 = time.time()
        taps, cycles = run_mfcc(x, nfr, nf, nc, stall)
        np.savez_compressed(os.path.join(HERE, "rtl_%s.npz" % name), samples=np.asarray(x, dtype=np.int16),
                            nfilters=nf, nceptrums=nc,
                            **{k: v.astype(np.int32) for k, v in taps.items()})
        print("%-18s %2d frames  %7d cycles  %.1f s   out[0][:6]=%s" % (
            name, nfr, cycles, time.time() - t0, taps["out"][0][:6].tolist()))
    if not only or "stages" in only:
        sv = stage_vectors()
        np.savez_compressed(os.path.join(HERE, "rtl_stage_vectors.npz"), **sv)
        print("stage vectors:", {k: v.shape for k, v in sv.items()})
        fv = frame_vectors()
        np.savez_compressed(os.path.join(HERE, "rtl_frame_vectors.npz"), **fv)
        print("frame vectors:", {k: v.shape for k, v in fv.items()})
        zv = size_vectors()
        np.savez_compressed(os.path.join(HERE, "rtl_size_vectors.npz"), **zv)
        print("size vectors:", {k: v.shape for k, v in zv.items()})
        np.savez_compressed(os.path.join(HERE, "rtl_param_vectors.npz"), **param_vectors())


if __name__ == "__main__":
    main()
