"""TEST INFRASTRUCTURE — CPU oracle (numpy / Python ints) for the MFCC hot path.

This is a restatement of the arithmetic of the reference's nMigen RTL
(/root/reference/mfcc/core/*.py, mfcc/misc/fft.py, mfcc/misc/mul.py), written
from the hardware description, one function per stage.  It is deliberately plain
and slow.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import it; the product (mfcc_b200/) never does and has no CPU fallback.

It builds its own tables (it does not import mfcc_b200.tables) so that the table
code of the product is checked against an independent construction.

Pinning (see DESIGN.md "Oracle"): the window curve and the mel filter points are
checked against the values printed in the reference's notebooks
(tests/golden/notebook_pins.json); the whole chain is checked against vectors made
by executing the reference's own source files under a small interpreter of the
nMigen constructs they use (oracle/hdl_shim, tests/golden/rtl_*.npz).

nMigen rules used throughout: assignment to a narrower signal keeps the low bits;
``>>`` on a signed value is an arithmetic shift; a slice is an unsigned bit field.
"""
from __future__ import annotations

import math

import numpy as np
from scipy.signal import get_window


def wrap(v, bits):
    """Two's-complement reinterpretation of the low `bits` bits (signed signal assign)."""
    v = np.asarray(v, dtype=np.int64)
    m = (1 << bits) - 1
    s = 1 << (bits - 1)
    return ((v + s) & m) - s


# ----------------------------------------------------------------------------- a1
def preemph(x, prev=0):
    """Preemph, mfcc/core/preemph.py:18-28.
    ``source.data = sink.data + (odata >> 5) - odata`` with odata the previous
    accepted sample (0 after reset); 16-bit signed output."""
    x = np.asarray(x, dtype=np.int64)
    od = np.empty_like(x)
    if x.shape[-1]:
        od[..., 0] = prev
        od[..., 1:] = x[..., :-1]
    return wrap(x + (od >> 5) - od, 16)


# ----------------------------------------------------------------------------- a2
def num_frames(n, nfft=512, hop=170):
    return 0 if n < nfft else (n - nfft) // hop + 1


def frame(y, nfft=512, hop=None, n_frames=None, windowlen=None):
    """Frame(windowlen, stepsize, nfft), mfcc/core/frame.py:49-155: frame k is
    y[hop*k : hop*k+windowlen] followed by nfft-windowlen zeros (the `padding` mux,
    frame.py:77, :120).  MFCC instantiates it with windowlen = nfft, stepsize = nfft//3
    (mfcc/core/mfcc.py:41-44), the defaults here.  With n_frames given, the stream is
    extended with zeros (what the reference hosts feed after the end of the audio:
    mfcc/core/mfcc.py:150-153, software/main.c:135-144)."""
    y = np.asarray(y, dtype=np.int64)
    hop = nfft // 3 if hop is None else hop
    wl = nfft if windowlen is None else windowlen
    k = num_frames(len(y), wl, hop) if n_frames is None else n_frames
    need = hop * (k - 1) + wl if k > 0 else 0
    if need > len(y):
        y = np.concatenate([y, np.zeros(need - len(y), dtype=np.int64)])
    out = np.zeros((k, nfft), dtype=np.int64)
    if k:
        out[:, :wl] = y[hop * np.arange(k)[:, None] + np.arange(wl)[None, :]]
    return out


# ----------------------------------------------------------------------------- a3
def window_coeffs(nfft=512, precision=8):
    """calc_coeffs, mfcc/core/window.py:22-43."""
    maxheight = 2 ** (precision + 1) - 1
    win = get_window("hamm", nfft, fftbins=True)
    winfull = (win * maxheight).astype(int)
    mem = np.copy(winfull[: nfft // 4][1::2])
    off_fst = int(mem[0])
    mem -= off_fst
    assert max(mem) < 2 ** precision
    off_lst = int(2 * (winfull[nfft // 4] - off_fst))
    return [int(v) for v in mem], off_fst, off_lst


def window_curve(nfft=512, precision=8, frames=1):
    """Sequence of `curve` values the multiplier sees, mfcc/core/window.py:53-123,
    emulated register by register over `frames` consecutive frames (count, point_r
    carried across frames exactly as the hardware does)."""
    mem, off_fst, off_lst = window_coeffs(nfft, precision)
    nb = int(math.log2(nfft))
    cw = precision + 1
    out = []
    count = 0
    point_r = 0
    for _ in range(frames * nfft):
        bit_msb = (count >> (nb - 1)) & 1        # count[-1]
        bit_dir = (count >> (nb - 2)) & 1        # count[-2]
        bits_addr = (count >> 1) & ((1 << (nb - 3)) - 1)   # count[1:-2]
        bit_odd = count & 1
        addr = bits_addr
        if bit_dir:
            addr = (~addr) & ((1 << (nb - 3)) - 1)
        data = mem[addr]
        if bit_msb ^ bit_dir:
            point = (off_lst - data) & ((1 << cw) - 1)
        else:
            point = data
        if not bit_odd:
            curve = (off_fst + ((point + point_r) >> 1)) & ((1 << cw) - 1)
        else:
            curve = (off_fst + point) & ((1 << cw) - 1)
            point_r = point
        out.append(curve)
        count = 0 if count == nfft - 1 else count + 1      # last resets the counter
    return np.array(out, dtype=np.int64).reshape(frames, nfft)


def window(frames, nfft=512, precision=8):
    """``source.data = mul.o.c[-width:]`` with c = signed16 * unsigned(precision+1),
    (precision+1+16) bits signed: the top 16 bits, i.e. an arithmetic shift by
    precision+1 (mfcc/core/window.py:74-75, :84; mfcc/misc/mul.py:11-13)."""
    curve = window_curve(nfft, precision, 1)[0]
    prod = np.asarray(frames, dtype=np.int64) * curve
    return wrap(prod >> (precision + 1), 16)


# ----------------------------------------------------------------------------- a4-a7
def twiddle_rom(size, width=16):
    """TwiddleROM.init, mfcc/misc/fft.py:28-36."""
    p = np.linspace(start=0, stop=np.pi / 2, num=int(size // 4), endpoint=False)
    z = np.round((1 << (width - 2)) * np.exp(-1j * p))
    return [int(v.real) for v in z], [int(v.imag) for v in z]


def twiddle(size, k, width=16):
    """TwiddleROM read for address k in [0, size/2), mfcc/misc/fft.py:43-59."""
    re, im = twiddle_rom(size, width)
    q = size // 4
    sel = k >= q            # rp_addr[-1]
    i = k % q               # rp_addr[:-1]
    if sel:
        return im[i], -re[i]
    return re[i], im[i]


def butterfly(x0r, x0i, x1r, x1i, twr, twi, width=16):
    """Butterfly(width, bias_width=width-2, scale_bit=1), mfcc/misc/fft.py:93-96,
    :140-192, written with the three multipliers of the hardware."""
    bw = width - 2
    bias = (1 << (bw - 1)) - 1
    w33 = 2 * width + 1
    add0 = x1r + x1i                              # s2_add_0
    mul0 = wrap(add0 * twr, w33)                  # s3_mul_0
    mul0 = wrap(mul0 + bias, w33)                 # s4_mul_0
    add1 = twr + twi                              # s4_add_1
    sub0 = twr - twi                              # s4_sub_0
    mul1 = wrap(x1i * add1, w33)                  # s5_mul_1
    mul2 = wrap(x1r * sub0, w33)                  # s5_mul_2
    sub1 = wrap(mul0 - mul1, w33)                 # s6_sub_1 = Re(x1 w) + bias
    sub2 = wrap(mul0 - mul2, w33)                 # s6_sub_2 = Im(x1 w) + bias
    # [bias_width:] is an unsigned slice of a 33-bit signal
    tr = (sub1 & ((1 << w33) - 1)) >> bw
    ti = (sub2 & ((1 << w33) - 1)) >> bw
    y0r = wrap((x0r + tr) >> 1, width)
    y0i = wrap((x0i + ti) >> 1, width)
    y1r = wrap((x0r - tr) >> 1, width)
    y1i = wrap((x0i - ti) >> 1, width)
    return y0r, y0i, y1r, y1i


def _bitrev(n, bits):
    r = 0
    for _ in range(bits):
        r = (r << 1) | (n & 1)
        n >>= 1
    return r


def fft_int(xr, xi, size):
    """FFT(size, 16, 16, 16), mfcc/misc/fft.py:349-484, as the in-place
    decimation-in-time loop its Scheduler (:197-346) is equivalent to: bit-reversed
    load (:413-418), stage s pairs positions p and p + 2**s with twiddle address
    k * (size/2 >> s), k = p mod 2**s (:254-258, :310-316).  tests/test_oracle_fft.py
    checks this loop against fft_rtl_banks() below, which follows the three memory
    banks cycle by cycle.  xr, xi: [..., size] integer arrays.  Returns all `size`
    positions (the stream adapters only read the first half or less)."""
    nb = int(math.log2(size))
    rev = np.array([_bitrev(n, nb) for n in range(size)])
    vr = np.array(np.asarray(xr, dtype=np.int64)[..., rev])
    vi = np.array(np.asarray(xi, dtype=np.int64)[..., rev])
    tw = [twiddle(size, k) for k in range(size // 2)]
    for s in range(nb):
        half = 1 << s
        p0 = np.array([g + k for g in range(0, size, 2 * half) for k in range(half)])
        p1 = p0 + half
        kk = np.array([k for g in range(0, size, 2 * half) for k in range(half)])
        twr = np.array([tw[k * ((size // 2) >> s)][0] for k in kk], dtype=np.int64)
        twi = np.array([tw[k * ((size // 2) >> s)][1] for k in kk], dtype=np.int64)
        y0r, y0i, y1r, y1i = butterfly(vr[..., p0], vi[..., p0], vr[..., p1], vi[..., p1], twr, twi)
        vr[..., p0], vi[..., p0], vr[..., p1], vi[..., p1] = y0r, y0i, y1r, y1i
    return vr, vi


def fft_rtl_banks(xr, xi, size, strict=True):
    """One transform, following FFT + Scheduler of mfcc/misc/fft.py bank by bank and
    cycle by cycle: INIT writes (:438-450), the consume/produce counters and the
    x/y bank multiplexers (:251-308), the twiddle address stepping (:310-316, :331)
    and the 1 + 1 + 7 cycle read/butterfly latency (:140-192, :344).  Reads see a
    write made in an earlier cycle; from 64 points up a location is never read in the
    cycle it is written, which this function asserts when `strict`, so the transparency of
    the read ports is irrelevant there.  Returns the `size/2` values read through FFT.o
    (:426-435, o.addr < size/2 selects mem0).

    A stage lasts size/2 cycles and a butterfly's results land 8 cycles after its read.
    At 32 points some locations are read in the very cycle they are written: the read
    ports are nMigen's default transparent ones (:385-393), they return the new value
    (writes are applied before reads below) and the result is still the transform.  Below
    32 points the next stage reads locations the previous one has not written yet: a
    genuine hazard.  With strict=False this model reproduces what the RTL then delivers
    (tests/golden/rtl_size_vectors.npz, sizes 8 and 16); it is not a Fourier transform,
    and fft_int() does not follow it there."""
    nb = int(math.log2(size))
    half = size // 2
    mem0 = [(0, 0)] * half
    mem1 = [(0, 0)] * half
    mem2 = [(0, 0)] * half
    # INIT state: i_addr_rev = bit reversal; bit 0 picks mem0 or mem1+mem2
    for n in range(size):
        r = _bitrev(n, nb)
        v = (int(xr[n]), int(xi[n]))
        if r & 1:
            mem1[r >> 1] = v
            mem2[r >> 1] = v
        else:
            mem0[r >> 1] = v
    tapbits = nb - 1
    nstage = nb                       # stage counts 0 .. log2(size/2)
    # Pipeline: cycle c presents (tap, stage) = consume; data returns next cycle
    # (sync read port), x_mem_sel is registered alongside; o.stb one cycle after BUSY;
    # the butterfly adds 7 register stages; the write happens in the cycle y appears.
    total = half * nstage
    lat = 1 + 7                       # read latency + butterfly depth
    pending = {}                      # cycle -> (produce index, y0, y1)
    trom_addr = 0
    last_write_cycle = {}
    for c in range(total + lat + 1):
        # writes of this cycle (produce side)
        if c in pending:
            pidx, y0, y1 = pending.pop(c)
            ptap = pidx % half
            pstage = pidx // half
            y_sel = (ptap >> pstage) & 1 if pstage < tapbits else 0
            d0, d12 = (y1, y0) if y_sel else (y0, y1)
            mem0[ptap] = d0
            if pstage & 1:
                mem2[ptap] = d12
            else:
                mem1[ptap] = d12
            last_write_cycle[(0, ptap)] = c
            last_write_cycle[(2 if pstage & 1 else 1, ptap)] = c
        # read issued this cycle (consume side)
        if c < total:
            tap = c % half
            stage = c // half
            a0 = tap
            a1 = ((tap << 1) ^ (1 << stage)) >> 1
            sel0 = ((tap >> (stage - 1)) & 1) if stage != 0 else 0
            sel1 = stage & 1
            for bank, adr in ((0, a0), (1 if sel1 else 2, a1)):
                assert not strict or last_write_cycle.get((bank, adr), -1) < c, "read/write hazard"
            other = mem1[a1] if sel1 else mem2[a1]
            if sel0:
                x0, x1 = other, mem0[a0]
            else:
                x0, x1 = mem0[a0], other
            twr, twi = twiddle(size, trom_addr % half)
            y = butterfly(x0[0], x0[1], x1[0], x1[1], twr, twi)
            y = [int(v) for v in y]
            pending[c + lat] = (c, (y[0], y[1]), (y[2], y[3]))
            # trom_addr_step: bit `stage` of the bit-reversed word set (:310-311)
            step = 1 << (nb - 1 - stage)
            trom_addr = (trom_addr + step) % half
    assert not pending
    return np.array([v[0] for v in mem0]), np.array([v[1] for v in mem0])


def fft_stream(frames, nfft=512):
    """FftStream, mfcc/core/fft_stream.py:30-38, :24, :28: real input, imag 0,
    bins 0 .. nfft/2-1.  Returns (re, im) [K, nfft/2]."""
    frames = np.asarray(frames, dtype=np.int64)
    vr, vi = fft_int(frames, np.zeros_like(frames), nfft)
    return vr[..., : nfft // 2], vi[..., : nfft // 2]


# ----------------------------------------------------------------------------- a8
def power_spectrum(re, im, width=16, width_output=30):
    """PowerSpectrum, mfcc/core/pow2.py:32 (sum into an unsigned 2*width signal),
    :64 (top width_output bits)."""
    re = np.asarray(re, dtype=np.int64)
    im = np.asarray(im, dtype=np.int64)
    s = (re * re + im * im) & ((1 << (2 * width)) - 1)
    return s >> (2 * width - width_output)


# ----------------------------------------------------------------------------- a9
def freq_to_mel(freq):
    return 2595.0 * np.log10(1.0 + freq / 700.0)


def met_to_freq(mels):
    return 700.0 * (10.0 ** (mels / 2595.0) - 1.0)


def get_filter_points(fmin, fmax, mel_filter_num, FFT_size, sample_rate=44100):
    """mfcc/core/filterbank.py:15-20."""
    mels = np.linspace(freq_to_mel(fmin), freq_to_mel(fmax), num=mel_filter_num + 2)
    freqs = met_to_freq(mels)
    return np.floor((FFT_size + 1) / sample_rate * freqs).astype(int)


def calc_filters(tab, wsize=9):
    """mfcc/core/filterbank.py:22-34."""
    out = []
    max_acc = 1 << (2 * wsize)
    for i in range(len(tab) - 1):
        diff = int(tab[i + 1]) - int(tab[i]) - 1
        out.append((max_acc // diff) - 1 if diff else max_acc - 1)
    return out


def filterbank(power, width=30, width_output=16, gain=18, sample_rate=16e3, nfft=512,
               ntap=32, state=None):
    """FilterBank, mfcc/core/filterbank.py:65-144, register by register.
    power: [K, nbins] (nbins values per frame, `last` on the final one).
    Python ints carry the 60-bit weight accumulator and the two accumulators of
    `maxvalrange` bits.  Registers persist from frame to frame as in the hardware.
    Returns [K, ntap]."""
    power = np.asarray(power)
    K, nbins = power.shape
    wm = width
    points = get_filter_points(0, sample_rate / 2, ntap, nfft, sample_rate=sample_rate)
    steps = calc_filters(points, wsize=wm)
    maxvalrange = int(math.log2(points[-1] - points[-3])) + width + wm
    rmask = (1 << maxvalrange) - 1
    amask = (1 << (2 * wm)) - 1
    lo = maxvalrange - (gain + width_output)        # o_regb[-(gain+wo):][:wo]
    st = state if state is not None else {"acc": 0, "adr": 0, "rega": 0, "regb": 0}
    out = np.zeros((K, ntap), dtype=np.int64)
    for k in range(K):
        nout = 0
        for b in range(nbins):
            d = int(power[k, b])
            last = b == nbins - 1
            # input side (:90-115)
            w = st["acc"] >> wm
            highest = w == (1 << wm) - 1
            adr = st["adr"]
            c = d * w                                 # Multiplier(width, width_mul)
            if highest or last:
                st["adr"] = 0 if last else adr + 1
                st["acc"] = 0
            else:
                st["acc"] = (st["acc"] + steps[adr]) & amask
            # output side (:119-142); o_data is taken from o_regb before the update
            o_data = (st["regb"] >> lo) & ((1 << width_output) - 1)
            if highest or last:
                if adr != 0:
                    out[k, nout] = o_data
                    nout += 1
                st["regb"] = (st["rega"] + (d << wm)) & rmask
                st["rega"] = 0
            else:
                st["rega"] = (st["rega"] + c) & rmask
                st["regb"] = (st["regb"] + (d << wm) - c) & rmask
        assert nout == ntap, (nout, ntap)
    return out


# ----------------------------------------------------------------------------- a10
def log2fix(m, width=16, width_output=15):
    """Log2Fix -> Log2FixCalc, mfcc/core/log.py:107-139 and :19-104, state by state."""
    precision = width_output - math.ceil(math.log2(width))       # :114
    flat = np.asarray(m, dtype=np.int64).reshape(-1)
    res = np.zeros_like(flat)
    for i, v in enumerate(flat):
        v = int(v)
        x = (1 << precision) if v == 0 else (v << precision)     # :123-126
        b = 1 << (precision - 1)                                 # START
        o = 0
        while x >> (precision + 1):                              # SHIFT-RIGHT :57-62
            x >>= 1
            o += 1 << precision
        cnt = precision - 1                                      # :65
        z = x
        while cnt != 0:                                          # CALC-1 / CALC-2
            c = (z & ((1 << (precision + 1)) - 1)) ** 2          # Multiplier(p+1, p+1)
            if (c >> (2 * precision + 1)) & 1:
                z = c >> (precision + 1)
                o += b
            else:
                z = c >> precision
            cnt -= 1
            b >>= 1
        res[i] = o & ((1 << width_output) - 1)                   # :133
    return res.reshape(np.shape(m))


# ----------------------------------------------------------------------------- a11
def dct_stream(L, nfft=32):
    """DCTStream(width=16, nfft), mfcc/core/dct_stream.py:23-37: every input is
    written four times into a 4*nfft point FFT, v[2n] = 0, v[~(2n)] = L[n],
    v[~(2n+1)] = 0, v[2n+1] = L[n]; real part of bins 0 .. nfft-1."""
    L = np.asarray(L, dtype=np.int64)
    size = 4 * nfft
    v = np.zeros(L.shape[:-1] + (size,), dtype=np.int64)
    for n in range(nfft):
        for c in range(4):
            cf = 4 * n + c
            hi = cf >> 1
            trig = (cf & 1) ^ ((cf >> 1) & 1)
            addr = (~hi) & (size - 1) if trig else hi
            v[..., addr] = L[..., n] if (cf & 1) else 0
    vr, _ = fft_int(v, np.zeros_like(v), size)
    return vr[..., :nfft]


# ----------------------------------------------------------------------------- a12
def discard(c, first=0, count=16):
    """Discard, mfcc/misc/discard.py:27-55."""
    return np.asarray(c)[..., first: first + count]


# ----------------------------------------------------------------------------- a14
def mfcc(x, nfft=512, samplerate=16e3, nfilters=16, nceptrums=16, n_frames=None, taps=False):
    """MFCC.elaborate chain, mfcc/core/mfcc.py:38-104.  x: int16 samples of one
    stream after reset.  Returns [K, nceptrums] int16 (and every stage if taps)."""
    x = np.asarray(x, dtype=np.int64)
    if n_frames is not None:
        # the hosts keep feeding zeros into the *input* of the chain, so the first
        # padded sample still carries the pre-emphasis tail of the last real one
        need = (nfft // 3) * (n_frames - 1) + nfft if n_frames > 0 else 0
        if need > len(x):
            x = np.concatenate([x, np.zeros(need - len(x), dtype=np.int64)])
    y = preemph(x)
    fr = frame(y, nfft, nfft // 3, n_frames)
    win = window(fr, nfft, 8)
    re, im = fft_stream(win, nfft)
    pw = power_spectrum(re, im, 16, 30)
    mel = filterbank(pw, 30, 16, 18, samplerate, nfft, nfilters)
    lg = log2fix(mel, 16, 15)
    dc = dct_stream(lg, nfilters)
    out = discard(dc, 0, nceptrums).astype(np.int16)
    if taps:
        return out, dict(preemph=y, frame=fr, window=win, fft_r=re, fft_i=im, power=pw,
                         filter=mel, log=lg, dct=dc)
    return out
