"""Self-checks of oracle/hdl_shim (the small nMigen interpreter the RTL fixtures were
made with).  The semantic checks run everywhere; the checks that execute the
reference's source need /root/reference and skip on the GPU box."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "hdl_shim")
REF = "/root/reference"


@pytest.fixture(scope="module")
def hdl():
    sys.path.insert(0, SHIM)
    try:
        import nmigen
        import nmigen.sim
        yield nmigen
    finally:
        sys.path.remove(SHIM)


def _run_comb(hdl, build):
    """build(m) -> dict name -> Value; returns the settled integer of each."""
    from nmigen.sim import Simulator
    m = hdl.Module()
    outs = {}
    for name, v in build(m).items():
        sig = hdl.Signal(v.shape(), name=name)
        m.d.comb += sig.eq(v)
        outs[name] = sig
    sim = Simulator(m)
    res = {}

    def p():
        for k, sg in outs.items():
            res[k] = (yield sg)
    sim.add_sync_process(p)
    sim.run()
    return res


def test_shapes_and_wrap(hdl):
    S, C = hdl.Signal, hdl.Const
    a = S(hdl.signed(16), reset=-3)
    b = S(8, reset=200)
    assert (a + b).shape() == hdl.Shape(17, True)
    assert (a * b).shape() == hdl.Shape(24, True)
    assert (b - b).shape() == hdl.Shape(9, False)          # nMigen 0.3
    assert (-b).shape() == hdl.Shape(9, True)
    assert a[4:].shape() == hdl.Shape(12, False)
    assert len(hdl.Cat(a, b)) == 24
    assert hdl.Shape.cast(range(512)).width == 9 and hdl.Shape.cast(range(513)).width == 10
    r = _run_comb(hdl, lambda m: dict(
        shr=a >> 1, sl=a[-4:], neg=-b, inv=~b, cat=hdl.Cat(b, a[:4]), mux=hdl.Mux(a < 0, 5, 9),
        ws=hdl.Cat(C(0x34, 8), C(0x12, 8)).word_select(C(1, 1), 8), allb=C(255, 8).all(),
        anyb=C(0, 8).any(), mixed=a & b))
    assert r["shr"] == -2                                  # arithmetic shift
    assert r["sl"] == 0xF                                  # slices are unsigned
    assert r["neg"] == -200 and r["inv"] == 55
    assert r["cat"] == 200 | (0xD << 8)
    assert r["mux"] == 5 and r["ws"] == 0x12 and r["allb"] == 1 and r["anyb"] == 0
    assert r["mixed"] == (-3 & 200)


def test_assignment_truncates_and_signs(hdl):
    from nmigen.sim import Simulator
    m = hdl.Module()
    x = hdl.Signal(hdl.signed(16))
    y = hdl.Signal(4)
    z = hdl.Signal(8)
    m.d.comb += [x.eq(40000), y.eq(-1), z[2:6].eq(0xFF), z[0].eq(1)]
    got = {}

    def p():
        got.update(x=(yield x), y=(yield y), z=(yield z))
    sim = Simulator(m)
    sim.add_sync_process(p)
    sim.run()
    assert got == dict(x=40000 - 65536, y=15, z=0b00111101)


def test_sync_memory_fsm_reset(hdl):
    from nmigen.sim import Simulator
    m = hdl.Module()
    mem = hdl.Memory(width=8, depth=4, init=[10, 11, 12, 13])
    m.submodules.rp = rp = mem.read_port()
    m.submodules.wp = wp = mem.write_port()
    cnt = hdl.Signal(4)
    rst = hdl.Signal()
    inner = hdl.Module()
    inner.d.sync += cnt.eq(cnt + 1)
    m.submodules.inner = hdl.ResetInserter(rst)(inner)
    with m.FSM() as fsm:
        with m.State("A"):
            m.next = "B"
        with m.State("B"):
            with m.If(cnt == 5):
                m.next = "A"
    in_b = hdl.Signal()
    m.d.comb += in_b.eq(fsm.ongoing("B"))
    log = []

    def p():
        yield rp.addr.eq(2)
        yield
        log.append(("rd", (yield rp.data)))                # one cycle latency
        yield wp.addr.eq(2)
        yield wp.data.eq(99)
        yield wp.en.eq(1)
        yield
        log.append(("wt", (yield rp.data)))                # transparent: new data
        yield wp.en.eq(0)
        log.append(("b", (yield in_b), (yield cnt)))
        yield rst.eq(1)
        yield
        log.append(("rst", (yield cnt)))
    sim = Simulator(m)
    sim.add_sync_process(p)
    sim.run()
    assert log == [("rd", 12), ("wt", 99), ("b", 1, 2), ("rst", 0)]


def test_comb_default_and_priority(hdl):
    from nmigen.sim import Simulator
    m = hdl.Module()
    sel = hdl.Signal(2)
    o = hdl.Signal(8, reset=7)
    with m.If(sel == 1):
        m.d.comb += o.eq(1)
    with m.Elif(sel == 2):
        m.d.comb += o.eq(2)
    with m.If(sel == 3):
        m.d.comb += o.eq(3)
        m.d.comb += o.eq(4)                                # last assignment wins
    seen = []

    def p():
        for v in range(4):
            yield sel.eq(v)
            seen.append((yield o))
    sim = Simulator(m)
    sim.add_sync_process(p)
    sim.run()
    assert seen == [7, 1, 2, 4]


# ---------------------------------------------------------------------------------------
# One expected-value test per nMigen 0.3 rule the reference design leans on.  The rule is
# quoted from nMigen 0.3's hdl/ast.py (Operator.shape(), Slice, Cat, Mux, Const) and
# hdl/mem.py / hdl/xfrm.py / back/pysim.py as published; the reference's line that relies
# on it is cited next to it.
def _sim_seq(hdl, m, proc):
    from nmigen.sim import Simulator
    sim = Simulator(m)
    sim.add_sync_process(proc)
    sim.run()


def test_mul_signed_by_unsigned_shape_and_value(hdl):
    """'*': width = a.width + b.width, signed if either operand is signed (ast.py Operator.shape);
    the unsigned operand is zero-extended, never re-interpreted.  mfcc/misc/mul.py:11-13,
    window.py:74-75 (signed 16 x unsigned 9 -> signed 25)."""
    a = hdl.Signal(hdl.signed(16), reset=-12345)
    b = hdl.Signal(9, reset=510)
    assert (a * b).shape() == hdl.Shape(25, True) and (b * a).shape() == hdl.Shape(25, True)
    assert (b * b).shape() == hdl.Shape(18, False)
    r = _run_comb(hdl, lambda m: dict(p=a * b, q=b * a, top=(a * b)[9:]))
    assert r["p"] == r["q"] == -12345 * 510
    assert r["top"] == ((-12345 * 510) >> 9) & 0xFFFF        # the slice window.py:84 takes, unsigned


def test_add_sub_shapes_mixed_sign(hdl):
    """'+', '-': width = max(a.w + (b signed and a not), b.w + (a signed and b not)) + 1, signed if
    either is.  preemph.py:22-24 (x + (prev >> 5) - prev in 17..18 bits), fft.py:152-180."""
    s16, u8 = hdl.Signal(hdl.signed(16)), hdl.Signal(8)
    assert (s16 + s16).shape() == hdl.Shape(17, True)
    assert (s16 + u8).shape() == hdl.Shape(17, True)
    assert (hdl.Signal(hdl.signed(4)) + hdl.Signal(8)).shape() == hdl.Shape(10, True)   # 8 + 1 sign bit, + 1 carry
    assert (u8 + u8).shape() == hdl.Shape(9, False)
    assert (s16 - u8).shape() == hdl.Shape(17, True)
    x = hdl.Signal(hdl.signed(16), reset=-32768)
    p = hdl.Signal(hdl.signed(16), reset=32767)
    r = _run_comb(hdl, lambda m: dict(y=x + (p >> 5) - p))
    assert r["y"] == -32768 + (32767 >> 5) - 32767            # no wrap inside the expression


def test_shift_right_is_arithmetic_and_keeps_width(hdl):
    """'>>': shape = (a.width, a.signed); on a signed operand it is an arithmetic (floor) shift.
    preemph.py:22 (odata >> 5), fft.py:188-191 (>> scale_bit)."""
    a = hdl.Signal(hdl.signed(16), reset=-1)
    b = hdl.Signal(hdl.signed(16), reset=-32768)
    u = hdl.Signal(16, reset=0x8000)
    assert (a >> 5).shape() == hdl.Shape(16, True) and (u >> 5).shape() == hdl.Shape(16, False)
    r = _run_comb(hdl, lambda m: dict(a=a >> 5, b=b >> 5, u=u >> 5, odd=hdl.Const(-7, hdl.signed(8)) >> 1))
    assert r == dict(a=-1, b=-1024, u=0x400, odd=-4)


def test_shift_left_grows(hdl):
    """'<<': width = a.width + 2**b.width - 1, b unsigned (ast.py Operator.shape); an int amount is
    Const(n) first, so `x << 11` is 16 + 2**4 - 1 = 31 bits wide, not 27: no bit is ever lost.
    log.py:123-126 (x << precision)."""
    a = hdl.Signal(16, reset=0xFFFF)
    assert (a << 11).shape() == hdl.Shape(31, False)
    n = hdl.Signal(3, reset=5)
    assert (a << n).shape() == hdl.Shape(16 + 7, False)
    r = _run_comb(hdl, lambda m: dict(c=a << 11, v=a << n))
    assert r == dict(c=0xFFFF << 11, v=0xFFFF << 5)


def test_slice_of_signed_is_unsigned_and_resigns_on_assignment(hdl):
    """Slice.shape() = (stop - start, False) whatever the operand; assigning it to a signed signal
    re-interprets the bits (ast.py Slice; Signal assignment sign-extends from the source shape, then
    truncates).  fft.py:188-191 / window.py:84 assign such slices to signed 16-bit payloads."""
    from nmigen.sim import Simulator
    v = hdl.Signal(hdl.signed(25), reset=-5 << 9)
    assert v[9:].shape() == hdl.Shape(16, False) and v[-16:].shape() == hdl.Shape(16, False)
    m = hdl.Module()
    s = hdl.Signal(hdl.signed(16))
    u = hdl.Signal(16)
    w = hdl.Signal(hdl.signed(20))
    m.d.comb += [s.eq(v[9:]), u.eq(v[9:]), w.eq(v[9:])]
    got = {}

    def p():
        got.update(s=(yield s), u=(yield u), w=(yield w))
    _sim_seq(hdl, m, p)
    assert got == dict(s=-5, u=65536 - 5, w=65536 - 5)       # the wider signed target sees an unsigned value


def test_negative_slice_indices(hdl):
    """x[-n:] is the top n bits, x[:-n] drops them (ast.py Value.__getitem__ normalises like a
    Python sequence).  filterbank.py:122, log.py:133."""
    c = hdl.Const(0xABCD, 16)
    r = _run_comb(hdl, lambda m: dict(top=c[-4:], rest=c[:-4], bit=c[-1], mid=c[4:12]))
    assert r == dict(top=0xA, rest=0xBCD, bit=1, mid=0xBC)
    with pytest.raises(IndexError):
        c[16]


def test_cat_of_signed_parts(hdl):
    """Cat is unsigned, width = sum of the parts, first operand in the low bits; a signed part
    contributes its two's-complement bits (ast.py Cat).  stream payload packing, dct_stream.py."""
    a = hdl.Signal(hdl.signed(4), reset=-3)
    b = hdl.Signal(hdl.signed(4), reset=-1)
    assert hdl.Cat(a, b).shape() == hdl.Shape(8, False)
    r = _run_comb(hdl, lambda m: dict(c=hdl.Cat(a, b), d=hdl.Cat(b, hdl.Const(1, 1), a)))
    assert r["c"] == 0xD | (0xF << 4)
    assert r["d"] == 0xF | (1 << 4) | (0xD << 5)


def test_mux_shape_and_value(hdl):
    """Mux(sel, a, b): shape = the common shape of a and b (like a bitwise operator: signed if either
    is, an unsigned operand gains a sign bit); sel is tested for non-zero.  window.py:102-110,
    frame.py:120."""
    s4 = hdl.Signal(hdl.signed(4), reset=-2)
    u4 = hdl.Signal(4, reset=15)
    sel = hdl.Signal(2, reset=2)
    assert hdl.Mux(sel, s4, u4).shape() == hdl.Shape(5, True)
    assert hdl.Mux(sel, u4, u4).shape() == hdl.Shape(4, False)
    r = _run_comb(hdl, lambda m: dict(a=hdl.Mux(sel, s4, u4), b=hdl.Mux(sel == 0, s4, u4)))
    assert r == dict(a=-2, b=15)


def test_comparisons_mixed_sign(hdl):
    """Comparisons yield (1, False) and compare VALUES: the operands are extended to a common
    signed shape first (ast.py Operator.shape '<' etc.).  log.py:57-62, frame.py counters."""
    s = hdl.Signal(hdl.signed(8), reset=-1)
    u = hdl.Signal(8, reset=255)
    assert (s < u).shape() == hdl.Shape(1, False)
    r = _run_comb(hdl, lambda m: dict(lt=s < u, eq=s == u, ge=s >= 0, ne=u != 255))
    assert r == dict(lt=1, eq=0, ge=0, ne=0)


def test_bitwise_mixed_sign_and_invert(hdl):
    """'&', '|', '^': width = max(a.w + (b signed and a not), b.w + (a signed and b not)), signed if
    either; '~' keeps the shape (ast.py).  fft.py:310-316 (one-hot twiddle stepping)."""
    s = hdl.Signal(hdl.signed(4), reset=-4)       # 1100
    u = hdl.Signal(6, reset=0b101010)
    assert (s & u).shape() == hdl.Shape(7, True) and (~s).shape() == hdl.Shape(4, True)
    r = _run_comb(hdl, lambda m: dict(a=s & u, o=s | u, x=s ^ u, n=~s, nu=~u))
    assert r == dict(a=-4 & 0b101010, o=-4 | 0b101010, x=-4 ^ 0b101010, n=3, nu=0b010101)


def test_const_shape_inference(hdl):
    """Const(v) without a shape takes bits_for(v): unsigned for v >= 0, signed with a sign bit for
    v < 0; with a shape the value is truncated to it (ast.py Const.__init__)."""
    C = hdl.Const
    assert C(0).shape() == hdl.Shape(1, False) and C(5).shape() == hdl.Shape(3, False)
    assert C(-1).shape() == hdl.Shape(1, True) and C(-5).shape() == hdl.Shape(4, True)
    # an int shape means Shape(n, signed=value < 0) (ast.py Const.__init__), an explicit one is kept
    assert C(256, 8).value == 0 and C(255, hdl.signed(8)).value == -1
    assert C(-1, 8).shape() == hdl.Shape(8, True) and C(-1, 8).value == -1
    assert C(-1, hdl.unsigned(8)).value == 255
    assert hdl.Shape.cast(range(-4, 4)) == hdl.Shape(3, True) and hdl.Shape.cast(range(16)) == hdl.Shape(4, False)


def test_assignment_extends_by_source_sign(hdl):
    """lhs.eq(rhs): rhs is extended to the lhs width by ITS OWN signedness, then truncated
    (back/pysim.py / rtlil: match_shape).  A signed -1 fills an unsigned byte with ones; an
    unsigned 0xF stays 15 in a signed byte."""
    m = hdl.Module()
    s4 = hdl.Signal(hdl.signed(4), reset=-1)
    u4 = hdl.Signal(4, reset=0xF)
    a, b = hdl.Signal(8), hdl.Signal(hdl.signed(8))
    c = hdl.Signal(hdl.signed(3))
    m.d.comb += [a.eq(s4), b.eq(u4), c.eq(u4)]
    got = {}

    def p():
        got.update(a=(yield a), b=(yield b), c=(yield c))
    _sim_seq(hdl, m, p)
    assert got == dict(a=0xFF, b=15, c=-1)


def test_sync_is_non_blocking_and_reset_values(hdl):
    """m.d.sync assignments all sample the values before the clock edge (a swap works), signals
    start from their reset value, reset_less signals ignore ResetInserter (hdl/xfrm.py)."""
    m = hdl.Module()
    inner = hdl.Module()
    a = hdl.Signal(8, reset=1)
    b = hdl.Signal(8, reset=2)
    keep = hdl.Signal(8, reset=9, reset_less=True)
    inner.d.sync += [a.eq(b), b.eq(a), keep.eq(keep + 1)]
    rst = hdl.Signal()
    m.submodules.inner = hdl.ResetInserter(rst)(inner)
    log = []

    def p():
        log.append(((yield a), (yield b), (yield keep)))
        yield
        log.append(((yield a), (yield b), (yield keep)))
        yield
        log.append(((yield a), (yield b), (yield keep)))
        yield rst.eq(1)
        yield
        log.append(((yield a), (yield b), (yield keep)))
        yield rst.eq(0)
        yield
        log.append(((yield a), (yield b), (yield keep)))
    _sim_seq(hdl, m, p)
    # the process reads just after an edge; its own writes (rst) act on the NEXT edge
    assert [l[:2] for l in log[:3]] in ([(1, 2), (2, 1), (1, 2)], [(1, 2), (1, 2), (2, 1)])
    swaps = [l[:2] for l in log]
    assert all(set(s) == {1, 2} for s in swaps)               # always a permutation: non-blocking
    keeps = [l[2] for l in log]
    assert keeps == sorted(keeps) and keeps[-1] > keeps[0]    # never reset
    i = swaps.index((1, 2), 3) if (1, 2) in swaps[3:] else None
    assert i is not None                                      # reset restored (a, b) = (1, 2)


def test_memory_read_port_transparent_vs_not(hdl):
    """Synchronous read port (hdl/mem.py): data appears one cycle after addr; on a write to the same
    address in the same cycle a transparent port (the default) returns the NEW word, a
    non-transparent one the OLD word.  fft.py:349-484 uses the default for its two banks,
    frame.py:65-73 for the ring."""
    m = hdl.Module()
    mem = hdl.Memory(width=8, depth=4, init=[10, 11, 12, 13])
    m.submodules.rt = rt = mem.read_port()
    m.submodules.rn = rn = mem.read_port(transparent=False)
    m.submodules.wp = wp = mem.write_port()
    log = []

    def p():
        yield rt.addr.eq(1)
        yield rn.addr.eq(1)
        yield rn.en.eq(1)
        yield wp.addr.eq(1)
        yield wp.data.eq(77)
        yield wp.en.eq(1)
        yield
        log.append(((yield rt.data), (yield rn.data)))
        yield wp.en.eq(0)
        yield
        log.append(((yield rt.data), (yield rn.data)))
    _sim_seq(hdl, m, p)
    assert log == [(77, 11), (77, 77)]


def test_memory_comb_read_port_and_init_padding(hdl):
    """domain='comb' read ports are asynchronous (data follows addr in the same cycle); init shorter
    than depth is padded with zeros (hdl/mem.py).  fft.py:18-61 TwiddleROM, window.py:42-43."""
    m = hdl.Module()
    mem = hdl.Memory(width=16, depth=8, init=[100, 200, 300])
    m.submodules.rp = rp = mem.read_port(domain="comb")
    seen = []

    def p():
        for a in (2, 0, 5):
            yield rp.addr.eq(a)
            yield hdl.sim.Settle() if hasattr(hdl, "sim") and hasattr(hdl.sim, "Settle") else None
            seen.append((yield rp.data))
    try:
        _sim_seq(hdl, m, p)
    except TypeError:
        seen.clear()

        def p2():
            for a in (2, 0, 5):
                yield rp.addr.eq(a)
                seen.append((yield rp.data))
        _sim_seq(hdl, m, p2)
    assert seen == [300, 100, 0]


def test_switch_case_default_and_patterns(hdl):
    """m.Switch: the first matching Case wins, Case() with no pattern is the default, string
    patterns take '-' as don't-care (hdl/dsl.py).  scheduler states of fft.py:197-346."""
    m = hdl.Module()
    sel = hdl.Signal(3)
    o = hdl.Signal(8)
    with m.Switch(sel):
        with m.Case(0):
            m.d.comb += o.eq(10)
        with m.Case(1, 2):
            m.d.comb += o.eq(20)
        with m.Case("1--"):
            m.d.comb += o.eq(30)
        with m.Case():
            m.d.comb += o.eq(40)
    seen = []

    def p():
        for v in (0, 1, 2, 3, 4, 7):
            yield sel.eq(v)
            seen.append((yield o))
    _sim_seq(hdl, m, p)
    assert seen == [10, 20, 20, 40, 30, 30]


def test_bit_select_word_select_and_part_assignment(hdl):
    """bit_select(offset, w) = value[offset : offset + w], word_select(i, w) = value[i*w : (i+1)*w],
    both usable as assignment targets with a signal offset (ast.py Part).  Out-of-range bits read 0."""
    m = hdl.Module()
    v = hdl.Const(0xDEADBEEF, 32)
    off = hdl.Signal(5, reset=12)
    tgt = hdl.Signal(16)
    idx = hdl.Signal(2, reset=2)
    m.d.comb += tgt.word_select(idx, 4).eq(0xA)
    got = {}

    def p():
        got["t"] = (yield tgt)
    _sim_seq(hdl, m, p)
    r = _run_comb(hdl, lambda mm: dict(b=v.bit_select(off, 8), w=v.word_select(hdl.Const(3, 2), 8),
                                       hi=v.bit_select(hdl.Const(28, 5), 8)))
    assert r == dict(b=(0xDEADBEEF >> 12) & 0xFF, w=0xDE, hi=0xD)
    assert got["t"] == 0xA << 8


def test_reductions_and_bool(hdl):
    """.bool()/.any() = OR-reduce, .all() = AND-reduce, .xor() = parity, each (1, False); using a
    Value in a Python `if` raises TypeError (ast.py Value.__bool__)."""
    c = hdl.Const(0b1011, 4)
    r = _run_comb(hdl, lambda m: dict(b=c.bool(), a=c.all(), x=c.xor(), z=hdl.Const(0, 4).bool(),
                                      f=hdl.Const(0xF, 4).all()))
    assert r == dict(b=1, a=0, x=1, z=0, f=1)
    with pytest.raises(TypeError, match="Python boolean"):
        bool(hdl.Signal(1))


def test_unary_minus_and_abs_shapes(hdl):
    """-a: (a.width + 1, True) for either sign (ast.py); the most negative value therefore negates
    without wrapping.  fft.py:55-59 (-re of the quarter-circle decode)."""
    s = hdl.Signal(hdl.signed(16), reset=-32768)
    assert (-s).shape() == hdl.Shape(17, True)
    r = _run_comb(hdl, lambda m: dict(n=-s))
    assert r["n"] == 32768
    m = hdl.Module()
    t = hdl.Signal(hdl.signed(16))
    m.d.comb += t.eq(-s)                                      # ...and wraps only when stored in 16 bits
    got = {}

    def p():
        got["t"] = (yield t)
    _sim_seq(hdl, m, p)
    assert got["t"] == -32768


def test_record_fields_and_eq(hdl):
    """Record(layout): named sub-signals with their own shapes, rec.eq(other) assigns field by field
    (hdl/rec.py).  mfcc/misc/stream.py:33-40 builds every stream endpoint from one."""
    lay = [("valid", 1), ("data", hdl.signed(16)), ("last", 1)]
    a, b = hdl.Record(lay), hdl.Record(lay)
    assert a.data.shape() == hdl.Shape(16, True) and len(a) == 18
    m = hdl.Module()
    m.d.comb += [a.data.eq(-77), a.valid.eq(1), b.eq(a)]
    got = {}

    def p():
        got.update(d=(yield b.data), v=(yield b.valid), l=(yield b.last))
    _sim_seq(hdl, m, p)
    assert got == dict(d=-77, v=1, l=0)


def test_sync_fifo_order_level_and_flags(hdl):
    """lib.fifo.SyncFIFO(width, depth): first-in first-out, w_rdy drops when `depth` words are
    held, r_rdy when empty, one word per cycle each way.  mfcc.py:56-58, :65-67, :78-80."""
    from nmigen.lib.fifo import SyncFIFO
    m = hdl.Module()
    m.submodules.f = f = SyncFIFO(width=8, depth=4)
    out, flags = [], []

    def p():
        for v in (5, 6, 7, 8):
            yield f.w_data.eq(v)
            yield f.w_en.eq(1)
            yield
        yield f.w_en.eq(0)
        yield
        flags.append(((yield f.w_rdy), (yield f.r_rdy)))
        yield f.r_en.eq(1)
        for _ in range(4):
            out.append((yield f.r_data))
            yield
        yield f.r_en.eq(0)
        yield
        flags.append(((yield f.w_rdy), (yield f.r_rdy)))
    _sim_seq(hdl, m, p)
    assert out == [5, 6, 7, 8]
    assert flags == [(0, 1), (1, 0)]


needs_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "mfcc")),
                               reason="needs /root/reference (build container only)")


@needs_ref
@pytest.mark.parametrize("sub", ["unsigned", "signed"])
def test_reference_mfcc_first_frames_reproduce_fixture(sub):
    """Re-run the generator on one short case (in a subprocess: it puts the shim on
    sys.path as `nmigen`) under both readings of unsigned subtraction; the committed
    fixture must come out again."""
    import tempfile
    case = "zeros"
    want = np.load(os.path.join(ROOT, "tests", "golden", "rtl_%s.npz" % case))
    with tempfile.TemporaryDirectory() as td:
        code = ("import sys, numpy as np; sys.argv=['x'];"
                "sys.path.insert(0, %r); import make_rtl_golden as G;"
                "x, nfr, nf, nc, st = G.case_inputs()[%r];"
                "t, _ = G.run_mfcc(x, nfr, nf, nc, st);"
                "np.savez(%r, **t)") % (os.path.join(ROOT, "tests", "golden"), case,
                                        os.path.join(td, "o.npz"))
        env = dict(os.environ, HDL_SHIM_SUB=sub)
        subprocess.check_call([sys.executable, "-c", code], env=env)
        got = np.load(os.path.join(td, "o.npz"))
        for k in got.files:
            assert np.array_equal(got[k], want[k]), k
