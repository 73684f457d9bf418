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
