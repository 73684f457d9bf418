"""The reference's OWN host receiver (software/cepstrum.c + serial.c, compiled unmodified
into oracle/_ref/libcepstrum_ref.so by `make -C oracle ref`) as the checker of row f3:
it parses the bytes our UART framing produces and its voice-activity gate is compared
with ours, column by column.  A pipe stands in for the serial port."""
import numpy as np
import pytest

from oracle import ref_host as H
from oracle import ref_recv as RR

pytestmark = pytest.mark.skipif(not RR.available(), reason="oracle/_ref not built and /root/reference absent")


def columns(steps, nc, seed):
    rng = np.random.default_rng(seed)
    cols = rng.integers(-1500, 1500, (steps, nc)).astype(np.int16)
    cols[120:200, 0] = 1900                                   # a burst just over the threshold: 31 * 1900^2 > 1e8
    cols[260:300, 0] = 32767                                   # sum overflows a C int
    cols[330:, 0] = -1790                                      # just under: 31 * 1790^2 < 1e8
    return cols


def test_reference_receiver_parses_our_framing():
    rng = np.random.default_rng(1)
    for nc in (16, 32):
        cols = rng.integers(-32768, 32768, (93, nc)).astype(np.int16)
        cols[5, 3] = np.int16(-23206)                          # 0xa55a inside the payload
        got = RR.get_window(H.uart_frame(cols), nc, 93)
        assert np.array_equal(got, cols)
    # the receiver drops bytes until it sees the magic: junk before the first column is ignored
    cols = rng.integers(-3000, 3000, (10, 16)).astype(np.int16)
    junk = bytes([0x00, 0xa5, 0x11, 0x5a, 0xa5, 0xa5, 0x00, 0x7f])
    assert np.array_equal(RR.get_window(junk + H.uart_frame(cols), 16, 10), cols)


def test_reference_gate_matches_restatement():
    nc, nf, steps = 16, 93, 420
    cols = columns(steps, nc, 2)
    rx = RR.Receiver(nc, nf)
    try:
        assert rx.push(H.uart_frame(cols[:nf])) == nf          # first fill: the whole window
        seen = set()
        for t in range(nf, steps):
            assert rx.push(H.uart_frame(cols[t:t + 1])) == 1
            hist = cols[t + 1 - nf:t + 1]                      # oldest column first
            _p, want = H.vad_power(hist, nc)
            assert rx.active() == want, t
            seen.add(want)
        assert seen == {True, False}
    finally:
        rx.close()


@pytest.mark.gpu
def test_cuda_framing_and_gate_against_reference_receiver():
    import torch
    from mfcc_b200 import post
    nc, nf, steps = 16, 93, 420
    cols = columns(steps, nc, 3)
    cd = torch.from_numpy(cols).to("cuda:0")
    wire = post.uart_frame(cd).cpu().numpy().tobytes()         # what the GPU puts on the line
    per = 2 * (nc + 1)
    assert np.array_equal(RR.get_window(wire[:per * nf], nc, nf), cols[:nf])
    gate = post.VoiceGate(1, nc, nf, device="cuda:0")
    rx = RR.Receiver(nc, nf)
    try:
        for t in range(nf):
            gate.step(cd[t:t + 1])
        assert rx.push(wire[:per * nf]) == nf
        for t in range(nf, steps):
            assert rx.push(wire[per * t:per * (t + 1)]) == 1
            _p, a = gate.step(cd[t:t + 1])
            assert bool(a[0]) == rx.active(), t
    finally:
        rx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("capture", [False, True])
def test_live_session_against_reference_receiver(wav, capture):
    """mfcc_b200.LiveSession = the mic build end to end (MFCC step + framing + gate) for
    several streams; stream 0's serial bytes go into the reference's own receiver, whose
    columns and gate decisions must match, and the columns must equal the batch path."""
    import torch
    import mfcc_b200
    S, nc, nf, steps = 3, 16, 93, 200
    loud = np.clip(wav[20000:20000 + 170 * steps].astype(np.int32) * 3, -32768, 32767).astype(np.int16)
    x = np.stack([loud, wav[60000:60000 + 170 * steps], np.zeros(170 * steps, dtype=np.int16)])
    ses = mfcc_b200.LiveSession(S, nceptrums=nc, device="cuda:0", capture=capture)
    cols, wire, act = [], [], []
    for t in range(steps):
        c, w, _p, a = ses.push(torch.from_numpy(np.ascontiguousarray(x[:, 170 * t:170 * (t + 1)])).to("cuda:0"))
        torch.cuda.synchronize()
        cols.append(c.cpu().numpy().copy()); wire.append(w.cpu().numpy().copy()); act.append(a.cpu().numpy().copy())
    cols = np.stack(cols, axis=1)                                # [S, steps, nc]
    # the streaming state after reset holds 342 zeros: same columns as the batch path on that stream
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=nc, device="cuda:0")
    xb = np.concatenate([np.zeros((S, 342), dtype=np.int16), x], axis=1)
    want = op(torch.from_numpy(xb).to("cuda:0")).cpu().numpy()
    assert np.array_equal(cols, want[:, :steps])
    rx = RR.Receiver(nc, nf)
    try:
        assert rx.push(b"".join(wire[t][0].tobytes() for t in range(nf))) == nf
        assert np.array_equal(rx.store.reshape(nf, nc), cols[0, :nf])
        seen = set()
        for t in range(nf, steps):
            assert rx.push(wire[t][0].tobytes()) == 1
            assert bool(act[t][0]) == rx.active(), t
            seen.add(bool(act[t][0]))
        assert True in seen                                      # the loud stream trips the gate
    finally:
        rx.close()
    assert not np.stack(act)[:, 2].any()                         # the silent stream never does
