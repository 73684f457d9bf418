"""Rows f2 / f3: lifter, standardisation, UART framing and the voice-activity gate
against the reference's own code (tests/golden/post_golden.npz, made by
tests/golden/make_post_golden.py from software/lift.py and sklearn) and against the
restatement in oracle/ref_host.py."""
import os

import numpy as np
import pytest

from oracle import ref_host as H

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "post_golden.npz"))


def test_oracle_lifter_equals_reference_script(gold):
    got = H.to_int16(H.lifter(gold["cepstra"]))
    assert np.array_equal(got, gold["lift_i16"])
    assert np.allclose(H.lifter(gold["cepstra"]), gold["lift_f64"], rtol=0, atol=0)
    assert np.array_equal(H.lifter(gold["cepstra"], L=0), gold["cepstra"])


def test_oracle_scale_equals_sklearn(gold):
    got = H.scale_time(gold["cepstra"])
    assert np.allclose(got, gold["scale_f64"], rtol=0, atol=1e-12)
    # the int16 cast may differ only where the float value sits on an integer
    a, b = H.to_int16(got), gold["scale_i16"]
    near = np.abs(gold["scale_f64"] - np.round(gold["scale_f64"])) < 1e-9
    assert np.array_equal(a[~near], b[~near])
    const = np.tile(np.array([[5, -3]], dtype=np.int16), (10, 1))
    assert not H.scale_time(const).any()                    # zero deviation: divides by 1


def test_oracle_uart_frame_and_vad():
    cols = np.array([[1, -2, 0x1234], [-32768, 32767, 0]], dtype=np.int16)
    b = H.uart_frame(cols)
    assert b.hex() == "a55a0001fffe1234" + "a55a80007fff0000"
    w = np.zeros((93, 16), dtype=np.int16)
    w[31:62, 0] = 2000                                     # the middle third, coefficient 0
    w[:, 1:] = 30000                                       # other coefficients never count
    w[:31, 0] = w[62:, 0] = 32000                          # outside the middle third
    p, act = H.vad_power(w)
    assert p == 31 * 2000 * 2000 and act
    w[31:62, 0] = 1000
    assert H.vad_power(w) == (31 * 1000 * 1000, False)
    w[31:62, 0] = 32767                                    # C int overflow: wraps
    assert H.vad_power(w)[0] == (31 * 32767 ** 2 + 2 ** 31) % 2 ** 32 - 2 ** 31


@pytest.mark.gpu
def test_cuda_lifter_scale_frame(gold, wav):
    import torch
    import mfcc_b200
    from mfcc_b200 import post
    dev = "cuda:0"
    c = torch.from_numpy(gold["cepstra"]).to(dev)
    assert np.array_equal(post.lifter(c).cpu().numpy(), gold["lift_i16"])
    assert np.array_equal(post.lifter(c, L=0).cpu().numpy(), gold["cepstra"])
    got = post.scale(c).cpu().numpy()
    near = np.abs(gold["scale_f64"] - np.round(gold["scale_f64"])) < 1e-9
    assert np.array_equal(got[~near], gold["scale_i16"][~near]) and near.mean() < 0.01
    # batched, float input, constant column
    xb = torch.stack([c, c.flip(0)]).float()
    xb[1, :, 3] = 7.0
    gb = post.scale(xb).cpu().numpy()
    assert np.array_equal(gb[0][~near], gold["scale_i16"][~near]) and not gb[1, :, 3].any()
    want1 = H.to_int16(H.scale_time(xb[1].cpu().numpy()))
    f1 = H.scale_time(xb[1].cpu().numpy())
    ok = np.abs(f1 - np.round(f1)) > 1e-9
    assert np.array_equal(gb[1][ok], want1[ok])
    cols = c[:50, :16].contiguous()
    assert post.uart_frame(cols).cpu().numpy().tobytes() == H.uart_frame(cols.cpu().numpy())
    # on top of the real path: cepstra of the bundled clip, lifted
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device=dev)
    cep = op(torch.from_numpy(wav[:30000]).to(dev))
    assert np.array_equal(post.lifter(cep).cpu().numpy(), H.to_int16(H.lifter(cep.cpu().numpy())))


@pytest.mark.gpu
def test_cuda_voice_gate_matches_host_ring():
    import torch
    from mfcc_b200 import post
    rng = np.random.default_rng(9)
    S, nc, nfr, steps = 6, 16, 93, 240
    cols = rng.integers(-1500, 1500, (steps, S, nc)).astype(np.int16)
    cols[60:130, 2, 0] = 2000                              # stream 2 speaks for a while
    cols[:, 4, 0] = 32767                                  # stream 4: sum overflows a C int
    gate = post.VoiceGate(S, nc, nfr, device="cuda:0")
    hist = np.zeros((S, nfr, nc), dtype=np.int16)          # oldest column first
    seen_active = np.zeros(S, dtype=bool)
    for t in range(steps):
        p, a = gate.step(torch.from_numpy(cols[t]).to("cuda:0"))
        hist = np.concatenate([hist[:, 1:], cols[t][:, None, :]], axis=1)
        for s in range(S):
            wp, wa = H.vad_power(hist[s], nc)
            assert int(p[s]) == wp and bool(a[s]) == wa, (t, s)
        seen_active |= a.cpu().numpy()
    assert seen_active[2] and not seen_active[0]
    gate.reset()
    assert int(gate.step(torch.zeros((S, nc), dtype=torch.int16, device="cuda:0"))[0].abs().sum()) == 0
    with pytest.raises(ValueError):
        post.VoiceGate(4, 16, 92)
