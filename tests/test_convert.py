"""Row f1: the wav2mfcc directory converter (mfcc_b200/convert.py) against the
reference host tool's protocol restated in oracle/ref_host.py."""
import os
import wave

import numpy as np
import pytest

from mfcc_b200 import convert
from mfcc_b200.config import MFCCConfig
from oracle import ref_host as H


def write_wav(path, x, rate=16000, width=2, channels=1):
    with wave.open(path, "wb") as w:
        w.setnchannels(channels)
        w.setsampwidth(width)
        w.setframerate(rate)
        w.writeframes(np.asarray(x).astype("<i2" if width == 2 else "u1").tobytes())


@pytest.mark.parametrize("n", [0, 1, 100, 511, 512, 513, 681, 682, 683, 852, 5000])
def test_host_round_count(n):
    stream, rounds = H.host_rounds(np.arange(n) % 1000)
    assert rounds == MFCCConfig().num_frames_host(n)
    assert len(stream) == 512 + 170 * (rounds - 1)
    assert np.array_equal(stream[:n], (np.arange(n) % 1000).astype(np.int16)) and not stream[n:].any()


def test_read_wav_checks(tmp_path):
    x = (np.arange(2000) * 7 % 3000 - 1500).astype(np.int16)
    p = str(tmp_path / "ok.wav")
    write_wav(p, x)
    assert np.array_equal(convert.read_wav(p), x)
    write_wav(str(tmp_path / "rate.wav"), x, rate=8000)
    with pytest.raises(ValueError, match="Unexpected samplerate: 8000"):
        convert.read_wav(str(tmp_path / "rate.wav"))
    write_wav(str(tmp_path / "w8.wav"), np.abs(x) % 255, width=1)
    with pytest.raises(ValueError, match="Unexpected samplesize: 1"):
        convert.read_wav(str(tmp_path / "w8.wav"))
    write_wav(str(tmp_path / "st.wav"), x, channels=2)
    with pytest.raises(ValueError, match="channel"):
        convert.read_wav(str(tmp_path / "st.wav"))
    (tmp_path / "junk.wav").write_bytes(b"not a wav file at all")
    with pytest.raises(ValueError, match="Unexpected format"):
        convert.read_wav(str(tmp_path / "junk.wav"))


def test_find_and_plan(tmp_path):
    (tmp_path / "a" / "b").mkdir(parents=True)
    for name in ("z.wav", "a/x.wav", "a/b/y.wav", "a/skip.txt", "a/wav"):
        (tmp_path / name).write_bytes(b"")
    got = [os.path.relpath(p, tmp_path) for p in convert.find_wavs(str(tmp_path))]
    assert got == ["a/b/y.wav", "a/x.wav", "z.wav"]
    assert convert.mfcc_path("/d/clip.wav") == "/d/clip.mfcc"
    lengths = [16000, 100, 48000, 16000, 0, 32000]
    b = convert.plan_batches(lengths, batch_bytes=2 * 2 * 16000)
    assert sorted(i for g in b for i in g) == list(range(6))
    for g in b:
        assert len(g) * max(max(lengths[i] for i in g), 8) * 2 <= 2 * 2 * 16000 or len(g) == 1


@pytest.mark.gpu
def test_convert_dir_matches_host_protocol(tmp_path, wav):
    import mfcc_b200
    from oracle import ref_c as RC
    cfg = RC.RefConfig(nfilters=32, nceptrums=32)
    rng = np.random.default_rng(11)
    clips = {"s/one.wav": wav[:16000], "s/two.wav": wav[40000:40000 + 512 + 170 * 9],
             "t/u/short.wav": wav[:300], "t/exact.wav": wav[1000:1000 + 512],
             "t/noise.wav": rng.integers(-32768, 32768, 7777).astype(np.int16),
             "empty.wav": np.zeros(0, dtype=np.int16), "long.wav": wav[:70001]}
    for rel, x in clips.items():
        os.makedirs(os.path.dirname(str(tmp_path / rel)), exist_ok=True)
        write_wav(str(tmp_path / rel), x)
    write_wav(str(tmp_path / "t" / "bad_rate.wav"), wav[:1000], rate=44100)
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device="cuda:0")
    done, rejected = convert.convert_dir(str(tmp_path), op=op, batch_bytes=100000)
    assert len(done) == len(clips) and len(rejected) == 1 and "44100" in rejected[0][1]
    for rel, x in clips.items():
        want = H.wav2mfcc(x, lambda s, k: RC.run(cfg, s, n_frames=k))
        got = np.fromfile(str(tmp_path / rel)[:-3] + "mfcc", dtype=np.int16).reshape(-1, 32)
        assert got.shape == want.shape == (MFCCConfig().num_frames_host(len(x)), 32)
        assert np.array_equal(got, want), rel
    assert not os.path.exists(str(tmp_path / "t" / "bad_rate.mfcc"))


@pytest.mark.gpu
def test_convert_dir_librosa_writes_spec_and_sklearn(tmp_path, wav):
    """software/genlibrosa.py:15-28: <dir>/*/*.wav -> .spec and .sklearn, int16 [32][frames]."""
    from oracle import ref_librosa as L
    rng = np.random.default_rng(12)
    clips = {"a/one.wav": wav[:16000], "a/two.wav": wav[50000:66000], "b/three.wav": wav[90000:95000],
             "b/noise.wav": np.clip(rng.normal(0, 3000, 16000), -32768, 32767).astype(np.int16)}
    for rel, x in clips.items():
        os.makedirs(os.path.dirname(str(tmp_path / rel)), exist_ok=True)
        write_wav(str(tmp_path / rel), x)
    write_wav(str(tmp_path / "top_level_is_not_globbed.wav"), wav[:1000])
    write_wav(str(tmp_path / "b" / "stereo.wav"), wav[:1000], channels=2)
    done, rejected = convert.convert_dir_librosa(str(tmp_path), device="cuda:0")
    assert len(done) == len(clips) and len(rejected) == 1
    for rel, x in clips.items():
        base = str(tmp_path / rel)[:-4]
        K = 1 + len(x) // 170                                   # 95 for one second: NFRAMES of software/view.py:37
        spec = np.fromfile(base + ".spec", dtype=np.int16).reshape(32, K)
        skl = np.fromfile(base + ".sklearn", dtype=np.int16).reshape(32, K)
        ref = L.mfcc(x)                                          # [K][32] float64
        for got, want in ((spec, ref.T), (skl, H.scale_time(ref).T)):
            tr = np.trunc(want)
            near = np.abs(want - np.round(want)) < 1e-2          # float32 noise may cross a truncation boundary
            assert np.array_equal(got[~near], tr[~near].astype(np.int16)), rel
            assert np.abs(got - tr).max() <= 1
    assert 1 + 16000 // 170 == 95
    assert not os.path.exists(str(tmp_path / "top_level_is_not_globbed.spec"))


# ---------------------------------------------------------------- the reference's own tool
from oracle import ref_wav2mfcc as RW  # noqa: E402

needs_ref_tool = pytest.mark.skipif(not RW.available(), reason="oracle/_ref not built and /root/reference absent")


def _tool_clips(wav):
    rng = np.random.default_rng(21)
    return {"s/one.wav": wav[:4000], "s/two.wav": wav[40000:40000 + 512 + 170 * 9],     # ends on a round boundary
            "t/u/short.wav": wav[:300], "t/exact.wav": wav[1000:1000 + 512],
            "t/noise.wav": rng.integers(-32768, 32768, 2777).astype(np.int16)}


def _run_reference_tool(root, wav):
    from oracle import ref_c as RC
    cfg = RC.RefConfig(nfilters=32, nceptrums=32)
    clips = _tool_clips(wav)
    for rel, x in clips.items():
        os.makedirs(os.path.dirname(os.path.join(root, rel)), exist_ok=True)
        write_wav(os.path.join(root, rel), x)
    model = RW.FpgaModel(lambda s, k: RC.run(cfg, s, n_frames=k))
    assert RW.run_tool(root, model) == 0
    return clips, model, cfg


@needs_ref_tool
def test_reference_tool_agrees_with_restated_protocol(tmp_path, wav):
    """software/main.c itself (compiled, USB replaced by a model of the FPGA top backed by
    the C oracle) writes the .mfcc files; the restated host protocol (oracle/ref_host.py)
    and the frame-count formula must give the same bytes."""
    from oracle import ref_c as RC
    clips, model, cfg = _run_reference_tool(str(tmp_path), wav)
    assert model.resets == 1 + len(clips)                     # mfcc_open + one per file (main.c:45, :112)
    rounds = 0
    for rel, x in clips.items():
        got = np.fromfile(str(tmp_path / rel)[:-3] + "mfcc", dtype=np.int16).reshape(-1, 32)
        want = H.wav2mfcc(x, lambda s, k: RC.run(cfg, s, n_frames=k))
        assert got.shape == (MFCCConfig().num_frames_host(len(x)), 32), rel
        assert np.array_equal(got, want), rel
        rounds += got.shape[0]
    assert model.words_in == model.resets + sum(512 + 170 * (MFCCConfig().num_frames_host(len(x)) - 1)
                                                for x in clips.values())


@needs_ref_tool
@pytest.mark.gpu
def test_gpu_converter_writes_the_reference_tools_files(tmp_path, wav):
    """Same directory through software/main.c (+ oracle-backed FPGA model) and through
    mfcc_b200.convert on the GPU: byte-identical .mfcc files."""
    import mfcc_b200
    ref_root, gpu_root = str(tmp_path / "ref"), str(tmp_path / "gpu")
    clips, _model, _cfg = _run_reference_tool(ref_root, wav)
    for rel, x in clips.items():
        os.makedirs(os.path.dirname(os.path.join(gpu_root, rel)), exist_ok=True)
        write_wav(os.path.join(gpu_root, rel), x)
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device="cuda:0")
    done, rejected = convert.convert_dir(gpu_root, op=op)
    assert len(done) == len(clips) and not rejected
    for rel in clips:
        a = open(os.path.join(ref_root, rel)[:-3] + "mfcc", "rb").read()
        b = open(os.path.join(gpu_root, rel)[:-3] + "mfcc", "rb").read()
        assert a == b and len(a) > 0, rel


# ---------------------------------------------------------------- the C host tool over the C ABI
def _build_c_tool():
    import subprocess
    host = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "host")
    subprocess.check_call(["make", "-C", host, "-s"])
    return os.path.join(host, "wav2mfcc")


def test_c_host_tool_builds_and_fails_loudly_without_a_gpu(tmp_path):
    """host/wav2mfcc.c: the reference's tool in its own language on top of include/mfcc_b200.h.
    It must link against the library with nothing but the header, print the reference's
    usage line, and report the CUDA error instead of producing files when there is no GPU."""
    import subprocess
    import torch
    exe = _build_c_tool()
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and r.stdout.startswith("Usage:") and "<wavdir>" in r.stdout   # main.c:250-253
    if not torch.cuda.is_available():
        write_wav(str(tmp_path / "a.wav"), np.zeros(1000, dtype=np.int16))
        r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
        assert r.returncode == 2 and "mfcc_b200_create" in r.stdout
        assert not os.path.exists(str(tmp_path / "a.mfcc"))


@needs_ref_tool
@pytest.mark.gpu
def test_c_host_tool_writes_the_reference_tools_files(tmp_path, wav):
    """Same directory through the reference's software/main.c (oracle-backed FPGA model) and
    through host/wav2mfcc (plain C over the C ABI, GPU): byte-identical .mfcc files; files
    the reference rejects are rejected with its messages."""
    import subprocess
    exe = _build_c_tool()
    ref_root, c_root = str(tmp_path / "ref"), str(tmp_path / "c")
    clips, _model, _cfg = _run_reference_tool(ref_root, wav)
    clips = dict(clips)
    for rel, x in clips.items():
        os.makedirs(os.path.dirname(os.path.join(c_root, rel)), exist_ok=True)
        write_wav(os.path.join(c_root, rel), x)
    write_wav(os.path.join(c_root, "t", "rate.wav"), wav[:1000], rate=8000)
    write_wav(os.path.join(c_root, "t", "stereo.wav"), wav[:1000], channels=2)
    write_wav(os.path.join(c_root, "empty.wav"), np.zeros(0, dtype=np.int16))
    r = subprocess.run([exe, c_root], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Unexpected samplerate: 8000" in r.stdout and "Unexpected channel count: 2" in r.stdout
    assert "%d files converted, 2 rejected" % (len(clips) + 1) in r.stdout
    for rel in clips:
        a = open(os.path.join(ref_root, rel)[:-3] + "mfcc", "rb").read()
        b = open(os.path.join(c_root, rel)[:-3] + "mfcc", "rb").read()
        assert a == b and len(a) > 0, rel
    assert os.path.getsize(os.path.join(c_root, "empty.mfcc")) == 64          # one zero-filled round
    assert not os.path.exists(os.path.join(c_root, "t", "rate.mfcc"))


@pytest.mark.gpu
def test_c_mic_tool_feeds_the_reference_receiver(tmp_path):
    """host/mic2mfcc.c (plain C over mfcc_b200_step_host): its byte stream is what the mic
    build sends (magic.py:27-39), column k = the batch path's frame k, and the reference's
    own receiver (software/cepstrum.c:15-68, restated in oracle/ref_recv.py) parses it."""
    import subprocess
    from oracle import ref_np, ref_recv as RR
    tool = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "host", "mic2mfcc")
    if not os.path.exists(tool):
        pytest.fail("host/mic2mfcc not built (python -c 'import __graft_entry__ as g; g.build()')")
    rng = np.random.default_rng(77)
    x = rng.integers(-20000, 20000, size=512 + 170 * 99 + 57, dtype=np.int16)
    r = subprocess.run([tool, "16"], input=x.astype("<i2").tobytes(), capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr.decode()
    wire = np.frombuffer(r.stdout, dtype=np.uint8)
    want = ref_np.mfcc(x, nfilters=32, nceptrums=16, n_frames=100)
    assert wire.size == 100 * 34
    words = wire.reshape(100, 17, 2)
    got = ((words[:, :, 0].astype(np.uint16) << 8) | words[:, :, 1]).astype(np.uint16)
    assert (got[:, 0] == 0xa55a).all()
    np.testing.assert_array_equal(got[:, 1:].view(np.int16), want)
    # through the receiver: it fills its 93-column window from the start of the stream
    if not RR.available():
        pytest.fail("oracle/_ref/libcepstrum_ref.so missing: build() compiles it where /root/reference exists")
    win = RR.get_window(wire.tobytes(), 16, 93)
    np.testing.assert_array_equal(win, want[:93])
