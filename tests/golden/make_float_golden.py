#!/usr/bin/env python
"""Golden vectors for the fp32 mode, made by EXECUTING the reference's float prototype.

    python tests/golden/make_float_golden.py        (build container only)

notebook/MFCC-INT.ipynb is the reference's float model of the path.  Its
code cells 2-10 are executed here verbatim (cell 0's imports are provided by hand,
plotting is stubbed out: matplotlib is not installed) on

  * the bundled clip ../f2bjrop1.0.wav (the notebook's commented-out default input), and
  * a seeded synthetic clip (speech-level Gaussian noise plus three tones),

and the per-stage arrays they leave in the namespace are stored in
tests/golden/notebook_float.npz (float64, a 16-frame slice of every stage plus the full
cepstrum matrix of each clip).
"""
import json
import os

import numpy as np
import scipy
import scipy.fft
import scipy.fftpack
from scipy.signal import get_window

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


class _Plot:
    """Stands in for matplotlib.pyplot: every call is accepted and ignored."""

    def __getattr__(self, name):
        return lambda *a, **k: self


def run_notebook(audio, sample_rate=16000):
    nb = json.load(open(os.path.join(REF, "notebook", "MFCC-INT.ipynb")))
    ns = dict(np=np, scipy=scipy, fft=scipy.fftpack, get_window=get_window, plt=_Plot(),
              audio=audio, sample_rate=sample_rate, print=lambda *a, **k: None)
    for idx in range(2, 11):
        cell = nb["cells"][idx]
        assert cell["cell_type"] == "code"
        src = "".join(cell["source"])
        with np.errstate(divide="ignore", invalid="ignore"):
            exec(compile(src, "MFCC-INT.ipynb[cell %d]" % idx, "exec"), ns)
    return ns


def main():
    wav = np.load(os.path.join(HERE, "f2bjrop1.npz"))["samples"]
    rng = np.random.default_rng(20260002)
    n = 512 + 170 * 63
    t = np.arange(n)
    synth = (rng.normal(0, 2000, n) + 6000 * np.sin(2 * np.pi * 440 / 16000 * t)
             + 3000 * np.sin(2 * np.pi * 2500 / 16000 * t) + 1500 * np.sin(2 * np.pi * 6100 / 16000 * t))
    synth = np.clip(synth, -32768, 32767).astype(np.int16)
    out = {}
    for name, audio in (("wav", wav), ("synth", synth)):
        ns = run_notebook(audio)
        sl = slice(100, 116) if name == "wav" else slice(0, 16)
        if name != "wav":                       # the clip itself is tests/golden/f2bjrop1.npz
            out[name + "_samples"] = audio
        out[name + "_slice"] = np.array([sl.start, sl.stop])
        out[name + "_emphasis"] = np.asarray(ns["audio_emphasis"], dtype=np.float64)[:4096]
        out[name + "_window"] = np.asarray(ns["audio_wins"])[sl]
        out[name + "_fft"] = np.asarray(ns["audio_fft"])[sl]
        out[name + "_power"] = np.asarray(ns["audio_power"])[sl]
        out[name + "_mel"] = np.transpose(ns["audio_filtered"])[sl]
        out[name + "_log"] = np.transpose(ns["audio_log"])[sl]
        out[name + "_dct"] = np.asarray(ns["audio_dct"])              # [K][32], every frame
        out[name + "_filters"] = np.asarray(ns["filters"])
        out[name + "_filter_points"] = np.asarray(ns["filter_points"])
        print(name, "frames", ns["audio_dct"].shape, "finite", np.isfinite(ns["audio_dct"]).all(),
              "first", np.round(ns["audio_dct"][0][:4], 3))
    np.savez_compressed(os.path.join(HERE, "notebook_float.npz"), **out)


if __name__ == "__main__":
    main()
