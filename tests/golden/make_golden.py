#!/usr/bin/env python
"""Generate the committed fixtures under tests/golden/ from /root/reference.

Run in the build container (the GPU box has no /root/reference):
    python tests/golden/make_golden.py

Writes
  notebook_pins.json   values PRINTED by the reference's notebooks (exact pins):
                       the rebuilt 512-entry Hamming curve, its 64-entry quarter table
                       and offsets (notebook/MFCC.ipynb cell 17), the 34 mel filter
                       points (notebook/MFCC.ipynb cell 28, MFCC-INT.ipynb cell 8), the
                       frame count for 16000 samples (MFCC.ipynb cell 9 output)
  f2bjrop1.npz         the bundled clip f2bjrop1.0.wav as int16 samples (config C1)
  stimuli.json         the fixed stimulus vectors of the reference's own benches
                       (fft_stream.py:82 / fft.py:490, dct_stream.py:81, log.py:149,
                       filterbank.py:151); they carry no expected outputs upstream
"""
import json
import os
import re

import numpy as np
from scipy.io import wavfile

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def cell_text(nb, idx):
    d = json.load(open(os.path.join(REF, "notebook", nb)))
    c = d["cells"][idx]
    return "".join("".join(o["text"]) for o in c.get("outputs", []) if "text" in o)


def ints(s):
    return [int(v) for v in re.findall(r"-?\d+", s)]


def main():
    pins = {}
    t = cell_text("MFCC.ipynb", 17)
    pins["window_after"] = ints(re.search(r"after:\s*(\d+)", t).group(0))[0]
    small = re.search(r"small \[([^\]]*)\]", t).group(1)
    pins["window_small"] = ints(small)
    pins["window_offsetlast"] = ints(re.search(r"offsetlast\s*(\d+)", t).group(0))[0]
    tail = t[t.index("offsetlast"):]
    curve = ints(tail[tail.index("["):])
    assert len(curve) == 512, len(curve)
    pins["window_curve"] = curve
    pins["source_window"] = "notebook/MFCC.ipynb cell 17 stdout"

    t = cell_text("MFCC-INT.ipynb", 8)
    pts = ints(t[t.index("["): t.index("]") + 1])
    assert len(pts) == 34
    pins["mel_points_32"] = pts
    pins["source_mel"] = "notebook/MFCC-INT.ipynb cell 8 stdout (first printed array)"

    d = json.load(open(os.path.join(REF, "notebook", "MFCC.ipynb")))
    for i, c in enumerate(d["cells"]):
        txt = "".join("".join(o.get("text", "")) for o in c.get("outputs", []) if "text" in o)
        m = re.search(r"before 16000\s+number of frames (\d+)", txt)
        if m:
            pins["frames_in_16000"] = int(m.group(1))
            pins["source_frames"] = "notebook/MFCC.ipynb cell %d stdout" % i
            break
    json.dump(pins, open(os.path.join(HERE, "notebook_pins.json"), "w"), indent=0)

    sr, audio = wavfile.read(os.path.join(REF, "f2bjrop1.0.wav"))
    assert sr == 16000 and audio.dtype == np.int16 and audio.ndim == 1
    np.savez_compressed(os.path.join(HERE, "f2bjrop1.npz"), samples=audio, samplerate=sr)

    def hexlist(path, lineno):
        line = open(os.path.join(REF, path)).read().splitlines()[lineno - 1]
        return [int(v, 16) for v in re.findall(r"0x[0-9a-fA-F]+", line)]

    stim = {
        "fft512_hex": hexlist("mfcc/core/fft_stream.py", 82),
        "dct16_hex": hexlist("mfcc/core/dct_stream.py", 81),
        "log2_value": 2207315,
        "filterbank_const": 1234,
    }
    assert len(stim["fft512_hex"]) == 512 and len(stim["dct16_hex"]) == 16
    assert stim["fft512_hex"] == hexlist("mfcc/misc/fft.py", 490)
    json.dump(stim, open(os.path.join(HERE, "stimuli.json"), "w"))
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
