"""Batch directory converter: every ``*.wav`` under a directory -> a headerless ``.mfcc``
file next to it, the workflow of the reference host tool ``wav2mfcc``
(/root/reference/software/main.c:100-177 ``mfcc_convert``, :206-247 directory walk).

File format (software/main.c:162-165; read back by software/lift.py:35-36 and
software/view.py:24-25): native-endian int16 ``[rows][nceptrums]``, no header.  Row
count is the host protocol's: the tool sends 512 samples, then 170 per round, zero-fills
the round in which the file ends and stops after it (main.c:128-166), i.e.
``(N - 512) // 170 + 2`` rows for N >= 512 samples and one row otherwise
(``MFCCConfig.num_frames_host``).

Where the reference streams one file at a time through the FPGA, this converter packs
many files into one launch: rows are zero-extended to the longest clip of the batch
(zeros after the end of the audio are exactly what the host tool feeds), all rows get
the longest clip's frame count, and each file keeps its own first rows.

    python -m mfcc_b200.convert <wavdir> [--nceptrums 32] [--nfilters 32] [--batch-mb 256]
"""
from __future__ import annotations

import argparse
import os
import sys
import wave

import numpy as np

SAMPLERATE = 16000          # software/main.c:22 (SAMPLERATE), checked at :87-92


def read_wav(path: str) -> np.ndarray:
    """16 kHz mono 16-bit PCM only, the checks of ``mfcc_wav_open`` (software/main.c:66-92);
    anything else raises ValueError with the tool's wording."""
    try:
        w = wave.open(path, "rb")
    except (wave.Error, EOFError) as e:
        # the wave module only opens PCM: software/main.c:73-78 "Unexpected format"
        raise ValueError("Unexpected format: %s (%s)" % (e, path)) from None
    with w:
        if w.getsampwidth() != 2:
            raise ValueError("Unexpected samplesize: %d (%s)" % (w.getsampwidth(), path))
        if w.getframerate() != SAMPLERATE:
            raise ValueError("Unexpected samplerate: %d (%s)" % (w.getframerate(), path))
        if w.getnchannels() != 1:
            raise ValueError("Unexpected channel count: %d (%s)" % (w.getnchannels(), path))
        data = w.readframes(w.getnframes())
    return np.frombuffer(data, dtype="<i2").astype(np.int16, copy=False)


def find_wavs(root: str):
    """Files whose name ends in ``.wav``, directories walked recursively
    (software/main.c:217-241), in sorted order."""
    out = []
    for d, _dirs, files in os.walk(root):
        out += [os.path.join(d, f) for f in files if f.endswith(".wav")]
    return sorted(out)


def mfcc_path(wav_path: str) -> str:
    return wav_path[:-3] + "mfcc"                            # software/main.c:227-228


def plan_batches(lengths, batch_bytes: int):
    """Group clip indices (sorted by length, so little padding is computed) into batches
    whose padded sample matrix stays under batch_bytes."""
    order = sorted(range(len(lengths)), key=lambda i: lengths[i])
    batches, cur, cur_max = [], [], 0
    for i in order:
        n = max(lengths[i], 8)
        if cur and (len(cur) + 1) * max(cur_max, n) * 2 > batch_bytes:
            batches.append(cur)
            cur, cur_max = [], 0
        cur.append(i)
        cur_max = max(cur_max, n)
    if cur:
        batches.append(cur)
    return batches


def convert_clips(op, clips):
    """clips: list of int16 arrays.  Returns the list of ``[rows_i, nceptrums]`` arrays the
    host tool would have written, computed in one launch through ``op.run_host``."""
    import torch
    cfg = op.cfg
    n_max = max(max(len(c) for c in clips), 8)
    n_pad = (n_max + 7) & ~7
    rows = [cfg.num_frames_host(len(c)) for c in clips]
    x = torch.zeros((len(clips), n_pad), dtype=torch.int16).pin_memory()
    xn = x.numpy()
    for i, c in enumerate(clips):
        xn[i, :len(c)] = c
    out = torch.empty((len(clips), max(rows), op.nceptrums), dtype=op.out_dtype).pin_memory()
    op.run_host(x, n_frames=max(rows), out=out)
    on = out.numpy()
    return [on[i, :rows[i]].copy() for i in range(len(clips))]


def convert_dir(root: str, op=None, nfilters: int = 32, nceptrums: int = 32,
                batch_bytes: int = 256 << 20, device=None, verbose: bool = False):
    """Convert every .wav under `root`.  Returns [(wav, mfcc, rows)] for the converted
    files and [(wav, reason)] for the rejected ones (the tool prints and moves on)."""
    if op is None:
        from .mfcc import MFCC
        # wav2mfcc is built with 32 filters and 32 coefficients (mfcc/targets/wav2mfcc.py:19)
        op = MFCC(nfilters=nfilters, nceptrums=nceptrums, device=device)
    wavs = find_wavs(root)
    clips, names, rejected = [], [], []
    for p in wavs:
        try:
            clips.append(read_wav(p))
            names.append(p)
        except ValueError as e:
            rejected.append((p, str(e)))
            if verbose:
                print(e)
    done = []
    for batch in plan_batches([len(c) for c in clips], batch_bytes):
        res = convert_clips(op, [clips[i] for i in batch])
        for i, r in zip(batch, res):
            dst = mfcc_path(names[i])
            r.tofile(dst)                                    # native endian, no header
            done.append((names[i], dst, r.shape[0]))
            if verbose:
                print(names[i], dst)
    return sorted(done), rejected


def read_wav_any_rate(path: str):
    """(samples int16, rate): ``librosa.load(wav, sr=None)`` keeps the file's rate
    (software/genlibrosa.py:21); mono 16-bit PCM only, like the rest of the tool chain."""
    try:
        w = wave.open(path, "rb")
    except (wave.Error, EOFError) as e:
        raise ValueError("Unexpected format: %s (%s)" % (e, path)) from None
    with w:
        if w.getsampwidth() != 2 or w.getnchannels() != 1:
            raise ValueError("Unexpected format: %d bytes/sample, %d channels (%s)"
                             % (w.getsampwidth(), w.getnchannels(), path))
        rate = w.getframerate()
        data = w.readframes(w.getnframes())
    return np.frombuffer(data, dtype="<i2").astype(np.int16, copy=False), rate


def convert_dir_librosa(root: str, ncepstrums: int = 32, hop_length: int = 170, device=None,
                        batch_bytes: int = 256 << 20, verbose: bool = False):
    """The loop of software/genlibrosa.py:15-28 on the GPU: for every ``root/*/*.wav``
    write ``<name>.spec`` (``librosa.feature.mfcc(..., hop_length=170, n_mfcc=32)`` cast to
    int16) and ``<name>.sklearn`` (the same features standardised over time, cast to
    int16), both laid out ``[n_mfcc][frames]`` as the script's ``tofile`` leaves them.

    Clips of equal length and rate share a launch (the reflect padding at the end of a
    clip depends on its length, so rows of one launch must be equally long)."""
    import glob
    import torch
    from . import post
    from .librosa_mode import LibrosaMFCC
    names = sorted(glob.glob(os.path.join(root, "*", "*.wav")))          # genlibrosa.py:15
    groups, rejected = {}, []
    for p in names:
        try:
            x, rate = read_wav_any_rate(p)
            if len(x) < 1:
                raise ValueError("empty file (%s)" % p)
            groups.setdefault((rate, len(x)), []).append((p, x))
        except ValueError as e:
            rejected.append((p, str(e)))
            if verbose:
                print(e)
    done, ops = [], {}
    for (rate, n), items in sorted(groups.items()):
        op = ops.get(rate)
        if op is None:
            op = ops[rate] = LibrosaMFCC(sr=rate, n_mfcc=ncepstrums, hop_length=hop_length, device=device)
        per = max(1, batch_bytes // max(2 * n, 1))
        for i in range(0, len(items), per):
            part = items[i:i + per]
            xh = torch.from_numpy(np.stack([x for _p, x in part])).pin_memory()
            feat = op(xh.to(op.device, non_blocking=True))               # [B, K, n_mfcc] float32
            spec = post.to_int16(feat).cpu().numpy()
            skl = post.scale(feat).cpu().numpy()
            for j, (p, _x) in enumerate(part):
                base = os.path.splitext(p)[0]
                np.ascontiguousarray(spec[j].T).tofile(base + ".spec")       # genlibrosa.py:27
                np.ascontiguousarray(skl[j].T).tofile(base + ".sklearn")     # genlibrosa.py:28
                done.append((p, base + ".spec", base + ".sklearn", spec.shape[1]))
                if verbose:
                    print(p, base + ".spec")
    return sorted(done), rejected


def main(argv=None):
    ap = argparse.ArgumentParser(description="wav2mfcc on the GPU: *.wav -> *.mfcc")
    ap.add_argument("wavdir")
    ap.add_argument("--nfilters", type=int, default=32)
    ap.add_argument("--nceptrums", type=int, default=32)
    ap.add_argument("--batch-mb", type=int, default=256)
    ap.add_argument("--librosa", action="store_true",
                    help="write .spec/.sklearn like software/genlibrosa.py instead of .mfcc")
    a = ap.parse_args(argv)
    if not os.path.isdir(a.wavdir):
        print("Usage: %s <wavdir>" % sys.argv[0])            # software/main.c:250-253
        return 1
    if a.librosa:
        done, rejected = convert_dir_librosa(a.wavdir, ncepstrums=a.nceptrums, batch_bytes=a.batch_mb << 20,
                                             verbose=True)
        print("%d files converted, %d rejected" % (len(done), len(rejected)))
        return 0
    done, rejected = convert_dir(a.wavdir, nfilters=a.nfilters, nceptrums=a.nceptrums,
                                 batch_bytes=a.batch_mb << 20, verbose=True)
    print("%d files converted, %d rejected" % (len(done), len(rejected)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
