"""``LiveSession``: the microphone build end to end, for many streams at once.

The FPGA top of the mic build (/root/reference/mfcc/targets/mic2mfcc.py:19-30, :53-74)
is ``MFCC(nfft=512, nfilters=32, nceptrums=16)`` fed by an I2S microphone, followed by
``MagicInserter`` and a UART sender; its host (software/cepstrum.c:92-183) keeps a
93-column window per stream and gates on ``cepstrum_eval_power``.  One ``push(hop)`` here
does what one 10.6 ms hop does there, for every stream: the streaming MFCC step, the
serial framing of the new columns, and the voice-activity gate.  The three launches of a
push can be captured once in a CUDA graph (``capture=True``).
"""
from __future__ import annotations

import torch

from . import post
from .mfcc import MFCC


class LiveSession:
    def __init__(self, n_streams, nfilters=32, nceptrums=16, nframes=post.NFRAMES_WINDOW,
                 threshold=post.POWER_THRESHOLD, device=None, capture=False):
        self.n_streams = n_streams
        self.mfcc = MFCC(nfft=512, nfilters=nfilters, nceptrums=nceptrums, device=device)   # mic2mfcc.py:19
        self.device = self.mfcc.device
        self.gate = post.VoiceGate(n_streams, nceptrums, nframes, threshold, device=self.device)
        self.mfcc.reset(n_streams)
        self.hop = torch.zeros((n_streams, self.mfcc.cfg.hop), dtype=torch.int16, device=self.device)
        self.cols = torch.zeros((n_streams, nceptrums), dtype=torch.int16, device=self.device)
        self.wire = None
        self._graph = None
        if capture:
            st = torch.cuda.Stream(device=self.device)
            with torch.cuda.stream(st):
                for _ in range(2):                       # warm up outside the capture
                    self._launch()
                st.synchronize()
                self.reset()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    self._launch()
            self._graph, self._stream = g, st
            self.reset()

    def _launch(self):
        self.mfcc.step(self.hop, out=self.cols)
        self.wire = post.uart_frame(self.cols)            # magic + big-endian words, magic.py:27-39
        self.power, self.active = self.gate.step(self.cols)

    def reset(self):
        """The soft reset of every stream: MFCC state, window and gate back to power-on."""
        self.mfcc.reset()
        self.gate.reset()
        return self

    def push(self, hop):
        """hop: int16 ``[n_streams, 170]`` on the device (or pinned host memory).  Returns
        (cols int16 ``[S, nceptrums]``, wire uint8 ``[S, 2*(nceptrums+1)]``, power int32
        ``[S]``, active bool ``[S]``); the tensors are reused by the next push."""
        if hop.shape != self.hop.shape or hop.dtype != torch.int16:
            raise ValueError("hop must be int16 [%d, %d]" % tuple(self.hop.shape))
        if self._graph is not None:
            with torch.cuda.stream(self._stream):
                self.hop.copy_(hop, non_blocking=True)
                self._graph.replay()
            self._stream.synchronize()
        else:
            self.hop.copy_(hop, non_blocking=True)
            self._launch()
        return self.cols, self.wire, self.power, self.active
