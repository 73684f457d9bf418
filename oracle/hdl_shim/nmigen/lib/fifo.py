"""TEST INFRASTRUCTURE -- synchronous FIFOs with the port names of nmigen.lib.fifo.

Behavioural stand-ins: first-word-fall-through FIFOs that keep element order.  The
buffered variant of nMigen has one more cycle of latency; every consumer in the
reference waits for `valid`, so latency does not reach the data.
"""
from .. import Elaboratable, Memory, Module, Signal

__all__ = ["FIFOInterface", "SyncFIFO", "SyncFIFOBuffered", "AsyncFIFO", "AsyncFIFOBuffered"]


class FIFOInterface:
    def __init__(self, *, width, depth, fwft=True):
        self.width, self.depth, self.fwft = width, depth, fwft
        self.w_data = Signal(width, reset_less=True)
        self.w_rdy = Signal()
        self.w_en = Signal()
        self.w_level = Signal(range(depth + 1))
        self.r_data = Signal(width, reset_less=True)
        self.r_rdy = Signal()
        self.r_en = Signal()
        self.r_level = Signal(range(depth + 1))


class SyncFIFO(Elaboratable, FIFOInterface):
    def __init__(self, *, width, depth, fwft=True):
        if not fwft:
            raise NotImplementedError("only first-word-fall-through FIFOs are modelled")
        super().__init__(width=width, depth=depth, fwft=fwft)
        self.level = Signal(range(depth + 1))

    def elaborate(self, platform):
        m = Module()
        if self.depth == 0:
            return m
        storage = Memory(width=self.width, depth=self.depth)
        m.submodules.wp = wp = storage.write_port()
        m.submodules.rp = rp = storage.read_port(domain="comb")
        produce = Signal(range(self.depth))
        consume = Signal(range(self.depth))
        do_write = Signal()
        do_read = Signal()
        m.d.comb += [
            self.w_rdy.eq(self.level != self.depth),
            self.r_rdy.eq(self.level != 0),
            self.w_level.eq(self.level),
            self.r_level.eq(self.level),
            do_write.eq(self.w_rdy & self.w_en),
            do_read.eq(self.r_rdy & self.r_en),
            wp.addr.eq(produce), wp.data.eq(self.w_data), wp.en.eq(do_write),
            rp.addr.eq(consume), self.r_data.eq(rp.data),
        ]
        with m.If(do_write):
            m.d.sync += produce.eq(produce + 1)
            with m.If(produce == self.depth - 1):
                m.d.sync += produce.eq(0)
        with m.If(do_read):
            m.d.sync += consume.eq(consume + 1)
            with m.If(consume == self.depth - 1):
                m.d.sync += consume.eq(0)
        with m.If(do_write & ~do_read):
            m.d.sync += self.level.eq(self.level + 1)
        with m.If(do_read & ~do_write):
            m.d.sync += self.level.eq(self.level - 1)
        return m


class SyncFIFOBuffered(SyncFIFO):
    pass


class AsyncFIFO:
    def __init__(self, *a, **k):
        raise NotImplementedError("AsyncFIFO is outside the simulated subset (one clock domain)")


AsyncFIFOBuffered = AsyncFIFO
