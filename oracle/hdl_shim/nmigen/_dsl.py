"""TEST INFRASTRUCTURE -- Module / control flow / Memory of the small nMigen interpreter
(see _ast.py).  Follows the documented behaviour of nMigen 0.3's hdl.dsl, hdl.mem and
hdl.xfrm for the constructs the reference design uses."""
from __future__ import annotations

from collections import OrderedDict
from contextlib import contextmanager

from ._ast import Assign, Const, Shape, Signal, Statement, Value, bits_for

__all__ = ["Elaboratable", "Module", "Memory", "ResetInserter", "EnableInserter", "DomainRenamer",
           "ClockDomain", "Instance", "Fragment"]


class Elaboratable:
    pass


class ClockDomain:
    def __init__(self, name=None, *, clk_edge="pos", reset_less=False, async_reset=False, local=False):
        self.name = name or "sync"
        self.clk = Signal(name=self.name + "_clk")
        self.rst = None if reset_less else Signal(name=self.name + "_rst")


class Instance:
    def __init__(self, *a, **k):
        raise NotImplementedError("Instance is outside the simulated subset")


class Fragment:
    pass


# ----------------------------------------------------------------------------- statements
class _If(Statement):
    """An If/Elif/Else chain: branches = [(cond or None, [stmts])]."""

    def __init__(self):
        self.branches = []


class _Switch(Statement):
    def __init__(self, test):
        self.test = test
        self.cases = []          # [(patterns tuple or None for default, [stmts])]


def _flat_stmts(x):
    if isinstance(x, Statement):
        yield x
    elif isinstance(x, (list, tuple)) or hasattr(x, "__iter__"):
        for e in x:
            yield from _flat_stmts(e)
    else:
        raise TypeError("Only assignments may be appended to d.<domain>, not {!r}".format(x))


class _DomainProxy:
    def __init__(self, module, domain):
        self._m, self._d = module, domain

    def __iadd__(self, stmts):
        for s in _flat_stmts(stmts):
            if not isinstance(s, Assign):
                raise TypeError("Only assignments may be appended to d.<domain>")
            s.domain = self._d
            self._m._stack[-1].append(s)
        return self


class _D:
    def __init__(self, module):
        object.__setattr__(self, "_m", module)

    def __getattr__(self, name):
        return _DomainProxy(self._m, None if name == "comb" else name)

    def __getitem__(self, name):
        return self.__getattr__(name)

    def __setattr__(self, name, value):
        if not isinstance(value, _DomainProxy):
            raise AttributeError("Cannot assign 'd.{}' attribute; did you mean 'd.{} +='?".format(name, name))

    def __setitem__(self, name, value):
        self.__setattr__(name, value)


class _Submodules:
    def __init__(self, module):
        object.__setattr__(self, "_m", module)

    def __iadd__(self, mods):
        for m in (mods if isinstance(mods, (list, tuple)) else [mods]):
            self._m._subs.append((None, m))
        return self

    def __setattr__(self, name, value):
        self._m._subs.append((name, value))

    def __setitem__(self, name, value):
        self._m._subs.append((name, value))

    def __getattr__(self, name):
        for n, v in self._m._subs:
            if n == name:
                return v
        raise AttributeError(name)


class FSM:
    def __init__(self, state, encoding, decoding):
        self.state, self.encoding, self.decoding = state, encoding, decoding

    def ongoing(self, name):
        if name not in self.encoding:
            self.encoding[name] = len(self.encoding)
        return _Ongoing(self, name)


class _Ongoing(Value):
    """state == encoding[name], resolved when compiled (the encoding may still grow)."""

    def __init__(self, fsm, name):
        self.fsm, self.name = fsm, name

    def shape(self):
        return Shape(1, False)

    def _deps(self, acc):
        acc.add(self.fsm.state)

    def _src(self, rd):
        return "(1 if %s == %d else 0)" % (rd(self.fsm.state), self.fsm.encoding[self.name])


class Module(Elaboratable):
    def __init__(self):
        self._top = []
        self._stack = [self._top]
        self._subs = []
        self._ctx = [None]
        self._last_if = [None]       # per nesting depth: the _If a following Elif/Else extends
        self.d = _D(self)
        self.submodules = _Submodules(self)
        self.domains = _Submodules(self)
        self._fsm = None
        self._fsm_state_stack = []

    def __setattr__(self, name, value):
        if name == "next":
            if self._fsm is None or not self._fsm_state_stack:
                raise SyntaxError("Only assignment to `m.next` inside an FSM state is permitted")
            fsm = self._fsm
            if value not in fsm.encoding:
                fsm.encoding[value] = len(fsm.encoding)
            a = _NextAssign(fsm, value)
            self._stack[-1].append(a)
            return
        object.__setattr__(self, name, value)

    # -- If / Elif / Else
    @contextmanager
    def _branch(self, node, cond):
        body = []
        node.branches.append((cond, body))
        self._stack.append(body)
        self._last_if.append(None)
        try:
            yield
        finally:
            self._stack.pop()
            self._last_if.pop()

    def _cond(self, cond):
        cond = Value.cast(cond)
        return cond.bool() if len(cond) != 1 else cond

    def If(self, cond):
        node = _If()
        self._stack[-1].append(node)
        self._last_if[-1] = node
        return self._branch(node, self._cond(cond))

    def Elif(self, cond):
        node = self._last_if[-1]
        if node is None:
            raise SyntaxError("Elif without preceding If")
        return self._branch(node, self._cond(cond))

    def Else(self):
        node = self._last_if[-1]
        if node is None:
            raise SyntaxError("Else without preceding If/Elif")
        ctx = self._branch(node, None)
        self._last_if[-1] = None
        return ctx

    # -- Switch / Case
    @contextmanager
    def Switch(self, test):
        node = _Switch(Value.cast(test))
        self._stack[-1].append(node)
        self._last_if[-1] = None
        self._ctx.append(node)
        try:
            yield
        finally:
            self._ctx.pop()

    @contextmanager
    def Case(self, *patterns):
        node = self._ctx[-1]
        if not isinstance(node, _Switch):
            raise SyntaxError("Case is not permitted outside of Switch")
        body = []
        node.cases.append((patterns if patterns else None, body))
        self._stack.append(body)
        self._last_if.append(None)
        try:
            yield
        finally:
            self._stack.pop()
            self._last_if.pop()

    def Default(self):
        return self.Case()

    # -- FSM / State
    @contextmanager
    def FSM(self, reset=None, domain="sync", name="fsm"):
        state = Signal(name=name + "_state")
        fsm = FSM(state, OrderedDict(), OrderedDict())
        fsm.domain = domain
        fsm.reset_name = reset
        node = _Switch(state)
        node.fsm = fsm
        self._stack[-1].append(node)
        self._last_if[-1] = None
        prev = self._fsm
        self._fsm = fsm
        self._ctx.append(node)
        try:
            yield fsm
        finally:
            self._ctx.pop()
            self._fsm = prev
            fsm.finalize = True

    @contextmanager
    def State(self, name):
        node = self._ctx[-1]
        fsm = self._fsm
        if fsm is None or getattr(node, "fsm", None) is not fsm:
            raise SyntaxError("FSM State is not permitted outside of FSM")
        if name not in fsm.encoding:
            fsm.encoding[name] = len(fsm.encoding)
        if fsm.reset_name is None:
            fsm.reset_name = name          # the first state defined is the reset state
        body = []
        node.cases.append((_StateName(fsm, name), body))
        self._stack.append(body)
        self._last_if.append(None)
        self._fsm_state_stack.append(name)
        try:
            yield
        finally:
            self._fsm_state_stack.pop()
            self._stack.pop()
            self._last_if.pop()

    def elaborate(self, platform):
        return self


class _StateName:
    def __init__(self, fsm, name):
        self.fsm, self.name = fsm, name


class _NextAssign(Assign):
    """m.next = "STATE": a sync assignment to the FSM state register."""

    def __init__(self, fsm, name):
        self.fsm, self.name = fsm, name
        self.lhs = fsm.state
        self.domain = fsm.domain

    @property
    def rhs(self):
        return Const(self.fsm.encoding[self.name], max(1, bits_for(len(self.fsm.encoding) - 1)))


# ----------------------------------------------------------------------------- memory
class Memory:
    def __init__(self, *, width, depth, init=None, name=None, attrs=None, simulate=True):
        self.width, self.depth = int(width), int(depth)
        self.name = name or "$memory"
        self.attrs = OrderedDict(() if attrs is None else attrs)
        self._init = []
        self.init = init
        self.read_ports, self.write_ports = [], []

    @property
    def init(self):
        return self._init

    @init.setter
    def init(self, new_init):
        vals = [] if new_init is None else [int(v) for v in new_init]
        if len(vals) > self.depth:
            raise ValueError("Memory initialization value count exceed memory depth")
        m = (1 << self.width) - 1
        self._init = [v & m for v in vals]

    def read_port(self, *, src_loc_at=0, **kwargs):
        return ReadPort(self, **kwargs)

    def write_port(self, *, src_loc_at=0, **kwargs):
        return WritePort(self, **kwargs)


class ReadPort(Elaboratable):
    def __init__(self, memory, *, domain="sync", transparent=True):
        if domain == "comb" and not transparent:
            raise ValueError("Read port cannot be simultaneously asynchronous and non-transparent")
        self.memory, self.domain, self.transparent = memory, domain, transparent
        self.addr = Signal(range(memory.depth), name=memory.name + "_r_addr")
        self.data = Signal(memory.width, name=memory.name + "_r_data")
        if domain != "comb" and not transparent:
            self.en = Signal(name=memory.name + "_r_en", reset=1)
        else:
            self.en = Const(1)
        # sync transparent port: the address is registered, data = mem[latched addr]
        # after the write of the same edge (hdl/mem.py ReadPort.elaborate, 0.3)
        self._latch = Signal(range(memory.depth), name=memory.name + "_r_latch")
        if memory.init:
            self.data.reset = memory.init[0]
        memory.read_ports.append(self)

    def elaborate(self, platform):
        return self


class WritePort(Elaboratable):
    def __init__(self, memory, *, domain="sync", granularity=None):
        if granularity not in (None, memory.width):
            raise NotImplementedError("write granularity is outside the simulated subset")
        self.memory, self.domain = memory, domain
        self.addr = Signal(range(memory.depth), name=memory.name + "_w_addr")
        self.data = Signal(memory.width, name=memory.name + "_w_data")
        self.en = Signal(name=memory.name + "_w_en")
        memory.write_ports.append(self)

    def elaborate(self, platform):
        return self


# ----------------------------------------------------------------------------- transformers
class _Transformed(Elaboratable):
    def __init__(self, inner, kind, controls):
        object.__setattr__(self, "_inner", inner)
        object.__setattr__(self, "_kind", kind)
        object.__setattr__(self, "_controls", controls)

    def __getattr__(self, name):
        return getattr(self._inner, name)

    def __setattr__(self, name, value):
        setattr(self._inner, name, value)

    def elaborate(self, platform):
        return self


class ResetInserter:
    def __init__(self, controls):
        self.controls = controls if isinstance(controls, dict) else {"sync": controls}

    def __call__(self, elaboratable):
        return _Transformed(elaboratable, "reset", self.controls)


class EnableInserter:
    def __init__(self, controls):
        self.controls = controls if isinstance(controls, dict) else {"sync": controls}

    def __call__(self, elaboratable):
        return _Transformed(elaboratable, "enable", self.controls)


class DomainRenamer:
    def __init__(self, domains):
        self.domains = domains

    def __call__(self, elaboratable):
        raise NotImplementedError("DomainRenamer is outside the simulated subset")
