"""TEST INFRASTRUCTURE -- cycle simulator of the small nMigen interpreter (see _ast.py).

The design is flattened, every driven signal gets one piece of straight-line Python
(its statements in program order, so later assignments win as in the HDL), comb signals
are evaluated once per settle in dependency order, sync signals are computed from the
settled values and committed together at the clock edge.  One clock domain ("sync").
"""
from __future__ import annotations

from ._ast import Assign, Const, Signal, Value, _wrap_src
from ._dsl import (Module, ReadPort, WritePort, _If, _Switch, _Transformed, _StateName,
                   _NextAssign, bits_for)
from ._ast import Shape

__all__ = ["Simulator", "Delay", "Settle", "Tick", "Passive", "Active"]


class Delay:
    def __init__(self, interval=None):
        self.interval = interval


class Settle:
    pass


class Tick:
    def __init__(self, domain="sync"):
        self.domain = domain


class Passive:
    pass


class Active:
    pass


class _Emit:
    """Collects the code of assignments to ONE target signal."""

    def __init__(self, target, rd0, lines, indent):
        self.target, self.rd0, self.lines, self.indent = target, rd0, lines, indent

    def rd(self, sig):
        return "t" if sig is self.target else self.rd0(sig)

    def __call__(self, sig, src):
        if sig is self.target:
            self.lines.append("%st = %s" % (self.indent, src))


class _Design:
    def __init__(self, top):
        self.frags = []            # (statement list, [reset signals])
        self.rports, self.wports = [], []
        self.mems = {}
        self.sig_index = {}
        self.sigs = []
        self.fsms = []
        self._collect(top, [])
        self._drivers()

    # -- flatten
    def _collect(self, e, resets):
        if isinstance(e, _Transformed):
            if e._kind != "reset":
                raise NotImplementedError("EnableInserter is outside the simulated subset")
            self._collect(e._inner, resets + [e._controls["sync"]])
        elif isinstance(e, ReadPort):
            self.rports.append(e)
            self.mems.setdefault(id(e.memory), e.memory)
        elif isinstance(e, WritePort):
            self.wports.append(e)
            self.mems.setdefault(id(e.memory), e.memory)
        elif isinstance(e, Module):
            self.frags.append((e._top, list(resets)))
            for _name, sub in e._subs:
                self._collect(sub, resets)
        else:
            m = e.elaborate(None)
            if m is e:
                raise TypeError("elaborate() returned the object itself")
            self._collect(m, resets)

    def idx(self, sig):
        i = self.sig_index.get(sig)
        if i is None:
            i = self.sig_index[sig] = len(self.sigs)
            self.sigs.append(sig)
        return i

    # -- who drives what
    def _walk(self, stmts, fn):
        for st in stmts:
            if isinstance(st, Assign):
                fn(st)
            elif isinstance(st, _If):
                for _c, body in st.branches:
                    self._walk(body, fn)
            elif isinstance(st, _Switch):
                fsm = getattr(st, "fsm", None)
                if fsm is not None and fsm not in self.fsms:
                    self.fsms.append(fsm)
                for _p, body in st.cases:
                    self._walk(body, fn)

    def _drivers(self):
        self.domain = {}           # signal -> None (comb) | "sync"
        self.resets = {}           # signal -> set of reset control signals
        self.users = {}            # signal -> list of fragment indices that drive it
        for fi, (stmts, resets) in enumerate(self.frags):
            def fn(st, fi=fi, resets=resets):
                tg = set()
                st.lhs._lhs_signals(tg)
                for s in tg:
                    d = st.domain
                    if s in self.domain and self.domain[s] != d:
                        raise RuntimeError("signal %r driven from two domains" % s)
                    self.domain[s] = d
                    self.resets.setdefault(s, [])
                    for r in resets:
                        if not any(r is q for q in self.resets[s]):
                            self.resets[s].append(r)
                    u = self.users.setdefault(s, [])
                    if fi not in u:
                        u.append(fi)
            self._walk(stmts, fn)
        for fsm in self.fsms:
            w = max(1, bits_for(len(fsm.encoding) - 1))
            fsm.state._shape = Shape(w, False)
            fsm.state.reset = fsm.encoding[fsm.reset_name]

    # -- code for one target signal
    def _gen(self, stmts, target, rd0, lines, indent, deps):
        """Append the code of `stmts` restricted to `target`; return True if anything was emitted."""
        any_ = False
        for st in stmts:
            if isinstance(st, Assign):
                tg = set()
                st.lhs._lhs_signals(tg)
                if target in tg:
                    st.rhs._deps(deps)
                    st.lhs._lhs_deps(deps)        # offsets of dynamic selects
                    st.lhs._lhs(st.rhs._src(rd0), _Emit(target, rd0, lines, indent))
                    any_ = True
            elif isinstance(st, _If):
                sub, hit = [], False
                for bi, (cond, body) in enumerate(st.branches):
                    kw = "if" if bi == 0 else "elif"
                    if cond is None:
                        sub.append("%selse:" % indent)
                    else:
                        sub.append("%s%s %s:" % (indent, kw, cond._src(rd0)))
                    n = len(sub)
                    if self._gen(body, target, rd0, sub, indent + " ", deps):
                        hit = True
                    if len(sub) == n:
                        sub.append("%s pass" % indent)
                if hit:
                    for cond, _b in st.branches:
                        if cond is not None:
                            cond._deps(deps)
                    lines.extend(sub)
                    any_ = True
            elif isinstance(st, _Switch):
                sub, hit = ["%s_v = %s" % (indent, st.test._src(rd0))], False
                first = True
                for pats, body in st.cases:
                    if pats is None:
                        cond = "True"
                    elif isinstance(pats, _StateName):
                        cond = "_v == %d" % pats.fsm.encoding[pats.name]
                    else:
                        cs = []
                        for p in pats:
                            if isinstance(p, str):
                                p = "".join(p.split())
                                msk = int(p.replace("0", "1").replace("-", "0"), 2)
                                val = int(p.replace("-", "0"), 2)
                                cs.append("(_v & %d) == %d" % (msk, val))
                            else:
                                cs.append("_v == %d" % int(p))
                        cond = " or ".join(cs)
                    sub.append("%s%s %s:" % (indent, "if" if first else "elif", cond))
                    first = False
                    n = len(sub)
                    if self._gen(body, target, rd0, sub, indent + " ", deps):
                        hit = True
                    if len(sub) == n:
                        sub.append("%s pass" % indent)
                if hit:
                    st.test._deps(deps)
                    lines.extend(sub)
                    any_ = True
        return any_

    def compile(self):
        rd0 = lambda sig: "s[%d]" % self.idx(sig)
        comb_code, comb_deps, sync_lines, commit = {}, {}, [], []
        for sig, dom in list(self.domain.items()):
            lines, deps = [], set()
            i = self.idx(sig)
            if dom is None:
                lines.append(" t = %d" % sig.reset_value())
            else:
                lines.append(" t = s[%d]" % i)
            for fi in self.users[sig]:
                self._gen(self.frags[fi][0], sig, rd0, lines, " ", deps)
            if dom is None:
                lines.append(" s[%d] = t" % i)
                deps.discard(sig)
                comb_code[sig], comb_deps[sig] = lines, deps
            else:
                if not sig.reset_less:
                    for r in self.resets[sig]:
                        lines.append(" if s[%d]: t = %d" % (self.idx(r), sig.reset_value()))
                lines.append(" n%d = t" % i)
                sync_lines += lines
                commit.append(" s[%d] = n%d" % (i, i))
        # memory ports
        mem_ids = {mid: k for k, mid in enumerate(self.mems)}
        self.mem_data = [list(m.init) + [0] * (m.depth - len(m.init)) for m in self.mems.values()]
        for rp in self.rports:
            k, d = mem_ids[id(rp.memory)], rp.memory.depth
            di = self.idx(rp.data)
            if rp.domain == "comb":
                comb_code[rp.data] = [" s[%d] = M[%d][min(s[%d], %d)]" % (di, k, self.idx(rp.addr), d - 1)]
                comb_deps[rp.data] = {rp.addr}
            elif rp.transparent:
                li = self.idx(rp._latch)
                sync_lines.append(" n%d = s[%d]" % (li, self.idx(rp.addr)))
                commit.append(" s[%d] = n%d" % (li, li))
                comb_code[rp.data] = [" s[%d] = M[%d][min(s[%d], %d)]" % (di, k, li, d - 1)]
                comb_deps[rp.data] = {rp._latch}
            else:
                en = rp.en._src(rd0)
                sync_lines.append(" n%d = M[%d][min(s[%d], %d)] if %s else s[%d]"
                                  % (di, k, self.idx(rp.addr), d - 1, en, di))
                commit.append(" s[%d] = n%d" % (di, di))
        for wp in self.wports:
            k, d = mem_ids[id(wp.memory)], wp.memory.depth
            commit.append(" if s[%d] and s[%d] < %d: W.append((%d, s[%d], s[%d]))"
                          % (self.idx(wp.en), self.idx(wp.addr), d, k, self.idx(wp.addr), self.idx(wp.data)))
        # the write list must be built from PRE-edge values: emit it before the commits
        wlines = [c for c in commit if "W.append" in c]
        commit = [c for c in commit if "W.append" not in c]
        # comb: dependency order
        order, state = [], {}

        def visit(sig, chain):
            st = state.get(sig)
            if st == 2 or sig not in comb_code:
                return
            if st == 1:
                raise RuntimeError("combinational loop through %r" % ([c.name for c in chain],))
            state[sig] = 1
            for dsig in sorted(comb_deps[sig], key=lambda x: self.idx(x)):
                visit(dsig, chain + [dsig])
            state[sig] = 2
            order.append(sig)

        for sig in list(comb_code):
            visit(sig, [sig])
        src = ["def comb(s, M):"]
        for sig in order:
            src += comb_code[sig]
        src.append(" return None")
        src.append("def sync(s, M, W):")
        src += sync_lines + wlines + commit
        src.append(" for k, a, d in W: M[k][a] = d")
        src.append(" del W[:]")
        self.usub_wraps = [0]
        uw = self.usub_wraps

        def _usub(r, mask):
            if r < 0:
                uw[0] += 1
            return r & mask

        def _setpart(cur, off, wmask, rhs, nmask):
            return (cur & ~(wmask << off)) | ((rhs & wmask) << off)

        env = {"_usub": _usub, "_setpart": _setpart, "min": min, "bin": bin}
        self.source = "\n".join(src)
        exec(compile(self.source, "<hdl_shim>", "exec"), env)
        self.comb, self.sync = env["comb"], env["sync"]
        self.env = env
        # widths may have changed for FSM states: (re)build the initial values
        self.values = [sg.reset_value() for sg in self.sigs]


class _Proc:
    def __init__(self, gen):
        self.gen, self.passive, self.done, self.started = gen, False, False, False


class Simulator:
    def __init__(self, fragment, **kwargs):
        self._d = _Design(fragment)
        self._d.compile()
        self._s = self._d.values
        self._M = self._d.mem_data
        self._W = []
        self._procs = []
        self._dirty = True
        self.cycles = 0

    # -- API of nmigen.sim.Simulator
    def add_clock(self, period, *, phase=None, domain="sync", if_exists=False):
        pass

    def add_sync_process(self, process, *, domain="sync"):
        self._procs.append(_Proc(process()))

    def add_process(self, process):
        self._procs.append(_Proc(process()))

    class _Vcd:
        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    def write_vcd(self, *a, **k):
        return Simulator._Vcd()

    # -- internals
    def _settle(self):
        if self._dirty:
            self._d.comb(self._s, self._M)
            self._dirty = False

    def _sig_index(self, sig):
        n = len(self._d.sigs)
        i = self._d.idx(sig)
        if i >= n:
            self._s.append(sig.reset_value())
        return i

    def _eval(self, value):
        self._settle()
        if isinstance(value, Signal):
            return self._s[self._sig_index(value)]
        src = value._src(lambda sg: "s[%d]" % self._sig_index(sg))
        return eval(src, self._d.env, {"s": self._s})

    def _exec(self, stmt):
        lhs, rhs = stmt.lhs, stmt.rhs
        v = rhs.value if isinstance(rhs, Const) else self._eval(rhs)
        tg = set()
        lhs._lhs_signals(tg)
        for sig in tg:
            if sig in self._d.domain:
                raise RuntimeError("process drives %r, which the design drives too" % sig)
            i = self._sig_index(sig)
            lines = []
            em = _Emit(sig, lambda sg: "s[%d]" % self._sig_index(sg), lines, "")
            lhs._lhs("(%d)" % v, em)
            loc = {"s": self._s, "t": self._s[i]}
            for ln in lines:
                exec(ln, self._d.env, loc)
            self._s[i] = loc["t"]
        self._dirty = True

    def _resume(self, p):
        """Run a process up to its next clock tick."""
        resp = None
        try:
            while True:
                cmd = p.gen.send(resp) if p.started else next(p.gen)
                p.started = True
                resp = None
                if cmd is None or isinstance(cmd, Tick):
                    return
                if isinstance(cmd, (Delay, Settle)):
                    self._settle()
                elif isinstance(cmd, Passive):
                    p.passive = True
                elif isinstance(cmd, Active):
                    p.passive = False
                elif isinstance(cmd, Assign):
                    self._exec(cmd)
                elif isinstance(cmd, (list, tuple)):
                    for c in cmd:
                        self._exec(c)
                elif isinstance(cmd, Value):
                    resp = self._eval(cmd)
                else:
                    raise TypeError("process yielded %r" % (cmd,))
        except StopIteration:
            p.done = True

    def _tick(self):
        self._settle()
        self._d.sync(self._s, self._M, self._W)
        self._dirty = True
        self._settle()
        self.cycles += 1

    def run(self):
        for p in self._procs:
            self._resume(p)
        while any(not p.done and not p.passive for p in self._procs):
            self._tick()
            for p in self._procs:
                if not p.done:
                    self._resume(p)

    def run_until(self, deadline, *, run_passive=False):
        raise NotImplementedError

    def reset(self):
        raise NotImplementedError
