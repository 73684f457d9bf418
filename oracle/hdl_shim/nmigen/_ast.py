"""TEST INFRASTRUCTURE -- value/statement model of the small nMigen interpreter.

Written from the documented behaviour of nMigen 0.3's ``hdl.ast`` (the version the
reference imports: it uses ``nmigen.sim``).  It exists so that the reference's OWN
source files (/root/reference/mfcc/core/*.py, mfcc/misc/*.py) can be executed here
without the nmigen package, to produce golden vectors (tests/golden/make_rtl_golden.py).
Nothing in the product imports it.

Every Value evaluates to an exact Python integer consistent with its shape
(non-negative if unsigned, two's complement value if signed); truncation happens on
assignment, slicing and concatenation, as in the HDL.
"""
from __future__ import annotations

import builtins
import os
from collections import OrderedDict

__all__ = ["Shape", "signed", "unsigned", "Value", "Const", "C", "Signal", "Cat", "Repl", "Mux",
           "Array", "Statement", "Assign", "bits_for", "ClockSignal", "ResetSignal"]


def bits_for(n, require_sign_bit=False):
    if n > 0:
        r = n.bit_length()
    else:
        require_sign_bit = True
        r = (-n - 1).bit_length() if n < 0 else 0
    if require_sign_bit:
        r += 1
    return r


class Shape:
    def __init__(self, width=1, signed=False):
        if not isinstance(width, int) or width < 0:
            raise TypeError("Width must be a non-negative integer, not {!r}".format(width))
        self.width, self.signed = width, bool(signed)

    def __iter__(self):
        return iter((self.width, self.signed))

    def __eq__(self, other):
        if isinstance(other, tuple) and len(other) == 2:
            return (self.width, self.signed) == other
        return isinstance(other, Shape) and (self.width, self.signed) == (other.width, other.signed)

    def __hash__(self):
        return hash((self.width, self.signed))

    def __repr__(self):
        return ("signed(%d)" if self.signed else "unsigned(%d)") % self.width

    @staticmethod
    def cast(obj):
        if isinstance(obj, Shape):
            return obj
        if isinstance(obj, bool):
            raise TypeError("bool is not a shape")
        if isinstance(obj, int):
            return Shape(obj)
        if isinstance(obj, tuple):
            w, s = obj
            return Shape(w, s)
        if isinstance(obj, range):
            if len(obj) == 0:
                return Shape(0, obj.start < 0)
            sgn = obj.start < 0 or (obj.stop - obj.step) < 0
            w = max(bits_for(obj.start, sgn), bits_for(obj.stop - obj.step, sgn))
            return Shape(w, sgn)
        raise TypeError("Object {!r} cannot be used as value shape".format(obj))


def signed(width):
    return Shape(width, True)


def unsigned(width):
    return Shape(width, False)


def _mask(w):
    return (1 << w) - 1


def _wrap_src(src, shape):
    """Python source that reinterprets the low bits of `src` in `shape`."""
    w, s = shape.width, shape.signed
    if w == 0:
        return "0"
    if s:
        h = 1 << (w - 1)
        return "((((%s) + %d) & %d) - %d)" % (src, h, _mask(w), h)
    return "((%s) & %d)" % (src, _mask(w))


def _bitwise_shape(a, b):
    """Shape two operands are extended to before a bitwise/arith op (ast.py, 0.3)."""
    if not a.signed and not b.signed:
        return Shape(max(a.width, b.width), False)
    if a.signed and b.signed:
        return Shape(max(a.width, b.width), True)
    if not a.signed and b.signed:
        return Shape(max(a.width + 1, b.width), True)
    return Shape(max(a.width, b.width + 1), True)


class Value:
    @staticmethod
    def cast(obj):
        if isinstance(obj, Value):
            return obj
        if isinstance(obj, (bool, int)) or hasattr(obj, "__index__"):
            return Const(int(obj))
        raise TypeError("Object {!r} cannot be converted to a Value".format(obj))

    # -- to be provided by subclasses
    def shape(self):
        raise NotImplementedError

    def _src(self, rd):
        """Python expression for the value; rd(signal) gives the expression of a signal."""
        raise NotImplementedError

    def _deps(self, acc):
        """Add every Signal / memory read by this value to `acc`."""
        raise NotImplementedError

    def _lhs(self, rhs, emit):
        raise TypeError("Value {!r} cannot be assigned to".format(self))

    def _lhs_signals(self, acc):
        raise TypeError("Value {!r} cannot be assigned to".format(self))

    def _lhs_deps(self, acc):
        """Signals READ when this value is an assignment target (dynamic offsets)."""

    # -- python protocol
    def __bool__(self):
        raise TypeError("Attempted to convert nMigen value to Python boolean")

    __hash__ = object.__hash__

    def __len__(self):
        return self.shape().width

    def __iter__(self):
        return iter([self[i] for i in range(len(self))])

    def __getitem__(self, key):
        n = len(self)
        if isinstance(key, int):
            if key not in range(-n, n):
                raise IndexError("Cannot index {} bits into {}-bit value".format(key, n))
            if key < 0:
                key += n
            return Slice(self, key, key + 1)
        if isinstance(key, builtins.slice):
            start, stop, step = key.indices(n)
            if step != 1:
                return Cat(self[i] for i in range(start, stop, step))
            return Slice(self, start, max(stop, start))
        raise TypeError("Cannot index value with {!r}".format(key))

    def __invert__(self):
        return Operator("~", [self])

    def __neg__(self):
        return Operator("-", [self])

    def __pos__(self):
        return self

    def __add__(self, o):
        return Operator("+", [self, o])

    def __radd__(self, o):
        return Operator("+", [o, self])

    def __sub__(self, o):
        return Operator("-", [self, o])

    def __rsub__(self, o):
        return Operator("-", [o, self])

    def __mul__(self, o):
        return Operator("*", [self, o])

    def __rmul__(self, o):
        return Operator("*", [o, self])

    def __and__(self, o):
        return Operator("&", [self, o])

    def __rand__(self, o):
        return Operator("&", [o, self])

    def __or__(self, o):
        return Operator("|", [self, o])

    def __ror__(self, o):
        return Operator("|", [o, self])

    def __xor__(self, o):
        return Operator("^", [self, o])

    def __rxor__(self, o):
        return Operator("^", [o, self])

    def __lshift__(self, o):
        o = Value.cast(o)
        if o.shape().signed:
            raise TypeError("Shift amount must be unsigned")
        return Operator("<<", [self, o])

    def __rlshift__(self, o):
        return Value.cast(o).__lshift__(self)

    def __rshift__(self, o):
        o = Value.cast(o)
        if o.shape().signed:
            raise TypeError("Shift amount must be unsigned")
        return Operator(">>", [self, o])

    def __rrshift__(self, o):
        return Value.cast(o).__rshift__(self)

    def __eq__(self, o):
        return Operator("==", [self, o])

    def __ne__(self, o):
        return Operator("!=", [self, o])

    def __lt__(self, o):
        return Operator("<", [self, o])

    def __le__(self, o):
        return Operator("<=", [self, o])

    def __gt__(self, o):
        return Operator(">", [self, o])

    def __ge__(self, o):
        return Operator(">=", [self, o])

    def __abs__(self):
        return Mux(self >= 0, self, -self) if self.shape().signed else self

    # -- nMigen methods
    def eq(self, value):
        return Assign(self, value)

    def bool(self):
        return Operator("b", [self])

    def any(self):
        return Operator("r|", [self])

    def all(self):
        return Operator("r&", [self])

    def xor(self):
        return Operator("r^", [self])

    def implies(self, conclusion):
        return ~self.bool() | Value.cast(conclusion).bool()

    def as_unsigned(self):
        return Operator("u", [self])

    def as_signed(self):
        return Operator("s", [self])

    def bit_select(self, offset, width):
        offset = Value.cast(offset)
        if isinstance(offset, Const):
            return self[offset.value:offset.value + width]
        return Part(self, offset, width, stride=1)

    def word_select(self, offset, width):
        offset = Value.cast(offset)
        if isinstance(offset, Const):
            return self[offset.value * width:(offset.value + 1) * width]
        return Part(self, offset, width, stride=width)

    def matches(self, *patterns):
        terms = []
        for p in patterns:
            if isinstance(p, str):
                p = "".join(p.split())
                msk = int(p.replace("0", "1").replace("-", "0"), 2)
                val = int(p.replace("-", "0"), 2)
                terms.append((self & msk) == val)
            else:
                terms.append(self == p)
        if not terms:
            return Const(0)
        r = terms[0]
        for t in terms[1:]:
            r = r | t
        return r

    def shift_left(self, amount):
        return Cat(Const(0, amount), self) if amount >= 0 else self.shift_right(-amount)

    def shift_right(self, amount):
        return self[amount:] if amount >= 0 else self.shift_left(-amount)


class Const(Value):
    def __init__(self, value, shape=None):
        value = int(value)
        if shape is None:
            shape = Shape(bits_for(value), value < 0)
        elif isinstance(shape, int):
            shape = Shape(shape, value < 0)
        else:
            shape = Shape.cast(shape)
        self._shape = shape
        w = shape.width
        v = value & _mask(w) if w else 0
        if shape.signed and w and v >> (w - 1):
            v -= 1 << w
        self.value = v

    @property
    def width(self):
        return self._shape.width

    @property
    def signed(self):
        return self._shape.signed

    def shape(self):
        return self._shape

    def _src(self, rd):
        return "(%d)" % self.value if self.value < 0 else "%d" % self.value

    def _deps(self, acc):
        pass

    def __repr__(self):
        return "(const %d'%sd%d)" % (self.width, "s" if self.signed else "", self.value)


C = Const

_sig_counter = [0]


class Signal(Value):
    def __init__(self, shape=None, *, name=None, reset=0, reset_less=False, attrs=None,
                 decoder=None, src_loc_at=0):
        self._shape = Shape(1, False) if shape is None else Shape.cast(shape)
        self.name = name or "$sig%d" % _sig_counter[0]
        _sig_counter[0] += 1
        self.reset = int(reset)
        self.reset_less = bool(reset_less)
        self.attrs = OrderedDict(() if attrs is None else attrs)
        self.decoder = decoder

    @property
    def width(self):
        return self._shape.width

    @property
    def signed(self):
        return self._shape.signed

    @classmethod
    def like(cls, other, *, name=None, name_suffix=None, src_loc_at=0, **kwargs):
        kw = dict(shape=Value.cast(other).shape(), name=name)
        if isinstance(other, Signal):
            kw.update(reset=other.reset, reset_less=other.reset_less, attrs=other.attrs,
                      decoder=other.decoder)
            if name is None:
                kw["name"] = other.name + (name_suffix or "$like")
        kw.update(kwargs)
        return cls(**kw)

    def shape(self):
        return self._shape

    def reset_value(self):
        return Const(self.reset, self._shape).value

    def _src(self, rd):
        return rd(self)

    def _deps(self, acc):
        acc.add(self)

    def _lhs(self, rhs, emit):
        emit(self, _wrap_src(rhs, self._shape))

    def _lhs_signals(self, acc):
        acc.add(self)

    def __repr__(self):
        return "(sig %s)" % self.name


class ClockSignal(Value):
    def __init__(self, domain="sync"):
        self.domain = domain

    def shape(self):
        return Shape(1)


class ResetSignal(Value):
    def __init__(self, domain="sync", allow_reset_less=False):
        self.domain = domain

    def shape(self):
        return Shape(1)


class Operator(Value):
    def __init__(self, op, operands):
        self.operator = op
        self.operands = [Value.cast(o) for o in operands]

    def shape(self):
        op, sh = self.operator, [o.shape() for o in self.operands]
        if len(sh) == 1:
            a = sh[0]
            if op in ("~",):
                return Shape(a.width, a.signed)
            if op == "-":
                return Shape(a.width + 1, True)
            if op in ("b", "r|", "r&", "r^"):
                return Shape(1, False)
            if op == "u":
                return Shape(a.width, False)
            if op == "s":
                return Shape(a.width, True)
        else:
            a, b = sh
            if op in ("+", "-"):
                # nMigen 0.3: width of the common extension plus one; an unsigned
                # difference stays unsigned (wraps), Operator.shape in hdl/ast.py
                c = _bitwise_shape(a, b)
                return Shape(c.width + 1, c.signed)
            if op == "*":
                return Shape(a.width + b.width, a.signed or b.signed)
            if op in ("&", "|", "^"):
                return _bitwise_shape(a, b)
            if op in ("==", "!=", "<", "<=", ">", ">="):
                return Shape(1, False)
            if op == "<<":
                return Shape(a.width + (1 << b.width) - 1, a.signed)
            if op == ">>":
                return Shape(a.width, a.signed)
        raise NotImplementedError(op)

    def _deps(self, acc):
        for o in self.operands:
            o._deps(acc)

    def _src(self, rd):
        op = self.operator
        s = [o._src(rd) for o in self.operands]
        sh = [o.shape() for o in self.operands]
        if len(s) == 1:
            a, = s
            w = sh[0].width
            if op == "~":
                return "(~%s)" % a if sh[0].signed else "((~%s) & %d)" % (a, _mask(w))
            if op == "-":
                return "(-%s)" % a
            if op in ("b", "r|"):
                return "(1 if %s else 0)" % a
            if op == "r&":
                return "(1 if (%s & %d) == %d else 0)" % (a, _mask(w), _mask(w))
            if op == "r^":
                return "(bin(%s & %d).count('1') & 1)" % (a, _mask(w))
            if op == "u":
                return "(%s & %d)" % (a, _mask(w))
            if op == "s":
                return _wrap_src(a, Shape(w, True))
        else:
            a, b = s
            if op in ("+", "*", "&", "|", "^"):
                return "(%s %s %s)" % (a, op, b)
            if op == "-":
                r = "(%s - %s)" % (a, b)
                if not self.shape().signed and os.environ.get("HDL_SHIM_SUB") != "signed":
                    # unsigned - unsigned wraps at the result width (nMigen 0.3; later
                    # releases make it signed -- HDL_SHIM_SUB=signed selects that, and
                    # tests/test_hdl_shim.py checks the design gives the same vectors)
                    return "_usub(%s, %d)" % (r, _mask(self.shape().width))
                return r
            if op in ("==", "!=", "<", "<=", ">", ">="):
                return "(1 if %s %s %s else 0)" % (a, op, b)
            if op == "<<":
                return "(%s << %s)" % (a, b)
            if op == ">>":
                return "(%s >> %s)" % (a, b)
        raise NotImplementedError(op)

    def __repr__(self):
        return "(%s %s)" % (self.operator, " ".join(map(repr, self.operands)))


def Mux(sel, val1, val0):
    sel = Value.cast(sel)
    if len(sel) != 1:
        sel = sel.bool()
    return _Mux(sel, Value.cast(val1), Value.cast(val0))


class _Mux(Value):
    def __init__(self, sel, a, b):
        self.sel, self.a, self.b = sel, a, b

    def shape(self):
        return _bitwise_shape(self.a.shape(), self.b.shape())

    def _deps(self, acc):
        self.sel._deps(acc)
        self.a._deps(acc)
        self.b._deps(acc)

    def _src(self, rd):
        return "(%s if %s else %s)" % (self.a._src(rd), self.sel._src(rd), self.b._src(rd))


class Slice(Value):
    def __init__(self, value, start, stop):
        self.value, self.start, self.stop = Value.cast(value), start, stop

    def shape(self):
        return Shape(self.stop - self.start, False)

    def _deps(self, acc):
        self.value._deps(acc)

    def _src(self, rd):
        w = self.stop - self.start
        if w == 0:
            return "0"
        return "((%s >> %d) & %d)" % (self.value._src(rd), self.start, _mask(w))

    def _lhs(self, rhs, emit):
        w = self.stop - self.start
        if w == 0:
            return
        m = _mask(w) << self.start
        cur = self.value._src(emit.rd)
        self.value._lhs("((%s & ~%d) | (((%s) << %d) & %d))" % (cur, m, rhs, self.start, m), emit)

    def _lhs_signals(self, acc):
        self.value._lhs_signals(acc)

    def _lhs_deps(self, acc):
        self.value._lhs_deps(acc)

    def __repr__(self):
        return "(slice %r %d:%d)" % (self.value, self.start, self.stop)


class Part(Value):
    def __init__(self, value, offset, width, stride=1):
        self.value, self.offset = Value.cast(value), Value.cast(offset)
        self.width, self.stride = width, stride

    def shape(self):
        return Shape(self.width, False)

    def _deps(self, acc):
        self.value._deps(acc)
        self.offset._deps(acc)

    def _src(self, rd):
        n = len(self.value)
        return "(((%s & %d) >> ((%s & %d) * %d)) & %d)" % (
            self.value._src(rd), _mask(n), self.offset._src(rd), _mask(len(self.offset)),
            self.stride, _mask(self.width))

    def _lhs(self, rhs, emit):
        n = len(self.value)
        cur = self.value._src(emit.rd)
        off = "((%s & %d) * %d)" % (self.offset._src(emit.rd0), _mask(len(self.offset)), self.stride)
        self.value._lhs("_setpart(%s, %s, %d, %s, %d)" % (cur, off, _mask(self.width), rhs, _mask(n)),
                        emit)

    def _lhs_signals(self, acc):
        self.value._lhs_signals(acc)

    def _lhs_deps(self, acc):
        self.offset._deps(acc)
        self.value._lhs_deps(acc)


class Cat(Value):
    def __init__(self, *args):
        self.parts = [Value.cast(x) for x in _flatten(args)]

    def shape(self):
        return Shape(sum(len(p) for p in self.parts), False)

    def _deps(self, acc):
        for p in self.parts:
            p._deps(acc)

    def _src(self, rd):
        terms, off = [], 0
        for p in self.parts:
            w = len(p)
            if w:
                t = "(%s & %d)" % (p._src(rd), _mask(w))
                terms.append(t if off == 0 else "(%s << %d)" % (t, off))
            off += w
        return "(%s)" % " | ".join(terms) if terms else "0"

    def _lhs(self, rhs, emit):
        off = 0
        for p in self.parts:
            w = len(p)
            if w:
                p._lhs("(((%s) >> %d) & %d)" % (rhs, off, _mask(w)), emit)
            off += w

    def _lhs_signals(self, acc):
        for p in self.parts:
            p._lhs_signals(acc)

    def _lhs_deps(self, acc):
        for p in self.parts:
            p._lhs_deps(acc)

    def __repr__(self):
        return "(cat %s)" % " ".join(map(repr, self.parts))


def _flatten(x):
    if isinstance(x, Value) or not hasattr(x, "__iter__"):
        yield x
    else:
        for e in x:
            yield from _flatten(e)


def Repl(value, count):
    value = Value.cast(value)
    return Cat(*[value for _ in range(count)])


class Array(list):
    """Indexable by a Value: a mux over the elements (only read access is modelled)."""

    def __getitem__(self, idx):
        if isinstance(idx, Value):
            return _ArrayProxy(self, idx)
        return list.__getitem__(self, idx)


class _ArrayProxy(Value):
    def __init__(self, elems, index):
        self.elems = [Value.cast(e) for e in elems]
        self.index = index

    def shape(self):
        sh = self.elems[0].shape()
        for e in self.elems[1:]:
            sh = _bitwise_shape(sh, e.shape())
        return sh

    def _deps(self, acc):
        self.index._deps(acc)
        for e in self.elems:
            e._deps(acc)

    def _src(self, rd):
        n = len(self.elems)
        return "(%s)[min(%s, %d)]" % (", ".join(e._src(rd) for e in self.elems) + ",",
                                      self.index._src(rd), n - 1)


class Statement:
    pass


class Assign(Statement):
    def __init__(self, lhs, rhs):
        self.lhs = lhs if isinstance(lhs, Value) else Value.cast(lhs)
        self.rhs = Value.cast(rhs)
        self.domain = None

    def __repr__(self):
        return "(eq %r %r)" % (self.lhs, self.rhs)
