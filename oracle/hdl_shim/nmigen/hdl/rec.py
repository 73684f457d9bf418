"""TEST INFRASTRUCTURE -- Record / Layout of the small nMigen interpreter (see ../_ast.py)."""
from __future__ import annotations

from collections import OrderedDict
from enum import Enum
from functools import reduce

from .._ast import Shape, Signal, Value, Cat

__all__ = ["Direction", "DIR_NONE", "DIR_FANOUT", "DIR_FANIN", "Layout", "Record"]


class Direction(Enum):
    NONE = "none"
    FANOUT = "fanout"
    FANIN = "fanin"


DIR_NONE, DIR_FANOUT, DIR_FANIN = Direction.NONE, Direction.FANOUT, Direction.FANIN


class Layout:
    @staticmethod
    def cast(obj, *, src_loc_at=0):
        return obj if isinstance(obj, Layout) else Layout(obj)

    def __init__(self, fields, *, src_loc_at=0):
        self.fields = OrderedDict()
        for field in fields:
            if not isinstance(field, tuple) or len(field) not in (2, 3):
                raise TypeError("Field {!r} has invalid layout".format(field))
            if len(field) == 2:
                name, shape = field
                direction = DIR_NONE
                if isinstance(shape, list):
                    shape = Layout.cast(shape)
            else:
                name, shape, direction = field
                if not isinstance(direction, Direction):
                    raise TypeError("Field {!r} has invalid direction".format(field))
            if not isinstance(shape, Layout):
                shape = Shape.cast(shape)
            if name in self.fields:
                raise NameError("Field {!r} has a name that is already present".format(field))
            self.fields[name] = (shape, direction)

    def __getitem__(self, item):
        return self.fields[item]

    def __iter__(self):
        for name, (shape, dir) in self.fields.items():
            yield (name, shape, dir)

    def __eq__(self, other):
        return isinstance(other, Layout) and self.fields == other.fields


class Record(Value):
    @staticmethod
    def like(other, *, name=None, name_suffix=None, src_loc_at=0):
        fields = OrderedDict()
        for fname, f in other.fields.items():
            if isinstance(f, Record):
                fields[fname] = Record.like(f, name=None if name is None else name + "__" + fname)
            else:
                fields[fname] = Signal.like(f, name=(name or "like") + "__" + fname)
        return Record(other.layout, name=name, fields=fields)

    def __init__(self, layout, *, name=None, fields=None, src_loc_at=0):
        self.name = name or "rec"
        self.layout = Layout.cast(layout)
        self.fields = OrderedDict()
        for fname, (fshape, fdir) in self.layout.fields.items():
            if fields is not None and fname in fields:
                self.fields[fname] = fields[fname]
            elif isinstance(fshape, Layout):
                self.fields[fname] = Record(fshape, name=self.name + "__" + fname)
            else:
                self.fields[fname] = Signal(fshape, name=self.name + "__" + fname)

    def __getattr__(self, name):
        try:
            return self.__dict__["fields"][name]
        except KeyError:
            raise AttributeError("Record '{}' does not have a field '{}'".format(
                self.__dict__.get("name"), name)) from None

    def __getitem__(self, item):
        if isinstance(item, str):
            try:
                return self.fields[item]
            except KeyError:
                raise AttributeError("Record does not have a field '{}'".format(item)) from None
        return Value.__getitem__(self, item)

    def _as_cat(self):
        return Cat(*self.fields.values())

    def shape(self):
        return Shape(sum(len(f) for f in self.fields.values()), False)

    def _src(self, rd):
        return self._as_cat()._src(rd)

    def _deps(self, acc):
        for f in self.fields.values():
            f._deps(acc)

    def _lhs(self, rhs, emit):
        self._as_cat()._lhs(rhs, emit)

    def _lhs_signals(self, acc):
        for f in self.fields.values():
            f._lhs_signals(acc)

    def _lhs_deps(self, acc):
        for f in self.fields.values():
            f._lhs_deps(acc)

    def connect(self, *subordinates, include=None, exclude=None):
        def rec_name(r):
            return r.name

        for field in include or {}:
            if field not in self.fields:
                raise AttributeError("Cannot include field '{}' because it is not present in "
                                     "record '{}'".format(field, rec_name(self)))
        for field in exclude or {}:
            if field not in self.fields:
                raise AttributeError("Cannot exclude field '{}' because it is not present in "
                                     "record '{}'".format(field, rec_name(self)))
        stmts = []
        for field in self.fields:
            if include is not None and field not in include:
                continue
            if exclude is not None and field in exclude:
                continue
            shape, direction = self.layout[field]
            if not isinstance(shape, Layout) and direction == DIR_NONE:
                raise TypeError("Cannot connect field '{}' of {} because it does not have "
                                "a direction".format(field, rec_name(self)))
            item = self.fields[field]
            subord_items = []
            for s in subordinates:
                if field not in s.fields:
                    raise AttributeError("Cannot connect field '{}' of record '{}' to subordinate "
                                         "record '{}' because the subordinate record does not "
                                         "have this field".format(field, rec_name(self), rec_name(s)))
                subord_items.append(s.fields[field])
            if isinstance(shape, Layout):
                sub_inc = include[field] if include and field in include and isinstance(include, dict) else None
                sub_exc = exclude[field] if exclude and field in exclude and isinstance(exclude, dict) else None
                stmts += item.connect(*subord_items, include=sub_inc, exclude=sub_exc)
            elif direction == DIR_FANOUT:
                stmts += [sub.eq(item) for sub in subord_items]
            elif direction == DIR_FANIN:
                stmts += [item.eq(reduce(lambda a, b: a | b, subord_items))]
        return stmts

    def __repr__(self):
        return "(rec %s %s)" % (self.name, " ".join(self.fields))
