"""pyspark.ml.param-style parameter plumbing.

Mirrors the parts of pyspark.ml.param.Params the reference's idiom touches: keyword-only
constructors (ref:L135, L146, L162), setX/getX accessors, setParams, getOrDefault, isSet,
explainParams, copy.  Validation errors surface as IllegalArgumentException, the Python face of
the JVM exception pyspark raises for the same mistakes.
"""
from __future__ import annotations

import copy as _copy
from typing import Any, Callable, Dict, Optional


class IllegalArgumentException(ValueError):
    """Stand-in for pyspark.errors.IllegalArgumentException (bad Param value)."""


class Param:
    def __init__(self, name: str, doc: str, validator: Optional[Callable[[Any], bool]] = None, converter: Optional[Callable[[Any], Any]] = None):
        self.name, self.doc, self.validator, self.converter = name, doc, validator, converter

    def __repr__(self) -> str:
        return f"Param(name={self.name!r}, doc={self.doc!r})"


def _to_int(v):
    if isinstance(v, bool) or not isinstance(v, (int,)) and not (isinstance(v, float) and v == int(v)):
        try:
            import numpy as np
            if isinstance(v, np.integer):
                return int(v)
        except ImportError:
            pass
        raise TypeError(f"Could not convert {v!r} to int")
    return int(v)


def _to_float(v):
    if isinstance(v, bool):
        raise TypeError(f"Could not convert {v!r} to float")
    try:
        return float(v)
    except (TypeError, ValueError):
        raise TypeError(f"Could not convert {v!r} to float")


def _to_str(v):
    if not isinstance(v, str):
        raise TypeError(f"Could not convert {v!r} to string type")
    return v


def _to_bool(v):
    if not isinstance(v, bool):
        raise TypeError(f"Boolean Param requires value of type bool. Found {type(v)}.")
    return v


def _to_list_str(v):
    if isinstance(v, str) or not all(isinstance(x, str) for x in v):
        raise TypeError(f"Could not convert {v!r} to list of strings")
    return list(v)


class TypeConverters:
    toInt = staticmethod(_to_int)
    toFloat = staticmethod(_to_float)
    toString = staticmethod(_to_str)
    toBoolean = staticmethod(_to_bool)
    toListString = staticmethod(_to_list_str)


class Params:
    """Base of every Estimator / Transformer / Evaluator here."""

    _params: Dict[str, Param] = {}

    def __init__(self):
        self._paramMap: Dict[str, Any] = {}
        self._defaultParamMap: Dict[str, Any] = {}
        self.uid = f"{type(self).__name__}_{id(self) & 0xFFFFFFFFFFFF:012x}"

    # -- declaration helpers ---------------------------------------------------------------
    @classmethod
    def _declare(cls, *params: Param) -> Dict[str, Param]:
        merged = dict(getattr(cls, "_params", {}))
        for p in params:
            merged[p.name] = p
        return merged

    def _setDefault(self, **kw):
        for k, v in kw.items():
            self._defaultParamMap[k] = v
        return self

    def _set(self, **kw):
        for name, value in kw.items():
            if name not in self._params:
                raise AttributeError(f"{type(self).__name__} has no Param {name!r}")
            p = self._params[name]
            if value is not None and p.converter is not None:
                value = p.converter(value)
            if value is not None and p.validator is not None and not p.validator(value):
                raise IllegalArgumentException(f"{self.uid} parameter {name} given invalid value {value!r}.")
            self._paramMap[name] = value
        return self

    # -- pyspark surface -------------------------------------------------------------------
    @property
    def params(self):
        return [self._params[k] for k in sorted(self._params)]

    def hasParam(self, name: str) -> bool:
        return name in self._params

    def getParam(self, name: str) -> Param:
        return self._params[name]

    def isSet(self, param) -> bool:
        return _pname(param) in self._paramMap

    def hasDefault(self, param) -> bool:
        return _pname(param) in self._defaultParamMap

    def isDefined(self, param) -> bool:
        return self.isSet(param) or self.hasDefault(param)

    def getOrDefault(self, param):
        n = _pname(param)
        if n in self._paramMap:
            return self._paramMap[n]
        if n in self._defaultParamMap:
            return self._defaultParamMap[n]
        raise KeyError(f"Param {n} is neither set nor has a default")

    def set(self, param, value):
        return self._set(**{_pname(param): value})

    def clear(self, param):
        self._paramMap.pop(_pname(param), None)

    def extractParamMap(self, extra: Optional[dict] = None) -> dict:
        m = dict(self._defaultParamMap)
        m.update(self._paramMap)
        if extra:
            m.update({_pname(k): v for k, v in extra.items()})
        return m

    def explainParam(self, param) -> str:
        p = self._params[_pname(param)]
        bits = []
        if self.hasDefault(p):
            bits.append(f"default: {self._defaultParamMap[p.name]}")
        if self.isSet(p):
            bits.append(f"current: {self._paramMap[p.name]}")
        return f"{p.name}: {p.doc} ({', '.join(bits) if bits else 'undefined'})"

    def explainParams(self) -> str:
        return "\n".join(self.explainParam(p) for p in self.params)

    def copy(self, extra: Optional[dict] = None):
        that = _copy.copy(self)
        that._paramMap = dict(self._paramMap)
        that._defaultParamMap = dict(self._defaultParamMap)
        if extra:
            that._set(**{_pname(k): v for k, v in extra.items()})
        return that

    def _copyValues(self, to: "Params") -> "Params":
        for n, v in self._defaultParamMap.items():
            if n in to._params and n not in to._defaultParamMap:
                to._defaultParamMap[n] = v
        for n, v in self._paramMap.items():
            if n in to._params:
                to._paramMap[n] = v
        return to


def _pname(param) -> str:
    return param.name if isinstance(param, Param) else str(param)


def accessor(name: str):
    """Build (setX, getX) for Param `name` with pyspark's capitalisation."""
    cap = name[0].upper() + name[1:]

    def setter(self, value):
        return self._set(**{name: value})

    def getter(self):
        return self.getOrDefault(name)

    setter.__name__, getter.__name__ = "set" + cap, "get" + cap
    setter.__doc__ = f"Sets the value of :py:attr:`{name}`."
    getter.__doc__ = f"Gets the value of {name} or its default value."
    return setter, getter


def add_accessors(cls, names, setters=True):
    for n in names:
        s, g = accessor(n)
        if setters:
            setattr(cls, s.__name__, s)
        setattr(cls, g.__name__, g)
    return cls


def java_string_hash(s: str) -> int:
    """java.lang.String.hashCode (the JVM default seed is ClassName.hashCode.toLong)."""
    h = 0
    for ch in s:
        h = (31 * h + ord(ch)) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h
