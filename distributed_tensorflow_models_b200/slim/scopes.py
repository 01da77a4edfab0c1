"""Argument scopes: default keyword arguments for decorated layer functions.

Same user-facing contract as the reference's ``inception/slim/scopes.py`` (``arg_scope`` :85-135,
``add_arg_scope`` :138-157, ``has_arg_scope`` :160-170):

    with arg_scope([ops.conv2d, ops.fc], stddev=0.01, weight_decay=5e-4):
        with arg_scope([ops.conv2d], padding='VALID') as sc:
            net = ops.conv2d(x, 64, [11, 11])          # stddev, weight_decay and padding applied
    with arg_scope(sc):                                  # a captured scope can be re-entered
        ...

Implementation: one stack of frames per thread (there is no TF graph to hang it on); a frame maps the
decorated function object's id-key to the kwargs that are in force.  Explicit call-site kwargs win.
"""
from __future__ import annotations

import contextlib
import functools
import inspect
import threading
from typing import Any, Callable, Dict, Iterable, Tuple, Union

_tls = threading.local()
_DECORATED: set = set()

ScopeKey = Tuple[str, str]
Scope = Dict[ScopeKey, Dict[str, Any]]


def _stack() -> list:
    st = getattr(_tls, "stack", None)
    if st is None:
        st = [{}]
        _tls.stack = st
    return st


def _key(func: Callable) -> ScopeKey:
    return (func.__module__, func.__name__)


def current_arg_scope() -> Scope:
    return _stack()[-1]


@contextlib.contextmanager
def arg_scope(list_ops_or_scope: Union[Iterable[Callable], Scope], **kwargs):
    """Set defaults for the listed (decorated) ops, or re-enter a previously captured scope dict."""
    st = _stack()
    if isinstance(list_ops_or_scope, dict):
        if kwargs:
            raise ValueError("kwargs must be empty when a captured scope (dict) is re-entered")
        frame = {k: dict(v) for k, v in list_ops_or_scope.items()}
    else:
        if not isinstance(list_ops_or_scope, (list, tuple)):
            raise TypeError("arg_scope expects a list/tuple of ops or a captured scope dict")
        frame = {k: dict(v) for k, v in st[-1].items()}
        for op in list_ops_or_scope:
            if not has_arg_scope(op):
                raise ValueError(f"{_key(op)} is not decorated with @add_arg_scope")
            merged = dict(frame.get(_key(op), {}))
            merged.update(kwargs)
            frame[_key(op)] = merged
    st.append(frame)
    try:
        yield frame
    finally:
        st.pop()


def add_arg_scope(func: Callable) -> Callable:
    """Decorator: the function picks up defaults from the enclosing ``arg_scope`` frames."""
    key = _key(func)
    try:
        positional = [n for n, p in inspect.signature(func).parameters.items()
                      if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
    except (TypeError, ValueError):
        positional = []

    @functools.wraps(func)
    def with_scoped_args(*args, **kwargs):
        defaults = current_arg_scope().get(key)
        if defaults:
            merged = dict(defaults)
            for name in positional[:len(args)]:      # an argument given positionally wins over the scope's default
                merged.pop(name, None)
            merged.update(kwargs)
            kwargs = merged
        return func(*args, **kwargs)

    _DECORATED.add(key)
    with_scoped_args._dtf_arg_scope_key = key  # type: ignore[attr-defined]
    return with_scoped_args


def has_arg_scope(func: Callable) -> bool:
    return _key(func) in _DECORATED
