"""Public-surface parity with the reference, checked against its source tree when it is present (read-only, parsed with
``ast`` - nothing is imported or executed from it): every function / class / method resolves through the ``elephas.*``
import paths, accepts the reference's parameter names and has the same default values."""
import ast
import importlib
import inspect
import os

import pytest

REF = '/root/reference/elephas'
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason='reference tree not available on this machine')


def _modules():
    for dp, _, files in os.walk(REF):
        for f in sorted(files):
            if f.endswith('.py'):
                rel = os.path.relpath(os.path.join(dp, f), os.path.dirname(REF))
                mod = rel[:-3].replace(os.sep, '.')
                yield (mod[:-9] if mod.endswith('.__init__') else mod), ast.parse(open(os.path.join(dp, f)).read())


def _params(fn):
    return [a.arg for a in fn.args.args + fn.args.kwonlyargs if a.arg not in ('self', 'cls')]


def _defaults(fn):
    out = {}
    pos = fn.args.args
    for a, d in zip(pos[len(pos) - len(fn.args.defaults):], fn.args.defaults):
        out[a.arg] = d
    for a, d in zip(fn.args.kwonlyargs, fn.args.kw_defaults):
        if d is not None:
            out[a.arg] = d
    return out


def _check_callable(where, node, obj, problems):
    try:
        sig = inspect.signature(obj)
    except (TypeError, ValueError):
        return
    has_kwargs = any(p.kind == p.VAR_KEYWORD for p in sig.parameters.values())
    for name in _params(node):
        if name not in sig.parameters and not has_kwargs:
            problems.append('%s: parameter %r missing' % (where, name))
    for name, dflt in _defaults(node).items():
        p = sig.parameters.get(name)
        if p is None or p.default is inspect.Parameter.empty:
            continue
        try:
            want = ast.literal_eval(dflt)
        except (ValueError, SyntaxError):
            continue
        if p.default != want:
            problems.append('%s: default of %r is %r, reference %r' % (where, name, p.default, want))


def test_every_reference_name_parameter_and_default_is_there():
    problems = []
    for mod, tree in _modules():
        m = importlib.import_module(mod)
        for node in tree.body:
            if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and not hasattr(m, node.name):
                problems.append('%s.%s missing' % (mod, node.name))
                continue
            if isinstance(node, ast.FunctionDef):
                _check_callable('%s.%s' % (mod, node.name), node, getattr(m, node.name), problems)
            elif isinstance(node, ast.ClassDef):
                cls = getattr(m, node.name)
                for item in node.body:
                    if not isinstance(item, ast.FunctionDef):
                        continue
                    if not hasattr(cls, item.name):
                        problems.append('%s.%s.%s missing' % (mod, node.name, item.name))
                    else:
                        _check_callable('%s.%s.%s' % (mod, node.name, item.name), item, getattr(cls, item.name), problems)
    assert not problems, '\n'.join(problems)
