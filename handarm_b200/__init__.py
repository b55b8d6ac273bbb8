"""Importable alias of the ``isaacgym-hand-arm_b200/`` package directory.

The product directory carries the reference's name (with hyphens), which Python cannot import; this
shim makes ``import handarm_b200`` resolve every submodule from that directory.
"""
import os as _os

_impl = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "isaacgym-hand-arm_b200")
__path__ = [_impl]
with open(_os.path.join(_impl, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_impl, "__init__.py"), "exec"))
