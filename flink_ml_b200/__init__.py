"""Importable alias of the package directory ``flink-ml_b200/`` (a hyphen is not a legal
module name, so this shim points ``flink_ml_b200`` at that directory)."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "flink-ml_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _os, _f
