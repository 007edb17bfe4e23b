"""Importable alias of the ``smartphone-based-rti_b200/`` package directory (whose name
contains hyphens): ``import rti_b200`` / ``import rti_b200.analysis`` resolve there."""
import os as _os

_pkg = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "smartphone-based-rti_b200")
__path__ = [_pkg]
with open(_os.path.join(_pkg, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_pkg, "__init__.py"), "exec"))
del _f
