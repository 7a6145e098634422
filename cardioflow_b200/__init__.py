"""Importable alias of the package directory ``cardiac-segmentation-optical-flow_b200/`` (a hyphen is not a
valid module name).  ``import cardioflow_b200`` executes that package's ``__init__`` with this module's
``__path__`` pointing at it, so ``cardioflow_b200.corr`` etc. resolve to the real files."""
import os as _os

_REAL = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "cardiac-segmentation-optical-flow_b200")
__path__ = [_REAL]
with open(_os.path.join(_REAL, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_REAL, "__init__.py"), "exec"))
