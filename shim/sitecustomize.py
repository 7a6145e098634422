"""Runs at interpreter start when ``shim/`` is on PYTHONPATH: registers the cardioflow_b200 import hook (cheap, stdlib only).
Chains to any other ``sitecustomize`` further down the path so that an existing one keeps working."""
import importlib.util
import os
import sys

import cardioflow_dropin

cardioflow_dropin.install()

_here = os.path.dirname(os.path.abspath(__file__))
for _d in sys.path:
    _cand = os.path.join(_d or ".", "sitecustomize.py")
    if os.path.isfile(_cand) and os.path.abspath(os.path.dirname(_cand)) != _here:
        _spec = importlib.util.spec_from_file_location("_chained_sitecustomize", _cand)
        _spec.loader.exec_module(importlib.util.module_from_spec(_spec))
        break
