"""Drop-in for the reference's ``nnunet/lib/raft_initial.py`` (imported at
nnunet/network_architecture/SegFlowGaussian.py:42: ``from nnunet.lib.raft_initial import CorrBlock, BasicUpdateBlock,
coords_grid``; the file itself was never committed upstream).  Loaded under that module name by
``shim/cardioflow_dropin.py``.  ``CorrBlock`` and ``coords_grid`` come from cardioflow_b200; every other name
(``BasicUpdateBlock``: conv-GRU, cuDNN work outside the hot path) passes through from the user's own RAFT copy: the
``raft_initial.py`` found in the ``nnunet.lib`` package of their tree, or the module named by ``CARDIOFLOW_RAFT_MODULE``."""
import importlib
import os

from cardioflow_b200 import CorrBlock, coords_grid  # noqa: F401


def __getattr__(name):
    import cardioflow_dropin
    up = cardioflow_dropin.upstream_module(__name__)
    if up is None and os.environ.get("CARDIOFLOW_RAFT_MODULE"):
        up = importlib.import_module(os.environ["CARDIOFLOW_RAFT_MODULE"])
    if up is not None and hasattr(up, name):
        return getattr(up, name)
    raise AttributeError(
        f"{name!r} is not part of the optical-flow hot path cardioflow_b200 replaces, and no upstream "
        "nnunet/lib/raft_initial.py (nor a module named by CARDIOFLOW_RAFT_MODULE) provides it")
