"""Drop-in for the reference's missing ``nnunet/lib/raft.py`` (``from nnunet.lib.raft import CorrVolume``,
nnunet/network_architecture/SegFlowGaussian.py:41).  ``CorrVolume`` comes from cardioflow_b200 (PARITY UNPINNED: the upstream
definition was never committed -- cardioflow_b200/local_corr.py states the semantics chosen); every other name passes through from
the user's own file of that name when their tree has one."""
from cardioflow_b200 import CorrVolume  # noqa: F401


def __getattr__(name):
    import cardioflow_dropin
    up = cardioflow_dropin.upstream_module(__name__)
    if up is not None and hasattr(up, name):
        return getattr(up, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r} (and the tree has no nnunet/lib/raft.py providing it)")
