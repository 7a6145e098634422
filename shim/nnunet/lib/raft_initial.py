"""Drop-in for the reference's missing ``nnunet/lib/raft_initial.py`` (imported at
nnunet/network_architecture/SegFlowGaussian.py:42): CorrBlock and coords_grid come from cardioflow_b200.
``BasicUpdateBlock`` (conv-GRU, cuDNN work) is outside the hot path and must come from the user's RAFT copy."""
from cardioflow_b200 import CorrBlock, coords_grid  # noqa: F401
