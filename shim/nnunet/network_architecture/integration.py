"""Drop-in for nnunet/network_architecture/integration.py: same class names, same forward signatures, same ``grid``
buffer, CUDA kernels underneath.  Loaded under that module name by ``shim/cardioflow_dropin.py``;
any other name is passed through from the reference's own file."""
from cardioflow_b200 import SpatialTransformer, SpatialTransformerContour, VecInt  # noqa: F401


def __getattr__(name):
    import cardioflow_dropin
    up = cardioflow_dropin.upstream_module(__name__)
    if up is not None and hasattr(up, name):
        return getattr(up, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
