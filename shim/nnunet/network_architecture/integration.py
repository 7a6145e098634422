"""Drop-in for nnunet/network_architecture/integration.py: same class names, same forward signatures,
same ``grid`` buffer, CUDA kernels underneath."""
from cardioflow_b200 import SpatialTransformer, SpatialTransformerContour, VecInt  # noqa: F401
