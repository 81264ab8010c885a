"""Network definitions.

``RefCNN``   plain PyTorch model with the reference's topology, layouts and init — the numerical
             oracle for the kernels and the model of the NCCL + elementwise baseline.
``B200CNN``  the same network as an nn.Module whose layers are the hand-written sm_100a kernels
             (``ops.functional``); the ``Engine`` is the fused, graph-captured form of it.
"""
from .reference import RefCNN, load_dense_params, dense_params, dense_grads  # noqa: F401
from .b200 import B200CNN  # noqa: F401
