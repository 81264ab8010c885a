"""Kernel wrappers.

``functional`` — the sm_100a kernels as torch.autograd Functions (conv2d_relu, linear_act,
softmax_xent, max_pool2d, sgd_update_); ``cpu`` — the native fp64 CPU engine (numerical oracle).
"""
from . import cpu  # noqa: F401
