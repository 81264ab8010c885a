"""ctypes wrapper of the native fp64 CPU engine (csrc/cpu/net.cpp).

API mirrors the reference's Layer_* functions (/root/reference/cnn.c:249-342): build a layer list,
``set_inputs`` (forward), ``learn_outputs`` (loss gradient + backward), ``error_total``, ``update``.
"""
import ctypes as C

import numpy as np

from .. import _native as N

QUIRK_SHARED_CONV_SLICE, QUIRK_EARLY_UPDATE, QUIRK_SOFTMAX_M1 = 1, 2, 4
QUIRKS_REFERENCE = 7


def _dp(a):
    return a.ctypes.data_as(N.c_double_p)


class CpuNet:
    def __init__(self, quirks=0, seed=0):
        self.lib = N.cpu()
        self.handle = self.lib.b200cnn_cpu_net_create(quirks)
        self.lib.b200cnn_cpu_net_seed(self.handle, seed)

    def __del__(self):
        try:
            if self.handle:
                self.lib.b200cnn_cpu_net_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    # constructors ------------------------------------------------------------------------------
    def add_input(self, depth, width, height):
        return self.lib.b200cnn_cpu_net_add_input(self.handle, depth, width, height)

    def add_conv(self, depth, width, height, kernsize, padding, stride, std=0.1):
        return self.lib.b200cnn_cpu_net_add_conv(self.handle, depth, width, height, kernsize, padding, stride, std)

    def add_full(self, nnodes, std=0.1):
        return self.lib.b200cnn_cpu_net_add_full(self.handle, nnodes, std)

    def add_pool(self, kernsize, stride):
        return self.lib.b200cnn_cpu_net_add_pool(self.handle, kernsize, stride)

    @classmethod
    def from_layers(cls, layers, seed=0, std=0.1, quirks=0):
        """layers: [(type, depth, width, height, k, pad, stride)] as used by ``Engine``."""
        net = cls(quirks=quirks, seed=seed)
        for (t, d, w, h, k, pad, stride) in layers:
            if t == N.LAYER_INPUT:
                net.add_input(d, w, h)
            elif t == N.LAYER_CONV:
                net.add_conv(d, w, h, k, pad, stride, std)
            elif t == N.LAYER_FULL:
                net.add_full(d * w * h, std)
            elif t == N.LAYER_POOL:
                net.add_pool(k, stride)
            else:
                raise ValueError(f"unknown layer type {t}")
        return net

    @classmethod
    def mnist(cls, seed=0, quirks=0):
        from ..engine import mnist_layers

        return cls.from_layers(mnist_layers(), seed=seed, quirks=quirks)

    # info --------------------------------------------------------------------------------------
    @property
    def num_layers(self):
        return self.lib.b200cnn_cpu_net_num_layers(self.handle)

    @property
    def num_params(self):
        return self.lib.b200cnn_cpu_net_num_params(self.handle)

    def layer_info(self, layer):
        info = (C.c_int * 8)()
        nb = self.lib.b200cnn_cpu_net_layer_info(self.handle, layer, info)
        keys = ("type", "depth", "width", "height", "kernsize", "padding", "stride", "nweights")
        d = dict(zip(keys, list(info)))
        d["nbiases"] = nb
        return d

    # Layer_* equivalents -------------------------------------------------------------------------
    def set_inputs(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1)
        self.lib.b200cnn_cpu_net_set_inputs(self.handle, _dp(x))

    def get_outputs(self):
        n = self.layer_info(self.num_layers - 1)
        y = np.empty(n["depth"] * n["width"] * n["height"], dtype=np.float64)
        self.lib.b200cnn_cpu_net_get_outputs(self.handle, _dp(y))
        return y

    def learn_outputs(self, target):
        t = np.ascontiguousarray(target, dtype=np.float64).reshape(-1)
        self.lib.b200cnn_cpu_net_learn_outputs(self.handle, _dp(t))

    def error_total(self):
        return self.lib.b200cnn_cpu_net_error_total(self.handle)

    # mini-batch schedule (csrc/cpu/net_batch.cpp) ------------------------------------------------
    @property
    def batch_supported(self):
        return bool(self.lib.b200cnn_cpu_net_batch_supported(self.handle))

    def learn_batch(self, xs, targets):
        """n x (set_inputs + learn_outputs) in one call; returns the per-sample error_total() values."""
        xs = np.ascontiguousarray(xs, dtype=np.float64)
        n = xs.shape[0]
        xs = xs.reshape(n, -1)
        t = np.ascontiguousarray(targets, dtype=np.float64).reshape(n, -1)
        err = np.empty(n, dtype=np.float64)
        self.lib.b200cnn_cpu_net_learn_batch(self.handle, _dp(xs), _dp(t), n, _dp(err))
        return err

    def forward_batch(self, xs):
        xs = np.ascontiguousarray(xs, dtype=np.float64)
        n = xs.shape[0]
        xs = xs.reshape(n, -1)
        info = self.layer_info(self.num_layers - 1)
        ys = np.empty((n, info["depth"] * info["width"] * info["height"]), dtype=np.float64)
        self.lib.b200cnn_cpu_net_forward_batch(self.handle, _dp(xs), n, _dp(ys))
        return ys

    def update(self, rate):
        self.lib.b200cnn_cpu_net_update(self.handle, rate)

    def rand(self):
        return self.lib.b200cnn_cpu_net_rand(self.handle)

    # flat views ---------------------------------------------------------------------------------
    def get_params(self):
        out = np.empty(self.num_params, dtype=np.float64)
        self.lib.b200cnn_cpu_net_get_params(self.handle, _dp(out))
        return out

    def set_params(self, flat):
        a = np.ascontiguousarray(flat, dtype=np.float64).reshape(-1)
        assert a.size == self.num_params
        self.lib.b200cnn_cpu_net_set_params(self.handle, _dp(a))

    def get_updates(self):
        out = np.empty(self.num_params, dtype=np.float64)
        self.lib.b200cnn_cpu_net_get_updates(self.handle, _dp(out))
        return out

    def layer_array(self, layer, which="outputs"):
        idx = {"outputs": 0, "gradients": 1, "errors": 2}[which]
        info = self.layer_info(layer)
        out = np.empty(info["depth"] * info["width"] * info["height"], dtype=np.float64)
        self.lib.b200cnn_cpu_net_layer_array(self.handle, layer, idx, _dp(out))
        return out

    def save(self, path):
        return self.lib.b200cnn_cpu_net_save(self.handle, str(path).encode())

    def load(self, path):
        return self.lib.b200cnn_cpu_net_load(self.handle, str(path).encode())


def glibc_rand(seed, n):
    out = (C.c_int * n)()
    N.cpu().b200cnn_glibc_rand(seed, n, out)
    return list(out)
