"""Python face of the native training engine (csrc/cuda/engine.cu).

One ``Engine.train_step`` = forward + backward + update of a whole mini-batch, enqueued by native
code (a single CUDA-graph launch once captured).  Replaces the reference's per-sample loop in
``main`` (/root/reference/cnn.c:451-474, CUDAcnn.cu:451-480).
"""
import ctypes as C

import numpy as np
import torch

from . import _native as N


def mnist_layers():
    """The reference's hard-coded topology (/root/reference/cnn.c:416-428)."""
    return [
        (N.LAYER_INPUT, 1, 28, 28, 0, 0, 1),
        (N.LAYER_CONV, 16, 14, 14, 3, 1, 2),
        (N.LAYER_CONV, 32, 7, 7, 3, 1, 2),
        (N.LAYER_FULL, 200, 1, 1, 0, 0, 1),
        (N.LAYER_FULL, 200, 1, 1, 0, 0, 1),
        (N.LAYER_FULL, 10, 1, 1, 0, 0, 1),
    ]


def pool_layers():
    """conv -> ReLU -> max-pool variant (the topology the project description names)."""
    return [
        (N.LAYER_INPUT, 1, 28, 28, 0, 0, 1),
        (N.LAYER_CONV, 16, 28, 28, 3, 1, 1),
        (N.LAYER_POOL, 16, 14, 14, 2, 0, 2),
        (N.LAYER_CONV, 32, 14, 14, 3, 1, 1),
        (N.LAYER_POOL, 32, 7, 7, 2, 0, 2),
        (N.LAYER_FULL, 200, 1, 1, 0, 0, 1),
        (N.LAYER_FULL, 200, 1, 1, 0, 0, 1),
        (N.LAYER_FULL, 10, 1, 1, 0, 0, 1),
    ]


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


class Engine:
    """Device-resident replica of a network plus its training step.

    Parameters are fp32; GEMM-shaped kernels multiply in TF32 on the tcgen05 path and in fp32 on the
    CUDA-core path, both accumulate in fp32.
    """

    def __init__(self, layers=None, max_batch=256, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("Engine needs a CUDA device (there is no CPU/PyTorch fallback; use ops.cpu.CpuNet)")
        self.lib = N.cuda()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.layers = list(layers if layers is not None else mnist_layers())
        self.max_batch = int(max_batch)
        arr = (N.LayerDesc * len(self.layers))(*[N.LayerDesc(*l) for l in self.layers])
        with torch.cuda.device(self.device):
            self.handle = self.lib.b200cnn_engine_create(arr, len(self.layers), self.max_batch)
        if not self.handle:
            raise RuntimeError("engine_create: " + self.lib.b200cnn_last_error().decode())
        self.num_params = self.lib.b200cnn_engine_num_params(self.handle)
        self.flat_floats = self.lib.b200cnn_engine_flat_floats(self.handle)
        self.num_classes = self.layers[-1][1]
        self.in_features = self.layers[0][1] * self.layers[0][2] * self.layers[0][3]
        self._metrics_host = torch.empty(4, dtype=torch.float32).pin_memory()
        self._comm = None

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200cnn_engine_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters ---------------------------------------------------------------------------
    def param_layout(self):
        """[(w_offset, w_size, b_offset, b_size)] per layer in floats inside the flat buffer."""
        n = len(self.layers)
        off = (C.c_long * (2 * n))()
        sz = (C.c_long * (2 * n))()
        self.lib.b200cnn_engine_param_layout(self.handle, off, sz)
        return [(off[2 * i], sz[2 * i], off[2 * i + 1], sz[2 * i + 1]) for i in range(n)]

    def set_params(self, dense):
        """dense: 1-D float array in layer order, weights then biases per layer (the CPU engine's order)."""
        a = np.ascontiguousarray(np.asarray(dense, dtype=np.float32).reshape(-1))
        assert a.size == self.num_params, (a.size, self.num_params)
        with torch.cuda.device(self.device):
            N.check(self.lib.b200cnn_engine_set_params(self.handle, a.ctypes.data_as(N.c_float_p)), "set_params")

    def get_params(self):
        out = np.empty(self.num_params, dtype=np.float32)
        with torch.cuda.device(self.device):
            N.check(self.lib.b200cnn_engine_get_params(self.handle, out.ctypes.data_as(N.c_float_p)), "get_params")
        return out

    def get_grads(self):
        out = np.empty(self.num_params, dtype=np.float32)
        with torch.cuda.device(self.device):
            N.check(self.lib.b200cnn_engine_get_grads(self.handle, out.ctypes.data_as(N.c_float_p)), "get_grads")
        return out

    def init_params(self, seed=0, std=0.1):
        """Reference initialisation: std * nrnd() from the glibc-compatible stream (cnn.c:49, :323-340)."""
        from .ops.cpu import CpuNet

        net = CpuNet.from_layers(self.layers, seed=seed, std=std)
        self.set_params(net.get_params().astype(np.float32))
        return net

    # ---- steps --------------------------------------------------------------------------------
    def _img_ptr(self, images):
        assert images.is_cuda and images.is_contiguous()
        is_u8 = 1 if images.dtype == torch.uint8 else 0
        if not is_u8:
            assert images.dtype == torch.float32
        batch = images.shape[0]
        assert images.numel() == batch * self.in_features, (tuple(images.shape), self.in_features)
        return C.c_void_p(images.data_ptr()), is_u8, batch

    def forward(self, images, labels=None, stream=None):
        """Runs the forward pass; returns nothing (read ``layer_output`` / ``metrics``)."""
        ptr, is_u8, batch = self._img_ptr(images)
        lp = C.c_void_p(labels.data_ptr()) if labels is not None else C.c_void_p(0)
        if labels is not None:
            assert labels.dtype == torch.uint8 and labels.is_cuda and labels.numel() == batch
        N.check(self.lib.b200cnn_engine_forward(self.handle, ptr, is_u8, lp, batch, _stream_ptr(stream)), "forward")

    def backward(self, stream=None):
        N.check(self.lib.b200cnn_engine_backward(self.handle, _stream_ptr(stream)), "backward")

    def sgd(self, lr, stream=None):
        N.check(self.lib.b200cnn_engine_sgd(self.handle, float(lr), _stream_ptr(stream)), "sgd")

    def train_step(self, images, labels, lr, use_graph=True, stream=None):
        """forward + backward + update.  ``lr`` multiplies the batch-SUM gradient (pass rate / batch)."""
        ptr, is_u8, batch = self._img_ptr(images)
        assert labels.dtype == torch.uint8 and labels.is_cuda and labels.numel() == batch
        N.check(
            self.lib.b200cnn_engine_train_step(
                self.handle, ptr, is_u8, C.c_void_p(labels.data_ptr()), batch, float(lr), 1 if use_graph else 0,
                _stream_ptr(stream)),
            "train_step")

    def train_step_host(self, images, labels, lr, metrics_out=None, stream=None):
        """One step fed from pinned HOST tensors (uint8 images [B, features], uint8 labels [B]).

        The engine copies them to a device staging slot on its own copy stream (double-buffered, so issuing steps back
        to back overlaps the next copy with the current step), replays the captured step, and — if ``metrics_out`` (a
        pinned float32[4] tensor) is given — copies {sum CE, sum MSE, #correct, #samples} back.  Nothing blocks;
        synchronise the stream (or an event recorded after this call) before reading ``metrics_out``."""
        assert not images.is_cuda and images.is_pinned() and not labels.is_cuda and labels.is_pinned()
        assert labels.dtype == torch.uint8 and images.is_contiguous() and labels.is_contiguous()
        is_u8 = 1 if images.dtype == torch.uint8 else 0
        batch = images.shape[0]
        assert images.numel() == batch * self.in_features and labels.numel() == batch
        mp = C.c_void_p(metrics_out.data_ptr()) if metrics_out is not None else C.c_void_p(0)
        N.check(self.lib.b200cnn_engine_train_step_host(self.handle, C.c_void_p(images.data_ptr()), is_u8, C.c_void_p(labels.data_ptr()),
                                                        batch, float(lr), mp, _stream_ptr(stream)), "train_step_host")

    def forward_backward(self, images, labels, use_graph=True, stream=None):
        """forward + backward as one (graph-captured) launch; gradients are left in the gradient buffer."""
        ptr, is_u8, batch = self._img_ptr(images)
        assert labels.dtype == torch.uint8 and labels.is_cuda and labels.numel() == batch
        N.check(self.lib.b200cnn_engine_forward_backward(self.handle, ptr, is_u8, C.c_void_p(labels.data_ptr()), batch,
                                                         1 if use_graph else 0, _stream_ptr(stream)), "forward_backward")

    @property
    def launches_per_step(self):
        return self.lib.b200cnn_engine_launches_per_step(self.handle)

    @property
    def conv_path(self):
        """'generic' (CHW kernels on the CUDA cores), 'tcgen05' (NHWC chain, reference topology) or 'tcgen05_pool'
        (the same for conv -> pool -> conv -> pool) under the current backend setting."""
        return ("generic", "tcgen05", "tcgen05_pool")[self.lib.b200cnn_engine_conv_path(self.handle)]

    # ---- results ------------------------------------------------------------------------------
    def metrics_tensor(self):
        """Device tensor view [4] = {sum CE, sum MSE, #correct, #samples} since the last reset."""
        ptr = self.lib.b200cnn_engine_metrics(self.handle)
        return _tensor_from_ptr(ptr, 4, self.device)

    def reset_metrics(self, stream=None):
        N.check(self.lib.b200cnn_engine_reset_metrics(self.handle, _stream_ptr(stream)), "reset_metrics")

    def read_metrics(self):
        """Synchronising device->host read (16 bytes) of the metrics."""
        self._metrics_host.copy_(self.metrics_tensor(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        ce, mse, correct, count = self._metrics_host.tolist()
        return {"ce_sum": ce, "mse_sum": mse, "correct": correct, "count": count}

    def layer_output(self, layer, batch):
        """Activations of `layer` from the last forward pass as [batch, c*h*w] (CHW order whatever the device layout).
        On the pooled topology's tensor-core path a training step does not store the un-pooled activations of the first
        convolution (fused with its pool): layer 1 is available after forward() only."""
        n = self.layers[layer][1] * self.layers[layer][2] * self.layers[layer][3]
        out = np.empty((batch, n), dtype=np.float32)
        torch.cuda.synchronize()
        N.check(self.lib.b200cnn_engine_layer_output(self.handle, layer, batch, out.ctypes.data_as(N.c_float_p)), "layer_output")
        return out

    def predict(self, images):
        """Class probabilities [batch, classes] as a numpy array."""
        self.forward(images)
        return self.layer_output(len(self.layers) - 1, images.shape[0])

    # ---- data parallel -------------------------------------------------------------------------
    def attach_comm(self, comm):
        """Run parameters/gradients inside ``comm``'s NVLink-mapped region and fuse the update per layer."""
        N.check(self.lib.b200cnn_engine_attach_comm(self.handle, comm.handle if comm is not None else None), "attach_comm")
        self._comm = comm


class _CudaArrayView:
    """Minimal __cuda_array_interface__ carrier so torch can wrap a raw device pointer without copying."""

    def __init__(self, ptr, nelem, typestr="<f4"):
        self.__cuda_array_interface__ = {
            "shape": (nelem,), "typestr": typestr, "data": (int(ptr), False), "version": 3, "strides": None,
        }


def _tensor_from_ptr(ptr, nelem, device, typestr="<f4"):
    with torch.cuda.device(device):
        return torch.as_tensor(_CudaArrayView(ptr, nelem, typestr), device=device)
