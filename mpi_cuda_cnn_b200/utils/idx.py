"""IDX (MNIST) files and the synthetic MNIST-shaped generator, through the native library.

Format per /root/reference/cnn.c:352-383: bytes 00 00 08 <ndims>, big-endian uint32 dims, uint8 payload.
"""
import ctypes as C
import os

import numpy as np

from .. import _native as N


def read_idx(path):
    lib = N.cpu()
    nd = C.c_int(0)
    dims = (C.c_uint32 * 4)()
    if lib.b200cnn_idx_read(str(path).encode(), C.byref(nd), dims, None, 0) != 0:
        raise OSError(f"cannot read IDX file {path}")
    shape = tuple(dims[i] for i in range(nd.value))
    out = np.empty(shape, dtype=np.uint8)
    if lib.b200cnn_idx_read(str(path).encode(), C.byref(nd), dims, out.ctypes.data_as(N.c_u8_p), out.size) != 0:
        raise OSError(f"cannot read IDX payload of {path}")
    return out


def write_idx(path, array):
    a = np.ascontiguousarray(array, dtype=np.uint8)
    dims = (C.c_uint32 * a.ndim)(*a.shape)
    if N.cpu().b200cnn_idx_write(str(path).encode(), a.ndim, dims, a.ctypes.data_as(N.c_u8_p)) != 0:
        raise OSError(f"cannot write IDX file {path}")


def make_synthetic(n, seed=1, h=28, w=28, nclass=10):
    """(images uint8 [n,h,w], labels uint8 [n]): class prototype + noise, learnable."""
    images = np.empty((n, h, w), dtype=np.uint8)
    labels = np.empty((n,), dtype=np.uint8)
    N.cpu().b200cnn_make_synthetic(n, h, w, nclass, seed, images.ctypes.data_as(N.c_u8_p), labels.ctypes.data_as(N.c_u8_p))
    return images, labels


def write_synthetic_dataset(directory, ntrain=1000, ntest=200):
    """Writes the four MNIST-named IDX files; returns their paths in the reference's argv order."""
    os.makedirs(directory, exist_ok=True)
    names = ["train-images-idx3-ubyte", "train-labels-idx1-ubyte", "t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"]
    paths = [os.path.join(directory, n) for n in names]
    x, y = make_synthetic(ntrain, seed=1)
    write_idx(paths[0], x)
    write_idx(paths[1], y)
    x, y = make_synthetic(ntest, seed=2)
    write_idx(paths[2], x)
    write_idx(paths[3], y)
    return paths
