"""ctypes bindings of the two native libraries built in-tree by ``make`` / ``__graft_entry__.build``.

``cpu``  -> lib/libb200cnn_cpu.so   (include/b200cnn_cpu.h)
``cuda`` -> lib/libb200cnn_cuda.so  (include/b200cnn.h, include/b200cnn_comm.h)

The CUDA library is loaded lazily and LOUDLY: on a machine with a GPU a missing or unloadable
library is an error, never a silent fallback to PyTorch ops.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_DIR = os.path.join(_HERE, "lib")

c_float_p = C.POINTER(C.c_float)
c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)
c_u8_p = C.POINTER(C.c_uint8)
c_u32_p = C.POINTER(C.c_uint32)
c_long_p = C.POINTER(C.c_long)
vp = C.c_void_p


class ConvShape(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("n", "c", "h", "w", "oc", "oh", "ow", "k", "pad", "stride")]


class LayerDesc(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("type", "depth", "width", "height", "k", "pad", "stride")]


LAYER_INPUT, LAYER_FULL, LAYER_CONV, LAYER_POOL = 0, 1, 2, 3
ACT_NONE, ACT_TANH, ACT_RELU = 0, 1, 2
BACKEND_AUTO, BACKEND_SIMT, BACKEND_TCGEN05 = 0, 1, 2
ALGO_AUTO, ALGO_ONE_SHOT, ALGO_TWO_SHOT, ALGO_NVLS, ALGO_LL_ONE_SHOT, ALGO_LL_TWO_SHOT = 0, 1, 2, 3, 4, 5


def build(targets=("cpu", "cuda"), quiet=True):
    """Run ``make`` for the native libraries (nvcc cross-compiles sm_100a without a GPU)."""
    cmd = ["make", "-C", _ROOT, "-j", str(os.cpu_count() or 4)] + list(targets)
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("native build failed:\n" + res.stdout[-8000:])
    if not quiet:
        print(res.stdout[-2000:])


def _sig(lib, name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_cpu = None
_cuda = None


def cpu():
    global _cpu
    if _cpu is not None:
        return _cpu
    path = os.path.join(LIB_DIR, "libb200cnn_cpu.so")
    if not os.path.exists(path):
        build(("cpu",))
    lib = C.CDLL(path)
    i, d, sz = C.c_int, C.c_double, C.c_size_t
    _sig(lib, "b200cnn_cpu_net_create", vp, C.c_uint)
    _sig(lib, "b200cnn_cpu_net_destroy", None, vp)
    _sig(lib, "b200cnn_cpu_net_seed", None, vp, C.c_uint)
    _sig(lib, "b200cnn_cpu_net_add_input", i, vp, i, i, i)
    _sig(lib, "b200cnn_cpu_net_add_conv", i, vp, i, i, i, i, i, i, d)
    _sig(lib, "b200cnn_cpu_net_add_full", i, vp, i, d)
    _sig(lib, "b200cnn_cpu_net_add_pool", i, vp, i, i)
    _sig(lib, "b200cnn_cpu_net_num_layers", i, vp)
    _sig(lib, "b200cnn_cpu_net_layer_info", i, vp, i, c_int_p)
    _sig(lib, "b200cnn_cpu_net_num_params", sz, vp)
    _sig(lib, "b200cnn_cpu_net_set_inputs", None, vp, c_double_p)
    _sig(lib, "b200cnn_cpu_net_get_outputs", None, vp, c_double_p)
    _sig(lib, "b200cnn_cpu_net_learn_outputs", None, vp, c_double_p)
    _sig(lib, "b200cnn_cpu_net_error_total", d, vp)
    _sig(lib, "b200cnn_cpu_net_batch_supported", i, vp)
    _sig(lib, "b200cnn_cpu_net_learn_batch", None, vp, c_double_p, c_double_p, i, c_double_p)
    _sig(lib, "b200cnn_cpu_net_forward_batch", None, vp, c_double_p, i, c_double_p)
    _sig(lib, "b200cnn_cpu_net_update", None, vp, d)
    _sig(lib, "b200cnn_cpu_net_get_params", None, vp, c_double_p)
    _sig(lib, "b200cnn_cpu_net_set_params", None, vp, c_double_p)
    _sig(lib, "b200cnn_cpu_net_get_updates", None, vp, c_double_p)
    _sig(lib, "b200cnn_cpu_net_layer_array", i, vp, i, i, c_double_p)
    _sig(lib, "b200cnn_cpu_net_save", i, vp, C.c_char_p)
    _sig(lib, "b200cnn_cpu_net_load", i, vp, C.c_char_p)
    _sig(lib, "b200cnn_glibc_rand", None, C.c_uint, i, c_int_p)
    _sig(lib, "b200cnn_cpu_net_rand", i, vp)
    _sig(lib, "b200cnn_make_synthetic", None, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint64, c_u8_p, c_u8_p)
    _sig(lib, "b200cnn_idx_write", i, C.c_char_p, i, c_u32_p, c_u8_p)
    _sig(lib, "b200cnn_idx_read", i, C.c_char_p, c_int_p, c_u32_p, c_u8_p, sz)
    _cpu = lib
    return lib


def cuda_lib_path():
    return os.path.join(LIB_DIR, "libb200cnn_cuda.so")


def cuda():
    """The sm_100a library.  Raises if it is missing — there is no PyTorch fallback."""
    global _cuda
    if _cuda is not None:
        return _cuda
    path = cuda_lib_path()
    if not os.path.exists(path):
        build(("cuda",))
    try:
        lib = C.CDLL(path)
    except OSError as exc:  # pragma: no cover - depends on the box
        raise RuntimeError(f"cannot load {path}: {exc}; the GPU path has no fallback") from exc
    i, f, sz, u64, u32 = C.c_int, C.c_float, C.c_size_t, C.c_uint64, C.c_uint32
    cs = C.POINTER(ConvShape)
    _sig(lib, "b200cnn_last_error", C.c_char_p)
    _sig(lib, "b200cnn_build_info", C.c_char_p)
    _sig(lib, "b200cnn_set_backend", i, i)
    _sig(lib, "b200cnn_get_backend", i)
    _sig(lib, "b200cnn_conv2d_fwd", i, vp, i, vp, vp, vp, cs, i, vp)
    _sig(lib, "b200cnn_conv2d_bwd_data", i, vp, vp, vp, i, vp, cs, vp)
    _sig(lib, "b200cnn_conv2d_bwd_weight_scratch_bytes", sz, cs)
    _sig(lib, "b200cnn_conv2d_bwd_weight", i, vp, vp, i, vp, vp, i, vp, cs, vp)
    _sig(lib, "b200cnn_conv2d_bwd_weight_noalloc_capacity", sz)
    _sig(lib, "b200cnn_conv2d_bwd_weight_noalloc", i, vp, vp, i, vp, vp, i, cs, vp)
    _sig(lib, "b200cnn_conv3x3_fwd_nhwc", i, vp, vp, vp, vp, i, cs, i, vp)
    _sig(lib, "b200cnn_conv3x3s2_bwd_data_nhwc", i, vp, i, vp, vp, i, vp, cs, vp)
    _sig(lib, "b200cnn_conv3x3_fwd_nhwc_to_nhwc", i, vp, vp, vp, vp, cs, i, vp)
    _sig(lib, "b200cnn_conv3x3s1_bwd_data_nhwc", i, vp, vp, vp, i, vp, cs, vp)
    _sig(lib, "b200cnn_maxpool_nhwc_fwd", i, vp, vp, vp, i, i, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_maxpool_nhwc_bwd", i, vp, i, vp, vp, i, vp, i, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes", sz, cs)
    _sig(lib, "b200cnn_conv12_fwd", i, vp, vp, vp, vp, vp, vp, vp, i, cs, cs, vp)
    _sig(lib, "b200cnn_conv1_pool_fwd", i, vp, i, vp, vp, vp, vp, vp, cs, i, vp)
    _sig(lib, "b200cnn_conv1_pool_bwd_weight", i, vp, i, vp, vp, vp, i, vp, vp, i, vp, cs, vp)
    _sig(lib, "b200cnn_conv2d_bwd_weight_nhwc", i, vp, i, vp, i, vp, vp, i, vp, cs, vp)
    _sig(lib, "b200cnn_linear_fwd", i, vp, vp, vp, vp, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_linear_bwd_data", i, vp, vp, vp, i, vp, i, i, i, i, vp)
    _sig(lib, "b200cnn_linear_bwd_weight", i, vp, vp, vp, vp, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_softmax_xent", i, vp, vp, vp, vp, vp, i, i, i, vp)
    _sig(lib, "b200cnn_linear_softmax_xent", i, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, vp)
    _sig(lib, "b200cnn_mlp_head", i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_fc_chain", i, vp, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_fc_mid", i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_maxpool_fwd", i, vp, vp, vp, i, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_maxpool_bwd", i, vp, vp, vp, i, vp, i, i, i, i, i, i, vp)
    _sig(lib, "b200cnn_sgd_update", i, vp, vp, sz, f, i, vp)
    _sig(lib, "b200cnn_synthetic_batch", i, vp, vp, i, i, i, i, u64, u64, vp)
    _sig(lib, "b200cnn_gather_batch", i, vp, vp, u32, i, vp, vp, i, u64, u64, u32, u32, vp)
    # engine
    _sig(lib, "b200cnn_engine_create", vp, C.POINTER(LayerDesc), i, i)
    _sig(lib, "b200cnn_engine_destroy", None, vp)
    _sig(lib, "b200cnn_engine_num_params", sz, vp)
    _sig(lib, "b200cnn_engine_flat_floats", sz, vp)
    _sig(lib, "b200cnn_engine_param_layout", None, vp, c_long_p, c_long_p)
    _sig(lib, "b200cnn_engine_bind_buffers", i, vp, vp, vp)
    _sig(lib, "b200cnn_engine_params", vp, vp)
    _sig(lib, "b200cnn_engine_grads", vp, vp)
    _sig(lib, "b200cnn_engine_set_params", i, vp, c_float_p)
    _sig(lib, "b200cnn_engine_get_params", i, vp, c_float_p)
    _sig(lib, "b200cnn_engine_get_grads", i, vp, c_float_p)
    _sig(lib, "b200cnn_engine_forward", i, vp, vp, i, vp, i, vp)
    _sig(lib, "b200cnn_engine_backward", i, vp, vp)
    _sig(lib, "b200cnn_engine_sgd", i, vp, f, vp)
    _sig(lib, "b200cnn_engine_train_step", i, vp, vp, i, vp, i, f, i, vp)
    _sig(lib, "b200cnn_engine_train_step_host", i, vp, vp, i, vp, i, f, vp, vp)
    _sig(lib, "b200cnn_engine_forward_backward", i, vp, vp, i, vp, i, i, vp)
    _sig(lib, "b200cnn_engine_metrics", vp, vp)
    _sig(lib, "b200cnn_engine_reset_metrics", i, vp, vp)
    _sig(lib, "b200cnn_engine_layer_output", i, vp, i, i, c_float_p)
    _sig(lib, "b200cnn_engine_launches_per_step", i, vp)
    _sig(lib, "b200cnn_engine_conv_path", i, vp)
    _sig(lib, "b200cnn_engine_attach_comm", i, vp, vp)
    # comm
    _sig(lib, "b200cnn_comm_region_bytes", sz, sz)
    _sig(lib, "b200cnn_comm_params_offset", sz, sz)
    _sig(lib, "b200cnn_comm_grads_offset", sz, sz, i)
    _sig(lib, "b200cnn_comm_create", vp, i, i, C.POINTER(vp), vp, sz)
    _sig(lib, "b200cnn_comm_destroy", None, vp)
    _sig(lib, "b200cnn_comm_rank", i, vp)
    _sig(lib, "b200cnn_comm_world", i, vp)
    _sig(lib, "b200cnn_comm_has_multicast", i, vp)
    _sig(lib, "b200cnn_comm_params", vp, vp)
    _sig(lib, "b200cnn_comm_grads", vp, vp, i)
    _sig(lib, "b200cnn_comm_set_algo", i, vp, i)
    _sig(lib, "b200cnn_comm_error", i, vp)
    _sig(lib, "b200cnn_comm_allreduce_sgd", i, vp, sz, sz, f, i, i, i, vp)
    _sig(lib, "b200cnn_comm_allreduce", i, vp, sz, sz, i, i, i, vp)
    _sig(lib, "b200cnn_comm_broadcast_params", i, vp, i, vp)
    _sig(lib, "b200cnn_comm_barrier", i, vp, vp)
    _sig(lib, "b200cnn_comm_wait_weights", i, vp, vp)
    _sig(lib, "b200cnn_comm_emulate_ll_scratch_bytes", sz, i)
    _sig(lib, "b200cnn_comm_emulate_ll", i, i, C.POINTER(vp), sz, sz, sz, f, i, i, i, i, i, vp, vp)
    _sig(lib, "b200cnn_comm_emulate_allreduce_sgd", i, i, C.POINTER(vp), sz, sz, sz, f, i, i, i, vp, vp)
    _sig(lib, "b200cnn_comm_emulate_allreduce_sgd_slot", i, i, C.POINTER(vp), sz, sz, sz, f, i, i, i, i, vp, vp)
    _sig(lib, "b200cnn_ipc_alloc", vp, sz)
    _sig(lib, "b200cnn_ipc_free", i, vp)
    _sig(lib, "b200cnn_ipc_get_handle", i, vp, C.c_char_p)
    _sig(lib, "b200cnn_ipc_open", vp, C.c_char_p)
    _sig(lib, "b200cnn_ipc_close", i, vp)
    _cuda = lib
    return lib


def check(rc, what=""):
    """Turn a non-zero return code of the CUDA library into an exception carrying its message."""
    if rc != 0:
        msg = cuda().b200cnn_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what or 'b200cnn'} failed (code {rc}): {msg}")
