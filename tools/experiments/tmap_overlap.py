"""Does cuTensorMapEncodeTiled accept a dim-1 stride SMALLER than the dim-0 extent (overlapping rows)?
rows of 32 floats that start every 16 floats: row p = pixels (p, p + 1) of a 16-channel NHWC tensor."""
import ctypes as C
import torch

torch.zeros(1, device="cuda")
lib = C.CDLL("libcuda.so.1")
buf = torch.arange(16 * 64, dtype=torch.float32, device="cuda")
tm = (C.c_uint8 * 128)()
def enc(rank, dims, strides, box, swz):
    d = (C.c_uint64 * rank)(*dims); s = (C.c_uint64 * (rank - 1))(*strides); b = (C.c_uint32 * rank)(*box); e = (C.c_uint32 * rank)(*([1] * rank))
    return lib.cuTensorMapEncodeTiled(C.byref(tm), 7, rank, C.c_void_p(buf.data_ptr()), d, s, b, e, 0, swz, 2, 0)
# CU_TENSOR_MAP_DATA_TYPE_FLOAT32 = 7; swizzle: NONE 0, 32B 1, 64B 2, 128B 3, 128B_ATOM_32B 4
for swz in (0, 3, 4):
    print("swizzle", swz, "overlap  dims (32, 48) stride 64 B ->", enc(2, (32, 48), (64,), (32, 8), swz))
    print("swizzle", swz, "regular  dims (32, 24) stride 128 B ->", enc(2, (32, 24), (128,), (32, 8), swz))
    print("swizzle", swz, "4-D overlap (32, 14, 14, 4) strides (64, 1024, 14336) ->", enc(4, (32, 14, 14, 4), (64, 1024, 14336), (32, 14, 14, 1), swz))
