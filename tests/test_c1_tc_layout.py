"""Index algebra of the opt-in tensor-core weight gradient of the first convolution (csrc/cuda/tc_conv1_wgrad.cu),
replayed in numpy on CPU: A = dz seen as [pixel pairs][(parity, oc)], B = [pixel pairs][(parity, tap | one | 0)],
D = A^T B, dW = sum of the two same-parity blocks — against autograd.  (The kernel itself still has to be run on
hardware; this pins the formulation it implements, element for element.)"""
import numpy as np
import torch
import torch.nn.functional as F


def _build_b_rows(x_u8):
    """The B tile exactly as the kernel's builder warps fill it: row (g, q), column lane = e * 16 + j."""
    n = x_u8.shape[0]
    rows = np.zeros((n * 98, 32), dtype=np.float64)
    for g in range(n):
        for q in range(98):
            for lane in range(32):
                e, j = lane >> 4, lane & 15
                v = 0.0
                if j < 9:
                    ky, kx = j // 3, j % 3
                    y, px = q // 7, 2 * (q % 7) + e
                    iy, ix = 2 * y - 1 + ky, 2 * px - 1 + kx
                    if 0 <= iy < 28 and 0 <= ix < 28:
                        v = x_u8[g, iy, ix] / 255.0
                elif j == 9:
                    v = 1.0
                rows[g * 98 + q, lane] = v
    return rows


def test_pair_row_formulation_matches_autograd():
    rng = np.random.default_rng(0)
    n = 3
    x = rng.integers(0, 256, size=(n, 28, 28), dtype=np.uint8)
    dz = rng.standard_normal((n, 16, 14, 14))
    w = torch.zeros(16, 1, 3, 3, dtype=torch.float64, requires_grad=True)
    F.conv2d(torch.from_numpy(x.astype(np.float64) / 255.0)[:, None], w, None, 2, 1).backward(torch.from_numpy(dz))
    want_w, want_b = w.grad.numpy().reshape(16, 9), dz.sum((0, 2, 3))

    dz_nhwc = np.ascontiguousarray(dz.transpose(0, 2, 3, 1))  # [n][14][14][16]
    a_rows = dz_nhwc.reshape(n * 98, 32)                      # the TMA view: two adjacent pixels per 128-byte row
    b_rows = _build_b_rows(x)
    d = a_rows.T @ b_rows                                     # D[(e, oc)][(e', j)]
    # epilogue: row (e, oc) keeps columns (e, 0..9); the two parities are added
    part = np.stack([d[e * 16:(e + 1) * 16, e * 16:e * 16 + 10] for e in (0, 1)])  # [e][oc][10]
    got = part[0] + part[1]
    assert np.allclose(got[:, :9], want_w, rtol=1e-12, atol=1e-12)
    assert np.allclose(got[:, 9], want_b, rtol=1e-12, atol=1e-12)
    assert np.all(b_rows[:, 10:16] == 0) and np.all(b_rows[:, 26:32] == 0)


def test_swizzled_tile_addresses_are_a_permutation_and_conflict_free():
    """Byte offset of element (row r, column c) in the 128B-swizzle-with-32B-atoms tile used for both operands:
    r * 128 + (((c >> 3) ^ (r & 3)) << 5) + (c & 7) * 4 — a bijection onto the tile, 32 distinct banks per row."""
    offs = np.array([[r * 128 + ((((c >> 3) ^ (r & 3))) << 5) + (c & 7) * 4 for c in range(32)] for r in range(200)])
    assert sorted(offs.reshape(-1).tolist()) == list(range(0, 200 * 128, 4))
    banks = (offs // 4) % 32
    assert all(len(set(row)) == 32 for row in banks.tolist())
