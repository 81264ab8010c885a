"""Index algebra of the tensor-core convolution kernels added in round 2, replayed in numpy on CPU (the kernels themselves are
tested on hardware in tests/test_gpu_ops.py; this pins the formulations they implement, element for element):

* csrc/cuda/tc_conv_halo.cu: ONE load of a band with a one-pixel halo, padded to width 16; tap (ky, kx) is the same tile
  read from row (ky * 16 + kx) on — forward and, with mirrored taps, backward-data;
* csrc/cuda/tc_conv_wgrad.cu, kMode 1: the weight gradient over even- / odd-aligned pixel pairs with the outputs split by
  the parity of x, including the out-of-bounds pairs that stand for the zero padding and the tile rows the reduction reads;
* csrc/cuda/tc_conv12.cu: the first convolution's pixels scattered into the tap tiles of the second (stride-2) convolution.
"""
import numpy as np
import torch
import torch.nn.functional as F

H = W = 14
PW, BR = 16, 8  # padded width, image rows per band (PW * BR = 128 GEMM rows)


def _halo_tile(img_hwc, band):
    """What the TMA box (C, PW, BR + 2) started at (0, -1, band * BR - 1) delivers: zero outside the image."""
    c = img_hwc.shape[2]
    tile = np.zeros(((BR + 2) * PW + 2 * PW + 2, c))  # + the rows an out-of-tile read of a dropped column may touch
    for r in range(BR + 2):
        for q in range(PW):
            y, x = band * BR - 1 + r, q - 1
            if 0 <= y < H and 0 <= x < W:
                tile[r * PW + q] = img_hwc[y, x]
    return tile


def _halo_conv(img_hwc, w_taps, mirrored):
    """out[y][x][:] = sum_taps tile[m + shift(tap)] @ w_taps[tap]  for GEMM row m = yl * PW + x, columns x >= W dropped."""
    no = w_taps.shape[2]
    out = np.zeros((H, W, no))
    for band in range((H + BR - 1) // BR):
        tile = _halo_tile(img_hwc, band)
        acc = np.zeros((128, no))
        for t in range(9):
            ky, kx = t // 3, t % 3
            shift = (2 - ky) * PW + (2 - kx) if mirrored else ky * PW + kx
            acc += tile[shift:shift + 128] @ w_taps[t]
        for m in range(128):
            yl, x = m // PW, m % PW
            y = band * BR + yl
            if x < W and y < H:
                out[y, x] = acc[m]
    return out


def test_halo_forward_and_backward_data_match_conv2d():
    rng = np.random.default_rng(1)
    x = rng.standard_normal((16, H, W))
    w = rng.standard_normal((32, 16, 3, 3))
    want = F.conv2d(torch.from_numpy(x)[None], torch.from_numpy(w), None, 1, 1)[0].numpy()
    w_taps = np.stack([w[:, :, t // 3, t % 3].T for t in range(9)])  # [tap][c][oc]
    got = _halo_conv(np.ascontiguousarray(x.transpose(1, 2, 0)), w_taps, mirrored=False)
    assert np.abs(got.transpose(2, 0, 1) - want).max() < 1e-10

    dz = rng.standard_normal((32, H, W))
    want_dx = F.conv_transpose2d(torch.from_numpy(dz)[None], torch.from_numpy(w), None, 1, 1)[0].numpy()
    wb_taps = np.stack([w[:, :, t // 3, t % 3] for t in range(9)])  # [tap][oc][c]
    got_dx = _halo_conv(np.ascontiguousarray(dz.transpose(1, 2, 0)), wb_taps, mirrored=True)
    assert np.abs(got_dx.transpose(2, 0, 1) - want_dx).max() < 1e-10


def _pair_rows(x_nhwc, odd, j0, ky):
    """Rows (img, y, i) of the box (32 floats, 7 pairs, 14 rows, n) started at (0, j0, ky - 1, 0) of the even- or odd-aligned
    pair view of x [n][14][14][16]; the odd view has 6 pairs per row, everything out of bounds is zero."""
    n = x_nhwc.shape[0]
    rows = np.zeros((n * H * 7, 32))
    for g in range(n):
        for y in range(H):
            for i in range(7):
                j, yy = j0 + i, y + ky - 1
                if 0 <= yy < H and 0 <= j < (6 if odd else 7):
                    px = 2 * j + (1 if odd else 0)
                    rows[(g * H + y) * 7 + i] = x_nhwc[g, yy, px:px + 2].reshape(32)
    return rows


def test_even_odd_pair_weight_gradient_matches_autograd():
    rng = np.random.default_rng(2)
    n = 3
    x = rng.standard_normal((n, 16, H, W))
    dz = rng.standard_normal((n, 32, H, W))
    w = torch.zeros(32, 16, 3, 3, dtype=torch.float64, requires_grad=True)
    F.conv2d(torch.from_numpy(x), w, None, 1, 1).backward(torch.from_numpy(dz))
    want = w.grad.numpy()

    x_nhwc = np.ascontiguousarray(x.transpose(0, 2, 3, 1))
    dz_nhwc = np.ascontiguousarray(dz.transpose(0, 2, 3, 1))
    got = np.zeros((32, 16, 3, 3))
    for ky in range(3):
        for parity in range(2):
            # B: dz through the 5-D view (32 oc, parity, 7, 14, n): rows (img, y, i) of the pixels x = 2 i + parity
            b_rows = dz_nhwc[:, :, parity::2, :].reshape(n * H * 7, 32)
            if parity == 0:  # x = 2i: [kx0 | kx1'] = odd pair i - 1, [kx1 | kx2] = even pair i
                a0, a1 = _pair_rows(x_nhwc, True, -1, ky), _pair_rows(x_nhwc, False, 0, ky)
                row_of = {0: 0, 1: 32, 2: 48}
            else:            # x = 2i + 1: [kx0 | kx1] = even pair i, [kx1' | kx2] = odd pair i
                a0, a1 = _pair_rows(x_nhwc, False, 0, ky), _pair_rows(x_nhwc, True, 0, ky)
                row_of = {0: 0, 1: 16, 2: 48}
            tile = np.concatenate([a0, a1], axis=1).T @ b_rows  # D[64 rows (M)][32 oc]
            for kx in range(3):
                got[:, :, ky, kx] += tile[row_of[kx]:row_of[kx] + 16].T
    assert np.abs(got - want).max() < 1e-9


def test_conv12_scatter_into_tap_tiles_matches_conv2d():
    """csrc/cuda/tc_conv12.cu: every pixel (y, x) of the first convolution's output is written into row (j, i) of the tap
    tile (ky, kx) wherever y = 2j - 1 + ky and x = 2i - 1 + kx (1, 2 or 4 tiles); rows whose source lies in the zero
    padding are cleared beforehand.  The nine tile GEMMs must add up to the stride-2 convolution."""
    rng = np.random.default_rng(3)
    a1 = rng.standard_normal((16, 14, 14))
    w2 = rng.standard_normal((32, 16, 3, 3))
    want = F.conv2d(torch.from_numpy(a1)[None], torch.from_numpy(w2), None, 2, 1)[0].numpy()
    tiles = np.full((9, 49, 16), np.nan)  # NaN = never written
    for tap in range(9):
        for row in range(49):
            j, i = row // 7, row % 7
            if 2 * j - 1 + tap // 3 < 0 or 2 * i - 1 + tap % 3 < 0:
                tiles[tap, row] = 0.0
    writes = np.zeros((14, 14), dtype=int)
    for y in range(14):
        for x in range(14):
            for ky in range(3):
                jy = y + 1 - ky
                if jy % 2 or jy < 0 or jy >= 14:
                    continue
                for kx in range(3):
                    ix = x + 1 - kx
                    if ix % 2 or ix < 0 or ix >= 14:
                        continue
                    row = (jy // 2) * 7 + ix // 2
                    assert np.isnan(tiles[ky * 3 + kx, row]).all(), "a tile row is written once"
                    tiles[ky * 3 + kx, row] = a1[:, y, x]
                    writes[y, x] += 1
    assert not np.isnan(tiles).any(), "every row of every tap tile is either padding or some pixel"
    assert set(np.unique(writes)) <= {1, 2, 4}
    got = np.zeros((49, 32))
    for tap in range(9):
        got += tiles[tap] @ w2[:, :, tap // 3, tap % 3].T
    assert np.abs(got.T.reshape(32, 7, 7) - want).max() < 1e-10
