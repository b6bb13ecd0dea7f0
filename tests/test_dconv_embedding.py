"""The index arithmetic of csrc/dconv.cu, restated in NumPy and checked against the oracle on the CPU.

The kernel embeds the output grid of a k x k stride-1 convolution in a padded grid (Hs = Ho + kh - 1 rows of
Wp = Wo + kw - 1 columns per image, linearised over the batch), loads "slabs" of consecutive positions, and reads tap (r, s)
of every output of a tile as the SAME slab shifted by r * Wp + s positions.  This file rebuilds exactly that -- position ->
pixel decode, zero padding positions, tiles of 128 positions with their slab of 128 + (kh - 1) Wp + kw - 1 positions, the
constant shift per tap, which positions are real outputs, and the mirrored taps / swapped roles of the data gradient --
and compares the result with oracle/ops.py.  It says nothing about the CUDA code (tests/test_gpu_dconv.py does); it pins
the geometry the CUDA code implements.
"""
import numpy as np
import pytest

from oracle import ops

BM = 128


def embedded_conv(inp, wt, pad_top, pad_left, out_hw):
    """out[b, h, w, n] = sum_{r, s, c} inp[b, h - pad_top + r, w - pad_left + s, c] * wt[r, s, c, n], the dconv.cu way."""
    B, Hi, Wi, Ci = inp.shape
    kh, kw, _, N = wt.shape
    Ho, Wo = out_hw
    Hs, Wp = Ho + kh - 1, Wo + kw - 1
    positions = B * Hs * Wp
    slab_positions = BM + (kh - 1) * Wp + kw - 1
    out = np.zeros((B, Ho, Wo, N))
    written = np.zeros((B, Ho, Wo), dtype=int)

    def decode(q):
        g, ww = divmod(q, Wp)
        img, hh = divmod(g, Hs)
        return img, hh, ww

    for tile in range(-(-positions // BM)):
        q0 = tile * BM
        slab = np.zeros((slab_positions, Ci))               # what the loader warps put into shared memory
        for j in range(slab_positions):
            img, hh, ww = decode(q0 + j)
            row, col = hh - pad_top, ww - pad_left
            if img < B and 0 <= row < Hi and 0 <= col < Wi:
                slab[j] = inp[img, row, col]
        acc = np.zeros((BM, N))                              # the TMEM accumulator: 128 rows = 128 consecutive positions
        for r in range(kh):
            for s in range(kw):
                shift = r * Wp + s                           # the descriptor's start address moves by 16 bytes per position
                acc += slab[shift:shift + BM] @ wt[r, s]
        for i in range(BM):                                  # the epilogue keeps the rows that are real outputs
            img, hh, ww = decode(q0 + i)
            if img < B and hh < Ho and ww < Wo:
                out[img, hh, ww] = acc[i]
                written[img, hh, ww] += 1
    assert (written == 1).all()                              # every output produced exactly once
    return out


CASES = [
    # x shape, kernel, out channels, padding
    ((2, 6, 5, 3), (3, 3), 4, "same"),
    ((2, 7, 6, 2), (3, 3), 3, "valid"),
    ((3, 5, 4, 2), (2, 2), 2, "same"),          # even kernel: SAME pads bottom / right only
    ((1, 9, 3, 2), (3, 1), 5, "same"),          # no padded columns
    ((2, 4, 9, 1), (1, 5), 2, "valid"),
    ((1, 6, 6, 2), (5, 5), 2, "same"),
]


@pytest.mark.parametrize("xs,k,kout,padding", CASES)
def test_forward_through_shifted_slabs(xs, k, kout, padding):
    rng = np.random.default_rng(sum(xs) + k[0])
    x, w = rng.standard_normal(xs), rng.standard_normal(k + (xs[3], kout))
    want = ops.conv2d_fwd(x, w, None, (1, 1), padding)
    ho, pt, _ = ops.same_or_valid(xs[1], k[0], 1, padding)
    wo, pl, _ = ops.same_or_valid(xs[2], k[1], 1, padding)
    got = embedded_conv(x, w, pt, pl, (ho, wo))
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("xs,k,kout,padding", CASES)
def test_data_gradient_is_the_same_kernel_over_dz(xs, k, kout, padding):
    """dc_plan(product = 1): input = dz on the Ho x Wo grid, output grid H x W, padding kh - 1 - pt / kw - 1 - pl, and
    the prep kernel's B tiles hold w[taps - 1 - tap][n][k] (mirrored taps, channels and filters swapped)."""
    rng = np.random.default_rng(7 * sum(xs) + k[1])
    x, w = rng.standard_normal(xs), rng.standard_normal(k + (xs[3], kout))
    ho, pt, _ = ops.same_or_valid(xs[1], k[0], 1, padding)
    wo, pl, _ = ops.same_or_valid(xs[2], k[1], 1, padding)
    dz = rng.standard_normal((xs[0], ho, wo, kout))
    want, _, _ = ops.conv2d_bwd(x, w, dz, (1, 1), padding)
    w_mirrored = w[::-1, ::-1].transpose(0, 1, 3, 2)          # [r'][s'][k][c] = w[kh - 1 - r'][kw - 1 - s'][c][k]
    got = embedded_conv(dz, w_mirrored, k[0] - 1 - pt, k[1] - 1 - pl, (xs[1], xs[2]))
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)
