"""Index arithmetic of the round-2 kernels restated in numpy and checked exhaustively against its
definition (no GPU): what the CUDA code computes with masks, shifts and compile-time offsets must be
the plain formula of the reference it replaces.

* harmonic_rows.cuh: read index of fraction m/2^s for lane j, operand i of a 1024-bin tile
  = P0 + j M16 + ((i M16 + 8) >> 4)  against  (k m + 2^(s-1)) >> s  (kernels.cu:45-93: (int)(k m / 2^s + 0.5))
* fft.cuh LoadResampled: plane address (i0 & plane_mask) n1 + (i0 >> plane_shift) against the two-phase
  definition (plane 2c + parity of the transposed pair layout)
* fft.cuh StorePass1: one base + j * step against the tiled index (k & ~twm) n2 + (k & twm) + j2 tw
"""
import numpy as np

FRACTIONS = [(1, 1), (3, 2), (1, 2), (1, 3), (3, 3), (5, 3), (7, 3)] + [(m, 4) for m in range(1, 16, 2)]


def test_harmonic_rows_offsets_are_the_reference_read_index():
    W = 1024
    j = np.arange(W // 16)[:, None]
    i = np.arange(16)[None, :]
    for m, s in FRACTIONS:
        m16 = m << (4 - s)
        c = (i * m16 + 8) >> 4
        assert c.max() == (15 * m16 + 8) >> 4 <= 15
        for tile in (0, 1, 2, 3, 7, 100, 1023, 4095, 4096, 8191, 32767):
            k0 = tile * W
            k = k0 + 16 * j + i
            want = (k.astype(np.int64) * m + (1 << (s - 1))) >> s
            p0 = (k0 >> 4) * m16
            assert p0 % 4 == 0  # 16-byte aligned start of the staged range
            assert np.array_equal(p0 + j * m16 + c, want), (m, s, tile)
            # ... and it is the reference's float expression
            ref = (k.astype(np.float64) * m / (1 << s) + 0.5).astype(np.int64)
            assert np.array_equal(want, ref)
        # words reserved per range hold the last word any lane reads
        words = (((W * m16) >> 4) + 1 + 3) & ~3
        assert (W // 16 - 1) * m16 + c.max() < words


def test_phase_plane_address_is_the_two_phase_layout():
    rng = np.random.default_rng(5)
    for lg_n2, n1 in ((8, 256), (9, 512), (11, 2048), (11, 4096), (12, 4096)):
        n2 = 1 << lg_n2
        M = n1 * n2
        i0 = np.concatenate([rng.integers(0, 2 * M - 1, 20000), [0, 1, 2 * M - 2, 2 * n2 - 1, 2 * n2]]).astype(np.uint32)
        plane_mask, plane_shift = (2 << lg_n2) - 1, lg_n2 + 1
        got = (i0 & plane_mask).astype(np.int64) * n1 + (i0 >> plane_shift)
        e = i0 >> 1                       # pair that holds sample i0 (even shift) or starts one sample before the wanted pair
        col, row, par = e & (n2 - 1), e >> lg_n2, i0 & 1
        want = (2 * col + par).astype(np.int64) * n1 + row
        assert np.array_equal(got, want)
        assert got.max() < 2 * M


def test_pass1_store_base_plus_step_is_the_tiled_index():
    for n1, n2, r3, tw in ((2048, 2048, 8, 4), (4096, 2048, 16, 4), (4096, 4096, 16, 2), (1024, 512, 4, 8), (512, 512, 2, 16),
                           (256, 256, 1, 32), (256, 512, 1, 16)):
        G, F = 16 * r3, 32 // r3
        stride = 256 if r3 > 1 else 16
        R = r3 if r3 > 1 else 16
        twm = (tw - 1) & 31
        for c0 in (0, F, n2 - F):
            for f in range(F):
                j2 = c0 + f
                for k0 in list(range(0, G, 7)) + [G - 1]:
                    jj = np.arange(R)
                    k = k0 + stride * jj
                    want = j2 * tw + (k & ~twm) * n2 + (k & twm)
                    if stride & twm == 0:
                        got = (j2 * tw + (k0 & ~twm) * n2 + (k0 & twm)) + stride * n2 * jj
                        assert np.array_equal(got, want), (n1, n2, r3, tw, k0)
                        assert want.max() < 2 ** 32
                    else:
                        assert r3 == 1 and tw == 32  # the 16-point stage with 32-row tiles keeps the general form


def test_flagged_bins_are_spread_over_the_warp_once_each():
    """hp_emit_rows (search.cu): bit i of lane l's mask = bin 16 l + i; the flagged bins of a warp are enumerated as
    sixteen ballots, word i first; the n-th goes to lane n % 32 of round n / 32 (found with popc + fns).  Every
    flagged bin must be evaluated exactly once."""
    rng = np.random.default_rng(11)

    def fns(word, k):  # position of the k-th (1-based) set bit of a 32-bit word, as __fns(word, 0, k)
        for pos in range(32):
            if (word >> pos) & 1:
                k -= 1
                if k == 0:
                    return pos
        return -1

    for density in (0.0, 0.01, 0.2, 0.9, 1.0):
        masks = [int(sum((1 << i) for i in range(16) if rng.random() < density)) for _ in range(32)]
        flagged = {16 * l + i for l in range(32) for i in range(16) if (masks[l] >> i) & 1}
        bal = [sum((((masks[l] >> i) & 1) << l) for l in range(32)) for i in range(16)]
        total = sum(bin(b).count("1") for b in bal)
        assert total == len(flagged)
        seen = []
        for n0 in range(0, total, 32):
            for lane in range(32):
                n = n0 + lane
                if n >= total:
                    continue
                c, idx = 0, -1
                for i in range(16):
                    pc = bin(bal[i]).count("1")
                    if idx < 0 and n < c + pc:
                        idx = 16 * fns(bal[i], n - c + 1) + i
                    c += pc
                seen.append(idx)
        assert sorted(seen) == sorted(flagged)
