"""Shared-memory layout and fast paths of one stable radix pass (csrc/lab5_sort.cu), restated in
numpy.  It checks the reasoning behind three choices - not the CUDA code, which
tests/test_gpu_kernels.py compares with torch.sort:

  * key_slot(pos) = pos ^ ((pos >> 5) & 31) is a permutation of the tile, keeps every run of
    32 consecutive positions on 32 different banks (the output loop), and spreads positions a
    multiple of 32 apart (what a warp writes in one step when the input is already ordered);
  * a warp whose 512 keys share one digit per half can take its tile positions from the
    digit's base alone and still produces the stable order;
  * a pass whose digit is the same for every key is the identity permutation, and the scan
    kernel's test for it (my bin occupied, the cyclically next one empty, my count == total)
    fires for exactly those histograms.
"""
import numpy as np

TILE, WARPS, ROWS = 8192, 16, 16


def key_slot(pos):
    return pos ^ ((pos >> 5) & 31)


def test_key_slot_is_a_bank_friendly_permutation():
    pos = np.arange(TILE)
    slot = key_slot(pos)
    assert np.array_equal(np.sort(slot), pos)
    # the output loop: thread t of a warp reads position base + t
    for base in range(0, TILE, 32):
        assert len(set(key_slot(base + np.arange(32)) % 32)) == 32
    # ordered input: the 32 lanes of a warp place keys of 32 consecutive digits, each digit
    # holding 32 keys of the tile, at the same rank j within their digit
    for d0 in (0, 1, 77, 224):
        for j in (0, 5, 31):
            lanes = (d0 + np.arange(32)) * 32 + j
            assert len(set(lanes % 32)) == 1                 # without the swizzle: one bank
            assert len(set(key_slot(lanes) % 32)) == 32      # with it: all different


def stable_tile_positions(digits):
    """Tile position of every key of a tile in the stable digit order (the definition)."""
    order = np.argsort(digits, kind="stable")
    pos = np.empty_like(order)
    pos[order] = np.arange(len(digits))
    return pos


def test_uniform_warp_positions_match_the_stable_order():
    rng = np.random.default_rng(5)
    # memory order of a tile: warp w holds keys [w*512, (w+1)*512), row i = 32 consecutive keys
    digits = rng.integers(0, 256, TILE)
    digits[3 * 512:4 * 512] = 9                      # warp 3: one digit
    digits[7 * 512:7 * 512 + 256] = 200              # warp 7: one digit per half
    digits[7 * 512 + 256:8 * 512] = 9
    digits[12 * 512:13 * 512] = 200                  # warp 12: shares a digit with warp 7's first half
    want = stable_tile_positions(digits)
    # what the kernel does: per digit, per warp counts in warp order give the warp's base
    counts = np.zeros((WARPS, 256), dtype=np.int64)
    for w in range(WARPS):
        counts[w] = np.bincount(digits[w * 512:(w + 1) * 512], minlength=256)
    digit_start = np.concatenate([[0], np.cumsum(counts.sum(axis=0))[:-1]])
    warp_base = digit_start[None, :] + np.cumsum(counts, axis=0) - counts   # [warp][digit]
    for w in (3, 7, 12):
        d = digits[w * 512:(w + 1) * 512]
        d_lo, d_hi = d[0], d[-1]
        assert (d[:256] == d_lo).all() and (d[256:] == d_hi).all()
        base_lo = warp_base[w][d_lo]
        base_hi = base_lo + 256 if d_hi == d_lo else warp_base[w][d_hi]
        got = np.concatenate([base_lo + np.arange(256), base_hi + np.arange(256)])
        assert np.array_equal(got, want[w * 512:(w + 1) * 512]), w


def one_bin_flag(hist):
    """sort_scan_kernel's test, thread by thread."""
    total = hist.sum()
    flag = False
    for d in range(256):
        if hist[d] != 0 and hist[(d + 1) % 256] == 0 and hist[d] == total:
            flag = True
    return flag


def test_one_bin_flag_fires_exactly_for_single_valued_digits():
    rng = np.random.default_rng(6)
    for d in (0, 1, 128, 255):
        h = np.zeros(256, dtype=np.int64)
        h[d] = 12345
        assert one_bin_flag(h)
    assert not one_bin_flag(np.zeros(256, dtype=np.int64))            # no keys at all
    for _ in range(200):
        h = np.zeros(256, dtype=np.int64)
        k = int(rng.integers(2, 6))
        h[rng.choice(256, k, replace=False)] = rng.integers(1, 1000, k)
        assert not one_bin_flag(h)
    h = np.zeros(256, dtype=np.int64)
    h[255], h[0] = 5, 1                                               # cyclic neighbour occupied
    assert not one_bin_flag(h)
    # and such a pass really is the identity: a stable sort by a constant digit moves nothing
    assert np.array_equal(stable_tile_positions(np.full(TILE, 42)), np.arange(TILE))
