"""The destination table of the multi-GPU partition kernel (csrc/mgpu_lab5.cu), restated in
numpy: the 4096-bucket table over the splitter range plus the exact fallback for ambiguous
buckets must give, for every key, the destination the definition gives (number of splitter
composites <= the key's composite).  This checks the algorithm — edge cases of the range
mapping that random GPU inputs rarely hit: spans below 4096, duplicate splitter keys, keys at
0 and 2^32-1, one rank — not the CUDA code itself (tests/test_gpu_multirank_single_gpu.py)."""
import numpy as np
import pytest

LUT_BITS = 12
AMBIGUOUS = 0x80


def build_table(s_hi: np.ndarray, world: int):
    """s_hi[k], k = 1..world-1: ordered key part of the splitters (sorted); index 0 unused."""
    base = int(s_hi[1]) if world > 1 else 0xFFFFFFFF
    span = int(s_hi[world - 1]) - base if world > 1 else 0
    shift = 0 if span < (1 << LUT_BITS) else 32 - LUT_BITS - (32 - span.bit_length())   # 32 - bits - clz
    bkt = np.arange(1 << LUT_BITS, dtype=np.uint64)
    lo = np.uint64(base) + (bkt << np.uint64(shift))
    hi = lo + np.uint64((1 << shift) - 1)
    sp = s_hi[1:world].astype(np.uint64)
    below = (sp[None, :] < lo[:, None]).sum(axis=1)
    upto = (sp[None, :] <= hi[:, None]).sum(axis=1)
    table = (below | np.where(upto != below, AMBIGUOUS, 0)).astype(np.uint8)
    return table, base, shift


def dest_via_table(x: np.ndarray, low: np.ndarray, s_hi, s_lo, world):
    table, base, shift = build_table(s_hi, world)
    t = ((x.astype(np.int64) - base) & 0xFFFFFFFF) >> shift          # 32-bit wrap-around, like the kernel
    inside = t < (1 << LUT_BITS)
    e = table[np.where(inside, t, 0)].astype(np.int64)
    e = np.where(inside, e, world - 1)
    e = np.where(x.astype(np.int64) < base, 0, e)
    amb = (e & AMBIGUOUS) != 0
    exact = exact_dest(x, low, s_hi, s_lo, world)
    return np.where(amb, exact, e), float(amb.mean())


def exact_dest(x, low, s_hi, s_lo, world):
    comp = (x.astype(np.uint64) << np.uint64(32)) | low.astype(np.uint64)
    sp = (s_hi[1:world].astype(np.uint64) << np.uint64(32)) | s_lo[1:world].astype(np.uint64)
    return (sp[None, :] <= comp[:, None]).sum(axis=1)


def splitters_from(rng, world, kind):
    if kind == "uniform":
        hi = np.sort(rng.integers(0, 2 ** 32, world - 1, dtype=np.uint64))
    elif kind == "narrow":                      # span < 4096: shift 0
        start = int(rng.integers(0, 2 ** 32 - 5000))
        hi = np.sort(start + rng.integers(0, 3000, world - 1, dtype=np.uint64))
    elif kind == "duplicates":                  # heavy duplicates: several splitters share a key
        vals = rng.integers(0, 2 ** 32, 2, dtype=np.uint64)
        hi = np.sort(rng.choice(vals, world - 1))
    elif kind == "extremes":
        hi = np.sort(rng.choice(np.array([0, 1, 2 ** 31, 2 ** 32 - 2, 2 ** 32 - 1], dtype=np.uint64), world - 1))
    else:
        raise ValueError(kind)
    lo = rng.integers(0, 2 ** 32, world - 1, dtype=np.uint64)
    order = np.lexsort((lo, hi))                # composites sorted
    s_hi = np.concatenate([[0], hi[order]]).astype(np.uint64)
    s_lo = np.concatenate([[0], lo[order]]).astype(np.uint64)
    return s_hi, s_lo


@pytest.mark.parametrize("kind", ["uniform", "narrow", "duplicates", "extremes"])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_table_plus_fallback_equals_definition(world, kind):
    rng = np.random.default_rng(world * 100 + len(kind))
    for _ in range(20):
        s_hi, s_lo = splitters_from(rng, world, kind)
        n = 4000
        x = rng.integers(0, 2 ** 32, n, dtype=np.uint64)
        # plant keys on and next to every splitter key and at the ends of the key range
        sk = s_hi[1:world].astype(np.int64)
        special = (np.concatenate([sk, sk + 1, sk - 1, [0, 1, 2 ** 32 - 1, 2 ** 32 - 2]]) % (2 ** 32)).astype(np.uint64)
        x[: special.size] = special
        if kind == "narrow":                    # most keys inside the narrow range
            x[special.size:] = s_hi[1] + rng.integers(0, 4000, n - special.size, dtype=np.uint64)
            x &= np.uint64(0xFFFFFFFF)
        low = rng.integers(0, 2 ** 32, n, dtype=np.uint64)
        low[: world - 1] = s_lo[1:world]         # exact composite ties as well
        got, amb_frac = dest_via_table(x, low, s_hi, s_lo, world)
        want = exact_dest(x, low, s_hi, s_lo, world)
        assert np.array_equal(got, want)
        assert got.min() >= 0 and got.max() <= world - 1
        if kind == "uniform":                    # the table resolves nearly every random key by itself
            assert amb_frac < 0.02


def test_single_rank_everything_goes_to_rank_zero():
    x = np.array([0, 1, 2 ** 31, 2 ** 32 - 1], dtype=np.uint64)
    low = np.zeros(4, dtype=np.uint64)
    s = np.zeros(1, dtype=np.uint64)
    got, amb = dest_via_table(x, low, s, s, 1)
    assert np.array_equal(got, np.zeros(4, dtype=np.int64)) and amb == 0.0
