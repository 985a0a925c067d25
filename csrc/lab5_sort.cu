// lab5: large-array ascending sort (int32 / uint32 / float32 / uint8 keys).
// The reference ships only lab5/data/{int10,float10,uchar10} and no program, so
// this lab is designed here; no CUB/Thrust — every kernel is in this file.
//
// B200 design: least-significant-digit radix sort, 8-bit digits, "onesweep":
//   1. one histogram kernel reads the keys once and builds all four digit
//      histograms (privatised per warp in shared memory);
//   2. per pass ONE kernel ranks and scatters: tiles of 8192 keys take a ticket,
//      count their digits, publish the counts, rank their keys inside each warp
//      (lanes sharing a digit are found through a shared-memory mask word;
//      MATCH.ANY is ADU-bound on B200 and ballots ALU-bound, see profiles/),
//      resolve their global offsets by decoupled look-back over earlier tiles
//      (single 64-bit status|pass|value words, so no fences and one zero-fill
//      serves all four passes), and write the keys out in coalesced runs.
// Traffic: 1 read + 4 x (read + write) of the keys = 9 x 4 B per key, plus
// ~2 KB of look-back words per 32 KB tile.  uint8 keys are a counting sort
// (histogram + fill).
//
// What bounds it (profiles/ncu_prof_lab5_v5a.md): the L1/shared-memory data pipe
// moves ONE 128-byte wavefront per cycle per SM, and a warp access indexed by 32
// random digits costs 3.5-5 wavefronts (bank conflicts).  A stable pass needs
// five such accesses per 32 keys (count, mask OR, mask+position read, position
// write, key placement), an unstable one two; with loads, stores and the per-tile
// scan that is ~36 / ~24 wavefronts per 32 keys against 10.6 cycles per 32 keys of
// HBM time.  DRAM sits at 20-30%; the kernels are shared-memory-bound, like CUB's.
#include <cstdlib>
#include <type_traits>

#include "common.cuh"

namespace b200 {

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
#ifndef B200_SORT_THREADS
#define B200_SORT_THREADS 512
#endif
constexpr int kSortThreads = B200_SORT_THREADS;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kItems = 16;
constexpr int kTile = kSortThreads * kItems;  // 8192 keys
#ifndef B200_SORT_LOOK
#define B200_SORT_LOOK 8   /* 16 measured 6% slower */
#endif

// order-preserving bijection key bits -> unsigned
template <int KT>
__device__ __forceinline__ uint32_t to_ordered(uint32_t k) {
  if (KT == B200_KEY_I32) return k ^ 0x80000000u;
  if (KT == B200_KEY_F32) return (k & 0x80000000u) ? ~k : (k | 0x80000000u);
  return k;
}

// --------------------------------------------------------------- histograms
constexpr int kHistThreads = 256;
template <int KT>
__global__ void __launch_bounds__(kHistThreads)
    sort_hist_kernel(const uint32_t* __restrict__ keys, size_t n,
                     const unsigned long long* __restrict__ n_dev,
                     unsigned long long* __restrict__ hist /* [4][256] */) {
  if (n_dev) n = (size_t)*n_dev;  // length decided on the device (multi-GPU bucket)
  __shared__ uint32_t sh[kHistThreads / 32][4][kRadix];  // 32 KB, one copy per warp
  uint32_t* flat = &sh[0][0][0];
  for (int i = threadIdx.x; i < (kHistThreads / 32) * 4 * kRadix; i += kHistThreads) flat[i] = 0;
  __syncthreads();
  uint32_t(*my)[kRadix] = sh[threadIdx.x >> 5];
  auto add = [&](uint32_t k) {
    const uint32_t o = to_ordered<KT>(k);
    atomicAdd(&my[0][o & 0xff], 1u);
    atomicAdd(&my[1][(o >> 8) & 0xff], 1u);
    atomicAdd(&my[2][(o >> 16) & 0xff], 1u);
    atomicAdd(&my[3][o >> 24], 1u);
  };
  const size_t n_vec = (((uintptr_t)keys) & 15u) == 0 ? n / 4 : 0;
  const size_t stride = (size_t)gridDim.x * kHistThreads;
  // bounded trip count per block keeps the 32-bit shared counters from overflowing
  for (size_t i = (size_t)blockIdx.x * kHistThreads + threadIdx.x; i < n_vec; i += stride) {
    const uint4 v = ld_stream_u4(keys + 4 * i);
    add(v.x), add(v.y), add(v.z), add(v.w);
  }
  for (size_t i = n_vec * 4 + (size_t)blockIdx.x * kHistThreads + threadIdx.x; i < n; i += stride)
    add(keys[i]);
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * kRadix; i += kHistThreads) {
    unsigned long long s = 0;
    for (int wv = 0; wv < kHistThreads / 32; ++wv) s += sh[wv][i >> 8][i & 0xff];
    if (s) atomicAdd(hist + i, s);
  }
}

// Same histograms for large arrays, with counters laid out so that the atomics of a warp never
// share a bank: counter (digit position p, bin b, lane l) lives at word (p * 256 + b) * 32 + l,
// i.e. every lane owns one bank column.  A warp-wide shared-memory atomic indexed by 32 random
// digits costs ~5 wavefronts of the L1 data pipe (bank conflicts), which made the kernel above
// pipe-bound at 113 us for 2^26 keys (profiles/ncu_prof_lab5_v5a.md: 92% of the pipe); here it
// costs one.  128 KB of counters, one 1024-thread CTA per SM.
constexpr int kHist2Threads = 1024;
constexpr int kHist2Smem = 4 * kRadix * 32 * 4;  // 131072 bytes
template <int KT>
__global__ void __launch_bounds__(kHist2Threads, 1)
    sort_hist_lanes_kernel(const uint32_t* __restrict__ keys, size_t n,
                           const unsigned long long* __restrict__ n_dev,
                           unsigned long long* __restrict__ hist /* [4][256] */) {
  if (n_dev) n = (size_t)*n_dev;
  extern __shared__ __align__(16) uint8_t hist_smem_raw[];
  uint32_t* cnt = reinterpret_cast<uint32_t*>(hist_smem_raw);  // [4][256][32]
  const int tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < kHist2Smem / 16; i += kHist2Threads) reinterpret_cast<uint4*>(cnt)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  const uint32_t base = smem_u32(cnt) + 4u * lane;
  auto add = [&](uint32_t k) {
    const uint32_t o = to_ordered<KT>(k);
    // address = base + (p * 256 + byte p of o) * 128
    asm volatile(
        "{\n\t.reg .b32 d, a;\n\t"
        "prmt.b32 d, %0, 0, 0x4440;\n\t mad.lo.u32 a, d, 128, %1;\n\t red.shared.add.u32 [a], 1;\n\t"
        "prmt.b32 d, %0, 0, 0x4441;\n\t mad.lo.u32 a, d, 128, %1;\n\t red.shared.add.u32 [a+32768], 1;\n\t"
        "prmt.b32 d, %0, 0, 0x4442;\n\t mad.lo.u32 a, d, 128, %1;\n\t red.shared.add.u32 [a+65536], 1;\n\t"
        "prmt.b32 d, %0, 0, 0x4443;\n\t mad.lo.u32 a, d, 128, %1;\n\t red.shared.add.u32 [a+98304], 1;\n\t}" ::"r"(o),
        "r"(base)
        : "memory");
  };
  const size_t n_vec = (((uintptr_t)keys) & 15u) == 0 ? n / 4 : 0;
  const size_t stride = (size_t)gridDim.x * kHist2Threads;
  size_t i = (size_t)blockIdx.x * kHist2Threads + tid;
  for (; i + stride < n_vec; i += 2 * stride) {  // two loads in flight per thread
    const uint4 v = ld_stream_u4(keys + 4 * i), w = ld_stream_u4(keys + 4 * (i + stride));
    add(v.x), add(v.y), add(v.z), add(v.w), add(w.x), add(w.y), add(w.z), add(w.w);
  }
  for (; i < n_vec; i += stride) {
    const uint4 v = ld_stream_u4(keys + 4 * i);
    add(v.x), add(v.y), add(v.z), add(v.w);
  }
  for (size_t j = n_vec * 4 + (size_t)blockIdx.x * kHist2Threads + tid; j < n; j += stride) add(keys[j]);
  __syncthreads();
  // thread t sums the 32 lane columns of counter row t, starting at its own column so that the
  // lanes of a warp (consecutive rows) read 32 different banks
  unsigned long long sum = 0;
#pragma unroll 8
  for (int j = 0; j < 32; ++j) sum += cnt[tid * 32 + ((j + lane) & 31)];
  if (sum) atomicAdd(hist + tid, sum);
}

// exclusive scan of each 256-bin histogram, in place: hist[p][d] -> first output
// index of digit d in pass p
// `one_bin[p]` (optional) is set when a single digit value holds every key of pass p: that pass
// is then the identity permutation and the pass kernel only copies.
__global__ void sort_scan_kernel(unsigned long long* __restrict__ hist, int passes,
                                 uint32_t* __restrict__ one_bin) {
  __shared__ unsigned long long wsum[kRadix / 32];
  for (int p = 0; p < passes; ++p) {
    unsigned long long v = hist[p * kRadix + threadIdx.x];
    if (one_bin && v != 0 && hist[p * kRadix + ((threadIdx.x + 1) & (kRadix - 1))] == 0) {
      // my bin is occupied and the next one (cyclically) is not: all keys are here iff every
      // other bin is empty too, i.e. iff exactly one thread gets this far and sees the total
      unsigned long long total = 0;
      for (int d = 0; d < kRadix; ++d) total += hist[p * kRadix + d];
      if (v == total) one_bin[p] = 1u;
    }
    __syncthreads();  // the reads above come before the in-place writes below
    unsigned long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
      if ((threadIdx.x & 31) >= o) inc += t;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    unsigned long long base = 0;
    for (int wv = 0; wv < (int)(threadIdx.x >> 5); ++wv) base += wsum[wv];
    hist[p * kRadix + threadIdx.x] = base + inc - v;
    __syncthreads();
  }
}

// ------------------------------------------------------------------ onesweep
constexpr unsigned long long kStatusShift = 62, kPassShift = 58;
constexpr unsigned long long kValueMask = (1ull << kPassShift) - 1;
constexpr unsigned long long kStAggregate = 1, kStInclusive = 2;
constexpr int kOneBinFlags = 8;  // tickets[8 + pass]: pass is an identity permutation (sort_scan_kernel)
constexpr int kGroupTiles = 16;  // look-back groups: [tiles][256] entries, then [groups][256] twice

__device__ __forceinline__ unsigned long long ld_lookback(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_lookback(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ------------------------------------------------------------------ onesweep pass
// One ticketed tile per CTA:
//   * EARLY COUNT: every key first bumps its (warp, digit) counter, so the tile's digit
//     counts exist before any ranking; the digit threads turn them into tile positions and
//     publish the tile aggregate right away (successor tiles never wait for our ranking);
//   * ranking that writes each key straight to its tile position (no rank registers, no
//     second staging loop): lanes with equal digits find each other through ONE 64-bit
//     shared word per (warp, digit) - low half = lane mask of the current row (atomicOr),
//     high half = next free tile position - read with one LDS.64 and rewritten by the
//     group's highest lane with one STS.64 (which also clears the mask for the next row);
//   * the FIRST pass has no earlier order to keep: it counts on CTA-wide counters (one bank
//     column per lane: conflict-free), turns them into cursors and places every key with one
//     atomicAdd on its (digit, lane) cursor;
//   * keys kept in their ordered (unsigned-comparable) form between passes: the first
//     pass converts on load, the last one converts back on store, the two in the middle
//     extract a digit with a single byte-permute;
//   * 40 registers, no spills, three CTAs (48 warps) per SM.
// Tried on this kernel and dropped (B200, 2^26 uniform keys, device ms for the whole sort):
// 8 or 16 of a thread's 16 rows matched with ballots instead of the mask word 1.38 / 1.42
// against 1.33 (the ALU pipe fills up as the L1 pipe empties); two CTAs per SM at 64
// registers 1.48; the first-generation kernel (ranks kept in registers, keys staged in a
// second loop, per-pass to_ordered) 1.43.
// WIDE: the array has 2^32 keys or more, so output indices need 64 bits
template <bool WIDE>
struct PassSmemT {
  using Index = typename std::conditional<WIDE, unsigned long long, uint32_t>::type;
  uint32_t keys[kTile];                         // 32 KB: keys in digit order
  unsigned long long pair[kSortWarps][kRadix];  // 32 KB: {lane mask of the row, next tile position}
  Index global_off[kRadix];                     // index in `out` of (digit, tile position 0), modulo 2^bits
  uint32_t scan_tmp[kRadix / 32];
  uint32_t tile_id;
  uint32_t one_bin;  // this pass is an identity permutation (sort_scan_kernel found one digit value)
};
#ifndef B200_SORT_PASS_BLOCKS
#define B200_SORT_PASS_BLOCKS 3
#endif

// one 64-bit shared-memory access for a {mask, position} word (never cached in registers)
__device__ __forceinline__ uint2 lds_pair(const unsigned long long* p) {
  uint2 v;
  asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void sts_pair(unsigned long long* p, uint32_t lo, uint32_t hi) {
  asm volatile("st.volatile.shared.v2.u32 [%0], {%1, %2};" ::"r"(smem_u32(p)), "r"(lo), "r"(hi) : "memory");
}
// counter[digit(key)] += 1 with the digit and the address formed inside one statement, so the
// compiler cannot hoist sixteen addresses above the loop (it spilled the KEYS to do that)
template <int STRIDE, int OFFSET>  // bytes per table entry, byte offset of the counter inside it
__device__ __forceinline__ void count_digit(uint32_t key, uint32_t sel, uint32_t table_smem) {
  asm volatile(
      "{\n\t.reg .b32 d, a;\n\t"
      "prmt.b32 d, %0, 0, %1;\n\t"
      "mad.lo.u32 a, d, %3, %2;\n\t"
      "red.shared.add.u32 [a+%4], 1;\n\t}" ::"r"(key), "r"(sel), "r"(table_smem), "n"(STRIDE), "n"(OFFSET)
      : "memory");
}
// position = cursor[digit(key)]++ on a table of STRIDE-byte entries
template <int STRIDE>
__device__ __forceinline__ uint32_t take_slot(uint32_t key, uint32_t sel, uint32_t table_smem) {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .b32 d, a;\n\t"
      "prmt.b32 d, %1, 0, %2;\n\t"
      "mad.lo.u32 a, d, %4, %3;\n\t"
      "atom.shared.add.u32 %0, [a], 1;\n\t}"
      : "=r"(r)
      : "r"(key), "r"(sel), "r"(table_smem), "n"(STRIDE)
      : "memory");
  return r;
}

// Where the key with tile position `pos` lives in PassSmem::keys: the low five bits are XORed with
// the next five.  A tile of already ordered input has (about) equally many keys of each digit, so
// the keys a warp places in one step go to positions a multiple of 32 apart - one bank without
// the swizzle (sorted 2^26 keys: 1.44 ms against 1.21 for random ones).  Runs of 32 consecutive
// positions stay conflict-free, so the output loop does not care.
__device__ __forceinline__ uint32_t key_slot(uint32_t pos) { return pos ^ ((pos >> 5) & 31u); }

__device__ __forceinline__ void pr_hi_add(unsigned long long* pr, uint32_t d, uint32_t v) {
  atomicAdd(reinterpret_cast<uint32_t*>(&pr[d]) + 1, v);
}

template <int KT>
__device__ __forceinline__ uint32_t from_ordered(uint32_t o) {
  if (KT == B200_KEY_I32) return o ^ 0x80000000u;
  if (KT == B200_KEY_F32) return (o & 0x80000000u) ? (o ^ 0x80000000u) : ~o;
  return o;
}

template <int KT, bool FIRST, bool LAST, bool FULL, bool WIDE>
__device__ __forceinline__ void pass_tile(PassSmemT<WIDE>& S, const uint32_t* __restrict__ in,
                                          uint32_t* __restrict__ out, size_t tile_base,
                                          uint32_t n_valid, uint32_t tile, int pass,
                                          const unsigned long long* __restrict__ digit_base,
                                          unsigned long long* __restrict__ lookback,
                                          unsigned long long* __restrict__ group_lb) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t sel = 0x4440u | (uint32_t)pass;  // byte `pass` of the key -> bits 0..7
  const uint32_t w_off = warp * (kItems * 32) + lane;
  unsigned long long* pr = S.pair[warp];
  // first pass: CTA-wide counters, one bank column per lane - counter (digit d, lane l) is word
  // d * 32 + l of the 32 KB that the other passes use for their per-warp pairs - so a warp's
  // atomics never conflict (a warp access indexed by 32 random digits costs ~5 wavefronts)
  uint32_t* cta_cnt = reinterpret_cast<uint32_t*>(&S.pair[0][0]);

  // ---- load (warp-striped: logical order == memory order) + early count
  uint32_t key[kItems];
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const uint32_t o = w_off + i * 32;
    uint32_t k = (FULL || o < n_valid) ? ld_stream_u32(in + tile_base + o) : 0xffffffffu;
    if (FIRST) k = to_ordered<KT>(k);
    key[i] = k;
  }
  const uint32_t tbl = FIRST ? smem_u32(cta_cnt) + 4u * lane : smem_u32(pr);
  // A warp whose 512 keys share ONE digit, or one digit per half (low-entropy or sorted input:
  // the upper digits of a tile are constant over long runs) would send every atomic of every
  // row to one shared word, where they serialise (2^26 equal keys: 1.59 ms against 1.21 for
  // random ones).  Such a warp counts and ranks without atomics.  Random digits pay two votes
  // per tile for the test.
  bool warp_uniform = false;
  uint32_t d_lo = 0, d_hi = 0;  // digit of rows [0, kItems/2) and of rows [kItems/2, kItems)
  if (!FIRST && FULL) {
    d_lo = __byte_perm(key[0], 0, sel);
    d_hi = __byte_perm(key[kItems - 1], 0, sel);
    // both shuffles outside the condition: `&&` would skip the second one on some lanes
    const uint32_t lo0 = __shfl_sync(0xffffffffu, d_lo, 0), hi0 = __shfl_sync(0xffffffffu, d_hi, 0);
    if (__all_sync(0xffffffffu, (d_lo ^ lo0 | d_hi ^ hi0) == 0)) {
      uint32_t diff = 0;
#pragma unroll
      for (int i = 1; i < kItems - 1; ++i)
        diff |= __byte_perm(key[i], 0, sel) ^ (i < kItems / 2 ? d_lo : d_hi);
      warp_uniform = __all_sync(0xffffffffu, diff == 0);
    }
  }
  if (warp_uniform) {
    if (lane == 0) {
      pr_hi_add(pr, d_lo, kItems * 16u);
      pr_hi_add(pr, d_hi, kItems * 16u);
    }
  } else {
#pragma unroll
    for (int i = 0; i < kItems; ++i)
      if (FULL || (w_off + i * 32) < n_valid) {
        if (FIRST) count_digit<128, 0>(key[i], sel, tbl);  // (digit, lane) counter: 32 words per digit
        else count_digit<8, 4>(key[i], sel, tbl);          // high half of the warp's {mask, position} pair
      }
  }
  __syncthreads();

  // ---- digit threads: counts -> tile positions, publish the aggregate, scan the digits
  uint32_t count = 0, start = 0;
  const unsigned long long tag = ((unsigned long long)pass << kPassShift);
  if (tid < kRadix) {
    uint32_t* col = reinterpret_cast<uint32_t*>(&S.pair[0][tid]) + 1;
    if (FIRST) {  // my row of 32 lane columns, walked from column `lane` on: 32 different banks per warp
#pragma unroll 8
      for (int j = 0; j < 32; ++j) count += cta_cnt[tid * 32 + ((j + lane) & 31)];
    } else {
#pragma unroll
      for (int wv = 0; wv < kSortWarps; ++wv) count += col[wv * kRadix * 2];
    }
    if (!FIRST)  // successors look back at this (the first pass reserves with atomics instead)
      st_lookback(lookback + (size_t)tile * kRadix + tid, (kStAggregate << kStatusShift) | tag | count);
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) S.scan_tmp[warp] = inc;
    asm volatile("bar.sync 1, 256;" ::: "memory");  // the kRadix digit threads only
    uint32_t base = 0;
    for (int wv = 0; wv < warp; ++wv) base += S.scan_tmp[wv];
    start = base + inc - count;
    if (FIRST) {  // counts -> cursors: each lane column gets its own sub-range of the digit's run
      uint32_t run = start;
#pragma unroll 8
      for (int j = 0; j < 32; ++j) {
        uint32_t* c = &cta_cnt[tid * 32 + ((j + lane) & 31)];
        const uint32_t v = *c;
        *c = run;
        run += v;
      }
    } else {
      uint32_t run = start;
#pragma unroll
      for (int wv = 0; wv < kSortWarps; ++wv) {
        const uint32_t c = col[wv * kRadix * 2];
        col[wv * kRadix * 2] = run;  // first tile position of (warp wv, digit tid)
        run += c;
      }
    }
  }
  __syncthreads();

  // ---- rank + place
  if (FIRST) {
    // no earlier order to keep: a key takes the next free position of its digit
#pragma unroll
    for (int i = 0; i < kItems; ++i)
      if (FULL || (w_off + i * 32) < n_valid) S.keys[key_slot(take_slot<128>(key[i], sel, tbl))] = key[i];
  } else if (warp_uniform) {
    // each half is one group in memory order: positions base + row * 32 + lane
    const uint32_t base_lo = lds_pair(&pr[d_lo]).y;
    const uint32_t base_hi = d_hi == d_lo ? base_lo + kItems * 16u : lds_pair(&pr[d_hi]).y;
#pragma unroll
    for (int i = 0; i < kItems; ++i)
      S.keys[key_slot((i < kItems / 2 ? base_lo + i * 32 : base_hi + (i - kItems / 2) * 32) + lane)] = key[i];
  } else {
    const uint32_t lane_bit = 1u << lane;
    const uint32_t le_mask = 0xffffffffu >> (31 - lane);
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
      const uint32_t d = __byte_perm(key[i], 0, sel);
      const bool valid = FULL || (w_off + i * 32) < n_valid;
      if (valid) atomicOr(reinterpret_cast<uint32_t*>(&pr[d]), lane_bit);
      __syncwarp();
      uint2 v = make_uint2(0u, 0u);  // {lanes of this row with my digit, next free position}
      if (valid) v = lds_pair(&pr[d]);
      __syncwarp();
      const uint32_t peers = v.x, pos = v.y;
      // on the group's highest lane `next` is the position after the whole group
      const uint32_t next = pos + __popc(peers & le_mask);
      if (valid) {
        if ((peers >> lane) == 1u) sts_pair(&pr[d], 0u, next);
        S.keys[key_slot(next - 1)] = key[i];
      }
      __syncwarp();
    }
  }

  // ---- digit threads: where the tile's keys of each digit go in `out`
  if (FIRST) {
    // The first pass keeps no order, so the ORDER OF THE TILES inside a digit's output range
    // does not matter either: each tile simply reserves its room with one atomicAdd per digit
    // on the (scanned) histogram entry, instead of the look-back below, which accounted for
    // 41% of this kernel's instructions and most of its barrier stall (ncu, v5b).
    if (tid < kRadix) {
      using Index = typename PassSmemT<WIDE>::Index;
      unsigned long long first = 0;
      if (count) first = atomicAdd(const_cast<unsigned long long*>(digit_base) + tid, (unsigned long long)count);
      S.global_off[tid] = (Index)first - (Index)start;
    }
  } else if (tid < kRadix) {
    // decoupled look-back, two levels: entries of my own group of kGroupTiles tiles first, then
    // one entry per earlier group (the last tile of a group publishes the group's aggregate,
    // later its inclusive sum)
    unsigned long long excl = 0;
    constexpr int kLook = B200_SORT_LOOK;
    auto walk = [&](const unsigned long long* arr, int64_t from, int64_t lo) -> bool {
      bool inclusive = false;
      for (int64_t t = from; t >= lo;) {
        unsigned long long e[kLook];
#pragma unroll
        for (int j = 0; j < kLook; ++j)
          e[j] = (t - j >= lo) ? ld_lookback(arr + (size_t)(t - j) * kRadix + tid) : 0ull;
        bool done = false;
#pragma unroll
        for (int j = 0; j < kLook; ++j) {
          if (done || t < lo) continue;
          const unsigned long long st = e[j] >> kStatusShift;
          if (st == 0 || ((e[j] >> kPassShift) & 0xf) != (unsigned long long)pass) {
            done = true;
            continue;
          }
          excl += e[j] & kValueMask;
          --t;
          if (st == kStInclusive) {
            inclusive = true;
            t = lo - 1;
            done = true;
          }
        }
      }
      return inclusive;
    };
    const int64_t grp = tile / kGroupTiles, grp_begin = grp * kGroupTiles;
    const bool group_first = (int64_t)tile == grp_begin;
    const bool group_end = (tile % kGroupTiles) == kGroupTiles - 1;
    // third array: what precedes a whole group, found ONCE per group by its first tile (the
    // oldest, so the first to get here); the other fifteen wait for that word instead of
    // walking the group level themselves - that walk was most of the ~60 probes per tile
    unsigned long long* group_excl = group_lb + (size_t)((gridDim.x + kGroupTiles - 1) / kGroupTiles) * kRadix;
    if (!walk(lookback, (int64_t)tile - 1, grp_begin)) {
      if (group_end)
        st_lookback(group_lb + (size_t)grp * kRadix + tid,
                    (kStAggregate << kStatusShift) | tag | (excl + count));
      if (group_first) {
        walk(group_lb, grp - 1, 0);  // excl was 0: now the sum of all earlier groups
        st_lookback(group_excl + (size_t)grp * kRadix + tid, (kStInclusive << kStatusShift) | tag | excl);
      } else {
        unsigned long long e;
        do {
          e = ld_lookback(group_excl + (size_t)grp * kRadix + tid);
        } while ((e >> kStatusShift) == 0 || ((e >> kPassShift) & 0xf) != (unsigned long long)pass);
        excl += e & kValueMask;
      }
    }
    if (group_end)
      st_lookback(group_lb + (size_t)grp * kRadix + tid,
                  (kStInclusive << kStatusShift) | tag | (excl + count));
    st_lookback(lookback + (size_t)tile * kRadix + tid,
                (kStInclusive << kStatusShift) | tag | (excl + count));
    // modular index arithmetic (32 bits unless WIDE): (index of the digit's first key) -
    // (its tile position) may wrap below zero, the sum in the output loop wraps back
    using Index = typename PassSmemT<WIDE>::Index;
    S.global_off[tid] = (Index)(digit_base[tid] + excl) - (Index)start;
  }
  __syncthreads();

  // ---- coalesced runs out
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const uint32_t p = tid + i * kSortThreads;
    if (FULL || p < n_valid) {
      const uint32_t k = S.keys[key_slot(p)];
      const uint32_t d = __byte_perm(k, 0, sel);
      out[(typename PassSmemT<WIDE>::Index)(S.global_off[d] + p)] = LAST ? from_ordered<KT>(k) : k;
    }
  }
}

template <int KT, bool FIRST, bool LAST, bool WIDE>
__global__ void __launch_bounds__(kSortThreads, B200_SORT_PASS_BLOCKS)
    sort_pass_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, size_t n,
                     const unsigned long long* __restrict__ n_dev, int pass,
                     const unsigned long long* __restrict__ digit_base /*[256]*/,
                     unsigned long long* __restrict__ lookback /*[tiles][256] then [groups][256]*/,
                     uint32_t* __restrict__ ticket) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  PassSmemT<WIDE>& S = *reinterpret_cast<PassSmemT<WIDE>*>(smem_raw);
  const int tid = threadIdx.x;
  if (tid == 0) {  // the flag's load latency hides behind the ticket's
    const uint32_t one_bin = ticket[kOneBinFlags];
    S.tile_id = atomicAdd(ticket, 1u);
    S.one_bin = one_bin;
  }
  {  // each warp clears 2 KB of the counter words (its own pairs; a slice of the first pass's table)
    uint4* z = reinterpret_cast<uint4*>(S.pair[tid >> 5]);
#pragma unroll
    for (int j = 0; j < kRadix * 8 / 16 / 32; ++j) z[(tid & 31) + 32 * j] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  const uint32_t tile = S.tile_id;
  const size_t tile_base = (size_t)tile * kTile;
  unsigned long long* group_lb = lookback + (size_t)gridDim.x * kRadix;
  if (n_dev) n = (size_t)*n_dev;
  if (tile_base >= n) return;  // grid was sized for the capacity, not the actual length
  const uint32_t n_valid = (uint32_t)min((size_t)kTile, n - tile_base);
  if (S.one_bin) {
    // every key has the same digit in this pass (small-range or low-entropy keys): keep the order
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
      const uint32_t p = tid + i * kSortThreads;
      if (p < n_valid) {
        uint32_t k = ld_stream_u32(in + tile_base + p);
        if (FIRST) k = to_ordered<KT>(k);
        out[tile_base + p] = LAST ? from_ordered<KT>(k) : k;
      }
    }
    return;
  }
  if (n_valid == kTile)
    pass_tile<KT, FIRST, LAST, true, WIDE>(S, in, out, tile_base, n_valid, tile, pass, digit_base, lookback,
                                         group_lb);
  else
    pass_tile<KT, FIRST, LAST, false, WIDE>(S, in, out, tile_base, n_valid, tile, pass, digit_base,
                                          lookback, group_lb);
}

// ------------------------------------------------------------------ uint8 keys
__global__ void __launch_bounds__(kHistThreads)
    sort_u8_hist_kernel(const uint8_t* __restrict__ keys, size_t n,
                        unsigned long long* __restrict__ hist) {
  __shared__ uint32_t sh[kHistThreads / 32][kRadix];
  for (int i = threadIdx.x; i < (kHistThreads / 32) * kRadix; i += kHistThreads)
    (&sh[0][0])[i] = 0;
  __syncthreads();
  uint32_t* my = sh[threadIdx.x >> 5];
  const size_t n_vec = (((uintptr_t)keys) & 15u) == 0 ? n / 16 : 0;
  const size_t stride = (size_t)gridDim.x * kHistThreads;
  for (size_t i = (size_t)blockIdx.x * kHistThreads + threadIdx.x; i < n_vec; i += stride) {
    const uint4 v = ld_stream_u4(keys + 16 * i);
    const uint32_t wds[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      atomicAdd(&my[wds[k] & 0xff], 1u);
      atomicAdd(&my[(wds[k] >> 8) & 0xff], 1u);
      atomicAdd(&my[(wds[k] >> 16) & 0xff], 1u);
      atomicAdd(&my[wds[k] >> 24], 1u);
    }
  }
  for (size_t i = n_vec * 16 + (size_t)blockIdx.x * kHistThreads + threadIdx.x; i < n; i += stride)
    atomicAdd(&my[keys[i]], 1u);
  __syncthreads();
  for (int i = threadIdx.x; i < kRadix; i += kHistThreads) {
    unsigned long long s = 0;
    for (int wv = 0; wv < kHistThreads / 32; ++wv) s += sh[wv][i];
    if (s) atomicAdd(hist + i, s);
  }
}

// out[i] = the digit whose [start, end) range holds i; starts come from the scan
__global__ void __launch_bounds__(256)
    sort_u8_fill_kernel(uint8_t* __restrict__ out, size_t n,
                        const unsigned long long* __restrict__ starts) {
  __shared__ unsigned long long st[kRadix + 1];
  st[threadIdx.x] = starts[threadIdx.x];
  if (threadIdx.x == 0) st[kRadix] = n;
  __syncthreads();
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    int lo = 0, hi = kRadix;  // largest d with st[d] <= i
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (st[mid] <= i) lo = mid; else hi = mid;
    }
    out[i] = (uint8_t)lo;
  }
}

struct SortScratch {
  unsigned long long* hist;      // [4][256]
  uint32_t* tickets;             // [4] (+pad)
  unsigned long long* lookback;  // [tiles][256]
  size_t tiles;
  size_t bytes;
};

static SortScratch carve(void* scratch, size_t n) {
  SortScratch s;
  uint8_t* p = (uint8_t*)scratch;
  s.hist = (unsigned long long*)p;
  p += 4 * kRadix * sizeof(unsigned long long);
  s.tickets = (uint32_t*)p;
  p += 64;
  s.lookback = (unsigned long long*)p;
  s.tiles = (n + kTile - 1) / kTile;
  p += s.tiles * kRadix * sizeof(unsigned long long);
  p += 2 * ((s.tiles + kGroupTiles - 1) / kGroupTiles) * kRadix * sizeof(unsigned long long);  // group entries + group prefixes
  s.bytes = (size_t)(p - (uint8_t*)scratch);
  return s;
}

using PassKernel = void (*)(const uint32_t*, uint32_t*, size_t, const unsigned long long*, int,
                            const unsigned long long*, unsigned long long*, uint32_t*);

template <int KT, bool WIDE>
static int sort32_passes(const uint32_t* input, uint32_t* keys, uint32_t* tmp, size_t n,
                         const unsigned long long* n_dev, const SortScratch& sc, cudaStream_t st) {
  const PassKernel pass_first = sort_pass_kernel<KT, true, false, WIDE>;
  const PassKernel pass_mid = sort_pass_kernel<KT, false, false, WIDE>;
  const PassKernel pass_last = sort_pass_kernel<KT, false, true, WIDE>;
  constexpr int smem = (int)sizeof(PassSmemT<WIDE>);
  static PerDeviceOnce attr_once;  // one per (key type, index width): the kernels differ
  if (int rc = attr_once.run("sort", [&]() -> int {
        for (PassKernel k : {pass_first, pass_mid, pass_last}) {
          B200_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
          B200_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        }
        return 0;
      }))
    return rc;
  // pass 0 reads the caller's input (which may be a separate, untouched buffer),
  // then the passes ping-pong tmp <-> keys and the fourth one lands in `keys`
  const uint32_t* src = input;
  uint32_t* dst = tmp;
  for (int pass = 0; pass < 4; ++pass) {
    const PassKernel k = pass == 0 ? pass_first : pass == 3 ? pass_last : pass_mid;
    k<<<(unsigned)sc.tiles, kSortThreads, smem, st>>>(src, dst, n, n_dev, pass, sc.hist + pass * kRadix,
                                                     sc.lookback, sc.tickets + pass);
    B200_LAUNCH_OK();
    src = dst;
    dst = (dst == tmp) ? keys : tmp;
  }
  return 0;  // four passes: the result is in `keys`
}

template <int KT>
static int sort32(const uint32_t* input, uint32_t* keys, uint32_t* tmp, size_t n,
                  const unsigned long long* n_dev, const SortScratch& sc, cudaStream_t st) {
  const int sms = device_info().sm_count;
  {
    size_t want = (n / 4 + kHistThreads - 1) / kHistThreads + 1;
    // >= sms*8 blocks and <= 2^22 keys per block (32-bit shared counters stay exact)
    size_t min_blocks = (n >> 22) + 1;
    size_t blocks = want < (size_t)sms * 8 ? want : (size_t)sms * 8;
    if (blocks < min_blocks) blocks = min_blocks;
    if (n >= (size_t(1) << 21)) {  // enough keys to amortise 128 KB of counters per CTA
      static PerDeviceOnce hist_once;
      if (int rc = hist_once.run("sort", [&]() -> int {
            B200_CUDA_OK(cudaFuncSetAttribute(sort_hist_lanes_kernel<KT>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, kHist2Smem));
            return 0;
          }))
        return rc;
      sort_hist_lanes_kernel<KT><<<(unsigned)sms, kHist2Threads, kHist2Smem, st>>>(input, n, n_dev, sc.hist);
    } else {
      sort_hist_kernel<KT><<<(unsigned)blocks, kHistThreads, 0, st>>>(input, n, n_dev, sc.hist);
    }
    B200_LAUNCH_OK();
  }
  sort_scan_kernel<<<1, kRadix, 0, st>>>(sc.hist, 4, sc.tickets + kOneBinFlags);
  B200_LAUNCH_OK();
  // n is the capacity (the upper bound on the length): below 2^32 every output index fits 32 bits
  if (n < (size_t(1) << 32)) return sort32_passes<KT, false>(input, keys, tmp, n, n_dev, sc, st);
  return sort32_passes<KT, true>(input, keys, tmp, n, n_dev, sc, st);
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_sort_scratch_bytes(size_t n, int key_type) {
  (void)key_type;
  return carve(nullptr, n).bytes;
}

// n is an upper bound (grid, scratch); when n_dev is non-null the actual length is
// read from it on the device, so a producer kernel can decide it without a host sync.
// `input` (optional, 32-bit keys only) is read instead of `keys` by the first pass
// and left untouched, so a caller can sort out of place without a copy.
extern "C" int b200_sort_ex(const void* input, void* keys, void* tmp, size_t n,
                            const unsigned long long* n_dev, int key_type, void* scratch,
                            size_t scratch_bytes, b200_stream_t s);

extern "C" int b200_sort(void* keys, void* tmp, size_t n, int key_type, void* scratch,
                         size_t scratch_bytes, b200_stream_t s) {
  return b200_sort_ex(nullptr, keys, tmp, n, nullptr, key_type, scratch, scratch_bytes, s);
}

extern "C" int b200_sort_devn(void* keys, void* tmp, size_t n, const unsigned long long* n_dev,
                              int key_type, void* scratch, size_t scratch_bytes, b200_stream_t s) {
  return b200_sort_ex(nullptr, keys, tmp, n, n_dev, key_type, scratch, scratch_bytes, s);
}

extern "C" int b200_sort_ex(const void* input, void* keys, void* tmp, size_t n,
                            const unsigned long long* n_dev, int key_type, void* scratch,
                            size_t scratch_bytes, b200_stream_t s) {
  if (n == 0) return 0;
  if (!input) input = keys;
  cudaStream_t st = as_stream(s);
  SortScratch sc = carve(scratch, n);
  B200_REQUIRE(scratch && scratch_bytes >= sc.bytes, "sort: scratch too small (%zu < %zu bytes)",
               scratch_bytes, sc.bytes);
  B200_REQUIRE((((uintptr_t)scratch) & 15u) == 0, "sort: scratch must be 16-byte aligned");
  B200_REQUIRE(n < (size_t(1) << 40), "sort: n too large");
  B200_CUDA_OK(cudaMemsetAsync(scratch, 0, sc.bytes, st));
  const int sms = device_info().sm_count;
  switch (key_type) {
    case B200_KEY_I32:
      return sort32<B200_KEY_I32>((const uint32_t*)input, (uint32_t*)keys, (uint32_t*)tmp, n, n_dev, sc, st);
    case B200_KEY_U32:
      return sort32<B200_KEY_U32>((const uint32_t*)input, (uint32_t*)keys, (uint32_t*)tmp, n, n_dev, sc, st);
    case B200_KEY_F32:
      return sort32<B200_KEY_F32>((const uint32_t*)input, (uint32_t*)keys, (uint32_t*)tmp, n, n_dev, sc, st);
    case B200_KEY_U8: {
      B200_REQUIRE(n_dev == nullptr && input == keys,
                   "sort: device-side length / separate input are for 32-bit keys only");
      size_t want = (n / 16 + kHistThreads - 1) / kHistThreads + 1;
      size_t min_blocks = (n >> 24) + 1;
      size_t blocks = want < (size_t)sms * 8 ? want : (size_t)sms * 8;
      if (blocks < min_blocks) blocks = min_blocks;
      sort_u8_hist_kernel<<<(unsigned)blocks, kHistThreads, 0, st>>>((const uint8_t*)keys, n,
                                                                     sc.hist);
      B200_LAUNCH_OK();
      sort_scan_kernel<<<1, kRadix, 0, st>>>(sc.hist, 1, nullptr);
      B200_LAUNCH_OK();
      size_t fb = (n + 255) / 256;
      if (fb > (size_t)sms * 16) fb = (size_t)sms * 16;
      sort_u8_fill_kernel<<<(unsigned)fb, 256, 0, st>>>((uint8_t*)keys, n, sc.hist);
      B200_LAUNCH_OK();
      return 0;
    }
    default:
      B200_REQUIRE(false, "sort: unknown key type %d", key_type);
  }
  return 0;
}

// ------------------------------------------------------------- no-allocation sort
// b200_sort_noalloc: the radix sort above for callers that must not allocate.
//
// b200_sort needs three pieces of device memory besides the keys: the digit
// histograms, the tile tickets + look-back words, and a second key buffer to
// ping-pong with.  Its callers obtain them with cudaMalloc (or from torch).  Some
// callers cannot: a plug-in running inside a host application that has already
// handed all device memory to its own pool, where a stray cudaMalloc fails or
// starves the host; a latency-critical loop that cannot afford cudaMalloc /
// cudaFree because both synchronise the whole device and stall every other
// stream; code under a "no allocation after start-up" rule.  For them this entry
// point keeps all of that memory in ONE file-scope device array that is part of
// the library image: it exists on every GPU that loads the library, from load to
// process exit, and costs B200_SORT_NOALLOC_SCRATCH_BYTES (about 4.3 MiB) there.
//
//   * at most B200_SORT_NOALLOC_MAX_KEYS (2^20) keys per call; larger n is
//     rejected with an error, nothing is launched;
//   * `input` is only read, `output` receives the sorted keys (they may be the
//     same buffer); the caller owns both;
//   * the first call on each device looks up the array's device address with
//     cudaGetSymbolAddress (the name is only a host-side handle in host code) and
//     caches it; make that first call before starting any stream capture;
//   * the array is shared by every call on that GPU and still holds the previous
//     call's words, so each call zero-fills the part it relies on (histograms,
//     tickets, look-back) on its stream before launching;
//   * because the array is shared, two calls whose kernels can overlap on the
//     same GPU (different streams, or different host threads) race on it: run
//     them on one stream or order them with events.
// Calls nothing that allocates: no cudaMalloc/cudaFree, no torch, no library
// routine with a workspace.  (No counterpart in the reference.)
#include <mutex>

namespace b200 {

constexpr size_t kNoAllocMaxKeys = B200_SORT_NOALLOC_MAX_KEYS;
constexpr size_t kNoAllocBookBytes =
    (4 * kRadix * sizeof(unsigned long long) + 64 +
     ((kNoAllocMaxKeys + kTile - 1) / kTile) * kRadix * sizeof(unsigned long long) +
     2 * (((kNoAllocMaxKeys + kTile - 1) / kTile + kGroupTiles - 1) / kGroupTiles) * kRadix *
         sizeof(unsigned long long) +
     255) / 256 * 256;
constexpr size_t kNoAllocBytes = kNoAllocBookBytes + kNoAllocMaxKeys * sizeof(uint32_t);
static_assert(kNoAllocBytes == B200_SORT_NOALLOC_SCRATCH_BYTES || B200_SORT_THREADS != 512,
              "header constant out of date");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// Device address of g_noalloc_scratch on the CURRENT device, resolved once per device.
static int noalloc_scratch(unsigned char** out) {
  static std::mutex mu;
  static unsigned char* cached[64] = {nullptr};
  int dev = -1;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "sort_noalloc: unexpected device ordinal %d", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[dev]) {
    void* p = nullptr;
    B200_CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    B200_REQUIRE(p != nullptr, "sort_noalloc: scratch symbol has no device address");
    cached[dev] = (unsigned char*)p;
  }
  *out = cached[dev];
  return 0;
}

}  // namespace b200

extern "C" int b200_sort_noalloc(const void* input, void* output, size_t n, int key_type,
                                 b200_stream_t s) {
  B200_REQUIRE(n <= kNoAllocMaxKeys,
               "sort_noalloc: n = %zu exceeds the built-in scratch (at most %zu keys); use b200_sort "
               "with caller-provided scratch for larger arrays",
               n, (size_t)kNoAllocMaxKeys);
  B200_REQUIRE(key_type == B200_KEY_I32 || key_type == B200_KEY_U32 || key_type == B200_KEY_F32 ||
                   key_type == B200_KEY_U8,
               "sort_noalloc: unknown key type %d", key_type);
  if (n == 0) return 0;
  B200_REQUIRE(input && output, "sort_noalloc: null key buffer");
  unsigned char* base = nullptr;
  if (int rc = noalloc_scratch(&base)) return rc;
  cudaStream_t st = as_stream(s);
  SortScratch sc = carve(base, n);  // histograms | tickets | look-back, then the key buffer
  uint32_t* tmp = (uint32_t*)(base + kNoAllocBookBytes);
  // the array still holds the previous call's counters and look-back states
  B200_CUDA_OK(cudaMemsetAsync(base, 0, sc.bytes, st));
  switch (key_type) {
    case B200_KEY_I32:
      return sort32<B200_KEY_I32>((const uint32_t*)input, (uint32_t*)output, tmp, n, nullptr, sc, st);
    case B200_KEY_U32:
      return sort32<B200_KEY_U32>((const uint32_t*)input, (uint32_t*)output, tmp, n, nullptr, sc, st);
    case B200_KEY_F32:
      return sort32<B200_KEY_F32>((const uint32_t*)input, (uint32_t*)output, tmp, n, nullptr, sc, st);
    default: {  // B200_KEY_U8: histogram of the input, then the output is filled run by run
      const int sms = device_info().sm_count;
      size_t blocks = (n / 16 + kHistThreads - 1) / kHistThreads + 1;
      if (blocks > (size_t)sms * 8) blocks = (size_t)sms * 8;
      sort_u8_hist_kernel<<<(unsigned)blocks, kHistThreads, 0, st>>>((const uint8_t*)input, n, sc.hist);
      B200_LAUNCH_OK();
      sort_scan_kernel<<<1, kRadix, 0, st>>>(sc.hist, 1, nullptr);
      B200_LAUNCH_OK();
      size_t fb = (n + 255) / 256;
      if (fb > (size_t)sms * 16) fb = (size_t)sms * 16;
      sort_u8_fill_kernel<<<(unsigned)fb, 256, 0, st>>>((uint8_t*)output, n, sc.hist);
      B200_LAUNCH_OK();
      return 0;
    }
  }
}

extern "C" int b200_preload_lab5(int key_type) {
  auto passes = [](auto kt) -> int {
    constexpr int KT = decltype(kt)::value;
    if (int rc = preload_kernel(sort_hist_kernel<KT>)) return rc;
    if (int rc = preload_kernel(sort_hist_lanes_kernel<KT>)) return rc;
    if (int rc = preload_kernel(sort_pass_kernel<KT, true, false, false>)) return rc;
    if (int rc = preload_kernel(sort_pass_kernel<KT, false, false, false>)) return rc;
    return preload_kernel(sort_pass_kernel<KT, false, true, false>);
  };
  if (int rc = preload_kernel(sort_scan_kernel)) return rc;
  switch (key_type) {
    case B200_KEY_I32: return passes(std::integral_constant<int, B200_KEY_I32>());
    case B200_KEY_U32: return passes(std::integral_constant<int, B200_KEY_U32>());
    case B200_KEY_F32: return passes(std::integral_constant<int, B200_KEY_F32>());
    case B200_KEY_U8:
      if (int rc = preload_kernel(sort_u8_hist_kernel)) return rc;
      return preload_kernel(sort_u8_fill_kernel);
    default: B200_REQUIRE(false, "preload: unknown key type %d", key_type);
  }
  return 0;
}
