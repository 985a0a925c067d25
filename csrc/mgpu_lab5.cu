// lab5, multi-GPU: sample sort whose key exchange is fused into the partitioning
// kernel — every CTA classifies a tile of keys against the splitters, reserves
// room in each destination GPU's bucket with one remote atomic per destination
// and stores its keys straight into the peers' buckets over NVLink / NVSwitch.
// No send buffers, no counts exchange, no collective call.
//
//   1. splitters_kernel (1 CTA per rank): stratified sample of the local keys,
//      pushed as unique 64-bit composites (ordered key | source rank | index) into
//      every peer's workspace, flag exchange, then each rank bitonic-sorts the
//      identical sample set and keeps every S-th composite as a splitter.  The
//      composites are unique, so equal keys are split between neighbouring ranks
//      and heavily duplicated inputs stay balanced.
//   1b. count_reserve_kernel: one read of the keys counts how many go to each rank;
//      a single remote atomic per destination reserves this rank's segment there.
//   2. partition_scatter_kernel: the fused partition + all-to-all described above
//      (tiles sub-allocate the reserved segments with local atomics);
//      its last CTA (ticket) fences at system scope, signals all peers, waits for
//      all peers and publishes the number of keys that landed in the local bucket.
//   3. local radix sort of the bucket, its length read on the device (lab5_sort.cu).
// The phase masks exist so that virtual ranks can be exercised on one GPU without
// kernels waiting on each other.  (No counterpart in the reference.)
#include <cstdlib>

#include "common.cuh"
#include "symm.cuh"

namespace b200 {

template <int KT>
__device__ __forceinline__ uint32_t to_ordered_mg(uint32_t k) {
  if (KT == B200_KEY_I32) return k ^ 0x80000000u;
  if (KT == B200_KEY_F32) return (k & 0x80000000u) ? ~k : (k | 0x80000000u);
  return k;
}

__device__ __forceinline__ unsigned long long make_composite(uint32_t ordered, int src,
                                                             unsigned long long idx, int idx_bits) {
  return ((unsigned long long)ordered << 32) | ((unsigned long long)src << idx_bits) | idx;
}

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16, x *= 0x7feb352du, x ^= x >> 15, x *= 0x846ca68bu, x ^= x >> 16;
  return x;
}

// Destination rank of a key: the number of splitters <= its composite.  The 32-bit
// ordered keys decide almost always; the (source, index) tie-break is consulted only
// when some lane of the warp holds a key equal to a splitter's key.
__device__ __forceinline__ int dest_of(uint32_t ordered, uint32_t low, const uint32_t* s_hi,
                                       const uint32_t* s_lo, int world) {
  int d = 0;
  bool tie = false;
#pragma unroll
  for (int k = 1; k < kMaxWorld; ++k) {
    if (k < world) {
      d += s_hi[k] < ordered ? 1 : 0;
      tie |= s_hi[k] == ordered;
    }
  }
  if (__any_sync(0xffffffffu, tie)) {
#pragma unroll
    for (int k = 1; k < kMaxWorld; ++k)
      if (k < world) d += (s_hi[k] == ordered && s_lo[k] <= low) ? 1 : 0;
  }
  return d;
}

constexpr int kSplitThreads = 1024;

template <int KT>
__global__ void __launch_bounds__(kSplitThreads)
    sort_splitters_kernel(const uint32_t* __restrict__ keys, unsigned long long n_local,
                          int samples_per_rank, int idx_bits, SymmPeers peers, int rank, int world,
                          uint64_t epoch, uint32_t seed, int phase_mask,
                          unsigned long long timeout_ns) {
  extern __shared__ unsigned long long sm[];  // world * samples_per_rank composites, padded to 2^k
  SymmWorkspace* me = peers.ws[rank];
  const int tid = threadIdx.x;
  const int total = world * samples_per_rank;
  int padded = 1;  // the bitonic network needs a power of two: any world size is padded with
  while (padded < total) padded <<= 1;  // composites that sort last

  if (phase_mask & 1) {
    if (tid == 0) {
      me->sort_cursor = 0;  // before anyone can learn my samples, hence before any scatter
      me->sort_recv_count = 0;
      me->sort_blocks_done = 0;
    }
    // stratified sample: one pseudo-random key out of each of S equal strata
    for (int j = tid; j < samples_per_rank; j += kSplitThreads) {
      unsigned long long comp = ~0ull;  // sorts last when the shard is empty
      if (n_local > 0) {
        const unsigned long long lo = n_local * (unsigned long long)j / samples_per_rank;
        const unsigned long long hi = n_local * (unsigned long long)(j + 1) / samples_per_rank;
        const unsigned long long idx =
            hi > lo ? lo + hash32(seed ^ (uint32_t)(j * 0x9e3779b9u)) % (hi - lo) : min(lo, n_local - 1);
        comp = make_composite(to_ordered_mg<KT>(keys[idx]), rank, idx, idx_bits);
      }
      for (int p = 0; p < world; ++p)
        peers.ws[p]->sort_samples[rank * samples_per_rank + j] = comp;  // remote store
    }
    __threadfence_system();
    __syncthreads();
    if (tid < world) {
      __threadfence_system();
      signal_peer(peers, tid, rank, epoch, kFlagSortSamples);
    }
  }
  if (phase_mask & 2) {
    if (tid < world) wait_peer(peers, rank, tid, epoch, timeout_ns, kFlagSortSamples);
    __syncthreads();
    // bitonic sort of the composites in shared memory
    for (int i = tid; i < padded; i += kSplitThreads) {
      unsigned long long v = ~0ull;
      if (i < total)
        asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(me->sort_samples + i) : "memory");
      sm[i] = v;
    }
    __syncthreads();
    for (int k = 2; k <= padded; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int i = tid; i < padded; i += kSplitThreads) {
          const int ixj = i ^ j;
          if (ixj > i) {
            const unsigned long long a = sm[i], b = sm[ixj];
            const bool up = (i & k) == 0;
            if ((a > b) == up) {
              sm[i] = b;
              sm[ixj] = a;
            }
          }
        }
        __syncthreads();
      }
    }
    if (tid < world) me->sort_splitters[tid] = tid == 0 ? 0ull : sm[tid * samples_per_rank];
  }
}

// ------------------------------------------------------- count + segment reservation
// One cheap read of the keys tells this rank how many of them go to each destination;
// the last CTA then reserves ONE contiguous segment per destination with a single
// remote atomic on that rank's bucket cursor.  Optional (B200LABS_SORT_RESERVE=1):
// measured on 2 GPUs, 2^30 keys each, the count pass costs 2.0 ms and saves 0.6 ms
// of per-tile remote atomics in the scatter, so the default path skips it; it is
// kept because it makes the bucket layout (segments in rank order) deterministic.
constexpr int kCountThreads = 256;
constexpr int kCountItems = 16;

template <int KT>
__global__ void __launch_bounds__(kCountThreads)
    sort_count_reserve_kernel(const uint32_t* __restrict__ keys, unsigned long long n_local,
                              const unsigned long long* __restrict__ splitters, int idx_bits,
                              unsigned long long capacity, SymmPeers peers, int rank, int world) {
  __shared__ uint32_t s_hi[kMaxWorld], s_lo[kMaxWorld];
  __shared__ uint32_t s_cnt[kCountThreads / 32][kMaxWorld];
  __shared__ bool is_last;
  SymmWorkspace* me = peers.ws[rank];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < kMaxWorld) {
    const unsigned long long sp = tid < world ? splitters[tid] : ~0ull;
    s_hi[tid] = (uint32_t)(sp >> 32);
    s_lo[tid] = (uint32_t)sp;
  }
  __syncthreads();
  uint32_t c[kMaxWorld];
#pragma unroll
  for (int d = 0; d < kMaxWorld; ++d) c[d] = 0;
  const uint32_t src_bits = idx_bits >= 32 ? 0u : ((uint32_t)rank << idx_bits);
  // per-thread tally in eight 8-bit fields of one 64-bit word (one shift + add per key
  // instead of eight compares), spilled into c[] before any field can reach 256
  unsigned long long packed = 0;
  int in_packed = 0;
  auto flush = [&]() {
#pragma unroll
    for (int k = 0; k < kMaxWorld; ++k) c[k] += (uint32_t)(packed >> (8 * k)) & 0xffu;
    packed = 0;
    in_packed = 0;
  };
  // 16-byte loads, two per thread in flight; trip counts are warp-uniform so the
  // vote inside dest_of always sees the whole warp
  const unsigned long long n_vec = (((uintptr_t)keys) & 15u) == 0 ? n_local / 4 : 0;
  const unsigned long long stride = (unsigned long long)gridDim.x * kCountThreads;
  const unsigned long long rounds = (n_vec + stride - 1) / stride;
  for (unsigned long long r = 0; r < rounds; r += 2) {
    const unsigned long long i0 = r * stride + (unsigned long long)blockIdx.x * kCountThreads + tid;
    const unsigned long long i1 = i0 + stride;
    uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
    const bool ok0 = i0 < n_vec, ok1 = (r + 1 < rounds) && i1 < n_vec;
    if (ok0) v0 = ld_stream_u4(keys + 4 * i0);
    if (ok1) v1 = ld_stream_u4(keys + 4 * i1);
    const uint32_t k0[4] = {v0.x, v0.y, v0.z, v0.w}, k1[4] = {v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int d0 = dest_of(to_ordered_mg<KT>(k0[j]), src_bits | (uint32_t)(4 * i0 + j), s_hi, s_lo, world);
      const int d1 = dest_of(to_ordered_mg<KT>(k1[j]), src_bits | (uint32_t)(4 * i1 + j), s_hi, s_lo, world);
      packed += (ok0 ? 1ull : 0ull) << (8 * d0);
      packed += (ok1 ? 1ull : 0ull) << (8 * d1);
    }
    in_packed += 8;
    if (in_packed > 240) flush();
  }
  flush();
  {
    const unsigned long long tail0 = n_vec * 4, left = n_local - tail0;
    const unsigned long long tr = (left + stride - 1) / stride;
    for (unsigned long long r = 0; r < tr; ++r) {
      const unsigned long long i = tail0 + r * stride + (unsigned long long)blockIdx.x * kCountThreads + tid;
      const bool ok = i < n_local;
      const uint32_t key = ok ? keys[i] : 0u;
      const int d = dest_of(to_ordered_mg<KT>(key), src_bits | (uint32_t)i, s_hi, s_lo, world);
#pragma unroll
      for (int k = 0; k < kMaxWorld; ++k) c[k] += (ok && d == k) ? 1u : 0u;
    }
  }
#pragma unroll
  for (int d = 0; d < kMaxWorld; ++d) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c[d] += __shfl_down_sync(0xffffffffu, c[d], o);
    if (lane == 0) s_cnt[warp][d] = c[d];
  }
  __syncthreads();
  if (tid < world) {
    unsigned long long t = 0;
    for (int wv = 0; wv < kCountThreads / 32; ++wv) t += s_cnt[wv][tid];
    if (t) atomicAdd(&me->sort_send_count[tid], t);
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = atomicAdd(&me->sort_count_blocks_done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (tid < world) {
    unsigned long long cnt;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(cnt) : "l"(&me->sort_send_count[tid]) : "memory");
    me->sort_send_count[tid] = 0;
    // one remote atomic per destination reserves my whole segment in its bucket
    unsigned long long base = cnt ? atomicAdd_system(&peers.ws[tid]->sort_cursor, cnt) : 0ull;
    if (base + cnt > capacity) {
      atomicOr(&me->error, (unsigned long long)kErrSortOverflow);
      base = ~0ull >> 1;  // poison: the scatter drops these keys
    }
    me->sort_local_cursor[tid] = base;
  }
  if (tid == 0) me->sort_count_blocks_done = 0;
}

// ---------------------------------------------------------- fused partition+exchange
// 512 threads x 2 CTAs per SM.  (256 x 4 was a few percent faster with all eight "ranks"
// on one GPU, but slower over real NVLink — 3.25 vs 2.9 ms at 8 GPUs: twice the remote
// reservations and half the run length per destination.)
#ifndef B200_PART_BLOCKS
#define B200_PART_BLOCKS 2
#endif
constexpr int kPartThreads = 512;
constexpr int kPartWarps = kPartThreads / 32;
constexpr int kPartItems = 16;
constexpr int kPartTile = kPartThreads * kPartItems;  // 8192 keys
constexpr int kDestLutBits = 12;
constexpr uint32_t kDestAmbiguous = 0x80u, kNoDest = 8u;

struct PartTargets {
  uint32_t* bucket[kMaxWorld];              // destination buckets (peer-mapped or local)
  unsigned long long* cursor[kMaxWorld];    // their fill cursors (remote atomics)
};

// One tile = 8192 keys, 16 per thread.  The kernel was ALU-bound when it ranked keys with
// ballots (111 Gkeys/s, profiles/ncu_prof_mgpu.md), so the per-key work is now mostly
// table look-ups:
//  * destination: the key range spanned by the splitters is cut into 4096 equal buckets; a
//    shared-memory table gives the destination of a bucket, or "ambiguous" when a splitter
//    key lies inside it — only those keys run the exact comparison against the 64-bit
//    composites;
//  * slot: the partition need not be stable (the bucket is sorted afterwards), so a thread
//    counts its own keys per destination in packed 8-bit register counters, a warp scan of
//    those counters plus a scan over the warps gives the thread's first slot per
//    destination, and a key's slot is that plus its rank among the thread's own keys
//    (private shared-memory counters instead of registers were measured slower: LSU-bound);
//  * the destination byte is staged beside the key, so the output loop does not search.
template <int KT, int BLOCKS>
__global__ void __launch_bounds__(kPartThreads, BLOCKS)
    sort_partition_scatter_kernel(const uint32_t* __restrict__ keys, unsigned long long n_local,
                                  const unsigned long long* __restrict__ splitters, int idx_bits,
                                  PartTargets tg, unsigned long long capacity, SymmPeers peers,
                                  int rank, int world, uint64_t epoch, int phase_mask,
                                  unsigned long long timeout_ns) {
  __shared__ uint32_t s_keys[kPartTile];
  __shared__ uint8_t s_dst[kPartTile];
  __shared__ uint8_t dest_lut[1 << kDestLutBits];
  // per-warp totals and run starts, two 16-bit fields per word: (d0,d1) (d2,d3) (d4,d5) (d6,d7)
  __shared__ __align__(16) uint32_t warp_tot[kPartWarps][kMaxWorld / 2];
  __shared__ __align__(16) uint16_t warp_off[kPartWarps][kMaxWorld];
  __shared__ uint32_t tile_total[kMaxWorld];
  __shared__ uint32_t* s_bias[kMaxWorld + 1];  // bucket + reserved base - tile start; null = dropped
  __shared__ uint32_t s_hi[kMaxWorld], s_lo[kMaxWorld];
  __shared__ uint32_t* s_bucket[kMaxWorld];            // dynamically indexed: keep out of
  __shared__ unsigned long long* s_cursor[kMaxWorld];  // param space (no local-memory copy)
  __shared__ bool is_last;
  SymmWorkspace* me = peers.ws[rank];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (phase_mask & 1) {
    if (tid < kMaxWorld) {
      const unsigned long long sp = tid < world ? splitters[tid] : ~0ull;
      s_hi[tid] = (uint32_t)(sp >> 32);
      s_lo[tid] = (uint32_t)sp;
    }
#pragma unroll
    for (int k = 0; k < kMaxWorld; ++k)
      if (tid == k) {
        s_bucket[k] = tg.bucket[k];
        s_cursor[k] = tg.cursor[k];
      }
    if (tid == 0) s_bias[kMaxWorld] = nullptr;
    __syncthreads();
    // ---- destination table over [first splitter key, last splitter key]
    const uint32_t lut_base = world > 1 ? s_hi[1] : 0xffffffffu;
    const uint32_t lut_span = world > 1 ? s_hi[world - 1] - lut_base : 0u;
    const int lut_shift = lut_span < (1u << kDestLutBits) ? 0 : 32 - kDestLutBits - __clz(lut_span);
    for (int bkt = tid; bkt < (1 << kDestLutBits); bkt += kPartThreads) {
      const unsigned long long lo = (unsigned long long)lut_base + ((unsigned long long)bkt << lut_shift);
      const unsigned long long hi = lo + ((1ull << lut_shift) - 1);
      uint32_t below = 0, upto = 0;
#pragma unroll
      for (int k = 1; k < kMaxWorld; ++k)
        if (k < world) {
          below += (unsigned long long)s_hi[k] < lo ? 1u : 0u;
          upto += (unsigned long long)s_hi[k] <= hi ? 1u : 0u;
        }
      dest_lut[bkt] = (uint8_t)(below | (upto != below ? kDestAmbiguous : 0u));
    }
    __syncthreads();
    const uint32_t src_bits = idx_bits >= 32 ? 0u : ((uint32_t)rank << idx_bits);
    const unsigned long long n_tiles = (n_local + kPartTile - 1) / kPartTile;
    // persistent: a CTA walks tiles blockIdx.x, +gridDim.x, ... and fences its remote
    // stores ONCE at the end (a system-scope fence per tile cost ~10 us of NVLink
    // round trip per tile and made this kernel 3x slower)
    for (unsigned long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const unsigned long long tile_base = tile * kPartTile;
      const uint32_t n_valid = (uint32_t)min((unsigned long long)kPartTile, n_local - tile_base);

      uint32_t key[kPartItems];
      const uint32_t w_off = warp * (kPartItems * 32) + lane;
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t o = w_off + i * 32;
        key[i] = o < n_valid ? ld_stream_u32(keys + tile_base + o) : 0u;
      }
      // ---- destinations from the table; meta byte per key: destination (8 = none)
      uint32_t meta[kPartItems / 4] = {};
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t o = w_off + i * 32;
        const uint32_t x = to_ordered_mg<KT>(key[i]);
        const uint32_t t = (x - lut_base) >> lut_shift;
        uint32_t e = dest_lut[t < (1u << kDestLutBits) ? t : 0u];
        e = t < (1u << kDestLutBits) ? e : (uint32_t)(world - 1);  // above every splitter
        e = x < lut_base ? 0u : e;                                  // below every splitter
        if (e & kDestAmbiguous) {
          // a splitter key lies in this key's table bucket (a fraction of a percent of random
          // keys): count the splitters whose 64-bit composite is <= mine
          const uint32_t low = src_bits | (uint32_t)(tile_base + o);
          e = 0;
#pragma unroll
          for (int k = 1; k < kMaxWorld; ++k)
            if (k < world) e += (s_hi[k] < x || (s_hi[k] == x && s_lo[k] <= low)) ? 1u : 0u;
        }
        e = o < n_valid ? e : kNoDest;
        meta[i >> 2] |= e << (8 * (i & 3));
      }
      // ---- count per destination in packed 8-bit registers (destinations 0-3 / 4-7);
      // rank among my own keys -> meta bits 4-7
      uint32_t c_lo = 0, c_hi = 0;
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t d = (meta[i >> 2] >> (8 * (i & 3))) & 0xfu;
        const uint32_t sh = (d & 3u) << 3;
        const bool hi = (d & 4u) != 0;
        const uint32_t r = ((hi ? c_hi : c_lo) >> sh) & 0xffu;
        const uint32_t inc = d < kNoDest ? (1u << sh) : 0u;
        c_lo += hi ? 0u : inc;
        c_hi += hi ? inc : 0u;
        meta[i >> 2] |= (d < kNoDest ? r : 0u) << (8 * (i & 3) + 4);
      }
      // ---- warp scan of the counters, widened to 16 bits
      const uint2 c = make_uint2(c_lo, c_hi);
      uint32_t f[4] = {__byte_perm(c.x, 0, 0x4140), __byte_perm(c.x, 0, 0x4342),
                       __byte_perm(c.y, 0, 0x4140), __byte_perm(c.y, 0, 0x4342)};
      uint32_t excl[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) excl[k] = f[k];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, f[k], o);
          if (lane >= o) f[k] += t;
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) excl[k] = f[k] - excl[k];
      __syncthreads();  // previous tile's s_keys / s_dst / warp_off / s_bias fully consumed
      if (lane == 31) *reinterpret_cast<uint4*>(warp_tot[warp]) = make_uint4(f[0], f[1], f[2], f[3]);
      __syncthreads();
      // ---- per destination: runs of the warps, tile total, reservation in the bucket
      unsigned long long b = 0;
      uint32_t run = 0;
      if (tid < kMaxWorld) {
        const int word = tid >> 1, shf = (tid & 1) * 16;
#pragma unroll
        for (int wv = 0; wv < kPartWarps; ++wv) {
          warp_off[wv][tid] = (uint16_t)run;
          run += (warp_tot[wv][word] >> shf) & 0xffffu;
        }
        tile_total[tid] = run;
        // reserve room in the destination bucket.  The (mostly remote) atomic's result is
        // not touched until the keys are staged, so its NVLink round trip overlaps staging.
        if (tid < world && run) {
          if (phase_mask & 8)
            // my segment in that rank's bucket was reserved by sort_count_reserve_kernel:
            // sub-allocate it with a LOCAL atomic
            b = atomicAdd(&me->sort_local_cursor[tid], (unsigned long long)run);
          else
            b = atomicAdd_system(s_cursor[tid], (unsigned long long)run);
        }
      }
      __syncthreads();
      uint32_t start = 0;
      if (tid < kMaxWorld) {
#pragma unroll
        for (int k = 0; k < kMaxWorld; ++k) start += k < tid ? tile_total[k] : 0u;
#pragma unroll
        for (int wv = 0; wv < kPartWarps; ++wv) warp_off[wv][tid] = (uint16_t)(warp_off[wv][tid] + start);
      }
      __syncthreads();
      // ---- stage keys (and their destination bytes) in destination order
      {
        const uint4 wo = *reinterpret_cast<const uint4*>(warp_off[warp]);
        const uint32_t g0 = wo.x + excl[0], g1 = wo.y + excl[1], g2 = wo.z + excl[2], g3 = wo.w + excl[3];
#pragma unroll
        for (int i = 0; i < kPartItems; ++i) {
          const uint32_t m = (meta[i >> 2] >> (8 * (i & 3))) & 0xffu;
          const uint32_t d = m & 0xfu;
          if (d < kNoDest) {
            const uint32_t g = (d & 4u) ? ((d & 2u) ? g3 : g2) : ((d & 2u) ? g1 : g0);
            const uint32_t pos = ((g >> ((d & 1u) << 4)) & 0xffffu) + (m >> 4);
            s_keys[pos] = key[i];
            s_dst[pos] = (uint8_t)d;
          }
        }
      }
      if (tid < kMaxWorld) {
        if (tid < world && b + run > capacity) {
          // overflow (or, with bit 8, a reservation that failed earlier): this tile's keys
          // for that destination are dropped and the error word says so
          if (!(phase_mask & 8)) atomicOr(&me->error, (unsigned long long)kErrSortOverflow);
          b = ~0ull;
        }
        s_bias[tid] = (tid < world && b != ~0ull) ? s_bucket[tid] + b - start : nullptr;
      }
      __syncthreads();
      // ---- runs of equal destination leave as contiguous (mostly remote) stores
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t p = tid + i * kPartThreads;
        if (p < n_valid) {
          uint32_t* dst = s_bias[s_dst[p]];
          if (dst) dst[p] = s_keys[p];
        }
      }
    }
    // ---- ticket: last CTA tells every peer that this rank has finished writing
    __threadfence_system();
    __syncthreads();
    if (tid == 0) is_last = atomicAdd(&me->sort_blocks_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    if (!(phase_mask & 4)) {  // bit 2 set: local-only partition (NCCL baseline), no signalling
      if (tid == 0) me->sort_blocks_done = 0;
      __threadfence_system();
      if (tid < world) signal_peer(peers, tid, rank, epoch, kFlagSortDone);
    } else if (tid == 0) {
      me->sort_blocks_done = 0;
    }
  }
  if ((phase_mask & 2) && !(phase_mask & 4)) {
    if (tid < world) wait_peer(peers, rank, tid, epoch, timeout_ns, kFlagSortDone);
    __syncthreads();
    if (tid == 0) {
      unsigned long long c;
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(c) : "l"(&me->sort_cursor) : "memory");
      me->sort_recv_count = c < capacity ? c : capacity;
    }
  }
}

// ---- second form of the fused partition + exchange ----------------------------------------
// Same job and same interface as the kernel above, built like the radix sort's first pass
// (lab5_sort.cu): this step keeps no order, so a tile COUNTS its keys per destination on
// (destination, lane) counters - one bank column per lane, conflict-free - turns the counters
// into cursors and PLACES every key with one shared-memory atomic on its (destination, lane)
// cursor.  No per-thread packed counters, no warp scans, three CTA barriers per tile instead of
// five, keys loaded as 16-byte vectors (a partition need not see them in memory order), and two
// counter sets used in turn so that clearing one never waits for the other.
struct Part2Smem {
  uint32_t keys[kPartTile];            // keys in destination order
  uint8_t dst[kPartTile];              // their destinations
  uint8_t dest_lut[1 << kDestLutBits];
  uint32_t cnt[2][kMaxWorld][32];      // [set][destination][lane]: counts, then cursors
  uint32_t* bias[kMaxWorld + 1];       // bucket + reserved base - tile start; null = dropped
  uint32_t hi[kMaxWorld], lo[kMaxWorld];
  uint32_t* bucket[kMaxWorld];
  unsigned long long* cursor[kMaxWorld];
  bool is_last;
};

template <int KT, int BLOCKS>
__global__ void __launch_bounds__(kPartThreads, BLOCKS)
    sort_partition_scatter2_kernel(const uint32_t* __restrict__ keys, unsigned long long n_local,
                                   const unsigned long long* __restrict__ splitters, int idx_bits,
                                   PartTargets tg, unsigned long long capacity, SymmPeers peers,
                                   int rank, int world, uint64_t epoch, int phase_mask,
                                   unsigned long long timeout_ns) {
  extern __shared__ __align__(16) uint8_t part_smem_raw[];
  Part2Smem& S = *reinterpret_cast<Part2Smem*>(part_smem_raw);
  SymmWorkspace* me = peers.ws[rank];
  const int tid = threadIdx.x, lane = tid & 31;

  if (phase_mask & 1) {
    if (tid < kMaxWorld) {
      const unsigned long long sp = tid < world ? splitters[tid] : ~0ull;
      S.hi[tid] = (uint32_t)(sp >> 32);
      S.lo[tid] = (uint32_t)sp;
      S.bucket[tid] = tg.bucket[tid];
      S.cursor[tid] = tg.cursor[tid];
    }
    if (tid == 0) S.bias[kMaxWorld] = nullptr;
    for (int i = tid; i < 2 * kMaxWorld * 32; i += kPartThreads) (&S.cnt[0][0][0])[i] = 0;
    __syncthreads();
    // ---- destination table over [first splitter key, last splitter key] (as above)
    const uint32_t lut_base = world > 1 ? S.hi[1] : 0xffffffffu;
    const uint32_t lut_span = world > 1 ? S.hi[world - 1] - lut_base : 0u;
    const int lut_shift = lut_span < (1u << kDestLutBits) ? 0 : 32 - kDestLutBits - __clz(lut_span);
    for (int bkt = tid; bkt < (1 << kDestLutBits); bkt += kPartThreads) {
      const unsigned long long lo = (unsigned long long)lut_base + ((unsigned long long)bkt << lut_shift);
      const unsigned long long hi = lo + ((1ull << lut_shift) - 1);
      uint32_t below = 0, upto = 0;
#pragma unroll
      for (int k = 1; k < kMaxWorld; ++k)
        if (k < world) {
          below += (unsigned long long)S.hi[k] < lo ? 1u : 0u;
          upto += (unsigned long long)S.hi[k] <= hi ? 1u : 0u;
        }
      S.dest_lut[bkt] = (uint8_t)(below | (upto != below ? kDestAmbiguous : 0u));
    }
    __syncthreads();
    const uint32_t src_bits = idx_bits >= 32 ? 0u : ((uint32_t)rank << idx_bits);
    const unsigned long long n_tiles = (n_local + kPartTile - 1) / kPartTile;
    const bool aligned = (((uintptr_t)keys) & 15u) == 0;
    int set = 0;
    // persistent: a CTA walks tiles blockIdx.x, +gridDim.x, ... and fences its remote stores ONCE
    // at the end (a system-scope fence per tile costs an NVLink round trip per tile)
    for (unsigned long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, set ^= 1) {
      const unsigned long long tile_base = tile * kPartTile;
      const uint32_t n_valid = (uint32_t)min((unsigned long long)kPartTile, n_local - tile_base);
      // ---- load: thread t owns elements (j * 512 + t) * 4 + k of the tile
      uint32_t key[kPartItems];
#pragma unroll
      for (int j = 0; j < kPartItems / 4; ++j) {
        const uint32_t o = (j * kPartThreads + tid) * 4;
        if (aligned && o + 3 < n_valid) {
          const uint4 v = ld_stream_u4(keys + tile_base + o);
          key[4 * j] = v.x, key[4 * j + 1] = v.y, key[4 * j + 2] = v.z, key[4 * j + 3] = v.w;
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) key[4 * j + k] = o + k < n_valid ? ld_stream_u32(keys + tile_base + o + k) : 0u;
        }
      }
      // ---- destinations (table; exact comparison only next to a splitter), 4 bits per key
      unsigned long long meta = 0;
      const uint32_t cnt_base = smem_u32(&S.cnt[set][0][0]) + 4u * lane;
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t o = ((i >> 2) * kPartThreads + tid) * 4 + (i & 3);
        const uint32_t x = to_ordered_mg<KT>(key[i]);
        const uint32_t t = (x - lut_base) >> lut_shift;
        uint32_t e = S.dest_lut[t < (1u << kDestLutBits) ? t : 0u];
        e = t < (1u << kDestLutBits) ? e : (uint32_t)(world - 1);  // above every splitter
        e = x < lut_base ? 0u : e;                                  // below every splitter
        if (e & kDestAmbiguous) {
          const uint32_t low = src_bits | (uint32_t)(tile_base + o);
          e = 0;
#pragma unroll
          for (int k = 1; k < kMaxWorld; ++k)
            if (k < world) e += (S.hi[k] < x || (S.hi[k] == x && S.lo[k] <= low)) ? 1u : 0u;
        }
        e = o < n_valid ? e : kNoDest;
        meta |= (unsigned long long)e << (4 * i);
        if (e < kNoDest) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(cnt_base + e * 128u) : "memory");
      }
      __syncthreads();
      // ---- per destination: total, reservation in the destination's bucket, cursors
      unsigned long long b = 0;
      uint32_t total = 0, start = 0;
      if (tid < kMaxWorld) {
        uint32_t* row = S.cnt[set][tid];
        const int rot = 4 * tid;  // the eight threads walk their rows through different banks
#pragma unroll 8
        for (int j = 0; j < 32; ++j) total += row[(j + rot) & 31];
        // The (mostly remote) atomic's result is not touched until the keys are placed, so its
        // NVLink round trip overlaps the placement.
        if (tid < world && total) {
          if (phase_mask & 8)
            b = atomicAdd(&me->sort_local_cursor[tid], (unsigned long long)total);
          else
            b = atomicAdd_system(S.cursor[tid], (unsigned long long)total);
        }
        uint32_t inc = total;
#pragma unroll
        for (int o = 1; o < kMaxWorld; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffu, inc, o);
          if (tid >= o) inc += t;
        }
        start = inc - total;
        uint32_t run = start;
#pragma unroll 8
        for (int j = 0; j < 32; ++j) {
          uint32_t* c = &row[(j + rot) & 31];
          const uint32_t v = *c;
          *c = run;
          run += v;
        }
      } else if (tid >= kPartThreads - kMaxWorld * 32) {
        // the other counter set (last used by the previous tile's placement) is cleared here
        (&S.cnt[set ^ 1][0][0])[tid - (kPartThreads - kMaxWorld * 32)] = 0;
      }
      __syncthreads();
      // ---- place: one shared-memory atomic per key on its (destination, lane) cursor
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t e = (uint32_t)(meta >> (4 * i)) & 0xfu;
        if (e < kNoDest) {
          uint32_t pos;
          asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(pos) : "r"(cnt_base + e * 128u) : "memory");
          S.keys[pos] = key[i];
          S.dst[pos] = (uint8_t)e;
        }
      }
      if (tid < kMaxWorld) {
        if (tid < world && b + total > capacity) {
          // overflow (or, with bit 8, a reservation that failed earlier): this tile's keys for that
          // destination are dropped and the error word says so (the host grows the buckets and re-runs)
          if (!(phase_mask & 8)) atomicOr(&me->error, (unsigned long long)kErrSortOverflow);
          b = ~0ull;
        }
        S.bias[tid] = (tid < world && b != ~0ull) ? S.bucket[tid] + b - start : nullptr;
      }
      __syncthreads();
      // ---- runs of equal destination leave as contiguous (mostly remote) stores
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const uint32_t p = tid + i * kPartThreads;
        if (p < n_valid) {
          uint32_t* dst = S.bias[S.dst[p]];
          if (dst) dst[p] = S.keys[p];
        }
      }
      // no barrier here: the next tile counts on the OTHER counter set, and its placement into
      // S.keys / S.dst comes after two more barriers, i.e. after every thread has left this loop
    }
    // ---- ticket: last CTA tells every peer that this rank has finished writing
    __threadfence_system();
    __syncthreads();
    if (tid == 0) S.is_last = atomicAdd(&me->sort_blocks_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!S.is_last) return;
    if (!(phase_mask & 4)) {  // bit 2 set: local-only partition (NCCL baseline), no signalling
      if (tid == 0) me->sort_blocks_done = 0;
      __threadfence_system();
      if (tid < world) signal_peer(peers, tid, rank, epoch, kFlagSortDone);
    } else if (tid == 0) {
      me->sort_blocks_done = 0;
    }
  }
  if ((phase_mask & 2) && !(phase_mask & 4)) {
    if (tid < world) wait_peer(peers, rank, tid, epoch, timeout_ns, kFlagSortDone);
    __syncthreads();
    if (tid == 0) {
      unsigned long long c;
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(c) : "l"(&me->sort_cursor) : "memory");
      me->sort_recv_count = c < capacity ? c : capacity;
    }
  }
}

}  // namespace b200

using namespace b200;

static int idx_bits_for(int world) {
  int src_bits = 0;
  while ((1 << src_bits) < world) ++src_bits;
  return 32 - src_bits;
}

// Step 1 of the multi-GPU sort: sample, exchange, choose splitters (left in
// ws[rank]->sort_splitters).  samples_per_rank * world <= 8192; any world size up to 8.
extern "C" int b200_sort_mgpu_splitters(const void* keys, size_t n_local, int key_type,
                                        int samples_per_rank, const void* ws_ptrs, int rank,
                                        int world, uint64_t epoch, int phase_mask,
                                        b200_stream_t s) {
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world,
               "sort_mgpu: bad rank/world %d/%d", rank, world);
  const int total = world * samples_per_rank;
  B200_REQUIRE(samples_per_rank > 0 && total <= kSortMaxSamplesTotal,
               "sort_mgpu: world*samples_per_rank must be at most %d (got %d)", kSortMaxSamplesTotal,
               total);
  const int idx_bits = idx_bits_for(world);
  B200_REQUIRE(idx_bits >= 32 || n_local <= (1ull << idx_bits), "sort_mgpu: shard too large");
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  size_t padded = 1;
  while (padded < (size_t)total) padded <<= 1;
  const size_t smem = padded * sizeof(unsigned long long);
  cudaStream_t st = as_stream(s);
#define LAUNCH_SPLIT(KT)                                                                       \
  do {                                                                                         \
    B200_CUDA_OK(cudaFuncSetAttribute(sort_splitters_kernel<KT>,                               \
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));    \
    sort_splitters_kernel<KT><<<1, kSplitThreads, smem, st>>>(                                 \
        (const uint32_t*)keys, n_local, samples_per_rank, idx_bits, peers, rank, world, epoch, \
        0x5bd1e995u, phase_mask, kDefaultTimeoutNs);                                           \
  } while (0)
  switch (key_type) {
    case B200_KEY_I32: LAUNCH_SPLIT(B200_KEY_I32); break;
    case B200_KEY_U32: LAUNCH_SPLIT(B200_KEY_U32); break;
    case B200_KEY_F32: LAUNCH_SPLIT(B200_KEY_F32); break;
    default: B200_REQUIRE(false, "sort_mgpu: 32-bit key types only (got %d)", key_type);
  }
#undef LAUNCH_SPLIT
  B200_LAUNCH_OK();
  return 0;
}

// Step 2: fused partition + all-to-all.  bucket_ptrs / cursor_ptrs: HOST arrays
// [world] of device addresses (peer-mapped buckets and their uint64 fill cursors;
// pass cursor_ptrs = NULL to use each rank's workspace cursor).  splitters: device
// uint64[world] (normally &ws[rank]->sort_splitters).  phase_mask: 1 scatter+signal,
// 2 wait+publish count, 4 (with 1) local partition only, no signalling.
extern "C" int b200_sort_mgpu_scatter(const void* keys, size_t n_local, int key_type,
                                      const void* splitters, void* const* bucket_ptrs,
                                      void* const* cursor_ptrs, size_t capacity,
                                      const void* ws_ptrs, void* const* ws_host_ptrs, int rank,
                                      int world, uint64_t epoch, int phase_mask,
                                      b200_stream_t s) {
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world,
               "sort_mgpu: bad rank/world %d/%d", rank, world);
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  PartTargets tg;
  for (int p = 0; p < kMaxWorld; ++p) {
    tg.bucket[p] = p < world ? (uint32_t*)bucket_ptrs[p] : nullptr;
    if (p >= world)
      tg.cursor[p] = nullptr;
    else if (cursor_ptrs)
      tg.cursor[p] = (unsigned long long*)cursor_ptrs[p];
    else
      tg.cursor[p] = &((SymmWorkspace*)ws_host_ptrs[p])->sort_cursor;
  }
  const int idx_bits = idx_bits_for(world);
  B200_REQUIRE(n_local > 0, "sort_mgpu: empty shards are not supported");
  // CTAs per SM the kernel is compiled for: 2 (64 registers) or 3 (40 registers, a few spills,
  // more tiles in flight to hide the remote reservations); B200LABS_PART_BLOCKS overrides
  static const int part_blocks = [] {
    const char* e = getenv("B200LABS_PART_BLOCKS");
    const int v = e ? atoi(e) : B200_PART_BLOCKS;
    return v == 3 ? 3 : 2;
  }();
  unsigned grid = 1u;
  if (phase_mask & 1) {
    const size_t tiles = (n_local + kPartTile - 1) / kPartTile;
    const size_t resident = (size_t)device_info().sm_count * part_blocks;
    grid = (unsigned)(tiles < resident ? tiles : resident);
  }
  cudaStream_t st = as_stream(s);
  // B200LABS_PART_IMPL=1 selects the first form of the kernel (packed register counters + warp scans)
  static const int part_impl = [] {
    const char* e = getenv("B200LABS_PART_IMPL");
    return e ? atoi(e) : 2;
  }();
  static_assert(sizeof(Part2Smem) <= 48 * 1024, "Part2Smem must fit the default dynamic shared memory limit");
#define LAUNCH_PART(KT)                                                                         \
  if (part_impl == 1)                                                                           \
    (part_blocks == 3 ? sort_partition_scatter_kernel<KT, 3> : sort_partition_scatter_kernel<KT, 2>)<<<grid, kPartThreads, 0, st>>>( \
        (const uint32_t*)keys, n_local, (const unsigned long long*)splitters, idx_bits, tg,     \
        capacity, peers, rank, world, epoch, phase_mask, kDefaultTimeoutNs);                    \
  else                                                                                          \
    (part_blocks == 3 ? sort_partition_scatter2_kernel<KT, 3> : sort_partition_scatter2_kernel<KT, 2>)<<<grid, kPartThreads, sizeof(Part2Smem), st>>>( \
        (const uint32_t*)keys, n_local, (const unsigned long long*)splitters, idx_bits, tg,     \
        capacity, peers, rank, world, epoch, phase_mask, kDefaultTimeoutNs)
  switch (key_type) {
    case B200_KEY_I32: LAUNCH_PART(B200_KEY_I32); break;
    case B200_KEY_U32: LAUNCH_PART(B200_KEY_U32); break;
    case B200_KEY_F32: LAUNCH_PART(B200_KEY_F32); break;
    default: B200_REQUIRE(false, "sort_mgpu: 32-bit key types only (got %d)", key_type);
  }
#undef LAUNCH_PART
  B200_LAUNCH_OK();
  return 0;
}

// Step 1b: count my keys per destination and reserve one segment in each rank's bucket
// (run after the splitters step, before a scatter launched with phase bit 8).
extern "C" int b200_sort_mgpu_count(const void* keys, size_t n_local, int key_type,
                                    const void* splitters, size_t capacity, const void* ws_ptrs,
                                    int rank, int world, b200_stream_t s) {
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world,
               "sort_mgpu: bad rank/world %d/%d", rank, world);
  B200_REQUIRE(n_local > 0, "sort_mgpu: empty shards are not supported");
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  const int idx_bits = idx_bits_for(world);
  size_t blocks = (n_local + (size_t)kCountThreads * kCountItems - 1) / ((size_t)kCountThreads * kCountItems);
  const size_t cap = (size_t)device_info().sm_count * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = as_stream(s);
#define LAUNCH_COUNT(KT)                                                                        \
  sort_count_reserve_kernel<KT><<<(unsigned)blocks, kCountThreads, 0, st>>>(                    \
      (const uint32_t*)keys, n_local, (const unsigned long long*)splitters, idx_bits, capacity, \
      peers, rank, world)
  switch (key_type) {
    case B200_KEY_I32: LAUNCH_COUNT(B200_KEY_I32); break;
    case B200_KEY_U32: LAUNCH_COUNT(B200_KEY_U32); break;
    case B200_KEY_F32: LAUNCH_COUNT(B200_KEY_F32); break;
    default: B200_REQUIRE(false, "sort_mgpu: 32-bit key types only (got %d)", key_type);
  }
#undef LAUNCH_COUNT
  B200_LAUNCH_OK();
  return 0;
}

// Offsets of workspace fields, so Python can read them without mirroring the struct.
extern "C" size_t b200_symm_offset(const char* field) {
  if (!strcmp(field, "error")) return offsetof(SymmWorkspace, error);
  if (!strcmp(field, "sort_cursor")) return offsetof(SymmWorkspace, sort_cursor);
  if (!strcmp(field, "sort_splitters")) return offsetof(SymmWorkspace, sort_splitters);
  if (!strcmp(field, "sort_recv_count")) return offsetof(SymmWorkspace, sort_recv_count);
  if (!strcmp(field, "sort_samples")) return offsetof(SymmWorkspace, sort_samples);
  if (!strcmp(field, "flags")) return offsetof(SymmWorkspace, flags);
  return (size_t)-1;
}
