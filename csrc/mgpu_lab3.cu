// lab3, multi-GPU: class statistics all-reduced INSIDE the statistics kernel.
//
// The image is row-sharded; every rank gathers the sample pixels that fall into
// its band and accumulates the ten exact integer moments per class.  The last CTA
// to finish (ticket counter) pushes this rank's 32x10 partial sums straight into
// every peer's slot over NVLink (plain remote stores), fences at system scope and
// raises the peers' flags; it then waits for all peers' flags, adds the slots in
// rank order and derives mean + inverse covariance — one launch per rank, no
// collective library call.  Integer sums are associative, so 1, 2, 4 or 8 GPUs
// produce bit-identical parameters.
// phase_mask: 1 = accumulate + push + signal, 2 = wait + reduce + finalize, 3 = both
// (the product path).  The split exists so that several virtual ranks can be
// exercised on ONE GPU without kernels that wait on each other.
// (ref: the reference computes these statistics on the host, lab3/src/main.cu:102-152,
// and has no multi-GPU path.)
#include <cstdlib>

#include "common.cuh"
#include "lab3_math.cuh"
#include "symm.cuh"

namespace b200 {

constexpr int kStatThreads = 256;

__device__ __forceinline__ long long ld_cg_s64(const long long* p) {
  long long v;
  asm volatile("ld.global.cg.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kStatThreads)
    lab3_stats_allreduce_kernel(const uint32_t* __restrict__ img, int w, int64_t y0,
                                int64_t h_local, const int2* __restrict__ pts,
                                const int* __restrict__ class_off, int nc, SymmPeers peers,
                                int rank, int world, uint64_t epoch, double* __restrict__ params,
                                long long* __restrict__ moments_out, int phase_mask,
                                unsigned long long timeout_ns,
                                unsigned long long* __restrict__ mc_acc /* NVLS multicast view */,
                                long long* __restrict__ local_acc /* this GPU's copy of it */,
                                int par_override /* >= 0: accumulator half to use */) {
  SymmWorkspace* me = peers.ws[rank];
  __shared__ unsigned long long red[kStatThreads / 32][kLab3M];
  __shared__ long long total[kLab3C][kLab3M];
  __shared__ bool is_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int par = par_override >= 0 ? par_override : (int)(epoch & 1);

  if (phase_mask & 1) {
    // ---- accumulate this CTA's share of class blockIdx.y --------------------
    const int c = blockIdx.y;
    const int begin = class_off[c], end = class_off[c + 1];
    unsigned long long m[kLab3M];
#pragma unroll
    for (int k = 0; k < kLab3M; ++k) m[k] = 0;
    for (int i = begin + blockIdx.x * kStatThreads + tid; i < end; i += gridDim.x * kStatThreads) {
      const int2 p = __ldg(pts + i);
      const int64_t yl = (int64_t)p.y - y0;
      if (yl < 0 || yl >= h_local || p.x < 0 || p.x >= w) continue;
      const uint32_t px = __ldg(img + yl * w + p.x);
      const unsigned r = px & 0xff, g = (px >> 8) & 0xff, b = (px >> 16) & 0xff;
      m[0] += 1;
      m[1] += r, m[2] += g, m[3] += b;
      m[4] += r * r, m[5] += r * g, m[6] += r * b;
      m[7] += g * g, m[8] += g * b, m[9] += b * b;
    }
#pragma unroll
    for (int k = 0; k < kLab3M; ++k) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m[k] += __shfl_down_sync(0xffffffffu, m[k], o);
    }
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < kLab3M; ++k) red[warp][k] = m[k];
    }
    __syncthreads();
    if (tid < kLab3M) {
      unsigned long long s = 0;
      for (int wv = 0; wv < kStatThreads / 32; ++wv) s += red[wv][tid];
      if (s) atomicAdd((unsigned long long*)&me->lab3_acc[c][tid], s);
    }
    // ---- ticket: the last CTA of the grid carries on -------------------------
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      const unsigned int n_blocks = gridDim.x * gridDim.y;
      is_last = atomicAdd(&me->lab3_blocks_done, 1u) == n_blocks - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (mc_acc) {
      // ---- NVLS: one multimem.red per value adds my partial sum into EVERY GPU's copy
      // of the accumulator (the NVSwitch performs the broadcast-reduction).  The other
      // parity's accumulator is zeroed first: nobody can add to it before passing this
      // epoch's barrier, which needs the signal sent below.
      for (int i = tid; i < kLab3C * kLab3M; i += kStatThreads) local_acc[(par ^ 1) * kLab3C * kLab3M + i] = 0;
      __threadfence_system();
      __syncthreads();
      for (int i = tid; i < nc * kLab3M; i += kStatThreads) {
        long long* acc = &me->lab3_acc[0][0] + i;
        const long long v = ld_cg_s64(acc);
        *acc = 0;
        if (v)
          asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(
                           mc_acc + par * kLab3C * kLab3M + i),
                       "l"(v)
                       : "memory");
      }
    } else {
      // ---- push my partial sums into every peer's slot (remote stores) ---------
      for (int i = tid; i < nc * kLab3M; i += kStatThreads) {
        long long* acc = &me->lab3_acc[0][0] + i;
        const long long v = ld_cg_s64(acc);
        *acc = 0;  // ready for the next call
        for (int p = 0; p < world; ++p) (&peers.ws[p]->lab3_slots[par][rank][0][0])[i] = v;
      }
    }
    if (tid == 0) me->lab3_blocks_done = 0;
    __threadfence_system();
    __syncthreads();
    if (tid < world) {
      __threadfence_system();
      signal_peer(peers, tid, rank, epoch, kFlagLab3);
    }
  }

  if (phase_mask & 2) {
    // ---- wait for every rank's push, reduce in rank order, finalize ---------
    if (tid < world) wait_peer(peers, rank, tid, epoch, timeout_ns, kFlagLab3);
    __syncthreads();
    for (int i = tid; i < nc * kLab3M; i += kStatThreads) {
      long long s = 0;
      if (local_acc) {
        s = ld_cg_s64(local_acc + par * kLab3C * kLab3M + i);  // already reduced by the switch
      } else {
        for (int p = 0; p < world; ++p) s += ld_cg_s64(&me->lab3_slots[par][p][0][0] + i);
      }
      (&total[0][0])[i] = s;
      if (moments_out) moments_out[i] = s;
    }
    __syncthreads();
    if (tid < nc) lab3_finalize_class(total[tid], params + (size_t)tid * B200_LAB3_PARAMS);
  }
}

// ---- colour table built cooperatively: 1/world of the 2^24 colours per rank ----
// Each rank classifies its slice of the colour cube (a range of B values) and stores
// the class bytes into EVERY rank's table over NVLink, so the table build scales with
// the GPU count instead of being repeated on every GPU; the last CTA fences, signals
// all peers and waits for them, after which the local table is complete and the image
// pass can gather from it.
struct LutTargets {
  uint8_t* lut[kMaxWorld];  // every rank's table, peer-mapped
};

// Rank r classifies the colours with B in [r*256/world, (r+1)*256/world) — one CTA
// per (G, 64 R values), hoisted chain as in lut_build_block — and stores every class
// byte into all tables (a warp writes 32 consecutive bytes = one sector per table).
__global__ void __launch_bounds__(kLutThreads)
    lab3_lut_build_allgather_kernel(LutTargets tg, int nc, const double* __restrict__ params,
                                    SymmPeers peers, int rank, int world, uint64_t epoch,
                                    int phase_mask, unsigned long long timeout_ns) {
  extern __shared__ __align__(16) uint8_t lut_smem_raw[];
  LutSmem& S = *reinterpret_cast<LutSmem*>(lut_smem_raw);
  __shared__ uint8_t* s_lut[kMaxWorld];
  __shared__ bool is_last;
  SymmWorkspace* me = peers.ws[rank];
  const int tid = threadIdx.x;
  if (phase_mask & 1) {
    if (tid < kMaxWorld) s_lut[tid] = tid < world ? tg.lut[(rank + tid) % world] : nullptr;  // own copy first
    const int per_rank = 256 / world;
    const int g = blockIdx.x / (256 / kLutR), r0 = (blockIdx.x % (256 / kLutR)) * kLutR;
    lut_build_block(S, nc, params, g, r0, rank * per_rank, (rank + 1) * per_rank,
                    [&](uint32_t id, uint8_t cls) {
                      for (int p = 0; p < world; ++p) s_lut[p][id] = cls;
                    });
    __threadfence_system();
    __syncthreads();
    if (tid == 0) is_last = atomicAdd(&me->lab3_lut_blocks_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    if (tid == 0) me->lab3_lut_blocks_done = 0;
    __threadfence_system();
    if (tid < world) signal_peer(peers, tid, rank, epoch, kFlagLab3Lut);
  }
  if (phase_mask & 2) {
    if (tid < world) wait_peer(peers, rank, tid, epoch, timeout_ns, kFlagLab3Lut);
  }
}

// Second form of the cooperative build: colour table AND cell table (lut_build_cells_cta),
// every piece leaving as ONE 16-byte store per destination - or, with `mc_lut` (the NVLS
// multicast address of the ranks' table buffers), as one multimem.st that the NVSwitch
// replicates into all of them, so a rank's slice crosses its NVLink once instead of once
// per peer.  Layout of a table buffer: 2^24 colour bytes, then 2^17 cell bytes.  Rank r
// owns the B chunks r, r + world, ... (any world size up to 8).
__device__ __forceinline__ void st_v4(uint8_t* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_v4(uint8_t* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}

__global__ void __launch_bounds__(kLutThreads)
    lab3_lut_build2_allgather_kernel(LutTargets tg, uint8_t* mc_lut, int nc, const double* __restrict__ params,
                                     SymmPeers peers, int rank, int world, uint64_t epoch, int phase_mask,
                                     unsigned long long timeout_ns) {
  extern __shared__ __align__(16) uint8_t lut_smem_raw[];
  Lut2Smem& S = *reinterpret_cast<Lut2Smem*>(lut_smem_raw);
  __shared__ uint8_t* s_lut[kMaxWorld];
  __shared__ bool is_last;
  SymmWorkspace* me = peers.ws[rank];
  const int tid = threadIdx.x;
  if (phase_mask & 1) {
    if (tid < kMaxWorld) s_lut[tid] = tid < world ? tg.lut[(rank + tid) % world] : nullptr;  // own copy first
    __syncthreads();
    const int rblock = blockIdx.x & 3, gcell = (blockIdx.x >> 2) & 63;
    const int chunk = rank + world * (int)(blockIdx.x >> 8);
    auto put = [&](size_t off, const uint4& v) {
      if (mc_lut) {
        multimem_st_v4(mc_lut + off, v);
      } else {
        for (int p = 0; p < world; ++p) st_v4(s_lut[p] + off, v);
      }
    };
    lut_build_cells_cta(
        S, nc, params, gcell, rblock * kLutR, chunk * kLut2B, [&](uint32_t id, const uint4& v) { put(id, v); },
        [&](uint32_t idx, const uint4& v) { put((size_t(1) << 24) + idx, v); });
    __threadfence_system();
    __syncthreads();
    if (tid == 0) is_last = atomicAdd(&me->lab3_lut_blocks_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    if (tid == 0) me->lab3_lut_blocks_done = 0;
    __threadfence_system();
    if (tid < world) signal_peer(peers, tid, rank, epoch, kFlagLab3Lut);
  }
  if (phase_mask & 2) {
    if (tid < world) wait_peer(peers, rank, tid, epoch, timeout_ns, kFlagLab3Lut);
  }
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_lab3_lut2_bytes(void) { return (size_t(1) << 24) + (size_t(1) << 17); }

extern "C" int b200_lab3_lut_build2_allgather(void* const* lut_ptrs, void* mc_lut, int nc,
                                              const double* params, const void* ws_ptrs, int rank,
                                              int world, uint64_t epoch, int phase_mask,
                                              b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kLab3C, "lab3: class count %d outside 1..%d", nc, kLab3C);
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world,
               "lab3: bad rank/world %d/%d", rank, world);
  B200_REQUIRE(mc_lut || lut_ptrs, "lab3: table build needs the peers' tables or their multicast address");
  LutTargets tg;
  for (int p = 0; p < kMaxWorld; ++p) tg.lut[p] = (lut_ptrs && p < world) ? (uint8_t*)lut_ptrs[p] : nullptr;
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  static PerDeviceOnce attr_once;
  if (int rc = attr_once.run("lab3_mgpu", [&]() -> int {
        B200_CUDA_OK(cudaFuncSetAttribute(lab3_lut_build2_allgather_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Lut2Smem)));
        return 0;
      }))
    return rc;
  // world <= 8 = the number of B chunks, so every rank owns at least one
  const unsigned grid = (phase_mask & 1) ? (unsigned)lut2_chunks_of(rank, world) * 256u : 1u;
  lab3_lut_build2_allgather_kernel<<<grid, kLutThreads, sizeof(Lut2Smem), as_stream(s)>>>(
      tg, (uint8_t*)mc_lut, nc, params, peers, rank, world, epoch, phase_mask, kDefaultTimeoutNs);
  B200_LAUNCH_OK();
  return 0;
}

// Cooperative colour-table build (see kernel comment).  lut_ptrs: HOST array [world]
// of the peer-mapped 16 MiB tables.  After the call (stream order) the local table
// is complete; follow with b200_lab3_apply_lut.  phase_mask as for the statistics.
extern "C" int b200_lab3_lut_build_allgather(void* const* lut_ptrs, int nc, const double* params,
                                             const void* ws_ptrs, int rank, int world,
                                             uint64_t epoch, int phase_mask, b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kLab3C, "lab3: class count %d outside 1..%d", nc, kLab3C);
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && (world & (world - 1)) == 0 && rank >= 0 &&
                   rank < world,
               "lab3: lut all-gather needs a power-of-two world <= %d (rank %d, world %d)",
               kMaxWorld, rank, world);
  LutTargets tg;
  for (int p = 0; p < kMaxWorld; ++p) tg.lut[p] = p < world ? (uint8_t*)lut_ptrs[p] : nullptr;
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  B200_REQUIRE((256 / world) % kLutBGroup == 0, "lab3: world %d leaves fewer than 8 B values per rank", world);
  static PerDeviceOnce attr_once;
  if (int rc = attr_once.run("lab3_mgpu", [&]() -> int {
    B200_CUDA_OK(cudaFuncSetAttribute(lab3_lut_build_allgather_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LutSmem)));
        return 0;
      }))
    return rc;
  const unsigned grid = (phase_mask & 1) ? 256u * (256 / kLutR) : 1u;
  lab3_lut_build_allgather_kernel<<<grid, kLutThreads, sizeof(LutSmem), as_stream(s)>>>(
      tg, nc, params, peers, rank, world, epoch, phase_mask, kDefaultTimeoutNs);
  B200_LAUNCH_OK();
  return 0;
}

// CTAs per class of the statistics kernel (B200LABS_LAB3_STATS_GRID overrides)
static int stats_grid_x() {
  static const int v = [] {
    const char* e = getenv("B200LABS_LAB3_STATS_GRID");
    const int g = e ? atoi(e) : 16;  // 64 / 16 / 8 CTAs per class: 0.0367 / 0.0342 / 0.0318 ms at 4096 points per class; 16 keeps headroom for 2^19
    return g < 1 ? 1 : g > 1024 ? 1024 : g;
  }();
  return v;
}

// Fused statistics + all-reduce + finalize.  ws_ptrs: device array [world] of the
// peer-mapped SymmWorkspace of every rank; epoch must grow by one per call (the
// same value on every rank).  moments_out (optional) receives the reduced integers.
extern "C" int b200_lab3_stats_allreduce(const void* img, int w, int64_t y0, int64_t h_local,
                                         const int* pts, const int* class_off, int nc,
                                         const void* ws_ptrs, int rank, int world,
                                         uint64_t epoch, double* params, long long* moments_out,
                                         int phase_mask, b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kLab3C, "lab3: class count %d outside 1..%d", nc, kLab3C);
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world,
               "lab3: bad rank/world %d/%d", rank, world);
  B200_REQUIRE(phase_mask >= 1 && phase_mask <= 3, "lab3: phase_mask must be 1, 2 or 3");
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  dim3 grid = (phase_mask & 1) ? dim3(stats_grid_x(), nc) : dim3(1, 1);
  lab3_stats_allreduce_kernel<<<grid, kStatThreads, 0, as_stream(s)>>>(
      (const uint32_t*)img, w, y0, h_local, (const int2*)pts, class_off, nc, peers, rank, world,
      epoch, params, moments_out, phase_mask, kDefaultTimeoutNs, nullptr, nullptr, -1);
  B200_LAUNCH_OK();
  return 0;
}

// Same, with the all-reduce done by NVLS: `mc_acc` is the MULTICAST address of a
// symmetric int64[2][32][10] accumulator (zero-initialised, e.g. from
// torch.distributed._symmetric_memory), `local_acc` this GPU's unicast address of the
// same buffer.  Every rank adds its partial sums with multimem.red — the NVSwitch
// applies them to all GPUs' copies — then the usual flag barrier, and each rank reads
// its own copy.  The two halves of the accumulator alternate: `call_parity` must flip
// on every call (identically on all ranks); the half not in use is zeroed by the call.
// Product path only (no phase split): it needs real NVLS hardware.
extern "C" int b200_lab3_stats_nvls(const void* img, int w, int64_t y0, int64_t h_local,
                                    const int* pts, const int* class_off, int nc,
                                    const void* ws_ptrs, void* mc_acc, void* local_acc, int rank,
                                    int world, uint64_t epoch, int call_parity, double* params,
                                    long long* moments_out, b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kLab3C, "lab3: class count %d outside 1..%d", nc, kLab3C);
  B200_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world,
               "lab3: bad rank/world %d/%d", rank, world);
  B200_REQUIRE(mc_acc && local_acc, "lab3: NVLS path needs the multicast and the local address");
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  lab3_stats_allreduce_kernel<<<dim3(stats_grid_x(), nc), kStatThreads, 0, as_stream(s)>>>(
      (const uint32_t*)img, w, y0, h_local, (const int2*)pts, class_off, nc, peers, rank, world,
      epoch, params, moments_out, 3, kDefaultTimeoutNs, (unsigned long long*)mc_acc,
      (long long*)local_acc, call_parity & 1);
  B200_LAUNCH_OK();
  return 0;
}
