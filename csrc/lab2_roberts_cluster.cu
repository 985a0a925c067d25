// lab2, TMA + thread-block-cluster variant: CTA pairs share tile edges through
// distributed shared memory (DSMEM).
//
// Same persistent TMA pipeline as lab2_roberts_tma.cu, but the two CTAs of a
// cluster walk horizontally adjacent tiles in lock-step and every tile is loaded as
// an exactly 256-px-wide (1 KB pitch, fully aligned) box.  The one column the
// stencil needs from beyond a tile's right edge is the FIRST column of the
// neighbour's tile, which is already sitting in the neighbour CTA's shared memory:
// the left CTA reads it from there with ld.shared::cluster instead of fetching a
// 4-px apron through L2.  (The right CTA of the pair has no in-cluster neighbour on
// its right and loads a 4-px-wide sliver box instead.)
//
// Cross-CTA protocol per pipeline stage, all waits local, all signals remote
// (mbarrier.arrive on a mapa'd shared::cluster address):
//   right CTA: its tile has landed     -> arrive on left.nbr_ready[s]
//   left  CTA: wait nbr_ready[s], read the column over DSMEM,
//              8 reader warps          -> arrive on right.nbr_done[s]
//   right CTA: producer waits nbr_done[s] (and its own empty[s]) before refilling.
// A cluster barrier before exit keeps both CTAs' shared memory alive until the
// partner has finished reading it.
// Measured on B200 (16384^2): 3.8 TB/s against 6.4 TB/s for the plain TMA kernel —
// the apron it saves is 1.5% of the L2->SM traffic, while the per-stage handshake
// couples the two SMs' pipelines.  It is bit-exact and kept as the
// B200_ROBERTS_CLUSTER variant (cluster launch, mapa, ld.shared::cluster, remote
// mbarrier arrives), but AUTO never picks it.
// ref: lab2/src/main.cu:15-52 (same arithmetic, see roberts_math.cuh).
#include <cuda.h>

#include "common.cuh"
#include "roberts_math.cuh"

namespace b200 {

namespace cl {
constexpr int kTileW = 256, kTileH = 32;
constexpr int kBoxH = kTileH + 1;
constexpr int kPitch = kTileW * 4;                 // 1024 B
constexpr int kMainBytes = kPitch * kBoxH;         // 33792
constexpr int kSliverBytes = 16 * kBoxH;           // 4 px x 33 rows
constexpr int kStageStride = 34816;                // main + sliver, 1 KB aligned
constexpr int kStages = 4;
constexpr int kConsumerWarps = 16;
constexpr int kSubW = 128, kSubH = 4;
constexpr int kOutBytes = kSubW * 4 * kSubH;       // 2048
constexpr int kThreads = (kConsumerWarps + 1) * 32;
constexpr int kSmemBytes = kStages * kStageStride + kConsumerWarps * 2 * kOutBytes + 4 * kStages * 8 + 128;
}  // namespace cl

struct ClusterParams {
  const uint32_t* below;
  int64_t h_local;
  int w;
  int pairs_x;   // tile pairs per tile row
  int n_pairs;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(cl::kThreads, 1)
    roberts_cluster_kernel(const __grid_constant__ CUtensorMap main_map,
                           const __grid_constant__ CUtensorMap sliver_map,
                           const __grid_constant__ CUtensorMap out_map, ClusterParams P) {
  using namespace cl;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* in_tiles = smem;
  uint8_t* out_stage = smem + kStages * kStageStride;
  uint64_t* full = (uint64_t*)(out_stage + kConsumerWarps * 2 * kOutBytes);
  uint64_t* empty = full + kStages;
  uint64_t* nbr_ready = empty + kStages;  // used in the LEFT CTA (rank 0)
  uint64_t* nbr_done = nbr_ready + kStages;  // used in the RIGHT CTA (rank 1)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kConsumerWarps);
      mbar_init(&nbr_ready[s], 1);
      mbar_init(&nbr_done[s], kConsumerWarps / 2);
    }
    mbar_fence_init();
  }
  // barrier inits must be visible cluster-wide before anyone signals remotely
  cluster_arrive();
  cluster_wait();

  if (warp == kConsumerWarps) {
    // ============================ producer ===================================
    if (lane == 0) {
      prefetch_tmap(&main_map);
      int it = 0;
      for (int p = cluster_id; p < P.n_pairs; p += n_clusters, ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        if (rank == 1) {  // the left CTA may still be reading my first column
          while (!mbar_try_wait_cluster(&nbr_done[s], ph ^ 1)) {
          }
        }
        const int ty = p / P.pairs_x;
        const int tx = 2 * (p - ty * P.pairs_x) + (int)rank;
        uint8_t* dst = in_tiles + s * kStageStride;
        mbar_expect_tx(&full[s], rank == 1 ? kMainBytes + kSliverBytes : kMainBytes);
        tma_load_2d(dst, &main_map, tx * kTileW, ty * kTileH, &full[s]);
        if (rank == 1) tma_load_2d(dst + kMainBytes, &sliver_map, (tx + 1) * kTileW, ty * kTileH, &full[s]);
      }
    }
  } else {
    // ============================ consumers ==================================
    prefetch_tmap(&out_map);
    const int chalf = warp & 1, rgrp = warp >> 1;
    uint8_t* my_out = out_stage + warp * 2 * kOutBytes;
    int it = 0, n_stores = 0;
    for (int p = cluster_id; p < P.n_pairs; p += n_clusters, ++it) {
      const int s = it % kStages;
      const uint32_t ph = (it / kStages) & 1;
      const int ty = p / P.pairs_x;
      const int tx = 2 * (p - ty * P.pairs_x) + (int)rank;
      const int x = tx * kTileW + chalf * kSubW + lane * 4;
      const int64_t y0 = (int64_t)ty * kTileH + rgrp * kSubH;
      const bool has_right = x + 4 < P.w;
      const uint8_t* stage = in_tiles + s * kStageStride;
      const uint8_t* tile = stage + (rgrp * kSubH) * kPitch + (chalf * kSubW + lane * 4) * 4;
      const bool outside = y0 >= P.h_local || tx * kTileW + chalf * kSubW >= P.w;

      const bool needs_below = !outside && y0 + kSubH >= P.h_local;
      uint4 below_px = make_uint4(0, 0, 0, 0);
      uint32_t below_edge = 0;
      if (needs_below) {
        if (x < P.w) below_px = ld_stream_u4(P.below + x);
        if (lane == 31 && has_right) below_edge = ld_stream_u32(P.below + x + 4);
      }

      mbar_wait_relaxed(&full[s], ph);
      // right CTA: tell the left one that my tile (its edge column) has landed
      if (rank == 1 && warp == 0 && lane == 0)
        mbar_arrive_cluster(dsmem_addr(smem_u32(&nbr_ready[s]), 0));

      uint4 px[kSubH + 1];
      uint32_t edge[kSubH + 1];
#pragma unroll
      for (int r = 0; r <= kSubH; ++r) px[r] = *(const uint4*)(tile + r * kPitch);
      if (chalf == 0) {
        // the column right of the left half is the first column of the right half
#pragma unroll
        for (int r = 0; r <= kSubH; ++r)
          edge[r] = (lane == 31) ? *(const uint32_t*)(tile + r * kPitch + 16) : 0u;
      } else if (rank == 1) {
        // right CTA, right half: the 4-px sliver loaded next to the tile
#pragma unroll
        for (int r = 0; r <= kSubH; ++r)
          edge[r] = (lane == 31) ? *(const uint32_t*)(stage + kMainBytes + (rgrp * kSubH + r) * 16) : 0u;
      } else {
        // left CTA, right half: first column of the partner's tile, over DSMEM
        while (!mbar_try_wait_cluster(&nbr_ready[s], ph)) {
        }
        const uint32_t remote = dsmem_addr(smem_u32(stage + (rgrp * kSubH) * kPitch), 1);
#pragma unroll
        for (int r = 0; r <= kSubH; ++r) edge[r] = (lane == 31) ? ld_dsmem_u32(remote + r * kPitch) : 0u;
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(dsmem_addr(smem_u32(&nbr_done[s]), 1));
      }
      if (needs_below) {
#pragma unroll
        for (int r = 1; r <= kSubH; ++r)
          if (y0 + r == P.h_local) {
            px[r] = below_px;
            edge[r] = below_edge;
          }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      if (outside) continue;  // (the handshakes above still ran: the partner relies on them)

      if (lane == 0) tma_store_wait_read<1>();
      __syncwarp();
      uint4* ob = (uint4*)(my_out + (n_stores & 1) * kOutBytes) + lane;
      float c0 = luma_rgba(px[0].x), c1 = luma_rgba(px[0].y), c2 = luma_rgba(px[0].z),
            c3 = luma_rgba(px[0].w);
      float c4 = __shfl_down_sync(0xffffffffu, c0, 1);
      if (lane == 31) c4 = luma_rgba(edge[0]);
      if (!has_right) c4 = c3;
#pragma unroll
      for (int r = 1; r <= kSubH; ++r) {
        const float n0 = luma_rgba(px[r].x), n1 = luma_rgba(px[r].y), n2 = luma_rgba(px[r].z),
                    n3 = luma_rgba(px[r].w);
        float n4 = __shfl_down_sync(0xffffffffu, n0, 1);
        if (lane == 31) n4 = luma_rgba(edge[r]);
        if (!has_right) n4 = n3;
        uint4 o;
        o.x = roberts_pack(c0, c1, n0, n1, px[r - 1].x);
        o.y = roberts_pack(c1, c2, n1, n2, px[r - 1].y);
        o.z = roberts_pack(c2, c3, n2, n3, px[r - 1].z);
        o.w = roberts_pack(c3, c4, n3, n4, px[r - 1].w);
        ob[(r - 1) * (kSubW / 4)] = o;
        c0 = n0, c1 = n1, c2 = n2, c3 = n3, c4 = n4;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(&out_map, my_out + (n_stores & 1) * kOutBytes, tx * kTileW + chalf * kSubW, (int)y0);
        tma_store_commit();
      }
      ++n_stores;  // staging halves alternate per store ISSUED (skipped tiles store nothing)
    }
    if (lane == 0) tma_store_wait<0>();
  }
  // nobody leaves while the partner may still read this CTA's shared memory or
  // signal its barriers
  __syncwarp();
  cluster_arrive();
  cluster_wait();
}

// ---- host side ---------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool roberts_cluster_supported(const void* in, const void* out, int w, int64_t h_local) {
  return w % 4 == 0 && w >= cl::kTileW && h_local >= 1 && h_local < (1ll << 31) &&
         ((((uintptr_t)in) | ((uintptr_t)out)) & 15u) == 0;
}

int roberts_cluster_launch(const void* in, void* out, int w, int64_t h_local,
                           const void* below_row, cudaStream_t st) {
  using namespace cl;
  static EncodeTiledFn2 enc = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn2)p;
  }();
  B200_REQUIRE(enc != nullptr, "roberts_cluster: cuTensorMapEncodeTiled is not available");
  CUtensorMap main_map, sliver_map, out_map;
  const cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h_local};
  const cuuint64_t strides[1] = {(cuuint64_t)w * 4};
  const cuuint32_t estr[2] = {1, 1};
  auto make = [&](CUtensorMap* m, void* base, cuuint32_t bw, cuuint32_t bh, CUtensorMapL2promotion l2) {
    const cuuint32_t box[2] = {bw, bh};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, base, dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, l2,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r = make(&main_map, (void*)in, kTileW, kBoxH, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
  B200_REQUIRE(r == CUDA_SUCCESS, "roberts_cluster: encode(main) failed with CUresult %d", (int)r);
  r = make(&sliver_map, (void*)in, 4, kBoxH, CU_TENSOR_MAP_L2_PROMOTION_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "roberts_cluster: encode(sliver) failed with CUresult %d", (int)r);
  r = make(&out_map, out, kSubW, kSubH, CU_TENSOR_MAP_L2_PROMOTION_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "roberts_cluster: encode(out) failed with CUresult %d", (int)r);

  ClusterParams P;
  P.below = (const uint32_t*)below_row;
  P.h_local = h_local;
  P.w = w;
  const int tiles_x = (w + kTileW - 1) / kTileW;
  P.pairs_x = (tiles_x + 1) / 2;
  const int64_t n_pairs = (int64_t)P.pairs_x * ((h_local + kTileH - 1) / kTileH);
  B200_REQUIRE(n_pairs < (1ll << 30), "roberts_cluster: too many tiles");
  P.n_pairs = (int)n_pairs;
  static PerDeviceOnce attr_once;
  if (int rc = attr_once.run("roberts_cluster", [&]() -> int {
    B200_CUDA_OK(cudaFuncSetAttribute(roberts_cluster_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        return 0;
      }))
    return rc;
  int64_t clusters = device_info().sm_count / 2;
  if (clusters > n_pairs) clusters = n_pairs;
  roberts_cluster_kernel<<<(int)clusters * 2, kThreads, kSmemBytes, st>>>(main_map, sliver_map,
                                                                          out_map, P);
  B200_LAUNCH_OK();
  return 0;
}

}  // namespace b200
