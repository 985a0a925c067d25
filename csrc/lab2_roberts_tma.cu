// lab2, TMA variant: persistent CTAs, one per SM.  A producer warp keeps a ring
// of kStages TMA tile loads in flight (cp.async.bulk.tensor -> UTMALDG); sixteen
// consumer warps wait on the tile's mbarrier, run the stencil out of shared
// memory and hand their 128 px x 4 row result block to the TMA store engine
// (UTMASTG) from a private double-buffered staging area, so no CTA-wide barrier
// is ever taken: the only synchronisation is the full/empty mbarrier pair per
// stage and the per-thread bulk-group wait of the storing lane.
//
// A tile is 256 x 32 output pixels and is loaded as a self-contained
// (256+4) x (32+1) box (element type u64 = 2 px so the 260-px row fits the
// 256-element box limit): the +1 row / +1 column the stencil looks at come with
// the tile, zero-filled by TMA outside the image and patched to clamp-to-edge /
// the neighbour GPU's halo row by the consumer.  TMA stores clip partial tiles.
// Tuning notes (each measured on B200, 16384^2 and 65536^2): tiles in row-major
// order beat a panel order meant to keep halo rows in L2 (-13%); 128-px warp blocks
// beat 124-px ones that need no edge column but lose sector-aligned stores (-25%);
// converting the edge column once per block and cvt.rzi.sat.u8 for clamp+truncate
// lowered the instruction count but not the time.  ncu: 76% issue slots busy at
// 6 TB/s, no DRAM re-reads (1.074 GB read for 1.074 GB of input).
// ref: lab2/src/main.cu:15-52 (same arithmetic, see roberts_math.cuh).
#include <cuda.h>

#include <cstdlib>

#include "common.cuh"
// ALU / FMA pipe balance (ncu: alu 52%, fma 39% of peak).  Additions that can be spelled
// fma(x, 1, c) with the same single rounding can be issued to the FMA pipe instead
// (B200_LUMA_SUB_ON_FMA_PIPE, roberts_pack<true>).  Measured two ways on 65536^2:
//   * one launch at a time, synchronised on both sides (`tools/lab2_layout.py`, interleaved on one
//     box): all additions on the ALU pipe 5.94 ms; red "- 2^23" on the FMA pipe 5.78; red and blue
//     5.75; red + the final round-toward-zero addition 5.74 (but 1-2% slower on bands up to 4 GiB);
//   * bench.py's back-to-back steps, where the 16 GiB band keeps the GPU at its power cap: 5.81-5.88
//     for every one of these variants, in whatever order they ran (the drift of a warming GPU, 1-2%,
//     is larger than any difference between them), and 5.94-6.00 for red + rz.
// The sustained figure is the one that counts, so the default stays with plain additions.
// Packed fp32 (FFMA2 / FADD2 / FMUL2, B200_TMA_PACKED: two pixels per instruction, 27% fewer
// instructions, bit-identical) is slower in both modes: 6.26 ms / 0.411 ms at 16384^2 against 0.341.
#ifndef B200_LUMA_SUB_ON_FMA_PIPE
#define B200_LUMA_SUB_ON_FMA_PIPE 0
#endif
#include "roberts_math.cuh"

namespace b200 {

// Tile shape knobs.  Measured on B200 (GB/s at 65536^2 / 16384^2): the default 32-row tiles,
// 4-row warp blocks, 4 stages, 2 staging blocks 5867 / 6632; 64-row tiles with 8-row blocks
// (less halo, less repeated luminance), 2 stages, 1 staging block 5566 / 5492; 32-row tiles
// with 8-row blocks (8 consumer warps) 5444 / 6003; 48-row tiles, 8-row blocks, 3 stages
// 5486 / 5426.  Fewer, fatter warps lose more to latency than they save in instructions.
#ifndef B200_TMA_TILE_H
#define B200_TMA_TILE_H 32
#endif
#ifndef B200_TMA_SUB_H
#define B200_TMA_SUB_H 4
#endif
#ifndef B200_TMA_STAGES
#define B200_TMA_STAGES 4
#endif
#ifndef B200_TMA_OUTBUFS
#define B200_TMA_OUTBUFS 2
#endif
#ifndef B200_TMA_PACKED
#define B200_TMA_PACKED 0  // packed fp32 arithmetic (FFMA2 / FADD2 / FMUL2), two pixels per instruction: slower
#endif
constexpr int kTileW = 256;               // output px per tile row
constexpr int kTileH = B200_TMA_TILE_H;   // output rows per tile
constexpr int kBoxW = kTileW + 4;         // loaded px per row (260 -> 130 u64)
constexpr int kBoxH = kTileH + 1;         // loaded rows
constexpr int kPitch = kBoxW * 4;         // 1040 B, 16-byte multiple
constexpr int kStageBytes = kPitch * kBoxH;  // 34320
constexpr int kStageStride = (kStageBytes + 127) / 128 * 128;  // 128-byte aligned slot
constexpr int kStages = B200_TMA_STAGES;
constexpr int kSubW = 128, kSubH = B200_TMA_SUB_H;  // per-warp output block
constexpr int kConsumerWarps = 2 * (kTileH / kSubH);  // 2 column halves x row groups of kSubH
constexpr int kOutBufs = B200_TMA_OUTBUFS;  // staging blocks per warp handed to the TMA store in turn
constexpr int kOutBytes = kSubW * 4 * kSubH;  // 2048
constexpr int kTmaThreads = (kConsumerWarps + 1) * 32;
constexpr int kSmemBytes =
    kStages * kStageStride + kConsumerWarps * kOutBufs * kOutBytes + 2 * kStages * 8 + 128;

struct TmaParams {
  const uint32_t* below;  // row after the band (maybe on a peer GPU), never null
  int64_t h_local;
  int w;
  int tiles_x;
  int64_t n_tiles;
  int l2_hints;  // evict-first for streamed rows, evict-last for the shared halo row
};

__global__ void __launch_bounds__(kTmaThreads, 1)
    roberts_tma_kernel(const __grid_constant__ CUtensorMap in_map,
                       const __grid_constant__ CUtensorMap halo_map,
                       const __grid_constant__ CUtensorMap out_map, TmaParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* in_tiles = smem;
  uint8_t* out_stage = smem + kStages * kStageStride;
  uint64_t* full = (uint64_t*)(out_stage + kConsumerWarps * kOutBufs * kOutBytes);
  uint64_t* empty = full + kStages;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kConsumerWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ===================== producer warp: one lane feeds the ring =============
    if (lane == 0) {
      prefetch_tmap(&in_map);
      prefetch_tmap(&halo_map);
      // The tile's 33rd row is the first row of the tile below, which another CTA loads
      // ~67 MB of traffic later on a 65536-px-wide image; ncu shows 4% extra DRAM reads
      // from that.  The row is fetched by its own TMA op so that L2 eviction hints can
      // be attached (B200LABS_TMA_L2HINTS): evict-first on the streamed rows + evict-last
      // on the halo row measured 12% SLOWER than no hints, evict-last on the halo row
      // alone 1-2% slower, so the default is none (the two-op load itself costs nothing).
      const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();
      int it = 0;
      for (int64_t t = blockIdx.x; t < P.n_tiles; t += gridDim.x, ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        const int tx = (int)(t % P.tiles_x);
        const int64_t ty = t / P.tiles_x;
        mbar_expect_tx(&full[s], kStageBytes);
        // coordinates: innermost in u64 elements (2 px), then rows
        uint8_t* dst = in_tiles + s * kStageStride;
        const int cx = tx * (kTileW / 2), cy = (int)(ty * kTileH);
        if (P.l2_hints == 1) {
          tma_load_2d_hint(dst, &in_map, cx, cy, &full[s], pol_stream);
          tma_load_2d_hint(dst + kTileH * kPitch, &halo_map, cx, cy + kTileH, &full[s], pol_keep);
        } else if (P.l2_hints == 2) {
          tma_load_2d(dst, &in_map, cx, cy, &full[s]);
          tma_load_2d_hint(dst + kTileH * kPitch, &halo_map, cx, cy + kTileH, &full[s], pol_keep);
        } else {
          tma_load_2d(dst, &in_map, cx, cy, &full[s]);
          tma_load_2d(dst + kTileH * kPitch, &halo_map, cx, cy + kTileH, &full[s]);
        }
      }
    }
    return;
  }

  // ========================= consumer warps ====================================
  prefetch_tmap(&out_map);
  const int chalf = warp & 1;       // which 128-px half of the tile
  const int rgrp = warp >> 1;       // rows rgrp*4 .. rgrp*4+3
  uint8_t* my_out = out_stage + warp * kOutBufs * kOutBytes;
  int it = 0;
  // The staging half is picked by the number of stores this warp has ISSUED, not by the
  // tile counter: a warp skips tiles whose block lies outside the image without storing,
  // and "wait until at most one store group is still reading" only protects the buffer
  // used two STORES ago.
  int n_stores = 0;
  for (int64_t t = blockIdx.x; t < P.n_tiles; t += gridDim.x, ++it) {
    const int s = it % kStages;
    const uint32_t ph = (it / kStages) & 1;
    const int tx = (int)(t % P.tiles_x);
    const int64_t ty = t / P.tiles_x;
    const int x = tx * kTileW + chalf * kSubW + lane * 4;  // global column of this lane
    const int64_t y0 = ty * kTileH + rgrp * kSubH;         // global row of this warp's block
    const bool has_right = x + 4 < P.w;
    const uint8_t* tile = in_tiles + s * kStageStride + (rgrp * kSubH) * kPitch +
                          (chalf * kSubW + lane * 4) * 4;

    if (y0 >= P.h_local || tx * kTileW + chalf * kSubW >= P.w) {
      // block entirely outside the image (partial tile): just release the stage
      mbar_wait(&full[s], ph);
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      continue;
    }
    // the halo row of the band's last row is not in the tensor: fetch it early
    const bool needs_below = y0 + kSubH >= P.h_local;
    uint4 below_px = make_uint4(0, 0, 0, 0);
    uint32_t below_edge = 0;
    if (needs_below) {
      if (x < P.w) below_px = ld_stream_u4(P.below + x);
      if (lane == 31 && has_right) below_edge = ld_stream_u32(P.below + x + 4);
    }

    mbar_wait_relaxed(&full[s], ph);

    uint4 px[kSubH + 1];
    uint32_t edge[kSubH + 1];
#pragma unroll
    for (int r = 0; r <= kSubH; ++r) {
      px[r] = *(const uint4*)(tile + r * kPitch);
      edge[r] = (lane == 31) ? *(const uint32_t*)(tile + r * kPitch + 16) : 0u;
    }
    if (needs_below) {
#pragma unroll
      for (int r = 1; r <= kSubH; ++r)
        if (y0 + r == P.h_local) {
          px[r] = below_px;
          edge[r] = below_edge;
        }
    }
    // all shared-memory reads of this stage are in registers: release it
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);

    // this staging block was handed to the TMA store kOutBufs stores ago
    if (lane == 0) tma_store_wait_read<kOutBufs - 1>();
    __syncwarp();
    uint4* ob = (uint4*)(my_out + (n_stores % kOutBufs) * kOutBytes) + lane;

#if B200_TMA_PACKED
    // Two pixels per instruction: a lane's pixels 0 and 2 travel in one register pair, 1 and 3
    // in another, and {2, 4} (4 = the right neighbour's pixel 0) in a third, so the four
    // outputs are two calls whose operands are whole pairs.
    f32x2 ca = luma_rgba2(px[0].x, px[0].z), cb = luma_rgba2(px[0].y, px[0].w), cc;
    {
      float c0, c2, c1, c3;
      unpack2(ca, c0, c2);
      unpack2(cb, c1, c3);
      float c4 = __shfl_down_sync(0xffffffffu, c0, 1);
      if (lane == 31) c4 = luma_rgba(edge[0]);
      if (!has_right) c4 = c3;
      cc = pack2(c2, c4);
    }
#pragma unroll
    for (int r = 1; r <= kSubH; ++r) {
      const f32x2 na = luma_rgba2(px[r].x, px[r].z), nb = luma_rgba2(px[r].y, px[r].w);
      float n0, n2, n1, n3;
      unpack2(na, n0, n2);
      unpack2(nb, n1, n3);
      float n4 = __shfl_down_sync(0xffffffffu, n0, 1);
      if (lane == 31) n4 = luma_rgba(edge[r]);
      if (!has_right) n4 = n3;
      const f32x2 nc = pack2(n2, n4);
      uint4 o;
      roberts_pack2(ca, cb, na, nb, px[r - 1].x, px[r - 1].z, o.x, o.z);
      roberts_pack2(cb, cc, nb, nc, px[r - 1].y, px[r - 1].w, o.y, o.w);
      ob[(r - 1) * (kSubW / 4)] = o;
      ca = na, cb = nb, cc = nc;
    }
#else
    float c0 = luma_rgba(px[0].x), c1 = luma_rgba(px[0].y), c2 = luma_rgba(px[0].z),
          c3 = luma_rgba(px[0].w);
    float c4 = __shfl_down_sync(0xffffffffu, c0, 1);
    if (lane == 31) c4 = luma_rgba(edge[0]);
    if (!has_right) c4 = c3;
#pragma unroll
    for (int r = 1; r <= kSubH; ++r) {
      float n0 = luma_rgba(px[r].x), n1 = luma_rgba(px[r].y), n2 = luma_rgba(px[r].z),
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
#endif
    fence_proxy_async_smem();  // generic-proxy writes -> visible to the TMA engine
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(&out_map, my_out + (n_stores % kOutBufs) * kOutBytes, tx * kTileW + chalf * kSubW, (int)y0);
      tma_store_commit();
    }
    ++n_stores;
  }
  if (lane == 0) tma_store_wait<0>();
}

// ---- host side ---------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) !=
            cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

bool roberts_tma_supported(const void* in, const void* out, int w, int64_t h_local) {
  return w % 4 == 0 && w >= kTileW && h_local >= 1 && h_local < (1ll << 31) &&
         ((((uintptr_t)in) | ((uintptr_t)out)) & 15u) == 0;
}

int roberts_tma_launch(const void* in, void* out, int w, int64_t h_local, const void* below_row,
                       cudaStream_t st) {
  EncodeTiledFn enc = encode_fn();
  B200_REQUIRE(enc != nullptr, "roberts_tma: cuTensorMapEncodeTiled is not available");
  CUtensorMap in_map, halo_map, out_map;
  {
    cuuint64_t dims[2] = {(cuuint64_t)(w / 2), (cuuint64_t)h_local};
    cuuint64_t strides[1] = {(cuuint64_t)w * 4};
    cuuint32_t box[2] = {kBoxW / 2, kTileH};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&in_map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, (void*)in, dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "roberts_tma: encode(in) failed with CUresult %d", (int)r);
    cuuint32_t hbox[2] = {kBoxW / 2, 1};  // the row below the tile, fetched with its own L2 policy
    r = enc(&halo_map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, (void*)in, dims, strides, hbox, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "roberts_tma: encode(halo) failed with CUresult %d", (int)r);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h_local};
    cuuint64_t strides[1] = {(cuuint64_t)w * 4};
    cuuint32_t box[2] = {kSubW, kSubH};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&out_map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, out, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "roberts_tma: encode(out) failed with CUresult %d", (int)r);
  }
  TmaParams P;
  P.below = (const uint32_t*)below_row;
  P.h_local = h_local;
  P.w = w;
  P.tiles_x = (w + kTileW - 1) / kTileW;
  P.n_tiles = (int64_t)P.tiles_x * ((h_local + kTileH - 1) / kTileH);
  static const int hints = [] {
    const char* e = getenv("B200LABS_TMA_L2HINTS");  // 0 none (default), 1 both, 2 halo row only
    return e ? atoi(e) : 0;
  }();
  P.l2_hints = hints;
  static PerDeviceOnce attr_once;
  if (int rc = attr_once.run("roberts_tma", [&]() -> int {
        B200_CUDA_OK(cudaFuncSetAttribute(roberts_tma_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        return 0;
      }))
    return rc;
  int64_t grid = device_info().sm_count;
  if (grid > P.n_tiles) grid = P.n_tiles;
  roberts_tma_kernel<<<(int)grid, kTmaThreads, kSmemBytes, st>>>(in_map, halo_map, out_map, P);
  B200_LAUNCH_OK();
  return 0;
}

int roberts_tma_preload() { return preload_kernel(roberts_tma_kernel); }

}  // namespace b200
