// lab2: Roberts-cross edge filter.  ref: lab2/src/main.cu:15-52 — one texture
// fetch quartet per pixel, 2-D grid-stride, fixed <<<(16,16),(32,32)>>>.
//
// B200 design (no texture unit, no re-reads):
//  * STREAM  — each warp owns a 128-pixel column strip and slides down it.  A
//    lane holds 4 pixels (one 128-bit load per row), the luminance of the
//    previous row stays in registers, the x+1 neighbour comes from the next lane
//    by shuffle.  Every input byte is loaded once and its luminance is computed
//    once (the reference computes it four times); rows are prefetched U deep so
//    each warp keeps U x 512 B in flight.
//  * TMA     — persistent CTAs, mbarrier ring of self-contained TMA tile loads
//    (tile + its one-row / one-column stencil apron in a single box), output
//    staged in shared memory and written back with per-warp TMA stores
//    (lab2_roberts_tma.cu).  Fastest on large images: the AUTO choice.
//  * CLUSTER — the TMA pipeline with CTA pairs (thread-block clusters) that read
//    each other's tile edge column over DSMEM (lab2_roberts_cluster.cu).
//  * GENERIC — one pixel per thread, any width/alignment, any launch geometry
//    (the harness's block/grid sweep).
// All variants take `below_row`: the row that follows the band.  For a row-
// sharded image it points into the NEXT GPU's memory (peer-mapped), so the halo
// is pulled over NVLink by the same loads that stream the band.
#include "common.cuh"
#include "roberts_math.cuh"

namespace b200 {

int roberts_tma_launch(const void* in, void* out, int w, int64_t h_local, const void* below_row,
                       cudaStream_t st);  // lab2_roberts_tma.cu
bool roberts_tma_supported(const void* in, const void* out, int w, int64_t h_local);
int roberts_tma_preload();
int roberts_cluster_launch(const void* in, void* out, int w, int64_t h_local,
                           const void* below_row, cudaStream_t st);  // lab2_roberts_cluster.cu
bool roberts_cluster_supported(const void* in, const void* out, int w, int64_t h_local);

constexpr int kRsThreads = 256;
constexpr int kRsWarps = kRsThreads / 32;
constexpr int kStripPx = 128;  // 32 lanes x 4 px

struct RowSrc {
  const uint32_t* in;
  const uint32_t* below;  // may be a peer-GPU pointer; never null (clamp folded in)
  int64_t h_local;
  int w;
  __device__ __forceinline__ const uint32_t* row(int64_t y) const {
    return y < h_local ? in + y * (int64_t)w : below;
  }
};

template <int U>
__global__ void __launch_bounds__(kRsThreads)
    roberts_stream_kernel(RowSrc src, uint32_t* __restrict__ out, int n_strips, int rows_per_task,
                          int64_t n_tasks) {
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = (int64_t)gridDim.x * kRsWarps;
  const int w = src.w;
  for (int64_t task = (int64_t)blockIdx.x * kRsWarps + (threadIdx.x >> 5); task < n_tasks;
       task += warps_total) {
    const int strip = (int)(task % n_strips);
    const int64_t ya = (task / n_strips) * rows_per_task;
    const int64_t yb = min(ya + rows_per_task, src.h_local);
    const int x = strip * kStripPx + lane * 4;
    const bool active = x < w;           // w % 4 == 0: the whole 4-px group is in or out
    const bool has_right = x + 4 < w;    // a pixel exists to the right of this lane's group
    const bool edge_load = has_right && lane == 31;  // it lives in the next strip

    uint4 cur = make_uint4(0, 0, 0, 0);
    uint32_t cur_e = 0;
    {
      const uint32_t* r = src.row(ya);
      if (active) cur = ld_stream_u4(r + x);
      if (edge_load) cur_e = ld_stream_u32(r + x + 4);
    }
    float yc0 = luma_rgba(cur.x), yc1 = luma_rgba(cur.y), yc2 = luma_rgba(cur.z),
          yc3 = luma_rgba(cur.w);
    float yc4 = __shfl_down_sync(0xffffffffu, yc0, 1);
    if (edge_load) yc4 = luma_rgba(cur_e);
    if (!has_right) yc4 = yc3;

    for (int64_t y = ya; y < yb; y += U) {
      uint4 nx[U];
      uint32_t ne[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        nx[u] = make_uint4(0, 0, 0, 0);
        ne[u] = 0;
        if (y + u < yb) {
          const uint32_t* r = src.row(y + u + 1);
          if (active) nx[u] = ld_stream_u4(r + x);
          if (edge_load) ne[u] = ld_stream_u32(r + x + 4);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (y + u < yb) {  // warp-uniform
          float yn0 = luma_rgba(nx[u].x), yn1 = luma_rgba(nx[u].y), yn2 = luma_rgba(nx[u].z),
                yn3 = luma_rgba(nx[u].w);
          float yn4 = __shfl_down_sync(0xffffffffu, yn0, 1);
          if (edge_load) yn4 = luma_rgba(ne[u]);
          if (!has_right) yn4 = yn3;
          uint4 o;
          o.x = roberts_pack(yc0, yc1, yn0, yn1, cur.x);
          o.y = roberts_pack(yc1, yc2, yn1, yn2, cur.y);
          o.z = roberts_pack(yc2, yc3, yn2, yn3, cur.z);
          o.w = roberts_pack(yc3, yc4, yn3, yn4, cur.w);
          if (active) st_stream_u4(out + (y + u) * (int64_t)w + x, o);
          cur = nx[u];
          yc0 = yn0, yc1 = yn1, yc2 = yn2, yc3 = yn3, yc4 = yn4;
        }
      }
    }
  }
}

// One pixel per thread, 2-D grid-stride, clamped coordinates; any geometry.
__global__ void roberts_generic_kernel(RowSrc src, uint32_t* __restrict__ out) {
  const int w = src.w;
  const int64_t h = src.h_local;
  const int sx = blockDim.x * gridDim.x;
  const int sy = blockDim.y * gridDim.y;
  for (int64_t y = blockDim.y * blockIdx.y + threadIdx.y; y < h; y += sy) {
    const uint32_t* r0 = src.row(y);
    const uint32_t* r1 = src.row(y + 1);
    for (int x = blockDim.x * blockIdx.x + threadIdx.x; x < w; x += sx) {
      const int x1 = min(x + 1, w - 1);
      uint32_t p00 = __ldg(r0 + x), p10 = __ldg(r0 + x1);
      uint32_t p01 = __ldg(r1 + x), p11 = __ldg(r1 + x1);
      out[y * (int64_t)w + x] =
          roberts_pack(luma_rgba(p00), luma_rgba(p10), luma_rgba(p01), luma_rgba(p11), p00);
    }
  }
}

static RowSrc make_src(const void* in, int w, int64_t h_local, const void* below_row) {
  RowSrc s;
  s.in = (const uint32_t*)in;
  s.h_local = h_local;
  s.w = w;
  // no row below: clamp-to-edge == re-read the band's own last row
  s.below = below_row ? (const uint32_t*)below_row : s.in + (h_local - 1) * (int64_t)w;
  return s;
}

static bool stream_supported(const void* in, const void* out, const void* below, int w) {
  return (w % 4 == 0) && ((((uintptr_t)in) | ((uintptr_t)out) | ((uintptr_t)below)) & 15u) == 0;
}

static int roberts_stream_launch(const RowSrc& src, void* out, cudaStream_t st) {
  const int sms = device_info().sm_count;
  const int n_strips = (src.w + kStripPx - 1) / kStripPx;
  // enough tasks to fill every resident warp ~4 times over, but long enough
  // strips that the one re-read halo row per task stays negligible
  const int64_t resident_warps = (int64_t)sms * 4 * kRsWarps;
  int64_t rows_per_task = 128;
  while (rows_per_task > 8 &&
         n_strips * ((src.h_local + rows_per_task - 1) / rows_per_task) < resident_warps * 4)
    rows_per_task >>= 1;
  const int64_t n_tasks = n_strips * ((src.h_local + rows_per_task - 1) / rows_per_task);
  int64_t blocks = (n_tasks + kRsWarps - 1) / kRsWarps;
  if (blocks > (int64_t)sms * 4) blocks = (int64_t)sms * 4;
  roberts_stream_kernel<4><<<(int)blocks, kRsThreads, 0, st>>>(
      src, (uint32_t*)out, n_strips, (int)rows_per_task, n_tasks);
  B200_LAUNCH_OK();
  return 0;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_roberts(const void* in, void* out, int w, int64_t h_local,
                            const void* below_row, int variant, b200_stream_t s) {
  B200_REQUIRE(w > 0 && h_local > 0, "roberts: empty image %d x %lld", w, (long long)h_local);
  B200_REQUIRE(in != out, "roberts: in-place operation is not supported");
  cudaStream_t st = as_stream(s);
  RowSrc src = make_src(in, w, h_local, below_row);
  const bool vec_ok = stream_supported(in, out, src.below, w);
  if (variant == B200_ROBERTS_AUTO) {
    // measured on B200 (profiles/): TMA 0.94 of the copy roofline on large images,
    // STREAM 0.67; tiny images are launch-latency bound either way
    if (roberts_tma_supported(in, out, w, h_local) && (int64_t)w * h_local >= (1 << 16))
      variant = B200_ROBERTS_TMA;
    else
      variant = vec_ok ? B200_ROBERTS_STREAM : B200_ROBERTS_GENERIC;
  }
  switch (variant) {
    case B200_ROBERTS_STREAM:
      B200_REQUIRE(vec_ok, "roberts: STREAM variant needs w %% 4 == 0 and 16-byte aligned buffers");
      return roberts_stream_launch(src, out, st);
    case B200_ROBERTS_TMA:
      B200_REQUIRE(roberts_tma_supported(in, out, w, h_local),
                   "roberts: TMA variant needs w %% 4 == 0, 16-byte aligned buffers, w >= 256");
      return roberts_tma_launch(in, out, w, h_local, src.below, st);
    case B200_ROBERTS_CLUSTER:
      B200_REQUIRE(roberts_cluster_supported(in, out, w, h_local),
                   "roberts: CLUSTER variant needs w %% 4 == 0, 16-byte aligned buffers, w >= 256");
      return roberts_cluster_launch(in, out, w, h_local, src.below, st);
    case B200_ROBERTS_GENERIC: {
      dim3 block(32, 8);
      const int64_t gy = (h_local + 7) / 8;
      dim3 grid((unsigned)((w + 31) / 32 < 1024 ? (w + 31) / 32 : 1024),
                (unsigned)(gy < 1024 ? gy : 1024));
      roberts_generic_kernel<<<grid, block, 0, st>>>(src, (uint32_t*)out);
      B200_LAUNCH_OK();
      return 0;
    }
    default:
      B200_REQUIRE(false, "roberts: unknown variant %d", variant);
  }
  return 0;
}

extern "C" int b200_roberts_geom(const void* in, void* out, int w, int h, int bx, int by, int gx,
                                 int gy, b200_stream_t s) {
  B200_REQUIRE(w > 0 && h > 0, "roberts_geom: empty image %d x %d", w, h);
  B200_REQUIRE(bx > 0 && by > 0 && gx > 0 && gy > 0,
               "roberts_geom: launch dims must be positive (block %d x %d, grid %d x %d)", bx, by,
               gx, gy);
  RowSrc src = make_src(in, w, h, nullptr);
  roberts_generic_kernel<<<dim3(gx, gy), dim3(bx, by), 0, as_stream(s)>>>(src, (uint32_t*)out);
  B200_LAUNCH_OK();
  return 0;
}

extern "C" int b200_preload_lab2(void) {
  if (int rc = preload_kernel(roberts_generic_kernel)) return rc;
  if (int rc = preload_kernel(roberts_stream_kernel<4>)) return rc;
  return roberts_tma_preload();
}
