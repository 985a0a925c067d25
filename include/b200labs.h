/*
 * b200labs.h — C ABI of libb200labs, the native layer of the B200 lab suite.
 *
 * Every kernel entry point takes raw device pointers plus a cudaStream_t (passed
 * as void* so the header is usable from plain C and from ctypes), enqueues work
 * on that stream and returns 0 on success or a non-zero code; the text of the
 * last failure on the calling thread is available from b200_last_error().
 * Nothing here synchronises the stream unless the comment says so.
 *
 * "ref:" comments name the file:line of KoryakovDmitry/cuda-mpi-openmp whose
 * behaviour the entry point reproduces (relative to the reference root).
 */
#ifndef B200LABS_H
#define B200LABS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200_stream_t; /* cudaStream_t */

/* ---------------------------------------------------------------- generic */
const char* b200_last_error(void);
int b200_version(void);
/* Number of kernels this library has launched in the calling process. */
uint64_t b200_launch_count(void);
/* Device properties in the eight-line layout of ref: gpu_info/src/main.cu:9-16.
 * Writes at most cap bytes (NUL-terminated) and returns the full length. */
int b200_gpu_info(int device, char* buf, int cap, int extended);
/* Writes > L2-size bytes so the next kernel starts cold (bench hygiene). */
int b200_flush_l2(void* scratch, size_t bytes, b200_stream_t s);

/* Load the GPU code of one lab's kernels now instead of at their first launch (CUDA loads
 * modules lazily; the load otherwise lands inside whatever is timing that first launch, which
 * is what the reference's single-shot cudaEvent timing measures: ref lab2/src/to_plot.cu:101-109).
 * Launches nothing.  The device must be current. */
int b200_preload_lab1(void);
int b200_preload_lab2(void);
int b200_preload_lab3(void);
int b200_preload_lab5(int key_type);

/* ------------------------------------------------------------------- lab1 */
/* c[i] = a[i] - b[i] over doubles.  ref: lab1/src/main.cu:22-29 */
int b200_vecsub_f64(const double* a, const double* b, double* c, size_t n,
                    b200_stream_t s);
/* Same result, launched as <<<grid, block>>> exactly as given (the harness's
 * launch-geometry sweep).  ref: lab1/src/to_plot.cu:33-40,72 */
int b200_vecsub_f64_geom(const double* a, const double* b, double* c, size_t n,
                         int grid, int block, b200_stream_t s);

/* ------------------------------------------------------------------- lab2 */
/* Roberts-cross edge filter over RGBA8 pixels, clamp-to-edge, alpha kept.
 * ref: lab2/src/main.cu:15-52 (kernel), :68-87 (clamp/point texture).
 *
 * in/out are row-major w x h_local bands of a taller image.  below_row is the
 * row that follows the band (w pixels): NULL means the band ends the image and
 * the last row is clamped; otherwise it may point into this GPU's memory or
 * into a peer GPU's mapped memory (the row-sharded multi-GPU path reads the
 * neighbour's first row through it from inside the kernel).
 * variant: B200_ROBERTS_AUTO picks by shape. */
enum {
  B200_ROBERTS_AUTO = 0,
  B200_ROBERTS_STREAM = 1, /* register sliding window, 128-bit accesses      */
  B200_ROBERTS_TMA = 2,    /* persistent TMA tile pipeline, TMA stores          */
  B200_ROBERTS_GENERIC = 3, /* one pixel per thread, any alignment           */
  B200_ROBERTS_CLUSTER = 4  /* TMA pipeline, CTA pairs share tile edges (DSMEM) */
};
int b200_roberts(const void* in, void* out, int w, int64_t h_local,
                 const void* below_row, int variant, b200_stream_t s);
/* Launch-geometry sweep build: block (bx,by), grid (gx,gy) honoured as given.
 * ref: lab2/src/to_plot.cu:57-64,106 */
int b200_roberts_geom(const void* in, void* out, int w, int h, int bx, int by,
                      int gx, int gy, b200_stream_t s);

/* ------------------------------------------------------------------- lab3 */
#define B200_LAB3_MAX_CLASSES 32
#define B200_LAB3_MOMENTS 10 /* n, SR,SG,SB, SRR,SRG,SRB,SGG,SGB,SBB */
#define B200_LAB3_PARAMS 12  /* avg[3], inv_cov[3][3] per class      */

/* Exact integer moments of the sample pixels of each class.  pts holds (x,y)
 * int pairs, class c owns pts[class_off[c] .. class_off[c+1]).  Only points
 * whose y lies in [y0, y0+h_local) are gathered from img (a band that starts
 * at global row y0); moments is int64[nc][10] and is ACCUMULATED into (zero it
 * first).  New: the reference does this on the host, lab3/src/main.cu:102-117. */
int b200_lab3_moments(const void* img, int w, int64_t y0, int64_t h_local,
                      const int* pts, const int* class_off, int nc,
                      long long* moments, b200_stream_t s);
/* moments -> params[nc][12] = mean and inverse covariance (double), same
 * determinant/adjugate formulas as ref: lab3/src/main.cu:117,137-150. */
int b200_lab3_finalize(const long long* moments, int nc, double* params,
                       b200_stream_t s);
enum { B200_CLASSIFY_AUTO = 0, B200_CLASSIFY_DIRECT = 1, B200_CLASSIFY_LUT = 2 };
/* bytes of scratch b200_lab3_classify needs for this variant/size (0 = none) */
size_t b200_lab3_classify_scratch_bytes(int64_t npix, int variant);
/* In place: alpha := argmin_c Mahalanobis(pixel, class c), lowest index wins
 * ties.  params is device double[nc][12].  ref: lab3/src/main.cu:40-76 */
int b200_lab3_classify(void* img, int64_t npix, int nc, const double* params,
                       int variant, void* scratch, size_t scratch_bytes,
                       b200_stream_t s);
/* Launch-geometry sweep build <<<blocks, threads>>>. ref: lab3/src/to_plot.cu:76-81,162 */
int b200_lab3_classify_geom(void* img, int64_t npix, int nc,
                            const double* params, int blocks, int threads,
                            b200_stream_t s);

/* ------------------------------------------------------------------- lab5 */
enum { B200_KEY_I32 = 0, B200_KEY_F32 = 1, B200_KEY_U8 = 2, B200_KEY_U32 = 3 };
size_t b200_sort_scratch_bytes(size_t n, int key_type);
/* Ascending sort of n keys.  keys is overwritten with the sorted result; tmp
 * is a second buffer of the same size; scratch from b200_sort_scratch_bytes.
 * (The reference ships only the lab5/data files, no program: this lab is new.) */
int b200_sort(void* keys, void* tmp, size_t n, int key_type, void* scratch,
              size_t scratch_bytes, b200_stream_t s);

/* As b200_sort, with two options: `input` (may be NULL) is read by the first pass
 * instead of `keys` and left untouched (out-of-place sort without a copy); `n_dev`
 * (may be NULL) points to the actual length on the device, n then being only the
 * upper bound that sizes grids and scratch (the multi-GPU bucket's fill count). */
int b200_sort_ex(const void* input, void* keys, void* tmp, size_t n,
                 const unsigned long long* n_dev, int key_type, void* scratch,
                 size_t scratch_bytes, b200_stream_t s);
int b200_sort_devn(void* keys, void* tmp, size_t n, const unsigned long long* n_dev,
                   int key_type, void* scratch, size_t scratch_bytes, b200_stream_t s);

/* Sort for callers that cannot allocate device memory (no counterpart in the
 * reference).  Same result as b200_sort, but the histograms, tickets, look-back
 * words and the second key buffer live in a device array built into the library
 * (B200_SORT_NOALLOC_SCRATCH_BYTES on every GPU that loads it) instead of in
 * caller-provided scratch.  Meant for code that runs where cudaMalloc is off
 * limits: inside a host application whose own pool already owns the GPU's memory,
 * in a latency-critical loop that cannot take cudaMalloc/cudaFree's device-wide
 * synchronisation, or under a no-allocation-after-start-up rule.
 *  - n must not exceed B200_SORT_NOALLOC_MAX_KEYS: larger n returns an error.
 *  - input is read only; output receives the sorted keys; they may alias.
 *  - the first call on a device resolves the array's address
 *    (cudaGetSymbolAddress); issue it before any stream capture begins.
 *  - all calls on a GPU share the array: calls that may overlap there (other
 *    streams or threads) race — keep them on one stream or order them. */
#define B200_SORT_NOALLOC_MAX_KEYS (1u << 20)
#define B200_SORT_NOALLOC_SCRATCH_BYTES 4497664u
int b200_sort_noalloc(const void* input, void* output, size_t n, int key_type,
                      b200_stream_t s);

/* ------------------------------------------------- peer memory / multi-GPU */
/* One process per GPU.  Buffers are cudaMalloc'ed here, exported as 64-byte CUDA
 * IPC handles, exchanged by the caller (torch.distributed) and opened on every
 * peer; kernels then dereference the mapped peer addresses over NVLink.
 * (The reference has no multi-GPU code: everything in this section is new.) */
int b200_dev_malloc(void** p, size_t bytes); /* zero-filled */
int b200_dev_free(void* p);
int b200_ipc_export(void* p, void* handle64);
int b200_ipc_open(const void* handle64, void** p);
int b200_ipc_close(void* p);
int b200_peer_can_access(int dev, int peer);
int b200_enable_peer_access(int peer);
/* size of the per-rank symmetric workspace (flags, lab3 slots, sort samples/cursor)
 * and byte offsets of its fields ("error", "sort_cursor", "sort_splitters",
 * "sort_recv_count", "sort_samples", "flags"). */
size_t b200_symm_workspace_bytes(void);
size_t b200_symm_offset(const char* field);

/* ---- VMM + NVSwitch multicast runtime (csrc/vmm.cu; no counterpart in the reference) ------
 * Physical memory from cuMemCreate, shareable between processes as a POSIX file descriptor,
 * and multicast objects (NVLS): one multimem.* instruction on the multicast address reaches
 * every bound GPU's copy.  Sequence for a multicast buffer over `world` ranks:
 *   granularity(world) -> create -> map (local address) -> [rank 0] mc_create + export_fd ->
 *   fds to the peers (unix socket, SCM_RIGHTS) -> import_fd -> EVERY rank mc_add_device ->
 *   process barrier -> mc_bind(own handle) -> process barrier -> map(mc handle). */
int b200_vmm_query(int* vmm_supported, int* multicast_supported);
int b200_vmm_granularity(int world, size_t* gran_out);
int b200_vmm_create(size_t bytes, uint64_t* handle_out);
int b200_vmm_release(uint64_t handle);
int b200_vmm_export_fd(uint64_t handle, int* fd_out);
int b200_vmm_import_fd(int fd, uint64_t* handle_out);
int b200_vmm_map(uint64_t handle, size_t bytes, void** va_out);
int b200_vmm_unmap(void* va, size_t bytes);
int b200_mc_create(int num_devices, size_t bytes, uint64_t* mc_out);
int b200_mc_add_device(uint64_t mc);
int b200_mc_bind(uint64_t mc, uint64_t mem_handle, size_t bytes);
int b200_mc_unbind(uint64_t mc, size_t bytes);
/* Self-tests of the runtime.  pingpong: two ranks bounce a value between one uint64 flag on
 * each side (`peer_flag` is the other rank's flag through a peer mapping); out2_u64[0] = ns
 * for `iters` round trips, [1] = 1 on timeout; relaxed = 0 uses st.release.sys/ld.acquire.sys,
 * 1 the relaxed forms.  multimem: mode 0 multimem.red adds
 * value*(i+1) into every copy of element i, mode 1 multimem.ld_reduce writes the sum of all
 * copies to out[i], mode 2 multimem.st stores value+i into every copy. */
int b200_selftest_pingpong(void* my_flag, void* peer_flag, int initiator, int iters, uint64_t base,
                           int relaxed, void* out2_u64, b200_stream_t s);
/* Kernel-driven copy between this GPU (`near_buf`) and a peer mapping (`far_buf`): mode 0 / 1 stores
 * of 4 / 16 bytes per thread, 2 / 3 loads of 4 / 16 bytes, 4 TMA bulk stores of 16 KB out of
 * shared memory.  n_words (32-bit) must be a multiple of 4096. */
int b200_selftest_peer_copy(void* far_buf, void* near_buf, size_t n_words, int mode, int blocks,
                            b200_stream_t s);
int b200_selftest_multimem(void* mc_ptr, void* out, size_t n, int mode, uint64_t value,
                           b200_stream_t s);
/* Stream-ordered barrier over peer memory: signal every rank, wait for every rank.
 * ws_ptrs: device array [world] of the ranks' workspace addresses; epoch grows by
 * one per call, identically on all ranks. */
int b200_symm_barrier(const void* ws_ptrs, int rank, int world, uint64_t epoch,
                      b200_stream_t s);

/* lab3 fused: class moments of this rank's band, all-reduced across ranks inside
 * the kernel (remote stores into the peers' slots + system-scope flags), then mean
 * and inverse covariance.  phase_mask 3 = whole thing (product path); 1 / 2 run the
 * push and the wait+reduce halves separately (single-GPU testing of virtual ranks). */
int b200_lab3_stats_allreduce(const void* img, int w, int64_t y0, int64_t h_local,
                              const int* pts, const int* class_off, int nc,
                              const void* ws_ptrs, int rank, int world, uint64_t epoch,
                              double* params, long long* moments_out, int phase_mask,
                              b200_stream_t s);

/* As above, but the sums are all-reduced by NVLS: every rank issues multimem.red on the
 * multicast address `mc_acc` of a symmetric, zero-initialised int64[2][32][10] buffer
 * (in-switch reduction lands in every GPU's copy); `local_acc` is this GPU's own
 * address of the same buffer; call_parity flips on every call.  Needs NVSwitch multicast (e.g. buffers from
 * torch.distributed._symmetric_memory). */
int b200_lab3_stats_nvls(const void* img, int w, int64_t y0, int64_t h_local, const int* pts,
                         const int* class_off, int nc, const void* ws_ptrs, void* mc_acc,
                         void* local_acc, int rank, int world, uint64_t epoch, int call_parity,
                         double* params, long long* moments_out, b200_stream_t s);

/* lab3 colour table built cooperatively: each rank classifies 1/world of the 2^24
 * colours and stores the class bytes into every rank's 16 MiB table (lut_ptrs: HOST
 * array [world] of peer-mapped tables); then the image pass against the finished
 * table.  Same decisions as b200_lab3_classify, the build just scales with N. */
int b200_lab3_lut_build_allgather(void* const* lut_ptrs, int nc, const double* params,
                                  const void* ws_ptrs, int rank, int world, uint64_t epoch,
                                  int phase_mask, b200_stream_t s);
/* coarse_scratch: b200_lab3_coarse_bytes() bytes (16-byte aligned) or NULL.  When given
 * (and the image has >= 2^22 pixels) a 2^17-entry table of 4x4x8 colour cells is derived
 * from `lut`, kept in shared memory, and only pixels in cells that straddle a class
 * boundary still gather from the 16 MiB table.  Same output either way. */
int b200_lab3_apply_lut(void* img, int64_t npix, const void* lut, void* coarse_scratch,
                        b200_stream_t s);
size_t b200_lab3_coarse_bytes(void);
/* Second form of the cooperative build: colour table AND cell table from one kernel, every
 * piece leaving as a 16-byte store.  A table buffer is b200_lab3_lut2_bytes() long (2^24
 * colour bytes, then 2^17 cell bytes).  lut_ptrs: HOST array [world] of the peer-mapped
 * buffers; mc_lut (may be NULL): NVLS multicast address of the same buffers - when given,
 * each piece is ONE multimem.st that the NVSwitch replicates to every GPU.  Any world <= 8.
 * Follow with b200_lab3_apply_lut2 on the local buffer.  (No counterpart in the reference.) */
size_t b200_lab3_lut2_bytes(void);
int b200_lab3_lut_build2_allgather(void* const* lut_ptrs, void* mc_lut, int nc, const double* params,
                                   const void* ws_ptrs, int rank, int world, uint64_t epoch,
                                   int phase_mask, b200_stream_t s);
int b200_lab3_apply_lut2(void* img, int64_t npix, const void* lut_and_cells, b200_stream_t s);

/* lab5 fused sample sort, step 1: sample + exchange + splitters (one CTA). */
int b200_sort_mgpu_splitters(const void* keys, size_t n_local, int key_type,
                             int samples_per_rank, const void* ws_ptrs, int rank, int world,
                             uint64_t epoch, int phase_mask, b200_stream_t s);
/* step 1b: count this rank's keys per destination and reserve one contiguous segment
 * in every destination bucket (one remote atomic each).  Follow with the scatter
 * launched with phase bit 8 set, which then needs no remote atomics at all. */
int b200_sort_mgpu_count(const void* keys, size_t n_local, int key_type, const void* splitters,
                         size_t capacity, const void* ws_ptrs, int rank, int world,
                         b200_stream_t s);
/* step 2: partition fused with the all-to-all — keys are stored straight into the
 * destination ranks' buckets (bucket_ptrs: HOST array [world] of peer-mapped device
 * addresses), space reserved with one remote atomic per tile and destination.
 * phase_mask: 1 scatter+signal, 2 wait+publish count, 4 (with 1) local only,
 * 8 segments were reserved by b200_sort_mgpu_count (tiles use local atomics). */
int b200_sort_mgpu_scatter(const void* keys, size_t n_local, int key_type,
                           const void* splitters, void* const* bucket_ptrs,
                           void* const* cursor_ptrs, size_t capacity, const void* ws_ptrs,
                           void* const* ws_host_ptrs, int rank, int world, uint64_t epoch,
                           int phase_mask, b200_stream_t s);

/* ------------------------------------------------------------- CPU oracles */
/* Plain C++ restatements used by the tests and the CPU binaries. */
void b200_cpu_vecsub_f64(const double* a, const double* b, double* c, size_t n);
/* FMA-faithful to the contraction nvcc applies to ref: lab2/src/main.cu:30-40 */
void b200_cpu_roberts(const void* in, void* out, int w, int h);
/* Two-pass host statistics in input order, ref: lab3/src/main.cu:102-150 */
void b200_cpu_lab3_stats_ref(const void* img, int w, const int* pts,
                             const int* class_off, int nc, double* params);
/* Integer-moment statistics (what the device path computes). */
void b200_cpu_lab3_moments(const void* img, int w, int64_t y0, int64_t h_local,
                           const int* pts, const int* class_off, int nc,
                           long long* moments);
void b200_cpu_lab3_finalize(const long long* moments, int nc, double* params);
void b200_cpu_lab3_classify(void* img, int64_t npix, int nc,
                            const double* params);
void b200_cpu_sort(void* keys, size_t n, int key_type);

#ifdef __cplusplus
}
#endif
#endif /* B200LABS_H */
