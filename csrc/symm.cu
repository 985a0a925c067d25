// Peer-memory runtime: the plumbing under the fused multi-GPU kernels.
//
// One process per GPU.  Each process cudaMalloc's its symmetric buffers, exports
// CUDA IPC handles, and maps every peer's buffer into its own address space; a
// kernel on GPU i then dereferences GPU j's memory directly over NVLink 5 /
// NVSwitch (plain ld/st/atom/red on the mapped address).  The handles travel
// between processes through torch.distributed (parallel/symm.py) — that is the
// only thing the "communication library" is used for on the product path.
//
// Cross-GPU ordering uses one uint64 flag per (receiver, sender) pair holding a
// monotonically increasing epoch: the sender publishes its data, fences at
// system scope and stores the epoch with st.release.sys; the receiver spins with
// ld.acquire.sys.  Every spin has a wall-clock timeout (globaltimer) that raises
// an error word instead of hanging the GPU.
// (No counterpart in the reference: it has no multi-GPU code at all.)
#include "common.cuh"
#include "symm.cuh"

namespace b200 {

__global__ void symm_barrier_kernel(SymmPeers peers, int rank, int world, uint64_t epoch,
                                    unsigned long long timeout_ns) {
  // thread p signals peer p, then waits for peer p's signal
  const int p = threadIdx.x;
  if (p < world) {
    __threadfence_system();
    signal_peer(peers, p, rank, epoch);
    wait_peer(peers, rank, p, epoch, timeout_ns);
  }
}

}  // namespace b200

using namespace b200;

// ---- self-test: what a kernel gets out of NVLink, by access width ------------------------
// mode 0: 4-byte stores (a warp writes 128 contiguous bytes), 1: 16-byte stores, 2: 4-byte
// loads, 3: 16-byte loads, 4: 16 KB cp.async.bulk (TMA) stores out of shared memory.
// `far` is a peer mapping, `near` local memory; n in 32-bit words, a multiple of 4096.
namespace b200 {
__global__ void __launch_bounds__(256) peer_copy_kernel(uint32_t* far, uint32_t* near, size_t n, int mode) {
  __shared__ __align__(128) uint32_t stage[4096];
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
  if (mode == 0) {
    for (size_t i = tid; i < n; i += nth) far[i] = near[i];
  } else if (mode == 1) {
    for (size_t i = tid; i < n / 4; i += nth) reinterpret_cast<uint4*>(far)[i] = reinterpret_cast<const uint4*>(near)[i];
  } else if (mode == 2) {
    for (size_t i = tid; i < n; i += nth) near[i] = far[i];
  } else if (mode == 3) {
    for (size_t i = tid; i < n / 4; i += nth) reinterpret_cast<uint4*>(near)[i] = reinterpret_cast<const uint4*>(far)[i];
  } else {
    for (size_t c = blockIdx.x; c < n / 4096; c += gridDim.x) {
      for (int j = threadIdx.x; j < 1024; j += 256)
        reinterpret_cast<uint4*>(stage)[j] = reinterpret_cast<const uint4*>(near + c * 4096)[j];
      fence_proxy_async_smem();
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 16384;" ::"l"(far + c * 4096),
                     "r"(smem_u32(stage))
                     : "memory");
        tma_store_commit();
        tma_store_wait_read<0>();  // the staging block may be refilled
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) tma_store_wait<0>();
  }
}
}  // namespace b200

extern "C" {

int b200_dev_malloc(void** p, size_t bytes) {
  B200_CUDA_OK(cudaMalloc(p, bytes ? bytes : 256));
  B200_CUDA_OK(cudaMemset(*p, 0, bytes ? bytes : 256));
  return 0;
}
int b200_dev_free(void* p) {
  B200_CUDA_OK(cudaFree(p));
  return 0;
}
int b200_ipc_export(void* p, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  return 0;
}
int b200_ipc_open(const void* handle64, void** p) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA_OK(cudaIpcOpenMemHandle(p, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int b200_ipc_close(void* p) {
  B200_CUDA_OK(cudaIpcCloseMemHandle(p));
  return 0;
}
int b200_peer_can_access(int dev, int peer) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) return 0;
  return ok;
}
// same-process peers (several GPUs driven by one process, e.g. tests)
int b200_enable_peer_access(int peer) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  B200_CUDA_OK(e);
  return 0;
}

// Signal every peer and wait for every peer: a full barrier on the stream.
// ws_ptrs: device array [world] of each rank's workspace base (SymmWorkspace).
int b200_symm_barrier(const void* ws_ptrs, int rank, int world, uint64_t epoch,
                      b200_stream_t s) {
  B200_REQUIRE(world >= 1 && world <= kMaxWorld, "symm_barrier: world %d outside 1..%d", world,
               kMaxWorld);
  SymmPeers peers{(SymmWorkspace* const*)ws_ptrs};
  symm_barrier_kernel<<<1, 32, 0, as_stream(s)>>>(peers, rank, world, epoch, kDefaultTimeoutNs);
  B200_LAUNCH_OK();
  return 0;
}

int b200_selftest_peer_copy(void* far, void* near, size_t n_words, int mode, int blocks, b200_stream_t s) {
  B200_REQUIRE(mode >= 0 && mode <= 4 && n_words % 4096 == 0 && blocks > 0, "peer_copy selftest: bad arguments");
  peer_copy_kernel<<<blocks, 256, 0, as_stream(s)>>>((uint32_t*)far, (uint32_t*)near, n_words, mode);
  B200_LAUNCH_OK();
  return 0;
}

size_t b200_symm_workspace_bytes(void) { return sizeof(SymmWorkspace); }

}  // extern "C"
