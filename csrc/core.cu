// libb200labs core: error text, launch accounting, device query, L2 flush.
#include <cstdarg>
#include <mutex>

#include "common.cuh"

namespace b200 {

std::atomic<uint64_t> g_launches{0};

char* tls_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tls_error_buf(), 512, fmt, ap);
  va_end(ap);
}

const DeviceInfo& device_info() {
  static DeviceInfo cache[64];
  static std::once_flag once[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  std::call_once(once[dev], [dev] {
    DeviceInfo d;
    cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&d.l2_bytes, cudaDevAttrL2CacheSize, dev);
    cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&d.cc_minor, cudaDevAttrComputeCapabilityMinor, dev);
    if (d.sm_count <= 0) d.sm_count = 148;
    cache[dev] = d;
  });
  return cache[dev];
}

__global__ void flush_kernel(uint4* p, size_t n16) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n16; i += stride) p[i] = make_uint4(i, 0, 0, 0);
}

}  // namespace b200

using namespace b200;

extern "C" {

const char* b200_last_error(void) { return tls_error_buf(); }
int b200_version(void) { return 100; }
uint64_t b200_launch_count(void) { return g_launches.load(); }

int b200_gpu_info(int device, char* buf, int cap, int extended) {
  cudaDeviceProp p;
  B200_CUDA_OK(cudaGetDeviceProperties(&p, device));
  // The first eight lines keep the labels of ref: gpu_info/src/main.cu:9-16.
  int n = snprintf(buf, cap,
                   "Compute capability : %d.%d\n"
                   "Name : %s\n"
                   "Total Global Memory : %zu\n"
                   "Shared memory per block : %zu\n"
                   "Max threads per block : (%d, %d, %d)\n"
                   "Max block : (%d, %d, %d)\n"
                   "Total constant memory : %zu\n"
                   "Multiprocessors count : %d\n",
                   p.major, p.minor, p.name, p.totalGlobalMem, p.sharedMemPerBlock,
                   p.maxThreadsDim[0], p.maxThreadsDim[1], p.maxThreadsDim[2], p.maxGridSize[0],
                   p.maxGridSize[1], p.maxGridSize[2], p.totalConstMem, p.multiProcessorCount);
  if (extended && n < cap) {
    int cluster = 0, peer_n = 0, ndev = 0;
    cudaDeviceGetAttribute(&cluster, cudaDevAttrClusterLaunch, device);
    cudaGetDeviceCount(&ndev);
    for (int d = 0; d < ndev; ++d) {
      int ok = 0;
      if (d != device && cudaDeviceCanAccessPeer(&ok, device, d) == cudaSuccess && ok) ++peer_n;
    }
    n += snprintf(buf + n, cap - n,
                  "L2 cache size : %d\n"
                  "Shared memory per block (opt-in) : %zu\n"
                  "Shared memory per multiprocessor : %zu\n"
                  "Registers per multiprocessor : %d\n"
                  "Memory bus width : %d\n"
                  "Cluster launch : %d\n"
                  "Devices : %d\n"
                  "Peer-accessible devices : %d\n",
                  p.l2CacheSize, p.sharedMemPerBlockOptin, p.sharedMemPerMultiprocessor,
                  p.regsPerMultiprocessor, p.memoryBusWidth, cluster, ndev, peer_n);
  }
  return n;
}

int b200_flush_l2(void* scratch, size_t bytes, b200_stream_t s) {
  size_t n16 = bytes / 16;
  if (n16 == 0) return 0;
  flush_kernel<<<device_info().sm_count * 4, 256, 0, as_stream(s)>>>((uint4*)scratch, n16);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
