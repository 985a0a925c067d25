// Virtual-memory-management (VMM) and NVSwitch-multicast runtime.
//
// The IPC path (symm.cu) maps cudaMalloc memory into peers; that memory cannot be
// bound to a multicast object.  This file is the second allocator of the runtime:
// physical memory made with cuMemCreate (exportable as a POSIX file descriptor),
// mapped wherever the caller wants, and — on NVSwitch systems — bound to a
// multicast object so that ONE multimem.* instruction reaches every GPU's copy
// (NVLS: `multimem.red` adds inside the switch, `multimem.st` broadcasts,
// `multimem.ld_reduce` reads the sum of all copies).
//
// Process model: one process per GPU.  File descriptors travel between the
// processes over unix sockets (SCM_RIGHTS) — parallel/vmm.py does that part;
// everything that touches the driver is here.  The driver API is reached through
// cudaGetDriverEntryPoint so the library still loads on a box without libcuda
// (the CPU test tier loads it for the oracles).
// (No counterpart in the reference: it has no multi-GPU code at all.)
#include <cuda.h>

#include "common.cuh"

namespace b200 {
namespace {

struct DriverApi {
  CUresult (*getErrorString)(CUresult, const char**);
  CUresult (*deviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*memGetAllocationGranularity)(size_t*, const CUmemAllocationProp*,
                                          CUmemAllocationGranularity_flags);
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                        unsigned long long);
  CUresult (*memRelease)(CUmemGenericAllocationHandle);
  CUresult (*memExportToShareableHandle)(void*, CUmemGenericAllocationHandle,
                                         CUmemAllocationHandleType, unsigned long long);
  CUresult (*memImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*,
                                           CUmemAllocationHandleType);
  CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*memAddressFree)(CUdeviceptr, size_t);
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle,
                     unsigned long long);
  CUresult (*memUnmap)(CUdeviceptr, size_t);
  CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*multicastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*multicastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*multicastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle,
                               size_t, size_t, unsigned long long);
  CUresult (*multicastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*multicastGetGranularity)(size_t*, const CUmulticastObjectProp*,
                                      CUmulticastGranularity_flags);
  bool ok;
};

template <typename F>
bool load(const char* name, F& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p) {
    cudaGetLastError();
    fn = nullptr;
    return false;
  }
  fn = (F)p;
  return true;
}

const DriverApi& drv() {
  static DriverApi api = [] {
    DriverApi a{};
    bool ok = true;
    ok &= load("cuGetErrorString", a.getErrorString);
    ok &= load("cuDeviceGetAttribute", a.deviceGetAttribute);
    ok &= load("cuMemGetAllocationGranularity", a.memGetAllocationGranularity);
    ok &= load("cuMemCreate", a.memCreate);
    ok &= load("cuMemRelease", a.memRelease);
    ok &= load("cuMemExportToShareableHandle", a.memExportToShareableHandle);
    ok &= load("cuMemImportFromShareableHandle", a.memImportFromShareableHandle);
    ok &= load("cuMemAddressReserve", a.memAddressReserve);
    ok &= load("cuMemAddressFree", a.memAddressFree);
    ok &= load("cuMemMap", a.memMap);
    ok &= load("cuMemUnmap", a.memUnmap);
    ok &= load("cuMemSetAccess", a.memSetAccess);
    // multicast entry points may be missing on an old driver: that only disables NVLS
    load("cuMulticastCreate", a.multicastCreate);
    load("cuMulticastAddDevice", a.multicastAddDevice);
    load("cuMulticastBindMem", a.multicastBindMem);
    load("cuMulticastUnbind", a.multicastUnbind);
    load("cuMulticastGetGranularity", a.multicastGetGranularity);
    a.ok = ok;
    return a;
  }();
  return api;
}

const char* cu_text(CUresult r) {
  const char* s = nullptr;
  if (drv().getErrorString && drv().getErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown driver error";
}

#define B200_CU_OK(call)                                                                   \
  do {                                                                                     \
    CUresult r_ = (call);                                                                  \
    if (r_ != CUDA_SUCCESS) {                                                              \
      set_error("%s:%d: %s -> %d (%s)", __FILE__, __LINE__, #call, (int)r_, cu_text(r_)); \
      return 1;                                                                            \
    }                                                                                      \
  } while (0)

int current_device(int* dev) {
  B200_CUDA_OK(cudaFree(nullptr));  // make sure the primary context exists and is current
  B200_CUDA_OK(cudaGetDevice(dev));
  return 0;
}

CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp p{};
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

CUmulticastObjectProp mc_prop(int num_devices, size_t bytes) {
  CUmulticastObjectProp p{};
  p.numDevices = (unsigned)num_devices;
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int attr(int dev, CUdevice_attribute a) {
  int v = 0;
  if (!drv().deviceGetAttribute || drv().deviceGetAttribute(&v, a, dev) != CUDA_SUCCESS) return 0;
  return v;
}

}  // namespace
}  // namespace b200

namespace b200 {

// ---- self-tests: ping-pong latency over peer memory, multimem red / ld_reduce / st ----
// Two ranks bounce an increasing value between one flag on each side; the initiator
// reports the elapsed time for `iters` round trips.  out[0] = ns, out[1] = 1 on timeout.
// RELAXED = false: st.release.sys / ld.acquire.sys (what the fused kernels use to publish
// data); true: st.relaxed.sys / ld.relaxed.sys (the bare NVLink round trip).
template <bool RELAXED>
__global__ void pingpong_kernel(uint64_t* my_flag, uint64_t* peer_flag, int initiator, int iters,
                                uint64_t base, unsigned long long* out,
                                unsigned long long timeout_ns) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  auto put = [](uint64_t* p, uint64_t v) {
    if (RELAXED)
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else
      st_release_sys_u64(p, v);
  };
  auto get = [](const uint64_t* p) { return RELAXED ? ld_relaxed_sys_u64(p) : ld_acquire_sys_u64(p); };
  const uint64_t t0 = globaltimer_ns();
  bool ok = true;
  for (int i = 1; i <= iters && ok; ++i) {
    if (initiator) put(peer_flag, base + i);
    while (get(my_flag) < base + i)
      if (globaltimer_ns() - t0 > timeout_ns) {
        ok = false;
        break;
      }
    if (!initiator && ok) put(peer_flag, base + i);
  }
  out[0] = globaltimer_ns() - t0;
  out[1] = ok ? 0 : 1;
}

// mode 0: multimem.red   — add (rank+1)*(i+1) into every GPU's copy of element i
// mode 1: multimem.ld_reduce — out[i] = sum over all GPUs' copies of element i
// mode 2: multimem.st    — write `value + i` into every GPU's copy of element i
__global__ void multimem_selftest_kernel(unsigned long long* mc, unsigned long long* out, size_t n,
                                         int mode, unsigned long long value) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    if (mode == 0) {
      asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(mc + i),
                   "l"(value * (unsigned long long)(i + 1))
                   : "memory");
    } else if (mode == 1) {
      unsigned long long v;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];"
                   : "=l"(v)
                   : "l"(mc + i)
                   : "memory");
      out[i] = v;
    } else {
      asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(mc + i), "l"(value + i)
                   : "memory");
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_selftest_pingpong(void* my_flag, void* peer_flag, int initiator, int iters, uint64_t base,
                           int relaxed, void* out2_u64, b200_stream_t s) {
  B200_REQUIRE(iters >= 1, "pingpong: iters %d < 1", iters);
  auto kern = relaxed ? pingpong_kernel<true> : pingpong_kernel<false>;
  kern<<<1, 32, 0, as_stream(s)>>>((uint64_t*)my_flag, (uint64_t*)peer_flag, initiator, iters, base,
                                   (unsigned long long*)out2_u64, 20ull * 1000 * 1000 * 1000);
  B200_LAUNCH_OK();
  return 0;
}

int b200_selftest_multimem(void* mc_ptr, void* out, size_t n, int mode, uint64_t value,
                           b200_stream_t s) {
  B200_REQUIRE(mode >= 0 && mode <= 2, "multimem selftest: mode %d outside 0..2", mode);
  B200_REQUIRE(mode != 1 || out != nullptr, "multimem selftest: ld_reduce needs an output buffer");
  if (n == 0) return 0;
  const int threads = 256;
  const int blocks = (int)((n + threads - 1) / threads < 1184 ? (n + threads - 1) / threads : 1184);
  multimem_selftest_kernel<<<blocks, threads, 0, as_stream(s)>>>(
      (unsigned long long*)mc_ptr, (unsigned long long*)out, n, mode, value);
  B200_LAUNCH_OK();
  return 0;
}

// What the current device can do: fd-exportable VMM allocations, multicast objects.
int b200_vmm_query(int* vmm_supported, int* multicast_supported) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  const bool api = drv().ok;
  const int vmm = api && attr(dev, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED) &&
                  attr(dev, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED);
  const int mc = vmm && drv().multicastCreate && drv().multicastBindMem &&
                 attr(dev, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED);
  if (vmm_supported) *vmm_supported = vmm;
  if (multicast_supported) *multicast_supported = mc;
  return 0;
}

// Size multiple that satisfies both the allocator and (when world > 1 and the device
// supports it) a multicast object over `world` devices.
int b200_vmm_granularity(int world, size_t* gran_out) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  B200_REQUIRE(drv().ok, "vmm: driver entry points unavailable");
  CUmemAllocationProp p = alloc_prop(dev);
  size_t g = 0;
  B200_CU_OK(drv().memGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  int mc = 0;
  b200_vmm_query(nullptr, &mc);
  if (mc && world > 1 && drv().multicastGetGranularity) {
    CUmulticastObjectProp mp = mc_prop(world, g);
    size_t mg = 0;
    B200_CU_OK(drv().multicastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mg > g) g = mg;
  }
  *gran_out = g;
  return 0;
}

// Physical memory on the current device; `bytes` must be a multiple of the granularity.
int b200_vmm_create(size_t bytes, uint64_t* handle_out) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  B200_REQUIRE(drv().ok, "vmm: driver entry points unavailable");
  CUmemAllocationProp p = alloc_prop(dev);
  CUmemGenericAllocationHandle h = 0;
  B200_CU_OK(drv().memCreate(&h, bytes, &p, 0));
  *handle_out = (uint64_t)h;
  return 0;
}

int b200_vmm_release(uint64_t handle) {
  B200_CU_OK(drv().memRelease((CUmemGenericAllocationHandle)handle));
  return 0;
}

// A new file descriptor naming the allocation (or multicast object); the caller closes it.
int b200_vmm_export_fd(uint64_t handle, int* fd_out) {
  int fd = -1;
  B200_CU_OK(drv().memExportToShareableHandle(&fd, (CUmemGenericAllocationHandle)handle,
                                              CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

// The handle behind a descriptor received from another process (memory or multicast).
int b200_vmm_import_fd(int fd, uint64_t* handle_out) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  CUmemGenericAllocationHandle h = 0;
  B200_CU_OK(drv().memImportFromShareableHandle(&h, (void*)(uintptr_t)fd,
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle_out = (uint64_t)h;
  return 0;
}

// Reserve an address range, map the handle there, give the current device read/write access.
int b200_vmm_map(uint64_t handle, size_t bytes, void** va_out) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  size_t gran = 0;
  if (b200_vmm_granularity(1, &gran)) return 1;
  CUdeviceptr va = 0;
  B200_CU_OK(drv().memAddressReserve(&va, bytes, gran, 0, 0));
  CUresult r = drv().memMap(va, bytes, 0, (CUmemGenericAllocationHandle)handle, 0);
  if (r != CUDA_SUCCESS) {
    drv().memAddressFree(va, bytes);
    B200_CU_OK(r);
  }
  CUmemAccessDesc d{};
  d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  d.location.id = dev;
  d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = drv().memSetAccess(va, bytes, &d, 1);
  if (r != CUDA_SUCCESS) {
    drv().memUnmap(va, bytes);
    drv().memAddressFree(va, bytes);
    B200_CU_OK(r);
  }
  *va_out = (void*)va;
  return 0;
}

int b200_vmm_unmap(void* va, size_t bytes) {
  B200_CU_OK(drv().memUnmap((CUdeviceptr)va, bytes));
  B200_CU_OK(drv().memAddressFree((CUdeviceptr)va, bytes));
  return 0;
}

// ---- multicast (NVLS) ---------------------------------------------------------
// Order of operations across the ranks (parallel/vmm.py enforces it):
//   rank 0: b200_mc_create -> export fd -> send to the others -> they import it
//   every rank: b200_mc_add_device      (all must finish before ANY bind)
//   every rank: b200_mc_bind(own memory handle)
//   every rank: b200_vmm_map(mc handle) -> the multicast address
int b200_mc_create(int num_devices, size_t bytes, uint64_t* mc_out) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  B200_REQUIRE(drv().multicastCreate != nullptr, "multicast: driver entry points unavailable");
  CUmulticastObjectProp p = mc_prop(num_devices, bytes);
  CUmemGenericAllocationHandle h = 0;
  B200_CU_OK(drv().multicastCreate(&h, &p));
  *mc_out = (uint64_t)h;
  return 0;
}

int b200_mc_add_device(uint64_t mc) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  B200_CU_OK(drv().multicastAddDevice((CUmemGenericAllocationHandle)mc, dev));
  return 0;
}

int b200_mc_bind(uint64_t mc, uint64_t mem_handle, size_t bytes) {
  B200_CU_OK(drv().multicastBindMem((CUmemGenericAllocationHandle)mc, 0,
                                    (CUmemGenericAllocationHandle)mem_handle, 0, bytes, 0));
  return 0;
}

int b200_mc_unbind(uint64_t mc, size_t bytes) {
  int dev = 0;
  if (current_device(&dev)) return 1;
  B200_CU_OK(drv().multicastUnbind((CUmemGenericAllocationHandle)mc, dev, 0, bytes));
  return 0;
}

}  // extern "C"
