// File <-> device transfers for the contract binaries, pipelined.
//
// The reference reads the whole file into pageable memory, copies it to the GPU with one
// blocking cudaMemcpy, and does the reverse on the way out (ref: lab2/src/main.cu:61-66,
// :106-123): four full passes over host memory, nothing overlapped.  Here a few I/O
// threads each own two pinned staging buffers and a CUDA stream: a thread preads its next
// chunk of the file straight into pinned memory while the DMA of its previous chunk is
// still in flight, so the page-cache copy, the PCIe transfer and (on the way out) the
// write into the output file all overlap.  The output file is created and its pages are
// allocated by a helper thread WHILE the CUDA context is being created, which is the
// other large fixed cost of a short-lived process.
//
// Small files (below kPipeMin) keep the plain path: pinning memory costs more than it saves.
#pragma once
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <thread>
#include <vector>

#include "cli_common.h"

namespace cli {

constexpr size_t kPipeChunk = 8u << 20;   // bytes per DMA
constexpr int kPipeThreads = 4;           // I/O threads, two pinned buffers each
constexpr size_t kPipeMin = 32u << 20;    // payload bytes from which the pipeline pays off

inline void pread_full(int fd, void* dst, size_t len, off_t off, const char* what) {
  uint8_t* p = (uint8_t*)dst;
  while (len) {
    ssize_t got = pread(fd, p, len, off);
    if (got <= 0) die("'%s': truncated data", what);
    p += got, off += got, len -= (size_t)got;
  }
}
inline void pwrite_full(int fd, const void* src, size_t len, off_t off, const char* what) {
  const uint8_t* p = (const uint8_t*)src;
  while (len) {
    ssize_t put = pwrite(fd, p, len, off);
    if (put <= 0) die("short write to '%s'", what);
    p += put, off += put, len -= (size_t)put;
  }
}

// true when `a` and `b` name the same existing file (then reading must finish before writing)
inline bool same_file(const std::string& a, const std::string& b) {
  struct stat sa, sb;
  if (stat(a.c_str(), &sa) != 0 || stat(b.c_str(), &sb) != 0) return false;
  return sa.st_dev == sb.st_dev && sa.st_ino == sb.st_ino;
}

// Creates the output file and allocates its pages in the background (tmpfs / page cache
// allocation is a third of the reference's wall time on a 1 GiB image).
struct OutputFile {
  int fd = -1;
  std::string path;
  std::thread worker;
  void begin(const std::string& p, const void* header, size_t header_bytes, size_t total_bytes) {
    path = p;
    fd = open(p.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) die("cannot open output file '%s'", p.c_str());
    std::vector<uint8_t> head((const uint8_t*)header, (const uint8_t*)header + header_bytes);
    worker = std::thread([this, head, total_bytes] {
      pwrite_full(fd, head.data(), head.size(), 0, path.c_str());
      (void)!posix_fallocate(fd, 0, (off_t)total_bytes);  // best effort: only a head start
    });
  }
  void ready() {
    if (worker.joinable()) worker.join();
  }
  void finish() {
    ready();
    if (fd >= 0 && close(fd) != 0) die("short write to '%s'", path.c_str());
    fd = -1;
  }
};

#ifdef __CUDACC__
struct PinnedRing {
  uint8_t* base = nullptr;
  PinnedRing() { CSC(cudaHostAlloc((void**)&base, kPipeChunk * 2 * kPipeThreads, cudaHostAllocDefault)); }
  ~PinnedRing() { cudaFreeHost(base); }
  uint8_t* buf(int thread, int half) { return base + (size_t)(thread * 2 + half) * kPipeChunk; }
};

// file[off, off + bytes) -> d_dst
inline void file_to_device(int fd, off_t off, void* d_dst, size_t bytes, PinnedRing& ring, const char* what) {
  const size_t n_chunks = (bytes + kPipeChunk - 1) / kPipeChunk;
  int dev = 0;
  CSC(cudaGetDevice(&dev));
  auto body = [&](int t) {
    CSC(cudaSetDevice(dev));
    cudaStream_t st;
    cudaEvent_t ev[2];
    CSC(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& e : ev) CSC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    size_t round = 0;
    for (size_t k = t; k < n_chunks; k += kPipeThreads, ++round) {
      const int half = (int)(round & 1);
      if (round >= 2) CSC(cudaEventSynchronize(ev[half]));  // this buffer's previous DMA is done
      const size_t o = k * kPipeChunk, len = std::min(kPipeChunk, bytes - o);
      pread_full(fd, ring.buf(t, half), len, off + (off_t)o, what);
      CSC(cudaMemcpyAsync((uint8_t*)d_dst + o, ring.buf(t, half), len, cudaMemcpyHostToDevice, st));
      CSC(cudaEventRecord(ev[half], st));
    }
    CSC(cudaStreamSynchronize(st));
    for (auto& e : ev) cudaEventDestroy(e);
    cudaStreamDestroy(st);
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < kPipeThreads; ++t) pool.emplace_back(body, t);
  body(0);
  for (auto& th : pool) th.join();
}

// d_src -> file[off, off + bytes)
inline void device_to_file(int fd, off_t off, const void* d_src, size_t bytes, PinnedRing& ring,
                           const char* what) {
  const size_t n_chunks = (bytes + kPipeChunk - 1) / kPipeChunk;
  int dev = 0;
  CSC(cudaGetDevice(&dev));
  auto body = [&](int t) {
    CSC(cudaSetDevice(dev));
    cudaStream_t st;
    cudaEvent_t ev[2];
    CSC(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& e : ev) CSC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    size_t round = 0;
    bool have_prev = false;
    size_t prev_o = 0, prev_len = 0;
    int prev_half = 0;
    auto flush_prev = [&] {  // the copy of the previous chunk has landed: write it out
      if (!have_prev) return;
      CSC(cudaEventSynchronize(ev[prev_half]));
      pwrite_full(fd, ring.buf(t, prev_half), prev_len, off + (off_t)prev_o, what);
    };
    for (size_t k = t; k < n_chunks; k += kPipeThreads, ++round) {
      const int half = (int)(round & 1);
      const size_t o = k * kPipeChunk, len = std::min(kPipeChunk, bytes - o);
      // buffer `half` was written out two rounds ago (flush_prev below), so it is free
      CSC(cudaMemcpyAsync(ring.buf(t, half), (const uint8_t*)d_src + o, len, cudaMemcpyDeviceToHost, st));
      CSC(cudaEventRecord(ev[half], st));
      flush_prev();  // overlaps with the copy just enqueued
      have_prev = true, prev_o = o, prev_len = len, prev_half = half;
    }
    flush_prev();
    for (auto& e : ev) cudaEventDestroy(e);
    cudaStreamDestroy(st);
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < kPipeThreads; ++t) pool.emplace_back(body, t);
  body(0);
  for (auto& th : pool) th.join();
}
#endif  // __CUDACC__

}  // namespace cli
