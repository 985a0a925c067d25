// Symmetric per-rank workspace and the device-side signalling primitives shared
// by the fused multi-GPU kernels (lab2 halo needs none of this: its halo row is
// read-only input; lab3 all-reduce and lab5 all-to-all use the slots below).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kMaxWorld = 8;
constexpr unsigned long long kDefaultTimeoutNs = 20ull * 1000 * 1000 * 1000;  // 20 s
constexpr int kLab3C = B200_LAB3_MAX_CLASSES, kLab3M = B200_LAB3_MOMENTS;
constexpr int kSortMaxSamplesTotal = 8192;  // world * samples_per_rank, sorted in one CTA

enum SymmFlag {
  kFlagBarrier = 0, kFlagLab3 = 1, kFlagSortSamples = 2, kFlagSortDone = 3, kFlagLab3Lut = 4, kNumFlags = 5
};
enum SymmError { kErrTimeout = 1, kErrSortOverflow = 2 };

// One of these lives on every rank at a symmetric (peer-mapped) address.  Fields
// marked [remote] are written by peers over NVLink, the rest only locally.
struct SymmWorkspace {
  uint64_t flags[kNumFlags][kMaxWorld];  // [remote] flags[f][sender] = last epoch signalled
  unsigned long long error;              // sticky error word (SymmError bits)
  // ---- lab3: in-kernel all-reduce of the class moments
  long long lab3_acc[kLab3C][kLab3M];                 // this rank's partial sums
  unsigned int lab3_blocks_done;
  unsigned int lab3_lut_blocks_done;
  long long lab3_slots[2][kMaxWorld][kLab3C][kLab3M];  // [remote] slots[epoch&1][sender]
  // ---- lab5: sample exchange + bucket cursor for the fused all-to-all
  unsigned long long sort_cursor;        // [remote atomics] next free index of my bucket
  unsigned int sort_blocks_done;
  unsigned int pad1;
  unsigned long long sort_samples[kSortMaxSamplesTotal];  // [remote] composites, by sender
  unsigned long long sort_splitters[kMaxWorld];
  unsigned long long sort_recv_count;    // keys that landed in my bucket (device-side n)
  // exact two-phase reservation: how many of MY keys go to each rank, and where my
  // segment inside that rank's bucket begins / how far I have filled it
  unsigned long long sort_send_count[kMaxWorld];
  unsigned long long sort_local_cursor[kMaxWorld];
  unsigned int sort_count_blocks_done;
  unsigned int pad2;
};

struct SymmPeers {
  SymmWorkspace* const* ws;  // device array [world]: every rank's workspace, peer-mapped
};

#if defined(__CUDACC__)
__device__ __forceinline__ void signal_peer(const SymmPeers& peers, int peer, int me, uint64_t epoch,
                                            int which = kFlagBarrier) {
  st_release_sys_u64(&peers.ws[peer]->flags[which][me], epoch);
}
// spin until flags[which][sender] on MY workspace reaches epoch; false on timeout
__device__ __forceinline__ bool wait_peer(const SymmPeers& peers, int me, int sender,
                                          uint64_t epoch, unsigned long long timeout_ns,
                                          int which = kFlagBarrier) {
  const uint64_t* f = &peers.ws[me]->flags[which][sender];
  const uint64_t t0 = globaltimer_ns();
  while (ld_acquire_sys_u64(f) < epoch) {
    if (globaltimer_ns() - t0 > timeout_ns) {
      atomicOr(&peers.ws[me]->error, (unsigned long long)kErrTimeout);
      return false;
    }
    __nanosleep(200);
  }
  return true;
}
#endif

}  // namespace b200
