// slc_peer.cuh -- the peer-memory ordering primitives shared by slc_peer.cu (stand-alone barrier
// kernel, group management) and the product kernels that fuse signal and wait into their own launch
// (slc_spmv.cu).  See slc_peer.cu for the protocol.
#pragma once

#include "slc_common.cuh"

namespace slc {
constexpr int MAX_PEERS = 16;
constexpr unsigned long long PEER_TIMEOUT_NS = 5ull * 1000 * 1000 * 1000;
// words of slc_peer_group::state (device memory of the owning rank)
enum : int {
  PEER_EPOCH = 0,         // epochs completed so far; the product in flight is PEER_EPOCH + 1
  PEER_TICKET = 1,        // blocks of the running fused product that have finished
  PEER_STAT_WAITS = 2,    // number of timed waits (one per waiting block)
  PEER_STAT_WAIT_SUM = 3, // nanoseconds spent in them, summed
  PEER_STAT_WAIT_MAX = 4, // and the longest one
  PEER_STATE_WORDS = 8,
};
}  // namespace slc

struct slc_peer_group {
  int32_t nparts = 0, self = 0;
  int device = 0;
  unsigned long long *flags = nullptr;             // this rank's array: slot p is written by rank p
  unsigned long long *peer_flags[slc::MAX_PEERS];  // every rank's array as mapped here
  bool imported[slc::MAX_PEERS] = {};              // mapping came from cudaIpcOpenMemHandle
  unsigned long long **table_dev = nullptr;        // device copy of peer_flags
  unsigned long long *state = nullptr;             // device: epoch, ticket, wait statistics
  int *timeout_host = nullptr;                     // mapped pinned host word: epoch of the first wait that gave up
  // Ranks this one exchanges operand data with in either direction (it reads their slice, or they
  // read its slice).  A product waits for exactly these; all peers until slc_peer_group_set_deps.
  uint32_t deps_mask = 0;
  bool connected = false;
};

namespace slc {

// What a kernel needs to signal and wait (passed by value).
struct PeerSync {
  unsigned long long *const *flags;  // device array [nparts]: every rank's flag array as mapped here
  unsigned long long *state;         // this rank's state words
  int *timeout_host;
  int32_t nparts, self;
  uint32_t signal_mask;              // peers told "my slice is written" (all: a store is fire-and-forget)
};

int peer_check(slc_peer_group *g, const char *fn);  // null / unconnected / earlier timeout -> error
PeerSync peer_sync_of(slc_peer_group *g);

__device__ __forceinline__ unsigned long long peer_globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// epoch of the product in flight: stable for the whole launch (advanced only by its last block)
__device__ __forceinline__ unsigned long long peer_epoch(const PeerSync &ps) {
  unsigned long long e;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(e) : "l"(ps.state + PEER_EPOCH));
  return e + 1;
}

// "everything this GPU wrote before this kernel (stream order) is visible to whoever observes `epoch`"
__device__ __forceinline__ void peer_signal_one(const PeerSync &ps, int p, unsigned long long epoch) {
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(ps.flags[p] + ps.self), "l"(epoch) : "memory");
}

// Spin (bounded) until peer p has published `epoch`.  Returns false on timeout, after recording the
// epoch in the group's host-visible word.
__device__ __forceinline__ bool peer_wait_one(const PeerSync &ps, int p, unsigned long long epoch) {
  const unsigned long long *mine = ps.flags[ps.self] + p;
  const unsigned long long t0 = peer_globaltimer_ns();
  bool ok = true;
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
    if (v >= epoch) break;
    if (peer_globaltimer_ns() - t0 > PEER_TIMEOUT_NS) {
      *(volatile int *)ps.timeout_host = (int)epoch;
      ok = false;
      break;
    }
    __nanosleep(40);
  }
  const unsigned long long waited = peer_globaltimer_ns() - t0;
  atomicAdd(ps.state + PEER_STAT_WAITS, 1ull);
  atomicAdd(ps.state + PEER_STAT_WAIT_SUM, waited);
  atomicMax(ps.state + PEER_STAT_WAIT_MAX, waited);
  return ok;
}

__device__ __forceinline__ void peer_advance(const PeerSync &ps, unsigned long long epoch) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(ps.state + PEER_EPOCH), "l"(epoch) : "memory");
}

// Called by every block of a fused product when it is done: the last one advances the epoch.
__device__ __forceinline__ void peer_block_done(const PeerSync &ps, unsigned nblocks) {
  __syncthreads();  // every thread of the block has read the epoch it needed
  if (threadIdx.x == 0) {
    const unsigned long long t = atomicAdd(ps.state + PEER_TICKET, 1ull);
    if (t + 1 == nblocks) {
      const unsigned long long epoch = peer_epoch(ps);
      ps.state[PEER_TICKET] = 0;
      __threadfence();
      peer_advance(ps, epoch);
    }
  }
}

}  // namespace slc
