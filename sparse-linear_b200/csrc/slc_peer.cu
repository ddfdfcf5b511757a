// slc_peer.cu -- a barrier between the GPUs of one NVSwitch domain, written over peer memory.
//
// The row-partitioned SpMV gathers remote operand entries directly from the peers' HBM (DistX in
// slc_spmv.cu), so the only communication step left is ordering: rank r may start its product once
// every peer has finished writing its slice of x.  An NCCL all-reduce does that but costs tens of
// microseconds per step, comparable to the product itself at 8 GPUs (~55 us).  This barrier is one
// small kernel: every rank stores its epoch into its slot of every peer's flag array (release,
// system scope, over NVLink) and then spins on its own array (acquire) until all slots reached the
// epoch.  The spin is bounded (SLC_PEER_TIMEOUT_NS): a missing peer raises an error flag instead of
// hanging the GPU.
#include "slc_common.cuh"

namespace slc {
constexpr int MAX_PEERS = 16;
constexpr unsigned long long PEER_TIMEOUT_NS = 5ull * 1000 * 1000 * 1000;
}  // namespace slc

struct slc_peer_group {
  int32_t nparts = 0, self = 0;
  unsigned long long epoch = 0;
  unsigned long long *flags = nullptr;             // this rank's array: slot p is written by rank p
  unsigned long long *peer_flags[slc::MAX_PEERS];  // every rank's array as mapped here
  int *timeout_flag = nullptr;                     // device: set when a spin gave up
  unsigned long long **table_dev = nullptr;        // device copy of peer_flags (for fused kernels)
  bool connected = false;
};

namespace slc {

struct PeerTable {
  unsigned long long *flags[MAX_PEERS];
};

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// mode bit 0: signal (store my epoch into every peer's array); bit 1: wait (until every slot of my
// own array reached the epoch).  Split in two, the signal goes out before the interior rows of a
// product and the wait sits just before its boundary rows, so the peers' latency hides behind work.
__global__ void k_peer_barrier(PeerTable peers, int32_t nparts, int32_t self, unsigned long long epoch,
                               int mode, int *__restrict__ timeout_flag) {
  const int p = threadIdx.x;
  if (p >= nparts) return;
  if (mode & 1) {
    // everything this GPU wrote before the barrier (earlier in stream order) is visible to a peer
    // that observes the epoch
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peers.flags[p] + self), "l"(epoch) : "memory");
  }
  if (!(mode & 2)) return;
  const unsigned long long *mine = peers.flags[self] + p;
  const unsigned long long t0 = global_ns();
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
    if (v >= epoch) break;
    if (global_ns() - t0 > PEER_TIMEOUT_NS) {
      atomicExch(timeout_flag, 1);
      break;
    }
    __nanosleep(100);
  }
}

}  // namespace slc

using namespace slc;

extern "C" {

int slc_peer_group_create(int32_t nparts, int32_t self, slc_peer_group **out, void *handle72) {
  if (!out || !handle72) return fail(SLC_ERR_INVALID, "slc_peer_group_create: null argument");
  *out = nullptr;
  if (nparts < 1 || nparts > MAX_PEERS || self < 0 || self >= nparts)
    return fail(SLC_ERR_INVALID, "slc_peer_group_create: 1..%d ranks supported", MAX_PEERS);
  slc_peer_group *g = new (std::nothrow) slc_peer_group();
  if (!g) return fail(SLC_ERR_NOMEM, "out of host memory");
  g->nparts = nparts;
  g->self = self;
  for (int p = 0; p < MAX_PEERS; ++p) g->peer_flags[p] = nullptr;
  // cudaMalloc (not the stream-ordered pool): pool memory cannot be exported through legacy IPC
  cudaError_t e = cudaMalloc((void **)&g->flags, MAX_PEERS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc((void **)&g->timeout_flag, sizeof(int));
  if (e == cudaSuccess) e = cudaMemset(g->flags, 0, MAX_PEERS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(g->timeout_flag, 0, sizeof(int));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(g->flags);
    cudaFree(g->timeout_flag);
    delete g;
    return fail(SLC_ERR_CUDA, "slc_peer_group_create: %s", cudaGetErrorString(e));
  }
  int rc = slc_ipc_export(g->flags, handle72);
  if (rc != SLC_OK) {
    cudaFree(g->flags);
    cudaFree(g->timeout_flag);
    delete g;
    return rc;
  }
  g->peer_flags[self] = g->flags;
  *out = g;
  return SLC_OK;
}

int slc_peer_group_connect(slc_peer_group *g, const void *handles72) {
  if (!g || !handles72) return fail(SLC_ERR_INVALID, "slc_peer_group_connect: null argument");
  for (int p = 0; p < g->nparts; ++p) {
    if (p == g->self) continue;
    void *ptr = nullptr;
    SLC_TRY(slc_ipc_import((const char *)handles72 + (size_t)p * 72, &ptr));
    g->peer_flags[p] = (unsigned long long *)ptr;
  }
  cudaError_t e = cudaMalloc((void **)&g->table_dev, MAX_PEERS * sizeof(unsigned long long *));
  if (e == cudaSuccess)
    e = cudaMemcpy(g->table_dev, g->peer_flags, MAX_PEERS * sizeof(unsigned long long *), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return fail(SLC_ERR_CUDA, "slc_peer_group_connect: %s", cudaGetErrorString(e));
  g->connected = true;
  return SLC_OK;
}

static int peer_launch(slc_peer_group *g, int mode, const char *fn) {
  if (!g) return fail(SLC_ERR_INVALID, "%s: null group", fn);
  if (g->nparts == 1) return SLC_OK;
  if (!g->connected) return fail(SLC_ERR_INVALID, "%s: group is not connected", fn);
  PeerTable t;
  for (int p = 0; p < MAX_PEERS; ++p) t.flags[p] = g->peer_flags[p];
  if (mode & 1) g->epoch += 1;
  k_peer_barrier<<<1, 32, 0, cur_stream()>>>(t, g->nparts, g->self, g->epoch, mode, g->timeout_flag);
  SLC_LAUNCHED();
  return SLC_OK;
}

int slc_peer_barrier(slc_peer_group *g) { return peer_launch(g, 3, "slc_peer_barrier"); }
int slc_peer_signal(slc_peer_group *g) { return peer_launch(g, 1, "slc_peer_signal"); }
int slc_peer_wait(slc_peer_group *g) { return peer_launch(g, 2, "slc_peer_wait"); }

int slc_peer_group_timed_out(slc_peer_group *g, int *flag) {
  if (!g || !flag) return fail(SLC_ERR_INVALID, "slc_peer_group_timed_out: null argument");
  SLC_CUDA(cudaMemcpy(flag, g->timeout_flag, sizeof(int), cudaMemcpyDeviceToHost));
  return SLC_OK;
}

void slc_peer_group_free(slc_peer_group *g) {
  if (!g) return;
  for (int p = 0; p < g->nparts; ++p)
    if (p != g->self && g->peer_flags[p]) slc_ipc_close(g->peer_flags[p]);
  cudaFree(g->flags);
  cudaFree(g->timeout_flag);
  cudaFree(g->table_dev);
  delete g;
}

}  // extern "C"

namespace slc {
// For kernels that fuse the barrier (k_spmv_sell_dist): start a new epoch and expose the tables.
int peer_sync_args(slc_peer_group *g, unsigned long long *const **flags, int **timeout_flag,
                   unsigned long long *epoch, int32_t *nparts, int32_t *self) {
  if (!g) return fail(SLC_ERR_INVALID, "peer sync: null group");
  *flags = nullptr;
  if (g->nparts == 1) return SLC_OK;
  if (!g->connected) return fail(SLC_ERR_INVALID, "peer sync: group is not connected");
  g->epoch += 1;
  *flags = g->table_dev;
  *timeout_flag = g->timeout_flag;
  *epoch = g->epoch;
  *nparts = g->nparts;
  *self = g->self;
  return SLC_OK;
}
}  // namespace slc
