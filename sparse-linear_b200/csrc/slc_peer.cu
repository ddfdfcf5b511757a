// slc_peer.cu -- ordering between the GPUs of one NVSwitch domain, written over peer memory.
//
// The row-partitioned products gather remote operand entries directly from the peers' HBM (DistX in
// slc_spmv.cu), so the only communication step left is ordering: rank r may read a peer's slice of x
// once that peer has finished writing it.  An NCCL all-reduce does that but costs tens of
// microseconds per step, comparable to the product itself at 8 GPUs (~45 us).  Here every rank owns a
// small flag array in its HBM that every peer maps (CUDA IPC between processes, plain peer access
// inside one process); "signal" stores this rank's epoch into its slot of the peers' arrays (release,
// system scope, over NVLink) and "wait" spins on the rank's own array (acquire) until the slots of
// the peers it depends on reached the epoch.
//
// Round 2 changes, from the round-1 review:
//   * the epoch lives in device memory and is advanced by the kernels themselves (the last block of
//     a fused product, or the wait kernel), so a product is a pure function of device state: K
//     products can be captured in one CUDA graph and the graph relaunched;
//   * a rank waits only for the peers it reads from or is read by (PeerDeps: for a z-slab stencil the
//     two neighbours), not for all N-1 -- a step is no longer a global rendezvous;
//   * the spin is bounded (PEER_TIMEOUT_NS) and a timeout is sticky in a word of mapped host memory
//     that every later call checks, so a product that gave up is reported, not silently wrong;
//   * waits are timed (globaltimer) into a small statistics block (slc_peer_group_stats).
#include "slc_peer.cuh"

namespace slc {

__global__ void k_peer_barrier(PeerSync ps, uint32_t wait_mask, int mode) {
  const int p = threadIdx.x;
  const unsigned long long epoch = peer_epoch(ps);
  if (p < ps.nparts && p != ps.self) {
    if ((mode & 1) && ((mode & 4) || ((ps.signal_mask >> p) & 1u))) peer_signal_one(ps, p, epoch);
    if ((mode & 2) && ((mode & 4) || ((wait_mask >> p) & 1u))) peer_wait_one(ps, p, epoch);
  }
  if (mode & 2) {
    __syncthreads();
    if (p == 0) peer_advance(ps, epoch);
  }
}

static int enable_peer(int from_dev, int to_dev) {
  if (from_dev == to_dev) return SLC_OK;
  int can = 0;
  SLC_CUDA(cudaDeviceCanAccessPeer(&can, from_dev, to_dev));
  if (!can) return fail(SLC_ERR_CUDA, "device %d cannot map device %d's memory", from_dev, to_dev);
  int prev = 0;
  SLC_CUDA(cudaGetDevice(&prev));
  SLC_CUDA(cudaSetDevice(from_dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(to_dev, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); e = cudaSuccess; }
  cudaSetDevice(prev);
  SLC_CUDA(e);
  return SLC_OK;
}

static int group_alloc(int32_t nparts, int32_t self, slc_peer_group **out) {
  slc_peer_group *g = new (std::nothrow) slc_peer_group();
  if (!g) return fail(SLC_ERR_NOMEM, "out of host memory");
  g->nparts = nparts;
  g->self = self;
  cudaGetDevice(&g->device);
  for (int p = 0; p < MAX_PEERS; ++p) g->peer_flags[p] = nullptr;
  g->deps_mask = nparts >= 32 ? 0xffffffffu : ((1u << nparts) - 1u);
  g->deps_mask &= ~(1u << self);
  // cudaMalloc (not the stream-ordered pool): pool memory cannot be exported through legacy IPC
  cudaError_t e = cudaMalloc((void **)&g->flags, MAX_PEERS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc((void **)&g->state, PEER_STATE_WORDS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc((void **)&g->table_dev, MAX_PEERS * sizeof(unsigned long long *));
  if (e == cudaSuccess) e = cudaHostAlloc((void **)&g->timeout_host, sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable);
  if (e == cudaSuccess) e = cudaMemset(g->flags, 0, MAX_PEERS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(g->state, 0, PEER_STATE_WORDS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    cudaFree(g->flags);
    cudaFree(g->state);
    cudaFree(g->table_dev);
    if (g->timeout_host) cudaFreeHost(g->timeout_host);
    delete g;
    return fail(SLC_ERR_CUDA, "slc_peer_group_create: %s", cudaGetErrorString(e));
  }
  *g->timeout_host = 0;
  g->peer_flags[self] = g->flags;
  *out = g;
  return SLC_OK;
}

static int group_publish_table(slc_peer_group *g) {
  SLC_CUDA(cudaMemcpy(g->table_dev, g->peer_flags, MAX_PEERS * sizeof(unsigned long long *), cudaMemcpyHostToDevice));
  g->connected = true;
  return SLC_OK;
}

int peer_check(slc_peer_group *g, const char *fn) {
  if (!g) return fail(SLC_ERR_INVALID, "%s: null group", fn);
  if (g->nparts > 1 && !g->connected) return fail(SLC_ERR_INVALID, "%s: group is not connected", fn);
  const int t = *(volatile int *)g->timeout_host;
  if (t) return fail(SLC_ERR_CUDA, "%s: a peer did not signal epoch %d within %.1f s; the rows that needed it were "
                     "filled with NaN", fn, t, (double)PEER_TIMEOUT_NS * 1e-9);
  return SLC_OK;
}

PeerSync peer_sync_of(slc_peer_group *g) {
  PeerSync ps;
  ps.flags = g->table_dev;
  ps.state = g->state;
  ps.timeout_host = g->timeout_host;
  ps.nparts = g->nparts;
  ps.self = g->self;
  ps.signal_mask = g->nparts >= 32 ? 0xffffffffu : ((1u << g->nparts) - 1u);
  ps.signal_mask &= ~(1u << g->self);
  return ps;
}

static int peer_launch(slc_peer_group *g, int mode, const char *fn) {
  SLC_TRY(peer_check(g, fn));
  if (g->nparts == 1) return SLC_OK;
  k_peer_barrier<<<1, 32, 0, cur_stream()>>>(peer_sync_of(g), g->deps_mask, mode);
  SLC_LAUNCHED();
  return SLC_OK;
}

}  // namespace slc

using namespace slc;

extern "C" {

int slc_peer_group_create(int32_t nparts, int32_t self, slc_peer_group **out, void *handle72) {
  if (!out || !handle72) return fail(SLC_ERR_INVALID, "slc_peer_group_create: null argument");
  *out = nullptr;
  if (nparts < 1 || nparts > MAX_PEERS || self < 0 || self >= nparts)
    return fail(SLC_ERR_INVALID, "slc_peer_group_create: 1..%d ranks supported", MAX_PEERS);
  slc_peer_group *g = nullptr;
  SLC_TRY(group_alloc(nparts, self, &g));
  int rc = slc_ipc_export(g->flags, handle72);
  if (rc != SLC_OK) {
    slc_peer_group_free(g);
    return rc;
  }
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
    g->imported[p] = true;
  }
  return group_publish_table(g);
}

int slc_peer_group_create_local(int32_t ndev, const int32_t *devices, slc_peer_group **groups) {
  if (!groups) return fail(SLC_ERR_INVALID, "slc_peer_group_create_local: null argument");
  if (ndev < 1 || ndev > MAX_PEERS) return fail(SLC_ERR_INVALID, "slc_peer_group_create_local: 1..%d devices supported", MAX_PEERS);
  int prev = 0;
  SLC_CUDA(cudaGetDevice(&prev));
  for (int p = 0; p < ndev; ++p) groups[p] = nullptr;
  int rc = SLC_OK;
  for (int p = 0; p < ndev && rc == SLC_OK; ++p) {
    const int dev = devices ? devices[p] : p;
    if (cudaSetDevice(dev) != cudaSuccess) { (void)cudaGetLastError(); rc = fail(SLC_ERR_CUDA, "slc_peer_group_create_local: no device %d", dev); break; }
    rc = group_alloc(ndev, p, &groups[p]);
  }
  for (int p = 0; p < ndev && rc == SLC_OK; ++p)
    for (int q = 0; q < ndev && rc == SLC_OK; ++q)
      if (p != q) rc = enable_peer(groups[p]->device, groups[q]->device);
  for (int p = 0; p < ndev && rc == SLC_OK; ++p) {
    for (int q = 0; q < ndev; ++q) groups[p]->peer_flags[q] = groups[q]->flags;
    cudaSetDevice(groups[p]->device);
    rc = group_publish_table(groups[p]);
  }
  cudaSetDevice(prev);
  if (rc != SLC_OK)
    for (int p = 0; p < ndev; ++p) { slc_peer_group_free(groups[p]); groups[p] = nullptr; }
  return rc;
}

int slc_peer_group_set_deps(slc_peer_group *g, uint32_t mask) {
  if (!g) return fail(SLC_ERR_INVALID, "slc_peer_group_set_deps: null group");
  const uint32_t all = g->nparts >= 32 ? 0xffffffffu : ((1u << g->nparts) - 1u);
  g->deps_mask = mask & all & ~(1u << g->self);
  return SLC_OK;
}

int slc_peer_barrier(slc_peer_group *g) { return peer_launch(g, 3, "slc_peer_barrier"); }
int slc_peer_barrier_all(slc_peer_group *g) { return peer_launch(g, 7, "slc_peer_barrier_all"); }
int slc_peer_signal(slc_peer_group *g) { return peer_launch(g, 1, "slc_peer_signal"); }
int slc_peer_wait(slc_peer_group *g) { return peer_launch(g, 2, "slc_peer_wait"); }

int slc_peer_group_timed_out(slc_peer_group *g, int *flag) {
  if (!g || !flag) return fail(SLC_ERR_INVALID, "slc_peer_group_timed_out: null argument");
  *flag = *(volatile int *)g->timeout_host != 0;
  return SLC_OK;
}

int slc_peer_group_stats(slc_peer_group *g, int64_t *out4, int reset) {
  if (!g || !out4) return fail(SLC_ERR_INVALID, "slc_peer_group_stats: null argument");
  int prev = 0;
  cudaGetDevice(&prev);
  if (prev != g->device) cudaSetDevice(g->device);
  unsigned long long h[PEER_STATE_WORDS];
  cudaError_t e = cudaMemcpy(h, g->state, sizeof h, cudaMemcpyDeviceToHost);
  static_assert(PEER_STAT_WAIT_SUM == PEER_STAT_WAITS + 1 && PEER_STAT_WAIT_MAX == PEER_STAT_WAITS + 2,
                "the three wait statistics are cleared as one run of words");
  if (e == cudaSuccess && reset) e = cudaMemset(g->state + PEER_STAT_WAITS, 0, 3 * sizeof(unsigned long long));
  if (prev != g->device) cudaSetDevice(prev);
  SLC_CUDA(e);
  out4[0] = (int64_t)h[PEER_EPOCH];
  out4[1] = (int64_t)h[PEER_STAT_WAITS];
  out4[2] = (int64_t)h[PEER_STAT_WAIT_SUM];
  out4[3] = (int64_t)h[PEER_STAT_WAIT_MAX];
  return SLC_OK;
}

void slc_peer_group_free(slc_peer_group *g) {
  if (!g) return;
  int prev = 0;
  cudaGetDevice(&prev);
  if (prev != g->device) cudaSetDevice(g->device);
  for (int p = 0; p < g->nparts; ++p)
    if (g->imported[p] && g->peer_flags[p]) slc_ipc_close(g->peer_flags[p]);
  cudaFree(g->flags);
  cudaFree(g->state);
  cudaFree(g->table_dev);
  if (g->timeout_host) cudaFreeHost(g->timeout_host);
  if (prev != g->device) cudaSetDevice(prev);
  delete g;
}

}  // extern "C"
