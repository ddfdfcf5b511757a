// slc_spmm_window.cu -- SpMM for banded matrices with short rows (2-D stencils: the Complex Double
// Laplacian times FEAST's 64-column subspace block): the window of X a row block needs is staged in
// shared memory by the TMA engine, one column per pipeline stage.
//
// Reference: mulM / FEAST multiplyWork (Data/Matrix/Sparse.hs:473-498, Feast.hs:203-208): every
// column of the block is an independent mulV, terms added in increasing column order.
//
// Why: a row block of R consecutive rows of a banded matrix gathers from ONE contiguous window
// [wlo, whi) of a column of X (R + 2*bandwidth entries).  k_spmm_short fetches that window through
// per-thread gathers: it needs registers for every load in flight (ncu, round 1: 94-156 registers,
// 12-20 warps per SM, 10 exposed-latency loads per warp) and re-fetches the overlap from L2 once per
// 128-row block (3.2x amplification on the 1024x1024 grid).  Here
//   * one producer lane issues ONE cp.async.bulk per (row block, column): window bytes travel
//     L2 -> shared memory with no register, no LSU instruction and no L1 allocation, arriving on an
//     mbarrier (complete_tx); NSTAGE columns are in flight per SM whatever the consumers are doing;
//   * 512 consumer threads own RPT = 2 rows each (a block is 1024 rows, so the window overlap is
//     fetched once per 1024 rows), keep the rows' (column - wlo, value) pairs in registers for all k
//     columns, read X from shared memory with conflict-free 64/128-bit loads, add in the
//     reference's order (bit-identical) and stream Y out with coalesced stores;
//   * full[] / empty[] mbarriers form the ring; no __syncthreads in the column loop.
// Every mbarrier wait is bounded (WINDOW_SPIN_LIMIT probes): a protocol bug sets an error flag and
// produces a wrong answer that the parity tests catch, it cannot hang the GPU.
#include <cstdlib>

#include "slc_common.cuh"

namespace slc {

constexpr int WIN_CONSUMERS = 512;             // consumer threads
constexpr int WIN_THREADS = WIN_CONSUMERS + 32;  // + one producer warp
constexpr int WIN_RPT = 2;                     // rows per consumer thread
constexpr int WIN_ROWS = WIN_CONSUMERS * WIN_RPT;
constexpr int WIN_SELL_C = 32;
constexpr uint32_t WINDOW_SPIN_LIMIT = 1u << 22;

// ---- mbarrier / bulk-copy primitives (PTX ISA 8.x, sm_90+) -------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// bounded wait; returns false (and raises *err) if the phase never completed
__device__ __forceinline__ bool mbar_wait(uint64_t *bar, uint32_t parity, int *err) {
  for (uint32_t spin = 0; spin < WINDOW_SPIN_LIMIT; ++spin)
    if (mbar_try_wait(bar, parity)) return true;
  *(volatile int *)err = 1;  // mapped host word: a plain store, no PCIe atomic
  return false;
}
// global -> shared bulk copy by the TMA engine; completes `bytes` on the mbarrier.  src, dst and bytes
// are multiples of 16.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---- cluster forms (two CTAs share the overlap of their windows through TMA multicast) --------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t *bar, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(rank));
  // .relaxed: the arrive only says "this warp has finished READING the stage" (its loads have completed: their
  // values went into the sums that were stored).  With the default .release.cluster the arrive first waited for
  // the warp's streaming stores of Y to be acknowledged, a microsecond per column: 1.09 ms against 0.47 ms for
  // the kernel without clusters.
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_wait_cluster(uint64_t *bar, uint32_t parity, int *err) {
  for (uint32_t spin = 0; spin < WINDOW_SPIN_LIMIT; ++spin)
    if (mbar_try_wait_cluster(bar, parity)) return true;
  *(volatile int *)err = 1;
  return false;
}
// one read of L2, delivered to the same shared-memory offset of every CTA in `mask`; each destination's
// mbarrier (same offset) is credited with the bytes
__device__ __forceinline__ void bulk_g2s_multicast(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar,
                                                   uint16_t mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}

__device__ __forceinline__ double lds_t(const double *p) { return *p; }
__device__ __forceinline__ double2 lds_t(const double2 *p) { return *p; }

// ---- the kernel -----------------------------------------------------------------------------------------

// products of one column for this thread's rows, X read from the staged window
template <class T, int LMAX, bool FULL>
__device__ __forceinline__ void window_column(const int32_t (&c)[WIN_RPT][LMAX], const T (&a)[WIN_RPT][LMAX],
                                              const int32_t (&len)[WIN_RPT], const bool (&live)[WIN_RPT],
                                              const T *__restrict__ xs, T *__restrict__ ycol,
                                              const int32_t (&row)[WIN_RPT], const YInit<T> &yi) {
#pragma unroll
  for (int i = 0; i < WIN_RPT; ++i) {
    if (!FULL && !live[i]) continue;
    T acc = yi(ycol + row[i]);
#pragma unroll
    for (int u = 0; u < LMAX; ++u)
      if (FULL || u < len[i]) acc = Num<T>::add(Num<T>::mul(a[i][u], lds_t(xs + c[i][u])), acc);
    st_stream(ycol + row[i], acc);
  }
}

template <class T, int LMAX>
__global__ void __launch_bounds__(WIN_THREADS, 1)
k_spmm_window(int32_t nrows, int32_t ncols, int32_t nslices, int32_t nstage, int32_t wcap,
              const int64_t *__restrict__ slice_ptr, const int32_t *__restrict__ rowlen,
              const int32_t *__restrict__ colidx, const T *__restrict__ vals,
              const int32_t *__restrict__ slice_cmin, const int32_t *__restrict__ slice_cmax,
              const T *__restrict__ X, int64_t ldx, T *__restrict__ Y, int64_t ldy, int32_t k,
              const YInit<T> yi, int *__restrict__ err) {
  extern __shared__ __align__(128) unsigned char win_smem[];
  T *stage0 = reinterpret_cast<T *>(win_smem);                                              // [nstage][wcap]
  uint64_t *full = reinterpret_cast<uint64_t *>(win_smem + (size_t)nstage * wcap * sizeof(T));  // [nstage]
  uint64_t *empty = full + nstage;                                                          // [nstage]
  int32_t *s_rng = reinterpret_cast<int32_t *>(empty + nstage);                             // lo, width
  const int tid = threadIdx.x, lane = tid & 31;
  const int32_t row0 = blockIdx.x * WIN_ROWS;

  // the block's window of columns = union of its 32 slices' ranges, 16-byte aligned in X
  if (tid < 32) {
    const int32_t s = row0 / WIN_SELL_C + tid;
    int32_t lo = 0x7fffffff, hi = -1;
    if (s < nslices) {
      lo = __ldg(slice_cmin + s);
      hi = __ldg(slice_cmax + s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (tid == 0) {
      constexpr int32_t EPV = 16 / sizeof(T);  // elements per 16 bytes: 2 (Double) or 1 (Complex)
      if (hi < lo) {                           // no entry in the whole block: stage one vector anyway
        lo = 0;
        hi = EPV - 1;
      }
      lo = lo / EPV * EPV;
      int32_t end = (hi + EPV) / EPV * EPV;    // exclusive, rounded up; the host guarantees ncols % EPV == 0
      end = min(end, ncols);
      s_rng[0] = lo;
      s_rng[1] = end - lo;
      for (int st = 0; st < nstage; ++st) {
        mbar_init(full + st, 1);
        mbar_init(empty + st, WIN_CONSUMERS);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  __syncthreads();
  const int32_t wlo = s_rng[0], width = s_rng[1];
  if (width > wcap || width <= 0) {  // cannot happen with the host's plan; never index past the stage
    if (tid == 0) *(volatile int *)err = 2;
    return;
  }

  if (tid >= WIN_CONSUMERS) {
    // ---- producer: one lane feeds the ring -------------------------------------------------------
    if (lane != 0) return;
    const uint32_t bytes = (uint32_t)width * (uint32_t)sizeof(T);
    int32_t st = 0, use = 0;  // stage index, how many times the ring has wrapped
    for (int32_t j = 0; j < k; ++j) {
      if (use > 0 && !mbar_wait(empty + st, (uint32_t)((use - 1) & 1), err)) return;
      mbar_arrive_expect_tx(full + st, bytes);
      bulk_g2s(stage0 + (size_t)st * wcap, X + (int64_t)j * ldx + wlo, bytes, full + st);
      if (++st == nstage) {
        st = 0;
        ++use;
      }
    }
    return;
  }

  // ---- consumers ---------------------------------------------------------------------------------------
  const uint64_t pol_s = policy_evict_first();
  int32_t row[WIN_RPT], len[WIN_RPT], c[WIN_RPT][LMAX];
  bool live[WIN_RPT];
  T a[WIN_RPT][LMAX];
  bool mine_full = true;
#pragma unroll
  for (int i = 0; i < WIN_RPT; ++i) {
    row[i] = row0 + i * WIN_CONSUMERS + tid;
    live[i] = row[i] < nrows;
    int64_t base = 0;
    len[i] = 0;
    if (live[i]) {
      base = __ldg(slice_ptr + (row[i] >> 5)) + lane;
      len[i] = __ldg(rowlen + row[i]);
    }
#pragma unroll
    for (int u = 0; u < LMAX; ++u) {
      const bool ok = u < len[i];
      c[i][u] = ok ? ld_stream(colidx + base + (int64_t)u * WIN_SELL_C, pol_s) - wlo : 0;
      a[i][u] = ok ? ld_stream(vals + base + (int64_t)u * WIN_SELL_C, pol_s) : Num<T>::zero();
    }
    mine_full = mine_full && live[i] && len[i] == LMAX;
  }
  const bool full_warp = __all_sync(0xffffffffu, mine_full);
  int32_t st = 0;
  uint32_t parity = 0;
  for (int32_t j = 0; j < k; ++j) {
    if (!mbar_wait(full + st, parity, err)) return;
    const T *xs = stage0 + (size_t)st * wcap;
    T *ycol = Y + (int64_t)j * ldy;
    if (full_warp) window_column<T, LMAX, true>(c, a, len, live, xs, ycol, row, yi);
    else window_column<T, LMAX, false>(c, a, len, live, xs, ycol, row, yi);
    mbar_arrive(empty + st);  // this thread is done reading the stage
    if (++st == nstage) {
      st = 0;
      parity ^= 1u;
    }
  }
}

// ---- the same product with two CTAs per cluster sharing their windows ------------------------------------
// Two neighbouring row blocks of a banded matrix need windows that overlap in all but 2 x 1024 entries
// (cfg 2: 3072 entries each, 4096 together).  k_spmm_window reads every window from L2 on its own -- a = 3
// times the block's rows per column -- and ncu put that kernel at the L2 -> SM ceiling.  Here
// the two CTAs of a cluster stage the UNION of their windows, the same layout in both: each producer reads
// one half of it from L2 once and the TMA engine delivers it to both CTAs (multicast), so a column costs
// 4096 entries of L2 bandwidth per 2048 rows, a = 2.
// Protocol per stage: full[st] of a CTA counts its own producer's arrive and is credited with the bytes of
// BOTH halves (its own multicast and the peer's land in it); empty[st] of a CTA counts one arrive per consumer
// warp of BOTH CTAs, because its producer overwrites the stage in both.  The cluster is synchronised once
// after the barriers are initialised and once before exit (a CTA may not leave while the peer can still
// write to it).  Every wait is bounded; a thread that gives up skips to the final synchronisation.
template <class T, int LMAX>
__global__ void __launch_bounds__(WIN_THREADS, 1)
k_spmm_window2(int32_t nrows, int32_t ncols, int32_t nslices, int32_t nstage, int32_t wcap,
               const int64_t *__restrict__ slice_ptr, const int32_t *__restrict__ rowlen,
               const int32_t *__restrict__ colidx, const T *__restrict__ vals,
               const int32_t *__restrict__ slice_cmin, const int32_t *__restrict__ slice_cmax,
               const T *__restrict__ X, int64_t ldx, T *__restrict__ Y, int64_t ldy, int32_t k,
               const YInit<T> yi, int *__restrict__ err, int mode) {
  extern __shared__ __align__(128) unsigned char win_smem[];
  T *stage0 = reinterpret_cast<T *>(win_smem);                                              // [nstage][wcap]
  uint64_t *full = reinterpret_cast<uint64_t *>(win_smem + (size_t)nstage * wcap * sizeof(T));  // [nstage]
  uint64_t *empty = full + nstage;                                                          // [nstage]
  int32_t *s_rng = reinterpret_cast<int32_t *>(empty + nstage);                             // lo, width
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t crank = cluster_ctarank();          // 0 or 1; blockIdx.x = 2 * pair + crank
  const int32_t row0 = blockIdx.x * WIN_ROWS;
  const int32_t pair_row0 = (int32_t)(blockIdx.x - crank) * WIN_ROWS;
  constexpr int32_t EPV = 16 / sizeof(T);  // elements per 16 bytes: 2 (Double) or 1 (Complex)

  // the pair's window = union of its 64 slices' ranges, 16-byte aligned in X (both CTAs compute the same)
  if (tid < 32) {
    int32_t lo = 0x7fffffff, hi = -1;
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int32_t s = pair_row0 / WIN_SELL_C + q * 32 + tid;
      if (s < nslices) {
        lo = min(lo, __ldg(slice_cmin + s));
        hi = max(hi, __ldg(slice_cmax + s));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (tid == 0) {
      if (hi < lo) {
        lo = 0;
        hi = EPV - 1;
      }
      lo = lo / EPV * EPV;
      int32_t end = (hi + EPV) / EPV * EPV;
      end = min(end, ncols);
      s_rng[0] = lo;
      s_rng[1] = end - lo;
      for (int st = 0; st < nstage; ++st) {
        mbar_init(full + st, 1);
        mbar_init(empty + st, 2 * (WIN_CONSUMERS / 32));
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  __syncthreads();
  cluster_sync_all();  // the peer's barriers exist before anything is sent to them
  const int32_t wlo = s_rng[0], width = s_rng[1];
  const bool sane = width <= wcap && width > 0;  // the same verdict in both CTAs
  if (!sane && tid == 0) *(volatile int *)err = 2;

  if (sane && tid >= WIN_CONSUMERS) {
    // ---- producer: one lane feeds this CTA's half of the pair's window to both CTAs ----------------------
    if (lane == 0) {
      const int32_t h0 = ((width / 2) + EPV - 1) / EPV * EPV;  // first half, whole 16-byte units
      const int32_t off = crank ? h0 : 0, cnt = crank ? width - h0 : h0;
      const uint32_t bytes_all = (uint32_t)width * (uint32_t)sizeof(T);
      int32_t st = 0, use = 0;
      for (int32_t j = 0; j < k; ++j) {
        if (use > 0 && !mbar_wait_cluster(empty + st, (uint32_t)((use - 1) & 1), err)) break;
        mbar_arrive_expect_tx(full + st, bytes_all);
        if (mode == 1) {  // diagnosis: no sharing, this CTA fetches the whole union for itself
          bulk_g2s(stage0 + (size_t)st * wcap, X + (int64_t)j * ldx + wlo, bytes_all, full + st);
        } else if (cnt > 0) {
          const int32_t pieces = mode == 2 ? 4 : 1;
          const int32_t per = ((cnt + pieces - 1) / pieces + EPV - 1) / EPV * EPV;
          for (int32_t o = 0; o < cnt; o += per)
            bulk_g2s_multicast(stage0 + (size_t)st * wcap + off + o, X + (int64_t)j * ldx + wlo + off + o,
                               (uint32_t)min(per, cnt - o) * (uint32_t)sizeof(T), full + st, (uint16_t)3);
        }
        if (++st == nstage) {
          st = 0;
          ++use;
        }
      }
    }
  } else if (sane) {
    // ---- consumers ---------------------------------------------------------------------------------------
    const uint64_t pol_s = policy_evict_first();
    int32_t row[WIN_RPT], len[WIN_RPT], c[WIN_RPT][LMAX];
    bool live[WIN_RPT];
    T a[WIN_RPT][LMAX];
    bool mine_full = true;
#pragma unroll
    for (int i = 0; i < WIN_RPT; ++i) {
      row[i] = row0 + i * WIN_CONSUMERS + tid;
      live[i] = row[i] < nrows;
      int64_t base = 0;
      len[i] = 0;
      if (live[i]) {
        base = __ldg(slice_ptr + (row[i] >> 5)) + lane;
        len[i] = __ldg(rowlen + row[i]);
      }
#pragma unroll
      for (int u = 0; u < LMAX; ++u) {
        const bool ok = u < len[i];
        c[i][u] = ok ? ld_stream(colidx + base + (int64_t)u * WIN_SELL_C, pol_s) - wlo : 0;
        a[i][u] = ok ? ld_stream(vals + base + (int64_t)u * WIN_SELL_C, pol_s) : Num<T>::zero();
      }
      mine_full = mine_full && live[i] && len[i] == LMAX;
    }
    const bool full_warp = __all_sync(0xffffffffu, mine_full);
    int32_t st = 0;
    uint32_t parity = 0;
    for (int32_t j = 0; j < k; ++j) {
      // warp-uniform verdict, so the whole warp leaves the loop together
      const bool landed = __all_sync(0xffffffffu, mbar_wait_cluster(full + st, parity, err));
      if (!landed) break;
      const T *xs = stage0 + (size_t)st * wcap;
      T *ycol = Y + (int64_t)j * ldy;
      if (full_warp) window_column<T, LMAX, true>(c, a, len, live, xs, ycol, row, yi);
      else window_column<T, LMAX, false>(c, a, len, live, xs, ycol, row, yi);
      __syncwarp();
      if (lane == 0) {  // this warp is done reading the stage: tell both producers
        mbar_arrive_cluster(empty + st, crank);
        mbar_arrive_cluster(empty + st, crank ^ 1u);
      }
      if (++st == nstage) {
        st = 0;
        parity ^= 1u;
      }
    }
  }
  cluster_sync_all();  // nobody leaves while the peer may still multicast into, or arrive on, this CTA
}

// ---- host side ------------------------------------------------------------------------------------------

// widest window over the row blocks of WIN_ROWS rows (one thread per block)
__global__ void k_window_max(int32_t nslices, const int32_t *__restrict__ slice_cmin,
                             const int32_t *__restrict__ slice_cmax, int32_t *__restrict__ out, int SPB) {
  const int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if ((int64_t)b * SPB >= nslices) return;
  int32_t lo = 0x7fffffff, hi = -1;
  for (int s = b * SPB; s < min(nslices, (b + 1) * SPB); ++s) {
    lo = min(lo, slice_cmin[s]);
    hi = max(hi, slice_cmax[s]);
  }
  if (hi >= lo) atomicMax(out, hi - lo + 1);
}

// Error words of the staged kernels: one word per device ordinal in one block of mapped pinned host memory
// (a kernel writes its device's word over PCIe only when a barrier wait gives up), so reading a word needs
// neither a device of its own nor a synchronising copy.  A launch is handed the word of the device it runs on.
// Who reads them: the host-operand products (slc_*_mulm, slc_*_mulm_sharded, slc_sharded_synchronize) after
// their final stream synchronisation, through spmm_window_check -- a product whose kernel gave up returns
// SLC_ERR_CUDA instead of a silently stale block, and the word is cleared so the next product starts clean;
// and slc_spmm_window_error, for callers of the asynchronous *_dev forms, which only reads.
static std::atomic<int *> g_window_err{nullptr};

static int window_err_base(int **out) {
  int *p = g_window_err.load(std::memory_order_acquire);
  if (!p) {
    int *fresh = nullptr;
    SLC_CUDA(cudaHostAlloc((void **)&fresh, MAX_DEVICES * sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable));
    for (int d = 0; d < MAX_DEVICES; ++d) fresh[d] = 0;
    int *expected = nullptr;
    if (g_window_err.compare_exchange_strong(expected, fresh, std::memory_order_acq_rel)) p = fresh;
    else { cudaFreeHost(fresh); p = expected; }
  }
  *out = p;
  return SLC_OK;
}

static int window_err_flag(int **out) {
  int *base = nullptr;
  SLC_TRY(window_err_base(&base));
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) dev = 0;
  *out = base + dev;
  return SLC_OK;
}

int spmm_window_err_word(int **out) { return window_err_flag(out); }

int spmm_window_error(int *flag) {
  *flag = 0;
  int *p = g_window_err.load(std::memory_order_acquire);
  if (!p) return SLC_OK;
  SLC_CUDA(cudaStreamSynchronize(cur_stream()));
  int any = 0;
  for (int d = 0; d < MAX_DEVICES; ++d) any |= ((volatile int *)p)[d];
  *flag = any;
  return SLC_OK;
}

// After the caller has synchronised the stream(s) of the current device: fail if a staged block product
// gave up there since the last check, and clear the word.  SLC_DEBUG_WINDOW_FAULT=<code> raises the word
// from the host first (tests of this reporting path; a real give-up needs a protocol bug).
int spmm_window_check(const char *who) {
  int *p = g_window_err.load(std::memory_order_acquire);
  const int fault = env_int("SLC_DEBUG_WINDOW_FAULT", 0);
  if (!p && !fault) return SLC_OK;
  int *word = nullptr;
  SLC_TRY(window_err_flag(&word));
  if (fault) *(volatile int *)word = fault;
  const int code = *(volatile int *)word;
  if (!code) return SLC_OK;
  *(volatile int *)word = 0;
  int dev = 0;
  (void)cudaGetDevice(&dev);
  return fail(SLC_ERR_CUDA,
              "%s: the TMA-staged block product on device %d %s; the result block is invalid", who, dev,
              code == 2 ? "met a column window wider than its plan" : "gave up waiting on a shared-memory barrier");
}

template <class T, int LMAX>
static int launch_window(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, int32_t nstage,
                         int32_t wcap, const YInit<T> &yi) {
  const SellPlan &sp = m->sell;
  const size_t smem = (size_t)nstage * wcap * sizeof(T) + (size_t)nstage * 2 * sizeof(uint64_t) + 16;
  auto kern = k_spmm_window<T, LMAX>;
  static PerDevice<size_t> smem_set;  // per instantiation, per device
  SLC_CUDA(ensure_dyn_smem(kern, smem_set, smem));
  int *err_flag = nullptr;
  SLC_TRY(window_err_flag(&err_flag));
  const unsigned nb = (unsigned)ceil_div(m->nrows, WIN_ROWS);
  kern<<<nb, WIN_THREADS, smem, cur_stream()>>>(m->nrows, m->ncols, sp.nslices, nstage, wcap, sp.slice_ptr, sp.rowlen,
                                                sp.colidx, (const T *)sp.vals, sp.slice_cmin, sp.slice_cmax, X, ldx, Y,
                                                ldy, k, yi, err_flag);
  SLC_LAUNCHED();
  return SLC_OK;
}

template <class T, int LMAX>
static int launch_window2(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, int32_t nstage,
                          int32_t wcap, const YInit<T> &yi) {
  const SellPlan &sp = m->sell;
  const size_t smem = (size_t)nstage * wcap * sizeof(T) + (size_t)nstage * 2 * sizeof(uint64_t) + 16;
  auto kern = k_spmm_window2<T, LMAX>;
  static PerDevice<size_t> smem_set;  // per instantiation, per device
  SLC_CUDA(ensure_dyn_smem(kern, smem_set, smem));
  int *err_flag = nullptr;
  SLC_TRY(window_err_flag(&err_flag));
  const unsigned nb = (unsigned)ceil_div(m->nrows, WIN_ROWS);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((nb + 1) / 2 * 2, 1, 1);  // whole clusters; a CTA past the last row only feeds its peer
  cfg.blockDim = dim3(WIN_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = cur_stream();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  SLC_CUDA(cudaLaunchKernelEx(&cfg, kern, m->nrows, m->ncols, sp.nslices, nstage, wcap, (const int64_t *)sp.slice_ptr,
                              (const int32_t *)sp.rowlen, (const int32_t *)sp.colidx, (const T *)sp.vals,
                              (const int32_t *)sp.slice_cmin, (const int32_t *)sp.slice_cmax, X, ldx, Y, ldy, k, yi,
                              err_flag, env_int("SLC_WIN2_MODE", 0)));
  SLC_LAUNCHED();
  return SLC_OK;
}

// widest window over groups of `spb` slices, computed once per matrix and group size
static int window_max_of(slc_matrix *m, int spb, int32_t *out) {
  SellPlan &sp = m->sell;
  cudaStream_t s = cur_stream();
  DevBuf<int32_t> w;
  SLC_TRY(w.alloc(1));
  SLC_CUDA(cudaMemsetAsync(w.p, 0, sizeof(int32_t), s));
  const int32_t nblocks = (int32_t)ceil_div(sp.nslices, spb);
  k_window_max<<<(unsigned)ceil_div(nblocks, 128), 128, 0, s>>>(sp.nslices, sp.slice_cmin, sp.slice_cmax, w.p, spb);
  SLC_LAUNCHED();
  int32_t h = 0;
  SLC_CUDA(cudaMemcpyAsync(&h, w.p, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  *out = h;
  return SLC_OK;
}

// Runs the product on the window kernel when the matrix and the operands qualify: sliced-ELL plan,
// rows of at most 8 entries, every 1024-row block's window of X fits a pipeline stage (at least 3
// stages in 200 KB of shared memory), X 16-byte aligned with an even column stride for Double.
// *done says whether it did.
template <class T>
int spmm_window_try(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, YInit<T> yi, bool *done) {
  *done = false;
  static const char *env = getenv("SLC_SPMM_WINDOW");
  if (env && env[0] == '0') return SLC_OK;
  SellPlan &sp = m->sell;
  if (m->spmv_active != SPMV_SELL || !sp.slice_cmin || m->max_rowlen < 1 || m->max_rowlen > 8 || k < 1)
    return SLC_OK;
  constexpr int EPV = 16 / (int)sizeof(T);
  if (((uintptr_t)X & 15) != 0 || m->ncols % EPV != 0 || ldx % EPV != 0) return SLC_OK;
  if (sp.window_max < 0) SLC_TRY(window_max_of(m, WIN_ROWS / WIN_SELL_C, &sp.window_max));  // once per matrix
  // SLC_SPMM_CLUSTER (0): 1 = two CTAs per cluster stage the union of their windows, each half read from L2 once
  // and multicast to both (k_spmm_window2), when the union still leaves three stages and is smaller than the two
  // windows it replaces.  Measured on cfg 2 (round 2, profiles/bench/r2n_cfg2_*.json): 0.515 ms with multicast,
  // 0.512 ms with the same kernel fetching both halves itself (SLC_WIN2_MODE=1, a third MORE L2 reads than one
  // CTA per window), 0.472 ms for k_spmm_window.  L2 read volume is not what bounds this product -- the shared
  // memory of an SM takes in 4096 instead of 3072 entries per column either way -- and the microarchitecture
  // notes say as much (multicast only pays from cluster size 8 up).  So one CTA per window stays the default.
  if (env_int("SLC_SPMM_CLUSTER", 0) != 0 && k >= 2 && m->nrows > WIN_ROWS) {
    if (sp.window_max2 < 0) SLC_TRY(window_max_of(m, 2 * WIN_ROWS / WIN_SELL_C, &sp.window_max2));
    const int32_t wcap2 = (int32_t)ceil_div((int64_t)sp.window_max2 + 2 * (EPV - 1), 16) * 16;
    const size_t stage2 = (size_t)wcap2 * sizeof(T);
    int32_t nstage2 = (int32_t)std::min<size_t>(8, ((size_t)200 * 1024) / stage2);
    if (sp.window_max2 > 0 && nstage2 >= 3 && (int64_t)sp.window_max2 * 4 <= (int64_t)sp.window_max * 7) {
      nstage2 = std::min(nstage2, std::max(k, 2));
      switch (m->max_rowlen) {
#define SLC_WIN2_CASE(L_) \
  case L_:                \
    SLC_TRY((launch_window2<T, L_>(m, X, ldx, Y, ldy, k, nstage2, wcap2, yi))); \
    break
        SLC_WIN2_CASE(1);
        SLC_WIN2_CASE(2);
        SLC_WIN2_CASE(3);
        SLC_WIN2_CASE(4);
        SLC_WIN2_CASE(5);
        SLC_WIN2_CASE(6);
        SLC_WIN2_CASE(7);
        SLC_WIN2_CASE(8);
#undef SLC_WIN2_CASE
        default:
          return SLC_OK;
      }
      *done = true;
      return SLC_OK;
    }
  }
  // the kernel aligns the window start down and its end up to 16 bytes: up to 2*(EPV-1) extra elements
  const int32_t wcap = (int32_t)ceil_div((int64_t)sp.window_max + 2 * (EPV - 1), 16) * 16;
  const size_t stage_bytes = (size_t)wcap * sizeof(T);
  const size_t budget = 200 * 1024;
  int32_t nstage = (int32_t)std::min<size_t>(8, budget / stage_bytes);
  if (sp.window_max <= 0 || nstage < 3) return SLC_OK;
  nstage = std::min(nstage, std::max(k, 2));
  switch (m->max_rowlen) {
#define SLC_WIN_CASE(L_) \
  case L_:               \
    SLC_TRY((launch_window<T, L_>(m, X, ldx, Y, ldy, k, nstage, wcap, yi))); \
    break
    SLC_WIN_CASE(1);
    SLC_WIN_CASE(2);
    SLC_WIN_CASE(3);
    SLC_WIN_CASE(4);
    SLC_WIN_CASE(5);
    SLC_WIN_CASE(6);
    SLC_WIN_CASE(7);
    SLC_WIN_CASE(8);
#undef SLC_WIN_CASE
    default:
      return SLC_OK;
  }
  *done = true;
  return SLC_OK;
}

template int spmm_window_try<double>(slc_matrix *, const double *, int64_t, double *, int64_t, int32_t, YInit<double>, bool *);
template int spmm_window_try<double2>(slc_matrix *, const double2 *, int64_t, double2 *, int64_t, int32_t, YInit<double2>, bool *);

}  // namespace slc

extern "C" int slc_spmm_window_error(int *flag) {
  if (!flag) return slc::fail(SLC_ERR_INVALID, "slc_spmm_window_error: null argument");
  return slc::spmm_window_error(flag);
}
