// slc_spmv.cu -- SpMV (mulV / axpy / axpy_) and SpMM (FEAST multiplyWork / mulM).
//
// Reference semantics (Data/Matrix/Sparse.hs:433-471): walk the CSC columns in order and do
// y[r] <- a*x[c] + y[r].  Seen from one output row that is: start from y0 (0 for mulV) and add
// the row's terms in increasing column order, product rounded, then sum rounded.  The kernels
// below gather along the row-major mirror instead of scattering along columns; wherever one
// thread owns a whole row it adds the terms in exactly that order, so the result is bit-identical
// to the reference; where a row is split across lanes the order is a fixed tree (deterministic,
// within the 1e-12 tolerance).
//
// Kernel families, chosen per matrix from its row-length statistics (slc_set_spmv_kernel
// overrides):
//   SELL-32        rows grouped in slices of 32, stored column-major inside the slice, one thread
//                  per row: index/value streams and the x gather are both lane-contiguous.  Used
//                  when padding costs < 12.5 % (stencils).
//   CSR balanced   fixed chunks of nonzeros per block, products staged in shared memory, rows
//                  reduced from there; rows spanning chunks are stitched by a fix-up kernel.
//                  Used for irregular (power-law) matrices.
//   CSR warp/row   one warp per row with a shuffle tree; kept as a simple cross-check.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "slc_common.cuh"
#include "slc_peer.cuh"

namespace slc {

// =================================================================================================
// plans
// =================================================================================================

constexpr int SELL_C = 32;
constexpr int32_t SELL_IRREGULAR = (int32_t)0x80000000;  // kbase: this (slice, k) keeps per-lane indices
constexpr int BAL_THREADS = 256;
constexpr int BAL_ITEMS = 8;
constexpr int BAL_CHUNK = BAL_THREADS * BAL_ITEMS;  // nonzeros per block
constexpr int MAX_PARTS = 16;

// ---- how a kernel finds x[c] ---------------------------------------------------------------------
// LocalX: the whole operand is one array in this GPU's HBM.
// DistX:  the operand is partitioned by contiguous column ranges over up to 16 GPUs of one NVSwitch
//         domain; parts[p] is this process's mapping of GPU p's slice (its own allocation for
//         p == self, a CUDA-IPC peer mapping otherwise).  A gather of a remote column is a plain
//         ld.global on the peer pointer: it travels over NVLink, is cached in this SM's L1 and
//         bypasses the local L2.  The halo "exchange" of a row-partitioned SpMV is therefore fused
//         into the product kernel: no pack kernel, no send/recv, no halo buffer.
template <class T>
struct LocalX {
  const T *x;
  __device__ __forceinline__ T operator()(int32_t c, uint64_t pol) const { return ld_gather(x + c, pol); }
};

// The partition table lives in device memory (slc_matrix::dist_table): the kernel parameter stays
// three words, and the owner search -- taken only by rows on a partition boundary -- indexes
// global memory instead of forcing the parameter block into local memory.
struct DistTable {
  int32_t nparts;
  int32_t splits[MAX_PARTS + 1];
  const void *parts[MAX_PARTS];
  int64_t ld[MAX_PARTS];  // SpMM: leading dimension of every part's block (0 for vectors)
};

template <class T>
struct DistX {
  const T *self_x;  // this rank's own slice
  int32_t self_lo, self_hi;
  const DistTable *tbl;
  __device__ __forceinline__ T operator()(int32_t c, uint64_t pol) const {
    if (c >= self_lo && c < self_hi) return ld_gather(self_x + (c - self_lo), pol);
    int q = 0;
    const int np = tbl->nparts;
#pragma unroll 1
    while (q + 1 < np && c >= tbl->splits[q + 1]) ++q;
    return ld_peer(static_cast<const T *>(tbl->parts[q]) + (c - tbl->splits[q]));
  }
};

// plain pointer behind an accessor (null when the operand is partitioned)
template <class T> static inline const T *local_ptr(const LocalX<T> &xa) { return xa.x; }
template <class T> static inline const T *local_ptr(const DistX<T> &) { return nullptr; }

// The same two accessors for a column-major block X (SpMM): col0(c, ld) is the address of X[c, 0] and
// ld the distance to X[c, 1] -- every part keeps its own leading dimension.
template <class T>
struct LocalXM {
  const T *X;
  int64_t ldx;
  __device__ __forceinline__ const T *col0(int32_t c, int64_t &ld) const {
    ld = ldx;
    return X + c;
  }
};

template <class T>
struct DistXM {
  const T *self_X;  // this rank's own row block of X
  int64_t self_ld;
  int32_t self_lo, self_hi;
  const DistTable *tbl;
  // remote: the entry lives in a peer's HBM (read it with ld_peer, not through the read-only path)
  __device__ __forceinline__ const T *col0(int32_t c, int64_t &ld, bool &remote) const {
    ld = self_ld;
    const T *p = self_X + (c - self_lo);
    remote = c < self_lo || c >= self_hi;
    if (remote) {
      int q = 0;
      const int np = tbl->nparts;
#pragma unroll 1
      while (q + 1 < np && c >= tbl->splits[q + 1]) ++q;
      p = static_cast<const T *>(tbl->parts[q]) + (c - tbl->splits[q]);
      ld = tbl->ld[q];
    }
    return p;
  }
};

// one warp per slice: row lengths and the slice width (elements = width * 32)
__global__ void k_sell_widths(int32_t nrows, int32_t nslices, const int32_t *__restrict__ rowptr,
                              int32_t *__restrict__ rowlen, int64_t *__restrict__ slice_elems,
                              int32_t *__restrict__ max_len) {
  const int32_t slice = (int32_t)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (slice >= nslices) return;
  const int32_t row = slice * SELL_C + lane;
  int32_t len = 0;
  if (row < nrows) {
    len = rowptr[row + 1] - rowptr[row];
    rowlen[row] = len;
  }
  int32_t w = len;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
  if (lane == 0) {
    slice_elems[slice] = (int64_t)w * SELL_C;
    atomicMax(max_len, w);
  }
  if (slice == 0 && lane == 0) slice_elems[nslices] = 0;
}

template <class T>
__global__ void k_sell_fill(int32_t nrows, int32_t nslices, const int32_t *__restrict__ rowptr,
                            const int32_t *__restrict__ colidx, const T *__restrict__ vals,
                            const int64_t *__restrict__ slice_ptr, int32_t *__restrict__ s_col,
                            T *__restrict__ s_val, int32_t *__restrict__ slice_cmin,
                            int32_t *__restrict__ slice_cmax, int32_t *__restrict__ kbase,
                            unsigned long long *__restrict__ unit_entries) {
  const int32_t slice = (int32_t)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (slice >= nslices) return;
  const int32_t row = slice * SELL_C + lane;
  const int64_t base = slice_ptr[slice];
  const int32_t width = (int32_t)((slice_ptr[slice + 1] - base) / SELL_C);
  int32_t start = 0, len = 0;
  if (row < nrows) {
    start = rowptr[row];
    len = rowptr[row + 1] - start;
  }
  int32_t cmin = 0x7fffffff, cmax = -1;
  unsigned long long units = 0;
  for (int32_t k = 0; k < width; ++k) {
    const int64_t at = base + (int64_t)k * SELL_C + lane;
    int32_t c = 0;
    if (k < len) {
      c = colidx[start + k];
      s_col[at] = c;
      s_val[at] = vals[start + k];
      cmin = min(cmin, c);
      cmax = max(cmax, c);
    } else {  // padding is never read by the product kernels (they stop at rowlen)
      s_col[at] = 0;
      s_val[at] = Num<T>::zero();
    }
    // unit stride: every lane that has an entry k sits at column (common base) + lane
    const unsigned have = __ballot_sync(0xffffffffu, k < len);
    int32_t kb = 0;
    if (have) {
      const int32_t mine = c - lane;
      const int32_t first = __shfl_sync(0xffffffffu, mine, __ffs(have) - 1);
      const bool same = __all_sync(0xffffffffu, k >= len || mine == first);
      kb = same ? first : SELL_IRREGULAR;
      if (same) units += __popc(have);
    }
    if (lane == 0) kbase[base / SELL_C + k] = kb;
  }
  if (lane == 0 && units) atomicAdd(unit_entries, units);
  // column range of the slice: lets a distributed product find the rows that need no peer
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cmin = min(cmin, __shfl_xor_sync(0xffffffffu, cmin, o));
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
  }
  if (lane == 0) {
    slice_cmin[slice] = cmin;
    slice_cmax[slice] = cmax;
  }
}

// first row whose start is >= chunk start (rows that begin inside chunk b are owned by b)
__global__ void k_bal_first_owned(int32_t nrows, int64_t nnz, int32_t nchunks,
                                  const int32_t *__restrict__ rowptr, int32_t *__restrict__ first_owned) {
  const int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nchunks) return;
  if (b == nchunks) {
    first_owned[b] = nrows;
    return;
  }
  const int64_t start = (int64_t)b * BAL_CHUNK;
  int32_t lo = 0, hi = nrows;  // lower_bound over rowptr[0..nrows)
  while (lo < hi) {
    int32_t mid = lo + ((hi - lo) >> 1);
    if ((int64_t)rowptr[mid] < start) lo = mid + 1;
    else hi = mid;
  }
  first_owned[b] = lo;
  (void)nnz;
}

template <class T>
static int build_sell(slc_matrix *m) {
  SellPlan &sp = m->sell;
  if (sp.built) return SLC_OK;
  cudaStream_t s = cur_stream();
  const int32_t nslices = (int32_t)ceil_div(m->nrows, SELL_C);
  DevBuf<int64_t> elems, sptr;
  DevBuf<int32_t> rowlen, maxlen;
  SLC_TRY(elems.alloc((size_t)nslices + 1));
  SLC_TRY(sptr.alloc((size_t)nslices + 1));
  SLC_TRY(rowlen.alloc((size_t)m->nrows));
  SLC_TRY(maxlen.alloc(1));
  SLC_CUDA(cudaMemsetAsync(maxlen.p, 0, sizeof(int32_t), s));
  SLC_CUDA(cudaMemsetAsync(elems.p, 0, ((size_t)nslices + 1) * sizeof(int64_t), s));
  if (nslices) {
    k_sell_widths<<<(unsigned)ceil_div((int64_t)nslices * 32, 256), 256, 0, s>>>(
        m->nrows, nslices, m->rowptr, rowlen.p, elems.p, maxlen.p);
    SLC_LAUNCHED();
  }
  SLC_TRY(exclusive_scan_i64(elems.p, sptr.p, (int64_t)nslices + 1));
  int64_t padded = 0;
  int32_t hmax = 0;
  SLC_CUDA(cudaMemcpyAsync(&padded, sptr.p + nslices, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaMemcpyAsync(&hmax, maxlen.p, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  m->max_rowlen = hmax;
  sp.nslices = nslices;
  sp.padded = padded;
  sp.slice_ptr = sptr.take();
  sp.rowlen = rowlen.take();
  sp.built = true;  // widths known; arrays filled lazily by fill_sell
  return SLC_OK;
}

template <class T>
static int fill_sell(slc_matrix *m) {
  SellPlan &sp = m->sell;
  if (sp.colidx) return SLC_OK;
  cudaStream_t s = cur_stream();
  DevBuf<int32_t> sc, cmin, cmax;
  DevBuf<T> sv;
  SLC_TRY(sc.alloc((size_t)sp.padded));
  SLC_TRY(sv.alloc((size_t)sp.padded));
  SLC_TRY(cmin.alloc((size_t)sp.nslices));
  SLC_TRY(cmax.alloc((size_t)sp.nslices));
  DevBuf<int32_t> kb;
  DevBuf<unsigned long long> units;
  SLC_TRY(kb.alloc((size_t)(sp.padded / SELL_C) + 16));  // + 16: a batch may look past the last slice
  SLC_TRY(units.alloc(1));
  SLC_CUDA(cudaMemsetAsync(kb.p, 0, ((size_t)(sp.padded / SELL_C) + 16) * sizeof(int32_t), s));
  SLC_CUDA(cudaMemsetAsync(units.p, 0, sizeof(unsigned long long), s));
  if (sp.nslices) {
    k_sell_fill<T><<<(unsigned)ceil_div((int64_t)sp.nslices * 32, 256), 256, 0, s>>>(
        m->nrows, sp.nslices, m->rowptr, m->colidx, (const T *)m->rvals, sp.slice_ptr, sc.p, sv.p, cmin.p,
        cmax.p, kb.p, units.p);
    SLC_LAUNCHED();
    unsigned long long h_units = 0;
    SLC_CUDA(cudaMemcpyAsync(&h_units, units.p, sizeof h_units, cudaMemcpyDeviceToHost, s));
    SLC_CUDA(cudaStreamSynchronize(s));
    sp.unit_entries = (int64_t)h_units;
    // Used when at least a quarter of the entries qualify and the matrix streams from HBM; a matrix
    // that stays in L2 gains nothing from fewer bytes and pays for the extra words (cfg 1: 16.7 µs with,
    // 14.5 µs without).  SLC_SELL_KBASE=0 keeps the per-entry indices, =2 compresses at any size (tests).
    const int kb_env = env_int("SLC_SELL_KBASE", 1);
    const bool streams = (size_t)sp.padded * (sizeof(T) + sizeof(int32_t)) > ((size_t)96 << 20);
    if (kb_env && (streams || kb_env == 2) && sp.unit_entries * 4 >= m->nnz) sp.kbase = kb.take();
  }
  sp.colidx = sc.take();
  sp.vals = sv.take();
  sp.slice_cmin = cmin.take();
  sp.slice_cmax = cmax.take();
  return SLC_OK;
}

static int build_balanced(slc_matrix *m) {
  BalancedPlan &bp = m->bal;
  if (bp.built) return SLC_OK;
  const int32_t nchunks = (int32_t)ceil_div(m->nnz, BAL_CHUNK);
  DevBuf<int32_t> fo;
  SLC_TRY(fo.alloc((size_t)nchunks + 1));
  k_bal_first_owned<<<(unsigned)ceil_div((int64_t)nchunks + 1, 256), 256, 0, cur_stream()>>>(
      m->nrows, m->nnz, nchunks, m->rowptr, fo.p);
  SLC_LAUNCHED();
  bp.nchunks = nchunks;
  bp.chunk_row = fo.take();
  bp.built = true;
  return SLC_OK;
}

__global__ void k_hot_remap(int64_t nnz, const int32_t *__restrict__ colidx, const int32_t *__restrict__ slot_of,
                            int32_t *__restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nnz) return;
  const int32_t c = colidx[e];
  const int32_t sl = __ldg(slot_of + c);
  out[e] = sl >= 0 ? (int32_t)(0x80000000u | (uint32_t)sl) : c;
}

// Hot columns of a balanced plan: the `cap` columns with the most entries (degrees come from the
// canonical CSC pointers; selection on the host, once per matrix).  Kept only when they hold at
// least a tenth of the nonzeros and x is much longer than the table.
static int build_hot(slc_matrix *m, int32_t cap) {
  BalancedPlan &bp = m->bal;
  if (bp.hot_k >= 0) return SLC_OK;
  bp.hot_k = 0;
  // read when the plan is built (not cached): SLC_BAL_HOT=0 switches the table off, SLC_BAL_HOT_K=n
  // caps it at n slots and drops the size thresholds (tests use it on small matrices)
  const int hot_env = env_int("SLC_BAL_HOT", 1);
  const int cap_env = env_int("SLC_BAL_HOT_K", 0);
  if (cap_env > 0) cap = std::min(cap, cap_env);
  if (!hot_env || cap < 1 || m->ncols <= cap) return SLC_OK;
  if (cap_env <= 0 && ((int64_t)m->ncols < (int64_t)cap * 16 || m->nnz < (1 << 20))) return SLC_OK;
  cudaStream_t s = cur_stream();
  std::vector<int32_t> cp((size_t)m->ncols + 1);
  SLC_CUDA(cudaMemcpyAsync(cp.data(), m->colptr, cp.size() * 4, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  std::vector<int32_t> order((size_t)m->ncols);
  for (int32_t c = 0; c < m->ncols; ++c) order[(size_t)c] = c;
  auto denser = [&](int32_t a, int32_t b) {
    const int32_t da = cp[(size_t)a + 1] - cp[(size_t)a], db = cp[(size_t)b + 1] - cp[(size_t)b];
    return da != db ? da > db : a < b;
  };
  std::nth_element(order.begin(), order.begin() + cap, order.end(), denser);
  order.resize((size_t)cap);
  std::sort(order.begin(), order.end());
  int64_t covered = 0;
  for (int32_t c : order) covered += cp[(size_t)c + 1] - cp[(size_t)c];
  bp.hot_share = (double)covered / (double)m->nnz;
  if (bp.hot_share < 0.10 && cap_env <= 0) return SLC_OK;
  std::vector<int32_t> slot_of((size_t)m->ncols, -1);
  for (int32_t i = 0; i < cap; ++i) slot_of[(size_t)order[(size_t)i]] = i;
  DevBuf<int32_t> d_slot, d_hot, d_idx;
  SLC_TRY(d_slot.alloc(slot_of.size()));
  SLC_TRY(d_hot.alloc((size_t)cap));
  SLC_TRY(d_idx.alloc((size_t)m->nnz));
  SLC_CUDA(cudaMemcpyAsync(d_slot.p, slot_of.data(), slot_of.size() * 4, cudaMemcpyHostToDevice, s));
  SLC_CUDA(cudaMemcpyAsync(d_hot.p, order.data(), (size_t)cap * 4, cudaMemcpyHostToDevice, s));
  k_hot_remap<<<(unsigned)ceil_div(m->nnz, 256), 256, 0, s>>>(m->nnz, m->colidx, d_slot.p, d_idx.p);
  SLC_LAUNCHED();
  SLC_CUDA(cudaStreamSynchronize(s));  // the host vectors go out of scope
  bp.hot_cols = d_hot.take();
  bp.colidx_hot = d_idx.take();
  bp.hot_k = cap;
  return SLC_OK;
}

// Build the mirror, measure the rows and settle on a kernel family.
template <class T>
static int ensure_plan(slc_matrix *m) {
  SLC_TRY(ensure_csr(m));
  SLC_TRY(build_sell<T>(m));  // cheap: gives padded size and max row length
  int kind = m->spmv_kind;
  if (kind == SPMV_AUTO) {
    // Sliced ELL streams `padded` entries instead of nnz; accept up to 12.5 % extra traffic.
    const bool regular = m->sell.padded <= m->nnz + m->nnz / 8 + 1024;
    kind = regular ? SPMV_SELL : SPMV_CSR_BALANCED;
  }
  if (kind == SPMV_SELL) SLC_TRY(fill_sell<T>(m));
  if (kind == SPMV_CSR_BALANCED || kind == SPMV_CSR_MERGE) SLC_TRY(build_balanced(m));
  m->spmv_active = kind;
  return SLC_OK;
}

// =================================================================================================
// SpMV kernels
// =================================================================================================

// ---- SELL-32, one thread per row ----------------------------------------------------------------
// Per k the 32 lanes of a warp read 32 consecutive indices (128 B) and 32 consecutive values
// (256/512 B); on a stencil the gathered x entries are consecutive too.  U entries are fetched
// before any is used, so each thread keeps U index + U value loads in flight.
template <class T, int U, bool ACC, class XA, bool KB = false>
__device__ __forceinline__ void sell_row(int32_t row, const int64_t *__restrict__ slice_ptr,
                                         const int32_t *__restrict__ rowlen,
                                         const int32_t *__restrict__ colidx, const T *__restrict__ vals,
                                         const XA &xa, T *__restrict__ y, const int32_t *__restrict__ kbase = nullptr,
                                         int pf = 0) {
  const uint64_t pol_s = policy_evict_first();
  const uint64_t pol_x = policy_evict_last();
  const int lane = threadIdx.x & 31;
  const int64_t sbase = __ldg(slice_ptr + (row >> 5));
  const int64_t base = sbase + lane;
  const int32_t len = __ldg(rowlen + row);
  const int32_t *ci = colidx + base;
  const T *va = vals + base;
  const int32_t *kb = KB ? kbase + (sbase >> 5) : nullptr;  // one word per (slice, k): same address for the warp
  T acc = ACC ? y[row] : Num<T>::zero();
  for (int32_t k = 0; k < len; k += U) {
    int32_t c[U];
    T a[U], xv[U];
    bool regular = false;
    if (KB) {
      // Compressed indices: one word per (slice, k) says "column = word + lane" for every row of the
      // slice, so a regular batch reads 4 U bytes of indices per warp instead of 128 U.  A batch with
      // an irregular (slice, k) in it reads its per-lane indices like the uncompressed kernel (the
      // branch is warp-uniform: every lane loads the same words).  Handing the words round by shuffle
      // from one coalesced load per slice measured slower (0.417 against 0.379 ms on cfg 5).
      int32_t b[U];
      regular = true;
      if (pf > 0) {  // pull the values `pf` batches ahead into L2 (the stream is perfectly predictable)
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (k + pf * U + u < len)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(va + (int64_t)(k + pf * U + u) * SELL_C));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        b[u] = __ldg(kb + k + u);  // the array is padded by 16 words past the last slice
        regular = regular && b[u] != SELL_IRREGULAR;
      }
      if (regular) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          a[u] = (k + u < len) ? ld_stream(va + (int64_t)(k + u) * SELL_C, pol_s) : Num<T>::zero();
          c[u] = b[u] + lane;
        }
      }
    }
    if (!regular) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const bool ok = k + u < len;
        c[u] = ok ? ld_stream(ci + (int64_t)(k + u) * SELL_C, pol_s) : 0;
        a[u] = ok ? ld_stream(va + (int64_t)(k + u) * SELL_C, pol_s) : Num<T>::zero();
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) xv[u] = (k + u < len) ? xa(c[u], pol_x) : Num<T>::zero();
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (k + u < len) acc = Num<T>::add(Num<T>::mul(a[u], xv[u]), acc);
  }
  st_stream(y + row, acc);
}

// PDL prologue of the SELL kernels: let the next kernel of the stream start launching, pull this row
// block's values towards L2 (the matrix is immutable, so this may overlap the previous kernel's tail),
// then wait for the previous kernel.  One 128-byte line per lane that starts one; only the blocks of the
// first wave prefetch (later blocks start in the steady state, where the one-batch-ahead prefetch of
// sell_row is enough).
template <class T>
__device__ __forceinline__ void sell_pdl_prologue(int32_t row, int32_t row_end, int32_t first_wave,
                                                  const int64_t *__restrict__ slice_ptr,
                                                  const int32_t *__restrict__ rowlen, const T *__restrict__ vals) {
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31;
  if ((int32_t)blockIdx.x < first_wave && row < row_end && (lane * sizeof(T)) % 128 == 0) {
    const T *va = vals + __ldg(slice_ptr + (row >> 5)) + lane;
    const int32_t len = __ldg(rowlen + row);
    for (int32_t k = 0; k < len; ++k) prefetch_l2(va + (int64_t)k * SELL_C);
  }
  pdl_wait();
}

template <class T, int U, bool ACC, class XA, bool KB = false, int MINB = 0>
__global__ void __launch_bounds__(256, MINB)
k_spmv_sell(int32_t row_begin, int32_t row_end, const int64_t *__restrict__ slice_ptr,
            const int32_t *__restrict__ rowlen, const int32_t *__restrict__ colidx,
            const T *__restrict__ vals, const XA xa, T *__restrict__ y, const int32_t *__restrict__ kbase = nullptr,
            int pf = 0, int first_wave = 0) {
  const int32_t row = row_begin + blockIdx.x * blockDim.x + threadIdx.x;  // row_begin % 32 == 0
  sell_pdl_prologue<T>(row, row_end, first_wave, slice_ptr, rowlen, vals);
  if (row >= row_end) return;
  sell_row<T, U, ACC, XA, KB>(row, slice_ptr, rowlen, colidx, vals, xa, y, kbase, pf);
}

// ---- SELL-32, row-partitioned over the GPUs of one box: ONE launch per product ---------------------
// The grid is  [low boundary blocks | high boundary blocks | interior blocks].  Interior rows (all
// columns inside this rank's slice of x) run the undistributed row body on a shifted base pointer.
// Boundary rows gather some columns from a peer's HBM over NVLink (DistX).  The ordering between
// ranks is fused in (slc_peer.cuh):
//   * block 0 publishes this rank's epoch into the peers' flag arrays (release, system scope):
//     "my slice of x is written" -- true when the kernel starts, by stream order;
//   * a boundary block first waits (acquire, bounded spin) until the peers ITS rows read from have
//     published the epoch: the rows below the interior wait for `wait_lo` (for a z-slab: rank-1), the
//     rows above it for `wait_hi` (rank+1).  Round 1 made every boundary block wait for all N-1 peers,
//     which turned each product into a global rendezvous (8 GPUs: 4.4x instead of 6x);
//   * the last block to finish advances the device-resident epoch, so the launch carries no
//     per-product host state and K products can sit in one CUDA graph.
// Boundary blocks have the lowest block indices, so they are resident from the start: their wait
// and their long-latency peer gathers overlap the interior blocks that fill the rest of the GPU.
// A wait that gives up (5 s) records the epoch in a host-visible word and its rows are filled with
// NaN: the next call on the group reports it (peer_check).
template <class T> __device__ __forceinline__ T poison();
template <> __device__ __forceinline__ double poison<double>() { return __longlong_as_double(0x7ff8000000000000ll); }
template <> __device__ __forceinline__ double2 poison<double2>() { return double2{poison<double>(), poison<double>()}; }

template <class T, int U, bool KB>
__global__ void __launch_bounds__(256, sizeof(T) == 8 ? (KB ? 8 : 6) : 4)
k_spmv_sell_dist(int32_t nrows, int32_t r0, int32_t r1, const int64_t *__restrict__ slice_ptr,
                 const int32_t *__restrict__ rowlen, const int32_t *__restrict__ colidx,
                 const T *__restrict__ vals, const DistX<T> xa, T *__restrict__ y, const PeerSync ps,
                 const uint32_t wait_lo, const uint32_t wait_hi, const int32_t *__restrict__ kbase,
                 const int32_t lead, const int32_t first_wave) {
  {  // PDL prologue on the rows this block would own if the grid were all interior (a prefetch hint only)
    const int32_t prow = min((int32_t)(blockIdx.x * 256 + threadIdx.x), nrows - 1);
    sell_pdl_prologue<T>(prow, nrows, first_wave, slice_ptr, rowlen, vals);
  }
  const int32_t nbl = (r0 + 255) >> 8;            // boundary blocks below the interior
  const int32_t nbh = (nrows - r1 + 255) >> 8;    // and above it
  // Block order: [`lead` interior blocks | low boundary | high boundary | the other interior blocks].
  // Blocks are dispatched in index order, so the boundary blocks -- which spin until the neighbours'
  // signal of this epoch has arrived -- become resident only after `lead` interior blocks have been
  // handed out.  With the boundary blocks first (round 2, first version) a neighbour that starts its
  // product d microseconds later left a quarter of this GPU's block slots spinning for d (measured at 8
  // GPUs: mean wait 24 us of a 52 us product); now a skew shorter than the lead costs nothing.  Block 0
  // (interior unless lead = 0) still sends the signals at once.
  const int32_t nbd = nbl + nbh;
  int32_t b = blockIdx.x;                          // index in the order [boundary | interior] used below
  if (b < lead) b += nbd;
  else if (b < lead + nbd) b -= lead;
  const bool boundary = b < nbd;
  __shared__ int s_timed_out;
  bool dead = false;
  if (ps.flags != nullptr) {
    if (blockIdx.x == 0 || boundary) {
      const unsigned long long epoch = peer_epoch(ps);
      const int p = threadIdx.x;
      if (threadIdx.x == 0) s_timed_out = 0;
      __syncthreads();
      if (p < ps.nparts && p != ps.self) {
        if (blockIdx.x == 0 && ((ps.signal_mask >> p) & 1u)) peer_signal_one(ps, p, epoch);
        const uint32_t wm = boundary ? (b < nbl ? wait_lo : wait_hi) : 0u;
        if (((wm >> p) & 1u) && !peer_wait_one(ps, p, epoch)) s_timed_out = 1;
      }
      __syncthreads();
      dead = s_timed_out != 0;
    }
  }
  if (boundary) {
    int32_t row, limit;
    if (b < nbl) {
      row = b * 256 + threadIdx.x;
      limit = r0;
    } else {
      row = r1 + (b - nbl) * 256 + threadIdx.x;
      limit = nrows;
    }
    if (row < limit) {
      if (dead) y[row] = poison<T>();
      else sell_row<T, 4, false, DistX<T>, KB>(row, slice_ptr, rowlen, colidx, vals, xa, y, kbase, 0);
    }
  } else {
    const int32_t row = r0 + (b - nbl - nbh) * 256 + threadIdx.x;
    if (row < r1)
      sell_row<T, U, false, LocalX<T>, KB>(row, slice_ptr, rowlen, colidx, vals, LocalX<T>{xa.self_x - xa.self_lo}, y,
                                           kbase, 1);
  }
  if (ps.flags != nullptr) peer_block_done(ps, gridDim.x);
}

// ---- CSR, one warp per row ----------------------------------------------------------------------
template <class T>
__device__ __forceinline__ T warp_sum(T v);
template <>
__device__ __forceinline__ double warp_sum<double>(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = __dadd_rn(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}
template <>
__device__ __forceinline__ double2 warp_sum<double2>(double2 v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v.x = __dadd_rn(v.x, __shfl_down_sync(0xffffffffu, v.x, o));
    v.y = __dadd_rn(v.y, __shfl_down_sync(0xffffffffu, v.y, o));
  }
  return v;
}

template <class T, bool ACC, class XA>
__global__ void __launch_bounds__(256)
k_spmv_csr_warp(int32_t nrows, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                const T *__restrict__ vals, const XA xa, T *__restrict__ y) {
  const int32_t row = (int32_t)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= nrows) return;
  const uint64_t pol_s = policy_evict_first();
  const uint64_t pol_x = policy_evict_last();
  const int32_t beg = rowptr[row], end = rowptr[row + 1];
  T acc = Num<T>::zero();
  for (int32_t p = beg + lane; p < end; p += 32) {
    const int32_t c = ld_stream(colidx + p, pol_s);
    const T a = ld_stream(vals + p, pol_s);
    acc = Num<T>::add(Num<T>::mul(a, xa(c, pol_x)), acc);
  }
  acc = warp_sum<T>(acc);
  if (lane == 0) y[row] = ACC ? Num<T>::add(acc, y[row]) : acc;
}

// ---- CSR, nnz-balanced --------------------------------------------------------------------------
// Block b owns nonzeros [b*CHUNK, (b+1)*CHUNK).  Phase 1: coalesced loads, products into shared
// memory.  Phase 2: the rows that START inside the chunk are reduced from shared memory (a thread
// per row when the chunk holds many rows, a warp per row when it holds few); the piece of a row
// that started in an earlier chunk goes to part_head[b], the piece of a row that runs past the
// chunk goes to part_tail[b].  k_spmv_bal_fixup then adds the pieces of each spanning row in order.
// The two phases of a chunk, shared by k_spmv_bal (one chunk per block) and k_spmv_bal_hot (a
// resident block walks many chunks).  tid = thread within the BAL_THREADS-wide group that owns chunk b.
template <class T, class XA>
__device__ __forceinline__ void bal_products(int32_t b, int tid, int64_t nnz, uint32_t zmask, int batch,
                                             const int32_t *__restrict__ colidx, const T *__restrict__ vals,
                                             const XA &xa, T *prod) {
  const int64_t start = (int64_t)b * BAL_CHUNK;
  const int64_t end = min(nnz, start + BAL_CHUNK);
  const int32_t cnt = (int32_t)(end - start);
  const uint64_t pol_s = policy_evict_first();
  const uint64_t pol_x = policy_evict_last();

  {
    int32_t c[BAL_ITEMS];
    T a[BAL_ITEMS];
#pragma unroll
    for (int i = 0; i < BAL_ITEMS; ++i) {
      const int32_t e = i * BAL_THREADS + tid;
      const bool ok = e < cnt;
      c[i] = ok ? ld_stream(colidx + start + e, pol_s) : 0;
      a[i] = ok ? ld_stream(vals + start + e, pol_s) : Num<T>::zero();
    }
    if (batch && cnt == BAL_CHUNK) {  // full chunk: the 8 gathers of a thread are issued together (probe/tie)
      T xv[BAL_ITEMS];
#pragma unroll
      for (int i = 0; i < BAL_ITEMS; ++i) xv[i] = xa(c[i], pol_x);
      uint32_t r = 0;
#pragma unroll
      for (int i = 0; i < BAL_ITEMS; ++i) r |= probe(xv[i]);
      tie(xv[0], r & zmask);
#pragma unroll
      for (int i = 0; i < BAL_ITEMS; ++i) prod[i * BAL_THREADS + tid] = Num<T>::mul(a[i], xv[i]);
    } else {
#pragma unroll
      for (int i = 0; i < BAL_ITEMS; ++i) {
        const int32_t e = i * BAL_THREADS + tid;
        if (e < cnt) prod[e] = Num<T>::mul(a[i], xa(c[i], pol_x));
      }
    }
  }
}

template <class T, bool ACC>
__device__ __forceinline__ void bal_reduce(int32_t b, int tid, int64_t nnz, const int32_t *__restrict__ rowptr,
                                           const int32_t *__restrict__ first_owned, const T *prod,
                                           T *__restrict__ y, T *__restrict__ part_head, T *__restrict__ part_tail) {
  const int64_t start = (int64_t)b * BAL_CHUNK;
  const int64_t end = min(nnz, start + BAL_CHUNK);

  const int32_t fo = first_owned[b];
  const int32_t fo_next = first_owned[b + 1];
  // piece of the row that began before this chunk
  const int64_t head_end = min((int64_t)rowptr[fo], end);  // rowptr[nrows] == nnz
  const int32_t head_cnt = (int32_t)(head_end - start);   // may be 0
  const int32_t nown = fo_next - fo;

  if (nown >= 64) {  // many short rows: thread per row, terms added in column order
    for (int32_t r = fo + tid; r < fo_next; r += BAL_THREADS) {
      const int64_t rb = rowptr[r], re = rowptr[r + 1];
      const int64_t se = min(re, end);
      const bool complete = re <= end;
      T acc = (ACC && complete) ? y[r] : Num<T>::zero();
      for (int64_t p = rb; p < se; ++p) acc = Num<T>::add(prod[p - start], acc);
      if (complete) y[r] = acc;
      else part_tail[b] = acc;
    }
    if (tid == BAL_THREADS - 1) {  // head piece, sequential (short by construction here)
      T acc = Num<T>::zero();
      for (int32_t p = 0; p < head_cnt; ++p) acc = Num<T>::add(prod[p], acc);
      part_head[b] = acc;
    }
  } else {  // few long rows: warp per segment, segment 0 is the head piece
    const int warp = tid >> 5, lane = tid & 31;
    for (int32_t sgm = warp; sgm <= nown; sgm += BAL_THREADS / 32) {
      int64_t sb, se;
      bool complete = false;
      int32_t r = -1;
      if (sgm == 0) {
        sb = start;
        se = head_end;
      } else {
        r = fo + sgm - 1;
        sb = rowptr[r];
        const int64_t re = rowptr[r + 1];
        se = min(re, end);
        complete = re <= end;
      }
      T acc = Num<T>::zero();
      for (int64_t p = sb + lane; p < se; p += 32) acc = Num<T>::add(prod[p - start], acc);
      acc = warp_sum<T>(acc);
      if (lane == 0) {
        if (sgm == 0) part_head[b] = acc;
        else if (complete) y[r] = ACC ? Num<T>::add(acc, y[r]) : acc;
        else part_tail[b] = acc;
      }
    }
  }
}

template <class T, bool ACC, class XA, int MINB = 0>
__global__ void __launch_bounds__(BAL_THREADS, MINB)
k_spmv_bal(int32_t nrows, int64_t nnz, uint32_t zmask, int batch, const int32_t *__restrict__ rowptr,
           const int32_t *__restrict__ colidx, const T *__restrict__ vals,
           const int32_t *__restrict__ first_owned, const XA xa, T *__restrict__ y,
           T *__restrict__ part_head, T *__restrict__ part_tail) {
  __shared__ T prod[BAL_CHUNK];
  bal_products<T, XA>(blockIdx.x, threadIdx.x, nnz, zmask, batch, colidx, vals, xa, prod);
  __syncthreads();
  bal_reduce<T, ACC>(blockIdx.x, threadIdx.x, nnz, rowptr, first_owned, prod, y, part_head, part_tail);
}

// Power-law matrices: a few thousand columns carry a large share of the nonzeros (R-MAT scale 24: the
// 16384 densest columns hold about a third), but they are scattered over x, so each one costs a whole
// 32-byte sector of L1 and the gathers miss (ncu, round 1: 4.6 % L1 hits, one L2 sector request per
// nonzero -- the request rate, not HBM, bounds k_spmv_bal).  Here a resident block of HOT_GROUPS x
// BAL_THREADS threads copies x at the plan's hot columns into shared memory once and then walks its
// chunks; colidx_hot holds (0x80000000 | slot) for a hot column, so those gathers are shared-memory
// reads.  Arithmetic and summation order are k_spmv_bal's.
// groups (chunks in flight) per resident block: 4 for Double; 3 for Complex Double, whose chunk body
// needs more than the 64 registers a 1024-thread block leaves
template <class T> struct HotGroups { static constexpr int value = sizeof(T) == 8 ? 4 : 3; };

template <class T>
struct HotX {
  const T *x;
  const T *hot;  // shared memory
  __device__ __forceinline__ T operator()(int32_t c, uint64_t pol) const {
    if (c < 0) return hot[c & 0x7fffffff];
    return ld_gather(x + c, pol);
  }
};

// The same table kept in global memory instead (mode 2): x at the hot columns is packed into a small
// contiguous array before the product, so the hot gathers share a few hundred L1 lines instead of one
// sector each; the product kernel is the ordinary k_spmv_bal.
template <class T>
struct HotPackedX {
  const T *x;
  const T *xhot;  // [hot_k], global
  __device__ __forceinline__ T operator()(int32_t c, uint64_t pol) const {
    if (c < 0) return __ldg(xhot + (c & 0x7fffffff));
    return ld_gather(x + c, pol);
  }
};

template <class T>
__global__ void k_hot_pack(int32_t hot_k, const int32_t *__restrict__ hot_cols, const T *__restrict__ x,
                           T *__restrict__ xhot) {
  const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < hot_k) xhot[i] = x[hot_cols[i]];
}

template <class T, bool ACC>
__global__ void __launch_bounds__(HotGroups<T>::value *BAL_THREADS, 1)
k_spmv_bal_hot(int32_t nrows, int64_t nnz, int32_t nchunks, uint32_t zmask, int batch,
               const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx_hot,
               const T *__restrict__ vals, const int32_t *__restrict__ first_owned, const T *__restrict__ x,
               const int32_t *__restrict__ hot_cols, int32_t hot_k, T *__restrict__ y, T *__restrict__ part_head,
               T *__restrict__ part_tail) {
  extern __shared__ __align__(16) unsigned char hot_smem[];
  const int HOT_GROUPS = blockDim.x / BAL_THREADS;  // <= HotGroups<T>::value
  T *prod_all = reinterpret_cast<T *>(hot_smem);               // [HOT_GROUPS][BAL_CHUNK]
  T *hot = prod_all + (size_t)HOT_GROUPS * BAL_CHUNK;         // [hot_k]
  const uint64_t pol_x = policy_evict_last();
  for (int32_t i = threadIdx.x; i < hot_k; i += blockDim.x) hot[i] = ld_gather(x + __ldg(hot_cols + i), pol_x);
  __syncthreads();
  const int grp = threadIdx.x / BAL_THREADS, tid = threadIdx.x % BAL_THREADS;
  T *prod = prod_all + (size_t)grp * BAL_CHUNK;
  const HotX<T> xa{x, hot};
  for (int64_t base = (int64_t)blockIdx.x * HOT_GROUPS; base < nchunks; base += (int64_t)gridDim.x * HOT_GROUPS) {
    const int32_t b = (int32_t)base + grp;
    const bool live = b < nchunks;
    if (live) bal_products<T, HotX<T>>(b, tid, nnz, zmask, batch, colidx_hot, vals, xa, prod);
    __syncthreads();
    if (live) bal_reduce<T, ACC>(b, tid, nnz, rowptr, first_owned, prod, y, part_head, part_tail);
    __syncthreads();  // prod is rewritten by the next chunk
  }
}

// One thread per chunk: if the last row owned by chunk b runs past the chunk, add its pieces.
template <class T, bool ACC>
__global__ void k_spmv_bal_fixup(int32_t nchunks, int64_t nnz, const int32_t *__restrict__ rowptr,
                                 const int32_t *__restrict__ first_owned, const T *__restrict__ part_head,
                                 const T *__restrict__ part_tail, T *__restrict__ y) {
  const int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nchunks) return;
  const int32_t fo = first_owned[b], fo_next = first_owned[b + 1];
  if (fo_next <= fo) return;
  const int32_t r = fo_next - 1;
  const int64_t re = rowptr[r + 1];
  const int64_t end = min(nnz, (int64_t)(b + 1) * BAL_CHUNK);
  if (re <= end) return;
  T total = part_tail[b];
  for (int32_t bb = b + 1; bb < nchunks; ++bb) {
    total = Num<T>::add(total, part_head[bb]);
    if (re <= min(nnz, (int64_t)(bb + 1) * BAL_CHUNK)) break;
  }
  y[r] = ACC ? Num<T>::add(total, y[r]) : total;
}

// ---- CSR, nnz-balanced, merge-style: registers and shuffles instead of a shared-memory round trip ----
// Same chunks, same part_head / part_tail / fix-up contract as k_spmv_bal.  Thread t of a block owns
// the 8 CONSECUTIVE nonzeros [8t, 8t+8) of the chunk: two 128-bit index loads and four (eight for
// Complex) 128-bit value loads, all issued together with the 8 gathers of x.  It finds the row of
// its first nonzero with one binary search over the chunk's own slice of rowptr, then walks: rows
// that begin and end inside its 8 nonzeros are written straight to y; the sum up to its first row
// end (`head`) and the sum after its last one (`tail`) are stitched to the neighbouring threads with
// a segmented scan over the block (shuffles inside a warp, 8 partials across warps).  ncu, round 1,
// on k_spmv_bal: the thread-per-row reduction read the staged products with ~16-way bank conflicts
// (row starts 16 doubles apart on average) and cost as many cycles per chunk as the chunk's whole
// HBM budget, and the gathers were issued two at a time.
// Order of additions inside a row: left to right inside a thread, scan tree across threads
// (deterministic; within the 1e-12 tolerance, like k_spmv_bal).
constexpr int MG_ITEMS = BAL_ITEMS;  // 8 nonzeros per thread, BAL_THREADS threads: the same BAL_CHUNK

template <class T>
struct Carry {
  T v;
  int f;  // 1: a row ended inside, v is what follows the last row end; 0: v is the whole range
};
template <class T>
__device__ __forceinline__ Carry<T> carry_join(const Carry<T> &a, const Carry<T> &b) {  // a, then b
  Carry<T> r;
  r.f = a.f | b.f;
  r.v = b.f ? b.v : Num<T>::add(a.v, b.v);
  return r;
}
__device__ __forceinline__ double shfl_up_t(double v, int o) { return __shfl_up_sync(0xffffffffu, v, o); }
__device__ __forceinline__ double2 shfl_up_t(double2 v, int o) {
  return double2{__shfl_up_sync(0xffffffffu, v.x, o), __shfl_up_sync(0xffffffffu, v.y, o)};
}
__device__ __forceinline__ void ld_stream8(const int32_t *p, int32_t (&c)[8], uint64_t pol) {
  const int4 a = ld_stream(reinterpret_cast<const int4 *>(p), pol);
  const int4 b = ld_stream(reinterpret_cast<const int4 *>(p) + 1, pol);
  c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w;
  c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
}
__device__ __forceinline__ void ld_stream8(const double *p, double (&a)[8], uint64_t pol) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double2 v = ld_stream(reinterpret_cast<const double2 *>(p) + i, pol);
    a[2 * i] = v.x;
    a[2 * i + 1] = v.y;
  }
}
__device__ __forceinline__ void ld_stream8(const double2 *p, double2 (&a)[8], uint64_t pol) {
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = ld_stream(p + i, pol);
}

template <class T, bool ACC, class XA>
__global__ void __launch_bounds__(BAL_THREADS)
k_spmv_merge(int32_t nrows, int64_t nnz, uint32_t zmask, const int32_t *__restrict__ rowptr,
             const int32_t *__restrict__ colidx, const T *__restrict__ vals,
             const int32_t *__restrict__ first_owned, const XA xa, T *__restrict__ y,
             T *__restrict__ part_head, T *__restrict__ part_tail) {
  __shared__ Carry<T> warp_carry[BAL_THREADS / 32];
  const int32_t b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t start = (int64_t)b * BAL_CHUNK;
  const int64_t end = min(nnz, start + BAL_CHUNK);
  const uint64_t pol_s = policy_evict_first();
  const uint64_t pol_x = policy_evict_last();
  const int32_t fo = first_owned[b], fo_next = first_owned[b + 1];

  const int64_t p0 = start + (int64_t)tid * MG_ITEMS;
  const int32_t cnt = (int32_t)max((int64_t)0, min((int64_t)MG_ITEMS, end - p0));
  int32_t c[MG_ITEMS];
  T a[MG_ITEMS], xv[MG_ITEMS];
  if (cnt == MG_ITEMS) {
    ld_stream8(colidx + p0, c, pol_s);
    ld_stream8(vals + p0, a, pol_s);
#pragma unroll
    for (int i = 0; i < MG_ITEMS; ++i) xv[i] = xa(c[i], pol_x);
    uint32_t r = 0;  // the 8 gathers are issued together (see probe/tie)
#pragma unroll
    for (int i = 0; i < MG_ITEMS; ++i) r |= probe(xv[i]);
    tie(xv[0], r & zmask);
  } else {
#pragma unroll
    for (int i = 0; i < MG_ITEMS; ++i) {
      const bool ok = i < cnt;
      c[i] = ok ? colidx[p0 + i] : 0;
      a[i] = ok ? vals[p0 + i] : Num<T>::zero();
      xv[i] = ok ? xa(c[i], pol_x) : Num<T>::zero();
    }
  }

  // rows of this chunk that hold no entry at all: the product is 0 (mulV) / leaves y alone (axpy)
  if (!ACC)
    for (int32_t r = fo + tid; r < fo_next; r += BAL_THREADS)
      if (rowptr[r + 1] == rowptr[r]) y[r] = Num<T>::zero();

  Carry<T> mine;
  mine.v = Num<T>::zero();
  mine.f = 0;
  T head = Num<T>::zero();
  int32_t head_row = -1;
  if (cnt > 0) {
    // row of p0: the last row that starts at or before p0 (it is not empty, it contains p0); it is
    // fo - 1 when no row of this chunk starts that early
    int32_t lo = fo, hi = fo_next;  // first q in [fo, fo_next] with rowptr[q] > p0
    while (lo < hi) {
      const int32_t mid = lo + ((hi - lo) >> 1);
      if ((int64_t)__ldg(rowptr + mid) <= p0) lo = mid + 1;
      else hi = mid;
    }
    int32_t r = lo - 1;
    int64_t row_end = __ldg(rowptr + r + 1);
    T sum = Num<T>::zero();
#pragma unroll
    for (int i = 0; i < MG_ITEMS; ++i) {
      if (i < cnt) {
        const int64_t p = p0 + i;
        if (p >= row_end) {  // row r ended at p: hand its sum over, move to the row that holds p
          if (!mine.f) {
            head = sum;
            head_row = r;
            mine.f = 1;
          } else {
            y[r] = ACC ? Num<T>::add(sum, y[r]) : sum;
          }
          sum = Num<T>::zero();
          do {
            ++r;
            row_end = __ldg(rowptr + r + 1);
          } while (row_end <= p);
        }
        sum = Num<T>::add(sum, Num<T>::mul(a[i], xv[i]));
      }
    }
    if (row_end == p0 + cnt) {  // the open row ends exactly where this thread's range does
      if (!mine.f) {
        head = sum;
        head_row = r;
        mine.f = 1;
      } else {
        y[r] = ACC ? Num<T>::add(sum, y[r]) : sum;
      }
      sum = Num<T>::zero();
    }
    mine.v = sum;
  }

  // inclusive segmented scan of the carries over the block
  Carry<T> inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Carry<T> up;
    up.v = shfl_up_t(inc.v, o);
    up.f = __shfl_up_sync(0xffffffffu, inc.f, o);
    if (lane >= o) inc = carry_join(up, inc);
  }
  if (lane == 31) warp_carry[warp] = inc;
  Carry<T> prev;  // exclusive inside the warp
  prev.v = shfl_up_t(inc.v, 1);
  prev.f = __shfl_up_sync(0xffffffffu, inc.f, 1);
  if (lane == 0) {
    prev.v = Num<T>::zero();
    prev.f = 0;
  }
  __syncthreads();
  Carry<T> before;  // everything in the warps below this one
  before.v = Num<T>::zero();
  before.f = 0;
#pragma unroll
  for (int w = 0; w < BAL_THREADS / 32; ++w)
    if (w < warp) before = carry_join(before, warp_carry[w]);
  const Carry<T> in = carry_join(before, prev);  // what flows into this thread

  if (mine.f) {  // the row that ends first in this thread: earlier threads' share + head
    const T total = Num<T>::add(in.v, head);
    if (head_row < fo) part_head[b] = total;  // it began in an earlier chunk: the fix-up kernel adds it
    else y[head_row] = ACC ? Num<T>::add(total, y[head_row]) : total;
  }
  if (tid == BAL_THREADS - 1) {
    const Carry<T> all = carry_join(in, mine);
    // all.f: the row still open at the end of the chunk began in it.  Otherwise no row ends in this
    // chunk and all of it is one piece of a long row -- its first piece if that row starts exactly at
    // the chunk start (then the chunk owns it and the fix-up kernel looks for it in part_tail).
    const bool owns_open_row = all.f || (fo < fo_next && (int64_t)rowptr[fo] == start);
    if (owns_open_row) part_tail[b] = all.v;
    else part_head[b] = all.v;
  }
}

template <class T, bool ACC>
__global__ void k_fill_zero_or_keep(int64_t n, T *__restrict__ y) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && !ACC) y[i] = Num<T>::zero();
}

// =================================================================================================
// SpMM kernels:  Y[:, j] <- A X[:, j], column-major X and Y
// =================================================================================================
// One thread per row and JT columns per thread; the row's (index, value) pairs are read once per
// column tile and reused for JT gathers.  Blocks are numbered tile-fastest, so the blocks that
// share a row block (and its slice of A and window of X) run at the same time and meet in L2.
template <class T, int JT, bool SELL>
__global__ void __launch_bounds__(128)
k_spmm_rows(int32_t nrows, int32_t ntiles, const int64_t *__restrict__ slice_ptr,
            const int32_t *__restrict__ rowlen, const int32_t *__restrict__ rowptr,
            const int32_t *__restrict__ colidx, const T *__restrict__ vals, const T *__restrict__ X,
            int64_t ldx, T *__restrict__ Y, int64_t ldy, int32_t k, const YInit<T> yi) {
  const int32_t rb = blockIdx.x / ntiles;
  const int32_t tile = blockIdx.x - rb * ntiles;
  const int32_t row = rb * blockDim.x + threadIdx.x;
  if (row >= nrows) return;
  const int32_t j0 = tile * JT;
  const uint64_t pol_x = policy_evict_last();
  int64_t base;
  int32_t len;
  int32_t stride;
  if (SELL) {
    base = __ldg(slice_ptr + (row >> 5)) + (threadIdx.x & 31);
    len = __ldg(rowlen + row);
    stride = SELL_C;
  } else {
    base = rowptr[row];
    len = rowptr[row + 1] - (int32_t)base;
    stride = 1;
  }
  T acc[JT];
#pragma unroll
  for (int j = 0; j < JT; ++j) acc[j] = (j0 + j < k) ? yi(Y + (int64_t)(j0 + j) * ldy + row) : Num<T>::zero();
  const T *xb = X + (int64_t)j0 * ldx;
  for (int32_t kk = 0; kk < len; ++kk) {
    const int32_t c = __ldg(colidx + base + (int64_t)kk * stride);
    const T a = __ldg(vals + base + (int64_t)kk * stride);
    T xv[JT];
#pragma unroll
    for (int j = 0; j < JT; ++j)
      xv[j] = (j0 + j < k) ? ld_gather(xb + (int64_t)j * ldx + c, pol_x) : Num<T>::zero();
#pragma unroll
    for (int j = 0; j < JT; ++j) acc[j] = Num<T>::add(Num<T>::mul(a, xv[j]), acc[j]);
  }
#pragma unroll
  for (int j = 0; j < JT; ++j)
    if (j0 + j < k) st_stream(Y + (int64_t)(j0 + j) * ldy + row, acc[j]);
}


// Short rows (<= LMAX entries, e.g. 5- and 7-point stencils): the row's (index, value) pairs are
// loaded once into registers and reused for a whole group of columns, so A is streamed once per
// column group instead of once per column tile, and the LMAX*JT gathers of a tile are all in flight
// before the first is used.  The kernel is issue-bound, not bandwidth-bound (ncu, round 1: 235
// instructions per (row, column) with LMAX = 8 on 5-entry rows), so LMAX is instantiated to the
// matrix's exact longest row and a warp whose 32 rows are all full-length runs a branch-free,
// predicate-free column loop; only boundary warps and the last partial column tile pay for predicates.
template <class T, int JT, int LMAX, bool FULL, bool PIPE>
__device__ __forceinline__ void spmm_short_columns(const int32_t (&c)[LMAX], const T (&a)[LMAX], int32_t len,
                                                   const T *__restrict__ X, int64_t ldx, T *__restrict__ Yrow,
                                                   int64_t ldy, int32_t jbeg, int32_t jend, uint64_t pol_x,
                                                   uint32_t zmask, const YInit<T> &yi) {
  // the LMAX*JT gathers of the column tile that starts at j0
  auto load = [&](T (&xv)[LMAX][JT], int32_t j0) {
    const T *xb = X + (int64_t)j0 * ldx;
#pragma unroll
    for (int u = 0; u < LMAX; ++u)
#pragma unroll
      for (int j = 0; j < JT; ++j)
        xv[u][j] = (FULL || u < len) ? ld_gather(xb + (int64_t)j * ldx + c[u], pol_x) : Num<T>::zero();
  };
  // products of the tile, in the reference's order per column; no product issues before every gather
  // of the tile has been issued (see probe/tie)
  auto finish = [&](T (&xv)[LMAX][JT], int32_t j0) {
    uint32_t r = 0;
#pragma unroll
    for (int u = 0; u < LMAX; ++u)
#pragma unroll
      for (int j = 0; j < JT; ++j) r |= probe(xv[u][j]);
    r &= zmask;
#pragma unroll
    for (int j = 0; j < JT; ++j) tie(xv[0][j], r);
#pragma unroll
    for (int j = 0; j < JT; ++j) {
      T acc = yi(Yrow + (int64_t)(j0 + j) * ldy);
#pragma unroll
      for (int u = 0; u < LMAX; ++u)
        if (FULL || u < len) acc = Num<T>::add(Num<T>::mul(a[u], xv[u][j]), acc);
      st_stream(Yrow + (int64_t)(j0 + j) * ldy, acc);
    }
  };
  int32_t j0 = jbeg;
  if (PIPE) {
    // two register buffers: the gathers of tile t+1 are issued before the products of tile t, so a
    // warp always has a tile in flight while it computes (ptxas keeps that order: checked in SASS)
    if (j0 + JT <= jend) {
      T x0[LMAX][JT], x1[LMAX][JT];
      load(x0, j0);
      for (;;) {
        const bool n1 = j0 + 2 * JT <= jend;
        if (n1) load(x1, j0 + JT);
        finish(x0, j0);
        j0 += JT;
        if (!n1) break;
        const bool n2 = j0 + 2 * JT <= jend;
        if (n2) load(x0, j0 + JT);
        finish(x1, j0);
        j0 += JT;
        if (!n2) break;
      }
    }
  } else {
    for (; j0 + JT <= jend; j0 += JT) {
      T xv[LMAX][JT];
      load(xv, j0);
      finish(xv, j0);
    }
  }
  for (; j0 < jend; ++j0) {  // last partial tile, one column at a time
    const T *xb = X + (int64_t)j0 * ldx;
    T acc = yi(Yrow + (int64_t)j0 * ldy);
#pragma unroll
    for (int u = 0; u < LMAX; ++u)
      if (FULL || u < len) acc = Num<T>::add(Num<T>::mul(a[u], ld_gather(xb + c[u], pol_x)), acc);
    st_stream(Yrow + (int64_t)j0 * ldy, acc);
  }
}

//
// Row mapping.  stride32 == 0: a block owns blockDim.x consecutive rows.  stride32 == S/32 > 0 (banded
// matrices whose rows reach S columns away, e.g. a 2-D grid of width S): the W warps of a block own
// 32-row slices S rows apart, block b of a chunk of W*S rows taking slices {b + w*S/32}.  The window
// of X that warp w gathers around column r+S is then the window warp w+1 gathers around its own
// diagonal, so it is fetched from L2 once per block (L1 hit for the others) instead of once per warp.
template <class T, int JT, int LMAX, bool SELL, bool PIPE>
__global__ void __launch_bounds__(256)
k_spmm_short(int32_t nrows, int32_t ngroups, int32_t jgroup, int32_t stride32, uint32_t zmask,
             const int64_t *__restrict__ slice_ptr,
             const int32_t *__restrict__ rowlen, const int32_t *__restrict__ rowptr,
             const int32_t *__restrict__ colidx, const T *__restrict__ vals, const T *__restrict__ X,
             int64_t ldx, T *__restrict__ Y, int64_t ldy, int32_t k, const YInit<T> yi) {
  const int32_t rb = blockIdx.x / ngroups;
  const int32_t grp = blockIdx.x - rb * ngroups;
  int32_t row;
  if (stride32 == 0) {
    row = rb * blockDim.x + threadIdx.x;
  } else {
    const int32_t chunk = rb / stride32, b = rb - chunk * stride32;
    const int64_t r64 = ((int64_t)chunk * (blockDim.x >> 5) * stride32 + (int64_t)(threadIdx.x >> 5) * stride32 + b) * 32 +
                        (threadIdx.x & 31);
    row = r64 < nrows ? (int32_t)r64 : nrows;
  }
  const bool live = row < nrows;
  const uint64_t pol_s = policy_evict_first();
  const uint64_t pol_x = policy_evict_last();
  int64_t base = 0;
  int32_t len = 0, stride = 1;
  if (live) {
    if (SELL) {
      base = __ldg(slice_ptr + (row >> 5)) + (threadIdx.x & 31);
      len = __ldg(rowlen + row);
      stride = SELL_C;
    } else {
      base = rowptr[row];
      len = rowptr[row + 1] - (int32_t)base;
    }
  }
  const bool full = __all_sync(0xffffffffu, live && len == LMAX);
  int32_t c[LMAX];
  T a[LMAX];
#pragma unroll
  for (int u = 0; u < LMAX; ++u) {
    const bool ok = u < len;
    c[u] = ok ? ld_stream(colidx + base + (int64_t)u * stride, pol_s) : 0;
    a[u] = ok ? ld_stream(vals + base + (int64_t)u * stride, pol_s) : Num<T>::zero();
  }
  if (!live) return;
  const int32_t jbeg = grp * jgroup;
  const int32_t jend = min(k, jbeg + jgroup);
  if (full)
    spmm_short_columns<T, JT, LMAX, true, PIPE>(c, a, len, X, ldx, Y + row, ldy, jbeg, jend, pol_x, zmask, yi);
  else
    spmm_short_columns<T, JT, LMAX, false, false>(c, a, len, X, ldx, Y + row, ldy, jbeg, jend, pol_x, zmask, yi);
}

// Longer rows (SELL plan, e.g. the 27-point stencil): a block owns blockDim.x consecutive rows and ALL
// k columns.  Each thread first copies its own row's (index, value) pairs from the SELL stream into
// shared memory (laid out [entry][thread]: conflict-free, thread-private, so no barrier), which makes
// this the only pass over A in HBM; entries past `kcap` (what fits in the block's shared memory) are
// re-read through L1 for every column tile.  The block then walks the columns JT at a time; inside
// a tile KU entries x JT columns = KU*JT gathers are issued together (probe/tie) and added in the
// reference's order, entry after entry, for every column.  ncu, round 1, on the per-tile-grid kernel
// this replaces (k_spmm_rows): A was streamed 16 times (11.4 GB DRAM reads for 6.7 GB compulsory)
// and half of the stall samples sat between an index load and the gather that needs it.
template <class T, int JT, int KU>
__global__ void __launch_bounds__(512)
k_spmm_panel(int32_t row_begin, int32_t row_end, int32_t kcap, uint32_t zmask, const int64_t *__restrict__ slice_ptr,
             const int32_t *__restrict__ rowlen, const int32_t *__restrict__ colidx,
             const T *__restrict__ vals, const T *__restrict__ X, int64_t ldx, T *__restrict__ Y,
             int64_t ldy, int32_t k, const int32_t *__restrict__ order, const YInit<T> yi) {
  extern __shared__ __align__(16) unsigned char panel_smem[];
  const int32_t nt = blockDim.x, tid = threadIdx.x;
  T *s_a = reinterpret_cast<T *>(panel_smem) + tid;                                       // [kcap][nt]
  int32_t *s_c = reinterpret_cast<int32_t *>(panel_smem + (size_t)kcap * nt * sizeof(T)) + tid;  // [kcap][nt]
  int32_t row = row_begin + blockIdx.x * nt + tid;  // row_begin % 32 == 0: lanes are slice rows
  if (order) {  // line-tiled blocks: every warp is told its slice (see panel_tiling)
    const int32_t sl = __ldg(order + (int64_t)blockIdx.x * (nt >> 5) + (tid >> 5));
    if (sl < 0) return;
    row = sl * SELL_C + (tid & 31);
  }
  if (row >= row_end) return;
  const uint64_t pol_s = policy_evict_first();
  const uint64_t pol_x = policy_evict_last();
  const int64_t base = __ldg(slice_ptr + (row >> 5)) + (tid & 31);
  const int32_t len = __ldg(rowlen + row);
  const int32_t nst = min(len, kcap);
  const int32_t *ci = colidx + base;
  const T *va = vals + base;
  for (int32_t e = 0; e < nst; ++e) {
    s_c[e * nt] = ld_stream(ci + (int64_t)e * SELL_C, pol_s);
    s_a[e * nt] = ld_stream(va + (int64_t)e * SELL_C, pol_s);
  }
  T *Yrow = Y + row;
  for (int32_t j0 = 0; j0 < k; j0 += JT) {
    const T *xb = X + (int64_t)j0 * ldx;
    T acc[JT];
#pragma unroll
    for (int j = 0; j < JT; ++j) acc[j] = (j0 + j < k) ? yi(Yrow + (int64_t)(j0 + j) * ldy) : Num<T>::zero();
    int32_t e = 0;
    if (j0 + JT <= k) {  // full column tile: batches of KU staged entries, no predicates
      for (; e + KU <= nst; e += KU) {
        int32_t c[KU];
        T a[KU], xv[KU][JT];
#pragma unroll
        for (int u = 0; u < KU; ++u) {
          c[u] = s_c[(e + u) * nt];
          a[u] = s_a[(e + u) * nt];
        }
#pragma unroll
        for (int u = 0; u < KU; ++u)
#pragma unroll
          for (int j = 0; j < JT; ++j) xv[u][j] = ld_gather(xb + (int64_t)j * ldx + c[u], pol_x);
        uint32_t r = 0;
#pragma unroll
        for (int u = 0; u < KU; ++u)
#pragma unroll
          for (int j = 0; j < JT; ++j) r |= probe(xv[u][j]);
        r &= zmask;
#pragma unroll
        for (int j = 0; j < JT; ++j) tie(xv[0][j], r);
#pragma unroll
        for (int u = 0; u < KU; ++u)
#pragma unroll
          for (int j = 0; j < JT; ++j) acc[j] = Num<T>::add(Num<T>::mul(a[u], xv[u][j]), acc[j]);
      }
    }
    for (; e < len; ++e) {  // leftovers: one entry at a time, any tile width
      int32_t c;
      T a;
      if (e < nst) {
        c = s_c[e * nt];
        a = s_a[e * nt];
      } else {
        c = __ldg(ci + (int64_t)e * SELL_C);
        a = __ldg(va + (int64_t)e * SELL_C);
      }
      T xv[JT];
#pragma unroll
      for (int j = 0; j < JT; ++j)
        xv[j] = (j0 + j < k) ? ld_gather(xb + (int64_t)j * ldx + c, pol_x) : Num<T>::zero();
#pragma unroll
      for (int j = 0; j < JT; ++j) acc[j] = Num<T>::add(Num<T>::mul(a, xv[j]), acc[j]);
    }
#pragma unroll
    for (int j = 0; j < JT; ++j)
      if (j0 + j < k) st_stream(Yrow + (int64_t)(j0 + j) * ldy, acc[j]);
  }
}

// Boundary rows of a row-partitioned SpMM: rows [0, r0) and [r1, nrows) of the local block, whose
// columns reach into a peer's rows of X.  One thread per row, JT columns per pass; the row's entries
// are re-read through L1 for every pass (these rows are a few per cent of the block -- two planes
// of a z-slab -- so the interior kernel's staging is not repeated here).  Remote entries are plain
// loads from the owner's HBM over NVLink, each part with its own leading dimension (DistXM).
template <class T, int JT>
__global__ void __launch_bounds__(128)
k_spmm_boundary(int32_t nrows, int32_t r0, int32_t r1, uint32_t zmask, const int64_t *__restrict__ slice_ptr,
                const int32_t *__restrict__ rowlen, const int32_t *__restrict__ colidx,
                const T *__restrict__ vals, const DistXM<T> xa, T *__restrict__ Y, int64_t ldy, int32_t k,
                const YInit<T> yi) {
  const int32_t nbl = (r0 + 127) >> 7;  // blocks below the interior
  int32_t row, limit;
  if ((int32_t)blockIdx.x < nbl) {
    row = blockIdx.x * 128 + threadIdx.x;
    limit = r0;
  } else {
    row = r1 + (blockIdx.x - nbl) * 128 + threadIdx.x;  // r1 % 32 == 0 unless r1 == nrows
    limit = nrows;
  }
  if (row >= limit) return;
  const uint64_t pol_x = policy_evict_last();
  const int64_t base = __ldg(slice_ptr + (row >> 5)) + (row & 31);
  const int32_t len = __ldg(rowlen + row);
  const int32_t *ci = colidx + base;
  const T *va = vals + base;
  // blockIdx.y = column tile: the two boundary planes are a few per cent of the rows, and a thread that
  // walked all k columns left 17 warps per SM waiting out one NVLink round trip after the other (a third
  // of the 8-GPU step); with a thread per (row, tile) k / JT times as many loads are in flight
  {
    const int32_t j0 = blockIdx.y * JT;
    T acc[JT];
#pragma unroll
    for (int j = 0; j < JT; ++j) acc[j] = (j0 + j < k) ? yi(Y + (int64_t)(j0 + j) * ldy + row) : Num<T>::zero();
    for (int32_t e = 0; e < len; ++e) {
      const int32_t c = __ldg(ci + (int64_t)e * SELL_C);
      const T a = __ldg(va + (int64_t)e * SELL_C);
      int64_t ld;
      bool remote;
      const T *xp = xa.col0(c, ld, remote);
      xp += (int64_t)j0 * ld;
      T xv[JT];
      if (remote) {
#pragma unroll
        for (int j = 0; j < JT; ++j) xv[j] = (j0 + j < k) ? ld_peer(xp + (int64_t)j * ld) : Num<T>::zero();
      } else {
#pragma unroll
        for (int j = 0; j < JT; ++j) xv[j] = (j0 + j < k) ? ld_gather(xp + (int64_t)j * ld, pol_x) : Num<T>::zero();
      }
      uint32_t r = 0;
#pragma unroll
      for (int j = 0; j < JT; ++j) r |= probe(xv[j]);
      r &= zmask;
#pragma unroll
      for (int j = 0; j < JT; ++j) tie(xv[j], r);
#pragma unroll
      for (int j = 0; j < JT; ++j) acc[j] = Num<T>::add(Num<T>::mul(a, xv[j]), acc[j]);
    }
#pragma unroll
    for (int j = 0; j < JT; ++j)
      if (j0 + j < k) st_stream(Y + (int64_t)(j0 + j) * ldy + row, acc[j]);
  }
}

// =================================================================================================
// dispatch
// =================================================================================================

template <class T, bool ACC, class XA>
static int launch_sell(slc_matrix *m, const XA &xa, T *y, int32_t row_begin = 0, int32_t row_end = -1) {
  const SellPlan &sp = m->sell;
  cudaStream_t s = cur_stream();
  if (row_end < 0) row_end = m->nrows;
  if (row_end <= row_begin) return SLC_OK;
  const unsigned nb = (unsigned)ceil_div(row_end - row_begin, 256);
  static const int u_env = env_int("SLC_SELL_U", 0);
  static const int pf_env = env_int("SLC_SELL_PREFETCH", 1);  // L2 prefetch distance in batches (compressed kernel)
  static const int cap_env = env_int("SLC_SELL_CAP32", 1);  // compressed U = 4 kernel held to 32 registers (8 blocks per SM)
  int u = u_env;
  // measured on B200, 27-point stencil: U=8 6690 GB/s, U=4 6464, U=16 5425 (register pressure)
  // with compressed indices: U=4 0.379 ms (40 registers, 6 blocks per SM), held to 32 registers (8 blocks)
  // 0.338 ms, plus an L2 prefetch one batch ahead 0.326 ms (two: 0.330, three: 0.334, five: 0.340); U=8 0.471 ms (64 registers); off 0.430 ms
  if (u == 0) u = (m->max_rowlen <= 4 || sp.kbase) ? 4 : 8;
  // SLC_PDL (0): 1 = programmatic dependent launch -- back-to-back products overlap the next kernel's first
  // wave (launch, values towards L2) with the previous kernel's tail.  Measured (round 2): no difference on
  // one GPU (0.3286 against 0.3280 ms on cfg 5) and 4 % slower at 8 GPUs (0.0555 against 0.0529 ms: the early
  // blocks of the next product sit on the slots the spinning boundary blocks of this one would take), so
  // plain stream order is the default.
  static const int pdl_env = env_int("SLC_PDL", 0);
  const bool pdl = pdl_env != 0;
  const int first_wave = pdl ? sm_count() * 8 : 0;
  const int32_t *no_kbase = nullptr;
  cudaError_t le = cudaSuccess;
#define SLC_SELL_CASE(UU)                                                                          \
  do {                                                                                             \
    if (sp.kbase && cap_env && UU == 4 && sizeof(T) == 8)                                          \
      le = launch_pdl(k_spmv_sell<T, 4, ACC, XA, true, 8>, nb, 256, 0, s, pdl, row_begin, row_end, sp.slice_ptr, \
                      sp.rowlen, sp.colidx, (const T *)sp.vals, xa, y, sp.kbase, pf_env, first_wave); \
    else if (sp.kbase)                                                                             \
      le = launch_pdl(k_spmv_sell<T, UU, ACC, XA, true>, nb, 256, 0, s, pdl, row_begin, row_end, sp.slice_ptr, \
                      sp.rowlen, sp.colidx, (const T *)sp.vals, xa, y, sp.kbase, pf_env, first_wave); \
    else                                                                                           \
      le = launch_pdl(k_spmv_sell<T, UU, ACC, XA>, nb, 256, 0, s, pdl, row_begin, row_end, sp.slice_ptr, \
                      sp.rowlen, sp.colidx, (const T *)sp.vals, xa, y, no_kbase, 0, first_wave);   \
  } while (0)
  if (u <= 4) SLC_SELL_CASE(4);
  else if (u <= 8) SLC_SELL_CASE(8);
  else SLC_SELL_CASE(16);
#undef SLC_SELL_CASE
  SLC_CUDA(le);
  SLC_LAUNCHED();
  return SLC_OK;
}

template <class T, bool ACC, class XA>
static int spmv_run(slc_matrix *m, const XA &xa, T *y) {
  if (m->nrows == 0) return SLC_OK;
  cudaStream_t s = cur_stream();
  if (m->nnz == 0) {
    if (!ACC) SLC_CUDA(cudaMemsetAsync(y, 0, (size_t)m->nrows * sizeof(T), s));
    return SLC_OK;
  }
  SLC_TRY(ensure_plan<T>(m));
  switch (m->spmv_active) {
    case SPMV_SELL:
      return launch_sell<T, ACC, XA>(m, xa, y);
    case SPMV_CSR_WARP: {
      k_spmv_csr_warp<T, ACC, XA><<<(unsigned)ceil_div((int64_t)m->nrows * 32, 256), 256, 0, s>>>(
          m->nrows, m->rowptr, m->colidx, (const T *)m->rvals, xa, y);
      SLC_LAUNCHED();
      return SLC_OK;
    }
    default: {
      const BalancedPlan &bp = m->bal;
      DevBuf<T> ph, pt;
      SLC_TRY(ph.alloc((size_t)bp.nchunks));
      SLC_TRY(pt.alloc((size_t)bp.nchunks));
      // measured on B200, R-MAT scale 24: shared-memory kernel 1.51 ms, merge-style kernel 1.95 ms
      // (both bound by the number of random 32-byte gathers an SM keeps in flight)
      static const int batch_env = env_int("SLC_BAL_BATCH", 0);
      if (m->spmv_active == SPMV_CSR_BALANCED && std::is_same<XA, LocalX<T>>::value) {
        // hot columns of x in shared memory: 128 KB of slots beside the four product buffers
        const int32_t cap = (int32_t)(((size_t)128 << 10) / sizeof(T));
        SLC_TRY(build_hot(m, cap));
        static const int mode_env = env_int("SLC_BAL_HOT_MODE", 2);  // 2 = packed table in global memory (default, -10 % on R-MAT scale 24), 1 = shared-memory table in resident blocks (lost its A/B)
        if (bp.hot_k > 0 && mode_env == 2) {
          DevBuf<T> xhot;
          SLC_TRY(xhot.alloc((size_t)bp.hot_k));
          k_hot_pack<T><<<(unsigned)ceil_div(bp.hot_k, 256), 256, 0, s>>>(bp.hot_k, bp.hot_cols, local_ptr(xa), xhot.p);
          SLC_LAUNCHED();
          static const int minb_env = env_int("SLC_BAL_MINB8", 0);  // hold the kernel to 32 registers (8 blocks per SM)
          if (minb_env && sizeof(T) == 8)
            k_spmv_bal<T, ACC, HotPackedX<T>, 8><<<(unsigned)bp.nchunks, BAL_THREADS, 0, s>>>(
                m->nrows, m->nnz, 0u, batch_env, m->rowptr, bp.colidx_hot, (const T *)m->rvals, bp.chunk_row,
                HotPackedX<T>{local_ptr(xa), xhot.p}, y, ph.p, pt.p);
          else
            k_spmv_bal<T, ACC, HotPackedX<T>><<<(unsigned)bp.nchunks, BAL_THREADS, 0, s>>>(
                m->nrows, m->nnz, 0u, batch_env, m->rowptr, bp.colidx_hot, (const T *)m->rvals, bp.chunk_row,
                HotPackedX<T>{local_ptr(xa), xhot.p}, y, ph.p, pt.p);
          SLC_LAUNCHED();
          k_spmv_bal_fixup<T, ACC><<<(unsigned)ceil_div(bp.nchunks, 256), 256, 0, s>>>(
              bp.nchunks, m->nnz, m->rowptr, bp.chunk_row, ph.p, pt.p, y);
          SLC_LAUNCHED();
          return SLC_OK;
        }
        if (bp.hot_k > 0) {
          static const int groups_env = env_int("SLC_BAL_HOT_GROUPS", 0);
          const int HOT_GROUPS = groups_env > 0 ? std::min(groups_env, HotGroups<T>::value) : HotGroups<T>::value;
          const size_t smem = ((size_t)HOT_GROUPS * BAL_CHUNK + (size_t)bp.hot_k) * sizeof(T);
          static PerDevice<size_t> smem_set;
          SLC_CUDA(ensure_dyn_smem(k_spmv_bal_hot<T, ACC>, smem_set, smem));
          const int per_sm = std::max<int>(1, (int)(((size_t)227 << 10) / (smem + 1024)));
          const int64_t grid = std::min<int64_t>(ceil_div(bp.nchunks, HOT_GROUPS), (int64_t)sm_count() * per_sm);
          k_spmv_bal_hot<T, ACC><<<(unsigned)grid, HOT_GROUPS * BAL_THREADS, smem, s>>>(
              m->nrows, m->nnz, bp.nchunks, 0u, batch_env, m->rowptr, bp.colidx_hot, (const T *)m->rvals, bp.chunk_row,
              local_ptr(xa), bp.hot_cols, bp.hot_k, y, ph.p, pt.p);
          SLC_LAUNCHED();
          k_spmv_bal_fixup<T, ACC><<<(unsigned)ceil_div(bp.nchunks, 256), 256, 0, s>>>(
              bp.nchunks, m->nnz, m->rowptr, bp.chunk_row, ph.p, pt.p, y);
          SLC_LAUNCHED();
          return SLC_OK;
        }
      }
      if (m->spmv_active == SPMV_CSR_BALANCED)
        k_spmv_bal<T, ACC, XA><<<(unsigned)bp.nchunks, BAL_THREADS, 0, s>>>(
            m->nrows, m->nnz, 0u, batch_env, m->rowptr, m->colidx, (const T *)m->rvals, bp.chunk_row, xa, y, ph.p,
            pt.p);
      else
        k_spmv_merge<T, ACC, XA><<<(unsigned)bp.nchunks, BAL_THREADS, 0, s>>>(
            m->nrows, m->nnz, 0u, m->rowptr, m->colidx, (const T *)m->rvals, bp.chunk_row, xa, y, ph.p, pt.p);
      SLC_LAUNCHED();
      k_spmv_bal_fixup<T, ACC><<<(unsigned)ceil_div(bp.nchunks, 256), 256, 0, s>>>(
          bp.nchunks, m->nnz, m->rowptr, bp.chunk_row, ph.p, pt.p, y);
      SLC_LAUNCHED();
      return SLC_OK;
    }
  }
}

// host copy of the per-slice column ranges (used to split distributed and pipelined products)
static int fetch_slice_ranges(slc_matrix *m) {
  SellPlan &sp = m->sell;
  if (!sp.h_cmin.empty() || sp.nslices == 0) return SLC_OK;
  sp.h_cmin.resize(sp.nslices);
  sp.h_cmax.resize(sp.nslices);
  cudaStream_t s = cur_stream();
  SLC_CUDA(cudaMemcpyAsync(sp.h_cmin.data(), sp.slice_cmin, (size_t)sp.nslices * 4, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaMemcpyAsync(sp.h_cmax.data(), sp.slice_cmax, (size_t)sp.nslices * 4, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  return SLC_OK;
}

template <class T, bool ACC>
static int spmv_dev(slc_matrix *m, const T *x, T *y) {
  return spmv_run<T, ACC, LocalX<T>>(m, LocalX<T>{x}, y);
}

template <class T>
static int check_handle(const slc_matrix *m, const char *fn) {
  if (!m) return fail(SLC_ERR_INVALID, "%s: null handle", fn);
  if (m->elem != Num<T>::elem) return fail(SLC_ERR_INVALID, "%s: element type mismatch", fn);
  return SLC_OK;
}

template <class T, bool ACC>
static int spmv_dev_entry(slc_matrix *m, const double *x, double *y, const char *fn) {
  SLC_TRY(check_handle<T>(m, fn));
  HandleLock lock(m);
  if ((m->ncols && !x) || (m->nrows && !y)) return fail(SLC_ERR_INVALID, "%s: null vector", fn);
  return spmv_dev<T, ACC>(m, (const T *)x, (T *)y);
}

// Host operands, large banded matrix: the product is cut into row chunks.  Chunk i needs x only up
// to the largest column its slices touch, so x is uploaded in pieces on one stream, chunk kernels
// run on the caller's stream as their piece lands, and finished y chunks go back on a third
// stream: H2D, compute and D2H overlap (PCIe is full duplex) instead of running back to back.
constexpr int PIPE_MAX_CHUNKS = 32;

// chunk i of `n` of the pipelined product covers rows [cut(i), cut(i+1)): equal pieces, except that with
// `taper` the last quarter is cut finer and finer, so that what is left after the last upload piece
// lands (one small kernel, one small download) is a few per cent of the call instead of 1/n of it
static void pipeline_cuts(int32_t nrows, int want, int taper, std::vector<int32_t> &cuts) {
  cuts.clear();
  cuts.push_back(0);
  auto round256 = [&](double r) { return (int32_t)std::min<int64_t>(nrows, (int64_t)(r / 256.0 + 0.5) * 256); };
  if (!taper || want < 4) {
    const int32_t per = (int32_t)(ceil_div(ceil_div(nrows, want), 256) * 256);
    for (int64_t r = per; r < nrows; r += per) cuts.push_back((int32_t)r);
  } else {
    // taper 2: the first pieces are small as well (1/64, 1/32, 1/16 of the rows) -- the downloads cannot
    // start before the first piece has been uploaded and multiplied, and with both directions of the
    // link busy for the rest of the call that lag is pure latency (trace, round 2: 0.31 ms with 8 equal
    // pieces of the 27-point stencil's 64 MB vectors)
    const double head = taper >= 2 ? 7.0 / 64.0 : 0.0;
    if (taper >= 2) {
      cuts.push_back(round256(nrows / 64.0));
      cuts.push_back(round256(3.0 * nrows / 64.0));
      cuts.push_back(round256(head * nrows));
    }
    const int big = want - 3;                       // equal pieces up to 7/8 of the rows ...
    const double body = 0.875 * nrows - head * nrows;
    for (int i = 1; i <= big; ++i) cuts.push_back(round256(head * nrows + body * i / big));
    cuts.push_back(round256(0.9375 * nrows));       // ... then 1/16, 1/32, 1/32
    cuts.push_back(round256(0.96875 * nrows));
  }
  std::vector<int32_t> uniq;
  for (int32_t c : cuts)
    if (uniq.empty() || c > uniq.back()) uniq.push_back(c);
  if (uniq.back() < nrows) uniq.push_back(nrows);
  cuts.swap(uniq);
}

template <class T, bool ACC>
static int spmv_host_pipelined(slc_matrix *m, const T *x, T *y, bool *done) {
  *done = false;
  const int chunks_env = env_int("SLC_E2E_CHUNKS", 8);  // read per call: scripts/e2e_probe.py sweeps it in one process
  const int diag = env_int("SLC_E2E_DIAG", 0);          // timing diagnosis: 1 = no download, 2 = no upload, 4 = no kernels
  const int align_env = env_int("SLC_E2E_ALIGN", 512);
  const int mode = env_int("SLC_E2E_MODE", 0);          // 0 = three streams chained by events, 1 = downloads follow their kernels
  const int taper = env_int("SLC_E2E_TAPER", 1);        // 1: finer chunks at the end, 2: at both ends (measured no better: pipeline_cuts)
  const int trace = env_int("SLC_E2E_TRACE", 0);        // print a timeline of the call (timed events) to stderr
  const int want = chunks_env < 1 ? 1 : (chunks_env > PIPE_MAX_CHUNKS ? PIPE_MAX_CHUNKS : chunks_env);
  if (want == 1 || m->nnz == 0 || m->nrows < (1 << 18)) return SLC_OK;
  SLC_TRY(ensure_plan<T>(m));
  if (m->spmv_active != SPMV_SELL) return SLC_OK;
  SLC_TRY(fetch_slice_ranges(m));
  SellPlan &sp = m->sell;
  std::vector<int32_t> cuts;
  pipeline_cuts(m->nrows, want, taper, cuts);
  const int nch = (int)cuts.size() - 1;
  if (nch > PIPE_MAX_CHUNKS) return SLC_OK;
  if (!m->pipe_in) {
    SLC_CUDA(cudaStreamCreateWithFlags(&m->pipe_in, cudaStreamNonBlocking));
    SLC_CUDA(cudaStreamCreateWithFlags(&m->pipe_out, cudaStreamNonBlocking));
    m->pipe_events.resize(2 * PIPE_MAX_CHUNKS + 1);
    for (auto &e : m->pipe_events) SLC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  cudaStream_t main_s = cur_stream();
  // the device copies of x and y live with the handle: no allocation per call
  const size_t need_x = (size_t)m->ncols * sizeof(T), need_y = (size_t)m->nrows * sizeof(T);
  if (m->pipe_x_bytes < need_x) {
    if (m->pipe_x) SLC_CUDA(cudaFree(m->pipe_x));
    m->pipe_x = nullptr;
    SLC_CUDA(cudaMalloc(&m->pipe_x, need_x));
    m->pipe_x_bytes = need_x;
  }
  if (m->pipe_y_bytes < need_y) {
    if (m->pipe_y) SLC_CUDA(cudaFree(m->pipe_y));
    m->pipe_y = nullptr;
    SLC_CUDA(cudaMalloc(&m->pipe_y, need_y));
    m->pipe_y_bytes = need_y;
  }
  T *dx = (T *)m->pipe_x, *dy = (T *)m->pipe_y;
  DevBuf<T> dcold;
  if (diag & 8) SLC_TRY(dcold.alloc((size_t)m->nrows));  // diagnosis: downloads read memory no kernel has written
  const T *dsrc = (diag & 8) ? dcold.p : dy;
  std::vector<cudaEvent_t> tev;  // trace: start, then (upload, kernel, download) per chunk
  auto mark = [&](cudaStream_t st) {
    if (!trace) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    tev.push_back(e);
  };
  cudaEvent_t ev_start = m->pipe_events[2 * PIPE_MAX_CHUNKS];
  mark(main_s);
  SLC_CUDA(cudaEventRecord(ev_start, main_s));
  SLC_CUDA(cudaStreamWaitEvent(m->pipe_in, ev_start, 0));
  SLC_CUDA(cudaStreamWaitEvent(m->pipe_out, ev_start, 0));
  int32_t sent = 0;  // x[0, sent) is on its way
  for (int i = 0; i < nch; ++i) {
    const int32_t r0 = cuts[i], r1 = cuts[i + 1];
    int32_t need = sent;
    for (int32_t sl = r0 / SELL_C; sl < (int32_t)ceil_div(r1, SELL_C); ++sl) need = std::max(need, sp.h_cmax[sl] + 1);
    // piece boundaries on 512-element (4 KB) multiples: both ends of every copy stay page-aligned
    if (align_env > 0) need = (int32_t)std::min<int64_t>(m->ncols, ceil_div(need, align_env) * (int64_t)align_env);
    if (i == nch - 1) need = m->ncols;
    if (need > sent) {
      if (!(diag & 2))
        SLC_CUDA(cudaMemcpyAsync(dx + sent, x + sent, (size_t)(need - sent) * sizeof(T), cudaMemcpyHostToDevice,
                                 m->pipe_in));
      sent = need;
    }
    if (ACC)
      SLC_CUDA(cudaMemcpyAsync(dy + r0, y + r0, (size_t)(r1 - r0) * sizeof(T), cudaMemcpyHostToDevice, m->pipe_in));
    mark(m->pipe_in);
    cudaEvent_t ev_in = m->pipe_events[i], ev_k = m->pipe_events[PIPE_MAX_CHUNKS + i];
    SLC_CUDA(cudaEventRecord(ev_in, m->pipe_in));
    if (mode == 1) {
      // chunk kernels alternate between the caller's stream and the side stream, and each chunk's
      // download follows its kernel on the same stream: one cross-stream dependency per chunk (upload
      // -> kernel) instead of two.
      cudaStream_t cs = (i & 1) ? m->pipe_out : main_s;
      SLC_CUDA(cudaStreamWaitEvent(cs, ev_in, 0));
      if (!(diag & 4)) {
        tls().stream = cs;
        const int rc = launch_sell<T, ACC, LocalX<T>>(m, LocalX<T>{dx}, dy, r0, r1);
        tls().stream = main_s;
        SLC_TRY(rc);
      }
      mark(cs);
      if (!(diag & 1))
        SLC_CUDA(cudaMemcpyAsync(y + r0, dsrc + r0, (size_t)(r1 - r0) * sizeof(T), cudaMemcpyDeviceToHost, cs));
      mark(cs);
      continue;
    }
    SLC_CUDA(cudaStreamWaitEvent(main_s, ev_in, 0));
    if (!(diag & 4)) SLC_TRY((launch_sell<T, ACC, LocalX<T>>(m, LocalX<T>{dx}, dy, r0, r1)));
    mark(main_s);
    SLC_CUDA(cudaEventRecord(ev_k, main_s));
    SLC_CUDA(cudaStreamWaitEvent(m->pipe_out, ev_k, 0));
    if (!(diag & 1))
      SLC_CUDA(cudaMemcpyAsync(y + r0, dsrc + r0, (size_t)(r1 - r0) * sizeof(T), cudaMemcpyDeviceToHost, m->pipe_out));
    mark(m->pipe_out);
  }
  // one wait: everything the call enqueued ends in the last download (or, in mode 1, on the two kernel streams)
  if (mode == 1) SLC_CUDA(cudaStreamSynchronize(main_s));
  SLC_CUDA(cudaStreamSynchronize(m->pipe_out));
  if (trace) {
    cudaStreamSynchronize(m->pipe_in);
    cudaStreamSynchronize(main_s);
    fprintf(stderr, "slc e2e trace (ms after the call started): chunk rows | upload done | kernel done | download done\n");
    for (int i = 0; i < nch && 3 * i + 3 < (int)tev.size(); ++i) {
      float a = 0, b = 0, c = 0;
      cudaEventElapsedTime(&a, tev[0], tev[3 * i + 1]);
      cudaEventElapsedTime(&b, tev[0], tev[3 * i + 2]);
      cudaEventElapsedTime(&c, tev[0], tev[3 * i + 3]);
      fprintf(stderr, "  %2d [%8d, %8d) | %7.3f | %7.3f | %7.3f\n", i, cuts[i], cuts[i + 1], a, b, c);
    }
    for (auto e : tev) cudaEventDestroy(e);
  }
  *done = true;
  return SLC_OK;
}

template <class T, bool ACC>
static int spmv_host_entry(slc_matrix *m, const double *x, double *y, const char *fn) {
  SLC_TRY(check_handle<T>(m, fn));
  HandleLock lock(m);
  if ((m->ncols && !x) || (m->nrows && !y)) return fail(SLC_ERR_INVALID, "%s: null vector", fn);
  // The chunked pipeline below overlaps upload, product and download of PINNED operands (the copies are
  // asynchronous DMA).  On pageable memory every cudaMemcpyAsync of it degenerates into the driver's own
  // staging (11 GB/s in, 19 out, profiles/r3_probe_host_link.txt), so large pageable vectors -- what a
  // Storable vector on the GHC heap is -- take the whole-vector path through the pinned ring instead.
  const bool pageable = host_needs_staging(x, (size_t)m->ncols * sizeof(T)) ||
                        host_needs_staging(y, (size_t)m->nrows * sizeof(T));
  if (!pageable) {
    bool done = false;
    SLC_TRY((spmv_host_pipelined<T, ACC>(m, (const T *)x, (T *)y, &done)));
    if (done) return SLC_OK;
  }
  cudaStream_t s = cur_stream();
  DevBuf<T> dx, dy;
  SLC_TRY(dx.alloc((size_t)m->ncols));
  SLC_TRY(dy.alloc((size_t)m->nrows));
  if (m->ncols) SLC_TRY(host_to_dev(dx.p, x, (size_t)m->ncols * sizeof(T), s));
  if (ACC && m->nrows) SLC_TRY(host_to_dev(dy.p, y, (size_t)m->nrows * sizeof(T), s));
  SLC_TRY((spmv_dev<T, ACC>(m, dx.p, dy.p)));
  if (m->nrows) SLC_TRY(dev_to_host(y, dy.p, (size_t)m->nrows * sizeof(T), s));
  SLC_CUDA(cudaStreamSynchronize(s));
  return SLC_OK;
}

// Validate a partition, put its table (splits, part pointers, leading dimensions) in HBM next to the
// matrix (refreshed only when it changes) and return the device copy.
template <class T>
static int dist_table_setup(slc_matrix *m, int32_t nparts, int32_t self, const int32_t *splits,
                            const double *const *parts, const int64_t *lds, const char *fn, const DistTable **out) {
  SLC_TRY(check_handle<T>(m, fn));  // callers hold the handle lock
  if (nparts < 1 || nparts > MAX_PARTS) return fail(SLC_ERR_INVALID, "%s: 1..%d partitions supported", fn, MAX_PARTS);
  if (self < 0 || self >= nparts || !splits || !parts) return fail(SLC_ERR_INVALID, "%s: bad partition argument", fn);
  if (splits[0] != 0 || splits[nparts] != m->ncols)
    return fail(SLC_ERR_DIM, "axpy_: column dimension %d does not match operand dimension %d", m->ncols,
                splits[nparts]);
  DistTable t;
  memset(&t, 0, sizeof t);
  for (int p = 0; p <= MAX_PARTS; ++p) t.splits[p] = m->ncols;
  for (int p = 0; p < nparts; ++p) {
    if (splits[p] > splits[p + 1]) return fail(SLC_ERR_INVALID, "%s: partition bounds decrease", fn);
    if (splits[p + 1] > splits[p] && !parts[p]) return fail(SLC_ERR_INVALID, "%s: null slice %d", fn, p);
    if (lds && lds[p] < splits[p + 1] - splits[p])
      return fail(SLC_ERR_DIM, "axpy_: column dimension %d does not match operand dimension %lld",
                  splits[p + 1] - splits[p], (long long)lds[p]);
    t.parts[p] = parts[p];
    t.splits[p] = splits[p];
    t.ld[p] = lds ? lds[p] : 0;
  }
  t.splits[nparts] = splits[nparts];
  t.nparts = nparts;
  static_assert(sizeof(DistTable) <= slc_matrix::DIST_TABLE_BYTES, "dist table cache too small");
  if (!m->dist_table) {
    DevBuf<char> buf;
    SLC_TRY(buf.alloc((size_t)slc_matrix::DIST_SLOTS * slc_matrix::DIST_TABLE_BYTES));
    m->dist_table = buf.take();
    for (bool &v : m->dist_slot_valid) v = false;
  }
  int slot = -1;
  for (int i = 0; i < slc_matrix::DIST_SLOTS; ++i)
    if (m->dist_slot_valid[i] && memcmp(m->dist_table_host[i], &t, sizeof t) == 0) slot = i;
  if (slot < 0) {  // not cached: take the next slot round-robin (stream order protects a table still in use)
    slot = m->dist_next;
    m->dist_next = (m->dist_next + 1) % slc_matrix::DIST_SLOTS;
    memcpy(m->dist_table_host[slot], &t, sizeof t);
    SLC_CUDA(cudaMemcpyAsync((char *)m->dist_table + (size_t)slot * slc_matrix::DIST_TABLE_BYTES,
                             m->dist_table_host[slot], sizeof t, cudaMemcpyHostToDevice, cur_stream()));
    m->dist_slot_valid[slot] = true;
  }
  *out = (const DistTable *)((char *)m->dist_table + (size_t)slot * slc_matrix::DIST_TABLE_BYTES);
  return SLC_OK;
}

// Sliced-ELL plan: the longest run of consecutive slices whose columns all fall inside [lo, hi), this
// rank's own slice of the operand (for a z-slab stencil: everything but the first and last plane).
// Rows [r0, r1) need no peer; cached per (lo, hi).
static int dist_interior(slc_matrix *m, int32_t lo, int32_t hi, int32_t *r0, int32_t *r1) {
  SellPlan &sp = m->sell;
  if (sp.interior_lo != lo || sp.interior_hi != hi || sp.interior_s0 < 0) {
    SLC_TRY(fetch_slice_ranges(m));
    int32_t best0 = 0, best1 = 0, run0 = 0;  // empty slices count as local
    for (int32_t sl = 0; sl <= sp.nslices; ++sl) {
      const bool local = sl < sp.nslices && (sp.h_cmax[sl] < 0 || (sp.h_cmin[sl] >= lo && sp.h_cmax[sl] < hi));
      if (!local) {
        if (sl - run0 > best1 - best0) { best0 = run0; best1 = sl; }
        run0 = sl + 1;
      }
    }
    sp.interior_s0 = best0;
    sp.interior_s1 = best1;
    sp.interior_lo = lo;
    sp.interior_hi = hi;
    // column ranges the rows below / above the interior reach: which peers each side has to wait for
    for (int side = 0; side < 2; ++side) {
      int32_t cmin = 0x7fffffff, cmax = -1;
      const int32_t a = side == 0 ? 0 : best1, b = side == 0 ? best0 : sp.nslices;
      for (int32_t sl = a; sl < b; ++sl)
        if (sp.h_cmax[sl] >= 0) { cmin = std::min(cmin, sp.h_cmin[sl]); cmax = std::max(cmax, sp.h_cmax[sl]); }
      sp.side_cmin[side] = cmin;
      sp.side_cmax[side] = cmax;
    }
  }
  *r0 = sp.interior_s0 * SELL_C;
  *r1 = std::min(m->nrows, sp.interior_s1 * SELL_C);
  return SLC_OK;
}

// parts other than `self` whose slice overlaps the column range [cmin, cmax]
static uint32_t parts_overlapping(int32_t nparts, int32_t self, const int32_t *splits, int32_t cmin, int32_t cmax) {
  uint32_t mask = 0;
  if (cmax < cmin) return 0;
  for (int p = 0; p < nparts; ++p)
    if (p != self && splits[p + 1] > splits[p] && splits[p] <= cmax && splits[p + 1] > cmin) mask |= 1u << p;
  return mask;
}

// y_local <- A_local x, x partitioned by column ranges over the GPUs of one box (see DistX)
template <class T>
static int spmv_dist_entry(slc_matrix *m, int32_t nparts, int32_t self, const int32_t *splits,
                           const double *const *x_parts, double *y, slc_peer_group *sync, const char *fn) {
  HandleLock lock(m);
  const DistTable *tbl = nullptr;
  SLC_TRY((dist_table_setup<T>(m, nparts, self, splits, x_parts, nullptr, fn, &tbl)));
  if (m->nrows && !y) return fail(SLC_ERR_INVALID, "%s: bad partition argument", fn);
  DistX<T> xa;
  xa.self_x = (const T *)x_parts[self];
  xa.self_lo = splits[self];
  xa.self_hi = splits[self + 1];
  xa.tbl = tbl;
  if (m->nnz > 0) SLC_TRY(ensure_plan<T>(m));
  if (sync) SLC_TRY(peer_check(sync, fn));
  if (m->nrows == 0 || m->nnz == 0 || m->spmv_active != SPMV_SELL) {
    if (sync) SLC_TRY(slc_peer_barrier(sync));  // separate barrier kernel, then the generic kernels
    return spmv_run<T, false, DistX<T>>(m, xa, (T *)y);
  }

  // the interior slices run the undistributed row body on a shifted base pointer, at full single-GPU
  // speed; only the boundary slices take the peer path
  SellPlan &sp = m->sell;
  int32_t r0 = 0, r1 = 0;
  SLC_TRY(dist_interior(m, xa.self_lo, xa.self_hi, &r0, &r1));
  // one launch: [boundary blocks | interior blocks], signal and waits fused in (see k_spmv_sell_dist)
  PeerSync ps;
  memset(&ps, 0, sizeof ps);
  uint32_t wait_lo = 0, wait_hi = 0;
  const int64_t nboundary = ceil_div(r0, 256) + ceil_div(m->nrows - r1, 256);
  if (sync && nparts > 1) {
    ps = peer_sync_of(sync);
    // each side waits for the peers its rows read, plus the ranks that only READ this rank's slice
    // (deps without a need): their signal of this epoch says they are done with the previous one,
    // which is what lets the caller rewrite the other operand buffer (see slc.h, "double buffering")
    const uint32_t need_lo = parts_overlapping(nparts, self, splits, sp.side_cmin[0], sp.side_cmax[0]);
    const uint32_t need_hi = parts_overlapping(nparts, self, splits, sp.side_cmin[1], sp.side_cmax[1]);
    const uint32_t extra = sync->deps_mask & ~(need_lo | need_hi);
    wait_lo = need_lo | extra;
    wait_hi = need_hi | extra;
    if (nboundary == 0 && extra) {  // nothing here reads a peer, but peers read this rank: plain barrier
      SLC_TRY(slc_peer_barrier(sync));
      ps.flags = nullptr;
    }
  }
  const int64_t ninterior = ceil_div(r1 - r0, 256);
  const int64_t nblocks = nboundary + ninterior;
  // share of the interior blocks handed out before the boundary blocks, per cent (see the kernel)
  static const int lead_env = env_int("SLC_DIST_LEAD", 40);
  const int32_t lead = (int32_t)(ninterior * std::max(0, std::min(100, lead_env)) / 100);
  cudaStream_t s = cur_stream();
  static const int u_env = env_int("SLC_SELL_U", 0);
  const int u = u_env ? u_env : ((m->max_rowlen <= 4 || sp.kbase) ? 4 : 8);
  static const int pdl_env = env_int("SLC_PDL", 0);  // see launch_sell
  const bool pdl = pdl_env != 0;
  const int32_t first_wave = pdl ? sm_count() * 8 : 0;
  cudaError_t le = cudaSuccess;
#define SLC_DIST_CASE(UU, KB_)                                                                                  \
  le = launch_pdl(k_spmv_sell_dist<T, UU, KB_>, (unsigned)nblocks, 256, 0, s, pdl, m->nrows, r0, r1, sp.slice_ptr, \
                  sp.rowlen, sp.colidx, (const T *)sp.vals, xa, (T *)y, ps, wait_lo, wait_hi, sp.kbase, lead, first_wave)
  if (sp.kbase) {
    if (u <= 4) SLC_DIST_CASE(4, true);
    else SLC_DIST_CASE(8, true);
  } else {
    if (u <= 4) SLC_DIST_CASE(4, false);
    else SLC_DIST_CASE(8, false);
  }
#undef SLC_DIST_CASE
  SLC_CUDA(le);
  SLC_LAUNCHED();
  return SLC_OK;
}

// Which peers the rows of this block read operand entries from (bit p = part p), for the caller that
// wants neighbour-only waits: every rank computes its mask, the masks are exchanged, and rank r's
// dependency set is  need(r) | {q : r in need(q)}  (slc_peer_group_set_deps).
template <class T>
static int dist_needs_entry(slc_matrix *m, int32_t nparts, int32_t self, const int32_t *splits, uint32_t *mask,
                            const char *fn) {
  SLC_TRY(check_handle<T>(m, fn));
  if (!mask || !splits || nparts < 1 || nparts > MAX_PARTS || self < 0 || self >= nparts)
    return fail(SLC_ERR_INVALID, "%s: bad partition argument", fn);
  *mask = 0;
  if (m->nnz == 0 || m->nrows == 0) return SLC_OK;
  SLC_TRY(ensure_plan<T>(m));
  if (m->spmv_active != SPMV_SELL) {  // irregular rows: the whole operand is gathered
    for (int p = 0; p < nparts; ++p)
      if (p != self && splits[p + 1] > splits[p]) *mask |= 1u << p;
    return SLC_OK;
  }
  int32_t r0 = 0, r1 = 0;
  SLC_TRY(dist_interior(m, splits[self], splits[self + 1], &r0, &r1));
  const SellPlan &sp = m->sell;
  *mask = parts_overlapping(nparts, self, splits, sp.side_cmin[0], sp.side_cmax[0]) |
          parts_overlapping(nparts, self, splits, sp.side_cmin[1], sp.side_cmax[1]);
  return SLC_OK;
}

// All-gather of the operand over peer memory: every block pulls 16-byte words of one part's slice
// (its own included) into this rank's full-length copy of x.  blockIdx.y = part.  Used when the rows
// of a rank touch columns all over the matrix (power-law graphs: SURVEY.md section 8e, config 4), where
// per-entry loads across NVLink would be one 32-byte sector per nonzero.
template <class T>
__global__ void __launch_bounds__(256) k_allgather_pull(const DistTable *tbl, T *xfull, int self) {
  // The parts are taken in the order self + 1, self + 2, ..., self (blocks are dispatched y-slowest, so
  // blockIdx.y is also an order in time): every rank starts on a different owner.  With the order 0, 1, 2, ...
  // all ranks pulled from the same one or two GPUs at any moment and those GPUs' NVLink egress, not the
  // pullers' ingress, set the pace (8 ranks, R-MAT scale 24: 0.5 of the 0.94 ms step).
  if ((int)blockIdx.y >= tbl->nparts) return;
  const int p = (int)((blockIdx.y + (unsigned)self + 1u) % (unsigned)tbl->nparts);
  const int64_t lo = tbl->splits[p], n = (int64_t)tbl->splits[p + 1] - lo;
  const T *src = static_cast<const T *>(tbl->parts[p]);
  T *dst = xfull + lo;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  constexpr int PER = 16 / sizeof(T) > 0 ? 16 / sizeof(T) : 1;  // elements per 16-byte word
  if (PER > 1 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
    const int64_t nw = n / PER;
    const double2 *s2 = reinterpret_cast<const double2 *>(src);
    double2 *d2 = reinterpret_cast<double2 *>(dst);
    int64_t i = tid;
    for (; i + 3 * nth < nw; i += 4 * nth) {  // four loads in flight per thread
      const double2 a = s2[i], b = s2[i + nth], c = s2[i + 2 * nth], d = s2[i + 3 * nth];
      d2[i] = a; d2[i + nth] = b; d2[i + 2 * nth] = c; d2[i + 3 * nth] = d;
    }
    for (; i < nw; i += nth) d2[i] = s2[i];
    for (int64_t e = nw * PER + tid; e < n; e += nth) dst[e] = src[e];
  } else {
    for (int64_t e = tid; e < n; e += nth) dst[e] = src[e];
  }
}

// y_local <- A_local x with x all-gathered first: barrier, pull kernel, then the undistributed product
// (merge-path balanced kernels for irregular rows) on the full local copy.
template <class T>
static int spmv_allgather_entry(slc_matrix *m, int32_t nparts, int32_t self, const int32_t *splits,
                                const double *const *x_parts, double *x_full, double *y, slc_peer_group *sync,
                                const char *fn) {
  HandleLock lock(m);
  const DistTable *tbl = nullptr;
  SLC_TRY((dist_table_setup<T>(m, nparts, self, splits, x_parts, nullptr, fn, &tbl)));
  if ((m->ncols && !x_full) || (m->nrows && !y)) return fail(SLC_ERR_INVALID, "%s: null vector", fn);
  if (sync) SLC_TRY(peer_check(sync, fn));
  static const int diag_env = env_int("SLC_AG_DIAG", 0);  // timing diagnosis: 1 = no pull, 2 = no product, 3 = no barrier
  if (sync && diag_env != 3) SLC_TRY(slc_peer_barrier(sync));
  if (m->ncols && diag_env != 1) {
    int64_t longest = 0;
    for (int p = 0; p < nparts; ++p) longest = std::max<int64_t>(longest, splits[p + 1] - splits[p]);
    const unsigned bx = (unsigned)std::max<int64_t>(1, std::min<int64_t>(ceil_div(longest, 256 * 8), 148 * 4));
    k_allgather_pull<T><<<dim3(bx, (unsigned)nparts), 256, 0, cur_stream()>>>(tbl, (T *)x_full, (int)self);
    SLC_LAUNCHED();
  }
  if (diag_env == 2) return SLC_OK;
  return spmv_dev<T, false>(m, (const T *)x_full, (T *)y);
}

// k_spmm_panel: block size, shared-memory budget and batch shape.  Defaults from the round-1 sweep on
// the 27-point stencil; the environment variables exist for that sweep.
// Line stride of a stencil-like matrix: the distance in rows between neighbouring grid lines, read off
// the middle row of the row-major mirror.
static int detect_line_stride(slc_matrix *m) {
  SellPlan &sp = m->sell;
  if (sp.line_stride >= 0) return SLC_OK;
  sp.line_stride = 0;
  if (!m->has_csr || m->nrows < 4096) return SLC_OK;
  const int32_t r = m->nrows / 2;
  int32_t rp[2];
  cudaStream_t s = cur_stream();
  SLC_CUDA(cudaMemcpyAsync(rp, m->rowptr + r, sizeof rp, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  const int32_t len = rp[1] - rp[0];
  if (len < 3 || len > 4096) return SLC_OK;
  std::vector<int32_t> cols((size_t)len);
  SLC_CUDA(cudaMemcpyAsync(cols.data(), m->colidx + rp[0], (size_t)len * 4, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  // runs of consecutive columns; the stride is the smallest distance between the centres of two
  // neighbouring runs (independent of where the row block starts in a partitioned matrix)
  int64_t best = 0, prev_centre = -1;
  for (size_t i = 0; i < cols.size();) {
    size_t j = i;
    while (j + 1 < cols.size() && cols[j + 1] == cols[j] + 1) ++j;
    const int64_t centre = ((int64_t)cols[i] + cols[j]) / 2;
    if (prev_centre >= 0 && (best == 0 || centre - prev_centre < best)) best = centre - prev_centre;
    prev_centre = centre;
    i = j + 1;
  }
  if (best >= 32 && best <= m->nrows / 64) sp.line_stride = (int32_t)best;
  return SLC_OK;
}

// Slice order of k_spmm_panel for a matrix with grid lines `stride` rows apart: a block takes `width`
// neighbouring slices from each of `warps / width` consecutive lines, so the rows of X that one warp
// fetches for the lines above and below it are the rows its neighbours in the block need too (a block
// of consecutive rows re-fetches every line three times from L2).  Line i is the slices
// [ceil(i stride / 32), ceil((i+1) stride / 32)); warps without a slice idle.  Any order is correct:
// the table only decides which rows share an SM.
static int panel_tiling(slc_matrix *m, int32_t s_begin, int32_t s_end, int32_t stride, int warps, int width) {
  SellPlan &sp = m->sell;
  const int32_t key[4] = {s_begin, s_end, stride, (warps << 8) | width};
  if (sp.panel_order && memcmp(key, sp.panel_key, sizeof key) == 0) return SLC_OK;
  const int lines_per = warps / width;
  auto line_first = [&](int64_t i) { return (int32_t)std::min<int64_t>(s_end, std::max<int64_t>(s_begin, ceil_div(i * stride, (int64_t)SELL_C))); };
  const int64_t line0 = ((int64_t)s_begin * SELL_C) / stride, line1 = ceil_div((int64_t)s_end * SELL_C, (int64_t)stride);
  std::vector<int32_t> order;
  for (int64_t l = line0; l < line1; l += lines_per) {
    int32_t widest = 0;
    for (int i = 0; i < lines_per; ++i) widest = std::max(widest, line_first(l + i + 1) - line_first(l + i));
    for (int32_t q = 0; q < widest; q += width)
      for (int i = 0; i < lines_per; ++i) {
        const int32_t a = line_first(l + i), b = line_first(l + i + 1);
        for (int w = 0; w < width; ++w) order.push_back(a + q + w < b ? a + q + w : -1);
      }
  }
  if (sp.panel_order) { dev_free(sp.panel_order); sp.panel_order = nullptr; }
  DevBuf<int32_t> buf;
  SLC_TRY(buf.alloc(order.size()));
  cudaStream_t s = cur_stream();
  SLC_CUDA(cudaMemcpyAsync(buf.p, order.data(), order.size() * 4, cudaMemcpyHostToDevice, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  sp.panel_order = buf.take();
  sp.panel_blocks = (int64_t)(order.size() / (size_t)warps);
  memcpy(sp.panel_key, key, sizeof key);
  return SLC_OK;
}

template <class T, int JT, int KU>
static int launch_panel_inst(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, int threads,
                             int32_t kcap, int32_t row_begin, int32_t row_end, const YInit<T> &yi) {
  SellPlan &sp = m->sell;
  const size_t smem = (size_t)kcap * threads * (sizeof(T) + sizeof(int32_t));
  static PerDevice<size_t> smem_set;  // per instantiation, per device
  SLC_CUDA(ensure_dyn_smem(k_spmm_panel<T, JT, KU>, smem_set, smem));
  if (row_end <= row_begin) return SLC_OK;
  int64_t nb = ceil_div(row_end - row_begin, threads);
  const int32_t *order = nullptr;
  // Line tiling lost its A/B on the 27-point stencil (7.25 ms against 6.29 ms for consecutive rows,
  // DESIGN.md 4.3), so it is opt-in: SLC_PANEL_STRIDE=-1 detects the stride, > 0 sets it.  Read on
  // every call (not cached) so a test can switch it on.
  const int stride_env = env_int("SLC_PANEL_STRIDE", 0);
  const int width_env = env_int("SLC_PANEL_WIDTH", 1);
  if (stride_env < 0) SLC_TRY(detect_line_stride(m));
  const int32_t stride = stride_env >= 0 ? stride_env : sp.line_stride;
  const int warps = threads / 32;
  const int width = (width_env >= 1 && warps % width_env == 0) ? width_env : 1;
  if (stride >= 32 && warps / width >= 2 && row_begin % SELL_C == 0) {
    SLC_TRY(panel_tiling(m, row_begin / SELL_C, (int32_t)ceil_div(row_end, SELL_C), stride, warps, width));
    order = sp.panel_order;
    nb = sp.panel_blocks;
  }
  k_spmm_panel<T, JT, KU><<<(unsigned)nb, threads, smem, cur_stream()>>>(
      row_begin, row_end, kcap, 0u, sp.slice_ptr, sp.rowlen, sp.colidx, (const T *)sp.vals, X, ldx, Y, ldy, k, order, yi);
  SLC_LAUNCHED();
  return SLC_OK;
}

template <class T>
static int launch_panel(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, const YInit<T> &yi,
                        int32_t row_begin = 0, int32_t row_end = -1, int64_t xlo = 0, int64_t xhi = -1) {
  if (row_end < 0) row_end = m->nrows;
  {  // 27-offset matrices: X staged in shared memory by the TMA engine (slc_spmm_brick.cu)
    bool done = false;
    SLC_TRY(spmm_brick_try(m, X, ldx, Y, ldy, k, yi, row_begin, row_end, xlo, xhi, &done));
    if (done) return SLC_OK;
  }
  static const int threads_env = env_int("SLC_PANEL_THREADS", 0);
  static const int smem_env = env_int("SLC_PANEL_SMEM_KB", 0);
  static const int jt_env = env_int("SLC_PANEL_JT", 0);
  static const int ku_env = env_int("SLC_PANEL_KU", 0);
  // measured on B200, 27-point stencil 200^3, k = 64 Double: 256 threads / 96 KB 6.32 ms,
  // 128 threads / 48 KB 8.69 ms, 64 threads 9.19 ms (a bigger row block re-fetches less of X from L2)
  int threads = threads_env ? threads_env : 256;
  threads = std::max(32, std::min(512, threads / 32 * 32));
  const int budget_kb = smem_env ? std::min(smem_env, 200) : (threads <= 128 ? 48 : threads <= 256 ? 96 : 192);
  const int32_t kcap =
      std::max(1, std::min<int32_t>(m->max_rowlen, (int32_t)((size_t)budget_kb * 1024 / ((size_t)threads * (sizeof(T) + 4)))));
  // batch depth: the one that leaves fewer single-entry leftovers on the longest row
  int ku = ku_env ? ku_env : ((m->max_rowlen % 3) < (m->max_rowlen % 4) ? 3 : 4);
  if (m->max_rowlen < 3) ku = 2;
  const bool wide = jt_env ? jt_env >= 8 : sizeof(T) == 8;  // 8 columns per tile for Double, 4 for Complex
#define SLC_PANEL(JT_, KU_) \
  return launch_panel_inst<T, JT_, KU_>(m, X, ldx, Y, ldy, k, threads, kcap, row_begin, row_end, yi)
  if (wide) {
    if (ku <= 2) SLC_PANEL(8, 2);
    if (ku == 3) SLC_PANEL(8, 3);
    SLC_PANEL(8, 4);
  } else {
    if (ku <= 2) SLC_PANEL(4, 2);
    if (ku == 3) SLC_PANEL(4, 3);
    SLC_PANEL(4, 4);
  }
#undef SLC_PANEL
}

// Y[:, 0:k] <- beta Y (an operand with no stored entry: the product term vanishes)
template <class T>
__global__ void k_scale_block(int32_t nrows, int32_t k, T *__restrict__ Y, int64_t ldy, T beta) {
  const int64_t n = (int64_t)nrows * k;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    T *p = Y + (i / nrows) * ldy + (i % nrows);
    *p = Num<T>::mul(beta, *p);
  }
}

template <class T>
static int spmm_dev(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, const YInit<T> &yi) {
  if (k == 0 || m->nrows == 0) return SLC_OK;
  cudaStream_t s = cur_stream();
  if (m->nnz == 0) {
    if (yi.mode == 0)
      SLC_CUDA(cudaMemset2DAsync(Y, (size_t)ldy * sizeof(T), 0, (size_t)m->nrows * sizeof(T), (size_t)k, s));
    else if (yi.mode == 2) {
      k_scale_block<T><<<(unsigned)std::min<int64_t>(ceil_div((int64_t)m->nrows * k, 256), 148 * 16), 256, 0, s>>>(
          m->nrows, k, Y, ldy, yi.beta);
      SLC_LAUNCHED();
    }
    return SLC_OK;
  }
  SLC_TRY(ensure_plan<T>(m));
  if (m->max_rowlen >= 1 && m->max_rowlen <= 8) {  // short rows: registers hold the row
    bool done = false;  // banded: X windows staged by TMA (slc_spmm_window.cu); else per-thread gathers
    SLC_TRY(spmm_window_try<T>(m, X, ldx, Y, ldy, k, yi, &done));
    if (done) return SLC_OK;
    static const int jg_env = env_int("SLC_SPMM_JG", 0);
    static const int stride_env = env_int("SLC_SPMM_STRIDE", -1);
    static const int threads_env = env_int("SLC_SPMM_THREADS", 0);
    const int32_t jgroup = jg_env > 0 ? jg_env : 64;
    const int32_t ngroups = (int32_t)ceil_div(k, jgroup);
    const int32_t stride = stride_env >= 0 ? stride_env : m->band_stride;
    const int32_t stride32 = (stride >= 64 && stride % 32 == 0) ? stride / 32 : 0;
    const int threads = threads_env ? threads_env : (stride32 ? 256 : 128);
    const int64_t row_blocks = stride32 ? ceil_div(m->nrows, (int64_t)(threads / 32) * stride) * stride32
                                        : ceil_div(m->nrows, threads);
    const int64_t nb = row_blocks * ngroups;
    if (nb > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "mulM: grid too large");
    const bool sell = m->spmv_active == SPMV_SELL;
    const SellPlan &sp = m->sell;
    constexpr int JTS = sizeof(T) == 16 ? 2 : 4;
    // PIPE: two register buffers of gathers per thread (SELL plans, rows of at most 6 entries: one
    // buffer is then at most 48 registers; see spmm_short_columns)
    static const int pipe_env = env_int("SLC_SPMM_PIPE", 0);  // measured: 0.535 ms with, 0.473 without (cfg 2)
    const bool pipe = sell && pipe_env != 0;
#define SLC_SHORT(L_, SELL_, PIPE_)                                                                  \
  k_spmm_short<T, JTS, L_, SELL_, (PIPE_) && (L_ * JTS * sizeof(T) / 4 <= 48)><<<(unsigned)nb, threads, 0, s>>>( \
      m->nrows, ngroups, jgroup, stride32, 0u, sp.slice_ptr, sp.rowlen, m->rowptr,                    \
      SELL_ ? sp.colidx : m->colidx, SELL_ ? (const T *)sp.vals : (const T *)m->rvals, X, ldx, Y, ldy, k, yi)
#define SLC_SHORT_L(L_)                                                                              \
  case L_:                                                                                           \
    if (pipe) SLC_SHORT(L_, true, true);                                                             \
    else if (sell) SLC_SHORT(L_, true, false);                                                       \
    else SLC_SHORT(L_, false, false);                                                                \
    break
    switch (m->max_rowlen < 8 ? m->max_rowlen : 8) {
      SLC_SHORT_L(1);
      SLC_SHORT_L(2);
      SLC_SHORT_L(3);
      SLC_SHORT_L(4);
      SLC_SHORT_L(5);
      SLC_SHORT_L(6);
      SLC_SHORT_L(7);
      SLC_SHORT_L(8);
    }
#undef SLC_SHORT_L
#undef SLC_SHORT
    SLC_LAUNCHED();
    return SLC_OK;
  }
  static const int panel_env = env_int("SLC_SPMM_PANEL", 1);
  if (m->spmv_active == SPMV_SELL && panel_env) return launch_panel<T>(m, X, ldx, Y, ldy, k, yi);
  constexpr int JT = sizeof(T) == 16 ? 4 : 8;
  const int32_t ntiles = (int32_t)ceil_div(k, JT);
  const int64_t nblocks = ceil_div(m->nrows, 128) * ntiles;
  if (nblocks > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "mulM: grid too large");
  if (m->spmv_active == SPMV_SELL) {
    const SellPlan &sp = m->sell;
    k_spmm_rows<T, JT, true><<<(unsigned)nblocks, 128, 0, s>>>(m->nrows, ntiles, sp.slice_ptr, sp.rowlen,
                                                                nullptr, sp.colidx, (const T *)sp.vals, X,
                                                                ldx, Y, ldy, k, yi);
  } else {
    k_spmm_rows<T, JT, false><<<(unsigned)nblocks, 128, 0, s>>>(m->nrows, ntiles, nullptr, nullptr,
                                                                 m->rowptr, m->colidx,
                                                                 (const T *)m->rvals, X, ldx, Y, ldy, k, yi);
  }
  SLC_LAUNCHED();
  return SLC_OK;
}

// Y_local <- A_local X, X partitioned by row blocks over the GPUs of one box (column-major, every
// part with its own leading dimension).  Sliced-ELL plans only (the stencils this is for): signal,
// interior rows on the undistributed panel kernel, wait, boundary rows through the peers' memory.
// Halo staging for the boundary rows of a row-partitioned block product: the columns [cmin, cmin + width)
// that one side's boundary rows reach -- a few grid planes, partly this rank's own rows of X, partly a
// neighbour's -- are copied for all k operand columns into one contiguous block W (leading dimension ldw).
// The remote part crosses NVLink as coalesced streams instead of one 8-byte load per stored entry, and the
// boundary rows then run on the same kernels as the interior (TMA-staged for 27-offset matrices) with W as
// their operand.  blockIdx.y = operand column.
template <class T>
__global__ void __launch_bounds__(256)
k_halo_stage(const DistTable *tbl, T *__restrict__ W, int64_t ldw, int32_t cmin, int32_t width) {
  const int np = tbl->nparts;
  const int32_t ncols = tbl->splits[np];
  const int64_t j = blockIdx.y;
  int q = 0;
  for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < width; i += gridDim.x * blockDim.x) {
    const int32_t c = cmin + i;
    T v = Num<T>::zero();  // columns outside the matrix (alignment slack) are never multiplied
    if (c >= 0 && c < ncols) {
      while (q + 1 < np && c >= tbl->splits[q + 1]) ++q;  // i only grows: the owner index only grows
      v = ld_peer(static_cast<const T *>(tbl->parts[q]) + j * tbl->ld[q] + (c - tbl->splits[q]));
    }
    W[j * ldw + i] = v;
  }
}

template <class T>
static int spmm_dist_entry(slc_matrix *m, int32_t nparts, int32_t self, const int32_t *splits,
                           const double *const *X_parts, const int64_t *ldx_parts, double *Y, int64_t ldy,
                           int32_t k, slc_peer_group *sync, const char *fn) {
  HandleLock lock(m);
  const DistTable *tbl = nullptr;
  if (!ldx_parts) return fail(SLC_ERR_INVALID, "%s: null leading dimensions", fn);
  if (sync) SLC_TRY(peer_check(sync, fn));
  SLC_TRY((dist_table_setup<T>(m, nparts, self, splits, X_parts, ldx_parts, fn, &tbl)));
  if (k < 0) return fail(SLC_ERR_INVALID, "%s: negative column count", fn);
  if (k > 0 && m->nrows && !Y) return fail(SLC_ERR_INVALID, "%s: null block", fn);
  if (ldy < m->nrows)
    return fail(SLC_ERR_DIM, "axpy_: row dimension %d does not match result dimension %lld", m->nrows,
                (long long)ldy);
  cudaStream_t s = cur_stream();
  if (k == 0 || m->nrows == 0 || m->nnz == 0) {
    if (sync) SLC_TRY(slc_peer_barrier(sync));
    if (k > 0 && m->nrows > 0)
      SLC_CUDA(cudaMemset2DAsync(Y, (size_t)ldy * sizeof(T), 0, (size_t)m->nrows * sizeof(T), (size_t)k, s));
    return SLC_OK;
  }
  SLC_TRY(ensure_plan<T>(m));
  if (m->spmv_active != SPMV_SELL) {
    // the distributed block product is written for the sliced layout only
    if (m->sell.padded > 2 * m->nnz + 1024)
      return fail(SLC_ERR_INVALID, "%s: the row-partitioned block product needs a sliced-ELL plan (padding %lld > nnz %lld)",
                  fn, (long long)(m->sell.padded - m->nnz), (long long)m->nnz);
    SLC_TRY(fill_sell<T>(m));
  }
  DistXM<T> xa;
  xa.self_X = (const T *)X_parts[self];
  xa.self_ld = ldx_parts[self];
  xa.self_lo = splits[self];
  xa.self_hi = splits[self + 1];
  xa.tbl = tbl;
  int32_t r0 = 0, r1 = 0;
  SLC_TRY(dist_interior(m, xa.self_lo, xa.self_hi, &r0, &r1));
  // Plans that are built lazily read words back from the device (a stream synchronisation).  Build them
  // here: a block without interior rows would otherwise meet the first of them in its boundary launch,
  // behind slc_peer_wait, and the host would sit on a wait kernel whose peers it has not launched yet.
  SLC_TRY(spmm_brick_prepare(m));
  if (env_int("SLC_PANEL_STRIDE", 0) < 0) SLC_TRY(detect_line_stride(m));
  if (sync) SLC_TRY(slc_peer_signal(sync));  // "my rows of X are written" (stream order)
  // interior rows: columns are global, so shift the local block's base by the first owned row
  SLC_TRY(launch_panel<T>(m, xa.self_X - xa.self_lo, xa.self_ld, (T *)Y, ldy, k, yinit_zero<T>(), r0, r1, xa.self_lo,
                          xa.self_hi));
  if (sync) SLC_TRY(slc_peer_wait(sync));
  // SLC_SPMM_HALO (1): boundary rows through a staged copy of the columns they reach (k_halo_stage) and the
  // interior kernels; 0 = k_spmm_boundary (per-entry loads from the owner's HBM) for every boundary row.
  // A side whose rows reach too far for a staged copy to pay keeps the per-entry kernel.
  bool staged_side[2] = {false, false};
  if (env_int("SLC_SPMM_HALO", 1) != 0) {
    for (int side = 0; side < 2; ++side) {
      const int32_t ra = side == 0 ? 0 : r1, rb = side == 0 ? r0 : m->nrows;
      const SellPlan &sp = m->sell;
      if (rb <= ra || sp.side_cmax[side] < sp.side_cmin[side]) { staged_side[side] = rb <= ra; continue; }
      const int32_t cmin = sp.side_cmin[side] & ~1;                 // even bounds and leading dimension: the
      const int32_t cend = (sp.side_cmax[side] + 2) & ~1;           // TMA-staged kernel wants 16-byte alignment
      const int64_t width = (int64_t)cend - cmin;
      if (width > 8 * (int64_t)(rb - ra) + 65536 || (size_t)width * k * sizeof(T) > ((size_t)1 << 30)) continue;
      DevBuf<T> w;
      SLC_TRY(w.alloc((size_t)width * k));
      k_halo_stage<T><<<dim3((unsigned)std::min<int64_t>(ceil_div(width, 256 * 4), 1024), (unsigned)k), 256, 0, s>>>(
          tbl, w.p, width, cmin, (int32_t)width);
      SLC_LAUNCHED();
      SLC_TRY(launch_panel<T>(m, w.p - cmin, width, (T *)Y, ldy, k, yinit_zero<T>(), ra, rb, cmin, cend));
      staged_side[side] = true;  // w is released in stream order behind the kernels that read it
    }
  }
  const int32_t b0 = staged_side[0] ? 0 : r0, b1 = staged_side[1] ? m->nrows : r1;  // rows left for the per-entry kernel
  const int64_t nb = ceil_div(b0, 128) + ceil_div(m->nrows - b1, 128);
  if (nb > 0) {
    const SellPlan &sp = m->sell;
    const int32_t r0 = b0, r1 = b1;
    constexpr int JT = sizeof(T) == 16 ? 4 : 8;
    k_spmm_boundary<T, JT><<<dim3((unsigned)nb, (unsigned)ceil_div(k, JT)), 128, 0, s>>>(m->nrows, r0, r1, 0u, sp.slice_ptr, sp.rowlen, sp.colidx,
                                                        (const T *)sp.vals, xa, (T *)Y, ldy, k, yinit_zero<T>());
    SLC_LAUNCHED();
  }
  return SLC_OK;
}

template <class T>
static int spmm_check(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                      const char *fn) {
  SLC_TRY(check_handle<T>(m, fn));
  if (k < 0) return fail(SLC_ERR_INVALID, "%s: negative column count", fn);
  if (k > 0 && ((m->ncols && !X) || (m->nrows && !Y))) return fail(SLC_ERR_INVALID, "%s: null block", fn);
  if (ldx < m->ncols)
    return fail(SLC_ERR_DIM, "axpy_: column dimension %d does not match operand dimension %lld", m->ncols,
                (long long)ldx);
  if (ldy < m->nrows)
    return fail(SLC_ERR_DIM, "axpy_: row dimension %d does not match result dimension %lld", m->nrows,
                (long long)ldy);
  return SLC_OK;
}

// beta of the C ABI -> start value of the accumulators: NULL or 0 -> mulM (Y is not read), 1 -> the
// reference's axpy_ into a block the caller filled (exact: y itself starts the sum), else beta * y
static YInit<double> yinit_of(const double *beta, double) {
  YInit<double> yi = yinit_zero<double>();
  if (beta && beta[0] != 0.0) { yi.mode = beta[0] == 1.0 ? 1 : 2; yi.beta = beta[0]; }
  return yi;
}
static YInit<double2> yinit_of(const double *beta, double2) {
  YInit<double2> yi = yinit_zero<double2>();
  if (beta && (beta[0] != 0.0 || beta[1] != 0.0)) {
    yi.mode = (beta[0] == 1.0 && beta[1] == 0.0) ? 1 : 2;
    yi.beta = double2{beta[0], beta[1]};
  }
  return yi;
}

template <class T>
static int spmm_dev_entry(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                          const double *beta, const char *fn) {
  SLC_TRY(spmm_check<T>(m, X, ldx, Y, ldy, k, fn));
  HandleLock lock(m);
  return spmm_dev<T>(m, (const T *)X, ldx, (T *)Y, ldy, k, yinit_of(beta, T()));
}

template <class T>
static int spmm_host_entry(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                           const double *beta, const char *fn) {
  SLC_TRY(spmm_check<T>(m, X, ldx, Y, ldy, k, fn));
  if (k == 0) return SLC_OK;
  HandleLock lock(m);
  const YInit<T> yi = yinit_of(beta, T());
  cudaStream_t s = cur_stream();
  // Device blocks are packed (leading dimension = ncols / nrows); the strided host blocks are
  // moved with 2-D copies so only the active columns cross PCIe.
  DevBuf<T> dX, dY;
  SLC_TRY(dX.alloc((size_t)m->ncols * k));
  SLC_TRY(dY.alloc((size_t)m->nrows * k));
  // Columns are independent products (FE:203-208 serves them one by one), so the block is cut into
  // column groups: group g+1 goes up on one stream while group g is multiplied on the caller's
  // stream and group g-1 comes back on a third -- PCIe is full duplex, so the upload and the
  // download overlap instead of running back to back.
  static const int groups_env = env_int("SLC_E2E_COLGROUPS", 8);
  const size_t bytes_total = ((size_t)m->ncols + (size_t)m->nrows) * (size_t)k * sizeof(T);
  int ngroups = std::max(1, std::min<int>(std::min(groups_env, PIPE_MAX_CHUNKS), k));
  if (bytes_total < ((size_t)8 << 20) || m->ncols == 0 || m->nrows == 0) ngroups = 1;
  if (ngroups > 1) {
    if (m->nnz > 0) SLC_TRY(ensure_plan<T>(m));
    if (!m->pipe_in) {
      SLC_CUDA(cudaStreamCreateWithFlags(&m->pipe_in, cudaStreamNonBlocking));
      SLC_CUDA(cudaStreamCreateWithFlags(&m->pipe_out, cudaStreamNonBlocking));
      m->pipe_events.resize(2 * PIPE_MAX_CHUNKS + 1);
      for (auto &e : m->pipe_events) SLC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    cudaEvent_t ev_start = m->pipe_events[2 * PIPE_MAX_CHUNKS];
    SLC_CUDA(cudaEventRecord(ev_start, s));
    SLC_CUDA(cudaStreamWaitEvent(m->pipe_in, ev_start, 0));
    SLC_CUDA(cudaStreamWaitEvent(m->pipe_out, ev_start, 0));
    const int32_t per = (int32_t)ceil_div(k, ngroups);
    int rc = SLC_OK;
    // Pageable blocks (what FEAST's MV.replicate'd work arrays are, Feast.hs:149-155): a cudaMemcpy2DAsync on
    // them is the driver's own staging at 11 GB/s in and 19 (2 into untouched pages) out.  Each column goes
    // through the pinned ring instead; the host is then the serial resource, so the order is: upload and
    // launch group g+1, then copy group g out while the device multiplies g+1.
    const bool pageable = host_needs_staging(X, (size_t)m->ncols * sizeof(T)) ||
                          host_needs_staging(Y, (size_t)m->nrows * sizeof(T));
    if (pageable) {
      const int G = (int)ceil_div(k, per);
      auto cols_of = [&](int g) { return std::min<int32_t>(per, k - g * per); };
      auto upload = [&](int g) -> int {
        const int32_t j0 = g * per;
        SLC_TRY(host_to_dev_2d(dX.p + (size_t)j0 * m->ncols, (size_t)m->ncols * sizeof(T), (const T *)X + (size_t)j0 * ldx,
                               (size_t)ldx * sizeof(T), (size_t)m->ncols * sizeof(T), (size_t)cols_of(g), m->pipe_in));
        if (yi.mode != 0)
          SLC_TRY(host_to_dev_2d(dY.p + (size_t)j0 * m->nrows, (size_t)m->nrows * sizeof(T), (const T *)Y + (size_t)j0 * ldy,
                                 (size_t)ldy * sizeof(T), (size_t)m->nrows * sizeof(T), (size_t)cols_of(g), m->pipe_in));
        SLC_CUDA(cudaEventRecord(m->pipe_events[g], m->pipe_in));
        return SLC_OK;
      };
      auto product = [&](int g) -> int {
        const int32_t j0 = g * per;
        SLC_CUDA(cudaStreamWaitEvent(s, m->pipe_events[g], 0));
        SLC_TRY(spmm_dev<T>(m, dX.p + (size_t)j0 * m->ncols, m->ncols, dY.p + (size_t)j0 * m->nrows, m->nrows, cols_of(g), yi));
        SLC_CUDA(cudaEventRecord(m->pipe_events[PIPE_MAX_CHUNKS + g], s));
        return SLC_OK;
      };
      auto download = [&](int g) -> int {
        const int32_t j0 = g * per;
        SLC_CUDA(cudaStreamWaitEvent(m->pipe_out, m->pipe_events[PIPE_MAX_CHUNKS + g], 0));
        SLC_TRY(dev_to_host_2d((T *)Y + (size_t)j0 * ldy, (size_t)ldy * sizeof(T), dY.p + (size_t)j0 * m->nrows,
                               (size_t)m->nrows * sizeof(T), (size_t)m->nrows * sizeof(T), (size_t)cols_of(g), m->pipe_out));
        return SLC_OK;
      };
      rc = upload(0);
      if (rc == SLC_OK) rc = product(0);
      for (int g = 0; g < G && rc == SLC_OK; ++g) {
        if (g + 1 < G) {
          rc = upload(g + 1);
          if (rc == SLC_OK) rc = product(g + 1);
        }
        if (rc == SLC_OK) rc = download(g);
      }
      cudaStreamSynchronize(m->pipe_out);
      cudaStreamSynchronize(m->pipe_in);
      SLC_CUDA(cudaStreamSynchronize(s));
      if (rc == SLC_OK) rc = spmm_window_check(fn);
      return rc;
    }
    for (int g = 0; g * per < k && rc == SLC_OK; ++g) {
      const int32_t j0 = g * per, kk = std::min<int32_t>(per, k - j0);
      const T *Xg = (const T *)X + (size_t)j0 * ldx;
      T *Yg = (T *)Y + (size_t)j0 * ldy;
      T *dXg = dX.p + (size_t)j0 * m->ncols, *dYg = dY.p + (size_t)j0 * m->nrows;
      cudaEvent_t ev_in = m->pipe_events[g], ev_k = m->pipe_events[PIPE_MAX_CHUNKS + g];
      SLC_CUDA(cudaMemcpy2DAsync(dXg, (size_t)m->ncols * sizeof(T), Xg, (size_t)ldx * sizeof(T),
                                 (size_t)m->ncols * sizeof(T), (size_t)kk, cudaMemcpyHostToDevice, m->pipe_in));
      if (yi.mode != 0)
        SLC_CUDA(cudaMemcpy2DAsync(dYg, (size_t)m->nrows * sizeof(T), Yg, (size_t)ldy * sizeof(T),
                                   (size_t)m->nrows * sizeof(T), (size_t)kk, cudaMemcpyHostToDevice, m->pipe_in));
      SLC_CUDA(cudaEventRecord(ev_in, m->pipe_in));
      SLC_CUDA(cudaStreamWaitEvent(s, ev_in, 0));
      rc = spmm_dev<T>(m, dXg, m->ncols, dYg, m->nrows, kk, yi);
      if (rc != SLC_OK) break;
      SLC_CUDA(cudaEventRecord(ev_k, s));
      SLC_CUDA(cudaStreamWaitEvent(m->pipe_out, ev_k, 0));
      SLC_CUDA(cudaMemcpy2DAsync(Yg, (size_t)ldy * sizeof(T), dYg, (size_t)m->nrows * sizeof(T),
                                 (size_t)m->nrows * sizeof(T), (size_t)kk, cudaMemcpyDeviceToHost, m->pipe_out));
    }
    SLC_CUDA(cudaStreamSynchronize(m->pipe_out));
    SLC_CUDA(cudaStreamSynchronize(m->pipe_in));
    SLC_CUDA(cudaStreamSynchronize(s));
    if (rc == SLC_OK) rc = spmm_window_check(fn);
    return rc;
  }
  if (m->ncols)
    SLC_CUDA(cudaMemcpy2DAsync(dX.p, (size_t)m->ncols * sizeof(T), X, (size_t)ldx * sizeof(T),
                               (size_t)m->ncols * sizeof(T), (size_t)k, cudaMemcpyHostToDevice, s));
  if (yi.mode != 0 && m->nrows)
    SLC_CUDA(cudaMemcpy2DAsync(dY.p, (size_t)m->nrows * sizeof(T), Y, (size_t)ldy * sizeof(T),
                               (size_t)m->nrows * sizeof(T), (size_t)k, cudaMemcpyHostToDevice, s));
  SLC_TRY(spmm_dev<T>(m, dX.p, m->ncols, dY.p, m->nrows, k, yi));
  if (m->nrows)
    SLC_CUDA(cudaMemcpy2DAsync(Y, (size_t)ldy * sizeof(T), dY.p, (size_t)m->nrows * sizeof(T),
                               (size_t)m->nrows * sizeof(T), (size_t)k, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  return spmm_window_check(fn);
}

}  // namespace slc

using namespace slc;

extern "C" {

int slc_prepare(slc_matrix *m) {
  if (!m) return fail(SLC_ERR_INVALID, "slc_prepare: null handle");
  HandleLock lock(m);
  if (m->nnz == 0) return SLC_OK;
  return m->elem == ELEM_D ? ensure_plan<double>(m) : ensure_plan<double2>(m);
}

int slc_set_spmv_kernel(slc_matrix *m, int kind) {
  if (!m) return fail(SLC_ERR_INVALID, "slc_set_spmv_kernel: null handle");
  if (kind < SPMV_AUTO || kind > SPMV_CSR_MERGE)
    return fail(SLC_ERR_INVALID, "slc_set_spmv_kernel: unknown kernel %d", kind);
  m->spmv_kind = kind;
  return SLC_OK;
}

int slc_plan_info(const slc_matrix *m, char *buf, int32_t buflen) {
  if (!m || !buf || buflen <= 0) return fail(SLC_ERR_INVALID, "slc_plan_info: bad argument");
  const char *names[] = {"unplanned", "sell32", "csr-warp", "csr-balanced", "csr-merge"};
  const double pad = m->nnz ? 100.0 * (double)(m->sell.padded - m->nnz) / (double)m->nnz : 0.0;
  const double unit = (m->sell.kbase && m->nnz) ? 100.0 * (double)m->sell.unit_entries / (double)m->nnz : 0.0;
  snprintf(buf, (size_t)buflen,
           "%s rows=%d cols=%d nnz=%lld max_rowlen=%d sell_padding=%.2f%% chunks=%d unit_stride=%.2f%% hot_cols=%d brick=%d",
           names[m->spmv_active], m->nrows, m->ncols, (long long)m->nnz, m->max_rowlen, pad, m->bal.nchunks, unit,
           m->bal.hot_k > 0 ? m->bal.hot_k : 0, m->brick.state == 1 ? 1 : 0);
  return SLC_OK;
}

int slc_d_mulv(slc_matrix *a, const double *x, double *y) {
  return spmv_host_entry<double, false>(a, x, y, "mulV");
}
int slc_z_mulv(slc_matrix *a, const double *x, double *y) {
  return spmv_host_entry<double2, false>(a, x, y, "mulV");
}
int slc_d_axpy(slc_matrix *a, const double *x, double *y) {
  return spmv_host_entry<double, true>(a, x, y, "axpy_");
}
int slc_z_axpy(slc_matrix *a, const double *x, double *y) {
  return spmv_host_entry<double2, true>(a, x, y, "axpy_");
}
int slc_d_mulv_dev(slc_matrix *a, const double *x, double *y) {
  return spmv_dev_entry<double, false>(a, x, y, "mulV");
}
int slc_z_mulv_dev(slc_matrix *a, const double *x, double *y) {
  return spmv_dev_entry<double2, false>(a, x, y, "mulV");
}
int slc_d_axpy_dev(slc_matrix *a, const double *x, double *y) {
  return spmv_dev_entry<double, true>(a, x, y, "axpy_");
}
int slc_z_axpy_dev(slc_matrix *a, const double *x, double *y) {
  return spmv_dev_entry<double2, true>(a, x, y, "axpy_");
}

int slc_d_mulv_dist_dev(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *x_parts, double *y_dev, slc_peer_group *sync) {
  return spmv_dist_entry<double>(a, nparts, self, splits, x_parts, y_dev, sync, "mulV");
}
int slc_z_mulv_dist_dev(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *x_parts, double *y_dev, slc_peer_group *sync) {
  return spmv_dist_entry<double2>(a, nparts, self, splits, x_parts, y_dev, sync, "mulV");
}

int slc_d_mulv_allgather_dev(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits,
                             const double *const *x_parts, double *x_full_dev, double *y_dev,
                             slc_peer_group *sync) {
  return spmv_allgather_entry<double>(a, nparts, self, splits, x_parts, x_full_dev, y_dev, sync, "mulV");
}
int slc_z_mulv_allgather_dev(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits,
                             const double *const *x_parts, double *x_full_dev, double *y_dev,
                             slc_peer_group *sync) {
  return spmv_allgather_entry<double2>(a, nparts, self, splits, x_parts, x_full_dev, y_dev, sync, "mulV");
}

int slc_d_mulm_dist_dev(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *X_parts, const int64_t *ldx_parts, double *Y, int64_t ldy, int32_t k,
                        slc_peer_group *sync) {
  return spmm_dist_entry<double>(a, nparts, self, splits, X_parts, ldx_parts, Y, ldy, k, sync, "mulM");
}
int slc_z_mulm_dist_dev(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *X_parts, const int64_t *ldx_parts, double *Y, int64_t ldy, int32_t k,
                        slc_peer_group *sync) {
  return spmm_dist_entry<double2>(a, nparts, self, splits, X_parts, ldx_parts, Y, ldy, k, sync, "mulM");
}
int slc_d_mulm(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k, const double *beta) {
  return spmm_host_entry<double>(a, X, ldx, Y, ldy, k, beta, "mulM");
}
int slc_z_mulm(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k, const double *beta) {
  return spmm_host_entry<double2>(a, X, ldx, Y, ldy, k, beta, "mulM");
}
int slc_d_mulm_dev(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                   const double *beta) {
  return spmm_dev_entry<double>(a, X, ldx, Y, ldy, k, beta, "mulM");
}
int slc_z_mulm_dev(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                   const double *beta) {
  return spmm_dev_entry<double2>(a, X, ldx, Y, ldy, k, beta, "mulM");
}
int slc_d_dist_needs(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits, uint32_t *mask) {
  HandleLock lock(a);
  return dist_needs_entry<double>(a, nparts, self, splits, mask, "slc_d_dist_needs");
}
int slc_z_dist_needs(slc_matrix *a, int32_t nparts, int32_t self, const int32_t *splits, uint32_t *mask) {
  HandleLock lock(a);
  return dist_needs_entry<double2>(a, nparts, self, splits, mask, "slc_z_dist_needs");
}

}  // extern "C"
