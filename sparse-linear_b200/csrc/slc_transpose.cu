// slc_transpose.cu -- transpose / ctrans / hermitian and the row-major mirror.
// Reference: transpose Data/Matrix/Sparse.hs:301-329 (counting sort over row indices, stable:
// inside each output column the new indices appear in increasing original-column order,
// values moved bit for bit), computePtrs :282-291, ctrans :371-375, hermitian :377-379.
//
// Device formulation: a stable LSD radix sort of (row index -> source position) over only the
// bits a row index needs, reading the row indices in place; the output pointers fall out of the
// last pass of the sort (a change of key between two neighbours is a row boundary: sort_buckets_u32)
// with no histogram, atomics or extra pass over the sorted keys; the column of every stored
// position is expanded once (k_expand_cols) and one gather kernel then moves column ids and values.
// (Round 1 searched colptr per entry inside the gather: 23 dependent loads each.)
#include "slc_common.cuh"

namespace slc {

// column that owns stored position p: the c with colptr[c] <= p < colptr[c+1]
__device__ __forceinline__ int32_t owner_of(const int32_t *__restrict__ ptr, int32_t n, int32_t p) {
  int32_t lo = 0, hi = n;  // first index in [0, n] with ptr[idx] > p
  while (lo < hi) {
    int32_t mid = lo + ((hi - lo) >> 1);
    if (__ldg(ptr + mid) <= p) lo = mid + 1;
    else hi = mid;
  }
  return lo - 1;
}

// colid[p] = the column that owns stored position p, one block per tile of EXPAND_TILE positions.
// k_tile_first_col records for every tile the column that owns its first position (a thread per
// column writes the tiles whose first position falls inside that column), so a block knows its
// column window without searching.  The window of colptr goes to shared memory; a thread then takes
// 16 consecutive positions: one search of the window for the first, a walk along it for the rest
// (empty columns are stepped over).  Round 1 searched all of colptr from global memory per thread,
// 23 dependent L2 loads each -- latency, not bytes, was the limit (ncu: 0.58 ms for 0.86 GB).
constexpr int EXPAND_ITEMS = 16;
constexpr int EXPAND_THREADS = 256;
constexpr int EXPAND_TILE = EXPAND_ITEMS * EXPAND_THREADS;
constexpr int EXPAND_WINDOW = EXPAND_TILE + 2;  // pointers of a window with no empty column, + 1

__global__ void __launch_bounds__(256)
k_tile_first_col(int32_t ncols, const int32_t *__restrict__ colptr, int32_t *__restrict__ tile_col) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncols) return;
  const int64_t lo = colptr[c], hi = colptr[c + 1];
  for (int64_t b = (lo + EXPAND_TILE - 1) / EXPAND_TILE; b * EXPAND_TILE < hi; ++b) tile_col[b] = (int32_t)c;
}

// The columns of positions p0 .. p0 + 15 (p0 = the calling thread's first position in the block's tile),
// as described above.  Every thread of the block must call it (it contains a barrier).
__device__ __forceinline__ void tile_columns(int64_t nnz, int32_t ncols, int64_t ntiles,
                                             const int32_t *__restrict__ colptr,
                                             const int32_t *__restrict__ tile_col, int32_t *win, int64_t p0,
                                             uint32_t (&out)[EXPAND_ITEMS]) {
  const int32_t c_lo = tile_col[blockIdx.x];
  const int32_t c_hi = (int64_t)blockIdx.x + 1 < ntiles ? tile_col[blockIdx.x + 1] : ncols - 1;
  const int32_t span = c_hi - c_lo + 2;  // pointers colptr[c_lo .. c_hi + 1]
  const bool staged = span <= EXPAND_WINDOW;
  if (staged)
    for (int32_t i = threadIdx.x; i < span; i += EXPAND_THREADS) win[i] = colptr[c_lo + i];
  __syncthreads();
  if (p0 >= nnz) return;
  if (staged) {
    int32_t lo = 0, hi = span - 1;  // first index in [0, span - 1] with win[idx] > p0; win[0] <= p0
    while (lo < hi) {
      const int32_t mid = lo + ((hi - lo) >> 1);
      if (win[mid] <= (int32_t)p0) lo = mid + 1;
      else hi = mid;
    }
    int32_t c = lo - 1;  // local column
    int32_t next = win[c + 1];
#pragma unroll
    for (int e = 0; e < EXPAND_ITEMS; ++e) {
      const int64_t p = p0 + e;
      while (p < nnz && next <= p) {  // p < nnz: some column of the window owns p
        ++c;
        next = win[c + 1];
      }
      out[e] = (uint32_t)(c_lo + c);
    }
  } else {  // a window with many empty columns: search colptr itself, between the tile's bounds
    int32_t c = c_lo + owner_of(colptr + c_lo, span - 1, (int32_t)p0);
    int32_t next = __ldg(colptr + c + 1);
#pragma unroll
    for (int e = 0; e < EXPAND_ITEMS; ++e) {
      const int64_t p = p0 + e;
      while (p < nnz && next <= p) {
        ++c;
        next = __ldg(colptr + c + 1);
      }
      out[e] = (uint32_t)c;
    }
  }
}

__global__ void __launch_bounds__(EXPAND_THREADS)
k_expand_cols(int64_t nnz, int32_t ncols, int64_t ntiles, const int32_t *__restrict__ colptr,
              const int32_t *__restrict__ tile_col, uint32_t *__restrict__ colid) {
  __shared__ int32_t win[EXPAND_WINDOW];
  const int64_t p0 = (int64_t)blockIdx.x * EXPAND_TILE + (int64_t)threadIdx.x * EXPAND_ITEMS;
  uint32_t out[EXPAND_ITEMS];
  tile_columns(nnz, ncols, ntiles, colptr, tile_col, win, p0, out);
  if (p0 >= nnz) return;
  if (p0 + EXPAND_ITEMS <= nnz) {
    uint4 *dst = reinterpret_cast<uint4 *>(colid + p0);  // p0 is a multiple of 16: 64-byte aligned
#pragma unroll
    for (int q = 0; q < EXPAND_ITEMS / 4; ++q)
      dst[q] = make_uint4(out[4 * q], out[4 * q + 1], out[4 * q + 2], out[4 * q + 3]);
  } else {
    for (int e = 0; p0 + e < nnz; ++e) colid[p0 + e] = out[e];
  }
}

// out[i] = (column, value) of source position perm[i]; 4 entries per thread, all loads of one level
// issued before the next level.
constexpr int GATHER_ITEMS = 4;
template <class T>
__global__ void __launch_bounds__(256)
k_transpose_gather(int64_t nnz, const uint32_t *__restrict__ colid, const T *__restrict__ vals,
                   const uint32_t *__restrict__ perm, bool conj, int32_t *__restrict__ out_idx,
                   T *__restrict__ out_vals) {
  const int64_t base = (int64_t)blockIdx.x * (256 * GATHER_ITEMS) + threadIdx.x;
  uint32_t p[GATHER_ITEMS];
#pragma unroll
  for (int u = 0; u < GATHER_ITEMS; ++u) {
    const int64_t i = base + u * 256;
    p[u] = i < nnz ? perm[i] : 0u;
  }
  uint32_t c[GATHER_ITEMS];
  T v[GATHER_ITEMS];
#pragma unroll
  for (int u = 0; u < GATHER_ITEMS; ++u) {
    c[u] = __ldg(colid + p[u]);
    v[u] = __ldg(vals + p[u]);
  }
#pragma unroll
  for (int u = 0; u < GATHER_ITEMS; ++u) {
    const int64_t i = base + u * 256;
    if (i < nnz) {
      out_idx[i] = (int32_t)c[u];
      out_vals[i] = conj ? Num<T>::conj(v[u]) : v[u];
    }
  }
}

// ---- bucket path (round 2): the reference's own shape ----------------------------------------------
// Sparse.hs:301-329 counts the entries of every row, prefix-sums the counts and places each entry at
// its row's cursor, walking the columns in order.  On the device the walk is parallel, so the cursor
// order inside a row is whatever the atomics gave; a row's entries are then sorted by column (unique
// inside a row), which restores exactly the stable order.  Taken when no row has more than
// TB_LIMIT entries (stencils, finite elements, banded matrices); the radix sort above stays for rows
// that are longer (power-law matrices).
//   k_row_rank    rank[p] = atomicAdd(cnt[row[p]], 1)            the only pass with atomics; 6 B per entry
//   scan          cnt -> out_ptr (final: a transpose has no duplicates to merge)
//   k_row_place   record {column, value} -> slot out_ptr[row] + rank; columns from the tile's window of
//                 colptr (tile_columns), one 16/32-byte record = one sector per entry
//   k_row_sort    128 rows per block, keys of the block's contiguous span in shared memory, insertion
//                 sort per thread (nearly sorted already: blocks run roughly in column order), coalesced
//                 write of column ids and values
constexpr int TB_LIMIT = 512;
constexpr int TB_ROWS = 128;
constexpr int TB_CAP = 4096;
constexpr int RANK_ITEMS = 4;

__global__ void __launch_bounds__(256)
k_row_rank(int64_t nnz, const int32_t *__restrict__ rowidx, int32_t *__restrict__ cnt, uint16_t *__restrict__ rank) {
  const int64_t base = (int64_t)blockIdx.x * (256 * RANK_ITEMS) + threadIdx.x;
  int32_t r[RANK_ITEMS];
#pragma unroll
  for (int u = 0; u < RANK_ITEMS; ++u) {
    const int64_t p = base + u * 256;
    r[u] = p < nnz ? rowidx[p] : -1;
  }
  int32_t k[RANK_ITEMS];
#pragma unroll
  for (int u = 0; u < RANK_ITEMS; ++u) k[u] = r[u] >= 0 ? atomicAdd(cnt + r[u], 1) : 0;
#pragma unroll
  for (int u = 0; u < RANK_ITEMS; ++u) {
    const int64_t p = base + u * 256;
    if (p < nnz) rank[p] = (uint16_t)min(k[u], 65535);  // 65535: a row far over TB_LIMIT, path not taken
  }
}

template <class T>
__global__ void __launch_bounds__(EXPAND_THREADS)
k_row_place(int64_t nnz, int32_t ncols, int64_t ntiles, const int32_t *__restrict__ colptr,
            const int32_t *__restrict__ tile_col, const int32_t *__restrict__ rowidx, const T *__restrict__ vals,
            bool conj, const int32_t *__restrict__ rowstart, const uint16_t *__restrict__ rank,
            AsmEnt<T> *__restrict__ ent) {
  __shared__ int32_t win[EXPAND_WINDOW];
  __shared__ uint32_t scol[EXPAND_TILE];
  const int64_t tile0 = (int64_t)blockIdx.x * EXPAND_TILE;
  {
    uint32_t out[EXPAND_ITEMS];
    tile_columns(nnz, ncols, ntiles, colptr, tile_col, win, tile0 + (int64_t)threadIdx.x * EXPAND_ITEMS, out);
    if (tile0 + (int64_t)threadIdx.x * EXPAND_ITEMS < nnz) {
#pragma unroll
      for (int e = 0; e < EXPAND_ITEMS; ++e) scol[threadIdx.x * EXPAND_ITEMS + e] = out[e];
    }
  }
  __syncthreads();
  // lanes now take consecutive positions: coalesced reads of rowidx, rank and vals
#pragma unroll 4
  for (int e = 0; e < EXPAND_ITEMS; ++e) {
    const int32_t l = e * EXPAND_THREADS + threadIdx.x;
    const int64_t p = tile0 + l;
    if (p < nnz) {
      const int32_t slot = __ldg(rowstart + rowidx[p]) + rank[p];
      AsmEnt<T> rec;
      rec.key = scol[l];
      const T v = vals[p];
      rec.val = conj ? Num<T>::conj(v) : v;
      ent[slot] = rec;
    }
  }
}

template <class T>
__global__ void __launch_bounds__(TB_ROWS)
k_row_sort(int32_t nrows, const int32_t *__restrict__ start, const AsmEnt<T> *__restrict__ ent,
           int32_t *__restrict__ out_idx, T *__restrict__ out_vals) {
  __shared__ uint32_t sk[TB_CAP];
  __shared__ uint16_t si[TB_CAP];
  const int t = threadIdx.x;
  const int32_t rend = min((int32_t)(blockIdx.x + 1) * TB_ROWS, nrows);
  int32_t r0 = blockIdx.x * TB_ROWS;
  while (r0 < rend) {
    const int32_t s0 = start[r0];
    const int32_t r = r0 + t;
    const int32_t e_t = r < rend ? start[r + 1] : 0;
    const int ncover = __syncthreads_count(r < rend && e_t - s0 <= TB_CAP);  // a prefix; >= 1 (TB_LIMIT <= TB_CAP)
    const int32_t len = start[r0 + ncover] - s0;
    for (int32_t i = t; i < len; i += TB_ROWS) {
      sk[i] = (uint32_t)ent[s0 + i].key;
      si[i] = (uint16_t)i;
    }
    __syncthreads();
    if (t < ncover) {
      const int32_t a = start[r] - s0, b = e_t - s0;
      for (int32_t i = a + 1; i < b; ++i) {
        const uint32_t key = sk[i];
        if (sk[i - 1] <= key) continue;
        const uint16_t idx = si[i];
        int32_t j = i;
        while (j > a && sk[j - 1] > key) {
          sk[j] = sk[j - 1];
          si[j] = si[j - 1];
          --j;
        }
        sk[j] = key;
        si[j] = idx;
      }
    }
    __syncthreads();
    for (int32_t i = t; i < len; i += TB_ROWS) {
      out_idx[s0 + i] = (int32_t)sk[i];
      out_vals[s0 + i] = ent[s0 + si[i]].val;
    }
    r0 += ncover;
    __syncthreads();  // sk / si are reused by the next pass
  }
}

// returns SLC_OK with *done = true when the bucket path produced the transpose, *done = false when a row is
// too long for it (nothing written)
template <class T>
static int transpose_buckets(int32_t nrows, int32_t ncols, int64_t nnz, const int32_t *colptr,
                             const int32_t *rowidx, const T *vals, bool conj, int32_t *out_ptr, int32_t *out_idx,
                             T *out_vals, bool *done) {
  cudaStream_t s = cur_stream();
  *done = false;
  DevBuf<int32_t> cnt;
  DevBuf<uint16_t> rank;
  DevBuf<unsigned long long> longest;
  SLC_TRY(cnt.alloc((size_t)nrows + 1));
  SLC_TRY(rank.alloc((size_t)nnz));
  SLC_TRY(longest.alloc(1));
  SLC_CUDA(cudaMemsetAsync(cnt.p, 0, ((size_t)nrows + 1) * sizeof(int32_t), s));
  SLC_CUDA(cudaMemsetAsync(longest.p, 0, sizeof(unsigned long long), s));
  k_row_rank<<<(unsigned)ceil_div(nnz, 256 * RANK_ITEMS), 256, 0, s>>>(nnz, rowidx, cnt.p, rank.p);
  SLC_LAUNCHED();
  SLC_TRY(max_i32_into(cnt.p, nrows, longest.p));
  unsigned long long hmax = 0;
  SLC_CUDA(cudaMemcpyAsync(&hmax, longest.p, sizeof hmax, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  if (hmax > (unsigned long long)TB_LIMIT) return SLC_OK;
  SLC_TRY(exclusive_scan_i32(cnt.p, out_ptr, (int64_t)nrows + 1));
  cnt.release();
  DevBuf<AsmEnt<T>> ent;
  SLC_TRY(ent.alloc((size_t)nnz));
  const int64_t ntiles = ceil_div(nnz, EXPAND_TILE);
  DevBuf<int32_t> tile_col;
  SLC_TRY(tile_col.alloc((size_t)ntiles));
  k_tile_first_col<<<(unsigned)ceil_div(ncols, 256), 256, 0, s>>>(ncols, colptr, tile_col.p);
  SLC_LAUNCHED();
  k_row_place<T><<<(unsigned)ntiles, EXPAND_THREADS, 0, s>>>(nnz, ncols, ntiles, colptr, tile_col.p, rowidx, vals,
                                                            conj, out_ptr, rank.p, ent.p);
  SLC_LAUNCHED();
  k_row_sort<T><<<(unsigned)ceil_div((int64_t)nrows, TB_ROWS), TB_ROWS, 0, s>>>(nrows, out_ptr, ent.p, out_idx,
                                                                                out_vals);
  SLC_LAUNCHED();
  *done = true;
  return SLC_OK;
}

template <class T>
int transpose_csc(int32_t nrows, int32_t ncols, int64_t nnz, const int32_t *colptr,
                  const int32_t *rowidx, const T *vals, bool conj, int32_t *out_ptr, int32_t *out_idx,
                  T *out_vals) {
  cudaStream_t s = cur_stream();
  if (nnz == 0) {
    SLC_CUDA(cudaMemsetAsync(out_ptr, 0, ((size_t)nrows + 1) * sizeof(int32_t), s));
    return SLC_OK;
  }
  // SLC_TR_BUCKET (0): 1 = count / place / sort-per-row when every row is short.  Measured on B200 (round 2,
  // profiles/bench/r2j_cfg*t_*.json): cfg 3 0.445 ms against 0.497 ms for the radix sort, but cfg 5 6.30
  // against 5.87 ms (a record of 16 bytes moved three times is no less traffic than 8-byte (key, position)
  // pairs moved three times plus one gather) and cfg 4 pays 3.6 ms of contended atomics before it finds out
  // that its rows are too long -- so the radix sort stays the default and this path is opt-in.
  if (env_int("SLC_TR_BUCKET", 0) != 0 && nrows > 0) {
    bool done = false;
    SLC_TRY(transpose_buckets<T>(nrows, ncols, nnz, colptr, rowidx, vals, conj, out_ptr, out_idx, out_vals, &done));
    if (done) return SLC_OK;
  }
  DevBuf<uint32_t> k0, k1, p0, p1;
  SLC_TRY(k0.alloc(nnz));
  SLC_TRY(k1.alloc(nnz));
  SLC_TRY(p0.alloc(nnz));
  SLC_TRY(p1.alloc(nnz));
  // keys = the row indices, read in place by the first pass; values = source positions
  uint32_t *skey = nullptr, *sperm = nullptr;
  SLC_TRY(sort_buckets_u32(reinterpret_cast<const uint32_t *>(rowidx), k0.p, k1.p, p0.p, p1.p, nnz, nrows, &skey,
                           &sperm, out_ptr));
  uint32_t *colid = (skey == k0.p) ? k1.p : k0.p;  // the key buffer the sort left free
  const int64_t ntiles = ceil_div(nnz, EXPAND_TILE);
  DevBuf<int32_t> tile_col;
  SLC_TRY(tile_col.alloc((size_t)ntiles));
  k_tile_first_col<<<(unsigned)ceil_div(ncols, 256), 256, 0, s>>>(ncols, colptr, tile_col.p);
  SLC_LAUNCHED();
  k_expand_cols<<<(unsigned)ntiles, EXPAND_THREADS, 0, s>>>(nnz, ncols, ntiles, colptr, tile_col.p, colid);
  SLC_LAUNCHED();
  k_transpose_gather<T><<<(unsigned)ceil_div(nnz, 256 * GATHER_ITEMS), 256, 0, s>>>(nnz, colid, vals, sperm, conj,
                                                                                  out_idx, out_vals);
  SLC_LAUNCHED();
  return SLC_OK;
}

template int transpose_csc<double>(int32_t, int32_t, int64_t, const int32_t *, const int32_t *,
                                   const double *, bool, int32_t *, int32_t *, double *);
template int transpose_csc<double2>(int32_t, int32_t, int64_t, const int32_t *, const int32_t *,
                                    const double2 *, bool, int32_t *, int32_t *, double2 *);

int ensure_csr(slc_matrix *m) {
  if (m->has_csr) return SLC_OK;
  DevBuf<int32_t> rp, ci;
  DevBuf<char> rv;
  SLC_TRY(rp.alloc((size_t)m->nrows + 1));
  SLC_TRY(ci.alloc((size_t)m->nnz));
  SLC_TRY(rv.alloc((size_t)m->nnz * elem_size(m->elem)));
  if (m->elem == ELEM_D)
    SLC_TRY(transpose_csc<double>(m->nrows, m->ncols, m->nnz, m->colptr, m->rowidx,
                                  (const double *)m->vals, false, rp.p, ci.p, (double *)rv.p));
  else
    SLC_TRY(transpose_csc<double2>(m->nrows, m->ncols, m->nnz, m->colptr, m->rowidx,
                                   (const double2 *)m->vals, false, rp.p, ci.p, (double2 *)rv.p));
  m->rowptr = rp.take();
  m->colidx = ci.take();
  m->rvals = rv.take();
  m->has_csr = true;
  return SLC_OK;
}

// ---- hermitian: ctrans m == m on every field ----------------------------------------------------

template <class T>
__global__ void k_compare(int64_t nnz, int32_t nptr, const int32_t *__restrict__ pa,
                          const int32_t *__restrict__ pb, const int32_t *__restrict__ ia,
                          const int32_t *__restrict__ ib, const T *__restrict__ va,
                          const T *__restrict__ vb, int *__restrict__ differ) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool bad = false;
  if (i < nptr) bad |= pa[i] != pb[i];
  if (i < nnz) bad |= (ia[i] != ib[i]) || !Num<T>::eq(va[i], vb[i]);
  if (bad) atomicOr(differ, 1);
}

template <class T>
static int transpose_entry(const slc_matrix *a, int conj, slc_matrix **at) {
  if (!a || !at) return fail(SLC_ERR_INVALID, "transpose: null argument");
  if (a->elem != Num<T>::elem) return fail(SLC_ERR_INVALID, "transpose: element type mismatch");
  *at = nullptr;
  slc_matrix *t = nullptr;
  SLC_TRY(matrix_alloc(a->elem, a->ncols, a->nrows, a->nnz, &t));
  int rc = transpose_csc<T>(a->nrows, a->ncols, a->nnz, a->colptr, a->rowidx, (const T *)a->vals,
                            conj != 0, t->colptr, t->rowidx, (T *)t->vals);
  if (rc != SLC_OK) {
    matrix_release(t);
    return rc;
  }
  *at = t;
  return SLC_OK;
}

template <class T>
static int hermitian_entry(const slc_matrix *a, int *result) {
  if (!a || !result) return fail(SLC_ERR_INVALID, "hermitian: null argument");
  if (a->elem != Num<T>::elem) return fail(SLC_ERR_INVALID, "hermitian: element type mismatch");
  *result = 0;
  if (a->nrows != a->ncols) return SLC_OK;  // dimensions are fields of the record too
  slc_matrix *t = nullptr;
  SLC_TRY(transpose_entry<T>(a, 1, &t));
  DevBuf<int> differ;
  int rc = differ.alloc(1);
  int h = 1;
  cudaStream_t s = cur_stream();
  if (rc == SLC_OK) {
    cudaMemsetAsync(differ.p, 0, sizeof(int), s);
    int64_t n = a->nnz > (int64_t)a->ncols + 1 ? a->nnz : (int64_t)a->ncols + 1;
    k_compare<T><<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(a->nnz, a->ncols + 1, a->colptr, t->colptr,
                                                            a->rowidx, t->rowidx, (const T *)a->vals,
                                                            (const T *)t->vals, differ.p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, differ.p, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) rc = fail(SLC_ERR_CUDA, "hermitian: CUDA error %s", cudaGetErrorString(e));
  }
  matrix_release(t);
  if (rc == SLC_OK) *result = h ? 0 : 1;
  return rc;
}

}  // namespace slc

using namespace slc;

extern "C" {
int slc_d_transpose(const slc_matrix *a, int conj, slc_matrix **at) {
  return transpose_entry<double>(a, conj, at);
}
int slc_z_transpose(const slc_matrix *a, int conj, slc_matrix **at) {
  return transpose_entry<double2>(a, conj, at);
}
int slc_d_hermitian(const slc_matrix *a, int *result) { return hermitian_entry<double>(a, result); }
int slc_z_hermitian(const slc_matrix *a, int *result) { return hermitian_entry<double2>(a, result); }
}
