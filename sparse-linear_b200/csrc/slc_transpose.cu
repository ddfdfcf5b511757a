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

__global__ void __launch_bounds__(EXPAND_THREADS)
k_expand_cols(int64_t nnz, int32_t ncols, int64_t ntiles, const int32_t *__restrict__ colptr,
              const int32_t *__restrict__ tile_col, uint32_t *__restrict__ colid) {
  __shared__ int32_t win[EXPAND_WINDOW];
  const int32_t c_lo = tile_col[blockIdx.x];
  const int32_t c_hi = (int64_t)blockIdx.x + 1 < ntiles ? tile_col[blockIdx.x + 1] : ncols - 1;
  const int32_t span = c_hi - c_lo + 2;  // pointers colptr[c_lo .. c_hi + 1]
  const bool staged = span <= EXPAND_WINDOW;
  if (staged)
    for (int32_t i = threadIdx.x; i < span; i += EXPAND_THREADS) win[i] = colptr[c_lo + i];
  __syncthreads();
  const int64_t p0 = (int64_t)blockIdx.x * EXPAND_TILE + (int64_t)threadIdx.x * EXPAND_ITEMS;
  if (p0 >= nnz) return;
  uint32_t out[EXPAND_ITEMS];
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

template <class T>
int transpose_csc(int32_t nrows, int32_t ncols, int64_t nnz, const int32_t *colptr,
                  const int32_t *rowidx, const T *vals, bool conj, int32_t *out_ptr, int32_t *out_idx,
                  T *out_vals) {
  cudaStream_t s = cur_stream();
  if (nnz == 0) {
    SLC_CUDA(cudaMemsetAsync(out_ptr, 0, ((size_t)nrows + 1) * sizeof(int32_t), s));
    return SLC_OK;
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
