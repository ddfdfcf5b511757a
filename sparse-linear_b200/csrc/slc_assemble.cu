// slc_assemble.cu -- triplet assembly (compress / fromTriples) and the builders the configs and
// FEAST use (lin / glin, kronecker).
//
// Reference: compress Data/Matrix/Sparse.hs:184-255 with dedupInPlace :257-280 and computePtrs
// :282-291; lin / glin :401-431; kronecker :597-638.
//
// compress on the device: bounds check (first offending position, rows before columns), then ONE
// stable LSD radix sort of the packed key  col * 2^rowbits + row  carrying the input position,
// over just the bits the two ranges need; runs of equal keys are duplicates and are summed left
// to right, i.e. in input order (the reference's unstable introsort leaves that order open; see
// oracle/oracle.h).  Explicit zeros and zero sums are kept.
#include "slc_common.cuh"

namespace slc {

// ---- compress -----------------------------------------------------------------------------------

__global__ void k_coo_check_pack(int64_t n, int32_t nrows, int32_t ncols, int rowbits,
                                 const int32_t *__restrict__ rows, const int32_t *__restrict__ cols,
                                 uint64_t *__restrict__ keys, unsigned long long *__restrict__ first_bad) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t r = rows[i], c = cols[i];
  const bool bad_r = r < 0 || r >= nrows;
  const bool bad_c = c < 0 || c >= ncols;
  if (bad_r) atomicMin(first_bad + 0, (unsigned long long)i);
  if (bad_c) atomicMin(first_bad + 1, (unsigned long long)i);
  keys[i] = (bad_r || bad_c) ? 0ull : (((uint64_t)(uint32_t)c << rowbits) | (uint32_t)r);
}

__global__ void k_head_flags(int64_t n, const uint64_t *__restrict__ keys, int32_t *__restrict__ head) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

// One thread per run of equal keys: xs[w] <- xs[w] + xs[r] for each later duplicate r
// (Sparse.hs:266-277), in sorted (= input) order.
template <class T>
__global__ void k_reduce_runs(int64_t n, int rowbits, const uint64_t *__restrict__ keys,
                              const uint32_t *__restrict__ perm, const int32_t *__restrict__ head,
                              const int32_t *__restrict__ seg, const T *__restrict__ vals,
                              int32_t *__restrict__ out_row, uint32_t *__restrict__ out_col,
                              T *__restrict__ out_val) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !head[i]) return;
  const uint64_t key = keys[i];
  T acc = vals[perm[i]];
  for (int64_t r = i + 1; r < n && keys[r] == key; ++r) acc = Num<T>::add(acc, vals[perm[r]]);
  const int32_t w = seg[i];
  out_row[w] = (int32_t)(key & ((1ull << rowbits) - 1ull));
  out_col[w] = (uint32_t)(key >> rowbits);
  out_val[w] = acc;
}

// colptr[c] = first output position whose column is >= c (same construction as the transpose)
__global__ void k_ptrs_from_sorted_u32(int64_t n, int32_t nkeys, const uint32_t *__restrict__ sorted,
                                       int32_t *__restrict__ ptr) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  int64_t prev = (i == 0) ? -1 : (int64_t)sorted[i - 1];
  int64_t cur = (i == n) ? (int64_t)nkeys : (int64_t)sorted[i];
  for (int64_t r = prev + 1; r <= cur; ++r) ptr[r] = (int32_t)i;
}

template <class T>
static int compress_entry(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows, const int32_t *cols,
                          const double *vals, slc_matrix **out, int64_t *bad_pos, bool on_device) {
  if (!out) return fail(SLC_ERR_INVALID, "compress: null output handle");
  *out = nullptr;
  if (bad_pos) *bad_pos = -1;
  if (nrows < 0 || ncols < 0 || n < 0) return fail(SLC_ERR_INVALID, "compress: negative size");
  if (n > 0 && (!rows || !cols || !vals)) return fail(SLC_ERR_INVALID, "compress: null triplet array");
  if (n > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "compress: more triplets than 32-bit pointers hold");
  cudaStream_t s = cur_stream();
  slc_matrix *m = nullptr;
  if (n == 0) {
    SLC_TRY(matrix_alloc(Num<T>::elem, nrows, ncols, 0, &m));
    cudaMemsetAsync(m->colptr, 0, ((size_t)ncols + 1) * sizeof(int32_t), s);
    *out = m;
    return SLC_OK;
  }
  const int rowbits = bits_for((uint32_t)nrows);
  const int colbits = bits_for((uint32_t)ncols);

  // triplets: lent host arrays are copied to the device first (Foreign.hs:33-37); device arrays
  // (the *_compress_dev entry points) are read in place and left intact
  DevBuf<int32_t> drows, dcols, head, seg;
  DevBuf<T> dvals;
  DevBuf<uint64_t> k0, k1;
  DevBuf<uint32_t> p0, p1;
  DevBuf<unsigned long long> bad;
  const int32_t *rows_d = rows, *cols_d = cols;
  const T *vals_d = reinterpret_cast<const T *>(vals);
  if (!on_device) {
    SLC_TRY(drows.alloc(n));
    SLC_TRY(dcols.alloc(n));
    SLC_TRY(dvals.alloc(n));
    SLC_CUDA(cudaMemcpyAsync(drows.p, rows, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    SLC_CUDA(cudaMemcpyAsync(dcols.p, cols, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    SLC_CUDA(cudaMemcpyAsync(dvals.p, vals, (size_t)n * sizeof(T), cudaMemcpyHostToDevice, s));
    rows_d = drows.p;
    cols_d = dcols.p;
    vals_d = dvals.p;
  }
  SLC_TRY(k0.alloc(n));
  SLC_TRY(k1.alloc(n));
  SLC_TRY(p0.alloc(n));
  SLC_TRY(p1.alloc(n));
  SLC_TRY(bad.alloc(2));
  SLC_CUDA(cudaMemsetAsync(bad.p, 0xff, 2 * sizeof(unsigned long long), s));
  const unsigned nb = (unsigned)ceil_div(n, 256);
  k_coo_check_pack<<<nb, 256, 0, s>>>(n, nrows, ncols, rowbits, rows_d, cols_d, k0.p, bad.p);
  SLC_LAUNCHED();
  unsigned long long hbad[2];
  SLC_CUDA(cudaMemcpyAsync(hbad, bad.p, sizeof hbad, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  if (hbad[0] != ~0ull) {  // rows are checked first (Sparse.hs:200-205)
    if (bad_pos) *bad_pos = (int64_t)hbad[0];
    return fail(SLC_ERR_ROW_BOUNDS, "compress: row index out of bounds (0,%d) at %lld", nrows,
                (long long)hbad[0]);
  }
  if (hbad[1] != ~0ull) {
    if (bad_pos) *bad_pos = (int64_t)hbad[1];
    return fail(SLC_ERR_COL_BOUNDS, "compress: column index out of bounds (0,%d) at %lld", ncols,
                (long long)hbad[1]);
  }
  drows.release();
  dcols.release();

  // the values carried by the sort are the input positions (generated by its first pass)
  uint64_t *skey = nullptr;
  uint32_t *sperm = nullptr;
  SLC_TRY(sort_pairs_u64(nullptr, k0.p, k1.p, p0.p, p1.p, n, rowbits + colbits, true, &skey, &sperm));

  SLC_TRY(head.alloc(n + 1));
  SLC_TRY(seg.alloc(n + 1));
  SLC_CUDA(cudaMemsetAsync(head.p + n, 0, sizeof(int32_t), s));
  k_head_flags<<<nb, 256, 0, s>>>(n, skey, head.p);
  SLC_LAUNCHED();
  SLC_TRY(exclusive_scan_i32(head.p, seg.p, n + 1));
  int32_t nnz_out = 0;
  SLC_CUDA(cudaMemcpyAsync(&nnz_out, seg.p + n, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));

  SLC_TRY(matrix_alloc(Num<T>::elem, nrows, ncols, nnz_out, &m));
  DevBuf<uint32_t> ocol;
  int rc = ocol.alloc((size_t)nnz_out);
  if (rc == SLC_OK) {
    k_reduce_runs<T><<<nb, 256, 0, s>>>(n, rowbits, skey, sperm, head.p, seg.p, vals_d, m->rowidx, ocol.p,
                                        (T *)m->vals);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    k_ptrs_from_sorted_u32<<<(unsigned)ceil_div((int64_t)nnz_out + 1, 256), 256, 0, s>>>(nnz_out, ncols,
                                                                                        ocol.p, m->colptr);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (cudaGetLastError() != cudaSuccess) rc = fail(SLC_ERR_CUDA, "compress: kernel launch failed");
  }
  if (rc != SLC_OK) {
    matrix_release(m);
    return rc;
  }
  *out = m;
  return SLC_OK;
}

// ---- lin: C = alpha A + beta B on the union pattern ---------------------------------------------
// One thread per column merges the two sorted row lists.  For every row of the union
// r = 0; r = r + alpha*a (if in A); r = r + beta*b (if in B)   -- glin, Sparse.hs:401-431.

__global__ void k_lin_count(int32_t ncols, const int32_t *__restrict__ ap, const int32_t *__restrict__ ai,
                            const int32_t *__restrict__ bp, const int32_t *__restrict__ bi,
                            int64_t *__restrict__ cnt) {
  const int32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > ncols) return;
  if (c == ncols) {
    cnt[c] = 0;
    return;
  }
  int32_t p = ap[c], pe = ap[c + 1], q = bp[c], qe = bp[c + 1];
  int64_t n = 0;
  while (p < pe && q < qe) {
    const int32_t ra = ai[p], rb = bi[q];
    p += ra <= rb;
    q += rb <= ra;
    ++n;
  }
  cnt[c] = n + (pe - p) + (qe - q);
}

template <class T>
__global__ void k_lin_fill(int32_t ncols, T alpha, T beta, const int32_t *__restrict__ ap,
                           const int32_t *__restrict__ ai, const T *__restrict__ av,
                           const int32_t *__restrict__ bp, const int32_t *__restrict__ bi,
                           const T *__restrict__ bv, const int32_t *__restrict__ cp,
                           int32_t *__restrict__ ci, T *__restrict__ cv) {
  const int32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncols) return;
  int32_t p = ap[c], pe = ap[c + 1], q = bp[c], qe = bp[c + 1];
  int32_t w = cp[c];
  while (p < pe || q < qe) {
    int32_t row;
    if (q >= qe || (p < pe && ai[p] <= bi[q])) row = ai[p];
    else row = bi[q];
    T r = Num<T>::zero();
    if (p < pe && ai[p] == row) {
      r = Num<T>::add(r, Num<T>::mul(alpha, av[p]));
      ++p;
    }
    if (q < qe && bi[q] == row) {
      r = Num<T>::add(r, Num<T>::mul(beta, bv[q]));
      ++q;
    }
    ci[w] = row;
    cv[w] = r;
    ++w;
  }
}

__global__ void k_narrow64(int64_t n, const int64_t *__restrict__ in, int32_t *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int32_t)in[i];
}

template <class T>
static int lin_entry(const double *alpha, const slc_matrix *a, const double *beta, const slc_matrix *b,
                     slc_matrix **out) {
  if (!alpha || !beta || !a || !b || !out) return fail(SLC_ERR_INVALID, "glin: null argument");
  *out = nullptr;
  if (a->elem != Num<T>::elem || b->elem != Num<T>::elem)
    return fail(SLC_ERR_INVALID, "glin: element type mismatch");
  if (a->nrows != b->nrows) return fail(SLC_ERR_DIM, "glin: row number mismatch");
  if (a->ncols != b->ncols) return fail(SLC_ERR_DIM, "glin: column number mismatch");
  cudaStream_t s = cur_stream();
  const int32_t nc = a->ncols;
  DevBuf<int64_t> cnt, ptr;
  SLC_TRY(cnt.alloc((size_t)nc + 1));
  SLC_TRY(ptr.alloc((size_t)nc + 1));
  const unsigned nb = (unsigned)ceil_div((int64_t)nc + 1, 128);
  k_lin_count<<<nb, 128, 0, s>>>(nc, a->colptr, a->rowidx, b->colptr, b->rowidx, cnt.p);
  SLC_LAUNCHED();
  SLC_TRY(exclusive_scan_i64(cnt.p, ptr.p, (int64_t)nc + 1));
  int64_t nnzc = 0;
  SLC_CUDA(cudaMemcpyAsync(&nnzc, ptr.p + nc, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  if (nnzc > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "glin: result exceeds 32-bit pointers");
  slc_matrix *c = nullptr;
  SLC_TRY(matrix_alloc(a->elem, a->nrows, nc, nnzc, &c));
  k_narrow64<<<nb, 128, 0, s>>>((int64_t)nc + 1, ptr.p, c->colptr);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (nc > 0) {
    const T al = *reinterpret_cast<const T *>(alpha), be = *reinterpret_cast<const T *>(beta);
    k_lin_fill<T><<<(unsigned)ceil_div(nc, 128), 128, 0, s>>>(nc, al, be, a->colptr, a->rowidx,
                                                              (const T *)a->vals, b->colptr, b->rowidx,
                                                              (const T *)b->vals, c->colptr, c->rowidx,
                                                              (T *)c->vals);
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
  if (cudaGetLastError() != cudaSuccess) {
    matrix_release(c);
    return fail(SLC_ERR_CUDA, "glin: kernel launch failed");
  }
  *out = c;
  return SLC_OK;
}

// ---- kronecker ----------------------------------------------------------------------------------
// Column (na, nb) -> na*ncolsB + nb holds, for (ra, a) in A[:, na] and (rb, b) in B[:, nb], both
// in order, row ra*nrowsB + rb with value b*a (Sparse.hs:620-630).

__global__ void k_kron_count(int32_t ac, int32_t bc, const int32_t *__restrict__ ap,
                             const int32_t *__restrict__ bp, int64_t *__restrict__ cnt) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nc = (int64_t)ac * bc;
  if (j > nc) return;
  if (j == nc) {
    cnt[j] = 0;
    return;
  }
  const int32_t na = (int32_t)(j / bc), nb = (int32_t)(j % bc);
  cnt[j] = (int64_t)(ap[na + 1] - ap[na]) * (bp[nb + 1] - bp[nb]);
}

template <class T>
__global__ void k_kron_fill(int32_t ac, int32_t bc, int32_t br, const int32_t *__restrict__ ap,
                            const int32_t *__restrict__ ai, const T *__restrict__ av,
                            const int32_t *__restrict__ bp, const int32_t *__restrict__ bi,
                            const T *__restrict__ bv, const int32_t *__restrict__ cp,
                            int32_t *__restrict__ ci, T *__restrict__ cv) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= (int64_t)ac * bc) return;
  const int32_t na = (int32_t)(j / bc), nb = (int32_t)(j % bc);
  int32_t w = cp[j];
  for (int32_t p = ap[na]; p < ap[na + 1]; ++p) {
    const int32_t ra = ai[p];
    const T a = av[p];
    for (int32_t q = bp[nb]; q < bp[nb + 1]; ++q) {
      ci[w] = ra * br + bi[q];
      cv[w] = Num<T>::mul(bv[q], a);
      ++w;
    }
  }
}

template <class T>
static int kron_entry(const slc_matrix *a, const slc_matrix *b, slc_matrix **out) {
  if (!a || !b || !out) return fail(SLC_ERR_INVALID, "kronecker: null argument");
  *out = nullptr;
  if (a->elem != Num<T>::elem || b->elem != Num<T>::elem)
    return fail(SLC_ERR_INVALID, "kronecker: element type mismatch");
  const int64_t nr = (int64_t)a->nrows * b->nrows, nc = (int64_t)a->ncols * b->ncols;
  const int64_t nz = a->nnz * b->nnz;
  if (nr > 0x7fffffffLL || nc > 0x7ffffffeLL || nz > 0x7fffffffLL)
    return fail(SLC_ERR_RANGE, "kronecker: result exceeds 32-bit indices");
  cudaStream_t s = cur_stream();
  DevBuf<int64_t> cnt, ptr;
  SLC_TRY(cnt.alloc((size_t)nc + 1));
  SLC_TRY(ptr.alloc((size_t)nc + 1));
  const unsigned nb = (unsigned)ceil_div(nc + 1, 256);
  k_kron_count<<<nb, 256, 0, s>>>(a->ncols, b->ncols, a->colptr, b->colptr, cnt.p);
  SLC_LAUNCHED();
  SLC_TRY(exclusive_scan_i64(cnt.p, ptr.p, nc + 1));
  slc_matrix *c = nullptr;
  SLC_TRY(matrix_alloc(a->elem, (int32_t)nr, (int32_t)nc, nz, &c));
  k_narrow64<<<nb, 256, 0, s>>>(nc + 1, ptr.p, c->colptr);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (nc > 0) {
    k_kron_fill<T><<<(unsigned)ceil_div(nc, 256), 256, 0, s>>>(a->ncols, b->ncols, b->nrows, a->colptr,
                                                               a->rowidx, (const T *)a->vals, b->colptr,
                                                               b->rowidx, (const T *)b->vals, c->colptr,
                                                               c->rowidx, (T *)c->vals);
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
  if (cudaGetLastError() != cudaSuccess) {
    matrix_release(c);
    return fail(SLC_ERR_CUDA, "kronecker: kernel launch failed");
  }
  *out = c;
  return SLC_OK;
}

}  // namespace slc

using namespace slc;

extern "C" {
int slc_d_compress(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows, const int32_t *cols,
                   const double *vals, slc_matrix **out, int64_t *bad_pos) {
  return compress_entry<double>(nrows, ncols, n, rows, cols, vals, out, bad_pos, false);
}
int slc_z_compress(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows, const int32_t *cols,
                   const double *vals, slc_matrix **out, int64_t *bad_pos) {
  return compress_entry<double2>(nrows, ncols, n, rows, cols, vals, out, bad_pos, false);
}
int slc_d_compress_dev(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows, const int32_t *cols,
                       const double *vals, slc_matrix **out, int64_t *bad_pos) {
  return compress_entry<double>(nrows, ncols, n, rows, cols, vals, out, bad_pos, true);
}
int slc_z_compress_dev(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows, const int32_t *cols,
                       const double *vals, slc_matrix **out, int64_t *bad_pos) {
  return compress_entry<double2>(nrows, ncols, n, rows, cols, vals, out, bad_pos, true);
}
int slc_d_lin(const double *alpha, const slc_matrix *a, const double *beta, const slc_matrix *b,
              slc_matrix **c) {
  return lin_entry<double>(alpha, a, beta, b, c);
}
int slc_z_lin(const double *alpha, const slc_matrix *a, const double *beta, const slc_matrix *b,
              slc_matrix **c) {
  return lin_entry<double2>(alpha, a, beta, b, c);
}
int slc_d_kronecker(const slc_matrix *a, const slc_matrix *b, slc_matrix **c) {
  return kron_entry<double>(a, b, c);
}
int slc_z_kronecker(const slc_matrix *a, const slc_matrix *b, slc_matrix **c) {
  return kron_entry<double2>(a, b, c);
}
}
