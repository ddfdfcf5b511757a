// slc_shardops.cu -- the device-side pieces of the transpose and the triplet assembly of a matrix that is
// spread over several GPUs ("bucket by destination block, one all-to-all-v, local
// sort").  Round 1 cut, pickled, shipped and stacked the pieces in host memory; here everything that
// touches an entry is a kernel and what crosses between ranks is raw device arrays:
//
//   slc_device_arrays        the CSC arrays of a handle as device pointers: a column range of a
//                            matrix is a slice of them, so "cutting a piece" is pointer arithmetic
//   slc_{d,z}_vstack_dev     row blocks of one matrix (device CSC arrays, e.g. the pieces a rank received),
//                            stacked top to bottom into one handle: column j of the result is the
//                            blocks' columns j one after the other (stable: block order is row order)
//   slc_{d,z}_coo_bucket_dev triplets bucketed by the rank that owns their column, input order kept
//                            inside a bucket (one stable radix pass over the owner id), columns rebased
//
// Reference semantics: transpose is stable (Data/Matrix/Sparse.hs:301-329) -- pieces stacked in rank
// order keep A's column order inside every row of A^T; compress sums duplicates in input order
// (oracle convention, Sparse.hs:184-291) -- buckets keep input order and are concatenated in rank order.
#include "slc_common.cuh"

namespace slc {

constexpr int VSTACK_MAX = 16;

struct VstackBlocks {
  const int32_t *ptr[VSTACK_MAX];   // [ncols+1] each
  const int32_t *idx[VSTACK_MAX];
  const void *val[VSTACK_MAX];
  int32_t row0[VSTACK_MAX];         // rows above the block
  int32_t n;
};

__global__ void k_vstack_counts(VstackBlocks b, int32_t ncols, int32_t *__restrict__ cnt) {
  const int32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j > ncols) return;
  int32_t c = 0;
  if (j < ncols)
    for (int q = 0; q < b.n; ++q) c += b.ptr[q][j + 1] - b.ptr[q][j];
  cnt[j] = c;
}

// one warp per output column: the blocks' segments are copied one after the other, lanes striding
template <class T>
__global__ void __launch_bounds__(256)
k_vstack_fill(VstackBlocks b, int32_t ncols, const int32_t *__restrict__ out_ptr, int32_t *__restrict__ out_idx,
              T *__restrict__ out_val) {
  const int32_t j = (int32_t)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (j >= ncols) return;
  int32_t dst = out_ptr[j];
  for (int q = 0; q < b.n; ++q) {
    const int32_t s0 = b.ptr[q][j], s1 = b.ptr[q][j + 1];
    const T *v = static_cast<const T *>(b.val[q]);
    for (int32_t e = s0 + lane; e < s1; e += 32) {
      out_idx[dst + (e - s0)] = b.idx[q][e] + b.row0[q];
      out_val[dst + (e - s0)] = v[e];
    }
    dst += s1 - s0;
  }
}

template <class T>
static int vstack_impl(int32_t nblocks, const int32_t *block_rows, int32_t ncols, const int32_t *const *ptrs,
                       const int32_t *const *idxs, const double *const *vals, slc_matrix **out) {
  if (!out) return fail(SLC_ERR_INVALID, "vstack: null output handle");
  *out = nullptr;
  if (nblocks < 1 || nblocks > VSTACK_MAX || ncols < 0 || !block_rows || !ptrs || !idxs || !vals)
    return fail(SLC_ERR_INVALID, "vstack: 1..%d blocks supported", VSTACK_MAX);
  VstackBlocks b;
  memset(&b, 0, sizeof b);
  b.n = nblocks;
  int64_t rows = 0;
  for (int q = 0; q < nblocks; ++q) {
    if (block_rows[q] < 0 || !ptrs[q]) return fail(SLC_ERR_INVALID, "vstack: bad block %d", q);
    b.ptr[q] = ptrs[q];
    b.idx[q] = idxs[q];
    b.val[q] = vals[q];
    b.row0[q] = (int32_t)rows;
    rows += block_rows[q];
  }
  if (rows > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "vstack: more rows than 32-bit indices hold");
  cudaStream_t s = cur_stream();
  DevBuf<int32_t> cnt;
  SLC_TRY(cnt.alloc((size_t)ncols + 1));
  k_vstack_counts<<<(unsigned)ceil_div((int64_t)ncols + 1, 256), 256, 0, s>>>(b, ncols, cnt.p);
  SLC_LAUNCHED();
  SLC_TRY(exclusive_scan_i32(cnt.p, cnt.p, (int64_t)ncols + 1));
  int32_t nnz = 0;
  SLC_CUDA(cudaMemcpyAsync(&nnz, cnt.p + ncols, sizeof nnz, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  if (nnz < 0) return fail(SLC_ERR_RANGE, "vstack: result does not fit 32-bit pointers");
  slc_matrix *m = nullptr;
  SLC_TRY(matrix_alloc(Num<T>::elem, (int32_t)rows, ncols, nnz, &m));
  SLC_CUDA(cudaMemcpyAsync(m->colptr, cnt.p, ((size_t)ncols + 1) * sizeof(int32_t), cudaMemcpyDeviceToDevice, s));
  if (ncols > 0 && nnz > 0) {
    k_vstack_fill<T><<<(unsigned)ceil_div((int64_t)ncols * 32, 256), 256, 0, s>>>(b, ncols, m->colptr, m->rowidx, (T *)m->vals);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) {
      matrix_release(m);
      return fail(SLC_ERR_CUDA, "vstack: launch failed: %s", cudaGetErrorString(e));
    }
  }
  *out = m;
  return SLC_OK;
}

// ---- triplets -> buckets by owner of the column --------------------------------------------------------

struct OwnerSplits {
  int32_t s[VSTACK_MAX + 1];
  int32_t n;
};

// owner of column c: the p with s[p] <= c < s[p+1]; columns outside [0, s[n]) go to the nearest block,
// whose compress then reports them (Sparse.hs:192-212)
__device__ __forceinline__ int owner_of_col(const OwnerSplits &sp, int32_t c) {
  int p = 0;
  while (p + 1 < sp.n && c >= sp.s[p + 1]) ++p;
  return p;
}

__global__ void k_owner_keys(int64_t n, const int32_t *__restrict__ cols, OwnerSplits sp, uint32_t *__restrict__ keys) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) keys[i] = (uint32_t)owner_of_col(sp, cols[i]);
}

__global__ void k_owner_counts(int64_t n, const uint32_t *__restrict__ sorted, int32_t nparts, int64_t *__restrict__ first) {
  // first[p] = first sorted position whose key is >= p, for p in [0, nparts]
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  const int64_t prev = i == 0 ? -1 : (int64_t)sorted[i - 1];
  const int64_t cur = i == n ? (int64_t)nparts : (int64_t)sorted[i];
  for (int64_t p = prev + 1; p <= cur; ++p) first[p] = i;
}

template <class T>
__global__ void k_bucket_gather(int64_t n, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ owner_sorted,
                                OwnerSplits sp, const int32_t *__restrict__ rows, const int32_t *__restrict__ cols,
                                const T *__restrict__ vals, int32_t *__restrict__ orows, int32_t *__restrict__ ocols,
                                T *__restrict__ ovals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t p = perm[i];
  orows[i] = rows[p];
  ocols[i] = cols[p] - sp.s[owner_sorted[i]];
  ovals[i] = vals[p];
}

template <class T>
static int coo_bucket_impl(int64_t n, const int32_t *rows, const int32_t *cols, const double *vals, int32_t nparts,
                           const int32_t *col_splits, int32_t *orows, int32_t *ocols, double *ovals, int64_t *counts) {
  if (n < 0 || nparts < 1 || nparts > VSTACK_MAX || !col_splits || !counts)
    return fail(SLC_ERR_INVALID, "coo_bucket: bad argument");
  if (n > 0 && (!rows || !cols || !vals || !orows || !ocols || !ovals)) return fail(SLC_ERR_INVALID, "coo_bucket: null array");
  if (n > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "coo_bucket: more triplets than 32-bit positions hold");
  for (int p = 0; p < nparts; ++p) counts[p] = 0;
  if (n == 0) return SLC_OK;
  OwnerSplits sp;
  sp.n = nparts;
  for (int p = 0; p <= VSTACK_MAX; ++p) sp.s[p] = col_splits[p <= nparts ? p : nparts];
  cudaStream_t s = cur_stream();
  DevBuf<uint32_t> ka, kb, va, vb;
  DevBuf<int64_t> first;
  SLC_TRY(ka.alloc((size_t)n));
  SLC_TRY(kb.alloc((size_t)n));
  SLC_TRY(va.alloc((size_t)n));
  SLC_TRY(vb.alloc((size_t)n));
  SLC_TRY(first.alloc((size_t)nparts + 1));
  k_owner_keys<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(n, cols, sp, ka.p);
  SLC_LAUNCHED();
  uint32_t *sk = nullptr, *sv = nullptr;
  // stable: equal owners keep their input order (values = input positions, generated by the first pass)
  SLC_TRY(sort_pairs_u32(nullptr, ka.p, kb.p, va.p, vb.p, n, bits_for((uint32_t)nparts), true, &sk, &sv));
  k_owner_counts<<<(unsigned)ceil_div(n + 1, 256), 256, 0, s>>>(n, sk, nparts, first.p);
  SLC_LAUNCHED();
  k_bucket_gather<T><<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(n, sv, sk, sp, rows, cols, (const T *)vals, orows, ocols,
                                                                 (T *)ovals);
  SLC_LAUNCHED();
  int64_t h[VSTACK_MAX + 1];
  SLC_CUDA(cudaMemcpyAsync(h, first.p, ((size_t)nparts + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  for (int p = 0; p < nparts; ++p) counts[p] = h[p + 1] - h[p];
  return SLC_OK;
}

}  // namespace slc

using namespace slc;

extern "C" {

int slc_device_arrays(const slc_matrix *m, const int32_t **colptr_dev, const int32_t **rowidx_dev, const void **vals_dev) {
  if (!m) return fail(SLC_ERR_INVALID, "slc_device_arrays: null handle");
  if (colptr_dev) *colptr_dev = m->colptr;
  if (rowidx_dev) *rowidx_dev = m->rowidx;
  if (vals_dev) *vals_dev = m->vals;
  return SLC_OK;
}

int slc_d_vstack_dev(int32_t nblocks, const int32_t *block_rows, int32_t ncols, const int32_t *const *colptr_dev,
                     const int32_t *const *rowidx_dev, const double *const *vals_dev, slc_matrix **out) {
  return vstack_impl<double>(nblocks, block_rows, ncols, colptr_dev, rowidx_dev, vals_dev, out);
}
int slc_z_vstack_dev(int32_t nblocks, const int32_t *block_rows, int32_t ncols, const int32_t *const *colptr_dev,
                     const int32_t *const *rowidx_dev, const double *const *vals_dev, slc_matrix **out) {
  return vstack_impl<double2>(nblocks, block_rows, ncols, colptr_dev, rowidx_dev, vals_dev, out);
}

int slc_d_coo_bucket_dev(int64_t n, const int32_t *rows_dev, const int32_t *cols_dev, const double *vals_dev, int32_t nparts,
                         const int32_t *col_splits, int32_t *out_rows_dev, int32_t *out_cols_dev, double *out_vals_dev,
                         int64_t *counts) {
  return coo_bucket_impl<double>(n, rows_dev, cols_dev, vals_dev, nparts, col_splits, out_rows_dev, out_cols_dev,
                                 out_vals_dev, counts);
}
int slc_z_coo_bucket_dev(int64_t n, const int32_t *rows_dev, const int32_t *cols_dev, const double *vals_dev, int32_t nparts,
                         const int32_t *col_splits, int32_t *out_rows_dev, int32_t *out_cols_dev, double *out_vals_dev,
                         int64_t *counts) {
  return coo_bucket_impl<double2>(n, rows_dev, cols_dev, vals_dev, nparts, col_splits, out_rows_dev, out_cols_dev,
                                  out_vals_dev, counts);
}

}  // extern "C"
