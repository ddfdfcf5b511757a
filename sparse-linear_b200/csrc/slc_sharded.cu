// slc_sharded.cu -- the multi-GPU products for a caller that is ONE host process.
//
// The reference's host is a single Haskell process (FEAST even serialises its native calls behind one
// lock, feast/src/Numeric/LinearAlgebra/Feast.hs:134), so the row-partitioned products must be
// reachable from one `foreign import ccall` without a process per GPU.  slc_{d,z}_upload_sharded cuts
// a host CSC matrix into contiguous row blocks, one per device; slc_{d,z}_mulv_sharded /
// slc_{d,z}_mulm_sharded are mulV (Data/Matrix/Sparse.hs:464-471) and multiplyWork (Feast.hs:203-208)
// over those blocks.  The data path is the one the process-per-GPU entry points use
// (slc_*_mulv_dist_dev, slc_*_mulm_dist_dev, slc_*_mulv_allgather_dev): halo entries are loaded from the
// peers' HBM over NVLink inside the product kernel, ordering is the peer-memory flag protocol of
// slc_peer.cu -- here over cudaDeviceEnablePeerAccess mappings instead of CUDA IPC -- and the operand
// slices are double-buffered so that the upload of product e+1 can never overtake a peer still
// reading product e.
//
// How the blocks are cut: the whole matrix goes to the first device once, its row-major mirror is
// built there (the transpose kernels), rows [r_p, r_p+1) of the mirror are one contiguous slice, which
// is copied device-to-device to GPU p and becomes that block's mirror; the block's canonical CSC is the
// transpose of the slice, computed on GPU p.  The host does no per-entry work.
#include <algorithm>
#include <mutex>

#include "slc_common.cuh"
#include "slc_peer.cuh"

struct slc_sharded {
  std::mutex mu;
  int elem = slc::ELEM_D;
  int32_t ndev = 0;
  int devices[slc::MAX_PEERS] = {};
  int32_t nrows = 0, ncols = 0;
  int64_t nnz = 0;
  int32_t row_splits[slc::MAX_PEERS + 1] = {};
  int32_t col_splits[slc::MAX_PEERS + 1] = {};
  slc_matrix *blocks[slc::MAX_PEERS] = {};
  slc_peer_group *groups[slc::MAX_PEERS] = {};
  cudaStream_t streams[slc::MAX_PEERS] = {};
  // operand slices (two buffers per device, alternated product by product), results, and for
  // irregular matrices a full-length copy of the operand per device (all-gather path)
  void *x[2][slc::MAX_PEERS] = {};
  void *y[slc::MAX_PEERS] = {};
  void *xfull[slc::MAX_PEERS] = {};
  int64_t cap_cols = 0;  // block columns the buffers hold (1 for vectors)
  int parity = 0;
  bool allgather = false;  // some block has no sliced-ELL plan: gather the whole operand first
};

namespace slc {

__global__ void k_rebase_ptr(int32_t n, int32_t base, int32_t *__restrict__ p) {
  const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] -= base;
}

// Run `body` with device `dev` current and `stream` as the library's stream; restore both after.
struct DeviceScope {
  int prev_dev = -1;
  cudaStream_t prev_stream;
  DeviceScope(int dev, cudaStream_t stream) : prev_stream(tls().stream) {
    cudaGetDevice(&prev_dev);
    if (prev_dev != dev) cudaSetDevice(dev);
    tls().stream = stream;
  }
  ~DeviceScope() {
    tls().stream = prev_stream;
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev_dev >= 0 && cur != prev_dev) cudaSetDevice(prev_dev);
  }
};

static void sharded_release(slc_sharded *h) {
  if (!h) return;
  for (int p = 0; p < h->ndev; ++p) {
    DeviceScope scope(h->devices[p], h->streams[p]);
    if (h->streams[p]) cudaStreamSynchronize(h->streams[p]);
    cudaFree(h->x[0][p]);
    cudaFree(h->x[1][p]);
    cudaFree(h->y[p]);
    cudaFree(h->xfull[p]);
    if (h->blocks[p]) matrix_release(h->blocks[p]);
  }
  for (int p = 0; p < h->ndev; ++p) slc_peer_group_free(h->groups[p]);
  for (int p = 0; p < h->ndev; ++p)
    if (h->streams[p]) {
      DeviceScope scope(h->devices[p], nullptr);
      cudaStreamDestroy(h->streams[p]);
    }
  delete h;
}

// contiguous row blocks with (almost) equal numbers of stored entries, bounds on multiples of 32 rows
static void balanced_row_splits(const std::vector<int32_t> &rowptr, int32_t nrows, int32_t nparts, int32_t *splits) {
  const int64_t total = nrows > 0 ? rowptr[(size_t)nrows] : 0;
  splits[0] = 0;
  for (int p = 1; p < nparts; ++p) {
    const int64_t target = total * p / nparts;
    int32_t r = (int32_t)(std::lower_bound(rowptr.begin(), rowptr.begin() + nrows + 1, (int32_t)std::min<int64_t>(target, 0x7fffffff)) - rowptr.begin());
    r = std::min(nrows, (r + 16) / 32 * 32);
    splits[p] = std::max(r, splits[p - 1]);
  }
  splits[nparts] = nrows;
}

template <class T>
static int upload_sharded_impl(int32_t ndev, const int32_t *devices, int32_t nrows, int32_t ncols,
                               const int32_t *colptr, const int32_t *rowidx, const double *vals, slc_sharded **out) {
  if (!out) return fail(SLC_ERR_INVALID, "upload_sharded: null output handle");
  *out = nullptr;
  if (ndev < 1 || ndev > MAX_PEERS) return fail(SLC_ERR_INVALID, "upload_sharded: 1..%d devices supported", MAX_PEERS);
  int have = 0;
  SLC_CUDA(cudaGetDeviceCount(&have));
  slc_sharded *h = new (std::nothrow) slc_sharded();
  if (!h) return fail(SLC_ERR_NOMEM, "out of host memory");
  h->elem = Num<T>::elem;
  h->nrows = nrows;
  h->ncols = ncols;
  for (int p = 0; p < ndev; ++p) {
    h->devices[p] = devices ? devices[p] : p;
    if (h->devices[p] < 0 || h->devices[p] >= have) {
      delete h;
      return fail(SLC_ERR_INVALID, "upload_sharded: no device %d (%d visible)", devices ? devices[p] : p, have);
    }
    for (int q = 0; q < p; ++q)
      if (h->devices[q] == h->devices[p]) {
        delete h;
        return fail(SLC_ERR_INVALID, "upload_sharded: device %d listed twice", h->devices[p]);
      }
  }
  int rc = SLC_OK;
  // streams first (sharded_release needs ndev to be the number of initialised slots)
  for (int p = 0; p < ndev && rc == SLC_OK; ++p) {
    DeviceScope scope(h->devices[p], nullptr);
    if (cudaStreamCreateWithFlags(&h->streams[p], cudaStreamNonBlocking) != cudaSuccess) {
      (void)cudaGetLastError();
      rc = fail(SLC_ERR_CUDA, "upload_sharded: cannot create a stream on device %d", h->devices[p]);
    }
    h->ndev = p + 1;
  }
  if (rc == SLC_OK) rc = slc_peer_group_create_local(ndev, h->devices, h->groups);

  // the whole matrix and its row-major mirror on the first device
  slc_matrix *full = nullptr;
  std::vector<int32_t> rowptr;
  if (rc == SLC_OK) {
    DeviceScope scope(h->devices[0], h->streams[0]);
    rc = Num<T>::elem == ELEM_D ? slc_d_upload(nrows, ncols, colptr, rowidx, vals, &full)
                                : slc_z_upload(nrows, ncols, colptr, rowidx, vals, &full);
    if (rc == SLC_OK) rc = ensure_csr(full);
    if (rc == SLC_OK) {
      h->nnz = full->nnz;
      rowptr.resize((size_t)nrows + 1);
      cudaError_t e = cudaMemcpyAsync(rowptr.data(), full->rowptr, rowptr.size() * sizeof(int32_t),
                                      cudaMemcpyDeviceToHost, h->streams[0]);
      if (e == cudaSuccess) e = cudaStreamSynchronize(h->streams[0]);
      if (e != cudaSuccess) rc = fail(SLC_ERR_CUDA, "upload_sharded: %s", cudaGetErrorString(e));
    }
  }
  if (rc == SLC_OK) {
    balanced_row_splits(rowptr, nrows, ndev, h->row_splits);
    if (nrows == ncols) {
      memcpy(h->col_splits, h->row_splits, sizeof h->col_splits);  // x partitioned like the rows
    } else {
      for (int p = 0; p <= ndev; ++p)
        h->col_splits[p] = p == ndev ? ncols : (int32_t)std::min<int64_t>(ncols, ((int64_t)ncols * p / ndev + 16) / 32 * 32);
    }
  }
  // block p: its slice of the mirror travels device to device, the canonical CSC is rebuilt there
  for (int p = 0; p < ndev && rc == SLC_OK; ++p) {
    const int32_t r0 = h->row_splits[p], r1 = h->row_splits[p + 1];
    const int64_t e0 = rowptr[(size_t)r0], e1 = rowptr[(size_t)r1];
    DeviceScope scope(h->devices[p], h->streams[p]);
    slc_matrix *blk = nullptr;
    rc = matrix_alloc(Num<T>::elem, r1 - r0, ncols, e1 - e0, &blk);
    if (rc != SLC_OK) break;
    h->blocks[p] = blk;
    DevBuf<int32_t> rp, ci;
    DevBuf<T> rv;
    rc = rp.alloc((size_t)(r1 - r0) + 1);
    if (rc == SLC_OK) rc = ci.alloc((size_t)(e1 - e0));
    if (rc == SLC_OK) rc = rv.alloc((size_t)(e1 - e0));
    if (rc != SLC_OK) break;
    cudaStream_t s = h->streams[p];
    cudaError_t e = cudaMemcpyPeerAsync(rp.p, h->devices[p], full->rowptr + r0, h->devices[0],
                                        ((size_t)(r1 - r0) + 1) * sizeof(int32_t), s);
    if (e == cudaSuccess && e1 > e0)
      e = cudaMemcpyPeerAsync(ci.p, h->devices[p], full->colidx + e0, h->devices[0], (size_t)(e1 - e0) * sizeof(int32_t), s);
    if (e == cudaSuccess && e1 > e0)
      e = cudaMemcpyPeerAsync(rv.p, h->devices[p], (const T *)full->rvals + e0, h->devices[0], (size_t)(e1 - e0) * sizeof(T), s);
    if (e != cudaSuccess) { rc = fail(SLC_ERR_CUDA, "upload_sharded: peer copy failed: %s", cudaGetErrorString(e)); break; }
    k_rebase_ptr<<<(unsigned)ceil_div((int64_t)(r1 - r0) + 1, 256), 256, 0, s>>>(r1 - r0 + 1, (int32_t)e0, rp.p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    // the slice is the CSC of the block's transpose (ncols x (r1-r0)); transposing it gives the block
    rc = transpose_csc<T>(ncols, r1 - r0, e1 - e0, rp.p, ci.p, rv.p, false, blk->colptr, blk->rowidx, (T *)blk->vals);
    if (rc != SLC_OK) break;
    blk->rowptr = rp.take();
    blk->colidx = ci.take();
    blk->rvals = rv.take();
    blk->has_csr = true;
    rc = slc_prepare(blk);
    if (rc == SLC_OK && cudaStreamSynchronize(s) != cudaSuccess) rc = fail(SLC_ERR_CUDA, "upload_sharded: device %d failed", h->devices[p]);
  }
  if (full) {
    DeviceScope scope(h->devices[0], h->streams[0]);
    matrix_release(full);
  }
  // dependency sets: need(p) from the plans, symmetrised
  if (rc == SLC_OK) {
    uint32_t need[MAX_PEERS] = {};
    for (int p = 0; p < ndev && rc == SLC_OK; ++p) {
      DeviceScope scope(h->devices[p], h->streams[p]);
      rc = Num<T>::elem == ELEM_D ? slc_d_dist_needs(h->blocks[p], ndev, p, h->col_splits, &need[p])
                                  : slc_z_dist_needs(h->blocks[p], ndev, p, h->col_splits, &need[p]);
      if (h->blocks[p]->nnz > 0 && h->blocks[p]->spmv_active != SPMV_SELL) h->allgather = true;
      if (rc == SLC_OK && cudaStreamSynchronize(h->streams[p]) != cudaSuccess) rc = fail(SLC_ERR_CUDA, "upload_sharded: device %d failed", h->devices[p]);
    }
    for (int p = 0; p < ndev && rc == SLC_OK; ++p) {
      uint32_t deps = h->allgather ? 0xffffffffu : need[p];
      for (int q = 0; q < ndev; ++q)
        if ((need[q] >> p) & 1u) deps |= 1u << q;
      rc = slc_peer_group_set_deps(h->groups[p], deps);
    }
  }
  if (rc != SLC_OK) {
    sharded_release(h);
    return rc;
  }
  *out = h;
  return SLC_OK;
}

// device buffers for block products of up to k columns
static int sharded_reserve(slc_sharded *h, int64_t k) {
  if (k <= h->cap_cols) return SLC_OK;
  const size_t el = elem_size(h->elem);
  // The buffers are freed and reallocated one by one: until the last allocation has succeeded the handle
  // holds none (a failure part-way leaves null pointers, and a later product must come back here to
  // allocate again instead of trusting the old capacity).
  h->cap_cols = 0;
  for (int p = 0; p < h->ndev; ++p) {
    DeviceScope scope(h->devices[p], h->streams[p]);
    SLC_CUDA(cudaStreamSynchronize(h->streams[p]));
    const size_t xn = (size_t)std::max<int64_t>(1, (int64_t)(h->col_splits[p + 1] - h->col_splits[p]) * k);
    const size_t yn = (size_t)std::max<int64_t>(1, (int64_t)(h->row_splits[p + 1] - h->row_splits[p]) * k);
    for (int b = 0; b < 2; ++b) {
      cudaFree(h->x[b][p]);
      h->x[b][p] = nullptr;
      SLC_CUDA(cudaMalloc(&h->x[b][p], xn * el));
    }
    cudaFree(h->y[p]);
    h->y[p] = nullptr;
    SLC_CUDA(cudaMalloc(&h->y[p], yn * el));
    if (h->allgather && !h->xfull[p]) SLC_CUDA(cudaMalloc(&h->xfull[p], (size_t)std::max(1, h->ncols) * el));
  }
  h->cap_cols = k;
  return SLC_OK;
}

template <class T>
static int sharded_check(slc_sharded *h, const char *fn) {
  if (!h) return fail(SLC_ERR_INVALID, "%s: null handle", fn);
  if (h->elem != Num<T>::elem) return fail(SLC_ERR_INVALID, "%s: element type mismatch", fn);
  return SLC_OK;
}

// enqueue the product of every block on its device's stream (operand = buffers of the current parity)
template <class T>
static int sharded_products(slc_sharded *h, int32_t k) {
  const double *parts[MAX_PEERS];
  int64_t lds[MAX_PEERS];
  for (int p = 0; p < h->ndev; ++p) {
    parts[p] = (const double *)h->x[h->parity][p];
    lds[p] = std::max<int64_t>(1, h->col_splits[p + 1] - h->col_splits[p]);
  }
  constexpr bool D = Num<T>::elem == ELEM_D;
  for (int p = 0; p < h->ndev; ++p) {
    DeviceScope scope(h->devices[p], h->streams[p]);
    int rc;
    if (k < 0) {
      if (h->allgather)
        rc = D ? slc_d_mulv_allgather_dev(h->blocks[p], h->ndev, p, h->col_splits, parts, (double *)h->xfull[p], (double *)h->y[p], h->groups[p])
               : slc_z_mulv_allgather_dev(h->blocks[p], h->ndev, p, h->col_splits, parts, (double *)h->xfull[p], (double *)h->y[p], h->groups[p]);
      else
        rc = D ? slc_d_mulv_dist_dev(h->blocks[p], h->ndev, p, h->col_splits, parts, (double *)h->y[p], h->groups[p])
               : slc_z_mulv_dist_dev(h->blocks[p], h->ndev, p, h->col_splits, parts, (double *)h->y[p], h->groups[p]);
    } else {
      const int64_t ldy = std::max<int64_t>(1, h->row_splits[p + 1] - h->row_splits[p]);
      rc = D ? slc_d_mulm_dist_dev(h->blocks[p], h->ndev, p, h->col_splits, parts, lds, (double *)h->y[p], ldy, k, h->groups[p])
             : slc_z_mulm_dist_dev(h->blocks[p], h->ndev, p, h->col_splits, parts, lds, (double *)h->y[p], ldy, k, h->groups[p]);
    }
    SLC_TRY(rc);
  }
  return SLC_OK;
}

static int sharded_sync(slc_sharded *h) {
  int rc = SLC_OK;
  for (int p = 0; p < h->ndev; ++p) {
    DeviceScope scope(h->devices[p], h->streams[p]);
    cudaError_t e = cudaStreamSynchronize(h->streams[p]);
    if (e != cudaSuccess && rc == SLC_OK) rc = fail(SLC_ERR_CUDA, "sharded product: device %d: %s", h->devices[p], cudaGetErrorString(e));
    // every device's word is read (and cleared) even after a failure, so nothing stale is left behind
    const int wrc = spmm_window_check("sharded product");
    if (rc == SLC_OK) rc = wrc;
  }
  if (rc == SLC_OK)
    for (int p = 0; p < h->ndev; ++p) SLC_TRY(peer_check(h->groups[p], "sharded product"));
  return rc;
}

// host operands: k < 0 is the vector form
template <class T>
static int sharded_host_product(slc_sharded *h, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                                const char *fn) {
  SLC_TRY(sharded_check<T>(h, fn));
  const int32_t kk = k < 0 ? 1 : k;
  if (k >= 0) {
    if (ldx < h->ncols)
      return fail(SLC_ERR_DIM, "axpy_: column dimension %d does not match operand dimension %lld", h->ncols, (long long)ldx);
    if (ldy < h->nrows)
      return fail(SLC_ERR_DIM, "axpy_: row dimension %d does not match result dimension %lld", h->nrows, (long long)ldy);
    if (k == 0) return SLC_OK;
    if (h->allgather)
      return fail(SLC_ERR_INVALID, "%s: the row-partitioned block product needs sliced-ELL plans (regular rows)", fn);
  }
  if ((h->ncols && !X) || (h->nrows && !Y)) return fail(SLC_ERR_INVALID, "%s: null operand", fn);
  std::lock_guard<std::mutex> lock(h->mu);
  SLC_TRY(sharded_reserve(h, kk));
  h->parity ^= 1;
  const T *Xh = (const T *)X;
  T *Yh = (T *)Y;
  for (int p = 0; p < h->ndev; ++p) {
    const int64_t n = h->col_splits[p + 1] - h->col_splits[p];
    if (n == 0) continue;
    DeviceScope scope(h->devices[p], h->streams[p]);
    SLC_CUDA(cudaMemcpy2DAsync(h->x[h->parity][p], (size_t)n * sizeof(T), Xh + h->col_splits[p], (size_t)ldx * sizeof(T),
                               (size_t)n * sizeof(T), (size_t)kk, cudaMemcpyHostToDevice, h->streams[p]));
  }
  SLC_TRY(sharded_products<T>(h, k));
  for (int p = 0; p < h->ndev; ++p) {
    const int64_t n = h->row_splits[p + 1] - h->row_splits[p];
    if (n == 0) continue;
    DeviceScope scope(h->devices[p], h->streams[p]);
    SLC_CUDA(cudaMemcpy2DAsync(Yh + h->row_splits[p], (size_t)ldy * sizeof(T), h->y[p], (size_t)n * sizeof(T),
                               (size_t)n * sizeof(T), (size_t)kk, cudaMemcpyDeviceToHost, h->streams[p]));
  }
  return sharded_sync(h);
}

template <class T>
static int sharded_dev_product(slc_sharded *h, int32_t k, const char *fn) {
  SLC_TRY(sharded_check<T>(h, fn));
  if (k == 0) return SLC_OK;
  if (k > 0 && h->allgather)
    return fail(SLC_ERR_INVALID, "%s: the row-partitioned block product needs sliced-ELL plans (regular rows)", fn);
  std::lock_guard<std::mutex> lock(h->mu);
  if ((k < 0 ? 1 : k) > h->cap_cols)
    return fail(SLC_ERR_INVALID, "%s: operand buffers hold %lld column(s); call slc_sharded_operand first", fn,
                (long long)h->cap_cols);
  h->parity ^= 1;  // the buffers slc_sharded_operand handed out become the operand of this product
  return sharded_products<T>(h, k);
}

}  // namespace slc

using namespace slc;

extern "C" {

int slc_d_upload_sharded(int32_t ndev, const int32_t *devices, int32_t nrows, int32_t ncols, const int32_t *colptr,
                         const int32_t *rowidx, const double *vals, slc_sharded **out) {
  return upload_sharded_impl<double>(ndev, devices, nrows, ncols, colptr, rowidx, vals, out);
}
int slc_z_upload_sharded(int32_t ndev, const int32_t *devices, int32_t nrows, int32_t ncols, const int32_t *colptr,
                         const int32_t *rowidx, const double *vals, slc_sharded **out) {
  return upload_sharded_impl<double2>(ndev, devices, nrows, ncols, colptr, rowidx, vals, out);
}

void slc_sharded_free(slc_sharded *h) { sharded_release(h); }

int slc_sharded_dims(const slc_sharded *h, int32_t *ndev, int32_t *nrows, int32_t *ncols, int64_t *nnz) {
  if (!h) return fail(SLC_ERR_INVALID, "slc_sharded_dims: null handle");
  if (ndev) *ndev = h->ndev;
  if (nrows) *nrows = h->nrows;
  if (ncols) *ncols = h->ncols;
  if (nnz) *nnz = h->nnz;
  return SLC_OK;
}

int slc_sharded_splits(const slc_sharded *h, int32_t *row_splits, int32_t *col_splits) {
  if (!h) return fail(SLC_ERR_INVALID, "slc_sharded_splits: null handle");
  for (int p = 0; p <= h->ndev; ++p) {
    if (row_splits) row_splits[p] = h->row_splits[p];
    if (col_splits) col_splits[p] = h->col_splits[p];
  }
  return SLC_OK;
}

int slc_sharded_block(const slc_sharded *h, int32_t part, slc_matrix **block, int32_t *device, void **stream) {
  if (!h || part < 0 || part >= h->ndev) return fail(SLC_ERR_INVALID, "slc_sharded_block: bad argument");
  if (block) *block = h->blocks[part];
  if (device) *device = h->devices[part];
  if (stream) *stream = (void *)h->streams[part];
  return SLC_OK;
}

int slc_sharded_operand(slc_sharded *h, int32_t part, int32_t k, void **x_dev, int64_t *ldx) {
  if (!h || part < 0 || part >= h->ndev || k < 1 || !x_dev) return fail(SLC_ERR_INVALID, "slc_sharded_operand: bad argument");
  std::lock_guard<std::mutex> lock(h->mu);
  SLC_TRY(sharded_reserve(h, k));
  *x_dev = h->x[h->parity ^ 1][part];
  if (ldx) *ldx = std::max<int64_t>(1, h->col_splits[part + 1] - h->col_splits[part]);
  return SLC_OK;
}

int slc_sharded_result(slc_sharded *h, int32_t part, void **y_dev, int64_t *ldy) {
  if (!h || part < 0 || part >= h->ndev || !y_dev) return fail(SLC_ERR_INVALID, "slc_sharded_result: bad argument");
  *y_dev = h->y[part];
  if (ldy) *ldy = std::max<int64_t>(1, h->row_splits[part + 1] - h->row_splits[part]);
  return SLC_OK;
}

int slc_sharded_synchronize(slc_sharded *h) {
  if (!h) return fail(SLC_ERR_INVALID, "slc_sharded_synchronize: null handle");
  std::lock_guard<std::mutex> lock(h->mu);
  return sharded_sync(h);
}

int slc_d_mulv_sharded(slc_sharded *a, const double *x, double *y) {
  return sharded_host_product<double>(a, x, a ? a->ncols : 0, y, a ? a->nrows : 0, -1, "mulV");
}
int slc_z_mulv_sharded(slc_sharded *a, const double *x, double *y) {
  return sharded_host_product<double2>(a, x, a ? a->ncols : 0, y, a ? a->nrows : 0, -1, "mulV");
}
int slc_d_mulm_sharded(slc_sharded *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k) {
  if (k < 0) return fail(SLC_ERR_INVALID, "mulM: negative column count");
  return sharded_host_product<double>(a, X, ldx, Y, ldy, k, "mulM");
}
int slc_z_mulm_sharded(slc_sharded *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k) {
  if (k < 0) return fail(SLC_ERR_INVALID, "mulM: negative column count");
  return sharded_host_product<double2>(a, X, ldx, Y, ldy, k, "mulM");
}
int slc_d_mulv_sharded_dev(slc_sharded *a) { return sharded_dev_product<double>(a, -1, "mulV"); }
int slc_z_mulv_sharded_dev(slc_sharded *a) { return sharded_dev_product<double2>(a, -1, "mulV"); }
int slc_d_mulm_sharded_dev(slc_sharded *a, int32_t k) {
  if (k < 0) return fail(SLC_ERR_INVALID, "mulM: negative column count");
  return sharded_dev_product<double>(a, k, "mulM");
}
int slc_z_mulm_sharded_dev(slc_sharded *a, int32_t k) {
  if (k < 0) return fail(SLC_ERR_INVALID, "mulM: negative column count");
  return sharded_dev_product<double2>(a, k, "mulM");
}

}  // extern "C"
