// slc_core.cu -- library state, handles, upload/download.
// Reference shapes: Matrix record Data/Matrix/Sparse.hs:67-76; withConstMatrix / fromForeign
// Data/Matrix/Sparse/Foreign.hs:24-88 (paths relative to /root/reference/sparse-linear/src).
#include "slc_common.cuh"
#include <vector>

namespace slc {

std::atomic<int64_t> g_launches{0};

ThreadState &tls() {
  static thread_local ThreadState st;
  return st;
}

int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  tls().last_error = buf;
  return code;
}

int sm_count() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    cached_dev = dev;
  }
  return cached;
}

int configure_pool_once() {
  // Keep freed blocks in the stream-ordered pool: repeated products and SpGEMM calls reuse
  // them instead of going back to the driver.
  static thread_local int done_dev = -1;
  int dev = 0;
  SLC_CUDA(cudaGetDevice(&dev));
  if (dev == done_dev) return SLC_OK;
  cudaMemPool_t pool;
  SLC_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
  uint64_t thr = UINT64_MAX;
  SLC_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
  done_dev = dev;
  return SLC_OK;
}

int matrix_alloc(int elem, int32_t nrows, int32_t ncols, int64_t nnz, slc_matrix **out) {
  SLC_TRY(configure_pool_once());
  slc_matrix *m = new (std::nothrow) slc_matrix();
  if (!m) return fail(SLC_ERR_NOMEM, "out of host memory");
  m->elem = elem;
  m->nrows = nrows;
  m->ncols = ncols;
  m->nnz = nnz;
  cudaGetDevice(&m->device);
  DevBuf<int32_t> cp, ri;
  DevBuf<char> v;
  int rc = cp.alloc((size_t)ncols + 1);
  if (rc == SLC_OK) rc = ri.alloc((size_t)nnz);
  if (rc == SLC_OK) rc = v.alloc((size_t)nnz * elem_size(elem));
  if (rc != SLC_OK) {
    delete m;
    return rc;
  }
  m->colptr = cp.take();
  m->rowidx = ri.take();
  m->vals = v.take();
  *out = m;
  return SLC_OK;
}

void matrix_release(slc_matrix *m) {
  if (!m) return;
  int prev = -1;
  cudaGetDevice(&prev);
  if (prev != m->device) cudaSetDevice(m->device);
  dev_free(m->colptr);
  dev_free(m->rowidx);
  dev_free(m->vals);
  dev_free(m->rowptr);
  dev_free(m->colidx);
  dev_free(m->rvals);
  dev_free(m->sell.slice_ptr);
  dev_free(m->sell.kbase);
  dev_free(m->sell.panel_order);
  dev_free(m->sell.rowlen);
  dev_free(m->sell.colidx);
  dev_free(m->sell.vals);
  dev_free(m->sell.slice_cmin);
  dev_free(m->sell.slice_cmax);
  dev_free(m->bal.chunk_row);
  dev_free(m->bal.hot_cols);
  dev_free(m->bal.colidx_hot);
  dev_free(m->dist_table);
  dev_free(m->brick.dia);
  dev_free(m->brick.mask);
  if (m->pipe_x) cudaFree(m->pipe_x);
  if (m->pipe_y) cudaFree(m->pipe_y);
  if (m->pipe_in) {
    cudaStreamSynchronize(m->pipe_in);
    cudaStreamSynchronize(m->pipe_out);
    for (auto e : m->pipe_events) cudaEventDestroy(e);
    cudaStreamDestroy(m->pipe_in);
    cudaStreamDestroy(m->pipe_out);
  }
  if (prev != m->device && prev >= 0) cudaSetDevice(prev);
  delete m;
}

// Format check of an uploaded matrix (the reference's checkMatrix): row indices in bounds and strictly
// increasing inside every column.  A descent rowidx[p-1] >= rowidx[p] is legal only where a column
// starts at p, so: one thread per entry counts the descents, one thread per non-empty column counts
// the descents that sit on its start, and the matrix is well formed iff the two counts agree.
// acc[0] = out-of-bounds flag, acc[1] = descents, acc[2] = descents at column starts.
// Grid-stride over the entries, counts kept in registers, reduced per warp and per block: one atomic per
// block (round 2 issued one per warp on a single address -- 4.5 ms for the 214 M entries of the 200^3
// stencil, where reading the indices takes 0.15 ms).
constexpr int VALIDATE_THREADS = 256;
__global__ void __launch_bounds__(VALIDATE_THREADS)
k_validate_entries(int32_t nrows, int64_t nnz, const int32_t *__restrict__ rowidx,
                   unsigned long long *__restrict__ acc) {
  __shared__ unsigned s_desc[VALIDATE_THREADS / 32];
  __shared__ unsigned s_oob[VALIDATE_THREADS / 32];
  unsigned descents = 0, oob = 0;
  const int64_t stride = (int64_t)gridDim.x * VALIDATE_THREADS;
  for (int64_t p = (int64_t)blockIdx.x * VALIDATE_THREADS + threadIdx.x; p < nnz; p += stride) {
    const int32_t r = rowidx[p];
    oob |= (unsigned)(r < 0 || r >= nrows);
    descents += (unsigned)(p > 0 && rowidx[p - 1] >= r);
  }
  descents = __reduce_add_sync(0xffffffffu, descents);
  oob = __reduce_or_sync(0xffffffffu, oob);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_desc[warp] = descents;
    s_oob[warp] = oob;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned d = 0, o = 0;
#pragma unroll
    for (int w = 0; w < VALIDATE_THREADS / 32; ++w) {
      d += s_desc[w];
      o |= s_oob[w];
    }
    if (o) atomicOr(acc, 1ull);
    if (d) atomicAdd(acc + 1, (unsigned long long)d);
  }
}

__global__ void k_validate_starts(int32_t ncols, const int32_t *__restrict__ colptr,
                                  const int32_t *__restrict__ rowidx, unsigned long long *__restrict__ acc) {
  const int32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  bool descent = false;
  if (c < ncols) {
    const int32_t p = colptr[c];
    descent = p > 0 && p < colptr[c + 1] && rowidx[p - 1] >= rowidx[p];
  }
  const unsigned d = __ballot_sync(0xffffffffu, descent);
  if ((threadIdx.x & 31) == 0 && d) atomicAdd(acc + 2, (unsigned long long)__popc(d));
}

template <class T>
static int upload_impl(int32_t nrows, int32_t ncols, const int32_t *colptr, const int32_t *rowidx,
                       const double *vals, slc_matrix **out) {
  if (!out) return fail(SLC_ERR_INVALID, "upload: null output handle");
  *out = nullptr;
  if (nrows < 0 || ncols < 0) return fail(SLC_ERR_INVALID, "upload: negative dimension");
  if (!colptr) return fail(SLC_ERR_INVALID, "upload: null column pointers");
  if (colptr[0] != 0) return fail(SLC_ERR_FORMAT, "upload: pointers do not start at 0");
  for (int32_t c = 0; c < ncols; ++c)
    if (colptr[c] > colptr[c + 1])
      return fail(SLC_ERR_FORMAT, "upload: pointers decrease at column %d", c);
  const int64_t nnz = colptr[ncols];
  if (nnz > 0 && (!rowidx || !vals)) return fail(SLC_ERR_INVALID, "upload: null index or value array");
  slc_matrix *m = nullptr;
  SLC_TRY(matrix_alloc(Num<T>::elem, nrows, ncols, nnz, &m));
  cudaStream_t s = cur_stream();
  cudaError_t e = cudaSuccess;
  int rc = host_to_dev(m->colptr, colptr, ((size_t)ncols + 1) * sizeof(int32_t), s);
  if (rc == SLC_OK && nnz) rc = host_to_dev(m->rowidx, rowidx, (size_t)nnz * sizeof(int32_t), s);
  if (rc == SLC_OK && nnz) rc = host_to_dev(m->vals, vals, (size_t)nnz * sizeof(T), s);
  DevBuf<unsigned long long> acc;
  unsigned long long hacc[3] = {0, 0, 0};
  if (rc == SLC_OK) rc = acc.alloc(3);
  if (e == cudaSuccess && rc == SLC_OK) e = cudaMemsetAsync(acc.p, 0, sizeof hacc, s);
  if (e == cudaSuccess && rc == SLC_OK && nnz) {
    // a few blocks per SM, each thread takes several entries
    const unsigned vblocks = (unsigned)std::min<int64_t>(ceil_div(nnz, VALIDATE_THREADS * 4), (int64_t)sm_count() * 16);
    k_validate_entries<<<vblocks, VALIDATE_THREADS, 0, s>>>(nrows, nnz, m->rowidx, acc.p);
    k_validate_starts<<<(unsigned)ceil_div(ncols, 256), 256, 0, s>>>(ncols, m->colptr, m->rowidx, acc.p);
    g_launches.fetch_add(2, std::memory_order_relaxed);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess && rc == SLC_OK)
    e = cudaMemcpyAsync(hacc, acc.p, sizeof hacc, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess && rc == SLC_OK) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess || rc != SLC_OK) {
    matrix_release(m);
    if (rc != SLC_OK) return rc;
    return fail(SLC_ERR_CUDA, "upload: CUDA error %s", cudaGetErrorString(e));
  }
  if (hacc[0] || hacc[1] != hacc[2]) {
    matrix_release(m);
    return fail(SLC_ERR_FORMAT, hacc[0] ? "upload: row index out of bounds"
                                        : "upload: row indices not increasing in a column");
  }
  *out = m;
  return SLC_OK;
}

template <class T>
static int download_impl(const slc_matrix *m, int32_t *colptr, int32_t *rowidx, double *vals) {
  if (!m) return fail(SLC_ERR_INVALID, "download: null handle");
  if (m->elem != Num<T>::elem) return fail(SLC_ERR_INVALID, "download: element type mismatch");
  cudaStream_t s = cur_stream();
  if (colptr) SLC_TRY(dev_to_host(colptr, m->colptr, ((size_t)m->ncols + 1) * sizeof(int32_t), s));
  if (rowidx && m->nnz) SLC_TRY(dev_to_host(rowidx, m->rowidx, (size_t)m->nnz * sizeof(int32_t), s));
  if (vals && m->nnz) SLC_TRY(dev_to_host(vals, m->vals, (size_t)m->nnz * sizeof(T), s));
  SLC_CUDA(cudaStreamSynchronize(s));
  return SLC_OK;
}

}  // namespace slc

using namespace slc;

extern "C" {

int slc_version(void) { return 100; }

const char *slc_last_error(void) { return tls().last_error.c_str(); }

int slc_device_count(int *count) {
  if (!count) return fail(SLC_ERR_INVALID, "slc_device_count: null argument");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) {
    *count = 0;
    (void)cudaGetLastError();
    return fail(SLC_ERR_CUDA, "slc_device_count: %s", cudaGetErrorString(e));
  }
  return SLC_OK;
}

int slc_set_device(int device) {
  SLC_CUDA(cudaSetDevice(device));
  return SLC_OK;
}

int slc_set_stream(void *cuda_stream) {
  tls().stream = (cudaStream_t)cuda_stream;
  return SLC_OK;
}

int slc_synchronize(void) {
  SLC_CUDA(cudaStreamSynchronize(cur_stream()));
  return SLC_OK;
}

int64_t slc_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int slc_host_alloc(void **ptr, int64_t bytes) {
  if (!ptr || bytes < 0) return fail(SLC_ERR_INVALID, "slc_host_alloc: bad argument");
  SLC_CUDA(cudaMallocHost(ptr, (size_t)(bytes ? bytes : 1)));
  return SLC_OK;
}
int slc_host_free(void *ptr) {
  if (ptr) SLC_CUDA(cudaFreeHost(ptr));
  return SLC_OK;
}
int slc_dev_alloc(void **ptr, int64_t bytes) {
  if (!ptr || bytes < 0) return fail(SLC_ERR_INVALID, "slc_dev_alloc: bad argument");
  SLC_CUDA(cudaMalloc(ptr, (size_t)(bytes ? bytes : 1)));
  return SLC_OK;
}
int slc_dev_free(void *ptr) {
  if (ptr) SLC_CUDA(cudaFree(ptr));
  return SLC_OK;
}
int slc_memcpy_h2d(void *dst_dev, const void *src_host, int64_t bytes) {
  if (bytes < 0) return fail(SLC_ERR_INVALID, "slc_memcpy_h2d: negative size");
  if (bytes) {
    SLC_CUDA(cudaMemcpyAsync(dst_dev, src_host, (size_t)bytes, cudaMemcpyHostToDevice, cur_stream()));
    SLC_CUDA(cudaStreamSynchronize(cur_stream()));
  }
  return SLC_OK;
}
int slc_memcpy_d2h(void *dst_host, const void *src_dev, int64_t bytes) {
  if (bytes < 0) return fail(SLC_ERR_INVALID, "slc_memcpy_d2h: negative size");
  if (bytes) {
    SLC_CUDA(cudaMemcpyAsync(dst_host, src_dev, (size_t)bytes, cudaMemcpyDeviceToHost, cur_stream()));
    SLC_CUDA(cudaStreamSynchronize(cur_stream()));
  }
  return SLC_OK;
}

int slc_memcpy_d2d(void *dst_dev, const void *src_dev, int64_t bytes) {
  if (bytes < 0) return fail(SLC_ERR_INVALID, "slc_memcpy_d2d: negative size");
  if (bytes) SLC_CUDA(cudaMemcpyAsync(dst_dev, src_dev, (size_t)bytes, cudaMemcpyDeviceToDevice, cur_stream()));
  return SLC_OK;
}

// ---- CUDA IPC: one process per GPU, peers map each other's x slices -------------------------------
// An IPC handle names a whole driver allocation, and small cudaMalloc requests are carved out of
// larger ones, so the exported blob is the 64-byte handle of the containing allocation followed by
// the 8-byte offset of the pointer inside it.
static int containing_allocation(const void *p, void **base) {
  typedef int (*range_fn)(unsigned long long *, size_t *, unsigned long long);
  static range_fn fn = nullptr;
  if (!fn) {
    void *f = nullptr;
    cudaDriverEntryPointQueryResult q;
    SLC_CUDA(cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q));
    if (!f) return fail(SLC_ERR_CUDA, "cuMemGetAddressRange is not available");
    fn = (range_fn)f;
  }
  unsigned long long b = 0;
  size_t sz = 0;
  if (fn(&b, &sz, (unsigned long long)(uintptr_t)p) != 0)
    return fail(SLC_ERR_INVALID, "slc_ipc_export: not a device allocation");
  *base = (void *)(uintptr_t)b;
  return SLC_OK;
}

int slc_ipc_export(const void *dev_ptr, void *handle72) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (!dev_ptr || !handle72) return fail(SLC_ERR_INVALID, "slc_ipc_export: null argument");
  void *base = nullptr;
  SLC_TRY(containing_allocation(dev_ptr, &base));
  cudaIpcMemHandle_t h;
  SLC_CUDA(cudaIpcGetMemHandle(&h, base));
  const int64_t off = (const char *)dev_ptr - (const char *)base;
  memcpy(handle72, &h, sizeof h);
  memcpy((char *)handle72 + sizeof h, &off, sizeof off);
  return SLC_OK;
}
// Two exported pointers of one peer can sit in the same driver allocation (small cudaMalloc blocks are
// sub-allocated), and a process may open an IPC handle only once: imports are therefore reference-counted per
// (device, handle), the second import reuses the first mapping, and the mapping is closed with its last user.
namespace {
struct IpcMapping {
  int device;
  cudaIpcMemHandle_t handle;
  void *base;
  int refs;
};
std::mutex g_ipc_mu;
std::vector<IpcMapping> g_ipc_open;
}  // namespace

int slc_ipc_import(const void *handle72, void **dev_ptr) {
  if (!dev_ptr || !handle72) return fail(SLC_ERR_INVALID, "slc_ipc_import: null argument");
  cudaIpcMemHandle_t h;
  int64_t off = 0;
  memcpy(&h, handle72, sizeof h);
  memcpy(&off, (const char *)handle72 + sizeof h, sizeof off);
  int dev = 0;
  SLC_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(g_ipc_mu);
  for (auto &mp : g_ipc_open)
    if (mp.device == dev && memcmp(&mp.handle, &h, sizeof h) == 0) {
      ++mp.refs;
      *dev_ptr = (char *)mp.base + off;
      return SLC_OK;
    }
  void *base = nullptr;
  SLC_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  g_ipc_open.push_back(IpcMapping{dev, h, base, 1});
  *dev_ptr = (char *)base + off;
  return SLC_OK;
}
int slc_ipc_close(void *dev_ptr) {
  if (!dev_ptr) return SLC_OK;
  void *base = nullptr;
  SLC_TRY(containing_allocation(dev_ptr, &base));
  std::lock_guard<std::mutex> lock(g_ipc_mu);
  for (size_t i = 0; i < g_ipc_open.size(); ++i)
    if (g_ipc_open[i].base == base) {
      if (--g_ipc_open[i].refs > 0) return SLC_OK;
      g_ipc_open.erase(g_ipc_open.begin() + (long)i);
      break;
    }
  SLC_CUDA(cudaIpcCloseMemHandle(base));
  return SLC_OK;
}

int slc_d_upload(int32_t nrows, int32_t ncols, const int32_t *colptr, const int32_t *rowidx,
                 const double *vals, slc_matrix **out) {
  return upload_impl<double>(nrows, ncols, colptr, rowidx, vals, out);
}
int slc_z_upload(int32_t nrows, int32_t ncols, const int32_t *colptr, const int32_t *rowidx,
                 const double *vals, slc_matrix **out) {
  return upload_impl<double2>(nrows, ncols, colptr, rowidx, vals, out);
}

void slc_free(slc_matrix *m) { matrix_release(m); }

int slc_dims(const slc_matrix *m, int32_t *nrows, int32_t *ncols, int64_t *nnz) {
  if (!m) return fail(SLC_ERR_INVALID, "slc_dims: null handle");
  if (nrows) *nrows = m->nrows;
  if (ncols) *ncols = m->ncols;
  if (nnz) *nnz = m->nnz;
  return SLC_OK;
}

int slc_elem_type(const slc_matrix *m) { return m ? m->elem : 0; }

int slc_d_download(const slc_matrix *m, int32_t *colptr, int32_t *rowidx, double *vals) {
  return download_impl<double>(m, colptr, rowidx, vals);
}
int slc_z_download(const slc_matrix *m, int32_t *colptr, int32_t *rowidx, double *vals) {
  return download_impl<double2>(m, colptr, rowidx, vals);
}

}  // extern "C"
