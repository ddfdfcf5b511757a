// slc_common.cuh -- shared host/device plumbing of libsparselinear_b200.so
// (handle layout, error reporting, stream-ordered buffers, element arithmetic).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>
#include <utility>
#include <new>
#include <string>
#include <vector>

#include "slc.h"

namespace slc {

// ---- per-thread state ---------------------------------------------------------------------

struct ThreadState {
  cudaStream_t stream = nullptr;
  std::string last_error;
};
ThreadState &tls();
inline cudaStream_t cur_stream() { return tls().stream; }

int fail(int code, const char *fmt, ...);  // records the message, returns code
extern std::atomic<int64_t> g_launches;

#define SLC_CUDA(expr)                                                                       \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ::slc::fail(_e == cudaErrorMemoryAllocation ? SLC_ERR_NOMEM : SLC_ERR_CUDA,     \
                         "%s: CUDA error %s at %s:%d", __func__, cudaGetErrorString(_e),     \
                         __FILE__, __LINE__);                                                \
  } while (0)

#define SLC_TRY(expr)                                                                        \
  do {                                                                                       \
    int _rc = (expr);                                                                        \
    if (_rc != SLC_OK) return _rc;                                                           \
  } while (0)

// Count a kernel launch and surface launch-configuration errors immediately.
#define SLC_LAUNCHED()                                                                       \
  do {                                                                                       \
    ::slc::g_launches.fetch_add(1, std::memory_order_relaxed);                               \
    SLC_CUDA(cudaGetLastError());                                                            \
  } while (0)

// ---- stream-ordered device buffer ------------------------------------------------------------

// The device's default memory pool keeps what is freed into it (release threshold = max), so the work
// buffers of one call are reused by the next instead of going back to the driver at the first
// synchronisation.  Cheap after the first call per thread and device.
int configure_pool_once();

// Owns one cudaMallocAsync allocation.  Freed on the current stream, so a buffer may be
// released right after the last kernel that reads it has been enqueued.
template <class U>
struct DevBuf {
  U *p = nullptr;
  size_t n = 0;
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept {
    if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
    return *this;
  }
  ~DevBuf() { release(); }
  int alloc(size_t count) {
    release();
    if (int rc = configure_pool_once(); rc != SLC_OK) return rc;
    n = count;
    size_t bytes = (count ? count : 1) * sizeof(U);
    cudaError_t e = cudaMallocAsync((void **)&p, bytes, cur_stream());
    if (e != cudaSuccess) {
      p = nullptr; n = 0;
      (void)cudaGetLastError();
      return fail(e == cudaErrorMemoryAllocation ? SLC_ERR_NOMEM : SLC_ERR_CUDA,
                  "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    return SLC_OK;
  }
  void release() {
    if (p) { cudaFreeAsync(p, cur_stream()); p = nullptr; n = 0; }
  }
  U *take() { U *q = p; p = nullptr; n = 0; return q; }
};

// ---- programmatic dependent launch (PDL) ----------------------------------------------------------
// A kernel launched with launch_pdl(..., pdl = true) may have its blocks scheduled while the previous
// kernel of the stream is still draining (after every block of that kernel has executed
// pdl_launch_dependents).  Such a block may only touch memory that no earlier kernel writes -- the
// matrix -- until it has passed pdl_wait(), which returns when the previous kernel has completed and
// its writes are visible.  Without the launch attribute both instructions do nothing.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t s,
                              bool pdl, Args &&...args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(block, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
#endif

// One record of the count / place / sort-per-bucket paths (compress: key = row << 32 | input position;
// transpose: key = column).  16 bytes for Double, 32 for Complex Double: one sector per record.
template <class T>
struct __align__(16) AsmEnt {
  uint64_t key;
  T val;
};

inline int env_int(const char *name, int dflt) {
  const char *v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

inline void dev_free(void *p) {
  if (p) cudaFreeAsync(p, cur_stream());
}

// Per-device cache of a value that CUDA keeps per device (e.g. the dynamic shared-memory limit set with
// cudaFuncSetAttribute): one slot per device ordinal, indexed by the calling thread's current device.
// A process-wide static would be wrong as soon as one process drives two GPUs (slc_set_device,
// slc_*_upload_sharded).  Slots are atomics, so two host threads may race to raise the same limit safely.
constexpr int MAX_DEVICES = 64;
template <class V>
struct PerDevice {
  std::atomic<V> slot[MAX_DEVICES];
  PerDevice() { for (auto &s : slot) s.store(V(), std::memory_order_relaxed); }
  std::atomic<V> &cur() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) dev = 0;
    return slot[dev];
  }
};

// Raise a kernel's dynamic shared-memory limit on the current device if it is below `bytes`.
template <class K>
inline cudaError_t ensure_dyn_smem(K kernel, PerDevice<size_t> &cache, size_t bytes) {
  std::atomic<size_t> &have = cache.cur();
  if (bytes <= have.load(std::memory_order_acquire)) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) have.store(bytes, std::memory_order_release);
  return e;
}

// ---- the handle ---------------------------------------------------------------------------------

enum ElemType : int { ELEM_D = 'd', ELEM_Z = 'z' };
enum SpmvKind : int {
  SPMV_AUTO = 0,
  SPMV_SELL = 1,
  SPMV_CSR_WARP = 2,
  SPMV_CSR_BALANCED = 3,  // nnz chunks, products staged in shared memory (k_spmv_bal)
  SPMV_CSR_MERGE = 4,     // nnz chunks, merge-style: registers + segmented scan (k_spmv_merge)
};

// Sliced-ELL plan (slices of 32 rows, column-major inside a slice, row per thread).
struct SellPlan {
  bool built = false;
  int32_t nslices = 0;
  int64_t padded = 0;        // stored elements including padding
  int64_t *slice_ptr = nullptr;  // [nslices+1] element offsets
  int32_t *rowlen = nullptr;     // [nrows]
  int32_t *colidx = nullptr;     // [padded]
  void *vals = nullptr;          // [padded]
  // Index compression: kbase[slice_ptr[s] / 32 + k] = c - lane when entry k of every row of slice s
  // that has one sits at column (that value + lane) -- a stencil away from the grid boundary -- so
  // the warp needs 4 bytes instead of 128 for those indices; SELL_IRREGULAR otherwise.  nullptr when
  // fewer than a quarter of the entries qualify.
  int32_t *kbase = nullptr;
  int64_t unit_entries = 0;       // stored entries covered by a unit-stride base
  int32_t *slice_cmin = nullptr;  // [nslices] smallest / largest column of each slice
  int32_t *slice_cmax = nullptr;
  // distributed products: host copy of the slice ranges and the cached interior run
  std::vector<int32_t> h_cmin, h_cmax;
  int32_t interior_s0 = -1, interior_s1 = -1, interior_lo = 0, interior_hi = 0;
  int32_t side_cmin[2] = {0x7fffffff, 0x7fffffff}, side_cmax[2] = {-1, -1};  // columns the rows below / above the interior reach
  // widest column window of a 1024-row block (k_spmm_window), -1 = not measured yet
  int32_t window_max = -1;
  int32_t window_max2 = -1;  // widest union window over pairs of neighbouring 1024-row blocks (cluster kernel)
  // k_spmm_panel: distance in rows between grid lines (the near off-diagonal band of a stencil),
  // -1 = not measured yet, 0 = none; and the cached slice order that tiles blocks over lines
  int32_t line_stride = -1;
  int32_t *panel_order = nullptr;   // [panel_blocks * warps] slice of every warp, -1 = idle
  int64_t panel_blocks = 0;
  int32_t panel_key[4] = {-1, -1, -1, -1};   // (first slice, last slice, stride, warps << 8 | width)
};

// nnz-balanced CSR plan: fixed-size chunks of nonzeros, first row of each chunk.
struct BalancedPlan {
  bool built = false;
  int32_t nchunks = 0;
  int32_t *chunk_row = nullptr;  // [nchunks+1] first row touching each chunk
  // hot columns (k_spmv_bal_hot): -1 = not looked at yet, 0 = not worth it
  int32_t hot_k = -1;
  int32_t *hot_cols = nullptr;    // [hot_k] column of every shared-memory slot, ascending
  int32_t *colidx_hot = nullptr;  // [nnz] mirror's column indices, hot ones as 0x80000000 | slot
  double hot_share = 0.0;         // fraction of the nonzeros that sit in hot columns
};

// Offset-slot plan for matrices whose rows all follow one 3 x 3 x 3 offset pattern (slc_spmm_brick.cu)
struct BrickPlan {
  int state = -1;            // -1 not looked at, 0 the matrix does not qualify, 1 built
  int32_t L = 0, P = 0;      // line and plane strides
  int64_t shift = 0;         // column of a row's centre entry minus the row (row blocks keep global columns)
  int64_t nrows_pad = 0;
  double *dia = nullptr;     // [27][nrows_pad] values by offset slot
  uint32_t *mask = nullptr;  // [nrows_pad] which slots a row has
};

}  // namespace slc

struct slc_matrix {
  // Serialises every call that may build a plan, touch the per-handle pipeline streams or refresh the
  // partition tables: two host threads (two Haskell sparks forcing `mulV a x` and `mulV a y`) may share
  // one handle.  Recursive because entry points nest (mulm_dist -> ensure_plan -> ensure_csr).
  std::recursive_mutex mu;
  int elem = slc::ELEM_D;
  int device = 0;
  int32_t nrows = 0, ncols = 0;
  int64_t nnz = 0;
  // canonical CSC (what the reference's Matrix record holds)
  int32_t *colptr = nullptr;
  int32_t *rowidx = nullptr;
  void *vals = nullptr;
  // row-major mirror = CSC of the transpose, built on demand by the transpose kernels
  bool has_csr = false;
  int32_t *rowptr = nullptr;
  int32_t *colidx = nullptr;
  void *rvals = nullptr;
  // row statistics (filled with the mirror)
  int32_t max_rowlen = 0;
  // banded matrices: the far off-diagonal distance (multiple of 32) that SpMM blocks interleave
  // their warps by, 0 = none detected (see k_spmm_short)
  int32_t band_stride = 0;
  // product plans
  int spmv_kind = slc::SPMV_AUTO;      // requested
  int spmv_active = slc::SPMV_AUTO;    // chosen
  slc::SellPlan sell;
  slc::BalancedPlan bal;
  slc::BrickPlan brick;
  // streams and events of the pipelined host-operand product (upload / compute / download overlap)
  cudaStream_t pipe_in = nullptr, pipe_out = nullptr;
  std::vector<cudaEvent_t> pipe_events;
  void *pipe_x = nullptr, *pipe_y = nullptr;  // device copies of the host operands (kept between calls)
  size_t pipe_x_bytes = 0, pipe_y_bytes = 0;
  // partition tables of the recent distributed products (device copies + host shadows).  A caller that
  // double-buffers its operand alternates between two tables; with a few cached slots neither costs a
  // host-to-device copy per product, and the device addresses stay fixed (CUDA-graph friendly).
  static constexpr int DIST_SLOTS = 4;
  void *dist_table = nullptr;                     // DIST_SLOTS tables of DIST_TABLE_BYTES each
  static constexpr int DIST_TABLE_BYTES = 384;
  bool dist_slot_valid[DIST_SLOTS] = {};
  int dist_next = 0;
  alignas(8) unsigned char dist_table_host[DIST_SLOTS][DIST_TABLE_BYTES];
};

namespace slc {

// lock a handle for the rest of the scope (null-safe; const handles are locked too: plans are caches)
struct HandleLock {
  std::recursive_mutex *mu;
  explicit HandleLock(const slc_matrix *m) : mu(m ? &const_cast<slc_matrix *>(m)->mu : nullptr) { if (mu) mu->lock(); }
  ~HandleLock() { if (mu) mu->unlock(); }
  HandleLock(const HandleLock &) = delete;
  HandleLock &operator=(const HandleLock &) = delete;
};

inline size_t elem_size(int elem) { return elem == ELEM_Z ? 16 : 8; }

int matrix_alloc(int elem, int32_t nrows, int32_t ncols, int64_t nnz, slc_matrix **out);
void matrix_release(slc_matrix *m);

// ---- element arithmetic -------------------------------------------------------------------------
// The reference multiplies, rounds, adds, rounds (GHC with -msse2 only,
// sparse-linear.cabal:33).  The __d*_rn intrinsics are never contracted into FMA, so a sum
// taken in the reference's order is bit-identical to it.

template <class T> struct Num;

template <> struct Num<double> {
  static constexpr int elem = ELEM_D;
  static __host__ __device__ __forceinline__ double zero() { return 0.0; }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double conj(double a) { return a; }
  static __device__ __forceinline__ bool eq(double a, double b) { return a == b; }
};

// Data.Complex: (x:+y)*(x':+y') = (x*x' - y*y') :+ (x*y' + y*x')
template <> struct Num<double2> {
  static constexpr int elem = ELEM_Z;
  static __host__ __device__ __forceinline__ double2 zero() { return double2{0.0, 0.0}; }
  static __device__ __forceinline__ double2 mul(double2 a, double2 b) {
    return double2{__dsub_rn(__dmul_rn(a.x, b.x), __dmul_rn(a.y, b.y)),
                   __dadd_rn(__dmul_rn(a.x, b.y), __dmul_rn(a.y, b.x))};
  }
  static __device__ __forceinline__ double2 add(double2 a, double2 b) {
    return double2{__dadd_rn(a.x, b.x), __dadd_rn(a.y, b.y)};
  }
  static __device__ __forceinline__ double2 conj(double2 a) { return double2{a.x, -a.y}; }
  static __device__ __forceinline__ bool eq(double2 a, double2 b) { return a.x == b.x && a.y == b.y; }
};

// Start value of an output entry of the block product: 0 (mulM), y (the reference's axpy_ into a block
// the caller filled, Sparse.hs:482-488) or beta * y.  A runtime mode, warp-uniform: one template
// instantiation serves all three.
template <class T>
struct YInit {
  int mode;  // 0: zero, 1: y, 2: beta * y
  T beta;
  __device__ __forceinline__ T operator()(const T *yp) const {
    if (mode == 0) return Num<T>::zero();
    const T y = *yp;
    return mode == 1 ? y : Num<T>::mul(beta, y);
  }
};
template <class T> inline YInit<T> yinit_zero() { YInit<T> yi; yi.mode = 0; yi.beta = Num<T>::zero(); return yi; }

// ---- cache-hinted memory access -----------------------------------------------------------------
// Matrix streams (indices, values) are read exactly once per product: keep them out of L1
// and mark them first to go in L2, so the gathered vector stays resident in the 126 MB L2.

__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ int32_t ld_stream(const int32_t *p, uint64_t pol) {
  int32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;"
               : "=r"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ double ld_stream(const double *p, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;"
               : "=d"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ double2 ld_stream(const double2 *p, uint64_t pol) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;"
               : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ int4 ld_stream(const int4 *p, uint64_t pol) {
  int4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.s32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
  return v;
}
// gathered operand: read-only path, allocate in L1, last to go in L2
__device__ __forceinline__ double ld_gather(const double *p, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ double2 ld_gather(const double2 *p, uint64_t pol) {
  double2 v;
  asm volatile("ld.global.nc.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;"
               : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
  return v;
}
// operand entry that lives in a peer GPU's HBM and is rewritten between products: a plain (coherent) load,
// not the read-only .nc path -- the data changes during the kernel's lifetime only before the acquire
// that orders this load, but .nc makes no promise at all for memory another GPU writes
__device__ __forceinline__ double ld_peer(const double *p) {
  double v;
  asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double2 ld_peer(const double2 *p) {
  double2 v;
  asm volatile("ld.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}
// streaming store of a result that is not read again by this kernel
__device__ __forceinline__ void st_stream(double *p, double v) {
  asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void st_stream(double2 *p, double2 v) {
  asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

// ---- keeping a batch of gathers in flight -------------------------------------------------------
// ptxas sinks independent loads below the arithmetic on earlier ones to reuse their registers, which
// turns a batch of N gathers into several exposed memory latencies (ncu, round 1: k_spmm_short
// stalled at 6 points per column tile, 3 loads in flight instead of 10).  Neither barriers nor
// volatile asm stop it: .nc loads move across both.  A true data dependence does: OR one word of
// every value of the batch (`probe`), mask it with a kernel argument that is always 0 (`zmask`, opaque
// to the compiler) and XOR the result into the first operand of every chain (`tie`).  Values are
// unchanged, and no product can issue before every load of the batch has been issued.
__device__ __forceinline__ uint32_t probe(double v) { return (uint32_t)__double2hiint(v); }
__device__ __forceinline__ uint32_t probe(double2 v) { return (uint32_t)__double2hiint(v.x); }
__device__ __forceinline__ void tie(double &v, uint32_t r) {
  v = __hiloint2double(__double2hiint(v) ^ (int)r, __double2loint(v));
}
__device__ __forceinline__ void tie(double2 &v, uint32_t r) { tie(v.x, r); }

// ---- small helpers ------------------------------------------------------------------------------

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
int sm_count();

// device primitives implemented in slc_prims.cu (radix sort, scans)
// tile shape of the radix kernels (a block takes RS_TILE consecutive keys, 256 digit values per pass)
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 keys per block
constexpr int RS_BINS = 256;
// offsets[d * ntiles + t], ntiles = ceil(n / RS_TILE): see slc_prims.cu
int tile_digit_offsets_u32(const uint32_t *keys, int64_t n, int shift, int32_t *offsets);
// *out <- max(*out, max(v[0 .. n)))  (values >= 0; *out is a device word the caller has initialised)
int max_i32_into(const int32_t *v, int64_t n, unsigned long long *out);
// *out += sum of v[0 .. n) (values >= 0), in 64 bits
int sum_i32_into(const int32_t *v, int64_t n, unsigned long long *out);
int exclusive_scan_i32(const int32_t *in, int32_t *out, int64_t n);  // out may alias in
int exclusive_scan_i64(const int64_t *in, int64_t *out, int64_t n);
// stable sort of (key, value) pairs on key bits [0, end_bit).  (ka, va) / (kb, vb) are ping-pong
// scratch; the first pass reads keys from `src` if given (left intact; needs iota_vals) else from ka,
// and takes values from va, or 0..n-1 when iota_vals.  *res_k / *res_v: the buffers holding the result.
int sort_pairs_u32(const uint32_t *src, uint32_t *ka, uint32_t *kb, uint32_t *va, uint32_t *vb, int64_t n,
                   int end_bit, bool iota_vals, uint32_t **res_k, uint32_t **res_v);
int sort_pairs_u64(const uint64_t *src, uint64_t *ka, uint64_t *kb, uint32_t *va, uint32_t *vb, int64_t n,
                   int end_bit, bool iota_vals, uint64_t **res_k, uint32_t **res_v);
// bucket sort: keys in [0, nkeys) read in place from src, values = the input positions; besides the
// sorted pairs it fills ptr_out[0 .. nkeys] = first sorted position of every key value (n > 0).
int sort_buckets_u32(const uint32_t *src, uint32_t *ka, uint32_t *kb, uint32_t *va, uint32_t *vb, int64_t n,
                     int32_t nkeys, uint32_t **res_k, uint32_t **res_v, int32_t *ptr_out);

// built in slc_transpose.cu
template <class T>
int transpose_csc(int32_t nrows, int32_t ncols, int64_t nnz, const int32_t *colptr,
                  const int32_t *rowidx, const T *vals, bool conj, int32_t *out_ptr,
                  int32_t *out_idx, T *out_vals);
int ensure_csr(slc_matrix *m);

// built in slc_spmm_window.cu: the TMA-staged SpMM for banded short-row matrices (*done = it ran)
template <class T>
int spmm_window_try(slc_matrix *m, const T *X, int64_t ldx, T *Y, int64_t ldy, int32_t k, YInit<T> yi, bool *done);

// slc_hostxfer.cu: copies between lent host arrays and HBM.  Pinned (or small) host memory goes straight to
// cudaMemcpyAsync; large pageable memory is moved through a ring of pinned buffers by a few host threads.
// host_to_dev: the source has been read when it returns.  dev_to_host: complete after the next
// synchronisation of the stream.
bool host_needs_staging(const void *host, size_t bytes);  // large and not pinned: the ring will carry it
int host_to_dev(void *dst_dev, const void *src_host, size_t bytes, cudaStream_t s);
int dev_to_host(void *dst_host, const void *src_dev, size_t bytes, cudaStream_t s);
// the same for `height` columns of `width` bytes (pitches in bytes), all columns in one pipeline
int host_to_dev_2d(void *dst_dev, size_t dpitch, const void *src_host, size_t spitch, size_t width, size_t height,
                   cudaStream_t s);
int dev_to_host_2d(void *dst_host, size_t dpitch, const void *src_dev, size_t spitch, size_t width, size_t height,
                   cudaStream_t s);

// slc_spmm_window.cu: after the current device's streams have been synchronised, SLC_ERR_CUDA (and the word
// cleared) if a TMA-staged block product gave up on a barrier there since the last check; `who` prefixes the message
int spmm_window_check(const char *who);

// built in slc_spmm_brick.cu: the TMA-staged SpMM for 27-offset matrices (*done = it ran).  X is indexed
// by column number; only its entries [xlo, xhi) exist (xhi < 0: all of [0, ncols)).
int spmm_brick_try(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                   const YInit<double> &yi, int32_t row_begin, int32_t row_end, int64_t xlo, int64_t xhi, bool *done);
inline int spmm_brick_try(slc_matrix *, const double2 *, int64_t, double2 *, int64_t, int32_t, const YInit<double2> &,
                          int32_t, int32_t, int64_t, int64_t, bool *done) {
  *done = false;  // Complex Double rows of 27 entries keep the panel kernel
  return SLC_OK;
}

// Builds the handle's offset-slot plan now if it has not been looked at yet (it reads a few words back from
// the device, i.e. synchronises the stream).  The row-partitioned products call it before they signal their
// peers: between the signal and the last boundary launch the host must never wait on the stream, or it
// would wait on a kernel whose peers this thread has not launched yet.
int spmm_brick_prepare(slc_matrix *m);

inline int bits_for(uint32_t n) {  // bits needed to represent values in [0, n)
  int b = 1;
  while (b < 32 && (1ull << b) < (uint64_t)n) ++b;
  return b;
}

}  // namespace slc
