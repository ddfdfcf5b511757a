// slc_noalloc.cu -- slc_coo_bucket_noalloc_dev: the triplet bucketing of slc_*_coo_bucket_dev for callers
// that must not allocate device memory.
//
// Who cannot allocate.  slc_*_coo_bucket_dev takes its sort buffers from the device's memory pool
// (cudaMallocAsync) on every call.  That is unusable for a host that (a) runs inside a fixed memory budget
// it has already handed out in full -- a solver that sized its own arena at start-up and leaves the driver
// nothing -- (b) drives the GPU from a thread where a driver allocation is forbidden or may block for an
// unbounded time, e.g. a Haskell `unsafe` foreign call that stops the runtime's garbage collector while
// it runs, or a latency-bound service loop, or (c) shares the device with another runtime whose allocator
// owns all free memory.  For those callers this file keeps the work area INSIDE the library image: a
// file-scope __device__ array that exists on every GPU from the moment the library is loaded there until
// the process exits.  The entry point below calls no cudaMalloc / cudaMallocAsync / cudaFree and no library
// routine that does; it launches three kernels, one 128-byte clear and one small device-to-host copy.
//
// What the work area costs and bounds.  NA_SCRATCH_BYTES (about 0.5 MiB) per GPU, always; in exchange one
// call handles at most SLC_NOALLOC_MAX_TRIPLETS (2^24) triplets and at most 16 parts.  More is rejected
// with SLC_ERR_RANGE -- use slc_*_coo_bucket_dev there.
//
// Sharing.  There is ONE work area per GPU, used by every call from every thread and stream, and it keeps
// whatever the previous call left.  The only cells that must start from zero (the per-part totals, which
// blocks add to with atomics) are cleared at the start of each call on the call's stream; the per-tile
// counters are overwritten before they are read.  Two calls whose kernels may be on the same GPU at the
// same time (different streams, or different host threads) would race on the work area: the caller must
// order them.  Calls on one stream are ordered by the stream.
//
// Set-up.  In host code the name of a __device__ array is only a stand-in for the symbol, not an address
// the GPU can use, so the first call on each device resolves the real address with cudaGetSymbolAddress
// (with that device current) and remembers it per device.  Make that first call outside any stream
// capture.
//
// Algorithm (no counterpart in the reference; same result as coo_bucket_impl in slc_shardops.cu, which
// sorts (owner, position) pairs): tiles of NA_TILE consecutive triplets,
//   k_na_tile_counts    per tile, how many triplets each part owns -> counter[tile][part]; totals[part] += ...
//   k_na_tile_offsets   one block, a warp per part: counter[tile][part] <- output position of the tile's
//                       first triplet of that part (parts laid out one after the other, tiles in order)
//   k_na_scatter        per tile, each triplet's rank among the tile's earlier triplets of the same part
//                       (warp match + per-warp counts), so input order is kept inside a part; rows, rebased
//                       columns and values go straight to their final position
#include "slc_common.cuh"

namespace slc {

constexpr int NA_PARTS = 16;
constexpr int NA_THREADS = 256;
constexpr int NA_WARPS = NA_THREADS / 32;
constexpr int NA_ROUNDS = 8;
constexpr int NA_TILE = NA_THREADS * NA_ROUNDS;                        // 2048 triplets per block
constexpr int64_t NA_MAX_TRIPLETS = (int64_t)1 << 24;                  // 16,777,216
constexpr int64_t NA_MAX_TILES = NA_MAX_TRIPLETS / NA_TILE;            // 8192
constexpr size_t NA_TOTALS_BYTES = NA_PARTS * sizeof(unsigned long long);  // 128
constexpr size_t NA_COUNTER_OFFSET = 256;                              // counters start on their own 256-byte line
constexpr size_t NA_SCRATCH_BYTES = NA_COUNTER_OFFSET + (size_t)NA_MAX_TILES * NA_PARTS * sizeof(uint32_t);  // 524,544

// The work area.  Host code below never passes this name to a kernel or a copy: it passes what
// cudaGetSymbolAddress returned for it on the current device (na_scratch()).
static __device__ __align__(256) unsigned char g_na_scratch[NA_SCRATCH_BYTES];

struct NaSplits {
  int32_t s[NA_PARTS + 1];
  int32_t n;
};

// the part p with s[p] <= c < s[p+1]; a column outside [0, s[n]) goes to the nearest part (the same rule as
// owner_of_col in slc_shardops.cu: the owner's compress then reports it, Sparse.hs:192-212)
__device__ __forceinline__ int na_owner(const NaSplits &sp, int32_t c) {
  int p = 0;
  while (p + 1 < sp.n && c >= sp.s[p + 1]) ++p;
  return p;
}

__global__ void __launch_bounds__(NA_THREADS)
k_na_tile_counts(int64_t n, const int32_t *__restrict__ cols, NaSplits sp, uint32_t *__restrict__ counter,
                 unsigned long long *__restrict__ totals) {
  __shared__ uint32_t hist[NA_PARTS];
  if (threadIdx.x < NA_PARTS) hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t tile0 = (int64_t)blockIdx.x * NA_TILE;
#pragma unroll
  for (int r = 0; r < NA_ROUNDS; ++r) {
    const int64_t i = tile0 + r * NA_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&hist[na_owner(sp, cols[i])], 1u);
  }
  __syncthreads();
  if (threadIdx.x < NA_PARTS) {
    const uint32_t h = hist[threadIdx.x];
    counter[(int64_t)blockIdx.x * NA_PARTS + threadIdx.x] = h;
    if (h) atomicAdd(totals + threadIdx.x, (unsigned long long)h);
  }
}

__global__ void __launch_bounds__(NA_PARTS * 32)
k_na_tile_offsets(int64_t ntiles, const unsigned long long *__restrict__ totals, uint32_t *__restrict__ counter) {
  const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned long long base = 0;
  for (int q = 0; q < p; ++q) base += totals[q];
  uint32_t running = (uint32_t)base;  // at most 2^24 positions
  for (int64_t t0 = 0; t0 < ntiles; t0 += 32) {
    const int64_t t = t0 + lane;
    const uint32_t c = t < ntiles ? counter[t * NA_PARTS + p] : 0u;
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    if (t < ntiles) counter[t * NA_PARTS + p] = running + incl - c;
    running += __shfl_sync(0xffffffffu, incl, 31);
  }
}

template <class T>
__global__ void __launch_bounds__(NA_THREADS)
k_na_scatter(int64_t n, const int32_t *__restrict__ rows, const int32_t *__restrict__ cols, const T *__restrict__ vals,
             NaSplits sp, const uint32_t *__restrict__ counter, int32_t *__restrict__ orows, int32_t *__restrict__ ocols,
             T *__restrict__ ovals) {
  __shared__ uint32_t run[NA_PARTS];            // next output position of every part, for this tile
  __shared__ uint32_t wcnt[NA_WARPS][NA_PARTS];  // this round: triplets of part p held by warp w
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid < NA_PARTS) run[tid] = counter[(int64_t)blockIdx.x * NA_PARTS + tid];
  const int64_t tile0 = (int64_t)blockIdx.x * NA_TILE;
  for (int r = 0; r < NA_ROUNDS; ++r) {
    if (tid < NA_WARPS * NA_PARTS) (&wcnt[0][0])[tid] = 0;
    __syncthreads();
    const int64_t i = tile0 + r * NA_THREADS + tid;
    const bool live = i < n;
    int32_t c = 0;
    int o = NA_PARTS;  // lanes past the end form their own group
    if (live) {
      c = cols[i];
      o = na_owner(sp, c);
    }
    const unsigned same = __match_any_sync(0xffffffffu, o);
    const int rank = __popc(same & ((1u << lane) - 1u));
    if (live && rank == 0) wcnt[warp][o] = (uint32_t)__popc(same);
    __syncthreads();
    if (live) {
      uint32_t dst = run[o] + (uint32_t)rank;
      for (int w = 0; w < warp; ++w) dst += wcnt[w][o];
      orows[dst] = rows[i];
      ocols[dst] = c - sp.s[o];
      ovals[dst] = vals[i];
    }
    __syncthreads();
    if (tid < NA_PARTS) {
      uint32_t s = 0;
#pragma unroll
      for (int w = 0; w < NA_WARPS; ++w) s += wcnt[w][tid];
      run[tid] += s;
    }
    __syncthreads();
  }
}

// Device address of the work area on the calling thread's current device, resolved on first use there.
static int na_scratch(unsigned char **out) {
  static std::mutex mu;
  static unsigned char *addr[MAX_DEVICES] = {};
  int dev = 0;
  SLC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) return fail(SLC_ERR_INVALID, "coo_bucket_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    unsigned char *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_na_scratch);
    if (e != cudaSuccess || !p) {
      (void)cudaGetLastError();
      return fail(SLC_ERR_CUDA, "coo_bucket_noalloc: cannot resolve the built-in work area on device %d: %s", dev,
                  cudaGetErrorString(e));
    }
    addr[dev] = p;
  }
  *out = addr[dev];
  return SLC_OK;
}

template <class T>
static int na_bucket(int64_t n, const int32_t *rows, const int32_t *cols, const void *vals, const NaSplits &sp,
                     int32_t *orows, int32_t *ocols, void *ovals, int64_t *counts) {
  unsigned char *scratch = nullptr;
  SLC_TRY(na_scratch(&scratch));
  unsigned long long *totals = reinterpret_cast<unsigned long long *>(scratch);
  uint32_t *counter = reinterpret_cast<uint32_t *>(scratch + NA_COUNTER_OFFSET);
  cudaStream_t s = cur_stream();
  const int64_t ntiles = ceil_div(n, NA_TILE);
  SLC_CUDA(cudaMemsetAsync(totals, 0, NA_TOTALS_BYTES, s));  // the previous call's totals are still there
  k_na_tile_counts<<<(unsigned)ntiles, NA_THREADS, 0, s>>>(n, cols, sp, counter, totals);
  SLC_LAUNCHED();
  k_na_tile_offsets<<<1, NA_PARTS * 32, 0, s>>>(ntiles, totals, counter);
  SLC_LAUNCHED();
  k_na_scatter<T><<<(unsigned)ntiles, NA_THREADS, 0, s>>>(n, rows, cols, static_cast<const T *>(vals), sp, counter, orows,
                                                          ocols, static_cast<T *>(ovals));
  SLC_LAUNCHED();
  unsigned long long h[NA_PARTS];
  SLC_CUDA(cudaMemcpyAsync(h, totals, NA_TOTALS_BYTES, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  for (int p = 0; p < sp.n; ++p) counts[p] = (int64_t)h[p];
  return SLC_OK;
}

}  // namespace slc

using namespace slc;

extern "C" int slc_coo_bucket_noalloc_dev(int elem, int64_t n, const int32_t *rows_dev, const int32_t *cols_dev,
                                          const void *vals_dev, int32_t nparts, const int32_t *col_splits,
                                          int32_t *out_rows_dev, int32_t *out_cols_dev, void *out_vals_dev,
                                          int64_t *counts) {
  if (elem != ELEM_D && elem != ELEM_Z) return fail(SLC_ERR_INVALID, "coo_bucket_noalloc: element type must be 'd' or 'z'");
  if (n < 0 || nparts < 1 || !col_splits || !counts) return fail(SLC_ERR_INVALID, "coo_bucket_noalloc: bad argument");
  if (nparts > NA_PARTS)
    return fail(SLC_ERR_RANGE, "coo_bucket_noalloc: %d parts, the built-in work area holds %d", nparts, NA_PARTS);
  if (n > NA_MAX_TRIPLETS)
    return fail(SLC_ERR_RANGE,
                "coo_bucket_noalloc: %lld triplets, the built-in work area holds %lld per call "
                "(use slc_*_coo_bucket_dev, which allocates)",
                (long long)n, (long long)NA_MAX_TRIPLETS);
  if (n > 0 && (!rows_dev || !cols_dev || !vals_dev || !out_rows_dev || !out_cols_dev || !out_vals_dev))
    return fail(SLC_ERR_INVALID, "coo_bucket_noalloc: null array");
  for (int p = 0; p < nparts; ++p) counts[p] = 0;
  if (n == 0) return SLC_OK;
  NaSplits sp;
  sp.n = nparts;
  for (int p = 0; p <= NA_PARTS; ++p) sp.s[p] = col_splits[p <= nparts ? p : nparts];
  if (elem == ELEM_D)
    return na_bucket<double>(n, rows_dev, cols_dev, vals_dev, sp, out_rows_dev, out_cols_dev, out_vals_dev, counts);
  return na_bucket<double2>(n, rows_dev, cols_dev, vals_dev, sp, out_rows_dev, out_cols_dev, out_vals_dev, counts);
}
