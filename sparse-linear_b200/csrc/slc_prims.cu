// slc_prims.cu -- device-wide building blocks, hand-written: exclusive prefix sums and the stable
// LSD radix sort of (key, value) pairs that transpose and triplet assembly are built on.
//
// Radix sort, 8 bits per pass, three kernels per pass:
//   k_radix_hist     per-tile digit histograms, stored digit-major  hist[digit][tile]
//   exclusive scan   over that whole table: entry (d, t) becomes the first output position of the
//                    keys of tile t with digit d (all smaller digits, then earlier tiles, come first)
//   k_radix_scatter  re-reads the tile; a warp owns a contiguous run of keys and ranks them in index
//                    order with __match_any_sync + per-warp digit counters in shared memory, so
//                    equal digits keep their input order (stable); the tile is reordered by digit
//                    in shared memory and then streamed to base[d] + position-in-run
// Only the bits the caller says are significant are sorted (a row index of a 2M-row matrix needs
// 21 bits = 3 passes, not 4).  All traffic is coalesced on the read side; on the write side a tile
// of 8192 keys lands in at most 256 contiguous runs.
#include "slc_common.cuh"

namespace slc {

// ---- exclusive scan --------------------------------------------------------------------------------

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <class U>
__device__ __forceinline__ U block_sum(U v, U *warp_buf) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) warp_buf[warp] = v;
  __syncthreads();
  U total = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) total += warp_buf[w];
  return total;
}

template <class U>
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_reduce(const U *__restrict__ in, U *__restrict__ tile_sums, int64_t n) {
  __shared__ U warp_buf[SCAN_THREADS / 32];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  U s = 0;
#pragma unroll
  for (int it = 0; it < SCAN_ITEMS; ++it) {
    const int64_t i = base + (int64_t)it * SCAN_THREADS + threadIdx.x;
    if (i < n) s += in[i];
  }
  const U total = block_sum<U>(s, warp_buf);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// out[i] = tile_offset + sum of in[tile start .. i); each thread owns SCAN_ITEMS consecutive items
// and reads all of them before writing, so out may alias in.
template <class U>
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_apply(const U *in, U *out, const U *__restrict__ tile_offsets, int64_t n) {
  __shared__ U warp_buf[SCAN_THREADS / 32];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  U v[SCAN_ITEMS];
  U s = 0;
#pragma unroll
  for (int it = 0; it < SCAN_ITEMS; ++it) {
    v[it] = (base + it < n) ? in[base + it] : (U)0;
    s += v[it];
  }
  // exclusive scan of the per-thread sums: inside the warp, then across warps
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  U incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const U t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_buf[warp] = incl;
  __syncthreads();
  U before = tile_offsets ? tile_offsets[blockIdx.x] : (U)0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w)
    if (w < warp) before += warp_buf[w];
  U run = before + (incl - s);
#pragma unroll
  for (int it = 0; it < SCAN_ITEMS; ++it) {
    if (base + it < n) out[base + it] = run;
    run += v[it];
  }
}

template <class U>
static int exclusive_scan_impl(const U *in, U *out, int64_t n) {
  if (n <= 0) return SLC_OK;
  cudaStream_t s = cur_stream();
  const int64_t ntiles = ceil_div(n, SCAN_TILE);
  if (ntiles == 1) {
    k_scan_apply<U><<<1, SCAN_THREADS, 0, s>>>(in, out, nullptr, n);
    SLC_LAUNCHED();
    return SLC_OK;
  }
  DevBuf<U> sums;
  SLC_TRY(sums.alloc((size_t)ntiles));
  k_scan_reduce<U><<<(unsigned)ntiles, SCAN_THREADS, 0, s>>>(in, sums.p, n);
  SLC_LAUNCHED();
  SLC_TRY(exclusive_scan_impl<U>(sums.p, sums.p, ntiles));  // in place, one level up
  k_scan_apply<U><<<(unsigned)ntiles, SCAN_THREADS, 0, s>>>(in, out, sums.p, n);
  SLC_LAUNCHED();
  return SLC_OK;
}

int exclusive_scan_i32(const int32_t *in, int32_t *out, int64_t n) {
  return exclusive_scan_impl<int32_t>(in, out, n);
}
int exclusive_scan_i64(const int64_t *in, int64_t *out, int64_t n) {
  return exclusive_scan_impl<int64_t>(in, out, n);
}

// ---- radix sort -------------------------------------------------------------------------------------

constexpr int RS_THREADS = 512;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 8192 keys per block
constexpr int RS_BINS = 256;

template <class K>
__global__ void __launch_bounds__(RS_THREADS)
k_radix_hist(const K *__restrict__ keys, int64_t n, int shift, int64_t ntiles, int32_t *__restrict__ hist) {
  __shared__ int32_t cnt[RS_BINS];
  if (threadIdx.x < RS_BINS) cnt[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * RS_TILE;
#pragma unroll
  for (int it = 0; it < RS_ITEMS; ++it) {
    const int64_t i = base + (int64_t)it * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&cnt[(int)((keys[i] >> shift) & (RS_BINS - 1))], 1);
  }
  __syncthreads();
  if (threadIdx.x < RS_BINS) hist[(int64_t)threadIdx.x * ntiles + blockIdx.x] = cnt[threadIdx.x];
}

// Shared memory of k_radix_scatter: the tile's keys and values in digit order, the per-warp digit
// counters, and two tables of 256.
template <class K>
constexpr size_t radix_scatter_smem() {
  return (size_t)RS_TILE * (sizeof(K) + sizeof(uint32_t)) + (size_t)(RS_WARPS + 2) * RS_BINS * sizeof(int32_t) +
         8 * sizeof(int32_t);
}

// Pass 1 of 2 inside the tile: every key gets its position in the tile's digit order -- keys of a
// smaller digit first, then earlier warps, then index order inside the warp (__match_any_sync rank),
// which keeps the sort stable -- and is parked there in shared memory.  Pass 2 streams the parked
// tile out: neighbouring threads hold neighbouring keys of the same digit, whose global positions
// are consecutive, so the writes are coalesced runs instead of one sector per key.
template <class K>
__global__ void __launch_bounds__(RS_THREADS)
k_radix_scatter(const K *__restrict__ keys_in, const uint32_t *__restrict__ vals_in, K *__restrict__ keys_out,
                uint32_t *__restrict__ vals_out, int64_t n, int shift, int64_t ntiles,
                const int32_t *__restrict__ hist_scanned) {
  extern __shared__ __align__(16) unsigned char rs_smem[];
  K *skeys = reinterpret_cast<K *>(rs_smem);                         // [RS_TILE]
  uint32_t *svals = reinterpret_cast<uint32_t *>(skeys + RS_TILE);   // [RS_TILE]
  int32_t *wcnt = reinterpret_cast<int32_t *>(svals + RS_TILE);      // [RS_WARPS][RS_BINS]
  int32_t *local_base = wcnt + RS_WARPS * RS_BINS;                   // [RS_BINS] start of digit d in the tile
  int32_t *gshift = local_base + RS_BINS;                            // [RS_BINS] global start - local start
  int32_t *wsum = gshift + RS_BINS;                                  // [8]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < RS_WARPS * RS_BINS; i += RS_THREADS) wcnt[i] = 0;
  int32_t tile_base = 0;
  if (tid < RS_BINS) tile_base = hist_scanned[(int64_t)tid * ntiles + blockIdx.x];
  __syncthreads();

  // warp `warp` owns keys [cb, cb + 32*RS_ITEMS), taken 32 at a time in index order
  const int64_t tb = (int64_t)blockIdx.x * RS_TILE;
  const int64_t cb = tb + (int64_t)warp * (32 * RS_ITEMS);
  const unsigned lt = (1u << lane) - 1u;
  K key[RS_ITEMS];
  uint32_t val[RS_ITEMS];
  int32_t rank[RS_ITEMS];
#pragma unroll
  for (int it = 0; it < RS_ITEMS; ++it) {
    const int64_t i = cb + it * 32 + lane;
    const bool valid = i < n;
    key[it] = valid ? keys_in[i] : (K)0;
    val[it] = valid ? vals_in[i] : 0u;
  }
  int32_t *mycnt = wcnt + warp * RS_BINS;
#pragma unroll
  for (int it = 0; it < RS_ITEMS; ++it) {
    const bool valid = cb + it * 32 + lane < n;
    const int d = (int)((key[it] >> shift) & (RS_BINS - 1));
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? d : RS_BINS);
    rank[it] = 0;
    if (valid) {
      const int leader = __ffs(peers) - 1;
      int32_t b = 0;
      if (lane == leader) {
        b = mycnt[d];
        mycnt[d] = b + __popc(peers);
      }
      b = __shfl_sync(peers, b, leader);
      rank[it] = b + __popc(peers & lt);
    }
    __syncwarp();
  }
  __syncthreads();
  int32_t total = 0, incl = 0;
  if (tid < RS_BINS) {  // counts per warp -> offset of each warp inside the tile's run of digit `tid`
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
      const int32_t c = wcnt[w * RS_BINS + tid];
      wcnt[w * RS_BINS + tid] = total;
      total += c;
    }
    incl = total;  // exclusive scan of the digit totals over the 256 threads
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
  }
  __syncthreads();
  if (tid < RS_BINS) {
    int32_t before = 0;
#pragma unroll
    for (int w = 0; w < RS_BINS / 32; ++w)
      if (w < warp) before += wsum[w];
    const int32_t lb = before + incl - total;
    local_base[tid] = lb;
    gshift[tid] = tile_base - lb;
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < RS_ITEMS; ++it) {
    if (cb + it * 32 + lane < n) {
      const int d = (int)((key[it] >> shift) & (RS_BINS - 1));
      const int32_t lp = local_base[d] + mycnt[d] + rank[it];
      skeys[lp] = key[it];
      svals[lp] = val[it];
    }
  }
  __syncthreads();
  const int32_t count = (int32_t)min((int64_t)RS_TILE, n - tb);
  for (int32_t i = tid; i < count; i += RS_THREADS) {
    const K kk = skeys[i];
    const int d = (int)((kk >> shift) & (RS_BINS - 1));
    const int64_t pos = (int64_t)gshift[d] + i;
    keys_out[pos] = kk;
    vals_out[pos] = svals[i];
  }
}

template <class K>
static int sort_pairs_impl(K *keys_in, K *keys_out, uint32_t *vals_in, uint32_t *vals_out, int64_t n,
                           int end_bit) {
  if (n <= 0) return SLC_OK;
  if (n > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "radix sort: more than 2^31-1 items");
  cudaStream_t s = cur_stream();
  const int max_bits = (int)sizeof(K) * 8;
  if (end_bit < 1) end_bit = 1;
  if (end_bit > max_bits) end_bit = max_bits;
  const int passes = (end_bit + 7) / 8;
  const int64_t ntiles = ceil_div(n, RS_TILE);
  DevBuf<int32_t> hist;
  SLC_TRY(hist.alloc((size_t)ntiles * RS_BINS));
  SLC_CUDA(cudaFuncSetAttribute(k_radix_scatter<K>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)radix_scatter_smem<K>()));
  K *kin = keys_in, *kout = keys_out;
  uint32_t *vin = vals_in, *vout = vals_out;
  for (int p = 0; p < passes; ++p) {
    const int shift = p * 8;
    k_radix_hist<K><<<(unsigned)ntiles, RS_THREADS, 0, s>>>(kin, n, shift, ntiles, hist.p);
    SLC_LAUNCHED();
    SLC_TRY(exclusive_scan_i32(hist.p, hist.p, ntiles * RS_BINS));
    k_radix_scatter<K><<<(unsigned)ntiles, RS_THREADS, radix_scatter_smem<K>(), s>>>(kin, vin, kout, vout, n, shift,
                                                                                    ntiles, hist.p);
    SLC_LAUNCHED();
    K *tk = kin; kin = kout; kout = tk;
    uint32_t *tv = vin; vin = vout; vout = tv;
  }
  // after the swap the sorted data is in (kin, vin); the caller wants it in (keys_out, vals_out)
  if (kin != keys_out) {
    SLC_CUDA(cudaMemcpyAsync(keys_out, kin, (size_t)n * sizeof(K), cudaMemcpyDeviceToDevice, s));
    SLC_CUDA(cudaMemcpyAsync(vals_out, vin, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
  }
  return SLC_OK;
}

// Both sorts overwrite their input buffers (they are the ping-pong scratch).
int sort_pairs_u32(uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in, uint32_t *vals_out,
                   int64_t n, int end_bit) {
  return sort_pairs_impl<uint32_t>(keys_in, keys_out, vals_in, vals_out, n, end_bit);
}
int sort_pairs_u64(uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in, uint32_t *vals_out,
                   int64_t n, int end_bit) {
  return sort_pairs_impl<uint64_t>(keys_in, keys_out, vals_in, vals_out, n, end_bit);
}

}  // namespace slc
