// slc_prims.cu -- device-wide building blocks, hand-written: exclusive prefix sums and the stable
// LSD radix sort of (key, value) pairs that transpose and triplet assembly are built on.
//
// Radix sort, 8 bits per pass, three kernels per pass:
//   k_radix_hist     per-tile digit histograms, stored digit-major  hist[digit][tile]
//   exclusive scan   over that whole table: entry (d, t) becomes the first output position of the
//                    keys of tile t with digit d (all smaller digits, then earlier tiles, come first)
//   k_radix_scatter  re-reads the tile; a warp owns a contiguous run of keys and ranks them in index
//                    order (eight ballots find the lanes with the same digit) against per-warp digit
//                    counters in shared memory, so equal digits keep their input order (stable); the
//                    tile is reordered by digit in shared memory and then streamed to
//                    base[d] + position-in-run
// Only the bits the caller says are significant are sorted (a row index of a 2M-row matrix needs
// 21 bits = 3 passes, not 4).  All traffic is coalesced on the read side; on the write side a tile
// of 4096 keys lands in at most 256 contiguous runs.  The first pass can read its keys from a
// caller's array that must stay intact (a matrix's row indices) and take the positions 0..n-1 as
// values without reading them; the result is handed back by pointer, never copied.
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

__global__ void k_max_i32(int64_t n, const int32_t *__restrict__ v, unsigned long long *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int32_t m = i < n ? v[i] : 0;
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, (unsigned long long)m);
}

__global__ void __launch_bounds__(256) k_sum_i32(int64_t n, const int32_t *__restrict__ v, unsigned long long *__restrict__ out) {
  unsigned long long acc = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    acc += (unsigned long long)(uint32_t)v[i];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

int sum_i32_into(const int32_t *v, int64_t n, unsigned long long *out) {
  if (n <= 0) return SLC_OK;
  k_sum_i32<<<(unsigned)std::min<int64_t>(ceil_div(n, 256 * 8), 1184), 256, 0, cur_stream()>>>(n, v, out);
  SLC_LAUNCHED();
  return SLC_OK;
}

int max_i32_into(const int32_t *v, int64_t n, unsigned long long *out) {
  if (n <= 0) return SLC_OK;
  k_max_i32<<<(unsigned)ceil_div(n, 256), 256, 0, cur_stream()>>>(n, v, out);
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

constexpr int RS_WARPS = RS_THREADS / 32;   // RS_THREADS, RS_ITEMS, RS_TILE, RS_BINS: slc_common.cuh

__device__ __forceinline__ uint32_t ld_once(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint64_t ld_once(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}

// All RS_ITEMS keys of a thread are loaded before the first shared-memory atomic (ncu, round 1: with
// the load inside the loop every atomic waited for its own load, 16 exposed latencies per thread).
template <class K>
__global__ void __launch_bounds__(RS_THREADS)
k_radix_hist(const K *__restrict__ keys, int64_t n, int shift, int64_t ntiles, int32_t *__restrict__ hist) {
  __shared__ int32_t cnt[RS_BINS];
  cnt[threadIdx.x] = 0;  // RS_THREADS == RS_BINS
  const int64_t base = (int64_t)blockIdx.x * RS_TILE;
  K key[RS_ITEMS];
  if (base + RS_TILE <= n) {
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) key[it] = ld_once(keys + base + it * RS_THREADS + threadIdx.x);
    __syncthreads();
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) atomicAdd(&cnt[(int)((key[it] >> shift) & (RS_BINS - 1))], 1);
  } else {
    __syncthreads();
    for (int it = 0; it < RS_ITEMS; ++it) {
      const int64_t i = base + (int64_t)it * RS_THREADS + threadIdx.x;
      if (i < n) atomicAdd(&cnt[(int)((keys[i] >> shift) & (RS_BINS - 1))], 1);
    }
  }
  __syncthreads();
  hist[(int64_t)threadIdx.x * ntiles + blockIdx.x] = cnt[threadIdx.x];
}

// Shared memory of k_radix_scatter: the tile's keys and values in digit order, the per-warp digit
// counters, and two tables of 256.
template <class K>
constexpr size_t radix_scatter_smem() {
  return (size_t)RS_TILE * (sizeof(K) + sizeof(uint32_t)) + (size_t)(RS_WARPS + 2) * RS_BINS * sizeof(int32_t) +
         RS_WARPS * sizeof(int32_t);
}

// lanes of the warp whose digit equals this lane's.  __match_any_sync does this in one instruction,
// but its cost grows with the number of distinct values in the warp (ncu, round 1: 64 % of the
// scatter kernel's stall samples sat behind MATCH with ~30 distinct digits per warp); eight ballots
// cost the same whatever the data.
__device__ __forceinline__ unsigned digit_peers(int d) {
  unsigned peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const bool bit = (d >> b) & 1;
    const unsigned m = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
}

// Pass 1 of 2 inside the tile: every key gets its position in the tile's digit order -- keys of a
// smaller digit first, then earlier warps, then index order inside the warp, which keeps the sort
// stable -- and is parked there in shared memory.  Pass 2 streams the parked tile out: neighbouring
// threads hold neighbouring keys of the same digit, whose global positions are consecutive, so the
// writes are coalesced runs instead of one sector per key.  IOTA: the values are the input positions
// themselves (first pass of a sort that carries positions) and are not read.
// PTRS (last pass of a 32-bit sort whose keys are bucket numbers): the pass also emits the bucket
// pointers, ptr[r] = first sorted position whose key is >= r.  Inside a digit run of the tile two
// neighbours of the parked tile are neighbours in the output, so a change of key between them is a
// bucket boundary at a known position; the first key of each run has its predecessor in another
// tile and is left to k_ptrs_run_starts.  This replaces a separate pass over the sorted keys.
// Bucket pointers of a run of EMPTY buckets: ptr[lo .. hi] = value.  Short runs are written by the thread that
// found them; a long run (a column block of a banded matrix has a million empty rows in a row: one thread
// took 3.6 ms over them in the sharded transpose) is queued for k_ptr_fill_gaps, which fills it with a block.
constexpr int PTR_GAP_INLINE = 64;
constexpr int PTR_GAP_QUEUE = 4096;
__device__ __forceinline__ void ptr_fill(int32_t *__restrict__ ptr, int64_t lo, int64_t hi, int32_t value,
                                         int4 *__restrict__ gaps, int *__restrict__ gap_count) {
  if (hi - lo >= PTR_GAP_INLINE && gaps) {
    const int g = atomicAdd(gap_count, 1);
    if (g < PTR_GAP_QUEUE) {
      gaps[g] = make_int4((int)lo, (int)hi, value, 0);
      return;
    }
  }
  for (int64_t r = lo; r <= hi; ++r) ptr[r] = value;
}

__global__ void __launch_bounds__(256)
k_ptr_fill_gaps(const int4 *__restrict__ gaps, const int *__restrict__ gap_count, int32_t *__restrict__ ptr) {
  const int n = min(*gap_count, PTR_GAP_QUEUE);
  for (int g = blockIdx.x; g < n; g += gridDim.x) {
    const int4 q = gaps[g];
    for (int64_t r = (int64_t)q.x + threadIdx.x; r <= q.y; r += blockDim.x) ptr[r] = q.z;
  }
}

template <class K, bool IOTA, bool PTRS>
__global__ void __launch_bounds__(RS_THREADS, 3)
k_radix_scatter(const K *__restrict__ keys_in, const uint32_t *__restrict__ vals_in, K *__restrict__ keys_out,
                uint32_t *__restrict__ vals_out, int64_t n, int shift, int64_t ntiles,
                const int32_t *__restrict__ hist_scanned, uint32_t zmask, int32_t *__restrict__ ptr_out,
                int4 *__restrict__ gaps, int *__restrict__ gap_count) {
  extern __shared__ __align__(16) unsigned char rs_smem[];
  K *skeys = reinterpret_cast<K *>(rs_smem);                         // [RS_TILE]
  uint32_t *svals = reinterpret_cast<uint32_t *>(skeys + RS_TILE);   // [RS_TILE]
  int32_t *wcnt = reinterpret_cast<int32_t *>(svals + RS_TILE);      // [RS_WARPS][RS_BINS]
  int32_t *local_base = wcnt + RS_WARPS * RS_BINS;                   // [RS_BINS] start of digit d in the tile
  int32_t *gshift = local_base + RS_BINS;                            // [RS_BINS] global start - local start
  int32_t *wsum = gshift + RS_BINS;                                  // [RS_WARPS]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) wcnt[w * RS_BINS + tid] = 0;
  const int32_t tile_base = hist_scanned[(int64_t)tid * ntiles + blockIdx.x];  // digit `tid`

  // warp `warp` owns keys [cb, cb + 32*RS_ITEMS), taken 32 at a time in index order
  const int64_t tb = (int64_t)blockIdx.x * RS_TILE;
  const int64_t cb = tb + (int64_t)warp * (32 * RS_ITEMS);
  const bool full_tile = tb + RS_TILE <= n;
  const unsigned lt = (1u << lane) - 1u;
  K key[RS_ITEMS];
  int32_t rank[RS_ITEMS];
  if (full_tile) {
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) key[it] = ld_once(keys_in + cb + it * 32 + lane);
    // the values are only needed after the ranking: start them towards L2 now (one line per lane)
    if (!IOTA && lane < RS_ITEMS) asm volatile("prefetch.global.L2 [%0];" ::"l"(vals_in + cb + lane * 32));
  } else {
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) {
      const int64_t i = cb + it * 32 + lane;
      key[it] = i < n ? keys_in[i] : (K)0;
    }
  }
  __syncthreads();
  int32_t *mycnt = wcnt + warp * RS_BINS;
#pragma unroll
  for (int it = 0; it < RS_ITEMS; ++it) {
    const bool valid = full_tile || cb + it * 32 + lane < n;
    const int d = (int)((key[it] >> shift) & (RS_BINS - 1));
    unsigned peers = digit_peers(d);
    if (!full_tile) peers &= __ballot_sync(0xffffffffu, valid);
    rank[it] = 0;
    int32_t b = 0;
    const int leader = __ffs(peers) - 1;
    if (valid && lane == leader) {
      b = mycnt[d];
      mycnt[d] = b + __popc(peers);
    }
    b = __shfl_sync(0xffffffffu, b, leader & 31);
    rank[it] = b + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  // counts per warp -> offset of each warp inside the tile's run of digit `tid`
  int32_t total = 0;
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) {
    const int32_t c = wcnt[w * RS_BINS + tid];
    wcnt[w * RS_BINS + tid] = total;
    total += c;
  }
  int32_t incl = total;  // exclusive scan of the digit totals over the 256 threads
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  {
    int32_t before = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w)
      if (w < warp) before += wsum[w];
    const int32_t lb = before + incl - total;
    local_base[tid] = lb;
    gshift[tid] = tile_base - lb;
  }
  __syncthreads();
  if (full_tile) {
    uint32_t val[RS_ITEMS];
    if (!IOTA) {
#pragma unroll
      for (int it = 0; it < RS_ITEMS; ++it) val[it] = ld_once(vals_in + cb + it * 32 + lane);
      uint32_t r = 0;  // every load is issued before the first value is parked (see probe/tie)
#pragma unroll
      for (int it = 0; it < RS_ITEMS; ++it) r |= val[it];
      val[0] ^= r & zmask;
    }
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) {
      const int d = (int)((key[it] >> shift) & (RS_BINS - 1));
      const int32_t lp = local_base[d] + mycnt[d] + rank[it];
      skeys[lp] = key[it];
      svals[lp] = IOTA ? (uint32_t)(cb + it * 32 + lane) : val[it];
    }
  } else {
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) {
      const int64_t i = cb + it * 32 + lane;
      if (i < n) {
        const int d = (int)((key[it] >> shift) & (RS_BINS - 1));
        const int32_t lp = local_base[d] + mycnt[d] + rank[it];
        skeys[lp] = key[it];
        svals[lp] = IOTA ? (uint32_t)i : vals_in[i];
      }
    }
  }
  __syncthreads();
  const int32_t count = (int32_t)min((int64_t)RS_TILE, n - tb);
#pragma unroll 4
  for (int32_t i = tid; i < count; i += RS_THREADS) {
    const K kk = skeys[i];
    const int d = (int)((kk >> shift) & (RS_BINS - 1));
    const int64_t pos = (int64_t)gshift[d] + i;
    keys_out[pos] = kk;
    vals_out[pos] = svals[i];
    if constexpr (PTRS) {
      if (i > local_base[d]) {
        const K pk = skeys[i - 1];
        if (pk != kk) ptr_fill(ptr_out, (int64_t)pk + 1, (int64_t)kk, (int32_t)pos, gaps, gap_count);
      }
    }
  }
}

// The bucket boundaries k_radix_scatter<.., PTRS> could not see: the first key of every (digit, tile)
// run, whose predecessor was written by another tile, and the tail after the last key.  One thread
// per entry of the scanned histogram (digit-major, so the entries are in output order).
__global__ void __launch_bounds__(256)
k_ptrs_run_starts(int64_t n, int32_t nkeys, int64_t nentries, const int32_t *__restrict__ hist_scanned,
                  const uint32_t *__restrict__ sorted, int32_t *__restrict__ ptr, int4 *__restrict__ gaps,
                  int *__restrict__ gap_count) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nentries) return;
  const int64_t start = hist_scanned[j];
  const int64_t end = j + 1 < nentries ? (int64_t)hist_scanned[j + 1] : n;
  if (end <= start) return;
  const int64_t cur = sorted[start];
  const int64_t prev = start > 0 ? (int64_t)sorted[start - 1] : -1;
  if (prev != cur) ptr_fill(ptr, prev + 1, cur, (int32_t)start, gaps, gap_count);
  if (end == n) {
    const int64_t last = sorted[n - 1];
    ptr_fill(ptr, last + 1, nkeys, (int32_t)n, gaps, gap_count);
  }
}

// Stable sort of (key, value) pairs on key bits [0, end_bit).  (ka, va) and (kb, vb) are ping-pong
// scratch.  The first pass reads its keys from `src` when given (which is left untouched) and from ka
// otherwise; its values are va, or the positions 0..n-1 when iota_vals.  *res_k / *res_v say which
// scratch buffers hold the sorted pairs.
template <class K, bool IOTA, bool PTRS>
static void launch_scatter(const K *kin, const uint32_t *vin, K *kout, uint32_t *vout, int64_t n, int shift,
                           int64_t ntiles, const int32_t *hist, int32_t *ptr_out, cudaStream_t s,
                           int4 *gaps = nullptr, int *gap_count = nullptr) {
  k_radix_scatter<K, IOTA, PTRS><<<(unsigned)ntiles, RS_THREADS, radix_scatter_smem<K>(), s>>>(
      kin, vin, kout, vout, n, shift, ntiles, hist, 0u, ptr_out, gaps, gap_count);
}

// ptr_out (32-bit keys only): also produce ptr_out[0 .. nkeys], the first sorted position of every key
// value, fused into the last pass.
template <class K>
static int sort_pairs_impl(const K *src, K *ka, K *kb, uint32_t *va, uint32_t *vb, int64_t n, int end_bit,
                           bool iota_vals, K **res_k, uint32_t **res_v, int32_t *ptr_out = nullptr,
                           int32_t nkeys = 0) {
  *res_k = ka;
  *res_v = va;
  if (n <= 0) return SLC_OK;
  if (n > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "radix sort: more than 2^31-1 items");
  cudaStream_t s = cur_stream();
  const int max_bits = (int)sizeof(K) * 8;
  if (end_bit < 1) end_bit = 1;
  if (end_bit > max_bits) end_bit = max_bits;
  const int passes = (end_bit + 7) / 8;
  const int64_t ntiles = ceil_div(n, RS_TILE);
  DevBuf<int32_t> hist;
  DevBuf<int4> gaps;
  DevBuf<int> gap_count;
  SLC_TRY(hist.alloc((size_t)ntiles * RS_BINS));
  static PerDevice<size_t> smem_a, smem_b, smem_c, smem_d;  // per instantiation (K), per device
  SLC_CUDA(ensure_dyn_smem(k_radix_scatter<K, false, false>, smem_a, radix_scatter_smem<K>()));
  SLC_CUDA(ensure_dyn_smem(k_radix_scatter<K, true, false>, smem_b, radix_scatter_smem<K>()));
  constexpr bool CAN_PTRS = sizeof(K) == 4;
  if constexpr (CAN_PTRS) {
    SLC_CUDA(ensure_dyn_smem(k_radix_scatter<K, false, true>, smem_c, radix_scatter_smem<K>()));
    SLC_CUDA(ensure_dyn_smem(k_radix_scatter<K, true, true>, smem_d, radix_scatter_smem<K>()));
  } else if (ptr_out) {
    return fail(SLC_ERR_INVALID, "radix sort: bucket pointers need 32-bit keys");
  }
  const K *kin = src ? src : ka;
  const uint32_t *vin = va;
  K *kout = src ? ka : kb;
  uint32_t *vout = src ? va : vb;
  if (src && !iota_vals) return fail(SLC_ERR_INVALID, "radix sort: a separate key source needs positional values");
  for (int p = 0; p < passes; ++p) {
    const int shift = p * 8;
    k_radix_hist<K><<<(unsigned)ntiles, RS_THREADS, 0, s>>>(kin, n, shift, ntiles, hist.p);
    SLC_LAUNCHED();
    SLC_TRY(exclusive_scan_i32(hist.p, hist.p, ntiles * RS_BINS));
    const bool iota = p == 0 && iota_vals;
    const bool ptrs = ptr_out != nullptr && p == passes - 1;
    if constexpr (CAN_PTRS) {
      if (ptrs) {  // queue for the long runs of empty buckets (filled by k_ptr_fill_gaps below)
        SLC_TRY(gaps.alloc(PTR_GAP_QUEUE));
        SLC_TRY(gap_count.alloc(1));
        SLC_CUDA(cudaMemsetAsync(gap_count.p, 0, sizeof(int), s));
      }
      if (ptrs && iota)
        launch_scatter<K, true, true>(kin, nullptr, kout, vout, n, shift, ntiles, hist.p, ptr_out, s, gaps.p, gap_count.p);
      else if (ptrs)
        launch_scatter<K, false, true>(kin, vin, kout, vout, n, shift, ntiles, hist.p, ptr_out, s, gaps.p, gap_count.p);
    }
    if (!ptrs) {
      if (iota) launch_scatter<K, true, false>(kin, nullptr, kout, vout, n, shift, ntiles, hist.p, nullptr, s);
      else launch_scatter<K, false, false>(kin, vin, kout, vout, n, shift, ntiles, hist.p, nullptr, s);
    }
    SLC_LAUNCHED();
    if constexpr (CAN_PTRS) {
      if (ptrs) {
        const int64_t nentries = ntiles * RS_BINS;
        k_ptrs_run_starts<<<(unsigned)ceil_div(nentries, 256), 256, 0, s>>>(
            n, nkeys, nentries, hist.p, reinterpret_cast<const uint32_t *>(kout), ptr_out, gaps.p, gap_count.p);
        SLC_LAUNCHED();
        k_ptr_fill_gaps<<<256, 256, 0, s>>>(gaps.p, gap_count.p, ptr_out);
        SLC_LAUNCHED();
      }
    }
    *res_k = kout;
    *res_v = vout;
    kin = kout;
    vin = vout;
    kout = (kout == ka) ? kb : ka;
    vout = (vout == va) ? vb : va;
  }
  return SLC_OK;
}

// One histogram pass on its own: offsets[d * ntiles + t] = where the keys of tile t (RS_TILE consecutive
// keys) whose digit (key >> shift) & 255 is d start in an array partitioned by digit (tiles in order
// inside a digit).  Used by the coarse column partition of compress.
int tile_digit_offsets_u32(const uint32_t *keys, int64_t n, int shift, int32_t *offsets) {
  if (n <= 0) return SLC_OK;
  const int64_t ntiles = ceil_div(n, RS_TILE);
  k_radix_hist<uint32_t><<<(unsigned)ntiles, RS_THREADS, 0, cur_stream()>>>(keys, n, shift, ntiles, offsets);
  SLC_LAUNCHED();
  return exclusive_scan_i32(offsets, offsets, ntiles * RS_BINS);
}

int sort_pairs_u32(const uint32_t *src, uint32_t *ka, uint32_t *kb, uint32_t *va, uint32_t *vb, int64_t n,
                   int end_bit, bool iota_vals, uint32_t **res_k, uint32_t **res_v) {
  return sort_pairs_impl<uint32_t>(src, ka, kb, va, vb, n, end_bit, iota_vals, res_k, res_v);
}
int sort_buckets_u32(const uint32_t *src, uint32_t *ka, uint32_t *kb, uint32_t *va, uint32_t *vb, int64_t n,
                     int32_t nkeys, uint32_t **res_k, uint32_t **res_v, int32_t *ptr_out) {
  return sort_pairs_impl<uint32_t>(src, ka, kb, va, vb, n, bits_for((uint32_t)nkeys), true, res_k, res_v, ptr_out,
                                   nkeys);
}
int sort_pairs_u64(const uint64_t *src, uint64_t *ka, uint64_t *kb, uint32_t *va, uint32_t *vb, int64_t n,
                   int end_bit, bool iota_vals, uint64_t **res_k, uint32_t **res_v) {
  return sort_pairs_impl<uint64_t>(src, ka, kb, va, vb, n, end_bit, iota_vals, res_k, res_v);
}

}  // namespace slc
