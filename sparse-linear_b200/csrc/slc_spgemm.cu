// slc_spgemm.cu -- SpGEMM, C = A B, both CSC.
//
// Reference: mm (Data/Matrix/Sparse.hs:691-702) over the scatter/gather accumulator
// (Data/Vector/Sparse/ScatterGather.hs:37-147): for each column j of B, reset the workspace,
// for (k, b) in B[:, j] in order scatter A[:, k] with c <- c + a*b, then gather in row order.
// What the result must keep from that:
//   * structural product: a row is present as soon as one term landed on it; zeros stay;
//   * rows strictly increasing inside a column;
//   * the terms of one entry are added in increasing k, starting from 0.
//
// Device formulation (CSC Gustavson, column j of C owned by one thread group):
//   pass 0  products per column (upper bound on the column's length) -> size class
//   pass 1  symbolic: rows hashed into a shared-memory table sized by the class -> exact counts
//   scan    counts -> C.colptr
//   pass 2  numeric: same table plus a value per slot.  The group walks B[:, j] sequentially and
//           spreads A[:, k] over its lanes: rows inside one A column are distinct, so slots are
//           touched by one lane per step and steps are separated by a group barrier -- every
//           entry receives its terms in increasing k, exactly like the reference.  The occupied
//           slots are then compacted, bitonic-sorted by row, and written out.
//   Columns too long for shared memory use a per-block dense workspace in HBM (the reference's
//   own accumulator, with a stamp instead of a full reset and a scan limited to [min,max] row).
//   Columns of B with few entries (<= 8: every stencil, most of a power-law matrix) skip the hash
//   table altogether: ONE THREAD merges the <= 8 sorted columns of A that the column selects
//   (k_mm_merge).  The merge emits rows in increasing order and adds the terms of a row in list
//   order = increasing k, so structure and values are again the reference's, with no shared
//   memory, no atomics and no sort.
#include <cstdlib>

#include "slc_common.cuh"

namespace slc {

// classes: 0/1 thread-per-column merge (<= 4 / <= 8 entries in B[:, j], <= MERGE_UB products);
//          2: <= 64 products, 3: <= 512, 4: <= 2048 (hash tables of twice the bound); 5: dense fallback
constexpr int NCLASS = 6;
constexpr int CLS_MERGE4 = 0, CLS_MERGE8 = 1, CLS_HASH0 = 2, CLS_HASH1 = 3, CLS_HASH2 = 4, CLS_DENSE = 5;
constexpr int CLASS_UB0 = 64, CLASS_UB1 = 512, CLASS_UB2 = 2048;  // table = 2 * bound
constexpr int MERGE_UB = 256;

// ---- pass 0: products per column, classification ------------------------------------------------

// Columns are bucketed by class with a STABLE split (block histograms, one scan, ranks inside the
// block): inside a class the columns keep their ascending order, so the 32 columns a warp of the
// product kernels takes are consecutive wherever the class does not change -- which is what lets
// k_mm_merge stream a warp's slice of C out as one contiguous range.  Bin NCLASS collects the
// columns with no product at all.
constexpr int MM_BINS = NCLASS + 1;
constexpr int MM_SPLIT_THREADS = 256;

__device__ __forceinline__ int mm_class_of(int32_t nb, int64_t total, int merge_ok) {
  if (total == 0) return NCLASS;
  if (merge_ok && nb <= 4 && total <= MERGE_UB) return CLS_MERGE4;
  if (merge_ok && nb <= 8 && total <= MERGE_UB) return CLS_MERGE8;
  if (total <= CLASS_UB0) return CLS_HASH0;
  if (total <= CLASS_UB1) return CLS_HASH1;
  if (total <= CLASS_UB2) return CLS_HASH2;
  return CLS_DENSE;
}

// rank of this thread among the threads of its block that hold the same bin, and the block's count
// per bin (valid in counts[] after the call for threadIdx.x < MM_BINS)
__device__ __forceinline__ int32_t mm_block_rank(int bin, int32_t (&wcount)[MM_SPLIT_THREADS / 32][MM_BINS]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int32_t rank = 0;
#pragma unroll
  for (int c = 0; c < MM_BINS; ++c) {
    const unsigned m = __ballot_sync(0xffffffffu, bin == c);
    if (lane == 0) wcount[warp][c] = __popc(m);
    if (bin == c) rank = __popc(m & ((1u << lane) - 1u));
  }
  __syncthreads();
  if (bin >= 0)
    for (int w = 0; w < warp; ++w) rank += wcount[w][bin];
  return rank;
}

__global__ void __launch_bounds__(MM_SPLIT_THREADS)
k_mm_upper(int32_t bc, int32_t nblocks, const int32_t *__restrict__ bp, const int32_t *__restrict__ bi,
           const int32_t *__restrict__ ap, int32_t *__restrict__ ub, int32_t *__restrict__ klass,
           int32_t *__restrict__ block_counts, int merge_ok) {
  __shared__ int32_t wcount[MM_SPLIT_THREADS / 32][MM_BINS];
  const int32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  int bin = -1;
  if (j < bc) {
    int64_t total = 0;
    const int32_t q0 = bp[j], q1 = bp[j + 1];
    for (int32_t q = q0; q < q1; ++q) {
      const int32_t k = bi[q];
      total += ap[k + 1] - ap[k];
    }
    ub[j] = total > 0x7fffffffLL ? 0x7fffffff : (int32_t)total;
    bin = mm_class_of(q1 - q0, total, merge_ok);
    klass[j] = bin;
  }
  (void)mm_block_rank(bin, wcount);
  if (threadIdx.x < MM_BINS) {  // bin-major table: one scan turns it into output offsets
    int32_t t = 0;
    for (int w = 0; w < MM_SPLIT_THREADS / 32; ++w) t += wcount[w][threadIdx.x];
    block_counts[(int64_t)threadIdx.x * nblocks + blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(MM_SPLIT_THREADS)
k_mm_bucket(int32_t bc, int32_t nblocks, const int32_t *__restrict__ klass,
            const int32_t *__restrict__ block_offsets, int32_t *__restrict__ cols) {
  __shared__ int32_t wcount[MM_SPLIT_THREADS / 32][MM_BINS];
  const int32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  const int bin = j < bc ? klass[j] : -1;
  const int32_t rank = mm_block_rank(bin, wcount);
  if (bin >= 0) cols[block_offsets[(int64_t)bin * nblocks + blockIdx.x] + rank] = j;
}

// ---- thread-per-column merge --------------------------------------------------------------------
// Column j of C is the union of the columns A[:, k], k in B[:, j]: NB sorted lists.  The thread
// keeps one cursor per list in registers (all loops are unrolled over NB), repeatedly takes the
// smallest head row m, and adds a*b over the lists whose head is m in list order -- increasing k,
// starting from 0, c + a*b (Sparse.hs:699) -- then advances those lists.  NUMERIC = false only
// counts the distinct rows (symbolic pass).
constexpr int32_t ROW_END = 0x7fffffff;  // past every valid row index

// Output staging (numeric pass).  A thread emits its column one (row, value) at a time, so written
// directly each store instruction of a warp touches 32 different sectors (launch list, round 1: the
// numeric merge took 7x the symbolic one).  When the 32 columns of a warp are consecutive -- the
// bucket order keeps them so except at class boundaries -- their slice of C is one contiguous range:
// the lanes park their entries in the warp's shared-memory window at (position in C - start of the
// range) and the warp then streams the window out with fully coalesced stores.
constexpr int MERGE_STAGE = 1024;  // default entries per warp window (12 or 20 KB per warp); the kernel takes it as `stage`

template <class T, int NB, bool NUMERIC, int MINB = 0>
__global__ void __launch_bounds__(128, MINB)
k_mm_merge(int32_t ncols_in_class, const int32_t *__restrict__ cols, const int32_t *__restrict__ ap,
           const int32_t *__restrict__ ai, const T *__restrict__ av, const int32_t *__restrict__ bp,
           const int32_t *__restrict__ bi, const T *__restrict__ bv, int32_t *__restrict__ cnt,
           const int32_t *__restrict__ cp, int32_t *__restrict__ ci, T *__restrict__ cv, int stage) {
  extern __shared__ __align__(16) unsigned char merge_smem[];
  // grid-stride over tiles of blockDim.x columns: a grid of a few blocks per SM (run_merge) keeps the
  // number of resident warps, and with it the L1 working set of A columns, bounded
  for (int32_t tile = blockIdx.x; (int64_t)tile * blockDim.x < ncols_in_class; tile += gridDim.x) {
  const int32_t slot = tile * blockDim.x + threadIdx.x;
  const bool active = slot < ncols_in_class;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int32_t j = active ? cols[slot] : -1;
  const int32_t q0 = active ? bp[j] : 0;
  const int32_t nb = active ? bp[j + 1] - q0 : 0;
  // Every list keeps its next two (row, value) pairs in registers: head/hval is the entry at p, nxt/nval
  // the one at p + 1.  Consuming an entry shifts nxt into head and issues the loads for p + 2, so the
  // load a merge step starts is not needed before the same list advances again -- the chain of a step
  // holds no L2 round trip (the launch list of round 1 had this kernel at 4x the symbolic pass with 16
  // warps per SM waiting on the head loads).
  int32_t p[NB], pe[NB], head[NB], nxt[NB];
  T b[NB], hval[NB], nval[NB];
#pragma unroll
  for (int u = 0; u < NB; ++u) {
    p[u] = 0;
    pe[u] = 0;
    head[u] = ROW_END;
    nxt[u] = ROW_END;
    b[u] = Num<T>::zero();
    hval[u] = Num<T>::zero();
    nval[u] = Num<T>::zero();
    if (u < nb) {
      const int32_t k = __ldg(bi + q0 + u);
      p[u] = __ldg(ap + k);
      pe[u] = __ldg(ap + k + 1);
      if (NUMERIC) b[u] = __ldg(bv + q0 + u);
      if (p[u] < pe[u]) {
        head[u] = __ldg(ai + p[u]);
        if (NUMERIC) hval[u] = __ldg(av + p[u]);
      }
      if (p[u] + 1 < pe[u]) {
        nxt[u] = __ldg(ai + p[u] + 1);
        if (NUMERIC) nval[u] = __ldg(av + p[u] + 1);
      }
    }
  }
  int32_t count = 0;
  int32_t *co = nullptr;
  T *vo = nullptr;
  bool staged = false;
  int32_t wbase = 0, wtotal = 0;
  if (NUMERIC) {
    const int32_t base = active ? cp[j] : 0;
    const int32_t bend = active ? cp[j + 1] : 0;
    co = ci + base;
    vo = cv + base;
    if (stage) {  // warp-uniform decision
      const unsigned live = __ballot_sync(0xffffffffu, active);
      const int32_t jfirst = __shfl_sync(0xffffffffu, j, 0);
      const bool consecutive = __all_sync(0xffffffffu, !active || j == jfirst + lane);
      const int last = 31 - __clz(live | 1u);
      wbase = __shfl_sync(0xffffffffu, base, 0);
      wtotal = __shfl_sync(0xffffffffu, bend, last) - wbase;
      staged = live != 0 && consecutive && wtotal <= stage;
      if (staged) {
        int32_t *s_i = reinterpret_cast<int32_t *>(merge_smem) + (size_t)warp * stage;
        T *s_v = reinterpret_cast<T *>(merge_smem + (size_t)(blockDim.x >> 5) * stage * sizeof(int32_t)) +
                 (size_t)warp * stage;
        co = s_i + (base - wbase);
        vo = s_v + (base - wbase);
      }
    }
  }
  for (;;) {
    int32_t m = head[0];
#pragma unroll
    for (int u = 1; u < NB; ++u) m = min(m, head[u]);
    if (m == ROW_END) break;
    T acc = Num<T>::zero();
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      if (head[u] == m) {
        if (NUMERIC) acc = Num<T>::add(acc, Num<T>::mul(hval[u], b[u]));
        ++p[u];
        head[u] = nxt[u];
        if (NUMERIC) hval[u] = nval[u];
        const bool more = p[u] + 1 < pe[u];
        nxt[u] = more ? __ldg(ai + p[u] + 1) : ROW_END;
        if (NUMERIC) nval[u] = more ? __ldg(av + p[u] + 1) : Num<T>::zero();
      }
    }
    if (NUMERIC) {
      co[count] = m;
      vo[count] = acc;
    }
    ++count;
  }
  if (!NUMERIC) {
    if (active) cnt[j] = count;
    continue;
  }
  if (staged) {
    __syncwarp();
    const int32_t *s_i = reinterpret_cast<const int32_t *>(merge_smem) + (size_t)warp * stage;
    const T *s_v = reinterpret_cast<const T *>(merge_smem + (size_t)(blockDim.x >> 5) * stage * sizeof(int32_t)) +
                   (size_t)warp * stage;
    for (int32_t e = lane; e < wtotal; e += 32) {
      ci[wbase + e] = s_i[e];
      cv[wbase + e] = s_v[e];
    }
    __syncwarp();  // the window is reused by the next tile
  }
  }
}

template <class T, int NB, bool NUMERIC>
static int run_merge(int32_t n, const int32_t *cols, const slc_matrix *a, const slc_matrix *b, int32_t *cnt,
                     slc_matrix *c, int fit_cap = 0) {
  if (n == 0) return SLC_OK;
  static const char *stage_env = getenv("SLC_MM_STAGE");  // SLC_MM_STAGE=0: direct (uncoalesced) stores
  // SLC_MM_STAGE_CAP: entries of a warp's output window.  Default: what the product needs (fit_cap = 32 x the
  // longest result column, known after the symbolic pass), at most MERGE_STAGE -- a window no larger than the
  // product leaves more of the SM's memory to L1 (cfg 3: 800 entries instead of 1024, 0.726 -> 0.708 ms)
  static const int cap_env = env_int("SLC_MM_STAGE_CAP", 0);
  static const int carve_env = env_int("SLC_MM_CARVEOUT", -1);  // preferred shared-memory carve-out, per cent
  const int cap_want = cap_env > 0 ? cap_env : (fit_cap > 0 ? std::min(fit_cap, (int)MERGE_STAGE) : (int)MERGE_STAGE);
  const int cap = std::max(32, std::min(cap_want, 2048)) / 4 * 4;
  const int stage = (NUMERIC && !(stage_env && stage_env[0] == '0')) ? cap : 0;
  const size_t smem = stage ? (size_t)(128 / 32) * stage * (sizeof(int32_t) + sizeof(T)) : 0;
  // SLC_MM_MINB=5/6: numeric 8-list merge for Double held to 96 / 80 registers (5 / 6 blocks per SM)
  static const int minb_env = env_int("SLC_MM_MINB", 0);
  auto kern = k_mm_merge<T, NB, NUMERIC>;
  if (NUMERIC && NB == 8 && sizeof(T) == 8 && minb_env == 5) kern = k_mm_merge<T, NB, NUMERIC, 5>;
  if (NUMERIC && NB == 8 && sizeof(T) == 8 && minb_env == 6) kern = k_mm_merge<T, NB, NUMERIC, 6>;
  if (smem > 48 * 1024) SLC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (NUMERIC && carve_env >= 0)
    SLC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve_env));
  // SLC_MM_RESIDENT=B: at most B blocks per SM (0 = one block per tile)
  static const int resident_env = env_int("SLC_MM_RESIDENT", 0);
  const int resident = NUMERIC ? resident_env : 0;
  const int64_t tiles = ceil_div(n, 128);
  const int64_t grid = resident > 0 ? std::min<int64_t>(tiles, (int64_t)resident * 148) : tiles;
  kern<<<(unsigned)grid, 128, smem, cur_stream()>>>(
      n, cols, a->colptr, a->rowidx, (const T *)a->vals, b->colptr, b->rowidx, (const T *)b->vals, cnt,
      c ? c->colptr : nullptr, c ? c->rowidx : nullptr, c ? (T *)c->vals : nullptr, stage);
  SLC_LAUNCHED();
  return SLC_OK;
}

// ---- group helpers ------------------------------------------------------------------------------

template <int G>
__device__ __forceinline__ void group_sync(unsigned mask) {
  if (G <= 32) __syncwarp(mask);
  else __syncthreads();
}

__device__ __forceinline__ uint32_t hash_row(int32_t i, int log2h) {
  return ((uint32_t)i * 0x9E3779B1u) >> (32 - log2h);
}

// find or claim the slot of row i; returns the slot, *fresh = claimed now
__device__ __forceinline__ int table_insert(int32_t *keys, int hmask, int log2h, int32_t i, bool *fresh) {
  uint32_t h = hash_row(i, log2h);
  for (;;) {
    const int32_t old = atomicCAS(keys + h, -1, i);
    if (old == -1) {
      *fresh = true;
      return (int)h;
    }
    if (old == i) {
      *fresh = false;
      return (int)h;
    }
    h = (h + 1) & hmask;
  }
}

// ---- pass 1: symbolic ---------------------------------------------------------------------------
// G lanes per column, table of H row slots per group in shared memory.
template <int G, int H, int THREADS>
__global__ void __launch_bounds__(THREADS)
k_mm_symbolic(int32_t ncols_in_class, const int32_t *__restrict__ cols, const int32_t *__restrict__ ap,
              const int32_t *__restrict__ ai, const int32_t *__restrict__ bp,
              const int32_t *__restrict__ bi, int32_t *__restrict__ cnt) {
  constexpr int GROUPS = THREADS / G;
  constexpr int LOG2H = (H == 128) ? 7 : (H == 1024) ? 10 : 12;
  static_assert((1 << LOG2H) == H, "table size");
  extern __shared__ int32_t smem_i32[];
  const int g = threadIdx.x / G;
  const int gl = threadIdx.x % G;
  int32_t *keys = smem_i32 + (size_t)g * H;
  const unsigned lane = threadIdx.x & 31;
  const unsigned mask = G >= 32 ? 0xffffffffu : (((1u << G) - 1u) << (lane / G * G));
  const int32_t slot = blockIdx.x * GROUPS + g;
  const bool active = slot < ncols_in_class;  // uniform per group
  for (int s = gl; s < H; s += G) keys[s] = -1;
  group_sync<G>(mask);
  int32_t mine = 0;
  int32_t j = 0;
  if (active) {
    j = cols[slot];
    for (int32_t q = bp[j]; q < bp[j + 1]; ++q) {
      const int32_t k = __ldg(bi + q);
      const int32_t pe = __ldg(ap + k + 1);
      for (int32_t p = __ldg(ap + k) + gl; p < pe; p += G) {
        bool fresh;
        table_insert(keys, H - 1, LOG2H, __ldg(ai + p), &fresh);
        mine += fresh ? 1 : 0;
      }
    }
  }
  // group reduction of the per-lane claim counts
  if (G <= 32) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) mine += __shfl_xor_sync(mask, mine, o, G <= 32 ? G : 32);
    if (active && gl == 0) cnt[j] = mine;
  } else {
    __shared__ int32_t total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0 && mine) atomicAdd(&total, mine);
    __syncthreads();
    if (active && threadIdx.x == 0) cnt[j] = total;
  }
}

// ---- pass 2: numeric ----------------------------------------------------------------------------

// in-place bitonic sort of n_pad (power of two) 64-bit items by a group of G lanes
template <int G>
__device__ __forceinline__ void group_bitonic(uint64_t *a, int n_pad, int gl, unsigned mask) {
  for (int k = 2; k <= n_pad; k <<= 1) {
    for (int jj = k >> 1; jj > 0; jj >>= 1) {
      for (int t = gl; t < (n_pad >> 1); t += G) {
        const int lo = ((t & ~(jj - 1)) << 1) | (t & (jj - 1));
        const int hi = lo | jj;
        const bool up = (lo & k) == 0;
        const uint64_t x = a[lo], y = a[hi];
        if ((x > y) == up) {
          a[lo] = y;
          a[hi] = x;
        }
      }
      group_sync<G>(mask);
    }
  }
}

template <class T, int G, int H, int THREADS>
__global__ void __launch_bounds__(THREADS)
k_mm_numeric(int32_t ncols_in_class, const int32_t *__restrict__ cols, const int32_t *__restrict__ ap,
             const int32_t *__restrict__ ai, const T *__restrict__ av, const int32_t *__restrict__ bp,
             const int32_t *__restrict__ bi, const T *__restrict__ bv, const int32_t *__restrict__ cp,
             int32_t *__restrict__ ci, T *__restrict__ cv) {
  constexpr int GROUPS = THREADS / G;
  constexpr int LOG2H = (H == 128) ? 7 : (H == 1024) ? 10 : 12;
  static_assert((1 << LOG2H) == H, "table size");
  // layout per block: [GROUPS][H] values | [GROUPS][H/2] packed sort items | [GROUPS][H] keys |
  // counters.  A column of this class has at most H/2 products, hence at most H/2 rows.
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T *vals_all = reinterpret_cast<T *>(smem_raw);
  uint64_t *items_all = reinterpret_cast<uint64_t *>(vals_all + (size_t)GROUPS * H);
  int32_t *keys_all = reinterpret_cast<int32_t *>(items_all + (size_t)GROUPS * (H / 2));
  int32_t *counters = keys_all + (size_t)GROUPS * H;
  const int g = threadIdx.x / G;
  const int gl = threadIdx.x % G;
  T *vals = vals_all + (size_t)g * H;
  uint64_t *items = items_all + (size_t)g * (H / 2);
  int32_t *keys = keys_all + (size_t)g * H;
  const unsigned lane = threadIdx.x & 31;
  const unsigned mask = G >= 32 ? 0xffffffffu : (((1u << G) - 1u) << (lane / G * G));
  const int32_t slot = blockIdx.x * GROUPS + g;
  const bool active = slot < ncols_in_class;  // uniform per group
  for (int s = gl; s < H; s += G) {
    keys[s] = -1;
    vals[s] = Num<T>::zero();
  }
  if (gl == 0) counters[g] = 0;
  group_sync<G>(mask);
  if (!active && G <= 32) return;  // whole sub-warp group leaves together
  int32_t j = 0;
  if (active) {
    j = cols[slot];
    const int32_t qe = bp[j + 1];
    for (int32_t q = bp[j]; q < qe; ++q) {
      const int32_t k = __ldg(bi + q);
      const T b = bv[q];
      const int32_t pe = __ldg(ap + k + 1);
      for (int32_t p = __ldg(ap + k) + gl; p < pe; p += G) {
        bool fresh;
        const int s = table_insert(keys, H - 1, LOG2H, __ldg(ai + p), &fresh);
        vals[s] = Num<T>::add(vals[s], Num<T>::mul(av[p], b));  // c + a*b, Sparse.hs:699
      }
      group_sync<G>(mask);  // all terms of inner index k land before any of k+1
    }
    // compact the occupied slots: item = row << 32 | slot
    for (int s = gl; s < H; s += G) {
      const int32_t key = keys[s];
      if (key >= 0) {
        const int at = atomicAdd(counters + g, 1);
        items[at] = ((uint64_t)(uint32_t)key << 32) | (uint32_t)s;
      }
    }
  }
  group_sync<G>(mask);
  const int n = counters[g];
  int n_pad = 1;
  while (n_pad < n) n_pad <<= 1;
  for (int s = n + gl; s < n_pad; s += G) items[s] = ~0ull;
  group_sync<G>(mask);
  group_bitonic<G>(items, n_pad, gl, mask);
  if (active) {
    const int32_t base = cp[j];
    for (int e = gl; e < n; e += G) {
      const uint64_t it = items[e];
      ci[base + e] = (int32_t)(it >> 32);
      cv[base + e] = vals[(uint32_t)it];
    }
  }
}

// ---- dense fallback for very long columns -------------------------------------------------------
// One block walks its share of the class-3 columns.  stamp[i] == j+1 marks row i as present in
// the column being built, so no reset pass is needed between columns.
template <class T, bool NUMERIC>
__global__ void __launch_bounds__(256)
k_mm_dense(int32_t ncols_in_class, const int32_t *__restrict__ cols, int32_t ar,
           const int32_t *__restrict__ ap, const int32_t *__restrict__ ai, const T *__restrict__ av,
           const int32_t *__restrict__ bp, const int32_t *__restrict__ bi, const T *__restrict__ bv,
           int32_t *__restrict__ stamp_all, T *__restrict__ work_all, int32_t *__restrict__ cnt,
           const int32_t *__restrict__ cp, int32_t *__restrict__ ci, T *__restrict__ cv) {
  __shared__ int32_t s_min, s_max, s_count;
  __shared__ int32_t s_warp[8];
  int32_t *stamp = stamp_all + (size_t)blockIdx.x * ar;
  T *work = NUMERIC ? work_all + (size_t)blockIdx.x * ar : nullptr;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int32_t slot = blockIdx.x; slot < ncols_in_class; slot += gridDim.x) {
    const int32_t j = cols[slot];
    const int32_t id = j + 1;
    if (tid == 0) {
      s_min = 0x7fffffff;
      s_max = -1;
      s_count = 0;
    }
    __syncthreads();
    int32_t lmin = 0x7fffffff, lmax = -1;
    for (int32_t q = bp[j]; q < bp[j + 1]; ++q) {
      const int32_t k = bi[q];
      const int32_t pe = ap[k + 1];
      T b = Num<T>::zero();
      if (NUMERIC) b = bv[q];
      for (int32_t p = ap[k] + tid; p < pe; p += 256) {
        const int32_t i = ai[p];
        if (stamp[i] != id) {
          stamp[i] = id;
          if (NUMERIC) work[i] = Num<T>::zero();
        }
        if (NUMERIC) work[i] = Num<T>::add(work[i], Num<T>::mul(av[p], b));
        lmin = min(lmin, i);
        lmax = max(lmax, i);
      }
      __syncthreads();  // keep the terms of one entry in increasing inner index
    }
    atomicMin(&s_min, lmin);
    atomicMax(&s_max, lmax);
    __syncthreads();
    const int32_t lo = s_min, hi = s_max;
    // ordered compaction of the stamped rows in [lo, hi]
    int32_t written = 0;
    for (int32_t base_row = lo; base_row <= hi; base_row += 256) {
      const int32_t i = base_row + tid;
      const bool present = i <= hi && stamp[i] == id;
      const unsigned bal = __ballot_sync(0xffffffffu, present);
      if (lane == 0) s_warp[warp] = __popc(bal);
      __syncthreads();
      int32_t before = 0, total = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        const int32_t c = s_warp[w];
        before += (w < warp) ? c : 0;
        total += c;
      }
      if (NUMERIC && present) {
        const int32_t at = cp[j] + written + before + __popc(bal & ((1u << lane) - 1));
        ci[at] = i;
        cv[at] = work[i];
      }
      written += total;
      __syncthreads();
    }
    if (!NUMERIC && tid == 0) cnt[j] = written;
    __syncthreads();
  }
}

template <class T, int G, int H, int THREADS>
static int run_symbolic(int32_t n, const int32_t *cols, const slc_matrix *a, const slc_matrix *b,
                        int32_t *cnt) {
  if (n == 0) return SLC_OK;
  constexpr int GROUPS = THREADS / G;
  const size_t smem = (size_t)GROUPS * H * sizeof(int32_t);
  auto kern = k_mm_symbolic<G, H, THREADS>;
  if (smem > 48 * 1024)
    SLC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)ceil_div(n, GROUPS), THREADS, smem, cur_stream()>>>(n, cols, a->colptr, a->rowidx,
                                                                        b->colptr, b->rowidx, cnt);
  SLC_LAUNCHED();
  return SLC_OK;
}

template <class T, int G, int H, int THREADS>
static int run_numeric(int32_t n, const int32_t *cols, const slc_matrix *a, const slc_matrix *b,
                       slc_matrix *c) {
  if (n == 0) return SLC_OK;
  constexpr int GROUPS = THREADS / G;
  const size_t smem = (size_t)GROUPS * H * (sizeof(T) + sizeof(int32_t)) +
                      (size_t)GROUPS * (H / 2) * sizeof(uint64_t) + (size_t)GROUPS * sizeof(int32_t);
  auto kern = k_mm_numeric<T, G, H, THREADS>;
  if (smem > 48 * 1024)
    SLC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)ceil_div(n, GROUPS), THREADS, smem, cur_stream()>>>(
      n, cols, a->colptr, a->rowidx, (const T *)a->vals, b->colptr, b->rowidx, (const T *)b->vals,
      c->colptr, c->rowidx, (T *)c->vals);
  SLC_LAUNCHED();
  return SLC_OK;
}

template <class T>
static int mm_entry(const slc_matrix *a, const slc_matrix *b, slc_matrix **out) {
  if (!a || !b || !out) return fail(SLC_ERR_INVALID, "mm: null argument");
  *out = nullptr;
  if (a->elem != Num<T>::elem || b->elem != Num<T>::elem)
    return fail(SLC_ERR_INVALID, "mm: element type mismatch");
  if (a->ncols != b->nrows) return fail(SLC_ERR_DIM, "mm: inner dimension mismatch");
  cudaStream_t s = cur_stream();
  const int32_t ar = a->nrows, bc = b->ncols;

  // pass 0: products per column, class, stable split of the column ids by class
  DevBuf<int32_t> ub, klass, cols, cnt, btab;
  SLC_TRY(ub.alloc((size_t)bc));
  SLC_TRY(klass.alloc((size_t)bc));
  SLC_TRY(cols.alloc((size_t)bc));
  SLC_TRY(cnt.alloc((size_t)bc + 1));
  SLC_CUDA(cudaMemsetAsync(cnt.p, 0, ((size_t)bc + 1) * sizeof(int32_t), s));
  int32_t hcount[MM_BINS] = {}, hbase[MM_BINS + 1] = {};
  const char *merge_env = getenv("SLC_MM_MERGE");  // SLC_MM_MERGE=0: hash tables for every column
  const int merge_ok = !(merge_env && merge_env[0] == '0');
  if (bc > 0) {
    const int32_t nb = (int32_t)ceil_div(bc, MM_SPLIT_THREADS);
    SLC_TRY(btab.alloc((size_t)nb * MM_BINS));
    k_mm_upper<<<(unsigned)nb, MM_SPLIT_THREADS, 0, s>>>(bc, nb, b->colptr, b->rowidx, a->colptr, ub.p, klass.p,
                                                         btab.p, merge_ok);
    SLC_LAUNCHED();
    SLC_TRY(exclusive_scan_i32(btab.p, btab.p, (int64_t)nb * MM_BINS));
    // first offset of every bin = where its columns start in cols[]
    SLC_CUDA(cudaMemcpy2DAsync(hbase, sizeof(int32_t), btab.p, (size_t)nb * sizeof(int32_t), sizeof(int32_t),
                               MM_BINS, cudaMemcpyDeviceToHost, s));
    k_mm_bucket<<<(unsigned)nb, MM_SPLIT_THREADS, 0, s>>>(bc, nb, klass.p, btab.p, cols.p);
    SLC_LAUNCHED();
    SLC_CUDA(cudaStreamSynchronize(s));
    hbase[MM_BINS] = bc;
    for (int c = 0; c < MM_BINS; ++c) hcount[c] = hbase[c + 1] - hbase[c];
  }

  // dense-fallback workspace, bounded to ~4 GB
  DevBuf<int32_t> stamp;
  DevBuf<T> work;
  int dense_blocks = 0;
  if (hcount[CLS_DENSE] > 0) {
    const size_t per_block = (size_t)(ar > 0 ? ar : 1) * (sizeof(int32_t) + sizeof(T));
    size_t fit = ((size_t)4 << 30) / per_block;
    if (fit < 1) fit = 1;
    dense_blocks = (int)(fit < (size_t)(2 * sm_count()) ? fit : (size_t)(2 * sm_count()));
    if (dense_blocks > hcount[CLS_DENSE]) dense_blocks = hcount[CLS_DENSE];
    SLC_TRY(stamp.alloc((size_t)dense_blocks * ar));
    SLC_TRY(work.alloc((size_t)dense_blocks * ar));
    SLC_CUDA(cudaMemsetAsync(stamp.p, 0, (size_t)dense_blocks * ar * sizeof(int32_t), s));
  }

  // pass 1: symbolic
  SLC_TRY((run_merge<T, 4, false>(hcount[CLS_MERGE4], cols.p + hbase[CLS_MERGE4], a, b, cnt.p, nullptr)));
  SLC_TRY((run_merge<T, 8, false>(hcount[CLS_MERGE8], cols.p + hbase[CLS_MERGE8], a, b, cnt.p, nullptr)));
  SLC_TRY((run_symbolic<T, 8, 128, 128>(hcount[CLS_HASH0], cols.p + hbase[CLS_HASH0], a, b, cnt.p)));
  SLC_TRY((run_symbolic<T, 32, 1024, 128>(hcount[CLS_HASH1], cols.p + hbase[CLS_HASH1], a, b, cnt.p)));
  SLC_TRY((run_symbolic<T, 256, 4096, 256>(hcount[CLS_HASH2], cols.p + hbase[CLS_HASH2], a, b, cnt.p)));
  if (hcount[CLS_DENSE] > 0) {
    k_mm_dense<T, false><<<dense_blocks, 256, 0, s>>>(hcount[CLS_DENSE], cols.p + hbase[CLS_DENSE], ar, a->colptr,
                                                      a->rowidx, (const T *)a->vals, b->colptr, b->rowidx,
                                                      (const T *)b->vals, stamp.p, work.p, cnt.p, nullptr,
                                                      nullptr, nullptr);
    SLC_LAUNCHED();
    // the numeric pass reuses the stamps with the same ids: clear them
    SLC_CUDA(cudaMemsetAsync(stamp.p, 0, (size_t)dense_blocks * ar * sizeof(int32_t), s));
  }

  // counts -> pointers.  The total is summed in 64 bits beside a 32-bit scan, so an oversized product is
  // reported, not wrapped (round 1 widened the counts, scanned them in 64 bits and narrowed the result: three
  // more passes over the columns than this).
  DevBuf<int32_t> ptr32;
  DevBuf<unsigned long long> total;
  SLC_TRY(ptr32.alloc((size_t)bc + 1));
  SLC_TRY(total.alloc(2));
  SLC_CUDA(cudaMemsetAsync(total.p, 0, 2 * sizeof(unsigned long long), s));
  SLC_TRY(sum_i32_into(cnt.p, (int64_t)bc + 1, total.p));
  SLC_TRY(max_i32_into(cnt.p, (int64_t)bc, total.p + 1));  // longest result column: sizes the output windows
  SLC_TRY(exclusive_scan_i32(cnt.p, ptr32.p, (int64_t)bc + 1));
  unsigned long long hboth[2] = {0, 0};
  SLC_CUDA(cudaMemcpyAsync(hboth, total.p, sizeof hboth, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  const unsigned long long htotal = hboth[0];
  const int fit_cap = (int)std::min<unsigned long long>(((hboth[1] * 32 + 31) / 32) * 32, 1u << 20);
  if (htotal > 0x7fffffffULL) return fail(SLC_ERR_RANGE, "mm: product has %lld entries, more than 32-bit pointers hold", (long long)htotal);
  const int64_t nnzc = (int64_t)htotal;

  slc_matrix *c = nullptr;
  SLC_TRY(matrix_alloc(a->elem, ar, bc, nnzc, &c));
  if (cudaError_t e = cudaMemcpyAsync(c->colptr, ptr32.p, ((size_t)bc + 1) * sizeof(int32_t), cudaMemcpyDeviceToDevice, s);
      e != cudaSuccess) {
    matrix_release(c);
    return fail(SLC_ERR_CUDA, "mm: pointer copy failed: %s", cudaGetErrorString(e));
  }

  // pass 2: numeric
  int rc = run_merge<T, 4, true>(hcount[CLS_MERGE4], cols.p + hbase[CLS_MERGE4], a, b, nullptr, c, fit_cap);
  if (rc == SLC_OK) rc = run_merge<T, 8, true>(hcount[CLS_MERGE8], cols.p + hbase[CLS_MERGE8], a, b, nullptr, c, fit_cap);
  if (rc == SLC_OK) rc = run_numeric<T, 8, 128, 128>(hcount[CLS_HASH0], cols.p + hbase[CLS_HASH0], a, b, c);
  if (rc == SLC_OK) rc = run_numeric<T, 32, 1024, 128>(hcount[CLS_HASH1], cols.p + hbase[CLS_HASH1], a, b, c);
  if (rc == SLC_OK) rc = run_numeric<T, 256, 4096, 256>(hcount[CLS_HASH2], cols.p + hbase[CLS_HASH2], a, b, c);
  if (rc == SLC_OK && hcount[CLS_DENSE] > 0) {
    k_mm_dense<T, true><<<dense_blocks, 256, 0, s>>>(hcount[CLS_DENSE], cols.p + hbase[CLS_DENSE], ar, a->colptr, a->rowidx,
                                                     (const T *)a->vals, b->colptr, b->rowidx,
                                                     (const T *)b->vals, stamp.p, work.p, cnt.p, c->colptr,
                                                     c->rowidx, (T *)c->vals);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (cudaError_t e = cudaGetLastError(); e != cudaSuccess)
      rc = fail(SLC_ERR_CUDA, "mm: dense pass launch failed: %s", cudaGetErrorString(e));
  }
  if (rc != SLC_OK) {
    matrix_release(c);
    return rc;
  }
  *out = c;
  return SLC_OK;
}

}  // namespace slc

using namespace slc;

extern "C" {
int slc_d_mm(const slc_matrix *a, const slc_matrix *b, slc_matrix **c) { return mm_entry<double>(a, b, c); }
int slc_z_mm(const slc_matrix *a, const slc_matrix *b, slc_matrix **c) { return mm_entry<double2>(a, b, c); }
}
