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
#include <chrono>

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

// ---- compress, bucket path (round 2) ---------------------------------------------------------------
// When no column receives more than ASM_BUCKET_LIMIT triplets (finite-element and stencil assembly: a
// few dozen per column) the full sort is replaced by the reference's own shape -- count per column,
// prefix sum, place (computePtrs + the scatter of Sparse.hs:282-329) -- followed by a sort of each
// column's few triplets:
//   k_coo_check_count   bounds check + triplets per column; the value the atomic returns is the
//                       triplet's rank inside its bucket and is kept (2 B), so this is the only
//                       pass with atomics
//   exclusive scan      bucket starts
//   k_bucket_scatter    triplet i -> slot start[col] + rank[i] as ONE 16/32-byte record
//                       { row << 32 | i, value }: a single sector written per triplet, values carried
//                       along so nothing is gathered later.  The order inside a bucket is whatever
//                       the atomics gave; the next step removes that freedom.
//   k_bucket_reduce     128 columns per block: keys of the block's contiguous span staged in shared
//                       memory, each thread insertion-sorts its column by (row, i), sums runs of equal
//                       row in increasing i (= input order, the same convention as the sort path and
//                       the oracle) and the block writes its reduced entries contiguously
//   exclusive scan      distinct rows per column -> colptr
//   k_bucket_compact    block-contiguous copy into the result
// Traffic on n triplets -> m entries (Double): 10n + 18n + 32n (sector per record) + 16n + 12m + 24m against
// 6 passes x 20n B of the sort path.  Deterministic: (row, i) is a total order.
constexpr int ASM_BUCKET_LIMIT = 512;   // longest bucket the per-thread insertion sort is allowed to see
constexpr int ASM_COLS = 128;           // columns (= threads) per block
constexpr int ASM_CAP = 2048;           // staged records per pass (36 KB for Double, 52 KB for Complex Double)

// first_bad[2] is raised as soon as some bucket passes ASM_BUCKET_LIMIT (the atomic's return value shows
// it); blocks that start after that only check bounds, so a matrix with long columns (power-law) gives
// the path up after a fraction of a per cent of its triplets instead of paying for contended atomics.
__global__ void k_coo_check_count(int64_t n, int32_t nrows, int32_t ncols, const int32_t *__restrict__ rows,
                                  const int32_t *__restrict__ cols, int32_t *__restrict__ cnt,
                                  uint16_t *__restrict__ rank, unsigned long long *__restrict__ first_bad) {
  __shared__ int given_up;
  if (threadIdx.x == 0) given_up = *reinterpret_cast<volatile unsigned long long *>(first_bad + 2) != 0ull;
  __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t r = rows[i], c = cols[i];
  const bool bad_r = r < 0 || r >= nrows;
  const bool bad_c = c < 0 || c >= ncols;
  if (bad_r) atomicMin(first_bad + 0, (unsigned long long)i);
  if (bad_c) atomicMin(first_bad + 1, (unsigned long long)i);
  if (!bad_c && !given_up) {
    const int32_t k = atomicAdd(cnt + c, 1);
    rank[i] = (uint16_t)min(k, 65535);
    if (k >= ASM_BUCKET_LIMIT) *reinterpret_cast<volatile unsigned long long *>(first_bad + 2) = 1ull;
  }
}

template <class T>
__global__ void k_bucket_scatter(int64_t n, const int32_t *__restrict__ rows, const int32_t *__restrict__ cols,
                                 const T *__restrict__ vals, const int32_t *__restrict__ start,
                                 const uint16_t *__restrict__ rank, AsmEnt<T> *__restrict__ ent) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t slot = __ldg(start + cols[i]) + rank[i];
  AsmEnt<T> e;
  e.key = ((uint64_t)(uint32_t)rows[i] << 32) | (uint32_t)i;
  e.val = vals[i];
  ent[slot] = e;
}


// ---- coarse column partition in front of the bucket scatter -----------------------------------------
// With triplets in arbitrary order every record of k_bucket_scatter lands in a different sector of a
// 16n-byte array: on cfg 3 (29 M shuffled triplets, 467 MB of records) that one kernel was 1.1 of the
// path's 2.0 ms, DRAM read-modify-write of a sector per record.  When the records outgrow L2 the triplets
// are therefore first split by the top 8 bits of their column (one radix-style pass: per-tile digit
// counts from k_radix_hist, a scan, and this kernel), which costs one streaming pass and makes every later
// access local: a slice of the column range owns 1/256 of the records, a few MB that stay in L2 while
// they are filled.  The order inside a slice is irrelevant (the buckets are sorted by (row, i) later), so
// a tile ranks its triplets with shared-memory atomics and writes each record where it falls -- the 256
// write fronts advance contiguously, L2 merges the partial sectors.  The bounds check and the rank atomics
// of k_coo_check_count ride along.
// The partitioned triplets, one array per field so that a run of records is a run of full sectors.
template <class T>
struct CoarseRecs {
  uint64_t *key = nullptr;      // row << 32 | input position
  uint64_t *colrank = nullptr;  // column << 16 | rank inside the column's bucket
  T *val = nullptr;
};

constexpr int CS_HALF = RS_TILE / 2;            // a tile is staged in two halves of 2048 records
constexpr int CS_ITEMS = CS_HALF / RS_THREADS;  // 8 per thread and half

template <class T>
__global__ void __launch_bounds__(RS_THREADS, 3)  // 4 blocks (64 registers, spills): 600 us against 458
k_coarse_scatter(int64_t n, int cshift, int64_t ntiles, const int32_t *__restrict__ rows,
                 const int32_t *__restrict__ cols, const T *__restrict__ vals, const uint16_t *__restrict__ rank,
                 const int32_t *__restrict__ offsets, const CoarseRecs<T> rec,
                 const unsigned long long *__restrict__ first_bad) {
  extern __shared__ __align__(16) unsigned char cs_smem[];
  uint64_t *s_key = reinterpret_cast<uint64_t *>(cs_smem);  // [CS_HALF]
  uint64_t *s_cr = s_key + CS_HALF;                         // [CS_HALF]
  T *s_val = reinterpret_cast<T *>(s_cr + CS_HALF);         // [CS_HALF]
  __shared__ int32_t fill[RS_BINS];    // records of digit d in this half
  __shared__ int32_t lbase[RS_BINS];   // where digit d starts in the staged half
  __shared__ int32_t gshift[RS_BINS];  // global position of a staged record = gshift[d] + its staged index
  __shared__ int32_t wsum[RS_THREADS / 32];
  __shared__ int given_up;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // k_coo_check_count ran before: a raised word means an index out of bounds or a bucket over the limit --
  // the caller is about to drop this path, and a column may not be a valid digit source
  if (tid == 0)
    given_up = (first_bad[0] & first_bad[1]) != ~0ull || first_bad[2] != 0ull;
  int32_t gnext = offsets[(int64_t)tid * ntiles + blockIdx.x];  // thread `tid` keeps digit `tid`'s write front
  const int64_t tb = (int64_t)blockIdx.x * RS_TILE;
  __syncthreads();
  if (given_up) return;
  for (int h = 0; h < 2; ++h) {
    const int64_t hb = tb + (int64_t)h * CS_HALF;
    if (hb >= n) break;  // block-uniform
    fill[tid] = 0;
    __syncthreads();
    uint64_t key[CS_ITEMS], cr[CS_ITEMS];
    T v[CS_ITEMS];
    int32_t lp[CS_ITEMS];
#pragma unroll
    for (int it = 0; it < CS_ITEMS; ++it) {
      const int64_t i = hb + it * RS_THREADS + tid;
      const bool have = i < n;
      const int32_t r = have ? rows[i] : 0, c = have ? cols[i] : 0;
      const uint32_t k = have ? rank[i] : 0;
      v[it] = have ? vals[i] : Num<T>::zero();
      key[it] = ((uint64_t)(uint32_t)r << 32) | (uint32_t)i;
      cr[it] = ((uint64_t)(uint32_t)c << 16) | k;
      lp[it] = have ? 0 : -1;
    }
#pragma unroll
    for (int it = 0; it < CS_ITEMS; ++it)
      if (lp[it] >= 0) lp[it] = atomicAdd(&fill[(int)(((uint32_t)(cr[it] >> 16) >> cshift) & (RS_BINS - 1))], 1);
    __syncthreads();
    {  // exclusive scan of the digit counts over the 256 threads
      const int32_t mine = fill[tid];
      int32_t incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      if (lane == 31) wsum[warp] = incl;
      __syncthreads();
      int32_t before = 0;
#pragma unroll
      for (int w = 0; w < RS_THREADS / 32; ++w)
        if (w < warp) before += wsum[w];
      const int32_t lb = before + incl - mine;
      lbase[tid] = lb;
      gshift[tid] = gnext - lb;
      gnext += mine;
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < CS_ITEMS; ++it) {
      if (lp[it] >= 0) {
        const int j = lbase[(int)(((uint32_t)(cr[it] >> 16) >> cshift) & (RS_BINS - 1))] + lp[it];
        s_key[j] = key[it];
        s_cr[j] = cr[it];
        s_val[j] = v[it];
      }
    }
    __syncthreads();
    const int32_t count = (int32_t)min((int64_t)CS_HALF, n - hb);
#pragma unroll 4
    for (int32_t j = tid; j < count; j += RS_THREADS) {
      const uint64_t c2 = s_cr[j];
      const int64_t pos = (int64_t)gshift[(int)(((uint32_t)(c2 >> 16) >> cshift) & (RS_BINS - 1))] + j;
      rec.key[pos] = s_key[j];
      rec.colrank[pos] = c2;
      rec.val[pos] = s_val[j];
    }
    __syncthreads();  // the staging arrays and fill[] are reused by the second half
  }
}

template <class T>
constexpr size_t coarse_scatter_smem() { return (size_t)CS_HALF * (2 * sizeof(uint64_t) + sizeof(T)); }

template <class T>
__global__ void k_fine_scatter(int64_t n, const CoarseRecs<T> rec, const int32_t *__restrict__ start,
                               AsmEnt<T> *__restrict__ ent) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t cr = rec.colrank[i];
  const int32_t slot = __ldg(start + (int32_t)(cr >> 16)) + (int32_t)(cr & 0xffffu);
  AsmEnt<T> e;
  e.key = rec.key[i];
  e.val = rec.val[i];
  ent[slot] = e;
}

// Sort of up to ASM_REG keys with their values entirely in registers (bitonic network, every index
// static): used when all 32 columns of a warp are that short, which is the common case of stencil and
// low-order finite-element assembly.  The first version insertion-sorted every bucket in shared memory;
// with 128 threads walking segments a fixed even stride apart that ran into 8-way bank conflicts on
// every step (ncu launch list: 476 us on cfg 3 for a kernel whose traffic is 0.15 ms).
constexpr int ASM_REG = 16;

template <class T>
__device__ __forceinline__ void sort_network(uint64_t (&key)[ASM_REG], T (&val)[ASM_REG]) {
#pragma unroll
  for (int k = 2; k <= ASM_REG; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
      for (int i = 0; i < ASM_REG; ++i) {
        const int l = i ^ j;
        if (l > i) {
          const bool up = (i & k) == 0;
          const bool swap = up ? key[i] > key[l] : key[i] < key[l];
          const uint64_t ka = swap ? key[l] : key[i], kb = swap ? key[i] : key[l];
          const T va = swap ? val[l] : val[i], vb = swap ? val[i] : val[l];
          key[i] = ka; key[l] = kb;
          val[i] = va; val[l] = vb;
        }
      }
    }
  }
}

template <class T>
constexpr size_t bucket_reduce_smem() { return (size_t)ASM_CAP * (sizeof(uint64_t) + sizeof(T) + sizeof(uint16_t)); }

template <class T>
__global__ void __launch_bounds__(ASM_COLS, sizeof(T) == 8 ? 6 : 3)
k_bucket_reduce(int32_t ncols, const int32_t *__restrict__ start, const AsmEnt<T> *__restrict__ ent,
                int32_t *__restrict__ red_row, T *__restrict__ red_val, int32_t *__restrict__ ndist) {
  extern __shared__ __align__(16) unsigned char br_smem[];
  T *sv = reinterpret_cast<T *>(br_smem);                       // [ASM_CAP] values of the staged span
  uint64_t *sk = reinterpret_cast<uint64_t *>(sv + ASM_CAP);    // [ASM_CAP] keys
  uint16_t *si = reinterpret_cast<uint16_t *>(sk + ASM_CAP);    // [ASM_CAP] staged index of a sorted key (long buckets)
  __shared__ int32_t wsum[ASM_COLS / 32];
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int32_t cbeg = blockIdx.x * ASM_COLS;
  const int32_t cend = min(cbeg + ASM_COLS, ncols);
  int32_t c0 = cbeg;
  int32_t wr = start[cbeg];  // next free slot of the block's reduced output (block-uniform)
  while (c0 < cend) {
    const int32_t s0 = start[c0];
    const int32_t c = c0 + t;
    const int32_t e_t = c < cend ? start[c + 1] : 0;
    const int ncover = __syncthreads_count(c < cend && e_t - s0 <= ASM_CAP);  // a prefix: ends are monotone
    // (the launch guarantees every bucket <= ASM_BUCKET_LIMIT <= ASM_CAP, so ncover >= 1)
    const int32_t len = start[c0 + ncover] - s0;
    for (int32_t i = t; i < len; i += ASM_COLS) {
      const AsmEnt<T> e = ent[s0 + i];
      sk[i] = e.key;
      sv[i] = e.val;
      si[i] = (uint16_t)i;
    }
    __syncthreads();
    int32_t a = 0, b = 0, k = 0;
    if (t < ncover) {
      a = start[c] - s0;
      b = e_t - s0;
    }
    const bool in_regs = __all_sync(0xffffffffu, b - a <= ASM_REG);
    uint64_t key[ASM_REG];
    T val[ASM_REG];
    if (in_regs) {
#pragma unroll
      for (int j = 0; j < ASM_REG; ++j) {
        const bool have = a + j < b;
        key[j] = have ? sk[a + j] : ~0ull;  // padding sorts behind every real key (rows are < 2^31)
        val[j] = have ? sv[a + j] : Num<T>::zero();
      }
      sort_network<T>(key, val);
#pragma unroll
      for (int j = 0; j < ASM_REG; ++j)
        k += (a + j < b) && (j == 0 || (key[j] >> 32) != (key[j > 0 ? j - 1 : 0] >> 32));
    } else {
      for (int32_t i = a + 1; i < b; ++i) {
        const uint64_t kk = sk[i];
        const uint16_t idx = si[i];
        int32_t j = i;
        while (j > a && sk[j - 1] > kk) {
          sk[j] = sk[j - 1];
          si[j] = si[j - 1];
          --j;
        }
        sk[j] = kk;
        si[j] = idx;
      }
      for (int32_t i = a; i < b; ++i) k += (i == a) || ((sk[i] >> 32) != (sk[i - 1] >> 32));
    }
    if (t < ncover) ndist[c] = k;
    // exclusive prefix of k over the block
    int32_t incl = k;
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    int32_t before = 0, total = 0;
    for (int w = 0; w < ASM_COLS / 32; ++w) {
      const int32_t v = wsum[w];
      if (w < wid) before += v;
      total += v;
    }
    int32_t w_out = wr + before + incl - k;
    if (in_regs) {
      const int32_t n_mine = b - a;
      T acc = val[0];
#pragma unroll
      for (int j = 1; j <= ASM_REG; ++j) {
        // entry j - 1 closes a run when it is the last one or the next row differs
        const bool last = j == n_mine;
        const bool next_differs = j < ASM_REG && j < n_mine && (key[j < ASM_REG ? j : 0] >> 32) != (key[j - 1] >> 32);
        if (j - 1 < n_mine && (last || next_differs)) {
          red_row[w_out] = (int32_t)(key[j - 1] >> 32);
          red_val[w_out] = acc;
          ++w_out;
        }
        if (j < ASM_REG) acc = next_differs ? val[j] : Num<T>::add(acc, val[j]);
      }
    } else {
      for (int32_t i = a; i < b;) {
        const uint32_t row = (uint32_t)(sk[i] >> 32);
        T acc = sv[si[i]];
        for (++i; i < b && (uint32_t)(sk[i] >> 32) == row; ++i) acc = Num<T>::add(acc, sv[si[i]]);
        red_row[w_out] = (int32_t)row;
        red_val[w_out] = acc;
        ++w_out;
      }
    }
    wr += total;
    c0 += ncover;
    __syncthreads();  // the staging arrays and wsum are reused by the next pass
  }
}

template <class T>
__global__ void __launch_bounds__(ASM_COLS)
k_bucket_compact(int32_t ncols, const int32_t *__restrict__ start, const int32_t *__restrict__ colptr,
                 const int32_t *__restrict__ red_row, const T *__restrict__ red_val,
                 int32_t *__restrict__ out_row, T *__restrict__ out_val) {
  const int32_t cbeg = blockIdx.x * ASM_COLS;
  const int32_t cend = min(cbeg + ASM_COLS, ncols);
  const int32_t src = start[cbeg];
  const int32_t dst = colptr[cbeg];
  const int32_t cnt = colptr[cend] - dst;
  for (int32_t i = threadIdx.x; i < cnt; i += ASM_COLS) {
    out_row[dst + i] = red_row[src + i];
    out_val[dst + i] = red_val[src + i];
  }
}

// SLC_ASM_TRACE=1: host wall time of every phase of the bucket path (each closed by a stream
// synchronisation, so the phases do not overlap as they do in a normal call) on stderr
struct PhaseTrace {
  bool on;
  cudaStream_t s;
  std::chrono::steady_clock::time_point t;
  int mode;  // 1: synchronise at every mark (device time per phase); 2: no synchronisation (host time per section)
  explicit PhaseTrace(cudaStream_t st)
      : on(env_int("SLC_ASM_TRACE", 0) != 0), s(st), t(std::chrono::steady_clock::now()), mode(env_int("SLC_ASM_TRACE", 0)) {}
  void mark(const char *what) {
    if (!on) return;
    if (mode == 1) cudaStreamSynchronize(s);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[slc compress] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

// The work buffers of the bucket path, carved out of ONE stream-ordered allocation of a size that depends
// only on (n, ncols): the pool then hands the same block back call after call.  (With seven separate
// buffers of different lifetimes cudaMallocAsync kept splitting and regrowing blocks: 0.2 - 0.6 ms per call
// on cfg 3, more than the scans cost.)  Buffers whose lifetimes do not overlap share storage.
template <class T>
struct BucketSlab {
  DevBuf<char> buf;
  bool coarse = false;
  int32_t *cnt = nullptr;     // triplets per column; later: distinct rows per column and their prefix sums
  int32_t *start = nullptr;   // bucket starts
  uint16_t *rank = nullptr;   // (direct path) rank of a triplet inside its bucket
  int32_t *offsets = nullptr; // (coarse path) scanned per-tile digit counts
  CoarseRecs<T> rec;          // (coarse path) triplets split by column slice; dead once the records are placed
  AsmEnt<T> *ent = nullptr;   // records by bucket
  int32_t *red_row = nullptr; // rows of the reduced entries   } share storage with rank / rec, which are dead
  T *red_val = nullptr;       // values of the reduced entries } by the time k_bucket_reduce writes them
  int32_t *ndist() const { return cnt; }
  static size_t up(size_t b) { return (b + 255) & ~(size_t)255; }
  int alloc(int64_t n, int32_t ncols, bool with_coarse) {
    coarse = with_coarse;
    const size_t ptrs = up(((size_t)ncols + 1) * sizeof(int32_t));
    const size_t b_row = up((size_t)n * sizeof(int32_t)), b_val = up((size_t)n * sizeof(T));
    const size_t b_ent = up((size_t)n * sizeof(AsmEnt<T>));
    const size_t b_u64 = up((size_t)n * sizeof(uint64_t));
    const size_t b_rec = coarse ? 2 * b_u64 + b_val : 0;
    const size_t b_off = coarse ? up((size_t)ceil_div(n, RS_TILE) * RS_BINS * sizeof(int32_t)) : 0;
    const size_t b_share = b_row + b_val > b_rec ? b_row + b_val : b_rec;
    // direct path: the ranks live where the reduced rows go later; coarse path: the ranks are read while the
    // partitioned records are written, so they get their own 2n bytes
    const size_t b_rank = coarse ? up((size_t)n * sizeof(uint16_t)) : 0;
    SLC_TRY(buf.alloc(2 * ptrs + b_off + b_rank + b_share + b_ent));
    char *p = buf.p;
    cnt = reinterpret_cast<int32_t *>(p); p += ptrs;
    start = reinterpret_cast<int32_t *>(p); p += ptrs;
    offsets = reinterpret_cast<int32_t *>(p); p += b_off;
    rank = reinterpret_cast<uint16_t *>(p); p += b_rank;
    rec.key = reinterpret_cast<uint64_t *>(p);
    rec.colrank = reinterpret_cast<uint64_t *>(p + b_u64);
    rec.val = reinterpret_cast<T *>(p + 2 * b_u64);
    red_row = reinterpret_cast<int32_t *>(p);
    red_val = reinterpret_cast<T *>(p + b_row);
    p += b_share;
    ent = reinterpret_cast<AsmEnt<T> *>(p);
    return SLC_OK;
  }
};

template <class T>
static int compress_buckets(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows_d, const int32_t *cols_d,
                            const T *vals_d, BucketSlab<T> &w, slc_matrix **out) {
  cudaStream_t s = cur_stream();
  PhaseTrace tr(s);
  SLC_TRY(exclusive_scan_i32(w.cnt, w.start, (int64_t)ncols + 1));
  tr.mark("scan of bucket sizes");
  if (w.coarse)
    k_fine_scatter<T><<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(n, w.rec, w.start, w.ent);
  else
    k_bucket_scatter<T><<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(n, rows_d, cols_d, vals_d, w.start, w.rank, w.ent);
  SLC_LAUNCHED();
  tr.mark(w.coarse ? "k_fine_scatter" : "k_bucket_scatter");
  SLC_CUDA(cudaMemsetAsync(w.ndist() + ncols, 0, sizeof(int32_t), s));
  const unsigned nblk = (unsigned)ceil_div((int64_t)ncols, ASM_COLS);
  static PerDevice<size_t> smem_set;
  SLC_CUDA(ensure_dyn_smem(k_bucket_reduce<T>, smem_set, bucket_reduce_smem<T>()));
  k_bucket_reduce<T><<<nblk, ASM_COLS, bucket_reduce_smem<T>(), s>>>(ncols, w.start, w.ent, w.red_row, w.red_val,
                                                                     w.ndist());
  SLC_LAUNCHED();
  tr.mark("k_bucket_reduce");
  SLC_TRY(exclusive_scan_i32(w.ndist(), w.ndist(), (int64_t)ncols + 1));
  int32_t nnz_out = 0;
  SLC_CUDA(cudaMemcpyAsync(&nnz_out, w.ndist() + ncols, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  tr.mark("scan of distinct rows");
  slc_matrix *m = nullptr;
  SLC_TRY(matrix_alloc(Num<T>::elem, nrows, ncols, nnz_out, &m));
  tr.mark("matrix_alloc");
  cudaError_t e = cudaMemcpyAsync(m->colptr, w.ndist(), ((size_t)ncols + 1) * sizeof(int32_t), cudaMemcpyDeviceToDevice, s);
  if (e == cudaSuccess) {
    k_bucket_compact<T><<<nblk, ASM_COLS, 0, s>>>(ncols, w.start, w.ndist(), w.red_row, w.red_val, m->rowidx,
                                                  (T *)m->vals);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    matrix_release(m);
    return fail(SLC_ERR_CUDA, "compress: %s", cudaGetErrorString(e));
  }
  tr.mark("k_bucket_compact");
  *out = m;
  return SLC_OK;
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
    SLC_TRY(host_to_dev(drows.p, rows, (size_t)n * sizeof(int32_t), s));
    SLC_TRY(host_to_dev(dcols.p, cols, (size_t)n * sizeof(int32_t), s));
    SLC_TRY(host_to_dev(dvals.p, vals, (size_t)n * sizeof(T), s));
    rows_d = drows.p;
    cols_d = dcols.p;
    vals_d = dvals.p;
  }
  PhaseTrace tr0(s);
  SLC_TRY(bad.alloc(3));
  SLC_CUDA(cudaMemsetAsync(bad.p, 0xff, 2 * sizeof(unsigned long long), s));
  SLC_CUDA(cudaMemsetAsync(bad.p + 2, 0, sizeof(unsigned long long), s));
  const unsigned nb = (unsigned)ceil_div(n, 256);
  // SLC_ASM_BUCKET (1): count / place / sort-per-column path when every column is short; 0 = always sort
  const bool try_buckets = env_int("SLC_ASM_BUCKET", 1) != 0 && ncols > 0;
  BucketSlab<T> slab;
  if (try_buckets) {
    // SLC_ASM_COARSE (-1 = when the records outgrow L2): split the triplets by column slice first
    const int coarse_env = env_int("SLC_ASM_COARSE", -1);
    const bool coarse = coarse_env > 0 || (coarse_env < 0 && (size_t)n * sizeof(AsmEnt<T>) > ((size_t)96 << 20));
    tr0.mark("(before slab alloc)");
    SLC_TRY(slab.alloc(n, ncols, coarse));
    tr0.mark("slab alloc");
    SLC_CUDA(cudaMemsetAsync(slab.cnt, 0, ((size_t)ncols + 1) * sizeof(int32_t), s));
    k_coo_check_count<<<nb, 256, 0, s>>>(n, nrows, ncols, rows_d, cols_d, slab.cnt, slab.rank, bad.p);
    SLC_LAUNCHED();
    if (coarse) {  // enqueued before the flags are read back: the kernel itself stands down if they are raised
      const int cshift = colbits > 8 ? colbits - 8 : 0;
      const int64_t ntiles = ceil_div(n, RS_TILE);
      SLC_TRY(tile_digit_offsets_u32(reinterpret_cast<const uint32_t *>(cols_d), n, cshift, slab.offsets));
      static PerDevice<size_t> smem_set;
      SLC_CUDA(ensure_dyn_smem(k_coarse_scatter<T>, smem_set, coarse_scatter_smem<T>()));
      k_coarse_scatter<T><<<(unsigned)ntiles, RS_THREADS, coarse_scatter_smem<T>(), s>>>(
          n, cshift, ntiles, rows_d, cols_d, vals_d, slab.rank, slab.offsets, slab.rec, bad.p);
      SLC_LAUNCHED();
    }
  } else {
    SLC_TRY(k0.alloc(n));
    k_coo_check_pack<<<nb, 256, 0, s>>>(n, nrows, ncols, rowbits, rows_d, cols_d, k0.p, bad.p);
    SLC_LAUNCHED();
  }
  tr0.mark("count kernels enqueued");
  unsigned long long hbad[3];
  SLC_CUDA(cudaMemcpyAsync(hbad, bad.p, sizeof hbad, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  tr0.mark("bounds check + count");
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
  if (try_buckets && hbad[2] == 0ull)  // no bucket over the limit
    return compress_buckets<T>(nrows, ncols, n, rows_d, cols_d, vals_d, slab, out);
  if (try_buckets) {  // a long column: the sort path, whose keys have not been packed yet
    slab.buf.release();
    SLC_TRY(k0.alloc(n));
    k_coo_check_pack<<<nb, 256, 0, s>>>(n, nrows, ncols, rowbits, rows_d, cols_d, k0.p, bad.p);
    SLC_LAUNCHED();
  }
  SLC_TRY(k1.alloc(n));
  SLC_TRY(p0.alloc(n));
  SLC_TRY(p1.alloc(n));
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
