// slc_spmm_brick.cu -- block product Y[:, j] <- A X[:, j] for matrices whose rows all follow one
// 3 x 3 x 3 offset pattern (27-point stencils on a grid numbered x-fastest: BASELINE.json config 5), the
// FEAST-shaped op of Feast.hs:203-208 on the headline matrix.
//
// What round 1's k_spmm_panel ran into (ncu, profiles/r1g_ncu_full_spmm_panel_cfg5mm.txt): one gather of X
// per product through L1 -- 443 M warp requests of ~2.9 cache lines each (the +-1 and +-L neighbours are
// misaligned), 38 GB from L2 to L1 -- plus the column index and value of every product re-read from
// shared memory.  The ceiling was the L1 tag stage (74 % busy), not HBM.
//
// Round 2, first attempt (kept in git history, profiles/r2_ncu_brick_registers.txt): rows that are
// neighbours in y owned by one thread with all their values in registers, X through L1.  255 registers,
// 8 warps per SM, every L1 miss exposed: 8.6-9.5 ms against the panel kernel's 6.3 ms.  Dropped.
//
// This kernel (k_spmm_planes) takes the gathers off the L1 path instead:
//   * the matrix is detected to have every row's columns in  row + shift + {a P + b L + c : a, b, c in
//     {-1, 0, 1}}  (P, L = plane and line strides, read off one row and then VERIFIED for all rows by the
//     plan builder; anything else keeps the panel kernel).  The values are stored by offset slot,
//     dia[slot][row] -- no column indices at all -- plus one 27-bit presence mask per row (grid
//     boundaries drop entries; an absent entry must not be added, not even as 0 * x);
//   * a block owns PL_ROWS consecutive rows.  The entries of X they touch are three contiguous windows
//     (one per plane offset a) of PL_ROWS + 2 L + 2 entries: a producer lane has the TMA engine copy
//     them from L2 into shared memory (cp.async.bulk, completion on an mbarrier), JT columns per
//     pipeline stage, nstage stages in flight -- no register, LSU instruction or L1 line is spent on
//     fetching X, and the 27 reads per row hit shared memory at fixed offsets (t + b L + c): always
//     two wavefronts per warp, whatever the alignment;
//   * a thread keeps its row's 27 values in registers for all k columns (A is read from HBM once).
// Every (row, column) is owned by one thread which adds the row's terms in increasing column order
// (plane, then line, then point), product and sum rounded separately: bit-identical to the reference
// (Sparse.hs:433-453), checked against the oracle at full size.
// Measured and dropped (profiles/r2h_ncu_full_spmm_planes2_cfg5mm.txt): two rows per thread with 16-byte
// shared loads (72 instead of 108 wavefronts per 64 rows, but 128-151 registers and 8-12 consumer warps per
// SM: 4.30-5.26 ms against 4.29-4.35 ms); two or four columns per stage (6.7 ms: coarser hand-over between
// producer and consumers); 256-row blocks, two per SM (4.34 ms: no difference).
// Every mbarrier wait is bounded (as in slc_spmm_window.cu): a protocol bug raises the window error
// flag and gives a wrong answer that the parity tests catch; it cannot hang the GPU.
#include <algorithm>
#include <vector>

#include "slc_common.cuh"

namespace slc {

constexpr uint32_t BRICK_FULL = (1u << 27) - 1u;
// rows per block = consumer threads (+ one producer warp): 512 rows with one block per SM, or 256 rows with two
// (a block that starts has to fetch its rows' values before it computes; a second resident block covers that)
constexpr uint32_t PL_SPIN_LIMIT = 1u << 22;

// one thread per row: file the row's entries by offset slot; any entry off the pattern vetoes the plan
template <class T>
__global__ void k_brick_build(int32_t nrows, int64_t nrows_pad, int64_t shift, int32_t L, int32_t P,
                              const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                              const T *__restrict__ vals, T *__restrict__ dia, uint32_t *__restrict__ mask,
                              int *__restrict__ veto) {
  const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= nrows) return;
  uint32_t m = 0;
  bool bad = false;
  for (int32_t e = rowptr[row]; e < rowptr[row + 1]; ++e) {
    const int64_t t = (int64_t)colidx[e] - row - shift + P + L + 1;
    const int64_t a = t / P, rem = t % P, b = rem / L, c = rem % L;
    if (t < 0 || a > 2 || b > 2 || c > 2) {
      bad = true;
      break;
    }
    const int slot = (int)(a * 9 + b * 3 + c);
    dia[(int64_t)slot * nrows_pad + row] = vals[e];
    m |= 1u << slot;
  }
  mask[row] = m;
  if (bad) atomicExch(veto, 1);
}

// Reads one row from the middle of the mirror and, if it has the 27-offset shape, builds the slot
// storage and verifies every row against it.
static int brick_plan(slc_matrix *m) {
  BrickPlan &bp = m->brick;
  if (bp.state >= 0) return SLC_OK;
  bp.state = 0;
  if (m->elem != ELEM_D || !m->has_csr || m->max_rowlen != 27 || m->nrows < 1024) return SLC_OK;
  cudaStream_t s = cur_stream();
  // a row with all 27 entries near the middle of the matrix (the very middle row of an even grid sits on a
  // grid face): scan the pointers of a window of rows
  const int32_t w0 = m->nrows / 2, wn = std::min<int32_t>(8192, m->nrows - w0);
  std::vector<int32_t> win((size_t)wn + 1);
  SLC_CUDA(cudaMemcpyAsync(win.data(), m->rowptr + w0, win.size() * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  int32_t r = -1;
  int32_t rp[2] = {0, 0};
  for (int32_t i = 0; i < wn && r < 0; ++i)
    if (win[(size_t)i + 1] - win[(size_t)i] == 27) {
      r = w0 + i;
      rp[0] = win[(size_t)i];
      rp[1] = win[(size_t)i + 1];
    }
  if (r < 0) return SLC_OK;
  int32_t cols[27];
  SLC_CUDA(cudaMemcpyAsync(cols, m->colidx + rp[0], sizeof cols, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  const int64_t L = (int64_t)cols[16] - cols[13], P = (int64_t)cols[22] - cols[13];
  if (cols[14] - cols[13] != 1 || L < 4 || P < 3 * L || P % L != 0 || P > m->ncols) return SLC_OK;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b)
      for (int c = 0; c < 3; ++c)
        if ((int64_t)cols[a * 9 + b * 3 + c] - cols[13] != (a - 1) * P + (b - 1) * L + (c - 1)) return SLC_OK;
  const int64_t shift = (int64_t)cols[13] - r;  // column of the centre entry minus the row: 0 for a square matrix
  if (shift < 0) return SLC_OK;
  const int64_t nrows_pad = ceil_div(m->nrows, 32) * 32;
  DevBuf<double> dia;
  DevBuf<uint32_t> mask;
  DevBuf<int> veto;
  SLC_TRY(dia.alloc((size_t)27 * nrows_pad));
  SLC_TRY(mask.alloc((size_t)nrows_pad));
  SLC_TRY(veto.alloc(1));
  SLC_CUDA(cudaMemsetAsync(dia.p, 0, (size_t)27 * nrows_pad * sizeof(double), s));
  SLC_CUDA(cudaMemsetAsync(mask.p, 0, (size_t)nrows_pad * sizeof(uint32_t), s));
  SLC_CUDA(cudaMemsetAsync(veto.p, 0, sizeof(int), s));
  k_brick_build<double><<<(unsigned)ceil_div(m->nrows, 256), 256, 0, s>>>(
      m->nrows, nrows_pad, shift, (int32_t)L, (int32_t)P, m->rowptr, m->colidx, (const double *)m->rvals, dia.p,
      mask.p, veto.p);
  SLC_LAUNCHED();
  int h = 0;
  SLC_CUDA(cudaMemcpyAsync(&h, veto.p, sizeof h, cudaMemcpyDeviceToHost, s));
  SLC_CUDA(cudaStreamSynchronize(s));
  if (h) return SLC_OK;  // some row is off the pattern: the general kernels serve this matrix
  bp.L = (int32_t)L;
  bp.P = (int32_t)P;
  bp.shift = shift;
  bp.nrows_pad = nrows_pad;
  bp.dia = dia.take();
  bp.mask = mask.take();
  bp.state = 1;
  return SLC_OK;
}

// ---- mbarrier / bulk-copy primitives (see slc_spmm_window.cu) ------------------------------------------

__device__ __forceinline__ uint32_t pl_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pl_mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pl_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void pl_mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(pl_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pl_mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pl_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool pl_mbar_wait(uint64_t *bar, uint32_t parity, int *err) {
  for (uint32_t spin = 0; spin < PL_SPIN_LIMIT; ++spin) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(pl_smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return true;
  }
  *(volatile int *)err = 3;  // mapped host word: a plain store
  return false;
}
__device__ __forceinline__ void pl_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(pl_smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(pl_smem_u32(bar))
               : "memory");
}

// ---- the product ---------------------------------------------------------------------------------------
// Block b owns rows [row_begin + b PL_ROWS, ...) of the launch's range.  Window a (a = 0, 1, 2) holds the
// entries of X[:, j] with index in [lo_a, lo_a + W): lo_a = first row + shift + (a - 1) P - L - 1 rounded
// down to an even index (16-byte alignment of the copy), clipped to the entries of X that exist: [0, ncols),
// or the owned slice when the interior rows of a partitioned matrix run here.  Thread t reads slot
// (a, b, c) of its row at window index t + d_a + b L + c, d_a = the rounding.

// the 27 terms of one column tile, X read from the staged windows
template <int JT, bool FULL>
__device__ __forceinline__ void planes_tile(const double (&a)[27], const uint32_t msk, const double *__restrict__ xs,
                                            const int32_t W, const int32_t L, const int32_t (&off)[3], const int jn,
                                            double (&acc)[JT]) {
#pragma unroll
  for (int pa = 0; pa < 3; ++pa) {
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const double *xl = xs + (int64_t)pa * W + off[pa] + b * L;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int slot = pa * 9 + b * 3 + c;
        // rows with holes: a branch per slot.  The branch-free form (unconditional load and product,
        // predicated sum) compiled to FSEL pairs and measured 5.33 ms against 4.35 ms (round 2, pass 8)
        if (FULL || ((msk >> slot) & 1u)) {
#pragma unroll
          for (int j = 0; j < JT; ++j)
            if (FULL || j < jn) acc[j] = __dadd_rn(__dmul_rn(a[slot], xl[(int64_t)j * 3 * W + c]), acc[j]);
        }
      }
    }
  }
}

template <int JT, int PL_ROWS>
__global__ void __launch_bounds__(PL_ROWS + 32, 512 / PL_ROWS)
k_spmm_planes(int32_t row_begin, int32_t row_end, int64_t shift, int32_t L, int32_t P, int64_t nrows_pad, int64_t xlo,
              int64_t xhi, int32_t nstage, int32_t W, const double *__restrict__ dia, const uint32_t *__restrict__ mask,
              const double *__restrict__ X, int64_t ldx, double *__restrict__ Y, int64_t ldy, int32_t k,
              const YInit<double> yi, int *__restrict__ err) {
  extern __shared__ __align__(128) unsigned char pl_smem[];
  double *stage0 = reinterpret_cast<double *>(pl_smem);                                  // [nstage][JT][3][W]
  const size_t stage_elems = (size_t)JT * 3 * W;
  uint64_t *full = reinterpret_cast<uint64_t *>(pl_smem + (size_t)nstage * stage_elems * sizeof(double));  // [nstage]
  uint64_t *empty = full + nstage;                                                       // [nstage]
  const int tid = threadIdx.x;
  const int64_t row0 = (int64_t)row_begin + (int64_t)blockIdx.x * PL_ROWS;
  if (tid == 0) {
    for (int st = 0; st < nstage; ++st) {
      pl_mbar_init(full + st, 1);
      pl_mbar_init(empty + st, PL_ROWS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // window a: [lo[a], lo[a] + W) in X's row numbering, before clipping
  int64_t lo[3];
#pragma unroll
  for (int pa = 0; pa < 3; ++pa) {
    const int64_t l = row0 + shift + (int64_t)(pa - 1) * P - L - 1;
    lo[pa] = l - (l & 1);
  }
  const int32_t ntiles = (k + JT - 1) / JT;

  if (tid >= PL_ROWS) {
    // ---- producer: one lane feeds the ring -------------------------------------------------------
    if ((tid & 31) != 0) return;
    int64_t c0[3], c1[3];  // clipped copy range of every window
    uint32_t bytes_col = 0;
#pragma unroll
    for (int pa = 0; pa < 3; ++pa) {
      c0[pa] = max(lo[pa], xlo);  // [xlo, xhi): the entries of X that exist (even bounds)
      c1[pa] = min(lo[pa] + W, xhi);
      if (c1[pa] > c0[pa]) bytes_col += (uint32_t)(c1[pa] - c0[pa]) * (uint32_t)sizeof(double);
    }
    int32_t st = 0, use = 0;
    for (int32_t t = 0; t < ntiles; ++t) {
      const int32_t jn = min(JT, k - t * JT);
      if (use > 0 && !pl_mbar_wait(empty + st, (uint32_t)((use - 1) & 1), err)) return;
      pl_mbar_arrive_expect_tx(full + st, bytes_col * (uint32_t)jn);
      double *dst = stage0 + (size_t)st * stage_elems;
      for (int32_t j = 0; j < jn; ++j) {
        const double *xc = X + (int64_t)(t * JT + j) * ldx;
#pragma unroll
        for (int pa = 0; pa < 3; ++pa)
          if (c1[pa] > c0[pa])
            pl_bulk_g2s(dst + ((size_t)j * 3 + pa) * W + (c0[pa] - lo[pa]), xc + c0[pa],
                        (uint32_t)(c1[pa] - c0[pa]) * (uint32_t)sizeof(double), full + st);
      }
      if (bytes_col == 0) pl_mbar_arrive(full + st);  // nothing to copy: cannot happen for a row that has entries
      if (++st == nstage) {
        st = 0;
        ++use;
      }
    }
    return;
  }

  // ---- consumers ---------------------------------------------------------------------------------------
  const uint64_t pol_s = policy_evict_first();
  const int64_t row = row0 + tid;
  const bool live = row < row_end;
  const uint32_t msk = live ? __ldg(mask + row) : 0u;
  double a[27];
#pragma unroll
  for (int s = 0; s < 27; ++s) a[s] = live ? ld_stream(dia + (int64_t)s * nrows_pad + row, pol_s) : 0.0;
  int32_t off[3];
#pragma unroll
  for (int pa = 0; pa < 3; ++pa) off[pa] = (int32_t)(row + shift + (int64_t)(pa - 1) * P - L - 1 - lo[pa]);
  const bool full_warp = __all_sync(0xffffffffu, msk == BRICK_FULL);
  int32_t st = 0;
  uint32_t parity = 0;
  for (int32_t t = 0; t < ntiles; ++t) {
    const int32_t j0 = t * JT, jn = min(JT, k - j0);
    double acc[JT];
#pragma unroll
    for (int j = 0; j < JT; ++j) acc[j] = (live && j < jn) ? yi(Y + (int64_t)(j0 + j) * ldy + row) : 0.0;
    if (!pl_mbar_wait(full + st, parity, err)) return;
    const double *xs = stage0 + (size_t)st * stage_elems;
    if (full_warp && jn == JT) planes_tile<JT, true>(a, msk, xs, W, L, off, jn, acc);
    else planes_tile<JT, false>(a, msk, xs, W, L, off, jn, acc);
    pl_mbar_arrive(empty + st);  // this thread is done reading the stage
#pragma unroll
    for (int j = 0; j < JT; ++j)
      if (live && j < jn) st_stream(Y + (int64_t)(j0 + j) * ldy + row, acc[j]);
    if (++st == nstage) {
      st = 0;
      parity ^= 1u;
    }
  }
}

int spmm_window_err_word(int **out);  // slc_spmm_window.cu: the mapped host word both staged kernels raise

template <int JT, int PL_ROWS>
static int launch_planes(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                         const YInit<double> &yi, int32_t row_begin, int32_t row_end, int64_t xlo, int64_t xhi,
                         int32_t nstage, int32_t W) {
  const BrickPlan &bp = m->brick;
  const size_t smem = (size_t)nstage * JT * 3 * W * sizeof(double) + (size_t)nstage * 2 * sizeof(uint64_t);
  auto kern = k_spmm_planes<JT, PL_ROWS>;
  static PerDevice<size_t> smem_set;  // per instantiation, per device
  SLC_CUDA(ensure_dyn_smem(kern, smem_set, smem));
  int *err_flag = nullptr;
  SLC_TRY(spmm_window_err_word(&err_flag));
  const int64_t nb = ceil_div((int64_t)row_end - row_begin, PL_ROWS);
  kern<<<(unsigned)nb, PL_ROWS + 32, smem, cur_stream()>>>(row_begin, row_end, bp.shift, bp.L, bp.P, bp.nrows_pad, xlo, xhi,
                                                        nstage, W, bp.dia, bp.mask, X, ldx, Y, ldy, k, yi, err_flag);
  SLC_LAUNCHED();
  return SLC_OK;
}

int spmm_brick_prepare(slc_matrix *m) {
  const char *env = getenv("SLC_SPMM_BRICK");
  if ((env && env[0] == '0') || m->elem != ELEM_D) return SLC_OK;
  return brick_plan(m);
}

// Runs rows [row_begin, row_end) of the block product on the staged kernel when the matrix has the plan
// and the operands qualify (X 16-byte aligned, even column stride and bounds, windows that fit).  X is
// indexed by column number and only its entries [xlo, xhi) exist (xhi < 0: all of [0, ncols)).
int spmm_brick_try(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                   const YInit<double> &yi, int32_t row_begin, int32_t row_end, int64_t xlo, int64_t xhi, bool *done) {
  *done = false;
  if (xhi < 0) xhi = m->ncols;
  const char *env = getenv("SLC_SPMM_BRICK");  // read on every call so a test can switch it
  if (env && env[0] == '0') return SLC_OK;
  if (k < 1 || row_end <= row_begin) return SLC_OK;
  if (((uintptr_t)X & 15) != 0 || (ldx & 1) != 0 || (xlo & 1) != 0 || (xhi & 1) != 0) return SLC_OK;
  SLC_TRY(brick_plan(m));
  if (m->brick.state != 1) return SLC_OK;
  // SLC_PLANES_ROWS / _JT / _STAGES: rows per block (512 or 256), columns per stage, stages; read on every
  // call so a test can switch them.  Measured on the 200^3 stencil, k = 64: see DESIGN.md 4.3.
  const int rows_env = env_int("SLC_PLANES_ROWS", 512);
  const int jt_env = env_int("SLC_PLANES_JT", 1);
  const int stage_env = env_int("SLC_PLANES_STAGES", 0);
  const int rows = rows_env <= 256 ? 256 : 512;
  const int32_t W = (rows + 2 * m->brick.L + 4 + 1) / 2 * 2;  // window entries incl. alignment slack, even
  const size_t col_bytes = (size_t)3 * W * sizeof(double);
  const size_t budget = (rows == 256 ? 100 : 200) * 1024;      // per block: two blocks per SM at 256 rows
  int jt = jt_env >= 4 ? 4 : jt_env >= 2 ? 2 : 1;
  while (jt > 1 && budget / (col_bytes * jt) < 3) jt >>= 1;
  int32_t nstage = (int32_t)std::min<size_t>(8, budget / (col_bytes * jt));
  if (nstage < 2) return SLC_OK;  // lines too long for shared memory: the panel kernel serves the product
  if (stage_env >= 2) nstage = std::min(nstage, stage_env);
  nstage = std::min(nstage, std::max(2, (k + jt - 1) / jt));
#define SLC_PLANES_CASE(JT_, ROWS_) \
  SLC_TRY((launch_planes<JT_, ROWS_>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end, xlo, xhi, nstage, W)))
  if (rows == 256) {
    if (jt == 4) SLC_PLANES_CASE(4, 256);
    else if (jt == 2) SLC_PLANES_CASE(2, 256);
    else SLC_PLANES_CASE(1, 256);
  } else {
    if (jt == 4) SLC_PLANES_CASE(4, 512);
    else if (jt == 2) SLC_PLANES_CASE(2, 512);
    else SLC_PLANES_CASE(1, 512);
  }
#undef SLC_PLANES_CASE
  *done = true;
  return SLC_OK;
}

}  // namespace slc
