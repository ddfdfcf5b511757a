// slc_spmm_brick.cu -- block product Y[:, j] <- A X[:, j] for matrices whose rows all follow one
// 3 x 3 x 3 offset pattern (27-point stencils on a grid numbered x-fastest: BASELINE.json config 5), the
// FEAST-shaped op of Feast.hs:203-208 on the headline matrix.
//
// What round 1's k_spmm_panel ran into (ncu, profiles/r1g_ncu_full_spmm_panel_cfg5mm.txt): one gather of X
// per product -- 443 M warp requests of ~2.9 L1 lines each, 38 GB from L2 to L1, 3.5 x the algorithmic
// bytes -- and A re-read from shared memory for every product.  The ceiling was the L1 tag stage, not HBM.
// This kernel removes loads instead of scheduling them better:
//
//   * the matrix is detected to have every row's columns in  row + shift + {a P + b L + c : a, b, c in
//     {-1, 0, 1}}  (P, L = plane and line strides, read off one row and then VERIFIED for all rows by the
//     plan builder; anything else keeps the panel kernel).  The values are stored by offset slot,
//     dia[slot][row] -- no column indices at all -- plus one 27-bit presence mask per row (grid
//     boundaries drop entries; an absent entry must not be added, not even as 0 * x);
//   * a thread owns R rows that are neighbours in y (row, row + L, ...) and keeps their R x 27 values in
//     REGISTERS for all k columns (A is read from HBM exactly once and never again, from any memory);
//   * the rows of a thread share X: lines y-1 .. y+R of the three planes, 9 (R + 2) loads for 27 R
//     products (0.67 loads per product at R = 2, 0.56 at R = 3, against 1);
//   * a block is a brick of 4 planes x (2 R lines) x 32 grid points, so the (6 x (2R+2) x 34)-point box of X
//     it touches is fetched from L2 once (L1 serves the reuse between the block's warps): ~2.4 x the
//     algorithmic bytes of X instead of 9 x.
// Every (row, column) is still owned by one thread which adds the row's terms in increasing column
// order (plane, then line, then point), product and sum rounded separately: bit-identical to the
// reference (Sparse.hs:433-453), checked against the oracle at full size.
#include <algorithm>
#include <vector>

#include "slc_common.cuh"

namespace slc {

constexpr uint32_t BRICK_FULL = (1u << 27) - 1u;
constexpr int BRICK_BZ = 4;        // planes per block (warp & 3)
constexpr int BRICK_YG = 2;        // y-groups per block (warp >> 2)
constexpr int BRICK_THREADS = 32 * BRICK_BZ * BRICK_YG;

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
  static const int env = env_int("SLC_SPMM_BRICK", 1);
  if (!env || m->elem != ELEM_D || !m->has_csr || m->max_rowlen != 27 || m->nrows < 1024) return SLC_OK;
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

// ---- the product ---------------------------------------------------------------------------------------
// Thread (warp w, lane): plane z = z0 + (w & 3), lines y0 + (w >> 2) R + i for i < R, point x = x0 + lane.
// Global row index g = z P + y L + x; the stored row is g - shift (row blocks of a partitioned matrix keep
// their global column numbering).  Rows outside [row_begin, row_end) are left alone: a distributed product
// runs its boundary planes elsewhere.
template <int R, int JT, bool FULL>
__device__ __forceinline__ void brick_columns(const double (&a)[R][27], const uint32_t (&msk)[R], const bool (&live)[R],
                                              const int64_t g0, const int32_t L, const int32_t P, const int64_t ncols,
                                              const double *__restrict__ X, const int64_t ldx, double *__restrict__ Y,
                                              const int64_t ldy, const int64_t (&row)[R], const int32_t k,
                                              const YInit<double> &yi, const uint64_t pol_x) {
  for (int32_t j0 = 0; j0 < k; j0 += JT) {
    const int32_t jn = min(JT, k - j0);  // columns of this tile (JT except in the last one)
    double acc[R][JT];
#pragma unroll
    for (int i = 0; i < R; ++i)
#pragma unroll
      for (int j = 0; j < JT; ++j) acc[i][j] = (live[i] && j < jn) ? yi(Y + (int64_t)(j0 + j) * ldy + row[i]) : 0.0;
    const double *xt = X + (int64_t)j0 * ldx;
#pragma unroll
    for (int pa = 0; pa < 3; ++pa) {
#pragma unroll
      for (int bq = 0; bq < R + 2; ++bq) {  // line y0 + bq - 1 of plane z + pa - 1
        const int64_t col0 = g0 + (int64_t)(pa - 1) * P + (int64_t)(bq - 1) * L;  // the point straight below/above
        double xv[3][JT];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          int64_t col = col0 + (c - 1);
          if (!FULL) col = min(max(col, (int64_t)0), ncols - 1);  // an absent entry may point outside X: never dereferenced out of range
#pragma unroll
          for (int j = 0; j < JT; ++j) xv[c][j] = (FULL || j < jn) ? ld_gather(xt + (int64_t)j * ldx + col, pol_x) : 0.0;
        }
#pragma unroll
        for (int i = 0; i < R; ++i) {
          const int b = bq - i;  // this line is offset b - 1 of row i
          if (b < 0 || b > 2) continue;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int slot = pa * 9 + b * 3 + c;
            if (FULL || ((msk[i] >> slot) & 1u)) {
#pragma unroll
              for (int j = 0; j < JT; ++j) acc[i][j] = __dadd_rn(__dmul_rn(a[i][slot], xv[c][j]), acc[i][j]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < R; ++i)
#pragma unroll
      for (int j = 0; j < JT; ++j)
        if (live[i] && j < jn) st_stream(Y + (int64_t)(j0 + j) * ldy + row[i], acc[i][j]);
  }
}

template <int R, int JT>
__global__ void __launch_bounds__(BRICK_THREADS, 1)
k_spmm_brick(int32_t row_begin, int32_t row_end, int64_t shift, int32_t L, int32_t P, int32_t Ny, int64_t z_first,
             int32_t nbx, int32_t nby, int64_t nrows_pad, int64_t ncols, const double *__restrict__ dia,
             const uint32_t *__restrict__ mask, const double *__restrict__ X, int64_t ldx, double *__restrict__ Y,
             int64_t ldy, int32_t k, const YInit<double> yi) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int32_t bx = blockIdx.x % nbx, by = (blockIdx.x / nbx) % nby;
  const int64_t bz = blockIdx.x / ((int64_t)nbx * nby);
  const int64_t z = z_first + bz * BRICK_BZ + (w & (BRICK_BZ - 1));
  const int32_t y0 = (by * BRICK_YG + (w >> 2)) * R;
  const int32_t x = bx * 32 + lane;
  const int64_t g0 = z * P + (int64_t)y0 * L + x;
  const uint64_t pol_x = policy_evict_last();
  const uint64_t pol_s = policy_evict_first();
  double a[R][27];
  uint32_t msk[R];
  bool live[R];
  int64_t row[R];
  bool all_full = true;
#pragma unroll
  for (int i = 0; i < R; ++i) {
    row[i] = g0 + (int64_t)i * L - shift;
    live[i] = x < L && y0 + i < Ny && row[i] >= row_begin && row[i] < row_end;
    msk[i] = live[i] ? __ldg(mask + row[i]) : 0u;
    all_full = all_full && msk[i] == BRICK_FULL;
#pragma unroll
    for (int s = 0; s < 27; ++s) a[i][s] = live[i] ? ld_stream(dia + (int64_t)s * nrows_pad + row[i], pol_s) : 0.0;
  }
  bool any = false;
#pragma unroll
  for (int i = 0; i < R; ++i) any = any || live[i];
  if (!__any_sync(0xffffffffu, any)) return;
  // interior warps (every lane has all 27 entries in all its rows, and whole column tiles): no predicates
  if (__all_sync(0xffffffffu, all_full) && k % JT == 0)
    brick_columns<R, JT, true>(a, msk, live, g0, L, P, ncols, X, ldx, Y, ldy, row, k, yi, pol_x);
  else
    brick_columns<R, JT, false>(a, msk, live, g0, L, P, ncols, X, ldx, Y, ldy, row, k, yi, pol_x);
}

template <int R, int JT>
static int launch_brick(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                        const YInit<double> &yi, int32_t row_begin, int32_t row_end) {
  const BrickPlan &bp = m->brick;
  const int32_t Ny = bp.P / bp.L;
  const int64_t z_first = (row_begin + bp.shift) / bp.P, z_last = (row_end - 1 + bp.shift) / bp.P;
  const int32_t nbx = (int32_t)ceil_div(bp.L, 32), nby = (int32_t)ceil_div(Ny, BRICK_YG * R);
  const int64_t nbz = ceil_div(z_last - z_first + 1, BRICK_BZ);
  const int64_t nb = (int64_t)nbx * nby * nbz;
  if (nb > 0x7fffffffLL) return fail(SLC_ERR_RANGE, "mulM: grid too large");
  k_spmm_brick<R, JT><<<(unsigned)nb, BRICK_THREADS, 0, cur_stream()>>>(
      row_begin, row_end, bp.shift, bp.L, bp.P, Ny, z_first, nbx, nby, bp.nrows_pad, m->ncols, bp.dia, bp.mask, X, ldx, Y,
      ldy, k, yi);
  SLC_LAUNCHED();
  return SLC_OK;
}

// Runs rows [row_begin, row_end) of the block product on the brick kernel when the matrix has the plan.
int spmm_brick_try(slc_matrix *m, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
                   const YInit<double> &yi, int32_t row_begin, int32_t row_end, bool *done) {
  *done = false;
  SLC_TRY(brick_plan(m));
  if (m->brick.state != 1 || k < 1 || row_end <= row_begin) return SLC_OK;
  static const int r_env = env_int("SLC_BRICK_R", 2);
  static const int jt_env = env_int("SLC_BRICK_JT", 0);
  const int jt = jt_env ? jt_env : (r_env == 3 ? 2 : 4);
  if (r_env == 3 && jt <= 2) SLC_TRY((launch_brick<3, 2>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end)));
  else if (r_env == 3) SLC_TRY((launch_brick<3, 4>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end)));
  else if (r_env == 1) SLC_TRY((launch_brick<1, 8>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end)));
  else if (jt <= 2) SLC_TRY((launch_brick<2, 2>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end)));
  else if (jt >= 8) SLC_TRY((launch_brick<2, 8>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end)));
  else SLC_TRY((launch_brick<2, 4>(m, X, ldx, Y, ldy, k, yi, row_begin, row_end)));
  *done = true;
  return SLC_OK;
}

}  // namespace slc
