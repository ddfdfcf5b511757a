/*
 * oracle.h -- CPU restatement of sparse-linear's compressed-sparse hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it, and only as the
 * checker (or as the timed CPU baseline).  The product path (sparse-linear_b200/) never
 * links, imports or calls anything in this directory.
 *
 * Parity status: the reference is Haskell and no GHC exists in this image, so the
 * reference itself cannot be run here (oracle/_ref cannot be built).  The restatement
 * is pinned against every known-answer case the reference's tests hold for this path
 * (sparse-linear/tests/Sparse.hs:61-73 ctrans vectors; the QuickCheck properties at
 * Sparse.hs:23-165 and Test/LinearAlgebra.hs:29-67 re-run on the oracle) and is
 * cross-checked against scipy as an independent second opinion.  One item stays
 * "parity unpinned": the order in which duplicate triplets are summed in `compress`
 * (the reference uses an unstable introsort from vector-algorithms, version unpinned;
 * Data/Matrix/Sparse.hs:219,262,279).  The oracle fixes that order as input order.
 *
 * Every function cites the reference lines it follows.  Paths are relative to
 * /root/reference/sparse-linear/src unless stated otherwise.
 *
 * Conventions (Data/Matrix/Sparse/Foreign.hs:24-41): CSC with 32-bit indices,
 * colptr[ncols+1], rowidx[nnz] strictly increasing inside each column, values as
 * doubles; Complex Double is interleaved (re, im) so a `z` array of n elements is
 * 2n doubles.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t nrows, ncols;
  int64_t nnz;
  int32_t *colptr; /* ncols+1 */
  int32_t *rowidx; /* nnz */
  double *vals;    /* nnz (d) or 2*nnz (z) */
} orc_csc;

void orc_csc_free(orc_csc *m);

/* error codes shared by all entry points */
#define ORC_OK 0
#define ORC_ERR_DIM (-1)
#define ORC_ERR_ROW_BOUNDS (-2)
#define ORC_ERR_COL_BOUNDS (-3)
#define ORC_ERR_ALLOC (-4)

#define ORC_DECLARE(P)                                                                      \
  /* axpy_: y[r] <- a*x[c] + y[r], columns in order (Data/Matrix/Sparse.hs:433-453) */      \
  int orc_##P##_axpy(int32_t nrows, int32_t ncols, const int32_t *colptr,                   \
                     const int32_t *rowidx, const double *vals, const double *x,            \
                     double *y);                                                            \
  /* mulV: y <- 0; axpy_ (Sparse.hs:464-471) */                                             \
  int orc_##P##_mulv(int32_t nrows, int32_t ncols, const int32_t *colptr,                   \
                     const int32_t *rowidx, const double *vals, const double *x,            \
                     double *y);                                                            \
  /* SpMM as FEAST drives it: per column, set 0 then axpy_ (feast/.../Feast.hs:203-208;     \
     mulM Sparse.hs:473-498).  X, Y column-major with leading dimensions ldx, ldy. */       \
  int orc_##P##_mulm(int32_t nrows, int32_t ncols, const int32_t *colptr,                   \
                     const int32_t *rowidx, const double *vals, const double *X,            \
                     int64_t ldx, double *Y, int64_t ldy, int32_t k);                       \
  /* transpose / ctrans (Sparse.hs:301-329, 371-375) */                                     \
  int orc_##P##_transpose(int32_t nrows, int32_t ncols, const int32_t *colptr,              \
                          const int32_t *rowidx, const double *vals, int conj,              \
                          orc_csc *out);                                                    \
  /* hermitian m = ctrans m == m (Sparse.hs:377-379); returns 1/0 */                        \
  int orc_##P##_hermitian(int32_t nrows, int32_t ncols, const int32_t *colptr,              \
                          const int32_t *rowidx, const double *vals);                       \
  /* mm, literal accumulator: full reset + three full scans per column                      \
     (Sparse.hs:691-702, Data/Vector/Sparse/ScatterGather.hs:37-147) */                     \
  int orc_##P##_mm_literal(int32_t ar, int32_t ac, const int32_t *ap, const int32_t *ai,    \
                           const double *av, int32_t br, int32_t bc, const int32_t *bp,     \
                           const int32_t *bi, const double *bv, orc_csc *out);              \
  /* mm, output-identical fast accumulator (touched list instead of full scans) */         \
  int orc_##P##_mm(int32_t ar, int32_t ac, const int32_t *ap, const int32_t *ai,            \
                   const double *av, int32_t br, int32_t bc, const int32_t *bp,             \
                   const int32_t *bi, const double *bv, orc_csc *out);                      \
  /* compress / fromTriples (Sparse.hs:184-291,357-369).  On a bounds error *bad_pos        \
     receives the first offending position (rows are checked before columns). */           \
  int orc_##P##_compress(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows,      \
                         const int32_t *cols, const double *vals, orc_csc *out,             \
                         int64_t *bad_pos);                                                 \
  /* lin alpha A beta B (Sparse.hs:401-431): union pattern, zeros kept */                   \
  int orc_##P##_lin(const double *alpha, int32_t ar, int32_t ac, const int32_t *ap,         \
                    const int32_t *ai, const double *av, const double *beta, int32_t br,    \
                    int32_t bc, const int32_t *bp, const int32_t *bi, const double *bv,     \
                    orc_csc *out);                                                          \
  /* lin through the literal full-length accumulator (for the agreement test) */           \
  int orc_##P##_lin_literal(const double *alpha, int32_t ar, int32_t ac,                    \
                            const int32_t *ap, const int32_t *ai, const double *av,         \
                            const double *beta, int32_t br, int32_t bc, const int32_t *bp,  \
                            const int32_t *bi, const double *bv, orc_csc *out);             \
  /* kronecker (Sparse.hs:597-638) */                                                       \
  int orc_##P##_kronecker(int32_t ar, int32_t ac, const int32_t *ap, const int32_t *ai,     \
                          const double *av, int32_t br, int32_t bc, const int32_t *bp,      \
                          const int32_t *bi, const double *bv, orc_csc *out);

ORC_DECLARE(d)
ORC_DECLARE(z)

/* decompress: expand column pointers to one column id per entry (Sparse.hs:293-299) */
void orc_decompress(int32_t ncols, const int32_t *colptr, int32_t *cols_out);

/* format properties checked by the reference's checkMatrix
   (sparse-linear/tests/Test/LinearAlgebra.hs:40-67); returns 0 when all hold,
   else the 1-based number of the first failed property */
int orc_check_format(int32_t nrows, int32_t ncols, int64_t nnz, const int32_t *colptr,
                     const int32_t *rowidx);

#ifdef __cplusplus
}
#endif
#endif
