/*
 * oracle.c -- CPU restatement of sparse-linear's compressed-sparse hot path.
 * TEST INFRASTRUCTURE ONLY: see oracle.h for who may use it and for the parity status.
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off -shared -fPIC)
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

void orc_csc_free(orc_csc *m) {
  if (!m) return;
  free(m->colptr);
  free(m->rowidx);
  free(m->vals);
  m->colptr = NULL;
  m->rowidx = NULL;
  m->vals = NULL;
}

static int cmp_i32(const void *a, const void *b) {
  const int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return (x > y) - (x < y);
}

/* sort a list of distinct row indices ascending (the fast accumulator's stand-in for
   the reference's in-order scan of the whole workspace, ScatterGather.hs:120-132) */
static void orc_sort_i32(int32_t *a, int64_t n) {
  if (n <= 32) {
    for (int64_t i = 1; i < n; ++i) {
      int32_t v = a[i];
      int64_t j = i;
      while (j > 0 && a[j - 1] > v) {
        a[j] = a[j - 1];
        --j;
      }
      a[j] = v;
    }
  } else {
    qsort(a, (size_t)n, sizeof(int32_t), cmp_i32);
  }
}

/* ---- Double (Data/Complex/Enhanced.hs:35-41: conj = id) --------------------------- */
#define P d
#define T double
#define W 1
#define T_ZERO 0.0
#define T_ADD(a, b) ((a) + (b))
#define T_SUB(a, b) ((a) - (b))
#define T_MUL(a, b) ((a) * (b))
#define T_CONJ(a) (a)
#define T_EQ(a, b) ((a) == (b))
#include "oracle_impl.inc"
#undef P
#undef T
#undef W
#undef T_ZERO
#undef T_ADD
#undef T_SUB
#undef T_MUL
#undef T_CONJ
#undef T_EQ

/* ---- Complex Double (Enhanced.hs:43-49: conj = conjugate) -------------------------
   Data.Complex: (x:+y) * (x':+y') = (x*x' - y*y') :+ (x*y' + y*x'); + and - are
   componentwise; == compares both parts. */
typedef struct {
  double re, im;
} orc_zc;
static inline orc_zc zc_make(double re, double im) {
  orc_zc r;
  r.re = re;
  r.im = im;
  return r;
}
static inline orc_zc zc_add(orc_zc a, orc_zc b) { return zc_make(a.re + b.re, a.im + b.im); }
static inline orc_zc zc_sub(orc_zc a, orc_zc b) { return zc_make(a.re - b.re, a.im - b.im); }
static inline orc_zc zc_mul(orc_zc a, orc_zc b) {
  return zc_make(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re);
}
static inline orc_zc zc_conj(orc_zc a) { return zc_make(a.re, -a.im); }

#define P z
#define T orc_zc
#define W 2
#define T_ZERO zc_make(0.0, 0.0)
#define T_ADD(a, b) zc_add((a), (b))
#define T_SUB(a, b) zc_sub((a), (b))
#define T_MUL(a, b) zc_mul((a), (b))
#define T_CONJ(a) zc_conj(a)
#define T_EQ(a, b) ((a).re == (b).re && (a).im == (b).im)
#include "oracle_impl.inc"

/* decompress (Data/Matrix/Sparse.hs:293-299): set slice c of the output to c */
void orc_decompress(int32_t ncols, const int32_t *colptr, int32_t *cols_out) {
  for (int32_t c = 0; c < ncols; ++c)
    for (int32_t p = colptr[c]; p < colptr[c + 1]; ++p) cols_out[p] = c;
}

/* checkMatrix (sparse-linear/tests/Test/LinearAlgebra.hs:40-67) */
int orc_check_format(int32_t nrows, int32_t ncols, int64_t nnz, const int32_t *colptr,
                     const int32_t *rowidx) {
  for (int32_t c = 0; c < ncols; ++c)
    if (colptr[c] > colptr[c + 1]) return 1; /* nondecreasing pointers */
  if (colptr[0] != 0) return 2;
  if ((int64_t)colptr[ncols] != nnz) return 3; /* length values == last pointers */
  for (int32_t c = 0; c < ncols; ++c)
    for (int32_t p = colptr[c] + 1; p < colptr[c + 1]; ++p)
      if (rowidx[p - 1] >= rowidx[p]) return 5; /* increasing indices in slice */
  for (int64_t p = 0; p < nnz; ++p) {
    if (rowidx[p] < 0) return 6;
    if (rowidx[p] >= nrows) return 7;
  }
  return 0;
}
