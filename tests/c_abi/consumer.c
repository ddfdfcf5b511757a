/* consumer.c -- a plain-C11 caller of include/slc.h: no Python, no torch, no C++.
 *
 * It drives the C ABI with exactly the argument shapes the Haskell shim passes
 * (sparse-linear/src/Data/Matrix/Sparse/Foreign.hs:24-41: CInt dimensions, Ptr CInt colptr / rowidx,
 * Ptr a values, arrays lent for the call; results copied out into caller-owned buffers, :43-55; handles
 * released by slc_free as a finalizer, Umfpack/Internal.hs:110-113; int status + message, Umfpack.hs:64-67).
 *
 *   consumer <in.bin> <out.bin>
 *
 * in.bin is written by tests/test_c_abi.py (a matrix, vectors, a block, triplets, per element type);
 * out.bin receives every result, and the test compares them with the oracle.  The record format is
 * a flat sequence of arrays, each preceded by its int64 length in elements.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "slc.h"

static FILE *fin, *fout;

#define CHECK(call)                                                                      \
  do {                                                                                   \
    int rc_ = (call);                                                                    \
    if (rc_ != SLC_OK) {                                                                 \
      fprintf(stderr, "%s:%d: %s -> %d: %s\n", __FILE__, __LINE__, #call, rc_, slc_last_error()); \
      exit(2);                                                                           \
    }                                                                                    \
  } while (0)

static void *read_array(size_t elem, int64_t *n_out) {
  int64_t n = 0;
  if (fread(&n, sizeof n, 1, fin) != 1 || n < 0) { fprintf(stderr, "short input\n"); exit(3); }
  void *p = malloc((size_t)(n ? n : 1) * elem);
  if (!p || (n && fread(p, elem, (size_t)n, fin) != (size_t)n)) { fprintf(stderr, "short input\n"); exit(3); }
  if (n_out) *n_out = n;
  return p;
}
static void write_array(const void *p, size_t elem, int64_t n) {
  fwrite(&n, sizeof n, 1, fout);
  if (n) fwrite(p, elem, (size_t)n, fout);
}
static void write_i64(int64_t v) { write_array(&v, sizeof v, 1); }

/* one symbol per element type, selected the way the shim's class instance does */
typedef struct {
  int w; /* doubles per element */
  int (*upload)(int32_t, int32_t, const int32_t *, const int32_t *, const double *, slc_matrix **);
  int (*download)(const slc_matrix *, int32_t *, int32_t *, double *);
  int (*mulv)(slc_matrix *, const double *, double *);
  int (*axpy)(slc_matrix *, const double *, double *);
  int (*mulv_dev)(slc_matrix *, const double *, double *);
  int (*axpy_dev)(slc_matrix *, const double *, double *);
  int (*mulm)(slc_matrix *, const double *, int64_t, double *, int64_t, int32_t, const double *);
  int (*mulm_dev)(slc_matrix *, const double *, int64_t, double *, int64_t, int32_t, const double *);
  int (*mm)(const slc_matrix *, const slc_matrix *, slc_matrix **);
  int (*transpose)(const slc_matrix *, int, slc_matrix **);
  int (*hermitian)(const slc_matrix *, int *);
  int (*compress)(int32_t, int32_t, int64_t, const int32_t *, const int32_t *, const double *, slc_matrix **, int64_t *);
  int (*compress_dev)(int32_t, int32_t, int64_t, const int32_t *, const int32_t *, const double *, slc_matrix **, int64_t *);
  int (*lin)(const double *, const slc_matrix *, const double *, const slc_matrix *, slc_matrix **);
  int (*kronecker)(const slc_matrix *, const slc_matrix *, slc_matrix **);
  int (*upload_sharded)(int32_t, const int32_t *, int32_t, int32_t, const int32_t *, const int32_t *, const double *, slc_sharded **);
  int (*mulv_sharded)(slc_sharded *, const double *, double *);
  int (*mulm_sharded)(slc_sharded *, const double *, int64_t, double *, int64_t, int32_t);
} ops;

static const ops OPS_D = {1, slc_d_upload, slc_d_download, slc_d_mulv, slc_d_axpy, slc_d_mulv_dev, slc_d_axpy_dev,
                          slc_d_mulm, slc_d_mulm_dev, slc_d_mm, slc_d_transpose, slc_d_hermitian, slc_d_compress,
                          slc_d_compress_dev, slc_d_lin, slc_d_kronecker, slc_d_upload_sharded, slc_d_mulv_sharded,
                          slc_d_mulm_sharded};
static const ops OPS_Z = {2, slc_z_upload, slc_z_download, slc_z_mulv, slc_z_axpy, slc_z_mulv_dev, slc_z_axpy_dev,
                          slc_z_mulm, slc_z_mulm_dev, slc_z_mm, slc_z_transpose, slc_z_hermitian, slc_z_compress,
                          slc_z_compress_dev, slc_z_lin, slc_z_kronecker, slc_z_upload_sharded, slc_z_mulv_sharded,
                          slc_z_mulm_sharded};

/* copy a device matrix out into caller-owned buffers (fromForeign's copy mode) and record it */
static void emit_matrix(const ops *o, const slc_matrix *m) {
  int32_t nr = 0, nc = 0;
  int64_t nnz = 0;
  CHECK(slc_dims(m, &nr, &nc, &nnz));
  int32_t *cp = malloc(((size_t)nc + 1) * sizeof *cp), *ri = malloc((size_t)(nnz ? nnz : 1) * sizeof *ri);
  double *v = malloc((size_t)(nnz ? nnz : 1) * sizeof(double) * (size_t)o->w);
  CHECK(o->download(m, cp, ri, v));
  write_i64(nr);
  write_i64(nc);
  write_array(cp, sizeof *cp, (int64_t)nc + 1);
  write_array(ri, sizeof *ri, nnz);
  write_array(v, sizeof(double), nnz * o->w);
  free(cp);
  free(ri);
  free(v);
}

static void run(const ops *o, int expect_elem) {
  const int w = o->w;
  int64_t n;
  int32_t *dims = read_array(sizeof(int32_t), &n); /* nrows, ncols, k, ldx, ldy */
  const int32_t nr = dims[0], nc = dims[1], k = dims[2], ldx = dims[3], ldy = dims[4];
  int32_t *cp = read_array(sizeof(int32_t), NULL), *ri = read_array(sizeof(int32_t), NULL);
  double *v = read_array(sizeof(double), NULL);
  double *x = read_array(sizeof(double), NULL), *y0 = read_array(sizeof(double), NULL);
  double *X = read_array(sizeof(double), NULL), *Y0 = read_array(sizeof(double), NULL);
  int64_t nt = 0;
  int32_t *tr = read_array(sizeof(int32_t), &nt), *tc = read_array(sizeof(int32_t), NULL);
  double *tv = read_array(sizeof(double), NULL);
  const size_t ybytes = (size_t)nr * (size_t)w * sizeof(double), xbytes = (size_t)nc * (size_t)w * sizeof(double);

  /* handles: upload, dims, element type, download */
  slc_matrix *a = NULL;
  CHECK(o->upload(nr, nc, cp, ri, v, &a));
  if (slc_elem_type(a) != expect_elem) { fprintf(stderr, "element type\n"); exit(4); }
  CHECK(slc_prepare(a));
  char plan[256];
  CHECK(slc_plan_info(a, plan, (int32_t)sizeof plan));
  emit_matrix(o, a);

  /* mulV, axpy_ with host vectors */
  double *y = malloc(ybytes);
  CHECK(o->mulv(a, x, y));
  write_array(y, sizeof(double), (int64_t)nr * w);
  memcpy(y, y0, ybytes);
  CHECK(o->axpy(a, x, y));
  write_array(y, sizeof(double), (int64_t)nr * w);

  /* the same with device-resident vectors */
  void *dx = NULL, *dy = NULL;
  CHECK(slc_dev_alloc(&dx, (int64_t)xbytes));
  CHECK(slc_dev_alloc(&dy, (int64_t)ybytes));
  CHECK(slc_memcpy_h2d(dx, x, (int64_t)xbytes));
  CHECK(o->mulv_dev(a, dx, dy));
  CHECK(slc_synchronize());
  CHECK(slc_memcpy_d2h(y, dy, (int64_t)ybytes));
  write_array(y, sizeof(double), (int64_t)nr * w);
  CHECK(slc_memcpy_h2d(dy, y0, (int64_t)ybytes));
  CHECK(o->axpy_dev(a, dx, dy));
  CHECK(slc_memcpy_d2h(y, dy, (int64_t)ybytes));
  write_array(y, sizeof(double), (int64_t)nr * w);

  /* mulM: whole block; then a column range chosen by pointer offset, as sliceMat does; then the
     accumulating form (beta = 1) into a block the caller filled */
  const size_t Ybytes = (size_t)ldy * (size_t)k * (size_t)w * sizeof(double);
  double *Y = malloc(Ybytes);
  memset(Y, 0, Ybytes);
  CHECK(o->mulm(a, X, ldx, Y, ldy, k, NULL));
  write_array(Y, sizeof(double), (int64_t)ldy * k * w);
  memset(Y, 0, Ybytes);
  if (k > 2) CHECK(o->mulm(a, X + (size_t)ldx * w, ldx, Y + (size_t)ldy * w, ldy, k - 2, NULL));
  write_array(Y, sizeof(double), (int64_t)ldy * k * w);
  memcpy(Y, Y0, Ybytes);
  const double one[2] = {1.0, 0.0}, beta[2] = {-0.5, 0.25};
  CHECK(o->mulm(a, X, ldx, Y, ldy, k, one));
  write_array(Y, sizeof(double), (int64_t)ldy * k * w);
  memcpy(Y, Y0, Ybytes);
  CHECK(o->mulm(a, X, ldx, Y, ldy, k, beta));
  write_array(Y, sizeof(double), (int64_t)ldy * k * w);
  /* device-resident block */
  void *dX = NULL, *dY = NULL;
  CHECK(slc_dev_alloc(&dX, (int64_t)((size_t)ldx * (size_t)k * (size_t)w * sizeof(double))));
  CHECK(slc_dev_alloc(&dY, (int64_t)Ybytes));
  CHECK(slc_memcpy_h2d(dX, X, (int64_t)((size_t)ldx * (size_t)k * (size_t)w * sizeof(double))));
  CHECK(slc_memcpy_h2d(dY, Y0, (int64_t)Ybytes));
  CHECK(o->mulm_dev(a, dX, ldx, dY, ldy, k, one));
  CHECK(slc_memcpy_d2h(Y, dY, (int64_t)Ybytes));
  write_array(Y, sizeof(double), (int64_t)ldy * k * w);

  /* error convention: int status < 0 and a message with the reference function's prefix */
  int rc = o->mulm(a, X, nc - 1, Y, ldy, k, NULL);
  write_i64(rc);
  const char *msg = slc_last_error();
  write_array(msg, 1, (int64_t)strlen(msg));

  /* transpose, ctrans, hermitian */
  slc_matrix *t = NULL, *ct = NULL;
  CHECK(o->transpose(a, 0, &t));
  emit_matrix(o, t);
  CHECK(o->transpose(a, 1, &ct));
  emit_matrix(o, ct);
  int herm = -1;
  CHECK(o->hermitian(a, &herm));
  write_i64(herm);

  /* mm: A^T A (always conformable), lin: A + 2 A, kronecker of two small pieces */
  slc_matrix *c = NULL, *l = NULL, *kr = NULL;
  CHECK(o->mm(t, a, &c));
  emit_matrix(o, c);
  const double two[2] = {2.0, 0.0};
  CHECK(o->lin(one, a, two, a, &l));
  emit_matrix(o, l);
  CHECK(o->kronecker(t, ct, &kr));
  emit_matrix(o, kr);

  /* compress from host triplets, from device triplets, and the bounds error with its position */
  slc_matrix *asm1 = NULL, *asm2 = NULL, *bad = NULL;
  int64_t pos = -1;
  CHECK(o->compress(nr, nc, nt, tr, tc, tv, &asm1, &pos));
  emit_matrix(o, asm1);
  void *dtr = NULL, *dtc = NULL, *dtv = NULL;
  CHECK(slc_dev_alloc(&dtr, nt * 4));
  CHECK(slc_dev_alloc(&dtc, nt * 4));
  CHECK(slc_dev_alloc(&dtv, nt * 8 * w));
  CHECK(slc_memcpy_h2d(dtr, tr, nt * 4));
  CHECK(slc_memcpy_h2d(dtc, tc, nt * 4));
  CHECK(slc_memcpy_h2d(dtv, tv, nt * 8 * w));
  CHECK(o->compress_dev(nr, nc, nt, dtr, dtc, dtv, &asm2, &pos));
  emit_matrix(o, asm2);
  if (nt > 3) {
    const int32_t keep = tr[3];
    tr[3] = nr; /* out of bounds at position 3 */
    rc = o->compress(nr, nc, nt, tr, tc, tv, &bad, &pos);
    tr[3] = keep;
    write_i64(rc);
    write_i64(pos);
    msg = slc_last_error();
    write_array(msg, 1, (int64_t)strlen(msg));
  }

  /* one process, several GPUs (when the box has them): the sharded handle */
  int ndev = 0;
  CHECK(slc_device_count(&ndev));
  const int use = ndev >= 2 ? 2 : 1;
  slc_sharded *sh = NULL;
  CHECK(o->upload_sharded(use, NULL, nr, nc, cp, ri, v, &sh));
  int32_t sdev = 0, snr = 0, snc = 0;
  int64_t snnz = 0;
  CHECK(slc_sharded_dims(sh, &sdev, &snr, &snc, &snnz));
  if (sdev != use || snr != nr || snc != nc || snnz != cp[nc]) { fprintf(stderr, "sharded dims\n"); exit(5); }
  for (int rep = 0; rep < 3; ++rep) CHECK(o->mulv_sharded(sh, x, y)); /* both operand buffers get used */
  write_array(y, sizeof(double), (int64_t)nr * w);
  memset(Y, 0, Ybytes);
  rc = o->mulm_sharded(sh, X, ldx, Y, ldy, k);
  write_i64(rc); /* a matrix without a sliced-ELL plan refuses the block form: recorded, not fatal */
  write_array(Y, sizeof(double), (int64_t)ldy * k * w);
  slc_sharded_free(sh);
  slc_sharded_free(NULL);

  /* finalizers */
  slc_free(a); slc_free(t); slc_free(ct); slc_free(c); slc_free(l); slc_free(kr);
  slc_free(asm1); slc_free(asm2); slc_free(NULL);
  CHECK(slc_dev_free(dx)); CHECK(slc_dev_free(dy)); CHECK(slc_dev_free(dX)); CHECK(slc_dev_free(dY));
  CHECK(slc_dev_free(dtr)); CHECK(slc_dev_free(dtc)); CHECK(slc_dev_free(dtv));
  free(dims); free(cp); free(ri); free(v); free(x); free(y0); free(X); free(Y0); free(tr); free(tc); free(tv);
  free(y); free(Y);
}

int main(int argc, char **argv) {
  if (argc == 2 && strcmp(argv[1], "--version") == 0) { /* links and loads without a GPU */
    printf("%d\n", slc_version());
    return 0;
  }
  if (argc != 3) { fprintf(stderr, "usage: consumer in.bin out.bin | consumer --version\n"); return 1; }
  fin = fopen(argv[1], "rb");
  fout = fopen(argv[2], "wb");
  if (!fin || !fout) { perror("open"); return 1; }
  int ndev = 0;
  CHECK(slc_device_count(&ndev));
  if (ndev < 1) { fprintf(stderr, "no CUDA device\n"); return 1; }
  CHECK(slc_set_device(0));
  CHECK(slc_set_stream(NULL));
  const int64_t before = slc_launch_count();
  run(&OPS_D, 'd');
  run(&OPS_Z, 'z');
  write_i64(slc_launch_count() - before);
  fclose(fin);
  fclose(fout);
  return 0;
}
