/*
 * slc.h -- C ABI of libsparselinear_b200.so: the B200-native compressed-sparse hot
 * path of ttuegel/sparse-linear (SpMV, SpMM, SpGEMM, transpose, triplet assembly).
 *
 * This is the drop-in boundary.  The reference has no FFI for this path (every
 * function is pure Haskell), so the ABI copies the conventions of the reference's only
 * matrix marshalling code and of its two existing bindings, and every entry point
 * names the reference function it replaces.  Citations are relative to /root/reference.
 *
 *   matrix argument shape   (CInt nrows, CInt ncols, Ptr CInt colptr, Ptr CInt rowidx,
 *                            Ptr a vals), CSC, 32-bit indices
 *                            sparse-linear/src/Data/Matrix/Sparse/Foreign.hs:24-41
 *   Complex Double layout   interleaved (re, im) doubles, as UMFPACK's packed mode
 *                            suitesparse/src/Numeric/LinearAlgebra/Umfpack/Internal.hs:124-132
 *   ownership, in           arrays are lent for the call only (unsafeWith); every
 *                            upload copies to the device before returning   Foreign.hs:33-37
 *   ownership, out          results stay on the device behind an opaque handle; the
 *                            caller copies them out with slc_*_download into buffers it
 *                            owns (fromForeign's copy=True mode)             Foreign.hs:43-55
 *   handle lifetime         opaque pointer, freed by slc_free, whose address is usable
 *                            as a ForeignPtr finalizer ("&slc_free")         Umfpack/Internal.hs:110-113
 *   errors                  int status, 0 = ok, < 0 = error; slc_last_error() gives the
 *                            message, prefixed with the reference function's name
 *                            Umfpack.hs:64-67; Data/Matrix/Sparse.hs:438-445,694
 *   two element types       one symbol per type: slc_d_* (Double), slc_z_* (Complex
 *                            Double), selected by a class instance on the Haskell side
 *                            Umfpack/Internal.hs:61-67,117-162
 *
 * There is no CPU fallback: without a CUDA device every compute entry point returns
 * SLC_ERR_CUDA.  No torch types appear here; device pointers are plain pointers and the
 * stream is a cudaStream_t passed as void*.
 *
 * All `double *` arguments of slc_z_* functions point at interleaved complex data
 * (2 doubles per element).
 */
#ifndef SLC_H
#define SLC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

/* ---- status codes ------------------------------------------------------------------ */
#define SLC_OK 0
#define SLC_ERR_INVALID (-1)    /* null pointer, negative size, wrong element type */
#define SLC_ERR_DIM (-2)        /* dimension mismatch (axpy_/mm/glin messages) */
#define SLC_ERR_ROW_BOUNDS (-3) /* compress: row index out of bounds */
#define SLC_ERR_COL_BOUNDS (-4) /* compress: column index out of bounds */
#define SLC_ERR_CUDA (-5)       /* CUDA runtime failure, or no device */
#define SLC_ERR_NOMEM (-6)
#define SLC_ERR_FORMAT (-7)     /* malformed CSC handed to upload */
#define SLC_ERR_RANGE (-8)      /* result does not fit 32-bit indices */

/* Opaque device-resident CSC matrix (the Matrix record of Data/Matrix/Sparse.hs:67-76
   held in HBM, plus the row-major mirror and product plans built from it). */
typedef struct slc_matrix slc_matrix;

/* ---- library state ------------------------------------------------------------------ */

int slc_version(void);
/* Message of the calling thread's last failed call ("" if none). */
const char *slc_last_error(void);
int slc_device_count(int *count);
/* Select the CUDA device for the calling thread (cudaSetDevice). */
int slc_set_device(int device);
/* Stream used by every kernel and copy this thread issues through the library
   (a cudaStream_t; NULL = the legacy default stream). */
int slc_set_stream(void *cuda_stream);
int slc_synchronize(void);
/* Number of kernels this library has launched so far in this process. */
int64_t slc_launch_count(void);
/* Pinned host buffers for the host-pointer entry points (optional; any host memory
   works, pinned memory lets the copies run as DMA). */
int slc_host_alloc(void **ptr, int64_t bytes);
int slc_host_free(void *ptr);
/* Plain device buffers, for callers without their own allocator. */
int slc_dev_alloc(void **ptr, int64_t bytes);
int slc_dev_free(void *ptr);
int slc_memcpy_h2d(void *dst_dev, const void *src_host, int64_t bytes);
int slc_memcpy_d2h(void *dst_host, const void *src_dev, int64_t bytes);

/* ---- handles: Matrix (Data/Matrix/Sparse.hs:67-76), withConstMatrix / fromForeign
        (Data/Matrix/Sparse/Foreign.hs:24-88) ------------------------------------------- */

/* Copy a host CSC matrix into HBM.  Validates what checkMatrix checks
   (sparse-linear/tests/Test/LinearAlgebra.hs:40-67): colptr[0]==0, nondecreasing,
   0 <= rowidx < nrows, strictly increasing inside a column. */
int slc_d_upload(int32_t nrows, int32_t ncols, const int32_t *colptr, const int32_t *rowidx,
                 const double *vals, slc_matrix **out);
int slc_z_upload(int32_t nrows, int32_t ncols, const int32_t *colptr, const int32_t *rowidx,
                 const double *vals, slc_matrix **out);
/* Finalizer: releases the handle and everything it owns.  NULL is ignored. */
void slc_free(slc_matrix *m);
/* nrows, ncols, nonZero (Sparse.hs:115-117) */
int slc_dims(const slc_matrix *m, int32_t *nrows, int32_t *ncols, int64_t *nnz);
/* 'd' or 'z' (0 on a null handle) */
int slc_elem_type(const slc_matrix *m);
/* Copy the CSC arrays back: colptr[ncols+1], rowidx[nnz], vals[nnz] */
int slc_d_download(const slc_matrix *m, int32_t *colptr, int32_t *rowidx, double *vals);
int slc_z_download(const slc_matrix *m, int32_t *colptr, int32_t *rowidx, double *vals);
/* Build the row-major mirror and the product plan now instead of at the first product. */
int slc_prepare(slc_matrix *m);
/* Force the SpMV kernel family for this handle: 0 auto, 1 sliced-ELL (row per thread),
   2 CSR row-per-warp, 3 CSR nnz-balanced with products staged in shared memory, 4 CSR nnz-balanced
   merge-style (registers + segmented scan).  For benchmarking; results are unchanged up to
   summation order. */
int slc_set_spmv_kernel(slc_matrix *m, int kind);
/* One-line description of the plan ("sell32 slices=... pad=...%"), for logs. */
int slc_plan_info(const slc_matrix *m, char *buf, int32_t buflen);

/* ---- SpMV: mulV / axpy / axpy_ (Data/Matrix/Sparse.hs:433-471) ------------------------ */

/* y <- A x          (mulV, Sparse.hs:464-471)      x[ncols], y[nrows], host memory */
int slc_d_mulv(slc_matrix *a, const double *x, double *y);
int slc_z_mulv(slc_matrix *a, const double *x, double *y);
/* y <- A x + y      (axpy_/axpy, Sparse.hs:433-462) */
int slc_d_axpy(slc_matrix *a, const double *x, double *y);
int slc_z_axpy(slc_matrix *a, const double *x, double *y);
/* Same, operands resident in HBM; asynchronous on the current stream. */
int slc_d_mulv_dev(slc_matrix *a, const double *x_dev, double *y_dev);
int slc_z_mulv_dev(slc_matrix *a, const double *x_dev, double *y_dev);
int slc_d_axpy_dev(slc_matrix *a, const double *x_dev, double *y_dev);
int slc_z_axpy_dev(slc_matrix *a, const double *x_dev, double *y_dev);

/* ---- SpMM: FEAST's multiplyWork (feast/src/Numeric/LinearAlgebra/Feast.hs:203-208) and
        the unexported mulM (Data/Matrix/Sparse.hs:473-498) ------------------------------- */

/* Y[:, j] <- A X[:, j] for j in [0, k); X and Y column-major with leading dimensions
   ldx >= ncols and ldy >= nrows (in elements).  A column range is chosen by offsetting
   the pointers, as sliceMat does (Feast.hs:189-192). */
int slc_d_mulm(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k);
int slc_z_mulm(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k);
int slc_d_mulm_dev(slc_matrix *a, const double *X_dev, int64_t ldx, double *Y_dev, int64_t ldy,
                   int32_t k);
int slc_z_mulm_dev(slc_matrix *a, const double *X_dev, int64_t ldx, double *Y_dev, int64_t ldy,
                   int32_t k);

/* *flag <- nonzero if a launch of the TMA-staged block product (banded matrices with short rows) gave
   up waiting on one of its shared-memory barriers or met a window wider than planned: its result
   is then invalid.  A diagnostic for tests and benchmarks; 0 in normal operation. */
int slc_spmm_window_error(int *flag);

/* ---- SpGEMM: mm = (*) (Data/Matrix/Sparse.hs:691-702, ScatterGather.hs:37-147) -------- */

/* C <- A B.  Structural product: an entry exists wherever a term landed, zeros are
   kept, row indices strictly increasing per column, terms added in increasing inner
   index.  Error: "mm: inner dimension mismatch" (Sparse.hs:694). */
int slc_d_mm(const slc_matrix *a, const slc_matrix *b, slc_matrix **c);
int slc_z_mm(const slc_matrix *a, const slc_matrix *b, slc_matrix **c);

/* ---- transpose / ctrans / hermitian (Data/Matrix/Sparse.hs:301-329,371-379) ----------- */

/* At <- transpose A (conj = 0) or ctrans A (conj != 0).  Stable; values move bit for bit. */
int slc_d_transpose(const slc_matrix *a, int conj, slc_matrix **at);
int slc_z_transpose(const slc_matrix *a, int conj, slc_matrix **at);
/* *result <- 1 if ctrans A == A on every field (Sparse.hs:377-379, Eq at :78), else 0 */
int slc_d_hermitian(const slc_matrix *a, int *result);
int slc_z_hermitian(const slc_matrix *a, int *result);

/* ---- triplet assembly: compress / fromTriples / (><) (Sparse.hs:184-291,357-369) ------- */

/* COO -> CSC: bounds check (rows first, then columns; *bad_pos = first offending
   position, Sparse.hs:196-212), sort, sum duplicates in input order, keep explicit
   zeros.  rows/cols/vals are host arrays of n entries. */
int slc_d_compress(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows,
                   const int32_t *cols, const double *vals, slc_matrix **out, int64_t *bad_pos);
int slc_z_compress(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows,
                   const int32_t *cols, const double *vals, slc_matrix **out, int64_t *bad_pos);
/* Same, with the triplet arrays already in device memory (read in place, left intact). */
int slc_d_compress_dev(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows_dev,
                       const int32_t *cols_dev, const double *vals_dev, slc_matrix **out,
                       int64_t *bad_pos);
int slc_z_compress_dev(int32_t nrows, int32_t ncols, int64_t n, const int32_t *rows_dev,
                       const int32_t *cols_dev, const double *vals_dev, slc_matrix **out,
                       int64_t *bad_pos);

/* ---- builders the configs and FEAST use: lin / glin, kronecker
        (Data/Matrix/Sparse.hs:401-431,597-638) ------------------------------------------- */

/* C <- alpha A + beta B on the union pattern, zeros kept.  alpha, beta point at one
   element (1 double for d, 2 for z).  Errors: "glin: row number mismatch",
   "glin: column number mismatch" (Sparse.hs:408-409). */
int slc_d_lin(const double *alpha, const slc_matrix *a, const double *beta, const slc_matrix *b,
              slc_matrix **c);
int slc_z_lin(const double *alpha, const slc_matrix *a, const double *beta, const slc_matrix *b,
              slc_matrix **c);
int slc_d_kronecker(const slc_matrix *a, const slc_matrix *b, slc_matrix **c);
int slc_z_kronecker(const slc_matrix *a, const slc_matrix *b, slc_matrix **c);

/* ---- one process per GPU: row-partitioned SpMV with the operand partitioned the same way -----
   The reference has no distributed code; this is the multi-GPU form of mulV
   (Data/Matrix/Sparse.hs:464-471) BASELINE.json's north_star asks for.  Rank `self` holds the row
   block A[r0:r1, :] as an ordinary handle of shape (r1-r0) x ncols and the slice
   x[splits[self]:splits[self+1]] in its own HBM; every rank maps the other ranks' slices with CUDA
   IPC.  The product kernel gathers local columns from HBM and remote ("halo") columns straight from
   the peer's memory over NVLink, so there is no separate exchange step.  The caller orders the
   product after the peers have written their slices (one barrier per iteration). */

typedef struct slc_peer_group slc_peer_group;

/* 72-byte export blob (CUDA IPC handle of the containing allocation + offset) of a pointer into
   a device allocation made with slc_dev_alloc / cudaMalloc */
#define SLC_IPC_HANDLE_BYTES 72
int slc_ipc_export(const void *dev_ptr, void *handle72);
int slc_ipc_import(const void *handle72, void **dev_ptr);
int slc_ipc_close(void *dev_ptr);

/* Barrier between the ranks, written over peer memory (one tiny kernel; see csrc/slc_peer.cu).
   create: allocates this rank's flag array and exports it; connect: handles72 holds the nparts
   blobs in rank order (the caller gathers them, e.g. with torch.distributed); barrier: enqueued on
   the current stream, returns at once; every rank must call it the same number of times. */
int slc_peer_group_create(int32_t nparts, int32_t self, slc_peer_group **out, void *handle72);
int slc_peer_group_connect(slc_peer_group *g, const void *handles72);
int slc_peer_barrier(slc_peer_group *g);
/* The two halves of the barrier: signal = "my slice of x is written" (starts a new epoch),
   wait = "every peer has signalled this epoch". */
int slc_peer_signal(slc_peer_group *g);
int slc_peer_wait(slc_peer_group *g);
/* *flag <- 1 if a barrier gave up waiting for a peer (bounded spin), else 0 */
int slc_peer_group_timed_out(slc_peer_group *g, int *flag);
void slc_peer_group_free(slc_peer_group *g);
/* y_local <- A_local x.  splits[nparts+1] (host) are the column bounds of the slices, x_parts[nparts]
   (host array) the device addresses of the slices as mapped in THIS process; 1 <= nparts <= 16.
   With sync != NULL the call orders itself against the peers: signal, the rows that need no peer,
   wait, the boundary rows.  With sync == NULL the caller has already ordered the ranks. */
int slc_d_mulv_dist_dev(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *x_parts, double *y_dev, slc_peer_group *sync);
int slc_z_mulv_dist_dev(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *x_parts, double *y_dev, slc_peer_group *sync);

/* The same product with the operand all-gathered first: barrier (when sync != NULL), one kernel
   that pulls every part's slice over NVLink into x_full_dev[ncols] (this rank's scratch copy of x),
   then the undistributed product.  For matrices whose rows touch columns all over the operand
   (power-law graphs, BASELINE.json config 4), where the halo is the whole vector. */
int slc_d_mulv_allgather_dev(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits,
                             const double *const *x_parts, double *x_full_dev, double *y_dev,
                             slc_peer_group *sync);
int slc_z_mulv_allgather_dev(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits,
                             const double *const *x_parts, double *x_full_dev, double *y_dev,
                             slc_peer_group *sync);

/* Y_local <- A_local X, the block form (FEAST's multiplyWork over a row-partitioned subspace block).
   X is partitioned by the same row blocks: X_parts[p] is the device address, as mapped in THIS process,
   of part p's (splits[p+1]-splits[p]) x k column-major block with leading dimension ldx_parts[p].
   Same ordering contract as slc_*_mulv_dist_dev.  Sliced-ELL plans only. */
int slc_d_mulm_dist_dev(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *X_parts, const int64_t *ldx_parts, double *Y_dev,
                        int64_t ldy, int32_t k, slc_peer_group *sync);
int slc_z_mulm_dist_dev(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits,
                        const double *const *X_parts, const int64_t *ldx_parts, double *Y_dev,
                        int64_t ldy, int32_t k, slc_peer_group *sync);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* SLC_H */
