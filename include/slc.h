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
/* device to device on the current stream, asynchronous */
int slc_memcpy_d2d(void *dst_dev, const void *src_dev, int64_t bytes);

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

/* Y[:, j] <- A X[:, j] + beta Y[:, j] for j in [0, k); X and Y column-major with leading dimensions
   ldx >= ncols and ldy >= nrows (in elements).  A column range is chosen by offsetting
   the pointers, as sliceMat does (Feast.hs:189-192).
   beta points at one element (1 double for d, 2 for z); NULL or 0 is mulM / multiplyWork proper (Y is
   written, never read: the reference zeroes the destination first, Sparse.hs:482-484); 1 is the
   reference's axpy_ into a block the caller filled (Sparse.hs:485-488: every entry starts from y and
   adds its row's terms in column order, bit for bit); any other value starts from beta * y. */
int slc_d_mulm(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
               const double *beta);
int slc_z_mulm(slc_matrix *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k,
               const double *beta);
int slc_d_mulm_dev(slc_matrix *a, const double *X_dev, int64_t ldx, double *Y_dev, int64_t ldy,
                   int32_t k, const double *beta);
int slc_z_mulm_dev(slc_matrix *a, const double *X_dev, int64_t ldx, double *Y_dev, int64_t ldy,
                   int32_t k, const double *beta);

/* *flag <- nonzero if a launch of the TMA-staged block product (banded matrices with short rows) on any
   device gave up waiting on one of its shared-memory barriers or met a window wider than planned: its
   result is then invalid.  0 in normal operation.  The host-operand products (slc_*_mulm,
   slc_*_mulm_sharded, slc_sharded_synchronize) look at their device's word themselves after their final
   synchronisation: they return SLC_ERR_CUDA for such a product and clear the word, so the next product
   starts clean.  This call is for users of the asynchronous *_dev forms; it synchronises the current
   stream, reads, and clears nothing. */
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
   the peer's memory over NVLink, so there is no separate exchange step.

   Ordering contract.  With sync != NULL a product orders itself: it tells the peers "my slice is
   written" (true by stream order when the product starts) and its boundary rows wait until the peers
   they depend on have said the same.  What the caller owns is the opposite hazard: a rank must not
   overwrite a slice that a peer may still be reading.  The rule that closes it without any extra
   exchange is DOUBLE BUFFERING: keep two exported buffers per rank and alternate them product by
   product (pass the table of the buffers that hold the operand of THIS product).  Rank r writes
   buffer b again only after its own next product -- which waited for its dependency set's next
   signal, sent after they finished reading b -- so the write cannot overtake a reader.  A caller that
   keeps one buffer must separate "all peers done reading" from "I overwrite" itself
   (slc_peer_barrier after the product). */

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
/* The same with every rank signalling and waiting for every other rank, whatever dependencies
   slc_peer_group_set_deps installed: a full rendezvous of the GPUs on their streams (e.g. to line the
   devices up before a timed region; the products themselves only ever wait for the ranks they read). */
int slc_peer_barrier_all(slc_peer_group *g);
/* The two halves of the barrier: signal = "my slice of x is written",
   wait = "every rank of my dependency set has signalled this epoch" (and closes the epoch).  The epoch
   counter lives in device memory and is advanced by the kernels, so barriers and products can be
   captured in a CUDA graph and the graph relaunched. */
int slc_peer_signal(slc_peer_group *g);
int slc_peer_wait(slc_peer_group *g);
/* The same group for ndev devices driven by ONE process (devices == NULL: 0..ndev-1): peer access is
   enabled both ways, no IPC.  groups[ndev] receives one group per device, already connected. */
int slc_peer_group_create_local(int32_t ndev, const int32_t *devices, slc_peer_group **groups);
/* Dependency set of this rank, bit p = rank p: the ranks whose slices it reads OR that read its
   slice (slc_*_dist_needs of every rank, exchanged and symmetrised by the caller).  Products and
   barriers wait for exactly these; the default is every peer. */
int slc_peer_group_set_deps(slc_peer_group *g, uint32_t mask);
/* *flag <- 1 if a wait gave up on a peer (5 s bound).  The rows that depended on it were filled with
   NaN, and every later call on the group fails with SLC_ERR_CUDA. */
int slc_peer_group_timed_out(slc_peer_group *g, int *flag);
/* out4 <- {epochs completed, timed waits, nanoseconds waited in total, longest wait in ns};
   reset != 0 clears the wait statistics.  Synchronises the device. */
int slc_peer_group_stats(slc_peer_group *g, int64_t *out4, int reset);
void slc_peer_group_free(slc_peer_group *g);
/* *mask <- the parts (bit p) whose slices the rows of a_local read, given the partition. */
int slc_d_dist_needs(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits, uint32_t *mask);
int slc_z_dist_needs(slc_matrix *a_local, int32_t nparts, int32_t self, const int32_t *splits, uint32_t *mask);
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

/* ---- one process, several GPUs: the same row-partitioned products behind one handle ------------------
   SURVEY.md section 8(b): slc_T_upload_sharded / slc_T_mulv_sharded / slc_T_mulm_sharded.  The
   reference's host is one Haskell process, so this is the form its shim binds.  upload cuts the
   matrix into ndev contiguous row blocks with (almost) equal numbers of stored entries (bounds on
   multiples of 32 rows), one per device; a square matrix partitions x like its rows.  The products
   take whole host vectors / column-major blocks (each device moves its own slice over its own PCIe
   link, the slices are double-buffered inside the handle) and return when y is complete.  Regular
   matrices (sliced-ELL plans) read halo entries from the peers' HBM inside the product kernel;
   irregular ones all-gather x over NVLink first (mulv only).  Not re-entrant per handle (a mutex
   serialises callers). */
typedef struct slc_sharded slc_sharded;
/* devices == NULL: devices 0 .. ndev-1 */
int slc_d_upload_sharded(int32_t ndev, const int32_t *devices, int32_t nrows, int32_t ncols,
                         const int32_t *colptr, const int32_t *rowidx, const double *vals, slc_sharded **out);
int slc_z_upload_sharded(int32_t ndev, const int32_t *devices, int32_t nrows, int32_t ncols,
                         const int32_t *colptr, const int32_t *rowidx, const double *vals, slc_sharded **out);
/* Finalizer ("&slc_sharded_free").  NULL is ignored. */
void slc_sharded_free(slc_sharded *a);
int slc_sharded_dims(const slc_sharded *a, int32_t *ndev, int32_t *nrows, int32_t *ncols, int64_t *nnz);
/* row_splits[ndev+1], col_splits[ndev+1]: block p holds rows [row_splits[p], row_splits[p+1]) and owns
   x[col_splits[p] : col_splits[p+1]]; either pointer may be NULL */
int slc_sharded_splits(const slc_sharded *a, int32_t *row_splits, int32_t *col_splits);
/* block p as an ordinary handle ((r1-r0) x ncols, owned by the sharded handle), its device ordinal and
   the cudaStream_t its products run on; any output pointer may be NULL */
int slc_sharded_block(const slc_sharded *a, int32_t part, slc_matrix **block, int32_t *device, void **stream);
/* y <- A x (mulV), Y[:, j] <- A X[:, j] (multiplyWork): host operands, synchronous */
int slc_d_mulv_sharded(slc_sharded *a, const double *x, double *y);
int slc_z_mulv_sharded(slc_sharded *a, const double *x, double *y);
int slc_d_mulm_sharded(slc_sharded *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k);
int slc_z_mulm_sharded(slc_sharded *a, const double *X, int64_t ldx, double *Y, int64_t ldy, int32_t k);
/* Device-resident form.  slc_sharded_operand gives, for block `part`, the buffer (on that block's
   device) that the NEXT product reads as its slice of x / its (slice x k) column-major rows of X with
   leading dimension *ldx; fill it on the block's stream (or synchronise your own stream first), then
   call slc_*_mul{v,m}_sharded_dev: asynchronous, one launch sequence per device.  The result of block
   `part` is at slc_sharded_result; slc_sharded_synchronize waits for every device and reports a barrier
   that gave up.  Two operand buffers alternate per product, so refilling right after the call is safe. */
int slc_sharded_operand(slc_sharded *a, int32_t part, int32_t k, void **x_dev, int64_t *ldx);
int slc_sharded_result(slc_sharded *a, int32_t part, void **y_dev, int64_t *ldy);
int slc_d_mulv_sharded_dev(slc_sharded *a);
int slc_z_mulv_sharded_dev(slc_sharded *a);
int slc_d_mulm_sharded_dev(slc_sharded *a, int32_t k);
int slc_z_mulm_sharded_dev(slc_sharded *a, int32_t k);
int slc_sharded_synchronize(slc_sharded *a);

/* ---- pieces of the transpose / assembly of a matrix spread over several GPUs --------
   The exchange itself (one all-to-all-v of raw device arrays) is the caller's: NCCL between processes,
   peer copies inside one.  These are the device-side steps around it. */
/* The CSC arrays of a handle as device pointers (valid while the handle lives; read-only).  A column
   range [c0, c1) of the matrix is the slice colptr[c0..c1], rowidx/vals[colptr[c0] .. colptr[c1]). */
int slc_device_arrays(const slc_matrix *m, const int32_t **colptr_dev, const int32_t **rowidx_dev,
                      const void **vals_dev);
/* Row blocks of one matrix, listed top to bottom as device CSC arrays (pointers starting at 0, ncols+1
   each), stacked into one handle: column j of the result is the blocks' columns j one after the other,
   row indices shifted by the rows above.  1 <= nblocks <= 16; the arrays are read, not adopted. */
int slc_d_vstack_dev(int32_t nblocks, const int32_t *block_rows, int32_t ncols, const int32_t *const *colptr_dev,
                     const int32_t *const *rowidx_dev, const double *const *vals_dev, slc_matrix **out);
int slc_z_vstack_dev(int32_t nblocks, const int32_t *block_rows, int32_t ncols, const int32_t *const *colptr_dev,
                     const int32_t *const *rowidx_dev, const double *const *vals_dev, slc_matrix **out);
/* Triplets in HBM bucketed by the part that owns their column (col_splits[nparts+1], host): bucket p
   occupies counts[0]+..+counts[p-1] onwards in the output arrays, input order kept inside a bucket,
   columns rebased to the owner's block.  counts[nparts] is a host array; synchronises the stream. */
int slc_d_coo_bucket_dev(int64_t n, const int32_t *rows_dev, const int32_t *cols_dev, const double *vals_dev,
                         int32_t nparts, const int32_t *col_splits, int32_t *out_rows_dev, int32_t *out_cols_dev,
                         double *out_vals_dev, int64_t *counts);
int slc_z_coo_bucket_dev(int64_t n, const int32_t *rows_dev, const int32_t *cols_dev, const double *vals_dev,
                         int32_t nparts, const int32_t *col_splits, int32_t *out_rows_dev, int32_t *out_cols_dev,
                         double *out_vals_dev, int64_t *counts);

/* ---- the same bucketing for callers that cannot allocate device memory ---------------------------------
   No counterpart in the reference (it has no device code at all): this exists for hosts that must not
   trigger a device allocation inside the call -- a process that has already committed all of HBM to its
   own arena, a Haskell `unsafe` foreign call or a latency-bound loop that cannot afford a driver
   allocation that may block, a GPU shared with a runtime whose allocator holds every free byte.
   slc_*_coo_bucket_dev takes nnz-sized sort buffers from the device memory pool; this variant calls no
   cudaMalloc / cudaMallocAsync / cudaFree and instead works in an area built into the library image
   (about 0.5 MiB, resident on every GPU that loads the library until the process exits).

   Arguments and result are those of slc_d_coo_bucket_dev / slc_z_coo_bucket_dev (bit for bit the same
   arrays and counts); elem is 'd' or 'z' and says how wide a value is (8 or 16 bytes).
   Limits: n <= SLC_NOALLOC_MAX_TRIPLETS and nparts <= 16, otherwise SLC_ERR_RANGE with a message that
   names the limit, and nothing is written.
   The work area is shared: one per GPU for all threads and streams, holding the last call's contents (the
   call clears the cells it needs at zero).  Calls that may overlap on one GPU race on it; issue them on
   one stream or order them yourself.  The first call on each device also looks up the area's device
   address (cudaGetSymbolAddress), so make that first call before any stream capture begins.
   Synchronises the stream (counts is a host array), like the allocating form. */
#define SLC_NOALLOC_MAX_TRIPLETS 16777216 /* 2^24 */
int slc_coo_bucket_noalloc_dev(int elem, int64_t n, const int32_t *rows_dev, const int32_t *cols_dev,
                               const void *vals_dev, int32_t nparts, const int32_t *col_splits,
                               int32_t *out_rows_dev, int32_t *out_cols_dev, void *out_vals_dev,
                               int64_t *counts);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* SLC_H */
