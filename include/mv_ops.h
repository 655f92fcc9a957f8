/* mv_ops.h -- the operator API of the conjugate-gradient path, served by libcgb200.so.
 *
 * This header declares, symbol for symbol, the C interface the reference exposes in its own
 * mv_ops.h (reference mv_ops.h:17-42) plus the solver entry point of cg.c (reference cg.c:24,88).
 * A program written against the reference's header recompiles against this one unchanged and
 * links libcgb200.so instead of mv_ops.c/cg.c; the arithmetic then runs as sm_100a CUDA kernels.
 *
 * ABI notes
 *   - struct __mv_sparse is byte-for-byte the reference layout (32 bytes on LP64).  It describes
 *     either a CSR matrix (values[nnz], col_indices[nnz] ascending and unique per row, 0-based;
 *     row_ptr[size+1], 0-based) or a dense vector (col_indices == row_ptr == NULL, nnz == size).
 *   - All arrays named by the struct are HOST memory owned by the caller.  Device copies are
 *     kept by the library in a side table keyed by the struct's address (see cgb200.h).
 *   - Output arguments follow the reference's allocate-or-reuse rule: *out == NULL makes the
 *     library allocate struct and values; otherwise values is realloc'ed to the new size.
 *     out may alias an input (the reference calls vec_sub in place, cg.c:123).
 *   - Return codes are the reference's: int ops 0 / -1, dot_product -1.0 on NULL or size
 *     mismatch.  A CUDA failure also yields -1 / -1.0; cgb200_last_error() says why.
 */
#ifndef CGB200_MV_OPS_H
#define CGB200_MV_OPS_H

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef __cplusplus
extern "C" {
#endif

struct __mv_sparse {
  int size;          /* rows (matrix) or length (vector) */
  int nnz;           /* stored entries; == size for a vector */
  double *values;
  int *col_indices;  /* NULL for a vector */
  int *row_ptr;      /* NULL for a vector */
};

/* -- lifecycle ------------------------------------------------------------------------------ */

/* Zero-initialised struct.                                   replaces reference mv_ops.c:14-21 */
struct __mv_sparse *new_mv_struct(void);
/* Zero vector of the given length.                           replaces reference mv_ops.c:23-37 */
struct __mv_sparse *new_mv_struct_with_size(int size);
/* Frees the struct only (not its arrays), as the reference does, and additionally drops the
 * library's device mirror of it.                             replaces reference mv_ops.c:39-42 */
void free_mv_struct(struct __mv_sparse *obj);
/* Independent copy (values; column indices and row pointers when present); NULL -> NULL.
 *                                                            replaces reference mv_ops.c:44-74 */
struct __mv_sparse *mv_deep_copy(struct __mv_sparse *orig);

/* -- inspection ----------------------------------------------------------------------------- */

/* Same text as the reference (Size / NNZ / values pointer / one %f per value); brings the
 * values back from the device first when they live there.    replaces reference mv_ops.c:77-95 */
void print_sparse(struct __mv_sparse *obj);
/* Expands CSR row row_id into the dense array p_row[size]; -1 on NULL.  Implements the
 * mathematically correct row (stops at row_ptr[row_id+1]); the reference's scan can run into the
 * next row on block-diagonal matrices.        replaces reference mv_ops.c:99-113 */
int mat_get_row(struct __mv_sparse *mat_A, int row_id, double *p_row);

/* -- arithmetic (device kernels) ------------------------------------------------------------ */

/* sum_i a[i]*b[i]                                           replaces reference mv_ops.c:117-132 */
double dot_product(struct __mv_sparse *vec_a, struct __mv_sparse *vec_b);
/* *out = s * a                                              replaces reference mv_ops.c:134-158 */
int sv_mult(double s, struct __mv_sparse *vec_a, struct __mv_sparse **out);
/* *out = A * b; -1 when A->size != b->size                  replaces reference mv_ops.c:160-201
 * CSR rows: each row's stored entries are multiplied and added in ascending column order, unfused,
 * so *out has the reference's bits.  A matrix whose every row is full (row_ptr[i] == i*size) takes
 * the dense GEMV kernel, which accumulates with fma over a fixed tree: agrees with the reference
 * to <= 4 n eps |A||b| per row, not bit for bit.  A's arrays are re-validated on every call
 * (cgb200.h: CGB200_CACHE_TRACKED), so in-place edits of values[] are seen as the reference sees
 * them. */
int mv_mult(struct __mv_sparse *mat_A, struct __mv_sparse *vec_b, struct __mv_sparse **out);
/* *out = a + b                                              replaces reference mv_ops.c:203-230 */
int vec_add(struct __mv_sparse *vec_a, struct __mv_sparse *vec_b, struct __mv_sparse **out);
/* *out = a - b                                              replaces reference mv_ops.c:232-259 */
int vec_sub(struct __mv_sparse *vec_a, struct __mv_sparse *vec_b, struct __mv_sparse **out);

/* -- the solve loop ------------------------------------------------------------------------- */

/* Classic conjugate gradients from x0 = 0 with the reference's fixed-count semantics:
 * max_iter = M performs M+1 matrix-vector products and x/r updates and M direction updates,
 * no tolerance test.  *x receives a freshly allocated vector.  Returns 0 (the reference always
 * does); -1 on NULL arguments or a CUDA failure.  Unlike the reference, iterating past exact
 * convergence (r.r == 0, or p.Ap zero / not finite) stops instead of producing NaN (cgb200.h,
 * CGB200_EXACT / CGB200_BREAKDOWN); a negative p.Ap is divided by as the reference does, so
 * negative-definite systems solve like (-A, -b).  Structs should be released through
 * free_mv_struct / cgb200_free_all so that their device mirrors go with them.
 *                                                               replaces reference cg.c:88-141 */
int conj_grad(int max_iter, struct __mv_sparse *mat_A, struct __mv_sparse *vec_b,
              struct __mv_sparse **x);

#ifdef __cplusplus
}
#endif

#endif /* CGB200_MV_OPS_H */
