/* abi_driver.c -- a C host on the femx C ABI, the way a PETSc / Fortran host would drive it (no Python, no CUDA headers).
 *
 * Builds an n^3 hex8 cube, plans it, and checks the hand-off forms against each other:
 *   - femx_get_coo (the arrays of MatSetPreallocationCOO) is the row/column expansion of femx_get_pattern;
 *   - femx_get_pattern_mpiaij over the two halves of a 2-rank PETSC_DECIDE row split reproduces every row: diagonal
 *     block in local column numbers, off-diagonal block through garray (ascending, unique);
 *   - femx_assemble_K: rigid translations are in the null space (row sums per component vanish), values symmetric;
 *   - femx_set_async: the enqueue-only form returns the same bits after femx_synchronize.
 * Compiled and run by tests/test_c_abi.py:  gcc -std=c99 -Iinclude tests/c/abi_driver.c -Lfem_b200 -lfemx -lm
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "femx.h"

#define CHECK(call)                                                                              \
  do {                                                                                           \
    int rc_ = (call);                                                                            \
    if (rc_ != FEMX_OK) {                                                                        \
      fprintf(stderr, "%s:%d %s -> %d: %s\n", __FILE__, __LINE__, #call, rc_, femx_last_error(ctx)); \
      return 1;                                                                                  \
    }                                                                                            \
  } while (0)
#define REQUIRE(cond, msg)                                           \
  do {                                                               \
    if (!(cond)) {                                                   \
      fprintf(stderr, "%s:%d FAILED: %s\n", __FILE__, __LINE__, msg); \
      return 1;                                                      \
    }                                                                \
  } while (0)

int main(int argc, char **argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 5, n1 = n + 1;
  const int nels = n * n * n, nnds = n1 * n1 * n1;
  femx_ctx *ctx = NULL;
  if (femx_create(&ctx, 0, NULL) != FEMX_OK) {
    fprintf(stderr, "femx_create: %s\n", femx_last_error(NULL));
    return 2;
  }
  /* mesh: lexicographic nodes, ShapeFuncdHex node order (src/m_elems.F90:574-579), mildly sheared coordinates */
  int *etype = malloc(sizeof(int) * nels), *eptr = malloc(sizeof(int) * (nels + 1)), *eind = malloc(sizeof(int) * 8 * nels);
  int *emat = malloc(sizeof(int) * nels), *enl = calloc(nels, sizeof(int)), *nl2g = malloc(sizeof(int) * nnds);
  double *coords = malloc(sizeof(double) * 3 * nnds);
  for (int k = 0; k < n1; ++k)
    for (int j = 0; j < n1; ++j)
      for (int i = 0; i < n1; ++i) {
        const int q = (k * n1 + j) * n1 + i;
        coords[3 * q] = i + 0.1 * j; coords[3 * q + 1] = j + 0.07 * k; coords[3 * q + 2] = k + 0.05 * i;
        nl2g[q] = q + 1;
      }
  int e = 0;
  eptr[0] = 0;
  for (int k = 0; k < n; ++k)
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < n; ++i, ++e) {
        const int b = (k * n1 + j) * n1 + i + 1, up = n1 * n1;
        const int nd[8] = {b, b + 1, b + 1 + n1, b + n1, b + up, b + 1 + up, b + 1 + n1 + up, b + n1 + up};
        memcpy(eind + 8 * e, nd, sizeof nd);
        etype[e] = FEMX_HEX; emat[e] = 1; eptr[e + 1] = 8 * (e + 1);
      }
  const double E_nu[2] = {3.0e10, 0.25};
  CHECK(femx_set_mesh(ctx, 3, nels, etype, eptr, eind, emat, enl, nnds, coords, FEMX_ROW_MAJOR, nl2g, nnds));
  CHECK(femx_set_materials(ctx, 1, E_nu, 0, E_nu));
  CHECK(femx_set_bcs(ctx, 0, eptr, E_nu));
  CHECK(femx_build_plan(ctx));
  int64_t nrows = 0, nnz = 0;
  CHECK(femx_pattern_size(ctx, &nrows, &nnz));
  REQUIRE(nrows == 3 * (int64_t)nnds, "row count");

  int64_t *rowptr = malloc(sizeof(int64_t) * (nrows + 1));
  int32_t *colidx = malloc(sizeof(int32_t) * nnz), *ci = malloc(sizeof(int32_t) * nnz), *cj = malloc(sizeof(int32_t) * nnz);
  CHECK(femx_get_pattern(ctx, rowptr, colidx, FEMX_HOST));
  CHECK(femx_get_coo(ctx, ci, cj, FEMX_HOST));
  REQUIRE(rowptr[0] == 0 && rowptr[nrows] == nnz, "rowptr ends");
  for (int64_t r = 0; r < nrows; ++r)
    for (int64_t s = rowptr[r]; s < rowptr[r + 1]; ++s) {
      REQUIRE(ci[s] == r && cj[s] == colidx[s], "COO (i, j) = CSR expansion");
      REQUIRE(s == rowptr[r] || colidx[s] > colidx[s - 1], "columns ascend");
    }

  /* MPIAIJ split of a 2-rank PETSC_DECIDE row split (src/main.F90:154) */
  for (int rank = 0; rank < 2; ++rank) {
    const int64_t r0 = rank == 0 ? 0 : (nrows + 1) / 2, r1 = rank == 0 ? (nrows + 1) / 2 : nrows;
    int64_t nd = 0, no = 0, ng = 0;
    CHECK(femx_mpiaij_sizes(ctx, r0, r1, &nd, &no, &ng));
    REQUIRE(nd + no == rowptr[r1] - rowptr[r0], "diag + offdiag = rows' nonzeros");
    int64_t *drp = malloc(sizeof(int64_t) * (r1 - r0 + 1)), *orp = malloc(sizeof(int64_t) * (r1 - r0 + 1));
    int32_t *dc = malloc(sizeof(int32_t) * (nd + 1)), *oc = malloc(sizeof(int32_t) * (no + 1)), *ga = malloc(sizeof(int32_t) * (ng + 1));
    CHECK(femx_get_pattern_mpiaij(ctx, r0, r1, drp, dc, orp, oc, ga));
    for (int64_t g = 1; g < ng; ++g) REQUIRE(ga[g] > ga[g - 1], "garray ascends");
    for (int64_t g = 0; g < ng; ++g) REQUIRE(ga[g] < r0 || ga[g] >= r1, "garray holds off-diagonal columns only");
    for (int64_t r = r0; r < r1; ++r) {   /* merge the two blocks of the row back: must be the CSR row */
      int64_t pd = drp[r - r0], po = orp[r - r0];
      for (int64_t s = rowptr[r]; s < rowptr[r + 1]; ++s) {
        const int32_t col = colidx[s];
        if (col >= r0 && col < r1) { REQUIRE(pd < drp[r - r0 + 1] && dc[pd] == col - r0, "diagonal block column"); ++pd; }
        else { REQUIRE(po < orp[r - r0 + 1] && ga[oc[po]] == col, "off-diagonal column through garray"); ++po; }
      }
      REQUIRE(pd == drp[r - r0 + 1] && po == orp[r - r0 + 1], "row lengths of the split");
    }
    free(drp); free(orp); free(dc); free(oc); free(ga);
  }

  /* values: null space and symmetry */
  double *vals = malloc(sizeof(double) * nnz), *vals2 = malloc(sizeof(double) * nnz);
  CHECK(femx_assemble_K(ctx, vals, FEMX_HOST));
  double vmax = 0.0;
  for (int64_t s = 0; s < nnz; ++s) vmax = fmax(vmax, fabs(vals[s]));
  REQUIRE(vmax > 0.0 && isfinite(vmax), "values finite");
  for (int64_t r = 0; r < nrows; ++r) {
    double sum[3] = {0.0, 0.0, 0.0};
    for (int64_t s = rowptr[r]; s < rowptr[r + 1]; ++s) sum[colidx[s] % 3] += vals[s];
    REQUIRE(fabs(sum[0]) + fabs(sum[1]) + fabs(sum[2]) <= 1e-9 * vmax, "rigid translations in the null space");
  }
  for (int64_t r = 0; r < nrows; r += 7)
    for (int64_t s = rowptr[r]; s < rowptr[r + 1]; ++s) {
      const int32_t c = colidx[s];
      int64_t lo = rowptr[c], hi = rowptr[c + 1];
      while (hi - lo > 1) { const int64_t mid = (lo + hi) / 2; if (colidx[mid] <= r) lo = mid; else hi = mid; }
      REQUIRE(colidx[lo] == r && fabs(vals[lo] - vals[s]) <= 1e-12 * vmax, "K symmetric");
    }

  /* enqueue-only calls on device buffers */
  void *d_vals = NULL;
  CHECK(femx_device_alloc(ctx, &d_vals, (int64_t)sizeof(double) * nnz));
  CHECK(femx_set_async(ctx, 1));
  CHECK(femx_assemble_K(ctx, (double *)d_vals, FEMX_DEVICE));
  CHECK(femx_assemble_K(ctx, (double *)d_vals, FEMX_DEVICE));   /* back to back, nothing waits on the host */
  CHECK(femx_synchronize(ctx));
  CHECK(femx_set_async(ctx, 0));
  CHECK(femx_memcpy(ctx, vals2, d_vals, (int64_t)sizeof(double) * nnz, FEMX_HOST, FEMX_DEVICE));
  REQUIRE(memcmp(vals, vals2, sizeof(double) * nnz) == 0, "async form: same bits");
  REQUIRE(femx_last_kernel_ms(ctx) > 0.0, "device time after femx_synchronize");
  CHECK(femx_device_free(ctx, d_vals));
  CHECK(femx_destroy(ctx));
  printf("ABI DRIVER OK: hex8 %d^3, %lld rows, %lld nonzeros\n", n, (long long)nrows, (long long)nnz);
  return 0;
}
