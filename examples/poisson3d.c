/* Minimal C program on the public C ABI (include/cgb200.h): solve the 3D 7-point Poisson problem A x = 1 with CG.
 *
 *   gcc -std=c99 -Iinclude examples/poisson3d.c -o poisson3d -Lcuda_mpi_parallel_b200/lib -lcgb200 \
 *       -Wl,-rpath,$PWD/cuda_mpi_parallel_b200/lib -lm
 *   ./poisson3d 128          # on the GPU (device 0)
 *   ./poisson3d 32 --cpu     # the host oracle, no GPU needed
 *
 * It is the reference program's flow (CUDACG.cu:41-366: stage the problem, iterate to ||r|| < 1e-7, read x back)
 * with the operator given as a grid instead of three hard-coded CSR arrays.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "cgb200.h"

#define CHECK(call)                                                        \
  do {                                                                     \
    if ((call) != 0) {                                                     \
      fprintf(stderr, "%s failed: %s\n", #call, cgb_last_error());         \
      return EXIT_FAILURE;                                                 \
    }                                                                      \
  } while (0)

int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 64;
  const int on_cpu = argc > 2 && strcmp(argv[2], "--cpu") == 0;
  if (n < 2) {
    fprintf(stderr, "usage: poisson3d N [--cpu]\n");
    return EXIT_FAILURE;
  }
  cgb_stencil_t st;
  CHECK(cgb_stencil_default(CGB_KIND_POISSON3D7, n, n, n, &st));
  const size_t N = (size_t)n * n * n;
  double* x = (double*)malloc(N * sizeof(double));
  double* b = (double*)malloc(N * sizeof(double));
  if (!x || !b) return EXIT_FAILURE;
  CHECK(cgb_rhs_fill_box(&st, CGB_RHS_ONES, 0, 0, n, 0, n, b)); /* the whole grid as one box */

  int iters = 0;
  double rnorm = 0.0;
  if (on_cpu) {
    cgb_cpu_result_t res;
    CHECK(cgb_cpu_cg_stencil(&st, 1, 1, 0, 0, b, x, 1e-7, CGB_TOL_ABS, 2000, 0, NULL, &res));
    iters = res.iters;
    rnorm = res.rnorm;
  } else {
    cgb_partition_t part;
    cgb_ctx* ctx = NULL;
    cgb_solve_opts_t opts;
    cgb_solve_result_t res;
    CHECK(cgb_partition_make(&st, 1, 0, 1, 0, 0, &part)); /* one rank owns the whole grid */
    CHECK(cgb_ctx_create(&st, &part, 0, 2048, &ctx));
    CHECK(cgb_ctx_set_rhs(ctx, b, 0, NULL));
    cgb_solve_opts_default(&opts); /* tol 1e-7 absolute, maxit 2000: CUDACG.cu:244-245 */
    CHECK(cgb_solve(ctx, &opts, &res));
    CHECK(cgb_ctx_get_x(ctx, x, 0, NULL));
    CHECK(cgb_ctx_destroy(ctx));
    if (res.status < 0) {
      fprintf(stderr, "solve failed with status %d\n", res.status);
      return EXIT_FAILURE;
    }
    iters = res.iters;
    rnorm = res.rnorm;
    printf("device time %.3f ms, %lld kernel launches\n", res.device_ms, (long long)res.launches);
  }

  /* check the answer with the host operator: ||b - A x|| */
  double* Ax = (double*)malloc(N * sizeof(double));
  if (!Ax) return EXIT_FAILURE;
  CHECK(cgb_cpu_spmv_stencil(&st, x, Ax));
  double rr = 0.0;
  for (size_t i = 0; i < N; ++i) rr += (b[i] - Ax[i]) * (b[i] - Ax[i]);
  printf("%d^3 unknowns, %d iterations, ||r|| = %.3e (recomputed %.3e), x[centre] = %f\n", n, iters, rnorm, sqrt(rr),
         x[((size_t)(n / 2) * n + n / 2) * n + n / 2]);
  free(Ax);
  free(b);
  free(x);
  printf("Success\n");
  return 0;
}
