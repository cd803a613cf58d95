// Reference-STYLE baseline, written by this project (NOT upstream code, NOT upstream's numbers).
//
// It issues the same library-call sequence per iteration as the reference's loop
// (CUDACG.cu:269-352): cusparseSpMV (general CSR, fp64, int32 indices) -> cublasDdot ->
// 2x cublasDaxpy -> cublasDnrm2 -> cublasDscal -> cublasDaxpy, cuBLAS/cuSPARSE in host pointer
// mode on the default stream, i.e. with the reference's two blocking scalar read-backs per
// iteration. Two flavours:
//   faithful = 1 : also re-queries the SpMV workspace and cudaMalloc's it EVERY iteration, as
//                  CUDACG.cu:273-285 does (the leak is bounded here by freeing the previous buffer)
//   faithful = 0 : "hoisted" — workspace allocated once; the stronger bar to beat.
// The matrix is the assembled stencil operator of the benchmark grid, built on the device.
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cusparse.h>
#include <thrust/device_ptr.h>
#include <thrust/scan.h>

#include <cstdio>
#include <string>

namespace {

thread_local std::string g_err;

#define RS_CUDA(x)                                                                  \
  do {                                                                              \
    cudaError_t e_ = (x);                                                           \
    if (e_ != cudaSuccess) {                                                        \
      g_err = std::string(#x) + ": " + cudaGetErrorString(e_);                      \
      return -2;                                                                    \
    }                                                                               \
  } while (0)
#define RS_LIB(x, okv)                                                              \
  do {                                                                              \
    if ((x) != (okv)) {                                                             \
      g_err = std::string(#x) + " failed";                                          \
      return -2;                                                                    \
    }                                                                               \
  } while (0)

struct Grid {
  int kind, gx, gy, gz;
  double cc, cf, ce, cn;
};

__device__ __forceinline__ double coeff(const Grid& g, int dx, int dy, int dz) {
  int m = (dx != 0) + (dy != 0) + (dz != 0);
  if (m == 0) return g.cc;
  if (m == 1) return (g.kind == 1 && dz != 0) ? 0.0 : g.cf;
  if (g.kind != 3) return 0.0;
  return m == 2 ? g.ce : g.cn;
}

// pass 0: count entries per row into cnt[row+1]; pass 1: fill colind/val at rowptr[row].
__global__ void k_assemble(Grid g, long long n, int* rowptr, int* colind, double* val, int fill) {
  for (long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (long long)gridDim.x * blockDim.x) {
    int i = (int)(row % g.gx), j = (int)((row / g.gx) % g.gy), k = (int)(row / ((long long)g.gx * g.gy));
    int zr = g.kind >= 2 ? 1 : 0;
    int pos = fill ? rowptr[row] : 0;
    for (int dz = -zr; dz <= zr; ++dz)
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          double c = coeff(g, dx, dy, dz);
          if (c == 0.0) continue;
          int ii = i + dx, jj = j + dy, kk = k + dz;
          if (ii < 0 || jj < 0 || kk < 0 || ii >= g.gx || jj >= g.gy || kk >= g.gz) continue;
          if (fill) {
            colind[pos] = (int)(ii + (long long)g.gx * (jj + (long long)g.gy * kk));
            val[pos] = c;
          }
          ++pos;
        }
    if (!fill) rowptr[row + 1] = pos;
  }
}

__global__ void k_fill_rhs(double* b, long long n, unsigned long long seed) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long z = (unsigned long long)i ^ (seed * 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    b[i] = (double)(z >> 11) * (2.0 / 9007199254740992.0) - 1.0;
  }
}

}  // namespace

extern "C" const char* cgb_refstyle_last_error(void) { return g_err.c_str(); }

// Runs `warmup` untimed + `steps` timed iterations of the reference loop (no early exit: the
// residual test is evaluated like the reference does, but tol is 0). Returns ms for the timed
// iterations (CUDA events on the default stream) and the last ||r||.
// b_dev (may be NULL: built-in hash RHS): n doubles on the device. hist (may be NULL): host array that receives
// ||r||_2 after each of the warmup+steps iterations, so that the baseline's numerics can be pinned to the solver's.
extern "C" int cgb_refstyle_run2(int kind, int gx, int gy, int gz, int faithful, int warmup, int steps,
                                 const double* b_dev, double* hist, double* out_ms, double* out_rnorm, long long* out_nnz) {
  Grid g{kind, gx, gy, gz, 0, -1.0, -1.0, -1.0};
  g.cc = kind == 1 ? 4.0 : (kind == 2 ? 6.0 : 26.0);
  const long long n = (long long)gx * gy * gz;
  if (n >= (1LL << 31)) {
    g_err = "n must be < 2^31 for int32 CSR indices";
    return -1;
  }
  int *rowptr = nullptr, *colind = nullptr;
  double *val = nullptr, *b = nullptr, *x = nullptr, *r = nullptr, *p = nullptr, *Ap = nullptr;
  void* dBuffer = nullptr;
  RS_CUDA(cudaMalloc(&rowptr, (size_t)(n + 1) * sizeof(int)));
  RS_CUDA(cudaMemset(rowptr, 0, sizeof(int)));
  k_assemble<<<1184, 256>>>(g, n, rowptr, nullptr, nullptr, 0);
  RS_CUDA(cudaGetLastError());
  thrust::device_ptr<int> rp(rowptr);
  thrust::inclusive_scan(rp, rp + n + 1, rp);
  int nnz = 0;
  RS_CUDA(cudaMemcpy(&nnz, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost));
  if (nnz <= 0) {
    g_err = "nnz overflowed int32";
    return -1;
  }
  RS_CUDA(cudaMalloc(&colind, (size_t)nnz * sizeof(int)));
  RS_CUDA(cudaMalloc(&val, (size_t)nnz * sizeof(double)));
  k_assemble<<<1184, 256>>>(g, n, rowptr, colind, val, 1);
  RS_CUDA(cudaGetLastError());
  for (double** v : {&b, &x, &r, &p, &Ap}) RS_CUDA(cudaMalloc(v, (size_t)n * sizeof(double)));
  if (b_dev) RS_CUDA(cudaMemcpy(b, b_dev, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice));
  else k_fill_rhs<<<1184, 256>>>(b, n, 1);
  RS_CUDA(cudaMemset(x, 0, (size_t)n * sizeof(double)));  // the reference forgets this (CUDACG.cu:182-186)

  cublasHandle_t hb;
  cusparseHandle_t hs;
  RS_LIB(cublasCreate(&hb), CUBLAS_STATUS_SUCCESS);
  RS_LIB(cusparseCreate(&hs), CUSPARSE_STATUS_SUCCESS);
  cusparseSpMatDescr_t matA;
  cusparseDnVecDescr_t vecp, vecAp;
  RS_LIB(cusparseCreateCsr(&matA, n, n, nnz, rowptr, colind, val, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                           CUSPARSE_INDEX_BASE_ZERO, CUDA_R_64F), CUSPARSE_STATUS_SUCCESS);
  RS_LIB(cusparseCreateDnVec(&vecp, n, p, CUDA_R_64F), CUSPARSE_STATUS_SUCCESS);
  RS_LIB(cusparseCreateDnVec(&vecAp, n, Ap, CUDA_R_64F), CUSPARSE_STATUS_SUCCESS);
  const double one = 1.0, zero = 0.0;
  size_t bufferSize = 0;
  if (!faithful) {
    RS_LIB(cusparseSpMV_bufferSize(hs, CUSPARSE_OPERATION_NON_TRANSPOSE, &one, matA, vecp, &zero, vecAp, CUDA_R_64F,
                                   CUSPARSE_SPMV_ALG_DEFAULT, &bufferSize), CUSPARSE_STATUS_SUCCESS);
    RS_CUDA(cudaMalloc(&dBuffer, bufferSize ? bufferSize : 8));
  }
  const int nn = (int)n;
  double alpha, beta, rho, rhop, tmp;
  RS_LIB(cublasDcopy(hb, nn, b, 1, r, 1), CUBLAS_STATUS_SUCCESS);
  RS_LIB(cublasDcopy(hb, nn, b, 1, p, 1), CUBLAS_STATUS_SUCCESS);
  RS_LIB(cublasDnrm2(hb, nn, r, 1, &rho), CUBLAS_STATUS_SUCCESS);
  rho = rho * rho;
  cudaEvent_t e0, e1;
  RS_CUDA(cudaEventCreate(&e0));
  RS_CUDA(cudaEventCreate(&e1));
  for (int it = 0; it < warmup + steps; ++it) {
    if (it == warmup) {
      RS_CUDA(cudaDeviceSynchronize());
      RS_CUDA(cudaEventRecord(e0, 0));
    }
    if (faithful) {
      RS_LIB(cusparseSpMV_bufferSize(hs, CUSPARSE_OPERATION_NON_TRANSPOSE, &one, matA, vecp, &zero, vecAp, CUDA_R_64F,
                                     CUSPARSE_SPMV_ALG_DEFAULT, &bufferSize), CUSPARSE_STATUS_SUCCESS);
      if (dBuffer) RS_CUDA(cudaFree(dBuffer));  // the reference leaks here; freeing keeps long runs alive
      RS_CUDA(cudaMalloc(&dBuffer, bufferSize ? bufferSize : 8));
    }
    RS_LIB(cusparseSpMV(hs, CUSPARSE_OPERATION_NON_TRANSPOSE, &one, matA, vecp, &zero, vecAp, CUDA_R_64F,
                        CUSPARSE_SPMV_ALG_DEFAULT, dBuffer), CUSPARSE_STATUS_SUCCESS);
    RS_LIB(cublasDdot(hb, nn, p, 1, Ap, 1, &tmp), CUBLAS_STATUS_SUCCESS);
    alpha = rho / tmp;
    RS_LIB(cublasDaxpy(hb, nn, &alpha, p, 1, x, 1), CUBLAS_STATUS_SUCCESS);
    const double malpha = -alpha;
    RS_LIB(cublasDaxpy(hb, nn, &malpha, Ap, 1, r, 1), CUBLAS_STATUS_SUCCESS);
    rhop = rho;
    RS_LIB(cublasDnrm2(hb, nn, r, 1, &rho), CUBLAS_STATUS_SUCCESS);
    *out_rnorm = rho;
    if (hist) hist[it] = rho;
    rho = rho * rho;
    beta = rho / rhop;
    RS_LIB(cublasDscal(hb, nn, &beta, p, 1), CUBLAS_STATUS_SUCCESS);
    RS_LIB(cublasDaxpy(hb, nn, &one, r, 1, p, 1), CUBLAS_STATUS_SUCCESS);
  }
  RS_CUDA(cudaEventRecord(e1, 0));
  RS_CUDA(cudaEventSynchronize(e1));
  float ms = 0.f;
  RS_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  *out_ms = ms;
  *out_nnz = nnz;
  cusparseDestroySpMat(matA);
  cusparseDestroyDnVec(vecp);
  cusparseDestroyDnVec(vecAp);
  cusparseDestroy(hs);
  cublasDestroy(hb);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  for (void* v : {(void*)rowptr, (void*)colind, (void*)val, (void*)b, (void*)x, (void*)r, (void*)p, (void*)Ap, dBuffer})
    if (v) cudaFree(v);
  return 0;
}

extern "C" int cgb_refstyle_run(int kind, int gx, int gy, int gz, int faithful, int warmup, int steps, double* out_ms,
                                double* out_rnorm, long long* out_nnz) {
  return cgb_refstyle_run2(kind, gx, gy, gz, faithful, warmup, steps, nullptr, nullptr, out_ms, out_rnorm, out_nnz);
}
