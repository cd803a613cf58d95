// Reference-STYLE multi-GPU baseline "B4", written by this project (NOT upstream code: the reference is single-GPU,
// cudaSetDevice(0) at CUDACG.cu:87, and has no MPI/NCCL call at all).
//
// It is what a straightforward multi-GPU port of the reference's loop (CUDACG.cu:269-352) looks like when the
// communication is left to libraries: z-slab row blocks, one host thread per GPU, and per iteration
//   halo of p        : cudaMemcpyPeerAsync of the two boundary planes into the neighbours' ghost planes, then a
//                      blocking rendezvous (stream synchronise + host barrier) - the MPI_Sendrecv of such a port,
//   Ap = A p         : cusparseSpMV on the local CSR rows (columns index the local p with its two ghost planes),
//   p.Ap, ||r||^2    : cublasDdot into device memory, ncclAllReduce(sum), blocking read-back of the scalar
//                      (the reference blocks on both scalars too: CUDACG.cu:304, CUDACG.cu:328),
//   x, r, p updates  : cublasDaxpy / cublasDscal exactly as CUDACG.cu:314-347.
// This is the "path that only calls MPI or NCCL for the halo or dot products" that the product is measured against.
//
//   refstyle_b4 --kind 2 --grid 512 512 512 --gpus 8 --steps 100 --warmup 10 [--cc 6]
// prints one JSON line. RHS = the framework's hash-random generator (cgb_rhs_fill_box), so residuals are comparable.
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cusparse.h>
#include <nccl.h>
#include <pthread.h>
#include <thrust/device_ptr.h>
#include <thrust/scan.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "cgb200.h"

#define B4_CUDA(x)                                                                          \
  do {                                                                                      \
    cudaError_t e_ = (x);                                                                   \
    if (e_ != cudaSuccess) {                                                                \
      std::fprintf(stderr, "B4 rank %d: %s: %s\n", rank, #x, cudaGetErrorString(e_));       \
      std::exit(2);                                                                         \
    }                                                                                       \
  } while (0)
#define B4_OK(x, okv)                                                         \
  do {                                                                        \
    if ((x) != (okv)) {                                                       \
      std::fprintf(stderr, "B4 rank %d: %s failed\n", rank, #x);              \
      std::exit(2);                                                           \
    }                                                                         \
  } while (0)

namespace {

struct Grid {
  int kind, gx, gy, gz;
  double cc, cf, ce, cn;
};

__device__ __forceinline__ double coeff(const Grid& g, int dx, int dy, int dz) {
  const int m = (dx != 0) + (dy != 0) + (dz != 0);
  if (m == 0) return g.cc;
  if (m == 1) return g.cf;
  if (g.kind != 3) return 0.0;
  return m == 2 ? g.ce : g.cn;
}

// Local rows = planes [z0, z0+nzl). Column index = position in the local p, which carries one ghost plane below and
// one above the owned planes. pass 0 counts, pass 1 fills.
__global__ void k_assemble_local(Grid g, int z0, int nzl, int* rowptr, int* colind, double* val, int fill) {
  const long long plane = (long long)g.gx * g.gy, n = plane * nzl;
  for (long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(row % g.gx), j = (int)((row / g.gx) % g.gy), kl = (int)(row / plane);
    const int k = z0 + kl;
    int pos = fill ? rowptr[row] : 0;
    for (int dz = -1; dz <= 1; ++dz)
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          const double c = coeff(g, dx, dy, dz);
          if (c == 0.0) continue;
          const int ii = i + dx, jj = j + dy, kk = k + dz;
          if (ii < 0 || jj < 0 || kk < 0 || ii >= g.gx || jj >= g.gy || kk >= g.gz) continue;
          if (fill) {
            colind[pos] = (int)(ii + (long long)g.gx * (jj + (long long)g.gy * (kl + dz + 1)));
            val[pos] = c;
          }
          ++pos;
        }
    if (!fill) rowptr[row + 1] = pos;
  }
}

struct Shared {
  Grid g;
  int P, warmup, steps;
  pthread_barrier_t bar;
  std::vector<double*> p_base;      // per rank: local p incl. ghost planes (peer-accessible)
  std::vector<int> z0, nzl;
  std::vector<float> ms;
  std::vector<double> rnorm, r0norm;
  ncclComm_t* comms;
  long long nnz_total = 0;
};

void worker(Shared* sh, int rank) {
  const Grid g = sh->g;
  const int P = sh->P;
  B4_CUDA(cudaSetDevice(rank));
  for (int q = 0; q < P; ++q)
    if (q != rank) {
      cudaError_t e = cudaDeviceEnablePeerAccess(q, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) B4_CUDA(e);
      cudaGetLastError();
    }
  const int z0 = sh->z0[rank], nzl = sh->nzl[rank];
  const long long plane = (long long)g.gx * g.gy, n = plane * nzl;
  cudaStream_t st;
  B4_CUDA(cudaStreamCreate(&st));
  int *rowptr = nullptr, *colind = nullptr;
  double *val = nullptr, *b = nullptr, *x = nullptr, *r = nullptr, *pfull = nullptr, *Ap = nullptr, *dscal = nullptr;
  B4_CUDA(cudaMalloc(&rowptr, (size_t)(n + 1) * sizeof(int)));
  B4_CUDA(cudaMemset(rowptr, 0, sizeof(int)));
  k_assemble_local<<<1184, 256, 0, st>>>(g, z0, nzl, rowptr, nullptr, nullptr, 0);
  B4_CUDA(cudaStreamSynchronize(st));
  thrust::device_ptr<int> rp(rowptr);
  thrust::inclusive_scan(rp, rp + n + 1, rp);
  int nnz = 0;
  B4_CUDA(cudaMemcpy(&nnz, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost));
  B4_CUDA(cudaMalloc(&colind, (size_t)nnz * sizeof(int)));
  B4_CUDA(cudaMalloc(&val, (size_t)nnz * sizeof(double)));
  k_assemble_local<<<1184, 256, 0, st>>>(g, z0, nzl, rowptr, colind, val, 1);
  for (double** v : {&b, &x, &r, &Ap}) B4_CUDA(cudaMalloc(v, (size_t)n * sizeof(double)));
  B4_CUDA(cudaMalloc(&pfull, (size_t)(n + 2 * plane) * sizeof(double)));
  B4_CUDA(cudaMemset(pfull, 0, (size_t)(n + 2 * plane) * sizeof(double)));  // ghost planes at the domain ends stay 0
  B4_CUDA(cudaMalloc(&dscal, 2 * sizeof(double)));
  double* p = pfull + plane;
  {  // the framework's hash-random right-hand side (seed 1), generated on the host for this rank's planes
    cgb_stencil_t stc;
    // (a 2D grid nx x ny is stored as nx x 1 x ny by the framework and passed here as --grid nx 1 ny)
    if (g.kind == 1) cgb_stencil_default(g.kind, g.gx, g.gz, 1, &stc);
    else cgb_stencil_default(g.kind, g.gx, g.gy, g.gz, &stc);
    std::vector<double> hb((size_t)n);
    if (cgb_rhs_fill_box(&stc, CGB_RHS_RANDOM, 1, 0, g.gy, z0, nzl, hb.data()) != 0) std::exit(3);
    B4_CUDA(cudaMemcpy(b, hb.data(), (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
  }
  B4_CUDA(cudaMemset(x, 0, (size_t)n * sizeof(double)));
  sh->p_base[rank] = pfull;

  cublasHandle_t hb;
  cusparseHandle_t hs;
  B4_OK(cublasCreate(&hb), CUBLAS_STATUS_SUCCESS);
  B4_OK(cusparseCreate(&hs), CUSPARSE_STATUS_SUCCESS);
  B4_OK(cublasSetStream(hb, st), CUBLAS_STATUS_SUCCESS);
  B4_OK(cusparseSetStream(hs, st), CUSPARSE_STATUS_SUCCESS);
  cusparseSpMatDescr_t matA;
  cusparseDnVecDescr_t vecp, vecAp;
  B4_OK(cusparseCreateCsr(&matA, n, n + 2 * plane, nnz, rowptr, colind, val, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                          CUSPARSE_INDEX_BASE_ZERO, CUDA_R_64F), CUSPARSE_STATUS_SUCCESS);
  B4_OK(cusparseCreateDnVec(&vecp, n + 2 * plane, pfull, CUDA_R_64F), CUSPARSE_STATUS_SUCCESS);
  B4_OK(cusparseCreateDnVec(&vecAp, n, Ap, CUDA_R_64F), CUSPARSE_STATUS_SUCCESS);
  const double one = 1.0, zero = 0.0;
  size_t bufferSize = 0;
  void* dBuffer = nullptr;
  B4_OK(cusparseSpMV_bufferSize(hs, CUSPARSE_OPERATION_NON_TRANSPOSE, &one, matA, vecp, &zero, vecAp, CUDA_R_64F,
                                CUSPARSE_SPMV_ALG_DEFAULT, &bufferSize), CUSPARSE_STATUS_SUCCESS);
  B4_CUDA(cudaMalloc(&dBuffer, bufferSize ? bufferSize : 8));
  const int nn = (int)n;
  ncclComm_t comm = sh->comms[rank];

  // global sum of a local dot product: device-pointer cublasDdot -> ncclAllReduce -> blocking read-back
  auto global_dot = [&](const double* u, const double* v) -> double {
    B4_OK(cublasSetPointerMode(hb, CUBLAS_POINTER_MODE_DEVICE), CUBLAS_STATUS_SUCCESS);
    B4_OK(cublasDdot(hb, nn, u, 1, v, 1, dscal), CUBLAS_STATUS_SUCCESS);
    B4_OK(cublasSetPointerMode(hb, CUBLAS_POINTER_MODE_HOST), CUBLAS_STATUS_SUCCESS);
    if (P > 1) B4_OK(ncclAllReduce(dscal, dscal, 1, ncclDouble, ncclSum, comm, st), ncclSuccess);
    double h = 0.0;
    B4_CUDA(cudaMemcpyAsync(&h, dscal, sizeof(double), cudaMemcpyDeviceToHost, st));
    B4_CUDA(cudaStreamSynchronize(st));
    return h;
  };
  pthread_barrier_wait(&sh->bar);  // every p_base is published
  auto halo = [&]() {
    // my first plane -> lower neighbour's upper ghost; my last plane -> upper neighbour's lower ghost
    if (rank > 0)
      B4_CUDA(cudaMemcpyPeerAsync(sh->p_base[rank - 1] + plane * (1 + sh->nzl[rank - 1]), rank - 1, p, rank,
                                  (size_t)plane * sizeof(double), st));
    if (rank + 1 < P)
      B4_CUDA(cudaMemcpyPeerAsync(sh->p_base[rank + 1], rank + 1, p + plane * (nzl - 1), rank,
                                  (size_t)plane * sizeof(double), st));
    B4_CUDA(cudaStreamSynchronize(st));
    if (P > 1) pthread_barrier_wait(&sh->bar);  // both neighbours' planes have landed here
  };

  B4_OK(cublasDcopy(hb, nn, b, 1, r, 1), CUBLAS_STATUS_SUCCESS);   // CUDACG.cu:248
  B4_OK(cublasDcopy(hb, nn, b, 1, p, 1), CUBLAS_STATUS_SUCCESS);   // CUDACG.cu:255
  double rho = global_dot(r, r);                                    // CUDACG.cu:261-266
  sh->r0norm[rank] = std::sqrt(rho);
  cudaEvent_t e0, e1;
  B4_CUDA(cudaEventCreate(&e0));
  B4_CUDA(cudaEventCreate(&e1));
  double rn = 0.0;
  for (int it = 0; it < sh->warmup + sh->steps; ++it) {
    if (it == sh->warmup) {
      B4_CUDA(cudaStreamSynchronize(st));
      pthread_barrier_wait(&sh->bar);
      B4_CUDA(cudaEventRecord(e0, st));
    }
    halo();
    B4_OK(cusparseSpMV(hs, CUSPARSE_OPERATION_NON_TRANSPOSE, &one, matA, vecp, &zero, vecAp, CUDA_R_64F,
                       CUSPARSE_SPMV_ALG_DEFAULT, dBuffer), CUSPARSE_STATUS_SUCCESS);          // CUDACG.cu:288
    const double pAp = global_dot(p, Ap);                                                       // CUDACG.cu:304
    const double alpha = rho / pAp, malpha = -alpha;
    B4_OK(cublasDaxpy(hb, nn, &alpha, p, 1, x, 1), CUBLAS_STATUS_SUCCESS);                      // CUDACG.cu:314
    B4_OK(cublasDaxpy(hb, nn, &malpha, Ap, 1, r, 1), CUBLAS_STATUS_SUCCESS);                    // CUDACG.cu:320
    const double rhop = rho;
    rho = global_dot(r, r);                                                                     // CUDACG.cu:328
    rn = std::sqrt(rho);
    const double beta = rho / rhop;
    B4_OK(cublasDscal(hb, nn, &beta, p, 1), CUBLAS_STATUS_SUCCESS);                             // CUDACG.cu:342
    B4_OK(cublasDaxpy(hb, nn, &one, r, 1, p, 1), CUBLAS_STATUS_SUCCESS);                        // CUDACG.cu:347
  }
  B4_CUDA(cudaEventRecord(e1, st));
  B4_CUDA(cudaEventSynchronize(e1));
  B4_CUDA(cudaEventElapsedTime(&sh->ms[rank], e0, e1));
  sh->rnorm[rank] = rn;
  __atomic_fetch_add(&sh->nnz_total, (long long)nnz, __ATOMIC_RELAXED);
  pthread_barrier_wait(&sh->bar);
}

}  // namespace

int main(int argc, char** argv) {
  Shared sh;
  sh.g = Grid{2, 512, 512, 512, 0.0, -1.0, -1.0, -1.0};
  sh.P = 1;
  sh.warmup = 10;
  sh.steps = 100;
  double cc = 0.0;
  for (int a = 1; a < argc; ++a) {
    const std::string k = argv[a];
    if (k == "--kind" && a + 1 < argc) sh.g.kind = std::atoi(argv[++a]);
    else if (k == "--grid" && a + 3 < argc) { sh.g.gx = std::atoi(argv[a + 1]); sh.g.gy = std::atoi(argv[a + 2]); sh.g.gz = std::atoi(argv[a + 3]); a += 3; }
    else if (k == "--gpus" && a + 1 < argc) sh.P = std::atoi(argv[++a]);
    else if (k == "--steps" && a + 1 < argc) sh.steps = std::atoi(argv[++a]);
    else if (k == "--warmup" && a + 1 < argc) sh.warmup = std::atoi(argv[++a]);
    else if (k == "--cc" && a + 1 < argc) cc = std::atof(argv[++a]);
    else { std::fprintf(stderr, "unknown argument %s\n", argv[a]); return 1; }
  }
  sh.g.cc = cc != 0.0 ? cc : (sh.g.kind == 3 ? 26.0 : (sh.g.kind == 1 ? 4.0 : 6.0));
  int ndev = 0;
  cudaGetDeviceCount(&ndev);
  if (sh.P < 1 || sh.P > ndev || sh.P > sh.g.gz) { std::fprintf(stderr, "bad --gpus %d (devices: %d)\n", sh.P, ndev); return 1; }
  const int P = sh.P;
  sh.p_base.assign(P, nullptr);
  sh.ms.assign(P, 0.f);
  sh.rnorm.assign(P, 0.0);
  sh.r0norm.assign(P, 0.0);
  for (int q = 0; q < P; ++q) {  // same block distribution as the framework's z slabs
    const int base = sh.g.gz / P, rem = sh.g.gz % P;
    sh.nzl.push_back(base + (q < rem ? 1 : 0));
    sh.z0.push_back(q * base + (q < rem ? q : rem));
  }
  std::vector<ncclComm_t> comms(P);
  std::vector<int> devs(P);
  for (int q = 0; q < P; ++q) devs[q] = q;
  if (P > 1 && ncclCommInitAll(comms.data(), P, devs.data()) != ncclSuccess) { std::fprintf(stderr, "ncclCommInitAll failed\n"); return 2; }
  sh.comms = comms.data();
  pthread_barrier_init(&sh.bar, nullptr, P);
  std::vector<std::thread> th;
  for (int q = 0; q < P; ++q) th.emplace_back(worker, &sh, q);
  for (auto& t : th) t.join();
  float ms = 0.f;
  for (int q = 0; q < P; ++q) ms = sh.ms[q] > ms ? sh.ms[q] : ms;  // max over ranks
  const double n = (double)sh.g.gx * sh.g.gy * sh.g.gz, its = sh.steps / (ms * 1e-3);
  std::printf("{\"impl\": \"refstyle-b4\", \"value\": %.6f, \"unit\": \"CG iterations/s\", \"n_gpus\": %d, \"steps\": %d, \"warmup\": %d, "
              "\"ms_per_step\": %.6f, \"gdof_per_s\": %.6f, \"nnz\": %lld, \"r0norm\": %.17g, \"rnorm_after_steps\": %.17g, "
              "\"grid\": [%d, %d, %d], \"kind\": %d}\n",
              its, P, sh.steps, sh.warmup, ms / sh.steps, its * n / 1e9, sh.nnz_total, sh.r0norm[0], sh.rnorm[0], sh.g.gx, sh.g.gy,
              sh.g.gz, sh.g.kind);
  return 0;
}
