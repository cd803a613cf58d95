// Launch wrappers of the CG kernels (definitions in cg_kernels.cu, spmv_tma.cu, pipecg_kernels.cu).
#pragma once
#include <vector>

#include "dev_common.cuh"

namespace cgb {

enum SpmvMode {
  kSpmvAlpha = 0,  // classic CG: all-reduce p.Ap, alpha = rho / p.Ap stored on the device
  kSpmvDot = 1,    // all-reduce in.out, result only in Scalars::last_sum
  kSpmvPlain = 2,  // no reduction at all
  // TMA kernel only ("recompute" variant): trade a second stencil evaluation for 16 B/DOF
  kSpmvAlphaNoStore = 3,  // like kSpmvAlpha but Ap is never written
  kSpmvUpdate = 4,        // recompute Ap from p; x += alpha p; r -= alpha Ap; r.r -> test, beta
  kSpmvPUpdate = 5        // (k_pspmv_tma) p_new = r + beta p_old folded into the SpMV, p.Ap -> alpha
};

struct LaunchCfg {
  int grid_ew;    // persistent grid of the elementwise kernels
  int grid_spmv;  // persistent grid of the simple stencil SpMV
  int grid_csr;
  int grid_csr_stream;
};

constexpr int kCsrCap = 2048;  // nonzeros per row block of k_spmv_csr_stream (16 KB of products in shared memory)
// Cut rows [0,n) into blocks of <= 256 rows and <= kCsrCap nonzeros (a longer row gets a block of its own).
// Host only; returns the block boundaries (nblk+1 entries).
std::vector<int> csr_row_blocks(int n, const int* rowptr);

// Occupancy-derived persistent grid sizes for this device.
int query_launch_cfg(int device, LaunchCfg* cfg);
// Choose the SpMV work decomposition (ntx, nty, nmc, mchunk) for a layout.
void choose_spmv_tiling(RankDev& P, int grid_spmv);

// r = b, p = b, x = 0, push p's boundary rows, rho = all-reduce(b.b), thresh, iter = 0.
// bump_halo: the next kernel that pushes halos is not preceded by an x/r update (merged variant), so
// the version it will publish is announced here.
// x0_mode: r = b - (A x0 found in P.Ap), x kept (warm start).
cudaError_t launch_init(const RankDev& P, const LaunchCfg& c, double tol, int tol_mode, int maxit, int bump_halo,
                        int x0_mode, cudaStream_t s);
// exchange buffer 0 := x with halo push (first step of a warm start).
cudaError_t launch_stage_x0(const RankDev& P, const LaunchCfg& c, cudaStream_t s);
// out = A in (+ reduction according to mode); simple = register-marching kernel.
cudaError_t launch_spmv_simple(const RankDev& P, const LaunchCfg& c, const double* in, double* out,
                               int mode, cudaStream_t s);
// TMA-fed stencil SpMV (spmv_tma.cu). `map` points at a 128-byte CUtensorMap of the input vector.
bool spmv_tma_supported(const RankDev& P);
int spmv_tma_geometry(const RankDev& P, int nsm, int mode, int out[16]);
int spmv_tma_make_map(int kind, const cgb_layout_t& L, const double* vec, void* map_out, int halo_rows = 1,
                      int box_cols = 128);
// map_h: the input's 2-column halo-box map (rows wider than 512 cells). map_x / map_r (halo_rows = 0 maps of
// x and r) are needed for mode kSpmvUpdate only.
cudaError_t launch_spmv_tma(const RankDev& P, int device, const void* map, const void* map_h, const void* map_x,
                            const void* map_r, double* out, int mode, cudaStream_t s);
// "merged" variant: p-update folded into the SpMV (maps: r with halo rows, p_old, p_new buffers).
cudaError_t launch_pspmv_tma(const RankDev& P, int device, const void* const maps[6], const double* pold, double* pnew,
                             int newbuf, cudaStream_t s);
// Persistent cooperative kernel: `niter` whole iterations of the merged variant in one launch (grid barrier +
// all-reduce on the device). maps: r with halo rows, exch[0], exch[1], their three halo-box twins, x, r (band rows).
cudaError_t launch_cg_persist(const RankDev& P, int device, const void* d_maps, int cur, int niter, cudaStream_t s);
size_t cg_persist_maps_bytes();                                            // size of the device-resident map block
cudaError_t cg_persist_upload_maps(void* d_maps, const void* const maps[8]);  // host maps -> device block (synchronous)
bool cg_persist_supported(int device);
cudaError_t cg_persist_prepare(int device);  // one-time per-device set-up (allocates): call from the prepare path
cudaError_t launch_spmv_csr(const RankDev& P, const LaunchCfg& c, const double* in, double* out, int mode,
                            cudaStream_t s);
// x += alpha p; r -= alpha Ap; rho' = all-reduce(r.r); convergence test; beta.
cudaError_t launch_update(const RankDev& P, const LaunchCfg& c, const double* p, cudaStream_t s);
// p = r + beta p, boundary rows first with P2P push into the neighbours' ghost layers.
cudaError_t launch_pupdate(const RankDev& P, const LaunchCfg& c, double* p, int exch_buf, cudaStream_t s);

// Pipelined CG (pipecg_kernels.cu): wbuf = exchange buffer holding the current w.
cudaError_t launch_pipe_start(const RankDev& P, const LaunchCfg& c, int wbuf, cudaStream_t s);
cudaError_t launch_pipe_update(const RankDev& P, const LaunchCfg& c, int wbuf, cudaStream_t s);

// Sets a CUDA-graph conditional handle to (done == 0): tail of the WHILE-graph body.
cudaError_t launch_cond_guard(const RankDev& P, unsigned long long handle, cudaStream_t s);

cudaError_t launch_rhs_gen(const RankDev& P, const LaunchCfg& c, int rhs_kind, unsigned long long seed,
                           cudaStream_t s);
cudaError_t launch_pad(const cgb_layout_t& L, const double* compact, double* padded, cudaStream_t s);
cudaError_t launch_unpad(const cgb_layout_t& L, const double* padded, double* compact, cudaStream_t s);
cudaError_t launch_dot(const double* a, const double* b, long long n, double* out, void* scratch,
                       cudaStream_t s);

}  // namespace cgb
