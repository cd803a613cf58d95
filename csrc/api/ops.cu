// Standalone ops on caller-owned device memory. Each op drives the SAME kernel the solver uses,
// with a one-rank RankDev whose control block lives in the caller's scratch buffer, so the unit
// tests (tier T2) exercise the production code path.
#include <climits>
#include <cstring>

#include "../common/status.hpp"
#include "../kernels/cg_kernels.cuh"

using namespace cgb;

namespace {

struct DevCfg {
  bool ready = false;
  LaunchCfg cfg{};
};
DevCfg g_cfg[64];

int get_cfg(LaunchCfg** out) {
  int dev = 0;
  CGB_CUDA(cudaGetDevice(&dev));
  CGB_REQUIRE(dev >= 0 && dev < 64, "device index out of range");
  if (!g_cfg[dev].ready) {
    CGB_TRY(query_launch_cfg(dev, &g_cfg[dev].cfg));
    g_cfg[dev].ready = true;
  }
  *out = &g_cfg[dev].cfg;
  return CGB_OK;
}

// One-rank RankDev over scratch; uploads a fresh control block image.
int make_op_rank(const cgb_stencil_t* st, const cgb_layout_t* lay, void* scratch, const Scalars& s_img,
                 cudaStream_t stream, RankDev* out) {
  CGB_REQUIRE(scratch != nullptr, "op scratch is NULL (allocate cgb_op_scratch_bytes())");
  RankDev P{};
  P.ctrl = (Control*)scratch;
  P.peer_ctrl[0] = P.ctrl;
  P.partials = (double2*)((char*)scratch + kPartialsOff);
  P.nranks = 1;
  P.rank = 0;
  P.lay = *lay;
  if (st) P.st = *st;
  for (int d = 0; d < CGB_NDIR; ++d) P.nbr[d] = -1;
  CGB_CUDA(cudaMemsetAsync(scratch, 0, offsetof(Control, s), stream));
  CGB_CUDA(cudaMemcpyAsync(&P.ctrl->s, &s_img, sizeof(Scalars), cudaMemcpyHostToDevice, stream));
  *out = P;
  return CGB_OK;
}

Scalars blank_scalars() {
  Scalars s{};
  s.timeout_cycles = 1LL << 34;
  s.maxit = INT_MAX;
  s.thresh = -1.0;
  s.rho = 1.0;
  return s;
}

}  // namespace

extern "C" size_t cgb_op_scratch_bytes(void) { return kExchOff; }

extern "C" int cgb_op_dot(const double* a, const double* b, int64_t n, double* out, void* scratch, void* stream) {
  CGB_REQUIRE(a && b && out && scratch && n >= 0, "cgb_op_dot: bad argument");
  CGB_CUDA(launch_dot(a, b, n, out, scratch, (cudaStream_t)stream));
  return CGB_OK;
}

extern "C" int cgb_op_spmv_stencil(const cgb_stencil_t* st, const cgb_layout_t* lay, const double* p, double* Ap,
                                   double* dot_out, void* scratch, int impl, void* stream) {
  CGB_REQUIRE(st && lay && p && Ap, "cgb_op_spmv_stencil: NULL argument");
  CGB_REQUIRE(kind_is_stencil(st->kind), "cgb_op_spmv_stencil: not a stencil problem");
  CGB_REQUIRE(impl >= CGB_SPMV_AUTO && impl <= CGB_SPMV_TMA, "unknown spmv impl %d", impl);
  cudaStream_t s = (cudaStream_t)stream;
  LaunchCfg* cfg;
  CGB_TRY(get_cfg(&cfg));
  RankDev P;
  CGB_TRY(make_op_rank(st, lay, scratch, blank_scalars(), s, &P));
  choose_spmv_tiling(P, cfg->grid_spmv);
  const bool tma = impl != CGB_SPMV_SIMPLE && spmv_tma_supported(P);
  if (impl == CGB_SPMV_TMA && !tma) return fail(CGB_ERR_UNSUPPORTED, "TMA SpMV needs a 3D stencil problem");
  if (tma) {
    alignas(64) unsigned char map[128], map_h[128];
    CGB_TRY(spmv_tma_make_map(st->kind, *lay, p, map));
    CGB_TRY(spmv_tma_make_map(st->kind, *lay, p, map_h, 1, 2));
    int dev = 0;
    CGB_CUDA(cudaGetDevice(&dev));
    CGB_CUDA(launch_spmv_tma(P, dev, map, map_h, nullptr, nullptr, Ap, dot_out ? kSpmvDot : kSpmvPlain, s));
  } else {
    CGB_CUDA(launch_spmv_simple(P, *cfg, p, Ap, dot_out ? kSpmvDot : kSpmvPlain, s));
  }
  if (dot_out) CGB_CUDA(cudaMemcpyAsync(dot_out, &P.ctrl->s.last_sum[0], sizeof(double), cudaMemcpyDeviceToDevice, s));
  return CGB_OK;
}

extern "C" int cgb_op_spmv_csr(int32_t n, const int32_t* rowptr, const int32_t* colind, const double* val,
                               const double* x, double* y, double* dot_out, void* scratch, void* stream) {
  CGB_REQUIRE(n > 0 && rowptr && colind && val && x && y, "cgb_op_spmv_csr: bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  LaunchCfg* cfg;
  CGB_TRY(get_cfg(&cfg));
  cgb_layout_t lay{};
  lay.nx = n;
  lay.ny = lay.nz = lay.py = lay.pz = 1;
  lay.pitch = (n + 15) / 16 * 16;
  lay.vec_elems = lay.pitch;
  lay.owned = n;
  RankDev P;
  CGB_TRY(make_op_rank(nullptr, &lay, scratch, blank_scalars(), s, &P));
  P.csr_rowptr = rowptr;
  P.csr_colind = colind;
  P.csr_val = val;
  int nnz_last = 0;
  CGB_CUDA(cudaMemcpyAsync(&nnz_last, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  P.csr_avg_nnz = (nnz_last + n - 1) / n;
  CGB_CUDA(launch_spmv_csr(P, *cfg, x, y, dot_out ? kSpmvDot : kSpmvPlain, s));
  if (dot_out) CGB_CUDA(cudaMemcpyAsync(dot_out, &P.ctrl->s.last_sum[0], sizeof(double), cudaMemcpyDeviceToDevice, s));
  return CGB_OK;
}

extern "C" int cgb_op_fused_update(const cgb_layout_t* lay, double alpha, const double* p, const double* Ap,
                                   double* x, double* r, double* rr_out, void* scratch, void* stream) {
  CGB_REQUIRE(lay && p && Ap && x && r, "cgb_op_fused_update: NULL argument");
  cudaStream_t s = (cudaStream_t)stream;
  LaunchCfg* cfg;
  CGB_TRY(get_cfg(&cfg));
  Scalars img = blank_scalars();
  img.alpha = alpha;
  RankDev P;
  CGB_TRY(make_op_rank(nullptr, lay, scratch, img, s, &P));
  P.x = x;
  P.r = r;
  P.Ap = const_cast<double*>(Ap);
  P.hist = nullptr;
  P.hist_cap = 0;
  CGB_CUDA(launch_update(P, *cfg, p, s));
  if (rr_out) CGB_CUDA(cudaMemcpyAsync(rr_out, &P.ctrl->s.last_sum[0], sizeof(double), cudaMemcpyDeviceToDevice, s));
  return CGB_OK;
}

extern "C" int cgb_op_p_update(const cgb_layout_t* lay, double beta, const double* r, double* p, void* scratch,
                               void* stream) {
  CGB_REQUIRE(lay && r && p, "cgb_op_p_update: NULL argument");
  cudaStream_t s = (cudaStream_t)stream;
  LaunchCfg* cfg;
  CGB_TRY(get_cfg(&cfg));
  Scalars img = blank_scalars();
  img.beta = beta;
  RankDev P;
  CGB_TRY(make_op_rank(nullptr, lay, scratch, img, s, &P));
  P.r = const_cast<double*>(r);
  CGB_CUDA(launch_pupdate(P, *cfg, p, 0, s));
  return CGB_OK;
}

extern "C" int cgb_op_pad(const cgb_layout_t* lay, const double* compact, double* padded, void* stream) {
  CGB_REQUIRE(lay && compact && padded, "cgb_op_pad: NULL argument");
  CGB_CUDA(launch_pad(*lay, compact, padded, (cudaStream_t)stream));
  return CGB_OK;
}

extern "C" int cgb_op_unpad(const cgb_layout_t* lay, const double* padded, double* compact, void* stream) {
  CGB_REQUIRE(lay && padded && compact, "cgb_op_unpad: NULL argument");
  CGB_CUDA(launch_unpad(*lay, padded, compact, (cudaStream_t)stream));
  return CGB_OK;
}

// Band-kernel geometry for a local box (no GPU needed): out = {LX, LY, RB, HY, slab rows, boxes per row, x tiles,
// row bands, bands, z chunks, planes per chunk, work items, stage bytes, stages, grid, halo CTAs}. Returns 1 if the
// TMA kernels cover the layout, 0 otherwise. mode: 0 SpMV, 4 update, 5 merged p-update+SpMV.
extern "C" int cgb_debug_tma_geometry(const cgb_stencil_t* st, const cgb_layout_t* lay, int nbr_lo, int nbr_hi, int nsm,
                                      int mode, int* out16) {
  if (!st || !lay || !out16) return 0;
  RankDev P{};
  P.st = *st;
  P.lay = *lay;
  for (int d = 0; d < CGB_NDIR; ++d) P.nbr[d] = -1;
  if (nbr_lo) {
    P.nbr[dir_index(0, -1)] = 0;
    P.nbr_mask |= 1u << dir_index(0, -1);
  }
  if (nbr_hi) {
    P.nbr[dir_index(0, 1)] = 0;
    P.nbr_mask |= 1u << dir_index(0, 1);
  }
  return spmv_tma_geometry(P, nsm, mode, out16);
}

extern "C" int cgb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
