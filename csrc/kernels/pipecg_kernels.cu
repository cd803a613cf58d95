// Pipelined CG (Ghysels & Vanroose): ONE fused two-scalar reduction per iteration, overlapped with
// the SpMV. Absent from the reference (classic CG with two blocking reductions, CUDACG.cu:304,328);
// required by BASELINE.json config 4.
//
//   iteration:  KP1  q = A w                              (k_spmv / k_spmv_tma, no reduction)
//               KP2  consume (gamma, delta); alpha, beta; z = q + b z; s = w + b s; p = r + b p;
//                    x += a p; r -= a s; w' = w - a z  -> other exchange buffer + halo push;
//                    publish (gamma', delta') = (r.r, w'.r) WITHOUT waiting.
//
// The published 16-byte messages cross NVLink while KP1 of the next iteration runs; every CTA of
// the next KP2 completes the reduction (rank-ordered sum => identical scalars on all ranks).
// Stopping rule = the reference's (absolute ||r||_2 < tol, tested on the residual produced by
// the previous update, CUDACG.cu:333); detection lags one SpMV.
#include "cg_kernels.cuh"

namespace cgb {

namespace {

// Runs `body(idx, push_mask, i, j, k)` over all owned pairs: boundary rows first (then the halo
// flags are raised), interior afterwards.
template <class Body>
__device__ __forceinline__ void for_pairs_boundary_first(const RankDev& P, unsigned* cnt, u64 version, Body&& body) {
  const cgb_layout_t& L = P.lay;
  const UnitMap um{(L.nx + 1) / 2, L.ny};
  const unsigned stride = gridDim.x * kThreads;
  const unsigned tid0 = blockIdx.x * kThreads + threadIdx.x;
  if (P.nbr_mask) {
    if (threadIdx.x < CGB_NDIR) cnt[threadIdx.x] = 0u;
    block_sync();
    const BRows B = brows_make(L);
    const unsigned UA = (unsigned)B.total * (unsigned)um.pw;
    for (unsigned u = tid0; u < UA; u += stride) {
      unsigned brow = u / (unsigned)um.pw;
      int i = 2 * (int)(u - brow * (unsigned)um.pw), j, k;
      brow_to_jk(B, L, (int)brow, j, k);
      unsigned mask = row_push_mask(L, j, k) & P.nbr_mask;
      if (!mask) continue;
      body(lay_index(L, i, j, k), mask, i, j, k);
    }
    halo_commit(P, cnt, version);
  }
  const unsigned U = (unsigned)(L.ny * L.nz) * (unsigned)um.pw;
  for (unsigned u = tid0; u < U; u += stride) {
    int i, j, k;
    unit_to_ijk(um, u, i, j, k);
    if (P.nbr_mask && (row_push_mask(L, j, k) & P.nbr_mask)) continue;
    body(lay_index(L, i, j, k), 0u, i, j, k);
  }
}

// After k_init (r = b, x = 0) and w = A r: zero z/s/p, push w's boundary, publish (r.r, w.r).
__global__ void __launch_bounds__(kThreads) k_pipe_start(const __grid_constant__ RankDev P, int wbuf) {
  __shared__ double red_smem[64];
  __shared__ unsigned cnt[CGB_NDIR];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const u64 version = S.halo_seq + 1;
  const double* __restrict__ w = P.exch[wbuf];
  const double* __restrict__ r = P.r;
  double g = 0.0, d = 0.0;
  const double2 zero2 = make_double2(0.0, 0.0);
  for_pairs_boundary_first(P, cnt, version, [&](int64_t idx, unsigned mask, int i, int j, int k) {
    const double2 vw = ld2(w + idx), vr = ld2(r + idx);
    st2(P.z + idx, zero2);
    st2(P.s + idx, zero2);
    st2(P.pp + idx, zero2);
    g += vr.x * vr.x + vr.y * vr.y;
    d += vw.x * vr.x + vw.y * vr.y;
    if (mask) push_pair(P, wbuf, mask, i, j, k, vw, cnt);
  });
  if (grid_sum2_last(g, d, P.partials, &S.ticket_red, red_smem)) {
    allreduce2_publish(P, g, d, red_smem);
    if (threadIdx.x == 0) {
      S.cur = wbuf;
      S.halo_seq = version;
      S.gamma_old = 0.0;
      S.alpha_old = 0.0;
      S.iter = 0;
    }
  }
}

__global__ void __launch_bounds__(kThreads) k_pipe_update(const __grid_constant__ RankDev P, int wbuf) {
  __shared__ double red_smem[64];
  __shared__ unsigned cnt[CGB_NDIR];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const int it = S.iter;
  const u64 version = S.halo_seq + 1;
  double gamma, delta;
  const bool ok = allreduce2_consume(P, S.pub_seq, gamma, delta, red_smem);
  const double rn = sqrt(gamma);
  const double thresh = it == 0 ? (S.tol_mode == CGB_TOL_REL ? S.tol * rn : S.tol) : S.thresh;
  const bool bad = !ok || !isfinite(gamma) || !isfinite(delta);
  const bool conv = it > 0 && rn < thresh;
  const bool stop = bad || conv || it >= S.maxit;
  double alpha = 0.0, beta = 0.0, g = 0.0, d = 0.0;
  if (!stop) {
    if (it > 0) {
      beta = gamma / S.gamma_old;
      alpha = gamma / (delta - beta * gamma / S.alpha_old);
    } else {
      alpha = gamma / delta;
    }
    const double* __restrict__ q = P.Ap;
    const double* __restrict__ w = P.exch[wbuf];
    double* __restrict__ wn = P.exch[wbuf ^ 1];
    for_pairs_boundary_first(P, cnt, version, [&](int64_t idx, unsigned mask, int i, int j, int k) {
      const double2 vq = ld2(q + idx), vw = ld2(w + idx);
      double2 vz = ld2(P.z + idx), vs = ld2(P.s + idx), vp = ld2(P.pp + idx), vx = ld2(P.x + idx), vr = ld2(P.r + idx);
      vz.x = vq.x + beta * vz.x;  vz.y = vq.y + beta * vz.y;
      vs.x = vw.x + beta * vs.x;  vs.y = vw.y + beta * vs.y;
      vp.x = vr.x + beta * vp.x;  vp.y = vr.y + beta * vp.y;
      vx.x += alpha * vp.x;       vx.y += alpha * vp.y;
      vr.x -= alpha * vs.x;       vr.y -= alpha * vs.y;
      double2 vn;
      vn.x = vw.x - alpha * vz.x; vn.y = vw.y - alpha * vz.y;
      st2(P.z + idx, vz);
      st2(P.s + idx, vs);
      st2(P.pp + idx, vp);
      st2(P.x + idx, vx);
      st2(P.r + idx, vr);
      st2(wn + idx, vn);
      g += vr.x * vr.x + vr.y * vr.y;
      d += vn.x * vr.x + vn.y * vr.y;
      if (mask) push_pair(P, wbuf ^ 1, mask, i, j, k, vn, cnt);
    });
  }
  if (grid_sum2_last(g, d, P.partials, &S.ticket_red, red_smem)) {
    if (!stop) allreduce2_publish(P, g, d, red_smem);
    if (threadIdx.x == 0) {
      if (it == 0) {
        S.r0norm = rn;
        S.thresh = thresh;
      } else {
        if (it - 1 < P.hist_cap) P.hist[it - 1] = rn;  // ||r|| after iteration it-1
        if (P.hist_host && it - 1 < P.hist_host_cap) P.hist_host[it - 1] = rn;
      }
      S.rnorm = rn;
      if (!ok) S.done = CGB_ERR_TIMEOUT;
      else if (bad) S.done = CGB_ERR_BREAKDOWN;
      else if (conv) S.done = 1;
      else if (it >= S.maxit) S.done = 2;
      else {
        S.gamma_old = gamma;
        S.alpha_old = alpha;
        S.alpha = alpha;
        S.beta = beta;
        S.iter = it + 1;
        S.cur = wbuf ^ 1;
        S.halo_seq = version;
      }
    }
  }
}

}  // namespace

cudaError_t launch_pipe_start(const RankDev& P, const LaunchCfg& c, int wbuf, cudaStream_t s) {
  k_pipe_start<<<c.grid_ew, kThreads, 0, s>>>(P, wbuf);
  return cudaGetLastError();
}

cudaError_t launch_pipe_update(const RankDev& P, const LaunchCfg& c, int wbuf, cudaStream_t s) {
  k_pipe_update<<<c.grid_ew, kThreads, 0, s>>>(P, wbuf);
  return cudaGetLastError();
}

}  // namespace cgb
