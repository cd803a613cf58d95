// Classic (Hestenes-Stiefel) CG as three fused persistent kernels per iteration, replacing the
// seven library calls + two blocking scalar read-backs + one cudaMalloc per iteration of
// CUDACG.cu:269-352:
//
//   k_spmv     Ap = A p  and  p.Ap        (cusparseSpMV :288 + cublasDdot :304) -> alpha on device
//   k_update   x += a p, r -= a Ap, r.r   (cublasDaxpy :314,:320 + cublasDnrm2 :328) -> test, beta
//   k_pupdate  p = r + b p                (cublasDscal :342 + cublasDaxpy :347) + halo push
//
// Each producing kernel's last CTA finishes the cross-GPU sum itself (one-shot P2P all-reduce,
// dev_common.cuh), so consumers read plain device scalars and the host never sits in the loop.
#include "cg_kernels.cuh"

namespace cgb {

namespace {

// ---------------------------------------------------------------------------------------------
// k_init: r = b, p = b, x = 0 (the reference never initialises x on the device, CUDACG.cu:182-186),
// rho = b.b (CUDACG.cu:261-266 without the sqrt/square round trip), halo push of p0.
// ---------------------------------------------------------------------------------------------
// x0_mode = 1 (warm start / restart from a checkpointed x): r = b - A x0 where A x0 was left in P.Ap by
// a plain SpMV, and x is kept.
__global__ void __launch_bounds__(kThreads) k_init(const __grid_constant__ RankDev P, double tol, int tol_mode,
                                                    int maxit, int bump_halo, int x0_mode) {
  __shared__ double red_smem[64];
  __shared__ unsigned cnt[CGB_NDIR];
  Scalars& S = P.ctrl->s;
  const cgb_layout_t& L = P.lay;
  const u64 version = S.halo_seq + 1;
  if (threadIdx.x < CGB_NDIR) cnt[threadIdx.x] = 0u;
  block_sync();
  double* __restrict__ p = P.exch[0];
  const UnitMap um{(L.nx + 1) / 2, L.ny};
  const unsigned stride = gridDim.x * kThreads;
  const unsigned tid0 = blockIdx.x * kThreads + threadIdx.x;
  double acc = 0.0;
  const double2 zero2 = make_double2(0.0, 0.0);
  if (P.nbr_mask) {
    const BRows B = brows_make(L);
    const unsigned UA = (unsigned)B.total * (unsigned)um.pw;
    for (unsigned u = tid0; u < UA; u += stride) {
      unsigned brow = u / (unsigned)um.pw;
      int i = 2 * (int)(u - brow * (unsigned)um.pw), j, k;
      brow_to_jk(B, L, (int)brow, j, k);
      unsigned mask = row_push_mask(L, j, k) & P.nbr_mask;
      if (!mask) continue;
      int64_t idx = lay_index(L, i, j, k);
      double2 v = ld2(P.b + idx);
      if (x0_mode) {
        const double2 ax = ld2(P.Ap + idx);
        v.x -= ax.x;
        v.y -= ax.y;
      } else {
        st2(P.x + idx, zero2);
      }
      st2(P.r + idx, v);
      st2(p + idx, v);
      acc += v.x * v.x + v.y * v.y;
      push_pair(P, 0, mask, i, j, k, v, cnt);
    }
    halo_commit(P, cnt, version);
  }
  const unsigned U = (unsigned)(L.ny * L.nz) * (unsigned)um.pw;
  for (unsigned u = tid0; u < U; u += stride) {
    int i, j, k;
    unit_to_ijk(um, u, i, j, k);
    if (P.nbr_mask && (row_push_mask(L, j, k) & P.nbr_mask)) continue;
    int64_t idx = lay_index(L, i, j, k);
    double2 v = ld2(P.b + idx);
    if (x0_mode) {
      const double2 ax = ld2(P.Ap + idx);
      v.x -= ax.x;
      v.y -= ax.y;
    } else {
      st2(P.x + idx, zero2);
    }
    st2(P.r + idx, v);
    st2(p + idx, v);
    acc += v.x * v.x + v.y * v.y;
  }
  double a = acc, z = 0.0;
  if (grid_sum2_last(a, z, P.partials, &S.ticket_red, red_smem)) {
    if (P.csr_dist) ready_publish(P, version);  // distributed CSR: my rows of p0 are complete
    bool ok = allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0) {
      S.rho = a;
      S.rho_prev = a;
      S.r0norm = sqrt(a);
      S.rnorm = S.r0norm;
      S.tol = tol;
      S.tol_mode = tol_mode;
      S.thresh = tol_mode == CGB_TOL_REL ? tol * S.r0norm : tol;
      S.iter = 0;
      S.maxit = maxit;
      S.cur = 0;
      S.alpha = 0.0;
      S.beta = 0.0;
      S.halo_seq = version + (bump_halo ? 1 : 0);
      S.done = !ok ? (int)CGB_ERR_TIMEOUT : (isfinite(a) && a > 0.0 ? 0 : (a == 0.0 ? 1 : (int)CGB_ERR_BREAKDOWN));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_stage_x0: exchange buffer 0 := x (owned cells) with the boundary rows pushed to the neighbours,
// so that a plain SpMV can form A x0 for a warm start.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_stage_x0(const __grid_constant__ RankDev P) {
  __shared__ double red_smem[64];
  __shared__ unsigned cnt[CGB_NDIR];
  Scalars& S = P.ctrl->s;
  const cgb_layout_t& L = P.lay;
  const u64 version = S.halo_seq + 1;
  if (threadIdx.x < CGB_NDIR) cnt[threadIdx.x] = 0u;
  block_sync();
  double* __restrict__ p = P.exch[0];
  const UnitMap um{(L.nx + 1) / 2, L.ny};
  const unsigned stride = gridDim.x * kThreads;
  const unsigned tid0 = blockIdx.x * kThreads + threadIdx.x;
  if (P.nbr_mask) {
    const BRows B = brows_make(L);
    const unsigned UA = (unsigned)B.total * (unsigned)um.pw;
    for (unsigned u = tid0; u < UA; u += stride) {
      unsigned brow = u / (unsigned)um.pw;
      int i = 2 * (int)(u - brow * (unsigned)um.pw), j, k;
      brow_to_jk(B, L, (int)brow, j, k);
      unsigned mask = row_push_mask(L, j, k) & P.nbr_mask;
      if (!mask) continue;
      int64_t idx = lay_index(L, i, j, k);
      double2 v = ld2(P.x + idx);
      st2(p + idx, v);
      push_pair(P, 0, mask, i, j, k, v, cnt);
    }
    halo_commit(P, cnt, version);
  }
  const unsigned U = (unsigned)(L.ny * L.nz) * (unsigned)um.pw;
  for (unsigned u = tid0; u < U; u += stride) {
    int i, j, k;
    unit_to_ijk(um, u, i, j, k);
    if (P.nbr_mask && (row_push_mask(L, j, k) & P.nbr_mask)) continue;
    int64_t idx = lay_index(L, i, j, k);
    st2(p + idx, ld2(P.x + idx));
  }
  double a = 0.0, z = 0.0;
  if (grid_sum2_last(a, z, P.partials, &S.ticket_red, red_smem)) {  // only to elect the last CTA
    if (threadIdx.x == 0) {
      S.halo_seq = version;
      S.done = 0;  // a previous solve may have left its final status here; the SpMV must run
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_spmv: matrix-free stencil SpMV, register marching along the slowest owned axis.
// Per plane q a thread keeps, for its two x-adjacent cells, S0 (centre), S1 (sum of the four
// in-plane face neighbours) and, for 27-pt, S2 (sum of the four in-plane diagonal neighbours):
//   Ap_k = cc*S0_k + cf*(S1_k + S0_{k-1} + S0_{k+1}) + ce*(S2_k + S1_{k-1} + S1_{k+1})
//          + cn*(S2_{k-1} + S2_{k+1})
// so every plane is loaded once per chunk and reused for three outputs.
// ---------------------------------------------------------------------------------------------
struct PS {
  double s0[2], s1[2], s2[2];
};

template <int KIND>
__device__ __forceinline__ void plane_center(const double* in, int64_t idx, PS& o) {
  double2 c = ld2(in + idx);
  o.s0[0] = c.x;
  o.s0[1] = c.y;
}

template <int KIND>
__device__ __forceinline__ void plane_full(const double* in, int64_t idx, int i, int nx, int pitch,
                                           PS& o) {
  const bool has_l = i > 0, has_r = i + 2 < nx;
  const double2 c = ld2(in + idx);
  const double xm = has_l ? in[idx - 1] : 0.0;
  const double xp = has_r ? in[idx + 2] : 0.0;
  o.s0[0] = c.x;
  o.s0[1] = c.y;
  if (KIND == CGB_KIND_POISSON2D5) {  // one row per plane: only the x neighbours are in-plane
    o.s1[0] = xm + c.y;
    o.s1[1] = c.x + xp;
    return;
  }
  const double2 u = ld2(in + idx - pitch);
  const double2 d = ld2(in + idx + pitch);
  o.s1[0] = (xm + c.y) + (u.x + d.x);
  o.s1[1] = (c.x + xp) + (u.y + d.y);
  if (KIND == CGB_KIND_STENCIL3D27) {
    const double um = has_l ? in[idx - pitch - 1] : 0.0, up = has_r ? in[idx - pitch + 2] : 0.0;
    const double dm = has_l ? in[idx + pitch - 1] : 0.0, dp = has_r ? in[idx + pitch + 2] : 0.0;
    o.s2[0] = (um + u.y) + (dm + d.y);
    o.s2[1] = (u.x + up) + (d.x + dp);
  }
}

template <int KIND>
// `in` is deliberately NOT __restrict__: its ghost layers are written by peer GPUs while this kernel runs, so
// the loads must stay ordinary (coherent) global loads ordered by the acquire in halo_wait_all; with __restrict__
// the compiler turns them into ld.global.nc (LDG.CONSTANT), which the memory model does not order after an acquire.
__global__ void __launch_bounds__(kThreads) k_spmv(const __grid_constant__ RankDev P, const double* in, double* out,
                                                    int mode) {
  __shared__ double red_smem[64];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const cgb_layout_t& L = P.lay;
  constexpr bool k2d = KIND == CGB_KIND_POISSON2D5;
  constexpr int TXC = k2d ? 2 * kThreads : 64;
  constexpr int TY = k2d ? 1 : kThreads / 32;
  const int lx = k2d ? threadIdx.x : (threadIdx.x & 31);
  const int ly = k2d ? 0 : (threadIdx.x >> 5);
  const int pitch = L.pitch;
  const int64_t ms = (int64_t)pitch * L.py;
  const int mext = L.nz;
  const double cc = P.st.c_center, cf = P.st.c_face, ce = P.st.c_edge, cn = P.st.c_corner;
  const int nitems = P.ntx * P.nty * P.nmc;
  const int mlo = dir_index(0, -1), mhi = dir_index(0, 1);
  const bool nb_mlo = P.nbr[mlo] >= 0, nb_mhi = P.nbr[mhi] >= 0;
  const bool nb_ylo = !k2d && P.nbr[dir_index(-1, 0)] >= 0, nb_yhi = !k2d && P.nbr[dir_index(1, 0)] >= 0;
  double acc = 0.0;
  for (int pass = 0; pass < 2; ++pass) {
    if (pass == 1) {
      if (!P.nbr_mask) break;
      // interior chunks are done; ghost layers are needed from here on
      if (!halo_wait_all(P, S.halo_seq)) break;
    }
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const int tx_t = item % P.ntx;
      const int rest = item / P.ntx;
      const int ty_t = rest % P.nty;
      const int mc = rest / P.nty;
      const bool bnd = (mc == 0 && nb_mlo) || (mc == P.nmc - 1 && nb_mhi) || (ty_t == 0 && nb_ylo) ||
                       (ty_t == P.nty - 1 && nb_yhi);
      if (bnd != (pass == 1)) continue;
      const int i = tx_t * TXC + 2 * lx;
      const int j = ty_t * TY + ly;
      if (i >= L.nx || (!k2d && j >= L.ny)) continue;
      const int m0 = mc * P.mchunk, m1 = min(mext, m0 + P.mchunk);
      int64_t idx = lay_index(L, i, j, m0 - 1);
      PS prev, cur, next;
      if (KIND == CGB_KIND_STENCIL3D27) plane_full<KIND>(in, idx, i, L.nx, pitch, prev);
      else plane_center<KIND>(in, idx, prev);
      idx += ms;
      plane_full<KIND>(in, idx, i, L.nx, pitch, cur);
      const bool pair_ok = i + 1 < L.nx;
      for (int m = m0; m < m1; ++m) {
        if (KIND == CGB_KIND_STENCIL3D27 || m + 1 < m1) plane_full<KIND>(in, idx + ms, i, L.nx, pitch, next);
        else plane_center<KIND>(in, idx + ms, next);
        double a0 = cc * cur.s0[0] + cf * (cur.s1[0] + (prev.s0[0] + next.s0[0]));
        double a1 = cc * cur.s0[1] + cf * (cur.s1[1] + (prev.s0[1] + next.s0[1]));
        if (KIND == CGB_KIND_STENCIL3D27) {
          a0 += ce * (cur.s2[0] + (prev.s1[0] + next.s1[0])) + cn * (prev.s2[0] + next.s2[0]);
          a1 += ce * (cur.s2[1] + (prev.s1[1] + next.s1[1])) + cn * (prev.s2[1] + next.s2[1]);
        }
        if (pair_ok) st2(out + idx, make_double2(a0, a1));
        else out[idx] = a0;
        acc += cur.s0[0] * a0 + (pair_ok ? cur.s0[1] * a1 : 0.0);
        prev = cur;
        cur = next;
        idx += ms;
      }
    }
  }
  if (mode == kSpmvPlain) return;
  double a = acc, z = 0.0;
  if (grid_sum2_last(a, z, P.partials, &S.ticket_red, red_smem)) {
    allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0 && mode == kSpmvAlpha) {
      S.pAp = a;
      S.alpha = S.rho / a;  // CUDACG.cu:311 (no breakdown guard there; NaN is caught in k_update)
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_spmv_csr: general CSR (fp64 values, int32 0-based indices, CUDACG.cu:213-216), G lanes per row.
// ---------------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(kThreads) k_spmv_csr(const __grid_constant__ RankDev P,
                                                        const double* __restrict__ in, double* __restrict__ out,
                                                        int mode) {
  __shared__ double red_smem[64];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const int n = P.lay.nx;
  const int lane = threadIdx.x % G;
  const int gpb = kThreads / G;
  double acc = 0.0;
  // Row-block distributed matrix: entries of p owned by other ranks are loaded straight from the
  // owners' arenas over NVLink, once every rank has announced that its part of this p is complete.
  const bool dist = P.csr_dist != 0;
  if (dist && !ready_wait(P, S.halo_seq)) return;
  auto fetch = [&](int c) -> double {
    if (!dist) return in[c];
    int li;
    const int q = csr_owner(P.csr_n_global, P.nranks, c, li);
    return q == P.rank ? in[li] : P.peer_exch0[q][li];
  };
  // The trip count depends only on blockIdx, so every lane of a warp reaches the shuffles.
  for (int base = blockIdx.x * gpb; base < n; base += gridDim.x * gpb) {
    const int row = base + threadIdx.x / G;
    const bool valid = row < n;
    const int e0 = valid ? P.csr_rowptr[row] : 0, e1 = valid ? P.csr_rowptr[row + 1] : 0;
    double s = 0.0;
    for (int e = e0 + lane; e < e1; e += G) s += P.csr_val[e] * fetch(P.csr_colind[e]);
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o, G);
    if (valid && lane == 0) {
      out[row] = s;
      acc += in[row] * s;
    }
  }
  if (mode == kSpmvPlain) return;
  double a = acc, z = 0.0;
  if (grid_sum2_last(a, z, P.partials, &S.ticket_red, red_smem)) {
    allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0 && mode == kSpmvAlpha) {
      S.pAp = a;
      S.alpha = S.rho / a;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_spmv_csr_stream: general CSR for matrices with short rows (what cusparseSpMV does at CUDACG.cu:288-291).
// The host cuts the rows into blocks of <= kThreads rows and <= kCsrCap nonzeros. A CTA streams the block's
// val/colind - one contiguous range of the CSR arrays - with 128-bit / 64-bit loads, all of them independent
// (two entries per thread per step), multiplies by the gathered x and parks the products in shared memory;
// then one thread per row adds its products in order (deterministic). A row longer than kCsrCap forms a block
// of its own and is reduced by the whole CTA. DRAM traffic: 12 B per nonzero + 4 B of row offsets + y, with x
// mostly served by L2; there is no per-row shuffle tree and no dependent rowptr -> colind -> x chain per lane.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_spmv_csr_stream(const __grid_constant__ RankDev P,
                                                               const double* __restrict__ in, double* __restrict__ out,
                                                               int mode) {
  __shared__ double prod[kCsrCap];
  __shared__ double red_smem[64];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const int* __restrict__ rowptr = P.csr_rowptr;
  const int* __restrict__ col = P.csr_colind;
  const double* __restrict__ val = P.csr_val;
  const int* __restrict__ rowblk = P.csr_rowblk;
  double acc = 0.0;
  for (int b = blockIdx.x; b < P.csr_nblk; b += gridDim.x) {
    const int r0 = rowblk[b], r1 = rowblk[b + 1];
    const int e0 = rowptr[r0], e1 = rowptr[r1];
    if (e1 - e0 <= kCsrCap) {
      const int ea = e0 & ~1;  // pairs start at even entries: val pairs are 16-byte, colind pairs 8-byte aligned
      for (int e = ea + 2 * (int)threadIdx.x; e < e1; e += 2 * kThreads) {
        if (e >= e0 && e + 1 < e1) {
          const double2 v = ld2(val + e);
          const int2 c = *reinterpret_cast<const int2*>(col + e);
          prod[e - e0] = v.x * in[c.x];
          prod[e + 1 - e0] = v.y * in[c.y];
        } else {
          if (e >= e0) prod[e - e0] = val[e] * in[col[e]];
          if (e + 1 < e1) prod[e + 1 - e0] = val[e + 1] * in[col[e + 1]];
        }
      }
      block_sync();
      const int r = r0 + (int)threadIdx.x;
      if (r < r1) {
        const int a = rowptr[r] - e0, z = rowptr[r + 1] - e0;
        double s = 0.0;
        for (int k = a; k < z; ++k) s += prod[k];
        out[r] = s;
        acc += in[r] * s;
      }
      block_sync();  // prod is reused by the next block
    } else {
      // a single long row: every thread strides over it, fixed-order tree over the CTA
      double s = 0.0, z = 0.0;
      for (int e = e0 + (int)threadIdx.x; e < e1; e += kThreads) s += val[e] * in[col[e]];
      block_sum2(s, z, red_smem);
      if (threadIdx.x == 0) {
        out[r0] = s;
        acc += in[r0] * s;
      }
      block_sync();
    }
  }
  if (mode == kSpmvPlain) return;
  double a = acc, z = 0.0;
  if (grid_sum2_last(a, z, P.partials, &S.ticket_red, red_smem)) {
    allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0 && mode == kSpmvAlpha) {
      S.pAp = a;
      S.alpha = S.rho / a;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_update: x += alpha p ; r -= alpha Ap ; rho' = r.r ; then, on the last CTA, the reference's
// stopping test (absolute ||r||_2 < tol, after the x/r update and before the p update,
// CUDACG.cu:333) and beta (CUDACG.cu:339).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_update(const __grid_constant__ RankDev P, const double* __restrict__ p) {
  __shared__ double red_smem[64];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const cgb_layout_t& L = P.lay;
  const double alpha = S.alpha;
  const UnitMap um{(L.nx + 1) / 2, L.ny};
  const unsigned U = (unsigned)(L.ny * L.nz) * (unsigned)um.pw;
  const unsigned stride = gridDim.x * kThreads;
  double* __restrict__ x = P.x;
  double* __restrict__ r = P.r;
  const double* __restrict__ Ap = P.Ap;
  double acc = 0.0;
  for (unsigned u0 = blockIdx.x * kThreads + threadIdx.x; u0 < U; u0 += 4 * stride) {
    int64_t idx[4];
    double2 vp[4], va[4], vx[4], vr[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      unsigned u = u0 + q * stride;
      idx[q] = -1;
      if (u < U) {
        int i, j, k;
        unit_to_ijk(um, u, i, j, k);
        idx[q] = lay_index(L, i, j, k);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (idx[q] >= 0) {
        vp[q] = ld2(p + idx[q]);
        va[q] = ld2(Ap + idx[q]);
        vx[q] = ld2(x + idx[q]);
        vr[q] = ld2(r + idx[q]);
      }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (idx[q] >= 0) {
        vx[q].x += alpha * vp[q].x;
        vx[q].y += alpha * vp[q].y;
        vr[q].x -= alpha * va[q].x;
        vr[q].y -= alpha * va[q].y;
        st2(x + idx[q], vx[q]);
        st2(r + idx[q], vr[q]);
        acc += vr[q].x * vr[q].x + vr[q].y * vr[q].y;
      }
  }
  double a = acc, z = 0.0;
  if (grid_sum2_last(a, z, P.partials, &S.ticket_red, red_smem)) {
    bool ok = allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0 && ok) cg_after_rr(P, a);
  }
}

// ---------------------------------------------------------------------------------------------
// k_pupdate: p = r + beta p. Boundary rows are produced first and stored both locally and, with
// plain P2P stores, straight into the neighbours' ghost layers; the CTA that completes a
// direction raises the neighbour's flag (release, system scope) while the interior is still
// being updated, so the halo is in flight during the rest of this kernel and the interior part
// of the next SpMV.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_pupdate(const __grid_constant__ RankDev P, double* __restrict__ p,
                                                       int exch_buf) {
  __shared__ unsigned cnt[CGB_NDIR];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  const cgb_layout_t& L = P.lay;
  const double beta = S.beta;
  const u64 version = S.halo_seq;
  const double* __restrict__ r = P.r;
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
      int64_t idx = lay_index(L, i, j, k);
      double2 vr = ld2(r + idx), vp = ld2(p + idx);
      vp.x = vr.x + beta * vp.x;
      vp.y = vr.y + beta * vp.y;
      st2(p + idx, vp);
      push_pair(P, exch_buf, mask, i, j, k, vp, cnt);
    }
    halo_commit(P, cnt, version);
  }
  const unsigned U = (unsigned)(L.ny * L.nz) * (unsigned)um.pw;
  for (unsigned u0 = tid0; u0 < U; u0 += 4 * stride) {
    int64_t idx[4];
    double2 vr[4], vp[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      unsigned u = u0 + q * stride;
      idx[q] = -1;
      if (u < U) {
        int i, j, k;
        unit_to_ijk(um, u, i, j, k);
        if (!(P.nbr_mask && (row_push_mask(L, j, k) & P.nbr_mask))) idx[q] = lay_index(L, i, j, k);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (idx[q] >= 0) {
        vr[q] = ld2(r + idx[q]);
        vp[q] = ld2(p + idx[q]);
      }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (idx[q] >= 0) {
        vp[q].x = vr[q].x + beta * vp[q].x;
        vp[q].y = vr[q].y + beta * vp[q].y;
        st2(p + idx[q], vp[q]);
      }
  }
  if (P.csr_dist) {
    // distributed CSR has no geometric halo: the last CTA tells every rank that this rank's rows of
    // the new p are in memory; the peers' SpMV then loads what it needs by index
    __shared__ int s_last;
    block_sync();
    if (threadIdx.x == 0) {
      __threadfence();
      s_last = atomicAdd(&S.ticket_red, 1u) == gridDim.x - 1;
    }
    block_sync();
    if (s_last) {
      if (threadIdx.x == 0) S.ticket_red = 0u;
      ready_publish(P, version);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// helpers: RHS generation on the device, compact <-> padded copies, standalone dot
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_rhs_gen(const __grid_constant__ RankDev P, int rhs_kind,
                                                       unsigned long long seed) {
  const cgb_layout_t& L = P.lay;
  const unsigned U = (unsigned)(L.ny * L.nz) * (unsigned)L.nx;
  for (unsigned u = blockIdx.x * kThreads + threadIdx.x; u < U; u += gridDim.x * kThreads) {
    unsigned row = u / (unsigned)L.nx;
    int i = (int)(u - row * (unsigned)L.nx);
    int k = (int)(row / (unsigned)L.ny);
    int j = (int)(row - (unsigned)k * (unsigned)L.ny);
    P.b[lay_index(L, i, j, k)] = rhs_value_ijk(P.st, rhs_kind, seed, i, j + P.y0, k + P.z0);
  }
}

__global__ void __launch_bounds__(kThreads) k_pad(cgb_layout_t L, double* compact, double* padded, int to_padded) {
  const long long U = (long long)L.owned;
  for (long long u = (long long)blockIdx.x * kThreads + threadIdx.x; u < U; u += (long long)gridDim.x * kThreads) {
    long long row = u / L.nx;
    int i = (int)(u - row * L.nx);
    int k = (int)(row / L.ny);
    int j = (int)(row - (long long)k * L.ny);
    int64_t idx = lay_index(L, i, j, k);
    if (to_padded) padded[idx] = compact[u];
    else compact[u] = padded[idx];
  }
}

__global__ void __launch_bounds__(kThreads) k_dot(const double* __restrict__ a, const double* __restrict__ b,
                                                   long long n, double2* partials, unsigned int* ticket,
                                                   double* out) {
  __shared__ double red_smem[64];
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kThreads)
    acc += a[i] * b[i];
  double z = 0.0;
  if (grid_sum2_last(acc, z, partials, ticket, red_smem)) {
    if (threadIdx.x == 0) *out = acc;
  }
}

// Last node of a conditional-WHILE graph body: keep looping while the solve is still running.
__global__ void k_cond_guard(const __grid_constant__ RankDev P, cudaGraphConditionalHandle h) {
  if (threadIdx.x == 0) cudaGraphSetConditional(h, P.ctrl->s.done == 0 ? 1u : 0u);
}

template <class K>
int occ_grid(K kernel, int device, int cap) {
  int nsm = 0, occ = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, 0);
  if (occ < 1) occ = 1;
  int g = nsm * occ;
  return g > cap ? cap : g;
}

}  // namespace

int query_launch_cfg(int device, LaunchCfg* cfg) {
  cfg->grid_ew = occ_grid(k_update, device, kMaxGrid);
  int g = occ_grid(k_spmv<CGB_KIND_STENCIL3D27>, device, kMaxGrid);
  cfg->grid_spmv = g;
  cfg->grid_csr = occ_grid(k_spmv_csr<8>, device, kMaxGrid);
  cfg->grid_csr_stream = occ_grid(k_spmv_csr_stream, device, kMaxGrid);
  return 0;
}

std::vector<int> csr_row_blocks(int n, const int* rowptr) {
  std::vector<int> blk;
  blk.reserve((size_t)n / 128 + 2);
  blk.push_back(0);
  int r = 0;
  while (r < n) {
    const int e0 = rowptr[r];
    int r1 = r + 1;  // a block always takes at least one row, however long
    while (r1 < n && r1 - r < kThreads && rowptr[r1 + 1] - e0 <= kCsrCap) ++r1;
    blk.push_back(r1);
    r = r1;
  }
  return blk;
}

void choose_spmv_tiling(RankDev& P, int grid_spmv) {
  const cgb_layout_t& L = P.lay;
  if (P.st.kind == CGB_KIND_POISSON2D5) {
    P.ntx = (L.nx + 2 * kThreads - 1) / (2 * kThreads);
    P.nty = 1;
    int ext = L.nz, mc = 32;
    while (mc > 4 && (long long)P.ntx * ((ext + mc - 1) / mc) < 2LL * grid_spmv) mc >>= 1;
    P.mchunk = mc;
    P.nmc = (ext + mc - 1) / mc;
  } else {
    P.ntx = (L.nx + 63) / 64;
    P.nty = (L.ny + 7) / 8;
    int ext = L.nz, mc = 32;
    while (mc > 4 && (long long)P.ntx * P.nty * ((ext + mc - 1) / mc) < 2LL * grid_spmv) mc >>= 1;
    P.mchunk = mc;
    P.nmc = (ext + mc - 1) / mc;
  }
}

cudaError_t launch_stage_x0(const RankDev& P, const LaunchCfg& c, cudaStream_t s) {
  k_stage_x0<<<c.grid_ew, kThreads, 0, s>>>(P);
  return cudaGetLastError();
}

cudaError_t launch_init(const RankDev& P, const LaunchCfg& c, double tol, int tol_mode, int maxit, int bump_halo,
                        int x0_mode, cudaStream_t s) {
  k_init<<<c.grid_ew, kThreads, 0, s>>>(P, tol, tol_mode, maxit, bump_halo, x0_mode);
  return cudaGetLastError();
}

cudaError_t launch_spmv_simple(const RankDev& P, const LaunchCfg& c, const double* in, double* out, int mode,
                               cudaStream_t s) {
  const int nitems = P.ntx * P.nty * P.nmc;
  int grid = c.grid_spmv < nitems ? c.grid_spmv : nitems;
  if (grid < 1) grid = 1;
  switch (P.st.kind) {
    case CGB_KIND_POISSON2D5: k_spmv<CGB_KIND_POISSON2D5><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    case CGB_KIND_POISSON3D7: k_spmv<CGB_KIND_POISSON3D7><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    case CGB_KIND_STENCIL3D27: k_spmv<CGB_KIND_STENCIL3D27><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_spmv_csr(const RankDev& P, const LaunchCfg& c, const double* in, double* out, int mode,
                            cudaStream_t s) {
  if (P.csr_rowblk && P.csr_nblk > 0 && !P.csr_dist) {  // row blocks were built: the streaming kernel
    int grid = P.csr_nblk < c.grid_csr_stream ? P.csr_nblk : c.grid_csr_stream;
    k_spmv_csr_stream<<<grid, kThreads, 0, s>>>(P, in, out, mode);
    return cudaGetLastError();
  }
  const int n = P.lay.nx;
  // lanes per row ~ average nonzeros per row, rounded to a power of two
  const long long nnz_avg = P.csr_avg_nnz;
  int G = 2;
  while (G < 32 && G < nnz_avg) G <<= 1;
  int gpb = kThreads / G;
  int grid = (n + gpb - 1) / gpb;
  if (grid > c.grid_csr) grid = c.grid_csr;
  if (grid < 1) grid = 1;
  switch (G) {
    case 2: k_spmv_csr<2><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    case 4: k_spmv_csr<4><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    case 8: k_spmv_csr<8><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    case 16: k_spmv_csr<16><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
    default: k_spmv_csr<32><<<grid, kThreads, 0, s>>>(P, in, out, mode); break;
  }
  return cudaGetLastError();
}

cudaError_t launch_cond_guard(const RankDev& P, unsigned long long handle, cudaStream_t s) {
  k_cond_guard<<<1, 32, 0, s>>>(P, (cudaGraphConditionalHandle)handle);
  return cudaGetLastError();
}

cudaError_t launch_update(const RankDev& P, const LaunchCfg& c, const double* p, cudaStream_t s) {
  k_update<<<c.grid_ew, kThreads, 0, s>>>(P, p);
  return cudaGetLastError();
}

cudaError_t launch_pupdate(const RankDev& P, const LaunchCfg& c, double* p, int exch_buf, cudaStream_t s) {
  k_pupdate<<<c.grid_ew, kThreads, 0, s>>>(P, p, exch_buf);
  return cudaGetLastError();
}

cudaError_t launch_rhs_gen(const RankDev& P, const LaunchCfg& c, int rhs_kind, unsigned long long seed,
                           cudaStream_t s) {
  k_rhs_gen<<<c.grid_ew, kThreads, 0, s>>>(P, rhs_kind, seed);
  return cudaGetLastError();
}

cudaError_t launch_pad(const cgb_layout_t& L, const double* compact, double* padded, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(padded, 0, (size_t)L.vec_elems * sizeof(double), s);
  if (e != cudaSuccess) return e;
  long long blocks = (L.owned + kThreads - 1) / kThreads;
  int grid = (int)(blocks > 1184 * 4 ? 1184 * 4 : (blocks < 1 ? 1 : blocks));
  k_pad<<<grid, kThreads, 0, s>>>(L, const_cast<double*>(compact), padded, 1);
  return cudaGetLastError();
}

cudaError_t launch_unpad(const cgb_layout_t& L, const double* padded, double* compact, cudaStream_t s) {
  long long blocks = (L.owned + kThreads - 1) / kThreads;
  int grid = (int)(blocks > 1184 * 4 ? 1184 * 4 : (blocks < 1 ? 1 : blocks));
  k_pad<<<grid, kThreads, 0, s>>>(L, compact, const_cast<double*>(padded), 0);
  return cudaGetLastError();
}

cudaError_t launch_dot(const double* a, const double* b, long long n, double* out, void* scratch, cudaStream_t s) {
  Control* ctrl = reinterpret_cast<Control*>(scratch);
  double2* partials = reinterpret_cast<double2*>(reinterpret_cast<char*>(scratch) + kPartialsOff);
  cudaError_t e = cudaMemsetAsync(&ctrl->s.ticket_red, 0, sizeof(unsigned int), s);
  if (e != cudaSuccess) return e;
  long long blocks = (n + kThreads - 1) / kThreads;
  int grid = (int)(blocks > 1184 ? 1184 : (blocks < 1 ? 1 : blocks));
  k_dot<<<grid, kThreads, 0, s>>>(a, b, n, partials, &ctrl->s.ticket_red, out);
  return cudaGetLastError();
}

}  // namespace cgb
