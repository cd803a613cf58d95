// CPU oracle (test tiers T1/T7, BASELINE config 1).
//
// Plain fp64 restatement of the reference algorithm CUDACG.cu:248-352 — x0 = 0, r = p = b,
// rho = (||r||_2)^2 via a norm that is then squared (CUDACG.cu:261-266, 328-336), ABSOLUTE test
// ||r||_2 < tol placed after the x/r update and before the p update (CUDACG.cu:333) — for CSR
// matrices and for the matrix-free stencils. The stencil solver runs `nranks` emulated ranks in
// one process: every rank owns the same padded box the GPU path uses, halos are copied box to
// box and scalars are summed in rank order, so partition / halo / reduction index logic is
// exercised without a GPU ("fake fabric").
#include <cmath>
#include <cstring>
#include <vector>

#include "../common/status.hpp"
#include "../topo/partition.hpp"

using namespace cgb;

namespace {

struct CpuRank {
  cgb_partition_t part;
  cgb_layout_t lay;
  std::vector<double> x, r, p, Ap, b;  // classic
  std::vector<double> w, s, z, q;      // pipelined extras
};

double local_spmv(const cgb_stencil_t& st, const cgb_layout_t& L, const double* in, double* out,
                  const double* dot_with) {
  const int zr = kind_has_z_taps(st.kind) ? 1 : 0;
  struct Tap { int dx; int64_t off; double c; };
  std::vector<Tap> taps;
  for (int dz = -zr; dz <= zr; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        double co = stencil_coeff(st, dx, dy, dz);
        if (co != 0.0) taps.push_back({dx, dx + (int64_t)L.pitch * (dy + (int64_t)L.py * dz), co});
      }
  double acc = 0.0;
  for (int k = 0; k < L.nz; ++k)
    for (int j = 0; j < L.ny; ++j)
      for (int i = 0; i < L.nx; ++i) {
        int64_t c = lay_index(L, i, j, k);
        double s = 0.0;
        for (const Tap& t : taps) {
          if (i + t.dx < 0 || i + t.dx >= L.nx) continue;  // y/z neighbours come from ghost layers
          s += t.c * in[c + t.off];
        }
        out[c] = s;
        if (dot_with) acc += dot_with[c] * s;
      }
  return acc;
}

template <class F>
void for_owned(const cgb_layout_t& L, F&& f) {
  for (int k = 0; k < L.nz; ++k)
    for (int j = 0; j < L.ny; ++j) {
      int64_t base = lay_index(L, 0, j, k);
      for (int i = 0; i < L.nx; ++i) f(base + i);
    }
}

// Copy boundary rows of `sel(rank)` into the neighbours' ghost rows (what the GPU p-update
// kernel does with P2P stores).
template <class Sel>
void halo_exchange(std::vector<CpuRank>& R, Sel&& sel) {
  for (auto& me : R) {
    const cgb_layout_t& L = me.lay;
    for (int d = 0; d < CGB_NDIR; ++d) {
      int q = me.part.nbr[d];
      if (q < 0) continue;
      CpuRank& nb = R[q];
      const cgb_layout_t& N = nb.lay;
      int dy = dir_dy(d), dz = dir_dz(d);
      int j0 = dy < 0 ? 0 : (dy > 0 ? L.ny - 1 : 0), j1 = dy == 0 ? L.ny : j0 + 1;
      int k0 = dz < 0 ? 0 : (dz > 0 ? L.nz - 1 : 0), k1 = dz == 0 ? L.nz : k0 + 1;
      for (int k = k0; k < k1; ++k)
        for (int j = j0; j < j1; ++j) {
          int jn = dy < 0 ? N.ny : (dy > 0 ? -1 : j);
          int kn = dz < 0 ? N.nz : (dz > 0 ? -1 : k);
          std::memcpy(&sel(nb)[lay_index(N, 0, jn, kn)], &sel(me)[lay_index(L, 0, j, k)],
                      sizeof(double) * L.nx);
        }
    }
  }
}

int setup_ranks(const cgb_stencil_t& st, int nranks, int decomp, int Py, int Pz, const double* b,
                bool pipelined, std::vector<CpuRank>& R) {
  R.resize(nranks);
  for (int q = 0; q < nranks; ++q) {
    CGB_TRY(make_partition(st, nranks, q, decomp, Py, Pz, &R[q].part));
    CGB_TRY(make_layout(st, R[q].part, &R[q].lay));
    const cgb_layout_t& L = R[q].lay;
    size_t e = (size_t)L.vec_elems;
    R[q].x.assign(e, 0.0);
    R[q].r.assign(e, 0.0);
    R[q].p.assign(e, 0.0);
    R[q].Ap.assign(e, 0.0);
    R[q].b.assign(e, 0.0);
    if (pipelined) {
      R[q].w.assign(e, 0.0);
      R[q].s.assign(e, 0.0);
      R[q].z.assign(e, 0.0);
      R[q].q.assign(e, 0.0);
    }
    for (int k = 0; k < L.nz; ++k)
      for (int j = 0; j < L.ny; ++j)
        for (int i = 0; i < L.nx; ++i)
          R[q].b[lay_index(L, i, j, k)] =
              b[(int64_t)i + (int64_t)st.gx * ((j + R[q].part.y0) + (int64_t)st.gy * (k + R[q].part.z0))];
  }
  return CGB_OK;
}

void gather_x(const cgb_stencil_t& st, const std::vector<CpuRank>& R, double* x) {
  for (const auto& me : R) {
    const cgb_layout_t& L = me.lay;
    for (int k = 0; k < L.nz; ++k)
      for (int j = 0; j < L.ny; ++j)
        for (int i = 0; i < L.nx; ++i)
          x[(int64_t)i + (int64_t)st.gx * ((j + me.part.y0) + (int64_t)st.gy * (k + me.part.z0))] =
              me.x[lay_index(L, i, j, k)];
  }
}

int cg_classic(const cgb_stencil_t& st, std::vector<CpuRank>& R, double tol, int tol_mode, int maxit,
               double* hist, cgb_cpu_result_t* res) {
  double sum = 0.0;
  for (auto& me : R) {  // r = b, p = b  (CUDACG.cu:248-255)
    double acc = 0.0;
    for_owned(me.lay, [&](int64_t c) {
      me.r[c] = me.b[c];
      me.p[c] = me.b[c];
      acc += me.b[c] * me.b[c];
    });
    sum += acc;
  }
  double rn = std::sqrt(sum);  // CUDACG.cu:261
  double rho = rn * rn;        // CUDACG.cu:266
  res->r0norm = rn;
  res->rnorm = rn;
  res->iters = 0;
  res->converged = 0;
  const double thresh = tol_mode == CGB_TOL_REL ? tol * rn : tol;
  for (int it = 0; it < maxit; ++it) {
    halo_exchange(R, [](CpuRank& q) -> std::vector<double>& { return q.p; });
    double pAp = 0.0;
    for (auto& me : R) pAp += local_spmv(st, me.lay, me.p.data(), me.Ap.data(), me.p.data());
    double alpha = rho / pAp;  // CUDACG.cu:311, no breakdown guard (as in the reference)
    sum = 0.0;
    for (auto& me : R) {
      double acc = 0.0;
      for_owned(me.lay, [&](int64_t c) {
        me.x[c] += alpha * me.p[c];   // CUDACG.cu:314
        me.r[c] -= alpha * me.Ap[c];  // CUDACG.cu:320-321
        acc += me.r[c] * me.r[c];
      });
      sum += acc;
    }
    double rhop = rho;
    rn = std::sqrt(sum);  // CUDACG.cu:328
    if (hist) hist[it] = rn;
    res->iters = it + 1;
    res->rnorm = rn;
    if (rn < thresh) {  // CUDACG.cu:333
      res->converged = 1;
      break;
    }
    rho = rn * rn;             // CUDACG.cu:336
    double beta = rho / rhop;  // CUDACG.cu:339
    for (auto& me : R) for_owned(me.lay, [&](int64_t c) { me.p[c] = me.r[c] + beta * me.p[c]; });
  }
  return CGB_OK;
}

// Ghysels & Vanroose pipelined CG, unpreconditioned: one fused reduction (gamma, delta) per
// iteration, overlappable with q = A w. Same stopping semantics as the classic loop.
int cg_pipelined(const cgb_stencil_t& st, std::vector<CpuRank>& R, double tol, int tol_mode, int maxit,
                 double* hist, cgb_cpu_result_t* res) {
  for (auto& me : R) for_owned(me.lay, [&](int64_t c) { me.r[c] = me.b[c]; });
  halo_exchange(R, [](CpuRank& q) -> std::vector<double>& { return q.r; });
  for (auto& me : R) local_spmv(st, me.lay, me.r.data(), me.w.data(), nullptr);
  double gamma_old = 0.0, alpha_old = 0.0, thresh = 0.0;
  res->iters = 0;
  res->converged = 0;
  for (int it = 0;; ++it) {
    double gamma = 0.0, delta = 0.0;
    for (auto& me : R) {
      double g = 0.0, d = 0.0;
      for_owned(me.lay, [&](int64_t c) {
        g += me.r[c] * me.r[c];
        d += me.w[c] * me.r[c];
      });
      gamma += g;
      delta += d;
    }
    double rn = std::sqrt(gamma);
    if (it == 0) {
      res->r0norm = rn;
      thresh = tol_mode == CGB_TOL_REL ? tol * rn : tol;
    } else {
      if (hist) hist[it - 1] = rn;
      res->iters = it;
      if (rn < thresh) {
        res->converged = 1;
        res->rnorm = rn;
        break;
      }
    }
    res->rnorm = rn;
    if (it >= maxit) break;
    halo_exchange(R, [](CpuRank& q) -> std::vector<double>& { return q.w; });
    for (auto& me : R) local_spmv(st, me.lay, me.w.data(), me.q.data(), nullptr);
    double beta, alpha;
    if (it > 0) {
      beta = gamma / gamma_old;
      alpha = gamma / (delta - beta * gamma / alpha_old);
    } else {
      beta = 0.0;
      alpha = gamma / delta;
    }
    for (auto& me : R)
      for_owned(me.lay, [&](int64_t c) {
        me.z[c] = me.q[c] + beta * me.z[c];
        me.s[c] = me.w[c] + beta * me.s[c];
        me.p[c] = me.r[c] + beta * me.p[c];
        me.x[c] += alpha * me.p[c];
        me.r[c] -= alpha * me.s[c];
        me.w[c] -= alpha * me.z[c];
      });
    gamma_old = gamma;
    alpha_old = alpha;
  }
  return CGB_OK;
}

}  // namespace

extern "C" int cgb_cpu_cg_csr(int64_t n, const int64_t* rowptr, const int32_t* colind, const double* val,
                              const double* b, double* x, double tol, int tol_mode, int maxit, double* hist,
                              cgb_cpu_result_t* res) {
  CGB_REQUIRE(n > 0 && rowptr && colind && val && b && x && res, "cgb_cpu_cg_csr: bad argument");
  std::vector<double> r(b, b + n), p(b, b + n), Ap((size_t)n);
  std::fill(x, x + n, 0.0);
  double sum = 0.0;
  for (int64_t i = 0; i < n; ++i) sum += r[i] * r[i];
  double rn = std::sqrt(sum), rho = rn * rn;
  res->r0norm = rn;
  res->rnorm = rn;
  res->iters = 0;
  res->converged = 0;
  const double thresh = tol_mode == CGB_TOL_REL ? tol * rn : tol;
  for (int it = 0; it < maxit; ++it) {
    double pAp = 0.0;
    for (int64_t i = 0; i < n; ++i) {
      double s = 0.0;
      for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e) s += val[e] * p[colind[e]];
      Ap[i] = s;
      pAp += p[i] * s;
    }
    double alpha = rho / pAp;
    sum = 0.0;
    for (int64_t i = 0; i < n; ++i) {
      x[i] += alpha * p[i];
      r[i] -= alpha * Ap[i];
      sum += r[i] * r[i];
    }
    double rhop = rho;
    rn = std::sqrt(sum);
    if (hist) hist[it] = rn;
    res->iters = it + 1;
    res->rnorm = rn;
    if (rn < thresh) {
      res->converged = 1;
      break;
    }
    rho = rn * rn;
    double beta = rho / rhop;
    for (int64_t i = 0; i < n; ++i) p[i] = r[i] + beta * p[i];
  }
  return CGB_OK;
}

extern "C" int cgb_cpu_cg_stencil(const cgb_stencil_t* st, int nranks, int decomp, int Py, int Pz,
                                  const double* b, double* x, double tol, int tol_mode, int maxit, int variant,
                                  double* hist, cgb_cpu_result_t* res) {
  CGB_REQUIRE(st && b && x && res, "cgb_cpu_cg_stencil: NULL argument");
  CGB_REQUIRE(kind_is_stencil(st->kind), "cgb_cpu_cg_stencil: not a stencil problem");
  CGB_REQUIRE(variant == 0 || variant == 1, "variant must be 0 (classic) or 1 (pipelined)");
  std::vector<CpuRank> R;
  CGB_TRY(setup_ranks(*st, nranks, decomp, Py, Pz, b, variant == 1, R));
  CGB_TRY(variant == 0 ? cg_classic(*st, R, tol, tol_mode, maxit, hist, res)
                       : cg_pipelined(*st, R, tol, tol_mode, maxit, hist, res));
  gather_x(*st, R, x);
  return CGB_OK;
}

extern "C" int cgb_cpu_spmv_stencil(const cgb_stencil_t* st, const double* x, double* y) {
  CGB_REQUIRE(st && x && y, "cgb_cpu_spmv_stencil: NULL argument");
  std::vector<CpuRank> R;
  CGB_TRY(setup_ranks(*st, 1, 1, 0, 0, x, false, R));  // b := x
  local_spmv(*st, R[0].lay, R[0].b.data(), R[0].x.data(), nullptr);
  gather_x(*st, R, y);
  return CGB_OK;
}

// out = A in on ONE rank's padded box (ghost layers of `in` filled by the caller); returns in.out
// over owned cells through dot_out. Used by the multi-process CPU path (gloo) and by tests.
extern "C" int cgb_cpu_spmv_local(const cgb_stencil_t* st, const cgb_layout_t* lay, const double* in, double* out,
                                  double* dot_out) {
  CGB_REQUIRE(st && lay && in && out, "cgb_cpu_spmv_local: NULL argument");
  double d = local_spmv(*st, *lay, in, out, in);
  if (dot_out) *dot_out = d;
  return CGB_OK;
}
