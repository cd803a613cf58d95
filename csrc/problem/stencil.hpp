// Problem layer: structured stencil descriptors and right-hand-side generators, usable from host
// and device code. Generalises the literal 3x3 system of CUDACG.cu:74-117,137-140 to the grids
// named by BASELINE.json (2D 5-pt, 3D 7-pt, 3D 27-pt), keeping the matrix implicit.
#pragma once
#include <cmath>
#include <cstdint>

#include "cgb200.h"

#ifdef __CUDACC__
#define CGB_HD __host__ __device__ __forceinline__
#else
#define CGB_HD inline
#endif

namespace cgb {

// A 2D grid nx x ny is stored as nx x 1 x ny: one row per plane. It then shares everything with the
// 3D problems (z-slab decomposition = row blocks, ghost planes, kernels); it simply has no y taps.
CGB_HD bool kind_is_3d(int kind) { return kind == CGB_KIND_POISSON3D7 || kind == CGB_KIND_STENCIL3D27; }
CGB_HD bool kind_is_stencil(int kind) { return kind != CGB_KIND_CSR; }
CGB_HD bool kind_has_y_taps(int kind) { return kind_is_3d(kind); }
CGB_HD bool kind_has_z_taps(int kind) { return kind_is_stencil(kind); }

// Coefficient for an offset (dx,dy,dz) in {-1,0,1}^3; 0.0 if the offset is not part of the stencil.
CGB_HD double stencil_coeff(const cgb_stencil_t& st, int dx, int dy, int dz) {
  int m = (dx != 0) + (dy != 0) + (dz != 0);
  if (m == 0) return st.c_center;
  if (m == 1) {
    if (st.kind == CGB_KIND_POISSON2D5 && dy != 0) return 0.0;
    return st.c_face;
  }
  if (st.kind != CGB_KIND_STENCIL3D27) return 0.0;
  return m == 2 ? st.c_edge : st.c_corner;
}

CGB_HD uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// Uniform in [-1,1), a pure function of (seed, global index): identical on every rank count.
CGB_HD double hash_uniform(uint64_t seed, uint64_t gidx) {
  uint64_t h = splitmix64(gidx ^ splitmix64(seed));
  return (double)(h >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}

CGB_HD double manufactured_u(const cgb_stencil_t& st, int i, int j, int k) {
  if (i < 0 || j < 0 || k < 0 || i >= st.gx || j >= st.gy || k >= st.gz) return 0.0;
  const double pi = 3.14159265358979323846;
  double u = sin(pi * (double)(i + 1) / (double)(st.gx + 1));
  if (st.gy > 1) u *= sin(pi * (double)(j + 1) / (double)(st.gy + 1));
  if (st.gz > 1) u *= sin(pi * (double)(k + 1) / (double)(st.gz + 1));
  return u;
}

// b = A u_manufactured evaluated with the stencil (zero Dirichlet outside the grid).
CGB_HD double manufactured_b(const cgb_stencil_t& st, int i, int j, int k) {
  double s = 0.0;
  int zr = kind_has_z_taps(st.kind) ? 1 : 0;
  for (int dz = -zr; dz <= zr; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        double c = stencil_coeff(st, dx, dy, dz);
        if (c != 0.0) s += c * manufactured_u(st, i + dx, j + dy, k + dz);
      }
  return s;
}

CGB_HD double rhs_value_ijk(const cgb_stencil_t& st, int rhs_kind, uint64_t seed, int i, int j, int k) {
  if (rhs_kind == CGB_RHS_ONES) return 1.0;
  if (rhs_kind == CGB_RHS_RANDOM)
    return hash_uniform(seed, (uint64_t)i + (uint64_t)st.gx * ((uint64_t)j + (uint64_t)st.gy * (uint64_t)k));
  return manufactured_b(st, i, j, k);
}

}  // namespace cgb
