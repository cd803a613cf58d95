// Host side of the problem layer: defaults, CSR assembly of the stencil operators (used by the
// CSR kernels' tests and by the library-call baseline), the reference 3x3 system.
#include <algorithm>
#include <vector>

#include "../common/status.hpp"
#include "stencil.hpp"

using namespace cgb;

extern "C" int cgb_stencil_default(int kind, int gx, int gy, int gz, cgb_stencil_t* out) {
  CGB_REQUIRE(out != nullptr, "cgb_stencil_default: out is NULL");
  CGB_REQUIRE(kind >= CGB_KIND_CSR && kind <= CGB_KIND_STENCIL3D27, "unknown problem kind %d", kind);
  CGB_REQUIRE(gx >= 1 && gy >= 1 && gz >= 1, "grid extents must be >= 1 (got %d,%d,%d)", gx, gy, gz);
  if (kind == CGB_KIND_POISSON2D5) {
    // stored as nx x 1 x ny (see stencil.hpp); (nx, ny, 1) is accepted and rearranged
    if (gz == 1 && gy > 1) {
      gz = gy;
      gy = 1;
    }
    CGB_REQUIRE(gy == 1, "2D problem: pass (nx, ny, 1) or (nx, 1, ny)");
  }
  if (kind == CGB_KIND_CSR) CGB_REQUIRE(gy == 1 && gz == 1, "CSR problem is 1D: gy == gz == 1");
  cgb_stencil_t s{};
  s.kind = kind;
  s.gx = gx;
  s.gy = gy;
  s.gz = gz;
  switch (kind) {
    case CGB_KIND_POISSON2D5: s.c_center = 4.0; s.c_face = -1.0; break;
    case CGB_KIND_POISSON3D7: s.c_center = 6.0; s.c_face = -1.0; break;
    case CGB_KIND_STENCIL3D27: s.c_center = 26.0; s.c_face = s.c_edge = s.c_corner = -1.0; break;
    default: s.c_center = 1.0; break;
  }
  *out = s;
  return CGB_OK;
}

namespace {
template <class F>
void for_each_entry(const cgb_stencil_t& st, int64_t row, F&& f) {
  int i = (int)(row % st.gx);
  int j = (int)((row / st.gx) % st.gy);
  int k = (int)(row / ((int64_t)st.gx * st.gy));
  int zr = kind_has_z_taps(st.kind) ? 1 : 0;
  for (int dz = -zr; dz <= zr; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        double c = stencil_coeff(st, dx, dy, dz);
        if (c == 0.0) continue;
        int ii = i + dx, jj = j + dy, kk = k + dz;
        if (ii < 0 || jj < 0 || kk < 0 || ii >= st.gx || jj >= st.gy || kk >= st.gz) continue;
        f((int64_t)ii + (int64_t)st.gx * (jj + (int64_t)st.gy * kk), c);
      }
}
}  // namespace

extern "C" int64_t cgb_stencil_csr_nnz(const cgb_stencil_t* st, int64_t row0, int64_t nrows) {
  if (!st || !kind_is_stencil(st->kind)) return -1;
  int64_t nnz = 0;
  for (int64_t r = row0; r < row0 + nrows; ++r) for_each_entry(*st, r, [&](int64_t, double) { ++nnz; });
  return nnz;
}

extern "C" int cgb_stencil_to_csr(const cgb_stencil_t* st, int64_t row0, int64_t nrows, int64_t* rowptr,
                                  int32_t* colind, double* val) {
  CGB_REQUIRE(st && rowptr && colind && val, "cgb_stencil_to_csr: NULL argument");
  CGB_REQUIRE(kind_is_stencil(st->kind), "cgb_stencil_to_csr: not a stencil problem");
  int64_t n = (int64_t)st->gx * st->gy * st->gz;
  CGB_REQUIRE(row0 >= 0 && nrows >= 0 && row0 + nrows <= n, "row range out of bounds");
  CGB_REQUIRE(n < (int64_t)1 << 31, "column indices are int32: n must be < 2^31");
  int64_t pos = 0;
  for (int64_t r = 0; r < nrows; ++r) {
    rowptr[r] = pos;
    for_each_entry(*st, row0 + r, [&](int64_t c, double v) {
      colind[pos] = (int32_t)c;
      val[pos] = v;
      ++pos;
    });
  }
  rowptr[nrows] = pos;
  return CGB_OK;
}

// A=[[3,0,2],[0,2,0],[2,0,1]], b=(3.5,1.5,2.0), solution (0.5,0.75,1.0): the values documented in
// the comment block CUDACG.cu:74-82.
extern "C" void cgb_ref3x3(int32_t* n, int32_t* nnz, double* val, int32_t* rowptr, int32_t* colind,
                           double* b) {
  static const double A[3][3] = {{3, 0, 2}, {0, 2, 0}, {2, 0, 1}};
  int32_t pos = 0;
  for (int r = 0; r < 3; ++r) {
    if (rowptr) rowptr[r] = pos;
    for (int c = 0; c < 3; ++c)
      if (A[r][c] != 0.0) {
        if (val) val[pos] = A[r][c];
        if (colind) colind[pos] = c;
        ++pos;
      }
  }
  if (rowptr) rowptr[3] = pos;
  if (n) *n = 3;
  if (nnz) *nnz = pos;
  if (b) {
    b[0] = 3.5;
    b[1] = 1.5;
    b[2] = 2.0;
  }
}

extern "C" double cgb_rhs_value(const cgb_stencil_t* st, int rhs_kind, uint64_t seed, int64_t gidx) {
  int i = (int)(gidx % st->gx);
  int j = (int)((gidx / st->gx) % st->gy);
  int k = (int)(gidx / ((int64_t)st->gx * st->gy));
  return rhs_value_ijk(*st, rhs_kind, seed, i, j, k);
}

// Owned box [y0,y0+ny) x [z0,z0+nz) (full rows) of the generated right-hand side, compact with x fastest:
// the host-side twin of the device generator, so a rank can stage exactly the RHS the device would generate.
extern "C" int cgb_rhs_fill_box(const cgb_stencil_t* st, int rhs_kind, uint64_t seed, int y0, int ny, int z0, int nz,
                                double* out) {
  if (!st || !out || ny < 0 || nz < 0 || y0 < 0 || z0 < 0 || y0 + ny > st->gy || z0 + nz > st->gz) return CGB_ERR_INVALID;
  if (rhs_kind < CGB_RHS_ONES || rhs_kind > CGB_RHS_MANUFACTURED) return CGB_ERR_INVALID;
  int64_t o = 0;
  for (int k = z0; k < z0 + nz; ++k)
    for (int j = y0; j < y0 + ny; ++j)
      for (int i = 0; i < st->gx; ++i) out[o++] = rhs_value_ijk(*st, rhs_kind, seed, i, j, k);
  return CGB_OK;
}

extern "C" double cgb_manufactured_u(const cgb_stencil_t* st, int64_t gidx) {
  int i = (int)(gidx % st->gx);
  int j = (int)((gidx / st->gx) % st->gy);
  int k = (int)(gidx / ((int64_t)st->gx * st->gy));
  return manufactured_u(*st, i, j, k);
}
