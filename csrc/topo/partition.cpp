#include "partition.hpp"

#include <cmath>

#include "../common/status.hpp"

namespace cgb {

int make_partition(const cgb_stencil_t& st, int nranks, int rank, int decomp, int Py, int Pz,
                   cgb_partition_t* out) {
  CGB_REQUIRE(out != nullptr, "make_partition: out is NULL");
  CGB_REQUIRE(nranks >= 1 && nranks <= CGB_MAX_RANKS, "nranks must be in [1,%d] (got %d)", CGB_MAX_RANKS, nranks);
  CGB_REQUIRE(rank >= 0 && rank < nranks, "rank %d out of range [0,%d)", rank, nranks);
  const bool is3d = kind_is_3d(st.kind);
  if (st.kind == CGB_KIND_CSR) {
    // general CSR: contiguous row blocks, no geometric neighbours (remote entries are fetched by index)
    CGB_REQUIRE(nranks <= st.gx, "more ranks than rows");
    cgb_partition_t p{};
    p.nranks = nranks;
    p.rank = rank;
    p.Py = p.Pz = 1;
    block_range(st.gx, nranks, rank, &p.x0, &p.nx);
    p.ny = p.nz = 1;
    for (int d = 0; d < CGB_NDIR; ++d) p.nbr[d] = -1;
    *out = p;
    return CGB_OK;
  } else if (!is3d) {
    // 2D grids are nx x 1 x ny: only the row (z) direction can be split
    CGB_REQUIRE(Py == 0 || Py == 1, "2D problems are split along their rows only");
    CGB_REQUIRE(Pz == 0 || Pz == nranks, "2D problems need Pz == nranks");
    Py = 1;
    Pz = nranks;
  } else if (Py == 0 && Pz == 0) {
    if (decomp == 2) {
      Py = 1;
      for (int f = 1; f * f <= nranks; ++f)
        if (nranks % f == 0) Py = f;  // largest factor <= sqrt
      Pz = nranks / Py;
    } else {
      Py = 1;
      Pz = nranks;
    }
  } else {
    if (Py == 0) Py = nranks / Pz;
    if (Pz == 0) Pz = nranks / Py;
  }
  CGB_REQUIRE(Py >= 1 && Pz >= 1 && Py * Pz == nranks, "rank grid %dx%d does not match nranks=%d", Py, Pz, nranks);
  CGB_REQUIRE(Py <= st.gy && Pz <= st.gz, "rank grid %dx%d larger than the grid %dx%d", Py, Pz, st.gy, st.gz);

  cgb_partition_t p{};
  p.nranks = nranks;
  p.rank = rank;
  p.Py = Py;
  p.Pz = Pz;
  p.ry = rank % Py;
  p.rz = rank / Py;
  p.nx = st.gx;
  block_range(st.gy, Py, p.ry, &p.y0, &p.ny);
  block_range(st.gz, Pz, p.rz, &p.z0, &p.nz);
  for (int d = 0; d < CGB_NDIR; ++d) {
    p.nbr[d] = -1;
    int dy = dir_dy(d), dz = dir_dz(d);
    if (dy == 0 && dz == 0) continue;
    if (dy != 0 && dz != 0 && st.kind != CGB_KIND_STENCIL3D27) continue;  // edges only for 27-pt
    int qy = p.ry + dy, qz = p.rz + dz;
    if (qy < 0 || qy >= Py || qz < 0 || qz >= Pz) continue;
    p.nbr[d] = qy + Py * qz;
  }
  *out = p;
  return CGB_OK;
}

int make_layout(const cgb_stencil_t& st, const cgb_partition_t& part, cgb_layout_t* out) {
  CGB_REQUIRE(out != nullptr, "make_layout: out is NULL");
  cgb_layout_t L{};
  L.nx = part.nx;
  L.ny = part.ny;
  L.nz = part.nz;
  L.gy = kind_has_y_taps(st.kind) ? 1 : 0;
  L.gz = kind_has_z_taps(st.kind) ? 1 : 0;
  L.pitch = (L.nx + 15) / 16 * 16;
  L.py = L.ny + 2 * L.gy;
  L.pz = L.nz + 2 * L.gz;
  int64_t e = (int64_t)L.pitch * L.py * L.pz;
  L.vec_elems = (e + 31) / 32 * 32;
  L.owned = (int64_t)L.nx * L.ny * L.nz;
  *out = L;
  return CGB_OK;
}

}  // namespace cgb

extern "C" int cgb_partition_make(const cgb_stencil_t* st, int nranks, int rank, int decomp, int Py, int Pz,
                                  cgb_partition_t* out) {
  CGB_REQUIRE(st != nullptr, "cgb_partition_make: stencil is NULL");
  return cgb::make_partition(*st, nranks, rank, decomp, Py, Pz, out);
}

extern "C" int cgb_layout_make(const cgb_stencil_t* st, const cgb_partition_t* part, cgb_layout_t* out) {
  CGB_REQUIRE(st && part, "cgb_layout_make: NULL argument");
  return cgb::make_layout(*st, *part, out);
}
