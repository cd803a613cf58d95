// Topology layer: rank grid, owned index ranges, neighbour table and the padded local layout.
// The reference runs everything on device 0 with one monolithic CSR (CUDACG.cu:87, 213-216);
// this is the domain-decomposition replacement demanded by BASELINE.json ("1D/2D row-block").
#pragma once
#include <cstdint>

#include "../problem/stencil.hpp"
#include "cgb200.h"

namespace cgb {

CGB_HD int dir_index(int dy, int dz) { return (dy + 1) + 3 * (dz + 1); }
CGB_HD int dir_dy(int d) { return d % 3 - 1; }
CGB_HD int dir_dz(int d) { return d / 3 - 1; }
CGB_HD int dir_opposite(int d) { return 8 - d; }

// Block distribution of n items over P parts.
CGB_HD void block_range(int n, int P, int i, int* start, int* count) {
  int base = n / P, rem = n % P;
  *start = i * base + (i < rem ? i : rem);
  *count = base + (i < rem ? 1 : 0);
}

// Linear index of owned cell (i,j,k) in the padded box.
CGB_HD int64_t lay_index(const cgb_layout_t& L, int i, int j, int k) {
  return (int64_t)i + (int64_t)L.pitch * ((int64_t)(j + L.gy) + (int64_t)L.py * (int64_t)(k + L.gz));
}

int make_partition(const cgb_stencil_t& st, int nranks, int rank, int decomp, int Py, int Pz,
                   cgb_partition_t* out);
int make_layout(const cgb_stencil_t& st, const cgb_partition_t& part, cgb_layout_t* out);

}  // namespace cgb
