// k_spmv_tma: matrix-free stencil SpMV fed by TMA (cp.async.bulk.tensor) through an mbarrier ring.
//
// The L1-path kernel (cg_kernels.cu) is latency bound: each warp has one plane of loads in flight
// (ncu: 31 % warps active, 44 % of copy bandwidth). Here a producer warp streams (TX+4)x(TY+2)
// plane tiles of the input vector into a ring of shared-memory stages with 3-D tensor-map loads,
// far ahead of the eight consumer warps, so memory-level parallelism no longer depends on
// occupancy. Out-of-bounds box coordinates are zero-filled by the TMA unit, which gives the
// homogeneous Dirichlet condition on the x faces for free (no predicates in the inner loop); y/z
// faces are the ghost layers of the padded layout.
//
// Work is the linearised space (tile column, plane); every CTA owns one contiguous, equally sized
// range of it (static => deterministic partial sums), cut into segments at column changes. The
// ring never drains between segments. Segments that touch a neighbour's ghost layer are issued
// last, after the producer thread has acquired the halo flags (then fence.proxy.async, because
// the ghost data was written through the generic proxy by a peer GPU and is read by the async
// proxy here).
//
// Replaces cusparseSpMV + cublasDdot of the reference (CUDACG.cu:288-304) for stencil problems.
#include <cuda.h>

#include <mutex>

#include "../common/status.hpp"
#include "cg_kernels.cuh"

namespace cgb {

namespace {

constexpr int TX = 64;                      // cells per tile in x (32 lanes x 2)
constexpr int TY = 16;                      // rows per tile (8 warps x 2 adjacent rows)
constexpr int BOXX = TX + 4;                // x0-2 .. x0+TX+1 : keeps cell pairs 16-byte aligned
constexpr int BOXY = TY + 2;
constexpr int kStageBytes = ((BOXX * BOXY * 8 + 127) / 128) * 128;
constexpr int kStages = 8;
constexpr int kConsumers = 256;
constexpr int kTmaThreads = kConsumers + 32;  // + one producer warp
constexpr size_t kSmemBytes = (size_t)kStages * kStageBytes + 2 * kStages * sizeof(unsigned long long) + 128;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, int x, int y, int z, unsigned bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
      "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar)
      : "memory");
}

struct Seg {
  int col, k0, k1;   // tile column, planes [k0,k1)
  bool ghost;        // touches a neighbour's ghost layer
};

// Segments of CTA `c`: its share [lo,hi) of the (column, plane) space, cut at column changes,
// ordered so that ghost-touching segments come last. At most kMaxSeg per CTA by construction
// (the host sizes the grid so that a share never spans more than kMaxSeg-1 whole columns).
constexpr int kMaxSeg = 8;
__device__ __forceinline__ int make_segments(const RankDev& P, int ncols, int nz, Seg* out) {
  const long long T = (long long)ncols * nz;
  const long long lo = T * blockIdx.x / gridDim.x, hi = T * (blockIdx.x + 1) / gridDim.x;
  const bool nb_zlo = P.nbr[dir_index(0, -1)] >= 0, nb_zhi = P.nbr[dir_index(0, 1)] >= 0;
  const bool nb_ylo = P.nbr[dir_index(-1, 0)] >= 0, nb_yhi = P.nbr[dir_index(1, 0)] >= 0;
  Seg tmp[kMaxSeg];
  int n = 0;
  long long pos = lo;
  while (pos < hi && n < kMaxSeg) {
    int col = (int)(pos / nz), k0 = (int)(pos - (long long)col * nz);
    long long end = (long long)(col + 1) * nz;
    if (end > hi) end = hi;
    int k1 = (int)(end - (long long)col * nz);
    int ty = col / P.ntx;
    bool g = (k0 == 0 && nb_zlo) || (k1 == nz && nb_zhi) || (ty == 0 && nb_ylo) || (ty == P.nty - 1 && nb_yhi);
    tmp[n++] = Seg{col, k0, k1, g};
    pos = end;
  }
  int m = 0;
  for (int i = 0; i < n; ++i)
    if (!tmp[i].ghost) out[m++] = tmp[i];
  for (int i = 0; i < n; ++i)
    if (tmp[i].ghost) out[m++] = tmp[i];
  return n;
}

struct PS2 {  // plane sums of one cell pair
  double s0[2], s1[2], s2[2];
};

// Plane sums for the thread's two adjacent rows (ra = 2*wy, rb = ra+1) from one smem plane tile.
template <int KIND>
__device__ __forceinline__ void plane_from_smem(const double* tile, int wy, int lane, PS2& A, PS2& B) {
  const double* r0 = tile + (2 * wy) * BOXX + 2 * lane + 2;  // row above ra (tile row index = cell row + 1)
  const double* r1 = r0 + BOXX;
  const double* r2 = r1 + BOXX;
  const double* r3 = r2 + BOXX;
  const double2 c0 = *reinterpret_cast<const double2*>(r0);
  const double2 c1 = *reinterpret_cast<const double2*>(r1);
  const double2 c2 = *reinterpret_cast<const double2*>(r2);
  const double2 c3 = *reinterpret_cast<const double2*>(r3);
  const double l1 = r1[-1], q1 = r1[2], l2 = r2[-1], q2 = r2[2];
  A.s0[0] = c1.x; A.s0[1] = c1.y;
  B.s0[0] = c2.x; B.s0[1] = c2.y;
  A.s1[0] = (l1 + c1.y) + (c0.x + c2.x);
  A.s1[1] = (c1.x + q1) + (c0.y + c2.y);
  B.s1[0] = (l2 + c2.y) + (c1.x + c3.x);
  B.s1[1] = (c2.x + q2) + (c1.y + c3.y);
  if (KIND == CGB_KIND_STENCIL3D27) {
    const double l0 = r0[-1], q0 = r0[2], l3 = r3[-1], q3 = r3[2];
    A.s2[0] = (l0 + c0.y) + (l2 + c2.y);
    A.s2[1] = (c0.x + q0) + (c2.x + q2);
    B.s2[0] = (l1 + c1.y) + (l3 + c3.y);
    B.s2[1] = (c1.x + q1) + (c3.x + q3);
  }
}

template <int KIND>
__device__ __forceinline__ void emit(const RankDev& P, double* __restrict__ out, const PS2& pv, const PS2& cu,
                                     const PS2& nx, int i, int j, int k, double& acc) {
  const cgb_layout_t& L = P.lay;
  if (i >= L.nx || j >= L.ny) return;
  const double cc = P.st.c_center, cf = P.st.c_face;
  double a0 = cc * cu.s0[0] + cf * (cu.s1[0] + (pv.s0[0] + nx.s0[0]));
  double a1 = cc * cu.s0[1] + cf * (cu.s1[1] + (pv.s0[1] + nx.s0[1]));
  if (KIND == CGB_KIND_STENCIL3D27) {
    const double ce = P.st.c_edge, cn = P.st.c_corner;
    a0 += ce * (cu.s2[0] + (pv.s1[0] + nx.s1[0])) + cn * (pv.s2[0] + nx.s2[0]);
    a1 += ce * (cu.s2[1] + (pv.s1[1] + nx.s1[1])) + cn * (pv.s2[1] + nx.s2[1]);
  }
  const int64_t idx = lay_index(L, i, j, k);
  if (i + 1 < L.nx) {
    *reinterpret_cast<double2*>(out + idx) = make_double2(a0, a1);
    acc += cu.s0[0] * a0 + cu.s0[1] * a1;
  } else {
    out[idx] = a0;
    acc += cu.s0[0] * a0;
  }
}

template <int KIND>
__global__ void __launch_bounds__(kTmaThreads, 2) k_spmv_tma(const __grid_constant__ RankDev P,
                                                              const __grid_constant__ CUtensorMap tmap,
                                                              double* __restrict__ out, int mode) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ double red_smem[64];
  Scalars& S = P.ctrl->s;
  if (S.done != 0) return;
  unsigned char* stage_base = smem_raw;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)kStages * kStageBytes);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + kStages);
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, kConsumers / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const cgb_layout_t& L = P.lay;
  const int ncols = P.ntx * P.nty, nz = L.nz;
  Seg segs[kMaxSeg];
  const int nseg = make_segments(P, ncols, nz, segs);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double acc = 0.0;

  if (warp == kConsumers / 32) {
    // ===== producer warp: one elected lane issues every TMA load =====
    if (lane == 0) {
      unsigned n = 0;
      bool waited = false;
      for (int s = 0; s < nseg; ++s) {
        const Seg sg = segs[s];
        if (sg.ghost && !waited) {
          // ghost layers are needed from here on: acquire every neighbour's flag once
          for (int d = 0; d < CGB_NDIR; ++d)
            if (d != 4 && P.nbr[d] >= 0 && !spin_until_ge(&P.ctrl->halo_flag[d], S.halo_seq, S.timeout_cycles))
              S.done = CGB_ERR_TIMEOUT;
          asm volatile("fence.proxy.async;" ::: "memory");
          waited = true;
        }
        const int x0 = (sg.col % P.ntx) * TX - 2;
        const int y0 = (sg.col / P.ntx) * TY;  // padded row of cell row j0-1
        for (int q = sg.k0 - 1; q <= sg.k1; ++q, ++n) {
          const unsigned st = n % kStages, ph = (n / kStages) & 1u;
          mbar_wait(empty0 + 8 * st, ph ^ 1u);
          mbar_expect_tx(full0 + 8 * st, BOXX * BOXY * 8);
          tma_load_3d(smem_u32(stage_base + (size_t)st * kStageBytes), &tmap, x0, y0, q + 1, full0 + 8 * st);
        }
      }
    }
  } else {
    // ===== consumer warps: two adjacent rows x two adjacent cells per thread =====
    unsigned n = 0;
    for (int s = 0; s < nseg; ++s) {
      const Seg sg = segs[s];
      const int i = (sg.col % P.ntx) * TX + 2 * lane;
      const int ja = (sg.col / P.ntx) * TY + 2 * warp;
      PS2 pa, pb, ca, cb, na, nb;
      for (int q = sg.k0 - 1; q <= sg.k1; ++q, ++n) {
        const unsigned st = n % kStages, ph = (n / kStages) & 1u;
        mbar_wait(full0 + 8 * st, ph);
        plane_from_smem<KIND>(reinterpret_cast<const double*>(stage_base + (size_t)st * kStageBytes), warp, lane, na, nb);
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * st);
        if (q >= sg.k0 + 1) {
          emit<KIND>(P, out, pa, ca, na, i, ja, q - 1, acc);
          emit<KIND>(P, out, pb, cb, nb, i, ja + 1, q - 1, acc);
        }
        pa = ca; pb = cb;
        ca = na; cb = nb;
      }
    }
  }
  if (mode == kSpmvPlain) return;
  double a = acc, z = 0.0;
  if (grid_sum2_last_t<kTmaThreads>(a, z, P.partials, &S.ticket_red, red_smem)) {
    allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0 && mode == kSpmvAlpha) {
      S.pAp = a;
      S.alpha = S.rho / a;
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn get_encode() {
  static EncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeFn)p;
  });
  return fn;
}

}  // namespace

// Grid for a layout: 2 CTAs per SM, every CTA one contiguous share that spans at most kMaxSeg
// columns. Returns -1 if the layout cannot be covered (falls back to the L1-path kernel).
static long long tma_grid(const RankDev& P, int nsm) {
  const long long ncols = (long long)((P.lay.nx + TX - 1) / TX) * ((P.lay.ny + TY - 1) / TY);
  const long long T = ncols * P.lay.nz;
  long long grid = 2LL * nsm;
  if (grid > T) grid = T;
  const long long min_grid = (ncols + (kMaxSeg - 2)) / (kMaxSeg - 1);
  if (grid < min_grid) grid = min_grid;
  if (grid < 1) grid = 1;
  return grid > kMaxGrid ? -1 : grid;
}

bool spmv_tma_supported(const RankDev& P) { return kind_is_3d(P.st.kind) && tma_grid(P, 148) > 0; }

int spmv_tma_make_map(const cgb_layout_t& L, const double* vec, void* map_out) {
  EncodeFn enc = get_encode();
  if (!enc) return fail(CGB_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[3] = {(cuuint64_t)L.nx, (cuuint64_t)L.py, (cuuint64_t)L.pz};
  cuuint64_t strides[2] = {(cuuint64_t)L.pitch * 8, (cuuint64_t)L.pitch * L.py * 8};
  cuuint32_t box[3] = {BOXX, BOXY, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = enc((CUtensorMap*)map_out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(vec), dims, strides,
                   box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(CGB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return CGB_OK;
}

void spmv_tma_tiling(RankDev& P) {
  P.ntx = (P.lay.nx + TX - 1) / TX;
  P.nty = (P.lay.ny + TY - 1) / TY;
}

cudaError_t launch_spmv_tma(const RankDev& Pin, int device, const void* map, double* out, int mode, cudaStream_t s) {
  static bool attr_set[64] = {};
  if (device >= 0 && device < 64 && !attr_set[device]) {
    cudaError_t e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON3D7>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kSmemBytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_STENCIL3D27>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)kSmemBytes);
    if (e != cudaSuccess) return e;
    attr_set[device] = true;
  }
  RankDev P = Pin;
  spmv_tma_tiling(P);
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device);
  const long long grid = tma_grid(P, nsm);
  if (grid < 0) return cudaErrorInvalidConfiguration;
  const CUtensorMap& tm = *reinterpret_cast<const CUtensorMap*>(map);
  if (P.st.kind == CGB_KIND_POISSON3D7)
    k_spmv_tma<CGB_KIND_POISSON3D7><<<(int)grid, kTmaThreads, kSmemBytes, s>>>(P, tm, out, mode);
  else if (P.st.kind == CGB_KIND_STENCIL3D27)
    k_spmv_tma<CGB_KIND_STENCIL3D27><<<(int)grid, kTmaThreads, kSmemBytes, s>>>(P, tm, out, mode);
  else
    return cudaErrorInvalidValue;
  return cudaGetLastError();
}

}  // namespace cgb
