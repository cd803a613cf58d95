// k_spmv_tma: matrix-free stencil SpMV fed by TMA (cp.async.bulk.tensor) through an mbarrier ring.
//
// Design (second version; profiles/r1b_spmv_tma_coltile_256.md records why the first one, 64x16
// column tiles spread over unrelated planes, re-read its halos from DRAM 1.69x):
//
//  * A CTA owns a BAND of RB full-width rows and sweeps it along z. All CTAs of a z-chunk start
//    together and move at the same pace, so at any moment the chip streams one or two whole
//    planes: DRAM sees long contiguous runs and a band's two halo rows are L2 hits (the
//    neighbouring band's CTA fetched them a moment ago).
//  * A producer warp keeps a ring of kStages plane slabs (RB+2 rows x nx) in flight with 3-D
//    tensor-map loads; 16 consumer warps (2 cells x 2 adjacent rows per thread) turn each slab
//    into per-plane partial sums held in registers (dev: S0/S1/S2, see cg_kernels.cu) and emit
//    one output plane per slab. Memory-level parallelism no longer depends on occupancy.
//  * x faces: the band spans the whole row, so x-1 / x+nx are the Dirichlet zeros (predicated
//    for two lanes); y/z faces are ghost layers of the padded layout; rows past the padded box
//    are zero-filled by the TMA unit.
//  * Halo overlap: a chunk that touches only ONE z ghost plane sweeps towards it, so that plane
//    is the last slab it loads; the producer thread acquires the neighbours' flags right before
//    that load (then fence.proxy.async: the ghost data was written through the generic proxy by a
//    peer GPU and is read through the async proxy here). Interior planes are therefore computed
//    while the halo is still in flight.
//
//  * Rows wider than 512 cells are cut into 512-cell x tiles; a tile's two x-halo columns arrive
//    in two extra 2-column boxes per slab (zero-filled by the TMA unit outside the domain).
//
// Replaces cusparseSpMV + cublasDdot of the reference (CUDACG.cu:288-304) for 3D stencils; 2D
// problems and general CSR use the L1-path kernels.
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <utility>

#include "../common/status.hpp"
#include "cg_kernels.cuh"

namespace cgb {

namespace {

// cells per TMA box row: 128 for 3D problems; 256 for 2D grids, whose one-row slabs would otherwise
// need five times as many TMA instructions per output as the 3D bands
template <int KIND> struct BoxW { static constexpr int v = KIND == CGB_KIND_POISSON2D5 ? 256 : 128; };
inline int box_w(int kind) { return kind == CGB_KIND_POISSON2D5 ? 256 : 128; }
constexpr int kNC = 512;                 // consumer threads
constexpr int kNT = kNC + 32;            // + one producer warp
constexpr int kMaxStages = 8;
constexpr int kTileX3 = 512;            // cells per x tile when a 3D row is wider than one tile (2 rows/thread)
constexpr int kTileX2 = 1024;           // ... and for 2D grids (1 row/thread)

struct BandGeo {
  int LX, LY;        // consumer thread grid: LX pair-columns x LY row-pairs
  int RB;            // rows per band = 2*LY (1 for 2D grids: one row per plane)
  int HY;            // halo rows above/below a band in a p slab: 1, or 0 for 2D grids (no y taps)
  int prow;          // rows of a p slab = RB + 2*HY
  int NBX;           // 128-wide TMA boxes per slab row
  int ntx, nby;      // x tiles (of 2*LX cells) and row bands; nbands = ntx*nby
  int h_off;         // x-tiled layouts: byte offset of the two 2-column halo boxes inside a p slab
  int nbands;
  int zs, zchunk;    // z split
  int nitems;
  int p_bytes;       // p slab: NBX boxes of (RB+2) rows
  int xr_bytes;      // x (or r) slab of the update mode: NBX boxes of RB rows
  int stage_bytes;   // p_bytes (+ 2*xr_bytes in update mode)
  int nstages;
  int nhalo;         // merged kernel: extra CTAs (on otherwise idle SMs) that only push the boundary planes
  int nctas;         // CTAs that sweep bands (= grid - nhalo); items are dealt round-robin over these
  int wide;          // 2D grids whose rows are a multiple of 256 cells: the main maps describe a row as (256, nx/256)
                     // so that ONE TMA operation brings the whole x tile of a row (NBX boxes) instead of NBX
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  // "memory": without it the compiler is free to sink the init below the __syncthreads() that publishes it (seen
  // with the persistent kernel: the TMA lane armed barriers that thread 0 had not initialised yet)
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Wait for a phase of an mbarrier. Bounded like every other wait of this library: if the phase does not complete
// within the peer-wait timeout (a lost TMA transaction, a dead producer), the solve is failed with
// CGB_ERR_TIMEOUT and the wait - and every later one, which sees the failed state after its first unsuccessful
// try - falls through, so the kernel drains instead of hanging. `site` lands in Scalars::dbg for the post-mortem.
// Hot path: one asm block that retries try_wait up to 4096 times and costs no registers beyond its operands
// (the 27-point loops run at the register limit: a C-level loop with a clock and a counter there cost them 25 %).
__device__ __forceinline__ bool mbar_spin(unsigned bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .u32 c;\n"
      "mov.u32 c, 4096;\n"
      "SPIN_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "@p bra HIT_%=;\n"
      "sub.u32 c, c, 1;\n"
      "setp.ne.u32 p, c, 0;\n"
      "@p bra SPIN_%=;\n"
      "mov.u32 %0, 0;\n"
      "bra END_%=;\n"
      "HIT_%=:\n"
      "mov.u32 %0, 1;\n"
      "END_%=:\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity, Scalars& S, int site) {
  if (mbar_spin(bar, parity)) return;
  // cold: the barrier is late by thousands of polls
  const long long t0 = clock64();
  do {
    if (*(volatile int*)&S.done == CGB_ERR_TIMEOUT) return;
    if (clock64() - t0 > 4 * S.timeout_cycles) {
      if (atomicCAS(&S.dbg[0], 0, site) == 0) {  // the first wait that gave up tells the story
        S.dbg[1] = (int)blockIdx.x;
        S.dbg[2] = (int)threadIdx.x;
        S.dbg[3] = (int)bar;
        S.dbg[4] = (int)parity;
        S.dbg[5] = (int)*(volatile u64*)&S.gbar;
        S.dbg[6] = (int)*(volatile u64*)&S.sbar;
      }
      S.done = CGB_ERR_TIMEOUT;
      return;
    }
  } while (!mbar_spin(bar, parity));
}
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, int x, int y, int z, unsigned bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
      "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar)
      : "memory");
}

// A 2D grid is one row per plane: the second row of a thread's 2x2 patch does not exist.
template <int KIND> constexpr bool kTwoRows = KIND != CGB_KIND_POISSON2D5;

// Next slot of the stage ring (st) and its mbarrier phase parity (ph).
__device__ __forceinline__ void ring_next(unsigned& st, unsigned& ph, unsigned nstages) {
  if (++st == nstages) {
    st = 0;
    ph ^= 1u;
  }
}

struct PS2 {  // plane sums of one cell pair
  double s0[2], s1[2], s2[2];
};

// Byte offsets (within a slab) of a thread's centre pair / left / right neighbour in slab row 0.
struct Taps {
  int c, l, r;        // l,r < 0: Dirichlet zero
  int ls, rs;         // bytes between slab rows for the l / r tap (1024 in a main box, 16 in a halo box)
};

// Offsets of a consumer thread's taps. il = x offset inside the tile, i = global x.
template <int BW>
__device__ __forceinline__ Taps make_taps(const BandGeo& G, int il, int i, int nx) {
  constexpr int RS = BW * 8;
  const int boxb = G.prow * RS;
  const int b = il / BW, col = il % BW;
  Taps tp;
  tp.c = b * boxb + col * 8;
  tp.ls = tp.rs = RS;
  if (il == 0) {
    tp.l = G.ntx > 1 ? G.h_off + 8 : -1;  // column x0-1 = element 1 of the left halo box (zero-filled at x = -1)
    if (G.ntx > 1) tp.ls = 16;
  } else {
    tp.l = col == 0 ? (b - 1) * boxb + (BW - 1) * 8 : tp.c - 8;
  }
  if (G.ntx > 1 && il + 2 == 2 * G.LX) {
    tp.r = G.h_off + 128;                 // column x0+tile = element 0 of the right halo box
    tp.rs = 16;
  } else if (G.ntx == 1 && i + 2 >= nx) {
    tp.r = -1;
  } else {
    tp.r = col + 2 == BW ? (b + 1) * boxb : tp.c + 16;
  }
  // Opaque to the optimiser: otherwise ptxas may re-derive ls/rs (selects on ntx, il) at every tap of every plane
  // instead of keeping five integers, which costs ~100 instructions per plane in the 27-point loops.
  asm volatile("" : "+r"(tp.c), "+r"(tp.l), "+r"(tp.r), "+r"(tp.ls), "+r"(tp.rs));
  return tp;
}

// Plane sums for the thread's two adjacent rows from one slab. Slab row t holds band row t-1.
template <int KIND>
__device__ __forceinline__ void plane_from_slab(const unsigned char* slab, const Taps& tp, int ra, PS2& A, PS2& B) {
  constexpr int RS = BoxW<KIND>::v * 8;  // slab row stride in bytes
  const unsigned char* b0 = slab + ra * RS;
  auto ld2 = [&](int row) { return *reinterpret_cast<const double2*>(b0 + row * RS + tp.c); };
  auto ldl = [&](int row) { return tp.l >= 0 ? *reinterpret_cast<const double*>(slab + tp.l + (ra + row) * tp.ls) : 0.0; };
  auto ldr = [&](int row) { return tp.r >= 0 ? *reinterpret_cast<const double*>(slab + tp.r + (ra + row) * tp.rs) : 0.0; };
  if (KIND == CGB_KIND_POISSON2D5) {  // one row per plane, no halo rows: slab row 0 is the thread's row
    const double2 c = ld2(0);
    const double l = ldl(0), q = ldr(0);
    A.s0[0] = c.x; A.s0[1] = c.y;
    A.s1[0] = l + c.y;
    A.s1[1] = c.x + q;
    B = A;
    return;
  }
  const double2 c0 = ld2(0), c1 = ld2(1), c2 = ld2(2), c3 = ld2(3);
  const double l1 = ldl(1), q1 = ldr(1), l2 = ldl(2), q2 = ldr(2);
  A.s0[0] = c1.x; A.s0[1] = c1.y;
  B.s0[0] = c2.x; B.s0[1] = c2.y;
  A.s1[0] = (l1 + c1.y) + (c0.x + c2.x);
  A.s1[1] = (c1.x + q1) + (c0.y + c2.y);
  B.s1[0] = (l2 + c2.y) + (c1.x + c3.x);
  B.s1[1] = (c2.x + q2) + (c1.y + c3.y);
  if (KIND == CGB_KIND_STENCIL3D27) {
    const double l0 = ldl(0), q0 = ldr(0), l3 = ldl(3), q3 = ldr(3);
    A.s2[0] = (l0 + c0.y) + (l2 + c2.y);
    A.s2[1] = (c0.x + q0) + (c2.x + q2);
    B.s2[0] = (l1 + c1.y) + (l3 + c3.y);
    B.s2[1] = (c1.x + q1) + (c3.x + q3);
  }
}

// Merged variant: plane sums of p_new = r + beta*p_old evaluated on the fly from the r slab and
// the p_old slab (for a neighbour's ghost plane the "r slab" already holds p_new). Everything a
// thread needs — its 2x2 cells, the rows above/below and the x neighbours — is recomputed in
// registers, so no p_new slab has to be written back to shared memory and no CTA barrier is
// needed between the update and the stencil. Also returns the thread's own p_new pairs.
template <int KIND>
__device__ __forceinline__ void plane_from_two(const unsigned char* rslab, const unsigned char* pslab, const Taps& tp,
                                               int ra, double beta, bool ghost, PS2& A, PS2& B, double2& pna,
                                               double2& pnb) {
  constexpr int RS = BoxW<KIND>::v * 8;
  const unsigned char* r0 = rslab + ra * RS;
  const unsigned char* p0 = pslab + ra * RS;
  const double bt = ghost ? 0.0 : beta;  // p_old slab holds finite data (zeros or a previous p), so 0*p is exact
  auto pn2 = [&](int row) {
    const double2 vr = *reinterpret_cast<const double2*>(r0 + row * RS + tp.c);
    const double2 vp = *reinterpret_cast<const double2*>(p0 + row * RS + tp.c);
    return make_double2(vr.x + bt * vp.x, vr.y + bt * vp.y);
  };
  auto pnl = [&](int row) {
    const int o = tp.l + (ra + row) * tp.ls;
    return tp.l >= 0 ? *reinterpret_cast<const double*>(rslab + o) + bt * *reinterpret_cast<const double*>(pslab + o) : 0.0;
  };
  auto pnr = [&](int row) {
    const int o = tp.r + (ra + row) * tp.rs;
    return tp.r >= 0 ? *reinterpret_cast<const double*>(rslab + o) + bt * *reinterpret_cast<const double*>(pslab + o) : 0.0;
  };
  if (KIND == CGB_KIND_POISSON2D5) {
    const double2 c = pn2(0);
    const double l = pnl(0), q = pnr(0);
    pna = pnb = c;
    A.s0[0] = c.x; A.s0[1] = c.y;
    A.s1[0] = l + c.y;
    A.s1[1] = c.x + q;
    B = A;
    return;
  }
  const double2 c0 = pn2(0), c1 = pn2(1), c2 = pn2(2), c3 = pn2(3);
  const double l1 = pnl(1), q1 = pnr(1), l2 = pnl(2), q2 = pnr(2);
  pna = c1;
  pnb = c2;
  A.s0[0] = c1.x; A.s0[1] = c1.y;
  B.s0[0] = c2.x; B.s0[1] = c2.y;
  A.s1[0] = (l1 + c1.y) + (c0.x + c2.x);
  A.s1[1] = (c1.x + q1) + (c0.y + c2.y);
  B.s1[0] = (l2 + c2.y) + (c1.x + c3.x);
  B.s1[1] = (c2.x + q2) + (c1.y + c3.y);
  if (KIND == CGB_KIND_STENCIL3D27) {
    const double l0 = pnl(0), q0 = pnr(0), l3 = pnl(3), q3 = pnr(3);
    A.s2[0] = (l0 + c0.y) + (l2 + c2.y);
    A.s2[1] = (c0.x + q0) + (c2.x + q2);
    B.s2[0] = (l1 + c1.y) + (l3 + c3.y);
    B.s2[1] = (c1.x + q1) + (c3.x + q3);
  }
}

// Partial-sum formulation of the z sweep. With the in-plane sums S0 (centre), S1 (4 faces), S2 (4 diagonals)
// of plane k, define
//   C_k = cc*S0_k + cf*S1_k + ce*S2_k     plane k's own contribution to output plane k
//   T_k = cf*S0_k + ce*S1_k + cn*S2_k     plane k's contribution to output planes k-1 and k+1
// so that Ap_k = C_k + T_{k-1} + T_{k+1}. A thread therefore carries three values per cell along the sweep -
// part = C_k + T_{k-1}, T_k and S0_k - instead of the six (S0,S1,S2 of two planes) of the direct form; for the
// 5/7-point stencils T_k = cf*S0_k is not even stored. This is what keeps the 27-point kernels and the
// persistent kernel inside the 96 registers a 544-thread CTA can have, without spills in the inner loops.
struct Run2 {
  double part[2], t[2], s0[2];
};

template <int KIND>
__device__ __forceinline__ void plane_ct(const RankDev& P, const PS2& n, double c[2], double t[2]) {
  const double cc = P.st.c_center, cf = P.st.c_face;
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    if (KIND == CGB_KIND_STENCIL3D27) {
      const double ce = P.st.c_edge, cn = P.st.c_corner;
      c[e] = cc * n.s0[e] + (cf * n.s1[e] + ce * n.s2[e]);
      t[e] = cf * n.s0[e] + (ce * n.s1[e] + cn * n.s2[e]);
    } else {
      c[e] = cc * n.s0[e] + cf * n.s1[e];
      t[e] = cf * n.s0[e];
    }
  }
}

// The newly arrived plane becomes the current one.
template <int KIND>
__device__ __forceinline__ void run_advance(const RankDev& P, Run2& A, const PS2& n, const double c[2], const double t[2]) {
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    if (KIND == CGB_KIND_STENCIL3D27) {
      A.part[e] = c[e] + A.t[e];
      A.t[e] = t[e];
    } else {
      A.part[e] = c[e] + P.st.c_face * A.s0[e];
    }
    A.s0[e] = n.s0[e];
  }
}

// Output of the current plane once its upper neighbour's T has arrived.
// OUT: 0 store Ap and accumulate p.Ap; 1 accumulate p.Ap only; 2 x += alpha p, r -= alpha Ap,
// accumulate r.r (the x/r pairs of the emitted plane arrive TMA-staged in vx/vr).
template <int KIND, int OUT>
__device__ __forceinline__ void emit(const RankDev& P, double* __restrict__ out, const Run2& A, const double tn[2],
                                     int64_t idx, bool pair, double alpha, double2 vx, double2 vr, double& acc) {
  const double a0 = A.part[0] + tn[0];
  const double a1 = A.part[1] + tn[1];
  if (OUT == 2) {
    vx.x += alpha * A.s0[0];
    vr.x -= alpha * a0;
    if (pair) {
      vx.y += alpha * A.s0[1];
      vr.y -= alpha * a1;
      *reinterpret_cast<double2*>(P.x + idx) = vx;
      *reinterpret_cast<double2*>(P.r + idx) = vr;
      acc += vr.x * vr.x + vr.y * vr.y;
    } else {
      P.x[idx] = vx.x;
      P.r[idx] = vr.x;
      acc += vr.x * vr.x;
    }
    return;
  }
  if (OUT == 0) {
    if (pair) *reinterpret_cast<double2*>(out + idx) = make_double2(a0, a1);
    else out[idx] = a0;
  }
  acc += A.s0[0] * a0 + (pair ? A.s0[1] * a1 : 0.0);
}

struct Item {
  int band, k0, k1, dir;
  int x0, yb;        // first cell of the x tile, row-band index
  bool y_ghost;
};
__device__ __forceinline__ Item make_item(const RankDev& P, const BandGeo& G, int item, bool nb_zlo, bool nb_zhi,
                                          bool nb_ylo, bool nb_yhi) {
  Item it;
  it.band = item % G.nbands;
  const int zc = item / G.nbands;
  it.k0 = zc * G.zchunk;
  it.k1 = min(P.lay.nz, it.k0 + G.zchunk);
  const bool lo = it.k0 == 0 && nb_zlo, hi = it.k1 == P.lay.nz && nb_zhi;
  it.dir = (lo && !hi) ? -1 : 1;  // sweep towards the only ghost plane so that it is loaded last
  it.x0 = (it.band % G.ntx) * 2 * G.LX;
  it.yb = it.band / G.ntx;
  it.y_ghost = (it.yb == 0 && nb_ylo) || (it.yb == G.nby - 1 && nb_yhi);
  return it;
}

__device__ __forceinline__ u64 ld_acquire_gpu(const u64* p) {
  u64 v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(u64* p, u64 v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Persistent kernel: block until the scalars produced by reduction `seq` (alpha, or beta/done/...) have been
// published by the service CTA. One poller per warp; the rest of the warp waits at the __syncwarp.
__device__ __forceinline__ void wait_scalars(const RankDev& P, u64 seq) {
  Scalars& S = P.ctrl->s;
  if ((threadIdx.x & 31) == 0 && ld_acquire_gpu(&S.sbar) < seq) {
    const long long t0 = clock64();
    while (ld_acquire_gpu(&S.sbar) < seq) {
      if (clock64() - t0 > 4 * S.timeout_cycles) {  // the service CTA's own waits are bounded by timeout_cycles
        if (atomicCAS(&S.dbg[0], 0, 10) == 0) {
          S.dbg[1] = (int)blockIdx.x;
          S.dbg[2] = (int)threadIdx.x;
          S.dbg[3] = (int)seq;
          S.dbg[5] = (int)*(volatile u64*)&S.gbar;
          S.dbg[6] = (int)*(volatile u64*)&S.sbar;
        }
        S.done = CGB_ERR_TIMEOUT;
        break;
      }
    }
  }
  __syncwarp();
}

// Programmatic dependent launch (PDL): a kernel launched with the programmatic-stream-serialization attribute may
// start as soon as every CTA of its predecessor has executed launch_dependents (or exited); it must then execute
// pdl_wait() before it touches anything the predecessor wrote. The band kernels trigger at their very first
// instruction and wait after their private set-up (mbarrier init), so the next kernel of the iteration is already
// resident - set up and parked in the wait - on every SM the moment its predecessor's last CTA retires: the launch
// latency and the prologue leave the critical path of the iteration.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// State of one mbarrier ring: smem addresses of the full/empty barrier arrays and the number of slabs this
// thread's role has issued (producer lane) or consumed (consumer threads) so far. A persistent kernel keeps
// one Ring per phase alive across iterations; the per-iteration kernels start one at zero.
struct Ring {
  unsigned full0, empty0;
  unsigned n;
};

// Branch-free variant for the persistent kernel: EVERY thread executes the same straight-line code and a predicate
// selects the one initialising thread, so no warp diverges in front of the __syncthreads() that publishes the
// barriers (see k_cg_persist for why that matters there).
__device__ __forceinline__ void mbar_init_if(unsigned bar, unsigned count, bool doit) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.u32 p, %2, 0;\n"
      "@p mbarrier.init.shared::cta.b64 [%0], %1;\n"
      "}\n" ::"r"(bar),
      "r"(count), "r"((unsigned)doit)
      : "memory");
}
__device__ __forceinline__ void ring_init_uniform(unsigned long long* bars, int nstages) {
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + kMaxStages);
#pragma unroll
  for (int s = 0; s < kMaxStages; ++s) {
    const bool mine = threadIdx.x == 0 && s < nstages;
    mbar_init_if(full0 + 8 * s, 1, mine);
    mbar_init_if(empty0 + 8 * s, kNC / 32, mine);
  }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void ring_init(Ring& R, unsigned long long* bars, int nstages) {
  R.full0 = smem_u32(bars);
  R.empty0 = smem_u32(bars + kMaxStages);
  R.n = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < nstages; ++s) {
      mbar_init(R.full0 + 8 * s, 1);
      mbar_init(R.empty0 + 8 * s, kNC / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
}

// One sweep of the band SpMV over this CTA's items (all threads of the CTA call it; warp kNC/32 is the TMA
// producer, the others consume). acc receives this thread's share of p.Ap (OUT 0/1) or r.r (OUT 2).
// halo_version: the halo version the ghost layers of the input must have reached before they are loaded.
template <int KIND, int OUT>
__device__ __forceinline__ void phase_spmv(const RankDev& P, const BandGeo& G, const CUtensorMap* tmap,
                                           const CUtensorMap* tmap_h, const CUtensorMap* tmx, const CUtensorMap* tmr,
                                           double* __restrict__ out, double alpha, u64 halo_version,
                                           unsigned char* smem_raw, Ring& R, double& acc, bool lazy = false,
                                           u64 lazy_seq = 0) {
  // lazy (persistent kernel, OUT == 2): alpha is not known yet when the sweep starts - it is the result
  // of reduction lazy_seq, still crossing NVLink. The TMA lane does not need it at all and the consumers only
  // when they emit their first plane, two slabs into the sweep; they pick it up there.
  constexpr int BW = BoxW<KIND>::v;
  Scalars& S = P.ctrl->s;
  // Opaque copies of the thread/block index: in the persistent kernel this keeps everything derived from them
  // (tap offsets, row indices) local to the phase instead of hoisted out of the iteration loop and held in
  // registers (or spilled) across the other phase.
  unsigned tid = threadIdx.x, bid = blockIdx.x;
  asm volatile("" : "+r"(tid), "+r"(bid));
  const unsigned full0 = R.full0, empty0 = R.empty0;
  const unsigned kStages = (unsigned)G.nstages;
  const cgb_layout_t& L = P.lay;
  const bool nb_zlo = P.nbr[dir_index(0, -1)] >= 0, nb_zhi = P.nbr[dir_index(0, 1)] >= 0;
  const bool nb_ylo = P.nbr[dir_index(-1, 0)] >= 0, nb_yhi = P.nbr[dir_index(1, 0)] >= 0;
  const int warp = tid >> 5, lane = tid & 31;
  unsigned n = R.n;

  if (warp == kNC / 32) {
    // ===== producer warp: one lane issues every TMA load =====
    if (lane == 0) {
      bool waited = false;
      auto wait_flags = [&]() {
        if (waited) return;
        for (int d = 0; d < CGB_NDIR; ++d)
          if (d != 4 && P.nbr[d] >= 0 && !spin_until_ge(&P.ctrl->halo_flag[d], halo_version, S.timeout_cycles))
            S.done = CGB_ERR_TIMEOUT;
        asm volatile("fence.proxy.async;" ::: "memory");
        waited = true;
      };
      // ring position of step n, advanced without a division per step (kStages is a run-time value)
      unsigned st = n % kStages, ph = (n / kStages) & 1u;
      for (int item = bid < (unsigned)G.nctas ? (int)bid : G.nitems; item < G.nitems; item += G.nctas) {
        const Item it = make_item(P, G, item, nb_zlo, nb_zhi, nb_ylo, nb_yhi);
        if (it.y_ghost) wait_flags();
        const int y0 = it.yb * G.RB;  // padded row of band row -1
        const int nsteps = it.k1 - it.k0 + 2;
        const unsigned p_tx = (unsigned)(G.NBX * G.prow * BW * 8 + (G.ntx > 1 ? 2 * G.prow * 16 : 0));
        for (int s = 0; s < nsteps; ++s, ++n, ring_next(st, ph, kStages)) {
          const int q = it.dir > 0 ? it.k0 - 1 + s : it.k1 - s;
          if ((q == -1 && nb_zlo) || (q == L.nz && nb_zhi)) wait_flags();
          mbar_wait(empty0 + 8 * st, ph ^ 1u, S, 1);
          const bool with_xr = OUT == 2 && s >= 2;  // this step emits a plane: bring its x and r rows too
          mbar_expect_tx(full0 + 8 * st, p_tx + (with_xr ? 2u * (unsigned)G.xr_bytes : 0u));
          const unsigned dst = smem_u32(smem_raw + (size_t)st * G.stage_bytes);
          if (G.wide) tma_load_3d(dst, tmap, 0, it.x0 / BW, q + 1, full0 + 8 * st);
          else
            for (int b = 0; b < G.NBX; ++b)
              tma_load_3d(dst + b * G.prow * BW * 8, tmap, it.x0 + b * BW, y0, q + 1, full0 + 8 * st);
          if (G.ntx > 1) {  // the tile's x halo: two 2-column boxes (zero-filled outside the domain)
            tma_load_3d(dst + G.h_off, tmap_h, it.x0 - 2, y0, q + 1, full0 + 8 * st);
            tma_load_3d(dst + G.h_off + 128, tmap_h, it.x0 + 2 * G.LX, y0, q + 1, full0 + 8 * st);
          }
          if (with_xr) {
            const int ko = it.dir > 0 ? it.k0 + s - 2 : it.k1 - s + 1;
            if (G.wide) {
              tma_load_3d(dst + G.p_bytes, tmx, 0, it.x0 / BW, ko + 1, full0 + 8 * st);
              tma_load_3d(dst + G.p_bytes + G.xr_bytes, tmr, 0, it.x0 / BW, ko + 1, full0 + 8 * st);
            } else
              for (int b = 0; b < G.NBX; ++b) {
                tma_load_3d(dst + G.p_bytes + b * G.RB * BW * 8, tmx, it.x0 + b * BW, y0 + G.HY, ko + 1, full0 + 8 * st);
                tma_load_3d(dst + G.p_bytes + G.xr_bytes + b * G.RB * BW * 8, tmr, it.x0 + b * BW, y0 + G.HY, ko + 1,
                            full0 + 8 * st);
              }
          }
        }
      }
    }
  } else {
    // ===== consumer warps =====
    const int lx = tid % G.LX, ly = tid / G.LX;
    unsigned st = n % kStages, ph = (n / kStages) & 1u;
    for (int item = bid < (unsigned)G.nctas ? (int)bid : G.nitems; item < G.nitems; item += G.nctas) {
      const Item it = make_item(P, G, item, nb_zlo, nb_zhi, nb_ylo, nb_yhi);
      // A thread without cells (beyond the row end of the last x tile, or in a row pair the band does not have)
      // shadows thread 0: its loads stay inside the slab and the plane loop needs no branch around them; jend = 0
      // keeps it from storing or accumulating anything.
      const bool active = ly < G.LY && it.x0 + 2 * lx < L.nx;
      const int il = active ? 2 * lx : 0;
      const int ra = active ? 2 * ly : 0;
      const int xr_off = (il / BW) * (G.RB * BW * 8) + (ra * BW + il % BW) * 8;
      const int i = it.x0 + il;
      const Taps tp = make_taps<BW>(G, il, i, L.nx);
      const int ja = it.yb * G.RB + ra;
      const int jend = active ? min(L.ny, (it.yb + 1) * G.RB) : 0;
      const bool va = ja < jend, vb = kTwoRows<KIND> && ja + 1 < jend, pair = i + 1 < L.nx;
      const int nsteps = it.k1 - it.k0 + 2;
      const int64_t ps = (int64_t)L.pitch * L.py * it.dir;           // signed plane stride along the sweep
      // index of (i, ja) on the plane emitted at step 2, then advanced by ps per step
      int64_t idx = lay_index(L, i, ja, it.dir > 0 ? it.k0 : it.k1 - 1);
      Run2 A{}, B{};
      for (int s = 0; s < nsteps; ++s, ++n, ring_next(st, ph, kStages)) {
        mbar_wait(full0 + 8 * st, ph, S, 2);
        const unsigned char* stage = smem_raw + (size_t)st * G.stage_bytes;
        PS2 na, nb;
        plane_from_slab<KIND>(stage, tp, ra, na, nb);
        double ca[2], ta[2], cb[2] = {0, 0}, tb[2] = {0, 0};
        plane_ct<KIND>(P, na, ca, ta);
        if (kTwoRows<KIND>) plane_ct<KIND>(P, nb, cb, tb);
        double2 cxa = make_double2(0, 0), cxb = cxa, cra = cxa, crb = cxa;
        if (OUT == 2 && s >= 2) {  // x and r rows of the plane emitted now (TMA-staged)
          const unsigned char* xs = stage + G.p_bytes + xr_off;
          cxa = *reinterpret_cast<const double2*>(xs);
          cra = *reinterpret_cast<const double2*>(xs + G.xr_bytes);
          if (KIND != CGB_KIND_POISSON2D5) {
            cxb = *reinterpret_cast<const double2*>(xs + BW * 8);
            crb = *reinterpret_cast<const double2*>(xs + G.xr_bytes + BW * 8);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * st);
        if (s >= 2) {
          if (OUT == 2 && lazy) {
            wait_scalars(P, lazy_seq);
            alpha = __ldcg(&S.alpha);
            lazy = false;
          }
          if (va) emit<KIND, OUT>(P, out, A, ta, idx, pair, alpha, cxa, cra, acc);
          if (kTwoRows<KIND> && vb) emit<KIND, OUT>(P, out, B, tb, idx + L.pitch, pair, alpha, cxb, crb, acc);
          idx += ps;
        }
        run_advance<KIND>(P, A, na, ca, ta);
        if (kTwoRows<KIND>) run_advance<KIND>(P, B, nb, cb, tb);
      }
    }
  }
  R.n = n;
}

template <int KIND, int OUT>
__global__ void __launch_bounds__(kNT, 1) k_spmv_tma(const __grid_constant__ RankDev P,
                                                      const __grid_constant__ CUtensorMap tmap,
                                                      const __grid_constant__ CUtensorMap tmap_h,
                                                      const __grid_constant__ CUtensorMap tmx,
                                                      const __grid_constant__ CUtensorMap tmr,
                                                      const __grid_constant__ BandGeo G, double* __restrict__ out,
                                                      int mode) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ double red_smem[64];
  __shared__ __align__(8) unsigned long long bars[2 * kMaxStages];
  Scalars& S = P.ctrl->s;
  pdl_launch_dependents();
  Ring R;
  ring_init(R, bars, G.nstages);
  block_sync();
  pdl_wait();  // everything below reads what the previous kernel of the stream produced
  if (S.done != 0) return;
  double acc = 0.0;
  phase_spmv<KIND, OUT>(P, G, &tmap, &tmap_h, &tmx, &tmr, out, S.alpha, S.halo_seq, smem_raw, R, acc);
  if (mode == kSpmvPlain) return;
  double a = acc, z = 0.0;
  if (grid_sum2_last_t<kNT>(a, z, P.partials, &S.ticket_red, red_smem)) {
    const bool ok = allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0) {
      if (OUT == 2) {
        if (ok) cg_after_rr(P, a);
      } else if (mode == kSpmvAlpha || mode == kSpmvAlphaNoStore) {
        S.pAp = a;
        S.alpha = S.rho / a;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_pspmv_tma ("merged" variant, 64 B/DOF): the p-update is folded into the SpMV.
//   p_new = r + beta * p_old   on the band's RB+2 rows of every plane, evaluated in registers from
//                              TMA slabs of r and p_old: the halo rows and x neighbours are
//                              RECOMPUTED instead of waiting for the neighbouring band/thread,
//   p_new -> global (owned rows only, into the other exchange buffer), p_new.A p_new -> alpha.
// Boundary first: before the sweep all threads compute p_new on the rank's first/last plane and
// store it straight into the neighbours' ghost planes (P2P) and raise their flags; the sweep runs
// towards the ghost plane, so the neighbour's data is needed last. For a ghost plane the "r slab"
// is loaded from the neighbour-written ghost plane of the NEW buffer and used as is.
// Replaces cublasDscal + cublasDaxpy + cusparseSpMV + cublasDdot (CUDACG.cu:342-347, 288-304).
// z-slab decompositions only (no y neighbours).
// ---------------------------------------------------------------------------------------------
// All threads of the CTA. cnt: CGB_NDIR zeroed counters in shared memory (a __syncthreads() after zeroing them
// must precede the call). acc receives this thread's share of p_new . A p_new.
template <int KIND>
__device__ __forceinline__ void phase_pspmv(const RankDev& P, const BandGeo& G, const CUtensorMap* tm_r,
                                            const CUtensorMap* tm_pold, const CUtensorMap* tm_pnew,
                                            const CUtensorMap* tm_r_h, const CUtensorMap* tm_pold_h,
                                            const CUtensorMap* tm_pnew_h, const double* pold, double* pnew, int newbuf,
                                            double beta, u64 version, unsigned* cnt, unsigned char* smem_raw, Ring& R,
                                            double& acc, bool lazy = false, u64 lazy_seq = 0,
                                            int* done_seen = nullptr) {
  constexpr int BW = BoxW<KIND>::v;
  Scalars& S = P.ctrl->s;
  unsigned tid = threadIdx.x, bid = blockIdx.x;  // opaque copies, see phase_spmv
  asm volatile("" : "+r"(tid), "+r"(bid));
  const unsigned full0 = R.full0, empty0 = R.empty0;
  const unsigned kStages = (unsigned)G.nstages;
  const cgb_layout_t& L = P.lay;
  const bool nb_zlo = P.nbr[dir_index(0, -1)] >= 0, nb_zhi = P.nbr[dir_index(0, 1)] >= 0;
  const int boxb = G.prow * BW * 8;
  unsigned n = R.n;

  // One slab (r or ghost p_new, plus p_old) of step s of an item into its ring stage. TMA lane only.
  auto issue = [&](const Item& it, int s, unsigned st, unsigned ph) {
    const int y0 = it.yb * G.RB;
    const unsigned p_tx = (unsigned)(G.NBX * boxb + (G.ntx > 1 ? 2 * G.prow * 16 : 0));
    const int q = it.dir > 0 ? it.k0 - 1 + s : it.k1 - s;
    const bool ghost = (q == -1 && nb_zlo) || (q == L.nz && nb_zhi);
    mbar_wait(empty0 + 8 * st, ph ^ 1u, S, 3);
    mbar_expect_tx(full0 + 8 * st, 2u * p_tx);
    const unsigned dst = smem_u32(smem_raw + (size_t)st * G.stage_bytes);
    if (G.wide) {
      tma_load_3d(dst, ghost ? tm_pnew : tm_r, 0, it.x0 / BW, q + 1, full0 + 8 * st);
      tma_load_3d(dst + G.p_bytes, tm_pold, 0, it.x0 / BW, q + 1, full0 + 8 * st);
    } else
      for (int b = 0; b < G.NBX; ++b) {
        tma_load_3d(dst + b * boxb, ghost ? tm_pnew : tm_r, it.x0 + b * BW, y0, q + 1, full0 + 8 * st);
        tma_load_3d(dst + G.p_bytes + b * boxb, tm_pold, it.x0 + b * BW, y0, q + 1, full0 + 8 * st);
      }
    if (G.ntx > 1) {
      const int xl = it.x0 - 2, xr = it.x0 + 2 * G.LX;
      tma_load_3d(dst + G.h_off, ghost ? tm_pnew_h : tm_r_h, xl, y0, q + 1, full0 + 8 * st);
      tma_load_3d(dst + G.h_off + 128, ghost ? tm_pnew_h : tm_r_h, xr, y0, q + 1, full0 + 8 * st);
      tma_load_3d(dst + G.p_bytes + G.h_off, tm_pold_h, xl, y0, q + 1, full0 + 8 * st);
      tma_load_3d(dst + G.p_bytes + G.h_off + 128, tm_pold_h, xr, y0, q + 1, full0 + 8 * st);
    }
  };
  auto is_ghost_step = [&](const Item& it, int s) {
    const int q = it.dir > 0 ? it.k0 - 1 + s : it.k1 - s;
    return (q == -1 && nb_zlo) || (q == L.nz && nb_zhi);
  };

  // ---- persistent kernel: beta is the result of reduction lazy_seq, which may still be crossing NVLink. The TMA
  // lane first fills the ring with the opening slabs of its first item (they depend on local data only, which the
  // grid barrier has made visible), then everybody waits for the scalars. ----
  int prefetched = 0;
  if (lazy) {
    const int item0 = bid < (unsigned)G.nctas ? (int)bid : G.nitems;
    if (tid == kNC && item0 < G.nitems && !(P.dbg_flags & 1)) {
      const Item it = make_item(P, G, item0, nb_zlo, nb_zhi, false, false);
      const int nsteps = it.k1 - it.k0 + 2;
      unsigned st = n % kStages, ph = (n / kStages) & 1u;
      while (prefetched < nsteps && prefetched < (int)kStages && !is_ghost_step(it, prefetched)) {
        issue(it, prefetched, st, ph);
        ring_next(st, ph, kStages);
        ++prefetched;
      }
    }
    wait_scalars(P, lazy_seq);
    beta = __ldcg(&S.beta);
    version = __ldcg(&S.halo_seq);
    if (done_seen) *done_seen = __ldcg(&S.done);
  }

  // ---- boundary planes first: push p_new of my first/last plane into the neighbours' ghosts ----
  if (P.nbr_mask) {
    const unsigned pw = (unsigned)((L.nx + 1) / 2);
    const unsigned per_plane = (unsigned)L.ny * pw;
    const unsigned total = ((nb_zlo ? 1u : 0u) + (nb_zhi ? 1u : 0u)) * per_plane;
    // With spare SMs, dedicated CTAs (blockIdx >= nctas) push the halo while the band CTAs are
    // already sweeping; otherwise every CTA takes a share before it starts.
    const bool dedicated = G.nhalo > 0;
    const bool pusher = bid >= (unsigned)G.nctas && bid < (unsigned)(G.nctas + G.nhalo);
    const unsigned first = dedicated ? (pusher ? (bid - G.nctas) * kNT + tid : total)
                                     : bid * kNT + tid;
    const unsigned step = (dedicated ? (unsigned)G.nhalo : gridDim.x - (lazy ? 1u : 0u)) * kNT;  // lazy: minus the service CTA
    for (unsigned u0 = first; u0 < total; u0 += 4 * step) {  // 4 units in flight per thread
      int64_t src[4], dst[4];
      int dd[4];
      double2 vr[4], vp[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const unsigned u = u0 + t * step;
        src[t] = -1;
        if (u < total) {
          const unsigned pl = u / per_plane, v = u - pl * per_plane;
          const int j = (int)(v / pw), i = 2 * (int)(v - (unsigned)j * pw);
          const bool lo = pl == 0 && nb_zlo;
          const int k = lo ? 0 : L.nz - 1;
          dd[t] = lo ? dir_index(0, -1) : dir_index(0, 1);
          src[t] = lay_index(L, i, j, k);
          dst[t] = ghost_index(P, dd[t], i, j, k);
        }
      }
      // L2 loads: in a persistent kernel these cells were rewritten by other SMs one phase ago
#pragma unroll
      for (int t = 0; t < 4; ++t)
        if (src[t] >= 0) {
          vr[t] = __ldcg(reinterpret_cast<const double2*>(P.r + src[t]));
          vp[t] = __ldcg(reinterpret_cast<const double2*>(pold + src[t]));
        }
#pragma unroll
      for (int t = 0; t < 4; ++t)
        if (src[t] >= 0) {
          vp[t].x = vr[t].x + beta * vp[t].x;
          vp[t].y = vr[t].y + beta * vp[t].y;
          *reinterpret_cast<double2*>(P.nbr_exch[newbuf][dd[t]] + dst[t]) = vp[t];
          atomicAdd(&cnt[dd[t]], 1u);
        }
    }
    halo_commit(P, cnt, version);
  }

  const int warp = tid >> 5, lane = tid & 31;

  if (warp == kNC / 32) {
    if (lane == 0) {
      bool waited = false;
      n += (unsigned)prefetched;   // slabs of the first item that are already in flight
      unsigned st = n % kStages, ph = (n / kStages) & 1u;
      for (int item = bid < (unsigned)G.nctas ? (int)bid : G.nitems; item < G.nitems; item += G.nctas) {
        const Item it = make_item(P, G, item, nb_zlo, nb_zhi, false, false);
        const int nsteps = it.k1 - it.k0 + 2;
        int s = prefetched;
        prefetched = 0;
        for (; s < nsteps; ++s, ++n, ring_next(st, ph, kStages)) {
          if (is_ghost_step(it, s) && !waited) {
            for (int d = 0; d < CGB_NDIR; ++d)
              if (d != 4 && P.nbr[d] >= 0 && !spin_until_ge(&P.ctrl->halo_flag[d], version, S.timeout_cycles))
                S.done = CGB_ERR_TIMEOUT;
            asm volatile("fence.proxy.async;" ::: "memory");
            waited = true;
          }
          issue(it, s, st, ph);
        }
      }
    }
  } else {
    const int lx = tid % G.LX, ly = tid / G.LX;
    unsigned st = n % kStages, ph = (n / kStages) & 1u;
    for (int item = bid < (unsigned)G.nctas ? (int)bid : G.nitems; item < G.nitems; item += G.nctas) {
      const Item it = make_item(P, G, item, nb_zlo, nb_zhi, false, false);
      const bool active = ly < G.LY && it.x0 + 2 * lx < L.nx;   // idle threads shadow thread 0, see phase_spmv
      const int il = active ? 2 * lx : 0;
      const int ra = active ? 2 * ly : 0;
      const int i = it.x0 + il;
      const bool pair = i + 1 < L.nx;
      const Taps tp = make_taps<BW>(G, il, i, L.nx);
      const int ja = it.yb * G.RB + ra;
      const int jend = active ? min(L.ny, (it.yb + 1) * G.RB) : 0;
      const bool va = ja < jend, vb = kTwoRows<KIND> && ja + 1 < jend;
      const int nsteps = it.k1 - it.k0 + 2;
      const int64_t ps = (int64_t)L.pitch * L.py * it.dir;
      int64_t idx = lay_index(L, i, ja, it.dir > 0 ? it.k0 : it.k1 - 1);   // plane emitted at step 2
      int64_t gq = lay_index(L, i, ja, it.dir > 0 ? it.k0 - 1 : it.k1);    // plane loaded at step 0
      Run2 A{}, B{};
      for (int s = 0; s < nsteps; ++s, ++n, ring_next(st, ph, kStages)) {
        const int q = it.dir > 0 ? it.k0 - 1 + s : it.k1 - s;
        const bool ghost = (q == -1 && nb_zlo) || (q == L.nz && nb_zhi);
        const bool inb = s >= 1 && s <= nsteps - 2;   // the item's own planes: the two outer slabs belong to others
        mbar_wait(full0 + 8 * st, ph, S, 4);
        const unsigned char* rs = smem_raw + (size_t)st * G.stage_bytes;
        double2 pna = make_double2(0, 0), pnb = pna;
        PS2 na, nb;
        plane_from_two<KIND>(rs, rs + G.p_bytes, tp, ra, beta, ghost, na, nb, pna, pnb);
        double ca[2], ta[2], cb[2] = {0, 0}, tb[2] = {0, 0};
        plane_ct<KIND>(P, na, ca, ta);
        if (kTwoRows<KIND>) plane_ct<KIND>(P, nb, cb, tb);
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * st);
        if (inb) {  // my two rows of p_new on the plane just loaded
          if (va) {
            if (pair) *reinterpret_cast<double2*>(pnew + gq) = pna;
            else pnew[gq] = pna.x;
          }
          if (vb) {
            if (pair) *reinterpret_cast<double2*>(pnew + gq + L.pitch) = pnb;
            else pnew[gq + L.pitch] = pnb.x;
          }
        }
        gq += ps;
        if (s >= 2) {
          const double2 z2 = make_double2(0, 0);
          if (va) emit<KIND, 1>(P, nullptr, A, ta, idx, pair, 0.0, z2, z2, acc);
          if (kTwoRows<KIND> && vb) emit<KIND, 1>(P, nullptr, B, tb, idx + L.pitch, pair, 0.0, z2, z2, acc);
          idx += ps;
        }
        run_advance<KIND>(P, A, na, ca, ta);
        if (kTwoRows<KIND>) run_advance<KIND>(P, B, nb, cb, tb);
      }
    }
  }
  R.n = n;
}

template <int KIND>
__global__ void __launch_bounds__(kNT, 1) k_pspmv_tma(const __grid_constant__ RankDev P,
                                                       const __grid_constant__ CUtensorMap tm_r,
                                                       const __grid_constant__ CUtensorMap tm_pold,
                                                       const __grid_constant__ CUtensorMap tm_pnew,
                                                       const __grid_constant__ CUtensorMap tm_r_h,
                                                       const __grid_constant__ CUtensorMap tm_pold_h,
                                                       const __grid_constant__ CUtensorMap tm_pnew_h,
                                                       const __grid_constant__ BandGeo G,
                                                       const double* __restrict__ pold, double* __restrict__ pnew,
                                                       int newbuf) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ double red_smem[64];
  __shared__ __align__(8) unsigned long long bars[2 * kMaxStages];
  __shared__ unsigned cnt[CGB_NDIR];
  Scalars& S = P.ctrl->s;
  pdl_launch_dependents();
  Ring R;
  ring_init(R, bars, G.nstages);
  if (threadIdx.x < CGB_NDIR) cnt[threadIdx.x] = 0u;
  block_sync();
  pdl_wait();  // everything below reads what the previous kernel of the stream produced
  if (S.done != 0) return;
  double acc = 0.0;
  phase_pspmv<KIND>(P, G, &tm_r, &tm_pold, &tm_pnew, &tm_r_h, &tm_pold_h, &tm_pnew_h, pold, pnew, newbuf, S.beta,
                    S.halo_seq, cnt, smem_raw, R, acc);
  double a = acc, z = 0.0;
  if (grid_sum2_last_t<kNT>(a, z, P.partials, &S.ticket_red, red_smem)) {
    allreduce2_last_cta(P, a, z, red_smem);
    if (threadIdx.x == 0) {
      S.pAp = a;
      S.alpha = S.rho / a;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_cg_persist: the whole "merged" CG iteration - and as many iterations as the host asks for - in ONE
// cooperative launch, one CTA per SM. Per iteration:
//   phase A = phase_pspmv  (p = r + beta p, boundary planes pushed to the peers first, SpMV, p.Ap)
//   barrier  = grid-wide rendezvous fused with the cross-GPU all-reduce: the CTA that arrives last adds the
//              per-CTA partials in index order, exchanges the rank's sum with all peers through 8-byte
//              flag-in-data P2P stores, forms alpha and releases the other CTAs through one flag in L2
//   phase B = phase_spmv<2> (stencil again from p, x += alpha p, r -= alpha Ap, r.r)
//   barrier  = same, then the reference's stopping test and beta (cg_after_rr)
// Compared with two kernels per iteration this removes two grid drains, two launches and two kernel
// prologues from the critical path of every iteration (the per-iteration fixed cost that capped strong
// scaling, profiles/ROOFLINE.md), and it takes the host out of the loop completely: the reference blocks
// the host twice per iteration (CUDACG.cu:304, CUDACG.cu:328).
// Data written through the generic proxy in one phase (p_new, x, r) is read by TMA (async proxy) in the
// next: every thread fences (fence.proxy.async) before it arrives at the barrier and the TMA lane fences
// again after the acquire.
// ---------------------------------------------------------------------------------------------
// The eight tensor maps of a context, in DEVICE memory (written once at context creation): r with halo rows,
// the two exchange buffers, their 2-column halo-box twins, x and r (band rows). The persistent kernel reads its
// descriptors from here rather than from its parameter block.
struct PersistMaps {
  CUtensorMap rh, e[2], rh_h, e_h[2], x, r;
};
static_assert(sizeof(PersistMaps) == 8 * 128, "PersistMaps layout");

// Loop state of the persistent kernel lives in shared memory (accessed volatile), not in registers: the two
// phases run at the register limit of a 544-thread CTA (96) and anything carried across them would be
// spilled inside their inner loops.
struct PersistState {
  u64 seq;          // generation of the last grid barrier this CTA has passed
  unsigned nA, nB;  // slabs issued/consumed so far on the two mbarrier rings
  int it;           // iterations completed by this launch
  int stop;         // the solve is finished (or failed): every thread leaves the loop
};

constexpr int kPersistPartials = 256;  // per-barrier-parity block of per-CTA partials (grid <= SM count < 256)

// Local half of a barrier: publish this CTA's partial sum, arrive, and wait until every working CTA of this GPU has
// arrived - i.e. until everything the phase wrote (p_new, or x and r) is visible. The arriving CTAs do NOT wait
// for the cross-GPU sum: the service CTA picks the partials up from here (persist_service) and the next phase
// starts loading at once; its threads wait for the scalars only where they first need them.
__device__ __forceinline__ void persist_arrive(const RankDev& P, double a, volatile PersistState* ps, double* red_smem,
                                               int done_seen, bool count_iter) {
  Scalars& S = P.ctrl->s;
  double z = 0.0;
  asm volatile("fence.proxy.async;" ::: "memory");  // my generic-proxy stores vs TMA reads of the next phase
  block_sum2_t<kNT>(a, z, red_smem);
  if (threadIdx.x == 0) {
    const u64 seq = ps->seq + 1;
    ps->seq = seq;
    __stcg(&P.partials[(seq & 1) * kPersistPartials + blockIdx.x], make_double2(a, 0.0));
    __threadfence();
    const unsigned narrive = gridDim.x - 1;  // everybody but the service CTA
    if (atomicAdd(&S.ticket_red, 1u) == narrive - 1) {
      S.ticket_red = 0u;
      __threadfence();
      st_release_gpu(&S.gbar, seq);
    } else {
      const long long t0 = clock64();
      while (ld_acquire_gpu(&S.gbar) < seq) {
        if (clock64() - t0 > 4 * S.timeout_cycles) {
          if (atomicCAS(&S.dbg[0], 0, 11) == 0) {
            S.dbg[1] = (int)blockIdx.x;
            S.dbg[3] = (int)seq;
            S.dbg[5] = (int)*(volatile u64*)&S.gbar;
            S.dbg[6] = (int)*(volatile u64*)&S.sbar;
            S.dbg[7] = (int)*(volatile unsigned*)&S.ticket_red;
          }
          S.done = CGB_ERR_TIMEOUT;
          done_seen = CGB_ERR_TIMEOUT;
          break;
        }
      }
    }
    if (count_iter) ps->it = ps->it + 1;
    if (done_seen != 0) ps->stop = 1;
  }
  block_sync();  // reconverges warp 0 before the aligned barrier
}

// The service CTA (last block of the grid, one warp): for every barrier it waits until all working CTAs have
// arrived, adds their partials in index order (deterministic), exchanges the rank's sum with every peer through
// 8-byte flag-in-data P2P stores (rank-ordered sum: bit-identical scalars on all ranks), updates the CG scalars
// and publishes them (sbar). None of this is on the critical path of the working CTAs, which are already loading
// the next phase. Replaces the two blocking scalar read-backs of the reference (CUDACG.cu:304, CUDACG.cu:328).
__device__ __forceinline__ void persist_service(const RankDev& P, int niter) {
  Scalars& S = P.ctrl->s;
  const int lane = threadIdx.x;
  const unsigned nwork = gridDim.x - 1;
  // sbar, not gbar: the working CTAs may pass their first barrier (and advance gbar) before this CTA has even
  // started, while sbar only ever moves here. The two are equal whenever no launch is in flight.
  u64 seq = __ldcg(&S.sbar);
  bool done = __ldcg(&S.done) != 0;
  if (done) return;  // the working CTAs leave at once, too
  // This warp is the only writer of the CG scalars while the kernel runs, so it keeps them in registers: between
  // "all partials are in" and "alpha/beta published" there is no dependent load from memory left.
  double rho = S.rho;
  const double thresh = S.thresh;
  const int maxit = S.maxit;
  int iter = S.iter;
  u64 halo_seq = S.halo_seq, red_seq = S.red_seq;
  const long long timeout = S.timeout_cycles;
  for (int it = 0; it < niter; ++it) {
    for (int phase = 0; phase < 2; ++phase) {
      ++seq;
      bool alive = true;
      if (lane == 0) {
        const long long t0 = clock64();
        while (ld_acquire_gpu(&S.gbar) < seq)
          if (clock64() - t0 > 8 * timeout) {  // the working CTAs are gone (their waits are bounded, too)
            if (atomicCAS(&S.dbg[0], 0, 12) == 0) {
              S.dbg[3] = (int)seq;
              S.dbg[5] = (int)*(volatile u64*)&S.gbar;
              S.dbg[6] = (int)*(volatile u64*)&S.sbar;
            }
            S.done = CGB_ERR_TIMEOUT;
            alive = false;
            break;
          }
      }
      if (!__shfl_sync(0xffffffffu, (int)alive, 0)) {
        // keep the invariant sbar == gbar for the next launch, then give up
        if (lane == 0) st_release_gpu(&S.sbar, ld_acquire_gpu(&S.gbar));
        return;
      }
      unsigned long long t_local = 0;
      if (P.trace && lane == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_local));
      double a = 0.0;
      const double2* part = P.partials + (seq & 1) * kPersistPartials;
      for (unsigned i = lane; i < nwork; i += 32) a += __ldcg(&part[i]).x;   // fixed order per lane ...
      a = warp_sum(a);                                                         // ... and a fixed tree across lanes
      bool ok = true;
      if (P.nranks > 1) {
        const int slot = (int)(red_seq & (kRedSlots - 1));
        const unsigned tag = (unsigned)(red_seq + 1);
        double va = 0.0, vb = 0.0;
        if (lane < P.nranks) {
          ll_send2(&P.peer_ctrl[lane]->red_ll[slot][P.rank][0], a, 0.0, tag);
          ok = ll_recv2(&P.ctrl->red_ll[slot][lane][0], tag, timeout, va, vb);
        }
        ok = __all_sync(0xffffffffu, ok);
        a = 0.0;
        for (int q = 0; q < P.nranks; ++q) a += __shfl_sync(0xffffffffu, va, q);   // rank order on every rank
      }
      red_seq += 1;
      const int tslot = iter;
      int new_done = 0;
      if (lane == 0) {
        // ---- what the working CTAs are waiting for, first ----
        double rn = 0.0;
        if (!ok) {
          new_done = CGB_ERR_TIMEOUT;
        } else if (done) {
          // the speculative p.Ap of a finished solve (see persist_iteration): nothing to update
        } else if (phase == 0) {
          S.alpha = rho / a;   // CUDACG.cu:311
        } else {
          // the reference's test position and rule: absolute ||r||_2 < tol after the x/r update (CUDACG.cu:333)
          rn = sqrt(a);
          if (!isfinite(a)) {
            new_done = CGB_ERR_BREAKDOWN;
          } else if (rn < thresh) {
            new_done = 1;
          } else {
            S.beta = a / rho;  // CUDACG.cu:339
            if (iter + 1 >= maxit) new_done = 2;
            else S.halo_seq = ++halo_seq;
          }
        }
        if (new_done != 0) S.done = new_done;
        __threadfence();
        st_release_gpu(&S.sbar, seq);
        // ---- bookkeeping, off the critical path ----
        S.last_sum[0] = a;
        S.last_sum[1] = 0.0;
        S.red_seq = red_seq;
        if (ok && !done) {
          if (phase == 0) {
            S.pAp = a;
          } else {
            if (iter < P.hist_cap) P.hist[iter] = rn;
            if (P.hist_host && iter < P.hist_host_cap) P.hist_host[iter] = rn;  // zero-copy D2H of this step's result
            S.rnorm = rn;
            S.iter = iter + 1;
            if (new_done == 0 || new_done == 2) {
              S.rho_prev = rho;
              S.rho = a;
            }
          }
        }
        if (P.trace && !done && tslot < P.trace_cap) {
          unsigned long long t_rel;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_rel));
          P.trace[4 * tslot + 2 * phase] = t_local;
          P.trace[4 * tslot + 2 * phase + 1] = t_rel;
        }
        if (ok && !done && phase == 1) {
          if (new_done == 0 || new_done == 2) rho = a;
          iter += 1;
        }
      }
      const bool was_done = done;
      done = __shfl_sync(0xffffffffu, (int)(done || new_done != 0), 0) != 0;
      // finished as of the previous barrier: this was the speculative barrier A (see persist_iteration) and the
      // working CTAs leave after it, too
      if (phase == 0 && was_done) return;
    }
  }
}

// One iteration with the SpMV input in exchange buffer CUR (compile-time, so that the vector and tensor-map
// addresses stay constant-bank operands of the kernel's parameter block). Returns false when the loop must end.
template <int KIND, int CUR>
__device__ __forceinline__ bool persist_iteration(const RankDev& P, const PersistMaps& M, const BandGeo& GA,
                                                  const BandGeo& GB, int niter, volatile PersistState* ps,
                                                  unsigned long long* bars, unsigned* cnt, double* red_smem,
                                                  unsigned char* smem_raw) {
  constexpr int NB = CUR ^ 1;
  if (ps->stop || ps->it >= niter) return false;
  if (threadIdx.x < CGB_NDIR) cnt[threadIdx.x] = 0u;
  block_sync();
  if (threadIdx.x == kNC) asm volatile("fence.proxy.async;" ::: "memory");  // TMA lane, after the barrier's acquire
  __syncwarp();
  {
    // beta, the halo version and the stopping decision come from the r.r reduction of the previous iteration
    // (reduction ps->seq), which the service CTA may still be finishing: phase A starts its loads and waits for
    // them inside. If the solve turns out to be finished, this phase still runs to its barrier (its output is
    // never used) so that all CTAs and all ranks leave the loop at the same place.
    Ring R{smem_u32(bars), smem_u32(bars + kMaxStages), ps->nA};
    double acc = 0.0;
    int done_seen = 0;
    phase_pspmv<KIND>(P, GA, &M.rh, &M.e[CUR], &M.e[NB], &M.rh_h, &M.e_h[CUR], &M.e_h[NB], P.exch[CUR], P.exch[NB], NB,
                      0.0, 0, cnt, smem_raw, R, acc, true, ps->seq, &done_seen);
    if (threadIdx.x == 0) ps->nA = R.n;
    __syncwarp();
    persist_arrive(P, acc, ps, red_smem, done_seen, false);
  }
  if (ps->stop) return false;
  if (threadIdx.x == kNC) asm volatile("fence.proxy.async;" ::: "memory");
  __syncwarp();
  {
    Ring R{smem_u32(bars + 2 * kMaxStages), smem_u32(bars + 3 * kMaxStages), ps->nB};
    double acc = 0.0;
    phase_spmv<KIND, 2>(P, GB, &M.e[NB], &M.e_h[NB], &M.x, &M.r, nullptr, 0.0, __ldcg(&P.ctrl->s.halo_seq), smem_raw, R,
                        acc, true, ps->seq);
    if (threadIdx.x == 0) ps->nB = R.n;
    __syncwarp();
    persist_arrive(P, acc, ps, red_smem, 0, true);
  }
  return true;
}

template <int KIND>
__global__ void __launch_bounds__(kNT, 1) k_cg_persist(const __grid_constant__ RankDev P,
                                                        const PersistMaps* __restrict__ Mp,
                                                        const __grid_constant__ BandGeo GA,
                                                        const __grid_constant__ BandGeo GB, int cur0, int niter) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ double red_smem[64];
  __shared__ __align__(8) unsigned long long bars[4 * kMaxStages];
  __shared__ unsigned cnt[CGB_NDIR];
  __shared__ PersistState ps_mem;
  volatile PersistState* ps = &ps_mem;
  Scalars& S = P.ctrl->s;
  if (blockIdx.x == gridDim.x - 1) {  // the service CTA: one warp, everything else leaves
    if (threadIdx.x < 32) persist_service(P, niter);
    return;
  }
  // Nothing below diverges inside a warp before the first barrier: all threads run the same straight-line code
  // (predicated mbarrier inits, every thread stores the same initial loop state). With a one-thread block here,
  // ptxas emitted no reconvergence point for it (the function is full of early exits), lanes 1-31 of warp 0 ran
  // ahead of thread 0 and clobbered the uniform registers its stores still needed: the loop state never reached
  // shared memory and the kernel dead-locked (found on B200, only for some code layouts).
  ring_init_uniform(bars, GA.nstages);
  ring_init_uniform(bars + 2 * kMaxStages, GB.nstages);
  {
    // Every CTA reads the barrier generation before it can possibly move: it only advances once ALL working CTAs
    // of this launch have arrived at the first barrier. At entry sbar == gbar (the previous launch has completed).
    const u64 g0 = __ldcg(&S.gbar);
    const int stop0 = __ldcg(&S.done) != 0 ? 1 : 0;
    ps->seq = g0;       // (all threads store identical values)
    ps->nA = 0u;
    ps->nB = 0u;
    ps->it = 0;
    ps->stop = stop0;
  }
  block_sync();
  // `cur` selects one of two instantiations per iteration; the loop is kept reducible (no jump into its middle for
  // an odd entry parity), which ptxas needs to place its reconvergence points.
  int cur = cur0;
#pragma unroll 1
  for (;;) {
    const bool go = cur == 0 ? persist_iteration<KIND, 0>(P, *Mp, GA, GB, niter, ps, bars, cnt, red_smem, smem_raw)
                             : persist_iteration<KIND, 1>(P, *Mp, GA, GB, niter, ps, bars, cnt, red_smem, smem_raw);
    if (!go) break;
    cur ^= 1;
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

// 2D grids (one row per plane, no y padding) whose rows are whole 256-cell boxes and are stored without pitch padding.
bool wide_rows(int kind, const cgb_layout_t& L) {
  static const bool off = [] { const char* e = std::getenv("CGB_TMA_NARROW"); return e && e[0] == '1'; }();
  return !off && !kind_has_y_taps(kind) && L.py == 1 && L.nx % box_w(kind) == 0 && L.pitch == L.nx;
}

bool make_geo(const RankDev& P, int nsm, int mode, BandGeo* out, int* grid, size_t* smem) {
  const cgb_layout_t& L = P.lay;
  if (!kind_is_stencil(P.st.kind)) return false;
  BandGeo G{};
  const bool flat = !kind_has_y_taps(P.st.kind);      // 2D grid stored as nx x 1 x ny
  const int tile_max = flat ? kTileX2 : kTileX3;
  G.ntx = (L.nx + tile_max - 1) / tile_max;
  const int tile = G.ntx > 1 ? tile_max : L.nx;     // wide rows are cut into tiles
  G.LX = (((tile + 1) / 2) + 31) / 32 * 32;
  G.LY = flat ? 1 : kNC / G.LX;
  if (G.LY > 8) G.LY = 8;  // keeps a slab <= 18 rows
  G.RB = flat ? 1 : 2 * G.LY;
  G.HY = flat ? 0 : 1;
  G.prow = G.RB + 2 * G.HY;
  const int BW = box_w(P.st.kind);
  G.NBX = (tile + BW - 1) / BW;
  G.wide = wide_rows(P.st.kind, L) ? 1 : 0;
  G.nby = (L.ny + G.RB - 1) / G.RB;
  G.nbands = G.ntx * G.nby;
  G.h_off = G.NBX * G.prow * BW * 8;                  // halo boxes sit behind the main boxes
  G.p_bytes = G.h_off + (G.ntx > 1 ? 256 : 0);
  G.xr_bytes = G.NBX * G.RB * BW * 8;
  G.stage_bytes = mode == kSpmvPUpdate ? 2 * G.p_bytes : G.p_bytes + (mode == kSpmvUpdate ? 2 * G.xr_bytes : 0);
  G.nstages = (int)((200 * 1024) / G.stage_bytes);
  const int cap = flat ? kMaxStages
                       : (mode == kSpmvAlpha || mode == kSpmvDot || mode == kSpmvPlain || mode == kSpmvAlphaNoStore ? 6 : 4);
  if (G.nstages > cap) G.nstages = cap;
  if (G.nstages < 2) return false;
  const size_t sm = (size_t)G.nstages * G.stage_bytes;
  const int capacity = nsm;  // 544 threads x ~96 registers: one CTA per SM
  // z split. One round (every CTA sweeps one item) is the default; when that leaves many SMs idle or the last round
  // mostly empty (96 bands on 148 SMs, 512 bands = 3.46 rounds), a finer split with several items per CTA is chosen
  // from the length of the longest CTA: rounds x (planes per item + the 2 planes an item loads without emitting).
  // The finer split re-reads 2 planes per item, so it has to promise > 12 % before it replaces the default.
  const int zs_max = L.nz / 8 > 1 ? L.nz / 8 : 1;
  int zs = capacity / G.nbands;
  if (zs > zs_max) zs = zs_max;
  if (zs < 1) zs = 1;
  {
    auto cost = [&](int z) {
      const int zc = (L.nz + z - 1) / z, ze = (L.nz + zc - 1) / zc;
      const long long items = (long long)G.nbands * ze;
      return ((items + capacity - 1) / capacity) * (long long)(zc + 2);
    };
    static const int forced = [] { const char* e = std::getenv("CGB_ZSPLIT"); return e ? std::atoi(e) : 0; }();
    int best = zs;
    for (int z = 1; z <= zs_max && z <= 256; ++z)
      if (cost(z) < cost(best)) best = z;
    if (cost(best) * 100 < cost(zs) * 88) zs = best;
    if (forced > 0) zs = forced < zs_max ? forced : zs_max;
  }
  G.zchunk = (L.nz + zs - 1) / zs;
  G.zs = (L.nz + G.zchunk - 1) / G.zchunk;
  const bool nb_lo = P.nbr[dir_index(0, -1)] >= 0, nb_hi = P.nbr[dir_index(0, 1)] >= 0;
  bool halves = false;
  if (mode == kSpmvPUpdate && nb_lo && nb_hi && G.zs == 1 && L.nz >= 4) {
    halves = true;
    // A rank with neighbours on both sides would need one ghost plane at the very start of a
    // single upward sweep, i.e. before the neighbour has pushed it. Two half sweeps, each towards
    // its own ghost plane, need both ghost planes last (a CTA runs them back to back).
    G.zchunk = (L.nz + 1) / 2;
    G.zs = 2;
  }
  G.nitems = G.nbands * G.zs;
  int g = G.nitems < capacity ? G.nitems : capacity;
  if (halves && G.nbands <= capacity) g = G.nbands;  // same band, both halves, on one CTA
  if (g > kMaxGrid) return false;
  G.nctas = g;
  G.nhalo = 0;
  if (mode == kSpmvPUpdate && P.nbr_mask && g < capacity) {
    G.nhalo = capacity - g < 24 ? capacity - g : 24;
    g += G.nhalo;
  }
  *out = G;
  *grid = g;
  *smem = sm;
  return true;
}

}  // namespace

// Geometry of the band kernels for inspection/tests (host only, needs no GPU).
int spmv_tma_geometry(const RankDev& P, int nsm, int mode, int out[16]) {
  BandGeo G;
  int grid;
  size_t sm;
  if (!make_geo(P, nsm, mode, &G, &grid, &sm)) return 0;
  const int v[16] = {G.LX, G.LY, G.RB, G.HY, G.prow, G.NBX, G.ntx, G.nby, G.nbands, G.zs, G.zchunk, G.nitems,
                     G.stage_bytes, G.nstages, grid, G.nhalo};
  for (int k = 0; k < 16; ++k) out[k] = v[k];
  return 1;
}

bool spmv_tma_supported(const RankDev& P) {
  BandGeo G;
  int grid;
  size_t sm;
  return make_geo(P, 148, kSpmvAlpha, &G, &grid, &sm) && make_geo(P, 148, kSpmvUpdate, &G, &grid, &sm) &&
         make_geo(P, 148, kSpmvPUpdate, &G, &grid, &sm);
}

// halo_rows = 1: slabs of RB+2 rows (SpMV input); 0: exactly the band's RB rows (x / r of the update).
// box_cols = 128 for the main boxes, 2 for the x-halo boxes of tiled rows.
int spmv_tma_make_map(int kind, const cgb_layout_t& L, const double* vec, void* map_out, int halo_rows, int box_cols) {
  EncodeFn enc = get_encode();
  if (!enc) return fail(CGB_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  RankDev P{};
  P.lay = L;
  P.st.kind = kind;
  BandGeo G;
  int grid;
  size_t sm;
  if (!make_geo(P, 148, kSpmvAlpha, &G, &grid, &sm)) return fail(CGB_ERR_UNSUPPORTED, "layout not supported by the TMA SpMV");
  cuuint64_t dims[3] = {(cuuint64_t)L.nx, (cuuint64_t)L.py, (cuuint64_t)L.pz};
  cuuint64_t strides[2] = {(cuuint64_t)L.pitch * 8, (cuuint64_t)L.pitch * L.py * 8};
  if (box_cols != 2) box_cols = box_w(kind);  // main boxes; 2 = the x-halo boxes of tiled rows
  cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)(halo_rows ? G.prow : G.RB), 1};
  if (G.wide && box_cols != 2) {  // a row seen as nx/256 lines of 256 cells: one box = the x tile of one row
    dims[0] = (cuuint64_t)box_cols;
    dims[1] = (cuuint64_t)(L.nx / box_cols);
    strides[0] = (cuuint64_t)box_cols * 8;
    box[1] = (cuuint32_t)G.NBX;
  }
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = enc((CUtensorMap*)map_out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(vec), dims, strides,
                   box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(CGB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return CGB_OK;
}

namespace {
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = std::getenv("CGB_PDL");   // measured on B200 (profiles/ROOFLINE.md): no gain at 512^3, so opt-in
    return e && e[0] == '1';
  }();
  return on;
}
// <<<>>> with the programmatic-stream-serialization attribute (see pdl_wait above).
template <class... KArgs, class... Args>
cudaError_t launch_band(void (*kernel)(KArgs...), int grid, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kNT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
}  // namespace

cudaError_t launch_spmv_tma(const RankDev& P, int device, const void* map, const void* map_h, const void* map_x,
                            const void* map_r, double* out, int mode, cudaStream_t s) {
  static bool attr_set[64] = {};
  if (device >= 0 && device < 64 && !attr_set[device]) {
    const int lim = 220 * 1024;
    const auto at = cudaFuncAttributeMaxDynamicSharedMemorySize;
    cudaError_t e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON3D7, 0>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON3D7, 1>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON3D7, 2>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_STENCIL3D27, 0>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_STENCIL3D27, 1>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_STENCIL3D27, 2>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON2D5, 0>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON2D5, 1>, at, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_spmv_tma<CGB_KIND_POISSON2D5, 2>, at, lim);
    if (e != cudaSuccess) return e;
    attr_set[device] = true;
  }
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device);
  BandGeo G;
  int grid;
  size_t sm;
  if (!make_geo(P, nsm, mode, &G, &grid, &sm)) return cudaErrorInvalidConfiguration;
  if (mode == kSpmvUpdate && (!map_x || !map_r)) return cudaErrorInvalidValue;
  if (G.ntx > 1 && !map_h) return cudaErrorInvalidValue;
  const CUtensorMap& tm = *reinterpret_cast<const CUtensorMap*>(map);
  const CUtensorMap& th = *reinterpret_cast<const CUtensorMap*>(map_h ? map_h : map);
  const CUtensorMap& tx = *reinterpret_cast<const CUtensorMap*>(map_x ? map_x : map);
  const CUtensorMap& tr = *reinterpret_cast<const CUtensorMap*>(map_r ? map_r : map);
  const int kd = P.st.kind;
  if (!kind_is_stencil(kd)) return cudaErrorInvalidValue;
  cudaError_t e;
#define CGB_TMA_LAUNCH(OUT)                                                                                   \
  do {                                                                                                        \
    if (kd == CGB_KIND_POISSON3D7)                                                                            \
      e = launch_band(k_spmv_tma<CGB_KIND_POISSON3D7, OUT>, grid, sm, s, P, tm, th, tx, tr, G, out, mode);    \
    else if (kd == CGB_KIND_STENCIL3D27)                                                                      \
      e = launch_band(k_spmv_tma<CGB_KIND_STENCIL3D27, OUT>, grid, sm, s, P, tm, th, tx, tr, G, out, mode);   \
    else                                                                                                      \
      e = launch_band(k_spmv_tma<CGB_KIND_POISSON2D5, OUT>, grid, sm, s, P, tm, th, tx, tr, G, out, mode);    \
  } while (0)
  if (mode == kSpmvUpdate) CGB_TMA_LAUNCH(2);
  else if (mode == kSpmvAlphaNoStore) CGB_TMA_LAUNCH(1);
  else CGB_TMA_LAUNCH(0);
#undef CGB_TMA_LAUNCH
  return e;
}

cudaError_t launch_pspmv_tma(const RankDev& P, int device, const void* const maps[6], const double* pold, double* pnew,
                             int newbuf, cudaStream_t s) {
  static bool attr_set[64] = {};
  if (device >= 0 && device < 64 && !attr_set[device]) {
    const auto at = cudaFuncAttributeMaxDynamicSharedMemorySize;
    cudaError_t e = cudaFuncSetAttribute(k_pspmv_tma<CGB_KIND_POISSON3D7>, at, 220 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_pspmv_tma<CGB_KIND_STENCIL3D27>, at, 220 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_pspmv_tma<CGB_KIND_POISSON2D5>, at, 220 * 1024);
    if (e != cudaSuccess) return e;
    attr_set[device] = true;
  }
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device);
  BandGeo G;
  int grid;
  size_t sm;
  if (!make_geo(P, nsm, kSpmvPUpdate, &G, &grid, &sm)) return cudaErrorInvalidConfiguration;
  // maps: r, p_old, p_new (main boxes), then their 2-column halo-box twins (used when rows are tiled)
  const CUtensorMap* m[6];
  for (int k = 0; k < 6; ++k) m[k] = reinterpret_cast<const CUtensorMap*>(maps[k] ? maps[k] : maps[k % 3]);
  if (P.st.kind == CGB_KIND_POISSON3D7)
    return launch_band(k_pspmv_tma<CGB_KIND_POISSON3D7>, grid, sm, s, P, *m[0], *m[1], *m[2], *m[3], *m[4], *m[5], G, pold,
                       pnew, newbuf);
  if (P.st.kind == CGB_KIND_STENCIL3D27)
    return launch_band(k_pspmv_tma<CGB_KIND_STENCIL3D27>, grid, sm, s, P, *m[0], *m[1], *m[2], *m[3], *m[4], *m[5], G, pold,
                       pnew, newbuf);
  if (P.st.kind == CGB_KIND_POISSON2D5)
    return launch_band(k_pspmv_tma<CGB_KIND_POISSON2D5>, grid, sm, s, P, *m[0], *m[1], *m[2], *m[3], *m[4], *m[5], G, pold,
                       pnew, newbuf);
  return cudaErrorInvalidValue;
}

// Per-device one-time set-up of the persistent kernel. The caller has made `device` current.
cudaError_t cg_persist_prepare(int device) {
  static std::mutex mu;
  static bool attr_set[64] = {};
  if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (attr_set[device]) return cudaSuccess;
  const auto at = cudaFuncAttributeMaxDynamicSharedMemorySize;
  cudaError_t e = cudaFuncSetAttribute(k_cg_persist<CGB_KIND_POISSON3D7>, at, 220 * 1024);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_cg_persist<CGB_KIND_STENCIL3D27>, at, 220 * 1024);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_cg_persist<CGB_KIND_POISSON2D5>, at, 220 * 1024);
  if (e != cudaSuccess) return e;
  attr_set[device] = true;
  return cudaSuccess;
}

size_t cg_persist_maps_bytes() { return sizeof(PersistMaps); }

// Copy the eight host-built tensor maps (order: r with halo rows, exch[0], exch[1], their three 2-column halo-box
// twins, x, r band rows) into the context's device-resident PersistMaps block.
cudaError_t cg_persist_upload_maps(void* d_maps, const void* const maps[8]) {
  PersistMaps M;
  for (int k = 0; k < 8; ++k)
    if (!maps[k]) return cudaErrorInvalidValue;
  M.rh = *reinterpret_cast<const CUtensorMap*>(maps[0]);
  M.e[0] = *reinterpret_cast<const CUtensorMap*>(maps[1]);
  M.e[1] = *reinterpret_cast<const CUtensorMap*>(maps[2]);
  M.rh_h = *reinterpret_cast<const CUtensorMap*>(maps[3]);
  M.e_h[0] = *reinterpret_cast<const CUtensorMap*>(maps[4]);
  M.e_h[1] = *reinterpret_cast<const CUtensorMap*>(maps[5]);
  M.x = *reinterpret_cast<const CUtensorMap*>(maps[6]);
  M.r = *reinterpret_cast<const CUtensorMap*>(maps[7]);
  return cudaMemcpy(d_maps, &M, sizeof(M), cudaMemcpyHostToDevice);
}

// d_maps: the context's device-resident PersistMaps (cg_persist_upload_maps). The caller has made `device`
// current and has called cg_persist_prepare(device).
cudaError_t launch_cg_persist(const RankDev& P, int device, const void* d_maps, int cur, int niter, cudaStream_t s) {
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device);
  BandGeo GA, GB;
  int ga, gb;
  size_t sa, sb;
  // one SM is reserved for the service CTA (cross-GPU reductions off the critical path)
  if (!make_geo(P, nsm - 1, kSpmvPUpdate, &GA, &ga, &sa) || !make_geo(P, nsm - 1, kSpmvUpdate, &GB, &gb, &sb))
    return cudaErrorInvalidConfiguration;
  if (!d_maps) return cudaErrorInvalidValue;
  const int grid = (ga > gb ? ga : gb) + 1;  // working CTAs + the service CTA, one per SM: all co-resident
  const size_t sm = sa > sb ? sa : sb;
  RankDev Pc = P;
  const void* Mp = d_maps;
  void* args[] = {&Pc, &Mp, &GA, &GB, &cur, &niter};
  const void* fn = P.st.kind == CGB_KIND_POISSON3D7    ? (const void*)k_cg_persist<CGB_KIND_POISSON3D7>
                   : P.st.kind == CGB_KIND_STENCIL3D27 ? (const void*)k_cg_persist<CGB_KIND_STENCIL3D27>
                   : P.st.kind == CGB_KIND_POISSON2D5  ? (const void*)k_cg_persist<CGB_KIND_POISSON2D5>
                                                       : nullptr;
  if (!fn) return cudaErrorInvalidValue;
  // cooperative: the launch fails instead of dead-locking if the CTAs cannot all be resident at once
  return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kNT), args, sm, s);
}

bool cg_persist_supported(int device) {
  int coop = 0;
  return cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device) == cudaSuccess && coop != 0;
}

}  // namespace cgb
