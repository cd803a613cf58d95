// Device-side foundation shared by every kernel: the per-rank control block that lives at the
// start of the peer-mapped arena, system-scope flag primitives, the deterministic block/grid
// reduction and the one-shot P2P all-reduce executed by the last CTA of a producing kernel.
//
// Where the reference blocks the host twice per iteration to read a scalar back
// (cublasDdot CUDACG.cu:304, cublasDnrm2 CUDACG.cu:328, host pointer mode), here the scalar
// recurrences (alpha, beta, rho, convergence flag) stay in device memory and the cross-GPU sum
// is a handful of 8-byte P2P stores plus a release flag over NVLink.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../topo/partition.hpp"
#include "cgb200.h"

namespace cgb {

constexpr int kRedSlots = 4;          // ring of reduction slots (2 would do)
constexpr int kMaxGrid = 2048;        // upper bound on CTAs of any persistent kernel
constexpr int kThreads = 256;         // CTA size of every kernel in this file family
constexpr size_t kCtrlBytes = 8192;   // control block footprint at arena offset 0
constexpr size_t kPartialsOff = kCtrlBytes;
static_assert(kPartialsOff + (size_t)kMaxGrid * sizeof(double2) <= 65536, "partials overlap the exchanged vectors");
constexpr size_t kExchOff = 65536;    // first exchanged vector; identical on every rank

typedef unsigned long long u64;

struct Scalars {
  double rho;        // r.r of the current residual (global)
  double rho_prev;
  double pAp;
  double alpha, beta;
  double r0norm;
  double thresh;     // absolute threshold on ||r||_2 (tol or tol*||r0||)
  double rnorm;      // last ||r||_2
  double tol;
  double gamma_old, alpha_old, delta;  // pipelined CG recurrences
  double last_sum[2];                  // last reduction result (standalone ops read this)
  u64 pub_seq;       // pipelined CG: sequence number of the reduction published by the last update
  u64 red_seq;       // sequence number of the next all-reduce this rank will publish
  u64 halo_seq;      // version of the newest halo this rank has pushed / expects
  u64 gbar;          // persistent kernel: generation of the last completed grid barrier (never reset by a solve)
  u64 sbar;          // persistent kernel: generation of the last reduction whose scalars have been published
  long long timeout_cycles;
  int tol_mode;
  int iter;          // completed iterations
  int done;          // 0 running, 1 converged, 2 maxit, <0 cgb_status
  int maxit;
  int cur;           // pipelined CG: exchange buffer that holds the current w (diagnostic)
  unsigned int ticket_red;
  unsigned int ticket_halo[CGB_NDIR];
  int dbg[8];        // post-mortem of a timed-out wait: site code, block, thread, ... (cgb_ctx_debug_print)
};

// Peers write red_ll and halo_flag with P2P stores; everything from `s` on is rank-local.
// red_ll[slot][rank][w]: the two fp64 partials of `rank` as four 8-byte words, each carrying half
// of a double in its low 32 bits and the reduction's sequence tag in its high 32 bits. An 8-byte
// store is single-copy atomic, so a word is either the old or the new (value, tag) pair and no
// fence or separate flag is needed: the reader simply polls until all four tags match.
struct Control {
  u64 red_ll[kRedSlots][CGB_MAX_RANKS][4];
  u64 halo_flag[CGB_NDIR];  // [d]: newest version pushed into my ghost layer by neighbour nbr[d]
  u64 pad_[7];
  u64 ready_flag[CGB_MAX_RANKS];  // distributed CSR: [q] = newest version of p that rank q has completed
  Scalars s;
};
static_assert(sizeof(Control) <= kCtrlBytes, "control block overflow");

// Everything a kernel needs to know about this rank and its peers; passed by value.
struct RankDev {
  Control* ctrl;
  Control* peer_ctrl[CGB_MAX_RANKS];  // peer control blocks as mapped in this process
  int nranks, rank;
  cgb_layout_t lay;
  cgb_stencil_t st;
  int y0, z0;                     // global offset of the owned box
  int nbr[CGB_NDIR];
  unsigned nbr_mask;              // bit d set if nbr[d] >= 0
  double* nbr_exch[2][CGB_NDIR];  // neighbours' exchange buffers (mapped)
  int nbr_ny[CGB_NDIR], nbr_nz[CGB_NDIR], nbr_py[CGB_NDIR];
  unsigned halo_expect[CGB_NDIR];  // units to count before the flag for direction d is raised
  double* exch[2];                 // my exchanged vectors (SpMV input, ping-pong)
  double *x, *r, *Ap, *b, *s, *z, *pp;
  double* hist;
  int hist_cap;
  double* hist_host;  // optional mapped pinned host buffer: ||r|| of every iteration is also stored there
  int hist_host_cap;
  // optional device buffer (persistent kernel): per iteration 4 x %globaltimer ns - last CTA reached barrier A,
  // alpha released, last CTA reached barrier B, beta released. Gives the compute / rendezvous split of an iteration.
  unsigned long long* trace;
  int trace_cap;      // iterations the buffer can hold
  int dbg_flags;      // CGB_DEBUG_FLAGS: bit 0 = no phase-A prologue prefetch in the persistent kernel
  double2* partials;
  // CSR problems
  const int* csr_rowptr;
  const int* csr_colind;
  const double* csr_val;
  int csr_avg_nnz;   // ceil(nnz/n): picks the lanes-per-row of k_spmv_csr
  const int* csr_rowblk;  // row blocks of the streaming CSR kernel: block b = rows [rowblk[b], rowblk[b+1]); NULL: none
  int csr_nblk;
  // row-block distributed CSR: remote entries of p are loaded from the owners' exchange buffers
  int csr_dist;                          // 1 if the columns span several ranks
  int csr_n_global;
  double* peer_exch0[CGB_MAX_RANKS];     // exchange buffer 0 (p) of every rank, as mapped here
  // SpMV work decomposition (host-chosen)
  int ntx, nty, nmc, mchunk;
};

// ---- system-scope flag primitives (the MPI_Sendrecv / MPI_Allreduce replacement) ----

__device__ __forceinline__ u64 ld_acquire_sys(const u64* p) {
  u64 v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(u64* p, u64 v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_f64(double* p, double v) {
  asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

__device__ __forceinline__ u64 ld_relaxed_sys_u64(const u64* p) {
  u64 v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_u64(u64* p, u64 v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- flag-in-data ("LL") messages for the all-reduce ----
__device__ __forceinline__ void ll_send2(u64* dst4, double a, double b, unsigned tag) {
  const u64 ua = (u64)__double_as_longlong(a), ub = (u64)__double_as_longlong(b), t = (u64)tag << 32;
  st_relaxed_sys_u64(dst4 + 0, t | (ua & 0xffffffffull));
  st_relaxed_sys_u64(dst4 + 1, t | (ua >> 32));
  st_relaxed_sys_u64(dst4 + 2, t | (ub & 0xffffffffull));
  st_relaxed_sys_u64(dst4 + 3, t | (ub >> 32));
}
__device__ __forceinline__ bool ll_recv2(const u64* src4, unsigned tag, long long timeout_cycles, double& a, double& b) {
  const long long t0 = clock64();
  for (;;) {
    const u64 w0 = ld_relaxed_sys_u64(src4 + 0), w1 = ld_relaxed_sys_u64(src4 + 1);
    const u64 w2 = ld_relaxed_sys_u64(src4 + 2), w3 = ld_relaxed_sys_u64(src4 + 3);
    if ((unsigned)(w0 >> 32) == tag && (unsigned)(w1 >> 32) == tag && (unsigned)(w2 >> 32) == tag &&
        (unsigned)(w3 >> 32) == tag) {
      a = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
      b = __longlong_as_double((long long)((w2 & 0xffffffffull) | (w3 << 32)));
      return true;
    }
    if (clock64() - t0 > timeout_cycles) return false;
  }
}

// Spin until *p >= target. Bounded: a dead or diverged peer produces an error, not a hang.
__device__ __forceinline__ bool spin_until_ge(const u64* p, u64 target, long long timeout_cycles) {
  if (ld_acquire_sys(p) >= target) return true;
  const long long t0 = clock64();
  for (;;) {
    if (ld_acquire_sys(p) >= target) return true;
    if (clock64() - t0 > timeout_cycles) return false;
    __nanosleep(20);
  }
}

// __syncthreads() is an ALIGNED barrier: each warp has to execute it converged. After a block that only some lanes
// run (if (threadIdx.x == 0) ..., if (lane == 0) ...) ptxas usually - but not always - reconverges the warp first;
// where it does not, the lanes that skipped the block run through the barrier ahead of the others (observed in
// the persistent CG kernel). block_sync() makes the reconvergence explicit.
__device__ __forceinline__ void block_sync() {
  __syncwarp();
  __syncthreads();
}

// ---- deterministic reductions ----

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over a CTA of NT threads in a fixed tree; result valid in thread 0. smem: >= 64 doubles.
template <int NT>
__device__ __forceinline__ void block_sum2_t(double& a, double& b, double* smem) {
  constexpr int NW = NT / 32;
  static_assert(NT % 32 == 0 && NW <= 32, "block_sum2_t: 1..32 full warps");
  a = warp_sum(a);
  b = warp_sum(b);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  block_sync();  // smem reuse guard
  if (l == 0) {
    smem[w] = a;
    smem[32 + w] = b;
  }
  block_sync();
  if (w == 0) {
    a = l < NW ? smem[l] : 0.0;
    b = l < NW ? smem[32 + l] : 0.0;
    a = warp_sum(a);
    b = warp_sum(b);
  }
}
__device__ __forceinline__ void block_sum2(double& a, double& b, double* smem) { block_sum2_t<kThreads>(a, b, smem); }

// Every CTA contributes (a,b); returns true in ALL threads of the CTA that arrived last, with
// the grid-wide sums (fixed order => run-to-run bitwise reproducible) in (a,b) of thread 0.
template <int NT>
__device__ __forceinline__ bool grid_sum2_last_t(double& a, double& b, double2* partials, unsigned int* ticket,
                                                 double* smem) {
  __shared__ int s_last;
  block_sum2_t<NT>(a, b, smem);
  if (threadIdx.x == 0) {
    __stcg(&partials[blockIdx.x], make_double2(a, b));
    __threadfence();
    unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  block_sync();
  if (!s_last) return false;
  __threadfence();
  double sa = 0.0, sb = 0.0;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += NT) {
    double2 v = __ldcg(&partials[i]);
    sa += v.x;
    sb += v.y;
  }
  block_sum2_t<NT>(sa, sb, smem);
  a = sa;
  b = sb;
  if (threadIdx.x == 0) *ticket = 0u;
  return true;
}
__device__ __forceinline__ bool grid_sum2_last(double& a, double& b, double2* partials, unsigned int* ticket,
                                               double* smem) {
  return grid_sum2_last_t<kThreads>(a, b, partials, ticket, smem);
}

// One-shot all-reduce run by the last CTA of a producing kernel: thread t stores this rank's
// partial into slot[seq%4][rank] of rank t and raises the matching flag, then waits for rank
// t's contribution to arrive here. Sums are taken in rank order, so every rank obtains
// bit-identical scalars (mandatory: all ranks must take the same convergence branch).
// (a,b) in: this rank's sums in thread 0. out: global sums in thread 0. Returns false on timeout.
__device__ __forceinline__ bool allreduce2_last_cta(const RankDev& P, double& a, double& b, double* smem) {
  Scalars& S = P.ctrl->s;
  if (P.nranks == 1) {
    if (threadIdx.x == 0) {
      S.last_sum[0] = a;
      S.last_sum[1] = b;
      S.red_seq += 1;
    }
    return true;
  }
  __shared__ int s_ok;
  const u64 seq = S.red_seq;
  const int slot = (int)(seq & (kRedSlots - 1));
  block_sync();
  if (threadIdx.x == 0) {
    smem[0] = a;
    smem[1] = b;
    s_ok = 1;
  }
  block_sync();
  const int t = threadIdx.x;
  const unsigned tag = (unsigned)(seq + 1);
  if (t < P.nranks) {
    ll_send2(&P.peer_ctrl[t]->red_ll[slot][P.rank][0], smem[0], smem[1], tag);
    double va, vb;
    if (!ll_recv2(&P.ctrl->red_ll[slot][t][0], tag, S.timeout_cycles, va, vb)) s_ok = 0;
    smem[16 + 2 * t] = va;
    smem[17 + 2 * t] = vb;
  }
  block_sync();
  if (threadIdx.x == 0) {
    double sa = 0.0, sb = 0.0;
    for (int q = 0; q < P.nranks; ++q) {
      sa += smem[16 + 2 * q];
      sb += smem[17 + 2 * q];
    }
    a = sa;
    b = sb;
    S.last_sum[0] = sa;
    S.last_sum[1] = sb;
    S.red_seq = seq + 1;
    if (!s_ok) S.done = CGB_ERR_TIMEOUT;
  }
  return s_ok != 0;
}

// Split-phase variant for pipelined CG: the last CTA of the producing kernel only PUBLISHES
// (no wait), so the following SpMV runs while the 16-byte messages cross NVLink ...
__device__ __forceinline__ void allreduce2_publish(const RankDev& P, double a, double b, double* smem) {
  Scalars& S = P.ctrl->s;
  const u64 seq = S.red_seq;
  const int slot = (int)(seq & (kRedSlots - 1));
  block_sync();
  if (threadIdx.x == 0) {
    smem[0] = a;
    smem[1] = b;
  }
  block_sync();
  const int t = threadIdx.x;
  if (t < P.nranks) ll_send2(&P.peer_ctrl[t]->red_ll[slot][P.rank][0], smem[0], smem[1], (unsigned)(seq + 1));
  block_sync();
  if (threadIdx.x == 0) {
    S.pub_seq = seq;
    S.red_seq = seq + 1;
  }
}

// ... and every CTA of the consuming kernel completes it: wait for all ranks' contributions of
// `seq`, sum in rank order (bit-identical everywhere). All threads return the sums.
__device__ __forceinline__ bool allreduce2_consume(const RankDev& P, u64 seq, double& a, double& b, double* smem) {
  __shared__ int s_cok;
  const int slot = (int)(seq & (kRedSlots - 1));
  if (threadIdx.x == 0) s_cok = 1;
  block_sync();
  const int t = threadIdx.x;
  if (t < P.nranks) {
    double va, vb;
    if (!ll_recv2(&P.ctrl->red_ll[slot][t][0], (unsigned)(seq + 1), P.ctrl->s.timeout_cycles, va, vb)) s_cok = 0;
    smem[2 * t] = va;
    smem[2 * t + 1] = vb;
  }
  block_sync();
  double sa = 0.0, sb = 0.0;
  for (int q = 0; q < P.nranks; ++q) {
    sa += smem[2 * q];
    sb += smem[2 * q + 1];
  }
  a = sa;
  b = sb;
  const bool ok = s_cok != 0;
  block_sync();
  return ok;
}

// What the last CTA does once the global r.r of an x/r update is known: record ||r||, apply the
// reference's stopping test (absolute ||r||_2 < tol, after the x/r update and before the p update,
// CUDACG.cu:333), compute beta (CUDACG.cu:339), bump the iteration counter and announce the halo
// version the following p-update will publish. Thread 0 only.
__device__ __forceinline__ void cg_after_rr(const RankDev& P, double rr) {
  Scalars& S = P.ctrl->s;
  const double rn = sqrt(rr);
  const int it = S.iter;
  if (it < P.hist_cap) P.hist[it] = rn;
  if (P.hist_host && it < P.hist_host_cap) P.hist_host[it] = rn;  // zero-copy D2H of this step's result
  S.rnorm = rn;
  S.iter = it + 1;
  if (!isfinite(rr)) {
    S.done = CGB_ERR_BREAKDOWN;
  } else if (rn < S.thresh) {
    S.done = 1;
  } else {
    S.beta = rr / S.rho;
    S.rho_prev = S.rho;
    S.rho = rr;
    if (it + 1 >= S.maxit) S.done = 2;
    else S.halo_seq += 1;
  }
}

// ---- halo signalling ----

// Called by all threads of a CTA after it stored `cnt[d]` units into neighbour d's ghost layer.
// The CTA that completes direction d raises the neighbour's flag with `version`.
__device__ __forceinline__ void halo_commit(const RankDev& P, const unsigned* cnt_smem, u64 version) {
  block_sync();  // counters complete
  unsigned any = 0;
#pragma unroll
  for (int d = 0; d < CGB_NDIR; ++d) any |= cnt_smem[d];
  if (!any) return;        // CTA-uniform: this CTA pushed nothing
  __threadfence_system();  // my remote stores are performed before the counters move
  block_sync();
  const int d = threadIdx.x;
  if (d < CGB_NDIR && d != 4 && cnt_smem[d] != 0 && P.nbr[d] >= 0) {
    Scalars& S = P.ctrl->s;
    __threadfence();
    unsigned old = atomicAdd(&S.ticket_halo[d], cnt_smem[d]);
    if (old + cnt_smem[d] == P.halo_expect[d]) {
      S.ticket_halo[d] = 0u;
      __threadfence_system();
      st_release_sys(&P.peer_ctrl[P.nbr[d]]->halo_flag[dir_opposite(d)], version);
    }
  }
}

// All threads of the CTA return once every existing neighbour's halo of `version` has landed.
__device__ __forceinline__ bool halo_wait_all(const RankDev& P, u64 version) {
  __shared__ int s_hok;
  if (threadIdx.x == 0) s_hok = 1;
  block_sync();
  const int d = threadIdx.x;
  if (d < CGB_NDIR && d != 4 && P.nbr[d] >= 0) {
    if (!spin_until_ge(&P.ctrl->halo_flag[d], version, P.ctrl->s.timeout_cycles)) {
      s_hok = 0;
      P.ctrl->s.done = CGB_ERR_TIMEOUT;
    }
  }
  block_sync();
  return s_hok != 0;
}

// ---- distributed CSR: "my part of p, version v, is complete" ----

// Owner rank and local index of global row/column c under the block distribution of block_range().
__device__ __forceinline__ int csr_owner(int n, int nranks, int c, int& local) {
  const int base = n / nranks, rem = n % nranks, cut = rem * (base + 1);
  int q;
  if (c < cut) {
    q = c / (base + 1);
    local = c - q * (base + 1);
  } else {
    q = rem + (c - cut) / base;
    local = c - cut - (q - rem) * base;
  }
  return q;
}
// All threads of the LAST CTA of a kernel that wrote p (every other CTA has passed its ticket).
__device__ __forceinline__ void ready_publish(const RankDev& P, u64 version) {
  __threadfence_system();
  if ((int)threadIdx.x < P.nranks) st_release_sys(&P.peer_ctrl[threadIdx.x]->ready_flag[P.rank], version);
}
// All threads of a CTA that is about to read remote entries of p.
__device__ __forceinline__ bool ready_wait(const RankDev& P, u64 version) {
  __shared__ int s_rok;
  if (threadIdx.x == 0) s_rok = 1;
  block_sync();
  if ((int)threadIdx.x < P.nranks && !spin_until_ge(&P.ctrl->ready_flag[threadIdx.x], version, P.ctrl->s.timeout_cycles)) {
    s_rok = 0;
    P.ctrl->s.done = CGB_ERR_TIMEOUT;
  }
  block_sync();
  return s_rok != 0;
}

// ---- elementwise unit mapping: unit u -> pair of x-cells in owned row (j,k) ----

struct UnitMap {
  int pw;   // pairs per row
  int ny;
};
__device__ __forceinline__ void unit_to_ijk(const UnitMap& m, unsigned u, int& i, int& j, int& k) {
  unsigned row = u / (unsigned)m.pw;
  i = 2 * (int)(u - row * (unsigned)m.pw);
  k = (int)(row / (unsigned)m.ny);
  j = (int)(row - (unsigned)k * (unsigned)m.ny);
}

// Directions a row (j,k) has to be pushed to, as a bit mask over d (before masking with nbr_mask).
__device__ __forceinline__ unsigned row_push_mask(const cgb_layout_t& L, int j, int k) {
  // dy candidates: bit0 = -1, bit1 = 0, bit2 = +1
  unsigned my = 2u | (L.gy && j == 0 ? 1u : 0u) | (L.gy && j == L.ny - 1 ? 4u : 0u);
  unsigned mz = 2u | (L.gz && k == 0 ? 1u : 0u) | (L.gz && k == L.nz - 1 ? 4u : 0u);
  unsigned m = 0;
#pragma unroll
  for (int dz = 0; dz < 3; ++dz)
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
      if ((my >> dy & 1u) && (mz >> dz & 1u)) m |= 1u << (dy + 3 * dz);
  return m & ~(1u << 4);
}

// Index of my owned row (j,k), cell i, inside neighbour d's ghost layer.
__device__ __forceinline__ int64_t ghost_index(const RankDev& P, int d, int i, int j, int k) {
  const int dy = dir_dy(d), dz = dir_dz(d);
  const int jn = dy < 0 ? P.nbr_ny[d] : (dy > 0 ? -1 : j);
  const int kn = dz < 0 ? P.nbr_nz[d] : (dz > 0 ? -1 : k);
  return (int64_t)i + (int64_t)P.lay.pitch * ((int64_t)(jn + P.lay.gy) + (int64_t)P.nbr_py[d] * (int64_t)(kn + P.lay.gz));
}

// ---- boundary-row enumeration for the "boundary first" pass of the kernels that write the SpMV input ----
__device__ __forceinline__ double2 ld2(const double* p) { return *reinterpret_cast<const double2*>(p); }
__device__ __forceinline__ void st2(double* p, double2 v) { *reinterpret_cast<double2*>(p) = v; }

struct BRows {
  int nA, nB, nC, nD, klo, total;
};
__device__ __forceinline__ BRows brows_make(const cgb_layout_t& L) {
  BRows B;
  B.nA = L.gz ? L.ny : 0;                  // plane k = 0
  B.nB = (L.gz && L.nz > 1) ? L.ny : 0;    // plane k = nz-1
  B.klo = L.gz ? 1 : 0;
  int khi = L.gz ? L.nz - 1 : L.nz;
  B.nC = L.gy ? max(0, khi - B.klo) : 0;   // rows j = 0 of the remaining planes
  B.nD = (L.gy && L.ny > 1) ? B.nC : 0;    // rows j = ny-1
  B.total = B.nA + B.nB + B.nC + B.nD;
  return B;
}
__device__ __forceinline__ void brow_to_jk(const BRows& B, const cgb_layout_t& L, int b, int& j, int& k) {
  if (b < B.nA) { j = b; k = 0; return; }
  b -= B.nA;
  if (b < B.nB) { j = b; k = L.nz - 1; return; }
  b -= B.nB;
  if (b < B.nC) { j = 0; k = B.klo + b; return; }
  b -= B.nC;
  j = L.ny - 1;
  k = B.klo + b;
}

// Store a pair into every neighbour ghost layer the row belongs to; count units per direction.
__device__ __forceinline__ void push_pair(const RankDev& P, int exch_buf, unsigned mask, int i, int j, int k,
                                          double2 v, unsigned* cnt) {
  while (mask) {
    int d = __ffs(mask) - 1;
    mask &= mask - 1;
    st2(P.nbr_exch[exch_buf][d] + ghost_index(P, d, i, j, k), v);
    atomicAdd(&cnt[d], 1u);
  }
}

}  // namespace cgb
