// Per-rank GPU context: owns the peer-mappable arena (control block, reduction slots, exchanged
// vectors), the rank-local work vectors, the stream/events and the captured CUDA graphs.
// RAII replacement of the reference's pointer list + CLEANUP macro (CUDACG.cu:10-33, 43-72).
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <vector>

#include "../kernels/cg_kernels.cuh"
#include "cgb200.h"

struct cgb_ctx {
  int device = 0;
  cgb_stencil_t st{};
  cgb_partition_t part{};
  cgb_layout_t lay{};
  int hist_cap = 0;

  void* arena = nullptr;      // ctrl | partials | exch0 | exch1   (peer visible)
  size_t arena_bytes = 0;
  void* local = nullptr;      // x | r | Ap | b | hist              (rank local)
  size_t local_bytes = 0;
  void* extra = nullptr;      // s | z | pp                         (pipelined CG, on demand)
  double* compact_tmp = nullptr;
  int* d_rowptr = nullptr;    // CSR problems
  int* d_colind = nullptr;
  double* d_val = nullptr;
  int* d_rowblk = nullptr;    // row blocks of the streaming CSR SpMV (single-rank CSR problems)
  long long csr_nnz = 0;

  void* peer_base[CGB_MAX_RANKS] = {};
  bool peer_is_ipc[CGB_MAX_RANKS] = {};
  bool peers_dirty = true;

  cgb::RankDev P{};
  cgb::LaunchCfg cfg{};

  cudaStream_t own_stream = nullptr;
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_poll[2] = {nullptr, nullptr}, ev_poll_nb = nullptr;
  cgb::Scalars* h_scal = nullptr;  // pinned: 2 polling slots of cgb_solve + 1 of the non-blocking cgb_solve_poll
  int poll_slot = 0;
  cgb::Scalars poll_snapshot{};    // newest completed state seen by cgb_solve_poll
  bool poll_pending = false;       // a non-blocking poll copy is in flight (h_scal[2], ev_poll_nb)

  // per-solve state
  cgb_solve_opts_t opts{};
  bool active = false;
  bool use_persist = false;  // this solve runs on the persistent cooperative kernel (merged variant)
  cudaStream_t stream = nullptr;
  int enqueued = 0;
  int host_cur = 0;      // exchange buffer holding the SpMV input of the next enqueued iteration
  int iters_target = 0;  // iterations the host must enqueue for a full-length solve
  long long launches = 0;
  long long launches_at_begin = 0;
  std::map<long long, cudaGraphExec_t> graphs;  // key: variant/impl/length
  std::vector<cudaGraph_t> while_graphs;        // templates of the conditional-WHILE graphs (own the handles)
  alignas(64) unsigned char tmap[2][128] = {};  // CUtensorMap of exch[0], exch[1] (spmv_tma.cu)
  alignas(64) unsigned char tmap_x[128] = {};   // band-row maps of x and r (update mode)
  alignas(64) unsigned char tmap_r[128] = {};
  alignas(64) unsigned char tmap_rh[128] = {};  // r with halo rows (merged variant)
  alignas(64) unsigned char tmap_h[2][128] = {};  // 2-column x-halo boxes of exch[0], exch[1] (tiled rows)
  alignas(64) unsigned char tmap_rh_h[128] = {};  // ... and of r
  void* d_maps = nullptr;  // the same eight maps in device memory: descriptor source of the persistent kernel
  bool tma_ok = false;

  ~cgb_ctx();
};

namespace cgb {
int ctx_finalize_peers(cgb_ctx* c);
int ctx_enqueue_iteration(cgb_ctx* c, cudaStream_t s);  // one CG iteration, plain launches
void ctx_drop_graphs(cgb_ctx* c);
}  // namespace cgb
