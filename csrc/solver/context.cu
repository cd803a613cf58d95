// Context life cycle, fabric bootstrap and the solve drivers.
//
// Layers L1/L3/L4/L5 of the reference (CUDACG.cu:10-33 cleanup, :87-186 staging, :189-233
// handles, :244-352 solver) re-designed: everything is allocated once, the iteration runs from
// captured CUDA graphs with all scalars device-resident, and the only host involvement is a poll
// of the device convergence flag every `chunk` iterations, one chunk behind the GPU.
#include "context.hpp"

#include <nvtx3/nvToolsExt.h>  // header-only; ranges show up in Nsight timelines

#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../common/status.hpp"
#include "../topo/partition.hpp"

using namespace cgb;

namespace {

// Peer waits give up after ~10 s (at 2 GHz); CGB_TIMEOUT_MS overrides (tests use short values).
long long default_timeout_cycles() {
  const char* e = std::getenv("CGB_TIMEOUT_MS");
  const double ms = e ? std::atof(e) : 10000.0;
  return (long long)(ms * 2.0e6);
}
#define kDefaultTimeoutCycles default_timeout_cycles()

struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

cudaStream_t pick_stream(cgb_ctx* c, void* s) { return s ? (cudaStream_t)s : c->own_stream; }

int alloc_common(cgb_ctx* c) {
  CGB_CUDA(cudaSetDevice(c->device));
  const size_t vec_bytes = (size_t)c->lay.vec_elems * sizeof(double);
  c->arena_bytes = round_up(kExchOff + 2 * vec_bytes, 1 << 21);
  CGB_CUDA(cudaMalloc(&c->arena, c->arena_bytes));
  CGB_CUDA(cudaMemset(c->arena, 0, c->arena_bytes));
  const size_t hist_bytes = round_up((size_t)(c->hist_cap + 2) * sizeof(double), 256);
  c->local_bytes = 4 * vec_bytes + hist_bytes;
  CGB_CUDA(cudaMalloc(&c->local, c->local_bytes));
  CGB_CUDA(cudaMemset(c->local, 0, c->local_bytes));
  CGB_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  CGB_CUDA(cudaEventCreate(&c->ev_begin));
  CGB_CUDA(cudaEventCreate(&c->ev_end));
  CGB_CUDA(cudaEventCreateWithFlags(&c->ev_poll[0], cudaEventDisableTiming));
  CGB_CUDA(cudaEventCreateWithFlags(&c->ev_poll[1], cudaEventDisableTiming));
  CGB_CUDA(cudaEventCreateWithFlags(&c->ev_poll_nb, cudaEventDisableTiming));
  CGB_CUDA(cudaMallocHost(&c->h_scal, 3 * sizeof(Scalars)));

  RankDev& P = c->P;
  P = RankDev{};
  char* a = (char*)c->arena;
  P.ctrl = (Control*)a;
  P.partials = (double2*)(a + kPartialsOff);
  P.exch[0] = (double*)(a + kExchOff);
  P.exch[1] = (double*)(a + kExchOff + vec_bytes);
  char* l = (char*)c->local;
  P.x = (double*)l;
  P.r = (double*)(l + vec_bytes);
  P.Ap = (double*)(l + 2 * vec_bytes);
  P.b = (double*)(l + 3 * vec_bytes);
  P.hist = (double*)(l + 4 * vec_bytes);
  P.hist_cap = c->hist_cap;
  P.nranks = c->part.nranks;
  P.rank = c->part.rank;
  if (const char* e = std::getenv("CGB_DEBUG_FLAGS")) P.dbg_flags = std::atoi(e);
  P.lay = c->lay;
  P.st = c->st;
  P.y0 = c->part.y0;
  P.z0 = c->part.z0;
  for (int d = 0; d < CGB_NDIR; ++d) P.nbr[d] = -1;
  c->peer_base[c->part.rank] = c->arena;
  CGB_TRY(query_launch_cfg(c->device, &c->cfg));
  if (c->st.kind != CGB_KIND_CSR) choose_spmv_tiling(P, c->cfg.grid_spmv);
  c->tma_ok = false;
  if (spmv_tma_supported(P)) {
    int r0 = spmv_tma_make_map(c->st.kind, c->lay, P.exch[0], c->tmap[0]);
    int r1 = spmv_tma_make_map(c->st.kind, c->lay, P.exch[1], c->tmap[1]);
    int r2 = spmv_tma_make_map(c->st.kind, c->lay, P.x, c->tmap_x, 0);
    int r3 = spmv_tma_make_map(c->st.kind, c->lay, P.r, c->tmap_r, 0);
    int r4 = spmv_tma_make_map(c->st.kind, c->lay, P.r, c->tmap_rh, 1);
    int r5 = spmv_tma_make_map(c->st.kind, c->lay, P.exch[0], c->tmap_h[0], 1, 2);
    int r6 = spmv_tma_make_map(c->st.kind, c->lay, P.exch[1], c->tmap_h[1], 1, 2);
    int r7 = spmv_tma_make_map(c->st.kind, c->lay, P.r, c->tmap_rh_h, 1, 2);
    c->tma_ok = (r0 | r1 | r2 | r3 | r4 | r5 | r6 | r7) == CGB_OK;
    if (c->tma_ok) {  // the persistent kernel reads its descriptors from device memory
      CGB_CUDA(cudaMalloc(&c->d_maps, cg_persist_maps_bytes()));
      const void* maps[8] = {c->tmap_rh, c->tmap[0], c->tmap[1], c->tmap_rh_h, c->tmap_h[0], c->tmap_h[1], c->tmap_x, c->tmap_r};
      CGB_CUDA(cg_persist_upload_maps(c->d_maps, maps));
    }
  }
  // timeout for peer waits lives in the control block and survives solves
  Scalars init{};
  init.timeout_cycles = kDefaultTimeoutCycles;
  CGB_CUDA(cudaMemcpy(&P.ctrl->s, &init, sizeof(Scalars), cudaMemcpyHostToDevice));
  c->peers_dirty = true;
  return CGB_OK;
}

}  // namespace

cgb_ctx::~cgb_ctx() {
  cudaSetDevice(device);
  cgb::ctx_drop_graphs(this);
  for (int q = 0; q < CGB_MAX_RANKS; ++q)
    if (peer_base[q] && peer_is_ipc[q]) cudaIpcCloseMemHandle(peer_base[q]);
  if (arena) cudaFree(arena);
  if (local) cudaFree(local);
  if (extra) cudaFree(extra);
  if (d_maps) cudaFree(d_maps);
  if (compact_tmp) cudaFree(compact_tmp);
  if (d_rowptr) cudaFree(d_rowptr);
  if (d_colind) cudaFree(d_colind);
  if (d_val) cudaFree(d_val);
  if (d_rowblk) cudaFree(d_rowblk);
  if (h_scal) cudaFreeHost(h_scal);
  if (ev_begin) cudaEventDestroy(ev_begin);
  if (ev_end) cudaEventDestroy(ev_end);
  if (ev_poll[0]) cudaEventDestroy(ev_poll[0]);
  if (ev_poll[1]) cudaEventDestroy(ev_poll[1]);
  if (ev_poll_nb) cudaEventDestroy(ev_poll_nb);
  if (own_stream) cudaStreamDestroy(own_stream);
}

namespace cgb {

void ctx_drop_graphs(cgb_ctx* c) {
  for (auto& kv : c->graphs) cudaGraphExecDestroy(kv.second);
  c->graphs.clear();
  for (auto g : c->while_graphs) cudaGraphDestroy(g);
  c->while_graphs.clear();
}

// Resolve neighbour pointers/layouts from the attached peer arenas.
int ctx_finalize_peers(cgb_ctx* c) {
  if (!c->peers_dirty) return CGB_OK;
  RankDev& P = c->P;
  P.nbr_mask = 0;
  for (int q = 0; q < c->part.nranks; ++q) {
    if (!c->peer_base[q]) return fail(CGB_ERR_STATE, "rank %d: peer %d not attached", c->part.rank, q);
    P.peer_ctrl[q] = (Control*)c->peer_base[q];
    P.peer_exch0[q] = (double*)((char*)c->peer_base[q] + kExchOff);
  }
  const int pw = (c->lay.nx + 1) / 2;
  for (int d = 0; d < CGB_NDIR; ++d) {
    int q = c->part.nbr[d];
    P.nbr[d] = q;
    P.halo_expect[d] = 0;
    if (q < 0) continue;
    cgb_partition_t np;
    cgb_layout_t nl;
    CGB_TRY(make_partition(c->st, c->part.nranks, q, 0, c->part.Py, c->part.Pz, &np));
    CGB_TRY(make_layout(c->st, np, &nl));
    char* base = (char*)c->peer_base[q];
    P.nbr_exch[0][d] = (double*)(base + kExchOff);
    P.nbr_exch[1][d] = (double*)(base + kExchOff + (size_t)nl.vec_elems * sizeof(double));
    P.nbr_ny[d] = nl.ny;
    P.nbr_nz[d] = nl.nz;
    P.nbr_py[d] = nl.py;
    const int dy = dir_dy(d), dz = dir_dz(d);
    P.halo_expect[d] = (unsigned)((dy == 0 ? c->lay.ny : 1) * (dz == 0 ? c->lay.nz : 1)) * (unsigned)pw;
    P.nbr_mask |= 1u << d;
  }
  ctx_drop_graphs(c);
  c->peers_dirty = false;
  return CGB_OK;
}

}  // namespace cgb

// ---------------------------------------------------------------------------------------------
// creation / destruction
// ---------------------------------------------------------------------------------------------
extern "C" int cgb_ctx_create(const cgb_stencil_t* st, const cgb_partition_t* part, int device, int hist_cap,
                              cgb_ctx** out) {
  CGB_REQUIRE(st && part && out, "cgb_ctx_create: NULL argument");
  CGB_REQUIRE(kind_is_stencil(st->kind), "cgb_ctx_create: use cgb_ctx_create_csr for CSR problems");
  CGB_REQUIRE(hist_cap >= 1, "hist_cap must be >= 1");
  cgb_ctx* c = new cgb_ctx();
  c->device = device;
  c->st = *st;
  c->part = *part;
  c->hist_cap = hist_cap;
  int rc = make_layout(*st, *part, &c->lay);
  if (rc == 0) rc = alloc_common(c);
  if (rc != 0) {
    delete c;
    return rc;
  }
  *out = c;
  return CGB_OK;
}

// Host-side sanity of a CSR triple before it reaches the device: a bad offset or column index would make the
// SpMV read out of bounds (over peer mappings in the distributed case).
static int validate_csr(int64_t nrows, int64_t ncols, int64_t nnz, const int32_t* rowptr, const int32_t* colind) {
  CGB_REQUIRE(rowptr[0] == 0, "CSR: rowptr[0] must be 0 (got %d)", rowptr[0]);
  for (int64_t r = 0; r < nrows; ++r)
    CGB_REQUIRE(rowptr[r + 1] >= rowptr[r], "CSR: rowptr is not monotone at row %lld", (long long)r);
  CGB_REQUIRE((int64_t)rowptr[nrows] == nnz, "CSR: rowptr[n] = %d does not match nnz = %lld", rowptr[nrows], (long long)nnz);
  for (int64_t e = 0; e < nnz; ++e)
    CGB_REQUIRE(colind[e] >= 0 && (int64_t)colind[e] < ncols, "CSR: column index %d at entry %lld is outside [0,%lld)",
                colind[e], (long long)e, (long long)ncols);
  return CGB_OK;
}

extern "C" int cgb_ctx_create_csr(int32_t n, int64_t nnz, const int32_t* rowptr, const int32_t* colind,
                                  const double* val, int device, int hist_cap, cgb_ctx** out) {
  CGB_REQUIRE(n > 0 && nnz > 0 && rowptr && colind && val && out, "cgb_ctx_create_csr: bad argument");
  CGB_REQUIRE(nnz < INT_MAX, "int32 row offsets: nnz must be < 2^31");
  CGB_REQUIRE(hist_cap >= 1, "hist_cap must be >= 1");
  CGB_TRY(validate_csr(n, n, nnz, rowptr, colind));
  cgb_ctx* c = new cgb_ctx();
  c->device = device;
  c->hist_cap = hist_cap;
  int rc = cgb_stencil_default(CGB_KIND_CSR, n, 1, 1, &c->st);
  if (rc == 0) rc = make_partition(c->st, 1, 0, 1, 0, 0, &c->part);
  if (rc == 0) rc = make_layout(c->st, c->part, &c->lay);
  if (rc == 0) rc = alloc_common(c);
  auto stage = [&]() -> int {  // CUDACG.cu:120-134
    CGB_CUDA(cudaMalloc(&c->d_rowptr, (size_t)(n + 1) * sizeof(int)));
    CGB_CUDA(cudaMalloc(&c->d_colind, (size_t)nnz * sizeof(int)));
    CGB_CUDA(cudaMalloc(&c->d_val, (size_t)nnz * sizeof(double)));
    CGB_CUDA(cudaMemcpy(c->d_rowptr, rowptr, (size_t)(n + 1) * sizeof(int), cudaMemcpyHostToDevice));
    CGB_CUDA(cudaMemcpy(c->d_colind, colind, (size_t)nnz * sizeof(int), cudaMemcpyHostToDevice));
    CGB_CUDA(cudaMemcpy(c->d_val, val, (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice));
    return CGB_OK;
  };
  if (rc == 0) rc = stage();
  if (rc != 0) {
    delete c;
    return rc;
  }
  c->csr_nnz = nnz;
  c->P.csr_rowptr = c->d_rowptr;
  c->P.csr_colind = c->d_colind;
  c->P.csr_val = c->d_val;
  c->P.csr_avg_nnz = (int)((nnz + n - 1) / n);
  {  // row blocks of the streaming SpMV kernel
    const std::vector<int> blk = csr_row_blocks(n, rowptr);
    auto up = [&]() -> int {
      CGB_CUDA(cudaMalloc(&c->d_rowblk, blk.size() * sizeof(int)));
      CGB_CUDA(cudaMemcpy(c->d_rowblk, blk.data(), blk.size() * sizeof(int), cudaMemcpyHostToDevice));
      return CGB_OK;
    };
    rc = up();
    if (rc != 0) {
      delete c;
      return rc;
    }
    c->P.csr_rowblk = c->d_rowblk;
    c->P.csr_nblk = (int)blk.size() - 1;
  }
  *out = c;
  return CGB_OK;
}

extern "C" int cgb_ctx_create_csr_dist(int32_t n_global, const cgb_partition_t* part, int64_t nnz, const int32_t* rowptr,
                                       const int32_t* colind, const double* val, int device, int hist_cap,
                                       cgb_ctx** out) {
  CGB_REQUIRE(n_global > 0 && part && nnz > 0 && rowptr && colind && val && out, "cgb_ctx_create_csr_dist: bad argument");
  CGB_REQUIRE(nnz < INT_MAX, "int32 row offsets: nnz must be < 2^31");
  CGB_REQUIRE(part->nx >= 1 && part->x0 >= 0 && part->x0 + part->nx <= n_global, "row block out of range");
  const int n = part->nx;
  CGB_REQUIRE(hist_cap >= 1, "hist_cap must be >= 1");
  CGB_TRY(validate_csr(n, n_global, nnz, rowptr, colind));
  cgb_ctx* c = new cgb_ctx();
  c->device = device;
  c->hist_cap = hist_cap;
  int rc = cgb_stencil_default(CGB_KIND_CSR, n_global, 1, 1, &c->st);
  c->part = *part;
  if (rc == 0) rc = make_layout(c->st, c->part, &c->lay);
  if (rc == 0) rc = alloc_common(c);
  auto stage = [&]() -> int {
    CGB_CUDA(cudaMalloc(&c->d_rowptr, (size_t)(n + 1) * sizeof(int)));
    CGB_CUDA(cudaMalloc(&c->d_colind, (size_t)nnz * sizeof(int)));
    CGB_CUDA(cudaMalloc(&c->d_val, (size_t)nnz * sizeof(double)));
    CGB_CUDA(cudaMemcpy(c->d_rowptr, rowptr, (size_t)(n + 1) * sizeof(int), cudaMemcpyHostToDevice));
    CGB_CUDA(cudaMemcpy(c->d_colind, colind, (size_t)nnz * sizeof(int), cudaMemcpyHostToDevice));
    CGB_CUDA(cudaMemcpy(c->d_val, val, (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice));
    return CGB_OK;
  };
  if (rc == 0) rc = stage();
  if (rc != 0) {
    delete c;
    return rc;
  }
  c->csr_nnz = nnz;
  c->P.csr_rowptr = c->d_rowptr;
  c->P.csr_colind = c->d_colind;
  c->P.csr_val = c->d_val;
  c->P.csr_avg_nnz = (int)((nnz + n - 1) / n);
  c->P.csr_dist = part->nranks > 1 ? 1 : 0;
  c->P.csr_n_global = n_global;
  *out = c;
  return CGB_OK;
}

extern "C" int cgb_ctx_destroy(cgb_ctx* c) {
  delete c;
  return CGB_OK;
}

// ---------------------------------------------------------------------------------------------
// fabric
// ---------------------------------------------------------------------------------------------
extern "C" int cgb_ctx_ipc_export(cgb_ctx* c, void* handle_out) {
  CGB_REQUIRE(c && handle_out, "cgb_ctx_ipc_export: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == CGB_IPC_HANDLE_BYTES, "IPC handle size");
  CGB_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  CGB_CUDA(cudaIpcGetMemHandle(&h, c->arena));
  std::memcpy(handle_out, &h, sizeof(h));
  return CGB_OK;
}

extern "C" int cgb_ctx_peer_attach_ipc(cgb_ctx* c, int peer_rank, const void* handle) {
  CGB_REQUIRE(c && handle, "cgb_ctx_peer_attach_ipc: NULL argument");
  CGB_REQUIRE(peer_rank >= 0 && peer_rank < c->part.nranks, "peer rank %d out of range", peer_rank);
  if (peer_rank == c->part.rank) return CGB_OK;
  CGB_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  CGB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  c->peer_base[peer_rank] = p;
  c->peer_is_ipc[peer_rank] = true;
  c->peers_dirty = true;
  return CGB_OK;
}

extern "C" int cgb_ctx_peer_attach_ptr(cgb_ctx* c, int peer_rank, void* base) {
  CGB_REQUIRE(c && base, "cgb_ctx_peer_attach_ptr: NULL argument");
  CGB_REQUIRE(peer_rank >= 0 && peer_rank < c->part.nranks, "peer rank %d out of range", peer_rank);
  if (peer_rank == c->part.rank) return CGB_OK;
  CGB_CUDA(cudaSetDevice(c->device));
  cudaPointerAttributes at{};
  CGB_CUDA(cudaPointerGetAttributes(&at, base));
  if (at.type == cudaMemoryTypeDevice && at.device != c->device) {
    cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
      cudaGetLastError();
    } else if (e != cudaSuccess) {
      return fail(CGB_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d->%d): %s", c->device, at.device,
                  cudaGetErrorString(e));
    }
  }
  c->peer_base[peer_rank] = base;
  c->peer_is_ipc[peer_rank] = false;
  c->peers_dirty = true;
  return CGB_OK;
}

extern "C" void* cgb_ctx_arena_base(cgb_ctx* c) { return c ? c->arena : nullptr; }
extern "C" size_t cgb_ctx_arena_bytes(cgb_ctx* c) { return c ? c->arena_bytes : 0; }

extern "C" int cgb_ctx_fabric_reset(cgb_ctx* c) {
  CGB_REQUIRE(c, "cgb_ctx_fabric_reset: NULL context");
  CGB_CUDA(cudaSetDevice(c->device));
  CGB_CUDA(cudaDeviceSynchronize());
  CGB_CUDA(cudaMemset(c->arena, 0, kCtrlBytes));
  Scalars init{};
  init.timeout_cycles = kDefaultTimeoutCycles;
  CGB_CUDA(cudaMemcpy(&c->P.ctrl->s, &init, sizeof(Scalars), cudaMemcpyHostToDevice));
  // ghost layers must read as zero Dirichlet data until a neighbour pushes
  CGB_CUDA(cudaMemset((char*)c->arena + kExchOff, 0, 2 * (size_t)c->lay.vec_elems * sizeof(double)));
  c->active = false;
  return CGB_OK;
}

// ---------------------------------------------------------------------------------------------
// data movement
// ---------------------------------------------------------------------------------------------
extern "C" int cgb_ctx_set_rhs_generated(cgb_ctx* c, int rhs_kind, uint64_t seed, void* stream) {
  CGB_REQUIRE(c, "cgb_ctx_set_rhs_generated: NULL context");
  CGB_REQUIRE(kind_is_stencil(c->st.kind), "generated right-hand sides need a stencil problem");
  CGB_REQUIRE(rhs_kind >= CGB_RHS_ONES && rhs_kind <= CGB_RHS_MANUFACTURED, "unknown rhs kind %d", rhs_kind);
  CGB_CUDA(cudaSetDevice(c->device));
  CGB_CUDA(launch_rhs_gen(c->P, c->cfg, rhs_kind, seed, pick_stream(c, stream)));
  c->launches += 1;
  if (!stream) CGB_CUDA(cudaStreamSynchronize(c->own_stream));  // private stream: make it visible
  return CGB_OK;
}

// z, s, p of pipelined CG (three more rank-local vectors), allocated on first use.
static int ensure_extra(cgb_ctx* c) {
  if (c->extra) return CGB_OK;
  const size_t vec_bytes = (size_t)c->lay.vec_elems * sizeof(double);
  CGB_CUDA(cudaMalloc(&c->extra, 3 * vec_bytes));
  CGB_CUDA(cudaMemset(c->extra, 0, 3 * vec_bytes));
  char* e = (char*)c->extra;
  c->P.z = (double*)e;
  c->P.s = (double*)(e + vec_bytes);
  c->P.pp = (double*)(e + 2 * vec_bytes);
  cgb::ctx_drop_graphs(c);  // captured kernels hold a copy of P
  return CGB_OK;
}

static int ensure_compact_tmp(cgb_ctx* c) {
  if (!c->compact_tmp) CGB_CUDA(cudaMalloc(&c->compact_tmp, (size_t)c->lay.owned * sizeof(double)));
  return CGB_OK;
}

extern "C" int cgb_ctx_set_rhs(cgb_ctx* c, const double* src, int src_is_device, void* stream) {
  CGB_REQUIRE(c && src, "cgb_ctx_set_rhs: NULL argument");
  CGB_CUDA(cudaSetDevice(c->device));
  cudaStream_t s = pick_stream(c, stream);
  const double* dsrc = src;
  if (!src_is_device) {  // CUDACG.cu:143-153
    CGB_TRY(ensure_compact_tmp(c));
    CGB_CUDA(cudaMemcpyAsync(c->compact_tmp, src, (size_t)c->lay.owned * sizeof(double), cudaMemcpyHostToDevice, s));
    dsrc = c->compact_tmp;
  }
  CGB_CUDA(launch_pad(c->lay, dsrc, c->P.b, s));
  c->launches += 1;
  if (!stream) CGB_CUDA(cudaStreamSynchronize(c->own_stream));
  return CGB_OK;
}

extern "C" int cgb_ctx_set_x0(cgb_ctx* c, const double* src, int src_is_device, void* stream) {
  CGB_REQUIRE(c && src, "cgb_ctx_set_x0: NULL argument");
  CGB_CUDA(cudaSetDevice(c->device));
  cudaStream_t s = pick_stream(c, stream);
  const double* dsrc = src;
  if (!src_is_device) {
    CGB_TRY(ensure_compact_tmp(c));
    CGB_CUDA(cudaMemcpyAsync(c->compact_tmp, src, (size_t)c->lay.owned * sizeof(double), cudaMemcpyHostToDevice, s));
    dsrc = c->compact_tmp;
  }
  CGB_CUDA(launch_pad(c->lay, dsrc, c->P.x, s));
  c->launches += 1;
  if (!stream) CGB_CUDA(cudaStreamSynchronize(c->own_stream));
  return CGB_OK;
}

extern "C" int cgb_ctx_get_x(cgb_ctx* c, double* dst, int dst_is_device, void* stream) {
  CGB_REQUIRE(c && dst, "cgb_ctx_get_x: NULL argument");
  CGB_CUDA(cudaSetDevice(c->device));
  cudaStream_t s = pick_stream(c, stream);
  if (dst_is_device) {
    CGB_CUDA(launch_unpad(c->lay, c->P.x, dst, s));
    if (!stream) CGB_CUDA(cudaStreamSynchronize(c->own_stream));
  } else {  // CUDACG.cu:355
    CGB_TRY(ensure_compact_tmp(c));
    CGB_CUDA(launch_unpad(c->lay, c->P.x, c->compact_tmp, s));
    CGB_CUDA(cudaMemcpyAsync(dst, c->compact_tmp, (size_t)c->lay.owned * sizeof(double), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaStreamSynchronize(s));
  }
  c->launches += 1;
  return CGB_OK;
}

extern "C" int cgb_ctx_get_history(cgb_ctx* c, double* dst, int n) {
  CGB_REQUIRE(c && dst && n >= 0 && n <= c->hist_cap, "cgb_ctx_get_history: bad argument");
  CGB_CUDA(cudaSetDevice(c->device));
  if (n > 0) CGB_CUDA(cudaMemcpy(dst, c->P.hist, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
  return CGB_OK;
}

extern "C" int cgb_ctx_set_host_history(cgb_ctx* c, double* pinned, int cap) {
  CGB_REQUIRE(c, "cgb_ctx_set_host_history: NULL context");
  CGB_CUDA(cudaSetDevice(c->device));
  double* dev = nullptr;
  if (pinned) {
    CGB_REQUIRE(cap > 0, "cgb_ctx_set_host_history: cap must be positive");
    cudaError_t e = cudaHostGetDevicePointer((void**)&dev, pinned, 0);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(CGB_ERR_INVALID, "cgb_ctx_set_host_history: buffer is not mapped pinned host memory (%s)",
                  cudaGetErrorString(e));
    }
  }
  if (dev != c->P.hist_host || cap != c->P.hist_host_cap) {
    c->P.hist_host = dev;
    c->P.hist_host_cap = pinned ? cap : 0;
    cgb::ctx_drop_graphs(c);  // captured kernels hold a copy of P
  }
  return CGB_OK;
}

extern "C" int cgb_ctx_set_trace(cgb_ctx* c, unsigned long long* dev_buf, int cap_iters) {
  CGB_REQUIRE(c, "cgb_ctx_set_trace: NULL context");
  CGB_REQUIRE(!dev_buf || cap_iters > 0, "cgb_ctx_set_trace: cap must be positive");
  c->P.trace = dev_buf;
  c->P.trace_cap = dev_buf ? cap_iters : 0;
  cgb::ctx_drop_graphs(c);  // captured kernels hold a copy of P
  return CGB_OK;
}

// Device-resident solver state to stderr (post-mortem of a CGB_ERR_TIMEOUT; synchronises the device).
extern "C" int cgb_ctx_debug_print(cgb_ctx* c) {
  CGB_REQUIRE(c, "cgb_ctx_debug_print: NULL context");
  CGB_CUDA(cudaSetDevice(c->device));
  Scalars s;
  CGB_CUDA(cudaMemcpy(&s, &c->P.ctrl->s, sizeof(Scalars), cudaMemcpyDeviceToHost));
  std::fprintf(stderr,
               "[cgb rank %d] iter=%d done=%d maxit=%d rho=%.6e alpha=%.6e beta=%.6e rnorm=%.6e red_seq=%llu halo_seq=%llu "
               "gbar=%llu sbar=%llu ticket=%u dbg={site=%d block=%d thread=%d bar=%d parity=%d gbar=%d sbar=%d %d}\n",
               c->part.rank, s.iter, s.done, s.maxit, s.rho, s.alpha, s.beta, s.rnorm, s.red_seq, s.halo_seq, s.gbar, s.sbar,
               s.ticket_red, s.dbg[0], s.dbg[1], s.dbg[2], s.dbg[3], s.dbg[4], s.dbg[5], s.dbg[6], s.dbg[7]);
  return CGB_OK;
}

extern "C" int64_t cgb_ctx_launch_count(cgb_ctx* c) { return c ? c->launches : 0; }

// ---------------------------------------------------------------------------------------------
// solve drivers
// ---------------------------------------------------------------------------------------------
extern "C" void cgb_solve_opts_default(cgb_solve_opts_t* o) {
  if (!o) return;
  std::memset(o, 0, sizeof(*o));
  o->tol = 1e-7;                 // CUDACG.cu:245
  o->tol_mode = CGB_TOL_ABS;     // CUDACG.cu:333
  o->maxit = 2000;               // CUDACG.cu:244
  o->variant = CGB_VAR_AUTO;
  o->spmv_impl = CGB_SPMV_AUTO;
  o->use_graph = 1;
  o->chunk = 25;
  o->persistent = 0;
}

namespace cgb {

// variants whose SpMV input ping-pongs between the two exchange buffers
static bool uses_parity(const cgb_ctx* c) {
  return c->opts.variant == CGB_VAR_PIPELINED || c->opts.variant == CGB_VAR_MERGED2;
}
static bool merged2_ok(const cgb_ctx* c) { return c->tma_ok && c->part.Py == 1; }

// kernels per iteration of a variant (for the launch counter)
static int launches_per_iter(const cgb_ctx* c) {
  switch (c->opts.variant) {
    case CGB_VAR_FUSED: return 3;
    case CGB_VAR_PIPELINED: return 2;
    case CGB_VAR_MERGED2: return 2;
    default: return 3;
  }
}

// AUTO: the TMA band kernel for the 7-point stencil (measured faster, profiles/), the L1-path
// kernel for 27-point (the band kernel is issue-bound there) and for shapes it does not cover.
static bool use_tma(const cgb_ctx* c) {
  if (!c->tma_ok || c->opts.spmv_impl == CGB_SPMV_SIMPLE) return false;
  if (c->opts.spmv_impl == CGB_SPMV_TMA) return true;
  return c->st.kind == CGB_KIND_POISSON3D7 || c->st.kind == CGB_KIND_POISSON2D5;
}

static int enqueue_spmv(cgb_ctx* c, int buf, int mode, cudaStream_t s) {
  const RankDev& P = c->P;
  if (c->st.kind == CGB_KIND_CSR) CGB_CUDA(launch_spmv_csr(P, c->cfg, P.exch[buf], P.Ap, mode, s));
  else if (use_tma(c)) CGB_CUDA(launch_spmv_tma(P, c->device, c->tmap[buf], c->tmap_h[buf], nullptr, nullptr, P.Ap, mode, s));
  else CGB_CUDA(launch_spmv_simple(P, c->cfg, P.exch[buf], P.Ap, mode, s));
  return CGB_OK;
}

// `n` iterations of the merged variant in ONE cooperative launch (grid barrier + all-reduce on the device).
static int enqueue_persist(cgb_ctx* c, int n, cudaStream_t s) {
  CGB_CUDA(launch_cg_persist(c->P, c->device, c->d_maps, c->host_cur, n, s));
  c->host_cur ^= (n & 1);
  return CGB_OK;
}

int ctx_enqueue_iteration(cgb_ctx* c, cudaStream_t s) {
  const RankDev& P = c->P;
  if (c->use_persist) return enqueue_persist(c, 1, s);
  if (c->opts.variant == CGB_VAR_PIPELINED) {
    CGB_TRY(enqueue_spmv(c, c->host_cur, kSpmvPlain, s));            // q = A w
    CGB_CUDA(launch_pipe_update(P, c->cfg, c->host_cur, s));         // consumes (gamma,delta), publishes the next
    c->host_cur ^= 1;
    return CGB_OK;
  }
  if (c->opts.variant == CGB_VAR_MERGED2) {
    // 64 B/DOF, two kernels: (p-update + SpMV + p.Ap) then (stencil again + x/r update + r.r)
    const int cur = c->host_cur, nb = cur ^ 1;
    const void* maps[6] = {c->tmap_rh, c->tmap[cur], c->tmap[nb], c->tmap_rh_h, c->tmap_h[cur], c->tmap_h[nb]};
    CGB_CUDA(launch_pspmv_tma(P, c->device, maps, P.exch[cur], P.exch[nb], nb, s));
    CGB_CUDA(launch_spmv_tma(P, c->device, c->tmap[nb], c->tmap_h[nb], c->tmap_x, c->tmap_r, P.Ap, kSpmvUpdate, s));
    c->host_cur = nb;
    return CGB_OK;
  }
  double* p = P.exch[0];
  if (c->opts.variant == CGB_VAR_MERGED) {
    // 72 B/DOF: Ap is never stored; the update kernel re-evaluates the stencil from the p slabs
    CGB_CUDA(launch_spmv_tma(P, c->device, c->tmap[0], c->tmap_h[0], nullptr, nullptr, P.Ap, kSpmvAlphaNoStore, s));
    CGB_CUDA(launch_spmv_tma(P, c->device, c->tmap[0], c->tmap_h[0], c->tmap_x, c->tmap_r, P.Ap, kSpmvUpdate, s));
    CGB_CUDA(launch_pupdate(P, c->cfg, p, 0, s));
    return CGB_OK;
  }
  CGB_TRY(enqueue_spmv(c, 0, kSpmvAlpha, s));
  CGB_CUDA(launch_update(P, c->cfg, p, s));
  CGB_CUDA(launch_pupdate(P, c->cfg, p, 0, s));
  return CGB_OK;
}

static int get_graph(cgb_ctx* c, int len, cudaGraphExec_t* out) {
  long long key = ((long long)c->opts.variant << 48) | ((long long)c->opts.spmv_impl << 40) |
                  ((long long)c->host_cur << 36) | (long long)len;
  const int saved_cur = c->host_cur;
  auto it = c->graphs.find(key);
  if (it != c->graphs.end()) {
    *out = it->second;
    return CGB_OK;
  }
  cudaGraph_t g = nullptr;
  CGB_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
  int rc = CGB_OK;
  for (int i = 0; i < len && rc == CGB_OK; ++i) rc = ctx_enqueue_iteration(c, c->stream);
  cudaError_t e = cudaStreamEndCapture(c->stream, &g);
  c->host_cur = saved_cur;  // capturing executes nothing
  if (rc != CGB_OK) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  if (e != cudaSuccess) return fail(CGB_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
  cudaGraphExec_t ge = nullptr;
  e = cudaGraphInstantiate(&ge, g, 0);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) return fail(CGB_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e));
  cudaGraphUpload(ge, c->stream);  // keep the one-time upload out of the first timed launch
  c->graphs[key] = ge;
  *out = ge;
  return CGB_OK;
}

// One graph for a whole solve: a conditional WHILE node (CUDA >= 12.4) whose body is one
// iteration (two for the ping-pong variants, so that captured pointers stay valid) followed by a
// one-thread kernel that keeps the loop alive while the device-resident `done` flag is 0. The
// host launches it once and synchronises once: no polling, no kernels after convergence.
static int get_while_graph(cgb_ctx* c, cudaGraphExec_t* out) {
  const long long key = (1LL << 60) | ((long long)c->opts.variant << 48) | ((long long)c->opts.spmv_impl << 40) |
                        ((long long)c->host_cur << 36);
  auto it = c->graphs.find(key);
  if (it != c->graphs.end()) {
    *out = it->second;
    return CGB_OK;
  }
  cudaGraph_t g = nullptr;
  CGB_CUDA(cudaGraphCreate(&g, 0));
  cudaGraphConditionalHandle h;
  CGB_CUDA(cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault));
  cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
  np.conditional.handle = h;
  np.conditional.type = cudaGraphCondTypeWhile;
  np.conditional.size = 1;
  cudaGraphNode_t node;
  CGB_CUDA(cudaGraphAddNode(&node, g, nullptr, 0, &np));
  cudaGraph_t body = np.conditional.phGraph_out[0];
  const int saved_cur = c->host_cur;
  CGB_CUDA(cudaStreamBeginCaptureToGraph(c->stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
  int rc = CGB_OK;
  for (int i = 0; i < (uses_parity(c) ? 2 : 1) && rc == CGB_OK; ++i) rc = ctx_enqueue_iteration(c, c->stream);
  if (rc == CGB_OK && launch_cond_guard(c->P, (unsigned long long)h, c->stream) != cudaSuccess)
    rc = fail(CGB_ERR_CUDA, "launch of the WHILE guard failed");
  cudaGraph_t dummy = nullptr;
  cudaError_t e = cudaStreamEndCapture(c->stream, &dummy);
  c->host_cur = saved_cur;
  if (rc != CGB_OK) {
    cudaGraphDestroy(g);
    return rc;
  }
  if (e != cudaSuccess) {
    cudaGraphDestroy(g);
    return fail(CGB_ERR_CUDA, "cudaStreamEndCapture (WHILE body): %s", cudaGetErrorString(e));
  }
  cudaGraphExec_t ge = nullptr;
  e = cudaGraphInstantiate(&ge, g, 0);
  if (e != cudaSuccess) {
    cudaGraphDestroy(g);
    return fail(CGB_ERR_CUDA, "cudaGraphInstantiate (WHILE): %s", cudaGetErrorString(e));
  }
  cudaGraphUpload(ge, c->stream);
  c->while_graphs.push_back(g);
  c->graphs[key] = ge;
  *out = ge;
  return CGB_OK;
}

}  // namespace cgb

// Everything a solve needs that may allocate or synchronise a device: option checks, the extra
// vectors of pipelined CG, peer resolution and the iteration graphs. cgb_solve_begin calls it; a
// driver that runs SEVERAL ranks from one host thread must call it on every rank before the first
// cgb_solve_begin, because with peer access enabled a cudaMalloc on one device waits for the
// others, and those would already be spinning on this rank's first message.
extern "C" int cgb_solve_prepare(cgb_ctx* c, const cgb_solve_opts_t* o) {
  CGB_REQUIRE(c && o, "cgb_solve_prepare: NULL argument");
  NvtxRange nvtx_("cgb.solve_prepare");
  CGB_REQUIRE(o->variant >= CGB_VAR_FUSED && o->variant <= CGB_VAR_MERGED2, "unknown variant %d", o->variant);
  if (o->variant == CGB_VAR_MERGED && !c->tma_ok)
    return fail(CGB_ERR_UNSUPPORTED, "the recompute variant needs the TMA band kernels (stencil problems only)");
  if (c->P.csr_dist && (o->variant == CGB_VAR_PIPELINED || o->warm_start))
    return fail(CGB_ERR_UNSUPPORTED, "distributed CSR supports the fused variant from x0 = 0 only");
  if (o->variant == CGB_VAR_MERGED2 && !merged2_ok(c))
    return fail(CGB_ERR_UNSUPPORTED, "the merged variant needs the TMA band kernels (stencil problems) and a z-slab / row-block decomposition");
  CGB_CUDA(cudaSetDevice(c->device));
  c->opts = *o;
  if (c->opts.variant == CGB_VAR_AUTO) {
    // measured on B200 (profiles/ROOFLINE.md): merged > recompute > fused for every stencil, the 27-point one
    // included since its band kernels stopped spilling (partial-sum sweep)
    const bool band = c->tma_ok && kind_is_stencil(c->st.kind) && o->spmv_impl != CGB_SPMV_SIMPLE;
    c->opts.variant = !band ? CGB_VAR_FUSED : (merged2_ok(c) ? CGB_VAR_MERGED2 : CGB_VAR_MERGED);
  }
  if (c->opts.chunk < 1) c->opts.chunk = 25;
  if (c->opts.fixed_iters > 0) {
    c->opts.maxit = c->opts.fixed_iters;
    c->opts.tol = -1.0;  // ||r|| < -1 never fires: convergence test disabled
    c->opts.tol_mode = CGB_TOL_ABS;
  }
  CGB_REQUIRE(c->opts.maxit >= 1 && c->opts.maxit <= c->hist_cap, "maxit %d exceeds the context's hist_cap %d",
              c->opts.maxit, c->hist_cap);
  if (c->opts.spmv_impl == CGB_SPMV_TMA && !c->tma_ok)
    return fail(CGB_ERR_UNSUPPORTED, "TMA SpMV is not available for this problem (stencil problems only)");
  if (c->opts.variant == CGB_VAR_PIPELINED) CGB_TRY(ensure_extra(c));
  CGB_TRY(ctx_finalize_peers(c));
  c->stream = pick_stream(c, o->stream);
  CGB_REQUIRE(c->opts.persistent >= -1 && c->opts.persistent <= 1, "persistent must be -1, 0 or 1");
  const bool persist_ok = c->opts.variant == CGB_VAR_MERGED2 && cg_persist_supported(c->device);
  if (c->opts.persistent == 1 && !persist_ok)
    return fail(CGB_ERR_UNSUPPORTED, "the persistent kernel needs the merged variant and cooperative launch support");
  // persistent == 0 (auto): measured on B200, the persistent kernel ties with the two-kernel driver on one GPU
  // (its inner loops come out ~10 % longer from the compiler, which the saved launches and the hidden reductions
  // only just pay back), so auto keeps two kernels per iteration unless CGB_PERSISTENT=1 says otherwise.
  static const bool env_on = [] { const char* e = std::getenv("CGB_PERSISTENT"); return e && e[0] == '1'; }();
  c->use_persist = persist_ok && (c->opts.persistent == 1 || (c->opts.persistent == 0 && env_on));
  if (c->use_persist) {  // nothing to capture: one cooperative launch runs a whole chunk
    CGB_CUDA(cg_persist_prepare(c->device));
    return CGB_OK;
  }
  if (c->opts.use_graph == 2) {
    cudaGraphExec_t ge;
    const int saved = c->host_cur;
    c->host_cur = c->opts.variant == CGB_VAR_PIPELINED ? 1 : 0;  // parity right after begin
    int rc = get_while_graph(c, &ge);
    c->host_cur = saved;
    CGB_TRY(rc);
  } else if (c->opts.use_graph) {
    cudaGraphExec_t ge;
    const int saved = c->host_cur;
    for (int par = 0; par < (uses_parity(c) ? 2 : 1); ++par) {
      c->host_cur = par;
      CGB_TRY(get_graph(c, c->opts.chunk, &ge));
    }
    c->host_cur = saved;
  }
  return CGB_OK;
}

extern "C" int cgb_solve_begin(cgb_ctx* c, const cgb_solve_opts_t* o) {
  CGB_REQUIRE(c && o, "cgb_solve_begin: NULL argument");
  NvtxRange nvtx_("cgb.solve_begin");
  CGB_TRY(cgb_solve_prepare(c, o));
  c->launches_at_begin = c->launches;
  if (c->opts.warm_start) {
    // r0 = b - A x0: stage x0 as SpMV input (with halos), multiply, then the usual init on the residual
    CGB_CUDA(launch_stage_x0(c->P, c->cfg, c->stream));
    const RankDev& P = c->P;
    // kSpmvDot, not kSpmvPlain: the all-reduce at its end is the cross-rank barrier that keeps k_init's push of
    // p0 into the neighbours' exch[0] ghosts from overtaking a neighbour that is still reading x0 there.
    const int wmode = P.nranks > 1 ? kSpmvDot : kSpmvPlain;
    if (c->st.kind == CGB_KIND_CSR) CGB_CUDA(launch_spmv_csr(P, c->cfg, P.exch[0], P.Ap, wmode, c->stream));
    else if (use_tma(c)) CGB_CUDA(launch_spmv_tma(P, c->device, c->tmap[0], c->tmap_h[0], nullptr, nullptr, P.Ap, wmode, c->stream));
    else CGB_CUDA(launch_spmv_simple(P, c->cfg, P.exch[0], P.Ap, wmode, c->stream));
    c->launches += 2;
  }
  CGB_CUDA(launch_init(c->P, c->cfg, c->opts.tol, c->opts.tol_mode, c->opts.maxit,
                       c->opts.variant == CGB_VAR_MERGED2 ? 1 : 0, c->opts.warm_start ? 1 : 0, c->stream));
  c->launches += 1;
  c->host_cur = 0;
  c->iters_target = c->opts.maxit;
  if (c->opts.variant == CGB_VAR_PIPELINED) {
    // w0 = A r0 (r0 = b sits in exchange buffer 0 after k_init), then the first (gamma, delta)
    const RankDev& P = c->P;
    if (c->st.kind == CGB_KIND_CSR) CGB_CUDA(launch_spmv_csr(P, c->cfg, P.exch[0], P.exch[1], kSpmvPlain, c->stream));
    else if (use_tma(c)) CGB_CUDA(launch_spmv_tma(P, c->device, c->tmap[0], c->tmap_h[0], nullptr, nullptr, P.exch[1], kSpmvPlain, c->stream));
    else CGB_CUDA(launch_spmv_simple(P, c->cfg, P.exch[0], P.exch[1], kSpmvPlain, c->stream));
    CGB_CUDA(launch_pipe_start(P, c->cfg, 1, c->stream));
    c->launches += 2;
    c->host_cur = 1;
    c->iters_target = c->opts.maxit + 1;  // the last update only records ||r|| of iteration maxit
  }
  CGB_CUDA(cudaEventRecord(c->ev_begin, c->stream));
  c->poll_snapshot = Scalars{};
  c->poll_pending = false;
  c->enqueued = 0;
  c->active = true;
  return CGB_OK;
}

extern "C" int cgb_solve_iterate(cgb_ctx* c, int n) {
  CGB_REQUIRE(c, "cgb_solve_iterate: NULL context");
  NvtxRange nvtx_("cgb.solve_iterate");
  if (!c->active) return fail(CGB_ERR_STATE, "cgb_solve_iterate before cgb_solve_begin");
  CGB_CUDA(cudaSetDevice(c->device));
  const int chunk = c->opts.chunk;
  if (c->use_persist && n > 0) {
    CGB_TRY(enqueue_persist(c, n, c->stream));
    c->enqueued += n;
    c->launches += 1;
    return CGB_OK;
  }
  while (n > 0) {
    if (c->opts.use_graph && n >= chunk) {
      cudaGraphExec_t ge;
      CGB_TRY(get_graph(c, chunk, &ge));
      CGB_CUDA(cudaGraphLaunch(ge, c->stream));
      if (uses_parity(c)) c->host_cur ^= (chunk & 1);
      n -= chunk;
      c->enqueued += chunk;
      c->launches += (long long)chunk * launches_per_iter(c);
    } else {
      CGB_TRY(ctx_enqueue_iteration(c, c->stream));
      n -= 1;
      c->enqueued += 1;
      c->launches += launches_per_iter(c);
    }
  }
  return CGB_OK;
}

// n iterations, each followed by a stream-ordered 8-byte D2H copy of ||r||_2 into host_hist[i]
// (pinned). This is the reference's per-iteration residual read-back (CUDACG.cu:328) without its
// host stall: the host never waits, the values simply arrive in order.
extern "C" int cgb_solve_iterate_streamed(cgb_ctx* c, int n, double* host_hist) {
  CGB_REQUIRE(c && host_hist && n >= 0, "cgb_solve_iterate_streamed: bad argument");
  if (!c->active) return fail(CGB_ERR_STATE, "cgb_solve_iterate_streamed before cgb_solve_begin");
  CGB_CUDA(cudaSetDevice(c->device));
  // plain launches: a graph of length 1 would have to be captured and instantiated here, in the middle of a
  // solve, while the peers already spin on this rank (see cgb_solve_prepare)
  int rc = CGB_OK;
  for (int i = 0; i < n && rc == CGB_OK; ++i) {
    rc = ctx_enqueue_iteration(c, c->stream);
    if (rc == CGB_OK) {
      c->enqueued += 1;
      c->launches += c->use_persist ? 1 : launches_per_iter(c);
      cudaError_t e = cudaMemcpyAsync(host_hist + i, &c->P.ctrl->s.rnorm, sizeof(double), cudaMemcpyDeviceToHost, c->stream);
      if (e != cudaSuccess) rc = fail(CGB_ERR_CUDA, "cudaMemcpyAsync(rnorm): %s", cudaGetErrorString(e));
    }
  }
  return rc;
}

static void fill_result(cgb_ctx* c, const Scalars& s, cgb_solve_result_t* res) {
  res->iters = s.iter;
  res->status = s.done;
  res->r0norm = s.r0norm;
  res->rnorm = s.rnorm;
  res->launches = c->launches - c->launches_at_begin;
}

extern "C" int cgb_solve_poll(cgb_ctx* c, int wait, cgb_solve_result_t* res) {
  CGB_REQUIRE(c && res, "cgb_solve_poll: NULL argument");
  if (!c->active) return fail(CGB_ERR_STATE, "cgb_solve_poll before cgb_solve_begin");
  CGB_CUDA(cudaSetDevice(c->device));
  if (wait) {
    Scalars* slot = &c->h_scal[0];
    CGB_CUDA(cudaMemcpyAsync(slot, &c->P.ctrl->s, sizeof(Scalars), cudaMemcpyDeviceToHost, c->stream));
    CGB_CUDA(cudaStreamSynchronize(c->stream));  // stream-ordered: reflects everything enqueued so far
    c->poll_snapshot = *slot;
    c->poll_pending = false;
  } else {
    // non-blocking: harvest the previous snapshot if its copy has landed, start the next one, and return the
    // newest completed state (possibly one call old). The host never waits for the GPU here.
    Scalars* slot = &c->h_scal[2];
    if (c->poll_pending) {
      cudaError_t q = cudaEventQuery(c->ev_poll_nb);
      if (q == cudaSuccess) {
        c->poll_snapshot = *slot;
        c->poll_pending = false;
      } else if (q != cudaErrorNotReady) {
        return fail(CGB_ERR_CUDA, "cudaEventQuery: %s", cudaGetErrorString(q));
      } else {
        cudaGetLastError();
      }
    }
    if (!c->poll_pending) {
      CGB_CUDA(cudaMemcpyAsync(slot, &c->P.ctrl->s, sizeof(Scalars), cudaMemcpyDeviceToHost, c->stream));
      CGB_CUDA(cudaEventRecord(c->ev_poll_nb, c->stream));
      c->poll_pending = true;
    }
  }
  fill_result(c, c->poll_snapshot, res);
  res->device_ms = 0.0;
  return CGB_OK;
}

extern "C" int cgb_solve_finish(cgb_ctx* c, cgb_solve_result_t* res) {
  CGB_REQUIRE(c, "cgb_solve_finish: NULL context");
  NvtxRange nvtx_("cgb.solve_finish");
  if (!c->active) return fail(CGB_ERR_STATE, "cgb_solve_finish before cgb_solve_begin");
  CGB_CUDA(cudaSetDevice(c->device));
  CGB_CUDA(cudaEventRecord(c->ev_end, c->stream));
  Scalars* slot = &c->h_scal[0];
  CGB_CUDA(cudaMemcpyAsync(slot, &c->P.ctrl->s, sizeof(Scalars), cudaMemcpyDeviceToHost, c->stream));
  CGB_CUDA(cudaStreamSynchronize(c->stream));
  c->active = false;
  if (res) {
    fill_result(c, *slot, res);
    float ms = 0.f;
    CGB_CUDA(cudaEventElapsedTime(&ms, c->ev_begin, c->ev_end));
    res->device_ms = ms;
    if (slot->done == CGB_ERR_TIMEOUT) return fail(CGB_ERR_TIMEOUT, "rank %d: peer wait timed out on the device", c->part.rank);
  }
  return CGB_OK;
}

// Host loop: keeps one chunk queued ahead of the chunk whose convergence flag it inspects, so
// the GPU never idles while the host looks. Kernels enqueued after convergence exit at once.
extern "C" int cgb_solve(cgb_ctx* c, const cgb_solve_opts_t* o, cgb_solve_result_t* res) {
  CGB_REQUIRE(c && o && res, "cgb_solve: NULL argument");
  CGB_TRY(cgb_solve_begin(c, o));
  if (c->opts.use_graph == 2 && c->use_persist) {
    // the persistent kernel IS the device-side loop: one launch, one synchronise, nothing runs after convergence
    CGB_TRY(cgb_solve_iterate(c, c->iters_target));
    return cgb_solve_finish(c, res);
  }
  if (c->opts.use_graph == 2) {
    cudaGraphExec_t ge;
    CGB_TRY(get_while_graph(c, &ge));
    CGB_CUDA(cudaGraphLaunch(ge, c->stream));
    int rc = cgb_solve_finish(c, res);
    // kernels that actually ran: init (+2 for pipelined) + per trip (iterations + guard)
    const int per_trip = uses_parity(c) ? 2 : 1;
    const int trips = (res->iters + (c->opts.variant == CGB_VAR_PIPELINED ? 1 : 0) + per_trip - 1) / per_trip;
    res->launches = (c->launches - c->launches_at_begin) + (long long)trips * (per_trip * launches_per_iter(c) + 1);
    c->launches += (long long)trips * (per_trip * launches_per_iter(c) + 1);
    return rc;
  }
  const int chunk = c->opts.chunk, maxit = c->iters_target;
  int slot = 0;
  bool pending[2] = {false, false};
  auto enqueue_chunk = [&]() -> int {
    int n = std::min(chunk, maxit - c->enqueued);
    CGB_TRY(cgb_solve_iterate(c, n));
    CGB_CUDA(cudaMemcpyAsync(&c->h_scal[slot], &c->P.ctrl->s, sizeof(Scalars), cudaMemcpyDeviceToHost, c->stream));
    CGB_CUDA(cudaEventRecord(c->ev_poll[slot], c->stream));
    pending[slot] = true;
    slot ^= 1;
    return CGB_OK;
  };
  CGB_TRY(enqueue_chunk());
  for (;;) {
    if (c->enqueued < maxit) CGB_TRY(enqueue_chunk());
    int oldest = pending[slot] ? slot : slot ^ 1;  // slot now points at the older entry if both pending
    if (!pending[oldest]) break;
    CGB_CUDA(cudaEventSynchronize(c->ev_poll[oldest]));
    pending[oldest] = false;
    if (c->h_scal[oldest].done != 0) break;
    if (c->enqueued >= maxit && !pending[0] && !pending[1]) break;
  }
  return cgb_solve_finish(c, res);
}
