/*
 * cgb200.h — C ABI of the B200-native distributed conjugate-gradient framework.
 *
 * One shared library (libcgb200.so, built for sm_100a) exposes:
 *   - problem layer   : structured stencil descriptors, CSR builders, the reference 3x3 system
 *   - topology layer  : 1D/2D row-block partitions and the padded local layout
 *   - CPU oracle      : plain fp64 restatement of the reference algorithm (CSR and matrix-free,
 *                       partitioned "fake fabric" mode, classic and pipelined CG)
 *   - GPU context     : one per rank/GPU; owns a peer-mappable arena, kernels, CUDA graphs
 *   - comm fabric     : peer table built from cudaIpc handles (multi-process) or raw pointers
 *                       (single-process); no MPI/NCCL on any iteration path
 *   - standalone ops  : each hot kernel callable on caller-owned device pointers (unit tests,
 *                       Python `ops` API)
 *
 * Behavioural parity target: /root/reference/CUDACG.cu (the whole reference).  Citations in the
 * comments below are `CUDACG.cu:LINE`.
 *
 * All functions return 0 on success and a negative cgb_status on failure unless stated
 * otherwise; cgb_last_error() gives a thread-local human readable message.
 */
#ifndef CGB200_H_
#define CGB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CGB_MAX_RANKS 16   /* reduction slots per arena; one 8xB200 box needs 8 */
#define CGB_NDIR 9         /* (dy+1)+3*(dz+1); entry 4 is "self" and unused */

typedef enum cgb_status {
  CGB_OK = 0,
  CGB_ERR_INVALID = -1,   /* bad argument */
  CGB_ERR_CUDA = -2,      /* CUDA runtime/driver call failed */
  CGB_ERR_STATE = -3,     /* call out of order (e.g. iterate before begin) */
  CGB_ERR_TIMEOUT = -4,   /* a peer flag wait expired on the device */
  CGB_ERR_BREAKDOWN = -5, /* non-finite scalar in the recurrence */
  CGB_ERR_UNSUPPORTED = -6
} cgb_status;

/* ---- problem layer (replaces the literals of CUDACG.cu:74-117,137-140) ---- */

typedef enum cgb_kind {
  CGB_KIND_CSR = 0,        /* general CSR, fp64 values, int32 indices (CUDACG.cu:213-216) */
  CGB_KIND_POISSON2D5 = 1, /* 2D 5-point, homogeneous Dirichlet; the nx x ny grid is stored as nx x 1 x ny */
  CGB_KIND_POISSON3D7 = 2, /* 3D 7-point */
  CGB_KIND_STENCIL3D27 = 3 /* 3D 27-point */
} cgb_kind;

typedef struct cgb_stencil_t {
  int32_t kind;         /* cgb_kind */
  int32_t gx, gy, gz;   /* global grid (2D: gy = 1 and the rows are in gz; CSR: gx = n, gy = gz = 1) */
  double c_center;      /* diagonal coefficient */
  double c_face;        /* neighbours sharing a face */
  double c_edge;        /* 27-pt only: neighbours sharing an edge */
  double c_corner;      /* 27-pt only: neighbours sharing a corner */
} cgb_stencil_t;

/* Fills default coefficients: 5-pt (4,-1), 7-pt (6,-1), 27-pt (26,-1,-1,-1). */
int cgb_stencil_default(int kind, int gx, int gy, int gz, cgb_stencil_t* out);
/* Number of nonzeros of rows [row0,row0+nrows) of the assembled stencil matrix. */
int64_t cgb_stencil_csr_nnz(const cgb_stencil_t* st, int64_t row0, int64_t nrows);
/* Assemble rows [row0,row0+nrows) into 0-based CSR with *global* column indices (host arrays).
 * rowptr has nrows+1 entries and is relative to the first assembled nonzero. */
int cgb_stencil_to_csr(const cgb_stencil_t* st, int64_t row0, int64_t nrows, int64_t* rowptr,
                       int32_t* colind, double* val);
/* The reference's hard-coded system (CUDACG.cu:74-117,137-140): n=3, nnz=5.
 * Buffers must hold val[5], rowptr[4], colind[5], b[3]. */
void cgb_ref3x3(int32_t* n, int32_t* nnz, double* val, int32_t* rowptr, int32_t* colind,
                double* b);

typedef enum cgb_rhs_kind {
  CGB_RHS_ONES = 0,         /* b = 1 */
  CGB_RHS_RANDOM = 1,       /* counter-based hash of the global index and a seed, in [-1,1) */
  CGB_RHS_MANUFACTURED = 2  /* b = A u with u = prod sin(pi*(i+1)/(g+1)) (known solution) */
} cgb_rhs_kind;
/* Host generators: value of b / of the manufactured solution at a global linear index. */
double cgb_rhs_value(const cgb_stencil_t* st, int rhs_kind, uint64_t seed, int64_t gidx);
double cgb_manufactured_u(const cgb_stencil_t* st, int64_t gidx);
/* Fill `out` (gx*ny*nz doubles, x fastest) with the generated RHS of the box of full rows [y0,y0+ny) x [z0,z0+nz):
 * bit-identical to what cgb_ctx_set_rhs_generated produces on the device for the same (kind, seed). */
int cgb_rhs_fill_box(const cgb_stencil_t* st, int rhs_kind, uint64_t seed, int y0, int ny, int z0, int nz, double* out);

/* ---- topology layer (absent in the reference: cudaSetDevice(0), CUDACG.cu:87) ---- */

typedef struct cgb_partition_t {
  int32_t nranks, rank;
  int32_t Py, Pz;        /* rank grid; 2D problems and CSR: Pz == 1 */
  int32_t ry, rz;        /* my coordinates: rank = ry + Py*rz */
  int32_t nx;            /* never split */
  int32_t y0, ny;        /* owned y range */
  int32_t z0, nz;        /* owned z range */
  int32_t nbr[CGB_NDIR]; /* neighbour rank per direction, -1 if none/not needed by the stencil */
  int32_t x0;            /* general CSR only: first global row of this rank's row block (nx = rows owned) */
} cgb_partition_t;

/* Py,Pz == 0 picks: decomp=1 -> z slabs (row blocks for 2D grids), decomp=2 (3D only) -> the most square
 * Py x Pz factorisation. */
int cgb_partition_make(const cgb_stencil_t* st, int nranks, int rank, int decomp, int Py, int Pz,
                       cgb_partition_t* out);

/* Padded local box: one ghost layer in y (and z for 3D) that is either a neighbour's data or a
 * zero Dirichlet layer, rows padded to 128 bytes. Same index for every vector of a solve. */
typedef struct cgb_layout_t {
  int32_t nx, ny, nz;
  int32_t gy, gz;       /* ghost widths (0 or 1) */
  int32_t pitch;        /* row pitch in elements (multiple of 16) */
  int32_t py, pz;       /* padded extents: ny+2*gy, nz+2*gz */
  int64_t vec_elems;    /* elements per vector incl. padding (multiple of 32) */
  int64_t owned;        /* nx*ny*nz */
} cgb_layout_t;
int cgb_layout_make(const cgb_stencil_t* st, const cgb_partition_t* part, cgb_layout_t* out);

/* ---- CPU oracle (test tier T1/T7; BASELINE config 1) ---- */

typedef enum cgb_tol_mode { CGB_TOL_ABS = 0, CGB_TOL_REL = 1 } cgb_tol_mode;

typedef struct cgb_cpu_result_t {
  int32_t iters;       /* iterations executed (reference: loop index at break + 1) */
  int32_t converged;   /* 1 if the residual test fired, 0 if maxit was reached */
  double r0norm;       /* ||b|| */
  double rnorm;        /* last ||r||_2 */
} cgb_cpu_result_t;

/* Restates CUDACG.cu:248-352 on the host: x0=0, r=p=b, absolute ||r||_2 < tol tested after the
 * x/r update and before the p update. hist (may be NULL) receives ||r|| after each iteration. */
int cgb_cpu_cg_csr(int64_t n, const int64_t* rowptr, const int32_t* colind, const double* val,
                   const double* b, double* x, double tol, int tol_mode, int maxit, double* hist,
                   cgb_cpu_result_t* res);
/* Matrix-free, partitioned over `nranks` emulated ranks in one process: every rank owns a padded
 * local box, halos are copied between the boxes and dot products are summed in rank order.
 * b and x are global compact arrays (gx*gy*gz). variant: 0 classic, 1 pipelined (Ghysels-Vanroose). */
int cgb_cpu_cg_stencil(const cgb_stencil_t* st, int nranks, int decomp, int Py, int Pz,
                       const double* b, double* x, double tol, int tol_mode, int maxit, int variant,
                       double* hist, cgb_cpu_result_t* res);
/* out = A in on one rank's padded box (caller fills the ghost layers); *dot_out = in.out. */
int cgb_cpu_spmv_local(const cgb_stencil_t* st, const cgb_layout_t* lay, const double* in, double* out,
                       double* dot_out);
/* y = A x on global compact arrays. */
int cgb_cpu_spmv_stencil(const cgb_stencil_t* st, const double* x, double* y);

/* ---- GPU context ---- */

typedef struct cgb_ctx cgb_ctx;

typedef enum cgb_variant {
  CGB_VAR_FUSED = 0,      /* classic CG, 3 fused kernels/iter (spmv+dot, update+dot, p-update+halo) */
  CGB_VAR_MERGED = 1,     /* classic CG "recompute": Ap is never stored, the x/r update re-evaluates the
                             stencil from TMA slabs of p (72 instead of 88 B/DOF); any stencil problem (rows wider than
                             one tile are cut into x tiles) */
  CGB_VAR_PIPELINED = 2,  /* Ghysels-Vanroose, one fused 2-scalar reduction per iteration */
  CGB_VAR_AUTO = 3,       /* fastest measured variant that applies: MERGED2, else MERGED, else FUSED */
  CGB_VAR_MERGED2 = 4     /* classic CG, 2 kernels/iter, 64 B/DOF: p = r + beta p folded into the SpMV (halo rows
                             recomputed in shared memory, boundary planes pushed first) + the MERGED update
                             kernel; any stencil problem, z-slab (2D: row-block) decomposition */
} cgb_variant;

typedef enum cgb_spmv_impl {
  CGB_SPMV_AUTO = 0,
  CGB_SPMV_SIMPLE = 1,    /* register-marching, global loads through L1 */
  CGB_SPMV_TMA = 2        /* TMA (cp.async.bulk.tensor) plane ring + mbarrier pipeline */
} cgb_spmv_impl;

typedef struct cgb_solve_opts_t {
  double tol;            /* reference: 1e-7 (CUDACG.cu:245) */
  int32_t tol_mode;      /* reference: absolute (CUDACG.cu:333) */
  int32_t maxit;         /* reference: 2000 (CUDACG.cu:244) */
  int32_t variant;       /* cgb_variant */
  int32_t spmv_impl;     /* cgb_spmv_impl */
  int32_t use_graph;     /* 0: plain launches; 1: graphs of `chunk` iterations, host polls the device
                            flag one chunk behind; 2: ONE graph with a conditional WHILE node keyed on
                            the device flag (cgb_solve only): no polling, nothing runs after convergence */
  int32_t chunk;         /* iterations per graph launch / per host convergence poll */
  int32_t fixed_iters;   /* >0: run exactly this many iterations, convergence test disabled */
  int32_t warm_start;    /* 0: x0 = 0 like the reference (CUDACG.cu:177-179, 247); 1: start from the x the
                            context holds (cgb_ctx_set_x0 or the previous solve): r0 = b - A x0. This is
                            also the restart path of a checkpoint (x + iteration count). */
  int32_t persistent;    /* merged variant only. 1: run the iteration loop inside ONE persistent cooperative kernel per
                            chunk - grid barrier on the device, the all-reduces done by a service CTA off the critical
                            path, no kernel boundary inside an iteration; -1: two kernels per iteration (captured in
                            graphs); 0 (default): auto = -1 unless the environment says CGB_PERSISTENT=1 (on one B200
                            the two are within 2 % of each other, see profiles/ROOFLINE.md) */
  void* stream;          /* cudaStream_t to run on; NULL = context stream */
} cgb_solve_opts_t;
void cgb_solve_opts_default(cgb_solve_opts_t* o);

typedef struct cgb_solve_result_t {
  int32_t iters;
  int32_t status;        /* 1 converged, 2 maxit reached, 0 still running, <0 cgb_status */
  double r0norm;
  double rnorm;
  double device_ms;      /* CUDA-event time of the iteration loop (excludes init) */
  int64_t launches;      /* kernels launched by this solve */
} cgb_solve_result_t;

/* Stencil problem on `device`, this rank's box given by `part`. hist_cap bounds maxit. */
int cgb_ctx_create(const cgb_stencil_t* st, const cgb_partition_t* part, int device, int hist_cap,
                   cgb_ctx** out);
/* General CSR problem (single rank), host arrays are copied to the device
 * (CUDACG.cu:120-134). */
int cgb_ctx_create_csr(int32_t n, int64_t nnz, const int32_t* rowptr, const int32_t* colind,
                       const double* val, int device, int hist_cap, cgb_ctx** out);
/* Row-block distributed general CSR (goes beyond the reference, which is single-GPU): this rank owns the
 * rows [part.x0, part.x0 + part.nx) of an n_global x n_global matrix (cgb_partition_make with a CSR
 * descriptor); rowptr is local (0-based, nx+1 entries), colind holds GLOBAL column indices. The SpMV
 * reads remote entries of p directly from the owners' arenas over NVLink (peer loads), after a
 * release/acquire hand-over of "p version v is complete"; variant FUSED only. */
int cgb_ctx_create_csr_dist(int32_t n_global, const cgb_partition_t* part, int64_t nnz_local, const int32_t* rowptr,
                            const int32_t* colind, const double* val, int device, int hist_cap, cgb_ctx** out);
int cgb_ctx_destroy(cgb_ctx* c);

/* Fabric: every rank exports its arena, every rank attaches all peers it talks to. */
#define CGB_IPC_HANDLE_BYTES 64
int cgb_ctx_ipc_export(cgb_ctx* c, void* handle_out);
int cgb_ctx_peer_attach_ipc(cgb_ctx* c, int peer_rank, const void* handle);
int cgb_ctx_peer_attach_ptr(cgb_ctx* c, int peer_rank, void* peer_arena_base); /* same process */
void* cgb_ctx_arena_base(cgb_ctx* c);
size_t cgb_ctx_arena_bytes(cgb_ctx* c);
/* Zero flags/sequence numbers; call on all ranks (with a host barrier after) before first use. */
int cgb_ctx_fabric_reset(cgb_ctx* c);

/* Right-hand side. Generators run on the device (no H2D for stencil problems). */
int cgb_ctx_set_rhs_generated(cgb_ctx* c, int rhs_kind, uint64_t seed, void* stream);
/* b given as this rank's compact owned box (nx*ny*nz, x fastest); src may be host (pinned for
 * async) or device memory. */
int cgb_ctx_set_rhs(cgb_ctx* c, const double* src, int src_is_device, void* stream);
/* Initial guess for a warm start: this rank's compact owned box, host or device memory. */
int cgb_ctx_set_x0(cgb_ctx* c, const double* src, int src_is_device, void* stream);
/* Copy the owned part of x into a compact buffer (host or device). */
int cgb_ctx_get_x(cgb_ctx* c, double* dst, int dst_is_device, void* stream);
/* Copy residual history [0,n) to host memory. hist[i] = ||r|| after iteration i. */
int cgb_ctx_get_history(cgb_ctx* c, double* dst, int n);

/* Solve = begin (init kernel) + iterate* + finish. cgb_solve does all three. */
int cgb_solve(cgb_ctx* c, const cgb_solve_opts_t* o, cgb_solve_result_t* res);
int cgb_solve_begin(cgb_ctx* c, const cgb_solve_opts_t* o);
/* Does every allocating / device-synchronising part of a solve (extra vectors, graph capture) and
 * nothing else. cgb_solve_begin calls it. A driver that runs several ranks from ONE host thread
 * must call it on every rank before the first cgb_solve_begin: with peer access enabled, a
 * cudaMalloc on one device waits for the other devices, which would already be spinning on this
 * rank's first message. */
int cgb_solve_prepare(cgb_ctx* c, const cgb_solve_opts_t* o);
/* Enqueue up to n more iterations without synchronising the host. */
int cgb_solve_iterate(cgb_ctx* c, int n);
/* Host-visible residual history: `pinned` (cudaHostAlloc'ed / torch pinned memory, cap doubles, or NULL to
 * switch off) receives ||r||_2 of every iteration through a device-side store the moment the iteration's
 * reduction completes - the reference's per-iteration read-back (CUDACG.cu:328) with no host stall and no
 * extra stream operation. Set it before cgb_solve_prepare/begin (captured graphs embed the pointer). */
int cgb_ctx_set_host_history(cgb_ctx* c, double* pinned, int cap);
/* n iterations, each followed by a stream-ordered 8-byte device->host copy of ||r||_2 into
 * host_hist[i] (pinned memory): the reference's per-iteration residual read-back (CUDACG.cu:328)
 * without stalling the host. */
int cgb_solve_iterate_streamed(cgb_ctx* c, int n, double* host_hist);
/* wait != 0: stream-ordered snapshot of the device state, host blocks until it has landed.
 * wait == 0: never blocks: returns the newest snapshot whose copy has completed (zeros before the first one
 * lands) and starts the next copy; a host loop can poll between chunks without serialising with the GPU. */
int cgb_solve_poll(cgb_ctx* c, int wait, cgb_solve_result_t* res);
/* Finalise x for variants that defer the last update; must follow the last iterate. */
int cgb_solve_finish(cgb_ctx* c, cgb_solve_result_t* res);
/* Profiling aid (the reference has no timing at all; its cpuSecond(), CUDACG.cu:35-39, is never called): the
 * persistent kernel stores four %globaltimer readings (ns) per iteration into dev_buf[4*iter .. 4*iter+3] - last CTA
 * reached the p.Ap barrier, alpha released, last CTA reached the r.r barrier, beta released - which splits an
 * iteration into compute and rendezvous time. dev_buf: device memory for 4*cap_iters u64, or NULL to switch off. */
int cgb_ctx_set_trace(cgb_ctx* c, unsigned long long* dev_buf, int cap_iters);
/* Print the device-resident solver state (iteration, scalars, barrier generations, the site of a timed-out wait)
 * to stderr: post-mortem aid after CGB_ERR_TIMEOUT. Synchronises the device. */
int cgb_ctx_debug_print(cgb_ctx* c);
/* Kernels launched by this context so far (what bench.py reports as gpu_launches). */
int64_t cgb_ctx_launch_count(cgb_ctx* c);

/* ---- standalone ops on caller-owned device memory (test tier T2; Python ops API) ---- */

/* Scratch bytes any op below needs (partials + ticket), independent of the problem size. */
size_t cgb_op_scratch_bytes(void);
/* out[0] = sum a[i]*b[i], deterministic two-stage tree; scratch from cgb_op_scratch_bytes(). */
int cgb_op_dot(const double* a, const double* b, int64_t n, double* out, void* scratch,
               void* stream);
/* No counterpart in the reference. Same result as cgb_op_dot, but allocates nothing and takes no
 * scratch argument: for callers that may not call cudaMalloc where they need the dot product
 * (time-critical loops that cannot afford the device-wide synchronisation an allocation implies,
 * plug-ins whose host framework owns all device memory, out-of-memory/teardown paths). The
 * scratch is a 65,792-byte __device__ array inside the library, resident on every GPU that
 * loads it until the process exits. Limit: n <= cgb_op_dot_static_max_n() = 2^28 elements; larger
 * n returns CGB_ERR_INVALID. The first call on each device resolves the array's device address
 * (cudaGetSymbolAddress) and should be made before any stream capture. All calls on a device
 * share the array: calls that overlap on the GPU race and must be serialised by the caller. */
int cgb_op_dot_static(const double* a, const double* b, int64_t n, double* out, void* stream);
long long cgb_op_dot_static_max_n(void);
/* Ap = A p on a padded local box (ghost layers of p must be filled by the caller);
 * dot_out (may be NULL) = p.Ap over owned cells. impl: cgb_spmv_impl. */
int cgb_op_spmv_stencil(const cgb_stencil_t* st, const cgb_layout_t* lay, const double* p,
                        double* Ap, double* dot_out, void* scratch, int impl, void* stream);
/* y = A x for CSR, dot_out (may be NULL) = x.y. */
int cgb_op_spmv_csr(int32_t n, const int32_t* rowptr, const int32_t* colind, const double* val,
                    const double* x, double* y, double* dot_out, void* scratch, void* stream);
/* x += alpha p ; r -= alpha Ap ; rr_out = r.r   (CUDACG.cu:314-328 in one pass) */
int cgb_op_fused_update(const cgb_layout_t* lay, double alpha, const double* p, const double* Ap,
                        double* x, double* r, double* rr_out, void* scratch, void* stream);
/* p = r + beta p   (CUDACG.cu:339-347 in one pass) */
int cgb_op_p_update(const cgb_layout_t* lay, double beta, const double* r, double* p, void* scratch,
                    void* stream);
/* compact <-> padded copies of one local box */
int cgb_op_pad(const cgb_layout_t* lay, const double* compact, double* padded, void* stream);
int cgb_op_unpad(const cgb_layout_t* lay, const double* padded, double* compact, void* stream);

/* Geometry the TMA band kernels would use for a local box (host only, for tests and tuning): see csrc/api/ops.cu. */
int cgb_debug_tma_geometry(const cgb_stencil_t* st, const cgb_layout_t* lay, int nbr_lo, int nbr_hi, int nsm,
                           int mode, int* out16);

/* ---- misc ---- */
const char* cgb_version(void);
const char* cgb_last_error(void);
int cgb_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* CGB200_H_ */
