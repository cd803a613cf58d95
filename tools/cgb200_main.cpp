// cgb200 — command-line front end (layers L5/L6 of the reference: solve + output).
//
// With no arguments it solves the reference's own 3x3 system with the reference's controls
// (tol 1e-7 absolute, maxit 2000; CUDACG.cu:244-245) and prints exactly what CUDACG.cu prints
// (:361-365): one "%f" line per entry of x, then "Success". On failure it prints one message line
// and exits with status 1, like the CLEANUP(msg); return EXIT_FAILURE idiom (:88-91).
//
// Everything the reference hard-codes is a flag here:
//   --problem ref3x3|poisson2d5|poisson3d7|stencil3d27|csr:<file.mtx>     --n NX[,NY[,NZ]]
//   --gpus P  --decomp 1d|2d[:PYxPZ]   --variant auto|fused|recompute|merged|pipelined|cpu   --spmv auto|simple|tma
//   --tol T  --tol-mode abs|rel  --maxit N  --rhs ones|random[:seed]|manufactured
//   --print-x / --no-print-x   --x-out FILE   --hist-out FILE   --json-out FILE   --chunk N
// stdout stays diffable against the reference; history ("iter ||r||" as "%d %.17e") and the timing
// JSON go to files. Multi-GPU runs use one context per GPU inside this process (peer access +
// raw arena pointers); no MPI, no NCCL.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "cgb200.h"

namespace {

struct Args {
  std::string problem = "ref3x3", csr_file, variant = "auto", spmv = "auto", rhs = "ones", x_out, hist_out, json_out, x0_in;
  int n[3] = {0, 0, 0};
  int gpus = 1, decomp = 1, Py = 0, Pz = 0, maxit = 2000, chunk = 25, tol_mode = CGB_TOL_ABS;
  unsigned long long seed = 0;
  double tol = 1e-7;
  int print_x = -1;  // -1: default (on for small problems)
  int graph = 1;     // 0 off, 1 chunked graphs, 2 conditional WHILE graph (single GPU driver)
};

[[noreturn]] void die(const std::string& msg) {
  std::printf("%s\n", msg.c_str());  // the reference reports failures on stdout (CUDACG.cu:11)
  std::exit(EXIT_FAILURE);
}

void check(int rc, const char* what) {
  if (rc != 0) die(std::string(what) + " failed: " + cgb_last_error());
}

Args parse(int argc, char** argv) {
  Args a;
  auto need = [&](int& i) -> std::string {
    if (i + 1 >= argc) die(std::string("missing value for ") + argv[i]);
    return argv[++i];
  };
  for (int i = 1; i < argc; ++i) {
    std::string f = argv[i];
    if (f == "--problem") {
      a.problem = need(i);
      if (a.problem.rfind("csr:", 0) == 0) {
        a.csr_file = a.problem.substr(4);
        a.problem = "csr";
      }
    } else if (f == "--n") {
      std::string v = need(i);
      int k = 0;
      std::stringstream ss(v);
      std::string tok;
      while (std::getline(ss, tok, ',') && k < 3) a.n[k++] = std::atoi(tok.c_str());
    } else if (f == "--gpus") a.gpus = std::atoi(need(i).c_str());
    else if (f == "--decomp") {
      std::string v = need(i);
      a.decomp = v.rfind("2d", 0) == 0 ? 2 : 1;
      size_t c = v.find(':');
      if (c != std::string::npos) std::sscanf(v.c_str() + c + 1, "%dx%d", &a.Py, &a.Pz);
    } else if (f == "--tol") a.tol = std::atof(need(i).c_str());
    else if (f == "--tol-mode") a.tol_mode = need(i) == "rel" ? CGB_TOL_REL : CGB_TOL_ABS;
    else if (f == "--maxit") a.maxit = std::atoi(need(i).c_str());
    else if (f == "--chunk") a.chunk = std::atoi(need(i).c_str());
    else if (f == "--variant") a.variant = need(i);
    else if (f == "--graph") {
      std::string v = need(i);
      a.graph = v == "while" ? 2 : (v == "off" ? 0 : 1);
    }
    else if (f == "--spmv") a.spmv = need(i);
    else if (f == "--rhs") {
      a.rhs = need(i);
      size_t c = a.rhs.find(':');
      if (c != std::string::npos) {
        a.seed = std::strtoull(a.rhs.c_str() + c + 1, nullptr, 10);
        a.rhs = a.rhs.substr(0, c);
      }
    } else if (f == "--print-x") a.print_x = 1;
    else if (f == "--no-print-x") a.print_x = 0;
    else if (f == "--x-out") a.x_out = need(i);
    else if (f == "--x0") a.x0_in = need(i);  // warm start / restart from a file written by --x-out
    else if (f == "--hist-out") a.hist_out = need(i);
    else if (f == "--json-out") a.json_out = need(i);
    else if (f == "--help" || f == "-h") {
      std::printf("usage: cgb200 [--problem ref3x3|poisson2d5|poisson3d7|stencil3d27|csr:<file.mtx>] [--n NX[,NY[,NZ]]]\n"
                  "              [--gpus P] [--decomp 1d|2d[:PYxPZ]] [--variant auto|fused|recompute|merged|pipelined|cpu] [--spmv auto|simple|tma]\n"
                  "              [--tol T] [--tol-mode abs|rel] [--maxit N] [--rhs ones|random[:seed]|manufactured]\n"
                  "              [--print-x|--no-print-x] [--x-out F] [--hist-out F] [--json-out F] [--x0 F] [--chunk N] [--graph chunk|while|off]\n");
      std::exit(0);
    } else die("unknown option " + f);
  }
  return a;
}

// MatrixMarket coordinate real general|symmetric -> 0-based CSR.
void read_mtx(const std::string& path, std::vector<int>& rowptr, std::vector<int>& colind, std::vector<double>& val) {
  std::ifstream in(path);
  if (!in) die("cannot open " + path);
  std::string line;
  if (!std::getline(in, line)) die("empty matrix file");
  const bool sym = line.find("symmetric") != std::string::npos;
  while (std::getline(in, line))
    if (!line.empty() && line[0] != '%') break;
  long long nr = 0, nc = 0, nz = 0;
  std::sscanf(line.c_str(), "%lld %lld %lld", &nr, &nc, &nz);
  if (nr <= 0 || nr != nc) die("matrix must be square");
  struct E { int r, c; double v; };
  std::vector<E> es;
  es.reserve((size_t)(sym ? 2 * nz : nz));
  for (long long e = 0; e < nz; ++e) {
    long long r, c;
    double v;
    if (!(in >> r >> c >> v)) die("truncated matrix file");
    es.push_back({(int)r - 1, (int)c - 1, v});
    if (sym && r != c) es.push_back({(int)c - 1, (int)r - 1, v});
  }
  std::sort(es.begin(), es.end(), [](const E& a, const E& b) { return a.r != b.r ? a.r < b.r : a.c < b.c; });
  rowptr.assign((size_t)nr + 1, 0);
  for (const E& e : es) rowptr[(size_t)e.r + 1]++;
  for (long long r = 0; r < nr; ++r) rowptr[(size_t)r + 1] += rowptr[(size_t)r];
  colind.resize(es.size());
  val.resize(es.size());
  for (size_t e = 0; e < es.size(); ++e) {
    colind[e] = es[e].c;
    val[e] = es[e].v;
  }
}

int rhs_kind(const std::string& s) {
  if (s == "ones") return CGB_RHS_ONES;
  if (s == "random") return CGB_RHS_RANDOM;
  if (s == "manufactured") return CGB_RHS_MANUFACTURED;
  die("unknown --rhs " + s);
}

}  // namespace

int main(int argc, char** argv) {
  Args a = parse(argc, argv);
  std::vector<double> x, hist;
  cgb_solve_result_t res{};
  long long n_total = 0;
  const bool cpu = a.variant == "cpu";

  if (a.problem == "ref3x3" || a.problem == "csr") {
    std::vector<int> rowptr, colind;
    std::vector<double> val, b;
    if (a.problem == "ref3x3") {
      int n, nnz;
      rowptr.resize(4);
      colind.resize(5);
      val.resize(5);
      b.resize(3);
      cgb_ref3x3(&n, &nnz, val.data(), rowptr.data(), colind.data(), b.data());
    } else {
      read_mtx(a.csr_file, rowptr, colind, val);
      b.assign(rowptr.size() - 1, 1.0);
    }
    const int n = (int)rowptr.size() - 1;
    n_total = n;
    x.assign((size_t)n, 0.0);
    hist.assign((size_t)a.maxit + 1, 0.0);
    if (cpu) {
      std::vector<int64_t> rp(rowptr.begin(), rowptr.end());
      cgb_cpu_result_t cr{};
      check(cgb_cpu_cg_csr(n, rp.data(), colind.data(), val.data(), b.data(), x.data(), a.tol, a.tol_mode, a.maxit,
                           hist.data(), &cr), "cpu cg");
      res.iters = cr.iters;
      res.status = cr.converged ? 1 : 2;
      res.r0norm = cr.r0norm;
      res.rnorm = cr.rnorm;
    } else if (a.gpus > 1) {
      // Row-block distributed general CSR (beyond the reference, which is single-GPU): every rank owns a block of rows
      // with global column indices; the SpMV loads remote entries of p from the owners' arenas over NVLink.
      const int P = a.gpus;
      if (P > cgb_device_count()) die("--gpus exceeds the visible devices");
      cgb_stencil_t st;
      check(cgb_stencil_default(CGB_KIND_CSR, n, 1, 1, &st), "problem setup");
      std::vector<cgb_ctx*> ctx((size_t)P, nullptr);
      std::vector<cgb_partition_t> part((size_t)P);
      for (int r = 0; r < P; ++r) {
        check(cgb_partition_make(&st, P, r, 1, 0, 0, &part[(size_t)r]), "partition");
        const cgb_partition_t& p = part[(size_t)r];
        const int e0 = rowptr[(size_t)p.x0], e1 = rowptr[(size_t)(p.x0 + p.nx)];
        std::vector<int> rp((size_t)p.nx + 1);
        for (int i = 0; i <= p.nx; ++i) rp[(size_t)i] = rowptr[(size_t)(p.x0 + i)] - e0;
        check(cgb_ctx_create_csr_dist(n, &p, e1 - e0, rp.data(), colind.data() + e0, val.data() + e0, r, a.maxit + 1,
                                      &ctx[(size_t)r]), "device staging");
      }
      for (int r = 0; r < P; ++r)
        for (int q = 0; q < P; ++q)
          if (q != r) check(cgb_ctx_peer_attach_ptr(ctx[(size_t)r], q, cgb_ctx_arena_base(ctx[(size_t)q])), "peer attach");
      for (int r = 0; r < P; ++r) check(cgb_ctx_fabric_reset(ctx[(size_t)r]), "fabric reset");
      for (int r = 0; r < P; ++r)
        check(cgb_ctx_set_rhs(ctx[(size_t)r], b.data() + part[(size_t)r].x0, 0, nullptr), "memcpy from host to device (b)");
      cgb_solve_opts_t o;
      cgb_solve_opts_default(&o);
      o.tol = a.tol;
      o.tol_mode = a.tol_mode;
      o.maxit = a.maxit;
      o.chunk = a.chunk;
      o.variant = CGB_VAR_FUSED;
      o.use_graph = a.graph == 2 ? 1 : a.graph;
      for (int r = 0; r < P; ++r) check(cgb_solve_prepare(ctx[(size_t)r], &o), "solve prepare");
      for (int r = 0; r < P; ++r) check(cgb_solve_begin(ctx[(size_t)r], &o), "solve begin");
      int enq = 0;
      cgb_solve_result_t pr{};
      while (enq < a.maxit) {
        const int m = std::min(a.chunk, a.maxit - enq);
        for (int r = 0; r < P; ++r) check(cgb_solve_iterate(ctx[(size_t)r], m), "iterate");
        enq += m;
        check(cgb_solve_poll(ctx[0], enq - pr.iters > 2 * a.chunk ? 1 : 0, &pr), "poll");
        if (pr.status != 0) break;
      }
      double ms_max = 0.0;
      for (int r = P - 1; r >= 0; --r) {
        check(cgb_solve_finish(ctx[(size_t)r], r == 0 ? &res : &pr), "solve finish");
        ms_max = std::max(ms_max, r == 0 ? res.device_ms : pr.device_ms);
      }
      res.device_ms = ms_max;
      check(cgb_ctx_get_history(ctx[0], hist.data(), std::min(res.iters, a.maxit)), "history read-back");
      for (int r = 0; r < P; ++r) {
        check(cgb_ctx_get_x(ctx[(size_t)r], x.data() + part[(size_t)r].x0, 0, nullptr), "memcpy from device to host (x)");
        cgb_ctx_destroy(ctx[(size_t)r]);
      }
    } else {
      cgb_ctx* c = nullptr;
      check(cgb_ctx_create_csr(n, (int64_t)val.size(), rowptr.data(), colind.data(), val.data(), 0, a.maxit + 1, &c),
            "device staging");
      check(cgb_ctx_fabric_reset(c), "fabric reset");
      check(cgb_ctx_set_rhs(c, b.data(), 0, nullptr), "memcpy from host to device (b)");
      cgb_solve_opts_t o;
      cgb_solve_opts_default(&o);
      o.tol = a.tol;
      o.tol_mode = a.tol_mode;
      o.maxit = a.maxit;
      o.chunk = a.chunk;
      o.variant = a.variant == "pipelined" ? CGB_VAR_PIPELINED : CGB_VAR_FUSED;
      check(cgb_solve(c, &o, &res), "solve");
      check(cgb_ctx_get_x(c, x.data(), 0, nullptr), "memcpy from device to host (x)");
      check(cgb_ctx_get_history(c, hist.data(), std::min(res.iters, a.maxit)), "history read-back");
      cgb_ctx_destroy(c);
    }
    if (a.print_x < 0) a.print_x = 1;
  } else {
    int kind = a.problem == "poisson2d5" ? CGB_KIND_POISSON2D5
             : a.problem == "poisson3d7" ? CGB_KIND_POISSON3D7
             : a.problem == "stencil3d27" ? CGB_KIND_STENCIL3D27 : -1;
    if (kind < 0) die("unknown --problem " + a.problem);
    int nx = a.n[0] ? a.n[0] : 64, ny = a.n[1] ? a.n[1] : nx, nz = kind == CGB_KIND_POISSON2D5 ? 1 : (a.n[2] ? a.n[2] : nx);
    cgb_stencil_t st;
    check(cgb_stencil_default(kind, nx, ny, nz, &st), "problem setup");
    nx = st.gx;  // 2D grids come back as nx x 1 x ny
    ny = st.gy;
    nz = st.gz;
    n_total = (long long)nx * ny * nz;
    const int rk = rhs_kind(a.rhs);
    hist.assign((size_t)a.maxit + 2, 0.0);
    x.assign((size_t)n_total, 0.0);
    if (cpu) {
      std::vector<double> b((size_t)n_total);
      for (long long g = 0; g < n_total; ++g) b[(size_t)g] = cgb_rhs_value(&st, rk, a.seed, g);
      cgb_cpu_result_t cr{};
      check(cgb_cpu_cg_stencil(&st, a.gpus, a.decomp, a.Py, a.Pz, b.data(), x.data(), a.tol, a.tol_mode, a.maxit, 0,
                               hist.data(), &cr), "cpu cg");
      res.iters = cr.iters;
      res.status = cr.converged ? 1 : 2;
      res.r0norm = cr.r0norm;
      res.rnorm = cr.rnorm;
    } else {
      const int P = a.gpus;
      if (P < 1 || P > cgb_device_count()) die("--gpus exceeds the visible CUDA devices");
      std::vector<cgb_ctx*> ctx((size_t)P, nullptr);
      std::vector<cgb_partition_t> part((size_t)P);
      for (int r = 0; r < P; ++r) {
        check(cgb_partition_make(&st, P, r, a.decomp, a.Py, a.Pz, &part[(size_t)r]), "partition");
        check(cgb_ctx_create(&st, &part[(size_t)r], r, a.maxit + 2, &ctx[(size_t)r]), "device malloc");
      }
      for (int r = 0; r < P; ++r)
        for (int q = 0; q < P; ++q)
          if (q != r) check(cgb_ctx_peer_attach_ptr(ctx[(size_t)r], q, cgb_ctx_arena_base(ctx[(size_t)q])), "peer attach");
      for (int r = 0; r < P; ++r) check(cgb_ctx_fabric_reset(ctx[(size_t)r]), "fabric reset");
      for (int r = 0; r < P; ++r) check(cgb_ctx_set_rhs_generated(ctx[(size_t)r], rk, a.seed, nullptr), "rhs");
      if (!a.x0_in.empty()) {  // scatter the global initial guess to the ranks
        std::ifstream in(a.x0_in);
        if (!in) die("cannot open " + a.x0_in);
        std::vector<double> g((size_t)n_total);
        for (long long i = 0; i < n_total; ++i)
          if (!(in >> g[(size_t)i])) die("initial guess file is too short");
        for (int r = 0; r < P; ++r) {
          const cgb_partition_t& p = part[(size_t)r];
          std::vector<double> loc((size_t)p.nx * p.ny * p.nz);
          for (int k = 0; k < p.nz; ++k)
            for (int j = 0; j < p.ny; ++j)
              std::memcpy(&loc[(size_t)p.nx * ((size_t)j + (size_t)p.ny * (size_t)k)],
                          &g[(size_t)nx * ((size_t)(j + p.y0) + (size_t)ny * (size_t)(k + p.z0))], sizeof(double) * (size_t)p.nx);
          check(cgb_ctx_set_x0(ctx[(size_t)r], loc.data(), 0, nullptr), "memcpy from host to device (x0)");
        }
      }
      cgb_solve_opts_t o;
      cgb_solve_opts_default(&o);
      o.tol = a.tol;
      o.tol_mode = a.tol_mode;
      o.maxit = a.maxit;
      o.chunk = a.chunk;
      o.variant = a.variant == "pipelined" ? CGB_VAR_PIPELINED
                : a.variant == "fused" ? CGB_VAR_FUSED
                : a.variant == "recompute" ? CGB_VAR_MERGED
                : a.variant == "merged" ? CGB_VAR_MERGED2 : CGB_VAR_AUTO;
      o.use_graph = (P > 1 && a.graph == 2) ? 1 : a.graph;
      o.warm_start = a.x0_in.empty() ? 0 : 1;
      o.spmv_impl = a.spmv == "simple" ? CGB_SPMV_SIMPLE : a.spmv == "tma" ? CGB_SPMV_TMA : CGB_SPMV_AUTO;
      if (P == 1) {
        check(cgb_solve(ctx[0], &o, &res), "solve");
      } else {
        // one host thread drives all ranks: enqueue a chunk everywhere, then look at the flags
        for (int r = 0; r < P; ++r) check(cgb_solve_prepare(ctx[(size_t)r], &o), "solve prepare");
        for (int r = 0; r < P; ++r) check(cgb_solve_begin(ctx[(size_t)r], &o), "solve begin");
        const int target = a.maxit + (o.variant == CGB_VAR_PIPELINED ? 1 : 0);
        // The scalars are bit-identical on all ranks, so the host looks at rank 0 only, without blocking, and keeps
        // the GPUs up to two chunks ahead of what it has seen; no rank is ever synchronised inside the loop.
        int enq = 0;
        cgb_solve_result_t pr{};
        while (enq < target) {
          int m = std::min(a.chunk, target - enq);
          for (int r = 0; r < P; ++r) check(cgb_solve_iterate(ctx[(size_t)r], m), "iterate");
          enq += m;
          check(cgb_solve_poll(ctx[0], enq - pr.iters > 2 * a.chunk ? 1 : 0, &pr), "poll");
          if (pr.status != 0) break;
        }
        double ms_max = 0.0;  // device time of the loop: max over ranks
        for (int r = P - 1; r >= 0; --r) {
          check(cgb_solve_finish(ctx[(size_t)r], r == 0 ? &res : &pr), "solve finish");
          ms_max = std::max(ms_max, r == 0 ? res.device_ms : pr.device_ms);
        }
        res.device_ms = ms_max;
      }
      check(cgb_ctx_get_history(ctx[0], hist.data(), std::min(res.iters, a.maxit)), "history read-back");
      for (int r = 0; r < P; ++r) {  // ordered gather of the owned boxes
        const cgb_partition_t& p = part[(size_t)r];
        std::vector<double> loc((size_t)p.nx * p.ny * p.nz);
        check(cgb_ctx_get_x(ctx[(size_t)r], loc.data(), 0, nullptr), "memcpy from device to host (x)");
        for (int k = 0; k < p.nz; ++k)
          for (int j = 0; j < p.ny; ++j)
            std::memcpy(&x[(size_t)nx * ((size_t)(j + p.y0) + (size_t)ny * (size_t)(k + p.z0))],
                        &loc[(size_t)p.nx * ((size_t)j + (size_t)p.ny * (size_t)k)], sizeof(double) * (size_t)p.nx);
      }
      for (int r = 0; r < P; ++r) cgb_ctx_destroy(ctx[(size_t)r]);
    }
    if (a.print_x < 0) a.print_x = n_total <= 64 ? 1 : 0;
  }

  if (res.status < 0) die(std::string("solve failed: status ") + std::to_string(res.status));
  if (a.print_x)
    for (long long i = 0; i < n_total; ++i) std::printf("%f\n", x[(size_t)i]);  // CUDACG.cu:361-364
  if (!a.x_out.empty()) {
    FILE* f = std::fopen(a.x_out.c_str(), "w");
    if (!f) die("cannot open " + a.x_out);
    for (long long i = 0; i < n_total; ++i) std::fprintf(f, "%.17e\n", x[(size_t)i]);
    std::fclose(f);
  }
  if (!a.hist_out.empty()) {
    FILE* f = std::fopen(a.hist_out.c_str(), "w");
    if (!f) die("cannot open " + a.hist_out);
    for (int i = 0; i < res.iters && i < (int)hist.size(); ++i) std::fprintf(f, "%d %.17e\n", i + 1, hist[(size_t)i]);
    std::fclose(f);
  }
  if (!a.json_out.empty()) {
    FILE* f = std::fopen(a.json_out.c_str(), "w");
    if (!f) die("cannot open " + a.json_out);
    const double its = res.device_ms > 0 ? res.iters / (res.device_ms * 1e-3) : 0.0;
    std::fprintf(f,
                 "{\"problem\": \"%s\", \"n\": %lld, \"gpus\": %d, \"variant\": \"%s\", \"iters\": %d, \"status\": %d, "
                 "\"r0norm\": %.17e, \"rnorm\": %.17e, \"device_ms\": %.6f, \"iters_per_s\": %.3f, \"gdof_per_s\": %.4f, "
                 "\"launches\": %lld}\n",
                 a.problem.c_str(), n_total, a.gpus, a.variant.c_str(), res.iters, res.status, res.r0norm, res.rnorm,
                 res.device_ms, its, its * (double)n_total / 1e9, (long long)res.launches);
    std::fclose(f);
  }
  std::printf("Success\n");  // CUDACG.cu:365
  return 0;
}
