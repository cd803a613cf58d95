// Status / error plumbing. Replaces the reference's CLEANUP(msg)+return idiom (CUDACG.cu:10-33,
// 88-91): every failing call records a message and unwinds through RAII instead of a macro that
// frees a fixed list of locals.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>

#include "cgb200.h"

namespace cgb {

std::string& last_error_ref();
int fail(int code, const char* fmt, ...);

}  // namespace cgb

#define CGB_REQUIRE(cond, ...)                                         \
  do {                                                                 \
    if (!(cond)) return ::cgb::fail(CGB_ERR_INVALID, __VA_ARGS__);     \
  } while (0)

#define CGB_TRY(expr)                  \
  do {                                 \
    int _cgb_rc = (expr);              \
    if (_cgb_rc != 0) return _cgb_rc;  \
  } while (0)

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define CGB_CUDA(expr)                                                                     \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      return ::cgb::fail(CGB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                         __FILE__, __LINE__);                                              \
  } while (0)
#endif
