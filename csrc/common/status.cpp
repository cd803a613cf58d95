#include "status.hpp"

namespace cgb {

std::string& last_error_ref() {
  static thread_local std::string msg;
  return msg;
}

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return code;
}

}  // namespace cgb

extern "C" const char* cgb_last_error(void) { return cgb::last_error_ref().c_str(); }
extern "C" const char* cgb_version(void) { return "cgb200 0.1 (sm_100a)"; }
