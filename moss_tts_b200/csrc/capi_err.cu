#include "common.cuh"
#include <stdarg.h>
namespace mt {
static thread_local std::string g_err;
void set_error(const std::string& m) { g_err = m; }
int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  g_err = buf; return code;
}
const char* last_error() { return g_err.c_str(); }
}
