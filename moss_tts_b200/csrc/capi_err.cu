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

#include <stdlib.h>
namespace mt {
int exp_env(const char* name, int dflt) {
#ifdef MT_EXPERIMENTS
  const char* e = getenv(name);
  return e && e[0] ? atoi(e) : dflt;
#else
  (void)name;
  return dflt;
#endif
}
static int g_pdl_state = -1;   // -1 = first use (an MT_EXPERIMENTS build reads MT_NO_PDL); mt_debug_set_pdl overrides
bool pdl_enabled() {
  if (g_pdl_state < 0) g_pdl_state = exp_env("MT_NO_PDL", 0) ? 0 : 1;
  return g_pdl_state == 1;
}
void set_pdl(int on) { g_pdl_state = on ? 1 : 0; }
}

namespace mt {
static unsigned long long* g_trace_buf = nullptr;
static long long g_trace_cap = 0, g_trace_next = 0;
void set_trace(void* buf, long long words) {
  g_trace_buf = reinterpret_cast<unsigned long long*>(buf); g_trace_cap = buf ? words : 0; g_trace_next = 0;
}
long long trace_used() { return g_trace_next; }
unsigned long long* trace_slot(int words) {
  if (!g_trace_buf || g_trace_next + words > g_trace_cap) return nullptr;
  unsigned long long* p = g_trace_buf + g_trace_next;
  g_trace_next += words;
  return p;
}
}
