// Thread-local error message behind pi0_last_error(); every C-ABI entry returns 0 or a negative code.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace pi0 {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PI0_PDL");
    v = (e != nullptr && strcmp(e, "0") == 0) ? 0 : 1;
  }
  return v == 1;
}

}  // namespace pi0
