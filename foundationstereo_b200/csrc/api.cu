// Library-wide entry points: version, thread-local error text, launch counter.
#include "common.cuh"
#include <stdlib.h>

namespace fsb {
static thread_local char g_err[512] = "";
static thread_local uint64_t g_launches = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches += (uint64_t)n; }

static int g_pdl = -1;  // -1: read FSB_PDL from the environment on first use (default on)
bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("FSB_PDL");
    g_pdl = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return g_pdl != 0;
}
}  // namespace fsb

extern "C" {
int fsb_version(void) { return FSB_VERSION; }
const char* fsb_last_error(void) { return fsb::g_err; }
uint64_t fsb_launch_count(void) { return fsb::g_launches; }
size_t fsb_conv_args_size(void) { return sizeof(fsb_conv_args); }
int fsb_set_pdl(int enable) {
  const int old = fsb::pdl_enabled() ? 1 : 0;
  fsb::g_pdl = enable ? 1 : 0;
  return old;
}
}
