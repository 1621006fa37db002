// Error plumbing of libinksam: a per-thread message buffer behind inksam_last_error().
#include "common.h"

#include <cstdarg>
#include <cstdio>

namespace inksam {

static thread_local char g_err[1024] = "";

int set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return -1;
}

const char* get_error() { return g_err; }

}  // namespace inksam
