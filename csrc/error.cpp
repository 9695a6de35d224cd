// Thread-local last-error text for the C ABI.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

static thread_local char g_err[512] = "";

void detb200_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" __attribute__((visibility("default"))) const char* detb200_last_error() { return g_err; }
extern "C" __attribute__((visibility("default"))) int detb200_abi_version() { return 1; }
