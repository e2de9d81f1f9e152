#include "common.cuh"

#include <stdarg.h>
#include <string.h>

namespace llmt {

static thread_local char g_error[512] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();  // clear
    set_error("%s: %s", what, cudaGetErrorString(e));
    return 2;
  }
  return 0;
}

}  // namespace llmt

extern "C" const char* llmt_last_error() { return llmt::g_error; }

extern "C" int llmt_abi_version() { return 1; }
