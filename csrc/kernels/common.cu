#include "common.cuh"

#include <stdarg.h>
#include <string.h>

namespace llmt {

static thread_local char g_error[512] = {0};
static unsigned long long g_launches = 0;  // kernels launched by this library (all threads)

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED);
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

// Number of kernels this library has launched since load (bench.py reports the delta as `gpu_launches`).
extern "C" unsigned long long llmt_launch_count() { return llmt::g_launches; }
namespace llmt { void count_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); } }

namespace llmt {
static int g_update_ctas_per_sm = 8;
int update_ctas_per_sm() { return g_update_ctas_per_sm; }
}  // namespace llmt

extern "C" void llmt_set_update_ctas_per_sm(int ctas) {
  llmt::g_update_ctas_per_sm = ctas < 1 ? 1 : (ctas > 8 ? 8 : ctas);
}
