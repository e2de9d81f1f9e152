// Shared device/host helpers for the sm_100a kernel library.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace llmt {

// ---- error plumbing (C-ABI: functions return 0 on success, message via llmt_last_error) ----------
void set_error(const char* fmt, ...);
int check_launch(const char* what);
void count_launch();

#define LLMT_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      ::llmt::set_error(__VA_ARGS__);      \
      return 1;                            \
    }                                      \
  } while (0)

// cudaFuncSetAttribute is per device: launchers keep one "configured" flag per device ordinal
constexpr int kMaxDevices = 64;

// Sets the kernel's dynamic shared memory limit once per device.  `flags` is a static bool[kMaxDevices] owned by the
// calling launcher (one per kernel instantiation).  Returns 0 on success.
template <typename Kernel>
inline int ensure_smem_attr(Kernel kern, int bytes, bool* flags, const char* what) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("%s: device ordinal %d not supported (max %d)", what, dev, kMaxDevices);
    return 2;
  }
  if (!flags[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
      set_error("%s: cannot set %d B dynamic smem: %s", what, bytes, cudaGetErrorString(e));
      return 2;
    }
    flags[dev] = true;
  }
  return 0;
}

// ---- small device utilities ----------------------------------------------------------------------
constexpr int kWarp = 32;

struct __align__(16) bf16x8 {
  __nv_bfloat162 v[4];
};

__device__ __forceinline__ bf16x8 ld_bf16x8(const __nv_bfloat16* p) {
  bf16x8 r;
  *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(p);
  return r;
}
__device__ __forceinline__ void st_bf16x8(__nv_bfloat16* p, const bf16x8& r) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&r);
}
// streaming variants: data touched once, keep it out of L1
__device__ __forceinline__ bf16x8 ld_bf16x8_stream(const __nv_bfloat16* p) {
  bf16x8 r;
  uint4 u;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
               : "l"(p));
  *reinterpret_cast<uint4*>(&r) = u;
  return r;
}
__device__ __forceinline__ void unpack8(const bf16x8& r, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(r.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ bf16x8 pack8(const float (&f)[8]) {
  bf16x8 r;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}

// Attention dropout: 32 random bits for the four keys 4*kv_group .. 4*kv_group+3 of (head, query) — byte k decides
// key 4*kv_group + k.  A stateless integer hash (lowbias32 finaliser over a linear combination of the coordinates), so
// the forward and the backward kernel regenerate the same mask from (seed, head, query, key) with no stored state.
__device__ __forceinline__ uint32_t dropout_bits(uint32_t seed, int head, int q, int kv_group) {
  uint32_t x = seed + (uint32_t)head * 0xC2B2AE3Du + (uint32_t)q * 0x9E3779B1u + (uint32_t)kv_group * 0x85EBCA77u;
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; `red` is a __shared__ float[32]. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from the previous use
  if (lane == 0) red[warp] = v;
  __syncthreads();
  v = (lane < nwarp) ? red[lane] : 0.f;
  return warp_sum(v);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  v = (lane < nwarp) ? red[lane] : -INFINITY;
  return warp_max(v);
}

// CTAs per SM used by the optimizer-update kernels (llmt_set_update_ctas_per_sm): 8 when they own the GPU, 2 when
// they run in the background of the next forward pass so GEMM / attention CTAs can share every SM with them.
int update_ctas_per_sm();

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace llmt
