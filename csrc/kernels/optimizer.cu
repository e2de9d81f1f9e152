// Fused AdamW over a flat shard with fp32 master weights, plus the squared-gradient-norm
// reduction that feeds gradient clipping — both without a host round trip.
//
//   g      = grad * grad_scale * clip_coef           clip_coef = min(1, max_norm / (sqrt(norm_sq) + 1e-6))
//   master = master * (1 - lr*wd) - lr/bc1 * m' / (sqrt(v')/sqrt(bc2) + eps)
//   param  = bf16(master)
//
// Replaces MasterWeightsOptimizer.step (reference: src/llm_training/optim/master_weight_wrapper.py:63-80:
// per-parameter bf16->fp32 grad copy, torch AdamW, per-parameter fp32->bf16 copy), DeepSpeed's
// FusedAdam (config/examples/llama-3.1/llama-3.1-8b_pt_example.yaml:44) and the clip_grad_norm_
// pass (src/llm_training/lightning/strategy/fsdp2/fsdp2_precision.py:165-169).  One read of
// grad/master/m/v and one write of master/m/v/param per element: 22 B/element at bf16 grads.
#include "common.cuh"

namespace llmt {

struct AdamWArgs {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2, grad_scale, max_norm;
};

template <typename GradT, bool kWriteParam>
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
             const GradT* __restrict__ grad, __nv_bfloat16* __restrict__ param, int64_t n, AdamWArgs a,
             const float* __restrict__ norm_sq) {
  float gscale = a.grad_scale;
  if (norm_sq != nullptr && a.max_norm > 0.f) {
    const float norm = sqrtf(*norm_sq) * fabsf(a.grad_scale);
    gscale *= fminf(1.f, a.max_norm / (norm + 1e-6f));
  }
  const float decay = 1.f - a.lr * a.weight_decay;
  const float step_size = a.lr / a.bc1;
  const float inv_sqrt_bc2 = rsqrtf(a.bc2);
  const int64_t nvec = n >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    float4 p4 = reinterpret_cast<float4*>(master)[i];
    float4 m4 = reinterpret_cast<float4*>(m)[i];
    float4 v4 = reinterpret_cast<float4*>(v)[i];
    float g[4];
    if constexpr (sizeof(GradT) == 4) {
      const float4 g4 = reinterpret_cast<const float4*>(grad)[i];
      g[0] = g4.x; g[1] = g4.y; g[2] = g4.z; g[3] = g4.w;
    } else {
      const uint2 raw = reinterpret_cast<const uint2*>(grad)[i];
      const float2 a0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
      const float2 a1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
      g[0] = a0.x; g[1] = a0.y; g[2] = a1.x; g[3] = a1.y;
    }
    float* pp = reinterpret_cast<float*>(&p4);
    float* mm = reinterpret_cast<float*>(&m4);
    float* vv = reinterpret_cast<float*>(&v4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = g[j] * gscale;
      mm[j] = a.beta1 * mm[j] + (1.f - a.beta1) * gj;
      vv[j] = a.beta2 * vv[j] + (1.f - a.beta2) * gj * gj;
      const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + a.eps;
      pp[j] = pp[j] * decay - step_size * (mm[j] / denom);
    }
    reinterpret_cast<float4*>(master)[i] = p4;
    reinterpret_cast<float4*>(m)[i] = m4;
    reinterpret_cast<float4*>(v)[i] = v4;
    if constexpr (kWriteParam) {
      uint2 o;
      *reinterpret_cast<__nv_bfloat162*>(&o.x) = __floats2bfloat162_rn(pp[0], pp[1]);
      *reinterpret_cast<__nv_bfloat162*>(&o.y) = __floats2bfloat162_rn(pp[2], pp[3]);
      reinterpret_cast<uint2*>(param)[i] = o;
    }
  }
  // scalar tail
  for (int64_t i = (nvec << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float gj = (float)grad[i] * gscale;
    const float mj = a.beta1 * m[i] + (1.f - a.beta1) * gj;
    const float vj = a.beta2 * v[i] + (1.f - a.beta2) * gj * gj;
    const float pj = master[i] * decay - step_size * (mj / (sqrtf(vj) * inv_sqrt_bc2 + a.eps));
    m[i] = mj; v[i] = vj; master[i] = pj;
    if constexpr (kWriteParam) param[i] = __float2bfloat16(pj);
  }
}

// partial[b] = sum of squares of this CTA's slice; deterministic two-stage reduction.
template <typename T>
__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const T* __restrict__ x, int64_t n, float* __restrict__ partial) {
  __shared__ float red[32];
  float s = 0.f;
  if constexpr (sizeof(T) == 2) {
    const int64_t nvec = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
         i += (int64_t)gridDim.x * blockDim.x) {
      float f[8];
      unpack8(ld_bf16x8_stream(reinterpret_cast<const __nv_bfloat16*>(x) + i * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += f[j] * f[j];
    }
    for (int64_t i = (nvec << 3) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
      const float f = (float)x[i];
      s += f * f;
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
      const float f = (float)x[i];
      s += f * f;
    }
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

__global__ void __launch_bounds__(1024)
sumsq_final_kernel(const float* __restrict__ partial, int nparts, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) s += partial[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + s : s;
}

}  // namespace llmt

using namespace llmt;

extern "C" int llmt_sumsq_num_partials() { return num_sms() * 4; }

// out[0] (+)= sum(x^2). `partial` is fp32 scratch of llmt_sumsq_num_partials() floats.
extern "C" int llmt_sumsq(const void* x, int64_t n, int is_fp32, float* partial, float* out, int accumulate,
                          cudaStream_t stream) {
  const int grid = llmt_sumsq_num_partials();
  if (is_fp32)
    sumsq_partial_kernel<float><<<grid, 256, 0, stream>>>((const float*)x, n, partial);
  else
    sumsq_partial_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)x, n, partial);
  int rc = check_launch("sumsq_partial");
  if (rc) return rc;
  sumsq_final_kernel<<<1, 1024, 0, stream>>>(partial, grid, out, accumulate);
  return check_launch("sumsq_final");
}

// norm_sq: optional device pointer to the global squared grad norm (already all-reduced); with
// max_norm > 0 the clip coefficient is derived in-kernel. param may be null (fp32 training).
extern "C" int llmt_adamw(float* master, float* m, float* v, const void* grad, void* param_bf16, int64_t n,
                          int grad_is_fp32, float lr, float beta1, float beta2, float eps, float weight_decay,
                          int step, float grad_scale, float max_norm, const float* norm_sq, cudaStream_t stream) {
  LLMT_REQUIRE(step >= 1, "adamw: step must be >= 1");
  if (n == 0) return 0;
  LLMT_REQUIRE(((uintptr_t)master % 16 == 0) && ((uintptr_t)m % 16 == 0) && ((uintptr_t)v % 16 == 0) &&
                   ((uintptr_t)grad % 16 == 0) && ((uintptr_t)param_bf16 % 8 == 0),
               "adamw: buffers must be 16-byte aligned");
  AdamWArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));   // in double: 1 - beta^step cancels badly in fp32
  a.bc2 = (float)(1.0 - pow((double)beta2, (double)step));
  a.grad_scale = grad_scale; a.max_norm = max_norm;
  int64_t blocks = (n / 4 + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * update_ctas_per_sm();
  const int grid = (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
  auto P = (__nv_bfloat16*)param_bf16;
  if (grad_is_fp32) {
    if (P) adamw_kernel<float, true><<<grid, 256, 0, stream>>>(master, m, v, (const float*)grad, P, n, a, norm_sq);
    else   adamw_kernel<float, false><<<grid, 256, 0, stream>>>(master, m, v, (const float*)grad, P, n, a, norm_sq);
  } else {
    if (P) adamw_kernel<__nv_bfloat16, true><<<grid, 256, 0, stream>>>(master, m, v, (const __nv_bfloat16*)grad, P, n, a, norm_sq);
    else   adamw_kernel<__nv_bfloat16, false><<<grid, 256, 0, stream>>>(master, m, v, (const __nv_bfloat16*)grad, P, n, a, norm_sq);
  }
  return check_launch("adamw");
}
