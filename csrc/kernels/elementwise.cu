// SiLU*mul (SwiGLU gate), RoPE, embedding gather (+NEFTune noise), embedding scatter-add.
//
// Replaces: Liger's Triton silu_mul (reference: src/llm_training/ops/liger_kernel/swiglu_op.py:36-39),
// the ~8-kernel torch RoPE (src/llm_training/ops/rope_op.py:10-20 and the cos/sin gather at
// models/llama/llama_model.py:389-412), ATen embedding and the NEFTune forward hook
// (src/llm_training/lms/clm/clm.py:45-78).  All are bandwidth-bound: 16-byte vector accesses,
// grid sized to fill 148 SMs.
#include "common.cuh"

namespace llmt {

// ---------------------------------------------------------------------------------------------
// SiLU * mul.  gate / up / out are row-strided so gate and up can be the two halves of one
// [T, 2F] GEMM output and d_gate / d_up the two halves of one [T, 2F] buffer for the dgrad GEMM.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
silu_mul_fwd_kernel(const __nv_bfloat16* __restrict__ gate, const __nv_bfloat16* __restrict__ up,
                    __nv_bfloat16* __restrict__ out, int rows, int cols, int64_t in_stride,
                    int64_t out_stride) {
  // flat grid-stride walk over 16-byte vectors, TWO vectors per operand in flight per thread; 32-bit index math
  // (the launcher guarantees rows * cols / 8 < 2^31)
  const uint32_t vec_per_row = (uint32_t)cols >> 3;
  const uint32_t total = (uint32_t)rows * vec_per_row;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += 2 * stride) {
    const uint32_t i1 = i + stride;
    const bool two = i1 < total;
    const uint32_t r0 = i / vec_per_row, c0 = (i - r0 * vec_per_row) << 3;
    const uint32_t r1 = two ? i1 / vec_per_row : r0, c1 = two ? (i1 - r1 * vec_per_row) << 3 : c0;
    const bf16x8 ga = ld_bf16x8_stream(gate + (int64_t)r0 * in_stride + c0);
    const bf16x8 ua = ld_bf16x8_stream(up + (int64_t)r0 * in_stride + c0);
    const bf16x8 gb = ld_bf16x8_stream(gate + (int64_t)r1 * in_stride + c1);
    const bf16x8 ub = ld_bf16x8_stream(up + (int64_t)r1 * in_stride + c1);
    float g[8], u[8], o[8];
    unpack8(ga, g); unpack8(ua, u);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = __fdividef(g[j], 1.f + __expf(-g[j])) * u[j];
    st_bf16x8(out + (int64_t)r0 * out_stride + c0, pack8(o));
    if (two) {
      unpack8(gb, g); unpack8(ub, u);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = __fdividef(g[j], 1.f + __expf(-g[j])) * u[j];
      st_bf16x8(out + (int64_t)r1 * out_stride + c1, pack8(o));
    }
  }
}

// kAct: also write act = silu(gate) * up (contiguous [rows, cols]) — the recomputed input of the down projection's
// wgrad GEMM — in the same pass over gate / up, instead of a second kernel that reads them again.
template <bool kAct>
__global__ void __launch_bounds__(256)
silu_mul_bwd_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ gate,
                    const __nv_bfloat16* __restrict__ up, __nv_bfloat16* __restrict__ dgate,
                    __nv_bfloat16* __restrict__ dup, __nv_bfloat16* __restrict__ act, int rows, int cols,
                    int64_t in_stride, int64_t dout_stride, int64_t dgu_stride) {
  const int vec_per_row = cols >> 3;
  const int64_t total = (int64_t)rows * vec_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / vec_per_row;
    const int c = (int)(i - r * vec_per_row) << 3;
    float g[8], u[8], d[8], dg[8], du[8], a[8];
    unpack8(ld_bf16x8_stream(gate + r * in_stride + c), g);
    unpack8(ld_bf16x8_stream(up + r * in_stride + c), u);
    unpack8(ld_bf16x8_stream(dout + r * dout_stride + c), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sig = 1.f / (1.f + __expf(-g[j]));
      const float silu = g[j] * sig;
      du[j] = d[j] * silu;
      dg[j] = d[j] * u[j] * sig * (1.f + g[j] * (1.f - sig));
      if constexpr (kAct) a[j] = __fdividef(g[j], 1.f + __expf(-g[j])) * u[j];   // bit-identical to the forward kernel
    }
    st_bf16x8(dgate + r * dgu_stride + c, pack8(dg));
    st_bf16x8(dup + r * dgu_stride + c, pack8(du));
    if constexpr (kAct) st_bf16x8(act + r * (int64_t)cols + c, pack8(a));
  }
}

// ---------------------------------------------------------------------------------------------
// RoPE, rotate-half convention, in place on a packed [T, heads, D] view with a row stride (so it
// can run directly on the q|k part of the fused QKV GEMM output).  cos/sin are fp32 tables
// [max_pos, D/2] indexed by position id.  `sign` = +1 forward, -1 backward (inverse rotation).
//   out[:D/2] = x1*cos - x2*sin ; out[D/2:] = x2*cos + x1*sin
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rope_inplace_kernel(__nv_bfloat16* __restrict__ x, const int* __restrict__ pos,
                    const float* __restrict__ cos_t, const float* __restrict__ sin_t, int tokens, int heads,
                    int head_dim, int64_t row_stride, float sign) {
  const int half = head_dim >> 1;
  const int vec_per_head = half >> 3;  // 8-element vectors in one half
  const int64_t total = (int64_t)tokens * heads * vec_per_head;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % vec_per_head);
    const int64_t th = i / vec_per_head;
    const int h = (int)(th % heads);
    const int64_t t = th / heads;
    const int p = pos[t];
    __nv_bfloat16* base = x + t * row_stride + (int64_t)h * head_dim + v * 8;
    float x1[8], x2[8], o1[8], o2[8];
    unpack8(ld_bf16x8(base), x1);
    unpack8(ld_bf16x8(base + half), x2);
    const float4* c4 = reinterpret_cast<const float4*>(cos_t + (int64_t)p * half + v * 8);
    const float4* s4 = reinterpret_cast<const float4*>(sin_t + (int64_t)p * half + v * 8);
    const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
    const float c[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
    const float s[8] = {sa.x, sa.y, sa.z, sa.w, sb.x, sb.y, sb.z, sb.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sn = s[j] * sign;
      o1[j] = x1[j] * c[j] - x2[j] * sn;
      o2[j] = x2[j] * c[j] + x1[j] * sn;
    }
    st_bf16x8(base, pack8(o1));
    st_bf16x8(base + half, pack8(o2));
  }
}

// ---------------------------------------------------------------------------------------------
// Embedding gather with optional NEFTune noise:
//   out[t, :] = W[ids[t], :] + U(-1,1) * noise_scale[t]
// noise_scale[t] = alpha / sqrt(len(row of t) * d) for supervised (non-pad) tokens, 0 for pad; it
// is computed once per batch by the caller.  Noise comes from a counter-based Philox4x32-10
// keyed by (seed, element index) so results do not depend on the launch geometry.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
__device__ __forceinline__ float u32_to_pm1(uint32_t r) {  // uniform in (-1, 1)
  return ((float)(r >> 8) + 0.5f) * (2.0f / 16777216.0f) - 1.0f;
}

__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ weight,
                     __nv_bfloat16* __restrict__ out, const float* __restrict__ noise_scale, int tokens,
                     int dim, int vocab, uint64_t seed) {
  const int vec_per_row = dim >> 3;
  const int64_t total = (int64_t)tokens * vec_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / vec_per_row;
    const int c = (int)(i - t * vec_per_row) << 3;
    int64_t id = ids[t];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    bf16x8 w = ld_bf16x8(weight + id * dim + c);
    if (noise_scale != nullptr) {
      const float ns = noise_scale[t];
      if (ns != 0.f) {
        float f[8];
        unpack8(w, f);
        const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
        const uint4 r0 = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), 0u, 0u), key);
        const uint4 r1 = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), 1u, 0u), key);
        const uint32_t r[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] += u32_to_pm1(r[j]) * ns;
        w = pack8(f);
      }
    }
    st_bf16x8(out + t * dim + c, w);
  }
}

// dW[ids[t], :] += dout[t, :]  (fp32 or bf16 gradient buffer, atomics; padding_idx rows skipped)
template <typename GradT>
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dout,
                     GradT* __restrict__ dweight, int tokens, int dim, int vocab, int64_t padding_idx) {
  const int pair_per_row = dim >> 1;
  const int64_t total = (int64_t)tokens * pair_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / pair_per_row;
    const int c = (int)(i - t * pair_per_row) << 1;
    const int64_t id = ids[t];
    if (id < 0 || id >= vocab || id == padding_idx) continue;
    const __nv_bfloat162 d = *reinterpret_cast<const __nv_bfloat162*>(dout + t * dim + c);
    if constexpr (sizeof(GradT) == 4) {
      const float2 f = __bfloat1622float2(d);
      atomicAdd(reinterpret_cast<float*>(dweight) + id * dim + c, f.x);
      atomicAdd(reinterpret_cast<float*>(dweight) + id * dim + c + 1, f.y);
    } else {
      atomicAdd(reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(dweight) + id * dim + c), d);
    }
  }
}

// Order-independent variant: the caller passes the token ids sorted (stable) with the permutation that sorted them.
// The CTA of sorted position i owns the whole run of equal ids that STARTS at i: it sums the run's dout rows in
// fp32, in run order, and writes the bf16 row once.  No atomics, so the result does not depend on scheduling; rows
// no token selects are left as the caller initialised them (zero).
__global__ void __launch_bounds__(256)
embedding_bwd_sorted_kernel(const int64_t* __restrict__ sorted_ids, const int64_t* __restrict__ perm,
                            const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dweight, int tokens,
                            int dim, int vocab, int64_t padding_idx) {
  const int i = blockIdx.x;
  const int64_t id = sorted_ids[i];
  if (id < 0 || id >= vocab || id == padding_idx) return;
  if (i > 0 && sorted_ids[i - 1] == id) return;          // not the first token of its run
  for (int v = threadIdx.x; v < (dim >> 3); v += blockDim.x) {
    float acc[8];
    unpack8(ld_bf16x8(dout + perm[i] * dim + v * 8), acc);
    for (int j = i + 1; j < tokens && sorted_ids[j] == id; ++j) {
      float f[8];
      unpack8(ld_bf16x8(dout + perm[j] * dim + v * 8), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += f[e];
    }
    st_bf16x8(dweight + id * dim + v * 8, pack8(acc));
  }
}

static int ew_grid(int64_t work_items, int threads) {
  int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)num_sms() * 8;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

}  // namespace llmt

using namespace llmt;

extern "C" int llmt_silu_mul_fwd(const void* gate, const void* up, void* out, int rows, int cols,
                                 int64_t in_stride, int64_t out_stride, cudaStream_t stream) {
  LLMT_REQUIRE(cols % 8 == 0 && in_stride % 8 == 0 && out_stride % 8 == 0, "silu_mul: cols/strides must be multiples of 8");
  if (rows == 0) return 0;
  LLMT_REQUIRE((int64_t)rows * (cols / 8) < (1ll << 31), "silu_mul: too many elements for 32-bit vector indexing");
  silu_mul_fwd_kernel<<<ew_grid((int64_t)rows * cols / 16, 256), 256, 0, stream>>>(
      (const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (__nv_bfloat16*)out, rows, cols, in_stride, out_stride);
  return check_launch("silu_mul_fwd");
}

extern "C" int llmt_silu_mul_bwd(const void* dout, const void* gate, const void* up, void* dgate, void* dup,
                                 int rows, int cols, int64_t in_stride, int64_t dout_stride, int64_t dgu_stride,
                                 cudaStream_t stream) {
  LLMT_REQUIRE(cols % 8 == 0 && in_stride % 8 == 0 && dout_stride % 8 == 0 && dgu_stride % 8 == 0,
               "silu_mul_bwd: cols/strides must be multiples of 8");
  if (rows == 0) return 0;
  silu_mul_bwd_kernel<false><<<ew_grid((int64_t)rows * cols / 8, 256), 256, 0, stream>>>(
      (const __nv_bfloat16*)dout, (const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (__nv_bfloat16*)dgate,
      (__nv_bfloat16*)dup, nullptr, rows, cols, in_stride, dout_stride, dgu_stride);
  return check_launch("silu_mul_bwd");
}

// Backward of silu(gate) * up that ALSO returns the forward product act = silu(gate) * up (contiguous [rows, cols]):
// callers that recompute act for the down projection's wgrad get it from the pass that already reads gate / up.
extern "C" int llmt_silu_mul_bwd_act(const void* dout, const void* gate, const void* up, void* dgate, void* dup,
                                     void* act, int rows, int cols, int64_t in_stride, int64_t dout_stride,
                                     int64_t dgu_stride, cudaStream_t stream) {
  LLMT_REQUIRE(cols % 8 == 0 && in_stride % 8 == 0 && dout_stride % 8 == 0 && dgu_stride % 8 == 0,
               "silu_mul_bwd_act: cols/strides must be multiples of 8");
  if (rows == 0) return 0;
  silu_mul_bwd_kernel<true><<<ew_grid((int64_t)rows * cols / 8, 256), 256, 0, stream>>>(
      (const __nv_bfloat16*)dout, (const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (__nv_bfloat16*)dgate,
      (__nv_bfloat16*)dup, (__nv_bfloat16*)act, rows, cols, in_stride, dout_stride, dgu_stride);
  return check_launch("silu_mul_bwd_act");
}

extern "C" int llmt_rope_inplace(void* x, const int* pos, const float* cos_t, const float* sin_t, int tokens,
                                 int heads, int head_dim, int64_t row_stride, int backward, cudaStream_t stream) {
  LLMT_REQUIRE(head_dim % 16 == 0 && row_stride % 8 == 0, "rope: head_dim must be a multiple of 16 (got %d)", head_dim);
  if (tokens == 0 || heads == 0) return 0;
  rope_inplace_kernel<<<ew_grid((int64_t)tokens * heads * head_dim / 16, 256), 256, 0, stream>>>(
      (__nv_bfloat16*)x, pos, cos_t, sin_t, tokens, heads, head_dim, row_stride, backward ? -1.f : 1.f);
  return check_launch("rope_inplace");
}

extern "C" int llmt_embedding_fwd(const int64_t* ids, const void* weight, void* out, const float* noise_scale,
                                  int tokens, int dim, int vocab, uint64_t seed, cudaStream_t stream) {
  LLMT_REQUIRE(dim % 8 == 0, "embedding: dim must be a multiple of 8");
  if (tokens == 0) return 0;
  embedding_fwd_kernel<<<ew_grid((int64_t)tokens * dim / 8, 256), 256, 0, stream>>>(
      ids, (const __nv_bfloat16*)weight, (__nv_bfloat16*)out, noise_scale, tokens, dim, vocab, seed);
  return check_launch("embedding_fwd");
}

extern "C" int llmt_embedding_bwd(const int64_t* ids, const void* dout, void* dweight, int tokens, int dim,
                                  int vocab, int64_t padding_idx, int grad_is_fp32, cudaStream_t stream) {
  LLMT_REQUIRE(dim % 2 == 0, "embedding_bwd: dim must be even");
  if (tokens == 0) return 0;
  const int grid = ew_grid((int64_t)tokens * dim / 2, 256);
  if (grad_is_fp32)
    embedding_bwd_kernel<float><<<grid, 256, 0, stream>>>(ids, (const __nv_bfloat16*)dout, (float*)dweight, tokens,
                                                          dim, vocab, padding_idx);
  else
    embedding_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(ids, (const __nv_bfloat16*)dout,
                                                                  (__nv_bfloat16*)dweight, tokens, dim, vocab,
                                                                  padding_idx);
  return check_launch("embedding_bwd");
}

// dweight (bf16 [vocab, dim]) is fully overwritten: zero, then one row per distinct token id.
extern "C" int llmt_embedding_bwd_sorted(const int64_t* sorted_ids, const int64_t* perm, const void* dout,
                                         void* dweight, int tokens, int dim, int vocab, int64_t padding_idx,
                                         cudaStream_t stream) {
  LLMT_REQUIRE(dim % 8 == 0, "embedding_bwd: dim must be a multiple of 8");
  cudaError_t e = cudaMemsetAsync(dweight, 0, sizeof(__nv_bfloat16) * (size_t)vocab * dim, stream);
  LLMT_REQUIRE(e == cudaSuccess, "embedding_bwd: clearing the gradient failed: %s", cudaGetErrorString(e));
  if (tokens == 0) return 0;
  embedding_bwd_sorted_kernel<<<tokens, 256, 0, stream>>>(sorted_ids, perm, (const __nv_bfloat16*)dout,
                                                          (__nv_bfloat16*)dweight, tokens, dim, vocab, padding_idx);
  return check_launch("embedding_bwd_sorted");
}
