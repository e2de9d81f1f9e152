// Softmax cross-entropy / token log-prob over bf16 logits [T, V] (row-strided), fp32 math.
//
// Forward (one CTA per row, online softmax, 16-byte loads):
//     lse[t]      = logsumexp(logits[t, :])
//     tgt[t]      = logits[t, label[t]]            (0 when label == ignore_index)
//   loss_t = lse - tgt, logp_t = tgt - lse.
// Backward, IN PLACE on the logits buffer (no second [T, V] tensor):
//     logits[t, j] <- (softmax_j - [j == label]) * row_scale[t]
//   where row_scale[t] is a per-row fp32 factor computed on device by the caller
//   (mean-CE: grad_out / n_valid; DPO/ORPO: -dL/dlogp_t), 0 for ignored rows.
//
// Replaces Liger's Triton CE on fp32 logits (reference: src/llm_training/ops/liger_kernel/
// cross_entropy_op.py:29-33; logits.float() at lms/clm/clm.py:147 materialises 4.2 GB at T=8192)
// and the [B,T,V] log_softmax + gather of DPO/ORPO (lms/dpo/dpo.py:110-114, lms/orpo/orpo.py:89-93).
#include "common.cuh"

namespace llmt {

constexpr int kCeThreads = 512;

__global__ void __launch_bounds__(kCeThreads)
ce_fwd_kernel(const __nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
              float* __restrict__ lse_out, float* __restrict__ tgt_out, int vocab, int64_t row_stride,
              int64_t ignore_index, float softcap) {
  __shared__ float red[32];
  const int64_t row = blockIdx.x;
  const __nv_bfloat16* lr = logits + row * row_stride;
  const int nvec = vocab >> 3;
  float m = -INFINITY, s = 0.f;
  for (int v = threadIdx.x; v < nvec; v += kCeThreads) {
    float f[8];
    unpack8(ld_bf16x8(lr + (size_t)v * 8), f);
    float lm = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (softcap > 0.f) f[j] = softcap * tanhf(f[j] / softcap);
      lm = fmaxf(lm, f[j]);
    }
    const float nm = fmaxf(m, lm);
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += __expf(f[j] - nm);
    s = s * __expf(m - nm) + acc;
    m = nm;
  }
  for (int c = (nvec << 3) + threadIdx.x; c < vocab; c += kCeThreads) {  // tail (vocab % 8)
    float f = __bfloat162float(lr[c]);
    if (softcap > 0.f) f = softcap * tanhf(f / softcap);
    const float nm = fmaxf(m, f);
    s = s * __expf(m - nm) + __expf(f - nm);
    m = nm;
  }
  const float gm = block_max(m, red);
  s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
  const float gs = block_sum(s, red);
  if (threadIdx.x == 0) {
    const int64_t label = labels[row];
    lse_out[row] = gm + logf(gs);
    float t = 0.f;
    if (label != ignore_index && label >= 0 && label < vocab) {
      t = __bfloat162float(lr[label]);
      if (softcap > 0.f) t = softcap * tanhf(t / softcap);
    }
    tgt_out[row] = t;
  }
}

__global__ void __launch_bounds__(kCeThreads)
ce_bwd_inplace_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
                      const float* __restrict__ lse, const float* __restrict__ row_scale, int vocab,
                      int64_t row_stride, int64_t ignore_index, float softcap) {
  const int64_t row = blockIdx.x;
  __nv_bfloat16* lr = logits + row * row_stride;
  const int nvec = vocab >> 3;
  // ignored rows get a zero gradient; a label outside [0, vocab) (another rank's vocabulary shard under
  // vocab parallelism) simply has no one-hot term on this rank
  const int64_t label = labels[row];
  const float scale = (label != ignore_index) ? row_scale[row] : 0.f;
  const float l = lse[row];
  if (scale == 0.f) {  // ignored row: zero gradient
    bf16x8 z;
#pragma unroll
    for (int i = 0; i < 4; ++i) z.v[i] = __floats2bfloat162_rn(0.f, 0.f);
    for (int v = threadIdx.x; v < nvec; v += kCeThreads) st_bf16x8(lr + (size_t)v * 8, z);
    for (int c = (nvec << 3) + threadIdx.x; c < vocab; c += kCeThreads) lr[c] = __float2bfloat16(0.f);
    return;
  }
  for (int v = threadIdx.x; v < nvec; v += kCeThreads) {
    float f[8];
    unpack8(ld_bf16x8(lr + (size_t)v * 8), f);
    const int c0 = v << 3;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float z = f[j], dcap = 1.f;
      if (softcap > 0.f) {
        const float th = tanhf(z / softcap);
        z = softcap * th;
        dcap = 1.f - th * th;
      }
      float g = __expf(z - l);
      if (c0 + j == label) g -= 1.f;
      f[j] = g * scale * dcap;
    }
    st_bf16x8(lr + (size_t)v * 8, pack8(f));
  }
  for (int c = (nvec << 3) + threadIdx.x; c < vocab; c += kCeThreads) {
    float z = __bfloat162float(lr[c]), dcap = 1.f;
    if (softcap > 0.f) {
      const float th = tanhf(z / softcap);
      z = softcap * th;
      dcap = 1.f - th * th;
    }
    float g = __expf(z - l);
    if (c == label) g -= 1.f;
    lr[c] = __float2bfloat16(g * scale * dcap);
  }
}

// Small on-device reductions so the loss never needs a host round trip:
//   out[0] = sum_t (lse - tgt) over valid rows ; out[1] = number of valid rows
__global__ void __launch_bounds__(1024)
ce_reduce_kernel(const float* __restrict__ lse, const float* __restrict__ tgt, const int64_t* __restrict__ labels,
                 float* __restrict__ out, int rows, int64_t ignore_index) {
  __shared__ float red[32];
  float s = 0.f, n = 0.f;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    if (labels[r] != ignore_index) {
      s += lse[r] - tgt[r];
      n += 1.f;
    }
  }
  s = block_sum(s, red);
  n = block_sum(n, red);
  if (threadIdx.x == 0) {
    out[0] = s;
    out[1] = n;
  }
}

}  // namespace llmt

using namespace llmt;

extern "C" int llmt_ce_fwd(const void* logits, const int64_t* labels, float* lse, float* tgt, int rows, int vocab,
                           int64_t row_stride, int64_t ignore_index, float softcap, cudaStream_t stream) {
  LLMT_REQUIRE(row_stride % 8 == 0, "ce_fwd: row stride must be a multiple of 8 elements");
  if (rows == 0) return 0;
  ce_fwd_kernel<<<rows, kCeThreads, 0, stream>>>((const __nv_bfloat16*)logits, labels, lse, tgt, vocab, row_stride,
                                                 ignore_index, softcap);
  return check_launch("ce_fwd");
}

extern "C" int llmt_ce_bwd_inplace(void* logits, const int64_t* labels, const float* lse, const float* row_scale,
                                   int rows, int vocab, int64_t row_stride, int64_t ignore_index, float softcap,
                                   cudaStream_t stream) {
  LLMT_REQUIRE(row_stride % 8 == 0, "ce_bwd: row stride must be a multiple of 8 elements");
  if (rows == 0) return 0;
  ce_bwd_inplace_kernel<<<rows, kCeThreads, 0, stream>>>((__nv_bfloat16*)logits, labels, lse, row_scale, vocab,
                                                         row_stride, ignore_index, softcap);
  return check_launch("ce_bwd_inplace");
}

extern "C" int llmt_ce_reduce(const float* lse, const float* tgt, const int64_t* labels, float* out2, int rows,
                              int64_t ignore_index, cudaStream_t stream) {
  ce_reduce_kernel<<<1, 1024, 0, stream>>>(lse, tgt, labels, out2, rows, ignore_index);
  return check_launch("ce_reduce");
}
