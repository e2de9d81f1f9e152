// Packed causal attention, backward, on tcgen05 tensor cores.
//
// Replaces flash-attn's varlen backward (reference path: src/llm_training/ops/attention_op.py:606-618 under
// autograd).  One CTA owns 128 KV rows of one KV head and walks over every (query head of the group, 64-row
// query block) that can see them; dK / dV accumulate in TMEM across the whole walk (no atomics), dQ is
// reduced into an fp32 [T, Hq*D] accumulator with coalesced red.global.add.
//
// Everything is computed TRANSPOSED so that "thread r <-> TMEM lane r <-> KV row r":
//   S^T  = K Q^T            (M=128 kv, N=64 q,  K=D)    A=K  K-major, B=Q  K-major
//   dP^T = V dO^T           (M=128 kv, N=64 q,  K=D)    A=V  K-major, B=dO K-major
//   P^T  = exp2(S^T*c - L[q]) masked;  dS^T = P^T o (dP^T - delta[q]) * scale        (threads, bf16 -> smem)
//   dV  += P^T  dO          (M=128 kv, N=D,     K=64 q) A=P^T  K-major, B=dO MN-major
//   dK  += dS^T Q           (M=128 kv, N=D,     K=64 q) A=dS^T K-major, B=Q  MN-major
//   dQ^T = K^T dS^T         (M=D,      N=64 q,  K=128)  A=K  MN-major,  B=dS^T MN-major   -> red.add to dq_acc
// Eight compute warps: warp w owns TMEM lanes / kv rows 32*(w%4)..+32 and the q-column half w/4 of every 64-column
// step, so each SM sub-partition interleaves two warps' exp2 / convert / smem-store streams (one warp per scheduler
// left the MUFU and TMEM-load latencies exposed: 25% tensor-pipe activity in ncu).
// TMEM: S^T @0 (64) | dP^T @64 (64) | dQ^T @128, @192 (2 x 64, double-buffered so its drain runs under the next
// step's MMAs) | dK @256 (128) | dV @384 (128)  = 512 columns, 1 CTA / SM.
// Measured (B200, 8192 tokens, 32/8 heads, d=128): 2.22 ms per backward = 2.05x flash-attn 2.8.  Switching off the
// dQ reduce or the exp2 does not move the time: per 128x64 step the five MMAs read 213 KB of operands from shared
// memory and the P^T / dS^T / dQ staging / TMA traffic adds ~128 KB, i.e. ~2700 cycles at 128 B/clk against a
// 1280-cycle tensor floor — the kernel is shared-memory-bandwidth bound (next step: P^T as a TMEM A-operand,
// 128-column q steps).
// The head dimension is always padded to 128 here (TMA zero-fills the out-of-range columns).
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "sm100.cuh"

namespace llmt {
namespace attn {

using namespace sm100;

int make_tmap_thd(CUtensorMap* out, const void* ptr, int64_t tokens, int heads, int head_dim, int64_t row_stride,
                  int box_rows);
int make_tmap_f32_2d(CUtensorMap* out, const void* ptr, int64_t rows, int64_t cols, int box_cols, int box_rows);

constexpr int kBwdKV = 128;   // kv rows per CTA
constexpr int kBwdQ = 64;     // q rows per step
constexpr int kBwdD = 128;    // padded head dim
constexpr int kBwdThreads = 320;   // 8 compute warps (2 per TMEM lane quarter: q columns 0-31 / 32-63) + MMA + TMA
constexpr int kBwdComputeThreads = 256;
constexpr int kBwdMmaWarp = 8, kBwdTmaWarp = 9;

struct BwdCfg {
  static constexpr int kKBytes = kBwdKV * kBwdD * 2;        // 32 KB : 2 boxes [128][64]
  static constexpr int kQBytes = kBwdQ * kBwdD * 2;         // 16 KB : 2 boxes [64][64]
  static constexpr int kPBytes = kBwdKV * kBwdQ * 2;        // 16 KB : 1 box  [128][64]
  static constexpr int kKOff = 0;
  static constexpr int kVOff = kKOff + kKBytes;
  static constexpr int kQOff = kVOff + kKBytes;             // 2 stages
  static constexpr int kDOOff = kQOff + 2 * kQBytes;        // 2 stages
  static constexpr int kPOff = kDOOff + 2 * kQBytes;
  static constexpr int kDSOff = kPOff + kPBytes;
  static constexpr int kDQOff = kDSOff + kPBytes;           // fp32 [64 q][128 d] staging for the TMA reduce-add
  static constexpr int kDQBytes = kBwdQ * kBwdD * 4;        // 32 KB
  static constexpr int kMetaOff = kDQOff + kDQBytes;        // 3 x {L[64], delta[64], seg_start[64]}
  static constexpr int kMetaBytes = 3 * 3 * 64 * 4;      // 3-deep ring
  static constexpr int kBarOff = kMetaOff + kMetaBytes;
  // kv_full, qdo_full[2], qdo_empty[2], st_full, st_empty, pds_full, pds_empty, dq_full[2], dq_empty[2], acc_done
  static constexpr int kNumBars = 14;
  static constexpr int kSmemBytes = kBarOff + (kNumBars + 1) * 8 + 16;
  static constexpr int kColST = 0, kColDPT = 64, kColDQT = 128 /* two buffers: 128, 192 */, kColDK = 256, kColDV = 384;
};

// "TS" layout (the default kernel): P^T and dS^T live in TENSOR MEMORY as the A operands of the dV / dK MMAs
// (tcgen05.mma with A from TMEM), so per 128x64 step the shared-memory ports no longer carry the 16 KB P^T tile
// write, nor the 32 KB the two MMAs read back; dS^T is additionally written to smem because the dQ^T MMA needs it as
// its B operand.  dQ^T is single-buffered to make room:
//   dK @0 (128) | dV @128 (128) | S^T @256 (64) | dP^T @320 (64) | dQ^T @384 (64) | P^T bf16 @448 (32) | dS^T bf16 @480 (32)
struct BwdTsCfg {
  static constexpr int kKBytes = kBwdKV * kBwdD * 2;        // 32 KB : 2 boxes [128][64]
  static constexpr int kQBytes = kBwdQ * kBwdD * 2;         // 16 KB : 2 boxes [64][64]
  static constexpr int kDSBytes = kBwdKV * kBwdQ * 2;       // 16 KB : 1 box  [128][64]
  static constexpr int kKOff = 0;
  static constexpr int kVOff = kKOff + kKBytes;
  // Q / dO ring: THREE stages.  With two, the load of step it+1 could only start when the MMAs of step it-1 had
  // released their stage, and its ~1 us flight sat in front of S^T(it+1) every step (35% of the issuer's samples).
  static constexpr int kStages = 3;
  static constexpr int kQOff = kVOff + kKBytes;
  static constexpr int kDOOff = kQOff + kStages * kQBytes;
  static constexpr int kDSOff = kDOOff + kStages * kQBytes;
  static constexpr int kDQOff = kDSOff + kDSBytes;          // fp32 [64 q][128 d] staging for the TMA reduce-add
  static constexpr int kDQBytes = kBwdQ * kBwdD * 4;        // 32 KB
  static constexpr int kMetaOff = kDQOff + kDQBytes;        // 3 x {L[64], delta[64]}
  static constexpr int kMetaBytes = 3 * 2 * 64 * 4;
  static constexpr int kBarOff = kMetaOff + kMetaBytes;
  // kv_full, qdo_full[3], qdo_empty[3], s_full, s_empty, dp_full, dp_empty, pds_full, pds_empty, dq_full, dq_empty,
  // acc_done
  static constexpr int kNumBars = 16;
  static constexpr int kSmemBytes = kBarOff + (kNumBars + 1) * 8 + 16;
  static constexpr int kColDK = 0, kColDV = 128, kColST = 256, kColDPT = 320, kColDQT = 384, kColP = 448, kColDS = 480;
};

struct BwdParams {
  const int* seg_start;
  const int* seg_end;
  const float* lse;     // [Hq, T] log2 units
  const float* delta;   // [Hq, T]
  float* dq_acc;        // [T, Hq*D] fp32, zero-initialised
  int tokens, hq, hkv, head_dim;
  float scale, scale_log2;
  int window;
  // extras (attn_bwd_ts_kernel<true> only); same meaning as in the forward kernel
  float softcap_in, softcap_out;   // scale / softcap, softcap * log2(e); 0 = no soft-capping
  uint32_t drop_thr, seed;         // dropped when the element's random byte < drop_thr; 0 = no dropout
  float drop_scale;                // 1 / (1 - p)
  // deterministic mode: dq_order[h * n_qblk + qb] = kv block whose dQ contribution to (head h, q block qb) is next;
  // zero on entry.  nullptr = unordered reduce-adds.
  int* dq_order;
  int n_qblk;
  // attn_bwd_ts_kernel only.  splits > 1: blockIdx.x = kv_head * splits + z, and CTA z takes the (head, q block)
  // steps z, z + splits, ... of its kv block; the CTAs of one kv block then ADD their fp32 dK / dV tiles into dkv_acc
  // ([T, 2*Hkv*D], zero on entry) instead of storing bf16.  For few kv heads per GPU (tensor parallel shards) the
  // grid would otherwise be 64 CTAs whose longest one walks G * 128 steps alone.
  int splits;
};

__device__ __forceinline__ float bwd_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// delta[h, t] = sum_d O[t, h, d] * dO[t, h, d]   (one warp per (t, h))
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                  float* __restrict__ delta, int tokens, int hq, int head_dim) {
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (int64_t)tokens * hq) return;
  const int64_t t = w / hq;
  const int h = (int)(w - t * hq);
  const __nv_bfloat16* po = o + (t * hq + h) * head_dim;
  const __nv_bfloat16* pd = d_o + (t * hq + h) * head_dim;
  float s = 0.f;
  for (int c = lane * 8; c < head_dim; c += 256) {
    float a[8], b[8];
    unpack8(ld_bf16x8(po + c), a);
    unpack8(ld_bf16x8(pd + c), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j] * b[j];
  }
  s = warp_sum(s);
  if (lane == 0) delta[(size_t)h * tokens + t] = s;
}

// dqkv[t, 0 : Hq*D] = bf16(mul * dq_acc[t, :])
__global__ void __launch_bounds__(256)
attn_dq_convert_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, int tokens, int width,
                       int64_t out_stride, float mul) {
  const int vec_per_row = width >> 3;
  const int64_t total = (int64_t)tokens * vec_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / vec_per_row;
    const int c = (int)(i - t * vec_per_row) << 3;
    const float4 a = *reinterpret_cast<const float4*>(dq_acc + t * width + c);
    const float4 b = *reinterpret_cast<const float4*>(dq_acc + t * width + c + 4);
    const float f[8] = {a.x * mul, a.y * mul, a.z * mul, a.w * mul, b.x * mul, b.y * mul, b.z * mul, b.w * mul};
    st_bf16x8(dqkv + t * out_stride + c, pack8(f));
  }
}

__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmKV,    // qkv, box {64,1,128}
                const __grid_constant__ CUtensorMap tmQ,     // qkv, box {64,1,64}
                const __grid_constant__ CUtensorMap tmDO,    // dO [T][Hq][D], box {64,1,64}
                const __grid_constant__ CUtensorMap tmDKV,   // dqkv, box {64,1,128} (store)
                const __grid_constant__ CUtensorMap tmDQ,    // dq_acc fp32 [T][Hq*D], box {128, 64}, no swizzle
                BwdParams p) {
  using C = BwdCfg;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0) __trap();

  const uint32_t bars = sbase + C::kBarOff;
  const uint32_t kv_full = bars;
  auto qdo_full = [&](int s) { return bars + 8u * (1 + s); };
  auto qdo_empty = [&](int s) { return bars + 8u * (3 + s); };
  const uint32_t st_full = bars + 8u * 5, st_empty = bars + 8u * 6;
  const uint32_t pds_full = bars + 8u * 7, pds_empty = bars + 8u * 8;
  auto dq_full = [&](int b) { return bars + 8u * (9 + b); };
  auto dq_empty = [&](int b) { return bars + 8u * (11 + b); };
  const uint32_t acc_done = bars + 8u * 13;
  const uint32_t tmem_ptr_smem = bars + 8u * 14;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem + C::kBarOff + 8 * 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // launch order = longest job first: blockIdx.x (fast axis) walks the kv heads, blockIdx.y the kv tiles, so the
  // early tiles (which see the most queries) of EVERY head start before any light tile
  const int g = blockIdx.x;                       // kv head
  const int G = p.hq / p.hkv;
  const int k0 = (int)blockIdx.y * kBwdKV;
  const int k_last = min(k0 + kBwdKV, p.tokens) - 1;
  // query blocks that can see this kv tile: q >= k0, q < seg_end[k_last] (bounds are non-decreasing)
  int q_hi = p.seg_end[k_last] - 1;
  if (p.window > 0) q_hi = min(q_hi, k_last + p.window - 1);
  q_hi = min(q_hi, p.tokens - 1);
  const int i_lo = k0 / kBwdQ;
  const int nq = q_hi / kBwdQ - i_lo + 1;
  const int n_total = G * nq;

  if (warp == kBwdMmaWarp && lane == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(qdo_full(s), 1);
      mbar_init(qdo_empty(s), 1);
    }
    mbar_init(st_full, 1);
    mbar_init(st_empty, kBwdComputeThreads);
    mbar_init(pds_full, kBwdComputeThreads);
    mbar_init(pds_empty, 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(dq_full(b), 1);
      mbar_init(dq_empty(b), kBwdComputeThreads);
    }
    mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == kBwdTmaWarp) {
    if (lane == 0) {
      tma_prefetch_desc(&tmKV);
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmDO);
      tma_prefetch_desc(&tmDKV);
      tma_prefetch_desc(&tmDQ);
    }
    tmem_alloc<1>(tmem_ptr_smem, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == kBwdTmaWarp) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_arrive_expect_tx(kv_full, 2 * C::kKBytes);
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        tma_load_3d(sbase + C::kKOff + b * (kBwdKV * 128), &tmKV, kv_full, b * 64, p.hq + g, k0);
        tma_load_3d(sbase + C::kVOff + b * (kBwdKV * 128), &tmKV, kv_full, b * 64, p.hq + p.hkv + g, k0);
      }
      for (int it = 0; it < n_total; ++it) {
        const int st = it & 1;
        const int h = g * G + it / nq;
        const int q0 = (i_lo + it % nq) * kBwdQ;
        mbar_wait(qdo_empty(st), ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(qdo_full(st), 2 * C::kQBytes);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          tma_load_3d(sbase + C::kQOff + st * C::kQBytes + b * (kBwdQ * 128), &tmQ, qdo_full(st), b * 64, h, q0);
          tma_load_3d(sbase + C::kDOOff + st * C::kQBytes + b * (kBwdQ * 128), &tmDO, qdo_full(st), b * 64, h, q0);
        }
      }
    }
  } else if (warp == kBwdMmaWarp) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc_st = make_idesc_bf16(kBwdKV, kBwdQ, false, false);   // S^T, dP^T
      constexpr uint32_t idesc_dkv = make_idesc_bf16(kBwdKV, kBwdD, false, true);   // dV, dK
      constexpr uint32_t idesc_dq = make_idesc_bf16(kBwdD, kBwdQ, true, true);      // dQ^T
      const uint32_t ka = sbase + C::kKOff, va = sbase + C::kVOff;
      const uint32_t pa = sbase + C::kPOff, dsa = sbase + C::kDSOff;
      auto issue_st = [&](int it) {
        const int st = it & 1;
        const uint32_t qa = sbase + C::kQOff + st * C::kQBytes, doa = sbase + C::kDOOff + st * C::kQBytes;
#pragma unroll
        for (int kk = 0; kk < kBwdD / 16; ++kk) {
          const uint32_t aoff = (kk / 4) * (kBwdKV * 128) + (kk % 4) * 32;
          const uint32_t boff = (kk / 4) * (kBwdQ * 128) + (kk % 4) * 32;
          umma_f16<1>(tmem_base + C::kColST, make_smem_desc(ka + aoff, 0, 1024), make_smem_desc(qa + boff, 0, 1024),
                      idesc_st, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < kBwdD / 16; ++kk) {
          const uint32_t aoff = (kk / 4) * (kBwdKV * 128) + (kk % 4) * 32;
          const uint32_t boff = (kk / 4) * (kBwdQ * 128) + (kk % 4) * 32;
          umma_f16<1>(tmem_base + C::kColDPT, make_smem_desc(va + aoff, 0, 1024),
                      make_smem_desc(doa + boff, 0, 1024), idesc_st, kk != 0);
        }
        umma_commit(st_full);
      };
      mbar_wait(kv_full, 0);
      mbar_wait(qdo_full(0), 0);
      tc_fence_after();
      issue_st(0);
      for (int it = 0; it < n_total; ++it) {
        if (it + 1 < n_total) {
          mbar_wait(qdo_full((it + 1) & 1), ((it + 1) >> 1) & 1);
          mbar_wait(st_empty, it & 1);           // threads hold S^T / dP^T (it) in registers
          tc_fence_after();
          issue_st(it + 1);
        }
        mbar_wait(pds_full, it & 1);
        mbar_wait(dq_empty(it & 1), ((it >> 1) & 1) ^ 1);   // dQ^T(it-2) has left this TMEM buffer
        tc_fence_after();
        const int st = it & 1;
        const uint32_t qa = sbase + C::kQOff + st * C::kQBytes, doa = sbase + C::kDOOff + st * C::kQBytes;
#pragma unroll
        for (int kk = 0; kk < kBwdQ / 16; ++kk) {   // dV += P^T dO
          umma_f16<1>(tmem_base + C::kColDV, make_smem_desc(pa + kk * 32, 0, 1024),
                      make_smem_desc(doa + kk * 2048, kBwdQ * 128, 1024), idesc_dkv, (it | kk) != 0);
        }
#pragma unroll
        for (int kk = 0; kk < kBwdQ / 16; ++kk) {   // dK += dS^T Q
          umma_f16<1>(tmem_base + C::kColDK, make_smem_desc(dsa + kk * 32, 0, 1024),
                      make_smem_desc(qa + kk * 2048, kBwdQ * 128, 1024), idesc_dkv, (it | kk) != 0);
        }
#pragma unroll
        for (int kk = 0; kk < kBwdKV / 16; ++kk) {  // dQ^T = K^T dS^T
          umma_f16<1>(tmem_base + C::kColDQT + (it & 1) * 64, make_smem_desc(ka + kk * 2048, kBwdKV * 128, 1024),
                      make_smem_desc(dsa + kk * 2048, kBwdKV * 128, 1024), idesc_dq, kk != 0);
        }
        umma_commit(qdo_empty(st));
        umma_commit(pds_empty);
        umma_commit(dq_full(it & 1));
      }
      umma_commit(acc_done);
    }
  } else {
    // ===== compute warps: thread <-> kv row r, q-column half hf =====
    const int r = (warp & 3) * 32 + lane;
    const int hf = warp >> 2;
    const int c_base = hf * 32;                          // first q column of this thread's half
    const int kvi = k0 + r;
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    uint8_t* p_row = smem + C::kPOff + r * 128;
    uint8_t* ds_row = smem + C::kDSOff + r * 128;

    // Per-step metadata {L[64], delta[64], seg_start[64]} lives in a 3-deep smem ring: step it+2 is fetched
    // from global memory while step it is processed, so the loads never sit on the critical path.
    auto meta = [&](int it) { return reinterpret_cast<float*>(smem + C::kMetaOff) + (it % 3) * 192; };
    float pre0 = 0.f, pre1 = 0.f;                       // prefetched values (r < 64: L, delta; r >= 64: seg_start)
    auto meta_fetch = [&](int it) {
      if (hf != 0) return;                               // warps 0-3 fetch for everybody
      const int h = g * G + it / nq;
      const int q = (i_lo + it % nq) * kBwdQ + (r & 63);
      const bool ok = q < p.tokens;
      if (r < 64) {
        pre0 = ok ? p.lse[(size_t)h * p.tokens + q] : INFINITY;
        pre1 = ok ? p.delta[(size_t)h * p.tokens + q] : 0.f;
      } else {
        pre0 = __int_as_float(ok ? p.seg_start[q] : 0x7fffffff);
      }
    };
    auto meta_store = [&](int it) {
      if (hf != 0) return;
      float* m = meta(it);
      if (r < 64) { m[r] = pre0; m[64 + r] = pre1; }
      else        { m[128 + (r - 64)] = pre0; }
    };
    // P^T of step `it` from S^T in TMEM -> registers (fp32 bit patterns in sv)
    uint32_t sv[32];
    auto compute_p = [&](int it) {
      const int q0 = (i_lo + it % nq) * kBwdQ;
      const float* mL = meta(it);
      const int* mS = reinterpret_cast<const int*>(mL + 128);
      const bool need_mask = (q0 < k0 + kBwdKV - 1) || (mS[63] > k0) || (p.window > 0);
      mbar_wait(st_full, it & 1);
      tc_fence_after();
      tmem_ld_32x32b_x32(t_lane + C::kColST + c_base, sv);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        float pv = bwd_exp2(fmaf(__uint_as_float(sv[c]), p.scale_log2, -mL[c_base + c]));
        if (need_mask) {
          const int q = q0 + c_base + c;
          const bool valid = (mS[c_base + c] <= kvi) && (kvi <= q) && (p.window == 0 || q - kvi < p.window);
          pv = valid ? pv : 0.f;
        }
        sv[c] = __float_as_uint(pv);
      }
    };

    auto drain_dq = [&](int j, int meta_it) {
      uint32_t dq[32];
      if (j >= 0) {
        mbar_wait(dq_full(j & 1), (j >> 1) & 1);
        tc_fence_after();
        tmem_ld_32x32b_x32(t_lane + C::kColDQT + (j & 1) * 64 + c_base, dq);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(dq_empty(j & 1));
        if (threadIdx.x == 0) tma_store_wait_read<0>();   // previous reduce has finished reading the staging tile
      }
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (j >= 0) {
        float* stage = reinterpret_cast<float*>(smem + C::kDQOff);
#pragma unroll
        for (int c = 0; c < 32; ++c) stage[(c_base + c) * kBwdD + r] = __uint_as_float(dq[c]);
      }
      if (meta_it < n_total) meta_store(meta_it);         // ring slot meta_it % 3 was last read two steps ago
      fence_proxy_async_smem();
      asm volatile("bar.sync 2, 256;" ::: "memory");      // also publishes the metadata stored above
      if (j >= 0 && threadIdx.x == 0) {
        const int hj = g * G + j / nq;
        const int qj = (i_lo + j % nq) * kBwdQ;
        tma_reduce_add_2d(&tmDQ, sbase + C::kDQOff, hj * p.head_dim, qj);
        tma_store_commit();
      }
    };

    // prologue: metadata of steps 0 and 1, then P^T(0)
    meta_fetch(0);
    meta_store(0);
    if (n_total > 1) { meta_fetch(1); meta_store(1); }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    compute_p(0);

    for (int it = 0; it < n_total; ++it) {
      const int h = g * G + it / nq;
      const int q0 = (i_lo + it % nq) * kBwdQ;
      const float* mD = meta(it) + 64;
      if (it + 2 < n_total) meta_fetch(it + 2);           // global loads in flight during this whole step

      // ---- P^T(it) (registers) -> smem, dS^T(it) from dP^T -> smem
      mbar_wait(pds_empty, (it & 1) ^ 1);                 // dV/dK/dQ MMAs of the previous step have read the tiles
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(sv[ch * 8 + e]);
        *reinterpret_cast<bf16x8*>(p_row + (((hf * 4 + ch) ^ (r & 7)) << 4)) = pack8(f);
      }
      {
        uint32_t dp[32];
        tmem_ld_32x32b_x32(t_lane + C::kColDPT + c_base, dp);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(st_empty);                            // this thread's S^T and dP^T columns are in registers
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int c = ch * 8 + e;
            f[e] = __uint_as_float(sv[c]) * (__uint_as_float(dp[c]) - mD[c_base + c]) * p.scale;
          }
          *reinterpret_cast<bf16x8*>(ds_row + (((hf * 4 + ch) ^ (r & 7)) << 4)) = pack8(f);
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(pds_full);

      // ---- dQ^T(it-1): TMEM [lane = d][col = q] -> smem [q][d] fp32 -> one TMA reduce-add into dq_acc.  dQ^T is
      //      double-buffered in TMEM, so this drain (and its two CTA barriers) runs UNDER the MMAs of step it
      //      instead of between them; the barriers also publish the metadata of step it+2.
      drain_dq(it - 1, it + 2);

      // ---- while the tensor cores run S^T/dP^T(it+1) and dV/dK/dQ^T(it), build P^T(it+1)
      if (it + 1 < n_total) compute_p(it + 1);
    }
    drain_dq(n_total - 1, n_total + 2);
    if (threadIdx.x == 0) tma_store_wait<0>();

    // ---- dK / dV : TMEM -> bf16 -> smem (K / V tile regions) -> TMA store
    mbar_wait(acc_done, 0);
    tc_fence_after();
    {
      const int which = hf;                                // warps 0-3 drain dK, warps 4-7 drain dV
      uint8_t* stage = smem + (which == 0 ? C::kKOff : C::kVOff);
      const uint32_t col = which == 0 ? C::kColDK : C::kColDV;
#pragma unroll
      for (int c0 = 0; c0 < kBwdD; c0 += 32) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(t_lane + col + c0, o);
        tmem_ld_wait();
        uint8_t* box_row = stage + (c0 / 64) * (kBwdKV * 128) + r * 128;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(o[ch * 8 + e]);
          const int chunk = ((c0 % 64) / 8) + ch;
          *reinterpret_cast<bf16x8*>(box_row + ((chunk ^ (r & 7)) << 4)) = pack8(f);
        }
      }
    }
    fence_proxy_async_smem();
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x == 0) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        tma_store_3d(&tmDKV, sbase + C::kKOff + b * (kBwdKV * 128), b * 64, p.hq + g, k0);
        tma_store_3d(&tmDKV, sbase + C::kVOff + b * (kBwdKV * 128), b * 64, p.hq + p.hkv + g, k0);
      }
      tma_store_commit();
      tma_store_wait<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kBwdTmaWarp) tmem_dealloc<1>(tmem_base, 512);
}


__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ float bwd_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// kSplit = false compiles S (BwdParams::splits) to the constant 1, so the usual many-kv-head launch keeps the exact
// instruction stream it had before kv blocks could be shared.
template <bool kExtras, bool kSplit>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_ts_kernel(const __grid_constant__ CUtensorMap tmKV,    // qkv, box {64,1,128}
                   const __grid_constant__ CUtensorMap tmQ,     // qkv, box {64,1,64}
                   const __grid_constant__ CUtensorMap tmDO,    // dO [T][Hq][D], box {64,1,64}
                   const __grid_constant__ CUtensorMap tmDKV,   // dqkv, box {64,1,128} (store)
                   const __grid_constant__ CUtensorMap tmDQ,    // dq_acc fp32 [T][Hq*D], box {128, 64}, no swizzle
                   const __grid_constant__ CUtensorMap tmDKVacc,  // dkv_acc fp32 [T][2*Hkv*D], same box (splits > 1)
                   BwdParams p) {
  using C = BwdTsCfg;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0) __trap();

  const uint32_t bars = sbase + C::kBarOff;
  const uint32_t kv_full = bars;
  auto qdo_full = [&](int s) { return bars + 8u * (1 + s); };
  auto qdo_empty = [&](int s) { return bars + 8u * (4 + s); };
  const uint32_t s_full = bars + 8u * 7, s_empty = bars + 8u * 8;
  const uint32_t dp_full = bars + 8u * 9, dp_empty = bars + 8u * 10;
  const uint32_t pds_full = bars + 8u * 11, pds_empty = bars + 8u * 12;
  const uint32_t dq_full = bars + 8u * 13, dq_empty = bars + 8u * 14;
  const uint32_t acc_done = bars + 8u * 15;
  const uint32_t tmem_ptr_smem = bars + 8u * 16;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem + C::kBarOff + 8 * 16);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const int S = kSplit ? p.splits : 1;
  const int g = (int)blockIdx.x / S;              // kv head (fast axis: longest jobs of every head start first)
  const int z = (int)blockIdx.x - g * S;          // this CTA's share of the kv block's steps: z, z + S, ...
  const int G = p.hq / p.hkv;
  const int k0 = (int)blockIdx.y * kBwdKV;
  const int k_last = min(k0 + kBwdKV, p.tokens) - 1;
  int q_hi = p.seg_end[k_last] - 1;
  if (p.window > 0) q_hi = min(q_hi, k_last + p.window - 1);
  q_hi = min(q_hi, p.tokens - 1);
  const int i_lo = k0 / kBwdQ;
  const int nq = q_hi / kBwdQ - i_lo + 1;
  const int n_total = (G * nq - z + S - 1) / S;
  if (n_total <= 0) return;                       // more CTAs than steps (short kv block): nothing to add
  const int h0 = g * G + z / nq, qi0 = z % nq;    // first step of this CTA
  // every cursor over the (head, q block) steps moves S steps at a time
  auto advance = [&](int& h, int& qi) {
    qi += S;
    while (qi >= nq) { qi -= nq; ++h; }
  };

  if (warp == kBwdMmaWarp && lane == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(qdo_full(s), 1);
      mbar_init(qdo_empty(s), 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_empty, kBwdComputeThreads);
    mbar_init(dp_full, 1);
    mbar_init(dp_empty, kBwdComputeThreads);
    mbar_init(pds_full, kBwdComputeThreads);
    mbar_init(pds_empty, 1);
    mbar_init(dq_full, 1);
    mbar_init(dq_empty, kBwdComputeThreads);
    mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == kBwdTmaWarp) {
    if (lane == 0) {
      tma_prefetch_desc(&tmKV);
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmDO);
      tma_prefetch_desc(&tmDKV);
      tma_prefetch_desc(&tmDQ);
      if (S > 1) tma_prefetch_desc(&tmDKVacc);
    }
    tmem_alloc<1>(tmem_ptr_smem, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == kBwdTmaWarp) {
    // ===== TMA producer (whole warp in uniform control flow, one elected lane issues) =====
    if (elect_one()) {
      mbar_arrive_expect_tx(kv_full, 2 * C::kKBytes);
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        tma_load_3d(sbase + C::kKOff + b * (kBwdKV * 128), &tmKV, kv_full, b * 64, p.hq + g, k0);
        tma_load_3d(sbase + C::kVOff + b * (kBwdKV * 128), &tmKV, kv_full, b * 64, p.hq + p.hkv + g, k0);
      }
    }
    __syncwarp();
    int st = 0, h = h0, qi = qi0;
    uint32_t ph = 0;
#pragma unroll 1
    for (int it = 0; it < n_total; ++it) {
      const int q0 = (i_lo + qi) * kBwdQ;
      mbar_wait(qdo_empty(st), ph ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(qdo_full(st), 2 * C::kQBytes);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          tma_load_3d(sbase + C::kQOff + st * C::kQBytes + b * (kBwdQ * 128), &tmQ, qdo_full(st), b * 64, h, q0);
          tma_load_3d(sbase + C::kDOOff + st * C::kQBytes + b * (kBwdQ * 128), &tmDO, qdo_full(st), b * 64, h, q0);
        }
      }
      __syncwarp();
      if (++st == C::kStages) { st = 0; ph ^= 1; }
      advance(h, qi);
    }
  } else if (warp == kBwdMmaWarp) {
    // ===== MMA issuer: the WHOLE warp walks the loop (uniform control flow, waits included); one elected lane
    // issues.  Descriptors are built once; a k-step only adds a constant to the start-address field.
    constexpr uint32_t idesc_st = make_idesc_bf16(kBwdKV, kBwdQ, false, false);   // S^T, dP^T
    constexpr uint32_t idesc_dkv = make_idesc_bf16(kBwdKV, kBwdD, false, true);   // dV, dK (A from TMEM)
    constexpr uint32_t idesc_dq = make_idesc_bf16(kBwdD, kBwdQ, true, true);      // dQ^T
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t k_a = make_smem_desc(sbase + C::kKOff, 0, 1024);               // K  as A, K-major   (S^T)
    const uint64_t v_a = make_smem_desc(sbase + C::kVOff, 0, 1024);               // V  as A, K-major   (dP^T)
    const uint64_t kt_a = make_smem_desc(sbase + C::kKOff, kBwdKV * 128, 1024);   // K^T as A, MN-major (dQ^T)
    const uint64_t ds_b = make_smem_desc(sbase + C::kDSOff, kBwdKV * 128, 1024);  // dS^T as B, MN-major (dQ^T)
    const uint64_t q_b0 = make_smem_desc(sbase + C::kQOff, 0, 1024);              // Q  as B, K-major   (S^T), stage 0
    const uint64_t do_b0 = make_smem_desc(sbase + C::kDOOff, 0, 1024);            // dO as B, K-major   (dP^T)
    const uint64_t q_mn0 = make_smem_desc(sbase + C::kQOff, kBwdQ * 128, 1024);   // Q  as B, MN-major  (dK)
    const uint64_t do_mn0 = make_smem_desc(sbase + C::kDOOff, kBwdQ * 128, 1024); // dO as B, MN-major  (dV)
    constexpr uint64_t kStage = C::kQBytes >> 4;                                  // descriptor units per Q / dO stage
    // k-step offsets in descriptor units (16 bytes): K-major tiles hold 4 k-steps per 64-column box
    auto kmaj = [](int kk, int rows) -> uint64_t { return (uint64_t)((kk / 4) * (rows * 8) + (kk % 4) * 2); };
    auto issue_s = [&](int st) {
      const uint64_t q_b = q_b0 + (uint64_t)st * kStage;
#pragma unroll
      for (int kk = 0; kk < kBwdD / 16; ++kk)
        umma_f16<1>(tb + C::kColST, k_a + kmaj(kk, kBwdKV), q_b + kmaj(kk, kBwdQ), idesc_st, kk != 0);
      umma_commit(s_full);
    };
    auto issue_dp = [&](int st) {
      const uint64_t do_b = do_b0 + (uint64_t)st * kStage;
#pragma unroll
      for (int kk = 0; kk < kBwdD / 16; ++kk)
        umma_f16<1>(tb + C::kColDPT, v_a + kmaj(kk, kBwdKV), do_b + kmaj(kk, kBwdQ), idesc_st, kk != 0);
      umma_commit(dp_full);
    };
    mbar_wait(kv_full, 0);
    mbar_wait(qdo_full(0), 0);
    tc_fence_after();
    if (elect_one()) {
      issue_s(0);
      issue_dp(0);
    }
    __syncwarp();
    int st = 0, st_next = 1 % C::kStages;                 // ring slots of step it / it+1
    uint32_t ph_next = (C::kStages == 1) ? 1u : 0u;       // parity of qdo_full(st_next) for step it+1
#pragma unroll 1
    for (int it = 0; it < n_total; ++it) {
      if (it + 1 < n_total) {
        mbar_wait(qdo_full(st_next), ph_next);
        mbar_wait(s_empty, it & 1);             // S^T(it) is in the threads' registers
        tc_fence_after();
        if (elect_one()) issue_s(st_next);
        __syncwarp();
        mbar_wait(dp_empty, it & 1);            // dP^T(it) is in the threads' registers
        tc_fence_after();
        if (elect_one()) issue_dp(st_next);
        __syncwarp();
      }
      mbar_wait(pds_full, it & 1);              // P^T / dS^T (it) are in tensor memory, dS^T also in smem
      tc_fence_after();
      const uint64_t q_mn = q_mn0 + (uint64_t)st * kStage, do_mn = do_mn0 + (uint64_t)st * kStage;
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < kBwdQ / 16; ++kk)   // dV += P^T dO     (A = P^T from TMEM, 8 columns per k-step)
          umma_f16_ts<1>(tb + C::kColDV, tb + C::kColP + kk * 8, do_mn + (uint64_t)kk * 128, idesc_dkv,
                         (it | kk) != 0);
#pragma unroll
        for (int kk = 0; kk < kBwdQ / 16; ++kk)   // dK += dS^T Q    (A = dS^T from TMEM; scaled at the end)
          umma_f16_ts<1>(tb + C::kColDK, tb + C::kColDS + kk * 8, q_mn + (uint64_t)kk * 128, idesc_dkv,
                         (it | kk) != 0);
        umma_commit(qdo_empty(st));               // Q / dO of this step are not read again: release the ring slot
      }
      __syncwarp();
      mbar_wait(dq_empty, (it & 1) ^ 1);          // dQ^T(it-1) has left its TMEM buffer
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < kBwdKV / 16; ++kk)  // dQ^T = K^T dS^T
          umma_f16<1>(tb + C::kColDQT, kt_a + (uint64_t)kk * 128, ds_b + (uint64_t)kk * 128, idesc_dq, kk != 0);
        umma_commit(pds_empty);
        umma_commit(dq_full);
      }
      __syncwarp();
      st = st_next;
      if (++st_next == C::kStages) { st_next = 0; ph_next ^= 1; }
    }
    if (elect_one()) umma_commit(acc_done);
    __syncwarp();
  } else {
    // ===== compute warps: thread <-> kv row r, q-column half hf =====
    const int r = (warp & 3) * 32 + lane;
    const int hf = warp >> 2;
    const int c_base = hf * 32;
    const int kvi = k0 + r;
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    uint8_t* ds_row = smem + C::kDSOff + r * 128;

    // A kv row sees the queries q in [kvi, q_end): its own document from its position on, clipped by the window.
    // Both ends are constants of the thread, so masking a block is one unsigned compare per element and needs no
    // per-query metadata.  Blocks strictly inside every row's range skip it (CTA-uniform test below).
    int q_end = 0;
    if (kvi < p.tokens) {
      q_end = p.seg_end[kvi];
      if (p.window > 0) q_end = min(q_end, kvi + p.window);
    }
    const uint32_t valid_len = q_end > kvi ? (uint32_t)(q_end - kvi) : 0u;
    int q_end_first = p.seg_end[k0];                     // smallest q_end of the tile (bounds are non-decreasing)
    if (p.window > 0) q_end_first = min(q_end_first, k0 + p.window);
    const bool tail_tile = k0 + kBwdKV > p.tokens;

    // Per-step metadata {L[64], delta[64]} lives in a 3-deep smem ring: step it+2 is fetched from global memory
    // while step it is processed and published by the CTA barrier of the dQ drain.
    auto meta = [&](int it) { return reinterpret_cast<float*>(smem + C::kMetaOff) + (it % 3) * 128; };
    float pre = 0.f;
    // (head, q block) of steps it+2 [metadata fetch], it+1 [P^T] and it-1 [dQ^T drain], advanced once per iteration
    int f_h = h0, f_qi = qi0, n_qi = qi0, d_h = h0, d_qi = qi0;
    auto meta_fetch = [&]() {                            // fetches step (f_h, f_qi), then advances the cursor
      if (hf == 0) {                                     // warps 0-3 fetch for everybody
        const int q = (i_lo + f_qi) * kBwdQ + (r & 63);
        const bool ok = q < p.tokens;
        if (r < 64) pre = ok ? p.lse[(size_t)f_h * p.tokens + q] : INFINITY;
        else        pre = ok ? p.delta[(size_t)f_h * p.tokens + q] : 0.f;
      }
      advance(f_h, f_qi);
    };
    auto meta_store = [&](int it) {
      if (hf != 0) return;
      meta(it)[r] = pre;                                 // [0, 64): L, [64, 128): delta
    };

    uint32_t sv[32];                                     // P^T of the current step, fp32 bit patterns
    float tv[32];                                        // soft-capping only: tanh(s * scale / cap) of the same step
    int a_h = h0, a_qi = qi0;                            // (head, q block) of the step whose P^T / dS^T is being built

    // Deterministic mode: the kv blocks that touch one dQ tile add to it in ascending order.  The contributors of a
    // q block are a contiguous run of kv blocks (q_hi is non-decreasing in the kv block), so this CTA waits for
    // "next = my kv block" on every q block its predecessor also covers and publishes "next = mine + 1" once its own
    // reduce-add has COMPLETED (wait_group without .read).  CTAs are dispatched in block order and a CTA only ever
    // waits for a lower block index, so the wait cannot deadlock.
    int prev_hi_blk = -1;                                // last q block of kv block blockIdx.y - 1
    int* pend_order = nullptr;                           // thread 0: counter to publish after the reduce in flight
    if constexpr (kExtras) {
      if (p.dq_order != nullptr && blockIdx.y > 0) {
        int q_hi_prev = p.seg_end[k0 - 1] - 1;
        if (p.window > 0) q_hi_prev = min(q_hi_prev, k0 - 1 + p.window - 1);
        prev_hi_blk = min(q_hi_prev, p.tokens - 1) / kBwdQ;
      }
    }

    meta_fetch();
    meta_store(0);
    if (n_total > 1) { meta_fetch(); meta_store(1); }
    asm volatile("bar.sync 1, 256;" ::: "memory");

    // One loop instance of every phase (the body is ~1.5k instructions; peeled copies pushed the first version
    // past the instruction cache): iteration `it` finishes step it, drains dQ^T(it-1) and starts P^T(it+1).
#pragma unroll 1
    for (int it = -1; it <= n_total; ++it) {
      if (it >= 0 && it + 2 < n_total) meta_fetch();         // step it+2: global loads in flight during this whole step

      if (it >= 0 && it < n_total) {
        const float* mD = meta(it) + 64 + c_base;
        // ---- P^T(it): registers -> tensor memory (bf16 pairs: column c/2 holds q columns c, c+1)
        mbar_wait(pds_empty, (it & 1) ^ 1);               // dV / dK / dQ MMAs of the previous step are done
        if constexpr (!kExtras) {
          {
            uint32_t pk[16];
#pragma unroll
            for (int c = 0; c < 16; ++c)
              pk[c] = pack_bf16x2(__uint_as_float(sv[2 * c]), __uint_as_float(sv[2 * c + 1]));
            tmem_st_32x32b_x16(t_lane + C::kColP + hf * 16, pk);
          }
          // ---- dS^T(it) = P^T o (dP^T - delta)  (the softmax scale is applied to dK / dQ at the end)
          {
            uint32_t dp[32];
            mbar_wait(dp_full, it & 1);
            tc_fence_after();
            tmem_ld_32x32b_x32(t_lane + C::kColDPT + c_base, dp);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(dp_empty);
            uint32_t pk[16];
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              const float4 d4 = *reinterpret_cast<const float4*>(mD + ch * 4);
              const int c = ch * 4;
              pk[ch * 2] = pack_bf16x2(__uint_as_float(sv[c]) * (__uint_as_float(dp[c]) - d4.x),
                                       __uint_as_float(sv[c + 1]) * (__uint_as_float(dp[c + 1]) - d4.y));
              pk[ch * 2 + 1] = pack_bf16x2(__uint_as_float(sv[c + 2]) * (__uint_as_float(dp[c + 2]) - d4.z),
                                           __uint_as_float(sv[c + 3]) * (__uint_as_float(dp[c + 3]) - d4.w));
            }
            tmem_st_32x32b_x16(t_lane + C::kColDS + hf * 16, pk);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch)
              *reinterpret_cast<uint4*>(ds_row + (((hf * 4 + ch) ^ (r & 7)) << 4)) =
                  make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
          }
        } else {
          uint32_t keep = 0xffffffffu;                       // bit c: P^T column c survives dropout
          if constexpr (kExtras) {
            if (p.drop_thr != 0) {
              keep = 0;
              const int qb = (i_lo + a_qi) * kBwdQ + c_base;
  #pragma unroll
              for (int c = 0; c < 32; ++c) {
                const uint32_t rnd = dropout_bits(p.seed, a_h, qb + c, kvi >> 2);
                keep |= (((rnd >> (8 * (kvi & 3))) & 0xffu) >= p.drop_thr ? 1u : 0u) << c;
              }
            }
          }
          {
            uint32_t pk[16];
  #pragma unroll
            for (int c = 0; c < 16; ++c) {
              float p0 = __uint_as_float(sv[2 * c]), p1 = __uint_as_float(sv[2 * c + 1]);
              if constexpr (kExtras) {
                p0 = (keep >> (2 * c)) & 1u ? p0 : 0.f;      // dV sees the dropped probabilities (x 1/(1-p) at the end)
                p1 = (keep >> (2 * c + 1)) & 1u ? p1 : 0.f;
              }
              pk[c] = pack_bf16x2(p0, p1);
            }
            tmem_st_32x32b_x16(t_lane + C::kColP + hf * 16, pk);
          }
          // ---- dS^T(it) = P^T o (dP^T - delta)  (the softmax scale is applied to dK / dQ at the end)
          {
            uint32_t dp[32];
            mbar_wait(dp_full, it & 1);
            tc_fence_after();
            tmem_ld_32x32b_x32(t_lane + C::kColDPT + c_base, dp);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(dp_empty);
            uint32_t pk[16];
  #pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              const float4 d4 = *reinterpret_cast<const float4*>(mD + ch * 4);
              const float dl[4] = {d4.x, d4.y, d4.z, d4.w};
              float ds[4];
  #pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int c = ch * 4 + e;
                float dpv = __uint_as_float(dp[c]);
                if constexpr (kExtras) {
                  if (p.drop_thr != 0) dpv = (keep >> c) & 1u ? dpv * p.drop_scale : 0.f;
                }
                ds[e] = __uint_as_float(sv[c]) * (dpv - dl[e]);
                if constexpr (kExtras) {
                  if (p.softcap_in > 0.f) ds[e] *= 1.f - tv[c] * tv[c];    // d tanh
                }
              }
              pk[ch * 2] = pack_bf16x2(ds[0], ds[1]);
              pk[ch * 2 + 1] = pack_bf16x2(ds[2], ds[3]);
            }
            tmem_st_32x32b_x16(t_lane + C::kColDS + hf * 16, pk);
  #pragma unroll
            for (int ch = 0; ch < 4; ++ch)
              *reinterpret_cast<uint4*>(ds_row + (((hf * 4 + ch) ^ (r & 7)) << 4)) =
                  make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
          }
          advance(a_h, a_qi);
        }
        tmem_st_wait();
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(pds_full);
      }

      // ---- dQ^T(j = it-1): TMEM [lane = d][col = q] -> smem [q][d] fp32 -> one TMA reduce-add into dq_acc, under
      //      the MMAs of step it.  The CTA barriers also publish metadata(it+2).
      if (it >= 0) {
        const int j = it - 1;
        uint32_t dq[32];
        if (j >= 0) {
          mbar_wait(dq_full, j & 1);
          tc_fence_after();
          tmem_ld_32x32b_x32(t_lane + C::kColDQT + c_base, dq);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(dq_empty);
          if (threadIdx.x == 0) {
            bool published = false;
            if constexpr (kExtras) {
              if (pend_order != nullptr) {
                tma_store_wait<0>();                        // previous reduce-add has been performed in memory
                st_release_gpu_s32(pend_order, (int)blockIdx.y + 1);
                pend_order = nullptr;
                published = true;
              }
            }
            if (!published) tma_store_wait_read<0>();       // previous reduce has finished reading the staging tile
          }
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");
        if (j >= 0) {
          float* stage = reinterpret_cast<float*>(smem + C::kDQOff);
#pragma unroll
          for (int c = 0; c < 32; ++c) stage[(c_base + c) * kBwdD + r] = __uint_as_float(dq[c]);
        }
        if (it + 2 < n_total) meta_store(it + 2);           // ring slot (it+2) % 3 was last read two steps ago
        fence_proxy_async_smem();
        asm volatile("bar.sync 2, 256;" ::: "memory");
        if (j >= 0) {
          if (threadIdx.x == 0) {
            if constexpr (kExtras) {
              if (p.dq_order != nullptr) {
                int* ctr = p.dq_order + (size_t)d_h * p.n_qblk + (i_lo + d_qi);
                if (i_lo + d_qi <= prev_hi_blk) {
                  unsigned spins = 0;
                  while (ld_acquire_gpu_s32(ctr) != (int)blockIdx.y) {
                    __nanosleep(64);
                    if (++spins > (1u << 26)) __trap();     // > 4 s: the predecessor never came; fail, do not hang
                  }
                  fence_proxy_async_all();                  // the earlier blocks' adds precede this async-proxy add
                }
                pend_order = ctr;
              }
            }
            tma_reduce_add_2d(&tmDQ, sbase + C::kDQOff, d_h * p.head_dim, (i_lo + d_qi) * kBwdQ);
            tma_store_commit();
          }
          advance(d_h, d_qi);
        }
      }

      // ---- while the tensor cores run dP^T(it+1) and dV / dK / dQ^T(it): P^T(it+1) = exp2(S^T c - L) masked
      if (it + 1 < n_total) {
        const int nx = it + 1;
        const int q0 = (i_lo + n_qi) * kBwdQ;
        n_qi += S;
        while (n_qi >= nq) n_qi -= nq;
        const float* mL = meta(nx) + c_base;
        const bool need_mask = tail_tile || (q0 < k0 + kBwdKV - 1) || (q0 + kBwdQ > q_end_first);
        mbar_wait(s_full, nx & 1);
        tc_fence_after();
        tmem_ld_32x32b_x32(t_lane + C::kColST + c_base, sv);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(s_empty);                              // S^T(it+2) may overwrite the columns now
        const uint32_t off0 = (uint32_t)(q0 + c_base - kvi);
        if constexpr (kExtras) {
          const bool capped = p.softcap_in > 0.f;
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) {
            const float4 l4 = *reinterpret_cast<const float4*>(mL + ch * 4);
            const float l[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = ch * 4 + e;
              float pv;
              if (capped) {
                tv[c] = bwd_tanh(__uint_as_float(sv[c]) * p.softcap_in);
                pv = bwd_exp2(fmaf(tv[c], p.softcap_out, -l[e]));
              } else {
                pv = bwd_exp2(fmaf(__uint_as_float(sv[c]), p.scale_log2, -l[e]));
              }
              if (need_mask) pv = (off0 + (uint32_t)c < valid_len) ? pv : 0.f;
              sv[c] = __float_as_uint(pv);
            }
          }
        } else if (need_mask) {
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) {
            const float4 l4 = *reinterpret_cast<const float4*>(mL + ch * 4);
            const float l[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = ch * 4 + e;
              const float pv = bwd_exp2(fmaf(__uint_as_float(sv[c]), p.scale_log2, -l[e]));
              sv[c] = (off0 + (uint32_t)c < valid_len) ? __float_as_uint(pv) : 0u;
            }
          }
        } else {
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) {
            const float4 l4 = *reinterpret_cast<const float4*>(mL + ch * 4);
            const float l[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = ch * 4 + e;
              sv[c] = __float_as_uint(bwd_exp2(fmaf(__uint_as_float(sv[c]), p.scale_log2, -l[e])));
            }
          }
        }
      }
    }
    if (threadIdx.x == 0) {
      tma_store_wait<0>();
      if constexpr (kExtras) {
        if (pend_order != nullptr) st_release_gpu_s32(pend_order, (int)blockIdx.y + 1);
      }
    }

    // ---- dK (x softmax scale) / dV : TMEM -> bf16 -> smem (K / V tile regions) -> TMA store
    mbar_wait(acc_done, 0);
    tc_fence_after();
    if (S > 1) {
      // several CTAs share this kv block: fp32 tiles [128 kv][128 d] (dK over the K | V tile regions, dV over the
      // Q / dO ring: every operand has been consumed) are reduce-ADDED into dkv_acc, two 64-row boxes each
      const int which = hf;                                // warps 0-3 drain dK, warps 4-7 drain dV
      float* stage = reinterpret_cast<float*>(smem + (which == 0 ? C::kKOff : C::kQOff)) + r * kBwdD;
      const uint32_t col = which == 0 ? C::kColDK : C::kColDV;
      float mul = which == 0 ? p.scale : 1.f;
      if constexpr (kExtras) {
        if (which == 1 && p.drop_thr != 0) mul = p.drop_scale;
      }
#pragma unroll 1
      for (int c0 = 0; c0 < kBwdD; c0 += 32) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(t_lane + col + c0, o);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; c += 4)
          *reinterpret_cast<float4*>(stage + c0 + c) =
              make_float4(__uint_as_float(o[c]) * mul, __uint_as_float(o[c + 1]) * mul,
                          __uint_as_float(o[c + 2]) * mul, __uint_as_float(o[c + 3]) * mul);
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          tma_reduce_add_2d(&tmDKVacc, sbase + C::kKOff + b * (kBwdQ * kBwdD * 4), g * p.head_dim, k0 + b * kBwdQ);
          tma_reduce_add_2d(&tmDKVacc, sbase + C::kQOff + b * (kBwdQ * kBwdD * 4), (p.hkv + g) * p.head_dim,
                            k0 + b * kBwdQ);
        }
        tma_store_commit();
        tma_store_wait<0>();
      }
    } else {
    {
      const int which = hf;                                // warps 0-3 drain dK, warps 4-7 drain dV
      uint8_t* stage = smem + (which == 0 ? C::kKOff : C::kVOff);
      const uint32_t col = which == 0 ? C::kColDK : C::kColDV;
      float mul = which == 0 ? p.scale : 1.f;
      if constexpr (kExtras) {
        if (which == 1 && p.drop_thr != 0) mul = p.drop_scale;   // dV = (P o mask)^T dO / (1 - p)
      }
#pragma unroll 1
      for (int c0 = 0; c0 < kBwdD; c0 += 32) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(t_lane + col + c0, o);
        tmem_ld_wait();
        uint8_t* box_row = stage + (c0 / 64) * (kBwdKV * 128) + r * 128;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(o[ch * 8 + e]) * mul;
          const int chunk = ((c0 % 64) / 8) + ch;
          *reinterpret_cast<bf16x8*>(box_row + ((chunk ^ (r & 7)) << 4)) = pack8(f);
        }
      }
    }
    fence_proxy_async_smem();
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x == 0) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        tma_store_3d(&tmDKV, sbase + C::kKOff + b * (kBwdKV * 128), b * 64, p.hq + g, k0);
        tma_store_3d(&tmDKV, sbase + C::kVOff + b * (kBwdKV * 128), b * 64, p.hq + p.hkv + g, k0);
      }
      tma_store_commit();
      tma_store_wait<0>();
    }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kBwdTmaWarp) tmem_dealloc<1>(tmem_base, 512);
}

}  // namespace attn
}  // namespace llmt

using namespace llmt;
using namespace llmt::attn;

// How many CTAs should share one kv block's steps (see BwdParams::splits): enough to put ~2 CTAs on every SM when the
// plain grid (kv heads x kv blocks) is smaller than that, at most 8.  1 = no split (no dkv_acc needed).
extern "C" int llmt_attn_bwd_splits(int tokens, int hkv) {
  const int ctas = hkv * ((tokens + kBwdKV - 1) / kBwdKV);
  const int want = 2 * num_sms();
  if (ctas <= 0 || ctas >= want) return 1;
  const int s = (want + ctas - 1) / ctas;
  return s > 8 ? 8 : s;
}

// Backward of llmt_attn_fwd.  delta: fp32 [Hq, T] scratch; dq_acc: fp32 [T, Hq*D] scratch, must be ZERO on
// entry; dqkv: bf16 [T, (Hq+2Hkv)*D] (same row stride as qkv), fully overwritten.
// softcap / dropout_thr / seed: the values the forward call (llmt_attn_fwd_ex) was given.
extern "C" int llmt_attn_bwd_ex(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                float* dq_acc, void* dqkv, const int* seg_start, const int* seg_end, int tokens, int hq,
                                int hkv, int head_dim, int64_t qkv_stride, float scale, int window, float softcap,
                                int dropout_thr, uint32_t seed, int* dq_order, float* dkv_acc,
                                cudaStream_t stream) {
  LLMT_REQUIRE(tokens > 0 && hq > 0 && hkv > 0 && hq % hkv == 0, "attn_bwd: bad head counts %d/%d", hq, hkv);
  LLMT_REQUIRE(head_dim % 16 == 0 && head_dim <= 128, "attn_bwd: head_dim %d unsupported", head_dim);
  LLMT_REQUIRE(qkv_stride % 8 == 0, "attn_bwd: qkv stride must be a multiple of 8");
  {
    const int64_t warps = (int64_t)tokens * hq;
    const int64_t blocks = (warps * 32 + 255) / 256;
    attn_delta_kernel<<<(unsigned)blocks, 256, 0, stream>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout,
                                                            delta, tokens, hq, head_dim);
    int rc = check_launch("attn_delta");
    if (rc) return rc;
  }
  CUtensorMap tkv, tq, tdo, tdkv;
  int rc = make_tmap_thd(&tkv, qkv, tokens, hq + 2 * hkv, head_dim, qkv_stride, kBwdKV);
  if (rc) return rc;
  rc = make_tmap_thd(&tq, qkv, tokens, hq + 2 * hkv, head_dim, qkv_stride, kBwdQ);
  if (rc) return rc;
  rc = make_tmap_thd(&tdo, dout, tokens, hq, head_dim, (int64_t)hq * head_dim, kBwdQ);
  if (rc) return rc;
  rc = make_tmap_thd(&tdkv, dqkv, tokens, hq + 2 * hkv, head_dim, qkv_stride, kBwdKV);
  if (rc) return rc;
  CUtensorMap tdq;
  rc = make_tmap_f32_2d(&tdq, dq_acc, tokens, (int64_t)hq * head_dim, kBwdD, kBwdQ);
  if (rc) return rc;
  BwdParams p;
  p.seg_start = seg_start; p.seg_end = seg_end; p.lse = lse; p.delta = delta; p.dq_acc = dq_acc;
  p.tokens = tokens; p.hq = hq; p.hkv = hkv; p.head_dim = head_dim;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f; p.window = window;
  LLMT_REQUIRE(softcap >= 0.f && dropout_thr >= 0 && dropout_thr < 256, "attn_bwd: softcap must be >= 0 and the "
               "dropout threshold in [0, 255] (got %f, %d)", softcap, dropout_thr);
  p.softcap_in = softcap > 0.f ? scale / softcap : 0.f;
  p.softcap_out = softcap * 1.4426950408889634f;
  p.drop_thr = (uint32_t)dropout_thr;
  p.seed = seed;
  p.drop_scale = 256.f / (256.f - (float)dropout_thr);
  // dq_order != nullptr: bit-reproducible dQ (int scratch of hq * ceil(tokens / 64) entries, zeroed here)
  p.dq_order = dq_order;
  p.n_qblk = (tokens + kBwdQ - 1) / kBwdQ;
  if (dq_order != nullptr) {
    cudaError_t e = cudaMemsetAsync(dq_order, 0, sizeof(int) * (size_t)hq * p.n_qblk, stream);
    LLMT_REQUIRE(e == cudaSuccess, "attn_bwd: clearing the dQ order counters failed: %s", cudaGetErrorString(e));
  }
  const bool extras = softcap > 0.f || dropout_thr > 0 || dq_order != nullptr;
  // dkv_acc != nullptr (fp32 [T, 2*Hkv*D] scratch, zeroed here): the steps of a kv block may be shared by several
  // CTAs.  Not in deterministic mode: the CTAs' dK / dV tiles would be added in arrival order.
  p.splits = 1;
  CUtensorMap tdkvacc = tdq;
  if (dkv_acc != nullptr && dq_order == nullptr) {
    p.splits = llmt_attn_bwd_splits(tokens, hkv);
    if (p.splits > 1) {
      cudaError_t e = cudaMemsetAsync(dkv_acc, 0, sizeof(float) * (size_t)tokens * 2 * hkv * head_dim, stream);
      LLMT_REQUIRE(e == cudaSuccess, "attn_bwd: clearing the dK/dV accumulator failed: %s", cudaGetErrorString(e));
      rc = make_tmap_f32_2d(&tdkvacc, dkv_acc, tokens, (int64_t)2 * hkv * head_dim, kBwdD, kBwdQ);
      if (rc) return rc;
    }
  }
  dim3 grid(hkv, (tokens + kBwdKV - 1) / kBwdKV);
  // LLMT_ATTN_BWD=ss selects the first-generation kernel (every MMA operand in shared memory); default: P^T / dS^T in
  // tensor memory
  static const bool use_ss = [] { const char* e = getenv("LLMT_ATTN_BWD"); return e && e[0] == 's'; }();
  LLMT_REQUIRE(!(use_ss && extras), "attn_bwd: LLMT_ATTN_BWD=ss has no soft-capping / dropout / deterministic mode");
  if (use_ss) {
    static bool configured[kMaxDevices] = {};
    rc = ensure_smem_attr(attn_bwd_kernel, BwdCfg::kSmemBytes, configured, "attn_bwd");
    if (rc) return rc;
    p.splits = 1;
    attn_bwd_kernel<<<grid, kBwdThreads, BwdCfg::kSmemBytes, stream>>>(tkv, tq, tdo, tdkv, tdq, p);
  } else {
    grid.x = hkv * p.splits;
    static bool configured[4][kMaxDevices] = {};
    auto launch = [&](auto kernel, bool* flags) -> int {
      int r = ensure_smem_attr(kernel, BwdTsCfg::kSmemBytes, flags, "attn_bwd");
      if (r) return r;
      kernel<<<grid, kBwdThreads, BwdTsCfg::kSmemBytes, stream>>>(tkv, tq, tdo, tdkv, tdq, tdkvacc, p);
      return 0;
    };
    const bool split = p.splits > 1;
    if (extras && split) rc = launch(attn_bwd_ts_kernel<true, true>, configured[3]);
    else if (extras)     rc = launch(attn_bwd_ts_kernel<true, false>, configured[2]);
    else if (split)      rc = launch(attn_bwd_ts_kernel<false, true>, configured[1]);
    else                 rc = launch(attn_bwd_ts_kernel<false, false>, configured[0]);
    if (rc) return rc;
  }
  rc = check_launch("attn_bwd");
  if (rc) return rc;
  const int width = hq * head_dim;
  int64_t blocks = ((int64_t)tokens * width / 8 + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  // the TS kernel accumulates unscaled dS^T K; the softmax scale is folded into this pass
  attn_dq_convert_kernel<<<(unsigned)blocks, 256, 0, stream>>>(dq_acc, (__nv_bfloat16*)dqkv, tokens, width,
                                                               qkv_stride, use_ss ? 1.f : scale);
  rc = check_launch("attn_dq_convert");
  if (rc || use_ss || p.splits == 1) return rc;
  // shared kv blocks: dK | dV were accumulated in fp32 (softmax scale already applied)
  const int kv_width = 2 * hkv * head_dim;
  blocks = ((int64_t)tokens * kv_width / 8 + 255) / 256;
  if (blocks > cap) blocks = cap;
  attn_dq_convert_kernel<<<(unsigned)blocks, 256, 0, stream>>>(dkv_acc, (__nv_bfloat16*)dqkv + (int64_t)hq * head_dim,
                                                               tokens, kv_width, qkv_stride, 1.f);
  return check_launch("attn_dkv_convert");
}

extern "C" int llmt_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                             float* dq_acc, void* dqkv, const int* seg_start, const int* seg_end, int tokens, int hq,
                             int hkv, int head_dim, int64_t qkv_stride, float scale, int window,
                             cudaStream_t stream) {
  return llmt_attn_bwd_ex(qkv, out, dout, lse, delta, dq_acc, dqkv, seg_start, seg_end, tokens, hq, hkv, head_dim,
                          qkv_stride, scale, window, 0.f, 0, 0u, nullptr, nullptr, stream);
}
