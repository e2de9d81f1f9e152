// Packed causal attention, forward, on tcgen05 tensor cores.
//
//   O[t] = softmax_j( scale * q[t].k[j] ) v[j]     over   lo[t] <= j <= t
//   lo[t] = max(seg_start[t], t - window + 1)      (seg_start[t] = first token of t's document)
//
// Replaces flash_attn_varlen_func (reference: src/llm_training/ops/attention_op.py:606-618, flash-attn's
// mma.sync kernels) and the per-layer mask->cu_seqlens / unpad / pad traffic around it (:375-462).
// The token stream stays packed [T, (Hq+2Hkv)*D]; documents are expressed by per-token bounds, so the
// kernel tiles T uniformly and skips KV blocks outside [lo(tile), hi(tile)].
//
// One CTA = 128 query rows of one head; KV is streamed in blocks of 64 rows.
//   warps 0-3 : softmax — thread r owns query row r == TMEM lane r.  S block: tcgen05.ld -> mask ->
//               online max/sum (exp2, lazy O rescale) -> P as packed bf16 pairs written back into TENSOR MEMORY, over
//               the first 32 columns of the S block it was computed from (tcgen05.st)
//   warp  4   : tcgen05.mma issuer (one elected lane):  S = Q K^T (SS, both K-major),  O += P V (TS form: A = P from
//               TMEM, V is the MN-major B operand in smem)
//   warp  5   : TMA producer (lane 0): Q once, K/V double-buffered
// S is double-buffered in TMEM (2 x 64 columns) so S(j+1) is computed while softmax works on S(j); O lives in
// TMEM (DPAD columns).  P never touches shared memory: no 16 KB tile write + read back per step, no generic->async
// proxy fence (the tensor pipe executes in issue order: S(j+2), which overwrites P(j), is issued after PV(j)).
// ~97 KB smem and 256 TMEM columns per CTA -> two CTAs per SM, so one CTA's softmax overlaps the other's MMAs.
#include <cuda.h>

#include "common.cuh"
#include "sm100.cuh"

namespace llmt {
namespace gemm { int make_tmap_2d(CUtensorMap*, const void*, int, int64_t, int64_t, int64_t, int, int); }
namespace attn {

using namespace sm100;

constexpr int kBlockQ = 128;
constexpr int kBlockKV = 64;
constexpr int kFwdThreads = 192;

template <int DPAD>
struct FwdCfg {
  static constexpr int kBoxes = DPAD / 64;                    // 64-column (128-byte) boxes per row
  static constexpr int kQBytes = kBlockQ * DPAD * 2;          // 32 KB @128
  static constexpr int kKVBytes = kBlockKV * DPAD * 2;        // 16 KB @128, per stage
  static constexpr int kPBytes = kBlockQ * kBlockKV * 2;      // 16 KB
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + kQBytes;
  static constexpr int kVOff = kKOff + 2 * kKVBytes;
  static constexpr int kBarOff = kVOff + 2 * kKVBytes;
  static constexpr int kNumBars = 15;   // q_full, k/v full+empty x2, s full+empty x2, p_full, o_done
  static constexpr int kSmemBytes = kBarOff + (kNumBars + 1) * 8 + 16;
  static constexpr int kTmemCols = (128 + DPAD) <= 256 ? 256 : 512;
  static constexpr int kOCol = 128;                           // S0 @0, S1 @64, O @128
};

struct FwdParams {
  const int* seg_start;
  float* lse;          // [Hq, T], log2 units: m + log2(l)
  int tokens, hq, hkv, head_dim;
  float scale_log2;    // softmax scale * log2(e)
  int window;          // 0 = none
  // extras (attn_fwd_kernel<DPAD, true> only)
  float softcap_in;    // scale / softcap            (0 = no soft-capping)
  float softcap_out;   // softcap * log2(e)
  uint32_t drop_thr;   // an element is DROPPED when its random byte < drop_thr (p = drop_thr / 256; 0 = no dropout)
  uint32_t seed;
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int DPAD, bool kExtras>
__global__ void __launch_bounds__(kFwdThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV_q,   // box {64, 1, 128} over [T][heads][D]
                const __grid_constant__ CUtensorMap tmQKV_kv,  // box {64, 1, 64}
                const __grid_constant__ CUtensorMap tmO,       // box {64, 1, 128} over [T][Hq][D]
                FwdParams p) {
  using C = FwdCfg<DPAD>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0) __trap();

  const uint32_t bars = sbase + C::kBarOff;
  const uint32_t q_full = bars + 0;
  auto k_full = [&](int s) { return bars + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bars + 8u * (3 + s); };
  auto v_full = [&](int s) { return bars + 8u * (5 + s); };
  auto v_empty = [&](int s) { return bars + 8u * (7 + s); };
  auto s_full = [&](int s) { return bars + 8u * (9 + s); };
  auto s_empty = [&](int s) { return bars + 8u * (11 + s); };
  const uint32_t p_full = bars + 8u * 13;
  const uint32_t o_done = bars + 8u * 14;
  const uint32_t tmem_ptr_smem = bars + 8u * 15;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem + C::kBarOff + 8 * 15);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  // Heavy (late) query tiles first ACROSS heads: the head index is the fast grid axis, so the 128-step tiles of every
  // head are dispatched before any light tile (head-major order would start the last head's heaviest tile near the end
  // of the kernel: +13% makespan for a causal 8192-token launch).  The four query heads of a kv group run together.
  const int num_q_tiles = (p.tokens + kBlockQ - 1) / kBlockQ;
  const int qt = num_q_tiles - 1 - (int)blockIdx.y;
  const int head = blockIdx.x;
  const int kv_head = head / (p.hq / p.hkv);
  const int q0 = qt * kBlockQ;
  const int q_last = min(q0 + kBlockQ, p.tokens) - 1;

  // KV range of this tile: bounds are non-decreasing in t, so the tile minimum is the first row's bound
  auto lower_bound = [&](int t) {
    int lo = p.seg_start[t];
    if (p.window > 0) lo = max(lo, t - p.window + 1);
    return lo;
  };
  const int lo_first = lower_bound(q0);
  const int lo_last = lower_bound(q_last);
  const int jb_lo = lo_first / kBlockKV;
  const int jb_hi = q_last / kBlockKV;
  const int nb = jb_hi - jb_lo + 1;

  if (warp == 4 && lane == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(k_full(s), 1);
      mbar_init(k_empty(s), 1);
      mbar_init(v_full(s), 1);
      mbar_init(v_empty(s), 1);
      mbar_init(s_full(s), 1);
      mbar_init(s_empty(s), 128);
    }
    mbar_init(p_full, 128);
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 5) {
    if (lane == 0) {
      tma_prefetch_desc(&tmQKV_q);
      tma_prefetch_desc(&tmQKV_kv);
      tma_prefetch_desc(&tmO);
    }
    tmem_alloc<1>(tmem_ptr_smem, C::kTmemCols);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 5) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, C::kQBytes);
#pragma unroll
      for (int b = 0; b < C::kBoxes; ++b)
        tma_load_3d(sbase + C::kQOff + b * (kBlockQ * 128), &tmQKV_q, q_full, b * 64, head, q0);
      for (int j = 0; j < nb; ++j) {
        const int st = j & 1;
        const uint32_t par = ((j >> 1) & 1) ^ 1;
        const int kv0 = (jb_lo + j) * kBlockKV;
        mbar_wait(k_empty(st), par);
        mbar_arrive_expect_tx(k_full(st), C::kKVBytes);
#pragma unroll
        for (int b = 0; b < C::kBoxes; ++b)
          tma_load_3d(sbase + C::kKOff + st * C::kKVBytes + b * (kBlockKV * 128), &tmQKV_kv, k_full(st), b * 64,
                      p.hq + kv_head, kv0);
        mbar_wait(v_empty(st), par);
        mbar_arrive_expect_tx(v_full(st), C::kKVBytes);
#pragma unroll
        for (int b = 0; b < C::kBoxes; ++b)
          tma_load_3d(sbase + C::kVOff + st * C::kKVBytes + b * (kBlockKV * 128), &tmQKV_kv, v_full(st), b * 64,
                      p.hq + p.hkv + kv_head, kv0);
      }
    }
  } else if (warp == 4) {
    // ===== MMA issuer: the whole warp walks the loop in uniform control flow, one elected lane issues; smem
    // descriptors are built once and a k-step only adds a constant to the start-address field (a divergent
    // `if (lane == 0)` made ptxas wrap every tcgen05.mma in an election loop with vector->uniform moves).
    constexpr uint32_t idesc_s = make_idesc_bf16(kBlockQ, kBlockKV, false, false);
    constexpr uint32_t idesc_pv = make_idesc_bf16(kBlockQ, DPAD, false, true);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t q_a = make_smem_desc(sbase + C::kQOff, 0, 1024);                    // Q as A, K-major
    const uint64_t k_b0 = make_smem_desc(sbase + C::kKOff, 0, 1024);                   // K as B, K-major, stage 0
    const uint64_t v_b0 = make_smem_desc(sbase + C::kVOff, kBlockKV * 128, 1024);      // V as B, MN-major, stage 0
    constexpr uint64_t kStage = C::kKVBytes >> 4;
    auto kmaj = [](int kk, int rows) -> uint64_t { return (uint64_t)((kk / 4) * (rows * 8) + (kk % 4) * 2); };
    auto issue_s = [&](int j) {
      const int st = j & 1;
      const uint64_t k_b = k_b0 + (uint64_t)st * kStage;
#pragma unroll
      for (int kk = 0; kk < DPAD / 16; ++kk)
        umma_f16<1>(tb + (uint32_t)(st * kBlockKV), q_a + kmaj(kk, kBlockQ), k_b + kmaj(kk, kBlockKV), idesc_s,
                    kk != 0);
      umma_commit(k_empty(st));
      umma_commit(s_full(st));
    };
    mbar_wait(q_full, 0);
    mbar_wait(k_full(0), 0);
    tc_fence_after();
    if (elect_one()) issue_s(0);
    __syncwarp();
#pragma unroll 1
    for (int j = 0; j < nb; ++j) {
      if (j + 1 < nb) {
        const int st = (j + 1) & 1;
        const uint32_t par = ((j + 1) >> 1) & 1;
        mbar_wait(k_full(st), par);
        mbar_wait(s_empty(st), par ^ 1);
        tc_fence_after();
        if (elect_one()) issue_s(j + 1);
        __syncwarp();
      }
      const int st = j & 1;
      mbar_wait(v_full(st), (j >> 1) & 1);
      mbar_wait(p_full, j & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t v_b = v_b0 + (uint64_t)st * kStage;
#pragma unroll
        for (int kk = 0; kk < kBlockKV / 16; ++kk)
          // P: bf16 pairs in the S block's first columns, 8 columns per k-step; V: MN-major B, [kv rows][64 d]
          // boxes, 16 kv rows per k-step
          umma_f16_ts<1>(tb + C::kOCol, tb + (uint32_t)(st * kBlockKV + kk * 8), v_b + (uint64_t)kk * 128, idesc_pv,
                         (j | kk) != 0);
        umma_commit(v_empty(st));
        umma_commit(o_done);
      }
      __syncwarp();
    }
  } else {
    // ===== softmax warps: thread <-> query row =====
    const int row = warp * 32 + lane;
    const int q_idx = q0 + row;
    const bool row_ok = q_idx < p.tokens;
    const int lo_row = row_ok ? lower_bound(q_idx) : q_idx;
    const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
    float m = -INFINITY, l = 0.f;

    for (int j = 0; j < nb; ++j) {
      const int sb = j & 1;
      const int kv0 = (jb_lo + j) * kBlockKV;
      mbar_wait(s_full(sb), (j >> 1) & 1);
      tc_fence_after();
      uint32_t sv[64];
      {
        uint32_t (&a)[32] = *reinterpret_cast<uint32_t (*)[32]>(&sv[0]);
        uint32_t (&b)[32] = *reinterpret_cast<uint32_t (*)[32]>(&sv[32]);
        tmem_ld_32x32b_x32(t_lane + (uint32_t)(sb * kBlockKV), a);
        tmem_ld_32x32b_x32(t_lane + (uint32_t)(sb * kBlockKV + 32), b);
      }
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_empty(sb));

      // block needs masking if it touches the diagonal or a document / window boundary of this tile
      const bool need_mask = (kv0 + kBlockKV - 1 > q0) || (kv0 < lo_last);
      float mx = -INFINITY;
      if (need_mask) {
#pragma unroll
        for (int c = 0; c < 64; ++c) {
          const int kv = kv0 + c;
          float s = __uint_as_float(sv[c]);
          s = (kv >= lo_row && kv <= q_idx) ? s : -INFINITY;
          sv[c] = __float_as_uint(s);
          mx = fmaxf(mx, s);
        }
      } else {
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // four short max chains
#pragma unroll
        for (int c = 0; c < 64; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], __uint_as_float(sv[c]));
        mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      }
      if constexpr (kExtras) {
        // soft-capping (cap * tanh(s * scale / cap)) turns the scores into their final log2-domain values here;
        // masked entries stay -inf
        if (p.softcap_in > 0.f) {
          mx = -INFINITY;
#pragma unroll
          for (int c = 0; c < 64; ++c) {
            const float s = __uint_as_float(sv[c]);
            const float x = (s == -INFINITY) ? -INFINITY : p.softcap_out * fast_tanh(s * p.softcap_in);
            sv[c] = __float_as_uint(x);
            mx = fmaxf(mx, x);
          }
        }
      }
      const bool capped = kExtras && p.softcap_in > 0.f;
      const float sc = capped ? 1.f : p.scale_log2;         // capped scores are already scaled
      const float m_blk = mx * sc;
      // lazy rescale: keep the stale max unless it is exceeded by more than 2^8
      float alpha = 1.f;
      if (m_blk > m + 8.f || m == -INFINITY) {
        const float m_new = fmaxf(m, m_blk);
        alpha = (m == -INFINITY) ? 0.f : fast_exp2(m - m_new);
        m = m_new;
      }
      const float m_use = (m == -INFINITY) ? 0.f : m;
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};               // four short add chains instead of one of 64
      uint32_t pk[32];
#pragma unroll
      for (int c = 0; c < 64; c += 2) {
        float e0 = fast_exp2(fmaf(__uint_as_float(sv[c]), sc, -m_use));
        float e1 = fast_exp2(fmaf(__uint_as_float(sv[c + 1]), sc, -m_use));
        sum4[(c >> 1) & 3] += e0 + e1;                      // the normaliser is the UNdropped sum
        if constexpr (kExtras) {
          if (p.drop_thr != 0) {
            const uint32_t rnd = dropout_bits(p.seed, head, q_idx, (kv0 + c) >> 2);   // c is even: same 4-group
            e0 = (((rnd >> (8 * ((kv0 + c) & 3))) & 0xffu) < p.drop_thr) ? 0.f : e0;
            e1 = (((rnd >> (8 * ((kv0 + c + 1) & 3))) & 0xffu) < p.drop_thr) ? 0.f : e1;
          }
        }
        __nv_bfloat162 h = __floats2bfloat162_rn(e0, e1);
        pk[c >> 1] = *reinterpret_cast<uint32_t*>(&h);
      }
      l = l * alpha + ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));

      if (j > 0) {
        // Every phase of o_done is waited for, in order (a parity wait is only meaningful one phase behind).  PV(j-1)
        // was issued a whole softmax step ago, so this rarely blocks; only a changed row max needs it (O rescale).
        mbar_wait(o_done, (j - 1) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
          for (int c0 = 0; c0 < DPAD; c0 += 32) {
            uint32_t o[32];
            tmem_ld_32x32b_x32(t_lane + C::kOCol + c0, o);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
            tmem_st_32x32b_x32(t_lane + C::kOCol + c0, o);
          }
          tmem_st_wait();
        }
      }
      // P row -> tensor memory, over the S block it came from (bf16 pairs: column c holds keys 2c, 2c+1)
      tmem_st_32x32b_x32(t_lane + (uint32_t)(sb * kBlockKV), pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full);
    }

    // ---- epilogue: O / l -> bf16 -> smem (Q region) -> TMA store; LSE
    mbar_wait(o_done, (nb - 1) & 1);
    tc_fence_after();
    float inv_l = (l > 0.f) ? 1.f / l : 0.f;
    if constexpr (kExtras) {
      if (p.drop_thr != 0) inv_l *= 256.f / (256.f - (float)p.drop_thr);   // kept elements are scaled by 1 / (1 - p)
    }
    uint8_t* o_stage = smem + C::kQOff;
#pragma unroll
    for (int c0 = 0; c0 < DPAD; c0 += 32) {
      uint32_t o[32];
      tmem_ld_32x32b_x32(t_lane + C::kOCol + c0, o);
      tmem_ld_wait();
      uint8_t* box_row = o_stage + (c0 / 64) * (kBlockQ * 128) + row * 128;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(o[ch * 8 + e]) * inv_l;
        const int chunk = ((c0 % 64) / 8) + ch;
        *reinterpret_cast<bf16x8*>(box_row + ((chunk ^ (row & 7)) << 4)) = pack8(f);
      }
    }
    if (row_ok) p.lse[(size_t)head * p.tokens + q_idx] = (l > 0.f) ? m + log2f(l) : -INFINITY;
    fence_proxy_async_smem();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (threadIdx.x == 0) {
#pragma unroll
      for (int b = 0; b < C::kBoxes; ++b) tma_store_3d(&tmO, sbase + C::kQOff + b * (kBlockQ * 128), b * 64, head, q0);
      tma_store_commit();
      tma_store_wait<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc<1>(tmem_base, C::kTmemCols);
}

// ---- tensor maps over [T][heads][D] views -----------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_tmap_thd(CUtensorMap* out, const void* ptr, int64_t tokens, int heads, int head_dim, int64_t row_stride,
                  int box_rows) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      set_error("cuTensorMapEncodeTiled entry point not available");
      return 3;
    }
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  cuuint64_t dims[3] = {(cuuint64_t)head_dim, (cuuint64_t)heads, (cuuint64_t)tokens};
  cuuint64_t strides[2] = {(cuuint64_t)head_dim * 2, (cuuint64_t)row_stride * 2};
  cuuint32_t box[3] = {64, 1, (cuuint32_t)box_rows};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("attention tensor map encode failed (%d): tokens=%lld heads=%d d=%d stride=%lld", (int)r,
              (long long)tokens, heads, head_dim, (long long)row_stride);
    return 3;
  }
  return 0;
}


// fp32 row-major [rows][cols] tensor map without swizzle (used for the dQ reduce-add tile)
int make_tmap_f32_2d(CUtensorMap* out, const void* ptr, int64_t rows, int64_t cols, int box_cols, int box_rows) {
  CUtensorMap probe;
  (void)probe;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return 3;
  }
  EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(sym);
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("fp32 tensor map encode failed (%d): rows=%lld cols=%lld box=%dx%d", (int)r, (long long)rows,
              (long long)cols, box_cols, box_rows);
    return 3;
  }
  return 0;
}

template <int DPAD, bool kExtras>
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tkv, const CUtensorMap& to, const FwdParams& p,
               cudaStream_t stream) {
  using C = FwdCfg<DPAD>;
  auto kern = attn_fwd_kernel<DPAD, kExtras>;
  static bool configured[kMaxDevices] = {};
  if (int rc = ensure_smem_attr(kern, C::kSmemBytes, configured, "attn_fwd")) return rc;
  dim3 grid(p.hq, (p.tokens + kBlockQ - 1) / kBlockQ);
  kern<<<grid, kFwdThreads, C::kSmemBytes, stream>>>(tq, tkv, to, p);
  return check_launch("attn_fwd");
}

}  // namespace attn
}  // namespace llmt

using namespace llmt;
using namespace llmt::attn;

// qkv: bf16 [T, (Hq + 2*Hkv) * D] (row stride in elements), q heads | k heads | v heads.
// out: bf16 [T, Hq*D] contiguous.  lse: fp32 [Hq, T] (log2 units).  seg_start: int32 [T].
// softcap > 0: scores become softcap * tanh(score * scale / softcap) before the softmax.  dropout_thr in [0, 255]:
// attention probabilities are dropped with p = dropout_thr / 256 and the kept ones scaled by 1 / (1 - p); the mask is
// a pure function of (seed, head, query, key) — see dropout_bits() — so the backward kernel regenerates it.
extern "C" int llmt_attn_fwd_ex(const void* qkv, void* out, float* lse, const int* seg_start, int tokens, int hq,
                                int hkv, int head_dim, int64_t qkv_stride, float scale, int window, float softcap,
                                int dropout_thr, uint32_t seed, cudaStream_t stream) {
  LLMT_REQUIRE(tokens > 0 && hq > 0 && hkv > 0 && hq % hkv == 0, "attn_fwd: bad head counts %d/%d", hq, hkv);
  LLMT_REQUIRE(head_dim % 16 == 0 && head_dim <= 128, "attn_fwd: head_dim %d unsupported (multiple of 16, <= 128)",
               head_dim);
  LLMT_REQUIRE(qkv_stride % 8 == 0 && ((uintptr_t)qkv % 16 == 0) && ((uintptr_t)out % 16 == 0),
               "attn_fwd: qkv/out must be 16-byte aligned");
  CUtensorMap tq, tkv, to;
  int rc = make_tmap_thd(&tq, qkv, tokens, hq + 2 * hkv, head_dim, qkv_stride, kBlockQ);
  if (rc) return rc;
  rc = make_tmap_thd(&tkv, qkv, tokens, hq + 2 * hkv, head_dim, qkv_stride, kBlockKV);
  if (rc) return rc;
  rc = make_tmap_thd(&to, out, tokens, hq, head_dim, (int64_t)hq * head_dim, kBlockQ);
  if (rc) return rc;
  FwdParams p;
  p.seg_start = seg_start;
  p.lse = lse;
  p.tokens = tokens; p.hq = hq; p.hkv = hkv; p.head_dim = head_dim;
  LLMT_REQUIRE(softcap >= 0.f && dropout_thr >= 0 && dropout_thr < 256, "attn_fwd: softcap must be >= 0 and the "
               "dropout threshold in [0, 255] (got %f, %d)", softcap, dropout_thr);
  p.scale_log2 = scale * 1.4426950408889634f;
  p.window = window;
  p.softcap_in = softcap > 0.f ? scale / softcap : 0.f;
  p.softcap_out = softcap * 1.4426950408889634f;
  p.drop_thr = (uint32_t)dropout_thr;
  p.seed = seed;
  const bool extras = softcap > 0.f || dropout_thr > 0;
  if (head_dim <= 64) return extras ? launch_fwd<64, true>(tq, tkv, to, p, stream) : launch_fwd<64, false>(tq, tkv, to, p, stream);
  return extras ? launch_fwd<128, true>(tq, tkv, to, p, stream) : launch_fwd<128, false>(tq, tkv, to, p, stream);
}

extern "C" int llmt_attn_fwd(const void* qkv, void* out, float* lse, const int* seg_start, int tokens, int hq,
                             int hkv, int head_dim, int64_t qkv_stride, float scale, int window,
                             cudaStream_t stream) {
  return llmt_attn_fwd_ex(qkv, out, lse, seg_start, tokens, hq, hkv, head_dim, qkv_stride, scale, window, 0.f, 0, 0u,
                          stream);
}
