// bf16 GEMM on 5th-gen tensor cores:  C[M,N] = op(A)[M,K] * op(B)[K,N] (+ R[M,N]),  fp32 accumulation.
//
// This is the kernel behind every nn.Linear of the reference hot path (cuBLAS there; SURVEY §2.3 K18:
// models/llama/llama_model.py:459-462, :421-423, :55) — forward (both operands K-major), dgrad
// (B MN-major) and wgrad (A and B MN-major), with the residual add / gradient accumulation fused into
// the epilogue (R == C gives C += A*B).
//
// Design (B200, sm_100a):
//   * persistent, warp-specialised, one CTA per SM:  warp 0 = TMA producer, warp 1 = tcgen05.mma issuer
//     (one elected lane), warp 2 = TMEM allocator, warps 4..7 = epilogue (TMEM -> regs -> smem -> TMA store)
//   * operands arrive by TMA (cp.async.bulk.tensor, SWIZZLE_128B) into a multi-stage smem ring guarded by
//     mbarriers; accumulators live in TMEM (2 x 256 columns, so the epilogue of tile i overlaps the MMAs of
//     tile i+1)
//   * kCtaGroup == 2: a CTA pair (cluster of 2, cta_group::2) computes a 256x256 tile; each CTA loads its own
//     128 rows of A and HALF of the B tile, the MMA reads both halves -> half the L2->smem operand traffic
//     per flop of the 1-CTA form. Only the leader CTA issues MMAs; commits are multicast to both CTAs.
//   * BLOCK 128(x2) x 256 x 64, UMMA 128(256) x 256 x 16.
#include <cuda.h>

#include <mutex>
#include <type_traits>
#include <unordered_map>

#include "common.cuh"
#include "sm100.cuh"

namespace llmt {
namespace gemm {

using namespace sm100;

constexpr int BM = 128;       // rows of C per CTA
constexpr int BN = 256;       // cols of C per tile (UMMA N)
constexpr int BK = 64;        // K per pipeline stage (= one 128-byte swizzle row of bf16)
constexpr int UMMA_K = 16;
constexpr int kThreads = 256;
constexpr int kEpiWarp0 = 4;  // warps 4..7
constexpr int kAccStages = 2;
constexpr int kTmemCols = 512;
constexpr int kStagingBytes = 16384;  // 128 rows x 128 B
constexpr int kNumStaging = 2;

template <int kCtaGroup>
struct Cfg {
  static constexpr int kABytes = BM * BK * 2;                  // 16 KB
  static constexpr int kBRows = BN / kCtaGroup;                // B rows (N) held by one CTA
  static constexpr int kBBytes = kBRows * BK * 2;              // 32 KB or 16 KB
  static constexpr int kStageBytes = kABytes + kBBytes;        // 48 KB or 32 KB
  static constexpr int kStages = (kCtaGroup == 1) ? 4 : 6;     // 192 KB of operand ring either way
  static constexpr int kRingBytes = kStages * kStageBytes;
  static constexpr int kBarOffset = kRingBytes + kNumStaging * kStagingBytes;
  // full[kStages] empty[kStages] tmem_full[2] tmem_empty[2] res[2] + tmem ptr
  static constexpr int kNumBars = 2 * kStages + 2 * kAccStages + kNumStaging;
  static constexpr int kSmemBytes = kBarOffset + kNumBars * 8 + 16 + 1024;  // +1024: manual alignment slack
};

struct Params {
  int M, N, K;
  int has_residual;
  int rows_per_rank;   // dist variants: rows of the gathered A / scattered C owned by one rank
  int tm_rot;          // land-once gather: first m-block of the tile walk (this rank's own rows)
};

// Fused collective variants (tensor parallel paths):
//   kDistA : A is the row-wise concatenation of every rank's [rows_per_rank, K] block living in symmetric memory
//            (all-gather -> GEMM).
//            1: the TMA producer loads each A tile straight from its owner over NVLink; simple, but a remote row
//               panel crosses NVLink once per N tile.
//            2: "land once": eight extra warps per CTA pull the remote blocks over NVLink into a local [M, K]
//               buffer, block by block in the order the tile schedule consumes them (own rows first, then rank+1,
//               rank+2, ...), and bump a per-block arrival counter; the TMA producer waits on the counter of a
//               remote block before its first load.  Every byte crosses NVLink once, the copy hides under the
//               MMAs of the earlier blocks, and the gathered activation is left behind for the backward pass.
//            3: "weight gather" (ZeRO-3 parameter all-gather -> first consumer GEMM): B = W[N, K] is a row range of
//               a unit's flat parameter buffer whose slices live on their owner ranks (symmetric shard arena).
//               The pull warps land W in the local full buffer n-block by n-block; tiles are walked n-major so
//               the producer of n-block j only waits for the bytes of block j.
//   kDistC : the [M, N] result is reduce-scattered by rows; each tile is TMA-reduce-ADDED into the owner rank's
//            [rows_per_rank, N] buffer (GEMM -> reduce-scatter)
constexpr int kMaxDistRanks = 8;
struct DistMaps {
  CUtensorMap a[kMaxDistRanks];
  CUtensorMap c[kMaxDistRanks];
  // land-once all-gather (kDistA == 2): raw views for the pull warps + arrival counters
  const void* a_raw[kMaxDistRanks];   // rank r's [rows_per_rank, K] block (symmetric memory), lda elements per row
  void* a_full;                       // local [M, K] landing buffer, ld_full elements per row
  int* flags;                         // one counter per (BM * kCtaGroup)-row block of a_full (counts 4 KB pieces
                                      // delivered), zero on entry
  int64_t lda, ld_full;
  int rank, world;
  // weight gather (kDistA == 3): a_raw[r] = rank r's slice of the unit, a_full = the unit's full buffer;
  // W starts w_off elements into the flat buffer, every slice holds shard_numel elements
  int64_t w_off, shard_numel;
};
struct NoDist {};
constexpr int kPullUnroll = 8;                  // 16-byte loads in flight per lane of a pull warp
constexpr int kPieceVecs = 32 * kPullUnroll;    // vectors per pull work unit (4 KB)

// grouped rasterisation: consecutive tile ids walk down M inside a band of kGroupM m-tiles so that the
// ~148 tiles in flight share A and B panels in L2
__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& tm, int& tn) {
  constexpr int kGroupM = 8;
  const int per_group = kGroupM * num_n;
  const int g = tile / per_group;
  const int first_m = g * kGroupM;
  const int group_m = min(num_m - first_m, kGroupM);
  const int r = tile - g * per_group;
  tm = first_m + (r % group_m);
  tn = r / group_m;
}

template <int kCtaGroup, bool kAMn, bool kBMn, typename OutT, int kDistA = 0, bool kDistC = false>
__global__ void __launch_bounds__((kDistA >= 2 ? 2 : 1) * kThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmR, Params p,
                 const __grid_constant__ typename std::conditional<(kDistA != 0 || kDistC), DistMaps, NoDist>::type dist) {
  static_assert(!(kDistA != 0 && kDistA != 3 && kAMn), "gathered A must be K-major");
  static_assert(!(kDistA == 3 && kBMn), "gathered weights are K-major ([N, K] rows)");
  using C = Cfg<kCtaGroup>;
  constexpr int kOutChunkCols = 128 / (int)sizeof(OutT);  // columns per 128-byte staging row: 64 bf16 / 32 fp32
  constexpr int kNumChunks = BN / kOutChunkCols;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024-B alignment
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t bars = smem_base + C::kBarOffset;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (C::kStages + s); };
  auto tmem_full_bar = [&](int a) { return bars + 8u * (2 * C::kStages + a); };
  auto tmem_empty_bar = [&](int a) { return bars + 8u * (2 * C::kStages + kAccStages + a); };
  auto res_bar = [&](int b) { return bars + 8u * (2 * C::kStages + 2 * kAccStages + b); };
  const uint32_t tmem_ptr_smem = bars + 8u * C::kNumBars;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + C::kBarOffset + 8 * C::kNumBars);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (kCtaGroup == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;

  const int num_m = (p.M + BM * kCtaGroup - 1) / (BM * kCtaGroup);
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = (p.K + BK - 1) / BK;
  const int worker = blockIdx.x / kCtaGroup;
  const int num_workers = gridDim.x / kCtaGroup;
  // land-once gather: walk the m-blocks starting at this rank's own rows (the order the pull warps fill them)
  auto tile_mn = [&](int tile, int& tm, int& tn) {
    tile_coords(tile, num_m, num_n, tm, tn);
    if constexpr (kDistA == 2) {
      tm += p.tm_rot;
      if (tm >= num_m) tm -= num_m;
    }
    if constexpr (kDistA == 3) {          // n-major: all m-tiles of weight block 0, then block 1, ... (arrival order)
      tm = tile % num_m;
      tn = tile / num_m;
    }
  };

  // ---------------------------------------------------------------- one-time setup
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    if (p.has_residual) tma_prefetch_desc(&tmR);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < kAccStages; ++a) {
      mbar_init(tmem_full_bar(a), 1);
      mbar_init(tmem_empty_bar(a), 4 * kCtaGroup);  // one arrive per epilogue warp of every CTA in the group
    }
    for (int b = 0; b < kNumStaging; ++b) mbar_init(res_bar(b), 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc<kCtaGroup>(tmem_ptr_smem, kTmemCols);
    tmem_relinquish<kCtaGroup>();
  }
  tc_fence_before();
  if constexpr (kCtaGroup == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  // ---------------------------------------------------------------- roles
  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      // in a CTA pair all operand bytes are accounted on the LEADER's full barrier
      const uint32_t full0 = (kCtaGroup == 2) ? mapa(full_bar(0), 0) : full_bar(0);
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = worker; tile < num_tiles; tile += num_workers) {
        int tm, tn;
        tile_mn(tile, tm, tn);
        const int m0 = (tm * kCtaGroup + (int)cta_rank) * BM;
        const int n0 = tn * BN + (int)cta_rank * C::kBRows;
        [[maybe_unused]] const CUtensorMap* mapA = &tmA;
        [[maybe_unused]] int m_local = m0;
        if constexpr (kDistA == 3) {
          // weight rows of this n-block come from the landing buffer once every piece of the block has arrived
          const int rows_here = min(BN, p.N - tn * BN);
          const int target = (int)(((int64_t)rows_here * (p.K >> 3) + kPieceVecs - 1) / kPieceVecs);
          const int* flag = dist.flags + tn;
          uint32_t spins = 0;
          while (ld_acquire_gpu_s32(flag) < target) {
            if (++spins > (1u << 26)) __trap();
          }
          fence_proxy_async_all();
        }
        if constexpr (kDistA == 1 || kDistA == 2) {    // A tile lives on rank m0 / rows_per_rank
          const int owner = min(m0 / p.rows_per_rank, kMaxDistRanks - 1);
          if (kDistA == 1 || owner == dist.rank) {
            mapA = &dist.a[owner];
            m_local = m0 - owner * p.rows_per_rank;
          } else {
            // remote rows come from the landing buffer once the pull warps of every CTA have delivered the block
            const int* flag = dist.flags + tm;
            const int target = (int)(((int64_t)BM * kCtaGroup * (p.K >> 3) + kPieceVecs - 1) / kPieceVecs);
            uint32_t spins = 0;
            while (ld_acquire_gpu_s32(flag) < target) {
              if (++spins > (1u << 26)) __trap();
            }
            fence_proxy_async_all();                   // generic-proxy writes (pull warps) -> async-proxy reads (TMA)
          }
        }
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * C::kStageBytes;
          const uint32_t sb = sa + C::kABytes;
          const uint32_t fb = full0 + 8u * stage;
          const int k0 = kb * BK;
          if (kCtaGroup == 1 || is_leader) mbar_arrive_expect_tx(full_bar(stage), C::kStageBytes * kCtaGroup);
          if constexpr (kCtaGroup == 1) {
            if constexpr (kDistA == 1 || kDistA == 2) tma_load_2d(sa, mapA, fb, k0, m_local);
            else if constexpr (!kAMn) tma_load_2d(sa, &tmA, fb, k0, m0);
            else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tmA, fb, m0 + j * 64, k0);
            }
            if constexpr (!kBMn) tma_load_2d(sb, &tmB, fb, k0, n0);
            else {
#pragma unroll
              for (int j = 0; j < C::kBRows / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tmB, fb, n0 + j * 64, k0);
            }
          } else {
            if constexpr (kDistA == 1 || kDistA == 2) tma_load_2d_cg2(sa, mapA, fb, k0, m_local);
            else if constexpr (!kAMn) tma_load_2d_cg2(sa, &tmA, fb, k0, m0);
            else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j) tma_load_2d_cg2(sa + j * (BK * 128), &tmA, fb, m0 + j * 64, k0);
            }
            if constexpr (!kBMn) tma_load_2d_cg2(sb, &tmB, fb, k0, n0);
            else {
#pragma unroll
              for (int j = 0; j < C::kBRows / 64; ++j)
                tma_load_2d_cg2(sb + j * (BK * 128), &tmB, fb, n0 + j * 64, k0);
            }
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only) =====
    if (is_leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * kCtaGroup, BN, kAMn, kBMn);
      constexpr uint32_t kALbo = kAMn ? BK * 128 : 0, kBLbo = kBMn ? BK * 128 : 0;
      constexpr uint32_t kAKStep = kAMn ? UMMA_K * 128 : UMMA_K * 2;  // bytes to advance 16 along K
      constexpr uint32_t kBKStep = kBMn ? UMMA_K * 128 : UMMA_K * 2;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = worker; tile < num_tiles; tile += num_workers) {
        mbar_wait_cluster(tmem_empty_bar(acc), acc_phase ^ 1);  // epilogue(s) drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait_cluster(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * C::kStageBytes;
          const uint32_t sb = sa + C::kABytes;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t ad = make_smem_desc(sa + k * kAKStep, kALbo, 1024);
            const uint64_t bd = make_smem_desc(sb + k * kBKStep, kBLbo, 1024);
            umma_f16<kCtaGroup>(d_tmem, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          // frees this smem stage (in both CTAs of a pair) once the MMAs above have read it
          if constexpr (kCtaGroup == 1) umma_commit(empty_bar(stage));
          else umma_commit_cg2_mcast(empty_bar(stage), 0b11);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        if constexpr (kCtaGroup == 1) umma_commit(tmem_full_bar(acc));
        else umma_commit_cg2_mcast(tmem_full_bar(acc), 0b11);
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (kDistA == 3 && warp >= kThreads / 32) {
    // ===== pull warps (weight gather): W's flat element range, 4 KB pieces dealt round-robin to every pull warp =====
    if constexpr (kDistA == 3) {
      constexpr int kPullWarps = kThreads / 32;
      const int pw = warp - kThreads / 32;
      const int64_t gw = (int64_t)blockIdx.x * kPullWarps + pw, GW = (int64_t)gridDim.x * kPullWarps;
      const int64_t total_vecs = (int64_t)p.N * (p.K >> 3);
      const int64_t vecs_per_blk = (int64_t)BN * (p.K >> 3);               // multiple of kPieceVecs
      const int64_t pieces = (total_vecs + kPieceVecs - 1) / kPieceVecs;
      __nv_bfloat16* full = reinterpret_cast<__nv_bfloat16*>(dist.a_full);
      for (int64_t piece = gw; piece < pieces; piece += GW) {
        uint4 r[kPullUnroll];
        int64_t e[kPullUnroll];
#pragma unroll
        for (int u = 0; u < kPullUnroll; ++u) {
          const int64_t v = piece * kPieceVecs + u * 32 + lane;
          e[u] = -1;
          if (v < total_vecs) {
            e[u] = dist.w_off + (v << 3);                                   // element index in the flat buffer
            const int64_t owner = e[u] / dist.shard_numel;
            const int64_t local = e[u] - owner * dist.shard_numel;
            r[u] = ld_relaxed_sys_16(reinterpret_cast<const __nv_bfloat16*>(dist.a_raw[owner]) + local);
          }
        }
#pragma unroll
        for (int u = 0; u < kPullUnroll; ++u)
          if (e[u] >= 0) *reinterpret_cast<uint4*>(full + e[u]) = r[u];
        __syncwarp();
        if (lane == 0) {
          __threadfence();
          red_release_gpu_add_s32(dist.flags + (int)((piece * kPieceVecs) / vecs_per_blk), 1);
        }
      }
    }
  } else if (kDistA == 2 && warp >= kThreads / 32) {
    // ===== pull warps (land-once gather): remote blocks -> local landing buffer, one arrival per CTA and block =====
    if constexpr (kDistA == 2) {
      // Work unit = one "piece": kPieceVecs consecutive 16-byte vectors of a block, moved by ONE warp with
      // kPullUnroll loads in flight per lane.  Pieces are numbered over all remote blocks in consumption order
      // and dealt round-robin to every pull warp of the grid, so warps never meet at a CTA barrier and the
      // copy of block b+1 overlaps the tail of block b.  A block is complete when all of its pieces have
      // arrived (one counter increment per piece).
      constexpr int kPullWarps = kThreads / 32;
      constexpr int kBlkRows = BM * kCtaGroup;
      const int pw = warp - kThreads / 32;
      const int64_t gw = (int64_t)blockIdx.x * kPullWarps + pw, GW = (int64_t)gridDim.x * kPullWarps;
      const int bpr = p.rows_per_rank / kBlkRows;
      const int vec_per_row = p.K >> 3;
      const int64_t vecs = (int64_t)kBlkRows * vec_per_row;
      const int64_t ppb = (vecs + kPieceVecs - 1) / kPieceVecs;           // pieces per block
      const int64_t remote_pieces = (int64_t)(dist.world - 1) * bpr * ppb;
      const int64_t all_pieces = (int64_t)dist.world * bpr * ppb;         // own block last: copied, not flagged
      for (int64_t piece = gw; piece < all_pieces; piece += GW) {
        const int64_t blk_seq = piece / ppb;                              // position in consumption order
        const int64_t in_blk = piece - blk_seq * ppb;
        const int s = (int)(blk_seq / bpr) + 1, b = (int)(blk_seq % bpr);
        const int owner = (dist.rank + s) % dist.world;
        const __nv_bfloat16* src =
            reinterpret_cast<const __nv_bfloat16*>(dist.a_raw[owner]) + (int64_t)b * kBlkRows * dist.lda;
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(dist.a_full) +
                             ((int64_t)owner * p.rows_per_rank + (int64_t)b * kBlkRows) * dist.ld_full;
        uint4 r[kPullUnroll];
        int64_t off[kPullUnroll];
#pragma unroll
        for (int u = 0; u < kPullUnroll; ++u) {
          const int64_t v = in_blk * kPieceVecs + u * 32 + lane;
          off[u] = -1;
          if (v < vecs) {
            const int64_t row = v / vec_per_row;
            const int64_t col = (v - row * vec_per_row) << 3;
            off[u] = row * dist.ld_full + col;
            r[u] = ld_relaxed_sys_16(src + row * dist.lda + col);
          }
        }
#pragma unroll
        for (int u = 0; u < kPullUnroll; ++u)
          if (off[u] >= 0) *reinterpret_cast<uint4*>(dst + off[u]) = r[u];
        if (piece < remote_pieces) {
          __syncwarp();
          if (lane == 0) {
            __threadfence();
            red_release_gpu_add_s32(dist.flags + owner * bpr + b, 1);
          }
        }
      }
    }
  } else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + 4) {
    // ===== epilogue: TMEM -> registers -> (+R) -> swizzled smem -> TMA store =====
    const int ew = warp - kEpiWarp0;       // == warp % 4 -> TMEM lanes [32*ew, 32*ew+32)
    const int row = ew * 32 + lane;        // row of the 128-row CTA tile owned by this thread
    const bool store_leader = (warp == kEpiWarp0 && lane == 0);
    const uint32_t staging0 = smem_base + C::kRingBytes;
    uint8_t* staging_gen = smem_gen + C::kRingBytes;
    const uint32_t tmem_empty_target = (kCtaGroup == 2) ? mapa(tmem_empty_bar(0), 0) : tmem_empty_bar(0);
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t res_phase_bits = 0;
    int chunk_counter = 0;
    for (int tile = worker; tile < num_tiles; tile += num_workers) {
      int tm, tn;
      tile_mn(tile, tm, tn);
      const int m0 = (tm * kCtaGroup + (int)cta_rank) * BM;
      const int n0 = tn * BN;
      mbar_wait_cluster(tmem_full_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int ch = 0; ch < kNumChunks; ++ch, ++chunk_counter) {
        const int buf = chunk_counter & 1;
        const uint32_t sbuf = staging0 + buf * kStagingBytes;
        uint8_t* gbuf = staging_gen + buf * kStagingBytes;
        const int c0 = n0 + ch * kOutChunkCols;
        // (1) the TMA store that last used this staging buffer must have finished reading it
        if (store_leader) tma_store_wait_read<kNumStaging - 1>();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // (2) optional residual / accumulate source: TMA load of the R chunk into the staging buffer
        if (p.has_residual) {
          if (store_leader) {
            mbar_arrive_expect_tx(res_bar(buf), kStagingBytes);
            tma_load_2d(sbuf, &tmR, res_bar(buf), c0, m0);
          }
        }
        // (3) accumulator chunk -> registers (kept as raw fp32 bits)
        uint32_t v[32 * (kOutChunkCols / 32)];
#pragma unroll
        for (int q = 0; q < kOutChunkCols / 32; ++q) {
          uint32_t (&vq)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[q * 32]);
          tmem_ld_32x32b_x32(t_row + (uint32_t)(ch * kOutChunkCols + q * 32), vq);
        }
        tmem_ld_wait();
        if (ch == kNumChunks - 1) {
          // whole accumulator is in registers: hand the TMEM stage back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (kCtaGroup == 2) mbar_arrive_cluster(tmem_empty_target + 8u * acc);
            else mbar_arrive(tmem_empty_target + 8u * acc);
          }
        }
        if (p.has_residual) {
          mbar_wait(res_bar(buf), (res_phase_bits >> buf) & 1u);
          res_phase_bits ^= 1u << buf;
        }
        // (4) convert (+ residual) and write this thread's 128-byte row in the SWIZZLE_128B pattern
        uint8_t* grow = gbuf + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {  // 16-byte pieces of the row
          uint8_t* dst = grow + ((j ^ (row & 7)) << 4);
          if constexpr (sizeof(OutT) == 2) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j * 8 + e]);
            if (p.has_residual) {
              float r[8];
              unpack8(*reinterpret_cast<const bf16x8*>(dst), r);
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] += r[e];
            }
            *reinterpret_cast<bf16x8*>(dst) = pack8(f);
          } else {
            float4 o = make_float4(__uint_as_float(v[j * 4 + 0]), __uint_as_float(v[j * 4 + 1]),
                                   __uint_as_float(v[j * 4 + 2]), __uint_as_float(v[j * 4 + 3]));
            if (p.has_residual) {
              const float4 r = *reinterpret_cast<const float4*>(dst);
              o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w;
            }
            *reinterpret_cast<float4*>(dst) = o;
          }
        }
        // (5) publish to the async proxy, then one thread stores the 128 x chunk box
        fence_proxy_async_smem();
        asm volatile("bar.sync 2, 128;" ::: "memory");
        if (store_leader) {
          if constexpr (kDistC) {
            const int owner = min(m0 / p.rows_per_rank, kMaxDistRanks - 1);
            tma_reduce_add_2d(&dist.c[owner], sbuf, c0, m0 - owner * p.rows_per_rank);
          } else {
            tma_store_2d(&tmC, sbuf, c0, m0);
          }
          tma_store_commit();
        }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
    if (store_leader) tma_store_wait<0>();
  }

  // ---------------------------------------------------------------- teardown
  tc_fence_before();
  if constexpr (kCtaGroup == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<kCtaGroup>(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// 2-D row-major tensor [rows, cols] with leading dimension ld (elements); box = {box_cols, box_rows}.
int make_tmap_2d(CUtensorMap* out, const void* ptr, int elem_bytes, int64_t rows, int64_t cols, int64_t ld,
                 int box_cols, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return 3;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * elem_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): ptr=%p rows=%lld cols=%lld ld=%lld box=%dx%d", (int)r, ptr,
              (long long)rows, (long long)cols, (long long)ld, box_cols, box_rows);
    return 3;
  }
  return 0;
}

template <int kCtaGroup, bool kAMn, bool kBMn, typename OutT, int kDistA = 0, bool kDistC = false>
int launch(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c, const CUtensorMap& r, Params p,
           int max_ctas, cudaStream_t stream, const DistMaps* dist_maps = nullptr) {
  using C = Cfg<kCtaGroup>;
  auto kern = gemm_bf16_kernel<kCtaGroup, kAMn, kBMn, OutT, kDistA, kDistC>;
  // function attributes are per device: one flag per (instantiation, device)
  static bool configured[kMaxDevices] = {};
  static int resident_clusters[kMaxDevices] = {};
  if (int rc = ensure_smem_attr(kern, C::kSmemBytes, configured, "gemm")) return rc;
  int dev = 0;
  cudaGetDevice(&dev);
  const int num_m = (p.M + BM * kCtaGroup - 1) / (BM * kCtaGroup);
  const int num_n = (p.N + BN - 1) / BN;
  int workers = num_m * num_n;
  const int cap = max_ctas / kCtaGroup;
  if (workers > cap) workers = cap;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(workers * kCtaGroup);
  cfg.blockDim = dim3((kDistA >= 2 ? 2 : 1) * kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtaGroup;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e;
  if constexpr (kDistA >= 2) {
    // The pull warps of ALL CTAs feed the arrival counters every TMA producer waits on, so the whole grid has to be
    // resident at once.  Ask the driver how many clusters of this exact shape (threads, dynamic smem, cluster size)
    // can be co-scheduled on this device / partition and never launch more than that.
    if (resident_clusters[dev] == 0) {
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
      if (e != cudaSuccess || n < 1) {
        set_error("gemm: the gathering GEMM cannot be made resident on this device (%s, %d clusters)",
                  e != cudaSuccess ? cudaGetErrorString(e) : "ok", n);
        cudaGetLastError();
        return 2;
      }
      resident_clusters[dev] = n;
    }
    if (workers > resident_clusters[dev]) {
      workers = resident_clusters[dev];
      cfg.gridDim = dim3(workers * kCtaGroup);
    }
  }
  if constexpr (kDistA != 0 || kDistC) e = cudaLaunchKernelEx(&cfg, kern, a, b, c, r, p, *dist_maps);
  else e = cudaLaunchKernelEx(&cfg, kern, a, b, c, r, p, NoDist{});
  if (e != cudaSuccess) {
    set_error("gemm launch: %s", cudaGetErrorString(e));
    return 2;
  }
  count_launch();
  return 0;
}

}  // namespace gemm
}  // namespace llmt

using namespace llmt;
using namespace llmt::gemm;

// C[M,N] = op(A) * op(B) (+ R).  All matrices row-major with leading dimensions in elements.
//   a_mn_major = 0: A stored [M, K] (lda >= K)      1: A stored [K, M] (lda >= M)
//   b_mn_major = 0: B stored [N, K] (ldb >= K)      1: B stored [K, N] (ldb >= N)
//   out_fp32: C (and R) are fp32, else bf16.  R may be null; R == C accumulates in place.
//   cta_group: 1 or 2 (CTA pairs, 256x256 tiles).  max_ctas <= 0 -> all SMs.
extern "C" int llmt_gemm_bf16(const void* A, const void* B, void* Cout, const void* R, int M, int N, int K,
                              int64_t lda, int64_t ldb, int64_t ldc, int64_t ldr, int a_mn_major, int b_mn_major,
                              int out_fp32, int cta_group, int max_ctas, cudaStream_t stream) {
  LLMT_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
  LLMT_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements (16-byte rows)");
  LLMT_REQUIRE((ldc * (out_fp32 ? 4 : 2)) % 16 == 0, "gemm: ldc must give 16-byte aligned rows");
  LLMT_REQUIRE(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && ((uintptr_t)Cout % 16 == 0) &&
                   ((uintptr_t)R % 16 == 0),
               "gemm: operands must be 16-byte aligned");
  LLMT_REQUIRE(cta_group == 1 || cta_group == 2, "gemm: cta_group must be 1 or 2");
  if (max_ctas <= 0) max_ctas = num_sms();
  if (max_ctas < cta_group) max_ctas = cta_group;

  CUtensorMap ta, tb, tc, tr;
  int rc;
  if (!a_mn_major) rc = make_tmap_2d(&ta, A, 2, M, K, lda, BK, BM);
  else             rc = make_tmap_2d(&ta, A, 2, K, M, lda, 64, BK);
  if (rc) return rc;
  const int b_rows = BN / cta_group;
  if (!b_mn_major) rc = make_tmap_2d(&tb, B, 2, N, K, ldb, BK, b_rows);
  else             rc = make_tmap_2d(&tb, B, 2, K, N, ldb, 64, BK);
  if (rc) return rc;
  const int esz = out_fp32 ? 4 : 2;
  rc = make_tmap_2d(&tc, Cout, esz, M, N, ldc, 128 / esz, BM);
  if (rc) return rc;
  if (R) {
    rc = make_tmap_2d(&tr, R, esz, M, N, ldr, 128 / esz, BM);
    if (rc) return rc;
  } else {
    tr = tc;
  }
  Params p{M, N, K, R ? 1 : 0, 0, 0};

#define LLMT_GEMM_DISPATCH(CG, AMN, BMN)                                                        \
  return out_fp32 ? launch<CG, AMN, BMN, float>(ta, tb, tc, tr, p, max_ctas, stream)            \
                  : launch<CG, AMN, BMN, __nv_bfloat16>(ta, tb, tc, tr, p, max_ctas, stream)
  if (cta_group == 1) {
    if (!a_mn_major && !b_mn_major) { LLMT_GEMM_DISPATCH(1, false, false); }
    if (!a_mn_major && b_mn_major) { LLMT_GEMM_DISPATCH(1, false, true); }
    if (a_mn_major && b_mn_major) { LLMT_GEMM_DISPATCH(1, true, true); }
    LLMT_GEMM_DISPATCH(1, true, false);
  } else {
    if (!a_mn_major && !b_mn_major) { LLMT_GEMM_DISPATCH(2, false, false); }
    if (!a_mn_major && b_mn_major) { LLMT_GEMM_DISPATCH(2, false, true); }
    if (a_mn_major && b_mn_major) { LLMT_GEMM_DISPATCH(2, true, true); }
    LLMT_GEMM_DISPATCH(2, true, false);
  }
#undef LLMT_GEMM_DISPATCH
}


// Fused collective GEMMs for tensor / sequence parallelism.  Both operands K-major, bf16 output.
//   mode 1 (all-gather -> GEMM):  C[M, N] = concat_r(A_r)[M, K] * B[N, K]^T, A_r = a_ptrs[r] is rank r's
//          [rows_per_rank, K] block in symmetric memory (lda elements per row); C is a local buffer.
//   mode 2 (GEMM -> reduce-scatter): A local [M, K]; the [M, N] product is added, rows r*rows_per_rank.., into
//          c_ptrs[r] ([rows_per_rank, N], ldc elements per row) on rank r.  The targets must be initialised
//          (zeros or the residual) and a cross-rank barrier must separate initialisation, this kernel and the
//          first read of the result.
extern "C" int llmt_gemm_bf16_dist(int mode, const void* A, const void* const* a_ptrs, const void* B, void* Cout,
                                   void* const* c_ptrs, int world, int rows_per_rank, int M, int N, int K,
                                   int64_t lda, int64_t ldb, int64_t ldc, int b_mn_major, int cta_group,
                                   int max_ctas, cudaStream_t stream) {
  LLMT_REQUIRE(mode == 1 || mode == 2, "gemm_dist: mode must be 1 (AG->GEMM) or 2 (GEMM->RS)");
  LLMT_REQUIRE(world >= 1 && world <= kMaxDistRanks, "gemm_dist: world size %d not in [1, %d]", world, kMaxDistRanks);
  LLMT_REQUIRE(rows_per_rank > 0 && rows_per_rank % BM == 0 && (int64_t)rows_per_rank * world == M,
               "gemm_dist: rows_per_rank=%d must be a multiple of %d and rows_per_rank * world == M=%d",
               rows_per_rank, BM, M);
  LLMT_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "gemm_dist: leading dimensions must be multiples of 8");
  LLMT_REQUIRE(cta_group == 1 || cta_group == 2, "gemm_dist: cta_group must be 1 or 2");
  if (max_ctas <= 0) max_ctas = num_sms();
  static thread_local DistMaps maps;   // filled per call; copied into kernel parameters by the launch
  CUtensorMap ta, tb, tc;
  int rc;
  const int b_rows = BN / cta_group;
  if (!b_mn_major) rc = make_tmap_2d(&tb, B, 2, N, K, ldb, BK, b_rows);   // B stored [N, K]
  else             rc = make_tmap_2d(&tb, B, 2, K, N, ldb, 64, BK);       // B stored [K, N]
  if (rc) return rc;
  Params p{M, N, K, 0, rows_per_rank, 0};
  if (mode == 1) {
    for (int r = 0; r < kMaxDistRanks; ++r) {
      const void* src = a_ptrs[r < world ? r : world - 1];
      rc = make_tmap_2d(&maps.a[r], src, 2, rows_per_rank, K, lda, BK, BM);
      if (rc) return rc;
      maps.c[r] = maps.a[r];
    }
    rc = make_tmap_2d(&tc, Cout, 2, M, N, ldc, 64, BM);
    if (rc) return rc;
    ta = maps.a[0];
    if (b_mn_major) {
      if (cta_group == 1) return launch<1, false, true, __nv_bfloat16, 1, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
      return launch<2, false, true, __nv_bfloat16, 1, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
    }
    if (cta_group == 1) return launch<1, false, false, __nv_bfloat16, 1, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
    return launch<2, false, false, __nv_bfloat16, 1, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
  }
  rc = make_tmap_2d(&ta, A, 2, M, K, lda, BK, BM);
  if (rc) return rc;
  for (int r = 0; r < kMaxDistRanks; ++r) {
    void* dst = c_ptrs[r < world ? r : world - 1];
    rc = make_tmap_2d(&maps.c[r], dst, 2, rows_per_rank, N, ldc, 64, BM);
    if (rc) return rc;
    maps.a[r] = ta;
  }
  tc = maps.c[0];
  if (b_mn_major) {
    if (cta_group == 1) return launch<1, false, true, __nv_bfloat16, 0, true>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
    return launch<2, false, true, __nv_bfloat16, 0, true>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
  }
  if (cta_group == 1) return launch<1, false, false, __nv_bfloat16, 0, true>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
  return launch<2, false, false, __nv_bfloat16, 0, true>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
}

// All-gather -> GEMM, "land once":  C[M, N] = concat_r(A_r)[M, K] * op(B), and a_full[M, K] = concat_r(A_r).
// a_ptrs[r] is rank r's [rows_per_rank, K] block in symmetric memory (HOST array of `world` device pointers, lda
// elements per row).  Extra warps of the GEMM kernel copy the remote blocks over NVLink into a_full while the
// tensor cores work on the blocks that are already local; the TMA producer waits on per-block arrival counters
// (`flags`, int32 [M / (128 * cta_group)], zeroed here on `stream`).  A cross-rank barrier must precede the call
// (every A_r written) and follow it before A_r is overwritten.
extern "C" int llmt_gemm_bf16_ag(const void* const* a_ptrs, void* a_full, const void* B, void* Cout, int* flags,
                                 int rank, int world, int rows_per_rank, int M, int N, int K, int64_t lda,
                                 int64_t ld_full, int64_t ldb, int64_t ldc, int b_mn_major, int cta_group,
                                 int max_ctas, cudaStream_t stream) {
  LLMT_REQUIRE(world >= 1 && world <= kMaxDistRanks && rank >= 0 && rank < world,
               "gemm_ag: rank %d / world %d not supported (max %d ranks)", rank, world, kMaxDistRanks);
  LLMT_REQUIRE(cta_group == 1 || cta_group == 2, "gemm_ag: cta_group must be 1 or 2");
  const int blk = BM * cta_group;
  LLMT_REQUIRE(rows_per_rank > 0 && rows_per_rank % blk == 0 && (int64_t)rows_per_rank * world == M,
               "gemm_ag: rows_per_rank=%d must be a multiple of %d and rows_per_rank * world == M=%d", rows_per_rank,
               blk, M);
  LLMT_REQUIRE(K % 8 == 0 && lda % 8 == 0 && ld_full % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0,
               "gemm_ag: K and the leading dimensions must be multiples of 8");
  LLMT_REQUIRE((uintptr_t)a_full % 16 == 0 && (uintptr_t)flags % 4 == 0, "gemm_ag: misaligned buffers");
  if (max_ctas <= 0) max_ctas = num_sms();
  if (max_ctas > num_sms()) max_ctas = num_sms();   // every CTA must be resident: the producers wait on all pullers
  static thread_local DistMaps maps;
  CUtensorMap ta, tb, tc;
  int rc;
  const int b_rows = BN / cta_group;
  if (!b_mn_major) rc = make_tmap_2d(&tb, B, 2, N, K, ldb, BK, b_rows);
  else             rc = make_tmap_2d(&tb, B, 2, K, N, ldb, 64, BK);
  if (rc) return rc;
  for (int r = 0; r < kMaxDistRanks; ++r) {
    const void* src = a_ptrs[r < world ? r : world - 1];
    LLMT_REQUIRE((uintptr_t)src % 16 == 0, "gemm_ag: misaligned peer buffer");
    rc = make_tmap_2d(&maps.a[r], src, 2, rows_per_rank, K, lda, BK, BM);
    if (rc) return rc;
    maps.c[r] = maps.a[r];
    maps.a_raw[r] = src;
  }
  rc = make_tmap_2d(&ta, a_full, 2, M, K, ld_full, BK, BM);
  if (rc) return rc;
  rc = make_tmap_2d(&tc, Cout, 2, M, N, ldc, 64, BM);
  if (rc) return rc;
  maps.a_full = a_full;
  maps.flags = flags;
  maps.lda = lda;
  maps.ld_full = ld_full;
  maps.rank = rank;
  maps.world = world;
  cudaError_t e = cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)(M / blk), stream);
  LLMT_REQUIRE(e == cudaSuccess, "gemm_ag: flag reset failed: %s", cudaGetErrorString(e));
  Params p{M, N, K, 0, rows_per_rank, rank * (rows_per_rank / blk)};
  if (b_mn_major) {
    if (cta_group == 1) return launch<1, false, true, __nv_bfloat16, 2, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
    return launch<2, false, true, __nv_bfloat16, 2, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
  }
  if (cta_group == 1) return launch<1, false, false, __nv_bfloat16, 2, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
  return launch<2, false, false, __nv_bfloat16, 2, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
}

// ZeRO-3 parameter all-gather -> first consumer GEMM:  C[M, N] = A[M, K] * W[N, K]^T where W is the element range
// [w_off, w_off + N*K) of a sharding unit's flat parameter buffer.  shard_ptrs[r] (HOST array) is rank r's slice of
// that unit in the symmetric shard arena (shard_numel elements each); full_base is the unit's local full buffer,
// whose W range this kernel fills while it computes.  flags: int32 [ceil(N / 256)], reset here on `stream`.
// A cross-rank barrier must separate the last write of any slice from this call.
extern "C" int llmt_gemm_bf16_wgather(const void* A, const void* const* shard_ptrs, void* full_base, int64_t w_off,
                                      int64_t shard_numel, void* Cout, int* flags, int rank, int world, int M, int N,
                                      int K, int64_t lda, int64_t ldc, int cta_group, int max_ctas,
                                      cudaStream_t stream) {
  LLMT_REQUIRE(world >= 1 && world <= kMaxDistRanks && rank >= 0 && rank < world,
               "gemm_wgather: rank %d / world %d not supported (max %d ranks)", rank, world, kMaxDistRanks);
  LLMT_REQUIRE(cta_group == 1 || cta_group == 2, "gemm_wgather: cta_group must be 1 or 2");
  LLMT_REQUIRE(K % 8 == 0 && lda % 8 == 0 && ldc % 8 == 0 && w_off % 8 == 0 && shard_numel % 8 == 0 && shard_numel > 0,
               "gemm_wgather: K, leading dimensions, w_off and shard_numel must be multiples of 8");
  LLMT_REQUIRE(w_off >= 0 && w_off + (int64_t)N * K <= shard_numel * world, "gemm_wgather: W exceeds the unit");
  LLMT_REQUIRE((uintptr_t)full_base % 16 == 0 && (uintptr_t)A % 16 == 0 && (uintptr_t)Cout % 16 == 0,
               "gemm_wgather: misaligned buffers");
  if (max_ctas <= 0) max_ctas = num_sms();
  if (max_ctas > num_sms()) max_ctas = num_sms();
  static thread_local DistMaps maps;
  CUtensorMap ta, tb, tc;
  int rc;
  const int b_rows = BN / cta_group;
  void* w_local = reinterpret_cast<__nv_bfloat16*>(full_base) + w_off;
  rc = make_tmap_2d(&ta, A, 2, M, K, lda, BK, BM);
  if (rc) return rc;
  rc = make_tmap_2d(&tb, w_local, 2, N, K, K, BK, b_rows);
  if (rc) return rc;
  rc = make_tmap_2d(&tc, Cout, 2, M, N, ldc, 64, BM);
  if (rc) return rc;
  for (int r = 0; r < kMaxDistRanks; ++r) {
    maps.a[r] = ta;
    maps.c[r] = tc;
    maps.a_raw[r] = shard_ptrs[r < world ? r : world - 1];
    LLMT_REQUIRE((uintptr_t)maps.a_raw[r] % 16 == 0, "gemm_wgather: misaligned slice");
  }
  maps.a_full = full_base;
  maps.flags = flags;
  maps.lda = lda;
  maps.ld_full = K;
  maps.rank = rank;
  maps.world = world;
  maps.w_off = w_off;
  maps.shard_numel = shard_numel;
  cudaError_t e = cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)((N + BN - 1) / BN), stream);
  LLMT_REQUIRE(e == cudaSuccess, "gemm_wgather: flag reset failed: %s", cudaGetErrorString(e));
  Params p{M, N, K, 0, 0, 0};
  if (cta_group == 1) return launch<1, false, false, __nv_bfloat16, 3, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
  return launch<2, false, false, __nv_bfloat16, 3, false>(ta, tb, tc, tc, p, max_ctas, stream, &maps);
}
