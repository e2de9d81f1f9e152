// Collectives over NVLink 5 / NVSwitch peer memory, written for the flat-parameter sharding engine.
//
// Every rank owns buffers in a symmetric heap (same offsets on every GPU, mapped into all peers with
// torch.distributed._symmetric_memory); kernels below read / write the peers' copies directly with
// ld/st.global on the mapped addresses — all 7 peers are one NVSwitch hop away at full bandwidth, so the
// patterns are direct pull / push, not rings.
//
//   llmt_comm_barrier           flag barrier through the signal pads (st.release.sys / ld.acquire.sys)
//   llmt_comm_allgather_pull    full[r*S .. (r+1)*S) <- peer r's shard, for all r       (replaces C1)
//   llmt_comm_rs_sumsq          shard = mean_r peer_r.full[rank*S ..] (+ partial sum of squares) (C2 + C9)
//   llmt_comm_adamw_push        AdamW on the local fp32 slice, bf16 result STORED INTO EVERY PEER's flat
//                               parameter buffer (optimizer step + parameter all-gather in one kernel; C2/C17)
//   llmt_comm_allreduce_f32     one-shot all-reduce of a few floats (grad-norm; C7-C9)
//
// Replaces NCCL all-gather / reduce-scatter issued by FSDP2 hooks (reference:
// src/llm_training/models/llama/llama_model.py:263-268) and by DeepSpeed ZeRO-2
// (src/llm_training/lightning/strategy/deepspeed/deepspeed_strategy.py:39-44).
#include "common.cuh"

namespace llmt {
namespace comm {

constexpr int kMaxRanks = 16;

struct Peers {
  void* ptr[kMaxRanks];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_16(const void* p) {   // peer data: read once, keep out of L1
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
// volatile variant for buffers another GPU may have just written
__device__ __forceinline__ uint4 ld_peer_16_fresh(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w) : "memory");
}

// ---- barrier: rank r bumps slot[r] in every peer's pad to `epoch`, then waits for all slots of its own pad.
// Slots are uint32 at pad + slot_base; epoch must increase by 1 per use of the same slots.
__global__ void barrier_kernel(Peers pads, int rank, int world, int slot_base, uint32_t epoch) {
  const int t = threadIdx.x;
  __threadfence_system();
  if (t < world) {
    uint32_t* remote = reinterpret_cast<uint32_t*>(pads.ptr[t]) + slot_base + rank;
    st_release_sys(remote, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot_base + t;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (++spins > (1ull << 31)) __trap();   // a peer died: fail loudly instead of hanging the box
    }
  }
  __syncthreads();
}

// ---- all-gather by pulling every peer's slice into the local full buffer (16-byte units)
__global__ void __launch_bounds__(512)
allgather_pull_kernel(Peers full, int rank, int world, int64_t shard_vec16) {
  uint4* dst = reinterpret_cast<uint4*>(full.ptr[rank]);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int p = 1; p < world; ++p) {
    const int src_rank = (rank + p) % world;   // stagger: at any instant the ranks read from different peers
    const uint4* src = reinterpret_cast<const uint4*>(full.ptr[src_rank]) + (int64_t)src_rank * shard_vec16;
    uint4* out = dst + (int64_t)src_rank * shard_vec16;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < shard_vec16; i += stride)
      out[i] = ld_peer_16_fresh(src + i);
  }
}

// ---- ZeRO-3 parameter gather: dst[r * shard .. (r+1) * shard) = rank r's persistent slice, every rank's slice
// (this rank's included) read from the symmetric shard arena.  Same stagger as above.
__global__ void __launch_bounds__(512)
gather_shards_kernel(Peers shards, uint4* __restrict__ dst, int rank, int world, int64_t shard_vec16) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int p = 0; p < world; ++p) {
    const int src_rank = (rank + 1 + p) % world;       // own slice last
    const uint4* src = reinterpret_cast<const uint4*>(shards.ptr[src_rank]);
    uint4* out = dst + (int64_t)src_rank * shard_vec16;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < shard_vec16; i += stride)
      out[i] = ld_peer_16_fresh(src + i);
  }
}

// Same for the 16-byte vectors [lo, hi) of the unit's flat buffer EXCEPT [skip_lo, skip_hi): the skipped range is
// gathered by its first consumer (the weight-gathering GEMM, csrc/kernels/gemm.cu kDistA == 3).
__global__ void __launch_bounds__(512)
gather_range_kernel(Peers shards, uint4* __restrict__ dst, int64_t shard_vec16, int64_t lo, int64_t hi,
                    int64_t skip_lo, int64_t skip_hi) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < hi; v += stride) {
    if (v >= skip_lo && v < skip_hi) continue;
    const int64_t owner = v / shard_vec16;
    dst[v] = ld_peer_16_fresh(reinterpret_cast<const uint4*>(shards.ptr[owner]) + (v - owner * shard_vec16));
  }
}

// ---- reduce-scatter (mean) by pulling this rank's slice from every peer + sum of squares of the result
template <typename T>
__global__ void __launch_bounds__(512)
rs_sumsq_kernel(Peers grads, int rank, int world, int64_t shard_elems, T* __restrict__ out_shard, float inv_world,
                float* __restrict__ sumsq_partial) {
  __shared__ float red[32];
  constexpr int kPer = 16 / (int)sizeof(T);
  const int64_t nvec = shard_elems / kPer;
  const int64_t base = (int64_t)rank * nvec;
  float ss = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    // all peers' vectors are requested before any is consumed: up to 8 independent NVLink loads in flight per
    // thread (a dependent load per peer left the links latency-bound); summation order stays rank, rank+1, ...
#pragma unroll 1
    for (int p0 = 0; p0 < world; p0 += 8) {
      uint4 raw[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (p0 + q < world) {
          const int src_rank = (rank + p0 + q) % world;
          raw[q] = ld_peer_16_fresh(reinterpret_cast<const uint4*>(grads.ptr[src_rank]) + base + i);
        }
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (p0 + q < world) {
          if constexpr (sizeof(T) == 2) {
            float f[8];
            unpack8(*reinterpret_cast<const bf16x8*>(&raw[q]), f);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += f[e];
          } else {
            acc[0] += __uint_as_float(raw[q].x); acc[1] += __uint_as_float(raw[q].y);
            acc[2] += __uint_as_float(raw[q].z); acc[3] += __uint_as_float(raw[q].w);
          }
        }
      }
    }
    if constexpr (sizeof(T) == 2) {
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] *= inv_world;
      const bf16x8 packed = pack8(acc);
      float back[8];
      unpack8(packed, back);   // norm of what the optimizer will actually consume
#pragma unroll
      for (int e = 0; e < 8; ++e) ss += back[e] * back[e];
      reinterpret_cast<uint4*>(out_shard)[i] = *reinterpret_cast<const uint4*>(&packed);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) { acc[e] *= inv_world; ss += acc[e] * acc[e]; }
      reinterpret_cast<float4*>(out_shard)[i] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    }
  }
  ss = block_sum(ss, red);
  if (threadIdx.x == 0) sumsq_partial[blockIdx.x] = ss;
}

// ---- reduce-scatter (mean) INSIDE THE SWITCH: one multimem.ld_reduce per 16 bytes of this rank's slice returns
// the sum over every rank's copy (fp32 accumulation in the NVSwitch, result rounded to bf16), so each GPU receives
// shard bytes instead of world x shard bytes.  `mc` is the multicast address of the gradient buffer (same offset on
// every rank).  bf16 only.  The sum is rounded to bf16 before the 1/world scaling (exact for power-of-two worlds).
__global__ void __launch_bounds__(512)
rs_sumsq_mc_kernel(const uint4* __restrict__ mc, int rank, int64_t shard_elems, __nv_bfloat16* __restrict__ out_shard,
                   float inv_world, float* __restrict__ sumsq_partial) {
  __shared__ float red[32];
  const int64_t nvec = shard_elems / 8;
  const int64_t base = (int64_t)rank * nvec;
  float ss = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    uint4 raw;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w) : "l"(mc + base + i) : "memory");
    float acc[8];
    unpack8(*reinterpret_cast<const bf16x8*>(&raw), acc);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] *= inv_world;
    const bf16x8 packed = pack8(acc);
    float back[8];
    unpack8(packed, back);
#pragma unroll
    for (int e = 0; e < 8; ++e) ss += back[e] * back[e];
    reinterpret_cast<uint4*>(out_shard)[i] = *reinterpret_cast<const uint4*>(&packed);
  }
  ss = block_sum(ss, red);
  if (threadIdx.x == 0) sumsq_partial[blockIdx.x] = ss;
}

__global__ void __launch_bounds__(1024)
sumsq_accumulate_kernel(const float* __restrict__ partial, int n, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + s : s;
}

// ---- one-shot all-reduce (sum) of n <= 32 floats through a symmetric fp32 buffer [2][world][32] + barrier slots.
// The buffer half is chosen by epoch parity: a rank that has left all-reduce n may already be writing its row for
// n + 1 while a slower peer still reads the rows of n; it cannot reach n + 2 (same half as n) before that peer has
// signalled n + 1, which the peer only does after it finished reading n.
__global__ void allreduce_oneshot_kernel(Peers bufs, Peers pads, int rank, int world, int slot_base, uint32_t epoch,
                                         float* __restrict__ value, int n) {
  const int t = threadIdx.x;
  const int half = (epoch & 1u) * world * 32;
  // write my contribution into row `rank` of every peer's buffer
  if (t < world * n) {
    const int peer = t / n, j = t % n;
    float* remote = reinterpret_cast<float*>(bufs.ptr[peer]) + half + rank * 32 + j;
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(remote), "f"(value[j]) : "memory");
  }
  __threadfence_system();
  __syncthreads();
  if (t < world) {
    st_release_sys(reinterpret_cast<uint32_t*>(pads.ptr[t]) + slot_base + rank, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot_base + t;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (++spins > (1ull << 31)) __trap();
    }
  }
  __syncthreads();
  if (t < n) {
    const float* mine = reinterpret_cast<const float*>(bufs.ptr[rank]) + half;
    float s = 0.f;
    for (int p = 0; p < world; ++p) {
      float v;
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(mine + p * 32 + t) : "memory");
      s += v;
    }
    value[t] = s;
  }
}

// ---- AdamW on the local slice + push of the bf16 result into every rank's flat parameter buffer
struct AdamWArgs {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2, grad_scale, max_norm;
};

template <typename GradT>
__global__ void __launch_bounds__(256)
adamw_push_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                  const GradT* __restrict__ grad, Peers params, uint4* __restrict__ mc_params, int rank, int world,
                  int64_t shard_elems, AdamWArgs a, const float* __restrict__ norm_sq) {
  float gscale = a.grad_scale;
  if (norm_sq != nullptr && a.max_norm > 0.f) {
    const float norm = sqrtf(*norm_sq) * fabsf(a.grad_scale);
    gscale *= fminf(1.f, a.max_norm / (norm + 1e-6f));
  }
  const float decay = 1.f - a.lr * a.weight_decay;
  const float step_size = a.lr / a.bc1;
  const float inv_sqrt_bc2 = rsqrtf(a.bc2);
  const int64_t nvec = shard_elems >> 3;   // 8 elements per thread-iteration -> one 16-byte bf16 store per peer
  const int64_t base = (int64_t)rank * nvec;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float p[8], mm[8], vv[8], g[8];
    *reinterpret_cast<float4*>(&p[0]) = reinterpret_cast<float4*>(master)[2 * i];
    *reinterpret_cast<float4*>(&p[4]) = reinterpret_cast<float4*>(master)[2 * i + 1];
    *reinterpret_cast<float4*>(&mm[0]) = reinterpret_cast<float4*>(m)[2 * i];
    *reinterpret_cast<float4*>(&mm[4]) = reinterpret_cast<float4*>(m)[2 * i + 1];
    *reinterpret_cast<float4*>(&vv[0]) = reinterpret_cast<float4*>(v)[2 * i];
    *reinterpret_cast<float4*>(&vv[4]) = reinterpret_cast<float4*>(v)[2 * i + 1];
    if constexpr (sizeof(GradT) == 2) {
      unpack8(ld_bf16x8(reinterpret_cast<const __nv_bfloat16*>(grad) + i * 8), g);
    } else {
      *reinterpret_cast<float4*>(&g[0]) = reinterpret_cast<const float4*>(grad)[2 * i];
      *reinterpret_cast<float4*>(&g[4]) = reinterpret_cast<const float4*>(grad)[2 * i + 1];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = g[j] * gscale;
      mm[j] = a.beta1 * mm[j] + (1.f - a.beta1) * gj;
      vv[j] = a.beta2 * vv[j] + (1.f - a.beta2) * gj * gj;
      p[j] = p[j] * decay - step_size * (mm[j] / (sqrtf(vv[j]) * inv_sqrt_bc2 + a.eps));
    }
    reinterpret_cast<float4*>(master)[2 * i] = *reinterpret_cast<float4*>(&p[0]);
    reinterpret_cast<float4*>(master)[2 * i + 1] = *reinterpret_cast<float4*>(&p[4]);
    reinterpret_cast<float4*>(m)[2 * i] = *reinterpret_cast<float4*>(&mm[0]);
    reinterpret_cast<float4*>(m)[2 * i + 1] = *reinterpret_cast<float4*>(&mm[4]);
    reinterpret_cast<float4*>(v)[2 * i] = *reinterpret_cast<float4*>(&vv[0]);
    reinterpret_cast<float4*>(v)[2 * i + 1] = *reinterpret_cast<float4*>(&vv[4]);
    const bf16x8 packed = pack8(p);
    const uint4 raw = *reinterpret_cast<const uint4*>(&packed);
    if (mc_params != nullptr) {
      // NVSwitch multicast: ONE 16-byte store leaves this GPU, the switch replicates it into every rank's arena
      // (this rank's included) -> outbound traffic is shard bytes instead of (world - 1) x shard bytes
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_params + base + i),
                   "f"(__uint_as_float(raw.x)), "f"(__uint_as_float(raw.y)), "f"(__uint_as_float(raw.z)),
                   "f"(__uint_as_float(raw.w)) : "memory");
    } else {
#pragma unroll 1
      for (int q = 0; q < world; ++q) {
        const int dst_rank = (rank + q) % world;
        st_peer_16(reinterpret_cast<uint4*>(params.ptr[dst_rank]) + base + i, raw);
      }
    }
  }
}

static int grid_for(int64_t items, int threads, int per_sm) {
  int64_t blocks = (items + threads - 1) / threads;
  const int64_t cap = (int64_t)num_sms() * per_sm;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

static int fill_peers(Peers* out, const void* const* ptrs, int world) {
  if (world < 1 || world > kMaxRanks) {
    set_error("comm: world size %d not in [1, %d]", world, kMaxRanks);
    return 1;
  }
  for (int i = 0; i < kMaxRanks; ++i) out->ptr[i] = i < world ? const_cast<void*>(ptrs[i]) : nullptr;
  return 0;
}

// ---- vocabulary-parallel loss statistics: every rank holds {lse_local[rows], tgt_local[rows]} (fp32, symmetric
// memory) of its vocabulary shard; the global statistics are lse = logsumexp_r(lse_local_r), tgt = sum_r tgt_local_r
// (the label's logit lives on exactly one rank, the others hold 0).  One pass, 4 rows per thread, every peer's
// values loaded before any arithmetic.  Replaces all_gather + logsumexp + all_reduce (reference: loss_parallel(),
// src/llm_training/lms/clm/clm.py:114-120).
__global__ void __launch_bounds__(256)
vocab_loss_combine_kernel(Peers stats, int world, int rows, const int64_t* __restrict__ labels, int64_t ignore_index,
                          float* __restrict__ lse_out, float* __restrict__ nll_out) {
  const int nvec = rows >> 2;
  for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
    uint4 l[kMaxRanks], t[kMaxRanks];
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r) {
      if (r < world) {
        const float* base = reinterpret_cast<const float*>(stats.ptr[r]);
        l[r] = ld_peer_16_fresh(base + 4 * v);
        t[r] = ld_peer_16_fresh(base + rows + 4 * v);
      }
    }
    float lse[4], tgt[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float m = -INFINITY, sum = 0.f, tg = 0.f;
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < world) m = fmaxf(m, __uint_as_float((&l[r].x)[e]));
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < world) {
          sum += __expf(__uint_as_float((&l[r].x)[e]) - m);
          tg += __uint_as_float((&t[r].x)[e]);
        }
      }
      lse[e] = m + logf(sum);
      tgt[e] = tg;
    }
    float nll[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) nll[e] = labels[4 * v + e] != ignore_index ? lse[e] - tgt[e] : 0.f;
    *reinterpret_cast<float4*>(lse_out + 4 * v) = make_float4(lse[0], lse[1], lse[2], lse[3]);
    *reinterpret_cast<float4*>(nll_out + 4 * v) = make_float4(nll[0], nll[1], nll[2], nll[3]);
  }
}

// ---- vocabulary-parallel embedding: each token's row lives on exactly one rank (the owner of its id) and is needed
// by exactly one rank (the owner of its sequence shard), so the "masked lookup + reduce-scatter" of the reference
// (RowwiseParallel embedding, models/llama/llama_model.py:205-208) is a scatter: the id's owner stores the row
// straight into the sequence owner's [rows_per_rank, dim] buffer.  No zero-filled partial results, no reduction.
__global__ void __launch_bounds__(256)
vocab_embedding_scatter_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ weight, Peers out,
                               int tokens, int dim, int64_t v0, int v_local, int rows_per_rank, int64_t pad_local) {
  const int vec_per_row = dim >> 3;
  const int64_t total = (int64_t)tokens * vec_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / vec_per_row;
    const int c = (int)(i - t * vec_per_row) << 3;
    const int64_t id = ids[t] - v0;
    if (id < 0 || id >= v_local) continue;                       // another rank's row
    uint4 row = make_uint4(0u, 0u, 0u, 0u);                      // the padding row embeds to zero
    if (id != pad_local) row = *reinterpret_cast<const uint4*>(weight + id * dim + c);
    const int dst = (int)(t / rows_per_rank);
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out.ptr[dst]) + (t - (int64_t)dst * rows_per_rank) * dim + c;
    st_peer_16(o, row);
  }
}

}  // namespace comm
}  // namespace llmt

using namespace llmt;
using namespace llmt::comm;

// All `*_ptrs` arguments are HOST arrays of `world` device pointers (this rank's view of every peer's buffer).

extern "C" int llmt_comm_barrier(const void* const* signal_pads, int rank, int world, int slot_base, uint32_t epoch,
                                 cudaStream_t stream) {
  Peers pads;
  if (fill_peers(&pads, signal_pads, world)) return 1;
  barrier_kernel<<<1, 32, 0, stream>>>(pads, rank, world, slot_base, epoch);
  return check_launch("comm_barrier");
}

extern "C" int llmt_comm_allgather_pull(const void* const* full_ptrs, int rank, int world, int64_t shard_bytes,
                                        cudaStream_t stream) {
  LLMT_REQUIRE(shard_bytes % 16 == 0, "allgather_pull: shard must be a multiple of 16 bytes");
  Peers full;
  if (fill_peers(&full, full_ptrs, world)) return 1;
  if (world == 1 || shard_bytes == 0) return 0;
  const int64_t nvec = shard_bytes / 16;
  allgather_pull_kernel<<<grid_for(nvec, 512, 2), 512, 0, stream>>>(full, rank, world, nvec);
  return check_launch("comm_allgather_pull");
}

// dst_full[r * shard_bytes ..] <- shard_ptrs[r][0 .. shard_bytes) for every rank r (shard_ptrs: this rank's view of
// every peer's slice, HOST array).  The caller separates this from the writers of the slices with a barrier.
extern "C" int llmt_comm_gather_shards(const void* const* shard_ptrs, void* dst_full, int rank, int world,
                                       int64_t shard_bytes, cudaStream_t stream) {
  LLMT_REQUIRE(shard_bytes % 16 == 0 && (uintptr_t)dst_full % 16 == 0, "gather_shards: 16-byte alignment required");
  Peers shards;
  if (fill_peers(&shards, shard_ptrs, world)) return 1;
  if (shard_bytes == 0) return 0;
  const int64_t nvec = shard_bytes / 16;
  gather_shards_kernel<<<grid_for(nvec, 512, 2), 512, 0, stream>>>(shards, (uint4*)dst_full, rank, world, nvec);
  return check_launch("comm_gather_shards");
}

// dst_full bytes [lo, hi) minus [skip_lo, skip_hi) <- the owners' slices (all offsets in bytes, multiples of 16).
extern "C" int llmt_comm_gather_range(const void* const* shard_ptrs, void* dst_full, int world, int64_t shard_bytes,
                                      int64_t lo, int64_t hi, int64_t skip_lo, int64_t skip_hi, cudaStream_t stream) {
  LLMT_REQUIRE(shard_bytes % 16 == 0 && lo % 16 == 0 && hi % 16 == 0 && skip_lo % 16 == 0 && skip_hi % 16 == 0 &&
                   (uintptr_t)dst_full % 16 == 0, "gather_range: 16-byte alignment required");
  LLMT_REQUIRE(lo >= 0 && hi <= shard_bytes * world, "gather_range: range exceeds the unit");
  Peers shards;
  if (fill_peers(&shards, shard_ptrs, world)) return 1;
  if (hi <= lo) return 0;
  gather_range_kernel<<<grid_for((hi - lo) / 16, 512, 2), 512, 0, stream>>>(
      shards, (uint4*)dst_full, shard_bytes / 16, lo / 16, hi / 16, skip_lo / 16, skip_hi / 16);
  return check_launch("comm_gather_range");
}

extern "C" int llmt_comm_rs_num_partials() { return num_sms() * 2; }

// out_shard[i] = mean over ranks of grad_r[rank*shard_elems + i]; norm_sq[0] (+)= sum(out_shard^2).
// partial: fp32 scratch of llmt_comm_rs_num_partials() floats.
extern "C" int llmt_comm_rs_sumsq(const void* const* grad_ptrs, int rank, int world, int64_t shard_elems, int is_fp32,
                                  void* out_shard, float* partial, float* norm_sq, int accumulate_norm,
                                  cudaStream_t stream) {
  const int per = is_fp32 ? 4 : 8;
  LLMT_REQUIRE(shard_elems % per == 0, "rs_sumsq: shard must be a multiple of 16 bytes");
  Peers grads;
  if (fill_peers(&grads, grad_ptrs, world)) return 1;
  if (shard_elems == 0) return 0;
  const int grid = grid_for(shard_elems / per, 512, 2);
  if (is_fp32)
    rs_sumsq_kernel<float><<<grid, 512, 0, stream>>>(grads, rank, world, shard_elems, (float*)out_shard,
                                                     1.f / world, partial);
  else
    rs_sumsq_kernel<__nv_bfloat16><<<grid, 512, 0, stream>>>(grads, rank, world, shard_elems,
                                                             (__nv_bfloat16*)out_shard, 1.f / world, partial);
  int rc = check_launch("comm_rs_sumsq");
  if (rc) return rc;
  if (norm_sq != nullptr) {
    sumsq_accumulate_kernel<<<1, 1024, 0, stream>>>(partial, grid, norm_sq, accumulate_norm);
    return check_launch("comm_rs_sumsq_final");
  }
  return 0;
}

// value[0..n) <- sum over ranks (n <= 32). bufs: symmetric fp32 [world][32]; uses barrier slots like llmt_comm_barrier.
// Same result layout as llmt_comm_rs_sumsq, computed by in-switch reduction (NVLS): mc_grads is the MULTICAST address
// of the gradient buffer's unit (bf16).  Needs a symmetric allocation that has a multicast mapping; a cross-rank
// barrier must precede the call (every rank's gradients written) and follow it before the buffer is rewritten.
extern "C" int llmt_comm_rs_sumsq_mc(const void* mc_grads, int rank, int world, int64_t shard_elems, void* out_shard,
                                     float* partial, float* norm_sq, int accumulate_norm, cudaStream_t stream) {
  LLMT_REQUIRE(mc_grads != nullptr && (uintptr_t)mc_grads % 16 == 0, "rs_sumsq_mc: multicast pointer missing / misaligned");
  LLMT_REQUIRE(shard_elems % 8 == 0 && world >= 1 && rank >= 0 && rank < world, "rs_sumsq_mc: bad shard / rank");
  if (shard_elems == 0) return 0;
  const int grid = grid_for(shard_elems / 8, 512, 2);
  rs_sumsq_mc_kernel<<<grid, 512, 0, stream>>>((const uint4*)mc_grads, rank, shard_elems, (__nv_bfloat16*)out_shard,
                                               1.f / world, partial);
  int rc = check_launch("comm_rs_sumsq_mc");
  if (rc) return rc;
  if (norm_sq != nullptr) {
    sumsq_accumulate_kernel<<<1, 1024, 0, stream>>>(partial, grid, norm_sq, accumulate_norm);
    return check_launch("comm_rs_sumsq_mc_final");
  }
  return 0;
}

extern "C" int llmt_comm_allreduce_f32(const void* const* buf_ptrs, const void* const* signal_pads, int rank,
                                       int world, int slot_base, uint32_t epoch, float* value, int n,
                                       cudaStream_t stream) {
  LLMT_REQUIRE(n >= 1 && n <= 32 && world * n <= 512, "allreduce_f32: n must be in [1, 32]");
  Peers bufs, pads;
  if (fill_peers(&bufs, buf_ptrs, world) || fill_peers(&pads, signal_pads, world)) return 1;
  allreduce_oneshot_kernel<<<1, 512, 0, stream>>>(bufs, pads, rank, world, slot_base, epoch, value, n);
  return check_launch("comm_allreduce_f32");
}

// mc_params: multicast (NVSwitch) address of the same arena range, or NULL -> one unicast store per peer.
extern "C" int llmt_comm_adamw_push_mc(float* master, float* m, float* v, const void* grad_shard,
                                       const void* const* param_ptrs, void* mc_params, int rank, int world,
                                       int64_t shard_elems, int grad_is_fp32, float lr, float beta1, float beta2,
                                       float eps, float weight_decay, int step, float grad_scale, float max_norm,
                                       const float* norm_sq, cudaStream_t stream) {
  LLMT_REQUIRE(step >= 1, "adamw_push: step must be >= 1");
  LLMT_REQUIRE(shard_elems % 8 == 0, "adamw_push: shard must be a multiple of 8 elements");
  LLMT_REQUIRE((uintptr_t)mc_params % 16 == 0, "adamw_push: multicast pointer must be 16-byte aligned");
  Peers params;
  if (fill_peers(&params, param_ptrs, world)) return 1;
  if (shard_elems == 0) return 0;
  comm::AdamWArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));   // in double: 1 - beta^step cancels badly in fp32
  a.bc2 = (float)(1.0 - pow((double)beta2, (double)step));
  a.grad_scale = grad_scale; a.max_norm = max_norm;
  const int grid = grid_for(shard_elems / 8, 256, update_ctas_per_sm());
  if (grad_is_fp32)
    adamw_push_kernel<float><<<grid, 256, 0, stream>>>(master, m, v, (const float*)grad_shard, params,
                                                       (uint4*)mc_params, rank, world, shard_elems, a, norm_sq);
  else
    adamw_push_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(master, m, v, (const __nv_bfloat16*)grad_shard, params,
                                                               (uint4*)mc_params, rank, world, shard_elems, a,
                                                               norm_sq);
  return check_launch("comm_adamw_push");
}

extern "C" int llmt_comm_adamw_push(float* master, float* m, float* v, const void* grad_shard,
                                    const void* const* param_ptrs, int rank, int world, int64_t shard_elems,
                                    int grad_is_fp32, float lr, float beta1, float beta2, float eps,
                                    float weight_decay, int step, float grad_scale, float max_norm,
                                    const float* norm_sq, cudaStream_t stream) {
  return llmt_comm_adamw_push_mc(master, m, v, grad_shard, param_ptrs, nullptr, rank, world, shard_elems,
                                 grad_is_fp32, lr, beta1, beta2, eps, weight_decay, step, grad_scale, max_norm,
                                 norm_sq, stream);
}

// stats_ptrs: every rank's {lse_local[rows], tgt_local[rows]} fp32 block in symmetric memory (HOST array of this
// rank's mappings); the caller puts a barrier between the writers and this call.  rows % 4 == 0.
extern "C" int llmt_comm_vocab_loss_combine(const void* const* stats_ptrs, int world, int rows, const int64_t* labels,
                                            int64_t ignore_index, float* lse_out, float* nll_out,
                                            cudaStream_t stream) {
  LLMT_REQUIRE(rows % 4 == 0 && world >= 1 && world <= kMaxRanks, "vocab_loss_combine: rows=%d must be a multiple of 4, "
               "world=%d at most %d", rows, world, kMaxRanks);
  Peers stats;
  if (fill_peers(&stats, stats_ptrs, world)) return 1;
  if (rows == 0) return 0;
  vocab_loss_combine_kernel<<<grid_for(rows / 4, 256, 1), 256, 0, stream>>>(stats, world, rows, labels, ignore_index,
                                                                         lse_out, nll_out);
  return check_launch("comm_vocab_loss_combine");
}

// out_ptrs: every rank's [rows_per_rank, dim] bf16 output buffer in symmetric memory.  ids: all `tokens` ids (the
// same on every rank); weight: this rank's rows [v0, v0 + v_local) of the embedding table.  The caller follows the
// call with a barrier before anybody reads its output buffer.
extern "C" int llmt_comm_vocab_embedding_scatter(const int64_t* ids, const void* weight, const void* const* out_ptrs,
                                                 int world, int tokens, int dim, int64_t v0, int v_local,
                                                 int rows_per_rank, int64_t pad_local, cudaStream_t stream) {
  LLMT_REQUIRE(dim % 8 == 0 && rows_per_rank > 0 && (int64_t)rows_per_rank * world >= tokens,
               "vocab_embedding_scatter: dim=%d must be a multiple of 8 and %d ranks x %d rows must cover %d tokens", dim,
               world, rows_per_rank, tokens);
  Peers out;
  if (fill_peers(&out, out_ptrs, world)) return 1;
  if (tokens == 0) return 0;
  vocab_embedding_scatter_kernel<<<grid_for((int64_t)tokens * dim / 8, 256, 4), 256, 0, stream>>>(
      ids, (const __nv_bfloat16*)weight, out, tokens, dim, v0, v_local, rows_per_rank, pad_local);
  return check_launch("comm_vocab_embedding_scatter");
}
