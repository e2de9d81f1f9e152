// RMSNorm forward / backward for bf16 activations (fp32 math), rows x cols with cols % 8 == 0.
//
//   y = x * rsqrt(mean(x^2) + eps) * w
//
// Replaces Liger's Triton RMSNorm on the reference hot path
// (reference: src/llm_training/ops/liger_kernel/rms_norm_op.py:15-19, called from
// models/llama/llama_model.py:283).  Memory-bound: one pass over x, row kept in registers,
// 16-byte vector accesses, one CTA per row in a grid-stride loop.
//
// Backward:  dx = rstd * (g - xhat * mean(g * xhat)),  g = dy * w,  xhat = x * rstd
//            dw = sum_rows dy * xhat
// dw is produced in two stages: every CTA accumulates its rows into registers and writes one fp32
// partial row into caller-provided scratch [grid, cols]; a second kernel sums the partials.
#include "common.cuh"

namespace llmt {

constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 8;  // cols <= 256 * 8 * 8 = 16384

template <int MAXV>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                   __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out, int rows, int cols,
                   float eps) {
  __shared__ float red[32];
  const int nvec = cols >> 3;
  float wreg[MAXV][8];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) unpack8(ld_bf16x8(w + (size_t)v * 8), wreg[i]);
  }
  // the next row's vectors are requested before this row's reduction: one row of loads is always in flight behind the
  // block-wide sum (the un-pipelined loop reached 49% of the HBM copy bandwidth in ncu)
  bf16x8 nxt[MAXV];
  auto load_row = [&](int row) {
    const __nv_bfloat16* xr = x + (size_t)row * cols;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) nxt[i] = ld_bf16x8_stream(xr + (size_t)v * 8);
    }
  };
  if ((int)blockIdx.x < rows) load_row(blockIdx.x);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    float xv[MAXV][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        unpack8(nxt[i], xv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += xv[i][j] * xv[i][j];
      }
    }
    if (row + (int)gridDim.x < rows) load_row(row + gridDim.x);
    ss = block_sum(ss, red);
    const float rstd = rsqrtf(ss / (float)cols + eps);
    if (threadIdx.x == 0 && rstd_out != nullptr) rstd_out[row] = rstd;
    __nv_bfloat16* yr = y + (size_t)row * cols;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = xv[i][j] * rstd * wreg[i][j];
        st_bf16x8(yr + (size_t)v * 8, pack8(o));
      }
    }
  }
}

template <int MAXV>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                   const __nv_bfloat16* __restrict__ dres,      // nullable: gradient of the residual branch, added to dx
                   __nv_bfloat16* __restrict__ dx, float* __restrict__ dw_partial, int rows, int cols) {
  __shared__ float red[32];
  const int nvec = cols >> 3;
  float wreg[MAXV][8];
  float dwacc[MAXV][8];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) unpack8(ld_bf16x8(w + (size_t)v * 8), wreg[i]);
#pragma unroll
    for (int j = 0; j < 8; ++j) dwacc[i][j] = 0.f;
  }
  bf16x8 nx[MAXV], ndy[MAXV], nres[MAXV];
  auto load_row = [&](int row) {
    const size_t off = (size_t)row * cols;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        nx[i] = ld_bf16x8_stream(x + off + (size_t)v * 8);
        ndy[i] = ld_bf16x8_stream(dy + off + (size_t)v * 8);
        if (dres != nullptr) nres[i] = ld_bf16x8_stream(dres + off + (size_t)v * 8);
      }
    }
  };
  if ((int)blockIdx.x < rows) load_row(blockIdx.x);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t off = (size_t)row * cols;
    const float rstd = rstd_in[row];
    float xh[MAXV][8], g[MAXV][8];
    bf16x8 cres[MAXV];                                   // this row's dres: the prefetch below reuses nres
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        float dyv[8];
        unpack8(nx[i], xh[i]);
        unpack8(ndy[i], dyv);
        if (dres != nullptr) cres[i] = nres[i];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] *= rstd;
          g[i][j] = dyv[j] * wreg[i][j];
          dot += g[i][j] * xh[i][j];
          dwacc[i][j] += dyv[j] * xh[i][j];
        }
      }
    }
    if (row + (int)gridDim.x < rows) load_row(row + gridDim.x);   // in flight behind the block-wide sum below
    dot = block_sum(dot, red) / (float)cols;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - xh[i][j] * dot);
        if (dres != nullptr) {
          float r[8];
          unpack8(cres[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];           // fp32 add, one rounding
        }
        st_bf16x8(dx + off + (size_t)v * 8, pack8(o));
      }
    }
  }
  float* part = dw_partial + (size_t)blockIdx.x * cols;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      *reinterpret_cast<float4*>(part + (size_t)v * 8) =
          make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      *reinterpret_cast<float4*>(part + (size_t)v * 8 + 4) =
          make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
    }
  }
}

// dw[c] (+)= sum_p partial[p, c]; output bf16 or fp32.  One CTA per 32 columns: 16 row groups stride over the
// partial rows (coalesced 128-byte row segments), then a fixed-order smem reduction — the first version walked all
// (up to 592) partial rows serially in one thread per column and cost 3x the backward kernel it finishes.
constexpr int kColsumRows = 16;
template <typename OutT>
__global__ void __launch_bounds__(32 * kColsumRows)
colsum_partials_kernel(const float* __restrict__ partial, OutT* __restrict__ out, int nparts, int cols,
                       int accumulate) {
  __shared__ float tile[kColsumRows][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < cols) {
    float s0 = 0.f, s1 = 0.f;
    int p = threadIdx.y;
    for (; p + kColsumRows < nparts; p += 2 * kColsumRows) {
      s0 += partial[(size_t)p * cols + c];
      s1 += partial[(size_t)(p + kColsumRows) * cols + c];
    }
    if (p < nparts) s0 += partial[(size_t)p * cols + c];
    s = s0 + s1;
  }
  tile[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int y = 0; y < kColsumRows; ++y) t += tile[y][threadIdx.x];
    if (accumulate) t += (float)out[c];
    out[c] = (OutT)t;
  }
}

static int launch_colsum(const float* partial, void* dw, int nparts, int cols, int dw_is_fp32, int accumulate,
                         cudaStream_t stream) {
  const dim3 block(32, kColsumRows);
  const int grid = (cols + 31) / 32;
  if (dw_is_fp32)
    colsum_partials_kernel<float><<<grid, block, 0, stream>>>(partial, (float*)dw, nparts, cols, accumulate);
  else
    colsum_partials_kernel<__nv_bfloat16><<<grid, block, 0, stream>>>(partial, (__nv_bfloat16*)dw, nparts, cols,
                                                                     accumulate);
  return 0;
}

int norm_grid(int rows) {
  const int g = num_sms() * 4;
  return rows < g ? rows : g;
}

}  // namespace llmt

using namespace llmt;

extern "C" int llmt_rmsnorm_bwd_num_partials(int rows) { return norm_grid(rows); }

extern "C" int llmt_rmsnorm_fwd(const void* x, const void* w, void* y, float* rstd, int rows, int cols,
                                float eps, cudaStream_t stream) {
  LLMT_REQUIRE(cols % 8 == 0 && cols <= kNormThreads * 8 * kNormMaxVec,
               "rmsnorm: cols=%d must be a multiple of 8 and <= %d", cols, kNormThreads * 8 * kNormMaxVec);
  if (rows == 0) return 0;
  const int grid = norm_grid(rows);
  auto X = (const __nv_bfloat16*)x;
  auto W = (const __nv_bfloat16*)w;
  auto Y = (__nv_bfloat16*)y;
  const int nvec = cols / 8;
  if (nvec <= kNormThreads * 2)
    rmsnorm_fwd_kernel<2><<<grid, kNormThreads, 0, stream>>>(X, W, Y, rstd, rows, cols, eps);
  else if (nvec <= kNormThreads * 4)
    rmsnorm_fwd_kernel<4><<<grid, kNormThreads, 0, stream>>>(X, W, Y, rstd, rows, cols, eps);
  else
    rmsnorm_fwd_kernel<8><<<grid, kNormThreads, 0, stream>>>(X, W, Y, rstd, rows, cols, eps);
  return check_launch("rmsnorm_fwd");
}

// dw_partial: fp32 scratch [llmt_rmsnorm_bwd_num_partials(rows), cols]. dw_is_fp32 selects the output
// type of dw; accumulate_dw adds into the existing dw (gradient accumulation).
// dres (nullable, bf16 [rows, cols]): gradient arriving through the residual connection around the norm; the kernel
// returns dx = norm backward + dres, so the pre-norm block's input gradient needs no separate add pass.
extern "C" int llmt_rmsnorm_bwd_res(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                                    void* dx, void* dw, float* dw_partial, int rows, int cols, int dw_is_fp32,
                                    int accumulate_dw, cudaStream_t stream) {
  LLMT_REQUIRE(cols % 8 == 0 && cols <= kNormThreads * 8 * kNormMaxVec,
               "rmsnorm_bwd: cols=%d must be a multiple of 8 and <= %d", cols, kNormThreads * 8 * kNormMaxVec);
  if (rows == 0) return 0;
  const int grid = norm_grid(rows);
  auto DY = (const __nv_bfloat16*)dy;
  auto X = (const __nv_bfloat16*)x;
  auto W = (const __nv_bfloat16*)w;
  auto R = (const __nv_bfloat16*)dres;
  auto DX = (__nv_bfloat16*)dx;
  const int nvec = cols / 8;
  if (nvec <= kNormThreads * 2)
    rmsnorm_bwd_kernel<2><<<grid, kNormThreads, 0, stream>>>(DY, X, W, rstd, R, DX, dw_partial, rows, cols);
  else if (nvec <= kNormThreads * 4)
    rmsnorm_bwd_kernel<4><<<grid, kNormThreads, 0, stream>>>(DY, X, W, rstd, R, DX, dw_partial, rows, cols);
  else
    rmsnorm_bwd_kernel<8><<<grid, kNormThreads, 0, stream>>>(DY, X, W, rstd, R, DX, dw_partial, rows, cols);
  int rc = check_launch("rmsnorm_bwd");
  if (rc) return rc;
  launch_colsum(dw_partial, dw, grid, cols, dw_is_fp32, accumulate_dw, stream);
  return check_launch("rmsnorm_bwd_colsum");
}

extern "C" int llmt_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, void* dx,
                                void* dw, float* dw_partial, int rows, int cols, int dw_is_fp32,
                                int accumulate_dw, cudaStream_t stream) {
  return llmt_rmsnorm_bwd_res(dy, x, w, rstd, nullptr, dx, dw, dw_partial, rows, cols, dw_is_fp32, accumulate_dw,
                              stream);
}

// -------------------------------------------------------------------------------------------------------
// RMSNorm backward for callers that must not allocate.
//
// Some users of this library run where device allocation is off limits: a process whose device memory is
// handed out in full by a fixed plan or by another allocator (nothing is left for cudaMalloc, and a failed
// allocation would take the job down), a latency-critical loop that cannot afford the implicit device
// synchronisation of cudaMalloc / cudaFree, or a plain C / C++ host that links this .so without torch and has
// no scratch pool to lend us.  For them the dw partial sums live in the fixed device array below instead of a
// caller buffer; the call itself performs no cudaMalloc, cudaFree or other allocating call.  The array is part of the
// library image: every GPU that loads the library keeps kNoAllocScratchFloats * 4 bytes (4.625 MiB) of it until
// the process exits.
//
// Limits: cols <= kNoAllocMaxCols (8192); at most kNoAllocMaxParts (148) partial rows are used, so the grid is
// min(rows, 148) CTAs.  Larger `cols` are rejected with an error.
//
// The array is shared by every call on the device and keeps whatever the previous call left in it.  Nothing
// has to start at zero: each partial row that the column-sum reads is written in full by the same call first.
// Two calls that overlap on the GPU (different streams) would use the same partial rows and race; calls on
// one stream are ordered and safe.
//
// The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that
// first call before starting any stream capture.
// -------------------------------------------------------------------------------------------------------
namespace llmt {
constexpr int kNoAllocMaxCols = 8192;
constexpr int kNoAllocMaxParts = 148;
constexpr size_t kNoAllocScratchFloats = (size_t)kNoAllocMaxParts * kNoAllocMaxCols;
static __device__ __align__(256) float g_rmsnorm_noalloc_scratch[kNoAllocScratchFloats];

constexpr int kNoAllocMaxDevices = 64;
static float* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

// Host code only knows a stand-in for the array; the kernels need its real device address.
static int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices) {
    set_error("rmsnorm_bwd_noalloc: cannot identify the current device (%s)", cudaGetErrorString(e));
    return 2;
  }
  float* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* addr = nullptr;
    e = cudaGetSymbolAddress(&addr, g_rmsnorm_noalloc_scratch);
    if (e != cudaSuccess || addr == nullptr) {
      set_error("rmsnorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return 2;
    }
    p = static_cast<float*>(addr);
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return 0;
}
}  // namespace llmt

extern "C" int llmt_rmsnorm_bwd_noalloc_max_cols() { return llmt::kNoAllocMaxCols; }

extern "C" int llmt_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* rstd, void* dx,
                                        void* dw, int rows, int cols, int dw_is_fp32, int accumulate_dw,
                                        cudaStream_t stream) {
  LLMT_REQUIRE(cols % 8 == 0, "rmsnorm_bwd_noalloc: cols=%d must be a multiple of 8", cols);
  LLMT_REQUIRE(cols <= kNoAllocMaxCols,
               "rmsnorm_bwd_noalloc: cols=%d exceeds the built-in scratch (max %d columns); use llmt_rmsnorm_bwd "
               "with a caller-provided buffer",
               cols, kNoAllocMaxCols);
  if (rows == 0) return 0;
  float* partial = nullptr;
  int rc = noalloc_scratch_for_current_device(&partial);
  if (rc) return rc;
  const int grid = rows < kNoAllocMaxParts ? rows : kNoAllocMaxParts;
  auto DY = (const __nv_bfloat16*)dy;
  auto X = (const __nv_bfloat16*)x;
  auto W = (const __nv_bfloat16*)w;
  auto DX = (__nv_bfloat16*)dx;
  const int nvec = cols / 8;
  if (nvec <= kNormThreads * 2)
    rmsnorm_bwd_kernel<2><<<grid, kNormThreads, 0, stream>>>(DY, X, W, rstd, nullptr, DX, partial, rows, cols);
  else
    rmsnorm_bwd_kernel<4><<<grid, kNormThreads, 0, stream>>>(DY, X, W, rstd, nullptr, DX, partial, rows, cols);
  rc = check_launch("rmsnorm_bwd_noalloc");
  if (rc) return rc;
  launch_colsum(partial, dw, grid, cols, dw_is_fp32, accumulate_dw, stream);
  return check_launch("rmsnorm_bwd_noalloc_colsum");
}
