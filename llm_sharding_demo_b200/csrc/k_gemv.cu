// Bandwidth-bound decode kernels: the Conv1D GEMV family (c_attn / c_proj / c_fc / mlp
// c_proj with fused LayerNorm prologue and bias / GELU / residual / KV-scatter epilogues)
// and the fused ln_f + lm_head + greedy argmax.  Weights are stored [N, K] with K contiguous
// so that one warp streams one output row with 128-bit loads and reduces with shuffles.
#include "common.cuh"
#include "kernels.h"

namespace gp {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

// Stage x[B,K] into shared memory (zero rows beyond B) and, when ln_g != NULL, normalise each
// row in place: (x - mean) * rsqrt(var + eps) * g + b with the biased variance of
// nn.LayerNorm ([hf] modeling_gpt2.py:252,254,505).
template <int BT>
__device__ __forceinline__ void stage_x(float* xs, float* red, const float* __restrict__ x, int ldx, int B,
                                        int K, const float* __restrict__ ln_g,
                                        const float* __restrict__ ln_b, float eps) {
  const int tid = threadIdx.x;
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    for (int k = tid; k < K; k += kThreads) xs[b * K + k] = (b < B) ? x[(size_t)b * ldx + k] : 0.f;
  }
  __syncthreads();
  if (ln_g != nullptr) {
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      if (b >= B) break;
      float s = 0.f;
      for (int k = tid; k < K; k += kThreads) s += xs[b * K + k];
      const float mean = block_sum<kWarps>(s, red) / (float)K;
      float v = 0.f;
      for (int k = tid; k < K; k += kThreads) {
        float dlt = xs[b * K + k] - mean;
        v += dlt * dlt;
      }
      const float rstd = rsqrtf(block_sum<kWarps>(v, red) / (float)K + eps);
      for (int k = tid; k < K; k += kThreads)
        xs[b * K + k] = (xs[b * K + k] - mean) * rstd * ln_g[k] + ln_b[k];
    }
    __syncthreads();
  }
}

// acc[b] += sum_k w[row,k] * xs[b,k] for one weight row, all lanes cooperating.
template <typename WT, int BT, bool VEC>
__device__ __forceinline__ void row_dot(const WT* __restrict__ wrow, const float* xs, int K, int lane,
                                        float (&acc)[BT]) {
  if (VEC) {
#pragma unroll 4
    for (int k0 = lane * 8; k0 < K; k0 += 256) {
      float w[8];
      load8_stream(wrow + k0, w);
#pragma unroll
      for (int b = 0; b < BT; ++b) {
        const float4 x0 = *reinterpret_cast<const float4*>(xs + b * K + k0);
        const float4 x1 = *reinterpret_cast<const float4*>(xs + b * K + k0 + 4);
        acc[b] = fmaf(w[0], x0.x, acc[b]); acc[b] = fmaf(w[1], x0.y, acc[b]);
        acc[b] = fmaf(w[2], x0.z, acc[b]); acc[b] = fmaf(w[3], x0.w, acc[b]);
        acc[b] = fmaf(w[4], x1.x, acc[b]); acc[b] = fmaf(w[5], x1.y, acc[b]);
        acc[b] = fmaf(w[6], x1.z, acc[b]); acc[b] = fmaf(w[7], x1.w, acc[b]);
      }
    }
  } else {
    for (int k = lane; k < K; k += 32) {
      const float w = to_f(wrow[k]);
#pragma unroll
      for (int b = 0; b < BT; ++b) acc[b] = fmaf(w, xs[b * K + k], acc[b]);
    }
  }
}

template <typename WT>
__device__ __forceinline__ void gemv_epilogue(const GemvParams& p, int b, int n, float v) {
  if (p.bias) v += p.bias[n];
  if (p.mode == 1) {
    // c_attn: columns [0,H) are q, [H,2H) k, [2H,3H) v ([hf] modeling_gpt2.py:185-191)
    if (n < p.H) {
      p.q_out[(size_t)b * p.H + n] = v;
    } else {
      const int which = n >= 2 * p.H;
      const int c = n - p.H * (1 + which);
      const int head = c / p.d, dd = c - head * p.d;
      const int slot = p.slot0 + b;
      const int pos = p.lens[slot];
      WT* cache = reinterpret_cast<WT*>(which ? p.vcache : p.kcache);
      // a full slot is refused on the host (engine.cu: slot_room); never write past the slot's rows
      if (pos < p.max_ctx) cache[(((size_t)slot * p.nh + head) * p.max_ctx + pos) * p.d + dd] = from_f<WT>(v);
    }
    return;
  }
  if (p.flags & GP_EPI_GELU) v = gelu_new(v);
  if (p.residual) v += p.residual[(size_t)b * p.ldr + n];
  p.out[(size_t)b * p.ldo + n] = v;
}

template <typename WT, int BT, bool VEC>
__global__ void __launch_bounds__(kThreads) gemv_kernel(const GemvParams p) {
  extern __shared__ __align__(16) float xs[];
  __shared__ float red[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  stage_x<BT>(xs, red, p.x, p.ldx, p.B, p.K, p.ln_g, p.ln_b, p.eps);

  const WT* __restrict__ W = reinterpret_cast<const WT*>(p.Wt);
  const int stride = gridDim.x * kWarps;
  // two rows per warp in flight to double the bytes outstanding per lane
  for (int n0 = blockIdx.x * kWarps + warp; n0 < p.N; n0 += 2 * stride) {
    const int n1 = n0 + stride;
    float a0[BT], a1[BT];
#pragma unroll
    for (int b = 0; b < BT; ++b) a0[b] = a1[b] = 0.f;
    row_dot<WT, BT, VEC>(W + (size_t)n0 * p.K, xs, p.K, lane, a0);
    if (n1 < p.N) row_dot<WT, BT, VEC>(W + (size_t)n1 * p.K, xs, p.K, lane, a1);
    float v0 = 0.f, v1 = 0.f;
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      const float s0 = warp_sum(a0[b]), s1 = warp_sum(a1[b]);
      if (lane == b) { v0 = s0; v1 = s1; }
    }
    if (lane < p.B) {
      gemv_epilogue<WT>(p, lane, n0, v0);
      if (n1 < p.N) gemv_epilogue<WT>(p, lane, n1, v1);
    }
  }
}

// ------------------------------------------------------------------------------------------
// ln_f + tied lm_head + greedy argmax.  Every warp keeps the best (value, index) per batch row
// over the vocabulary rows it streams; the CTA writes one partial per batch row to `scratch`
// and the last CTA to arrive (ticket) reduces the partials, lowest index winning ties, so the
// result is deterministic.  Logits only leave registers if `logits_out` is given.
struct Best {
  float v;
  int i;
};
__device__ __forceinline__ void best_merge(Best& a, float v, int i) {
  if (v > a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; }
}

template <typename WT, int BT, bool VEC>
__global__ void __launch_bounds__(kThreads) lmhead_argmax_kernel(const LmheadParams p) {
  extern __shared__ __align__(16) float xs[];
  __shared__ float red[kWarps];
  __shared__ float sbv[kWarps][BT];
  __shared__ int sbi[kWarps][BT];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  stage_x<BT>(xs, red, p.x, p.ldx, p.B, p.H, p.ln_g, p.ln_b, p.eps);

  const WT* __restrict__ W = reinterpret_cast<const WT*>(p.wte);
  const int stride = gridDim.x * kWarps;
  Best best[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) { best[b].v = -INFINITY; best[b].i = 0x7fffffff; }
  for (int n0 = blockIdx.x * kWarps + warp; n0 < p.V; n0 += 2 * stride) {
    const int n1 = n0 + stride;
    float a0[BT], a1[BT];
#pragma unroll
    for (int b = 0; b < BT; ++b) a0[b] = a1[b] = 0.f;
    row_dot<WT, BT, VEC>(W + (size_t)n0 * p.H, xs, p.H, lane, a0);
    if (n1 < p.V) row_dot<WT, BT, VEC>(W + (size_t)n1 * p.H, xs, p.H, lane, a1);
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      const float s0 = warp_sum(a0[b]), s1 = warp_sum(a1[b]);
      best_merge(best[b], s0, n0);
      if (n1 < p.V) best_merge(best[b], s1, n1);
      if (p.logits_out != nullptr && lane == 0 && b < p.B) {
        p.logits_out[(size_t)b * p.ldl + n0] = s0;
        if (n1 < p.V) p.logits_out[(size_t)b * p.ldl + n1] = s1;
      }
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int b = 0; b < BT; ++b) { sbv[warp][b] = best[b].v; sbi[warp][b] = best[b].i; }
  }
  __syncthreads();
  unsigned* ticket = reinterpret_cast<unsigned*>(p.scratch);
  float* pv = reinterpret_cast<float*>(reinterpret_cast<char*>(p.scratch) + kLmheadScratchHeader);
  int* pi = reinterpret_cast<int*>(pv + (size_t)kLmheadMaxGrid * kGemvMaxBatch);
  if (threadIdx.x < BT) {
    const int b = threadIdx.x;
    Best r{sbv[0][b], sbi[0][b]};
#pragma unroll
    for (int w = 1; w < kWarps; ++w) best_merge(r, sbv[w][b], sbi[w][b]);
    pv[(size_t)blockIdx.x * kGemvMaxBatch + b] = r.v;
    pi[(size_t)blockIdx.x * kGemvMaxBatch + b] = r.i;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // final reduction: warp b reduces batch row b over the gridDim.x partials
  for (int b = warp; b < p.B; b += kWarps) {
    Best r{-INFINITY, 0x7fffffff};
    for (int g = lane; g < (int)gridDim.x; g += 32) {
      const float v = __ldcg(pv + (size_t)g * kGemvMaxBatch + b);
      const int i = __ldcg(pi + (size_t)g * kGemvMaxBatch + b);
      best_merge(r, v, i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(FULL_MASK, r.v, o);
      const int oi = __shfl_xor_sync(FULL_MASK, r.i, o);
      best_merge(r, ov, oi);
    }
    if (lane == 0) {
      p.ids_out[b] = r.i;
      if (p.lens != nullptr) {
        const int old_len = p.lens[p.slot0 + b];
        if (p.seq != nullptr) p.seq[(size_t)b * p.seq_stride + old_len + 1] = r.i;
        p.lens[p.slot0 + b] = old_len + 1;
      }
    }
  }
  if (threadIdx.x == 0) *ticket = 0u;  // ready for the next launch
}

template <typename K>
cudaError_t set_smem(K kernel) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvMaxSmem);
}

template <typename WT, int BT>
cudaError_t gemv_dispatch_vec(const GemvParams& p, int grid, size_t smem, cudaStream_t s) {
  const bool vec = (p.K % 8 == 0);
  if (vec) gemv_kernel<WT, BT, true><<<grid, kThreads, smem, s>>>(p);
  else gemv_kernel<WT, BT, false><<<grid, kThreads, smem, s>>>(p);
  return cudaGetLastError();
}
template <typename WT>
cudaError_t gemv_dispatch(const GemvParams& p, int grid, size_t smem, cudaStream_t s) {
  if (p.B <= 1) return gemv_dispatch_vec<WT, 1>(p, grid, smem, s);
  if (p.B <= 2) return gemv_dispatch_vec<WT, 2>(p, grid, smem, s);
  if (p.B <= 4) return gemv_dispatch_vec<WT, 4>(p, grid, smem, s);
  return gemv_dispatch_vec<WT, 8>(p, grid, smem, s);
}

template <typename WT, int BT>
cudaError_t lmhead_dispatch_vec(const LmheadParams& p, int grid, size_t smem, cudaStream_t s) {
  const bool vec = (p.H % 8 == 0);
  if (vec) lmhead_argmax_kernel<WT, BT, true><<<grid, kThreads, smem, s>>>(p);
  else lmhead_argmax_kernel<WT, BT, false><<<grid, kThreads, smem, s>>>(p);
  return cudaGetLastError();
}
template <typename WT>
cudaError_t lmhead_dispatch(const LmheadParams& p, int grid, size_t smem, cudaStream_t s) {
  if (p.B <= 1) return lmhead_dispatch_vec<WT, 1>(p, grid, smem, s);
  if (p.B <= 2) return lmhead_dispatch_vec<WT, 2>(p, grid, smem, s);
  if (p.B <= 4) return lmhead_dispatch_vec<WT, 4>(p, grid, smem, s);
  return lmhead_dispatch_vec<WT, 8>(p, grid, smem, s);
}

int padded_batch(int B) { return B <= 1 ? 1 : B <= 2 ? 2 : B <= 4 ? 4 : 8; }

}  // namespace

bool gemv_supported(int B, int K) {
  return B >= 1 && B <= kGemvMaxBatch && (size_t)padded_batch(B) * K * sizeof(float) <= (size_t)kGemvMaxSmem;
}

cudaError_t init_gemv_attrs() {
  cudaError_t e;
#define GP_SET(WT, BT)                                                        \
  if ((e = set_smem(gemv_kernel<WT, BT, true>)) != cudaSuccess) return e;     \
  if ((e = set_smem(gemv_kernel<WT, BT, false>)) != cudaSuccess) return e;    \
  if ((e = set_smem(lmhead_argmax_kernel<WT, BT, true>)) != cudaSuccess) return e; \
  if ((e = set_smem(lmhead_argmax_kernel<WT, BT, false>)) != cudaSuccess) return e;
  GP_SET(float, 1) GP_SET(float, 2) GP_SET(float, 4) GP_SET(float, 8)
  GP_SET(bf16, 1) GP_SET(bf16, 2) GP_SET(bf16, 4) GP_SET(bf16, 8)
#undef GP_SET
  return cudaSuccess;
}

cudaError_t launch_gemv(const GemvParams& p, int dtype, cudaStream_t s) {
  if (!gemv_supported(p.B, p.K) || p.N <= 0) return cudaErrorInvalidValue;
  const size_t smem = (size_t)padded_batch(p.B) * p.K * sizeof(float);
  // one output row per warp while that still fills the GPU, two beyond
  int grid = (p.N + kWarps - 1) / kWarps;
  const int cap = device_sm_count() * 8;
  if (grid > cap) grid = cap;
  return dtype == GP_BF16 ? gemv_dispatch<bf16>(p, grid, smem, s) : gemv_dispatch<float>(p, grid, smem, s);
}

cudaError_t launch_lmhead_argmax(const LmheadParams& p, int dtype, cudaStream_t s) {
  if (!gemv_supported(p.B, p.H) || p.V <= 0 || p.scratch == nullptr) return cudaErrorInvalidValue;
  const size_t smem = (size_t)padded_batch(p.B) * p.H * sizeof(float);
  int grid = (p.V + 2 * kWarps - 1) / (2 * kWarps);
  if (grid > kLmheadMaxGrid) grid = kLmheadMaxGrid;
  return dtype == GP_BF16 ? lmhead_dispatch<bf16>(p, grid, smem, s) : lmhead_dispatch<float>(p, grid, smem, s);
}

}  // namespace gp

// ------------------------------------------------------------------------------------------
// gp_op_lmhead_argmax_static: the fused ln_f + lm_head + argmax with NO caller- or
// library-allocated scratch.  Who needs this: a host that embeds the kernel in a decode loop
// where device allocation is off the table - the process's HBM has been handed in full to
// another allocator (a serving runtime or framework pool, so cudaMalloc fails or fragments
// it), or the loop is latency-critical and cannot afford the device-wide synchronisation a
// cudaMalloc/cudaFree implies, or the caller simply owns no allocator it could get the
// `scratch` argument of gp_op_lmhead_argmax from.  The cross-CTA partials and the arrival
// ticket live in the fixed array below instead; every GPU that loads this library carries
// its kStaticScratchBytes until the process exits.
#include <cstdio>
#include <mutex>
namespace gp {
namespace {
std::mutex g_static_scratch_mu;   // guards the per-device set-up below (two host threads may make the first call)
constexpr size_t kStaticScratchBytes =
    kLmheadScratchHeader + (size_t)kLmheadMaxGrid * kGemvMaxBatch * (sizeof(float) + sizeof(int));  // 76,032 B
static __device__ __align__(256) unsigned char g_lmhead_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
void* g_static_scratch_addr[kMaxDevices];  // device address of the array, per CUDA ordinal
bool g_static_scratch_ready[kMaxDevices];
}  // namespace
}  // namespace gp

extern "C" int gp_op_lmhead_argmax_static(const float* x, const float* gamma, const float* beta, const void* wte,
                                          int32_t* ids_out, float* logits_out, int32_t B, int32_t V, int32_t H,
                                          float eps, int32_t dtype, void* stream) {
  using namespace gp;
  auto fail = [&](int code, const char* msg, const char* detail) {
    char buf[400];
    snprintf(buf, sizeof buf, "gp_op_lmhead_argmax_static: %s%s", msg, detail);
    record_error(buf);  // readable through gp_last_error(NULL)
    return code;
  };
  if (!x || !wte || !ids_out) return fail(GP_EINVAL, "NULL argument", "");
  if ((gamma == nullptr) != (beta == nullptr))
    return fail(GP_EINVAL, "gamma and beta must both be given (ln_f is applied) or both be NULL (x is used as is)", "");
  if (dtype != GP_F32 && dtype != GP_BF16) return fail(GP_EINVAL, "dtype must be GP_F32 or GP_BF16", "");
  // the fixed array holds partials for at most kGemvMaxBatch rows; x must fit shared memory
  if (B < 1 || B > kGemvMaxBatch || V < 1 || H < 1 || !gemv_supported(B, H)) {
    char lim[160];
    snprintf(lim, sizeof lim, " (got B=%d V=%d H=%d; limits: 1 <= B <= %d and padded(B)*H*4 <= %d bytes)", B, V, H,
             kGemvMaxBatch, kGemvMaxSmem);
    return fail(GP_EINVAL, "size beyond what the static scratch serves", lim);
  }
  int dev = -1;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) return fail(GP_ECUDA, "cudaGetDevice: ", cudaGetErrorString(ce));
  if (dev < 0 || dev >= kMaxDevices) return fail(GP_EINVAL, "device ordinal out of range", "");
  {
    std::lock_guard<std::mutex> lock(g_static_scratch_mu);
    if (!g_static_scratch_ready[dev]) {
    // One-time set-up on this device (it is current).  In host code the array's NAME is only a
    // host-side handle, not a device address: handing it to a kernel would fault, or silently hit
    // the wrong memory where the GPU can reach host memory.  Ask the runtime for the real address.
    void* p = nullptr;
    ce = cudaGetSymbolAddress(&p, g_lmhead_static_scratch);
    if (ce != cudaSuccess || p == nullptr)
      return fail(GP_ECUDA, "cudaGetSymbolAddress: ", cudaGetErrorString(ce));
    if ((ce = init_gemv_attrs()) != cudaSuccess) return fail(GP_ECUDA, "kernel attribute set-up: ", cudaGetErrorString(ce));
    g_static_scratch_addr[dev] = p;
    g_static_scratch_ready[dev] = true;
    }
  }
  void* scratch = g_static_scratch_addr[dev];
  cudaStream_t s = (cudaStream_t)stream;
  // The array keeps whatever the previous call (or an aborted one) left; only the arrival
  // ticket has to start at zero, the partials are fully rewritten by every launch.
  ce = cudaMemsetAsync(scratch, 0, kLmheadScratchHeader, s);
  if (ce != cudaSuccess) return fail(GP_ECUDA, "cudaMemsetAsync: ", cudaGetErrorString(ce));
  LmheadParams p;
  p.x = x; p.ldx = H; p.ln_g = gamma; p.ln_b = beta; p.eps = eps; p.wte = wte; p.B = B; p.V = V; p.H = H;
  p.ids_out = ids_out; p.logits_out = logits_out; p.ldl = V; p.scratch = scratch;
  ce = launch_lmhead_argmax(p, dtype, s);
  if (ce != cudaSuccess) return fail(ce == cudaErrorInvalidValue ? GP_EINVAL : GP_ECUDA, "launch: ", cudaGetErrorString(ce));
  return GP_OK;
}
