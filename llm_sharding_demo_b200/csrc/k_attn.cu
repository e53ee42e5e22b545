// Causal attention over the device-resident KV cache (decode) and over a c_attn output
// (prefill).  Restates softmax(q k^T * d^-0.5 + causal) v of
// [hf] integrations/sdpa_attention.py:65,77,92-101 / modeling_gpt2.py:96-100,186-222.
// KV cache layout: [slot, head, max_ctx, d], d contiguous, so one (slot, head) is a single
// contiguous stream for the decode kernel.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace gp {
namespace {

constexpr int kDecThreads = 256;
constexpr int kDecWarps = kDecThreads / 32;

template <typename OT> __device__ __forceinline__ void store_out(OT* p, float v) { *p = from_f<OT>(v); }

// ---- decode, head_dim 64: 8 lanes x 8 elements cover one key/value row (128-bit loads for
// bf16), a warp covers 4 rows per load instruction, the CTA covers 32.
template <typename KT, typename OT>
__global__ void __launch_bounds__(kDecThreads)
attn_decode64_kernel(const float* __restrict__ q, const KT* __restrict__ kc, const KT* __restrict__ vc,
                     const int32_t* __restrict__ lens, int len_add, int slot0, OT* __restrict__ out, int nh,
                     int max_ctx, float scale) {
  constexpr int D = 64;
  extern __shared__ __align__(16) float sm[];
  float* s = sm;                      // [max_ctx] scores -> probabilities
  float* part = sm + max_ctx;         // [kDecWarps][D]
  __shared__ float red[kDecWarps];
  const int head = blockIdx.x, b = blockIdx.y, slot = slot0 + b;
  const int T = min(lens[slot] + len_add, max_ctx);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane >> 3, l8 = lane & 7;
  const int H = nh * D;
  const size_t base = ((size_t)slot * nh + head) * (size_t)max_ctx * D;
  const KT* __restrict__ K = kc + base;
  const KT* __restrict__ V = vc + base;

  float qv[8];
  load8(q + (size_t)b * H + head * D + l8 * 8, qv);
#pragma unroll
  for (int i = 0; i < 8; ++i) qv[i] *= scale;

  float lmax = -INFINITY;
  for (int t0 = 0; t0 < T; t0 += 32) {
    const int t = t0 + warp * 4 + sub;
    float dot = 0.f;
    if (t < T) {
      float kv[8];
      load8(K + (size_t)t * D + l8 * 8, kv);
#pragma unroll
      for (int i = 0; i < 8; ++i) dot = fmaf(qv[i], kv[i], dot);
    }
    dot += __shfl_xor_sync(FULL_MASK, dot, 1);
    dot += __shfl_xor_sync(FULL_MASK, dot, 2);
    dot += __shfl_xor_sync(FULL_MASK, dot, 4);
    if (t < T) {
      if (l8 == 0) s[t] = dot;
      lmax = fmaxf(lmax, dot);
    }
  }
  const float m = block_max<kDecWarps>(lmax, red);  // barriers make s[] visible
  float lsum = 0.f;
  for (int t = threadIdx.x; t < T; t += kDecThreads) {
    const float pexp = expf(s[t] - m);
    s[t] = pexp;
    lsum += pexp;
  }
  const float total = block_sum<kDecWarps>(lsum, red);

  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int t0 = 0; t0 < T; t0 += 32) {
    const int t = t0 + warp * 4 + sub;
    if (t < T) {
      float vv[8];
      load8(V + (size_t)t * D + l8 * 8, vv);
      const float pt = s[t];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = fmaf(pt, vv[i], acc[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(FULL_MASK, acc[i], 8);
    acc[i] += __shfl_xor_sync(FULL_MASK, acc[i], 16);
  }
  if (sub == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) part[warp * D + l8 * 8 + i] = acc[i];
  }
  __syncthreads();
  if (threadIdx.x < D) {
    float o = 0.f;
#pragma unroll
    for (int w = 0; w < kDecWarps; ++w) o += part[w * D + threadIdx.x];
    store_out(out + (size_t)b * H + head * D + threadIdx.x, o / total);
  }
}

// ---- decode, any head_dim (tiny-gpt2 has d = 1): one thread per key, scalar loads.
template <typename KT, typename OT>
__global__ void __launch_bounds__(128)
attn_decode_generic_kernel(const float* __restrict__ q, const KT* __restrict__ kc, const KT* __restrict__ vc,
                           const int32_t* __restrict__ lens, int len_add, int slot0, OT* __restrict__ out,
                           int nh, int d, int max_ctx, float scale) {
  extern __shared__ __align__(16) float sm[];
  float* s = sm;  // [max_ctx]
  __shared__ float red[4];
  const int head = blockIdx.x, b = blockIdx.y, slot = slot0 + b;
  const int T = min(lens[slot] + len_add, max_ctx);
  const int H = nh * d;
  const size_t base = ((size_t)slot * nh + head) * (size_t)max_ctx * d;
  const float* qh = q + (size_t)b * H + head * d;
  float lmax = -INFINITY;
  for (int t = threadIdx.x; t < T; t += 128) {
    float dot = 0.f;
    for (int i = 0; i < d; ++i) dot = fmaf(qh[i] * scale, to_f(kc[base + (size_t)t * d + i]), dot);
    s[t] = dot;
    lmax = fmaxf(lmax, dot);
  }
  const float m = block_max<4>(lmax, red);
  float lsum = 0.f;
  for (int t = threadIdx.x; t < T; t += 128) {
    const float pexp = expf(s[t] - m);
    s[t] = pexp;
    lsum += pexp;
  }
  const float total = block_sum<4>(lsum, red);
  for (int i = threadIdx.x; i < d; i += 128) {
    float o = 0.f;
    for (int t = 0; t < T; ++t) o = fmaf(s[t], to_f(vc[base + (size_t)t * d + i]), o);
    store_out(out + (size_t)b * H + head * d + i, o / total);
  }
}

// ---- prefill, head_dim 64: one CTA = 64 queries of one (sequence, head); two threads per
// query (each owns 32 of the 64 dims); K/V tiles of 64 keys staged in shared memory as fp32;
// online softmax in base 2, rescaling once per 16 keys.
constexpr int kPfRow = 72;  // floats per smem row: dims [0,32) at 0, [32,64) at 36 (bank skew)
template <typename AT>
__global__ void __launch_bounds__(128)
attn_prefill64_kernel(const AT* __restrict__ qkv, AT* __restrict__ out, int T, int nh, float scale_log2e) {
  constexpr int D = 64;
  __shared__ __align__(16) float Ks[64 * kPfRow];
  __shared__ __align__(16) float Vs[64 * kPfRow];
  const int qt = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int H = nh * D, ld = 3 * H;
  const int tid = threadIdx.x;
  const int qi_local = tid >> 1, half = tid & 1;
  const int qi = qt * 64 + qi_local;
  const bool q_ok = qi < T;
  const AT* base = qkv + (size_t)b * T * ld;

  float qr[32], o[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    qr[i] = q_ok ? to_f(base[(size_t)qi * ld + head * D + half * 32 + i]) * scale_log2e : 0.f;
    o[i] = 0.f;
  }
  float m = -INFINITY, l = 0.f;

  for (int kt = 0; kt <= qt; ++kt) {
    __syncthreads();
    // stage 64 keys x 64 dims of K and V (zero beyond T)
    for (int idx = tid; idx < 64 * 64; idx += 128) {
      const int j = idx >> 6, dd = idx & 63;
      const int kj = kt * 64 + j;
      float kval = 0.f, vval = 0.f;
      if (kj < T) {
        kval = to_f(base[(size_t)kj * ld + H + head * D + dd]);
        vval = to_f(base[(size_t)kj * ld + 2 * H + head * D + dd]);
      }
      const int col = dd < 32 ? dd : dd + 4;
      Ks[j * kPfRow + col] = kval;
      Vs[j * kPfRow + col] = vval;
    }
    __syncthreads();
    const int hoff = half * 36;
#pragma unroll 1
    for (int j0 = 0; j0 < 64; j0 += 16) {
      float sc[16];
      float cmax = -INFINITY;
#pragma unroll
      for (int jj = 0; jj < 16; ++jj) {
        const float* kr = Ks + (j0 + jj) * kPfRow + hoff;
        float dot = 0.f;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 k4 = *reinterpret_cast<const float4*>(kr + i);
          dot = fmaf(qr[i], k4.x, dot); dot = fmaf(qr[i + 1], k4.y, dot);
          dot = fmaf(qr[i + 2], k4.z, dot); dot = fmaf(qr[i + 3], k4.w, dot);
        }
        dot += __shfl_xor_sync(FULL_MASK, dot, 1);
        const int kj = kt * 64 + j0 + jj;
        const bool ok = q_ok && kj <= qi;  // causal; kj <= qi < T
        sc[jj] = ok ? dot : -INFINITY;
        cmax = fmaxf(cmax, sc[jj]);
      }
      const float m_new = fmaxf(m, cmax);
      if (m_new == -INFINITY) continue;  // nothing visible yet for this query
      const float corr = exp2f(m - m_new);
      l *= corr;
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] *= corr;
#pragma unroll
      for (int jj = 0; jj < 16; ++jj) {
        const float pj = exp2f(sc[jj] - m_new);
        l += pj;
        const float* vr = Vs + (j0 + jj) * kPfRow + hoff;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 v4 = *reinterpret_cast<const float4*>(vr + i);
          o[i] = fmaf(pj, v4.x, o[i]); o[i + 1] = fmaf(pj, v4.y, o[i + 1]);
          o[i + 2] = fmaf(pj, v4.z, o[i + 2]); o[i + 3] = fmaf(pj, v4.w, o[i + 3]);
        }
      }
      m = m_new;
    }
  }
  if (q_ok) {
    const float inv = 1.f / l;
    AT* op = out + ((size_t)b * T + qi) * H + head * D + half * 32;
#pragma unroll
    for (int i = 0; i < 32; ++i) op[i] = from_f<AT>(o[i] * inv);
  }
}

// ---- prefill, any head_dim: one CTA per (query, head, sequence), one thread per key.
template <typename AT>
__global__ void __launch_bounds__(128)
attn_prefill_generic_kernel(const AT* __restrict__ qkv, AT* __restrict__ out, int T, int nh, int d, float scale) {
  extern __shared__ __align__(16) float sm[];
  float* s = sm;  // [T]
  __shared__ float red[4];
  const int qi = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int H = nh * d, ld = 3 * H;
  const AT* base = qkv + (size_t)b * T * ld;
  const AT* qh = base + (size_t)qi * ld + head * d;
  const int Tk = qi + 1;
  float lmax = -INFINITY;
  for (int t = threadIdx.x; t < Tk; t += 128) {
    float dot = 0.f;
    for (int i = 0; i < d; ++i) dot = fmaf(to_f(qh[i]) * scale, to_f(base[(size_t)t * ld + H + head * d + i]), dot);
    s[t] = dot;
    lmax = fmaxf(lmax, dot);
  }
  const float m = block_max<4>(lmax, red);
  float lsum = 0.f;
  for (int t = threadIdx.x; t < Tk; t += 128) {
    const float pexp = expf(s[t] - m);
    s[t] = pexp;
    lsum += pexp;
  }
  const float total = block_sum<4>(lsum, red);
  for (int i = threadIdx.x; i < d; i += 128) {
    float o = 0.f;
    for (int t = 0; t < Tk; ++t) o = fmaf(s[t], to_f(base[(size_t)t * ld + 2 * H + head * d + i]), o);
    out[((size_t)b * T + qi) * H + head * d + i] = from_f<AT>(o / total);
  }
}

template <typename KT, typename OT>
cudaError_t decode_launch(const float* q, const void* kc, const void* vc, const int32_t* lens, int len_add,
                          int slot0, void* out, int B, int nh, int d, int max_ctx, cudaStream_t s) {
  const float scale = 1.0f / sqrtf((float)d);
  dim3 grid(nh, B);
  if (d == 64) {
    const size_t smem = (size_t)(max_ctx + kDecWarps * 64) * sizeof(float);
    attn_decode64_kernel<KT, OT><<<grid, kDecThreads, smem, s>>>(
        q, (const KT*)kc, (const KT*)vc, lens, len_add, slot0, (OT*)out, nh, max_ctx, scale);
  } else {
    const size_t smem = (size_t)max_ctx * sizeof(float);
    attn_decode_generic_kernel<KT, OT><<<grid, 128, smem, s>>>(
        q, (const KT*)kc, (const KT*)vc, lens, len_add, slot0, (OT*)out, nh, d, max_ctx, scale);
  }
  return cudaGetLastError();
}

template <typename AT>
cudaError_t prefill_launch(const void* qkv, void* out, int B, int T, int nh, int d, cudaStream_t s) {
  const float scale = 1.0f / sqrtf((float)d);
  if (d == 64) {
    dim3 grid((T + 63) / 64, nh, B);
    attn_prefill64_kernel<AT><<<grid, 128, 0, s>>>((const AT*)qkv, (AT*)out, T, nh, scale * 1.4426950408889634f);
  } else {
    dim3 grid(T, nh, B);
    attn_prefill_generic_kernel<AT><<<grid, 128, (size_t)T * sizeof(float), s>>>((const AT*)qkv, (AT*)out, T, nh,
                                                                                 d, scale);
  }
  return cudaGetLastError();
}

}  // namespace

cudaError_t init_attn_attrs() { return cudaSuccess; }  // every kernel here fits the 48 KB default

cudaError_t launch_attn_decode(const float* q, const void* kc, const void* vc, const int32_t* lens, int len_add,
                               int slot0, void* out, bool out_bf16, int B, int nh, int d, int max_ctx, int dtype,
                               cudaStream_t s) {
  if (B <= 0 || nh <= 0 || d <= 0 || max_ctx <= 0 || max_ctx > 8192) return cudaErrorInvalidValue;
  if (dtype == GP_BF16) {
    return out_bf16 ? decode_launch<bf16, bf16>(q, kc, vc, lens, len_add, slot0, out, B, nh, d, max_ctx, s)
                    : decode_launch<bf16, float>(q, kc, vc, lens, len_add, slot0, out, B, nh, d, max_ctx, s);
  }
  if (out_bf16) return cudaErrorInvalidValue;
  return decode_launch<float, float>(q, kc, vc, lens, len_add, slot0, out, B, nh, d, max_ctx, s);
}

cudaError_t launch_attn_prefill(const void* qkv, void* out, int B, int T, int nh, int d, int dtype,
                                cudaStream_t s) {
  if (B <= 0 || T <= 0 || nh <= 0 || d <= 0 || T > 8192) return cudaErrorInvalidValue;
  // bf16 storage, head_dim 64: the tensor-core kernel (k_attn_mma.cu); GP_ATTN_SIMT=1 keeps the SIMT kernel for A/B runs
  static const bool force_simt = getenv("GP_ATTN_SIMT") != nullptr;
  if (dtype == GP_BF16 && d == 64 && !force_simt) return launch_attn_prefill64_mma(qkv, out, B, T, nh, s);
  return dtype == GP_BF16 ? prefill_launch<bf16>(qkv, out, B, T, nh, d, s)
                          : prefill_launch<float>(qkv, out, B, T, nh, d, s);
}

}  // namespace gp
