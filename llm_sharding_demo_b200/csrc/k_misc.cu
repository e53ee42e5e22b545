// Embedding gather, stand-alone LayerNorm, the SIMT tile GEMM (generic shapes / fp32 mode),
// KV-cache scatter kernels and the weight re-layout kernel.
#include "common.cuh"
#include "kernels.h"

namespace gp {
namespace {

// K1: out[m,:] = wte[ids[m]] + wpe[pos(m)] (server.py:79-82).  pos(m) is, in order of
// precedence: pos[m]; lens[slot0+m] (decode: position = cache length); pos0 + m % T (prefill).
template <typename WT>
__global__ void __launch_bounds__(256)
embed_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ pos, int pos0, int T,
             const int32_t* __restrict__ lens, int slot0, const WT* __restrict__ wte,
             const WT* __restrict__ wpe, float* __restrict__ out, int H, int V, int n_pos) {
  const int m = blockIdx.x;
  int id = ids[m];
  id = min(max(id, 0), V - 1);  // host validates; clamp so a bad id can never fault
  int p = pos ? pos[m] : (lens ? lens[slot0 + m] : pos0 + (m % T));
  p = min(max(p, 0), n_pos - 1);
  const WT* te = wte + (size_t)id * H;
  const WT* pe = wpe + (size_t)p * H;
  float* o = out + (size_t)m * H;
  if (H % 8 == 0) {
    for (int k = threadIdx.x * 8; k < H; k += 256 * 8) {
      float a[8], b[8];
      load8(te + k, a);
      load8(pe + k, b);
      *reinterpret_cast<float4*>(o + k) = make_float4(a[0] + b[0], a[1] + b[1], a[2] + b[2], a[3] + b[3]);
      *reinterpret_cast<float4*>(o + k + 4) = make_float4(a[4] + b[4], a[5] + b[5], a[6] + b[6], a[7] + b[7]);
    }
  } else {
    for (int k = threadIdx.x; k < H; k += 256) o[k] = to_f(te[k]) + to_f(pe[k]);
  }
}

// K2: one warp per row, the row held in registers (H <= 2048, H % 4 == 0), 128-bit loads,
// shuffle reductions, two-pass variance.
constexpr int kLnMaxV = 16;  // float4 per lane -> H <= 2048
template <typename OT>
__global__ void __launch_bounds__(256)
layernorm_warp_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ g,
                      const float* __restrict__ bta, OT* __restrict__ out, int M, int H, float eps) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31;
  const int nv = H >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * ldx);
  float4 v[kLnMaxV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kLnMaxV; ++i) {
    const int c = i * 32 + lane;
    if (c < nv) {
      v[i] = xr[c];
      s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
  }
  const float mean = warp_sum(s) / (float)H;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kLnMaxV; ++i) {
    const int c = i * 32 + lane;
    if (c < nv) {
      const float a = v[i].x - mean, b = v[i].y - mean, cc = v[i].z - mean, d = v[i].w - mean;
      q += a * a + b * b + cc * cc + d * d;
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)H + eps);
  OT* orow = out + (size_t)row * H;
#pragma unroll
  for (int i = 0; i < kLnMaxV; ++i) {
    const int c = i * 32 + lane;
    if (c < nv) {
      const float4 gg = reinterpret_cast<const float4*>(g)[c];
      const float4 bb = reinterpret_cast<const float4*>(bta)[c];
      const float r0 = (v[i].x - mean) * rstd * gg.x + bb.x, r1 = (v[i].y - mean) * rstd * gg.y + bb.y;
      const float r2 = (v[i].z - mean) * rstd * gg.z + bb.z, r3 = (v[i].w - mean) * rstd * gg.w + bb.w;
      if (sizeof(OT) == 4) {
        reinterpret_cast<float4*>(orow)[c] = make_float4(r0, r1, r2, r3);
      } else {
        __nv_bfloat162 lo = __floats2bfloat162_rn(r0, r1), hi = __floats2bfloat162_rn(r2, r3);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        reinterpret_cast<uint2*>(orow)[c] = pk;
      }
    }
  }
}

// generic LayerNorm (odd / huge H): one warp per row, three passes over global memory
template <typename OT>
__global__ void __launch_bounds__(256)
layernorm_generic_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ g,
                         const float* __restrict__ bta, OT* __restrict__ out, int M, int H, float eps) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + (size_t)row * ldx;
  float s = 0.f;
  for (int k = lane; k < H; k += 32) s += xr[k];
  const float mean = warp_sum(s) / (float)H;
  float q = 0.f;
  for (int k = lane; k < H; k += 32) {
    const float dlt = xr[k] - mean;
    q += dlt * dlt;
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)H + eps);
  for (int k = lane; k < H; k += 32) out[(size_t)row * H + k] = from_f<OT>((xr[k] - mean) * rstd * g[k] + bta[k]);
}

// SIMT tile GEMM: out[M,N] = epi(A[M,K] * Wt[N,K]^T + bias) (+ residual).  64x64 tile, K step
// 16, 256 threads x (4x4) outputs.  Any M, N, K; the fallback for shapes tcgen05 does not take
// and the only GEMM of fp32 mode.
template <typename AT, typename OT>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const AT* __restrict__ A, int lda, const AT* __restrict__ Wt, const float* __restrict__ bias,
                 const float* __restrict__ residual, int ldr, OT* __restrict__ out, int ldo, int M, int N, int K,
                 uint32_t flags) {
  __shared__ float As[16][64 + 4];
  __shared__ float Ws[16][64 + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int lr = tid >> 2, lk = (tid & 3) * 4;  // each thread stages 4 consecutive k of one row
  for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = k0 + lk + i;
      const int am = m0 + lr, wn = n0 + lr;
      As[lk + i][lr] = (am < M && k < K) ? to_f(A[(size_t)am * lda + k]) : 0.f;
      Ws[lk + i][lr] = (wn < N && k < K) ? to_f(Wt[(size_t)wn * K + k]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 w4 = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w}, w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j];
      if (bias) v += bias[n];
      if (flags & GP_EPI_GELU) v = gelu_new(v);
      if (residual) v += residual[(size_t)m * ldr + n];
      out[(size_t)m * ldo + n] = from_f<OT>(v);
    }
  }
}

// prefill: copy the k and v thirds of qkv[B*T, 3H] into the cache at positions [0,T)
template <typename AT>
__global__ void __launch_bounds__(256)
kv_scatter_kernel(const AT* __restrict__ qkv, AT* __restrict__ kc, AT* __restrict__ vc, int T, int slot0, int H,
                  int nh, int d, int max_ctx) {
  const int m = blockIdx.x;  // token index b*T + t
  const int b = m / T, t = m - b * T;
  const AT* row = qkv + (size_t)m * 3 * H;
  for (int c = threadIdx.x; c < H; c += 256) {
    const int head = c / d, dd = c - head * d;
    const size_t dst = (((size_t)(slot0 + b) * nh + head) * max_ctx + t) * d + dd;
    kc[dst] = row[H + c];
    vc[dst] = row[2 * H + c];
  }
}

// batched decode on the GEMM path: qkv[B,3H] -> q fp32 + cache rows at position lens[slot]
template <typename AT>
__global__ void __launch_bounds__(256)
qkv_split_decode_kernel(const AT* __restrict__ qkv, float* __restrict__ q, AT* __restrict__ kc,
                        AT* __restrict__ vc, const int32_t* __restrict__ lens, int slot0, int H, int nh, int d,
                        int max_ctx) {
  const int b = blockIdx.x, slot = slot0 + b;
  const int pos = lens[slot];
  const AT* row = qkv + (size_t)b * 3 * H;
  for (int c = threadIdx.x; c < H; c += 256) {
    const int head = c / d, dd = c - head * d;
    const size_t dst = (((size_t)slot * nh + head) * max_ctx + pos) * d + dd;
    q[(size_t)b * H + c] = to_f(row[c]);
    if (pos < max_ctx) {   // a full slot is refused on the host; never write past the slot's rows
      kc[dst] = row[H + c];
      vc[dst] = row[2 * H + c];
    }
  }
}

__global__ void set_lens_kernel(int32_t* lens, int slot0, int B, int value, int add) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) lens[slot0 + i] = add ? lens[slot0 + i] + value : value;
}

// src fp32 [R,C] -> dst [C,R] (transpose) through a 32x33 shared tile, or a straight convert
template <typename WT>
__global__ void __launch_bounds__(256)
convert_transpose_kernel(const float* __restrict__ src, WT* __restrict__ dst, int R, int C) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int i = ty; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + tx;
    tile[i][tx] = (r < R && c < C) ? src[(size_t)r * C + c] : 0.f;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + tx;
    if (c < C && r < R) dst[(size_t)c * R + r] = from_f<WT>(tile[tx][i]);
  }
}
template <typename WT>
__global__ void __launch_bounds__(256)
convert_kernel(const float* __restrict__ src, WT* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    dst[i] = from_f<WT>(src[i]);
}

// greedy selection on materialised logits (batched decode through the GEMM path): one CTA per
// sequence, lowest index wins ties; optionally appends to seq and advances the cache length
__global__ void __launch_bounds__(512)
argmax_rows_kernel(const float* __restrict__ logits, int ldl, int V, int32_t* __restrict__ ids_out, int32_t* lens,
                   int slot0, int32_t* seq, int seq_stride) {
  __shared__ float sv[16];
  __shared__ int si[16];
  const int b = blockIdx.x;
  const float* row = logits + (size_t)b * ldl;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < V; i += 512) {
    const float v = row[i];
    if (v > bv) { bv = v; bi = i; }   // ascending i per thread: first maximum kept
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(FULL_MASK, bv, o);
    const int oi = __shfl_xor_sync(FULL_MASK, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 16; ++w)
      if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
    ids_out[b] = bi;
    if (lens != nullptr) {
      const int old_len = lens[slot0 + b];
      if (seq != nullptr) seq[(size_t)b * seq_stride + old_len + 1] = bi;
      lens[slot0 + b] = old_len + 1;
    }
  }
}

__global__ void __launch_bounds__(256)
gather_last_kernel(const float* __restrict__ src, float* __restrict__ dst, int T, int H) {
  const int b = blockIdx.x;
  const float* s = src + ((size_t)b * T + (T - 1)) * H;
  for (int k = threadIdx.x; k < H; k += 256) dst[(size_t)b * H + k] = s[k];
}

}  // namespace

cudaError_t launch_embed(const int32_t* ids, const int32_t* pos, int pos0, int T, const int32_t* lens, int slot0,
                         const void* wte, const void* wpe, float* out, int M, int H, int V, int n_pos, int dtype,
                         cudaStream_t s) {
  if (M <= 0) return cudaErrorInvalidValue;
  if (T <= 0) T = 1;
  if (dtype == GP_BF16)
    embed_kernel<bf16><<<M, 256, 0, s>>>(ids, pos, pos0, T, lens, slot0, (const bf16*)wte, (const bf16*)wpe, out,
                                         H, V, n_pos);
  else
    embed_kernel<float><<<M, 256, 0, s>>>(ids, pos, pos0, T, lens, slot0, (const float*)wte, (const float*)wpe,
                                          out, H, V, n_pos);
  return cudaGetLastError();
}

cudaError_t launch_layernorm(const float* x, int ldx, const float* g, const float* b, void* out, bool out_bf16,
                             int M, int H, float eps, cudaStream_t s) {
  if (M <= 0 || H <= 0) return cudaErrorInvalidValue;
  const int grid = (M + 7) / 8;
  const bool fast = (H % 4 == 0) && (ldx % 4 == 0) && H <= kLnMaxV * 128;
  if (fast) {
    if (out_bf16) layernorm_warp_kernel<bf16><<<grid, 256, 0, s>>>(x, ldx, g, b, (bf16*)out, M, H, eps);
    else layernorm_warp_kernel<float><<<grid, 256, 0, s>>>(x, ldx, g, b, (float*)out, M, H, eps);
  } else {
    if (out_bf16) layernorm_generic_kernel<bf16><<<grid, 256, 0, s>>>(x, ldx, g, b, (bf16*)out, M, H, eps);
    else layernorm_generic_kernel<float><<<grid, 256, 0, s>>>(x, ldx, g, b, (float*)out, M, H, eps);
  }
  return cudaGetLastError();
}

cudaError_t launch_gemm_simt(const void* A, int lda, const void* Wt, const float* bias, const float* residual,
                             int ldr, void* out, int ldo, int M, int N, int K, uint32_t flags, int dtype,
                             cudaStream_t s) {
  if (M <= 0 || N <= 0 || K <= 0) return cudaErrorInvalidValue;
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  const bool obf = flags & GP_OUT_BF16;
  if (dtype == GP_BF16) {
    if (obf) gemm_simt_kernel<bf16, bf16><<<grid, 256, 0, s>>>((const bf16*)A, lda, (const bf16*)Wt, bias, residual,
                                                               ldr, (bf16*)out, ldo, M, N, K, flags);
    else gemm_simt_kernel<bf16, float><<<grid, 256, 0, s>>>((const bf16*)A, lda, (const bf16*)Wt, bias, residual,
                                                            ldr, (float*)out, ldo, M, N, K, flags);
  } else {
    if (obf) return cudaErrorInvalidValue;
    gemm_simt_kernel<float, float><<<grid, 256, 0, s>>>((const float*)A, lda, (const float*)Wt, bias, residual, ldr,
                                                        (float*)out, ldo, M, N, K, flags);
  }
  return cudaGetLastError();
}

cudaError_t launch_kv_scatter(const void* qkv, void* kc, void* vc, int B, int T, int slot0, int H, int nh, int d,
                              int max_ctx, int dtype, cudaStream_t s) {
  if (dtype == GP_BF16)
    kv_scatter_kernel<bf16><<<B * T, 256, 0, s>>>((const bf16*)qkv, (bf16*)kc, (bf16*)vc, T, slot0, H, nh, d, max_ctx);
  else
    kv_scatter_kernel<float><<<B * T, 256, 0, s>>>((const float*)qkv, (float*)kc, (float*)vc, T, slot0, H, nh, d,
                                                   max_ctx);
  return cudaGetLastError();
}

cudaError_t launch_qkv_split_decode(const void* qkv, float* q, void* kc, void* vc, const int32_t* lens, int B,
                                    int slot0, int H, int nh, int d, int max_ctx, int dtype, cudaStream_t s) {
  if (dtype == GP_BF16)
    qkv_split_decode_kernel<bf16><<<B, 256, 0, s>>>((const bf16*)qkv, q, (bf16*)kc, (bf16*)vc, lens, slot0, H, nh, d,
                                                    max_ctx);
  else
    qkv_split_decode_kernel<float><<<B, 256, 0, s>>>((const float*)qkv, q, (float*)kc, (float*)vc, lens, slot0, H,
                                                     nh, d, max_ctx);
  return cudaGetLastError();
}

cudaError_t launch_set_lens(int32_t* lens, int slot0, int B, int value, int add, cudaStream_t s) {
  set_lens_kernel<<<(B + 127) / 128, 128, 0, s>>>(lens, slot0, B, value, add);
  return cudaGetLastError();
}

cudaError_t launch_convert(const float* src, void* dst, int R, int C, bool transpose, int dtype, cudaStream_t s) {
  if (transpose) {
    dim3 grid((C + 31) / 32, (R + 31) / 32);
    if (dtype == GP_BF16) convert_transpose_kernel<bf16><<<grid, 256, 0, s>>>(src, (bf16*)dst, R, C);
    else convert_transpose_kernel<float><<<grid, 256, 0, s>>>(src, (float*)dst, R, C);
  } else {
    const size_t n = (size_t)R * C;
    int grid = (int)((n + 255) / 256);
    if (grid > device_sm_count() * 16) grid = device_sm_count() * 16;
    if (dtype == GP_BF16) convert_kernel<bf16><<<grid, 256, 0, s>>>(src, (bf16*)dst, n);
    else convert_kernel<float><<<grid, 256, 0, s>>>(src, (float*)dst, n);
  }
  return cudaGetLastError();
}

cudaError_t launch_argmax_rows(const float* logits, int ldl, int B, int V, int32_t* ids_out, int32_t* lens, int slot0,
                               int32_t* seq, int seq_stride, cudaStream_t s) {
  argmax_rows_kernel<<<B, 512, 0, s>>>(logits, ldl, V, ids_out, lens, slot0, seq, seq_stride);
  return cudaGetLastError();
}

cudaError_t launch_gather_last(const float* src, float* dst, int B, int T, int H, cudaStream_t s) {
  gather_last_kernel<<<B, 256, 0, s>>>(src, dst, T, H);
  return cudaGetLastError();
}

int device_sm_count() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return cached;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) { cached = n; cached_dev = dev; }
  }
  return cached;
}

}  // namespace gp
