// Token selection of the reference coordinator (server.py:187-205): logits / temperature -> top-k
// (sorted) -> softmax over the k values -> one multinomial draw -> the token id.  One CTA per
// sequence; the logits row (V = 50257 floats, L2-resident right after the lm_head GEMM) is read
// five times: four 8-bit radix-select passes find the k-th largest value, one pass collects the
// candidates.  Ties are broken by the lower token id, so the candidate set is unique.
//
// The reference draws from torch's unseeded global generator, which cannot be reproduced; here the
// draw is a pure function of (seed, sequence row, position of the new token): Philox4x32-10 with key
// = seed and counter = (position, row, 0, 0), first output word -> u = (x >> 8) * 2^-24.  A request
// therefore samples the same continuation whatever the batch composition, launch geometry or call
// history, and the CPU oracle can restate it exactly.
#include "common.cuh"
#include "kernels.h"

namespace gp {
namespace {

constexpr int kSampleThreads = 1024;

__device__ __forceinline__ uint32_t order_key(float v) {   // larger float <=> larger key (NaN counts as smallest)
  if (v != v) return 0u;
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) { return __umulhi(a, b); }

// Philox4x32-10 (Salmon et al., SC'11), first word of the output block
__device__ __forceinline__ uint32_t philox_first(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c0;
}

__global__ void __launch_bounds__(kSampleThreads)
sample_topk_kernel(const float* __restrict__ logits, int ldl, int V, float temperature, int top_k, uint32_t seed_lo,
                   uint32_t seed_hi, int row0, int fixed_pos, int32_t* __restrict__ ids_out, int32_t* lens, int slot0,
                   int32_t* seq, int seq_stride) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t s_prefix, s_need, s_n_gt, s_n_eq;
  __shared__ float cv[kMaxTopK];
  __shared__ int ci[kMaxTopK];
  __shared__ float sv[kMaxTopK];
  __shared__ int si[kMaxTopK];
  __shared__ uint32_t s_scan[kSampleThreads / 32];
  const int b = blockIdx.x, tid = threadIdx.x;
  const float* row = logits + (size_t)b * ldl;
  const int k = min(min(top_k, kMaxTopK), V);

  // ---- radix select: the key of the k-th largest element, 8 bits per pass from the top
  uint32_t prefix = 0, need = (uint32_t)k;
  for (int shift = 24; shift >= 0; shift -= 8) {
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const uint32_t hi_mask = (shift == 24) ? 0u : (0xFFFFFFFFu << (shift + 8));
    for (int i = tid; i < V; i += kSampleThreads) {
      const uint32_t key = order_key(row[i]);
      if ((key & hi_mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      uint32_t acc = 0;
      int bin = 255;
      for (; bin > 0; --bin) {
        if (acc + hist[bin] >= need) break;
        acc += hist[bin];
      }
      s_prefix = prefix | ((uint32_t)bin << shift);
      s_need = need - acc;   // how many elements of this bin (and, at the end, of this key) are wanted
    }
    __syncthreads();
    prefix = s_prefix;
    need = s_need;
    __syncthreads();
  }
  const uint32_t kth = prefix;   // key of the k-th largest; `need` of the elements equal to it are wanted

  // ---- collect: everything above the threshold, plus the `need` lowest-index elements equal to it
  if (tid == 0) { s_n_gt = 0; s_n_eq = 0; }
  __syncthreads();
  uint32_t my_eq = 0;
  for (int i = tid; i < V; i += kSampleThreads) {
    const float v = row[i];
    const uint32_t key = order_key(v);
    if (key > kth) {
      const uint32_t at = atomicAdd(&s_n_gt, 1u);
      cv[at] = v; ci[at] = i;
    } else if (key == kth) {
      ++my_eq;
    }
  }
  if (my_eq) atomicAdd(&s_n_eq, my_eq);
  __syncthreads();
  const uint32_t n_gt = s_n_gt, n_eq = s_n_eq;   // n_gt = k - need
  if (n_eq == need) {
    // the usual case: no surplus ties, order does not matter (the candidates are sorted below)
    __syncthreads();
    if (tid == 0) s_n_eq = 0;
    __syncthreads();
    for (int i = tid; i < V; i += kSampleThreads) {
      const float v = row[i];
      if (order_key(v) == kth) {
        const uint32_t at = n_gt + atomicAdd(&s_n_eq, 1u);
        cv[at] = v; ci[at] = i;
      }
    }
  } else {
    // surplus ties at the threshold: walk the row in index order, block-wide exclusive scan per 1024 ids
    uint32_t taken = 0;   // uniform
    for (int base = 0; base < V && taken < need; base += kSampleThreads) {
      const int i = base + tid;
      const float v = (i < V) ? row[i] : 0.f;
      const bool tie = (i < V) && order_key(v) == kth;
      const uint32_t bal = __ballot_sync(FULL_MASK, tie);
      if ((tid & 31) == 0) s_scan[tid >> 5] = __popc(bal);
      __syncthreads();
      uint32_t before = 0, total = 0;
      for (int w = 0; w < kSampleThreads / 32; ++w) {
        const uint32_t c = s_scan[w];
        if (w < (tid >> 5)) before += c;
        total += c;
      }
      const uint32_t rank = taken + before + __popc(bal & ((1u << (tid & 31)) - 1u));
      if (tie && rank < need) { cv[n_gt + rank] = v; ci[n_gt + rank] = i; }
      taken += total;
      __syncthreads();
    }
  }
  __syncthreads();

  // ---- sort the k candidates: value descending, then token id ascending (rank by counting)
  if (tid < k) {
    const float v = cv[tid];
    const int id = ci[tid];
    int rank = 0;
    for (int j = 0; j < k; ++j) {
      const float vj = cv[j];
      const int ij = ci[j];
      rank += (vj > v || (vj == v && ij < id)) ? 1 : 0;
    }
    sv[rank] = v; si[rank] = id;
  }
  __syncthreads();

  // ---- softmax over the k scaled values and one inverse-CDF draw (sequential fp32 sums: the oracle's order)
  if (tid == 0) {
    int pos = fixed_pos;
    int old_len = 0;
    if (lens != nullptr) {
      old_len = lens[slot0 + b];
      pos = old_len + 1;
    }
    const float inv_t = 1.0f / temperature;
    const float m = sv[0] * inv_t;
    float sum = 0.f;
    for (int j = 0; j < k; ++j) sum += expf(sv[j] * inv_t - m);
    const uint32_t x = philox_first((uint32_t)pos, (uint32_t)(row0 + b), 0u, 0u, seed_lo, seed_hi);
    const float u = (float)(x >> 8) * (1.0f / 16777216.0f);
    const float target = u * sum;
    float acc = 0.f;
    int pick = k - 1;
    for (int j = 0; j < k; ++j) {
      acc += expf(sv[j] * inv_t - m);
      if (target < acc) { pick = j; break; }
    }
    const int id = si[pick];
    ids_out[b] = id;
    if (lens != nullptr) {
      if (seq != nullptr) seq[(size_t)b * seq_stride + old_len + 1] = id;
      lens[slot0 + b] = old_len + 1;
    }
  }
}

}  // namespace

cudaError_t launch_sample_topk(const float* logits, int ldl, int B, int V, float temperature, int top_k, uint64_t seed,
                               int row0, int fixed_pos, int32_t* ids_out, int32_t* lens, int slot0, int32_t* seq,
                               int seq_stride, cudaStream_t s) {
  if (B <= 0 || V <= 0 || top_k <= 0 || top_k > kMaxTopK || !(temperature > 0.f)) return cudaErrorInvalidValue;
  sample_topk_kernel<<<B, kSampleThreads, 0, s>>>(logits, ldl, V, temperature, top_k, (uint32_t)seed,
                                                  (uint32_t)(seed >> 32), row0, fixed_pos, ids_out, lens, slot0, seq,
                                                  seq_stride);
  return cudaGetLastError();
}

}  // namespace gp
