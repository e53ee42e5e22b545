// Token selection shared by the stand-alone sampling kernel (k_sample.cu) and the persistent decode kernel
// (k_mega2.cu): see k_sample.cu for the algorithm and its reference (server.py:187-205).
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace gp {

__device__ __forceinline__ uint32_t order_key(float v) {   // larger float <=> larger key (NaN counts as smallest)
  if (v != v) return 0u;
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Philox4x32-10 (Salmon et al., SC'11), first word of the output block
__device__ __forceinline__ uint32_t philox_first(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c0;
}

// NT threads take part: the whole CTA (__syncthreads) or, NAMED, the NT lowest threads of a larger CTA on barrier 1
template <int NT, bool NAMED> __device__ __forceinline__ void sample_sync() {
  if constexpr (NAMED) asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
  else __syncthreads();
}

constexpr int kSampleCand = 448;                                    // survivors of the optional pre-threshold
constexpr int kSampleScratchWords = 2 * kSampleCand + 4 + 4 * kMaxTopK + 32;   // shared memory sample_row needs (32-bit words)

// One draw from a row of V logits (global memory, read through L2).  Called by all NT threads (tid = their index);
// the token is returned in thread 0 only.  Out of line: the persistent kernel sits at its register cap.
//
// floor_value: a value known to have at least top_k elements >= it (-inf: none known).  The persistent kernel passes
// the top_k-th largest of its per-CTA maxima.  With it, one pass keeps the few elements >= floor_value and the
// candidates are ranked among those directly (value descending, token id ascending - the same unique order); the
// radix select below runs only when there is no floor or too many elements survive it (e.g. a constant row).
template <int NT, bool NAMED>
__device__ __noinline__ int sample_row(const float* row, int V, float temperature, int top_k, uint32_t seed_lo, uint32_t seed_hi,
                                       uint32_t position, uint32_t row_id, uint32_t* scratch, int tid, float floor_value) {
  uint32_t* hist = scratch;                               // radix path: 256 bins; floor path: kSampleCand (value, id) pairs
  float* cand_v = reinterpret_cast<float*>(scratch);
  int* cand_i = reinterpret_cast<int*>(scratch + kSampleCand);
  uint32_t* misc = scratch + 2 * kSampleCand;             // prefix, need, n_gt, n_eq
  float* cv = reinterpret_cast<float*>(misc + 4);
  int* ci = reinterpret_cast<int*>(misc + 4 + kMaxTopK);
  float* sv = reinterpret_cast<float*>(misc + 4 + 2 * kMaxTopK);
  int* si = reinterpret_cast<int*>(misc + 4 + 3 * kMaxTopK);
  uint32_t* s_scan = misc + 4 + 4 * kMaxTopK;
  const int k = min(min(top_k, kMaxTopK), V);

  if (!(floor_value > -INFINITY) && k <= NT && V >= 4 * NT) {
    // no floor from the caller: the k-th largest of the NT per-thread maxima is one (they are distinct elements)
    static_assert(NT <= kSampleScratchWords, "the per-thread maxima alias the front of the scratch area");
    float* tmax = reinterpret_cast<float*>(scratch);      // [NT], before anything else uses the scratch area
    float mx = -INFINITY;
    for (int i = tid; i < V; i += NT) mx = fmaxf(mx, __ldcg(row + i));
    tmax[tid] = mx;
    sample_sync<NT, NAMED>();
    int rank = 0;
    for (int j = 0; j < NT; ++j) {
      const float vj = tmax[j];
      rank += (vj > mx || (vj == mx && j < tid)) ? 1 : 0;
    }
    sample_sync<NT, NAMED>();
    if (rank == k - 1) misc[0] = __float_as_uint(mx);
    sample_sync<NT, NAMED>();
    floor_value = __uint_as_float(misc[0]);
    sample_sync<NT, NAMED>();
  }
  bool ranked = false;
  if (floor_value > -INFINITY) {
    if (tid == 0) misc[2] = 0;
    sample_sync<NT, NAMED>();
    auto keep = [&](float v, int i) {
      if (v >= floor_value) {
        const uint32_t at = atomicAdd(&misc[2], 1u);
        if (at < (uint32_t)kSampleCand) { cand_v[at] = v; cand_i[at] = i; }
      }
    };
    // 16-byte loads over the aligned middle of the row, four in flight per thread; scalars at both ends
    const int head = (int)((4u - ((uint32_t)(reinterpret_cast<uintptr_t>(row) >> 2) & 3u)) & 3u);
    const int n4 = (V - min(head, V)) >> 2;
    const float4* row4 = reinterpret_cast<const float4*>(row + head);
    if (tid < head && tid < V) keep(__ldcg(row + tid), tid);
    for (int i = head + 4 * n4 + tid; i < V; i += NT) keep(__ldcg(row + i), i);
    for (int i0 = tid; i0 < n4; i0 += 4 * NT) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = (i0 + u * NT < n4) ? __ldcg(row4 + i0 + u * NT) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = head + 4 * (i0 + u * NT);
        if (v[u].x >= floor_value || v[u].y >= floor_value || v[u].z >= floor_value || v[u].w >= floor_value) {
          keep(v[u].x, i); keep(v[u].y, i + 1); keep(v[u].z, i + 2); keep(v[u].w, i + 3);
        }
      }
    }
    sample_sync<NT, NAMED>();
    const int n = (int)misc[2];
    if (n >= k && n <= kSampleCand) {
      for (int c = tid; c < n; c += NT) {
        const float v = cand_v[c];
        const int id = cand_i[c];
        int rank = 0;
        for (int j = 0; j < n; ++j) {
          const float vj = cand_v[j];
          const int ij = cand_i[j];
          rank += (vj > v || (vj == v && ij < id)) ? 1 : 0;
        }
        if (rank < k) { sv[rank] = v; si[rank] = id; }
      }
      ranked = true;
    }
    sample_sync<NT, NAMED>();
  }
  if (!ranked) {
  // ---- radix select: the key of the k-th largest element, 8 bits per pass from the top
  uint32_t prefix = 0, need = (uint32_t)k;
  for (int shift = 24; shift >= 0; shift -= 8) {
    if (tid < 256) hist[tid] = 0;
    sample_sync<NT, NAMED>();
    const uint32_t hi_mask = (shift == 24) ? 0u : (0xFFFFFFFFu << (shift + 8));
    for (int i = tid; i < V; i += NT) {
      const uint32_t key = order_key(__ldcg(row + i));
      if ((key & hi_mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    sample_sync<NT, NAMED>();
    if (tid == 0) {
      uint32_t acc = 0;
      int bin = 255;
      for (; bin > 0; --bin) {
        if (acc + hist[bin] >= need) break;
        acc += hist[bin];
      }
      misc[0] = prefix | ((uint32_t)bin << shift);
      misc[1] = need - acc;   // how many elements of this bin (and, at the end, of this key) are wanted
    }
    sample_sync<NT, NAMED>();
    prefix = misc[0];
    need = misc[1];
    sample_sync<NT, NAMED>();
  }
  const uint32_t kth = prefix;   // key of the k-th largest; `need` of the elements equal to it are wanted

  // ---- collect: everything above the threshold, plus the `need` lowest-index elements equal to it
  if (tid == 0) { misc[2] = 0; misc[3] = 0; }
  sample_sync<NT, NAMED>();
  uint32_t my_eq = 0;
  for (int i = tid; i < V; i += NT) {
    const float v = __ldcg(row + i);
    const uint32_t key = order_key(v);
    if (key > kth) {
      const uint32_t at = atomicAdd(&misc[2], 1u);
      cv[at] = v; ci[at] = i;
    } else if (key == kth) {
      ++my_eq;
    }
  }
  if (my_eq) atomicAdd(&misc[3], my_eq);
  sample_sync<NT, NAMED>();
  const uint32_t n_gt = misc[2], n_eq = misc[3];   // n_gt = k - need
  if (n_eq == need) {
    // the usual case: no surplus ties, order does not matter (the candidates are sorted below)
    sample_sync<NT, NAMED>();
    if (tid == 0) misc[3] = 0;
    sample_sync<NT, NAMED>();
    for (int i = tid; i < V; i += NT) {
      const float v = __ldcg(row + i);
      if (order_key(v) == kth) {
        const uint32_t at = n_gt + atomicAdd(&misc[3], 1u);
        cv[at] = v; ci[at] = i;
      }
    }
  } else {
    // surplus ties at the threshold: walk the row in index order, block-wide exclusive scan per 1024 ids
    uint32_t taken = 0;   // uniform
    for (int base = 0; base < V && taken < need; base += NT) {
      const int i = base + tid;
      const float v = (i < V) ? __ldcg(row + i) : 0.f;
      const bool tie = (i < V) && order_key(v) == kth;
      const uint32_t bal = __ballot_sync(FULL_MASK, tie);
      if ((tid & 31) == 0) s_scan[tid >> 5] = __popc(bal);
      sample_sync<NT, NAMED>();
      uint32_t before = 0, total = 0;
      for (int w = 0; w < NT / 32; ++w) {
        const uint32_t c = s_scan[w];
        if (w < (tid >> 5)) before += c;
        total += c;
      }
      const uint32_t rank = taken + before + __popc(bal & ((1u << (tid & 31)) - 1u));
      if (tie && rank < need) { cv[n_gt + rank] = v; ci[n_gt + rank] = i; }
      taken += total;
      sample_sync<NT, NAMED>();
    }
  }
  sample_sync<NT, NAMED>();

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
  sample_sync<NT, NAMED>();

  }   // !ranked

  // ---- softmax over the k scaled values and one inverse-CDF draw.  The k weights are computed in parallel; their sum
  // and the scan stay sequential in fp32, candidate order - the oracle's arithmetic.
  const float inv_t = 1.0f / temperature;
  if (tid < k) cv[tid] = expf(sv[tid] * inv_t - sv[0] * inv_t);
  sample_sync<NT, NAMED>();
  int id = 0;
  if (tid == 0) {
    float sum = 0.f;
    for (int j = 0; j < k; ++j) sum += cv[j];
    const uint32_t x = philox_first(position, row_id, 0u, 0u, seed_lo, seed_hi);
    const float u = (float)(x >> 8) * (1.0f / 16777216.0f);
    const float target = u * sum;
    float acc = 0.f;
    int pick = k - 1;
    for (int j = 0; j < k; ++j) {
      acc += cv[j];
      if (target < acc) { pick = j; break; }
    }
    id = si[pick];
  }
  return id;
}

}  // namespace gp
