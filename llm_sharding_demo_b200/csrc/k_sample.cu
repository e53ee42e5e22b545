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
#include "sample.cuh"

namespace gp {
namespace {

constexpr int kSampleThreads = 1024;

__global__ void __launch_bounds__(kSampleThreads)
sample_topk_kernel(const float* __restrict__ logits, int ldl, int V, float temperature, int top_k, uint32_t seed_lo,
                   uint32_t seed_hi, int row0, int fixed_pos, int32_t* __restrict__ ids_out, int32_t* lens, int slot0,
                   int32_t* seq, int seq_stride) {
  __shared__ uint32_t scratch[kSampleScratchWords];
  const int b = blockIdx.x, tid = threadIdx.x;
  int pos = fixed_pos, old_len = 0;
  if (lens != nullptr) {
    old_len = lens[slot0 + b];
    pos = old_len + 1;
  }
  const int id = sample_row<kSampleThreads, false>(logits + (size_t)b * ldl, V, temperature, top_k, seed_lo, seed_hi, (uint32_t)pos,
                                                   (uint32_t)(row0 + b), scratch, tid, -INFINITY);
  if (tid == 0) {
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
