// Causal prefill attention on the tensor cores (bf16 storage, head_dim 64): the restatement of
// [hf] integrations/sdpa_attention.py:65-101 (softmax(q k^T / sqrt(d) + causal) v) for a whole prompt.
// One CTA = 64 queries of one (sequence, head), 4 warps x 16 query rows.  Per 64-key tile: S = Q K^T with
// mma.sync m16n8k16 (Q fragments live in registers for the whole kernel, K fragments via ldmatrix), an online softmax in
// base 2 on the accumulator fragments, then O += P V with P re-used straight from the S accumulators as the A operand
// (the C layout of one mma is the A layout of the next) and V fragments via ldmatrix.trans.  K/V tiles are staged with
// cp.async into a two-stage ring (rows 144 bytes apart: conflict-free ldmatrix).  Replaces the SIMT kernel of round 1
// (21 TFLOP/s); fp32 storage keeps that kernel.
#include "common.cuh"
#include "kernels.h"

namespace gp {
namespace {

constexpr int kD = 64;
constexpr int kTile = 64;             // keys per tile = queries per CTA
constexpr int kRowB = kD * 2 + 16;    // bytes per staged row

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool on) {
  const int bytes = on ? 16 : 0;   // src-size 0: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldsm4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

__global__ void __launch_bounds__(128)
attn_prefill64_mma_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, int T, int nh, float scale_log2e) {
  __shared__ __align__(128) uint8_t Ks[2][kTile * kRowB];
  __shared__ __align__(128) uint8_t Vs[2][kTile * kRowB];
  // heavy (late) query tiles first: the causal triangle makes tile qt cost qt + 1 key tiles
  const int qt = gridDim.x - 1 - blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int H = nh * kD, ld = 3 * H;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
  const bf16* base = qkv + (size_t)b * T * ld;
  const int q0 = qt * kTile + warp * 16;            // first query row of this warp

  // ---- Q fragments (A operand), 4 k-steps of 16 dims
  uint32_t qa[4][4];
  {
    const int r0 = q0 + gid, r1 = r0 + 8;
    const bf16* p0 = base + (size_t)min(r0, T - 1) * ld + head * kD + 2 * tig;
    const bf16* p1 = base + (size_t)min(r1, T - 1) * ld + head * kD + 2 * tig;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      qa[ks][0] = *reinterpret_cast<const uint32_t*>(p0 + 16 * ks);
      qa[ks][1] = *reinterpret_cast<const uint32_t*>(p1 + 16 * ks);
      qa[ks][2] = *reinterpret_cast<const uint32_t*>(p0 + 16 * ks + 8);
      qa[ks][3] = *reinterpret_cast<const uint32_t*>(p1 + 16 * ks + 8);
    }
  }
  // stage one 64-key tile of K and V: 64 rows x 8 chunks of 16 bytes each, 4 + 4 chunks per thread
  auto stage = [&](int kt, int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = tid + i * 128, row = c >> 3, ch = c & 7;
      const int kj = kt * kTile + row;
      const bf16* src = base + (size_t)min(kj, T - 1) * ld + head * kD + ch * 8;
      cp_async16(smem_u32(Ks[buf] + row * kRowB + ch * 16), src + H, kj < T);
      cp_async16(smem_u32(Vs[buf] + row * kRowB + ch * 16), src + 2 * H, kj < T);
    }
    cp_async_commit();
  };

  float o[8][4];
#pragma unroll
  for (int n = 0; n < 8; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;   // rows gid and gid + 8 of this warp's 16

  const int n_kt = qt + 1;
  stage(0, 0);
  for (int kt = 0; kt < n_kt; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < n_kt) { stage(kt + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const uint32_t kbase = smem_u32(Ks[buf]), vbase = smem_u32(Vs[buf]);
    // ---- S = Q K^T (16 x 64 per warp)
    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      // ldmatrix.x4: matrices (keys 8nt.., dims 16ks..+7), (.., +8..+15), (.., 16(ks+1)..), (.., +8)
      const uint32_t addr = kbase + (nt * 8 + (lane & 7)) * kRowB + (lane >> 3) * 16;
#pragma unroll
      for (int ks = 0; ks < 4; ks += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm4(addr + ks * 32, b0, b1, b2, b3);
        mma16816(s[nt], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b0, b1);
        mma16816(s[nt], qa[ks + 1][0], qa[ks + 1][1], qa[ks + 1][2], qa[ks + 1][3], b2, b3);
      }
    }
    // ---- scale, causal mask (only the diagonal tile and the tail beyond T need it), online softmax in base 2
    const int r0 = q0 + gid, r1 = r0 + 8;
    const bool need_mask = (kt == qt) || ((kt + 1) * kTile > T);
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = s[nt][e] * scale_log2e;
        if (need_mask) {
          const int kj = kt * kTile + nt * 8 + 2 * tig + (e & 1);
          const int qi = (e < 2) ? r0 : r1;
          if (kj > qi || kj >= T) v = -INFINITY;
        }
        s[nt][e] = v;
      }
      mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
      mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(FULL_MASK, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(FULL_MASK, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(FULL_MASK, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(FULL_MASK, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    // a row that has seen no key yet keeps m = -inf; exp2(-inf - (-inf)) must not produce NaN
    const float c0 = (mn0 == -INFINITY) ? 1.f : exp2f(m0 - mn0), c1 = (mn1 == -INFINITY) ? 1.f : exp2f(m1 - mn1);
    const float sub0 = (mn0 == -INFINITY) ? 0.f : mn0, sub1 = (mn1 == -INFINITY) ? 0.f : mn1;
    float ps0 = 0.f, ps1 = 0.f;
    uint32_t pa[8][2];   // P as bf16 pairs: [nt][0] = row r0, [nt][1] = row r1
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float p0 = exp2f(s[nt][0] - sub0), p1 = exp2f(s[nt][1] - sub0);
      const float p2 = exp2f(s[nt][2] - sub1), p3 = exp2f(s[nt][3] - sub1);
      ps0 += p0 + p1; ps1 += p2 + p3;
      pa[nt][0] = pack_bf16(p0, p1);
      pa[nt][1] = pack_bf16(p2, p3);
    }
    l0 = l0 * c0 + ps0; l1 = l1 * c1 + ps1;
    m0 = mn0; m1 = mn1;
#pragma unroll
    for (int nd = 0; nd < 8; ++nd) { o[nd][0] *= c0; o[nd][1] *= c0; o[nd][2] *= c1; o[nd][3] *= c1; }
    // ---- O += P V: k-step kk = keys 16kk..+15 (n-tiles 2kk, 2kk+1 of S), output n-tile nd = dims 8nd..+7
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const uint32_t a0 = pa[2 * kk][0], a1 = pa[2 * kk][1], a2 = pa[2 * kk + 1][0], a3 = pa[2 * kk + 1][1];
      // ldmatrix.x4.trans: matrices (keys 16kk..+7, dims 8nd..), (keys 16kk+8.., dims 8nd..), then the same for nd + 1
      const uint32_t addr = vbase + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * kRowB + (lane >> 4) * 16;
#pragma unroll
      for (int nd = 0; nd < 8; nd += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm4_t(addr + nd * 16, b0, b1, b2, b3);
        mma16816(o[nd], a0, a1, a2, a3, b0, b1);
        mma16816(o[nd + 1], a0, a1, a2, a3, b2, b3);
      }
    }
    __syncthreads();   // everyone is done with this buffer before the stage after next overwrites it
  }
  // ---- normalise and store: row sums live in the 4 lanes of a quad
  l0 += __shfl_xor_sync(FULL_MASK, l0, 1); l0 += __shfl_xor_sync(FULL_MASK, l0, 2);
  l1 += __shfl_xor_sync(FULL_MASK, l1, 1); l1 += __shfl_xor_sync(FULL_MASK, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  const int r0 = q0 + gid, r1 = r0 + 8;
#pragma unroll
  for (int nd = 0; nd < 8; ++nd) {
    if (r0 < T) *reinterpret_cast<uint32_t*>(out + ((size_t)b * T + r0) * H + head * kD + nd * 8 + 2 * tig) = pack_bf16(o[nd][0] * i0, o[nd][1] * i0);
    if (r1 < T) *reinterpret_cast<uint32_t*>(out + ((size_t)b * T + r1) * H + head * kD + nd * 8 + 2 * tig) = pack_bf16(o[nd][2] * i1, o[nd][3] * i1);
  }
}

}  // namespace

cudaError_t launch_attn_prefill64_mma(const void* qkv, void* out, int B, int T, int nh, cudaStream_t s) {
  dim3 grid((T + kTile - 1) / kTile, nh, B);
  attn_prefill64_mma_kernel<<<grid, 128, 0, s>>>(reinterpret_cast<const bf16*>(qkv), reinterpret_cast<bf16*>(out), T, nh,
                                                0.125f * 1.4426950408889634f);
  return cudaGetLastError();
}

}  // namespace gp
