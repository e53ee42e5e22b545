// Persistent decode kernel for the bandwidth-bound regime (a few sequences per step): ONE
// cooperative launch runs every phase of every layer of a cached decode step - and, for a
// single-stage engine, every step of the greedy loop - so no launch gaps remain and the weight
// stream never stops at a phase boundary.
//
//   * one CTA per SM; warp 16 is the PRODUCER: a single thread walks the (static) sequence of
//     weight row-chunks this CTA will need and streams them HBM -> shared memory with
//     cp.async.bulk (TMA bulk copy, mbarrier complete_tx) into a ring of 24 KB slots.  Weights
//     do not depend on activations, so the producer runs ahead across phases, layers and steps
//     up to the ring depth (~190 KB per SM in flight).
//   * warps 0..15 are CONSUMERS: per phase they stage the activation vector(s) from the L2-
//     resident exchange buffer into shared memory (LayerNorm fused), then each warp takes one
//     weight row of the current slot, dots it with 128-bit shared-memory loads, reduces with
//     shuffles and applies the epilogue (bias / gelu_new / residual / argmax candidate).
//   * phases exchange activations through small fp32 buffers in global memory (L2) and a
//     monotonic grid barrier (release add + acquire poll); 5 barriers per layer.
//   * decode attention is a phase of the same kernel: CTA (b, head) appends the new K/V row to
//     the cache and attends over it.
// Arithmetic per element is the same as the stand-alone GEMV / attention kernels (fp32
// accumulation of storage-rounded weights), restating [hf] modeling_gpt2.py:262-309.
#include "common.cuh"
#include "kernels.h"

namespace gp {
namespace {

constexpr int kNW = 16;                    // consumer warps
constexpr int kConsumers = kNW * 32;       // 512
constexpr int kMegaThreads = kConsumers + 32;
constexpr int kSlotBytes = 24 * 1024;
constexpr int kMaxSlots = 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// sum over the 512 consumer threads; `red` = kNW floats of shared memory
__device__ __forceinline__ float consumer_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  consumer_sync();
  if (l == 0) red[w] = v;
  consumer_sync();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < kNW; ++i) t += red[i];
  return t;
}
__device__ __forceinline__ float consumer_max(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  consumer_sync();
  if (l == 0) red[w] = v;
  consumer_sync();
  float t = red[0];
#pragma unroll
  for (int i = 1; i < kNW; ++i) t = fmaxf(t, red[i]);
  return t;
}

// 8 consecutive storage elements of a shared-memory weight row -> floats
__device__ __forceinline__ void lds8(const bf16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  unpack8(u, v);
}
__device__ __forceinline__ void lds8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

struct Ring {
  uint8_t* slots;
  uint64_t* full;
  uint64_t* empty;
  int n_slots;
  uint32_t it;  // chunks issued (producer) / consumed (each consumer warp) so far
};

__device__ __forceinline__ void rows_of(int cta, int G, int N, int& r0, int& r1) {
  r0 = (int)(((long long)cta * N) / G);
  r1 = (int)(((long long)(cta + 1) * N) / G);
}
__device__ __forceinline__ int rows_per_slot(int rowbytes) {
  int r = kSlotBytes / rowbytes;
  return r > kNW ? kNW : r;
}

// ---- producer: stream this CTA's rows [r0,r1) of W[N,K] through the ring
template <typename WT>
__device__ __forceinline__ void produce_phase(Ring& rg, const void* W, int N, int K, int cta, int G) {
  int r0, r1;
  rows_of(cta, G, N, r0, r1);
  const int rowbytes = K * (int)sizeof(WT);
  const int rps = rows_per_slot(rowbytes);
  const uint8_t* base = reinterpret_cast<const uint8_t*>(W);
  for (int r = r0; r < r1; r += rps) {
    const int n = min(rps, r1 - r);
    const int s = rg.it % rg.n_slots;
    const uint32_t ph = (rg.it / rg.n_slots) & 1;
    mbar_wait(&rg.empty[s], ph ^ 1);
    mbar_expect_tx(&rg.full[s], (uint32_t)(n * rowbytes));
    bulk_g2s(rg.slots + (size_t)s * kSlotBytes, base + (size_t)r * rowbytes, (uint32_t)(n * rowbytes), &rg.full[s]);
    ++rg.it;
  }
}

// ---- consumers: every warp walks the same chunk sequence; warp w owns row w of each chunk.
// pre(row) runs BEFORE the wait on the weight slot (it issues the bias / residual loads so their
// L2 latency overlaps the wait); epi(row, acc[], pre_value) is called by all lanes of the
// owning warp with the reduced sums.
template <typename WT, int BT, typename Pre, typename Epi>
__device__ __forceinline__ void consume_phase(Ring& rg, const float* xs, int N, int K, int cta, int G, Pre pre,
                                              Epi epi) {
  int r0, r1;
  rows_of(cta, G, N, r0, r1);
  const int rowbytes = K * (int)sizeof(WT);
  const int rps = rows_per_slot(rowbytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = r0; r < r1; r += rps) {
    const int n = min(rps, r1 - r);
    const int s = rg.it % rg.n_slots;
    const uint32_t ph = (rg.it / rg.n_slots) & 1;
    float pv = 0.f;
    if (warp < n) pv = pre(r + warp);
    mbar_wait(&rg.full[s], ph);
    if (warp < n) {
      const WT* wrow = reinterpret_cast<const WT*>(rg.slots + (size_t)s * kSlotBytes + (size_t)warp * rowbytes);
      float acc[BT];
#pragma unroll
      for (int b = 0; b < BT; ++b) acc[b] = 0.f;
#pragma unroll 4
      for (int k0 = lane * 8; k0 < K; k0 += 256) {
        float w[8];
        lds8(wrow + k0, w);
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
#pragma unroll
      for (int b = 0; b < BT; ++b) acc[b] = warp_sum(acc[b]);
      epi(r + warp, acc, pv);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&rg.empty[s]);
    ++rg.it;
  }
}

// Stage x[B,K] (fp32, written by other CTAs in the previous phase: read through L2) into shared
// memory.  Plain copy variant.
template <int BT>
__device__ __forceinline__ void stage_copy(float* xs, const float* x, int ldx, int B, int K) {
  const int tid = threadIdx.x;
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    if (b < B) {
      for (int k = tid * 4; k < K; k += kConsumers * 4)
        *reinterpret_cast<float4*>(xs + b * K + k) = __ldcg(reinterpret_cast<const float4*>(x + (size_t)b * ldx + k));
    } else {
      for (int k = tid; k < K; k += kConsumers) xs[b * K + k] = 0.f;
    }
  }
  consumer_sync();
}

// LayerNorm variant (K = H <= 2048): the row lives in registers, mean and the two-pass biased
// variance each cost one shared-memory exchange; three barriers in all.
constexpr int kLnPer = 4;  // elements per thread: H <= 4 * 512
template <int BT>
__device__ __forceinline__ void stage_ln(float* xs, float* red, const float* x, int ldx, int B, int K,
                                         const float* __restrict__ g, const float* __restrict__ bta, float eps) {
  const int tid = threadIdx.x, w = tid >> 5, l = tid & 31;
  float v[BT][kLnPer], gg[kLnPer], bb[kLnPer];
#pragma unroll
  for (int i = 0; i < kLnPer; ++i) {
    const int k = tid + i * kConsumers;
    gg[i] = (k < K) ? g[k] : 0.f;
    bb[i] = (k < K) ? bta[k] : 0.f;
#pragma unroll
    for (int b = 0; b < BT; ++b) v[b][i] = (b < B && k < K) ? __ldcg(x + (size_t)b * ldx + k) : 0.f;
  }
  float* redA = red;             // [BT][kNW]
  float* redB = red + BT * kNW;  // [BT][kNW]
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kLnPer; ++i) s += v[b][i];
    s = warp_sum(s);
    if (l == 0) redA[b * kNW + w] = s;
  }
  consumer_sync();
  float mean[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kNW; ++i) t += redA[b * kNW + i];
    mean[b] = t / (float)K;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kLnPer; ++i) {
      const int k = tid + i * kConsumers;
      const float dlt = (k < K) ? v[b][i] - mean[b] : 0.f;
      q += dlt * dlt;
    }
    q = warp_sum(q);
    if (l == 0) redB[b * kNW + w] = q;
  }
  consumer_sync();
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kNW; ++i) t += redB[b * kNW + i];
    const float rstd = rsqrtf(t / (float)K + eps);
#pragma unroll
    for (int i = 0; i < kLnPer; ++i) {
      const int k = tid + i * kConsumers;
      if (k < K) xs[b * K + k] = (b < B) ? (v[b][i] - mean[b]) * rstd * gg[i] + bb[i] : 0.f;
    }
  }
  consumer_sync();
}

struct Best {
  float v;
  int i;
};
__device__ __forceinline__ void best_merge(Best& a, float v, int i) {
  if (v > a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; }
}

// raw (still packed) 8-element chunk of a K/V cache row, so a prefetch costs few registers
template <typename WT> struct Raw8;
template <> struct Raw8<bf16> {
  uint4 u;
  __device__ __forceinline__ void load(const bf16* p) { u = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void get(float (&v)[8]) const { unpack8(u, v); }
};
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = *reinterpret_cast<const float4*>(p);
    b = *reinterpret_cast<const float4*>(p + 4);
  }
  __device__ __forceinline__ void get(float (&v)[8]) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};

// online-softmax state of one 8-lane group: running max, denominator, 8 output dims per lane
struct SoftState {
  float m, l, acc[8];
};
__device__ __forceinline__ void soft_merge(SoftState& a, float mo, float lo, const float (&ao)[8]) {
  const float mn = fmaxf(a.m, mo);
  if (mn == -INFINITY) return;
  const float c1 = expf(a.m - mn), c2 = expf(mo - mn);
  a.l = a.l * c1 + lo * c2;
#pragma unroll
  for (int i = 0; i < 8; ++i) a.acc[i] = a.acc[i] * c1 + ao[i] * c2;
  a.m = mn;
}

template <typename WT, int BT>
__global__ void __launch_bounds__(kMegaThreads, 1) decode_mega_kernel(const MegaParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int G = gridDim.x, cta = blockIdx.x;
  const int H = p.H, F = 4 * p.H, nh = p.nh, B = p.B;
  constexpr int D = 64;
  // ---- shared memory carve-up (mirrored by fixed_smem() on the host)
  uint8_t* ring_base = smem;
  float* xs = reinterpret_cast<float*>(smem + (size_t)p.n_slots * kSlotBytes);  // [BT][F]
  float* part = xs + (size_t)BT * F;                                            // [kNW][D]
  float* pm = part + kNW * D;                                                   // [kNW]
  float* pl = pm + kNW;                                                         // [kNW]
  float* red = pl + kNW;                                                        // [2][BT][kNW]
  float* sbv = red + 2 * BT * kNW;                                              // [kNW][BT]
  int* sbi = reinterpret_cast<int*>(sbv + kNW * BT);                            // [kNW][BT]
  int* s_ids = sbi + kNW * BT;                                                  // [8]
  int* s_pos = s_ids + 8;                                                       // [8]
  uint64_t* full = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(s_pos + 8) + 15) & ~(uintptr_t)15);
  uint64_t* empty = full + kMaxSlots;
  MegaLayer* s_layers = reinterpret_cast<MegaLayer*>(empty + kMaxSlots);        // [L]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < p.n_slots; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kNW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  {  // layer table -> shared memory (pointer chasing through global memory would sit on every phase)
    const int n_words = p.L * (int)(sizeof(MegaLayer) / 8);
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(p.layers);
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(s_layers);
    for (int i = tid; i < n_words; i += kMegaThreads) dst[i] = src[i];
  }
  __syncthreads();
  Ring rg{ring_base, full, empty, p.n_slots, 0u};

  if (warp == kNW) {
    // =============================== PRODUCER ===============================
    if (lane == 0) {
      for (int step = 0; step < p.n_steps; ++step) {
        for (int li = 0; li < p.L; ++li) {
          const MegaLayer& w = s_layers[li];
          produce_phase<WT>(rg, w.attn_w, 3 * H, H, cta, G);
          produce_phase<WT>(rg, w.proj_w, H, H, cta, G);
          produce_phase<WT>(rg, w.fc_w, F, H, cta, G);
          produce_phase<WT>(rg, w.mproj_w, H, F, cta, G);
        }
        if (p.has_head) produce_phase<WT>(rg, p.wte, p.V, H, cta, G);
      }
    }
    return;
  }

  // ================================= CONSUMERS =================================
  unsigned bar_target = 0;
  // grid barrier, split so that independent work (prefetches) can sit between arrive and wait
  auto grid_arrive = [&]() {
    consumer_sync();
    if (tid == 0) {
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.bar) : "memory");
      bar_target += (unsigned)G;
    }
  };
  auto grid_wait = [&]() {
    if (tid == 0) {
      while (ld_acquire(p.bar) < bar_target) {}
    }
    consumer_sync();
  };
  auto grid_sync = [&]() {
    grid_arrive();
    grid_wait();
  };
  // optional phase trace (GP_MEGA_TRACE): SM clock of CTA 0 at every phase boundary of the last step
  int n_stamp = 0;
  auto stamp = [&](int step) {
    if (p.trace != nullptr && cta == 0 && tid == 0 && step == p.n_steps - 1 && n_stamp < p.trace_cap)
      p.trace[n_stamp++] = clock64();
  };
  if (tid < 8) s_pos[tid] = (tid < B) ? p.lens[p.slot0 + tid] : 0;
  if (cta == 0 && tid == 0 && p.step_clock != nullptr) p.step_clock[p.n_steps] = globaltimer();  // start stamp
  consumer_sync();

  const int sub = lane >> 3, l8 = lane & 7;
  const int grp = warp * 4 + sub;                    // 0..63: key rows t = grp + 64 j belong to this 8-lane group
  const unsigned grp_mask = 0xFFu << (sub * 8);

  for (int step = 0; step < p.n_steps; ++step) {
    const float* hcur = p.hidden_in;  // residual stream feeding the first layer
    if (p.has_embed) {
      // ---- embed: h[b,:] = wte[id] + wpe[pos]; every CTA writes its column slice
      int c0, c1;
      rows_of(cta, G, H, c0, c1);
      const WT* wte = reinterpret_cast<const WT*>(p.wte);
      const WT* wpe = reinterpret_cast<const WT*>(p.wpe);
      for (int i = tid; i < B * (c1 - c0); i += kConsumers) {
        const int b = i / (c1 - c0), col = c0 + i % (c1 - c0);
        int id = (step == 0) ? __ldcg(p.ids + b) : s_ids[b];   // later steps: this CTA's own argmax result
        id = min(max(id, 0), p.V - 1);
        const int pos = min(s_pos[b], p.n_pos - 1);
        p.hbuf[(size_t)b * H + col] = to_f(wte[(size_t)id * H + col]) + to_f(wpe[(size_t)pos * H + col]);
      }
      grid_sync();
      hcur = p.hbuf;
    }
    stamp(step);
    for (int li = 0; li < p.L; ++li) {
      const MegaLayer& w = s_layers[li];
      const bool last = (li == p.L - 1);
      // ---- P1: ln_1 + c_attn -> qkv[b, 3H]
      stage_ln<BT>(xs, red, hcur, H, B, H, w.ln1_g, w.ln1_b, p.eps);
      stamp(step);
      consume_phase<WT, BT>(
          rg, xs, 3 * H, H, cta, G, [&](int n) { return (lane < B) ? w.attn_b[n] : 0.f; },
          [&](int n, const float(&acc)[BT], float pv) {
            if (lane < B) {
              float v = 0.f;
#pragma unroll
              for (int b = 0; b < BT; ++b) if (lane == b) v = acc[b];
              p.qkv[(size_t)lane * 3 * H + n] = v + pv;
            }
          });
      stamp(step);
      grid_arrive();
      // ---- P2: attention for (b, head) tasks.  Old K/V rows do not depend on this step, so the
      // first two rows of every 8-lane group are fetched while the barrier completes.
      const bool has_task = cta < B * nh;
      Raw8<WT> pk0, pk1, pv0, pv1;
      if (has_task) {
        const int b = cta / nh, head = cta - b * nh;
        const size_t base = ((size_t)(p.slot0 + b) * nh + head) * (size_t)p.max_ctx * D;
        const WT* Kc = reinterpret_cast<const WT*>(w.kc) + base;
        const WT* Vc = reinterpret_cast<const WT*>(w.vc) + base;
        const int pos = s_pos[b];
        if (grp < pos) {
          pk0.load(Kc + (size_t)grp * D + l8 * 8);
          pv0.load(Vc + (size_t)grp * D + l8 * 8);
        }
        if (grp + 64 < pos) {
          pk1.load(Kc + (size_t)(grp + 64) * D + l8 * 8);
          pv1.load(Vc + (size_t)(grp + 64) * D + l8 * 8);
        }
      }
      grid_wait();
      stamp(step);
      for (int task = cta; task < B * nh; task += G) {
        const int b = task / nh, head = task - b * nh;
        const size_t base = ((size_t)(p.slot0 + b) * nh + head) * (size_t)p.max_ctx * D;
        WT* Kc = reinterpret_cast<WT*>(w.kc) + base;
        WT* Vc = reinterpret_cast<WT*>(w.vc) + base;
        const int pos = min(s_pos[b], p.max_ctx - 1);
        const float* qrow = p.qkv + (size_t)b * 3 * H + head * D;
        float qv[8];
        {
          const float4 a = __ldcg(reinterpret_cast<const float4*>(qrow + l8 * 8));
          const float4 c4 = __ldcg(reinterpret_cast<const float4*>(qrow + l8 * 8 + 4));
          const float sc = 0.125f;  // 64^-0.5
          qv[0] = a.x * sc; qv[1] = a.y * sc; qv[2] = a.z * sc; qv[3] = a.w * sc;
          qv[4] = c4.x * sc; qv[5] = c4.y * sc; qv[6] = c4.z * sc; qv[7] = c4.w * sc;
        }
        // the new row goes into the cache for later steps (this step reads it from qkv directly)
        if (tid >= 64 && tid < 64 + 2 * D) {
          const int i = tid - 64;
          const float val = __ldcg(qrow + H + i);   // i < 64: k, else v (contiguous thirds: +H, +2H)
          if (i < D) Kc[(size_t)pos * D + i] = from_f<WT>(val);
          else Vc[(size_t)pos * D + (i - D)] = from_f<WT>(__ldcg(qrow + 2 * H + (i - D)));
        }
        SoftState st;
        st.m = -INFINITY; st.l = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) st.acc[i] = 0.f;
        for (int j = 0; grp + 64 * j <= pos; ++j) {
          const int t = grp + 64 * j;
          float kk[8], vv[8];
          if (t == pos) {
            // storage-rounded like every other cached row, so a step is independent of where the
            // row is read from
            const float4 k0 = __ldcg(reinterpret_cast<const float4*>(qrow + H + l8 * 8));
            const float4 k1 = __ldcg(reinterpret_cast<const float4*>(qrow + H + l8 * 8 + 4));
            const float4 v0 = __ldcg(reinterpret_cast<const float4*>(qrow + 2 * H + l8 * 8));
            const float4 v1 = __ldcg(reinterpret_cast<const float4*>(qrow + 2 * H + l8 * 8 + 4));
            const float kf[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
            const float vf[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) { kk[i] = to_f(from_f<WT>(kf[i])); vv[i] = to_f(from_f<WT>(vf[i])); }
          } else if (j == 0 && task == cta) {
            pk0.get(kk);
            pv0.get(vv);
          } else if (j == 1 && task == cta) {
            pk1.get(kk);
            pv1.get(vv);
          } else {
            load8(Kc + (size_t)t * D + l8 * 8, kk);
            load8(Vc + (size_t)t * D + l8 * 8, vv);
          }
          float dot = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) dot = fmaf(qv[i], kk[i], dot);
          dot += __shfl_xor_sync(grp_mask, dot, 1);
          dot += __shfl_xor_sync(grp_mask, dot, 2);
          dot += __shfl_xor_sync(grp_mask, dot, 4);
          const float mn = fmaxf(st.m, dot);
          const float c1 = expf(st.m - mn), pe = expf(dot - mn);
          st.l = st.l * c1 + pe;
#pragma unroll
          for (int i = 0; i < 8; ++i) st.acc[i] = st.acc[i] * c1 + pe * vv[i];
          st.m = mn;
        }
        __syncwarp();
        // merge the four groups of the warp, then the 16 warps through shared memory
#pragma unroll
        for (int o = 8; o <= 16; o <<= 1) {
          const float mo = __shfl_xor_sync(FULL_MASK, st.m, o), lo = __shfl_xor_sync(FULL_MASK, st.l, o);
          float ao[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) ao[i] = __shfl_xor_sync(FULL_MASK, st.acc[i], o);
          soft_merge(st, mo, lo, ao);
        }
        if (sub == 0) {
#pragma unroll
          for (int i = 0; i < 8; ++i) part[warp * D + l8 * 8 + i] = st.acc[i];
          if (l8 == 0) { pm[warp] = st.m; pl[warp] = st.l; }
        }
        consumer_sync();
        if (tid < D) {
          float M = pm[0];
#pragma unroll
          for (int ww = 1; ww < kNW; ++ww) M = fmaxf(M, pm[ww]);
          float Lsum = 0.f, o = 0.f;
#pragma unroll
          for (int ww = 0; ww < kNW; ++ww) {
            const float c = (pm[ww] == -INFINITY) ? 0.f : expf(pm[ww] - M);
            Lsum += pl[ww] * c;
            o += part[ww * D + tid] * c;
          }
          p.attn[(size_t)b * H + head * D + tid] = o / Lsum;
        }
        consumer_sync();
      }
      stamp(step);
      grid_sync();
      stamp(step);
      // ---- P3: attention c_proj + residual -> h2
      stage_copy<BT>(xs, p.attn, H, B, H);
      stamp(step);
      consume_phase<WT, BT>(
          rg, xs, H, H, cta, G,
          [&](int n) { return (lane < B) ? w.proj_b[n] + __ldcg(hcur + (size_t)lane * H + n) : 0.f; },
          [&](int n, const float(&acc)[BT], float pv) {
            if (lane < B) {
              float v = 0.f;
#pragma unroll
              for (int b = 0; b < BT; ++b) if (lane == b) v = acc[b];
              p.h2[(size_t)lane * H + n] = v + pv;
            }
          });
      stamp(step);
      grid_sync();
      stamp(step);
      // ---- P4: ln_2 + c_fc + gelu_new -> act
      stage_ln<BT>(xs, red, p.h2, H, B, H, w.ln2_g, w.ln2_b, p.eps);
      stamp(step);
      consume_phase<WT, BT>(
          rg, xs, F, H, cta, G, [&](int n) { return (lane < B) ? w.fc_b[n] : 0.f; },
          [&](int n, const float(&acc)[BT], float pv) {
            if (lane < B) {
              float v = 0.f;
#pragma unroll
              for (int b = 0; b < BT; ++b) if (lane == b) v = acc[b];
              p.act[(size_t)lane * F + n] = gelu_new(v + pv);
            }
          });
      stamp(step);
      grid_sync();
      stamp(step);
      // ---- P5: mlp c_proj + residual -> h (or the stage output)
      float* dst = (last && !p.has_head) ? p.hidden_out : p.hbuf;
      stage_copy<BT>(xs, p.act, F, B, F);
      stamp(step);
      consume_phase<WT, BT>(
          rg, xs, H, F, cta, G,
          [&](int n) { return (lane < B) ? w.mproj_b[n] + __ldcg(p.h2 + (size_t)lane * H + n) : 0.f; },
          [&](int n, const float(&acc)[BT], float pv) {
            if (lane < B) {
              float v = 0.f;
#pragma unroll
              for (int b = 0; b < BT; ++b) if (lane == b) v = acc[b];
              dst[(size_t)lane * H + n] = v + pv;
            }
          });
      stamp(step);
      grid_sync();
      stamp(step);
      hcur = p.hbuf;
    }
    if (p.has_head) {
      // ---- ln_f + lm_head + argmax
      stage_ln<BT>(xs, red, p.hbuf, H, B, H, p.lnf_g, p.lnf_b, p.eps);
      stamp(step);
      Best best[BT];
#pragma unroll
      for (int b = 0; b < BT; ++b) { best[b].v = -INFINITY; best[b].i = 0x7fffffff; }
      consume_phase<WT, BT>(
          rg, xs, p.V, H, cta, G, [&](int) { return 0.f; },
          [&](int n, const float(&acc)[BT], float) {
#pragma unroll
            for (int b = 0; b < BT; ++b) {
              best_merge(best[b], acc[b], n);
              if (p.logits_out != nullptr && lane == 0 && b < B) p.logits_out[(size_t)b * p.V + n] = acc[b];
            }
          });
      if (lane == 0) {
#pragma unroll
        for (int b = 0; b < BT; ++b) { sbv[warp * BT + b] = best[b].v; sbi[warp * BT + b] = best[b].i; }
      }
      consumer_sync();
      if (tid < BT) {
        Best r{sbv[tid], sbi[tid]};
        for (int ww = 1; ww < kNW; ++ww) best_merge(r, sbv[ww * BT + tid], sbi[ww * BT + tid]);
        p.part_v[(size_t)cta * kGemvMaxBatch + tid] = r.v;
        p.part_i[(size_t)cta * kGemvMaxBatch + tid] = r.i;
      }
      stamp(step);
      grid_sync();
      stamp(step);
      // every CTA reduces the G partials (fixed order + index tie-break => deterministic), so
      // every CTA knows the next token without another exchange
      for (int b = warp; b < B; b += kNW) {
        Best r{-INFINITY, 0x7fffffff};
        for (int g = lane; g < G; g += 32)
          best_merge(r, __ldcg(p.part_v + (size_t)g * kGemvMaxBatch + b), __ldcg(p.part_i + (size_t)g * kGemvMaxBatch + b));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(FULL_MASK, r.v, o);
          const int oi = __shfl_xor_sync(FULL_MASK, r.i, o);
          best_merge(r, ov, oi);
        }
        if (lane == 0) {
          s_ids[b] = r.i;
          if (cta == 0) {
            p.ids_out[b] = r.i;
            if (p.seq != nullptr) p.seq[(size_t)b * p.seq_stride + s_pos[b] + 1] = r.i;
          }
        }
      }
      consumer_sync();
    }
    // ---- bookkeeping: the cache grew by one position
    if (cta == 0 && tid < B) p.lens[p.slot0 + tid] = s_pos[tid] + 1;
    if (cta == 0 && tid == 0 && p.step_clock != nullptr) p.step_clock[step] = globaltimer();
    consumer_sync();
    if (tid < B) s_pos[tid] += 1;
    consumer_sync();
    // no barrier between steps: the argmax partial buffers are next written ~5L barriers from now
    // and every CTA carries the token id in its own shared memory
  }
}

template <typename WT, int BT>
cudaError_t launch_t(const MegaParams& p, int grid, size_t smem, cudaStream_t s) {
  void* args[] = {const_cast<MegaParams*>(&p)};
  return cudaLaunchCooperativeKernel((const void*)decode_mega_kernel<WT, BT>, dim3(grid), dim3(kMegaThreads), args,
                                     smem, s);
}
template <typename WT, int BT> cudaError_t set_attr() {
  return cudaFuncSetAttribute(decode_mega_kernel<WT, BT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
}

int padded(int B) { return B <= 1 ? 1 : (B <= 2 ? 2 : 4); }

size_t fixed_smem(int BT, int H, int L) {
  return (size_t)BT * 4 * H * 4 /*xs*/ + kNW * 64 * 4 /*part*/ + 2 * kNW * 4 /*pm,pl*/ + (size_t)2 * BT * kNW * 4 /*red*/ +
         (size_t)kNW * BT * 8 /*sbv,sbi*/ + 64 /*ids,pos*/ + 16 + 2 * kMaxSlots * 8 + (size_t)L * sizeof(MegaLayer) + 128;
}

}  // namespace

bool mega_supported(int B, int H, int nh, int L, int dtype) {
  if (B < 1 || B > 4 || H % 8 != 0 || nh <= 0 || H / nh != 64 || H % nh != 0 || H > kLnPer * kConsumers) return false;
  const size_t fixed = fixed_smem(padded(B), H, L);
  if (fixed + 2 * (size_t)kSlotBytes > 227 * 1024) return false;
  // one row of the widest matrix (K = 4H) must fit a slot
  return (size_t)4 * H * (dtype == GP_BF16 ? 2 : 4) <= (size_t)kSlotBytes;
}

cudaError_t init_mega_attrs() {
  cudaError_t e;
#define GP_SET(WT, BT) if ((e = set_attr<WT, BT>()) != cudaSuccess) return e;
  GP_SET(bf16, 1) GP_SET(bf16, 2) GP_SET(bf16, 4) GP_SET(float, 1) GP_SET(float, 2) GP_SET(float, 4)
#undef GP_SET
  return cudaSuccess;
}

cudaError_t launch_decode_mega(MegaParams p, int dtype, int num_sms, cudaStream_t s) {
  if (!mega_supported(p.B, p.H, p.nh, p.L, dtype) || p.n_steps < 1 || p.bar == nullptr) return cudaErrorInvalidValue;
  const int BT = padded(p.B);
  const size_t fixed = fixed_smem(BT, p.H, p.L);
  int n_slots = (int)((227 * 1024 - fixed) / kSlotBytes);
  if (n_slots > kMaxSlots) n_slots = kMaxSlots;
  p.n_slots = n_slots;
  const size_t smem = fixed + (size_t)n_slots * kSlotBytes;
  cudaError_t e = cudaMemsetAsync(p.bar, 0, sizeof(unsigned), s);
  if (e != cudaSuccess) return e;
  const int grid = num_sms;
  if (dtype == GP_BF16) {
    if (BT == 1) return launch_t<bf16, 1>(p, grid, smem, s);
    if (BT == 2) return launch_t<bf16, 2>(p, grid, smem, s);
    return launch_t<bf16, 4>(p, grid, smem, s);
  }
  if (BT == 1) return launch_t<float, 1>(p, grid, smem, s);
  if (BT == 2) return launch_t<float, 2>(p, grid, smem, s);
  return launch_t<float, 4>(p, grid, smem, s);
}

}  // namespace gp
