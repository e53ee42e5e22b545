// Persistent decode kernel for the bandwidth-bound regime (a few sequences per step): ONE
// cooperative launch runs every phase of every layer of a cached decode step - and, for a
// single-stage engine, every step of the greedy loop.  At batch 1 a GPT-2 124M step is ~60
// dependent phases of a few hundred KB each, so the cost is latency, not bandwidth; the design
// removes every latency that is not a data dependency:
//
//   * one CTA per SM; warp 15 is the PRODUCER: a single thread walks the (static) sequence of
//     weight row-chunks this CTA will need and streams them HBM -> shared memory with
//     cp.async.bulk (TMA bulk copy, mbarrier complete_tx) into a ring of six 32 KB slots, the
//     matching bias slice riding in the same transaction.  Weights do not depend on activations,
//     so the producer runs ahead across phases, layers and steps (~190 KB per SM in flight).
//   * warps 0..14 are CONSUMERS: per phase they collect the activation vector into shared memory
//     (LayerNorm fused, row held in registers), then each warp dots its one or two weight rows of the
//     current slot (conflict-free shared loads, shuffle reduction) and applies the epilogue.
//   * phases exchange activations through "flag-in-data" buffers in L2: every fp32 value travels
//     in one 64-bit word together with the epoch of the phase that produced it, so a reader can
//     verify each element by itself - no grid barrier, no memory fence, and (with one exception, the
//     argmax exchange at the end of a step) no "ready" counter:
//     a reader polls exactly the words it needs, every outstanding word of a thread in one batch of
//     predicated loads per round.  (A counter gate in front of the polling - one warp watching eight
//     "published" counters per buffer - was the first design; it only protected against an L2 hot spot
//     that clamped out-of-range addresses had created, and cost one extra L2 round trip per exchange.)
//     Reuse of a buffer one layer later is ordered by the data dependencies themselves.
//   * the residual stream slice a CTA owns stays in its shared memory across phases and layers;
//     decode attention is a phase of the same kernel: CTA (b, head) prefetches the cached K/V rows
//     while it waits for q, appends the new row and runs an online softmax per 8-lane group.
// Arithmetic per element is that of the stand-alone GEMV / attention kernels (fp32 accumulation
// of storage-rounded weights), restating [hf] modeling_gpt2.py:262-309.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace gp {
namespace {

typedef unsigned long long u64;

constexpr int kNW = 15;                    // consumer warps (+1 producer warp = 512 threads, 128 registers each)
constexpr int kConsumers = kNW * 32;       // 480
constexpr int kMegaThreads = kConsumers + 32;
constexpr int kGroups = kNW * 4;           // 8-lane groups of the attention phase
constexpr int kSlotData = 32 * 1024;       // weight rows of one chunk
constexpr int kMaxRows = 24;               // rows per chunk (a warp may own two)
constexpr int kSlotBias = 128;             // fp32 bias window of the chunk (<= 24 rows + alignment slack)
constexpr int kSlotBytes = kSlotData + kSlotBias;
constexpr int kMaxSlots = 6;
constexpr int kOwnMax = 16;                // residual rows one CTA owns (ceil(H/G)+1 <= 16)
constexpr int kLnPer = 4;                  // LayerNorm elements per thread: H <= 4 * 480
constexpr int kCopyPer = 16;               // plain staging elements per thread: K <= 16 * 480
constexpr int kFlagStride = 256;           // flag words per exchange buffer (>= CTAs, >= B*nh)

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
// The weights are a 250 MB-per-token stream through a 126 MB L2: left at normal priority they evict what a token
// re-reads (above all the KV cache rows) before it comes round again.  The stream is therefore marked evict-first
// and what the consumers load stays resident: +10 % at a context of 576, +3 % at 160.  The host sets the hint only
// when the stage's weights exceed what L2 could keep (MegaParams::l2_stream_hint; GP_MEGA_L2HINT overrides).
__device__ __forceinline__ uint64_t l2_stream_policy(bool evict_first) {
  uint64_t pol;
  if (evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory"); }
__device__ __forceinline__ u64 globaltimer() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// ---- flag-in-data exchange: {epoch:32 | fp32 bits:32} in one naturally aligned 64-bit word.
// A 64-bit access is single-copy atomic, so a reader that sees the epoch also sees the value.
__device__ __forceinline__ void ll_store(u64* p, float v, uint32_t epoch) {
  const u64 w = ((u64)epoch << 32) | (u64)__float_as_uint(v);
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ void ll_store_int(u64* p, int v, uint32_t epoch) {
  const u64 w = ((u64)epoch << 32) | (u64)(uint32_t)v;
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ u64 ll_load(const u64* p) {
  u64 w;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  return w;
}
__device__ __forceinline__ void red_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// predicated form (no branch, no access when `on` is false): lanes beyond the end of a row must not fall back
// to a clamped address - 148 CTAs x several warps re-reading ONE word turn it into an L2 hot spot
__device__ __forceinline__ u64 ll_load_if(const u64* p, bool on) {
  u64 w;
  asm volatile(
      "{\n"
      ".reg .pred q;\n"
      "setp.ne.u32 q, %2, 0;\n"
      "mov.u64 %0, 0;\n"
      "@q ld.relaxed.gpu.global.u64 %0, [%1];\n"
      "}\n"
      : "=l"(w)
      : "l"(p), "r"((uint32_t)on)
      : "memory");
  return w;
}
// system scope: words another GPU writes into this GPU's inbox over NVLink (and our writes into a peer's)
__device__ __forceinline__ void ll_store_sys(u64* p, uint32_t bits, uint32_t epoch) {
  const u64 w = ((u64)epoch << 32) | (u64)bits;
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ uint32_t ll_wait_sys(const u64* p, uint32_t epoch) {
  u64 w;
  do {
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  } while ((uint32_t)(w >> 32) != epoch);
  return (uint32_t)w;
}
__device__ __forceinline__ uint32_t ll_wait_bits(const u64* p, uint32_t epoch) {
  u64 w = ll_load(p);
  while ((uint32_t)(w >> 32) != epoch) w = ll_load(p);
  return (uint32_t)w;
}
// four consecutive weights of a row (8 bytes of bf16 / 16 bytes of fp32)
__device__ __forceinline__ void lds4(const bf16* p, float (&v)[4]) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  v[0] = bf16lo(u.x); v[1] = bf16hi(u.x); v[2] = bf16lo(u.y); v[3] = bf16hi(u.y);
}
__device__ __forceinline__ void lds4(const float* p, float (&v)[4]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}

struct Ring {
  uint8_t* slots;
  uint64_t* full;
  uint64_t* empty;
  int n_slots;
  int slot;       // next slot to fill (producer) / to read (each consumer warp)
  uint32_t par;   // its phase parity: flips every time `slot` wraps
  __device__ __forceinline__ void advance() {
    if (++slot == n_slots) { slot = 0; par ^= 1u; }
  }
};

// What a CTA needs to know about one kind of phase (0 qkv, 2 proj, 3 fc, 4 proj2, 5 head; 1 = attention
// has no matrix).  Computed once per launch into shared memory: the divisions behind it would otherwise
// sit on the dependent path of every one of the ~60 phases of a step.
struct KindInfo {
  int r0, r1;   // rows of the [N,K] matrix this CTA owns
  int rps;      // rows per ring slot
  int K, N;
  int pad[3];
};

// CTA `cta` of G owns rows [r0, r1) of an N-row matrix.  32-bit arithmetic on purpose (64-bit division
// is a software routine and this sits on every phase): (cta+1)*N < 2^31 for G <= 256, N <= 8M.
__device__ __forceinline__ void rows_of(int cta, int G, int N, int& r0, int& r1) {
  r0 = (int)(((unsigned)cta * (unsigned)N) / (unsigned)G);
  r1 = (int)(((unsigned)(cta + 1) * (unsigned)N) / (unsigned)G);
}
__device__ __forceinline__ int rows_per_slot(int rowbytes) {
  int r = kSlotData / rowbytes;
  return r > kMaxRows ? kMaxRows : r;
}

// ---- producer: stream this CTA's rows [r0,r1) of W[N,K] (+ their biases) through the ring
template <typename WT>
__device__ __forceinline__ void produce_phase(Ring& rg, const void* W, const float* bias, const KindInfo& ki, uint64_t policy) {
  const int rowbytes = ki.K * (int)sizeof(WT);
  const int rps = ki.rps, r1 = ki.r1;
  const uint8_t* base = reinterpret_cast<const uint8_t*>(W);
  for (int r = ki.r0; r < r1; r += rps) {
    const int n = min(rps, r1 - r);
    const int s = rg.slot;
    uint8_t* slot = rg.slots + (size_t)s * kSlotBytes;
    mbar_wait(&rg.empty[s], rg.par ^ 1u);
    // bias window: 16-byte aligned superset [r & ~3, (r + n + 3) & ~3) of the chunk's rows
    const int b0 = r & ~3, b1 = (r + n + 3) & ~3;
    const uint32_t bias_bytes = bias ? (uint32_t)(b1 - b0) * 4u : 0u;
    mbar_expect_tx(&rg.full[s], (uint32_t)(n * rowbytes) + bias_bytes);
    bulk_g2s(slot, base + (size_t)r * rowbytes, (uint32_t)(n * rowbytes), &rg.full[s], policy);
    if (bias) bulk_g2s(slot + kSlotData, bias + b0, bias_bytes, &rg.full[s], policy);
    rg.advance();
  }
}

// ======================================================================================
// Consumer side.  A step is ~60 phases that differ only in parameters; each of the five phases of a
// layer gets its own inlined, constant-propagated copy of the stage / rows routines (~120 KB of SASS
// in all).  Both alternatives were measured and lost: calling shared routines puts the parameter block
// on the local-memory stack (with the 227 KB shared-memory carve-out there is no L1 behind it: 2.5x
// slower), and ONE generic inlined copy driven by run-time parameters (64 KB) costs more per phase in
// selects and branches than it saves in instruction fetches (ncu: 2.5 % "no instruction" stalls).
// ======================================================================================
enum { SRC_LL = 0, SRC_PLAIN = 1, SRC_EMBED = 2 };
enum { EPI_QKV = 0, KIND_ATTN = 1, EPI_PROJ = 2, EPI_FC = 3, EPI_PROJ2 = 4, EPI_HEAD = 5 };   // = kind of the phase

struct Phase {
  // ---- input vector x[B,K]
  int src_kind; const u64* src_ll; const float* src_plain; int K; uint32_t src_epoch;
  const float* ln_g; const float* ln_b;              // LayerNorm affine, or NULL for a plain copy
  // ---- rows
  int N; int has_bias; int mode;
  u64* dst_ll; float* dst_plain; uint32_t dst_epoch;
  int dst_sys;                                       // dst_ll is a peer GPU's inbox: system-scope stores
  const u64* sentinel; uint32_t sentinel_epoch;      // inbox input: one word thread 0 watches before all poll
};

// TR: the phase trace is compiled in only for the kernels launched under GP_MEGA_TRACE - a stamp that is merely
// predicated off still costs ~85 cycles of dependent latency, and there are a dozen per phase.
template <typename WT, bool TR> struct Ctx {
  Ring rg;
  float *xs, *red, *hloc, *h2loc, *sbv; int* sbi; int *s_ids, *s_pos;
  const WT *wte, *wpe;
  int B, H, V, n_pos, G, cta, c0; float eps; float* logits_out;
  long long* trace; int trace_cap; int n_stamp; bool tracing;   // debug: per-CTA clock64 stamps (thread 0)
  // tag = kind * 8 + event (kind of the phase, 7 = step boundary), kept in the top byte
  __device__ __forceinline__ void stamp(int tag) {
    if constexpr (TR) {
      if (tracing && threadIdx.x == 0 && n_stamp < trace_cap)
        trace[n_stamp++] = (long long)(((u64)clock64() & 0x00FFFFFFFFFFFFFFull) | ((u64)tag << 56));
    }
  }
};

template <typename WT, bool TR>
__device__ __forceinline__ bool fetch_x(const Ctx<WT, TR>& cx, const Phase& ph, int b, int k, float& v, bool live = true) {
  if (ph.src_kind == SRC_LL) {
    u64 w;
    if (ph.sentinel != nullptr) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(ph.src_ll + (size_t)b * ph.K + k) : "memory");
    else w = ll_load_if(ph.src_ll + (size_t)b * ph.K + k, live);   // (b, k) are clamped by the caller; dead lanes skip the access
    v = __uint_as_float((uint32_t)w);
    return (uint32_t)(w >> 32) == ph.src_epoch;
  }
  if (ph.src_kind == SRC_PLAIN) {
    v = __ldcg(ph.src_plain + (size_t)b * ph.K + k);
    return true;
  }
  v = to_f(cx.wte[(size_t)cx.s_ids[b] * cx.H + k]) + to_f(cx.wpe[(size_t)min(cx.s_pos[b], cx.n_pos - 1) * cx.H + k]);
  return true;
}

// barrier (previous phase fully done in this CTA: its shared-memory operands may be overwritten); a stage whose
// input comes from the previous pipeline stage first lets one thread watch a single inbox word
__device__ __forceinline__ void stage_enter(const u64* sentinel = nullptr, uint32_t sentinel_epoch = 0) {
  consumer_sync();
  if (sentinel != nullptr) {
    if (threadIdx.x == 0) ll_wait_sys(sentinel, sentinel_epoch);
    consumer_sync();
  }
}

// One array of flag-in-data words -> shared memory.  All PER loads of a thread are issued back to back
// (predicated inside the asm, so no branch separates them) and only then checked; words whose epoch is
// still old are re-read.
template <int PER>
__device__ __forceinline__ void copy_ll(float* xrow, const u64* src, int k0, int K, uint32_t epoch) {
  const int tid = threadIdx.x;
  u64 w[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) w[i] = ll_load_if(src + k0 + tid + i * kConsumers, k0 + tid + i * kConsumers < K);
  unsigned pending = 0;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int k = k0 + tid + i * kConsumers;
    if (k < K) {
      if ((uint32_t)(w[i] >> 32) == epoch) xrow[k] = __uint_as_float((uint32_t)w[i]);
      else pending |= 1u << i;
    }
  }
  while (pending) {   // what is still missing is re-read as one batch of (predicated) loads per round
#pragma unroll
    for (int i = 0; i < PER; ++i) w[i] = ll_load_if(src + k0 + tid + i * kConsumers, (pending >> i) & 1u);
#pragma unroll
    for (int i = 0; i < PER; ++i)
      if (((pending >> i) & 1u) && (uint32_t)(w[i] >> 32) == epoch) {
        xrow[k0 + tid + i * kConsumers] = __uint_as_float((uint32_t)w[i]);
        pending &= ~(1u << i);
      }
  }
}

// Collect x[B,K] into shared memory; elements not yet visible are re-read until they are.
// LayerNorm variant (K = H <= 1920): the row lives in registers; mean and the two-pass biased
// variance ([hf] modeling_gpt2.py:252,254,505) each cost one shared-memory exchange.  Fully
// unrolled on purpose: every load of a thread is in flight before the first one is checked.
template <typename WT, int BT, bool TR>
__device__ __forceinline__ void stage_x(Ctx<WT, TR>& cx, const Phase& ph) {
  const int tid = threadIdx.x, w = tid >> 5, l = tid & 31, K = ph.K, B = cx.B;
  float* xs = cx.xs;
  if (ph.ln_g == nullptr) {
    // plain copy: always from a flag-in-data buffer (attention output, GELU output)
    stage_enter();
    cx.stamp(ph.mode * 8 + 0);
    // the B rows are adjacent both in the exchange buffer and in xs: one array of B * K words, so that the loads
    // of ALL rows are in flight together (row after row cost one L2 round trip per row)
    const int KT = B * K;
    const int per = (KT + kConsumers - 1) / kConsumers;
    if (per <= 2) {
      copy_ll<2>(xs, ph.src_ll, 0, KT, ph.src_epoch);
    } else if (per <= 7) {
      copy_ll<7>(xs, ph.src_ll, 0, KT, ph.src_epoch);       // K = 4 * 768: 6.4 words per thread
    } else {
      for (int k0 = 0; k0 < KT; k0 += 9 * kConsumers) copy_ll<9>(xs, ph.src_ll, k0, KT, ph.src_epoch);
    }
    for (int k = KT + tid; k < BT * K; k += kConsumers) xs[k] = 0.f;   // padding rows of a partial batch
    cx.stamp(ph.mode * 8 + 1);
    consumer_sync();
    return;
  }
  // element i of a thread is k = tid + 480 i; `i * 480 < K` is uniform over the CTA, so slots a narrow
  // model never uses (i >= 2 at H = 768) cost one branch instead of their loads and arithmetic
  float v[BT][kLnPer], gg[kLnPer], bb[kLnPer];
#pragma unroll
  for (int i = 0; i < kLnPer; ++i) {
    const int k = tid + i * kConsumers;
    gg[i] = 0.f; bb[i] = 0.f;
#pragma unroll
    for (int b = 0; b < BT; ++b) v[b][i] = 0.f;
    if (i * kConsumers < K) {
      gg[i] = (k < K) ? ph.ln_g[k] : 0.f;
      bb[i] = (k < K) ? ph.ln_b[k] : 0.f;
    }
  }
  stage_enter(ph.sentinel, ph.sentinel_epoch);
  cx.stamp(ph.mode * 8 + 0);
  unsigned pending = 0;
#pragma unroll
  for (int i = 0; i < kLnPer; ++i) {
    if (i * kConsumers < K) {
#pragma unroll
      for (int b = 0; b < BT; ++b) {
        const bool live = b < B && tid + i * kConsumers < K;
        const bool ok = fetch_x(cx, ph, min(b, B - 1), min(tid + i * kConsumers, K - 1), v[b][i], live);
        if (!live) v[b][i] = 0.f;
        else if (!ok) pending |= 1u << (b * kLnPer + i);
      }
    }
  }
  if (ph.src_kind == SRC_LL && ph.sentinel == nullptr) {
    // what is still missing is re-read as ONE batch of loads per round (one at a time it cost an L2 round trip
    // per word: 8 in a row at batch 4)
    while (pending) {
      const unsigned want = pending;   // the load predicates of a round must not depend on its checks
#pragma unroll
      for (int i = 0; i < kLnPer; ++i)
        if (i * kConsumers < K) {
          u64 wv[BT];   // the rows of one element slot together: BT loads in flight, then BT checks
#pragma unroll
          for (int b = 0; b < BT; ++b)
            wv[b] = ll_load_if(ph.src_ll + (size_t)b * K + tid + i * kConsumers, (want >> (b * kLnPer + i)) & 1u);
#pragma unroll
          for (int b = 0; b < BT; ++b) {
            const unsigned bit = 1u << (b * kLnPer + i);
            if ((want & bit) && (uint32_t)(wv[b] >> 32) == ph.src_epoch) {
              v[b][i] = __uint_as_float((uint32_t)wv[b]);
              pending &= ~bit;
            }
          }
        }
    }
  } else {
    while (pending) {
#pragma unroll
      for (int b = 0; b < BT; ++b)
#pragma unroll
        for (int i = 0; i < kLnPer; ++i)
          if (pending & (1u << (b * kLnPer + i))) {
            if (fetch_x(cx, ph, b, tid + i * kConsumers, v[b][i])) pending &= ~(1u << (b * kLnPer + i));
          }
    }
  }
  cx.stamp(ph.mode * 8 + 1);
  float* redA = cx.red;             // [BT][kNW]
  float* redB = cx.red + BT * kNW;  // [BT][kNW]
  // the BT butterflies are independent: one shuffle level of all rows at a time
  float part[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    part[b] = 0.f;
#pragma unroll
    for (int i = 0; i < kLnPer; ++i) part[b] += v[b][i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int b = 0; b < BT; ++b) part[b] += __shfl_xor_sync(FULL_MASK, part[b], o);
  if (l == 0) {
#pragma unroll
    for (int b = 0; b < BT; ++b) redA[b * kNW + w] = part[b];
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
      if (i * kConsumers < K) {
        const int k = tid + i * kConsumers;
        const float dlt = (k < K) ? v[b][i] - mean[b] : 0.f;
        q += dlt * dlt;
      }
    }
    part[b] = q;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int b = 0; b < BT; ++b) part[b] += __shfl_xor_sync(FULL_MASK, part[b], o);
  if (l == 0) {
#pragma unroll
    for (int b = 0; b < BT; ++b) redB[b * kNW + w] = part[b];
  }
  consumer_sync();
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kNW; ++i) t += redB[b * kNW + i];
    const float rstd = rsqrtf(t / (float)K + cx.eps);
#pragma unroll
    for (int i = 0; i < kLnPer; ++i) {
      const int k = tid + i * kConsumers;
      if (i * kConsumers < K && k < K) xs[b * K + k] = (b < B) ? (v[b][i] - mean[b]) * rstd * gg[i] + bb[i] : 0.f;
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

// This warp's rows of one chunk.  RB = 2: the warp owns two rows (first, first + 15) and walks them together
// - their loads and FMA chains interleave and every x vector pulled from shared memory serves both.
// Shared-memory bandwidth is what bounds this loop (15 warps x (1.5 KB of weights + 3 KB of x) per row), so the
// element-to-lane mapping is the conflict-free one: of every 256-element block a lane takes elements
// [4 lane, 4 lane + 4) and [128 + 4 lane, 128 + 4 lane + 4) - a quarter-warp then reads 128 contiguous bytes
// of x per float4 load (8 consecutive floats per lane would be a 2-way bank conflict on every x load).
template <typename WT, int BT, int RB, bool TR>
__device__ __forceinline__ void chunk_rows(Ctx<WT, TR>& cx, const Phase& ph, const uint8_t* slot, int first, int r,
                                           Best (&best)[BT]) {
  const int K = ph.K, B = cx.B;
  const int rowbytes = K * (int)sizeof(WT);
  const int lane = threadIdx.x & 31;
  const float* xs = cx.xs;
  float acc[RB][BT];
#pragma unroll
  for (int rI = 0; rI < RB; ++rI)
#pragma unroll
    for (int b = 0; b < BT; ++b) acc[rI][b] = 0.f;
  const WT* wrow[RB];
#pragma unroll
  for (int rI = 0; rI < RB; ++rI) wrow[rI] = reinterpret_cast<const WT*>(slot + (size_t)(first + rI * kNW) * rowbytes);
  constexpr int kUnrollK = (BT * RB == 1) ? 3 : (BT * RB == 2 ? 2 : 1);   // K = 768: exactly 3 trips
#pragma unroll kUnrollK
  for (int k0 = lane * 4; k0 < K; k0 += 256) {
    // K is a multiple of 8 only: in the last block the upper half may lie beyond the row (x counts as 0)
    const bool hi = k0 + 128 < K;
    float4 x0[BT], x1[BT];
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      x0[b] = *reinterpret_cast<const float4*>(xs + b * K + k0);
      x1[b] = hi ? *reinterpret_cast<const float4*>(xs + b * K + k0 + 128) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int rI = 0; rI < RB; ++rI) {
      float w0[4], w1[4];
      lds4(wrow[rI] + k0, w0);
      lds4(wrow[rI] + (hi ? k0 + 128 : k0), w1);
#pragma unroll
      for (int b = 0; b < BT; ++b) {
        float t = acc[rI][b];
        t = fmaf(w0[0], x0[b].x, t); t = fmaf(w0[1], x0[b].y, t); t = fmaf(w0[2], x0[b].z, t); t = fmaf(w0[3], x0[b].w, t);
        t = fmaf(w1[0], x1[b].x, t); t = fmaf(w1[1], x1[b].y, t); t = fmaf(w1[2], x1[b].z, t); t = fmaf(w1[3], x1[b].w, t);
        acc[rI][b] = t;
      }
    }
  }
  cx.stamp(ph.mode * 8 + 5);
  // the RB * BT butterflies are independent: one shuffle level of all of them at a time
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int rI = 0; rI < RB; ++rI)
#pragma unroll
      for (int b = 0; b < BT; ++b) acc[rI][b] += __shfl_xor_sync(FULL_MASK, acc[rI][b], o);
  cx.stamp(ph.mode * 8 + 6);
#pragma unroll
  for (int rI = 0; rI < RB; ++rI) {
    const int rr = first + rI * kNW;
    const float bv = ph.has_bias ? reinterpret_cast<const float*>(slot + kSlotData)[(r & 3) + rr] : 0.f;
    const int row = r + rr;
    if (ph.mode == EPI_HEAD) {
#pragma unroll
      for (int b = 0; b < BT; ++b) {
        best_merge(best[b], acc[rI][b], row);
        if (cx.logits_out != nullptr && lane == 0 && b < B) cx.logits_out[(size_t)b * cx.V + row] = acc[rI][b];
      }
    } else if (lane < B) {
      float v = 0.f;
#pragma unroll
      for (int b = 0; b < BT; ++b) if (lane == b) v = acc[rI][b];
      v += bv;
      const int own = lane * kOwnMax + (row - cx.c0);
      if (ph.mode == EPI_PROJ) { v += cx.hloc[own]; cx.h2loc[own] = v; }
      else if (ph.mode == EPI_FC) v = gelu_new(v);
      else if (ph.mode == EPI_PROJ2) { v += cx.h2loc[own]; cx.hloc[own] = v; }
      if (ph.dst_plain != nullptr) ph.dst_plain[(size_t)lane * ph.N + row] = v;
      else if (ph.dst_sys) ll_store_sys(ph.dst_ll + (size_t)lane * ph.N + row, __float_as_uint(v), ph.dst_epoch);
      else ll_store(ph.dst_ll + (size_t)lane * ph.N + row, v, ph.dst_epoch);
    }
  }
  cx.stamp(ph.mode * 8 + 7);
}

// Every warp walks the same chunk sequence.  Rows are dealt round-robin over the CTA's whole row range
// (row j of the range belongs to warp j % 15), not per chunk: with 21-row chunks a per-chunk deal gives
// six warps twice the work of the other nine in every chunk, and the warps only meet at the ring.
template <typename WT, int BT, bool TR>
__device__ __forceinline__ void run_rows(Ctx<WT, TR>& cx, const Phase& ph, const KindInfo& ki) {
  Ring& rg = cx.rg;
  const int r0 = ki.r0, r1 = ki.r1, rps = ki.rps;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Best best[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) { best[b].v = -INFINITY; best[b].i = 0x7fffffff; }
  int first = warp;   // index in the chunk of this warp's first row
  for (int r = r0; r < r1; r += rps) {
    const int n = min(rps, r1 - r);
    const int s = rg.slot;
    const uint8_t* slot = rg.slots + (size_t)s * kSlotBytes;
    if (lane == 0) mbar_wait(&rg.full[s], rg.par);   // one waiter per warp
    __syncwarp();
    if (r == r0) cx.stamp(ph.mode * 8 + 3);
    // a chunk holds at most 24 rows: a warp owns none, one or two of them
    static_assert(kMaxRows <= 2 * kNW, "a warp owns at most two rows of a chunk");
    if (first + kNW < n) chunk_rows<WT, BT, 2>(cx, ph, slot, first, r, best);
    else if (first < n) chunk_rows<WT, BT, 1>(cx, ph, slot, first, r, best);
    __syncwarp();
    if (lane == 0) mbar_arrive(&rg.empty[s]);
    rg.advance();
    // (row index in the range) % 15 == warp: the next chunk starts n rows further on
    first -= n % kNW;
    if (first < 0) first += kNW;
  }
  if (ph.mode == EPI_HEAD && lane == 0) {
#pragma unroll
    for (int b = 0; b < BT; ++b) { cx.sbv[warp * BT + b] = best[b].v; cx.sbi[warp * BT + b] = best[b].i; }
  }
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
  const float c1 = __expf(a.m - mn), c2 = __expf(mo - mn);
  a.l = a.l * c1 + lo * c2;
#pragma unroll
  for (int i = 0; i < 8; ++i) a.acc[i] = a.acc[i] * c1 + ao[i] * c2;
  a.m = mn;
}

struct AttnArgs {
  void *kc, *vc; const u64* qkv_ll; u64* attn_ll; uint32_t ep_qkv, ep_attn;
  int nh, slot0, max_ctx; float *part, *pm, *pl, *qkv_s;
};

// Decode attention for one (b, head) task.  Cached K/V rows do not depend on this step, so the
// first two rows of every 8-lane group are fetched before q is waited for.
template <typename WT, bool TR>
__device__ __forceinline__ void attention_task(Ctx<WT, TR>& cx, const AttnArgs& a, int task) {
  constexpr int D = 64;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, sub = lane >> 3, l8 = lane & 7;
  const int grp = warp * 4 + sub;                    // key rows t = grp + kGroups*j belong to this 8-lane group
  const unsigned grp_mask = 0xFFu << (sub * 8);
  const int H = cx.H, nh = a.nh;
  const int b = task / nh, head = task - b * nh;
  const size_t base = ((size_t)(a.slot0 + b) * nh + head) * (size_t)a.max_ctx * D;
  WT* Kc = reinterpret_cast<WT*>(a.kc) + base;
  WT* Vc = reinterpret_cast<WT*>(a.vc) + base;
  const int pos = cx.s_pos[b];
  Raw8<WT> pk0, pk1, pv0, pv1;
  if (grp < pos) {
    pk0.load(Kc + (size_t)grp * D + l8 * 8);
    pv0.load(Vc + (size_t)grp * D + l8 * 8);
  }
  if (grp + kGroups < pos) {
    pk1.load(Kc + (size_t)(grp + kGroups) * D + l8 * 8);
    pv1.load(Vc + (size_t)(grp + kGroups) * D + l8 * 8);
  }
  // a task needs only the 192 words of its head, which 192 threads poll directly
  stage_enter();
  cx.stamp(KIND_ATTN * 8 + 0);
  // q, k_new, v_new of this head: 192 values, one per thread, into shared memory
  if (tid < 3 * D) {
    const int which = tid / D, i = tid - which * D;
    const float val = __uint_as_float(ll_wait_bits(a.qkv_ll + (size_t)b * 3 * H + which * H + head * D + i, a.ep_qkv));
    if (which == 0) {
      a.qkv_s[i] = val * 0.125f;                     // 64^-0.5
    } else {
      // the new row goes into the cache for later steps, storage-rounded like every row
      const WT r = from_f<WT>(val);
      (which == 1 ? Kc : Vc)[(size_t)pos * D + i] = r;
      a.qkv_s[which * D + i] = to_f(r);
    }
  }
  consumer_sync();
  cx.stamp(KIND_ATTN * 8 + 1);
  float qv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) qv[i] = a.qkv_s[l8 * 8 + i];
  SoftState st;
  st.m = -INFINITY; st.l = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) st.acc[i] = 0.f;
  // software pipeline over this group's rows t = grp, grp+60, ...: two register buffers; a row's K/V
  // chunk is consumed and the buffer immediately refilled with the row two steps ahead
  auto step = [&](int t, Raw8<WT>& kb, Raw8<WT>& vb) {
    float kk[8], vv[8];
    if (t == pos) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { kk[i] = a.qkv_s[D + l8 * 8 + i]; vv[i] = a.qkv_s[2 * D + l8 * 8 + i]; }
    } else {
      kb.get(kk);
      vb.get(vv);
    }
    const int t2 = t + 2 * kGroups;
    if (t2 < pos) {
      kb.load(Kc + (size_t)t2 * D + l8 * 8);
      vb.load(Vc + (size_t)t2 * D + l8 * 8);
    }
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) dot = fmaf(qv[i], kk[i], dot);
    dot += __shfl_xor_sync(grp_mask, dot, 1);
    dot += __shfl_xor_sync(grp_mask, dot, 2);
    dot += __shfl_xor_sync(grp_mask, dot, 4);
    const float mn = fmaxf(st.m, dot);
    const float c1 = __expf(st.m - mn), pe = __expf(dot - mn);
    st.l = st.l * c1 + pe;
#pragma unroll
    for (int i = 0; i < 8; ++i) st.acc[i] = st.acc[i] * c1 + pe * vv[i];
    st.m = mn;
  };
  for (int t = grp; t <= pos; t += 2 * kGroups) {
    step(t, pk0, pv0);
    if (t + kGroups <= pos) step(t + kGroups, pk1, pv1);
  }
  __syncwarp();
  cx.stamp(KIND_ATTN * 8 + 2);
  // merge the four groups of the warp, then the 15 warps through shared memory
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
    for (int i = 0; i < 8; ++i) a.part[warp * D + l8 * 8 + i] = st.acc[i];
    if (l8 == 0) { a.pm[warp] = st.m; a.pl[warp] = st.l; }
  }
  consumer_sync();
  if (tid < D) {
    float M = a.pm[0];
#pragma unroll
    for (int ww = 1; ww < kNW; ++ww) M = fmaxf(M, a.pm[ww]);
    float Lsum = 0.f, o = 0.f;
#pragma unroll
    for (int ww = 0; ww < kNW; ++ww) {
      const float c = (a.pm[ww] == -INFINITY) ? 0.f : __expf(a.pm[ww] - M);
      Lsum += a.pl[ww] * c;
      o += a.part[ww * D + tid] * c;
    }
    ll_store(a.attn_ll + (size_t)b * H + head * D + tid, o / Lsum, a.ep_attn);
  }
}

template <typename WT, int BT, bool TR>
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
  float* hloc = red + 2 * BT * kNW;                                             // [BT][kOwnMax] residual h
  float* h2loc = hloc + BT * kOwnMax;                                           // [BT][kOwnMax] residual h2
  float* qkv_s = h2loc + BT * kOwnMax;                                          // [3*D] q, k_new, v_new of a task
  float* sbv = qkv_s + 3 * D;                                                   // [kNW][BT]
  int* sbi = reinterpret_cast<int*>(sbv + kNW * BT);                            // [kNW][BT]
  int* s_ids = sbi + kNW * BT;                                                  // [8]
  int* s_pos = s_ids + 8;                                                       // [8]
  uint64_t* full = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(s_pos + 8) + 15) & ~(uintptr_t)15);
  uint64_t* empty = full + kMaxSlots;
  int* s_pos0 = reinterpret_cast<int*>(empty + kMaxSlots);                      // [64] cache lengths at launch
  KindInfo* s_kind = reinterpret_cast<KindInfo*>(s_pos0 + 64);                  // [8] (6 used)
  MegaLayer* s_layers = reinterpret_cast<MegaLayer*>(s_kind + 8);               // [L]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // one monotonic counter gates the argmax exchange (see there); it lives in an unused row of the partials buffer
  uint32_t* const argmax_ctr = reinterpret_cast<uint32_t*>(p.part_ll + (size_t)14 * kFlagStride);
  if (tid == 0) {
    s_kind[7].r0 = (int)ld_relaxed_u32(argmax_ctr);   // its value when this launch starts: nobody bumps it before every CTA got here
    for (int s = 0; s < p.n_slots; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kNW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid >= 32 && tid < 38) {
    const int kind = tid - 32;
    KindInfo ki;
    ki.N = (kind == EPI_QKV) ? 3 * H : (kind == EPI_FC) ? F : (kind == EPI_HEAD) ? p.V : H;
    ki.K = (kind == EPI_PROJ2) ? F : H;
    rows_of(cta, G, ki.N, ki.r0, ki.r1);
    ki.rps = rows_per_slot(ki.K * (int)sizeof(WT));
    ki.pad[0] = ki.pad[1] = ki.pad[2] = 0;
    s_kind[kind] = ki;
  }
  {  // layer table -> shared memory (pointer chasing through global memory would sit on every phase)
    const int n_words = p.L * (int)(sizeof(MegaLayer) / 8);
    const u64* src = reinterpret_cast<const u64*>(p.layers);
    u64* dst = reinterpret_cast<u64*>(s_layers);
    for (int i = tid; i < n_words; i += kMegaThreads) dst[i] = src[i];
  }
  __syncthreads();

  if (warp == kNW) {
    // =============================== PRODUCER ===============================
    if (lane == 0) {
      Ring rg{ring_base, full, empty, p.n_slots, 0, 0u};
      const uint64_t policy = l2_stream_policy(p.l2_stream_hint != 0);
      for (int step = 0; step < p.n_steps * p.n_streams; ++step) {
        for (int li = 0; li < p.L; ++li) {
          const MegaLayer& w = s_layers[li];
#pragma unroll 1
          for (int kind = 0; kind < 5; ++kind) {
            if (kind == KIND_ATTN) continue;
            const void* W = (kind == EPI_QKV) ? w.attn_w : (kind == EPI_PROJ) ? w.proj_w : (kind == EPI_FC) ? w.fc_w : w.mproj_w;
            const float* bias = (kind == EPI_QKV) ? w.attn_b : (kind == EPI_PROJ) ? w.proj_b : (kind == EPI_FC) ? w.fc_b : w.mproj_b;
            produce_phase<WT>(rg, W, bias, s_kind[kind], policy);
          }
        }
        if (p.has_head) produce_phase<WT>(rg, p.wte, nullptr, s_kind[EPI_HEAD], policy);
      }
    }
    return;
  }

  // ================================= CONSUMERS =================================
  // optional phase trace (GP_MEGA_TRACE): every CTA records %globaltimer (ns, one clock for the whole GPU)
  Ctx<WT, TR> cx;
  cx.trace = p.trace ? p.trace + (size_t)cta * p.trace_cap : nullptr; cx.trace_cap = p.trace_cap; cx.n_stamp = 0;
  cx.tracing = false;
  const int n_items = p.n_steps * p.n_streams;   // item = (round, stream); one decode step of B sequences each
  if (tid < 64) s_pos0[tid] = (tid < p.n_streams * B) ? p.lens[p.slot0 + tid] : 0;
  if (tid < 8) { s_pos[tid] = 0; s_ids[tid] = 0; }
  if (cta == 0 && tid == 0 && p.step_clock != nullptr) p.step_clock[n_items] = globaltimer();  // start stamp
  consumer_sync();

  cx.rg = Ring{ring_base, full, empty, p.n_slots, 0, 0u};
  cx.xs = xs; cx.red = red; cx.hloc = hloc; cx.h2loc = h2loc; cx.sbv = sbv; cx.sbi = sbi; cx.s_ids = s_ids; cx.s_pos = s_pos;
  cx.wte = reinterpret_cast<const WT*>(p.wte); cx.wpe = reinterpret_cast<const WT*>(p.wpe);
  cx.B = B; cx.H = H; cx.V = p.V; cx.n_pos = p.n_pos; cx.G = G; cx.cta = cta; cx.eps = p.eps; cx.logits_out = p.logits_out;
  int c0, c1;                                        // residual rows this CTA owns (same split as N = H phases)
  rows_of(cta, G, H, c0, c1);
  cx.c0 = c0;
  for (int step = 0; step < n_items; ++step) {
    const int round = step / p.n_streams, strm = step - round * p.n_streams;
    const int slot_base = p.slot0 + strm * B;
    const uint32_t cep = p.comm_epoch0 + (uint32_t)round;   // epoch of this item's words on the stage boundary
    const uint32_t ep_step = p.epoch0 + (uint32_t)step * (uint32_t)(p.L + 1) * 8u;
    const u64* in_hid = p.inbox_hid ? p.inbox_hid + (size_t)strm * B * H : nullptr;
    // ---- position and (stage 0) token of every sequence of this item
    if (tid < B) {
      s_pos[tid] = min(s_pos0[strm * B + tid] + round, p.max_ctx - 1);
      if (p.has_embed) {
        int id = s_ids[tid];   // single-stage loop: this CTA's own argmax of the previous step
        if (p.inbox_ids != nullptr && round > 0) {
          // pipeline: the token the last stage produced for this stream in the previous round
          id = (int)ll_wait_sys(p.inbox_ids + strm * B + tid, cep - 1u);
          if (cta == 0 && p.tok_out != nullptr) p.tok_out[(size_t)(strm * B + tid) * p.tok_stride + (round - 1)] = id;
        } else if (round == 0) {
          id = __ldcg(p.ids + strm * B + tid);
        }
        s_ids[tid] = min(max(id, 0), p.V - 1);
      }
    }
    consumer_sync();
    // ---- the residual slice this CTA owns: embedding rows (stage 0) or the incoming hidden state
    for (int i = tid; i < B * (c1 - c0); i += kConsumers) {
      const int b = i / (c1 - c0), col = c0 + i % (c1 - c0);
      float hv;
      if (p.has_embed) hv = to_f(cx.wte[(size_t)s_ids[b] * H + col]) + to_f(cx.wpe[(size_t)min(s_pos[b], p.n_pos - 1) * H + col]);
      else if (in_hid != nullptr) hv = __uint_as_float(ll_wait_sys(in_hid + (size_t)b * H + col, cep));
      else hv = __ldcg(p.hidden_in + (size_t)b * H + col);
      hloc[b * kOwnMax + (col - c0)] = hv;
    }
    cx.tracing = (p.trace != nullptr && step == n_items - 1);
    cx.stamp(7 * 8 + 0);
    for (int li = 0; li < p.L; ++li) {
      const MegaLayer& w = s_layers[li];
      const bool last = (li == p.L - 1);
      const uint32_t ep = ep_step + (uint32_t)li * 8u;         // +1 qkv, +2 attn, +3 h2, +4 act, +5 h
      (void)last;
      Phase ph;
      // ---- P1: ln_1 + c_attn -> qkv[b, 3H]
      ph.src_kind = (li > 0) ? SRC_LL : (p.has_embed ? SRC_EMBED : (in_hid ? SRC_LL : SRC_PLAIN));
      ph.src_ll = p.h_ll; ph.src_plain = p.hidden_in; ph.K = H; ph.src_epoch = ep - 8u + 5u;
      ph.dst_sys = 0; ph.sentinel = nullptr; ph.sentinel_epoch = 0;
      if (li == 0 && !p.has_embed && in_hid != nullptr) {
        // input = what the previous pipeline stage wrote into this GPU's inbox
        ph.src_ll = in_hid; ph.src_epoch = cep;
        ph.sentinel = in_hid + (H - 1); ph.sentinel_epoch = cep;
      }
      ph.ln_g = w.ln1_g; ph.ln_b = w.ln1_b;
      ph.N = 3 * H; ph.has_bias = 1; ph.mode = EPI_QKV; ph.dst_ll = p.qkv_ll; ph.dst_plain = nullptr; ph.dst_epoch = ep + 1u;
      stage_x<WT, BT>(cx, ph);
      cx.stamp(ph.mode * 8 + 2);
      run_rows<WT, BT>(cx, ph, s_kind[EPI_QKV]);
      cx.stamp(ph.mode * 8 + 4);
      ph.sentinel = nullptr;
      // ---- P2: attention for the (b, head) tasks of this CTA
      if (cta < B * nh) {
        AttnArgs a;
        a.kc = w.kc; a.vc = w.vc; a.qkv_ll = p.qkv_ll; a.attn_ll = p.attn_ll; a.ep_qkv = ep + 1u; a.ep_attn = ep + 2u;
        a.nh = nh; a.slot0 = slot_base; a.max_ctx = p.max_ctx; a.part = part; a.pm = pm; a.pl = pl; a.qkv_s = qkv_s;
        for (int task = cta; task < B * nh; task += G) {
          attention_task<WT>(cx, a, task);
        }
      }
      cx.stamp(KIND_ATTN * 8 + 3);
      // ---- P3: attention c_proj + residual -> h2   (CTAs without an attention task publish qkv here)
      ph.src_kind = SRC_LL; ph.src_ll = p.attn_ll; ph.K = H; ph.src_epoch = ep + 2u;
      ph.ln_g = nullptr; ph.ln_b = nullptr;
      ph.N = H; ph.mode = EPI_PROJ; ph.dst_ll = p.h2_ll; ph.dst_epoch = ep + 3u;
      stage_x<WT, BT>(cx, ph);
      cx.stamp(ph.mode * 8 + 2);
      run_rows<WT, BT>(cx, ph, s_kind[EPI_PROJ]);
      cx.stamp(ph.mode * 8 + 4);
      // ---- P4: ln_2 + c_fc + gelu_new -> act
      ph.src_ll = p.h2_ll; ph.src_epoch = ep + 3u;
      ph.ln_g = w.ln2_g; ph.ln_b = w.ln2_b;
      ph.N = F; ph.mode = EPI_FC; ph.dst_ll = p.act_ll; ph.dst_epoch = ep + 4u;
      stage_x<WT, BT>(cx, ph);
      cx.stamp(ph.mode * 8 + 2);
      run_rows<WT, BT>(cx, ph, s_kind[EPI_FC]);
      cx.stamp(ph.mode * 8 + 4);
      // ---- P5: mlp c_proj + residual -> h (or the stage output)
      ph.src_ll = p.act_ll; ph.K = F; ph.src_epoch = ep + 4u;
      ph.ln_g = nullptr; ph.ln_b = nullptr;
      ph.N = H; ph.mode = EPI_PROJ2; ph.dst_ll = p.h_ll; ph.dst_epoch = ep + 5u;
      if (last && !p.has_head) {
        if (p.peer_hid != nullptr) {   // straight into the next stage's inbox over NVLink
          ph.dst_ll = p.peer_hid + (size_t)strm * B * H; ph.dst_epoch = cep; ph.dst_sys = 1;
        } else {
          ph.dst_plain = p.hidden_out;
        }
      }
      stage_x<WT, BT>(cx, ph);
      cx.stamp(ph.mode * 8 + 2);
      run_rows<WT, BT>(cx, ph, s_kind[EPI_PROJ2]);
      cx.stamp(ph.mode * 8 + 4);
    }
    if (p.has_head) {
      // ---- ln_f + lm_head + argmax
      const uint32_t ep = ep_step + (uint32_t)p.L * 8u;
      Phase ph;
      ph.src_kind = SRC_LL; ph.src_ll = p.h_ll; ph.src_plain = nullptr; ph.K = H; ph.src_epoch = ep - 8u + 5u;
      ph.ln_g = p.lnf_g; ph.ln_b = p.lnf_b;
      ph.N = p.V; ph.has_bias = 0; ph.mode = EPI_HEAD; ph.dst_ll = nullptr; ph.dst_plain = nullptr; ph.dst_epoch = 0;
      ph.dst_sys = 0; ph.sentinel = nullptr; ph.sentinel_epoch = 0;
      stage_x<WT, BT>(cx, ph);
      cx.stamp(ph.mode * 8 + 2);
      run_rows<WT, BT>(cx, ph, s_kind[EPI_HEAD]);
      consumer_sync();
      // partials are laid out [b][value|index][cta] so that the readers' loads coalesce
      if (tid < B) {
        Best r{sbv[tid], sbi[tid]};
        for (int ww = 1; ww < kNW; ++ww) best_merge(r, sbv[ww * BT + tid], sbi[ww * BT + tid]);
        ll_store(p.part_ll + ((size_t)tid * 2 + 0) * kFlagStride + cta, r.v, ep + 6u);
        ll_store_int(p.part_ll + ((size_t)tid * 2 + 1) * kFlagStride + cta, r.i, ep + 6u);
      }
      cx.stamp(7 * 8 + 2);
      // every CTA reduces all partials itself (index tie-break => deterministic): every CTA knows the next
      // token without another exchange.  Unlike the per-layer exchanges this one keeps a counter in front of the
      // data: 148 warps polling the same 19 lines that 148 producers are still writing delayed those writes
      // (+3 us per token measured); one lane per reading warp watches one counter line instead.
      if (warp == 0) {
        __syncwarp();
        if (lane == 0) red_add_u32(argmax_ctr, 1u);
      }
      for (int b = warp; b < B; b += kNW) {
        if (lane == 0) {
          const uint32_t target = (uint32_t)s_kind[7].r0 + (uint32_t)(step + 1) * (uint32_t)G;
          while ((int32_t)(ld_relaxed_u32(argmax_ctr) - target) < 0) {}
        }
        __syncwarp();
        Best r{-INFINITY, 0x7fffffff};
        // 148 CTAs = two batches of 2 x 4 words per lane.  Every word of a batch that is still missing is re-read
        // in ONE round of predicated loads (spinning on them one after the other cost a round trip per word:
        // +3 us per token once the counter gate in front of this loop was gone)
        const u64* pv = p.part_ll + ((size_t)b * 2 + 0) * kFlagStride;
        const u64* pi = p.part_ll + ((size_t)b * 2 + 1) * kFlagStride;
#pragma unroll 1
        for (int j0 = 0; j0 < G; j0 += 128) {
          u64 wv[4], wi[4];
          unsigned pend = 0;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j0 + j * 32 + lane < G) pend |= 3u << (2 * j);
          while (pend) {
            const unsigned want = pend;   // the predicates of a round must not depend on its checks
            u64 nv[4], ni[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int g = j0 + j * 32 + lane;
              nv[j] = ll_load_if(pv + g, (want >> (2 * j)) & 1u);
              ni[j] = ll_load_if(pi + g, (want >> (2 * j + 1)) & 1u);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (((want >> (2 * j)) & 1u) && (uint32_t)(nv[j] >> 32) == ep + 6u) { wv[j] = nv[j]; pend &= ~(1u << (2 * j)); }
              if (((want >> (2 * j + 1)) & 1u) && (uint32_t)(ni[j] >> 32) == ep + 6u) { wi[j] = ni[j]; pend &= ~(2u << (2 * j)); }
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j0 + j * 32 + lane < G) best_merge(r, __uint_as_float((uint32_t)wv[j]), (int)(uint32_t)wi[j]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(FULL_MASK, r.v, o);
          const int oi = __shfl_xor_sync(FULL_MASK, r.i, o);
          best_merge(r, ov, oi);
        }
        if (lane == 0) {
          if (cta == 0) {
            p.ids_out[strm * B + b] = r.i;
            if (p.seq != nullptr) p.seq[(size_t)(strm * B + b) * p.seq_stride + s_pos[b] + 1] = r.i;
            // pipeline: the token goes straight into stage 0's inbox
            if (p.peer_ids != nullptr) ll_store_sys(p.peer_ids + strm * B + b, (uint32_t)r.i, cep);
          }
          s_ids[b] = min(max(r.i, 0), p.V - 1);
        }
      }
      cx.stamp(7 * 8 + 1);
    }
    // ---- bookkeeping: the cache grew by one position
    if (cta == 0 && tid < B) p.lens[slot_base + tid] = s_pos[tid] + 1;
    if (cta == 0 && tid == 0 && p.step_clock != nullptr) p.step_clock[step] = globaltimer();
    consumer_sync();
  }
  // pipeline, stage 0: the tokens of the last round never come back as inputs; collect them here
  if (p.has_embed && p.inbox_ids != nullptr && p.tok_out != nullptr && cta == 0 && tid < p.n_streams * B) {
    const int id = (int)ll_wait_sys(p.inbox_ids + tid, p.comm_epoch0 + (uint32_t)p.n_steps - 1u);
    p.tok_out[(size_t)tid * p.tok_stride + (p.n_steps - 1)] = id;
  }
}

// traced kernels exist for bf16 storage only (the trace is a tuning aid for the bandwidth-bound configuration)
template <typename WT> constexpr bool kHasTraced = false;
template <> constexpr bool kHasTraced<bf16> = true;

template <typename WT, int BT>
cudaError_t launch_t(const MegaParams& p, int grid, size_t smem, cudaStream_t s) {
  void* args[] = {const_cast<MegaParams*>(&p)};
  const void* fn = (const void*)decode_mega_kernel<WT, BT, false>;
  if constexpr (kHasTraced<WT>) {
    if (p.trace != nullptr) fn = (const void*)decode_mega_kernel<WT, BT, true>;
  }
  return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kMegaThreads), args, smem, s);
}
template <typename WT, int BT> cudaError_t set_attr() {
  cudaError_t e = cudaFuncSetAttribute(decode_mega_kernel<WT, BT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if constexpr (kHasTraced<WT>) {
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(decode_mega_kernel<WT, BT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  }
  return e;
}

int padded(int B) { return B <= 1 ? 1 : (B <= 2 ? 2 : 4); }

size_t fixed_smem(int BT, int H, int L) {
  return (size_t)BT * 4 * H * 4 /*xs*/ + kNW * 64 * 4 /*part*/ + 2 * kNW * 4 /*pm,pl*/ + (size_t)2 * BT * kNW * 4 /*red*/ +
         (size_t)2 * BT * kOwnMax * 4 /*hloc,h2loc*/ + 3 * 64 * 4 /*qkv_s*/ + (size_t)kNW * BT * 8 /*sbv,sbi*/ +
         64 /*ids,pos*/ + 16 + 2 * kMaxSlots * 8 + 256 /*pos0*/ + 8 * sizeof(KindInfo) +
         (size_t)L * sizeof(MegaLayer) + 128;
}

}  // namespace

bool mega_supported(int B, int H, int nh, int L, int dtype, int num_sms) {
  if (B < 1 || B > 4 || H % 8 != 0 || nh <= 0 || H % nh != 0 || H / nh != 64) return false;
  if (H > kLnPer * kConsumers || 4 * H > kCopyPer * kConsumers) return false;
  if (num_sms < 1 || num_sms > kFlagStride || B * nh > kFlagStride) return false;
  if ((H + num_sms - 1) / num_sms + 1 > kOwnMax) return false;   // residual rows per CTA
  const size_t fixed = fixed_smem(padded(B), H, L);
  if (fixed + 2 * (size_t)kSlotBytes > 227 * 1024) return false;
  // one row of the widest matrix (K = 4H) must fit a slot
  return (size_t)4 * H * (dtype == GP_BF16 ? 2 : 4) <= (size_t)kSlotData;
}

uint32_t mega_epochs_per_launch(int L, int n_steps) { return (uint32_t)n_steps * (uint32_t)(L + 1) * 8u + 8u; }
size_t mega_part_words() { return ((size_t)kGemvMaxBatch * 2 + 2) * kFlagStride; }   // + the counter row and the sampled-token row of k_mega2.cu

cudaError_t init_mega_attrs() {
  cudaError_t e;
#define GP_SET(WT, BT) if ((e = set_attr<WT, BT>()) != cudaSuccess) return e;
  GP_SET(bf16, 1) GP_SET(bf16, 2) GP_SET(bf16, 4) GP_SET(float, 1) GP_SET(float, 2) GP_SET(float, 4)
#undef GP_SET
  return cudaSuccess;
}

cudaError_t launch_decode_mega(MegaParams p, int dtype, int num_sms, cudaStream_t s) {
  if (!mega_supported(p.B, p.H, p.nh, p.L, dtype, num_sms) || p.n_steps < 1 || p.epoch0 == 0 ||
      p.n_streams < 1 || p.n_streams * p.B > 64)
    return cudaErrorInvalidValue;
  const int BT = padded(p.B);
  const size_t fixed = fixed_smem(BT, p.H, p.L);
  int n_slots = (int)((227 * 1024 - fixed) / kSlotBytes);
  if (n_slots > kMaxSlots) n_slots = kMaxSlots;
  if (const char* env = getenv("GP_MEGA_SLOTS")) { int v = atoi(env); if (v >= 2 && v < n_slots) n_slots = v; }
  p.n_slots = n_slots;
  const size_t smem = fixed + (size_t)n_slots * kSlotBytes;
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
