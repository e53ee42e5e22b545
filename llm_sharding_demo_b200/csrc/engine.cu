// The pipeline-stage engine behind the C ABI of include/gpt2pipe.h: owns weights, KV cache
// and workspaces in HBM and sequences the sm_100a kernels for prefill, cached decode and the
// device-resident greedy loop.  Host-side C++ only; every kernel lives in k_*.cu / gemm_*.cu.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <algorithm>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "kernels.h"

using namespace gp;

namespace {
thread_local std::string g_err;

// Every entry point works on its engine's device but leaves the CALLER's current device as it found it: a host
// that owns several GPUs (or torch, which caches what it believes the current device to be) must not see it move.
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int dev) {
    err = cudaGetDevice(&prev);
    if (err != cudaSuccess) { prev = -1; return; }
    if (prev != dev) err = cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

void set_err(gp_engine* e, const char* fmt, ...);

struct LayerW {
  float *ln1_g = nullptr, *ln1_b = nullptr, *attn_b = nullptr, *proj_b = nullptr;
  float *ln2_g = nullptr, *ln2_b = nullptr, *fc_b = nullptr, *mproj_b = nullptr;
  void *attn_w = nullptr, *proj_w = nullptr, *fc_w = nullptr, *mproj_w = nullptr;  // [N,K] in cfg.dtype
  void *kc = nullptr, *vc = nullptr;  // [max_batch, nh, max_ctx, d]
  // LayerNorm folded into the following matrix for the bf16 persistent kernel: u = W.gamma, c = W.beta + bias
  float *attn_u = nullptr, *attn_c = nullptr, *fc_u = nullptr, *fc_c = nullptr;
};
}  // namespace

struct gp_engine {
  gp_config cfg{};
  int H = 0, nh = 0, d = 0, V = 0, F = 0, L = 0;
  size_t esz = 4;  // bytes per storage element
  std::vector<LayerW> layers;
  void *wte = nullptr, *wpe = nullptr;
  float *lnf_g = nullptr, *lnf_b = nullptr;
  float *lnf_u = nullptr, *lnf_c = nullptr;   // ln_f folded into the tied lm_head (bf16 persistent kernel)
  std::set<std::string> required, loaded;
  std::vector<void*> allocs;
  int32_t* lens = nullptr;
  // host mirror of lens[]: prefill sets it, decode advances it, kv_reset clears it.  It exists so that a step
  // into a full slot is refused before launch (the reference raises IndexError past wpe)
  std::vector<int32_t> hlens;
  // workspaces
  float* h = nullptr;       // [max_tokens, H]   residual stream
  void* xa = nullptr;       // [max_tokens, H]   LN / attention output feeding a GEMM
  void* qkv = nullptr;      // [max_tokens, 3H]
  void* act = nullptr;      // [max_tokens, 4H]
  float* q = nullptr;       // [max_batch, H]    decode query
  float* attn_f = nullptr;  // [max_batch, H]    decode attention output (GEMV path)
  float* act_f = nullptr;   // [max_batch, 4H]   decode GELU output (GEMV path)
  float* hlast = nullptr;   // [max_batch, H]
  float* splitk = nullptr;  // split-K partial tiles of the tcgen05 GEMM (decode-sized M)
  size_t splitk_floats = 0;
  float* logits_ws = nullptr;  // [max_batch, V] logits of the GEMM-path lm_head / input of the sampling kernel
  void* lm_scratch = nullptr;
  float* stage = nullptr;
  size_t stage_floats = 0;
  // generate state
  int32_t* seq = nullptr;      // [max_batch, seq_stride]
  int32_t* prompt = nullptr;   // [max_batch * n_positions]
  int32_t* mailbox = nullptr;  // [max_batch] next token ids
  int seq_stride = 0;
  cudaStream_t stream = nullptr;
  std::map<int, cudaGraphExec_t> decode_graphs;  // keyed by batch (+ kSampledKey when the sampler is on)
  // token selection: greedy (default) or the reference's temperature / top-k / multinomial (gp_set_sampler)
  bool sampler_on = false;
  float sampler_temperature = 1.f;
  int sampler_top_k = 0;
  uint64_t sampler_seed = 0;
  std::map<int, int64_t> decode_graph_launches;
  // gp_generate's prefill, captured per (batch, prompt length, sampler on): ~150 launches of a few us each
  std::map<int64_t, cudaGraphExec_t> prefill_graphs;
  std::map<int64_t, int64_t> prefill_graph_launches;
  std::vector<cudaEvent_t> events;
  // persistent decode kernel
  MegaLayer* mega_layers = nullptr;
  unsigned long long *qkv_ll = nullptr, *attn_ll = nullptr, *h2_ll = nullptr, *act_ll = nullptr, *h_ll = nullptr;
  unsigned long long* part_ll = nullptr;
  size_t ll_bytes[5] = {0, 0, 0, 0, 0};
  // pipeline stage boundary: this stage's inbox (peers write into it) and the mapped inbox of the peer it feeds
  unsigned char* inbox = nullptr;       // [1024 B ids words][64*H hidden words]
  unsigned char* peer_inbox = nullptr;  // next stage's inbox (stage 0's for the last stage), cudaIpcOpenMemHandle
  bool peer_is_ipc = false;
  uint32_t mega_epoch = 8;
  // watchdog words of the persistent kernel: ctrl[0] = abort (host -> kernel), ctrl[1] = items finished (kernel -> host);
  // written / read through a side stream and a pinned host word while the kernel runs
  int* ctrl = nullptr;
  int* ctrl_host = nullptr;
  cudaStream_t ctrl_stream = nullptr;
  int32_t* slot_map = nullptr;     // [64] KV slot of every sequence of a slot-indexed decode step
  bool coop_ok = true;             // the persistent kernels can be co-scheduled on this device / context
  unsigned long long* step_clock = nullptr;
  bool mega_uploaded = false;
  long long* trace = nullptr;
  int num_sms = 0;
  std::string err;
  int64_t launches = 0;
};

namespace {

void set_err(gp_engine* e, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  if (e) e->err = buf;
}

#define CK(e, call)                                                                          \
  do {                                                                                       \
    cudaError_t _c = (call);                                                                 \
    if (_c != cudaSuccess) {                                                                 \
      set_err(e, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_c), __FILE__, __LINE__); \
      return _c == cudaErrorMemoryAllocation ? GP_ENOMEM : GP_ECUDA;                         \
    }                                                                                        \
  } while (0)
// a kernel launch: counted, checked
#define LAUNCH(e, call)  \
  do {                   \
    CK(e, call);         \
    ++(e)->launches;     \
  } while (0)

int dev_alloc(gp_engine* e, void** p, size_t bytes) {
  if (bytes == 0) bytes = 16;
  CK(e, cudaMalloc(p, bytes));
  e->allocs.push_back(*p);
  return GP_OK;
}
#define ALLOC(e, ptr, bytes)                                    \
  do {                                                          \
    int _r = dev_alloc(e, reinterpret_cast<void**>(&(ptr)), (bytes)); \
    if (_r != GP_OK) return _r;                                 \
  } while (0)

bool bf(const gp_engine* e) { return e->cfg.dtype == GP_BF16; }

// GEMM dispatch: tcgen05 when the shape allows and the storage type is bf16, else SIMT tiles
cudaError_t gemm(gp_engine* e, const void* A, int lda, const void* Wt, const float* bias, const float* residual,
                 int ldr, void* out, int ldo, int M, int N, int K, uint32_t flags, cudaStream_t s) {
  if (bf(e) && !(e->cfg.flags & GP_FLAG_NO_TCGEN05) && gemm_tc_supported(M, N, K) && lda == K) {
    float* part = (M <= 256) ? e->splitk : nullptr;
    if (part && gemm_tc_splits(M, N, K, e->splitk_floats) > 1) ++e->launches;   // the split-K epilogue kernel
    return launch_gemm_tc(A, lda, Wt, bias, residual, ldr, out, ldo, M, N, K, flags, s, part, e->splitk_floats);
  }
  return launch_gemm_simt(A, lda, Wt, bias, residual, ldr, out, ldo, M, N, K, flags, e->cfg.dtype, s);
}

// ln_f + lm_head + argmax over rows x[B,H] in chunks of kGemvMaxBatch
constexpr int kSampledKey = 1 << 20;

int run_lm_head(gp_engine* e, const float* x, int ldx, int B, int slot0, float* logits, int32_t* next_ids,
                int32_t* lens, int32_t* seq, cudaStream_t s) {
  for (int c = 0; c < B; c += kGemvMaxBatch) {
    LmheadParams p;
    p.x = x + (size_t)c * ldx; p.ldx = ldx;
    p.ln_g = e->lnf_g; p.ln_b = e->lnf_b; p.eps = e->cfg.ln_eps;
    p.wte = e->wte;
    p.B = (B - c < kGemvMaxBatch) ? B - c : kGemvMaxBatch;
    p.V = e->V; p.H = e->H;
    p.ids_out = (next_ids ? next_ids : e->mailbox) + c;
    p.logits_out = logits ? logits + (size_t)c * e->V : nullptr; p.ldl = e->V;
    p.scratch = e->lm_scratch;
    p.lens = lens; p.slot0 = slot0 + c;
    p.seq = seq ? seq + (size_t)c * e->seq_stride : nullptr; p.seq_stride = e->seq_stride;
    LAUNCH(e, launch_lmhead_argmax(p, e->cfg.dtype, s));
  }
  return GP_OK;
}

// ln_f + lm_head + token selection (+ bookkeeping when `lens` is given) for rows x[B,H]: greedy argmax fused into
// the lm_head kernel, or - sampler on - logits into the workspace followed by the sampling kernel.
// fixed_pos: position of the new token when `lens` is NULL (prefill: T).
int select_tokens(gp_engine* e, const float* x, int ldx, int B, int slot0, float* logits, int32_t* next_ids,
                  int32_t* lens, int32_t* seq, int fixed_pos, cudaStream_t s) {
  if (!e->sampler_on) return run_lm_head(e, x, ldx, B, slot0, logits, next_ids, lens, seq, s);
  float* lg = logits ? logits : e->logits_ws;
  int32_t* ids = next_ids ? next_ids : e->mailbox;
  int r = run_lm_head(e, x, ldx, B, slot0, lg, ids, nullptr, nullptr, s);   // its argmax is overwritten below
  if (r != GP_OK) return r;
  LAUNCH(e, launch_sample_topk(lg, e->V, B, e->V, e->sampler_temperature, e->sampler_top_k, e->sampler_seed, slot0,
                               fixed_pos, ids, lens, slot0, seq, e->seq_stride, s));
  return GP_OK;
}

int check_ready(gp_engine* e) {
  if (e->loaded.size() != e->required.size()) {
    for (auto& n : e->required)
      if (!e->loaded.count(n)) {
        set_err(e, "weight '%s' has not been loaded (%zu of %zu present)", n.c_str(), e->loaded.size(),
                e->required.size());
        return GP_ESTATE;
      }
  }
  return GP_OK;
}

// GP_OK when every slot of [slot0, slot0+n) can take `steps` more tokens
int slot_room(gp_engine* e, const char* who, int slot0, int n, int steps) {
  for (int i = slot0; i < slot0 + n; ++i)
    if (e->hlens[i] + steps > e->cfg.max_ctx) {
      set_err(e, "%s: KV slot %d holds %d tokens; %d more would exceed max_ctx=%d", who, i, e->hlens[i], steps,
              e->cfg.max_ctx);
      return GP_ESTATE;
    }
  return GP_OK;
}

// half-block stage boundaries (GP_FLAG_BEGIN_HALF / GP_FLAG_END_HALF): which halves of block li this stage runs
bool has_attn_half(const gp_engine* e, int li) { return !(li == 0 && (e->cfg.flags & GP_FLAG_BEGIN_HALF)); }
bool has_mlp_half(const gp_engine* e, int li) { return !(li == e->L - 1 && (e->cfg.flags & GP_FLAG_END_HALF)); }
bool half_cut(const gp_engine* e) { return (e->cfg.flags & (GP_FLAG_BEGIN_HALF | GP_FLAG_END_HALF)) != 0; }

// second-generation persistent kernel (k_mega2.cu): bf16 storage, up to 8 sequences, tensor-core rows
bool use_mega2(const gp_engine* e, int B) {
  static const bool force_v1 = getenv("GP_MEGA_V1") != nullptr;
  return !force_v1 && e->cfg.dtype == GP_BF16 && mega2_supported(B, e->H, e->nh, e->L, e->cfg.dtype, e->num_sms);
}
bool use_mega(const gp_engine* e, int B) {
  if ((e->cfg.flags & GP_FLAG_NO_MEGA) || !e->coop_ok) return false;
  // the bf16 persistent kernel draws sampled tokens itself; the fp32 one fuses the greedy argmax only
  if (e->sampler_on) return use_mega2(e, B) && (!e->cfg.has_head || e->logits_ws != nullptr);
  return use_mega2(e, B) || (e->L > 0 && !half_cut(e) && mega_supported(B, e->H, e->nh, e->L, e->cfg.dtype, e->num_sms));
}

int upload_mega_layers(gp_engine* e) {
  if (e->mega_uploaded) return GP_OK;
  if (e->cfg.dtype == GP_BF16) {
    // per-row constants of the folded LayerNorms (k_mega2.cu, stage_ln): recomputed whenever a weight was (re)loaded
    const int H = e->H;
    for (int i = 0; i < e->L; ++i) {
      const LayerW& w = e->layers[i];
      if (has_attn_half(e, i)) CK(e, launch_ln_fold(w.attn_w, w.ln1_g, w.ln1_b, w.attn_b, w.attn_u, w.attn_c, 3 * H, H, e->stream));
      if (has_mlp_half(e, i)) CK(e, launch_ln_fold(w.fc_w, w.ln2_g, w.ln2_b, w.fc_b, w.fc_u, w.fc_c, e->F, H, e->stream));
    }
    if (e->cfg.has_head) CK(e, launch_ln_fold(e->wte, e->lnf_g, e->lnf_b, nullptr, e->lnf_u, e->lnf_c, e->V, H, e->stream));
  }
  std::vector<MegaLayer> h(e->L);
  for (int i = 0; i < e->L; ++i) {
    const LayerW& w = e->layers[i];
    h[i] = MegaLayer{w.attn_w, w.proj_w, w.fc_w, w.mproj_w, w.attn_b, w.proj_b, w.fc_b, w.mproj_b,
                     w.ln1_g,  w.ln1_b,  w.ln2_g, w.ln2_b,  w.kc,     w.vc,      w.attn_u, w.attn_c, w.fc_u, w.fc_c};
  }
  if (e->L > 0) CK(e, cudaMemcpyAsync(e->mega_layers, h.data(), sizeof(MegaLayer) * e->L, cudaMemcpyHostToDevice, e->stream));
  CK(e, cudaStreamSynchronize(e->stream));
  e->mega_uploaded = true;
  return GP_OK;
}

// one launch = n_steps cached decode steps of this stage (n_steps > 1 only for single-stage engines)
struct PipeArgs {
  int n_streams = 1; uint32_t comm_epoch0 = 0; int32_t* tok_out = nullptr; int tok_stride = 0; bool on = false;
  bool slot_indexed = false;   // KV slots come from e->slot_map instead of slot0 + index
};
size_t inbox_bytes(const gp_engine* e) { return 1024 + (size_t)64 * e->H * 8; }

int decode_mega(gp_engine* e, const int32_t* ids, const float* hidden_in, int B, int slot0, float* hidden_out,
                float* logits, int32_t* next_ids, int32_t* seq, int n_steps, bool clock, cudaStream_t s,
                const PipeArgs& pipe = PipeArgs()) {
  int r = upload_mega_layers(e);
  if (r != GP_OK) return r;
  const gp_config& c = e->cfg;
  MegaParams p;
  p.layers = e->mega_layers; p.L = e->L; p.H = e->H; p.nh = e->nh; p.V = e->V; p.n_pos = c.n_positions;
  p.max_ctx = c.max_ctx; p.eps = c.ln_eps; p.B = B; p.slot0 = slot0; p.has_embed = c.has_embed; p.has_head = c.has_head;
  p.wte = e->wte; p.wpe = e->wpe; p.lnf_g = e->lnf_g; p.lnf_b = e->lnf_b; p.lnf_u = e->lnf_u; p.lnf_c = e->lnf_c;
  p.ids = ids; p.hidden_in = hidden_in;
  p.hidden_out = hidden_out; p.lens = e->lens; p.qkv_ll = e->qkv_ll; p.attn_ll = e->attn_ll; p.h2_ll = e->h2_ll;
  p.act_ll = e->act_ll; p.h_ll = e->h_ll; p.part_ll = e->part_ll; p.ids_out = next_ids ? next_ids : e->mailbox;
  p.seq = seq; p.seq_stride = e->seq_stride; p.logits_out = logits; p.n_steps = n_steps;
  p.n_streams = pipe.n_streams;
  if (pipe.on) {
    p.comm_epoch0 = pipe.comm_epoch0; p.tok_out = pipe.tok_out; p.tok_stride = pipe.tok_stride;
    if (c.has_embed) p.inbox_ids = reinterpret_cast<const unsigned long long*>(e->inbox);
    else p.inbox_hid = reinterpret_cast<const unsigned long long*>(e->inbox + 1024);
    if (c.has_head) p.peer_ids = reinterpret_cast<unsigned long long*>(e->peer_inbox);
    else p.peer_hid = reinterpret_cast<unsigned long long*>(e->peer_inbox + 1024);
  }
  const bool v2 = use_mega2(e, B);
  if (pipe.slot_indexed && !v2) { set_err(e, "slot-indexed decode needs the bf16 persistent kernel (B <= 8, head_dim 64)"); return GP_EINVAL; }
  const uint32_t span = v2 ? mega2_epochs_per_launch(e->L, n_steps * pipe.n_streams) : mega_epochs_per_launch(e->L, n_steps * pipe.n_streams);
  if (e->mega_epoch > 0xFFFFFFFFu - span - 8u) {
    // epoch counter about to wrap: start over on zeroed exchange buffers
    CK(e, cudaMemsetAsync(e->qkv_ll, 0, e->ll_bytes[0], s)); CK(e, cudaMemsetAsync(e->attn_ll, 0, e->ll_bytes[1], s));
    CK(e, cudaMemsetAsync(e->h2_ll, 0, e->ll_bytes[2], s)); CK(e, cudaMemsetAsync(e->act_ll, 0, e->ll_bytes[3], s));
    CK(e, cudaMemsetAsync(e->h_ll, 0, e->ll_bytes[4], s)); CK(e, cudaMemsetAsync(e->part_ll, 0, mega_part_words() * 8, s));
    e->mega_epoch = 8;
  }
  p.epoch0 = e->mega_epoch;
  e->mega_epoch += span;
  {
    // Demote the weight stream in L2 only when it cannot stay resident anyway: a stage that holds a few layers of a
    // small model (124M on 8 stages: 21 MB) re-reads its weights from L2 for every stream of a round, and marking
    // them evict-first would send each of those reads back to HBM (measured: 31.8 k -> 27.2 k tok/s on 8 GPUs).
    const size_t esz = bf(e) ? 2 : 4;
    const size_t stage_weights = esz * ((size_t)e->L * (12 * (size_t)e->H * e->H + 13 * (size_t)e->H) +
                                        (c.has_head ? (size_t)e->V * e->H : 0));
    p.l2_stream_hint = stage_weights > ((size_t)96 << 20);
  }
  if (const char* env = getenv("GP_MEGA_L2HINT")) p.l2_stream_hint = atoi(env);   // 0: weight copies at normal L2 priority
  p.step_clock = clock ? e->step_clock : nullptr;
  const char* trace_path = getenv("GP_MEGA_TRACE");
  const bool trace = trace_path != nullptr && n_steps > 1;
  const int kTraceCap = use_mega2(e, B) ? 1024 : 512;   // the second-generation kernel stamps more events per phase
  if (trace) {  // debug aid: per-CTA %globaltimer at every phase boundary of the last step -> raw int64 file
    if (!e->trace) CK(e, cudaMalloc((void**)&e->trace, (size_t)e->num_sms * kTraceCap * sizeof(long long)));
    CK(e, cudaMemsetAsync(e->trace, 0, (size_t)e->num_sms * kTraceCap * sizeof(long long), s));
    p.trace = e->trace; p.trace_cap = kTraceCap;
  }
  if (v2) {
    p.slot_map = pipe.slot_indexed ? e->slot_map : nullptr;
    p.abort_flag = e->ctrl; p.progress = reinterpret_cast<unsigned int*>(e->ctrl + 1);
    p.ll_shift = kLLShiftMax;
    p.begin_half = (c.flags & GP_FLAG_BEGIN_HALF) ? 1 : 0; p.end_half = (c.flags & GP_FLAG_END_HALF) ? 1 : 0;
    if (e->sampler_on && c.has_head) {
      p.sampler_on = 1; p.sample_temperature = e->sampler_temperature; p.sample_top_k = e->sampler_top_k;
      p.sample_seed_lo = (uint32_t)e->sampler_seed; p.sample_seed_hi = (uint32_t)(e->sampler_seed >> 32);
      if (p.logits_out == nullptr) p.logits_out = e->logits_ws;   // the draw reads the logits back
    }
    {  // the host mirror of the cache lengths bounds the contexts of this launch (callers may already have advanced it)
      int longest = 0;
      for (int32_t v : e->hlens) longest = v > longest ? v : longest;
      p.ctx_hint = longest + n_steps < c.max_ctx ? longest + n_steps : c.max_ctx;
    }
    if (const char* env = getenv("GP_MEGA_LLSHIFT")) { int v = atoi(env); if (v >= 2 && v <= kLLShiftMax) p.ll_shift = v; }
    CK(e, cudaMemsetAsync(e->ctrl, 0, 2 * sizeof(int), s));
    cudaError_t le = launch_decode_mega2(p, e->num_sms, s);
    if (le == cudaErrorCooperativeLaunchTooLarge || le == cudaErrorLaunchOutOfResources) {
      cudaGetLastError();
      e->coop_ok = false;   // callers fall back to one kernel per op from now on
      set_err(e, "persistent kernel cannot be co-scheduled here (%s)", cudaGetErrorString(le));
      return GP_ESTATE;
    }
    LAUNCH(e, le);
  } else {
    cudaError_t le = launch_decode_mega(p, c.dtype, e->num_sms, s);
    if (le == cudaErrorCooperativeLaunchTooLarge || le == cudaErrorLaunchOutOfResources) {
      cudaGetLastError();
      e->coop_ok = false;
      set_err(e, "persistent kernel cannot be co-scheduled here (%s)", cudaGetErrorString(le));
      return GP_ESTATE;
    }
    LAUNCH(e, le);
  }
  if (trace) {
    std::vector<long long> t((size_t)e->num_sms * kTraceCap);
    CK(e, cudaStreamSynchronize(s));
    CK(e, cudaMemcpy(t.data(), e->trace, t.size() * sizeof(long long), cudaMemcpyDeviceToHost));
    char path[512];
    snprintf(path, sizeof path, trace_path, c.device);   // "%d" in GP_MEGA_TRACE = device ordinal (one file per pipeline stage)
    FILE* f = fopen(path, "wb");
    if (f) { fwrite(t.data(), sizeof(long long), t.size(), f); fclose(f); }
  }
  return GP_OK;
}

// ------------------------------------------------------------------------------- prefill
int prefill_impl(gp_engine* e, const int32_t* ids, const float* hidden_in, int B, int T, int slot0, int use_cache,
                 float* hidden_out, float* logits, int logits_all, int32_t* next_ids, int32_t* seq_for_generate,
                 cudaStream_t s) {
  const gp_config& c = e->cfg;
  const int H = e->H, F = e->F, M = B * T;
  if (B <= 0 || T <= 0) { set_err(e, "prefill: B and T must be positive (B=%d T=%d)", B, T); return GP_EINVAL; }
  if (T > c.n_positions) {
    set_err(e, "prefill: T=%d exceeds n_positions=%d (the reference would index past wpe)", T, c.n_positions);
    return GP_EINVAL;
  }
  if ((int64_t)M > (int64_t)c.max_tokens) { set_err(e, "prefill: B*T=%d exceeds max_tokens=%d", M, c.max_tokens); return GP_EINVAL; }
  if (use_cache && (slot0 < 0 || slot0 + B > c.max_batch || T > c.max_ctx)) {
    set_err(e, "prefill: slots [%d,%d) / T=%d outside cache (max_batch=%d max_ctx=%d)", slot0, slot0 + B, T,
            c.max_batch, c.max_ctx);
    return GP_EINVAL;
  }
  if (c.has_embed ? ids == nullptr : hidden_in == nullptr) { set_err(e, "prefill: missing input pointer"); return GP_EINVAL; }
  if (!c.has_head && hidden_out == nullptr) { set_err(e, "prefill: hidden_out is NULL"); return GP_EINVAL; }
  int r = check_ready(e);
  if (r != GP_OK) return r;

  const bool obf = bf(e);
  const uint32_t oflag = obf ? GP_OUT_BF16 : 0u;
  if (c.has_embed) {
    LAUNCH(e, launch_embed(ids, nullptr, 0, T, nullptr, 0, e->wte, e->wpe, e->h, M, H, e->V, c.n_positions, c.dtype, s));
  } else {
    CK(e, cudaMemcpyAsync(e->h, hidden_in, (size_t)M * H * sizeof(float), cudaMemcpyDeviceToDevice, s));
  }
  for (int li = 0; li < e->L; ++li) {
    LayerW& w = e->layers[li];
    if (has_attn_half(e, li)) {
      LAUNCH(e, launch_layernorm(e->h, H, w.ln1_g, w.ln1_b, e->xa, obf, M, H, c.ln_eps, s));
      LAUNCH(e, gemm(e, e->xa, H, w.attn_w, w.attn_b, nullptr, 0, e->qkv, 3 * H, M, 3 * H, H, oflag, s));
      if (use_cache) LAUNCH(e, launch_kv_scatter(e->qkv, w.kc, w.vc, B, T, slot0, H, e->nh, e->d, c.max_ctx, c.dtype, s));
      LAUNCH(e, launch_attn_prefill(e->qkv, e->xa, B, T, e->nh, e->d, c.dtype, s));
      LAUNCH(e, gemm(e, e->xa, H, w.proj_w, w.proj_b, e->h, H, e->h, H, M, H, H, 0u, s));
    }
    if (has_mlp_half(e, li)) {
      LAUNCH(e, launch_layernorm(e->h, H, w.ln2_g, w.ln2_b, e->xa, obf, M, H, c.ln_eps, s));
      LAUNCH(e, gemm(e, e->xa, H, w.fc_w, w.fc_b, nullptr, 0, e->act, F, M, F, H, oflag | GP_EPI_GELU, s));
      LAUNCH(e, gemm(e, e->act, F, w.mproj_w, w.mproj_b, e->h, H, e->h, H, M, H, F, 0u, s));
    }
  }
  if (use_cache) LAUNCH(e, launch_set_lens(e->lens, slot0, B, T, 0, s));
  if (!c.has_head) {
    CK(e, cudaMemcpyAsync(hidden_out, e->h, (size_t)M * H * sizeof(float), cudaMemcpyDeviceToDevice, s));
    return GP_OK;
  }
  if (logits && logits_all) {
    // ShardB semantics: logits for ALL T rows (server.py:101-102)
    LAUNCH(e, launch_layernorm(e->h, H, e->lnf_g, e->lnf_b, e->xa, obf, M, H, c.ln_eps, s));
    LAUNCH(e, gemm(e, e->xa, H, e->wte, nullptr, nullptr, 0, logits, e->V, M, e->V, H, 0u, s));
  }
  if (next_ids || seq_for_generate || (logits && !logits_all)) {
    LAUNCH(e, launch_gather_last(e->h, e->hlast, B, T, H, s));
    float* want_lg = (logits && !logits_all) ? logits : nullptr;
    const bool tc_ok = bf(e) && !(c.flags & GP_FLAG_NO_TCGEN05) && B > kGemvMaxBatch && e->logits_ws != nullptr &&
                       B <= c.max_batch && gemm_tc_supported(B, e->V, H) && (int64_t)B <= (int64_t)c.max_tokens;
    if (tc_ok) {
      // many sequences: ln_f -> tcgen05 GEMM over the tied head -> per-row selection, instead of streaming the 77 MB
      // head once per 8 rows
      float* lg = want_lg ? want_lg : e->logits_ws;
      int32_t* ids_dst = next_ids ? next_ids : e->mailbox;
      LAUNCH(e, launch_layernorm(e->hlast, H, e->lnf_g, e->lnf_b, e->xa, true, B, H, c.ln_eps, s));
      LAUNCH(e, gemm(e, e->xa, H, e->wte, nullptr, nullptr, 0, lg, e->V, B, e->V, H, 0u, s));
      if (e->sampler_on)
        LAUNCH(e, launch_sample_topk(lg, e->V, B, e->V, e->sampler_temperature, e->sampler_top_k, e->sampler_seed, slot0, T, ids_dst,
                                     nullptr, slot0, nullptr, e->seq_stride, s));
      else
        LAUNCH(e, launch_argmax_rows(lg, e->V, B, e->V, ids_dst, nullptr, slot0, nullptr, e->seq_stride, s));
    } else {
      // generate: the first new token goes to seq[b][len] with len (= T) unchanged
      r = select_tokens(e, e->hlast, H, B, slot0, want_lg, next_ids, nullptr, nullptr, T, s);
      if (r != GP_OK) return r;
    }
  }
  return GP_OK;
}

// -------------------------------------------------------------------------------- decode
int decode_impl(gp_engine* e, const int32_t* ids, const float* hidden_in, int B, int slot0, float* hidden_out,
                float* logits, int32_t* next_ids, int32_t* seq, cudaStream_t s) {
  const gp_config& c = e->cfg;
  const int H = e->H, F = e->F;
  if (B <= 0 || slot0 < 0 || slot0 + B > c.max_batch) {
    set_err(e, "decode: slots [%d,%d) outside cache (max_batch=%d)", slot0, slot0 + B, c.max_batch);
    return GP_EINVAL;
  }
  if (c.has_embed ? ids == nullptr : hidden_in == nullptr) { set_err(e, "decode: missing input pointer"); return GP_EINVAL; }
  if (!c.has_head && hidden_out == nullptr) { set_err(e, "decode: hidden_out is NULL"); return GP_EINVAL; }
  int r = check_ready(e);
  if (r != GP_OK) return r;

  if (use_mega(e, B)) {
    r = decode_mega(e, ids, hidden_in, B, slot0, hidden_out, logits, next_ids, seq, 1, false, s);
    if (r != GP_ESTATE || e->coop_ok) return r;
    // the cooperative launch was refused (MPS / MIG / SM-limited context): one kernel per op instead
  }

  // GEMV kernels: fp32 mode and odd shapes; in bf16 mode more than 4 rows go to the tcgen05 GEMMs
  const bool tc_ok = bf(e) && !(c.flags & GP_FLAG_NO_TCGEN05) && gemm_tc_supported(B, H, H);
  const bool use_gemv = gemv_supported(B, F) && !(tc_ok && B > 4);
  const bool obf = bf(e);
  const uint32_t oflag = obf ? GP_OUT_BF16 : 0u;

  const float* cur = hidden_in;  // current residual stream [B,H]
  if (e->L == 0 && !c.has_embed) {   // head-only stage: the lm_head below reads e->h
    CK(e, cudaMemcpyAsync(e->h, hidden_in, (size_t)B * H * sizeof(float), cudaMemcpyDeviceToDevice, s));
    cur = e->h;
  }
  if (c.has_embed) {
    LAUNCH(e, launch_embed(ids, nullptr, 0, 1, e->lens, slot0, e->wte, e->wpe, e->h, B, H, e->V, c.n_positions, c.dtype, s));
    cur = e->h;
  }
  for (int li = 0; li < e->L; ++li) {
    LayerW& w = e->layers[li];
    const bool last = (li == e->L - 1);
    const bool attn_half = has_attn_half(e, li), mlp_half = has_mlp_half(e, li);
    float* dst = (last && !c.has_head) ? hidden_out : e->h;          // where the stage's last residual add writes
    float* attn_dst = mlp_half ? e->h : dst;                          // the attention residual is the output of a cut block
    if (use_gemv) {
      if (attn_half) {
        GemvParams p;
        // ln_1 + c_attn, q -> e->q, k/v -> cache[pos = len]
        p.x = cur; p.ldx = H; p.Wt = w.attn_w; p.bias = w.attn_b; p.B = B; p.N = 3 * H; p.K = H;
        p.ln_g = w.ln1_g; p.ln_b = w.ln1_b; p.eps = c.ln_eps; p.mode = 1; p.q_out = e->q; p.kcache = w.kc;
        p.vcache = w.vc; p.lens = e->lens; p.slot0 = slot0; p.H = H; p.nh = e->nh; p.d = e->d; p.max_ctx = c.max_ctx;
        LAUNCH(e, launch_gemv(p, c.dtype, s));
        LAUNCH(e, launch_attn_decode(e->q, w.kc, w.vc, e->lens, 1, slot0, e->attn_f, false, B, e->nh, e->d, c.max_ctx, c.dtype, s));
        // attention c_proj + residual
        GemvParams p2;
        p2.x = e->attn_f; p2.ldx = H; p2.Wt = w.proj_w; p2.bias = w.proj_b; p2.residual = cur; p2.ldr = H;
        p2.out = attn_dst; p2.ldo = H; p2.B = B; p2.N = H; p2.K = H;
        LAUNCH(e, launch_gemv(p2, c.dtype, s));
        cur = attn_dst;
      }
      if (mlp_half) {
        // ln_2 + c_fc + gelu_new
        GemvParams p3;
        p3.x = cur; p3.ldx = H; p3.Wt = w.fc_w; p3.bias = w.fc_b; p3.out = e->act_f; p3.ldo = F; p3.B = B; p3.N = F;
        p3.K = H; p3.ln_g = w.ln2_g; p3.ln_b = w.ln2_b; p3.eps = c.ln_eps; p3.flags = GP_EPI_GELU;
        LAUNCH(e, launch_gemv(p3, c.dtype, s));
        // mlp c_proj + residual
        GemvParams p4;
        p4.x = e->act_f; p4.ldx = F; p4.Wt = w.mproj_w; p4.bias = w.mproj_b; p4.residual = cur; p4.ldr = H;
        p4.out = dst; p4.ldo = H; p4.B = B; p4.N = H; p4.K = F;
        LAUNCH(e, launch_gemv(p4, c.dtype, s));
        cur = dst;
      }
    } else {
      if (B > c.max_tokens) { set_err(e, "decode: B=%d exceeds max_tokens=%d", B, c.max_tokens); return GP_EINVAL; }
      if (attn_half) {
        LAUNCH(e, launch_layernorm(cur, H, w.ln1_g, w.ln1_b, e->xa, obf, B, H, c.ln_eps, s));
        LAUNCH(e, gemm(e, e->xa, H, w.attn_w, w.attn_b, nullptr, 0, e->qkv, 3 * H, B, 3 * H, H, oflag, s));
        LAUNCH(e, launch_qkv_split_decode(e->qkv, e->q, w.kc, w.vc, e->lens, B, slot0, H, e->nh, e->d, c.max_ctx, c.dtype, s));
        LAUNCH(e, launch_attn_decode(e->q, w.kc, w.vc, e->lens, 1, slot0, e->xa, obf, B, e->nh, e->d, c.max_ctx, c.dtype, s));
        LAUNCH(e, gemm(e, e->xa, H, w.proj_w, w.proj_b, cur, H, attn_dst, H, B, H, H, 0u, s));
        cur = attn_dst;
      }
      if (mlp_half) {
        LAUNCH(e, launch_layernorm(cur, H, w.ln2_g, w.ln2_b, e->xa, obf, B, H, c.ln_eps, s));
        LAUNCH(e, gemm(e, e->xa, H, w.fc_w, w.fc_b, nullptr, 0, e->act, F, B, F, H, oflag | GP_EPI_GELU, s));
        LAUNCH(e, gemm(e, e->act, F, w.mproj_w, w.mproj_b, cur, H, dst, H, B, H, F, 0u, s));
        cur = dst;
      }
    }
  }
  if (c.has_head && !use_gemv && tc_ok && e->logits_ws != nullptr && gemm_tc_supported(B, e->V, H)) {
    // batched decode: ln_f -> tcgen05 GEMM over the tied head -> per-row argmax (+ bookkeeping)
    float* lg = logits ? logits : e->logits_ws;
    LAUNCH(e, launch_layernorm(e->h, H, e->lnf_g, e->lnf_b, e->xa, true, B, H, c.ln_eps, s));
    LAUNCH(e, gemm(e, e->xa, H, e->wte, nullptr, nullptr, 0, lg, e->V, B, e->V, H, 0u, s));
    if (e->sampler_on)
      LAUNCH(e, launch_sample_topk(lg, e->V, B, e->V, e->sampler_temperature, e->sampler_top_k, e->sampler_seed, slot0, 0,
                                   next_ids ? next_ids : e->mailbox, e->lens, slot0, seq, e->seq_stride, s));
    else
      LAUNCH(e, launch_argmax_rows(lg, e->V, B, e->V, next_ids ? next_ids : e->mailbox, e->lens, slot0, seq,
                                   e->seq_stride, s));
  } else if (c.has_head) {
    // the last CTA of the (last) lm_head launch advances the cache lengths and appends to seq
    r = select_tokens(e, e->h, H, B, slot0, logits, next_ids, e->lens, seq, 0, s);
    if (r != GP_OK) return r;
  } else {
    LAUNCH(e, launch_set_lens(e->lens, slot0, B, 1, 1, s));
  }
  return GP_OK;
}

int parse_layer(const char* name, int* layer, const char** rest) {
  if (strncmp(name, "h.", 2) != 0) return 0;
  char* end = nullptr;
  long v = strtol(name + 2, &end, 10);
  if (end == name + 2 || *end != '.') return 0;
  *layer = (int)v;
  *rest = end + 1;
  return 1;
}

}  // namespace

namespace gp {
void record_error(const char* msg) { set_err(nullptr, "%s", msg); }
}  // namespace gp

// =========================================================================== C ABI
extern "C" {

int gp_version(void) { return GP_VERSION; }

const char* gp_last_error(const gp_engine* e) { return e ? e->err.c_str() : g_err.c_str(); }

int64_t gp_last_launch_count(const gp_engine* e) { return e ? e->launches : 0; }

int64_t gp_lmhead_scratch_bytes(int32_t B, int32_t V) {
  (void)B; (void)V;
  return kLmheadScratchHeader + (int64_t)kLmheadMaxGrid * kGemvMaxBatch * (sizeof(float) + sizeof(int));
}

int gp_engine_create(const gp_config* cfg, gp_engine** out) {
  if (!cfg || !out) { set_err(nullptr, "gp_engine_create: NULL argument"); return GP_EINVAL; }
  *out = nullptr;
  if (cfg->n_embd <= 0 || cfg->n_head <= 0 || cfg->n_embd % cfg->n_head != 0 || cfg->vocab_size <= 0 ||
      cfg->n_positions <= 0 || cfg->n_layer <= 0 || cfg->layer_begin < 0 || cfg->layer_end > cfg->n_layer ||
      cfg->layer_begin > cfg->layer_end || (cfg->layer_begin == cfg->layer_end && !(cfg->has_head && !cfg->has_embed)) ||
      cfg->max_batch <= 0 || cfg->max_ctx <= 0 ||
      cfg->max_ctx > cfg->n_positions || (cfg->dtype != GP_F32 && cfg->dtype != GP_BF16)) {
    set_err(nullptr, "gp_engine_create: invalid configuration");
    return GP_EINVAL;
  }
  {
    // half-block boundaries: a stage may start at the MLP half of its first block and / or stop after the attention
    // half of its last one; such a stage needs a block to cut and neither the embeddings before nor the head after it
    const bool bh = cfg->flags & GP_FLAG_BEGIN_HALF, eh = cfg->flags & GP_FLAG_END_HALF;
    const int nl = cfg->layer_end - cfg->layer_begin;
    if ((bh && (cfg->has_embed || nl < 1)) || (eh && (cfg->has_head || nl < 1)) || (bh && eh && nl < 2)) {
      set_err(nullptr, "gp_engine_create: GP_FLAG_BEGIN_HALF / GP_FLAG_END_HALF do not fit this stage (blocks [%d,%d), has_embed=%d, has_head=%d)",
              cfg->layer_begin, cfg->layer_end, cfg->has_embed, cfg->has_head);
      return GP_EINVAL;
    }
  }
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
    set_err(nullptr, "gp_engine_create: CUDA device %d unavailable (%s); this library has no CPU path", cfg->device,
            ce == cudaSuccess ? "no such ordinal" : cudaGetErrorString(ce));
    return GP_ECUDA;
  }
  gp_engine* e = new gp_engine();
  e->cfg = *cfg;
  if (e->cfg.max_tokens <= 0) e->cfg.max_tokens = cfg->max_batch * cfg->max_ctx;
  if (e->cfg.max_tokens < cfg->max_batch) e->cfg.max_tokens = cfg->max_batch;
  e->H = cfg->n_embd; e->nh = cfg->n_head; e->d = e->H / e->nh; e->V = cfg->vocab_size; e->F = 4 * e->H;
  e->L = cfg->layer_end - cfg->layer_begin;
  e->esz = cfg->dtype == GP_BF16 ? 2 : 4;
  e->seq_stride = cfg->n_positions + 1;
  e->hlens.assign(cfg->max_batch, 0);
  auto fail = [&](int rc) { gp_engine_destroy(e); return rc; };
#define TRY(x) do { int _r = (x); if (_r != GP_OK) return fail(_r); } while (0)
#define TRYCK(call) do { cudaError_t _c = (call); if (_c != cudaSuccess) { set_err(nullptr, "%s: %s", #call, cudaGetErrorString(_c)); return fail(GP_ECUDA); } } while (0)
  DeviceGuard dev_guard(cfg->device);
  TRYCK(dev_guard.err);
  TRYCK(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  TRYCK(init_gemv_attrs());
  TRYCK(init_attn_attrs());
  if (cfg->dtype == GP_BF16 && !(cfg->flags & GP_FLAG_NO_TCGEN05)) TRYCK(init_gemm_tc());

  const size_t H = e->H, F = e->F, V = e->V, esz = e->esz;
  auto alloc = [&](void** p, size_t bytes) -> int { return dev_alloc(e, p, bytes); };
  e->layers.resize(e->L);
  const size_t kv_elems = (size_t)cfg->max_batch * e->nh * cfg->max_ctx * e->d;
  for (int li = 0; li < e->L; ++li) {
    LayerW& w = e->layers[li];
    const std::string p = "h." + std::to_string(cfg->layer_begin + li) + ".";
    TRY(alloc((void**)&w.ln1_g, H * 4)); TRY(alloc((void**)&w.ln1_b, H * 4));
    TRY(alloc((void**)&w.ln2_g, H * 4)); TRY(alloc((void**)&w.ln2_b, H * 4));
    TRY(alloc((void**)&w.attn_b, 3 * H * 4)); TRY(alloc((void**)&w.proj_b, H * 4));
    TRY(alloc((void**)&w.fc_b, F * 4)); TRY(alloc((void**)&w.mproj_b, H * 4));
    TRY(alloc(&w.attn_w, 3 * H * H * esz)); TRY(alloc(&w.proj_w, H * H * esz));
    TRY(alloc(&w.fc_w, F * H * esz)); TRY(alloc(&w.mproj_w, H * F * esz));
    TRY(alloc(&w.kc, kv_elems * esz)); TRY(alloc(&w.vc, kv_elems * esz));
    if (cfg->dtype == GP_BF16) {
      TRY(alloc((void**)&w.attn_u, 3 * H * 4)); TRY(alloc((void**)&w.attn_c, 3 * H * 4));
      TRY(alloc((void**)&w.fc_u, F * 4)); TRY(alloc((void**)&w.fc_c, F * 4));
    }
    if (has_attn_half(e, li))
      for (const char* n : {"ln_1.weight", "ln_1.bias", "attn.c_attn.weight", "attn.c_attn.bias", "attn.c_proj.weight",
                            "attn.c_proj.bias"})
        e->required.insert(p + n);
    if (has_mlp_half(e, li))
      for (const char* n : {"ln_2.weight", "ln_2.bias", "mlp.c_fc.weight", "mlp.c_fc.bias", "mlp.c_proj.weight",
                            "mlp.c_proj.bias"})
        e->required.insert(p + n);
  }
  if (cfg->has_embed || cfg->has_head) {
    TRY(alloc(&e->wte, V * H * esz));
    e->required.insert("wte.weight");
  }
  if (cfg->has_embed) {
    TRY(alloc(&e->wpe, (size_t)cfg->n_positions * H * esz));
    e->required.insert("wpe.weight");
  }
  if (cfg->has_head) {
    TRY(alloc((void**)&e->lnf_g, H * 4)); TRY(alloc((void**)&e->lnf_b, H * 4));
    if (cfg->dtype == GP_BF16) { TRY(alloc((void**)&e->lnf_u, V * 4 + 16)); TRY(alloc((void**)&e->lnf_c, V * 4 + 16));   /* + 16: the kernel copies 16-byte aligned windows of these */ }
    e->required.insert("ln_f.weight"); e->required.insert("ln_f.bias");
    TRY(alloc(&e->lm_scratch, (size_t)gp_lmhead_scratch_bytes(kGemvMaxBatch, e->V)));
    TRYCK(cudaMemsetAsync(e->lm_scratch, 0, kLmheadScratchHeader, e->stream));
  }
  const size_t MT = e->cfg.max_tokens, MB = cfg->max_batch;
  TRY(alloc((void**)&e->lens, MB * sizeof(int32_t)));
  TRYCK(cudaMemsetAsync(e->lens, 0, MB * sizeof(int32_t), e->stream));
  TRY(alloc((void**)&e->h, MT * H * 4));
  TRY(alloc(&e->xa, MT * H * esz));
  TRY(alloc(&e->qkv, MT * 3 * H * esz));
  TRY(alloc(&e->act, MT * F * esz));
  TRY(alloc((void**)&e->q, MB * H * 4));
  TRY(alloc((void**)&e->attn_f, MB * H * 4));
  TRY(alloc((void**)&e->act_f, MB * F * 4));
  TRY(alloc((void**)&e->hlast, MB * H * 4));
  if (cfg->dtype == GP_BF16 && !(cfg->flags & GP_FLAG_NO_TCGEN05)) {
    e->splitk_floats = (size_t)8 * 256 * (size_t)(F > V ? F : (cfg->has_head ? V : F));
    if (e->splitk_floats > ((size_t)48 << 20)) e->splitk_floats = (size_t)48 << 20;   // 192 MB cap
    TRY(alloc((void**)&e->splitk, e->splitk_floats * 4));
  }
  if (cfg->has_head) TRY(alloc((void**)&e->logits_ws, MB * V * 4));   // GEMM-path lm_head and the sampler
  TRY(alloc((void**)&e->seq, MB * (size_t)e->seq_stride * sizeof(int32_t)));
  TRY(alloc((void**)&e->prompt, MB * (size_t)cfg->n_positions * sizeof(int32_t)));
  TRY(alloc((void**)&e->mailbox, MB * sizeof(int32_t)));
  TRY(alloc((void**)&e->mega_layers, sizeof(MegaLayer) * (e->L > 0 ? e->L : 1)));
  {
    // exchange buffers: dense [max_batch][width] words for the fp32 kernel; the bf16 kernel spreads the words of at
    // most 8 sequences over L2 slices (one 32-byte sector per (4 << (shift - 2)) words, k_mega2.cu ll_off)
    auto ll_bytes = [&](size_t width) {
      const size_t dense = MB * width * 8;
      const size_t rows = 8;   // the kernel addresses up to 8 sequences (and the attention partials behind 8 rows) whatever max_batch is
      const size_t spread = cfg->dtype == GP_BF16 ? ((rows * width + 3) / 4 << kLLShiftMax) * 8 + 64 : 0;
      return dense > spread ? dense : spread;
    };
    e->ll_bytes[0] = ll_bytes(3 * H);
    e->ll_bytes[1] = ll_bytes(H + (size_t)e->nh * 4 * 72);   // + the partials of split attention tasks (k_mega2.cu: kMaxSplit x kPartStride per head)
    e->ll_bytes[2] = ll_bytes(H); e->ll_bytes[3] = ll_bytes(F);
    e->ll_bytes[4] = ll_bytes(H);
    TRY(alloc((void**)&e->qkv_ll, e->ll_bytes[0])); TRYCK(cudaMemsetAsync(e->qkv_ll, 0, e->ll_bytes[0], e->stream));
    TRY(alloc((void**)&e->attn_ll, e->ll_bytes[1])); TRYCK(cudaMemsetAsync(e->attn_ll, 0, e->ll_bytes[1], e->stream));
    TRY(alloc((void**)&e->h2_ll, e->ll_bytes[2])); TRYCK(cudaMemsetAsync(e->h2_ll, 0, e->ll_bytes[2], e->stream));
    TRY(alloc((void**)&e->act_ll, e->ll_bytes[3])); TRYCK(cudaMemsetAsync(e->act_ll, 0, e->ll_bytes[3], e->stream));
    TRY(alloc((void**)&e->h_ll, e->ll_bytes[4])); TRYCK(cudaMemsetAsync(e->h_ll, 0, e->ll_bytes[4], e->stream));
  }
  TRY(alloc((void**)&e->part_ll, mega_part_words() * 8)); TRYCK(cudaMemsetAsync(e->part_ll, 0, mega_part_words() * 8, e->stream));
  TRY(alloc((void**)&e->step_clock, (size_t)(cfg->n_positions + 2) * 64 * sizeof(unsigned long long)));
  TRYCK(cudaDeviceGetAttribute(&e->num_sms, cudaDevAttrMultiProcessorCount, cfg->device));
  if (e->num_sms > 1024) e->num_sms = 1024;
  TRYCK(init_mega_attrs());
  TRYCK(init_mega2_attrs());
  TRY(alloc((void**)&e->ctrl, 2 * sizeof(int)));
  TRYCK(cudaMemsetAsync(e->ctrl, 0, 2 * sizeof(int), e->stream));
  TRY(alloc((void**)&e->slot_map, 64 * sizeof(int32_t)));
  TRYCK(cudaHostAlloc((void**)&e->ctrl_host, 2 * sizeof(int), cudaHostAllocDefault));
  TRYCK(cudaStreamCreateWithFlags(&e->ctrl_stream, cudaStreamNonBlocking));
  {
    // The persistent kernels spin on each other: the whole grid must be co-resident.  Under MPS / MIG / a green or
    // SM-limited context that may be impossible; then decode runs one kernel per op.
    int coop = 0;
    TRYCK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, cfg->device));
    e->coop_ok = coop != 0;
    if (e->coop_ok && cfg->dtype == GP_BF16 && mega2_supported(1, e->H, e->nh, e->L, cfg->dtype, e->num_sms))
      e->coop_ok = mega2_max_coresident(1, e->H, e->L) >= 1;
  }
  TRYCK(cudaStreamSynchronize(e->stream));
#undef TRY
#undef TRYCK
  *out = e;
  return GP_OK;
}

int gp_engine_destroy(gp_engine* e) {
  if (!e) return GP_OK;
  DeviceGuard dev_guard(e->cfg.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (auto& kv : e->decode_graphs) cudaGraphExecDestroy(kv.second);
  for (auto& kv : e->prefill_graphs) cudaGraphExecDestroy(kv.second);
  for (auto ev : e->events) cudaEventDestroy(ev);
  for (void* p : e->allocs) cudaFree(p);
  if (e->stage) cudaFree(e->stage);
  if (e->trace) cudaFree(e->trace);
  if (e->peer_inbox && e->peer_is_ipc) cudaIpcCloseMemHandle(e->peer_inbox);
  if (e->inbox) cudaFree(e->inbox);
  if (e->ctrl_host) cudaFreeHost(e->ctrl_host);
  if (e->ctrl_stream) cudaStreamDestroy(e->ctrl_stream);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
  return GP_OK;
}

int gp_load_weight(gp_engine* e, const char* name, const float* data_host, const int64_t* shape, int32_t ndim) {
  if (!e || !name || !data_host || !shape || ndim < 1 || ndim > 2) { set_err(e, "gp_load_weight: bad argument"); return GP_EINVAL; }
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  const int64_t H = e->H, F = e->F;
  void* dst = nullptr;
  int64_t R = shape[0], C = ndim == 2 ? shape[1] : 1, eR = 0, eC = 1;
  bool transpose = false, keep_f32 = false;
  std::string nm(name);
  int layer = -1;
  const char* rest = nullptr;
  if (nm == "wte.weight" && e->wte) { dst = e->wte; eR = e->V; eC = H; }
  else if (nm == "wpe.weight" && e->wpe) { dst = e->wpe; eR = e->cfg.n_positions; eC = H; }
  else if (nm == "ln_f.weight" && e->lnf_g) { dst = e->lnf_g; eR = H; keep_f32 = true; }
  else if (nm == "ln_f.bias" && e->lnf_b) { dst = e->lnf_b; eR = H; keep_f32 = true; }
  else if (parse_layer(name, &layer, &rest) && layer >= e->cfg.layer_begin && layer < e->cfg.layer_end) {
    LayerW& w = e->layers[layer - e->cfg.layer_begin];
    std::string r(rest);
    if (r == "ln_1.weight") { dst = w.ln1_g; eR = H; keep_f32 = true; }
    else if (r == "ln_1.bias") { dst = w.ln1_b; eR = H; keep_f32 = true; }
    else if (r == "ln_2.weight") { dst = w.ln2_g; eR = H; keep_f32 = true; }
    else if (r == "ln_2.bias") { dst = w.ln2_b; eR = H; keep_f32 = true; }
    else if (r == "attn.c_attn.bias") { dst = w.attn_b; eR = 3 * H; keep_f32 = true; }
    else if (r == "attn.c_proj.bias") { dst = w.proj_b; eR = H; keep_f32 = true; }
    else if (r == "mlp.c_fc.bias") { dst = w.fc_b; eR = F; keep_f32 = true; }
    else if (r == "mlp.c_proj.bias") { dst = w.mproj_b; eR = H; keep_f32 = true; }
    // Conv1D weights arrive [in,out] ([hf] pytorch_utils.py:97-123) and are stored [out,in]
    else if (r == "attn.c_attn.weight") { dst = w.attn_w; eR = H; eC = 3 * H; transpose = true; }
    else if (r == "attn.c_proj.weight") { dst = w.proj_w; eR = H; eC = H; transpose = true; }
    else if (r == "mlp.c_fc.weight") { dst = w.fc_w; eR = H; eC = F; transpose = true; }
    else if (r == "mlp.c_proj.weight") { dst = w.mproj_w; eR = F; eC = H; transpose = true; }
  }
  if (!dst) { set_err(e, "gp_load_weight: '%s' does not belong to this stage (layers [%d,%d) embed=%d head=%d)", name, e->cfg.layer_begin, e->cfg.layer_end, e->cfg.has_embed, e->cfg.has_head); return GP_EINVAL; }
  if (R != eR || C != eC) { set_err(e, "gp_load_weight: '%s' has shape [%lld,%lld], expected [%lld,%lld]", name, (long long)R, (long long)C, (long long)eR, (long long)eC); return GP_EINVAL; }
  const size_t n = (size_t)R * C;
  if (keep_f32) {
    // stream-ordered copy + sync: a plain cudaMemcpy from pageable memory may return before
    // the DMA lands and is only ordered against the legacy stream
    CK(e, cudaMemcpyAsync(dst, data_host, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(e, cudaStreamSynchronize(e->stream));
  } else {
    if (e->stage_floats < n) {
      if (e->stage) cudaFree(e->stage);
      e->stage = nullptr; e->stage_floats = 0;
      CK(e, cudaMalloc((void**)&e->stage, n * sizeof(float)));
      e->stage_floats = n;
    }
    CK(e, cudaMemcpyAsync(e->stage, data_host, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(e, launch_convert(e->stage, dst, (int)R, (int)C, transpose, e->cfg.dtype, e->stream));
    CK(e, cudaStreamSynchronize(e->stream));
  }
  e->loaded.insert(nm);
  e->mega_uploaded = false;   // derived constants (folded LayerNorms) follow the weights
  return GP_OK;
}

int gp_weights_ready(gp_engine* e) {
  if (!e) return GP_EINVAL;
  int r = check_ready(e);
  if (r == GP_OK && e->stage) { cudaFree(e->stage); e->stage = nullptr; e->stage_floats = 0; }
  return r;
}

int gp_kv_reset(gp_engine* e, int32_t slot, void* stream) {
  if (!e || slot >= e->cfg.max_batch) { set_err(e, "gp_kv_reset: bad slot"); return GP_EINVAL; }
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  cudaStream_t s = (cudaStream_t)stream;
  if (slot < 0) {
    CK(e, cudaMemsetAsync(e->lens, 0, e->cfg.max_batch * sizeof(int32_t), s));
    std::fill(e->hlens.begin(), e->hlens.end(), 0);
  } else {
    CK(e, cudaMemsetAsync(e->lens + slot, 0, sizeof(int32_t), s));
    e->hlens[slot] = 0;
  }
  return GP_OK;
}

int gp_kv_lengths(gp_engine* e, int32_t* lens_host, int32_t n, void* stream) {
  if (!e || !lens_host || n < 0 || n > e->cfg.max_batch) { set_err(e, "gp_kv_lengths: bad argument"); return GP_EINVAL; }
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  CK(e, cudaMemcpyAsync(lens_host, e->lens, n * sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CK(e, cudaStreamSynchronize((cudaStream_t)stream));
  return GP_OK;
}

int gp_stage_prefill(gp_engine* e, const int32_t* ids, const float* hidden_in, int32_t B, int32_t T, int32_t slot0,
                     int32_t use_cache, float* hidden_out, float* logits, int32_t logits_all, int32_t* next_ids,
                     void* stream) {
  if (!e) return GP_EINVAL;
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  e->launches = 0;
  int r = prefill_impl(e, ids, hidden_in, B, T, slot0, use_cache, hidden_out, logits, logits_all, next_ids, nullptr,
                       (cudaStream_t)stream);
  if (r == GP_OK && use_cache)
    for (int b = 0; b < B; ++b) e->hlens[slot0 + b] = T;
  return r;
}

int gp_stage_decode(gp_engine* e, const int32_t* ids, const float* hidden_in, int32_t B, int32_t slot0,
                    float* hidden_out, float* logits, int32_t* next_ids, void* stream) {
  if (!e) return GP_EINVAL;
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  e->launches = 0;
  if (B > 0 && slot0 >= 0 && slot0 + B <= e->cfg.max_batch) {
    int room = slot_room(e, "gp_stage_decode", slot0, B, 1);
    if (room != GP_OK) return room;
  }
  int r = decode_impl(e, ids, hidden_in, B, slot0, hidden_out, logits, next_ids, nullptr, (cudaStream_t)stream);
  if (r == GP_OK)
    for (int b = 0; b < B; ++b) ++e->hlens[slot0 + b];
  return r;
}

int gp_set_sampler(gp_engine* e, float temperature, int32_t top_k, uint64_t seed) {
  if (!e) return GP_EINVAL;
  if (!(temperature > 0.f) || top_k <= 0) {   // greedy
    e->sampler_on = false;
  } else {
    if (top_k > kMaxTopK) { set_err(e, "gp_set_sampler: top_k %d exceeds the supported maximum of %d", top_k, kMaxTopK); return GP_EINVAL; }
    if (!e->cfg.has_head) { set_err(e, "gp_set_sampler: this stage has no lm_head"); return GP_EINVAL; }
    e->sampler_on = true;
    e->sampler_temperature = temperature;
    e->sampler_top_k = top_k;
    e->sampler_seed = seed;
  }
  // captured steps carry the sampler's parameters as kernel arguments
  for (auto it = e->prefill_graphs.begin(); it != e->prefill_graphs.end();) {
    if (it->first & 1) {
      cudaGraphExecDestroy(it->second);
      e->prefill_graph_launches.erase(it->first);
      it = e->prefill_graphs.erase(it);
    } else {
      ++it;
    }
  }
  for (auto it = e->decode_graphs.begin(); it != e->decode_graphs.end();) {
    if (it->first >= kSampledKey) {
      cudaGraphExecDestroy(it->second);
      e->decode_graph_launches.erase(it->first);
      it = e->decode_graphs.erase(it);
    } else {
      ++it;
    }
  }
  return GP_OK;
}

int gp_generate(gp_engine* e, const int32_t* prompt_ids_host, int32_t B, int32_t P, int32_t n_new,
                int32_t* out_ids_host, float* timings_ms_host) {
  if (!e || !prompt_ids_host || !out_ids_host) { set_err(e, "gp_generate: NULL argument"); return GP_EINVAL; }
  const gp_config& c = e->cfg;
  if (!c.has_embed || !c.has_head) { set_err(e, "gp_generate needs a single-stage engine (has_embed && has_head)"); return GP_EINVAL; }
  if (B <= 0 || B > c.max_batch || P <= 0 || n_new < 0) { set_err(e, "gp_generate: bad B/P/n_new (%d,%d,%d)", B, P, n_new); return GP_EINVAL; }
  // positions 0 .. P+n_new-2 are run through the model; the reference has no such guard and
  // would raise IndexError past wpe
  if (P + n_new - 1 > c.max_ctx || P > c.max_ctx || (int64_t)B * P > c.max_tokens) {
    set_err(e, "gp_generate: prompt %d + %d new tokens exceeds the context of %d positions", P, n_new, c.max_ctx);
    return GP_EINVAL;
  }
  for (int64_t i = 0; i < (int64_t)B * P; ++i)
    if (prompt_ids_host[i] < 0 || prompt_ids_host[i] >= e->V) {
      set_err(e, "gp_generate: token id %d out of range [0,%d)", prompt_ids_host[i], e->V);
      return GP_EINVAL;
    }
  DeviceGuard dev_guard(c.device);
  CK(e, dev_guard.err);
  cudaStream_t s = e->stream;
  e->launches = 0;
  const int total = P + n_new;
  const int n_ev = 3 + (timings_ms_host ? n_new : 0);
  while ((int)e->events.size() < n_ev) {
    cudaEvent_t ev;
    CK(e, cudaEventCreate(&ev));
    e->events.push_back(ev);
  }
  CK(e, cudaMemcpyAsync(e->prompt, prompt_ids_host, (size_t)B * P * sizeof(int32_t), cudaMemcpyHostToDevice, s));
  CK(e, cudaMemcpy2DAsync(e->seq, e->seq_stride * sizeof(int32_t), prompt_ids_host, P * sizeof(int32_t),
                          P * sizeof(int32_t), B, cudaMemcpyHostToDevice, s));
  CK(e, cudaMemsetAsync(e->lens, 0, c.max_batch * sizeof(int32_t), s));
  std::fill(e->hlens.begin(), e->hlens.end(), 0);
  if (n_new > 0)
    for (int b = 0; b < B; ++b) e->hlens[b] = P + n_new - 1;
  CK(e, cudaEventRecord(e->events[0], s));
  if (n_new > 0) {
    // the prefill works on engine-owned buffers only, so its launch sequence is replayable for a given shape
    const bool graph_ok = !(c.flags & GP_FLAG_NO_GRAPH);
    const int64_t pkey = ((int64_t)B << 32) | ((int64_t)P << 1) | (e->sampler_on ? 1 : 0);
    auto pit = e->prefill_graphs.find(pkey);
    if (graph_ok && pit == e->prefill_graphs.end()) {
      if (e->prefill_graphs.size() >= 64) {   // a server sees many prompt lengths: bound the cache
        for (auto& kv : e->prefill_graphs) cudaGraphExecDestroy(kv.second);
        e->prefill_graphs.clear();
        e->prefill_graph_launches.clear();
      }
      const int64_t before = e->launches;
      CK(e, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
      int r = prefill_impl(e, e->prompt, nullptr, B, P, 0, 1, nullptr, nullptr, 0, e->mailbox, nullptr, s);
      cudaGraph_t g = nullptr;
      cudaError_t ce = cudaStreamEndCapture(s, &g);
      if (r != GP_OK) { if (g) cudaGraphDestroy(g); return r; }
      CK(e, ce);
      cudaGraphExec_t exec = nullptr;
      CK(e, cudaGraphInstantiate(&exec, g, 0));
      cudaGraphDestroy(g);
      e->prefill_graph_launches[pkey] = e->launches - before;
      e->launches = before;
      pit = e->prefill_graphs.emplace(pkey, exec).first;
    }
    if (graph_ok) {
      CK(e, cudaGraphLaunch(pit->second, s));
      e->launches += e->prefill_graph_launches[pkey];
    } else {
      int r = prefill_impl(e, e->prompt, nullptr, B, P, 0, 1, nullptr, nullptr, 0, e->mailbox, nullptr, s);
      if (r != GP_OK) return r;
    }
    // first new token: mailbox -> seq[:, P]
    CK(e, cudaMemcpy2DAsync(e->seq + P, e->seq_stride * sizeof(int32_t), e->mailbox, sizeof(int32_t), sizeof(int32_t), B,
                            cudaMemcpyDeviceToDevice, s));
  }
  CK(e, cudaEventRecord(e->events[1], s));
  const int steps = n_new > 1 ? n_new - 1 : 0;
  bool mega = steps > 0 && use_mega(e, B);
  if (mega) {
    // the whole greedy loop is ONE persistent launch; per-step times come from %globaltimer
    int r = decode_mega(e, e->mailbox, nullptr, B, 0, nullptr, nullptr, e->mailbox, e->seq, steps,
                        timings_ms_host != nullptr, s);
    if (r == GP_ESTATE && !e->coop_ok) mega = false;   // cooperative launch refused: one kernel per op below
    else if (r != GP_OK) return r;
  }
  if (mega) {
  } else if (steps > 0) {
    const bool use_graph = !(c.flags & GP_FLAG_NO_GRAPH);
    cudaGraphExec_t exec = nullptr;
    int64_t per_step = 0;
    if (use_graph) {
      const int gkey = B + (e->sampler_on ? kSampledKey : 0);
      auto it = e->decode_graphs.find(gkey);
      if (it == e->decode_graphs.end()) {
        const int64_t before = e->launches;
        CK(e, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        int r = decode_impl(e, e->mailbox, nullptr, B, 0, nullptr, nullptr, e->mailbox, e->seq, s);
        cudaGraph_t g = nullptr;
        cudaError_t ce = cudaStreamEndCapture(s, &g);
        if (r != GP_OK) { if (g) cudaGraphDestroy(g); return r; }
        CK(e, ce);
        per_step = e->launches - before;
        e->launches = before;
        CK(e, cudaGraphInstantiate(&exec, g, 0));
        cudaGraphDestroy(g);
        e->decode_graphs[gkey] = exec;
        e->decode_graph_launches[gkey] = per_step;
      } else {
        exec = it->second;
        per_step = e->decode_graph_launches[gkey];
      }
    }
    for (int i = 0; i < steps; ++i) {
      if (use_graph) {
        CK(e, cudaGraphLaunch(exec, s));
        e->launches += per_step;
      } else {
        int r = decode_impl(e, e->mailbox, nullptr, B, 0, nullptr, nullptr, e->mailbox, e->seq, s);
        if (r != GP_OK) return r;
      }
      if (timings_ms_host) CK(e, cudaEventRecord(e->events[3 + i], s));
    }
  }
  CK(e, cudaEventRecord(e->events[2], s));
  CK(e, cudaMemcpy2DAsync(out_ids_host, total * sizeof(int32_t), e->seq, e->seq_stride * sizeof(int32_t),
                          total * sizeof(int32_t), B, cudaMemcpyDeviceToHost, s));
  CK(e, cudaStreamSynchronize(s));
  if (timings_ms_host) {
    float ms = 0.f;
    CK(e, cudaEventElapsedTime(&ms, e->events[0], e->events[1]));
    timings_ms_host[0] = ms;
    CK(e, cudaEventElapsedTime(&ms, e->events[1], e->events[2]));
    timings_ms_host[1] = ms;
    for (int i = 0; i < n_new; ++i) timings_ms_host[2 + i] = 0.f;
    if (mega) {
      std::vector<unsigned long long> clk(steps + 1);
      CK(e, cudaMemcpy(clk.data(), e->step_clock, sizeof(unsigned long long) * (steps + 1), cudaMemcpyDeviceToHost));
      for (int i = 0; i < steps; ++i)
        timings_ms_host[2 + i] = (float)((double)(clk[i] - (i == 0 ? clk[steps] : clk[i - 1])) * 1e-6);
    } else {
      for (int i = 0; i < steps; ++i) {
        CK(e, cudaEventElapsedTime(&ms, i == 0 ? e->events[1] : e->events[3 + i - 1], e->events[3 + i]));
        timings_ms_host[2 + i] = ms;
      }
    }
  }
  return GP_OK;
}

// ------------------------------------------------------------------- pipeline stage boundary
int gp_comm_export(gp_engine* e, void* handle_host) {
  if (!e || !handle_host) { set_err(e, "gp_comm_export: NULL argument"); return GP_EINVAL; }
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  if (!e->inbox) {
    CK(e, cudaMalloc((void**)&e->inbox, inbox_bytes(e)));
    CK(e, cudaMemset(e->inbox, 0, inbox_bytes(e)));
    CK(e, cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t h;
  CK(e, cudaIpcGetMemHandle(&h, e->inbox));
  memcpy(handle_host, &h, sizeof h);
  return GP_OK;
}

int gp_comm_connect(gp_engine* e, const void* peer_handle_host) {
  if (!e || !peer_handle_host) { set_err(e, "gp_comm_connect: NULL argument"); return GP_EINVAL; }
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  if (e->peer_inbox && e->peer_is_ipc) { cudaIpcCloseMemHandle(e->peer_inbox); e->peer_inbox = nullptr; }
  cudaIpcMemHandle_t h;
  memcpy(&h, peer_handle_host, sizeof h);
  void* ptr = nullptr;
  CK(e, cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
  e->peer_inbox = reinterpret_cast<unsigned char*>(ptr);
  e->peer_is_ipc = true;
  return GP_OK;
}

int gp_comm_connect_local(gp_engine* e, gp_engine* peer) {
  if (!e || !peer) { set_err(e, "gp_comm_connect_local: NULL argument"); return GP_EINVAL; }
  if (!peer->inbox) {
    DeviceGuard peer_guard(peer->cfg.device);
    CK(e, peer_guard.err);
    CK(e, cudaMalloc((void**)&peer->inbox, inbox_bytes(peer)));
    CK(e, cudaMemset(peer->inbox, 0, inbox_bytes(peer)));
    CK(e, cudaDeviceSynchronize());
  }
  if (inbox_bytes(peer) != inbox_bytes(e)) { set_err(e, "gp_comm_connect_local: the two stages have different hidden sizes"); return GP_EINVAL; }
  DeviceGuard dev_guard(e->cfg.device);
  CK(e, dev_guard.err);
  if (peer->cfg.device != e->cfg.device) {
    int can = 0;
    CK(e, cudaDeviceCanAccessPeer(&can, e->cfg.device, peer->cfg.device));
    if (!can) { set_err(e, "gp_comm_connect_local: device %d cannot access device %d", e->cfg.device, peer->cfg.device); return GP_ECUDA; }
    cudaError_t pe = cudaDeviceEnablePeerAccess(peer->cfg.device, 0);
    if (pe == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    else CK(e, pe);
  }
  if (e->peer_inbox && e->peer_is_ipc) cudaIpcCloseMemHandle(e->peer_inbox);
  e->peer_inbox = peer->inbox;
  e->peer_is_ipc = false;
  return GP_OK;
}

int gp_stage_decode_slots(gp_engine* e, const int32_t* ids, const float* hidden_in, int32_t B, const int32_t* slots_host,
                          float* hidden_out, float* logits, int32_t* next_ids, void* stream) {
  if (!e || !slots_host) { set_err(e, "gp_stage_decode_slots: NULL argument"); return GP_EINVAL; }
  const gp_config& c = e->cfg;
  if (B < 1 || B > 8) { set_err(e, "gp_stage_decode_slots: B=%d outside [1,8]", B); return GP_EINVAL; }
  for (int i = 0; i < B; ++i) {
    if (slots_host[i] < 0 || slots_host[i] >= c.max_batch) { set_err(e, "gp_stage_decode_slots: slot %d outside [0,%d)", slots_host[i], c.max_batch); return GP_EINVAL; }
    for (int j = 0; j < i; ++j)
      if (slots_host[j] == slots_host[i]) { set_err(e, "gp_stage_decode_slots: slot %d listed twice", slots_host[i]); return GP_EINVAL; }
    int room = slot_room(e, "gp_stage_decode_slots", slots_host[i], 1, 1);
    if (room != GP_OK) return room;
  }
  if (c.has_embed ? ids == nullptr : hidden_in == nullptr) { set_err(e, "gp_stage_decode_slots: missing input pointer"); return GP_EINVAL; }
  if (!c.has_head && hidden_out == nullptr) { set_err(e, "gp_stage_decode_slots: hidden_out is NULL"); return GP_EINVAL; }
  DeviceGuard dev_guard(c.device);
  CK(e, dev_guard.err);
  int r = check_ready(e);
  if (r != GP_OK) return r;
  if (!use_mega(e, B) || !use_mega2(e, B)) {
    set_err(e, "gp_stage_decode_slots: needs the bf16 persistent kernel (head_dim 64, B <= 8, greedy selection)");
    return GP_EINVAL;
  }
  cudaStream_t s = (cudaStream_t)stream;
  e->launches = 0;
  CK(e, cudaMemcpyAsync(e->slot_map, slots_host, B * sizeof(int32_t), cudaMemcpyHostToDevice, s));
  PipeArgs pa;
  pa.slot_indexed = true;
  r = decode_mega(e, ids, hidden_in, B, 0, hidden_out, logits, next_ids, nullptr, 1, false, s, pa);
  if (r == GP_OK)
    for (int i = 0; i < B; ++i) ++e->hlens[slots_host[i]];
  return r;
}

int gp_pipeline_decode(gp_engine* e, const int32_t* first_ids, int32_t n_rounds, int32_t n_streams, int32_t B,
                       uint32_t comm_epoch0, int32_t* tokens_out, int32_t timeout_ms, float* ms_host, void* stream) {
  if (!e) return GP_EINVAL;
  const gp_config& c = e->cfg;
  if (c.has_embed && c.has_head) { set_err(e, "gp_pipeline_decode is for multi-stage pipelines; use gp_generate on a single stage"); return GP_EINVAL; }
  if (n_rounds < 1 || n_streams < 1 || B < 1 || n_streams * B > 64 || n_streams * B > c.max_batch || comm_epoch0 == 0) {
    set_err(e, "gp_pipeline_decode: bad n_rounds/n_streams/B (%d,%d,%d) for max_batch %d", n_rounds, n_streams, B, c.max_batch);
    return GP_EINVAL;
  }
  if (!e->inbox || !e->peer_inbox) { set_err(e, "gp_pipeline_decode: stage boundary not connected (gp_comm_export / gp_comm_connect)"); return GP_ESTATE; }
  if (c.has_embed && (!first_ids || !tokens_out)) { set_err(e, "gp_pipeline_decode: stage 0 needs first_ids and tokens_out"); return GP_EINVAL; }
  if (e->sampler_on && !use_mega2(e, B)) { set_err(e, "gp_pipeline_decode: sampled selection inside the device-resident pipeline needs bf16 storage (the fp32 kernel selects greedily); sample per step with gp_stage_decode"); return GP_ESTATE; }
  if (!use_mega(e, B)) { set_err(e, "gp_pipeline_decode: outside the persistent kernel's envelope (head_dim 64; B <= 8 in bf16 storage, <= 4 in fp32; cooperative launch available)"); return GP_EINVAL; }
  DeviceGuard dev_guard(c.device);
  CK(e, dev_guard.err);
  int r = check_ready(e);
  if (r != GP_OK) return r;
  r = slot_room(e, "gp_pipeline_decode", 0, n_streams * B, n_rounds);
  if (r != GP_OK) return r;
  cudaStream_t s = (cudaStream_t)stream;
  e->launches = 0;
  while ((int)e->events.size() < 2) {
    cudaEvent_t ev;
    CK(e, cudaEventCreate(&ev));
    e->events.push_back(ev);
  }
  PipeArgs pa;
  pa.on = true; pa.n_streams = n_streams; pa.comm_epoch0 = comm_epoch0; pa.tok_out = tokens_out; pa.tok_stride = n_rounds;
  CK(e, cudaEventRecord(e->events[0], s));
  r = decode_mega(e, first_ids, nullptr, B, 0, nullptr, nullptr, nullptr, nullptr, n_rounds, false, s, pa);
  if (r != GP_OK) return r;
  for (int i = 0; i < n_streams * B; ++i) e->hlens[i] += n_rounds;
  CK(e, cudaEventRecord(e->events[1], s));
  // Host watchdog.  The kernel waits on other GPUs; what is watched is PROGRESS (items finished, a device word the
  // kernel bumps), not elapsed time, so a long healthy run is never cut short.  When nothing moved for timeout_ms the
  // abort word is set through a side stream: every wait in the kernel then falls through, the launch runs to its end
  // within milliseconds on whatever data is there, and the call returns an error with the GPU idle and the engine
  // usable (KV slots hold garbage: gp_kv_reset before the next request).  The reference's counterpart is
  // `timeout=30` + raise_for_status (server.py:173-174,179-180).
  {
    struct timespec t_last; clock_gettime(CLOCK_MONOTONIC, &t_last);
    int last_prog = -1, polls = 0;
    bool aborted = false;
    for (;;) {
      cudaError_t q = cudaEventQuery(e->events[1]);
      if (q == cudaSuccess) break;
      if (q != cudaErrorNotReady) { CK(e, q); }
      struct timespec nap = {0, 50 * 1000};
      nanosleep(&nap, nullptr);
      if (timeout_ms <= 0 || (++polls & 31) != 0) continue;     // look at the progress word every ~2 ms
      struct timespec now; clock_gettime(CLOCK_MONOTONIC, &now);
      if (use_mega2(e, B)) {
        CK(e, cudaMemcpyAsync(e->ctrl_host, e->ctrl + 1, sizeof(int), cudaMemcpyDeviceToHost, e->ctrl_stream));
        CK(e, cudaStreamSynchronize(e->ctrl_stream));
        if (e->ctrl_host[0] != last_prog) { last_prog = e->ctrl_host[0]; t_last = now; }
      }
      const double idle = (now.tv_sec - t_last.tv_sec) * 1e3 + (now.tv_nsec - t_last.tv_nsec) * 1e-6;
      if (idle <= timeout_ms) continue;
      if (!use_mega2(e, B)) {
        set_err(e, "gp_pipeline_decode: not finished after %d ms (a peer stage is missing?); the fp32 persistent kernel cannot be "
                   "aborted - terminate the process", timeout_ms);
        return GP_ECUDA;
      }
      if (!aborted) {
        e->ctrl_host[1] = 1;
        CK(e, cudaMemcpyAsync(e->ctrl, e->ctrl_host + 1, sizeof(int), cudaMemcpyHostToDevice, e->ctrl_stream));
        CK(e, cudaStreamSynchronize(e->ctrl_stream));
        aborted = true;
        t_last = now;                                            // give the kernel timeout_ms more to drain
        continue;
      }
      set_err(e, "gp_pipeline_decode: the kernel did not drain after the abort request");
      return GP_ECUDA;
    }
    if (aborted) {
      for (int i = 0; i < n_streams * B; ++i) e->hlens[i] -= n_rounds;
      set_err(e, "gp_pipeline_decode: no progress for %d ms after %d of %d items (a peer stage is missing or stalled); the launch "
                 "was aborted, the GPU is idle, KV slots are undefined until gp_kv_reset", timeout_ms, last_prog < 0 ? 0 : last_prog,
              n_rounds * n_streams);
      return GP_ECUDA;
    }
  }
  if (ms_host) CK(e, cudaEventElapsedTime(ms_host, e->events[0], e->events[1]));
  return GP_OK;
}

// ------------------------------------------------------------------- op-level entry points
#define OPCK(call)                                                                   \
  do {                                                                               \
    cudaError_t _c = (call);                                                         \
    if (_c != cudaSuccess) {                                                         \
      set_err(nullptr, "%s failed: %s", #call, cudaGetErrorString(_c));              \
      return _c == cudaErrorInvalidValue ? GP_EINVAL : GP_ECUDA;                     \
    }                                                                                \
  } while (0)

static int op_init() {
  static thread_local int inited_dev = -1;
  int dev = 0;
  OPCK(cudaGetDevice(&dev));
  if (inited_dev != dev) {
    OPCK(init_gemv_attrs());
    OPCK(init_attn_attrs());
    OPCK(init_gemm_tc());
    inited_dev = dev;
  }
  return GP_OK;
}

int gp_op_embed(const int32_t* ids, const int32_t* pos, const void* wte, const void* wpe, float* out, int32_t M,
                int32_t H, int32_t dtype, void* stream) {
  if (!ids || !pos || !wte || !wpe || !out) { set_err(nullptr, "gp_op_embed: NULL argument"); return GP_EINVAL; }
  OPCK(launch_embed(ids, pos, 0, 1, nullptr, 0, wte, wpe, out, M, H, 0x7fffffff, 0x7fffffff, dtype, (cudaStream_t)stream));
  return GP_OK;
}

int gp_op_layernorm(const float* x, const float* gamma, const float* beta, void* out, int32_t M, int32_t H, float eps,
                    uint32_t flags, void* stream) {
  if (!x || !gamma || !beta || !out) { set_err(nullptr, "gp_op_layernorm: NULL argument"); return GP_EINVAL; }
  OPCK(launch_layernorm(x, H, gamma, beta, out, flags & GP_OUT_BF16, M, H, eps, (cudaStream_t)stream));
  return GP_OK;
}

int gp_op_gemv(const float* x, const void* Wt, const float* bias, const float* residual, float* out, int32_t B,
               int32_t N, int32_t K, const float* ln_gamma, const float* ln_beta, float eps, uint32_t flags,
               int32_t dtype, void* stream) {
  if (!x || !Wt || !out) { set_err(nullptr, "gp_op_gemv: NULL argument"); return GP_EINVAL; }
  if (!gemv_supported(B, K)) { set_err(nullptr, "gp_op_gemv: B=%d K=%d outside the GEMV envelope (B<=%d, B*K*4<=%d)", B, K, kGemvMaxBatch, kGemvMaxSmem); return GP_EINVAL; }
  int r = op_init();
  if (r != GP_OK) return r;
  GemvParams p;
  p.x = x; p.ldx = K; p.Wt = Wt; p.bias = bias; p.residual = residual; p.ldr = N; p.out = out; p.ldo = N;
  p.B = B; p.N = N; p.K = K; p.ln_g = ln_gamma; p.ln_b = ln_beta; p.eps = eps; p.flags = flags;
  OPCK(launch_gemv(p, dtype, (cudaStream_t)stream));
  return GP_OK;
}

int gp_op_gemm(const void* A, const void* Wt, const float* bias, const float* residual, void* out, int32_t M, int32_t N,
               int32_t K, uint32_t flags, int32_t dtype, void* stream) {
  if (!A || !Wt || !out) { set_err(nullptr, "gp_op_gemm: NULL argument"); return GP_EINVAL; }
  int r = op_init();
  if (r != GP_OK) return r;
  if (dtype == GP_BF16 && !(flags & GP_GEMM_SIMT)) {
    if (!gemm_tc_supported(M, N, K)) { set_err(nullptr, "gp_op_gemm: shape M=%d N=%d K=%d not supported by the tcgen05 kernel", M, N, K); return GP_EINVAL; }
    OPCK(launch_gemm_tc(A, K, Wt, bias, residual, N, out, N, M, N, K, flags, (cudaStream_t)stream));
  } else {
    OPCK(launch_gemm_simt(A, K, Wt, bias, residual, N, out, N, M, N, K, flags, dtype, (cudaStream_t)stream));
  }
  return GP_OK;
}

int gp_op_attn_decode(const float* q, const void* kcache, const void* vcache, const int32_t* ctx_len, float* out,
                      int32_t B, int32_t nh, int32_t d, int32_t max_ctx, int32_t dtype, void* stream) {
  if (!q || !kcache || !vcache || !ctx_len || !out) { set_err(nullptr, "gp_op_attn_decode: NULL argument"); return GP_EINVAL; }
  OPCK(launch_attn_decode(q, kcache, vcache, ctx_len, 0, 0, out, false, B, nh, d, max_ctx, dtype, (cudaStream_t)stream));
  return GP_OK;
}

int gp_op_attn_prefill(const void* qkv, void* out, int32_t B, int32_t T, int32_t nh, int32_t d, int32_t dtype,
                       void* stream) {
  if (!qkv || !out) { set_err(nullptr, "gp_op_attn_prefill: NULL argument"); return GP_EINVAL; }
  OPCK(launch_attn_prefill(qkv, out, B, T, nh, d, dtype, (cudaStream_t)stream));
  return GP_OK;
}

int gp_op_sample_topk(const float* logits, int32_t B, int32_t V, float temperature, int32_t top_k, uint64_t seed,
                      int32_t row0, int32_t position, int32_t* ids_out, void* stream) {
  if (!logits || !ids_out || B <= 0 || V <= 0) { record_error("gp_op_sample_topk: bad argument"); return GP_EINVAL; }
  if (!(temperature > 0.f) || top_k <= 0 || top_k > kMaxTopK) {
    record_error("gp_op_sample_topk: temperature must be > 0 and 1 <= top_k <= 64");
    return GP_EINVAL;
  }
  OPCK(launch_sample_topk(logits, V, B, V, temperature, top_k, seed, row0, position, ids_out, nullptr, 0, nullptr, 0,
                          (cudaStream_t)stream));
  return GP_OK;
}

int gp_op_lmhead_argmax(const float* x, const float* gamma, const float* beta, const void* wte, int32_t* ids_out,
                        float* logits_out, void* scratch, int32_t B, int32_t V, int32_t H, float eps, int32_t dtype,
                        void* stream) {
  if (!x || !wte || !ids_out || !scratch) { set_err(nullptr, "gp_op_lmhead_argmax: NULL argument"); return GP_EINVAL; }
  if ((gamma == nullptr) != (beta == nullptr)) {
    set_err(nullptr, "gp_op_lmhead_argmax: gamma and beta must both be given (ln_f is applied) or both be NULL (x is used as is)");
    return GP_EINVAL;
  }
  if (!gemv_supported(B, H)) { set_err(nullptr, "gp_op_lmhead_argmax: B=%d H=%d outside the GEMV envelope", B, H); return GP_EINVAL; }
  int r = op_init();
  if (r != GP_OK) return r;
  LmheadParams p;
  p.x = x; p.ldx = H; p.ln_g = gamma; p.ln_b = beta; p.eps = eps; p.wte = wte; p.B = B; p.V = V; p.H = H;
  p.ids_out = ids_out; p.logits_out = logits_out; p.ldl = V; p.scratch = scratch;
  OPCK(launch_lmhead_argmax(p, dtype, (cudaStream_t)stream));
  return GP_OK;
}

}  // extern "C"
