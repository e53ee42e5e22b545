// Internal host-side launch interface of the sm_100a kernels (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gpt2pipe.h"

namespace gp {

constexpr int kGemvMaxBatch = 8;        // rows one GEMV launch carries
constexpr int kMaxTopK = 64;            // candidates the sampling kernel can hold (the reference uses 40)
constexpr int kGemvMaxSmem = 200 * 1024;  // bytes of x staged in shared memory
constexpr int kLmheadMaxGrid = 1184;    // 148 SMs x 8 CTAs
constexpr int kLmheadScratchHeader = 256;
constexpr int kLLShiftMax = 7;          // exchange buffers of the bf16 persistent kernel: one 32-byte sector per 1 KB at most

// out[b,n] = epi(sum_k LN?(x)[b,k] Wt[n,k] + bias[n]) (+ residual[b,n]), or the c_attn
// variant (mode 1) whose epilogue scatters q to `q_out` and k/v into the KV cache.
struct GemvParams {
  const float* x = nullptr; int ldx = 0;
  const void* Wt = nullptr;
  const float* bias = nullptr;
  const float* residual = nullptr; int ldr = 0;
  float* out = nullptr; int ldo = 0;
  int B = 0, N = 0, K = 0;
  const float* ln_g = nullptr; const float* ln_b = nullptr; float eps = 1e-5f;
  uint32_t flags = 0;
  int mode = 0;                       // 0 plain, 1 qkv scatter
  float* q_out = nullptr;             // [B,H]
  void* kcache = nullptr; void* vcache = nullptr;  // [slots, nh, max_ctx, d]
  const int32_t* lens = nullptr; int slot0 = 0;
  int H = 0, nh = 0, d = 0, max_ctx = 0;
};
cudaError_t launch_gemv(const GemvParams& p, int dtype, cudaStream_t s);
bool gemv_supported(int B, int K);
cudaError_t init_gemv_attrs();

struct LmheadParams {
  const float* x = nullptr; int ldx = 0;   // [B,H] pre-ln_f hidden
  const float* ln_g = nullptr; const float* ln_b = nullptr; float eps = 1e-5f;
  const void* wte = nullptr;               // [V,H]
  int B = 0, V = 0, H = 0;
  int32_t* ids_out = nullptr;              // [B]
  float* logits_out = nullptr; int ldl = 0;  // [B,V] optional
  void* scratch = nullptr;                 // gp_lmhead_scratch_bytes
  // optional device-side bookkeeping done by the last CTA
  int32_t* lens = nullptr; int slot0 = 0;  // lens[slot0+b] += 1
  int32_t* seq = nullptr; int seq_stride = 0;  // seq[b*stride + old_len + 1] = id
};
cudaError_t launch_lmhead_argmax(const LmheadParams& p, int dtype, cudaStream_t s);

// ---- attention
cudaError_t launch_attn_decode(const float* q, const void* kc, const void* vc, const int32_t* lens,
                               int len_add, int slot0, void* out, bool out_bf16, int B, int nh, int d,
                               int max_ctx, int dtype, cudaStream_t s);
cudaError_t launch_attn_prefill(const void* qkv, void* out, int B, int T, int nh, int d, int dtype,
                                cudaStream_t s);
cudaError_t init_attn_attrs();
// causal prefill attention on mma.sync tensor cores: qkv bf16 [B,T,3H] -> out bf16 [B,T,H], head_dim 64 (k_attn_mma.cu)
cudaError_t launch_attn_prefill64_mma(const void* qkv, void* out, int B, int T, int nh, cudaStream_t s);

// ---- misc
cudaError_t launch_embed(const int32_t* ids, const int32_t* pos, int pos0, int T, const int32_t* lens,
                         int slot0, const void* wte, const void* wpe, float* out, int M, int H, int V,
                         int n_pos, int dtype, cudaStream_t s);
cudaError_t launch_layernorm(const float* x, int ldx, const float* g, const float* b, void* out, bool out_bf16,
                             int M, int H, float eps, cudaStream_t s);
cudaError_t launch_gemm_simt(const void* A, int lda, const void* Wt, const float* bias, const float* residual,
                             int ldr, void* out, int ldo, int M, int N, int K, uint32_t flags, int dtype,
                             cudaStream_t s);
// qkv [B*T, 3H] -> caches at positions [0,T) of slots slot0.. (prefill)
cudaError_t launch_kv_scatter(const void* qkv, void* kc, void* vc, int B, int T, int slot0, int H, int nh, int d,
                              int max_ctx, int dtype, cudaStream_t s);
// qkv [B, 3H] -> q fp32 [B,H] + caches at position lens[slot] (batched decode on the GEMM path)
cudaError_t launch_qkv_split_decode(const void* qkv, float* q, void* kc, void* vc, const int32_t* lens, int B,
                                    int slot0, int H, int nh, int d, int max_ctx, int dtype, cudaStream_t s);
cudaError_t launch_set_lens(int32_t* lens, int slot0, int B, int value, int add, cudaStream_t s);
// src fp32 [R,C] -> dst (dtype) [C,R] if transpose else [R,C]
cudaError_t launch_convert(const float* src, void* dst, int R, int C, bool transpose, int dtype, cudaStream_t s);
// gather rows: dst[b,:] = src[(b*T + T-1),:]  (last position of each sequence)
cudaError_t launch_gather_last(const float* src, float* dst, int B, int T, int H, cudaStream_t s);

// ---- persistent decode kernel (k_mega.cu)
struct MegaLayer {
  const void *attn_w, *proj_w, *fc_w, *mproj_w;              // [N,K] storage dtype
  const float *attn_b, *proj_b, *fc_b, *mproj_b, *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  void *kc, *vc;                                             // [slots, nh, max_ctx, 64]
  // k_mega2.cu: per-row constants of the LayerNorm folded into c_attn / c_fc: u = W.gamma, c = W.beta + bias
  const float *attn_u, *attn_c, *fc_u, *fc_c;
};
struct MegaParams {
  const MegaLayer* layers = nullptr; int L = 0;              // device array of this stage's layers
  int H = 0, nh = 0, V = 0, n_pos = 0, max_ctx = 0; float eps = 1e-5f;
  int B = 0, slot0 = 0, has_embed = 0, has_head = 0;
  const void *wte = nullptr, *wpe = nullptr; const float *lnf_g = nullptr, *lnf_b = nullptr;
  const int32_t* ids = nullptr;                              // [B] token to process (has_embed)
  const float* hidden_in = nullptr; float* hidden_out = nullptr;
  int32_t* lens = nullptr;
  // flag-in-data exchange buffers: one 64-bit word {epoch, fp32} per element
  unsigned long long *qkv_ll = nullptr, *attn_ll = nullptr, *h2_ll = nullptr, *act_ll = nullptr, *h_ll = nullptr;
  unsigned long long* part_ll = nullptr;                     // [8][2][256] argmax partials (value | index per CTA)
  uint32_t epoch0 = 0;                                       // first epoch of this launch (> every earlier one)
  int32_t* ids_out = nullptr; int32_t* seq = nullptr; int seq_stride = 0; float* logits_out = nullptr;
  int n_steps = 1;                                           // decode steps (rounds) per launch
  int n_streams = 1;                                         // sequences groups per round: item (round, stream) uses slots slot0 + stream*B ..
  // pipeline mode: the stage boundary as peer-memory writes + flag-in-data words
  const unsigned long long* inbox_hid = nullptr;             // local [n_streams][B][H]: hidden from the previous stage
  const unsigned long long* inbox_ids = nullptr;             // local [n_streams][B]: tokens from the last stage (stage 0)
  unsigned long long* peer_hid = nullptr;                    // next stage's inbox_hid (mapped peer memory)
  unsigned long long* peer_ids = nullptr;                    // stage 0's inbox_ids (mapped peer memory; last stage)
  uint32_t comm_epoch0 = 0;                                  // epoch of round 0 on the stage boundary
  int32_t* tok_out = nullptr; int tok_stride = 0;            // stage 0: tokens [n_streams*B][n_steps]
  unsigned long long* step_clock = nullptr;                  // [items+1] globaltimer ns: end of each item, then start
  int n_slots = 0;                                           // filled by the launcher
  int l2_stream_hint = 1;                                    // weight copies carry an L2 evict-first policy
  long long* trace = nullptr; int trace_cap = 0;             // debug: clock64 of CTA 0 per phase boundary
  // second-generation kernel (k_mega2.cu) only:
  int slot_data = 0;                                         // bytes of weight rows per ring slot (filled by the launcher)
  const int32_t* slot_map = nullptr;                         // [n_streams*B] KV slot of every sequence (NULL: slot0 + index)
  const int* abort_flag = nullptr;                           // device word the host sets to make every wait fall through
  unsigned int* progress = nullptr;                          // device word: items finished by this launch
  const float *lnf_u = nullptr, *lnf_c = nullptr;            // ln_f folded into the tied lm_head: [V] each
  int ll_shift = 2;                                          // spread of the exchange buffers: sector s at word s << ll_shift (2 = dense)
  int ablate = 0;                                            // timing experiments (GP_MEGA_ABLATE): parts of the kernel skipped, results wrong
  // token selection: 0 = greedy argmax; 1 = temperature / top-k / multinomial (k_sample.cu), needs logits_out
  int sampler_on = 0; float sample_temperature = 1.f; int sample_top_k = 0; uint32_t sample_seed_lo = 0, sample_seed_hi = 0;
  int begin_half = 0, end_half = 0;                          // the first block starts at its MLP half / the last block stops after its attention half
  int ctx_hint = 0;                                          // longest context this launch can reach (0: unknown, max_ctx is assumed)
  int split_min = 256, split_keys = 128;                     // attention tasks are split over CTAs from this context on, one CTA per split_keys keys
  int poll_ns = 0;                                           // back-off between polls of the exchange words (0: spin)
};
bool mega_supported(int B, int H, int nh, int L, int dtype, int num_sms);
uint32_t mega_epochs_per_launch(int L, int n_steps);
size_t mega_part_words();
cudaError_t init_mega_attrs();
cudaError_t launch_decode_mega(MegaParams p, int dtype, int num_sms, cudaStream_t s);
// k_mega2.cu: bf16 storage, B <= 8, tensor-core rows
bool mega2_supported(int B, int H, int nh, int L, int dtype, int num_sms);
uint32_t mega2_epochs_per_launch(int L, int n_items);
cudaError_t init_mega2_attrs();
int mega2_max_coresident(int B, int H, int L);   // CTAs per SM the runtime would co-schedule (0: launch impossible)
cudaError_t launch_decode_mega2(MegaParams p, int num_sms, cudaStream_t s);
// u[n] = W[n,:].gamma, c[n] = W[n,:].beta (+ bias[n]) for a bf16 matrix W[N,K]
cudaError_t launch_ln_fold(const void* W, const float* g, const float* b, const float* bias, float* u, float* c, int N, int K,
                           cudaStream_t s);

// SM count of the current device (queried once per device; 148 on B200)
int device_sm_count();

// stores `msg` as this thread's last error (what gp_last_error(NULL) returns); engine.cu
void record_error(const char* msg);

// ---- tcgen05 GEMM (gemm_tcgen05.cu)
bool gemm_tc_supported(int M, int N, int K);
cudaError_t init_gemm_tc();
// `partial` (optional, `partial_floats` fp32 of scratch) enables deterministic split-K for
// decode-sized problems; gemm_tc_splits tells how many kernels the call will launch (1 or 2)
int gemm_tc_splits(int M, int N, int K, size_t partial_floats);
cudaError_t launch_gemm_tc(const void* A, int lda, const void* Wt, const float* bias, const float* residual,
                           int ldr, void* out, int ldo, int M, int N, int K, uint32_t flags, cudaStream_t s,
                           float* partial = nullptr, size_t partial_floats = 0);
// per-row argmax of fp32 logits [B,V] (lowest index wins ties) + the decode bookkeeping of the lm_head kernel
// temperature / top-k / multinomial selection of server.py:187-205 (k_sample.cu); `lens`, `seq` as in launch_argmax_rows,
// the draw depends on (seed, row0 + b, position of the new token) only; position = lens + 1, or fixed_pos when lens is NULL
cudaError_t launch_sample_topk(const float* logits, int ldl, int B, int V, float temperature, int top_k, uint64_t seed,
                               int row0, int fixed_pos, int32_t* ids_out, int32_t* lens, int slot0, int32_t* seq,
                               int seq_stride, cudaStream_t s);
cudaError_t launch_argmax_rows(const float* logits, int ldl, int B, int V, int32_t* ids_out, int32_t* lens, int slot0,
                               int32_t* seq, int seq_stride, cudaStream_t s);

}  // namespace gp
