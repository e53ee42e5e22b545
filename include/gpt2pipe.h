/*
 * gpt2pipe.h — C ABI of libgpt2pipe.so: the B200 (sm_100a) pipeline-stage engine that
 * replaces the reference's CPU PyTorch shards and its HTTP/JSON stage boundary.
 *
 * The reference (kanchan-rihan/llm-sharding-demo) has no FFI of its own: its shard
 * compute is `ShardA.forward` / `ShardB.forward` (server.py:76-86, :97-103) called from
 * the FastAPI handlers (server.py:132-151) and its decode loop is `generate`
 * (server.py:154-210).  Each entry point below names the reference code it stands in
 * for; `INTEGRATION.md` shows the ctypes binding a maintainer adds to server.py.
 *
 * Conventions
 *  - every function returns GP_OK (0) or a negative GP_E* code; no exception crosses the
 *    ABI; `gp_last_error` returns the text of the last failure on this thread/engine;
 *  - pointers are DEVICE pointers unless the parameter name ends in `_host`;
 *  - `stream` is a `cudaStream_t` passed as `void*` (NULL = the legacy default stream);
 *  - an engine is bound to one CUDA device and is not re-entrant: one call at a time;
 *  - there is no CPU fallback anywhere: without a CUDA device creation fails.
 */
#ifndef GPT2PIPE_H_
#define GPT2PIPE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GP_VERSION 200

#define GP_OK 0
#define GP_EINVAL (-1)   /* bad argument / shape / state              */
#define GP_ECUDA (-2)    /* a CUDA runtime or driver call failed      */
#define GP_ENOMEM (-3)   /* device allocation failed                  */
#define GP_ESTATE (-4)   /* weights missing, cache overflow, ...      */

/* storage type of weights, KV cache and GEMM-path activations (accumulation is fp32) */
#define GP_F32 0
#define GP_BF16 1

/* gp_config.flags */
#define GP_FLAG_NO_TCGEN05 1u /* route every GEMM through the SIMT tile kernel      */
#define GP_FLAG_NO_GRAPH 2u   /* gp_generate launches kernels directly (no CUDA graph) */
#define GP_FLAG_NO_MEGA 4u    /* decode with one kernel per op instead of the persistent kernel */
/* Stage boundaries inside a block (finer load balance for S-stage pipelines; the reference cuts between blocks
 * only, server.py:63-64).  A GPT-2 block is x += attn(ln_1 x); x += mlp(ln_2 x): the residual stream between the two
 * halves is as good a hand-off as the one between blocks.  BEGIN_HALF: block layer_begin contributes only its MLP
 * half (its attention half ran on the previous stage, which carries END_HALF); END_HALF: block layer_end - 1
 * contributes only its attention half.  Such a stage loads only the weights of the halves it runs. */
#define GP_FLAG_BEGIN_HALF 16u
#define GP_FLAG_END_HALF 32u

/* epilogue / behaviour bits of the op-level entry points */
#define GP_EPI_GELU 1u   /* gelu_new (tanh form) after bias, before the residual add */
#define GP_OUT_BF16 2u   /* `out` is bf16 instead of fp32 (GEMM ops only)            */
#define GP_GEMM_SIMT 4u  /* gp_op_gemm: use the SIMT kernel instead of tcgen05       */

typedef struct gp_engine gp_engine;

/* One pipeline stage = blocks [layer_begin, layer_end) of the model, plus the embeddings
 * when has_embed (Shard A's wte/wpe, server.py:68-86) and ln_f + tied lm_head when
 * has_head (Shard B's tail, server.py:90-103).  Generalises split_gpt2_model
 * (server.py:51-105) from 2 to S stages. */
typedef struct gp_config {
  int32_t n_layer;       /* layers of the whole model                               */
  int32_t n_embd;        /* H                                                       */
  int32_t n_head;        /* nh; head_dim = H / nh                                   */
  int32_t vocab_size;    /* V                                                       */
  int32_t n_positions;   /* rows of wpe (1024 for GPT-2)                            */
  float ln_eps;          /* 1e-5                                                    */
  int32_t layer_begin;   /* first block of this stage                               */
  int32_t layer_end;     /* one past the last block of this stage                   */
  int32_t has_embed;     /* stage 0                                                 */
  int32_t has_head;      /* last stage                                              */
  int32_t dtype;         /* GP_F32 or GP_BF16                                       */
  int32_t max_batch;     /* KV-cache slots (sequences resident at once)             */
  int32_t max_ctx;       /* tokens per slot, <= n_positions                         */
  int32_t max_tokens;    /* most tokens (B*T) one prefill call may carry; 0 = max_batch*max_ctx */
  int32_t device;        /* CUDA ordinal                                            */
  uint32_t flags;        /* GP_FLAG_*                                               */
} gp_config;

int gp_version(void);

/* Text of the most recent failure (engine may be NULL for creation failures). */
const char* gp_last_error(const gp_engine* e);

/* Replaces model load + split_gpt2_model + `.to(device)` (server.py:40-42,108-110) for one
 * stage: allocates weights, KV cache and workspaces in HBM on cfg->device. */
int gp_engine_create(const gp_config* cfg, gp_engine** out);
int gp_engine_destroy(gp_engine* e);

/* Copies one fp32 HOST tensor into engine-owned device storage, converting to cfg.dtype and
 * re-laying Conv1D weights ([in,out], [hf] pytorch_utils.py:97-123) as [out,in] rows.
 * `name` uses GPT2Model.state_dict() naming without the `transformer.` prefix, e.g.
 * "wte.weight", "h.7.attn.c_attn.weight", "ln_f.bias".  Tensors outside this stage's range
 * are rejected with GP_EINVAL.  The caller may free `data_host` on return. */
int gp_load_weight(gp_engine* e, const char* name, const float* data_host, const int64_t* shape,
                   int32_t ndim);
/* GP_OK once every tensor this stage needs has been loaded, else GP_ESTATE. */
int gp_weights_ready(gp_engine* e);

/* Empties KV slot `slot` (or every slot when slot < 0). */
int gp_kv_reset(gp_engine* e, int32_t slot, void* stream);
/* Copies the per-slot cache lengths to the host (synchronises `stream`). */
int gp_kv_lengths(gp_engine* e, int32_t* lens_host, int32_t n, void* stream);

/* Full-sequence forward of this stage: ShardA.forward (server.py:76-86) when has_embed,
 * ShardB.forward (server.py:97-103) when has_head, a middle slice otherwise.
 *   ids         int32 [B,T]   (has_embed)           else NULL
 *   hidden_in   fp32  [B,T,H] (!has_embed)          else NULL
 *   hidden_out  fp32  [B,T,H] (!has_head)
 *   logits      fp32  [B,T,V] if logits_all else [B,V] (last position); may be NULL
 *   next_ids    int32 [B] greedy argmax of the last position (has_head); may be NULL
 * use_cache != 0 also stores K/V for positions [0,T) in slots [slot0, slot0+B) and sets
 * their lengths to T (the prefill of a cached decode); use_cache == 0 leaves the cache
 * untouched (the stateless /forward and /forward_b routes, server.py:132-151). */
int gp_stage_prefill(gp_engine* e, const int32_t* ids, const float* hidden_in, int32_t B, int32_t T,
                     int32_t slot0, int32_t use_cache, float* hidden_out, float* logits,
                     int32_t logits_all, int32_t* next_ids, void* stream);

/* One cached decode step of this stage for slots [slot0, slot0+B): the new token sits at
 * position = cache length (the KV-cache semantics of [hf] modeling_gpt2.py:155-199,569-582
 * that replace the reference's full recompute per token, server.py:169-181).
 *   ids fp: int32 [B] (has_embed) | hidden_in fp32 [B,H] (!has_embed)
 *   hidden_out fp32 [B,H] (!has_head) | logits fp32 [B,V] optional, next_ids int32 [B] (has_head)
 * Lengths advance by one.  All positions are read from device memory, so the launch
 * sequence is identical every step (CUDA-graph capturable). */
int gp_stage_decode(gp_engine* e, const int32_t* ids, const float* hidden_in, int32_t B,
                    int32_t slot0, float* hidden_out, float* logits, int32_t* next_ids,
                    void* stream);

/* gp_stage_decode for an arbitrary set of KV slots (continuous batching, SURVEY.md 8(f) N3: requests join and leave
 * between steps, so the sequences of a step do not sit in adjacent slots).  slots_host: B distinct slot numbers
 * (HOST memory); row b of ids / hidden_in / hidden_out / logits / next_ids belongs to slot slots_host[b].
 * bf16 storage, head_dim 64, 1 <= B <= 8, greedy selection.  The reference serves one request at a time
 * (server.py:154-169). */
int gp_stage_decode_slots(gp_engine* e, const int32_t* ids, const float* hidden_in, int32_t B,
                          const int32_t* slots_host, float* hidden_out, float* logits, int32_t* next_ids,
                          void* stream);

/* The coordinator loop of server.py:169-206 with greedy selection, run entirely on the
 * device by a single-stage engine (has_embed && has_head): prefill the prompts, then
 * n_new-1 cached decode steps; only token ids cross PCIe.
 *   prompt_ids_host int32 [B,P] -> out_ids_host int32 [B,P+n_new] (prompt + continuation)
 *   timings_ms_host optional float[2 + n_new]: [0] prefill ms, [1] decode ms (all steps),
 *                   [2+i] device time of decode step i (0 when steps are not timed singly) */
int gp_generate(gp_engine* e, const int32_t* prompt_ids_host, int32_t B, int32_t P, int32_t n_new,
                int32_t* out_ids_host, float* timings_ms_host);

/* Token selection.  Default: greedy.  temperature > 0 and top_k >= 1 switch the engine
 * to the reference coordinator's sampler (server.py:187-205, temperature 0.6 / top_k 40 there):
 * logits / temperature -> top-k (ties: lower token id first) -> softmax over the k values -> one
 * multinomial draw.  The reference draws from torch's unseeded global generator; here the draw is
 * a pure function of (seed, KV slot of the sequence, position of the new token) - Philox4x32-10,
 * key = seed, counter = (position, slot, 0, 0) - so a request is reproducible and independent of
 * batch composition.  Applies to gp_stage_prefill / gp_stage_decode / gp_generate on a stage with
 * the lm_head, and to gp_pipeline_decode on the last stage.  In bf16 storage the persistent decode
 * kernel draws the token itself (CTA b selects for sequence b from the logits the kernel has just
 * written, then publishes the token to the other CTAs / to stage 0); in fp32 storage sampled steps run
 * the per-op kernels and gp_pipeline_decode answers GP_ESTATE while the sampler is on.
 * temperature <= 0 or top_k <= 0 restores greedy.  top_k <= 64. */
int gp_set_sampler(gp_engine* e, float temperature, int32_t top_k, uint64_t seed);

/* ---- stage boundary on the device (replaces the HTTP/JSON hops of server.py:172-181).
 * One process per GPU.  Every stage owns an "inbox" in its HBM; the stage before it writes the
 * hidden state of each decode step straight into that inbox over NVLink (the last stage writes
 * the next token ids into stage 0's inbox), every fp32 / int32 value travelling in one 64-bit
 * word together with the round number, so the receiving kernel polls its own memory and no
 * host, NCCL call or fence is involved per token.
 *   gp_comm_export : creates the inbox and returns its 64-byte cudaIpcMemHandle_t
 *   gp_comm_connect: maps the inbox of the stage this one feeds ((rank+1) % stages)
 *   gp_pipeline_decode: ONE persistent launch per stage runs n_rounds decode steps of n_streams
 *     micro-batches (B sequences each, KV slots [m*B,(m+1)*B)); call it on every stage at the same
 *     time.  comm_epoch0 must be the same on all stages and larger than any used before
 *     (previous comm_epoch0 + n_rounds).  Stage 0: first_ids int32 [n_streams*B] (tokens the
 *     prefill produced) -> tokens_out int32 [n_streams*B, n_rounds].
 *     timeout_ms > 0 arms a watchdog on PROGRESS: when no item finished for timeout_ms (a peer stage is missing or
 *     stalled - the reference's `timeout=30`, server.py:173,179) the kernel is told to stop waiting, drains within
 *     milliseconds, and the call returns GP_ECUDA with the GPU idle; the KV slots are then undefined until
 *     gp_kv_reset.  (fp32 storage runs the first-generation kernel, which can only report the timeout.) */
int gp_comm_export(gp_engine* e, void* handle_host);
int gp_comm_connect(gp_engine* e, const void* peer_handle_host);
/* The same connection for a host that owns several GPUs in ONE process (CUDA IPC handles cannot be opened by the
 * process that exported them): `peer` is the engine this stage feeds; peer access is enabled when the two sit on
 * different devices.  `peer` must outlive `e`'s use of the connection. */
int gp_comm_connect_local(gp_engine* e, gp_engine* peer);
int gp_pipeline_decode(gp_engine* e, const int32_t* first_ids, int32_t n_rounds, int32_t n_streams, int32_t B,
                       uint32_t comm_epoch0, int32_t* tokens_out, int32_t timeout_ms, float* ms_host,
                       void* stream);

/* Number of kernel launches the last gp_stage_prefill / gp_stage_decode / gp_generate
 * issued (graph replays count the kernels inside the graph). */
int64_t gp_last_launch_count(const gp_engine* e);

/* ---------------------------------------------------------------------------------------
 * Op-level entry points: each hot-path kernel on caller-provided device buffers, for
 * parity tests against the oracle and for ncu.  `dtype` is the storage type of the weight
 * / cache operands.  K1..K8 refer to SURVEY.md §2.3.
 * ------------------------------------------------------------------------------------- */

/* K1: out[m,:] = wte[ids[m]] + wpe[pos[m]]  (server.py:79-82).  out fp32 [M,H]. */
int gp_op_embed(const int32_t* ids, const int32_t* pos, const void* wte, const void* wpe, float* out,
                int32_t M, int32_t H, int32_t dtype, void* stream);

/* K2: nn.LayerNorm over H ([hf] modeling_gpt2.py:252,254,505). x fp32 [M,H]; out fp32 or
 * bf16 (flags & GP_OUT_BF16). */
int gp_op_layernorm(const float* x, const float* gamma, const float* beta, void* out, int32_t M,
                    int32_t H, float eps, uint32_t flags, void* stream);

/* K3/K5/K6/K7 at decode: out[b,n] = epi(sum_k LN?(x)[b,k] * Wt[n,k] + bias[n]) (+ residual[b,n])
 * Conv1D ([hf] pytorch_utils.py:119-123) with W pre-transposed to Wt[N,K]; B <= 8.
 * ln_gamma/ln_beta non-NULL fuses the preceding LayerNorm into the prologue. */
int gp_op_gemv(const float* x, const void* Wt, const float* bias, const float* residual, float* out,
               int32_t B, int32_t N, int32_t K, const float* ln_gamma, const float* ln_beta,
               float eps, uint32_t flags, int32_t dtype, void* stream);

/* K3/K5/K6/K7 for prefill and batched decode: out[M,N] = epi(A[M,K] * Wt[N,K]^T + bias) (+res).
 * A and Wt are `dtype`; out is fp32 or bf16 (GP_OUT_BF16); residual fp32 [M,N].
 * tcgen05 + TMA when dtype == GP_BF16 and !(flags & GP_GEMM_SIMT), else the SIMT kernel. */
int gp_op_gemm(const void* A, const void* Wt, const float* bias, const float* residual, void* out,
               int32_t M, int32_t N, int32_t K, uint32_t flags, int32_t dtype, void* stream);

/* K4 decode: one query per sequence over its cached keys (length lens[b]+... see below).
 * q fp32 [B,H]; kcache/vcache `dtype` [B, nh, max_ctx, d]; ctx_len int32 [B] = number of
 * valid keys INCLUDING the current token; out fp32 [B,H]. */
int gp_op_attn_decode(const float* q, const void* kcache, const void* vcache, const int32_t* ctx_len,
                      float* out, int32_t B, int32_t nh, int32_t d, int32_t max_ctx, int32_t dtype,
                      void* stream);

/* K4 prefill: causal softmax(q k^T / sqrt(d)) v over qkv `dtype` [B,T,3H] (c_attn output,
 * [hf] modeling_gpt2.py:185-191); out `dtype` [B,T,H]. */
int gp_op_attn_prefill(const void* qkv, void* out, int32_t B, int32_t T, int32_t nh, int32_t d,
                       int32_t dtype, void* stream);

/* A10 (server.py:187-205), sampled: one token per row of logits fp32 [B,V] (row stride V) ->
 * ids_out int32 [B].  Row b draws with counter (position, row0 + b); see gp_set_sampler. */
int gp_op_sample_topk(const float* logits, int32_t B, int32_t V, float temperature, int32_t top_k,
                      uint64_t seed, int32_t row0, int32_t position, int32_t* ids_out, void* stream);

/* K8+K9: ln_f, tied lm_head and greedy argmax fused (server.py:101-102,187-205 with greedy
 * selection); lowest index wins ties.  x fp32 [B,H]; wte `dtype` [V,H]; ids_out int32 [B];
 * logits_out fp32 [B,V] or NULL.  `scratch` must hold gp_lmhead_scratch_bytes(B,V) bytes.
 * gamma and beta are given together (ln_f is applied to x first) or are both NULL (x is already normalised);
 * one without the other is GP_EINVAL. */
int64_t gp_lmhead_scratch_bytes(int32_t B, int32_t V);
int gp_op_lmhead_argmax(const float* x, const float* gamma, const float* beta, const void* wte,
                        int32_t* ids_out, float* logits_out, void* scratch, int32_t B, int32_t V,
                        int32_t H, float eps, int32_t dtype, void* stream);

/* Same computation as gp_op_lmhead_argmax for callers that cannot provide (or allocate)
 * scratch memory.  No counterpart in the reference: it exists for hosts that run this kernel
 * where device allocation is impossible or unwelcome - all HBM already belongs to another
 * allocator (a serving runtime's or framework's pool), or the decode loop cannot tolerate the
 * device-wide synchronisation of cudaMalloc/cudaFree, or the caller owns no allocator at all.
 * The partial results and arrival ticket live in a fixed 76,032-byte array inside the library
 * (resident on every GPU that loads it, until process exit); the call performs no cudaMalloc,
 * cudaFree or other allocating call.  Limits: 1 <= B <= 8 and padded(B)*H*4 <= 204,800 bytes of
 * shared memory; anything larger returns GP_EINVAL with the limits in gp_last_error(NULL).
 * The first call on each device looks up the array's device address
 * (cudaGetSymbolAddress) and sets kernel attributes: make that first call before starting any
 * stream capture on that device (concurrent first calls from several host threads are serialised
 * inside).  The looked-up address is cached per device ordinal for the life of the process: it does
 * not survive cudaDeviceReset, nor a second context created on the same ordinal.  All calls share the one array: it still holds the previous
 * call's partials (the ticket is cleared on `stream` before every launch), and two calls whose
 * kernels overlap on the GPU - different streams, or different host threads - race on it;
 * serialise them.  Arguments as gp_op_lmhead_argmax, minus `scratch`. */
int gp_op_lmhead_argmax_static(const float* x, const float* gamma, const float* beta, const void* wte,
                               int32_t* ids_out, float* logits_out, int32_t B, int32_t V, int32_t H,
                               float eps, int32_t dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GPT2PIPE_H_ */
