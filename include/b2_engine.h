/*
 * b2_engine.h -- C ABI of the B200-native model-server engine (libb2engine.so).
 *
 * The reference (redhat-et/aws-k8s-ansible-provisioner) has no FFI or plugin interface for
 * its hot path: the path is the forward pass of the model-server container that
 * llm-d-deploy.yaml:176-184 launches for the model named at llm-d-deploy.yaml:118, reached
 * over HTTP by llm-d-test.yaml:32-43,61-73.  The arithmetic contract is HF transformers'
 * Qwen3 forward (models/qwen3/modeling_qwen3.py, cited per entry point below as "HF:").
 * This header is the in-process boundary the C++ server (b2serve) and the Python host mirror
 * (aws-k8s-ansible-provisioner_b200/engine.py) both bind; INTEGRATION.md shows the ctypes stub.
 *
 * Conventions: plain pointers and sizes only.  Every function returns 0 on success or a
 * negative errno-style code; b2_last_error() returns the message for the calling thread.
 * "host" pointers are caller-owned CPU memory; "dev" pointers are CUDA device memory.
 * One engine per device; an engine is NOT thread-safe (the server serialises onto one GPU
 * thread).  There is no CPU fallback anywhere behind this ABI.
 */
#ifndef B2_ENGINE_H
#define B2_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 2

typedef struct b2_engine b2_engine;

/* Architecture + runtime sizing.  Field meanings follow HF Qwen3Config
 * (HF:models/qwen3/configuration_qwen3.py) and `vllm serve` flags the llm-d chart passes
 * (--max-model-len, --block-size, --max-num-seqs, --gpu-memory-utilization). */
typedef struct b2_config {
  int32_t abi_version;       /* must be B2_ABI_VERSION */
  int32_t num_layers;        /* 28 */
  int32_t hidden;            /* 1024 */
  int32_t num_q_heads;       /* 16 */
  int32_t num_kv_heads;      /* 8 */
  int32_t head_dim;          /* 128; narrower heads (Phi-3: 96) run zero-padded to 128 */
  int32_t intermediate;      /* 3072 */
  int32_t vocab;             /* 151936 */
  int32_t tie_embeddings;    /* 1: lm_head shares embed_tokens (HF:...modeling_qwen3.py:443) */
  float   rms_eps;           /* 1e-6 */
  float   rope_theta;        /* 1e6 */
  int32_t max_model_len;     /* longest prompt+output per sequence */
  int32_t block_size;        /* KV page size in tokens; 16 */
  int32_t num_blocks;        /* KV pool size in pages; 0 = derive from gpu_mem_fraction */
  float   gpu_mem_fraction;  /* used when num_blocks == 0 */
  int32_t max_seqs;          /* max concurrently resident sequences (decode batch ceiling) */
  int32_t max_batch_tokens;  /* max tokens in one prefill forward */
  int32_t use_cuda_graphs;   /* 1: decode steps replay a captured graph per batch bucket */
  const float* rope_inv_freq;/* optional host [rotary_dim/2] fp32 table (HF:...:129-131); NULL =
                                computed as theta^(-2i/d) in double, rounded to fp32 */
  int32_t arch;              /* B2_ARCH_*: which HF model family the tensors and the layer follow */
  int32_t rotary_dim;        /* width of the rotated prefix of each head; 0 = head_dim */
  int32_t max_positions;     /* OPT: rows of the learned position table minus 2; 0 = max_model_len */
  int32_t sliding_window;    /* Phi-3: every query attends to its last `sliding_window` keys only; 0 = all */
} b2_config;

/* Model families.  QWEN3: HF:models/qwen3/modeling_qwen3.py (QK-norm, separate q/k/v and gate/up
 * tensors, GQA).  PHI3: HF:models/phi3/modeling_phi3.py:49-64,208-272 (fused qkv_proj and
 * gate_up_proj tensors, no QK-norm, one query head per kv head or two, rotary over rotary_dim,
 * usually an untied lm_head, sliding-window attention when config.sliding_window is set).
 * OPT: HF:models/opt/modeling_opt.py:45-70,97-181,184-254,273-397 (learned positions with the +2
 * offset, LayerNorm with bias before attention and before the feed-forward, biased q/k/v/out and
 * fc1/fc2 projections, query pre-scaled by head_dim^-0.5, ReLU, tied lm_head; rms_eps carries the
 * LayerNorm eps; the do_layer_norm_before=False (350m) and word_embed_proj_dim != hidden variants are
 * not built). */
enum b2_arch { B2_ARCH_QWEN3 = 0, B2_ARCH_PHI3 = 1, B2_ARCH_OPT = 2 };

const char* b2_last_error(void);
int  b2_abi_version(void);

/* ---- engine lifetime ------------------------------------------------------------------- */
/* Creates the engine on CUDA device `device`.  If safetensors_path is non-NULL, every
 * HF-named tensor is read from that file (bf16) and the weights are finalised; otherwise
 * the caller uploads tensors with b2_engine_load_tensor and calls b2_engine_finalize.
 * Replaces: the model-server container start at llm-d-deploy.yaml:176-193. */
int  b2_engine_create(const b2_config* cfg, const char* safetensors_path, int device, b2_engine** out);
int  b2_engine_load_tensor(b2_engine* e, const char* hf_name, const void* host_bf16, size_t nbytes);
/* Reads every tensor of one safetensors file (BF16, F16 or F32; converted to bf16) into an engine created without
 * a path: call it once per shard of a sharded checkpoint (model-0000i-of-0000N.safetensors), then finalize. */
int  b2_engine_load_safetensors(b2_engine* e, const char* path);
int  b2_engine_finalize(b2_engine* e);
void b2_engine_free(b2_engine* e);
/* Info: KV pages total / free, bytes of weights resident. */
int  b2_engine_info(b2_engine* e, int64_t* num_blocks, int64_t* free_blocks, int64_t* weight_bytes);

/* ---- sequences (paged KV allocator; HF:...modeling_qwen3.py:270-271 cache update) ------- */
/* Reserves pages for up to reserve_tokens tokens (prompt + output) and returns a slot id. */
int  b2_seq_create(b2_engine* e, int32_t reserve_tokens, int32_t* seq_id);
int  b2_seq_free(b2_engine* e, int32_t seq_id);
int  b2_seq_len(b2_engine* e, int32_t seq_id);
/* Drops the tail of a sequence (new_len <= current length); its pages stay reserved.  Used to
 * re-run a measurement over the same context and by stop-string handling in the server. */
int  b2_seq_truncate(b2_engine* e, int32_t seq_id, int32_t new_len);

/* ---- forward passes --------------------------------------------------------------------- */
/* Prefill: appends ids[seq_offsets[i] .. seq_offsets[i+1]) to sequence seq_ids[i], runs the
 * forward over all appended tokens (HF:...modeling_qwen3.py:378-438), and returns the greedy
 * next token of every sequence (HF:...:504-505 lm_head on the last position + argmax, lowest
 * index on ties) in out_tokens (host, may be NULL).  The next token is NOT appended. */
int  b2_prefill(b2_engine* e, const int32_t* ids, const int32_t* seq_offsets, int32_t nseq,
                const int32_t* seq_ids, int32_t* out_tokens);

/* One decode step: appends in_tokens[i] to sequence seq_ids[i] (in_tokens == NULL: append the
 * token the previous forward produced for that sequence, kept on the device), runs one
 * forward, returns the greedy next tokens in out_tokens (host, may be NULL = no D2H copy). */
int  b2_decode_step(b2_engine* e, int32_t nseq, const int32_t* seq_ids, const int32_t* in_tokens,
                    int32_t* out_tokens);

/* `steps` decode steps back to back, token feedback on the device, no host round trip
 * between steps.  elapsed_ms (may be NULL) receives the CUDA-event time of the region on
 * the engine stream.  out_tokens (host, [steps][nseq], may be NULL) is copied at the end. */
int  b2_decode_run(b2_engine* e, int32_t nseq, const int32_t* seq_ids, int32_t steps,
                   int32_t* out_tokens, float* elapsed_ms);

/* Logits of the last forward's last positions, [nseq][vocab] fp32 on the host. */
int  b2_get_logits(b2_engine* e, int32_t nseq, float* out_logits);
/* Residual stream after layer `layer` (0-based; -1 = embeddings, num_layers = final norm
 * output) for the tokens of the last forward, [ntokens][hidden] as bf16 bits on the host.
 * Only filled when debug capture was enabled with b2_set_debug(e, 1). */
int  b2_set_debug(b2_engine* e, int32_t capture_hidden);
int  b2_get_hidden(b2_engine* e, int32_t layer, int32_t ntokens, uint16_t* out_bf16);

/* Measurement knobs that can change while the engine lives (bench.py's ablation leg, the two-lane tests); every
 * call drops the captured decode graphs, so the next step re-captures with the new setting.
 *   B2_TUNE_SKIP_MASK   bit k set = kernel class k (b2_kernel_class order) is left out of the step: timing only,
 *                       the outputs are then meaningless
 *   B2_TUNE_PREFIX_CACHE  1 = full pages of prompt tokens are kept under the hash of the token prefix they end, and
 *                       b2_prefill lets a fresh sequence whose prompt starts with the same tokens take those pages
 *                       instead of recomputing them (vLLM's --enable-prefix-caching; environment B2_PREFIX_CACHE).
 *                       Results are unchanged: the cached K/V are the bits the same prefill would write.  0 = off and
 *                       forget what was cached (the default)
 *   B2_TUNE_TWO_LANE    smallest padded decode batch that runs as two half-batch lanes (0 = never, the default;
 *                       environment B2_MB) */
enum b2_tuning_key { B2_TUNE_SKIP_MASK = 0, B2_TUNE_TWO_LANE = 1, B2_TUNE_PREFIX_CACHE = 2 };
int  b2_set_tuning(b2_engine* e, int32_t key, int64_t value);
/* Prefix cache counters since the engine was created: prompt tokens looked up, prompt tokens served from cached pages,
 * pages currently registered (any pointer may be NULL). */
int  b2_prefix_cache_stats(b2_engine* e, int64_t* queried_tokens, int64_t* hit_tokens, int64_t* cached_pages);

/* Sampling for the API surface (temperature / top-k / top-p over the last logits). */
typedef struct b2_sampling {
  float    temperature;  /* 0 = greedy */
  int32_t  top_k;        /* <=0 = off */
  float    top_p;        /* >=1 = off; taken over what top_k leaves (top-k first, then top-p, as vLLM) */
  uint64_t seed;         /* != 0: the draw depends on (seed, row parameters, logits) only -- a caller that wants a
                          * different draw per generated token advances the seed itself (b2serve adds the number of
                          * tokens generated so far); 0: no seed given, draws also mix in a launch counter */
} b2_sampling;
int  b2_sample(b2_engine* e, int32_t nseq, const b2_sampling* params, int32_t* out_tokens);
/* Per-row variant for a batch whose requests carry different sampling parameters: rows[i] indexes
 * the last forward's batch, params[i] applies to it (temperature 0 = greedy). */
int  b2_sample_rows(b2_engine* e, int32_t nrows, const int32_t* rows, const b2_sampling* params,
                    int32_t* out_tokens);
/* Overrides the device-resident "next token" of each listed sequence (after sampling on host
 * or with b2_sample) so the following b2_decode_step(in_tokens=NULL) consumes it. */
int  b2_set_next_tokens(b2_engine* e, int32_t nseq, const int32_t* seq_ids, const int32_t* tokens);

/* ---- measurement ------------------------------------------------------------------------ */
#define B2_NUM_KERNEL_CLASSES 12
/* Runs `steps` decode steps eagerly (no graph) with a CUDA event pair around every kernel
 * launch on the engine stream; ms_by_class / launches_by_class receive totals per class
 * (index = enum below).  Used by bench.py for the live roofline figure. */
enum b2_kernel_class {
  B2K_EMBED = 0, B2K_NORM = 1, B2K_GEMM_QKV = 2, B2K_ROPE_KV = 3, B2K_ATTN_DECODE = 4,
  B2K_CHAIN = 5, B2K_GEMM_O = 6, B2K_GEMM_GATEUP = 7, B2K_ACT = 8, B2K_GEMM_DOWN = 9,
  B2K_LMHEAD = 10, B2K_MISC = 11
};
int  b2_decode_profile(b2_engine* e, int32_t nseq, const int32_t* seq_ids, int32_t steps,
                       float* ms_by_class, int64_t* launches_by_class);
/* Times the dominant kernel on its own: the decode-attention pass of every layer, launched back to
 * back on the engine stream over the batch's real K/V pages (the batch of the last decode call),
 * `reps` times, between ONE pair of CUDA events.  ms_per_launch = elapsed / (reps * layers);
 * kv_bytes_per_launch = the K and V bytes one launch has to read (sum over sequences of
 * (len + 1) * kv_heads * head_dim * 2 tensors * 2 bytes).  Results of the pass are discarded.
 * bench.py's roofline figure comes from here: an event pair around every single launch (as
 * b2_decode_profile does) adds several microseconds of launch latency to a ~30 us kernel. */
int  b2_attn_profile(b2_engine* e, int32_t nseq, const int32_t* seq_ids, int32_t reps,
                     float* ms_per_launch, int64_t* kv_bytes_per_launch);
/* Number of kernel launches (graph nodes included) issued by this engine so far. */
int64_t b2_launch_count(b2_engine* e);

/* ---- kernel-level entry points (tests call each kernel on device buffers) --------------- */
/* All take a cudaStream_t as void*; NULL = default stream.  Shapes are row-major. */

/* HF:...modeling_qwen3.py:59-64 (+ residual add :327,333).  h [T][H] bf16 in/out; delta may be
 * NULL, bf16 [T][H] (delta_is_f32 = 0) or fp32 split partials [nsplit][T][H] (= 1) that are
 * summed in split order and rounded to bf16 before the add.  x = RMSNorm(h) * w. */
int b2k_add_rmsnorm(void* h_bf16, const void* delta, int delta_is_f32, int nsplit,
                    const void* w_bf16, void* x_bf16, int T, int H, float eps, void* stream);

/* HF:...:248-249,263-264 (QK-norm), :151-181 (RoPE), :270-271 (cache write).
 * qkv [T][(nq+2nkv)*hd] bf16 (or fp32 partials), positions/slots [T] int32,
 * cos_sin [max_pos][hd] bf16 = cos[hd/2] | sin[hd/2].  Writes q_out [T][nq*hd] and the K/V
 * pages (layout: see "KV page layout" below). */
int b2k_qknorm_rope_kv(const void* qkv, int qkv_is_f32, int nsplit, const void* qnorm_w,
                       const void* knorm_w, const void* cos_sin, const int32_t* positions,
                       const int32_t* slots, void* q_out, void* k_pages, void* v_pages, int T,
                       int nq, int nkv, int hd, int block_size, float eps, void* stream);

/* HF:...:81-83: act = silu(gate) * up, gu [T][2*I] (gate | up) bf16 or fp32 partials. */
int b2k_silu_mul(const void* gu, int gu_is_f32, int nsplit, void* act_bf16, int T, int I, void* stream);

/* y[T][N] = x[T][K] . W[N][K]^T, bf16 in, fp32 accumulate (HF nn.Linear, :236-247).
 * impl: 0 = SIMT GEMV path, 1 = tcgen05/TMEM path.  nsplit > 1: out is fp32 partials
 * [nsplit][T][N]; nsplit == 1 && out_is_f32 == 0: bf16 [T][N]. */
int b2k_linear(const void* x_bf16, const void* w_bf16, void* out, int out_is_f32, int nsplit,
               int T, int N, int K, int impl, void* stream);

/* The two fused projections of a decode step (csrc/gemm_pair.cu), T <= 64 tokens, K <= 1024 in steps of 128:
 *   b2k_gateup_swiglu: act[T][I] = bf16(bf16(silu(g)) * u) with g | u = bf16(x . Wgu^T), Wgu [2I][K] = gate rows then up
 *                      rows (HF:...:81-83); the [T][2I] intermediate is never written
 *   b2k_qkv_rope:      b2k_linear with Wqkv [(nq + 2 nkv) * 128][K] followed by b2k_qknorm_rope_kv, in one launch
 * Same rounding points as the separate kernels; the fp32 sum of a dot product is formed from two K halves. */
int b2k_gateup_swiglu(const void* x_bf16, const void* w_gu_bf16, void* act_bf16, int T, int I, int K, void* stream);
int b2k_qkv_rope(const void* x_bf16, const void* w_qkv_bf16, const void* qnorm_w, const void* knorm_w, const void* cos_sin,
                 const int32_t* positions, const int32_t* slots, void* q_out, void* k_pages, void* v_pages, int T, int nq,
                 int nkv, int K, int block_size, float eps, void* stream);

/* The decode step's norm-free layer chain (round 2): the two projections that feed the residual stream update it
 * themselves, and the two projections that read it normalise it as they load it, so that no RMSNorm kernel runs between
 * them (HF:models/qwen3/modeling_qwen3.py:339-351 for the adds, :50-62 for the norm).
 * b2k_linear_resid: h[T][N] (bf16, in place) = bf16(h + bf16(x[T][K] . W[N][K]^T)), T <= 64, N % 128 == 0, K % 512 == 0,
 *   K <= 3072; ssq[N/16][ssq_ld] (fp32) receives per token column the sum of squares of each 16-feature piece of the new h.
 * b2k_gateup_swiglu_norm / b2k_qkv_rope_norm: b2k_gateup_swiglu / b2k_qkv_rope with x = bf16(norm_w * bf16(h * rstd)) formed
 *   on the fly, rstd = rsqrt(sum of the `pieces` pieces of ssq / K + eps). */
/* Prefill-sized y[T][N] (bf16) = x[T][K] . W[N][K]^T by pairs of CTAs that share the token tile through TMA multicast (both
 * operands in this GPU's memory); T >= 512, N % 256 == 0, K % 64 == 0.  Bit-identical to b2k_linear's bf16 output. */
int b2k_linear_multicast(const void* x_bf16, const void* w_bf16, void* out_bf16, int T, int N, int K, void* stream);
/* Prefill-sized residual add in the projection's epilogue: h[T][N] (bf16, in place) = bf16(h + bf16(x[T][K] . W[N][K]^T)), T > 16. */
int b2k_linear_add(const void* x_bf16, const void* w_bf16, void* h_bf16, int T, int N, int K, void* stream);
int b2k_linear_resid(const void* x_bf16, const void* w_bf16, void* h_bf16, float* ssq, int T, int N, int K, int ssq_ld, void* stream);
int b2k_gateup_swiglu_norm(const void* h_bf16, const void* w_gu_bf16, const void* norm_w, const float* ssq, int pieces, int ssq_ld,
                           float eps, void* act_bf16, int T, int I, int K, void* stream);
int b2k_qkv_rope_norm(const void* h_bf16, const void* w_qkv_bf16, const void* norm_w, const float* ssq, int pieces, int ssq_ld,
                      float norm_eps, const void* qnorm_w, const void* knorm_w, const void* cos_sin, const int32_t* positions,
                      const int32_t* slots, void* q_out, void* k_pages, void* v_pages, int T, int nq, int nkv, int K, int block_size,
                      float eps, void* stream);

/* Diagnostics (tools/step_timeline.py): register a device buffer of uint64 in which thread 0 of every CTA of every
 * kernel on the decode path appends one 4-word record {class | blockIdx << 8 | smid << 32 | aux << 40, CTA start,
 * dependency wait returned, CTA end} in %globaltimer ns.  The caller sets buf[0] = 0 (record counter) and
 * buf[1] = capacity in records; records start at buf[4].  NULL switches the trace off.  class 100 = a projection
 * (aux = N/64 | (K/64) << 12 | second lane << 23); for the other kernels aux = 1 on the second lane of a two-lane step. */
int b2k_debug_step_trace(void* dev_u64_buf);
/* Diagnostics: register a device buffer of (148 * 8) uint64 in which every CTA of the tcgen05
 * projection kernel stamps %globaltimer at its phase boundaries (NULL switches the trace off). */
int b2k_debug_gemm_trace(void* dev_u64_buf);
/* Same for the decode chain kernel: (148 * 32) uint64; slots 0..7 = entry into step i, 16..23 = the
 * CTA finished step i's work (before the grid barrier), 8 = CTA start, 9 = dependency satisfied,
 * 10 = CTA done. */
int b2k_debug_chain_trace(void* dev_u64_buf);
int b2k_debug_fused_trace(void* dev_u64_buf);  /* same, for the fused four-step chain: 128 CTAs x 32 slots */

/* lm_head + greedy argmax fused (HF:...:504-505 + generate(do_sample=False)).
 * logits_bf16 may be NULL.  out_tokens [T] int32 (device). scratch: [T] uint64 (device). */
int b2k_lmhead_argmax(const void* x_bf16, const void* w_bf16, void* logits_bf16,
                      int32_t* out_tokens, void* scratch_u64, int T, int V, int K, int impl,
                      void* stream);

/* KV page layout shared by every kernel below: one allocation per layer and per K/V,
 * [page][kv_head][block_size = 16 tokens][head_dim = 128] bf16, i.e. a 4 KB slice per (page, kv head),
 * in which the 16-byte chunk c (8 dims) of token row r sits at chunk position c ^ (r & 7).
 * b2k_qknorm_rope_kv writes that layout; the two attention entry points read it. */

/* HF:integrations/sdpa_attention.py:40-104 at T == 1: paged GQA decode attention.
 * q [B][nq*hd]; pages as above; block_table [B][bt_stride]; seq_lens [B] (including the current
 * token; 0 = empty row, output zeros).  scratch_f32: b2k_attn_decode_scratch_floats(...) floats of
 * device memory for the partial records of the page-balanced attention pass.  out [B][nq*hd] bf16. */
int b2k_attn_decode(const void* q, const void* k_pages, const void* v_pages,
                    const int32_t* block_table, int bt_stride, const int32_t* seq_lens,
                    void* scratch_f32, void* out, int B, int nq, int nkv, int hd, int block_size,
                    void* stream);
size_t b2k_attn_decode_scratch_floats(int B, int nq, int nkv, int hd);

/* No counterpart in the reference (it has no kernels at all).  Same computation and arguments as
 * b2k_attn_decode, minus scratch_f32: for callers that cannot allocate device memory -- a server
 * whose KV pool already owns the free HBM, request-path code that cannot afford the device-wide
 * synchronisation of cudaMalloc/cudaFree, a plugin that is only handed input/output pointers.  The
 * partial records go to an array compiled into the library (B2K_ATTN_NOALLOC_MAX_FLOATS floats =
 * 8 MiB, resident on every GPU that loads the library until the process exits).  The call allocates
 * and frees nothing.  It fails with a message if b2k_attn_decode_scratch_floats(B, nq, nkv, hd)
 * exceeds that size: 5120 + (B * nkv + 2560) * 260 floats must fit, i.e. B * nkv <= 5486 (sequence, kv
 * head) units; the kernel itself handles at most 4096 units per launch, so that is the limit reached
 * first, and it is reported the same way.
 * The first call on each device looks up the array's device address (that device must be current)
 * and caches it; make that first call before starting any stream capture.
 * All calls share the one array and it keeps whatever the last call wrote.  The only words that must
 * start at zero are 16 KB of arrival counters, and every call clears them itself (a memset on the
 * caller's stream ahead of the kernel); everything else is written before it is read.  Two calls
 * whose kernels overlap on the GPU (different streams, or different host threads) race on the array
 * -- serialise them. */
#define B2K_ATTN_NOALLOC_MAX_FLOATS (2 * 1024 * 1024)
int b2k_attn_decode_noalloc(const void* q, const void* k_pages, const void* v_pages,
                            const int32_t* block_table, int bt_stride, const int32_t* seq_lens,
                            void* out, int B, int nq, int nkv, int hd, int block_size, void* stream);

/* Same contract at T > 1 (causal): ragged prefill over the paged cache.  q [T][nq*hd];
 * q_seq[T] / q_pos[T] int32 give the sequence row in block_table and the position of each
 * query token (tokens of one sequence contiguous and ascending).  Host arrays.
 * A test hook, not a serving call: it cudaMallocs a tile list, synchronises `stream` before it returns and frees
 * the list again (the engine's own prefill keeps its tile list in engine-owned buffers and does none of that). */
int b2k_attn_prefill(const void* q, const void* k_pages, const void* v_pages,
                     const int32_t* block_table_dev, int bt_stride, const int32_t* host_seq_first_tok,
                     const int32_t* host_seq_pos0, const int32_t* host_seq_ntok,
                     const int32_t* host_seq_row, int nseq, void* out, int nq, int nkv, int hd,
                     int block_size, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2_ENGINE_H */
