// kernels.h -- internal C++ launch API of the sm_100a kernels (engine.cu and capi.cu call these).
#pragma once
#include "../../include/b2_engine.h"
#include "common.cuh"

namespace b2 {

// step timeline (common.cuh): one setter per .cu file that holds kernels of the decode path
int step_trace_set_elementwise(unsigned long long* dev_buf);
int step_trace_set_gemm(unsigned long long* dev_buf);
int step_trace_set_attn(unsigned long long* dev_buf);

// An activation operand that is either bf16 [T][N] or fp32 split-K partials [nsplit][T][N]
// (summed in split order, then rounded to bf16 -- the rounding point of HF's nn.Linear).
struct ActIn {
  const void* ptr;
  int is_f32;
  int nsplit;
};
inline ActIn act_bf16(const void* p) { return ActIn{p, 0, 1}; }
inline ActIn act_f32(const void* p, int nsplit) { return ActIn{p, 1, nsplit}; }

// ---- elementwise.cu ----
int launch_embed(const int32_t* ids, const bf16* table, bf16* h, int T, int H, int vocab, unsigned long long* zero_buf, int nzero,
                 cudaStream_t s);
int launch_add_rmsnorm(bf16* h, ActIn delta, const bf16* w, bf16* x, int T, int H, float eps, cudaStream_t s, int tag = 0);
// tag: lane marker for the step timeline (common.cuh), no effect on the result
int launch_qknorm_rope_kv(ActIn qkv, const bf16* qw, const bf16* kw, const bf16* cos_sin, const int32_t* positions,
                          const int32_t* slots, bf16* q_out, bf16* k_pages, bf16* v_pages, int T, int nq, int nkv,
                          int hd, int block_size, float eps, cudaStream_t s, int rotary_dim = 128, const bf16* bias = nullptr,
                          float q_mul = 1.0f, int tag = 0);
// qw == kw == nullptr: no QK-norm (Phi-3).  rotary_dim: width of the rotated prefix of each head; heads narrower
// than 128 are stored zero-padded to 128 (see engine.cu, "internal geometry").  bias [(nq + 2 nkv) * 128]: added to the
// fp32 partials before the rounding; q_mul: the query is multiplied by it afterwards (OPT pre-scales q).

// ---- elementwise_opt.cu (OPT variant) ----
int launch_embed_pos(const int32_t* ids, const int32_t* positions, const bf16* table, const bf16* pos_table, bf16* h, int T, int H,
                     int vocab, int pos_rows, unsigned long long* zero_buf, int nzero, cudaStream_t s);
int launch_add_layernorm(bf16* h, ActIn delta, const bf16* delta_bias, const bf16* w, const bf16* b, bf16* x, int T, int H, float eps,
                         cudaStream_t s);
int launch_bias_relu(ActIn fc1, const bf16* bias, bf16* act, int T, int I, cudaStream_t s);
int launch_silu_mul(ActIn gu, bf16* act, int T, int I, cudaStream_t s, int tag = 0);
int launch_gather_rows(const bf16* src, const int32_t* rows, bf16* dst, int n, int H, cudaStream_t s);
int launch_argmax_rows(const bf16* logits, int32_t* out, int T, int V, cudaStream_t s);
int launch_argmax_finalize(const unsigned long long* packed, int32_t* out, int T, cudaStream_t s);
// decode bookkeeping on the device: positions/seq_lens += 1, slots from the block table, and
// the tokens produced by the previous step become this step's input ids.
int launch_decode_advance(int32_t* positions, int32_t* seq_lens, int32_t* slots, const int32_t* block_table,
                          int bt_stride, int block_size, int B, int32_t* frac_q16, long long l2_budget_bytes,
                          long long bytes_per_page_all_heads, cudaStream_t s);
int launch_logits_to_f32(const bf16* in, float* out, size_t n, cudaStream_t s);

// ---- gemm_simt.cu ----
// out = x[T][K] . w[N][K]^T ; bf16 out [T][N] or fp32 [T][N] (single split).
int launch_linear_simt(const bf16* x, const bf16* w, void* out, int out_is_f32, int T, int N, int K, cudaStream_t s);

// Ride-along L2 prefetch of the next attention pass's K/V pages.  The projection kernels are
// latency-bound and leave HBM idle, and B200's L2 (126 MB) holds a good part of one layer's pages,
// so each projection asks the TMA engine to pull a share of them into L2 while it waits for its own
// operands (cp.async.bulk.prefetch.L2; no SM resources, no completion to wait for).
struct KvPrefetch {
  const bf16* k_pages = nullptr;   // target layer's pools (null = no prefetch)
  const bf16* v_pages = nullptr;
  const int32_t* block_table = nullptr;
  const int32_t* seq_lens = nullptr;
  const int32_t* frac_q16 = nullptr;  // device: share of every sequence's pages to pull, in 1/65536
  int bt_stride = 0, B = 0, nkv = 0;
  int part = 0, nparts = 1;           // this launch takes items with (index % nparts == part)
};

// ---- gemm_tc.cu ----
enum TcEpilogue { TC_EPI_BF16 = 0, TC_EPI_F32_PARTIAL = 1, TC_EPI_ARGMAX = 2, TC_EPI_ARGMAX_ROWS = 3 };
struct TcGemmArgs {
  CUtensorMap tmap_w;  // W [N][K], box 64 x 128   (TC_EPI_ARGMAX_ROWS: the operands trade places -- this is X, box 64 x 128)
  CUtensorMap tmap_x;  // X [>=T rows][K], box 64 x bn  (TC_EPI_ARGMAX_ROWS: W [N][K], box 64 x bn)
  int bn;              // token tile: 16, 32, 64, 128 or 256
  int T, N, K;
  int nsplit;          // split-K factor (EPI_F32_PARTIAL when > 1)
  int epilogue;
  void* out;           // bf16 [T][N] | fp32 [nsplit][T][N] | bf16 logits (may be null) for ARGMAX
  unsigned long long* argmax_packed;  // [T] for TC_EPI_ARGMAX (must be zeroed before launch)
  KvPrefetch prefetch;                // optional ride-along K/V prefetch
  int small_smem = 0;                 // F32_PARTIAL, bn <= 64: short stage ring so the CTA fits beside a decode-attention CTA
  int x_row0 = 0;                     // first row of X (tmap_x) this launch reads: a micro-batch of a larger buffer
  const void* resid = nullptr;        // TC_EPI_BF16, bn >= 32: out = bf16(resid[T][N] + bf16(x . W^T)); may be `out` itself (in-place residual add)
  const void* x_raw = nullptr;        // X rows [0, T) as a plain pointer (row stride K), ALREADY offset to the launch's first row:
                                      // lets decode-sized launches fetch X with cp.async instead of the tensor map (gemm_tc.cu, XCP)
};
int launch_linear_tc(const TcGemmArgs& a, cudaStream_t s);
int tc_pick_bn(int T);
int tc_set_trace(unsigned long long* dev_buf);  // diagnostics: per-CTA phase timestamps (null = off)
// choose a split-K factor so that tiles * splits fills the SMs (each split gets >= 1 k-block)
int tc_pick_nsplit(int T, int N, int K, int num_sms);

// ---- gemm_pair.cu: a projection and the element-wise kernel behind it in one launch (clusters of two CTAs split K and
//      exchange accumulator halves through distributed shared memory); at most 64 tokens, K <= 1024 ----
bool pair_shapes_ok(int T, int K);
// gemm_mc.cu: prefill-sized bf16 projection, two CTAs per cluster share the token tile through TMA multicast
bool linear_mc_shapes_ok(int T, int N, int K);
int launch_linear_mc(const CUtensorMap& tm_w, const CUtensorMap& tm_x128, bf16* out, int T, int N, int K, int x_row0, cudaStream_t s);
int launch_qkv_rope_prefill(const CUtensorMap& tm_qkv, const CUtensorMap& tm_x128, const bf16* qw, const bf16* kw, const bf16* cos_sin,
                            const int32_t* positions, const int32_t* slots, bf16* q_out, bf16* k_pages, bf16* v_pages, int T, int nq, int nkv,
                            int K, int block_size, float eps, int x_row0, cudaStream_t s);
// gemm_gateup.cu: the prefill-sized counterpart (two accumulators per tile, SwiGLU in the epilogue)
bool gateup_prefill_shapes_ok(int T, int I, int K);
int launch_gateup_swiglu_prefill(const CUtensorMap& tm_w, const CUtensorMap& tm_x128, bf16* act, int T, int I, int K, int x_row0, cudaStream_t s);
// Normalise-on-load (see gemm_pair.cu): x is the residual stream, ssq [pieces][ld] the per-16-feature sums of squares of
// its rows (gemm_resid.cu writes them), gain the RMSNorm weight.  Default: x is used as given.
struct PairNorm {
  const float* ssq = nullptr;
  const bf16* gain = nullptr;
  int pieces = 0, ld = 0;
  float eps = 0.f;
  L2Prefetch next_w;  // optional: the next projection's weights, prefetched into L2 by this launch's idle lanes
};
int launch_gateup_swiglu(const CUtensorMap& tm_gu64, const bf16* x, bf16* act, int T, int I, int K, cudaStream_t s, int tag = 0,
                         const PairNorm& norm = PairNorm());
// gemm_resid.cu: h += x . W^T for the residual stream (O / down projection), split-K summed inside a cluster of 8;
// also writes the pieces of the next RMSNorm's statistic
bool resid_shapes_ok(int T, int N, int K);
int launch_linear_resid(const CUtensorMap& tm_w, const bf16* x, bf16* h, float* ssq, int T, int N, int K, int ssq_ld, cudaStream_t s, int tag = 0,
                        const L2Prefetch& next_w = L2Prefetch());
int step_trace_set_resid(unsigned long long* buf);
int launch_qkv_rope(const CUtensorMap& tm_qkv, const bf16* x, const bf16* qw, const bf16* kw, const bf16* cos_sin, const int32_t* positions,
                    const int32_t* slots, bf16* q_out, bf16* k_pages, bf16* v_pages, int T, int nq, int nkv, int K, int block_size,
                    float eps, cudaStream_t s, int tag = 0, const PairNorm& norm = PairNorm());
int step_trace_set_pair(unsigned long long* dev_buf);

// ---- chain.cu ----
// The eight dependent steps between two decode-attention passes as one persistent kernel
// (step numbers: 0 O-proj, 1 add+norm, 2 gate-up, 3 SwiGLU, 4 down, 5 add+norm, 6 QKV, 7 QK-norm/RoPE/KV).
struct ChainGemm {
  int N, K, nsplit, kps;  // kps = k-blocks (64 wide) per split
};
struct ChainParams {
  int T, bn, phase_begin, phase_end;  // runs steps [phase_begin, phase_end)
  ChainGemm g[4];                     // O, gate-up, down, QKV(next layer)
  float* partial;                     // split-K partials, reused by every GEMM step
  bf16 *h, *x, *act, *qrot;
  const bf16 *norm_w_a, *norm_w_b;    // gains of step 1 / step 5
  int norm_b_has_delta;               // 0: step 5 only normalises h (first call of a forward, after the embedding)
  const bf16 *qw, *kw, *cos_sin;
  const int32_t *positions, *slots;
  bf16 *k_pages, *v_pages;            // pools of the layer whose QKV step 6 computes
  int H, I, nq, nkv, block_size;
  float eps;
  unsigned long long* gbar;           // grid barrier arrival counter, zero before the first launch; every launch
                                      // that shares it must use the same grid size
};
int chain_set_trace(unsigned long long* dev_buf);  // diagnostics (null = off)
int launch_decode_chain(const CUtensorMap* w_maps4, const CUtensorMap* x_maps4, const ChainParams& p, cudaStream_t s);

// ---- chain_cluster.cu: the chain in four fused steps (0 O + residual, 1 RMSNorm + gate-up + SwiGLU,
//      2 down + residual, 3 RMSNorm + QKV of the next layer + QK-norm/RoPE/KV write) ----
struct FusedChainParams {
  int T, phase_begin, phase_end;      // runs steps [phase_begin, phase_end)
  bf16 *h, *act, *qrot;
  const bf16 *norm_w_a, *norm_w_b;    // gains folded into step 1 (post-attention norm) / step 3 (next layer's input norm)
  const bf16 *qw, *kw, *cos_sin;
  const int32_t *positions, *slots;
  bf16 *k_pages, *v_pages;            // pools of the layer whose QKV step 3 computes
  float *ss_a, *ss_b;                 // [32 slices][64 tokens] sums of squares left by step 0 / step 2
  int H, I, nq, nkv, block_size;
  float eps;
  unsigned long long* gbar;           // grid barrier arrival counter (its own: 128 CTAs)
};
int fused_chain_supported(int H, int I, int nq, int nkv, int hd, int T, int block_size);
int fused_chain_max_clusters();
int fused_chain_set_trace(unsigned long long* dev_buf);
// w_maps4: Wo (box 64x32), Wgu (64x96), Wd (64x32), Wqkv of the next layer (64x128); a_maps2: attention output, SwiGLU
// output (box 64x64)
int launch_decode_fused_chain(const CUtensorMap* w_maps4, const CUtensorMap* a_maps2, const FusedChainParams& p, cudaStream_t s);

// ---- attn_decode.cu ----
// KV page-slice layout ([16 tokens][128 dims] bf16 = 4 KB per (page, kv head)): the 16-byte chunk c
// (8 dims) of token row r is stored at chunk position c ^ (r & 7).  Every kernel that touches the
// pages goes through this one function.
__host__ __device__ inline int kv_swizzled_chunk(int row_in_page, int chunk) { return chunk ^ (row_in_page & 7); }
size_t attn_decode_scratch_floats(int B, int nq, int nkv, int hd);
// scratch counters must be zero at launch (attn_decode_zero_counters once; the kernel re-zeroes them)
int attn_decode_zero_counters(float* scratch, cudaStream_t s);
int launch_attn_decode(const bf16* q, const bf16* k_pages, const bf16* v_pages, const int32_t* block_table,
                       int bt_stride, const int32_t* seq_lens, float* scratch, bf16* out, int B, int nq, int nkv,
                       int hd, int block_size, cudaStream_t s, float softmax_scale = 0.f, int window = 0, int tag = 0);
// softmax_scale 0 = hd^-0.5; a padded head passes the scale of its real width.  window > 0: sliding-window attention,
// each query sees its last `window` keys (itself included).

// ---- attn_prefill.cu ----
struct PrefillTile {  // one 64-query tile of one sequence
  int32_t q_row0;     // first token row of the tile in the batch
  int32_t q_pos0;     // position of that token in its sequence
  int32_t nrows;      // valid rows (<= 64)
  int32_t bt_row;     // row of the sequence in the block table
};
int launch_attn_prefill(const bf16* q, const bf16* k_pages, const bf16* v_pages, const int32_t* block_table,
                        int bt_stride, const PrefillTile* tiles_dev, int ntiles, bf16* out, int nq, int nkv, int hd,
                        int block_size, cudaStream_t s, float softmax_scale = 0.f, int window = 0);

// ---- attn_prefill_tc.cu: the same pass on tcgen05 (S and P.V in TMEM, K/V pages through TMA); num_pages = size of the pool ----
int launch_attn_prefill_tc(const bf16* q, const bf16* k_pages, const bf16* v_pages, int64_t num_pages, const int32_t* block_table,
                           int bt_stride, const PrefillTile* tiles_dev, int ntiles, bf16* out, int nq, int nkv, int hd, int block_size,
                           cudaStream_t s, float softmax_scale = 0.f, int window = 0, int pair_mode = 0);

// ---- sampling.cu ----
struct SampleRow {  // one row to draw from (temperature <= 0: greedy arg-max, lowest index on ties)
  int row;
  float temperature;
  int top_k;
  float top_p;
  unsigned long long seed;
};
int launch_sample_rows(const bf16* logits, const SampleRow* rows_host, SampleRow* stage_dev, int32_t* out, int nrows, int V, uint64_t step,
                       cudaStream_t s);

int device_num_sms();

}  // namespace b2
