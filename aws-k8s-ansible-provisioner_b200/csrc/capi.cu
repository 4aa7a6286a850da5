// capi.cu -- kernel-level entry points of the C ABI (include/b2_engine.h, "b2k_*"): each one
// launches exactly one of the engine's kernels on caller-provided device buffers so the parity
// tests can check every kernel against the oracle in isolation.
#include <vector>

#include "../../include/b2_engine.h"
#include "kernels.h"

using namespace b2;

#define CAPI_TRY(expr)      \
  do {                      \
    int _r = (expr);        \
    if (_r != 0) return _r; \
  } while (0)

static ActIn mk_in(const void* p, int is_f32, int nsplit) { return is_f32 ? act_f32(p, nsplit) : act_bf16(p); }

extern "C" int b2k_add_rmsnorm(void* h, const void* delta, int delta_is_f32, int nsplit, const void* w, void* x, int T, int H,
                               float eps, void* stream) {
  B2_REQUIRE(h && w && x, "b2k_add_rmsnorm: null pointer");
  return launch_add_rmsnorm((bf16*)h, delta ? mk_in(delta, delta_is_f32, nsplit) : act_bf16(nullptr), (const bf16*)w, (bf16*)x, T, H,
                            eps, (cudaStream_t)stream);
}

extern "C" int b2k_qknorm_rope_kv(const void* qkv, int qkv_is_f32, int nsplit, const void* qnorm_w, const void* knorm_w,
                                  const void* cos_sin, const int32_t* positions, const int32_t* slots, void* q_out, void* k_pages,
                                  void* v_pages, int T, int nq, int nkv, int hd, int block_size, float eps, void* stream) {
  B2_REQUIRE(qkv && qnorm_w && knorm_w && cos_sin && positions && slots && q_out && k_pages && v_pages,
             "b2k_qknorm_rope_kv: null pointer");
  return launch_qknorm_rope_kv(mk_in(qkv, qkv_is_f32, nsplit), (const bf16*)qnorm_w, (const bf16*)knorm_w, (const bf16*)cos_sin,
                               positions, slots, (bf16*)q_out, (bf16*)k_pages, (bf16*)v_pages, T, nq, nkv, hd, block_size, eps,
                               (cudaStream_t)stream);
}

extern "C" int b2k_silu_mul(const void* gu, int gu_is_f32, int nsplit, void* act, int T, int I, void* stream) {
  B2_REQUIRE(gu && act, "b2k_silu_mul: null pointer");
  return launch_silu_mul(mk_in(gu, gu_is_f32, nsplit), (bf16*)act, T, I, (cudaStream_t)stream);
}

extern "C" int b2k_linear(const void* x, const void* w, void* out, int out_is_f32, int nsplit, int T, int N, int K, int impl,
                          void* stream) {
  B2_REQUIRE(x && w && out, "b2k_linear: null pointer");
  if (impl == 0) {
    B2_REQUIRE(nsplit == 1, "b2k_linear: the SIMT path has no split-K");
    return launch_linear_simt((const bf16*)x, (const bf16*)w, out, out_is_f32, T, N, K, (cudaStream_t)stream);
  }
  TcGemmArgs a;
  a.bn = tc_pick_bn(T);
  CAPI_TRY(b2_make_tmap_bf16(&a.tmap_w, w, (uint64_t)N, (uint64_t)K, (uint64_t)K, 128));
  CAPI_TRY(b2_make_tmap_bf16(&a.tmap_x, x, (uint64_t)T, (uint64_t)K, (uint64_t)K, (uint32_t)a.bn));
  a.T = T; a.N = N; a.K = K;
  a.nsplit = nsplit;
  a.out = out;
  a.x_raw = x;
  a.argmax_packed = nullptr;
  if (nsplit > 1) {
    B2_REQUIRE(out_is_f32, "b2k_linear: split-K output must be fp32 partials");
    a.epilogue = TC_EPI_F32_PARTIAL;
  } else {
    a.epilogue = out_is_f32 ? TC_EPI_F32_PARTIAL : TC_EPI_BF16;
  }
  return launch_linear_tc(a, (cudaStream_t)stream);
}

extern "C" int b2k_linear_multicast(const void* x, const void* w, void* out, int T, int N, int K, void* stream) {
  B2_REQUIRE(x && w && out, "b2k_linear_multicast: null pointer");
  CUtensorMap tw, tx;
  CAPI_TRY(b2_make_tmap_bf16(&tw, w, (uint64_t)N, (uint64_t)K, (uint64_t)K, 128));
  CAPI_TRY(b2_make_tmap_bf16(&tx, x, (uint64_t)T, (uint64_t)K, (uint64_t)K, 128));
  return launch_linear_mc(tw, tx, (bf16*)out, T, N, K, 0, (cudaStream_t)stream);
}

extern "C" int b2k_linear_add(const void* x, const void* w, void* h, int T, int N, int K, void* stream) {
  B2_REQUIRE(x && w && h, "b2k_linear_add: null pointer");
  TcGemmArgs a;
  a.bn = tc_pick_bn(T);
  B2_REQUIRE(a.bn >= 32, "b2k_linear_add: built for prefill-sized token counts (T > 16)");
  CAPI_TRY(b2_make_tmap_bf16(&a.tmap_w, w, (uint64_t)N, (uint64_t)K, (uint64_t)K, 128));
  CAPI_TRY(b2_make_tmap_bf16(&a.tmap_x, x, (uint64_t)T, (uint64_t)K, (uint64_t)K, (uint32_t)a.bn));
  a.T = T; a.N = N; a.K = K;
  a.nsplit = 1;
  a.out = h;
  a.resid = h;
  a.argmax_packed = nullptr;
  a.epilogue = TC_EPI_BF16;
  return launch_linear_tc(a, (cudaStream_t)stream);
}

extern "C" int b2k_debug_step_trace(void* dev_u64_buf) {
  unsigned long long* b = (unsigned long long*)dev_u64_buf;
  CAPI_TRY(step_trace_set_elementwise(b));
  CAPI_TRY(step_trace_set_gemm(b));
  CAPI_TRY(step_trace_set_pair(b));
  CAPI_TRY(step_trace_set_resid(b));
  return step_trace_set_attn(b);
}
extern "C" int b2k_gateup_swiglu(const void* x, const void* w_gu, void* act, int T, int I, int K, void* stream) {
  B2_REQUIRE(x && w_gu && act, "b2k_gateup_swiglu: null pointer");
  CUtensorMap tm;
  if (T > 64) {  // prefill-sized: gemm_gateup.cu
    CUtensorMap tx;
    CAPI_TRY(b2_make_tmap_bf16(&tm, w_gu, (uint64_t)2 * I, (uint64_t)K, (uint64_t)K, 128));
    CAPI_TRY(b2_make_tmap_bf16(&tx, x, (uint64_t)T, (uint64_t)K, (uint64_t)K, 128));
    return launch_gateup_swiglu_prefill(tm, tx, (bf16*)act, T, I, K, 0, (cudaStream_t)stream);
  }
  CAPI_TRY(b2_make_tmap_bf16(&tm, w_gu, (uint64_t)2 * I, (uint64_t)K, (uint64_t)K, 64));
  return launch_gateup_swiglu(tm, (const bf16*)x, (bf16*)act, T, I, K, (cudaStream_t)stream);
}

extern "C" int b2k_qkv_rope(const void* x, const void* w_qkv, const void* qnorm_w, const void* knorm_w, const void* cos_sin,
                            const int32_t* positions, const int32_t* slots, void* q_out, void* k_pages, void* v_pages, int T, int nq, int nkv,
                            int K, int block_size, float eps, void* stream) {
  B2_REQUIRE(x && w_qkv && q_out && k_pages && v_pages, "b2k_qkv_rope: null pointer");
  CUtensorMap tm;
  CAPI_TRY(b2_make_tmap_bf16(&tm, w_qkv, (uint64_t)(nq + 2 * nkv) * 128, (uint64_t)K, (uint64_t)K, 128));
  if (T > 64) {  // prefill-sized: gemm_mc.cu, ROPE variant
    CUtensorMap tx;
    CAPI_TRY(b2_make_tmap_bf16(&tx, x, (uint64_t)T, (uint64_t)K, (uint64_t)K, 128));
    return launch_qkv_rope_prefill(tm, tx, (const bf16*)qnorm_w, (const bf16*)knorm_w, (const bf16*)cos_sin, positions, slots, (bf16*)q_out,
                                   (bf16*)k_pages, (bf16*)v_pages, T, nq, nkv, K, block_size, eps, 0, (cudaStream_t)stream);
  }
  return launch_qkv_rope(tm, (const bf16*)x, (const bf16*)qnorm_w, (const bf16*)knorm_w, (const bf16*)cos_sin, positions, slots, (bf16*)q_out,
                         (bf16*)k_pages, (bf16*)v_pages, T, nq, nkv, K, block_size, eps, (cudaStream_t)stream);
}

extern "C" int b2k_linear_resid(const void* x, const void* w, void* h, float* ssq, int T, int N, int K, int ssq_ld, void* stream) {
  B2_REQUIRE(x && w && h && ssq, "b2k_linear_resid: null pointer");
  CUtensorMap tm;
  CAPI_TRY(b2_make_tmap_bf16(&tm, w, (uint64_t)N, (uint64_t)K, (uint64_t)K, 128));
  return launch_linear_resid(tm, (const bf16*)x, (bf16*)h, ssq, T, N, K, ssq_ld, (cudaStream_t)stream);
}

extern "C" int b2k_gateup_swiglu_norm(const void* h, const void* w_gu, const void* norm_w, const float* ssq, int pieces, int ssq_ld, float eps,
                                      void* act, int T, int I, int K, void* stream) {
  B2_REQUIRE(h && w_gu && norm_w && ssq && act, "b2k_gateup_swiglu_norm: null pointer");
  CUtensorMap tm;
  CAPI_TRY(b2_make_tmap_bf16(&tm, w_gu, (uint64_t)2 * I, (uint64_t)K, (uint64_t)K, 64));
  PairNorm pn;
  pn.ssq = ssq; pn.gain = (const bf16*)norm_w; pn.pieces = pieces; pn.ld = ssq_ld; pn.eps = eps;
  return launch_gateup_swiglu(tm, (const bf16*)h, (bf16*)act, T, I, K, (cudaStream_t)stream, 0, pn);
}

extern "C" int b2k_qkv_rope_norm(const void* h, const void* w_qkv, const void* norm_w, const float* ssq, int pieces, int ssq_ld, float norm_eps,
                                 const void* qnorm_w, const void* knorm_w, const void* cos_sin, const int32_t* positions, const int32_t* slots,
                                 void* q_out, void* k_pages, void* v_pages, int T, int nq, int nkv, int K, int block_size, float eps,
                                 void* stream) {
  B2_REQUIRE(h && w_qkv && norm_w && ssq && q_out && k_pages && v_pages, "b2k_qkv_rope_norm: null pointer");
  CUtensorMap tm;
  CAPI_TRY(b2_make_tmap_bf16(&tm, w_qkv, (uint64_t)(nq + 2 * nkv) * 128, (uint64_t)K, (uint64_t)K, 128));
  PairNorm pn;
  pn.ssq = ssq; pn.gain = (const bf16*)norm_w; pn.pieces = pieces; pn.ld = ssq_ld; pn.eps = norm_eps;
  return launch_qkv_rope(tm, (const bf16*)h, (const bf16*)qnorm_w, (const bf16*)knorm_w, (const bf16*)cos_sin, positions, slots, (bf16*)q_out,
                         (bf16*)k_pages, (bf16*)v_pages, T, nq, nkv, K, block_size, eps, (cudaStream_t)stream, 0, pn);
}

extern "C" int b2k_debug_gemm_trace(void* dev_u64_buf) { return tc_set_trace((unsigned long long*)dev_u64_buf); }
extern "C" int b2k_debug_chain_trace(void* dev_u64_buf) { return chain_set_trace((unsigned long long*)dev_u64_buf); }
extern "C" int b2k_debug_fused_trace(void* dev_u64_buf) { return fused_chain_set_trace((unsigned long long*)dev_u64_buf); }

extern "C" int b2k_lmhead_argmax(const void* x, const void* w, void* logits, int32_t* out_tokens, void* scratch_u64, int T, int V,
                                 int K, int impl, void* stream) {
  B2_REQUIRE(x && w && out_tokens, "b2k_lmhead_argmax: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  if (impl == 0) {
    B2_REQUIRE(logits != nullptr, "b2k_lmhead_argmax: the SIMT path needs a logits buffer");
    CAPI_TRY(launch_linear_simt((const bf16*)x, (const bf16*)w, logits, 0, T, V, K, s));
    return launch_argmax_rows((const bf16*)logits, out_tokens, T, V, s);
  }
  B2_REQUIRE(scratch_u64 != nullptr, "b2k_lmhead_argmax: the tcgen05 path needs T uint64 of scratch");
  B2_CUDA_OK(cudaMemsetAsync(scratch_u64, 0, sizeof(unsigned long long) * (size_t)T, s));
  TcGemmArgs a;
  a.bn = 128;  // tokens on MMA-M, 128 vocabulary rows per tile on MMA-N (see TC_EPI_ARGMAX_ROWS)
  CAPI_TRY(b2_make_tmap_bf16(&a.tmap_w, x, (uint64_t)T, (uint64_t)K, (uint64_t)K, 128));
  CAPI_TRY(b2_make_tmap_bf16(&a.tmap_x, w, (uint64_t)V, (uint64_t)K, (uint64_t)K, 128));
  a.T = T; a.N = V; a.K = K;
  a.nsplit = 1;
  a.epilogue = TC_EPI_ARGMAX_ROWS;
  a.out = logits;
  a.argmax_packed = (unsigned long long*)scratch_u64;
  CAPI_TRY(launch_linear_tc(a, s));
  return launch_argmax_finalize((const unsigned long long*)scratch_u64, out_tokens, T, s);
}

extern "C" size_t b2k_attn_decode_scratch_floats(int B, int nq, int nkv, int hd) {
  return attn_decode_scratch_floats(B, nq, nkv, hd);
}

extern "C" int b2k_attn_decode(const void* q, const void* k_pages, const void* v_pages, const int32_t* block_table, int bt_stride,
                               const int32_t* seq_lens, void* scratch_f32, void* out, int B, int nq, int nkv, int hd,
                               int block_size, void* stream) {
  B2_REQUIRE(q && k_pages && v_pages && block_table && seq_lens && out && scratch_f32, "b2k_attn_decode: null pointer");
  CAPI_TRY(attn_decode_zero_counters((float*)scratch_f32, (cudaStream_t)stream));  // callers hand in arbitrary memory
  return launch_attn_decode((const bf16*)q, (const bf16*)k_pages, (const bf16*)v_pages, block_table, bt_stride, seq_lens,
                            (float*)scratch_f32, (bf16*)out, B, nq, nkv, hd, block_size, (cudaStream_t)stream);
}

extern "C" int b2k_attn_prefill(const void* q, const void* k_pages, const void* v_pages, const int32_t* block_table_dev,
                                int bt_stride, const int32_t* host_seq_first_tok, const int32_t* host_seq_pos0,
                                const int32_t* host_seq_ntok, const int32_t* host_seq_row, int nseq, void* out, int nq, int nkv,
                                int hd, int block_size, void* stream) {
  B2_REQUIRE(q && k_pages && v_pages && block_table_dev && host_seq_first_tok && host_seq_pos0 && host_seq_ntok && host_seq_row &&
                 out,
             "b2k_attn_prefill: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<PrefillTile> tiles;
  for (int i = 0; i < nseq; ++i)
    for (int j = 0; j < host_seq_ntok[i]; j += 64)
      tiles.push_back(PrefillTile{host_seq_first_tok[i] + j, host_seq_pos0[i] + j, std::min(64, host_seq_ntok[i] - j),
                                  host_seq_row[i]});
  if (tiles.empty()) return 0;
  PrefillTile* d = nullptr;
  B2_CUDA_OK(cudaMalloc(&d, tiles.size() * sizeof(PrefillTile)));
  cudaError_t ce = cudaMemcpyAsync(d, tiles.data(), tiles.size() * sizeof(PrefillTile), cudaMemcpyHostToDevice, s);
  int rc = 0;
  if (ce != cudaSuccess) {
    b2_set_error("b2k_attn_prefill: tile upload failed: %s", cudaGetErrorString(ce));
    rc = -5;
  } else {
    const char* env = getenv("B2_PREFILL_TC");
    if (env == nullptr || *env != '0') {
      // this hook is not told the size of the page pool: the largest page its block table names bounds what can be read
      int max_row = 0;
      for (int i = 0; i < nseq; ++i) max_row = std::max(max_row, host_seq_row[i]);
      std::vector<int32_t> bth((size_t)(max_row + 1) * bt_stride);
      cudaStreamSynchronize(s);
      cudaMemcpy(bth.data(), block_table_dev, bth.size() * sizeof(int32_t), cudaMemcpyDeviceToHost);
      int64_t pages = 1;
      for (int32_t v : bth) pages = std::max<int64_t>(pages, (int64_t)v + 1);
      rc = launch_attn_prefill_tc((const bf16*)q, (const bf16*)k_pages, (const bf16*)v_pages, pages, block_table_dev, bt_stride, d,
                                  (int)tiles.size(), (bf16*)out, nq, nkv, hd, block_size, s);
    } else {
      rc = launch_attn_prefill((const bf16*)q, (const bf16*)k_pages, (const bf16*)v_pages, block_table_dev, bt_stride, d,
                               (int)tiles.size(), (bf16*)out, nq, nkv, hd, block_size, s);
    }
  }
  cudaStreamSynchronize(s);
  cudaFree(d);
  return rc;
}
