// engine.cu -- the model engine behind the C ABI (include/b2_engine.h): weights resident in
// HBM, paged KV pool + block allocator, the Qwen3 layer loop (HF:models/qwen3/
// modeling_qwen3.py:378-438, :305-334) as a stream of hand-written kernels, CUDA-graph replay
// of the decode step, and the measurement hooks bench.py uses.
#include <fcntl.h>
#include <math.h>
#include <stdarg.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <list>
#include <map>
#include <unordered_map>
#include <string>
#include <vector>

#include "../../include/b2_engine.h"
#include "json.hpp"
#include "kernels.h"

// ---------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";
void b2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
extern "C" const char* b2_last_error(void) { return g_err; }
int b2_pdl_early_mask() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2_PDL_EARLY");
    v = e ? atoi(e) : 7;
  }
  return v;
}
bool b2_pdl_trigger_early() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2_PDL_TRIGGER");
    v = (e && *e == '0') ? 0 : 1;
  }
  return v != 0;
}
// The next launch of this thread goes out WITHOUT the programmatic attribute (the first kernel after a cross-stream
// fork or join: its predecessors are ordinary full dependencies).
static thread_local int g_pdl_skip_next = 0;
void b2_pdl_skip_next() { g_pdl_skip_next = 1; }
bool b2_pdl_enabled() {
  if (g_pdl_skip_next) {
    g_pdl_skip_next = 0;
    return false;
  }
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2_PDL");
    v = (e && *e == '0') ? 0 : 1;
  }
  return v != 0;
}
extern "C" int b2_abi_version(void) { return B2_ABI_VERSION; }

using namespace b2;

namespace {

struct LayerW {
  bf16 *ln1 = nullptr, *ln2 = nullptr, *qnorm = nullptr, *knorm = nullptr;
  bf16 *wqkv = nullptr, *wo = nullptr, *wgu = nullptr, *wd = nullptr;
  // OPT only (biased projections, LayerNorm biases); wgu then holds fc1 [I][H] and wd holds fc2
  bf16 *bqkv = nullptr, *bo = nullptr, *ln1_b = nullptr, *ln2_b = nullptr, *b1 = nullptr, *b2 = nullptr;
  CUtensorMap tm_qkv, tm_o, tm_gu, tm_d;
  CUtensorMap tm_o32, tm_gu96, tm_d32;  // narrow-box maps of the fused chain (chain_cluster.cu)
  CUtensorMap tm_gu64;                  // 64-row box: gate rows and up rows of one SwiGLU tile (gemm_pair.cu)
};

struct Seq {
  bool active = false;
  int len = 0;        // tokens whose K/V are in the cache
  int reserved = 0;   // token capacity of the pages held
  int32_t next_token = 0;   // greedy token the last forward produced for this sequence (host copy)
  bool next_valid = false;  // false when only the device holds it (no D2H was requested)
  std::vector<int32_t> blocks;
  // prefix cache: chain[j] = hash of the first 16 (j + 1) prompt tokens (one entry per full page whose token ids the
  // host has seen), pending = ids since the last page boundary, chain_open = every token so far came from a prefill
  // call (a decode step's token lives on the device only, so the chain ends there)
  std::vector<uint64_t> chain;
  std::vector<int32_t> pending;
  bool chain_open = false;
};

enum XBuf { XB_X = 0, XB_ATTN = 1, XB_ACT = 2, XB_LAST = 3, XB_COUNT = 4 };
static const int kBnList[5] = {16, 32, 64, 128, 256};
static int bn_index(int bn) { return bn == 16 ? 0 : bn == 32 ? 1 : bn == 64 ? 2 : bn == 128 ? 3 : 4; }

// One stream of dependent launches over a contiguous block of batch rows.  A forward normally has a single lane
// (all rows, the engine's stream); a decode step over a large enough batch runs as TWO lanes -- two half-batches on
// two streams inside one step / one captured graph (see forward()).
struct Lane {
  cudaStream_t stream = nullptr;
  int row0 = 0;                 // first batch row of the lane
  float* partial = nullptr;     // split-K partials of this lane's projections
  size_t partial_floats = 0;
  float* attn_scratch = nullptr;
  bool beside_attention = false;  // projections use the short stage ring so their CTAs fit on an SM next to a
                                  // decode-attention CTA of the OTHER lane
};

struct GraphKey {
  int B, attn_split;
  bool operator<(const GraphKey& o) const { return B != o.B ? B < o.B : attn_split < o.attn_split; }
};

}  // namespace

struct b2_engine {
  b2_config cfg;
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  bool finalized = false;
  int q_size = 0, kv_size = 0, qkv_size = 0;
  // Internal geometry.  The attention kernels are built for 128-wide heads in GQA groups of two; a
  // model with narrower heads (Phi-3: 96) or one query head per kv head (MHA) is laid out to fit:
  // every head is zero-padded to 128 dims and, for MHA, every real query head is followed by an
  // all-zero partner head.  The padding lives in the WEIGHTS (zero rows of Wqkv, zero columns of Wo,
  // scattered once at load time), so activations, KV pages and kernels need no special cases and the
  // arithmetic on the real elements is untouched; the cost is wasted bytes, which only matters for the
  // parity configurations that use it.
  int nq = 0, nkv = 0, hd = 128;  // internal (padded) head counts / width handed to the kernels
  int qhead_step = 1;             // real query head h lives at internal head h * qhead_step
  bool native_geom = true;        // cfg geometry == internal geometry (Qwen3): weights copy straight in
  float attn_scale = 0.f;         // real head_dim^-0.5
  int rotary_dim = 128;

  // weights
  bf16* embed = nullptr;
  bf16* lm_head = nullptr;  // == embed when tied
  bf16* final_norm = nullptr;
  bf16* final_norm_b = nullptr;  // OPT
  bf16* pos_embed = nullptr;     // OPT: learned positions [pos_rows][H], row = position + 2
  int pos_rows = 0;
  int gu_rows = 0;               // rows of wgu: 2*I (gate; up) or I (OPT fc1)
  std::vector<LayerW> layers;
  CUtensorMap tm_lm_head;
  std::map<std::string, int> loaded;
  int64_t weight_bytes = 0;
  bf16* cos_sin = nullptr;  // [max_model_len][hd]: cos[hd/2] | sin[hd/2]

  // KV pool
  bf16* kv_pool = nullptr;
  int64_t num_blocks = 0;
  size_t layer_kv_elems = 0;  // elements of one layer's K (or V) pool
  std::vector<int32_t> free_blocks;
  // Prefix cache (B2_TUNE_PREFIX_CACHE; vLLM's --enable-prefix-caching): full pages of PROMPT tokens are registered under
  // the hash of the whole token prefix they end; a new sequence whose prompt starts with the same tokens takes those
  // pages instead of recomputing them (reference-counted; a page nobody holds any more stays registered on an LRU list
  // until the allocator needs it).  A hit is verified against the stored token ids and the parent page, not the hash alone.
  bool prefix_cache = false;
  std::vector<int32_t> blk_ref;       // sequences holding the page
  std::vector<uint64_t> blk_hash;     // 0 = not registered
  std::vector<int32_t> blk_parent;    // registered: the page holding the 16 tokens before it (-1 = first page)
  std::vector<int32_t> blk_tokens;    // registered: its 16 token ids
  std::unordered_map<uint64_t, int32_t> pc_map;
  std::list<int32_t> pc_lru;          // registered pages with no holder, oldest first
  std::vector<std::list<int32_t>::iterator> pc_lru_pos;
  std::vector<char> pc_in_lru;
  int64_t pc_queries = 0, pc_hits = 0;  // prompt tokens looked up / served from the cache
  std::vector<Seq> seqs;
  int bt_stride = 0;

  // activations (capacity rows = cap_tokens)
  int cap_tokens = 0;
  bf16 *h = nullptr, *x = nullptr, *qkv = nullptr, *qrot = nullptr, *attn = nullptr, *gu = nullptr, *act = nullptr,
       *proj = nullptr, *xlast = nullptr;
  float* partial = nullptr;
  size_t partial_floats = 0;
  float* attn_scratch = nullptr;
  size_t attn_scratch_floats = 0;
  CUtensorMap tm_x[XB_COUNT][5];

  // batch state on the device
  int32_t *d_tokens = nullptr, *d_positions = nullptr, *d_seq_lens = nullptr, *d_slots = nullptr, *d_block_table = nullptr,
          *d_last_rows = nullptr, *d_out_tokens = nullptr, *d_token_log = nullptr, *d_step = nullptr;
  unsigned long long* d_argmax = nullptr;
  PrefillTile* d_tiles = nullptr;
  int tiles_cap = 0;
  int log_cap_steps = 0;
  // pinned host staging
  int32_t* h_stage = nullptr;
  size_t h_stage_ints = 0;
  PrefillTile* h_tiles = nullptr;
  int32_t* h_out = nullptr;
  int32_t* h_tok = nullptr;
  cudaEvent_t ev_tok = nullptr;

  std::vector<int32_t> cur_batch;  // sequence ids resident in d_* (decode fast path)
  int cur_B = 0;                   // padded batch (graph bucket)
  int last_n = 0;                  // sequences in the last forward
  int last_tokens = 0;             // tokens in the last forward

  // debug capture
  int debug_flags = 0;
  bf16* dbg_hidden = nullptr;
  int dbg_cap_tokens = 0;
  bf16* logits = nullptr;  // [max_seqs][vocab], allocated on demand

  // graphs
  std::map<GraphKey, cudaGraphExec_t> graphs;
  std::map<GraphKey, int64_t> graph_nodes;
  int64_t launches = 0;

  // profiling
  bool profiling = false;
  std::vector<cudaEvent_t> ev_pool;
  std::vector<std::pair<int, int>> ev_used;  // (class, index of start event)
  int ev_next = 0;

  // tunables (env)
  int simt_max_tokens = 0;  // T <= this -> SIMT GEMV path
  int attn_split_override = 0;
  int force_gemm_impl = -1;
  unsigned skip_mask = 0;
  long long kv_l2_budget = 0;      // B2_KV_L2_MB: bytes of next-layer K/V the projections prefetch into L2 (0 = off)
  int32_t* d_pf_frac = nullptr;
  int32_t* d_sample_out = nullptr;
  SampleRow* d_sample_rows = nullptr;     // per-row sampling parameters of one b2_sample(_rows) call
  std::vector<SampleRow> h_sample_rows;
  unsigned long long* d_gbar = nullptr;  // grid-barrier arrival counter of the decode chain kernel
  float* d_ss = nullptr;                  // fused chain: 2 x [32][64] sums of squares
  unsigned long long* d_gbar2 = nullptr;  // fused chain: its grid-barrier counter
  bool fused_ok = false;                  // shapes + cluster occupancy allow the fused chain
  int o_small_smem = 0;                   // B2_O_SMALL=1: O projection with a short stage ring so that it is resident beside
                                          // attention's CTAs (measured slower: 1.869 vs 1.832 ms/step at C2 -- its weight
                                          // prefetch competes with the tail of the K/V stream)
  int use_chain = 0;  // B2_CHAIN: 0 separate kernels (default, fastest measured), 1 chain.cu, 2 chain_cluster.cu
  int prefill_tc = 1; // B2_PREFILL_TC: prefill attention on tcgen05 (attn_prefill_tc.cu); 0 = the mma.sync kernel (attn_prefill.cu)
  int l2_weight_prefetch = 0;  // B2_WPF: resid path, see forward_layers (measured slower: off)
  int prefill_resid_fused = 0;  // B2_RESID_FUSED: prefill O / down projections add into the residual stream in their epilogue (measured slower: off)
  int qkv_rope_prefill_fused = 1;  // B2_QKV_ROPE_FUSED: prefill QKV projection finishes norm / RoPE / page write itself
  int prefill_pairs = 1;  // B2_PREFILL_PAIRS: see prefill_tile_pairs()
  int prefill_mc = 1;  // B2_MC: prefill projections as multicast CTA pairs (gemm_mc.cu)
  int gu_prefill_fused = 1;  // B2_GU_FUSED: prefill-sized gate/up projection with SwiGLU in its epilogue (gemm_gateup.cu)
  int use_pair = 1;   // B2_PAIR: bit 0 gate/up + SwiGLU, bit 1 QKV + QK-norm/RoPE/KV write as fused cluster-pair launches (gemm_pair.cu)

  // two-lane decode (B2_MB, see forward()): second stream, its own partials and attention scratch, fork/join events
  int mb_min_batch = 0;  // padded decode batches of at least this many rows (and at most 128) run as two lanes; 0 = never
  cudaStream_t stream2 = nullptr;
  float* partial2 = nullptr;
  size_t partial2_floats = 0;
  float* attn_scratch2 = nullptr;
  float* ssq = nullptr;  // [hidden / 16][64]: pieces of the next RMSNorm's sum of squares (gemm_resid.cu -> gemm_pair.cu)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};

namespace {

template <typename T>
int dalloc(T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T));
  if (e != cudaSuccess) {
    b2_set_error("cudaMalloc(%zu bytes) failed: %s", n * sizeof(T), cudaGetErrorString(e));
    return -12;
  }
  *p = reinterpret_cast<T*>(q);
  return 0;
}
#define B2_TRY(expr)      \
  do {                    \
    int _r = (expr);      \
    if (_r != 0) return _r; \
  } while (0)

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

// kernel-launch bracket: counts launches and, in profile mode, times each one
struct KL {
  b2_engine* e;
  int cls;
  int ev = -1;
  bool skip() const { return (e->skip_mask >> cls) & 1u; }
  KL(b2_engine* e_, int cls_) : e(e_), cls(cls_) {
    if (skip()) return;
    e->launches++;
    if (e->profiling && e->ev_next + 2 <= (int)e->ev_pool.size()) {  // profiled steps are single-lane: the engine's stream
      ev = e->ev_next;
      e->ev_next += 2;
      cudaEventRecord(e->ev_pool[ev], e->stream);
    }
  }
  ~KL() {
    if (ev >= 0) {
      cudaEventRecord(e->ev_pool[ev + 1], e->stream);
      e->ev_used.emplace_back(cls, ev);
    }
  }
};

// Models with one query head per kv head run through the GQA-2 kernels with an all-zero partner head after every real one;
// the tcgen05 prefill attention then pairs two 64-token tiles per CTA instead of (real, zero) heads.
static bool prefill_tile_pairs(const b2_engine* e) { return e->prefill_tc && e->qhead_step == 2 && e->cfg.arch != B2_ARCH_OPT && e->prefill_pairs; }

int bucket_batch(int B) {
  static const int bk[] = {1, 2, 4, 8, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512};
  for (int v : bk)
    if (B <= v) return v;
  return ((B + 127) / 128) * 128;
}

int pick_attn_split(b2_engine* e, int B, int max_len) {
  // The decode-attention kernel balances pages over its persistent warps by itself; there is no
  // host-side split factor any more.  The slot of the graph key now tells the two step variants apart.
  (void)B; (void)max_len;
  return (e->debug_flags & 2) ? 2 : 1;
}

// ---- one Linear layer: picks SIMT GEMV or the tcgen05 kernel, bf16 or split-K partial output
// xptr / out_bf16 are already offset to the lane's first row; the X tensor map covers the whole buffer, so the
// tcgen05 kernel gets the lane's row offset as a coordinate.
ActIn linear(b2_engine* e, const Lane& ln, int cls, XBuf xb, const bf16* xptr, const bf16* w, const CUtensorMap& tm_w, bf16* out_bf16,
             int T, int N, int K, int* rc, const KvPrefetch& pf = KvPrefetch(), bool want_f32 = false, bool after_attention = false,
             bf16* resid = nullptr, bool* resid_done = nullptr) {
  // want_f32: hand the consumer fp32 sums even when K is not split (a bias is added before the rounding)
  // resid: the residual stream rows this projection's output is added to.  When the launch is a single-split tcgen05 GEMM
  // with bf16 output (prefill shapes) the add happens in its epilogue, in place, and *resid_done tells the caller that the
  // norm kernel behind it has no delta to add; otherwise nothing changes.
  *rc = 0;
  if (resid_done) *resid_done = false;
  int impl = e->force_gemm_impl >= 0 ? e->force_gemm_impl : (T <= e->simt_max_tokens ? 0 : 1);
  if (ln.beside_attention && impl == 1 && tc_pick_bn(T) <= 64) want_f32 = true;  // the short ring is built for the partial epilogue
  if (want_f32 && (size_t)T * N > ln.partial_floats) {
    b2_set_error("linear: %d x %d fp32 outputs exceed the partial buffer", T, N);
    *rc = -12;
    return act_bf16(nullptr);
  }
  if (impl == 0) {
    KL k(e, cls);
    if (want_f32) {
      if (!k.skip()) *rc = launch_linear_simt(xptr, w, ln.partial, 1, T, N, K, ln.stream);
      return act_f32(ln.partial, 1);
    }
    if (!k.skip()) *rc = launch_linear_simt(xptr, w, out_bf16, 0, T, N, K, ln.stream);
    return act_bf16(out_bf16);
  }
  TcGemmArgs a;
  a.tmap_w = tm_w;
  a.bn = tc_pick_bn(T);
  a.tmap_x = e->tm_x[xb][bn_index(a.bn)];
  a.T = T; a.N = N; a.K = K;
  a.argmax_packed = nullptr;
  a.prefetch = pf;
  a.x_row0 = ln.row0;
  a.x_raw = xptr;
  a.small_smem = ((after_attention && e->o_small_smem) || ln.beside_attention) ? 1 : 0;
  int ns = tc_pick_nsplit(T, N, K, e->num_sms);
  {  // experiments: B2_NS_<class index>=n forces the split factor of one projection class
    static int forced[B2_NUM_KERNEL_CLASSES];
    static bool init = false;
    if (!init) {
      for (int i = 0; i < B2_NUM_KERNEL_CLASSES; ++i) {
        char name[32];
        snprintf(name, sizeof name, "B2_NS_%d", i);
        forced[i] = env_int(name, 0);
      }
      init = true;
    }
    if (cls >= 0 && cls < B2_NUM_KERNEL_CLASSES && forced[cls] > 0) ns = std::min(forced[cls], K / 64);
  }
  while (ns > 1 && (size_t)ns * T * N > ln.partial_floats) --ns;
  if (ns > 1) {  // re-validate (every split non-empty)
    const int kb = K / 64, kps = (kb + ns - 1) / ns;
    ns = (kb + kps - 1) / kps;
  }
  a.nsplit = ns;
  KL k(e, cls);
  if (ns == 1 && !want_f32) {
    a.epilogue = TC_EPI_BF16;
    a.out = out_bf16;
    if (resid != nullptr && resid_done != nullptr && a.bn >= 32 && !k.skip()) {
      a.out = resid;
      a.resid = resid;
      *rc = launch_linear_tc(a, ln.stream);
      *resid_done = true;
      return act_bf16(nullptr);
    }
    if (e->prefill_mc && a.bn == 256 && linear_mc_shapes_ok(T, N, K) && !pf.k_pages) {
      // prefill shapes: pairs of CTAs share the token tile (gemm_mc.cu); same arithmetic, a third less L2 -> shared-memory traffic
      if (!k.skip()) *rc = launch_linear_mc(tm_w, e->tm_x[xb][bn_index(128)], out_bf16, T, N, K, ln.row0, ln.stream);
      return act_bf16(out_bf16);
    }
    if (!k.skip()) *rc = launch_linear_tc(a, ln.stream);
    return act_bf16(out_bf16);
  }
  a.epilogue = TC_EPI_F32_PARTIAL;
  a.out = ln.partial;
  if (!k.skip()) *rc = launch_linear_tc(a, ln.stream);
  return act_f32(ln.partial, ns);
}

Lane whole_batch_lane(b2_engine* e) {
  Lane ln;
  ln.stream = e->stream;
  ln.row0 = 0;
  ln.partial = e->partial;
  ln.partial_floats = e->partial_floats;
  ln.attn_scratch = e->attn_scratch;
  return ln;
}

void dbg_capture(b2_engine* e, int slot, const bf16* src, int T) {
  if (!(e->debug_flags & 1) || e->dbg_hidden == nullptr || T > e->dbg_cap_tokens) return;
  cudaMemcpyAsync(e->dbg_hidden + (size_t)slot * e->dbg_cap_tokens * e->cfg.hidden, src, (size_t)T * e->cfg.hidden * 2,
                  cudaMemcpyDeviceToDevice, e->stream);
}

// The forward pass over T tokens already described by d_tokens/d_positions/d_slots (and, for
// decode, d_seq_lens/d_block_table rows 0..T-1).  prefill_tiles > 0 selects the causal prefill
// attention; nlast = number of rows that get logits (decode: T; prefill: one per sequence,
// gathered through d_last_rows).
int lm_head_and_argmax(b2_engine* e, int nlast, const bf16* xl, XBuf xb);

// Decode forward (one token per sequence): embedding, then per layer ONE attention launch and ONE
// chain launch (chain.cu) instead of ten separate kernels.
int forward_decode_chain(b2_engine* e, int T) {
  const b2_config& c = e->cfg;
  const int H = c.hidden, I = c.intermediate, L = c.num_layers;
  {
    KL k(e, B2K_EMBED);
    if (!k.skip()) B2_TRY(launch_embed(e->d_tokens, e->embed, e->h, T, H, c.vocab, e->d_argmax, T, e->stream));
  }
  const int bn = tc_pick_bn(T), bi = bn_index(bn);
  auto gemm_cfg = [&](int N, int K) {
    ChainGemm g;
    g.N = N; g.K = K;
    // enough splits to give most SMs a tile, but at least 4 k-blocks per split: the weight tile of a
    // split is prefetched a whole step ahead, so fatter splits cost nothing and halve the partials
    int ns = tc_pick_nsplit(T, N, K, e->num_sms);
    const int kb = K / 64;
    if (kb / ns < 4) ns = std::max(1, kb / 4);
    while (ns > 1 && (size_t)ns * T * N > e->partial_floats) --ns;
    g.kps = (kb + ns - 1) / ns;
    g.nsplit = (kb + g.kps - 1) / g.kps;
    return g;
  };
  ChainParams p;
  memset(&p, 0, sizeof p);
  p.T = T; p.bn = bn;
  p.g[0] = gemm_cfg(H, e->q_size);
  p.g[1] = gemm_cfg(2 * I, H);
  p.g[2] = gemm_cfg(H, I);
  p.g[3] = gemm_cfg(e->qkv_size, H);
  p.partial = e->partial;
  p.h = e->h; p.x = e->x; p.act = e->act; p.qrot = e->qrot;
  p.cos_sin = e->cos_sin; p.positions = e->d_positions; p.slots = e->d_slots;
  p.H = H; p.I = I; p.nq = e->nq; p.nkv = e->nkv; p.block_size = c.block_size; p.eps = c.rms_eps;
  p.gbar = e->d_gbar;
  CUtensorMap xm[4] = {e->tm_x[XB_ATTN][bi], e->tm_x[XB_X][bi], e->tm_x[XB_ACT][bi], e->tm_x[XB_X][bi]};
  auto kv_of = [&](int l, bf16** kp, bf16** vp) {
    *kp = e->kv_pool + (size_t)(2 * l) * e->layer_kv_elems;
    *vp = *kp + e->layer_kv_elems;
  };
  {  // head of the forward: ln1 of layer 0 on the embeddings, QKV(0), QK-norm/RoPE/KV(0)
    LayerW& L0 = e->layers[0];
    CUtensorMap wm[4] = {L0.tm_o, L0.tm_gu, L0.tm_d, L0.tm_qkv};
    p.phase_begin = 5; p.phase_end = 8;
    p.norm_w_a = L0.ln2; p.norm_w_b = L0.ln1; p.norm_b_has_delta = 0;
    p.qw = L0.qnorm; p.kw = L0.knorm;
    kv_of(0, &p.k_pages, &p.v_pages);
    KL k(e, B2K_CHAIN);
    if (!k.skip()) B2_TRY(launch_decode_chain(wm, xm, p, e->stream));
  }
  for (int l = 0; l < L; ++l) {
    LayerW& Lw = e->layers[l];
    bf16 *kp, *vp;
    kv_of(l, &kp, &vp);
    {
      KL k(e, B2K_ATTN_DECODE);
      if (!k.skip())
        B2_TRY(launch_attn_decode(e->qrot, kp, vp, e->d_block_table, e->bt_stride, e->d_seq_lens, e->attn_scratch, e->attn, T,
                                  e->nq, e->nkv, e->hd, c.block_size, e->stream, e->attn_scale));
    }
    const bool last = (l + 1 == L);
    LayerW& Ln = e->layers[last ? l : l + 1];
    CUtensorMap wm[4] = {Lw.tm_o, Lw.tm_gu, Lw.tm_d, Ln.tm_qkv};
    p.phase_begin = 0; p.phase_end = last ? 6 : 8;
    p.norm_w_a = Lw.ln2; p.norm_w_b = last ? e->final_norm : Ln.ln1; p.norm_b_has_delta = 1;
    p.qw = Ln.qnorm; p.kw = Ln.knorm;
    kv_of(last ? l : l + 1, &p.k_pages, &p.v_pages);
    KL k(e, B2K_CHAIN);
    if (!k.skip()) B2_TRY(launch_decode_chain(wm, xm, p, e->stream));
  }
  return lm_head_and_argmax(e, T, e->x, XB_X);
}

// Decode through the fused four-step chain (chain_cluster.cu).  The head of the forward (embedding, layer 0's input
// norm, QKV and QK-norm/RoPE) and the tail (final norm, lm_head) use the separate kernels.
int forward_decode_fused(b2_engine* e, int T) {
  const b2_config& c = e->cfg;
  const int H = c.hidden, I = c.intermediate, L = c.num_layers;
  int rc = 0;
  {
    KL k(e, B2K_EMBED);
    if (!k.skip()) B2_TRY(launch_embed(e->d_tokens, e->embed, e->h, T, H, c.vocab, e->d_argmax, T, e->stream));
  }
  {
    KL k(e, B2K_NORM);
    if (!k.skip()) B2_TRY(launch_add_rmsnorm(e->h, act_bf16(nullptr), e->layers[0].ln1, e->x, T, H, c.rms_eps, e->stream));
  }
  {
    LayerW& L0 = e->layers[0];
    bf16* kp = e->kv_pool;
    bf16* vp = kp + e->layer_kv_elems;
    ActIn qkv = linear(e, whole_batch_lane(e), B2K_GEMM_QKV, XB_X, e->x, L0.wqkv, L0.tm_qkv, e->qkv, T, e->qkv_size, H, &rc);
    B2_TRY(rc);
    KL k(e, B2K_ROPE_KV);
    if (!k.skip()) B2_TRY(launch_qknorm_rope_kv(qkv, L0.qnorm, L0.knorm, e->cos_sin, e->d_positions, e->d_slots, e->qrot, kp, vp, T,
                                 e->nq, e->nkv, e->hd, c.block_size, c.rms_eps, e->stream, e->rotary_dim));
  }
  FusedChainParams p;
  memset(&p, 0, sizeof p);
  p.T = T;
  p.h = e->h; p.act = e->act; p.qrot = e->qrot;
  p.cos_sin = e->cos_sin; p.positions = e->d_positions; p.slots = e->d_slots;
  p.ss_a = e->d_ss; p.ss_b = e->d_ss + 32 * 64;
  p.H = H; p.I = I; p.nq = e->nq; p.nkv = e->nkv; p.block_size = c.block_size; p.eps = c.rms_eps;
  p.gbar = e->d_gbar2;
  const int bi = bn_index(64);
  CUtensorMap am[2] = {e->tm_x[XB_ATTN][bi], e->tm_x[XB_ACT][bi]};
  for (int l = 0; l < L; ++l) {
    LayerW& Lw = e->layers[l];
    bf16* kp = e->kv_pool + (size_t)(2 * l) * e->layer_kv_elems;
    bf16* vp = kp + e->layer_kv_elems;
    {
      KL k(e, B2K_ATTN_DECODE);
      if (!k.skip())
        B2_TRY(launch_attn_decode(e->qrot, kp, vp, e->d_block_table, e->bt_stride, e->d_seq_lens, e->attn_scratch, e->attn, T,
                                  e->nq, e->nkv, e->hd, c.block_size, e->stream, e->attn_scale));
    }
    const bool last = (l + 1 == L);
    LayerW& Ln = e->layers[last ? l : l + 1];
    CUtensorMap wm[4] = {Lw.tm_o32, Lw.tm_gu96, Lw.tm_d32, Ln.tm_qkv};
    p.phase_begin = 0; p.phase_end = last ? 3 : 4;
    p.norm_w_a = Lw.ln2; p.norm_w_b = Ln.ln1;
    p.qw = Ln.qnorm; p.kw = Ln.knorm;
    p.k_pages = e->kv_pool + (size_t)(2 * (last ? l : l + 1)) * e->layer_kv_elems;
    p.v_pages = p.k_pages + e->layer_kv_elems;
    KL k(e, B2K_CHAIN);
    if (!k.skip()) B2_TRY(launch_decode_fused_chain(wm, am, p, e->stream));
  }
  {
    KL k(e, B2K_NORM);  // the last down step has already added the residual: normalise only
    if (!k.skip()) B2_TRY(launch_add_rmsnorm(e->h, act_bf16(nullptr), e->final_norm, e->x, T, H, c.rms_eps, e->stream));
  }
  return lm_head_and_argmax(e, T, e->x, XB_X);
}

// OPT decoder (HF:models/opt/modeling_opt.py:363-391 OPTDecoder.forward, :203-254 OPTDecoderLayer.forward with
// do_layer_norm_before): same GEMM and attention kernels, the elementwise steps of elementwise_opt.cu.
int forward_opt(b2_engine* e, int T, int ntiles, int nlast) {
  const b2_config& c = e->cfg;
  const int H = c.hidden, I = c.intermediate;
  int rc = 0;
  const Lane ln = whole_batch_lane(e);
  {
    KL k(e, B2K_EMBED);
    if (!k.skip()) B2_TRY(launch_embed_pos(e->d_tokens, e->d_positions, e->embed, e->pos_embed, e->h, T, H, c.vocab, e->pos_rows,
                                           e->d_argmax, nlast, e->stream));
  }
  dbg_capture(e, 0, e->h, T);
  ActIn delta = act_bf16(nullptr);
  const bf16* delta_bias = nullptr;
  for (int l = 0; l < c.num_layers; ++l) {
    LayerW& L = e->layers[l];
    bf16* kp = e->kv_pool + (size_t)(2 * l) * e->layer_kv_elems;
    bf16* vp = kp + e->layer_kv_elems;
    {
      KL k(e, B2K_NORM);
      if (!k.skip()) B2_TRY(launch_add_layernorm(e->h, delta, delta_bias, L.ln1, L.ln1_b, e->x, T, H, c.rms_eps, e->stream));
    }
    if (l > 0) dbg_capture(e, l, e->h, T);
    ActIn qkv = linear(e, ln, B2K_GEMM_QKV, XB_X, e->x, L.wqkv, L.tm_qkv, e->qkv, T, e->qkv_size, H, &rc, KvPrefetch(), true);
    B2_TRY(rc);
    {
      // HF:...:151 "query_states = self.q_proj(hidden_states) * self.scaling", then attention with scaling 1.0
      KL k(e, B2K_ROPE_KV);
      if (!k.skip()) B2_TRY(launch_qknorm_rope_kv(qkv, nullptr, nullptr, e->cos_sin, e->d_positions, e->d_slots, e->qrot, kp, vp, T,
                                   e->nq, e->nkv, e->hd, c.block_size, c.rms_eps, e->stream, 0, L.bqkv, e->attn_scale));
    }
    if (ntiles > 0) {
      KL k(e, B2K_MISC);
      if (!k.skip()) B2_TRY(launch_attn_prefill(e->qrot, kp, vp, e->d_block_table, e->bt_stride, e->d_tiles, ntiles, e->attn,
                                 e->nq, e->nkv, e->hd, c.block_size, e->stream, 1.0f));
    } else {
      KL k(e, B2K_ATTN_DECODE);
      if (!k.skip())
        B2_TRY(launch_attn_decode(e->qrot, kp, vp, e->d_block_table, e->bt_stride, e->d_seq_lens, e->attn_scratch, e->attn, T,
                                  e->nq, e->nkv, e->hd, c.block_size, e->stream, 1.0f));
    }
    ActIn o = linear(e, ln, B2K_GEMM_O, XB_ATTN, e->attn, L.wo, L.tm_o, e->proj, T, H, e->q_size, &rc, KvPrefetch(), true);
    B2_TRY(rc);
    {
      KL k(e, B2K_NORM);
      if (!k.skip()) B2_TRY(launch_add_layernorm(e->h, o, L.bo, L.ln2, L.ln2_b, e->x, T, H, c.rms_eps, e->stream));
    }
    ActIn f1 = linear(e, ln, B2K_GEMM_GATEUP, XB_X, e->x, L.wgu, L.tm_gu, e->gu, T, I, H, &rc, KvPrefetch(), true);
    B2_TRY(rc);
    {
      KL k(e, B2K_ACT);
      if (!k.skip()) B2_TRY(launch_bias_relu(f1, L.b1, e->act, T, I, e->stream));
    }
    delta = linear(e, ln, B2K_GEMM_DOWN, XB_ACT, e->act, L.wd, L.tm_d, e->proj, T, H, I, &rc, KvPrefetch(), true);
    delta_bias = L.b2;
    B2_TRY(rc);
  }
  {
    KL k(e, B2K_NORM);
    if (!k.skip()) B2_TRY(launch_add_layernorm(e->h, delta, delta_bias, e->final_norm, e->final_norm_b, e->x, T, H, c.rms_eps, e->stream));
  }
  dbg_capture(e, c.num_layers, e->h, T);
  dbg_capture(e, c.num_layers + 1, e->x, T);
  const bf16* xl = e->x;
  XBuf xb = XB_X;
  if (ntiles > 0) {
    KL k(e, B2K_MISC);
    if (!k.skip()) B2_TRY(launch_gather_rows(e->x, e->d_last_rows, e->xlast, nlast, H, e->stream));
    xl = e->xlast;
    xb = XB_LAST;
  }
  return lm_head_and_argmax(e, nlast, xl, xb);
}

// The decoder layers and the final norm over the lane's rows [row0, row0 + T): everything between the embedding and
// the lm_head.  Qwen3 / Phi-3 (HF:models/qwen3/modeling_qwen3.py:423-434).
int forward_layers(b2_engine* e, const Lane& ln, int T, int ntiles) {
  const b2_config& c = e->cfg;
  const int H = c.hidden, I = c.intermediate;
  const size_t r0 = (size_t)ln.row0;
  cudaStream_t st = ln.stream;
  const int tag = ln.row0 != 0 ? 1 : 0;  // lane marker in the step timeline
  bf16* h = e->h + r0 * H;
  bf16* x = e->x + r0 * H;
  bf16* qkvb = e->qkv + r0 * e->qkv_size;
  bf16* qrot = e->qrot + r0 * e->q_size;
  bf16* attn = e->attn + r0 * e->q_size;
  bf16* gub = e->gu + r0 * 2 * I;
  bf16* act = e->act + r0 * I;
  bf16* proj = e->proj + r0 * H;
  const int32_t* positions = e->d_positions + r0;
  const int32_t* slots = e->d_slots + r0;
  const int32_t* block_table = e->d_block_table + r0 * e->bt_stride;
  const int32_t* seq_lens = e->d_seq_lens + r0;
  int rc = 0;
  ActIn delta = act_bf16(nullptr);
  // decode only: the projections of layer l prefetch shares of the K/V pages the next attention pass
  // will read (QKV: this layer's own; O, gate-up, down: layer l+1's)
  auto prefetch_for = [&](int layer, int part) -> KvPrefetch {
    KvPrefetch pf;
    if (ntiles > 0 || e->kv_l2_budget <= 0 || layer >= c.num_layers) return pf;
    pf.k_pages = e->kv_pool + (size_t)(2 * layer) * e->layer_kv_elems;
    pf.v_pages = pf.k_pages + e->layer_kv_elems;
    pf.block_table = block_table;
    pf.seq_lens = seq_lens;
    pf.frac_q16 = e->d_pf_frac;
    pf.bt_stride = e->bt_stride;
    pf.B = T;
    pf.nkv = e->nkv;
    pf.part = part;
    pf.nparts = 4;
    return pf;
  };
  for (int l = 0; l < c.num_layers; ++l) {
    LayerW& L = e->layers[l];
    bf16* kp = e->kv_pool + (size_t)(2 * l) * e->layer_kv_elems;
    bf16* vp = kp + e->layer_kv_elems;
    // decode-sized forwards: the fused cluster-pair launches of gemm_pair.cu take the place of (projection, element-wise
    // kernel) pairs where the model's shapes allow it
    const bool pair_ok = ntiles == 0 && !ln.beside_attention && e->force_gemm_impl != 0 && T > e->simt_max_tokens && pair_shapes_ok(T, H) &&
                         !(e->debug_flags & 1) && e->kv_l2_budget <= 0;
    const bool qkv_pair = pair_ok && (e->use_pair & 2) && c.arch == B2_ARCH_QWEN3 && e->native_geom && e->rotary_dim == 128 && L.qnorm != nullptr &&
                          !((e->skip_mask >> B2K_ROPE_KV) & 1u);
    const bool gu_pair = pair_ok && (e->use_pair & 1) && I % 64 == 0 && !((e->skip_mask >> B2K_ACT) & 1u);
    // ... and with both in place the O / down projections add into the residual stream themselves (gemm_resid.cu) and the
    // pair kernels normalise h as they load it: no norm kernel inside the layer loop (skip-mask ablations take the old path)
    const bool resid = qkv_pair && gu_pair && (e->use_pair & 4) && e->ssq != nullptr && H / 16 <= 64 && resid_shapes_ok(T, H, e->q_size) &&
                       resid_shapes_ok(T, H, I) && !((e->skip_mask >> B2K_NORM) & 1u) && !((e->skip_mask >> B2K_GEMM_O) & 1u) &&
                       !((e->skip_mask >> B2K_GEMM_DOWN) & 1u) && !((e->skip_mask >> B2K_GEMM_QKV) & 1u) && !((e->skip_mask >> B2K_GEMM_GATEUP) & 1u);
    PairNorm pn1, pn2;
    if (resid) {
      pn1.ssq = pn2.ssq = e->ssq;
      pn1.gain = L.ln1; pn2.gain = L.ln2;
      pn1.pieces = pn2.pieces = H / 16;
      pn1.ld = pn2.ld = 64;
      pn1.eps = pn2.eps = c.rms_eps;
    }
    // resid path: each launch's idle lanes pull the NEXT launch's weights into L2 (the fat CTAs of these kernels cannot be
    // resident beside each other, so a launch can no longer start its own weight stream while its predecessor runs)
    L2Prefetch pf_gu, pf_d, pf_qkv_next;
    if (resid && e->l2_weight_prefetch) {
      pf_gu.ptr = L.wgu; pf_gu.bytes = (unsigned long long)2 * I * H * 2;
      pf_d.ptr = L.wd; pf_d.bytes = (unsigned long long)H * I * 2;
      if (l + 1 < c.num_layers) { pf_qkv_next.ptr = e->layers[l + 1].wqkv; pf_qkv_next.bytes = (unsigned long long)e->qkv_size * H * 2; }
      pn2.next_w = pf_d;
    }
    if (!(resid && l > 0)) {  // layer 0 (h comes from the embedding kernel) and every layer of the stand-alone path
      KL k(e, B2K_NORM);
      if (!k.skip()) B2_TRY(launch_add_rmsnorm(h, delta, L.ln1, x, T, H, c.rms_eps, st, tag));
    }
    if (l > 0) dbg_capture(e, l, e->h, T);
    const bool qkv_prefill_fused = ntiles > 0 && e->qkv_rope_prefill_fused && e->prefill_mc && e->force_gemm_impl != 0 && c.arch == B2_ARCH_QWEN3 &&
                                   e->native_geom && e->rotary_dim == 128 && L.qnorm != nullptr && linear_mc_shapes_ok(T, e->qkv_size, H) &&
                                   !((e->skip_mask >> B2K_ROPE_KV) & 1u) && !((e->skip_mask >> B2K_GEMM_QKV) & 1u);
    if (qkv_prefill_fused) {
      KL k(e, B2K_GEMM_QKV);  // prefill: QK-norm + RoPE + page write in the QKV projection's epilogue (gemm_mc.cu)
      B2_TRY(launch_qkv_rope_prefill(L.tm_qkv, e->tm_x[XB_X][bn_index(128)], L.qnorm, L.knorm, e->cos_sin, positions, slots, qrot, kp, vp, T, e->nq,
                                     e->nkv, H, c.block_size, c.rms_eps, ln.row0, st));
    } else if (qkv_pair) {
      KL k(e, B2K_GEMM_QKV);
      if (!k.skip())
        B2_TRY(launch_qkv_rope(L.tm_qkv, (resid && l > 0) ? h : x, L.qnorm, L.knorm, e->cos_sin, positions, slots, qrot, kp, vp, T, e->nq, e->nkv, H,
                               c.block_size, c.rms_eps, st, tag, (resid && l > 0) ? pn1 : PairNorm()));
    } else {
      ActIn qkv = linear(e, ln, B2K_GEMM_QKV, XB_X, x, L.wqkv, L.tm_qkv, qkvb, T, e->qkv_size, H, &rc, prefetch_for(l, 3));
      B2_TRY(rc);
      KL k(e, B2K_ROPE_KV);
      if (!k.skip()) B2_TRY(launch_qknorm_rope_kv(qkv, L.qnorm, L.knorm, e->cos_sin, positions, slots, qrot, kp, vp, T,
                                   e->nq, e->nkv, e->hd, c.block_size, c.rms_eps, st, e->rotary_dim, nullptr, 1.0f, tag));
    }
    if (ntiles > 0) {
      KL k(e, B2K_MISC);
      if (!k.skip()) {
        if (e->prefill_tc)
          B2_TRY(launch_attn_prefill_tc(qrot, kp, vp, e->num_blocks, block_table, e->bt_stride, e->d_tiles, ntiles, attn, e->nq, e->nkv, e->hd,
                                        c.block_size, st, e->attn_scale, c.sliding_window, prefill_tile_pairs(e) ? 1 : 0));
        else
          B2_TRY(launch_attn_prefill(qrot, kp, vp, block_table, e->bt_stride, e->d_tiles, ntiles, attn,
                                     e->nq, e->nkv, e->hd, c.block_size, st, e->attn_scale, c.sliding_window));
      }
    } else {
      KL k(e, B2K_ATTN_DECODE);
      if (!k.skip())
        B2_TRY(launch_attn_decode(qrot, kp, vp, block_table, e->bt_stride, seq_lens, ln.attn_scratch, attn, T,
                                  e->nq, e->nkv, e->hd, c.block_size, st, e->attn_scale, c.sliding_window, tag));
    }
    if (resid) {
      {
        KL k(e, B2K_GEMM_O);
        B2_TRY(launch_linear_resid(L.tm_o, attn, h, e->ssq, T, H, e->q_size, 64, st, tag, pf_gu));
      }
      {
        KL k(e, B2K_GEMM_GATEUP);
        B2_TRY(launch_gateup_swiglu(L.tm_gu64, h, act, T, I, H, st, tag, pn2));
      }
      KL k(e, B2K_GEMM_DOWN);
      B2_TRY(launch_linear_resid(L.tm_d, act, h, e->ssq, T, H, I, 64, st, tag, pf_qkv_next));
      delta = act_bf16(nullptr);  // h is up to date: the next norm (layer 0's successor never runs one; the final norm) adds nothing
      continue;
    }
    // prefill-sized forwards, opt-in (B2_RESID_FUSED=1): the O and down projections add their output to the residual stream in
    // their epilogue (gemm_tc.cu, `resid`), and the norm kernel behind them only normalises: half its traffic.  Measured at C2:
    // the projection grows by 3.8 us (32 two-byte loads per thread and chunk, behind the accumulator read), the norm kernel
    // shrinks by 1.7 us: 44.9 against 42.4 ms for the prefill on the same box.  Off.
    bf16* fuse_resid = (ntiles > 0 && e->prefill_resid_fused && !((e->skip_mask >> B2K_NORM) & 1u)) ? h : nullptr;
    bool o_added = false;
    ActIn o = linear(e, ln, B2K_GEMM_O, XB_ATTN, attn, L.wo, L.tm_o, proj, T, H, e->q_size, &rc, prefetch_for(l + 1, 0), false,
                     /*after_attention=*/ntiles == 0, fuse_resid, &o_added);
    B2_TRY(rc);
    {
      KL k(e, B2K_NORM);
      if (!k.skip()) B2_TRY(launch_add_rmsnorm(h, o_added ? act_bf16(nullptr) : o, L.ln2, x, T, H, c.rms_eps, st, tag));
    }
    if (gu_pair) {
      KL k(e, B2K_GEMM_GATEUP);
      if (!k.skip()) B2_TRY(launch_gateup_swiglu(L.tm_gu64, x, act, T, I, H, st, tag));
    } else if (e->gu_prefill_fused && e->force_gemm_impl != 0 && gateup_prefill_shapes_ok(T, I, H) && !((e->skip_mask >> B2K_ACT) & 1u)) {
      KL k(e, B2K_GEMM_GATEUP);  // prefill-sized: SwiGLU in the projection's epilogue, no gu round trip through HBM
      if (!k.skip()) B2_TRY(launch_gateup_swiglu_prefill(L.tm_gu, e->tm_x[XB_X][bn_index(128)], act, T, I, H, ln.row0, st));
    } else {
      ActIn gu = linear(e, ln, B2K_GEMM_GATEUP, XB_X, x, L.wgu, L.tm_gu, gub, T, 2 * I, H, &rc, prefetch_for(l + 1, 1));
      B2_TRY(rc);
      KL k(e, B2K_ACT);
      if (!k.skip()) B2_TRY(launch_silu_mul(gu, act, T, I, st, tag));
    }
    bool d_added = false;
    delta = linear(e, ln, B2K_GEMM_DOWN, XB_ACT, act, L.wd, L.tm_d, proj, T, H, I, &rc, prefetch_for(l + 1, 2), false, false, fuse_resid, &d_added);
    B2_TRY(rc);
    if (d_added) delta = act_bf16(nullptr);
  }
  {
    KL k(e, B2K_NORM);
    if (!k.skip()) B2_TRY(launch_add_rmsnorm(h, delta, e->final_norm, x, T, H, c.rms_eps, st, tag));
  }
  return 0;
}

// Two-lane decode.  At batch 64 a decode step is two things that do not overlap when they depend on each other: the
// attention pass of a layer (HBM-bound, about half the step) and the eight small kernels between two attention
// passes (a chain of dependent launches that moves 31 MB of weights in ~30 us: latency-bound, HBM nearly idle).
// Sequences are independent, so the batch is cut into two halves that run as two dependent chains on two streams of
// the same step (and the same captured graph): while one half is in its latency-bound chain the other half streams
// its K/V pages.  Both attention kernels want every SM (129 KB of shared memory per CTA), so they take turns by
// themselves and the halves fall into opposite phases; the projections of a lane use the short stage ring (<= 82 KB)
// so that their CTAs are resident NEXT to the other lane's attention CTAs instead of queueing behind them.  The price
// is that the weights are streamed once per lane (L2 keeps part of the second pass).
bool two_lane_decode(const b2_engine* e, int T, int ntiles) {
  return ntiles == 0 && e->mb_min_batch > 0 && T >= e->mb_min_batch && T <= 128 && T % 2 == 0 && e->stream2 != nullptr &&
         !e->profiling && !(e->debug_flags & 1) && e->force_gemm_impl != 0 && (T / 2) > e->simt_max_tokens;
}

int forward(b2_engine* e, int T, int ntiles, int nlast, int attn_split) {
  const b2_config& c = e->cfg;
  const int H = c.hidden;
  (void)attn_split;
  if (c.arch == B2_ARCH_OPT) return forward_opt(e, T, ntiles, nlast);
  if (ntiles == 0 && e->use_chain == 2 && e->fused_ok && c.arch == B2_ARCH_QWEN3 && !(e->debug_flags & 1) &&
      e->force_gemm_impl != 0 && T > e->simt_max_tokens && T <= 64 && e->skip_mask == 0)
    return forward_decode_fused(e, T);
  if (ntiles == 0 && e->use_chain == 1 && c.arch == B2_ARCH_QWEN3 && !(e->debug_flags & 1) && e->force_gemm_impl != 0 && T > e->simt_max_tokens)
    return forward_decode_chain(e, T);
  {
    KL k(e, B2K_EMBED);
    if (!k.skip()) B2_TRY(launch_embed(e->d_tokens, e->embed, e->h, T, H, c.vocab, e->d_argmax, nlast, e->stream));
  }
  dbg_capture(e, 0, e->h, T);
  if (two_lane_decode(e, T, ntiles)) {
    Lane a = whole_batch_lane(e), b;
    a.beside_attention = true;
    b.stream = e->stream2;
    b.row0 = T / 2;
    b.partial = e->partial2;
    b.partial_floats = e->partial2_floats;
    b.attn_scratch = e->attn_scratch2;
    b.beside_attention = true;
    // fork: lane b starts once the embedding is done (a full dependency: no programmatic edge across streams)
    B2_CUDA_OK(cudaEventRecord(e->ev_fork, e->stream));
    B2_CUDA_OK(cudaStreamWaitEvent(e->stream2, e->ev_fork, 0));
    b2_pdl_skip_next();
    int rc_b = forward_layers(e, b, T - T / 2, 0);
    int rc_a = forward_layers(e, a, T / 2, 0);
    // join (always, so that a capture in progress is left with both streams merged)
    cudaError_t j1 = cudaEventRecord(e->ev_join, e->stream2);
    cudaError_t j2 = cudaStreamWaitEvent(e->stream, e->ev_join, 0);
    B2_TRY(rc_a);
    B2_TRY(rc_b);
    B2_CUDA_OK(j1);
    B2_CUDA_OK(j2);
    b2_pdl_skip_next();  // the lm_head waits for both lanes in full
  } else {
    B2_TRY(forward_layers(e, whole_batch_lane(e), T, ntiles));
  }
  dbg_capture(e, c.num_layers, e->h, T);
  dbg_capture(e, c.num_layers + 1, e->x, T);

  // lm_head on the last position of every sequence + greedy argmax
  const bf16* xl = e->x;
  XBuf xb = XB_X;
  if (ntiles > 0) {
    KL k(e, B2K_MISC);
    if (!k.skip()) B2_TRY(launch_gather_rows(e->x, e->d_last_rows, e->xlast, nlast, H, e->stream));
    xl = e->xlast;
    xb = XB_LAST;
  }
  return lm_head_and_argmax(e, nlast, xl, xb);
}

int lm_head_and_argmax(b2_engine* e, int nlast, const bf16* xl, XBuf xb) {
  const b2_config& c = e->cfg;
  const int H = c.hidden;
  bf16* logits = (e->debug_flags & 2) ? e->logits : nullptr;
  int impl = e->force_gemm_impl >= 0 ? e->force_gemm_impl : (nlast <= e->simt_max_tokens ? 0 : 1);
  if (impl == 0) {
    if (e->logits == nullptr) {
      b2_set_error("SIMT lm_head path needs the logits buffer (b2_set_debug(e, 2))");
      return -22;
    }
    {
      KL k(e, B2K_LMHEAD);
      if (!k.skip()) B2_TRY(launch_linear_simt(xl, e->lm_head, e->logits, 0, nlast, c.vocab, H, e->stream));
    }
    KL k2(e, B2K_MISC);
    if (!k2.skip()) B2_TRY(launch_argmax_rows(e->logits, e->d_out_tokens, nlast, c.vocab, e->stream));
  } else {
    TcGemmArgs a;
    // operands trade places for the fused arg-max: tokens on MMA-M (128-row box of X), 128 vocabulary
    // rows per tile on MMA-N
    a.bn = 128;
    a.tmap_w = e->tm_x[xb][bn_index(128)];
    a.tmap_x = e->tm_lm_head;
    a.T = nlast; a.N = c.vocab; a.K = H;
    a.nsplit = 1;
    a.epilogue = TC_EPI_ARGMAX_ROWS;
    a.out = logits;
    a.argmax_packed = e->d_argmax;
    {
      KL k(e, B2K_LMHEAD);
      if (!k.skip()) B2_TRY(launch_linear_tc(a, e->stream));
    }
    KL k2(e, B2K_MISC);
    if (!k2.skip()) B2_TRY(launch_argmax_finalize(e->d_argmax, e->d_out_tokens, nlast, e->stream));
  }
  return 0;
}

__global__ void publish_tokens_kernel(const int32_t* __restrict__ out_tokens, int32_t* __restrict__ next_in,
                                      int32_t* __restrict__ log, int32_t* step, int B, int log_cap_steps, int stride) {
  pdl_wait();
  pdl_launch_dependents();
  const int st = *step;  // single CTA
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    next_in[b] = out_tokens[b];
    // the log has `stride` (= max_seqs) columns; B is the padded bucket and may be wider
    if (st < log_cap_steps && b < stride) log[(size_t)st * stride + b] = out_tokens[b];
  }
  __syncthreads();
  if (threadIdx.x == 0) *step = st < 0x7fffffff ? st + 1 : st;  // saturates: a server steps for weeks without b2_decode_run resetting it
}

// decode step body recorded into a graph (or run eagerly): bookkeeping, forward, token feedback
int decode_body(b2_engine* e, int B, int attn_split) {
  {
    KL k(e, B2K_MISC);
    if (!k.skip()) B2_TRY(launch_decode_advance(e->d_positions, e->d_seq_lens, e->d_slots, e->d_block_table, e->bt_stride,
                                 e->cfg.block_size, B, e->d_pf_frac, e->kv_l2_budget,
                                 (long long)e->nkv * e->cfg.block_size * e->hd * 2 * 2, e->stream));
  }
  B2_TRY(forward(e, B, 0, B, attn_split));
  {
    KL k(e, B2K_MISC);
    if (!k.skip()) B2_CUDA_OK(b2_launch(publish_tokens_kernel, dim3(1), dim3(256), 0, e->stream, (const int32_t*)e->d_out_tokens, e->d_tokens,
                         e->d_token_log, e->d_step, B, e->log_cap_steps, e->cfg.max_seqs));
  }
  return 0;
}

// Captures (once per (batch bucket, attention split)) the decode step as a CUDA graph.  Capture
// does not advance any state: the eager warm-up that precedes it runs on saved bookkeeping.
int ensure_graph(b2_engine* e, int B, int attn_split) {
  GraphKey key{B, attn_split};
  if (e->graphs.count(key)) return 0;
  // Warm every kernel once outside capture (function attributes, module load), then restore the
  // bookkeeping so the real state is not advanced twice.  (The warm-up writes K/V for position
  // pos+1 with the same values the real step will write.)
  std::vector<int32_t> save(3 * (size_t)B + 1);
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  B2_CUDA_OK(cudaMemcpy(save.data(), e->d_positions, B * 4, cudaMemcpyDeviceToHost));
  B2_CUDA_OK(cudaMemcpy(save.data() + B, e->d_seq_lens, B * 4, cudaMemcpyDeviceToHost));
  B2_CUDA_OK(cudaMemcpy(save.data() + 2 * B, e->d_tokens, B * 4, cudaMemcpyDeviceToHost));
  B2_CUDA_OK(cudaMemcpy(save.data() + 3 * B, e->d_step, 4, cudaMemcpyDeviceToHost));
  const int64_t l0 = e->launches;
  B2_TRY(decode_body(e, B, attn_split));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  const int64_t nodes = e->launches - l0;
  e->launches = l0;
  B2_CUDA_OK(cudaMemcpy(e->d_positions, save.data(), B * 4, cudaMemcpyHostToDevice));
  B2_CUDA_OK(cudaMemcpy(e->d_seq_lens, save.data() + B, B * 4, cudaMemcpyHostToDevice));
  B2_CUDA_OK(cudaMemcpy(e->d_tokens, save.data() + 2 * B, B * 4, cudaMemcpyHostToDevice));
  B2_CUDA_OK(cudaMemcpy(e->d_step, save.data() + 3 * B, 4, cudaMemcpyHostToDevice));
  cudaGraph_t g = nullptr;
  B2_CUDA_OK(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
  int rc = decode_body(e, B, attn_split);
  cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
  e->launches = l0;
  if (rc != 0) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  B2_CUDA_OK(ce);
  cudaGraphExec_t ge = nullptr;
  B2_CUDA_OK(cudaGraphInstantiate(&ge, g, 0));
  cudaGraphDestroy(g);
  e->graphs[key] = ge;
  e->graph_nodes[key] = nodes;
  return 0;
}

// hidden-state capture needs eager launches; the logits buffer is just another kernel argument, so it
// only selects a different graph (attn_split slot of the key: 1 = argmax only, 2 = logits written too)
bool graphs_enabled(b2_engine* e) { return e->cfg.use_cuda_graphs && !e->profiling && !(e->debug_flags & 1); }

int run_decode_step(b2_engine* e, int B, int attn_split) {
  if (!graphs_enabled(e)) return decode_body(e, B, attn_split);
  B2_TRY(ensure_graph(e, B, attn_split));
  GraphKey key{B, attn_split};
  B2_CUDA_OK(cudaGraphLaunch(e->graphs[key], e->stream));
  e->launches += e->graph_nodes[key];
  return 0;
}

// Uploads the batch state of `seq_ids` (decode slow path).  Device positions hold len-1 so the
// advance kernel at the head of the step brings them to the position of the appended token.
int upload_decode_batch(b2_engine* e, int n, const int32_t* seq_ids, int B) {
  const int bs = e->cfg.block_size;
  size_t need = (size_t)3 * B + (size_t)B * e->bt_stride;
  if (need > e->h_stage_ints) {
    b2_set_error("decode batch of %d does not fit the staging buffer", B);
    return -12;
  }
  int32_t* pos = e->h_stage;
  int32_t* lens = pos + B;
  int32_t* slots = lens + B;
  int32_t* bt = slots + B;
  memset(e->h_stage, 0, need * sizeof(int32_t));
  for (int i = 0; i < B; ++i) {
    if (i < n) {
      Seq& s = e->seqs[seq_ids[i]];
      pos[i] = s.len - 1;
      lens[i] = s.len;
      slots[i] = -1;
      memcpy(bt + (size_t)i * e->bt_stride, s.blocks.data(), s.blocks.size() * sizeof(int32_t));
    } else {  // padding rows of the graph bucket: no KV write, empty attention
      pos[i] = -2;  // sentinel: decode_advance leaves padding rows inert
      lens[i] = 0;
      slots[i] = -1;
    }
  }
  (void)bs;
  B2_CUDA_OK(cudaMemcpyAsync(e->d_positions, pos, B * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_seq_lens, lens, B * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_slots, slots, B * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_block_table, bt, (size_t)B * e->bt_stride * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));  // staging buffer is reused
  e->cur_batch.assign(seq_ids, seq_ids + n);
  e->cur_B = B;
  return 0;
}

int check_seq(b2_engine* e, int id) {
  if (id < 0 || id >= (int)e->seqs.size() || !e->seqs[id].active) {
    b2_set_error("invalid sequence id %d", id);
    return -22;
  }
  return 0;
}

// ---- page allocator with the prefix cache in it ----
int64_t pages_available(const b2_engine* e) { return (int64_t)e->free_blocks.size() + (int64_t)e->pc_lru.size(); }

void pc_unregister(b2_engine* e, int32_t b) {
  if (e->blk_hash[b] == 0) return;
  auto it = e->pc_map.find(e->blk_hash[b]);
  if (it != e->pc_map.end() && it->second == b) e->pc_map.erase(it);
  e->blk_hash[b] = 0;
}

// a page for exclusive use: a free one, else the registered page that has gone unused the longest
int32_t take_page(b2_engine* e) {
  if (!e->free_blocks.empty()) {
    const int32_t b = e->free_blocks.back();
    e->free_blocks.pop_back();
    e->blk_ref[b] = 1;
    return b;
  }
  const int32_t b = e->pc_lru.front();
  e->pc_lru.pop_front();
  e->pc_in_lru[b] = 0;
  pc_unregister(e, b);
  e->blk_ref[b] = 1;
  return b;
}

void release_page(b2_engine* e, int32_t b) {
  if (--e->blk_ref[b] > 0) return;
  e->blk_ref[b] = 0;
  if (e->prefix_cache && e->blk_hash[b] != 0) {  // still worth something to a later prompt
    e->pc_lru.push_back(b);
    e->pc_lru_pos[b] = std::prev(e->pc_lru.end());
    e->pc_in_lru[b] = 1;
  } else {
    pc_unregister(e, b);
    e->free_blocks.push_back(b);
  }
}

uint64_t pc_hash(uint64_t parent, const int32_t* tok) {
  uint64_t h = parent ^ 0x9e3779b97f4a7c15ull;
  for (int i = 0; i < 16; ++i) {
    h ^= (uint64_t)(uint32_t)tok[i] + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
  }
  return h ? h : 1;  // 0 means "not registered"
}

// Leading pages of a fresh sequence's prompt that are already in the cache: the sequence drops its own pages for them
// and holds the cached ones.  At least one token is always left to compute (its logits are the call's result).
int pc_match(b2_engine* e, Seq& s, const int32_t* ids, int n) {
  const int bs = e->cfg.block_size;
  int matched = 0;
  uint64_t h = 0;
  int32_t parent = -1;
  const int max_pages = (n - 1) / bs;
  for (int j = 0; j < max_pages; ++j) {
    h = pc_hash(h, ids + (size_t)j * bs);
    auto it = e->pc_map.find(h);
    if (it == e->pc_map.end()) break;
    const int32_t b = it->second;
    if (e->blk_parent[b] != parent || memcmp(&e->blk_tokens[(size_t)b * bs], ids + (size_t)j * bs, bs * sizeof(int32_t)) != 0) break;
    if (e->blk_ref[b] == 0 && e->pc_in_lru[b]) {
      e->pc_lru.erase(e->pc_lru_pos[b]);
      e->pc_in_lru[b] = 0;
    }
    e->blk_ref[b] += 1;
    release_page(e, s.blocks[j]);  // the private page reserved for these positions is not needed
    s.blocks[j] = b;
    s.chain.push_back(h);
    parent = b;
    matched += bs;
  }
  return matched;
}

// After a prefill call: pages of this sequence that the call's tokens completed get registered.
void pc_register(b2_engine* e, Seq& s, const int32_t* ids, int n) {
  const int bs = e->cfg.block_size;
  if (!s.chain_open) return;
  for (int i = 0; i < n; ++i) {
    s.pending.push_back(ids[i]);
    if ((int)s.pending.size() < bs) continue;
    const int j = (int)s.chain.size();
    const uint64_t h = pc_hash(j > 0 ? s.chain[j - 1] : 0, s.pending.data());
    const int32_t b = s.blocks[j];
    if (e->blk_hash[b] == 0 && e->pc_map.find(h) == e->pc_map.end()) {
      e->blk_hash[b] = h;
      e->blk_parent[b] = j > 0 ? s.blocks[j - 1] : -1;
      // the parent must itself be the registered holder of its hash, or a later match could not verify the chain
      if (j > 0 && (e->blk_hash[s.blocks[j - 1]] != s.chain[j - 1])) e->blk_hash[b] = 0;
      if (e->blk_hash[b] != 0) {
        memcpy(&e->blk_tokens[(size_t)b * bs], s.pending.data(), bs * sizeof(int32_t));
        e->pc_map[h] = b;
      }
    }
    s.chain.push_back(h);
    s.pending.clear();
  }
}

int ensure_logits(b2_engine* e) {
  if (e->logits) return 0;
  // rows for the padded graph bucket: a captured step runs the lm_head over bucket_batch(n) rows, and max_seqs need
  // not be a bucket value (--max-num-seqs is free-form)
  return dalloc(&e->logits, (size_t)bucket_batch(std::max(e->cfg.max_seqs, 1)) * e->cfg.vocab);
}

int layer_tensor(b2_engine* e, const std::string& name, const void* host, size_t nbytes) {
  const b2_config& c = e->cfg;
  const int H = c.hidden, I = c.intermediate;
  auto put = [&](bf16* dst, size_t elems) -> int {
    if (nbytes != elems * 2) {
      b2_set_error("tensor %s: expected %zu bytes, got %zu", name.c_str(), elems * 2, nbytes);
      return -22;
    }
    B2_CUDA_OK(cudaMemcpy(dst, host, nbytes, cudaMemcpyHostToDevice));
    e->loaded[name] = 1;
    return 0;
  };
  // Per-head blocks of the attention projections go through the internal geometry (see b2_engine):
  // head h of the checkpoint lands at internal head h * step, its head_dim rows (or columns, for the
  // output projection) at the start of a 128-wide slot; everything else in the slot stays zero.
  const int hd_r = c.head_dim;
  auto put_head_rows = [&](bf16* dst, const uint16_t* src, int heads, int step, int slots) -> int {
    std::vector<uint16_t> stage((size_t)slots * 128 * H, 0);
    for (int h = 0; h < heads; ++h)
      memcpy(&stage[(size_t)h * step * 128 * H], src + (size_t)h * hd_r * H, (size_t)hd_r * H * 2);
    B2_CUDA_OK(cudaMemcpy(dst, stage.data(), stage.size() * 2, cudaMemcpyHostToDevice));
    return 0;
  };
  auto put_head_cols = [&](bf16* dst, const uint16_t* src, int heads, int step, int slots) -> int {
    const size_t in_cols = (size_t)heads * hd_r, out_cols = (size_t)slots * 128;
    std::vector<uint16_t> stage((size_t)H * out_cols, 0);
    for (int r = 0; r < H; ++r)
      for (int h = 0; h < heads; ++h)
        memcpy(&stage[(size_t)r * out_cols + (size_t)h * step * 128], src + (size_t)r * in_cols + (size_t)h * hd_r, (size_t)hd_r * 2);
    B2_CUDA_OK(cudaMemcpy(dst, stage.data(), stage.size() * 2, cudaMemcpyHostToDevice));
    return 0;
  };
  auto expect = [&](size_t elems) -> int {
    if (nbytes != elems * 2) {
      b2_set_error("tensor %s: expected %zu bytes, got %zu", name.c_str(), elems * 2, nbytes);
      return -22;
    }
    return 0;
  };
  const size_t q_r = (size_t)c.num_q_heads * hd_r, kv_r = (size_t)c.num_kv_heads * hd_r;  // checkpoint widths
  const uint16_t* src16 = reinterpret_cast<const uint16_t*>(host);
  if (c.arch == B2_ARCH_OPT) {
    // HF:models/opt/modeling_opt.py state dict: model.decoder.{embed_tokens,embed_positions,final_layer_norm,layers.N.*}
    auto put_head_vec = [&](bf16* dst, int heads, int step) -> int {  // bias of a per-head projection, zero in the padding
      std::vector<uint16_t> stage((size_t)heads * step * 128, 0);
      for (int h = 0; h < heads; ++h) memcpy(&stage[(size_t)h * step * 128], src16 + (size_t)h * hd_r, (size_t)hd_r * 2);
      B2_CUDA_OK(cudaMemcpy(dst, stage.data(), stage.size() * 2, cudaMemcpyHostToDevice));
      return 0;
    };
    auto mark = [&](int rc) {
      if (rc == 0) e->loaded[name] = 1;
      return rc;
    };
    const std::string dec = "model.decoder.";
    if (name == "lm_head.weight") {
      e->loaded[name] = 1;  // tied to embed_tokens
      return 0;
    }
    if (name.compare(0, dec.size(), dec) != 0) {
      b2_set_error("unknown tensor name %s", name.c_str());
      return -22;
    }
    const std::string r0 = name.substr(dec.size());
    if (r0 == "embed_tokens.weight") return put(e->embed, (size_t)c.vocab * H);
    if (r0 == "embed_positions.weight") return put(e->pos_embed, (size_t)e->pos_rows * H);
    if (r0 == "final_layer_norm.weight") return put(e->final_norm, H);
    if (r0 == "final_layer_norm.bias") return put(e->final_norm_b, H);
    const std::string lp = "layers.";
    size_t d2 = r0.find('.', lp.size());
    if (r0.compare(0, lp.size(), lp) != 0 || d2 == std::string::npos) {
      b2_set_error("unknown tensor name %s", name.c_str());
      return -22;
    }
    const int li = atoi(r0.substr(lp.size(), d2 - lp.size()).c_str());
    if (li < 0 || li >= c.num_layers) {
      b2_set_error("tensor %s: layer out of range", name.c_str());
      return -22;
    }
    LayerW& L = e->layers[li];
    const std::string rest = r0.substr(d2 + 1);
    bf16* wq = L.wqkv;
    bf16* wk = L.wqkv + (size_t)e->q_size * H;
    bf16* wv = L.wqkv + (size_t)(e->q_size + e->kv_size) * H;
    if (rest == "self_attn_layer_norm.weight") return put(L.ln1, H);
    if (rest == "self_attn_layer_norm.bias") return put(L.ln1_b, H);
    if (rest == "final_layer_norm.weight") return put(L.ln2, H);
    if (rest == "final_layer_norm.bias") return put(L.ln2_b, H);
    if (rest == "self_attn.q_proj.weight") { B2_TRY(expect(q_r * H)); return mark(put_head_rows(wq, src16, c.num_q_heads, e->qhead_step, e->nq)); }
    if (rest == "self_attn.k_proj.weight") { B2_TRY(expect(kv_r * H)); return mark(put_head_rows(wk, src16, c.num_kv_heads, 1, e->nkv)); }
    if (rest == "self_attn.v_proj.weight") { B2_TRY(expect(kv_r * H)); return mark(put_head_rows(wv, src16, c.num_kv_heads, 1, e->nkv)); }
    if (rest == "self_attn.out_proj.weight") { B2_TRY(expect((size_t)H * q_r)); return mark(put_head_cols(L.wo, src16, c.num_q_heads, e->qhead_step, e->nq)); }
    if (rest == "self_attn.q_proj.bias") { B2_TRY(expect(q_r)); return mark(put_head_vec(L.bqkv, c.num_q_heads, e->qhead_step)); }
    if (rest == "self_attn.k_proj.bias") { B2_TRY(expect(kv_r)); return mark(put_head_vec(L.bqkv + e->q_size, c.num_kv_heads, 1)); }
    if (rest == "self_attn.v_proj.bias") { B2_TRY(expect(kv_r)); return mark(put_head_vec(L.bqkv + e->q_size + e->kv_size, c.num_kv_heads, 1)); }
    if (rest == "self_attn.out_proj.bias") return put(L.bo, H);
    if (rest == "fc1.weight") return put(L.wgu, (size_t)I * H);
    if (rest == "fc1.bias") return put(L.b1, I);
    if (rest == "fc2.weight") return put(L.wd, (size_t)H * I);
    if (rest == "fc2.bias") return put(L.b2, H);
    b2_set_error("unknown tensor name %s", name.c_str());
    return -22;
  }
  if (name == "model.embed_tokens.weight") return put(e->embed, (size_t)c.vocab * H);
  if (name == "model.norm.weight") return put(e->final_norm, H);
  if (name == "lm_head.weight") {
    if (c.tie_embeddings) {
      e->loaded[name] = 1;  // tied: ignore the duplicate
      return 0;
    }
    return put(e->lm_head, (size_t)c.vocab * H);
  }
  const std::string pre = "model.layers.";
  if (name.compare(0, pre.size(), pre) != 0) {
    b2_set_error("unknown tensor name %s", name.c_str());
    return -22;
  }
  size_t dot = name.find('.', pre.size());
  if (dot == std::string::npos) {
    b2_set_error("bad tensor name %s", name.c_str());
    return -22;
  }
  int l = atoi(name.substr(pre.size(), dot - pre.size()).c_str());
  if (l < 0 || l >= c.num_layers) {
    b2_set_error("tensor %s: layer out of range", name.c_str());
    return -22;
  }
  LayerW& L = e->layers[l];
  std::string rest = name.substr(dot + 1);
  if (rest == "input_layernorm.weight") return put(L.ln1, H);
  if (rest == "post_attention_layernorm.weight") return put(L.ln2, H);
  bf16* wq = L.wqkv;
  bf16* wk = L.wqkv + (size_t)e->q_size * H;
  bf16* wv = L.wqkv + (size_t)(e->q_size + e->kv_size) * H;
  auto done = [&](int rc) {
    if (rc == 0) e->loaded[name] = 1;
    return rc;
  };
  if (c.arch == B2_ARCH_QWEN3) {
    if (rest == "self_attn.q_norm.weight") return put(L.qnorm, c.head_dim);
    if (rest == "self_attn.k_norm.weight") return put(L.knorm, c.head_dim);
    if (rest == "mlp.gate_proj.weight") return put(L.wgu, (size_t)I * H);
    if (rest == "mlp.up_proj.weight") return put(L.wgu + (size_t)I * H, (size_t)I * H);
    if (e->native_geom) {
      if (rest == "self_attn.q_proj.weight") return put(wq, (size_t)e->q_size * H);
      if (rest == "self_attn.k_proj.weight") return put(wk, (size_t)e->kv_size * H);
      if (rest == "self_attn.v_proj.weight") return put(wv, (size_t)e->kv_size * H);
      if (rest == "self_attn.o_proj.weight") return put(L.wo, (size_t)H * e->q_size);
    } else {
      if (rest == "self_attn.q_proj.weight") { B2_TRY(expect(q_r * H)); return done(put_head_rows(wq, src16, c.num_q_heads, e->qhead_step, e->nq)); }
      if (rest == "self_attn.k_proj.weight") { B2_TRY(expect(kv_r * H)); return done(put_head_rows(wk, src16, c.num_kv_heads, 1, e->nkv)); }
      if (rest == "self_attn.v_proj.weight") { B2_TRY(expect(kv_r * H)); return done(put_head_rows(wv, src16, c.num_kv_heads, 1, e->nkv)); }
      if (rest == "self_attn.o_proj.weight") { B2_TRY(expect((size_t)H * q_r)); return done(put_head_cols(L.wo, src16, c.num_q_heads, e->qhead_step, e->nq)); }
    }
  } else {  // Phi-3: fused projections in the checkpoint, HF:models/phi3/modeling_phi3.py:53-54,224,237-241
    if (rest == "self_attn.qkv_proj.weight") {
      B2_TRY(expect((q_r + 2 * kv_r) * H));
      B2_TRY(put_head_rows(wq, src16, c.num_q_heads, e->qhead_step, e->nq));
      B2_TRY(put_head_rows(wk, src16 + q_r * H, c.num_kv_heads, 1, e->nkv));
      return done(put_head_rows(wv, src16 + (q_r + kv_r) * H, c.num_kv_heads, 1, e->nkv));
    }
    if (rest == "self_attn.o_proj.weight") { B2_TRY(expect((size_t)H * q_r)); return done(put_head_cols(L.wo, src16, c.num_q_heads, e->qhead_step, e->nq)); }
    if (rest == "mlp.gate_up_proj.weight") return put(L.wgu, (size_t)2 * I * H);  // [gate; up], the order chunk(2) splits
  }
  if (rest == "mlp.down_proj.weight") return put(L.wd, (size_t)H * I);
  b2_set_error("unknown tensor name %s", name.c_str());
  return -22;
}

int load_safetensors(b2_engine* e, const char* path) {
  int fd = open(path, O_RDONLY);
  if (fd < 0) {
    b2_set_error("cannot open %s: %s", path, strerror(errno));
    return -2;
  }
  struct stat st;
  fstat(fd, &st);
  void* map = mmap(nullptr, st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (map == MAP_FAILED) {
    b2_set_error("mmap %s failed", path);
    return -5;
  }
  int rc = 0;
  try {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(map);
    uint64_t hlen = 0;
    memcpy(&hlen, p, 8);
    if (hlen + 8 > (uint64_t)st.st_size) throw std::runtime_error("header length exceeds file");
    b2json::Value hdr = b2json::parse(reinterpret_cast<const char*>(p + 8), (size_t)hlen);
    const uint8_t* data = p + 8 + hlen;
    std::vector<uint16_t> conv;
    for (auto& kv : *hdr.obj) {
      if (kv.first == "__metadata__") continue;
      const b2json::Value* dt = kv.second.find("dtype");
      const b2json::Value* off = kv.second.find("data_offsets");
      if (!dt || !off || off->size() != 2) throw std::runtime_error("malformed entry " + kv.first);
      uint64_t a = (uint64_t)off->at(0).as_int(), b = (uint64_t)off->at(1).as_int();
      if (b < a || 8 + hlen + b > (uint64_t)st.st_size) throw std::runtime_error("offsets out of range for " + kv.first);
      const void* src = data + a;
      size_t nbytes = b - a;
      if (dt->as_string() == "F32") {  // convert to bf16 (round to nearest even)
        size_t n = nbytes / 4;
        conv.resize(n);
        const uint32_t* f = reinterpret_cast<const uint32_t*>(src);
        for (size_t i = 0; i < n; ++i) {
          uint32_t u = f[i];
          conv[i] = (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
        }
        src = conv.data();
        nbytes = n * 2;
      } else if (dt->as_string() == "F16") {  // fp16 -> fp32 (exact) -> bf16 (round to nearest even)
        size_t n = nbytes / 2;
        conv.resize(n);
        const uint16_t* hsrc = reinterpret_cast<const uint16_t*>(src);
        for (size_t i = 0; i < n; ++i) {
          const uint32_t hbits = hsrc[i], sign = (hbits & 0x8000u) << 16, ex = (hbits >> 10) & 0x1fu, man = hbits & 0x3ffu;
          uint32_t u;
          if (ex == 0) {
            if (man == 0) u = sign;
            else {  // subnormal half: normalise
              int sh = 0;
              uint32_t m = man;
              while (!(m & 0x400u)) { m <<= 1; ++sh; }
              u = sign | ((uint32_t)(127 - 15 - sh + 1) << 23) | ((m & 0x3ffu) << 13);
            }
          } else if (ex == 31) u = sign | 0x7f800000u | (man << 13);
          else u = sign | ((ex + 112u) << 23) | (man << 13);
          conv[i] = (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
        }
        src = conv.data();
        nbytes = n * 2;
      } else if (dt->as_string() != "BF16") {
        throw std::runtime_error("tensor " + kv.first + ": dtype " + dt->as_string() + " not supported (BF16/F16/F32)");
      }
      if (kv.first.find("rotary_emb.inv_freq") != std::string::npos) continue;
      rc = layer_tensor(e, kv.first, src, nbytes);
      if (rc != 0) break;
    }
  } catch (const std::exception& ex) {
    b2_set_error("safetensors %s: %s", path, ex.what());
    rc = -22;
  }
  munmap(map, st.st_size);
  return rc;
}

}  // namespace

// =======================================================================================
// C ABI
// =======================================================================================
extern "C" int b2_engine_create(const b2_config* cfg, const char* safetensors_path, int device, b2_engine** out) {
  B2_REQUIRE(cfg != nullptr && out != nullptr, "b2_engine_create: null argument");
  B2_REQUIRE(cfg->abi_version == B2_ABI_VERSION, "b2_engine_create: abi_version %d != %d", cfg->abi_version, B2_ABI_VERSION);
  B2_REQUIRE(cfg->arch == B2_ARCH_QWEN3 || cfg->arch == B2_ARCH_PHI3 || cfg->arch == B2_ARCH_OPT,
             "arch %d not built (0 = Qwen3, 1 = Phi-3, 2 = OPT)", cfg->arch);
  B2_REQUIRE(cfg->arch != B2_ARCH_OPT || cfg->tie_embeddings, "OPT: the lm_head is tied to embed_tokens");
  B2_REQUIRE(cfg->sliding_window >= 0 && (cfg->sliding_window == 0 || cfg->arch == B2_ARCH_PHI3),
             "sliding_window %d: only the Phi-3 family has one", cfg->sliding_window);
  B2_REQUIRE(cfg->head_dim >= 8 && cfg->head_dim <= 128 && cfg->head_dim % 8 == 0, "head_dim %d not supported (multiple of 8, <= 128)",
             cfg->head_dim);
  B2_REQUIRE(cfg->num_kv_heads >= 1 && (cfg->num_q_heads == 2 * cfg->num_kv_heads || cfg->num_q_heads == cfg->num_kv_heads),
             "query heads per kv head must be 1 or 2 (nq=%d nkv=%d)", cfg->num_q_heads, cfg->num_kv_heads);
  {
    const int rd = cfg->arch == B2_ARCH_OPT ? 0 : (cfg->rotary_dim > 0 ? cfg->rotary_dim : cfg->head_dim);
    B2_REQUIRE(rd % 8 == 0 && rd <= cfg->head_dim, "rotary_dim %d must be a multiple of 8 and <= head_dim", rd);
    B2_REQUIRE(cfg->arch != B2_ARCH_QWEN3 || (rd == cfg->head_dim && cfg->head_dim == 128),
               "Qwen3: QK-norm is built for 128-wide, fully rotated heads (head_dim %d, rotary_dim %d)", cfg->head_dim, rd);
  }
  B2_REQUIRE(cfg->block_size == 16, "block_size %d not built (16 only)", cfg->block_size);
  B2_REQUIRE(cfg->hidden % 256 == 0 && cfg->intermediate % 256 == 0, "hidden/intermediate must be multiples of 256");
  B2_REQUIRE(cfg->max_seqs >= 1 && cfg->max_batch_tokens >= 1 && cfg->max_model_len >= 1, "bad sizing");
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev <= device) {
    b2_set_error("no CUDA device %d available (%s); this engine has no CPU path", device,
                 ce == cudaSuccess ? "device count too small" : cudaGetErrorString(ce));
    return -19;
  }
  B2_CUDA_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  B2_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  B2_REQUIRE(prop.major == 10, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);

  b2_engine* e = new b2_engine();
  e->cfg = *cfg;
  e->cfg.rope_inv_freq = nullptr;
  e->device = device;
  e->num_sms = prop.multiProcessorCount;
  const b2_config& c = e->cfg;
  e->hd = 128;
  e->nkv = c.num_kv_heads;
  e->nq = 2 * c.num_kv_heads;
  e->qhead_step = (c.num_q_heads == c.num_kv_heads) ? 2 : 1;
  e->native_geom = (c.head_dim == 128 && e->qhead_step == 1);
  e->attn_scale = (float)(1.0 / sqrt((double)c.head_dim));  // HF passes head_dim**-0.5 as a double
  e->rotary_dim = c.arch == B2_ARCH_OPT ? 0 : (c.rotary_dim > 0 ? c.rotary_dim : c.head_dim);
  e->gu_rows = c.arch == B2_ARCH_OPT ? c.intermediate : 2 * c.intermediate;
  e->pos_rows = c.arch == B2_ARCH_OPT ? (c.max_positions > 0 ? c.max_positions : c.max_model_len) + 2 : 0;
  e->q_size = e->nq * e->hd;
  e->kv_size = e->nkv * e->hd;
  e->qkv_size = e->q_size + 2 * e->kv_size;
  e->simt_max_tokens = env_int("B2_SIMT_MAX_TOKENS", 0);
  e->attn_split_override = env_int("B2_ATTN_SPLITS", 0);
  e->force_gemm_impl = env_int("B2_GEMM_IMPL", -1);
  e->skip_mask = (unsigned)env_int("B2_SKIP_MASK", 0);
  // 0 separate kernels (default: fastest measured), 1 eight-step chain kernel (chain.cu), 2 fused four-step cluster
  // chain (chain_cluster.cu); 1 and 2 are kept as measured experiments with parity tests
  e->use_chain = env_int("B2_CHAIN", 0);
  e->use_pair = env_int("B2_PAIR", 7);  // bit 2: O / down projections update the residual stream themselves (gemm_resid.cu)
  e->prefill_tc = env_int("B2_PREFILL_TC", 1);
  e->gu_prefill_fused = env_int("B2_GU_FUSED", 1);
  e->prefill_mc = env_int("B2_MC", 1);
  e->prefill_pairs = env_int("B2_PREFILL_PAIRS", 1);
  e->qkv_rope_prefill_fused = env_int("B2_QKV_ROPE_FUSED", 1);
  e->prefill_resid_fused = env_int("B2_RESID_FUSED", 0);
  e->l2_weight_prefetch = env_int("B2_WPF", 0);
  e->o_small_smem = env_int("B2_O_SMALL", 0);
  // Off by default: measured on B200 at C2 (ms/step) 0 MB 2.22, 40 MB 2.24, 72 MB 2.33, 100 MB 2.38 --
  // un-paced prefetch bursts queue ahead of the projections' own latency-critical weight loads.
  e->kv_l2_budget = (long long)env_int("B2_KV_L2_MB", 0) << 20;
  const int H = c.hidden, I = c.intermediate;
  int rc = 0;
#define E_TRY(expr)        \
  do {                     \
    rc = (expr);           \
    if (rc != 0) goto fail; \
  } while (0)
#define E_CUDA(expr)                                                                           \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      b2_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      rc = -5;                                                                                 \
      goto fail;                                                                               \
    }                                                                                          \
  } while (0)
  {
    E_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    // ---- weights ----
    E_TRY(dalloc(&e->embed, (size_t)c.vocab * H));
    if (c.tie_embeddings) e->lm_head = e->embed;
    else E_TRY(dalloc(&e->lm_head, (size_t)c.vocab * H));
    E_TRY(dalloc(&e->final_norm, (size_t)H));
    e->layers.resize(c.num_layers);
    for (auto& L : e->layers) {
      E_TRY(dalloc(&L.ln1, (size_t)H));
      E_TRY(dalloc(&L.ln2, (size_t)H));
      if (c.arch == B2_ARCH_QWEN3) {  // Phi-3 has no QK-norm: the null gains switch it off in the kernel
        E_TRY(dalloc(&L.qnorm, (size_t)c.head_dim));
        E_TRY(dalloc(&L.knorm, (size_t)c.head_dim));
      }
      E_TRY(dalloc(&L.wqkv, (size_t)e->qkv_size * H));
      E_TRY(dalloc(&L.wo, (size_t)H * e->q_size));
      E_TRY(dalloc(&L.wgu, (size_t)e->gu_rows * H));
      E_TRY(dalloc(&L.wd, (size_t)H * I));
      if (c.arch == B2_ARCH_OPT) {
        E_TRY(dalloc(&L.bqkv, (size_t)e->qkv_size));
        E_CUDA(cudaMemset(L.bqkv, 0, (size_t)e->qkv_size * 2));  // padded slots keep a zero bias
        E_TRY(dalloc(&L.bo, (size_t)H));
        E_TRY(dalloc(&L.ln1_b, (size_t)H));
        E_TRY(dalloc(&L.ln2_b, (size_t)H));
        E_TRY(dalloc(&L.b1, (size_t)I));
        E_TRY(dalloc(&L.b2, (size_t)H));
      }
    }
    if (c.arch == B2_ARCH_OPT) {
      E_TRY(dalloc(&e->final_norm_b, (size_t)H));
      E_TRY(dalloc(&e->pos_embed, (size_t)e->pos_rows * H));
    }
    e->weight_bytes = 2ll * ((int64_t)c.vocab * H * (c.tie_embeddings ? 1 : 2) + H +
                             (int64_t)c.num_layers * (2 * H + (c.arch == B2_ARCH_QWEN3 ? 2 * c.head_dim : 0) + (int64_t)e->qkv_size * H +
                                                      (int64_t)H * e->q_size + ((int64_t)e->gu_rows + I) * H) +
                             (c.arch == B2_ARCH_OPT ? (int64_t)e->pos_rows * H + H +
                                                          (int64_t)c.num_layers * (e->qkv_size + 4ll * H + I) : 0));
    // ---- RoPE table: cos/sin in fp32 from fp32 inv_freq * position, rounded to bf16
    //      (HF:models/qwen3/modeling_qwen3.py:132-148) ----
    {
      // table row = [cos 64 | sin 64]; a rotary width below 128 (Phi-3: 96) fills the first rotary_dim/2
      // entries of each half, HF:models/phi3/modeling_phi3.py:199-205
      const int rd = e->rotary_dim, half = rd / 2, row = 128;
      std::vector<float> inv(half);
      for (int i = 0; i < half; ++i) {
        if (cfg->rope_inv_freq) inv[i] = cfg->rope_inv_freq[i];
        else inv[i] = (float)(1.0 / pow((double)c.rope_theta, (double)(2 * i) / (double)rd));
      }
      std::vector<uint16_t> tab((size_t)c.max_model_len * row);
      auto to_bf16 = [](float f) -> uint16_t {
        uint32_t u;
        memcpy(&u, &f, 4);
        return (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
      };
      for (int p = 0; p < c.max_model_len; ++p) {
        for (int i = 0; i < 64; ++i) {
          tab[(size_t)p * row + i] = to_bf16(1.0f);
          tab[(size_t)p * row + 64 + i] = to_bf16(0.0f);
        }
        for (int i = 0; i < half; ++i) {
          float fr = (float)p * inv[i];  // fp32 product, as the fp32 matmul in HF
          tab[(size_t)p * row + i] = to_bf16((float)cos((double)fr));
          tab[(size_t)p * row + 64 + i] = to_bf16((float)sin((double)fr));
        }
      }
      E_TRY(dalloc(&e->cos_sin, tab.size()));
      E_CUDA(cudaMemcpy(e->cos_sin, tab.data(), tab.size() * 2, cudaMemcpyHostToDevice));
    }
    // ---- activations ----
    e->cap_tokens = ((std::max(c.max_batch_tokens, c.max_seqs) + 255) / 256) * 256;
    const size_t Tc = e->cap_tokens;
    E_TRY(dalloc(&e->h, Tc * H));
    E_TRY(dalloc(&e->x, Tc * H));
    E_TRY(dalloc(&e->qkv, Tc * e->qkv_size));
    E_TRY(dalloc(&e->qrot, Tc * e->q_size));
    E_TRY(dalloc(&e->attn, Tc * e->q_size));
    E_TRY(dalloc(&e->gu, Tc * 2 * I));
    E_TRY(dalloc(&e->act, Tc * I));
    E_TRY(dalloc(&e->proj, Tc * H));
    const size_t last_rows = ((size_t)(c.max_seqs + 255) / 256) * 256;
    E_TRY(dalloc(&e->xlast, last_rows * H));
    E_CUDA(cudaMemset(e->x, 0, Tc * H * 2));
    E_CUDA(cudaMemset(e->attn, 0, Tc * e->q_size * 2));
    E_CUDA(cudaMemset(e->act, 0, Tc * I * 2));
    E_CUDA(cudaMemset(e->xlast, 0, last_rows * H * 2));
    e->partial_floats = (size_t)16 << 20;  // 64 MB of fp32 split-K partials
    if (c.arch == B2_ARCH_OPT)  // every projection hands over fp32 sums (the bias goes in before the rounding), even unsplit
      e->partial_floats = std::max(e->partial_floats, Tc * (size_t)std::max(std::max(e->qkv_size, I), H));
    E_TRY(dalloc(&e->partial, e->partial_floats));
    e->attn_scratch_floats = attn_decode_scratch_floats(bucket_batch(c.max_seqs), e->nq, e->nkv, e->hd);
    E_TRY(dalloc(&e->attn_scratch, e->attn_scratch_floats));
    E_CUDA(cudaMemset(e->attn_scratch, 0, e->attn_scratch_floats * sizeof(float)));  // arrival counters start at zero
    // two-lane decode: B2_MB = smallest padded decode batch that is cut into two half-batch lanes (0 = off)
    // Off by default: measured slower (C2: 2.03 against 1.90 ms/step) -- DESIGN.md section 4 has the timeline.
    e->mb_min_batch = c.arch == B2_ARCH_OPT ? 0 : env_int("B2_MB", 0);
    if (c.arch != B2_ARCH_OPT && c.max_seqs >= 2) {
      E_CUDA(cudaStreamCreateWithFlags(&e->stream2, cudaStreamNonBlocking));
      E_CUDA(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
      E_CUDA(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
      // a lane has at most 64 rows: 8 splits x 64 x the widest projection
      e->partial2_floats = (size_t)8 * 64 * (size_t)std::max(std::max(e->qkv_size, 2 * I), H);
      E_TRY(dalloc(&e->partial2, e->partial2_floats));
      E_TRY(dalloc(&e->attn_scratch2, e->attn_scratch_floats));
      E_CUDA(cudaMemset(e->attn_scratch2, 0, e->attn_scratch_floats * sizeof(float)));
    }
    E_TRY(dalloc(&e->ssq, (size_t)(H / 16) * 64));
    for (int bi = 0; bi < 5; ++bi) {
      E_TRY(b2_make_tmap_bf16(&e->tm_x[XB_X][bi], e->x, Tc, H, H, kBnList[bi]));
      E_TRY(b2_make_tmap_bf16(&e->tm_x[XB_ATTN][bi], e->attn, Tc, e->q_size, e->q_size, kBnList[bi]));
      E_TRY(b2_make_tmap_bf16(&e->tm_x[XB_ACT][bi], e->act, Tc, I, I, kBnList[bi]));
      E_TRY(b2_make_tmap_bf16(&e->tm_x[XB_LAST][bi], e->xlast, last_rows, H, H, kBnList[bi]));
    }
    for (auto& L : e->layers) {
      E_TRY(b2_make_tmap_bf16(&L.tm_qkv, L.wqkv, e->qkv_size, H, H, 128));
      E_TRY(b2_make_tmap_bf16(&L.tm_o, L.wo, H, e->q_size, e->q_size, 128));
      E_TRY(b2_make_tmap_bf16(&L.tm_gu, L.wgu, e->gu_rows, H, H, 128));
      E_TRY(b2_make_tmap_bf16(&L.tm_o32, L.wo, H, e->q_size, e->q_size, 32));
      E_TRY(b2_make_tmap_bf16(&L.tm_gu96, L.wgu, e->gu_rows, H, H, 96));
      E_TRY(b2_make_tmap_bf16(&L.tm_gu64, L.wgu, e->gu_rows, H, H, 64));
      E_TRY(b2_make_tmap_bf16(&L.tm_d32, L.wd, H, I, I, 32));
      E_TRY(b2_make_tmap_bf16(&L.tm_d, L.wd, H, I, I, 128));
    }
    E_TRY(b2_make_tmap_bf16(&e->tm_lm_head, e->lm_head, c.vocab, H, H, 128));
    // ---- batch state ----
    e->bt_stride = (c.max_model_len + c.block_size - 1) / c.block_size;
    const size_t Bcap = bucket_batch(c.max_seqs);
    E_TRY(dalloc(&e->d_tokens, Tc));
    E_TRY(dalloc(&e->d_positions, Tc));
    E_TRY(dalloc(&e->d_slots, Tc));
    E_TRY(dalloc(&e->d_seq_lens, Bcap));
    E_TRY(dalloc(&e->d_block_table, Bcap * e->bt_stride));
    E_TRY(dalloc(&e->d_last_rows, Bcap));
    E_TRY(dalloc(&e->d_out_tokens, Bcap));
    E_TRY(dalloc(&e->d_argmax, Bcap));
    E_TRY(dalloc(&e->d_step, (size_t)1));
    E_TRY(dalloc(&e->d_pf_frac, (size_t)1));
    E_TRY(dalloc(&e->d_sample_out, Bcap));
    E_TRY(dalloc(&e->d_sample_rows, Bcap));
    E_TRY(dalloc(&e->d_gbar, (size_t)1));
    E_CUDA(cudaMemset(e->d_gbar, 0, 8));
    E_TRY(dalloc(&e->d_gbar2, (size_t)1));
    E_CUDA(cudaMemset(e->d_gbar2, 0, 8));
    // the fused chain spins on a grid-wide barrier: all of its 32 clusters must be resident at once
    e->fused_ok = e->native_geom && c.arch == B2_ARCH_QWEN3 &&
                  fused_chain_supported(H, I, e->nq, e->nkv, e->hd, 1, c.block_size) && fused_chain_max_clusters() >= 32;
    E_TRY(dalloc(&e->d_ss, (size_t)2 * 32 * 64));
    E_CUDA(cudaMemset(e->d_ss, 0, 2 * 32 * 64 * 4));
    E_CUDA(cudaMemset(e->d_pf_frac, 0, 4));
    e->log_cap_steps = 4096;
    E_TRY(dalloc(&e->d_token_log, (size_t)e->log_cap_steps * c.max_seqs));
    E_CUDA(cudaMemset(e->d_step, 0, 4));
    E_CUDA(cudaMemset(e->d_block_table, 0, Bcap * e->bt_stride * 4));
    e->tiles_cap = (int)(Tc / 64 + 2 * Bcap + 16);
    E_TRY(dalloc(&e->d_tiles, (size_t)e->tiles_cap));
    e->h_stage_ints = std::max((size_t)3 * Tc + Bcap * (size_t)e->bt_stride + 4 * Bcap, (size_t)1 << 16);
    E_CUDA(cudaMallocHost(&e->h_stage, e->h_stage_ints * sizeof(int32_t)));
    E_CUDA(cudaMallocHost(&e->h_tiles, (size_t)e->tiles_cap * sizeof(PrefillTile)));
    E_CUDA(cudaMallocHost(&e->h_out, (size_t)e->log_cap_steps * c.max_seqs * sizeof(int32_t)));
    E_CUDA(cudaMallocHost(&e->h_tok, Bcap * sizeof(int32_t)));
    E_CUDA(cudaEventCreateWithFlags(&e->ev_tok, cudaEventDisableTiming));
    e->seqs.resize(c.max_seqs);
    // ---- KV pool ----
    {
      const size_t blk_elems = (size_t)e->nkv * c.block_size * e->hd;
      int64_t nb = c.num_blocks;
      if (nb <= 0) {
        size_t free_b = 0, total_b = 0;
        E_CUDA(cudaMemGetInfo(&free_b, &total_b));
        const double frac = c.gpu_mem_fraction > 0 ? c.gpu_mem_fraction : 0.9;
        const double usable = (double)free_b - (1.0 - frac) * (double)total_b - (512.0 * 1024 * 1024);
        nb = (int64_t)(usable / ((double)blk_elems * 2 * 2 * c.num_layers));
        if (nb < 16) {
          b2_set_error("not enough free device memory for a KV pool (free %zu MiB)", free_b >> 20);
          rc = -12;
          goto fail;
        }
      }
      e->num_blocks = nb;
      e->layer_kv_elems = (size_t)nb * blk_elems;
      E_TRY(dalloc(&e->kv_pool, (size_t)2 * c.num_layers * e->layer_kv_elems));
      E_CUDA(cudaMemset(e->kv_pool, 0, (size_t)2 * c.num_layers * e->layer_kv_elems * 2));
      e->free_blocks.reserve(nb);
      for (int64_t b = nb - 1; b >= 1; --b) e->free_blocks.push_back((int32_t)b);  // page 0 = null page
      e->blk_ref.assign(nb, 0);
      e->blk_hash.assign(nb, 0);
      e->blk_parent.assign(nb, -1);
      e->blk_tokens.assign((size_t)nb * c.block_size, 0);
      e->pc_lru_pos.assign(nb, e->pc_lru.end());
      e->pc_in_lru.assign(nb, 0);
      e->prefix_cache = env_int("B2_PREFIX_CACHE", 0) != 0;
    }
    if (e->force_gemm_impl == 0 || e->simt_max_tokens > 0) E_TRY(ensure_logits(e));  // SIMT lm_head writes logits
    e->ev_pool.resize(2 * 400);
    for (auto& ev : e->ev_pool) E_CUDA(cudaEventCreate(&ev));
    if (safetensors_path) {
      E_TRY(load_safetensors(e, safetensors_path));
      E_TRY(b2_engine_finalize(e));
    }
  }
  *out = e;
  return 0;
fail:
  b2_engine_free(e);
  return rc;
#undef E_TRY
#undef E_CUDA
}

extern "C" int b2_engine_load_tensor(b2_engine* e, const char* hf_name, const void* host_bf16, size_t nbytes) {
  B2_REQUIRE(e && hf_name && host_bf16, "b2_engine_load_tensor: null argument");
  B2_CUDA_OK(cudaSetDevice(e->device));
  return layer_tensor(e, hf_name, host_bf16, nbytes);
}

extern "C" int b2_engine_load_safetensors(b2_engine* e, const char* path) {
  B2_REQUIRE(e && path, "b2_engine_load_safetensors: null argument");
  B2_CUDA_OK(cudaSetDevice(e->device));
  return load_safetensors(e, path);
}

extern "C" int b2_engine_finalize(b2_engine* e) {
  B2_REQUIRE(e, "b2_engine_finalize: null engine");
  const int arch = e->cfg.arch;
  const int per_layer = arch == B2_ARCH_QWEN3 ? 11 : arch == B2_ARCH_PHI3 ? 6 : 16;
  const int want = (arch == B2_ARCH_OPT ? 4 : 2) + e->cfg.num_layers * per_layer;
  int have = 0;
  for (auto& kv : e->loaded)
    if (kv.first != "lm_head.weight") ++have;
  if (!e->cfg.tie_embeddings) {
    B2_REQUIRE(e->loaded.count("lm_head.weight"), "weights incomplete: lm_head.weight missing (untied head)");
  }
  B2_REQUIRE(have == want, "weights incomplete: %d of %d tensors loaded", have, want);
  e->finalized = true;
  return 0;
}

extern "C" void b2_engine_free(b2_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (auto& kv : e->graphs) cudaGraphExecDestroy(kv.second);
  for (auto& ev : e->ev_pool)
    if (ev) cudaEventDestroy(ev);
  auto F = [](void* p) {
    if (p) cudaFree(p);
  };
  if (e->lm_head != e->embed) F(e->lm_head);
  F(e->embed); F(e->final_norm); F(e->cos_sin); F(e->final_norm_b); F(e->pos_embed);
  for (auto& L : e->layers) {
    F(L.ln1); F(L.ln2); F(L.qnorm); F(L.knorm); F(L.wqkv); F(L.wo); F(L.wgu); F(L.wd);
    F(L.bqkv); F(L.bo); F(L.ln1_b); F(L.ln2_b); F(L.b1); F(L.b2);
  }
  F(e->kv_pool); F(e->h); F(e->x); F(e->qkv); F(e->qrot); F(e->attn); F(e->gu); F(e->act); F(e->proj); F(e->xlast);
  F(e->partial); F(e->attn_scratch); F(e->d_tokens); F(e->d_positions); F(e->d_seq_lens); F(e->d_slots);
  F(e->d_block_table); F(e->d_last_rows); F(e->d_out_tokens); F(e->d_token_log); F(e->d_step); F(e->d_argmax); F(e->d_pf_frac); F(e->d_sample_out); F(e->d_sample_rows); F(e->d_gbar); F(e->d_gbar2); F(e->d_ss);
  F(e->d_tiles); F(e->dbg_hidden); F(e->logits);
  if (e->h_stage) cudaFreeHost(e->h_stage);
  if (e->h_tiles) cudaFreeHost(e->h_tiles);
  if (e->h_out) cudaFreeHost(e->h_out);
  if (e->h_tok) cudaFreeHost(e->h_tok);
  if (e->ev_tok) cudaEventDestroy(e->ev_tok);
  F(e->partial2); F(e->attn_scratch2); F(e->ssq);
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  if (e->stream2) cudaStreamDestroy(e->stream2);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

extern "C" int b2_engine_info(b2_engine* e, int64_t* num_blocks, int64_t* free_blocks, int64_t* weight_bytes) {
  B2_REQUIRE(e, "b2_engine_info: null engine");
  if (num_blocks) *num_blocks = e->num_blocks;
  if (free_blocks) *free_blocks = pages_available(e);  // cached pages nobody holds are reusable at once
  if (weight_bytes) *weight_bytes = e->weight_bytes;
  return 0;
}

extern "C" int b2_seq_create(b2_engine* e, int32_t reserve_tokens, int32_t* seq_id) {
  B2_REQUIRE(e && seq_id, "b2_seq_create: null argument");
  B2_REQUIRE(reserve_tokens >= 1 && reserve_tokens <= e->cfg.max_model_len, "reserve_tokens %d outside [1, max_model_len=%d]",
             reserve_tokens, e->cfg.max_model_len);
  const int bs = e->cfg.block_size;
  const int need = (reserve_tokens + bs - 1) / bs;
  if (pages_available(e) < need) {
    b2_set_error("KV pool exhausted: need %d pages, %lld free", need, (long long)pages_available(e));
    return -12;
  }
  int id = -1;
  for (int i = 0; i < (int)e->seqs.size(); ++i)
    if (!e->seqs[i].active) {
      id = i;
      break;
    }
  if (id < 0) {
    b2_set_error("all %d sequence slots in use", e->cfg.max_seqs);
    return -11;
  }
  Seq& s = e->seqs[id];
  s.active = true;
  s.len = 0;
  s.reserved = need * bs;
  s.blocks.clear();
  for (int i = 0; i < need; ++i) s.blocks.push_back(take_page(e));
  s.chain.clear();
  s.pending.clear();
  s.chain_open = e->prefix_cache;
  *seq_id = id;
  return 0;
}

extern "C" int b2_seq_free(b2_engine* e, int32_t seq_id) {
  B2_REQUIRE(e, "b2_seq_free: null engine");
  B2_TRY(check_seq(e, seq_id));
  Seq& s = e->seqs[seq_id];
  // leading pages first, so that on the LRU list a prefix outlives its continuations
  for (auto it = s.blocks.rbegin(); it != s.blocks.rend(); ++it) release_page(e, *it);
  s.blocks.clear();
  s.chain.clear();
  s.pending.clear();
  s.chain_open = false;
  s.active = false;
  s.len = 0;
  e->cur_batch.clear();
  return 0;
}

extern "C" int b2_seq_len(b2_engine* e, int32_t seq_id) {
  if (!e || check_seq(e, seq_id) != 0) return -22;
  return e->seqs[seq_id].len;
}

extern "C" int b2_seq_truncate(b2_engine* e, int32_t seq_id, int32_t new_len) {
  B2_REQUIRE(e, "b2_seq_truncate: null engine");
  B2_TRY(check_seq(e, seq_id));
  Seq& s = e->seqs[seq_id];
  B2_REQUIRE(new_len >= 0 && new_len <= s.len, "b2_seq_truncate: new_len %d outside [0, %d]", new_len, s.len);
  // Positions >= new_len will be written again.  Pages from there on that are registered in the prefix cache or held by
  // other sequences too must not be: the sequence swaps them for private pages (copying the kept part of the page
  // new_len falls into), and its own chain ends at the cut.
  const int bs = e->cfg.block_size;
  const int first = new_len / bs;
  int need = 0;
  for (int j = first; j < (int)s.blocks.size(); ++j)
    if (e->blk_hash[s.blocks[j]] != 0 || e->blk_ref[s.blocks[j]] > 1) ++need;
  if (need > 0) {
    B2_REQUIRE(pages_available(e) >= need, "b2_seq_truncate: no free pages to un-share %d cached pages", need);
    B2_CUDA_OK(cudaSetDevice(e->device));
    B2_CUDA_OK(cudaStreamSynchronize(e->stream));
    const size_t blk_elems = (size_t)e->nkv * bs * e->hd;
    for (int j = first; j < (int)s.blocks.size(); ++j) {
      const int32_t old_b = s.blocks[j];
      if (e->blk_hash[old_b] == 0 && e->blk_ref[old_b] <= 1) continue;
      const int32_t nb = take_page(e);
      if (j == first && new_len % bs != 0) {
        for (int l = 0; l < 2 * e->cfg.num_layers; ++l) {
          bf16* base = e->kv_pool + (size_t)l * e->layer_kv_elems;
          B2_CUDA_OK(cudaMemcpyAsync(base + (size_t)nb * blk_elems, base + (size_t)old_b * blk_elems, blk_elems * 2, cudaMemcpyDeviceToDevice,
                                     e->stream));
        }
      }
      s.blocks[j] = nb;
      release_page(e, old_b);
    }
    B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  }
  if ((int)s.chain.size() > first) s.chain.resize(first);
  s.pending.clear();
  s.chain_open = false;
  s.len = new_len;
  e->cur_batch.clear();
  return 0;
}

extern "C" int b2_prefill(b2_engine* e, const int32_t* ids, const int32_t* seq_offsets, int32_t nseq, const int32_t* seq_ids,
                          int32_t* out_tokens) {
  B2_REQUIRE(e && ids && seq_offsets && seq_ids && nseq >= 1, "b2_prefill: bad arguments");
  B2_REQUIRE(e->finalized, "b2_prefill: weights not finalized");
  B2_REQUIRE(nseq <= e->cfg.max_seqs, "b2_prefill: %d sequences > max_seqs %d", nseq, e->cfg.max_seqs);
  B2_CUDA_OK(cudaSetDevice(e->device));
  const int T_in = seq_offsets[nseq] - seq_offsets[0];
  B2_REQUIRE(T_in >= 1 && T_in <= e->cap_tokens && T_in <= e->cfg.max_batch_tokens, "b2_prefill: %d tokens exceed max_batch_tokens %d", T_in,
             e->cfg.max_batch_tokens);
  const int bs = e->cfg.block_size;
  // ---- validate, then let fresh sequences take what the prefix cache already holds of their prompt ----
  std::vector<int> skip(nseq, 0);
  for (int i = 0; i < nseq; ++i) {
    B2_TRY(check_seq(e, seq_ids[i]));
    Seq& s = e->seqs[seq_ids[i]];
    const int n = seq_offsets[i + 1] - seq_offsets[i];
    B2_REQUIRE(n >= 1, "b2_prefill: sequence %d has no tokens", i);
    B2_REQUIRE(s.len + n <= s.reserved, "b2_prefill: sequence %d would exceed its reservation (%d + %d > %d)", seq_ids[i], s.len, n,
               s.reserved);
  }
  for (int i = 0; i < nseq; ++i) {
    Seq& s = e->seqs[seq_ids[i]];
    const int n = seq_offsets[i + 1] - seq_offsets[i];
    if (e->prefix_cache && s.len == 0 && s.chain_open && s.chain.empty()) {
      skip[i] = pc_match(e, s, ids + seq_offsets[i], n);
      s.len = skip[i];  // K/V of those positions are in the pages just taken
      e->pc_queries += n;
      e->pc_hits += skip[i];
    }
  }
  int T = 0;
  for (int i = 0; i < nseq; ++i) T += seq_offsets[i + 1] - seq_offsets[i] - skip[i];
  int32_t* tok = e->h_stage;
  int32_t* pos = tok + T;
  int32_t* slots = pos + T;
  int32_t* last = slots + T;
  int32_t* bt = last + nseq;
  int ntiles = 0, row0 = 0;
  for (int i = 0; i < nseq; ++i) {
    Seq& s = e->seqs[seq_ids[i]];
    const int n = seq_offsets[i + 1] - seq_offsets[i] - skip[i];
    const int32_t* src = ids + seq_offsets[i] + skip[i];
    for (int j = 0; j < n; ++j) {
      const int p = s.len + j;
      tok[row0 + j] = src[j];
      pos[row0 + j] = p;
      slots[row0 + j] = s.blocks[p / bs] * bs + p % bs;
    }
    last[i] = row0 + n - 1;
    memset(bt + (size_t)i * e->bt_stride, 0, (size_t)e->bt_stride * 4);
    memcpy(bt + (size_t)i * e->bt_stride, s.blocks.data(), s.blocks.size() * 4);
    const int first_tile = ntiles;
    for (int j = 0; j < n; j += 64) {
      B2_REQUIRE(ntiles < e->tiles_cap, "b2_prefill: too many attention tiles");
      e->h_tiles[ntiles++] = PrefillTile{row0 + j, s.len + j, std::min(64, n - j), i};
    }
    if (prefill_tile_pairs(e) && ((ntiles - first_tile) & 1)) {  // attn_prefill_tc.cu, pair mode: tiles 2i and 2i + 1 belong together
      B2_REQUIRE(ntiles < e->tiles_cap, "b2_prefill: too many attention tiles");
      e->h_tiles[ntiles++] = PrefillTile{row0, s.len + n, 0, i};
    }
    row0 += n;
  }
  B2_CUDA_OK(cudaMemcpyAsync(e->d_tokens, tok, T * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_positions, pos, T * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_slots, slots, T * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_last_rows, last, nseq * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_block_table, bt, (size_t)nseq * e->bt_stride * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaMemcpyAsync(e->d_tiles, e->h_tiles, ntiles * sizeof(PrefillTile), cudaMemcpyHostToDevice, e->stream));
  if (e->debug_flags & 2) B2_TRY(ensure_logits(e));
  e->cur_batch.clear();
  B2_TRY(forward(e, T, ntiles, nseq, 1));
  // the produced tokens become the device-resident "next input" of these sequences
  B2_CUDA_OK(cudaMemcpyAsync(e->h_out, e->d_out_tokens, nseq * 4, cudaMemcpyDeviceToHost, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  for (int i = 0; i < nseq; ++i) {
    Seq& s = e->seqs[seq_ids[i]];
    const int n = seq_offsets[i + 1] - seq_offsets[i] - skip[i];
    s.len += n;
    s.next_token = e->h_out[i];
    s.next_valid = true;
    if (e->prefix_cache) pc_register(e, s, ids + seq_offsets[i] + skip[i], n);  // the forward that filled the pages is complete
  }
  if (out_tokens) memcpy(out_tokens, e->h_out, nseq * 4);
  // remember them host-side too so the first decode step can seed d_tokens
  e->last_n = nseq;
  e->last_tokens = T;
  return 0;
}

namespace {
// Prepares device state for a decode step over seq_ids; returns padded batch B.
int prepare_decode(b2_engine* e, int n, const int32_t* seq_ids, const int32_t* in_tokens, int* B_out, int* split_out) {
  B2_REQUIRE(e->finalized, "decode: weights not finalized");
  B2_REQUIRE(n >= 1 && n <= e->cfg.max_seqs, "decode: batch %d outside [1, max_seqs=%d]", n, e->cfg.max_seqs);
  int max_len = 0;
  for (int i = 0; i < n; ++i) {
    B2_TRY(check_seq(e, seq_ids[i]));
    Seq& s = e->seqs[seq_ids[i]];
    B2_REQUIRE(s.len >= 1, "decode: sequence %d has no prefilled tokens", seq_ids[i]);
    B2_REQUIRE(s.len + 1 <= s.reserved, "decode: sequence %d would exceed its reservation (%d tokens)", seq_ids[i], s.reserved);
    s.chain_open = false;  // generated tokens are not registered in the prefix cache
    max_len = std::max(max_len, s.len + 1);
  }
  const int B = e->cfg.use_cuda_graphs ? bucket_batch(n) : n;
  const bool same = (int)e->cur_batch.size() == n && e->cur_B == B && std::equal(seq_ids, seq_ids + n, e->cur_batch.begin());
  if (!same) {
    // The batch composition changed: the device-resident "next tokens" belong to another batch
    // order, so seed this batch from the host copies kept per sequence (or from in_tokens).
    if (in_tokens == nullptr) {
      for (int i = 0; i < n; ++i) {
        Seq& s = e->seqs[seq_ids[i]];
        B2_REQUIRE(s.next_valid, "decode: batch changed and sequence %d has no host copy of its next token; pass in_tokens",
                   seq_ids[i]);
        e->h_stage[i] = s.next_token;
      }
      for (int i = n; i < B; ++i) e->h_stage[i] = 0;
      B2_CUDA_OK(cudaMemcpyAsync(e->d_tokens, e->h_stage, B * 4, cudaMemcpyHostToDevice, e->stream));
      B2_CUDA_OK(cudaStreamSynchronize(e->stream));
    }
    B2_TRY(upload_decode_batch(e, n, seq_ids, B));
  }
  if (in_tokens != nullptr) {
    // dedicated pinned slot for the per-step token upload; the event keeps a later call from
    // overwriting it while the copy is still in flight (callers may skip the D2H sync)
    B2_CUDA_OK(cudaEventSynchronize(e->ev_tok));
    memcpy(e->h_tok, in_tokens, n * 4);
    for (int i = n; i < B; ++i) e->h_tok[i] = 0;
    B2_CUDA_OK(cudaMemcpyAsync(e->d_tokens, e->h_tok, B * 4, cudaMemcpyHostToDevice, e->stream));
    B2_CUDA_OK(cudaEventRecord(e->ev_tok, e->stream));
  }
  *B_out = B;
  *split_out = pick_attn_split(e, B, max_len);
  return 0;
}
}  // namespace

namespace {
int fetch_next_tokens(b2_engine* e, int nseq, const int32_t* seq_ids) {
  int32_t* tmp = e->h_stage;
  B2_CUDA_OK(cudaMemcpyAsync(tmp, e->d_out_tokens, nseq * 4, cudaMemcpyDeviceToHost, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  for (int i = 0; i < nseq; ++i) {
    e->seqs[seq_ids[i]].next_token = tmp[i];
    e->seqs[seq_ids[i]].next_valid = true;
  }
  return 0;
}
}  // namespace

extern "C" int b2_decode_step(b2_engine* e, int32_t nseq, const int32_t* seq_ids, const int32_t* in_tokens, int32_t* out_tokens) {
  B2_REQUIRE(e && seq_ids, "b2_decode_step: null argument");
  B2_CUDA_OK(cudaSetDevice(e->device));
  int B = 0, split = 1;
  B2_TRY(prepare_decode(e, nseq, seq_ids, in_tokens, &B, &split));
  if (e->debug_flags & 2) B2_TRY(ensure_logits(e));
  B2_TRY(run_decode_step(e, B, split));
  for (int i = 0; i < nseq; ++i) e->seqs[seq_ids[i]].len += 1;
  e->last_n = nseq;
  e->last_tokens = nseq;
  if (out_tokens) {
    B2_CUDA_OK(cudaMemcpyAsync(e->h_out, e->d_out_tokens, nseq * 4, cudaMemcpyDeviceToHost, e->stream));
    B2_CUDA_OK(cudaStreamSynchronize(e->stream));
    memcpy(out_tokens, e->h_out, nseq * 4);
    for (int i = 0; i < nseq; ++i) {
      e->seqs[seq_ids[i]].next_token = e->h_out[i];
      e->seqs[seq_ids[i]].next_valid = true;
    }
  } else {
    for (int i = 0; i < nseq; ++i) e->seqs[seq_ids[i]].next_valid = false;
  }
  return 0;
}

extern "C" int b2_decode_run(b2_engine* e, int32_t nseq, const int32_t* seq_ids, int32_t steps, int32_t* out_tokens,
                             float* elapsed_ms) {
  B2_REQUIRE(e && seq_ids && steps >= 1, "b2_decode_run: bad arguments");
  B2_REQUIRE(steps <= e->log_cap_steps, "b2_decode_run: steps %d > %d", steps, e->log_cap_steps);
  B2_CUDA_OK(cudaSetDevice(e->device));
  for (int i = 0; i < nseq; ++i) {
    B2_TRY(check_seq(e, seq_ids[i]));
    Seq& s = e->seqs[seq_ids[i]];
    B2_REQUIRE(s.len + steps <= s.reserved, "b2_decode_run: sequence %d reservation too small for %d steps", seq_ids[i], steps);
  }
  int B = 0, split = 1;
  B2_TRY(prepare_decode(e, nseq, seq_ids, nullptr, &B, &split));
  B2_CUDA_OK(cudaMemsetAsync(e->d_step, 0, 4, e->stream));
  // the graph is captured before the timed region; the event pair brackets exactly `steps` steps
  cudaEvent_t ev0 = e->ev_pool[e->ev_pool.size() - 2], ev1 = e->ev_pool[e->ev_pool.size() - 1];
  int done = 0;
  {
    int max_len = 0;
    for (int i = 0; i < nseq; ++i) max_len = std::max(max_len, e->seqs[seq_ids[i]].len + steps);
    split = pick_attn_split(e, B, max_len);  // one graph for the whole run
  }
  if (graphs_enabled(e)) B2_TRY(ensure_graph(e, B, split));
  B2_CUDA_OK(cudaEventRecord(ev0, e->stream));
  for (; done < steps; ++done) B2_TRY(run_decode_step(e, B, split));
  B2_CUDA_OK(cudaEventRecord(ev1, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  if (elapsed_ms) B2_CUDA_OK(cudaEventElapsedTime(elapsed_ms, ev0, ev1));
  for (int i = 0; i < nseq; ++i) e->seqs[seq_ids[i]].len += steps;
  e->last_n = nseq;
  e->last_tokens = nseq;
  B2_TRY(fetch_next_tokens(e, nseq, seq_ids));
  if (out_tokens) {
    B2_CUDA_OK(cudaMemcpy(e->h_out, e->d_token_log, (size_t)steps * e->cfg.max_seqs * 4, cudaMemcpyDeviceToHost));
    for (int st = 0; st < steps; ++st)
      memcpy(out_tokens + (size_t)st * nseq, e->h_out + (size_t)st * e->cfg.max_seqs, nseq * 4);
  }
  return 0;
}

extern "C" int b2_decode_profile(b2_engine* e, int32_t nseq, const int32_t* seq_ids, int32_t steps, float* ms_by_class,
                                 int64_t* launches_by_class) {
  B2_REQUIRE(e && seq_ids && steps >= 1 && ms_by_class && launches_by_class, "b2_decode_profile: bad arguments");
  B2_CUDA_OK(cudaSetDevice(e->device));
  for (int k = 0; k < B2_NUM_KERNEL_CLASSES; ++k) {
    ms_by_class[k] = 0.f;
    launches_by_class[k] = 0;
  }
  for (int st = 0; st < steps; ++st) {
    int B = 0, split = 1;
    B2_TRY(prepare_decode(e, nseq, seq_ids, nullptr, &B, &split));
    e->profiling = true;
    e->ev_next = 0;
    e->ev_used.clear();
    int rc = run_decode_step(e, B, split);
    e->profiling = false;
    B2_TRY(rc);
    B2_CUDA_OK(cudaStreamSynchronize(e->stream));
    for (auto& u : e->ev_used) {
      float ms = 0.f;
      B2_CUDA_OK(cudaEventElapsedTime(&ms, e->ev_pool[u.second], e->ev_pool[u.second + 1]));
      ms_by_class[u.first] += ms;
      launches_by_class[u.first] += 1;
    }
    for (int i = 0; i < nseq; ++i) e->seqs[seq_ids[i]].len += 1;
  }
  e->last_n = nseq;
  e->last_tokens = nseq;
  B2_TRY(fetch_next_tokens(e, nseq, seq_ids));
  return 0;
}

extern "C" int b2_attn_profile(b2_engine* e, int32_t nseq, const int32_t* seq_ids, int32_t reps, float* ms_per_launch,
                               int64_t* kv_bytes_per_launch) {
  B2_REQUIRE(e && seq_ids && reps >= 1 && ms_per_launch && kv_bytes_per_launch, "b2_attn_profile: bad arguments");
  B2_CUDA_OK(cudaSetDevice(e->device));
  B2_REQUIRE((int)e->cur_batch.size() == nseq && std::equal(seq_ids, seq_ids + nseq, e->cur_batch.begin()),
             "b2_attn_profile: run a decode step on exactly this batch first (the pass reuses its device-side page tables)");
  const b2_config& c = e->cfg;
  int64_t bytes = 0;
  for (int i = 0; i < nseq; ++i) bytes += (int64_t)(e->seqs[seq_ids[i]].len + 1) * e->nkv * e->hd * 2 * 2;
  const int B = e->cur_B;
  auto pass = [&]() -> int {
    for (int l = 0; l < c.num_layers; ++l) {
      bf16* kp = e->kv_pool + (size_t)(2 * l) * e->layer_kv_elems;
      bf16* vp = kp + e->layer_kv_elems;
      B2_TRY(launch_attn_decode(e->qrot, kp, vp, e->d_block_table, e->bt_stride, e->d_seq_lens, e->attn_scratch, e->attn, B,
                                e->nq, e->nkv, e->hd, c.block_size, e->stream, e->attn_scale));
    }
    return 0;
  };
  cudaEvent_t a, b;
  B2_CUDA_OK(cudaEventCreate(&a));
  B2_CUDA_OK(cudaEventCreate(&b));
  int rc = pass();  // warm-up; each layer's K/V (B x S x 4 KB) is a different region, so no pass finds its pages in L2
  if (rc == 0) rc = cudaEventRecord(a, e->stream) == cudaSuccess ? 0 : -5;
  for (int r = 0; r < reps && rc == 0; ++r) rc = pass();
  if (rc == 0) rc = cudaEventRecord(b, e->stream) == cudaSuccess ? 0 : -5;
  if (rc == 0) rc = cudaEventSynchronize(b) == cudaSuccess ? 0 : -5;
  float ms = 0.f;
  if (rc == 0) rc = cudaEventElapsedTime(&ms, a, b) == cudaSuccess ? 0 : -5;
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  if (rc != 0) {
    if (rc == -5) b2_set_error("b2_attn_profile: CUDA error: %s", cudaGetErrorString(cudaGetLastError()));
    return rc;
  }
  *ms_per_launch = ms / (float)(reps * c.num_layers);
  *kv_bytes_per_launch = bytes;
  return 0;
}

extern "C" int64_t b2_launch_count(b2_engine* e) { return e ? e->launches : 0; }

extern "C" int b2_prefix_cache_stats(b2_engine* e, int64_t* queried_tokens, int64_t* hit_tokens, int64_t* cached_pages) {
  B2_REQUIRE(e, "b2_prefix_cache_stats: null engine");
  if (queried_tokens) *queried_tokens = e->pc_queries;
  if (hit_tokens) *hit_tokens = e->pc_hits;
  if (cached_pages) *cached_pages = (int64_t)e->pc_map.size();
  return 0;
}

extern "C" int b2_set_tuning(b2_engine* e, int32_t key, int64_t value) {
  B2_REQUIRE(e, "b2_set_tuning: null engine");
  B2_CUDA_OK(cudaSetDevice(e->device));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  switch (key) {
    case B2_TUNE_SKIP_MASK: e->skip_mask = (unsigned)value; break;
    case B2_TUNE_PREFIX_CACHE:
      e->prefix_cache = value != 0;
      if (!e->prefix_cache) {  // forget everything: registered pages nobody holds go back to the free list
        for (int32_t b : e->pc_lru) {
          e->pc_in_lru[b] = 0;
          e->free_blocks.push_back(b);
        }
        e->pc_lru.clear();
        e->pc_map.clear();
        std::fill(e->blk_hash.begin(), e->blk_hash.end(), 0);
      }
      break;
    case B2_TUNE_TWO_LANE:
      B2_REQUIRE(value == 0 || e->stream2 != nullptr, "b2_set_tuning: this engine was created without a second lane (B2_MB=0 or OPT)");
      e->mb_min_batch = (int)value;
      break;
    default: b2_set_error("b2_set_tuning: unknown key %d", key); return -22;
  }
  for (auto& kv : e->graphs) cudaGraphExecDestroy(kv.second);  // captured with the old setting
  e->graphs.clear();
  e->graph_nodes.clear();
  return 0;
}

extern "C" int b2_set_debug(b2_engine* e, int32_t flags) {
  B2_REQUIRE(e, "b2_set_debug: null engine");
  B2_CUDA_OK(cudaSetDevice(e->device));
  e->debug_flags = flags;
  if ((flags & 1) && e->dbg_hidden == nullptr) {
    e->dbg_cap_tokens = std::min(e->cap_tokens, 1024);
    B2_TRY(dalloc(&e->dbg_hidden, (size_t)(e->cfg.num_layers + 2) * e->dbg_cap_tokens * e->cfg.hidden));
  }
  if (flags & 2) B2_TRY(ensure_logits(e));
  return 0;
}

extern "C" int b2_get_hidden(b2_engine* e, int32_t layer, int32_t ntokens, uint16_t* out_bf16) {
  B2_REQUIRE(e && out_bf16, "b2_get_hidden: null argument");
  B2_REQUIRE((e->debug_flags & 1) && e->dbg_hidden, "b2_get_hidden: hidden capture not enabled (b2_set_debug(e, 1))");
  B2_REQUIRE(layer >= -1 && layer <= e->cfg.num_layers, "b2_get_hidden: layer %d out of range", layer);
  B2_REQUIRE(ntokens >= 1 && ntokens <= e->dbg_cap_tokens && ntokens <= e->last_tokens, "b2_get_hidden: ntokens %d out of range", ntokens);
  B2_CUDA_OK(cudaSetDevice(e->device));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  const int slot = layer + 1;  // -1 -> embeddings, l -> residual after layer l, L -> final norm output
  B2_CUDA_OK(cudaMemcpy(out_bf16, e->dbg_hidden + (size_t)slot * e->dbg_cap_tokens * e->cfg.hidden,
                        (size_t)ntokens * e->cfg.hidden * 2, cudaMemcpyDeviceToHost));
  return 0;
}

extern "C" int b2_get_logits(b2_engine* e, int32_t nseq, float* out_logits) {
  B2_REQUIRE(e && out_logits, "b2_get_logits: null argument");
  B2_REQUIRE((e->debug_flags & 2) && e->logits, "b2_get_logits: logits capture not enabled (b2_set_debug(e, 2))");
  B2_REQUIRE(nseq >= 1 && nseq <= e->last_n, "b2_get_logits: nseq %d > sequences of the last forward (%d)", nseq, e->last_n);
  B2_CUDA_OK(cudaSetDevice(e->device));
  const size_t n = (size_t)nseq * e->cfg.vocab;
  float* tmp = nullptr;
  B2_TRY(dalloc(&tmp, n));
  int rc = launch_logits_to_f32(e->logits, tmp, n, e->stream);
  if (rc == 0) {
    cudaError_t ce = cudaMemcpyAsync(out_logits, tmp, n * 4, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
    if (ce != cudaSuccess) {
      b2_set_error("b2_get_logits: copy failed: %s", cudaGetErrorString(ce));
      rc = -5;
    }
  }
  cudaFree(tmp);
  return rc;
}

extern "C" int b2_set_next_tokens(b2_engine* e, int32_t nseq, const int32_t* seq_ids, const int32_t* tokens) {
  B2_REQUIRE(e && seq_ids && tokens, "b2_set_next_tokens: null argument");
  B2_REQUIRE((int)e->cur_batch.size() == nseq && std::equal(seq_ids, seq_ids + nseq, e->cur_batch.begin()),
             "b2_set_next_tokens: sequences must be the resident decode batch, in order");
  B2_CUDA_OK(cudaSetDevice(e->device));
  memcpy(e->h_stage, tokens, nseq * 4);
  B2_CUDA_OK(cudaMemcpyAsync(e->d_tokens, e->h_stage, nseq * 4, cudaMemcpyHostToDevice, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  return 0;
}

extern "C" int b2_sample(b2_engine* e, int32_t nseq, const b2_sampling* params, int32_t* out_tokens) {
  B2_REQUIRE(e && params && out_tokens, "b2_sample: null argument");
  B2_REQUIRE((e->debug_flags & 2) && e->logits, "b2_sample: logits capture not enabled (b2_set_debug(e, 2))");
  B2_REQUIRE(nseq >= 1 && nseq <= e->last_n, "b2_sample: nseq %d > sequences of the last forward (%d)", nseq, e->last_n);
  B2_CUDA_OK(cudaSetDevice(e->device));
  e->h_sample_rows.resize(nseq);
  for (int i = 0; i < nseq; ++i) {
    // one seed for the whole batch: row i draws from its own stream derived from (seed, i)
    unsigned long long sd = params->seed == 0 ? 0ull : (params->seed * 0x9E3779B97F4A7C15ull + (unsigned long long)i);
    if (params->seed != 0 && sd == 0) sd = 1;
    e->h_sample_rows[i] = SampleRow{i, params->temperature, params->top_k, params->top_p, sd};
  }
  B2_TRY(launch_sample_rows(e->logits, e->h_sample_rows.data(), e->d_sample_rows, e->d_out_tokens, nseq, e->cfg.vocab, (uint64_t)e->launches,
                            e->stream));
  e->launches++;
  B2_CUDA_OK(cudaMemcpyAsync(e->h_out, e->d_out_tokens, nseq * 4, cudaMemcpyDeviceToHost, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  memcpy(out_tokens, e->h_out, nseq * 4);
  return 0;
}

extern "C" int b2_sample_rows(b2_engine* e, int32_t nrows, const int32_t* rows, const b2_sampling* params, int32_t* out_tokens) {
  B2_REQUIRE(e && rows && params && out_tokens && nrows >= 1, "b2_sample_rows: bad arguments");
  B2_REQUIRE((e->debug_flags & 2) && e->logits, "b2_sample_rows: logits capture not enabled (b2_set_debug(e, 2))");
  B2_REQUIRE(nrows <= e->cfg.max_seqs, "b2_sample_rows: too many rows");
  B2_CUDA_OK(cudaSetDevice(e->device));
  e->h_sample_rows.resize(nrows);
  for (int i = 0; i < nrows; ++i) {
    B2_REQUIRE(rows[i] >= 0 && rows[i] < e->last_n, "b2_sample_rows: row %d outside the last forward (%d rows)", rows[i], e->last_n);
    e->h_sample_rows[i] = SampleRow{rows[i], params[i].temperature, params[i].top_k, params[i].top_p, params[i].seed};
  }
  // one launch for all rows (the host vector outlives the copy: the call synchronises below)
  B2_TRY(launch_sample_rows(e->logits, e->h_sample_rows.data(), e->d_sample_rows, e->d_sample_out, nrows, e->cfg.vocab, (uint64_t)e->launches,
                            e->stream));
  e->launches++;
  B2_CUDA_OK(cudaMemcpyAsync(e->h_out, e->d_sample_out, nrows * 4, cudaMemcpyDeviceToHost, e->stream));
  B2_CUDA_OK(cudaStreamSynchronize(e->stream));
  memcpy(out_tokens, e->h_out, nrows * 4);
  return 0;
}
