// attn_decode.cu -- paged GQA decode attention
// (HF:integrations/sdpa_attention.py:40-104 at T == 1; eager form HF:models/qwen3/
// modeling_qwen3.py:196-218).  This is the HBM-dominant kernel of batch-64 decode: per step and
// layer it reads every resident K and V page exactly once.
//
// Work distribution ("stream-K" over pages): the pages of all (sequence, kv head) units are laid
// end to end, and each of the grid's persistent warps (8 per SM) takes one contiguous run of that
// line, so every warp moves the same number of bytes whatever the mix of sequence lengths, and a
// warp's TMA pipeline runs without interruption across unit boundaries.  Where a run ends inside a
// unit the warp leaves a partial (m, l, O) record and bumps the unit's arrival counter; the warp that
// arrives last merges the unit's 1..few records IN RUN ORDER (so the result does not depend on who
// was last) and writes the bf16 output.  No second kernel, no floating-point atomics.
//
// Per warp: lane 0 issues two 4 KB bulk copies (cp.async.bulk, one K and one V page-slice of
// [16 tokens][128 dims]) per stage onto an mbarrier, three stages deep.  Pages are stored with the
// 16-byte chunks of every token row XOR-swizzled by (row & 7) (kv_swizzled_chunk in kernels.h), so
// the dense copy lands in shared memory already bank-conflict-free for ldmatrix.  Math per page:
//   S = K.Q^T     mma.sync m16n8k16: 16 tokens x (2 query heads of the GQA group, padded to 8)
//   online softmax in fp32; p = exp(s*scale - m) ROUNDED TO bf16 for P.V (what torch's CPU flash
//   kernel does for bf16 inputs), the row sum uses the unrounded fp32 p
//   O^T += V^T.P^T  mma.sync with V read through ldmatrix.trans; fp32 accumulate
// Final O * (1/l) is rounded once to bf16 in the combine pass.
#include <mutex>

#include "../../include/b2_engine.h"
#include "kernels.h"

namespace b2 {

B2_STEP_TRACE_DEFINE(step_trace_set_attn)

constexpr int AD_WARPS_MAX = 16;
constexpr int AD_BS = 16;    // tokens per KV page
constexpr int AD_HD = 128;   // head dim
constexpr int AD_TILE = AD_BS * AD_HD * 2;       // 4096 B: one page-slice of K (or V) for one kv head
constexpr int AD_STAGE = 2 * AD_TILE;            // K + V
constexpr int AD_MAX_B = 512;
constexpr int AD_MAX_CTAS = 160;
constexpr int AD_CNT = 4096;                     // per-unit arrival counters (ints) after the geometry header; zero between launches
constexpr int AD_META = 1024 + AD_CNT;           // ints at the head of the scratch: [0]=pages per warp (base), [599]=warps with one extra, [1..B+1]=page prefix
constexpr int AD_LAST_PAGE = 1 << 30;            // flag on a slice id: last page of its sequence
constexpr int AD_REC = 2 * AD_HD + 4;            // floats per partial record: O[2][128], m[2], l[2]

__device__ __forceinline__ void ad_ldsm_x4(uint32_t addr, uint32_t* r) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ad_ldsm_x4_t(uint32_t addr, uint32_t* r) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ad_mma(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct AdPos {  // a position on the page line
  int b, kvh, j, nb;
};

#ifndef B2_ATTN_TRACE_BY_SMID
#define B2_ATTN_TRACE_BY_SMID 1  // timing breadcrumbs indexed by SM id (tools/attn_timeline.py)
#endif
template <int AD_WARPS, int AD_STAGES>
__global__ void __launch_bounds__(AD_WARPS * 32, 1)
attn_decode_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k_pages, const bf16* __restrict__ v_pages,
                   const int32_t* __restrict__ block_table, int bt_stride, const int32_t* __restrict__ seq_lens,
                   float* __restrict__ scratch, bf16* __restrict__ out, int B, int nq, int nkv, float scale, int early,
                   int window, int tag) {
  extern __shared__ __align__(128) uint8_t smem[];
  // Sliding-window attention (Phi-3, HF:models/phi3/modeling_phi3.py:265 sliding_window): a query at position
  // S-1 sees the last `window` keys only, i.e. tokens lo(S) = max(0, S - window) .. S-1.  Whole pages below lo are
  // never enumerated (first_page), the partly covered one is masked.  window == 0: everything is visible.
  auto lo_of = [&](int S) { return (window > 0 && S > window) ? S - window : 0; };
  auto first_page_of = [&](int S) { return lo_of(S) / AD_BS; };
  auto pages_of = [&](int S) { return (S + AD_BS - 1) / AD_BS - first_page_of(S); };
  __shared__ __align__(8) uint64_t bars[AD_WARPS][AD_STAGES];
  __shared__ int s_prefix[AD_MAX_B + 1];  // pages of sequences before b (exclusive prefix)

  // Everything up to pdl_wait() below runs while the previous kernel (QK-norm/RoPE, which writes q and
  // the NEW token's K/V slot) is still executing: sequence lengths and block tables were written at the
  // top of the step, and every page except the last one of each sequence is at least a step old.
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  B2_STAMP_BEGIN();
  if (!early) pdl_wait();
  if (threadIdx.x == 0) {
    for (int w = 0; w < AD_WARPS; ++w)
      for (int s = 0; s < AD_STAGES; ++s) mbar_init(smem_u32(&bars[w][s]), 1);
    mbar_fence_init();
  }
  // ---- page prefix over sequences (warp 0: a slab of entries per lane, then a warp scan) ----
  if (warp == 0) {
    const int per = (B + 31) / 32;
    int local = 0;
    for (int i = 0; i < per; ++i) {
      const int b = lane * per + i;
      if (b < B) local += pages_of(seq_lens[b]);
    }
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - local;
    for (int i = 0; i < per; ++i) {
      const int b = lane * per + i;
      if (b < B) {
        s_prefix[b] = run;
        run += pages_of(seq_lens[b]);
      }
    }
    if (lane == 31) s_prefix[B] = incl;
  }
  __syncthreads();
  const int total_pages = s_prefix[B] * nkv;  // length of the page line
  // balanced partition of the page line: the first `rem` warps take base+1 pages, the rest base
  const int wtot = gridDim.x * AD_WARPS;
  const int base = total_pages / wtot, rem = total_pages - base * wtot;
  unsigned long long t_start = 0;
  if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  const int gw = blockIdx.x * AD_WARPS + warp;
  const int p0 = gw * base + min(gw, rem);
  const int ntiles = base + (gw < rem ? 1 : 0);
  if (ntiles <= 0 && blockIdx.x != 0) return;
  const bool dbg_time = (warp == 0 && lane == 0);

  auto locate = [&](int p) -> AdPos {  // page-line position -> (sequence, kv head, page in sequence)
    int lo = 0, hi = B - 1;
    while (lo < hi) {  // largest b with nkv * prefix[b] <= p
      const int mid = (lo + hi + 1) >> 1;
      if (s_prefix[mid] * nkv <= p) lo = mid; else hi = mid - 1;
    }
    AdPos r;
    r.b = lo;
    r.nb = s_prefix[lo + 1] - s_prefix[lo];
    const int rem = p - s_prefix[lo] * nkv;
    r.kvh = rem / r.nb;
    r.j = rem - r.kvh * r.nb;
    return r;
  };

  uint8_t* wsm = smem + (size_t)warp * AD_STAGES * AD_STAGE;
  int cached_group = -1, cached_slice = 0;
  auto slice_of = [&](int n) -> int {  // (physical page * nkv + kv head) of this warp's n-th tile; all lanes call
    const int g = n >> 5;
    if (g != cached_group) {
      const int nn = (g << 5) + lane;
      cached_slice = 0;
      if (nn < ntiles) {
        const AdPos r = locate(p0 + nn);
        cached_slice = block_table[(size_t)r.b * bt_stride + r.j + first_page_of(seq_lens[r.b])] * nkv + r.kvh;
        if (r.j == r.nb - 1) cached_slice |= AD_LAST_PAGE;  // holds the slot the previous kernel is writing
      }
      cached_group = g;
    }
    return __shfl_sync(0xffffffffu, cached_slice, n & 31);
  };
  auto issue = [&](int n, int slice) {  // lane 0 only
    const int stage = n % AD_STAGES;
    const uint32_t bar = smem_u32(&bars[warp][stage]);
    const uint32_t dst = smem_u32(wsm + stage * AD_STAGE);
    const size_t off = (size_t)(slice & ~AD_LAST_PAGE) * (AD_BS * AD_HD);
    mbar_arrive_expect_tx(bar, AD_STAGE);
    bulk_g2s(dst, k_pages + off, AD_TILE, bar);
    bulk_g2s(dst + AD_TILE, v_pages + off, AD_TILE, bar);
  };
  // start streaming: pages that do not hold a freshly written slot go out before the dependency wait
  int issued = 0;
  for (int n = 0; n < AD_STAGES; ++n) {
    const int sl = slice_of(n);
    if (n < ntiles && issued == n && !(sl & AD_LAST_PAGE)) {
      if (lane == 0) issue(n, sl);
      issued = n + 1;
    }
  }
  pdl_wait();
  pdl_launch_dependents();
  B2_STAMP_WAIT();
  if (blockIdx.x == 0) {
    // empty sequences (padding rows of a graph bucket) own no pages: their output rows are zeros
    for (int b = 0; b < B; ++b)
      if (s_prefix[b + 1] == s_prefix[b])
        for (int i = threadIdx.x; i < nq * AD_HD / 8; i += blockDim.x)
          reinterpret_cast<uint4*>(out + (size_t)b * nq * AD_HD)[i] = make_uint4(0, 0, 0, 0);
  }
  if (ntiles <= 0) return;
  for (int n = issued; n < AD_STAGES; ++n) {
    const int sl = slice_of(n);
    if (n < ntiles && lane == 0) issue(n, sl);
  }

  const int g = lane >> 2, t4 = lane & 3;
  AdPos cur = locate(p0);
  int S = seq_lens[cur.b];
  int lo = lo_of(S), fp = first_page_of(S);  // first visible token / first enumerated page of the current sequence
  uint32_t qb0[8], qb1[8];
  float oacc[8][4];
  float m_run[2], l_run[2];
  bool fresh = true;  // (re)load q and reset the running state at the next tile

  // Arrival + merge of the units this warp's run touched, deferred to the end of the run so that the fence and
  // the atomic round trips never stall the TMA pipeline mid-stream -- and done for ALL pending units together:
  // one fence, the arrivals of up to 8 units issued by different lanes at once, and in a merge the loads of up
  // to four records in flight before the first is used (a serial version cost ~1 us per unit at the very end
  // of the slowest warps, i.e. of the kernel).
  auto unit_span = [&](int unit, int& ub, int& ukvh, int& w0, int& w1) {  // which warps' runs touch this unit
    ub = unit / nkv;
    ukvh = unit - ub * nkv;
    const int unb = s_prefix[ub + 1] - s_prefix[ub];
    const int u0 = s_prefix[ub] * nkv + ukvh * unb;
    const int cut = rem * (base + 1);
    w0 = u0 < cut ? u0 / (base + 1) : rem + (u0 - cut) / base;
    const int ulast = u0 + unb - 1;
    w1 = ulast < cut ? ulast / (base + 1) : rem + (ulast - cut) / base;
  };
  auto merge_unit = [&](int unit) {  // last to arrive: merge in run order and write the output
    float* recs = scratch + AD_META;
    int ub, ukvh, w0, w1;
    unit_span(unit, ub, ukvh, w0, w1);
    const int hh = lane >> 4, d0 = (lane & 15) * 8;  // lane covers 8 of the 256 (head, dim) outputs
    float M = -INFINITY;
    for (int wb = w0; wb <= w1; wb += 4) {
      float mv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (wb + j <= w1) mv[j] = __ldcg(recs + (size_t)(unit + wb + j) * AD_REC + 2 * AD_HD + hh);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (wb + j <= w1) M = fmaxf(M, mv[j]);
    }
    float O[8], L = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) O[e] = 0.f;
    for (int wb = w0; wb <= w1; wb += 4) {
      float mv[4], lv[4];
      float4 a[4], b[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (wb + j <= w1) {
          const float* r2 = recs + (size_t)(unit + wb + j) * AD_REC;
          mv[j] = __ldcg(r2 + 2 * AD_HD + hh);
          lv[j] = __ldcg(r2 + 2 * AD_HD + 2 + hh);
          a[j] = __ldcg(reinterpret_cast<const float4*>(r2 + hh * AD_HD + d0));
          b[j] = __ldcg(reinterpret_cast<const float4*>(r2 + hh * AD_HD + d0 + 4));
        }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (wb + j <= w1) {  // run order: the sums do not depend on which warp does the merge
          const float f = expf(mv[j] - M);
          O[0] = fmaf(f, a[j].x, O[0]); O[1] = fmaf(f, a[j].y, O[1]); O[2] = fmaf(f, a[j].z, O[2]); O[3] = fmaf(f, a[j].w, O[3]);
          O[4] = fmaf(f, b[j].x, O[4]); O[5] = fmaf(f, b[j].y, O[5]); O[6] = fmaf(f, b[j].z, O[6]); O[7] = fmaf(f, b[j].w, O[7]);
          L = fmaf(f, lv[j], L);
        }
    }
    const float inv = L > 0.f ? __fdiv_rn(1.0f, L) : 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) O[e] *= inv;
    *reinterpret_cast<uint4*>(out + ((size_t)ub * nq + ukvh * 2 + hh) * AD_HD + d0) = pack8(O);
  };
  auto arrive_all = [&](const int* units, int n) {
    if (n <= 0) return;
    __threadfence();  // this warp's records are visible before its arrivals are
    __syncwarp();
    int my_unit = lane < n ? units[lane] : -1;
    bool last = false;
    if (my_unit >= 0) {
      int ub, ukvh, w0, w1;
      unit_span(my_unit, ub, ukvh, w0, w1);
      int* cnt = reinterpret_cast<int*>(scratch) + 1024 + my_unit;
      const int prev = atomicAdd(cnt, 1);
      if (prev == w1 - w0) {
        last = true;
        *cnt = 0;  // ready for the next launch
      }
    }
    unsigned todo = __ballot_sync(0xffffffffu, last);
    if (todo) __threadfence();  // acquire side: the other warps' records
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      merge_unit(__shfl_sync(0xffffffffu, my_unit, src));
    }
  };
  int pending[8];
  int npending = 0;

  for (int n = 0; n < ntiles; ++n) {
    if (fresh) {
      const bf16* qp = q + ((size_t)cur.b * nq + (size_t)cur.kvh * 2 + (g & 1)) * AD_HD;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        qb0[ks] = g < 2 ? *reinterpret_cast<const uint32_t*>(qp + ks * 16 + 2 * t4) : 0u;
        qb1[ks] = g < 2 ? *reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8 + 2 * t4) : 0u;
      }
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) oacc[dt][0] = oacc[dt][1] = oacc[dt][2] = oacc[dt][3] = 0.f;
      m_run[0] = m_run[1] = -INFINITY;
      l_run[0] = l_run[1] = 0.f;
      fresh = false;
    }
    const int stage = n % AD_STAGES;
    mbar_wait(smem_u32(&bars[warp][stage]), (uint32_t)((n / AD_STAGES) & 1));
    const uint32_t kb = smem_u32(wsm + stage * AD_STAGE);
    const uint32_t vb = kb + AD_TILE;

    // ---- S = K.Q^T : 16 tokens x 8 (2 real) heads ----
    float sacc[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const int row = (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        uint32_t a[4];
        ad_ldsm_x4(kb + row * (AD_HD * 2) + (((2 * ks + (lane >> 4)) ^ (row & 7)) << 4), a);
        ad_mma(sacc, a, qb0[ks], qb1[ks]);
      }
    }
    // lanes with t4 == 0 hold (token g | g+8) x (head 0 | 1); give every lane of the quad a copy
    float s[2][2];  // [token half][head]
    s[0][0] = __shfl_sync(0xffffffffu, sacc[0], lane & ~3);
    s[0][1] = __shfl_sync(0xffffffffu, sacc[1], lane & ~3);
    s[1][0] = __shfl_sync(0xffffffffu, sacc[2], lane & ~3);
    s[1][1] = __shfl_sync(0xffffffffu, sacc[3], lane & ~3);
    const int tok = (cur.j + fp) * AD_BS + g;
    const bool ok_a = tok < S && tok >= lo, ok_b = tok + 8 < S && tok + 8 >= lo;
    float p[2][2], corr[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float sa = ok_a ? s[0][h] * scale : -INFINITY;
      const float sb = ok_b ? s[1][h] * scale : -INFINITY;
      float mx = fmaxf(sa, sb);
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
      const float m_new = fmaxf(m_run[h], mx);  // finite: every page holds >= 1 valid token
      corr[h] = expf(m_run[h] - m_new);
      const float pa = expf(sa - m_new), pb = expf(sb - m_new);
      float ls = pa + pb;
      ls += __shfl_xor_sync(0xffffffffu, ls, 4);
      ls += __shfl_xor_sync(0xffffffffu, ls, 8);
      ls += __shfl_xor_sync(0xffffffffu, ls, 16);
      l_run[h] = l_run[h] * corr[h] + ls;
      m_run[h] = m_new;
      p[0][h] = pa;
      p[1][h] = pb;
    }
    // ---- P^T as the B operand: lane (g = head, t4) needs tokens 2t4, 2t4+1 (+8) of head g ----
    uint32_t pb0, pb1;
    {
      const int src = 8 * t4;  // quad that owns token 2*t4; the next quad (+4 lanes) owns token 2*t4+1
      const float a00 = __shfl_sync(0xffffffffu, p[0][0], src), a01 = __shfl_sync(0xffffffffu, p[0][0], src + 4);
      const float a10 = __shfl_sync(0xffffffffu, p[0][1], src), a11 = __shfl_sync(0xffffffffu, p[0][1], src + 4);
      const float b00 = __shfl_sync(0xffffffffu, p[1][0], src), b01 = __shfl_sync(0xffffffffu, p[1][0], src + 4);
      const float b10 = __shfl_sync(0xffffffffu, p[1][1], src), b11 = __shfl_sync(0xffffffffu, p[1][1], src + 4);
      pb0 = g == 0 ? pack_bf16x2(a00, a01) : g == 1 ? pack_bf16x2(a10, a11) : 0u;
      pb1 = g == 0 ? pack_bf16x2(b00, b01) : g == 1 ? pack_bf16x2(b10, b11) : 0u;
    }
    // ---- O^T = O^T * corr + V^T.P^T ----
    {
      const float c0 = corr[0], c1 = corr[1];  // accumulator columns are heads (2*t4, 2*t4+1): real for t4 == 0
      const int row = (lane & 7) + ((lane >> 4) & 1) * 8;
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) {
        oacc[dt][0] *= c0; oacc[dt][1] *= c1; oacc[dt][2] *= c0; oacc[dt][3] *= c1;
        uint32_t a[4];
        ad_ldsm_x4_t(vb + row * (AD_HD * 2) + (((2 * dt + ((lane >> 3) & 1)) ^ (row & 7)) << 4), a);
        ad_mma(oacc[dt], a, pb0, pb1);
      }
    }
    __syncwarp();
    {
      const int nxt = n + AD_STAGES;
      const int sl = slice_of(nxt);
      if (nxt < ntiles && lane == 0) issue(nxt, sl);
    }
    // ---- advance along the page line; leave a record when the unit (or the run) ends ----
    const int unit = cur.b * nkv + cur.kvh;
    bool unit_ends = (n + 1 == ntiles);
    if (++cur.j == cur.nb) {
      unit_ends = true;
      cur.j = 0;
      if (++cur.kvh == nkv) {
        cur.kvh = 0;
        do { ++cur.b; } while (cur.b < B && s_prefix[cur.b + 1] == s_prefix[cur.b]);
        if (cur.b < B) {
          cur.nb = s_prefix[cur.b + 1] - s_prefix[cur.b];
          S = seq_lens[cur.b];
          lo = lo_of(S);
          fp = first_page_of(S);
        }
      }
    }
    if (unit_ends) {
      float* recs = scratch + AD_META;
      float* rec = recs + (size_t)(unit + gw) * AD_REC;
      if (t4 == 0) {
#pragma unroll
        for (int dt = 0; dt < 8; ++dt) {
          rec[dt * 16 + g] = oacc[dt][0];
          rec[AD_HD + dt * 16 + g] = oacc[dt][1];
          rec[dt * 16 + 8 + g] = oacc[dt][2];
          rec[AD_HD + dt * 16 + 8 + g] = oacc[dt][3];
        }
      }
      if (lane == 0) {
        rec[2 * AD_HD + 0] = m_run[0];
        rec[2 * AD_HD + 1] = m_run[1];
        rec[2 * AD_HD + 2] = l_run[0];
        rec[2 * AD_HD + 3] = l_run[1];
      }
      if (npending == 8) {  // (never in practice: more than 8 unit ends in one run)
        arrive_all(pending, npending);
        npending = 0;
      }
      pending[npending++] = unit;
      fresh = true;
    }
  }
  arrive_all(pending, npending);
  B2_STAMP_END_AUX(B2K_ATTN_DECODE, (unsigned)tag);  // warp 0's end: all warps of a CTA run the same number of pages
  if (dbg_time) {  // timing breadcrumbs for tools/attn_timeline.py (unused words of the geometry header)
    unsigned long long t_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
    int* meta = reinterpret_cast<int*>(scratch);
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    const int slot = B2_ATTN_TRACE_BY_SMID ? (int)smid : (int)blockIdx.x;
    if (slot < 148) {
      meta[600 + 2 * slot] = (int)(t_start & 0x7fffffff);
      meta[601 + 2 * slot] = (int)(t_end & 0x7fffffff);
    }
  }
}

size_t attn_decode_scratch_floats(int B, int nq, int nkv, int hd) {
  (void)nq;
  (void)hd;
  return (size_t)AD_META + ((size_t)B * nkv + (size_t)AD_MAX_CTAS * AD_WARPS_MAX) * AD_REC;
}

// (warps per CTA, pipeline stages per warp).  Default 8 x 2 x 8 KB = 128 KB of K/V in flight per SM.  Measured on
// B200 at the bench shape (median / slowest CTA busy time, us): 8x2 28.5/36, 8x3 28.6/36, 16x1 28.1/35, 12x1 28.1/35,
// 10x1 28.5/35, 6x2 30.2/35, 8x1 30.1/37 -- nothing below 96 KB in flight keeps up, nothing above 128 KB helps, and
// the spread between the SMs of one launch is the same for all of them.
// B2_ATTN_CFG=<warps>x<stages> selects another built shape for experiments.
static int ad_grid() {
  static int sms = device_num_sms();
  static int forced = [] { const char* e = getenv("B2_ATTN_GRID"); return e ? atoi(e) : 0; }();  // experiments
  if (forced > 0 && forced <= AD_MAX_CTAS) return forced;
  return sms < AD_MAX_CTAS ? sms : AD_MAX_CTAS;
}

template <int W, int S>
static int launch_cfg(const bf16* q, const bf16* k_pages, const bf16* v_pages, const int32_t* block_table, int bt_stride,
                      const int32_t* seq_lens, float* scratch, bf16* out, int B, int nq, int nkv, float scale, int window,
                      cudaStream_t s, int tag) {
  constexpr int SMEM = W * S * AD_STAGE;
  static B2PerDeviceOnce attr_once, carve_once;
  if (attr_once.need()) B2_CUDA_OK(cudaFuncSetAttribute(attn_decode_kernel<W, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  b2_prefer_max_smem(attn_decode_kernel<W, S>, carve_once);
  B2_CUDA_OK(b2_launch(attn_decode_kernel<W, S>, dim3(ad_grid()), dim3(W * 32), (size_t)SMEM, s, q, k_pages, v_pages, block_table,
                       bt_stride, seq_lens, scratch, out, B, nq, nkv, scale, b2_pdl_early_mask() & 1, window, tag));
  return 0;
}

// scratch: attn_decode_scratch_floats() floats; its counter words (AD_CNT ints after the first 1024)
// must be zero when the kernel starts -- the kernel leaves them zero again when it is done.
int launch_attn_decode(const bf16* q, const bf16* k_pages, const bf16* v_pages, const int32_t* block_table,
                       int bt_stride, const int32_t* seq_lens, float* scratch, bf16* out, int B, int nq, int nkv,
                       int hd, int block_size, cudaStream_t s, float softmax_scale, int window, int tag) {
  if (B <= 0) return 0;
  B2_REQUIRE(window >= 0, "attn_decode: negative window");
  B2_REQUIRE(hd == AD_HD && block_size == AD_BS, "attn_decode: built for head_dim 128 / page 16 (got %d / %d)", hd, block_size);
  B2_REQUIRE(nq == 2 * nkv, "attn_decode: built for a GQA group of 2 (nq=%d nkv=%d)", nq, nkv);
  B2_REQUIRE(B <= AD_MAX_B && B * nkv <= AD_CNT, "attn_decode: batch %d x %d kv heads exceeds %d units", B, nkv, AD_CNT);
  B2_REQUIRE(scratch != nullptr, "attn_decode: scratch required");
  // HF passes head_dim**-0.5 as a double
  const float scale = softmax_scale > 0.f ? softmax_scale : (float)(1.0 / sqrt((double)hd));
  static int cfg = -1;
  if (cfg < 0) {
    const char* e = getenv("B2_ATTN_CFG");
    cfg = 82;
    if (e) { int w = 0, st = 0; if (sscanf(e, "%dx%d", &w, &st) == 2) cfg = w * 10 + st; }
  }
#define AD_CASE(W, S) case (W) * 10 + (S): return launch_cfg<W, S>(q, k_pages, v_pages, block_table, bt_stride, seq_lens, scratch, out, B, nq, nkv, scale, window, s, tag)
  switch (cfg) {
    AD_CASE(8, 3);
    AD_CASE(6, 4);
    AD_CASE(12, 2);
    AD_CASE(9, 3);
    AD_CASE(13, 2);
    AD_CASE(4, 6);
    AD_CASE(8, 2);
    AD_CASE(16, 1);
    AD_CASE(8, 1);
    AD_CASE(6, 2);
    AD_CASE(12, 1);
    AD_CASE(10, 1);
  }
#undef AD_CASE
  b2_set_error("attn_decode: B2_ATTN_CFG shape %d not built", cfg);
  return -22;
}

int attn_decode_zero_counters(float* scratch, cudaStream_t s) {
  B2_CUDA_OK(cudaMemsetAsync(reinterpret_cast<int*>(scratch) + 1024, 0, AD_CNT * sizeof(int), s));
  return 0;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant of b2k_attn_decode.
//
// Some users of this library are not in a position to hand the kernel a scratch buffer:
//   * a serving process that gave all remaining HBM to its KV pool at start-up (b2serve does, via
//     gpu_mem_fraction) -- a cudaMalloc on the request path can simply fail there;
//   * request-path code that must not stall: cudaMalloc/cudaFree synchronise the device, so one
//     call delays every stream of the process, not just the caller's;
//   * a plugin that only receives raw input/output pointers from a host framework and has no
//     allocator of its own to ask.
// For them the partial records live in the file-scope device array below.  It is part of the
// library image: every GPU that loads libb2engine.so holds B2K_ATTN_NOALLOC_MAX_FLOATS * 4 bytes
// (8 MiB) for it until the process exits, whether or not this entry point is ever called.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) float g_attn_noalloc_scratch[B2K_ATTN_NOALLOC_MAX_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};

// In host code `g_attn_noalloc_scratch` is only a handle for the symbol, not a device address: a
// kernel given `&g_attn_noalloc_scratch[0]` from here would fault (or, where the GPU can reach host
// memory, quietly use the wrong bytes).  The device address is looked up once per device, with that
// device current, and cached.
int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  B2_CUDA_OK(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < kMaxDevices, "b2k_attn_decode_noalloc: device index %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    B2_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_attn_noalloc_scratch));
    B2_REQUIRE(p != nullptr, "b2k_attn_decode_noalloc: no device address for the static scratch on device %d", dev);
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

}  // namespace b2

extern "C" int b2k_attn_decode_noalloc(const void* q, const void* k_pages, const void* v_pages, const int32_t* block_table,
                                       int bt_stride, const int32_t* seq_lens, void* out, int B, int nq, int nkv, int hd,
                                       int block_size, void* stream) {
  using namespace b2;
  B2_REQUIRE(q && k_pages && v_pages && block_table && seq_lens && out, "b2k_attn_decode_noalloc: null pointer");
  B2_REQUIRE(B >= 1 && nq >= 1 && nkv >= 1, "b2k_attn_decode_noalloc: bad sizes");
  const size_t need = attn_decode_scratch_floats(B, nq, nkv, hd);
  B2_REQUIRE(need <= (size_t)B2K_ATTN_NOALLOC_MAX_FLOATS,
             "b2k_attn_decode_noalloc: batch %d x %d kv heads needs %zu floats of scratch, the built-in array holds %d; "
             "use a smaller batch or b2k_attn_decode with your own buffer",
             B, nkv, need, (int)B2K_ATTN_NOALLOC_MAX_FLOATS);
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  // The only words that must start at zero are the per-unit arrival counters (16 KB); the kernel
  // leaves them at zero itself, but they are cleared here as well so that a call cut short (a fault
  // in an earlier launch) cannot poison the next one.  Records and geometry words are always written
  // before they are read within one call, so their old contents are never observed.
  rc = attn_decode_zero_counters(scratch, (cudaStream_t)stream);
  if (rc != 0) return rc;
  return launch_attn_decode((const bf16*)q, (const bf16*)k_pages, (const bf16*)v_pages, block_table, bt_stride, seq_lens, scratch,
                            (bf16*)out, B, nq, nkv, hd, block_size, (cudaStream_t)stream);
}
