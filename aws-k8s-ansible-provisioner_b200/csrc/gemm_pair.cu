// gemm_pair.cu -- two projections of a decode step with the element-wise kernel that used to follow them folded in:
//   gate/up projection + SwiGLU                      (HF:models/qwen3/modeling_qwen3.py:81-83)
//   QKV projection + QK-norm + RoPE + KV page write  (HF:...:236-249,263-271,151-181)
// for at most 64 tokens.  Same tcgen05 / TMEM / TMA machinery and the same swap-AB layout as gemm_tc.cu (weight rows on
// MMA-M, tokens on MMA-N); what is new is how split-K is reduced.
//
// Why: in the captured step (tools/step_timeline.py, profiles/r02_step_timeline_one_lane.txt) a projection cost
// ~1.4 us of hand-off + ~3.5 us of work and the element-wise kernel behind it another ~1.3 + ~2.3 us, most of which
// is writing fp32 split-K partials to L2, a kernel boundary, and reading them back.  The projections need split-K
// because one SM streams weights at ~60 GB/s at best: the 8-13 MB of a projection have to be spread over ~100 SMs.
//
// Here a CLUSTER OF TWO CTAs owns one 128-row weight tile; CTA r multiplies K-half r (all of its weight slice,
// <= 128 KB, is in flight before the dependency wait).  Then the pair exchanges halves of the fp32 accumulator through
// distributed shared memory -- CTA r keeps the tokens of half r and stores the other half straight into its peer's
// shared memory -- one cluster barrier, and each CTA finishes ITS tokens: add the two partials (two addends: the sum
// does not depend on the order), round to bf16 as nn.Linear does, apply the element-wise step, store.  No partials in
// global memory, no second kernel.  Rounding points are those of elementwise.cu.
//
//   SwiGLU tile  = 64 gate rows f0..f0+63 (TMEM lanes 0..63) + the 64 up rows of the same features (lanes 64..127)
//   QKV tile     = one 128-wide head
#include "kernels.h"

namespace b2 {

B2_STEP_TRACE_DEFINE(step_trace_set_pair)

namespace {

constexpr int GP_BM = 128, GP_BK = 64;
constexpr int GP_W_BYTES = GP_BM * GP_BK * 2;  // 16 KB of weights per k-block
constexpr int GP_MAX_KB = 8;                   // k-blocks per CTA (K/2 <= 512)
constexpr int GP_LOADERS = 256;                // X loader threads when the rows are normalised on load (warps 2..9), else 128 (warps 2..5)
constexpr int GP_THREADS = 64 + GP_LOADERS;    // largest block; warps 2..5 are also the epilogue

template <int BN>
struct GpCfg {
  static constexpr int X_BYTES = BN * GP_BK * 2;
  static constexpr int STAGE_BYTES = GP_W_BYTES + X_BYTES;
  static constexpr int HALF = BN / 2;                       // tokens a CTA finishes
  static constexpr int RED_BYTES = HALF * GP_BM * 4;        // peer's partial for my tokens: [token][row] fp32
  static constexpr int CS_BYTES = HALF * 128 * 2;           // QKV mode: the cos|sin table rows of my tokens
  static constexpr int SMEM_BYTES = GP_MAX_KB * STAGE_BYTES + RED_BYTES + CS_BYTES + 1024;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
};

enum { GP_SWIGLU = 0, GP_QKV_ROPE = 1 };

// Normalise-on-load: x is the RESIDUAL stream h (not a normalised copy) and ssq holds, per 16-feature piece, the sum of
// squares of each token's row (written by gemm_resid.cu); the X loader threads scale the rows as they stage them:
//   x = bf16(gain * bf16(h * rsqrt(mean(h^2) + eps)))          (HF:models/qwen3/modeling_qwen3.py:50-62)
// so that no norm kernel runs between the projection that updated h and this one.  ssq == nullptr: x is used as it is.
struct GpNorm {
  const float* ssq;   // [pieces][ld]
  const bf16* gain;   // [K]
  int pieces, ld;
  float eps;
  L2Prefetch next_w;  // weights of the projection after this one, to be pulled into L2 (independent of ssq)
};

struct GpRope {  // QKV mode
  const bf16 *qw, *kw, *cos_sin;
  const int32_t *positions, *slots;
  bf16 *q_out, *k_pages, *v_pages;
  int nq, nkv, block_size;
  float eps;
};

__device__ __forceinline__ uint32_t gp_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t gp_mapa(uint32_t local_smem, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem), "r"(rank));
  return r;
}
__device__ __forceinline__ void gp_st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// grid = 2 * m_tiles CTAs in clusters of 2; block = 320 threads: warp 0 TMA producer (weights), warp 1 MMA issuer,
// warps 2-9 bring in X (cp.async, or through registers when the rows are normalised on the way), warps 2-5 then
// exchange / finish the accumulator.
template <int BN, int MODE>
__global__ void __launch_bounds__(GP_THREADS, 1)
linear_pair_kernel(const __grid_constant__ CUtensorMap tmap_w, const bf16* __restrict__ x, bf16* __restrict__ out, int T, int N, int K,
                   int up_row0, int m_tiles, const GpRope rp, const GpNorm nm, int tag) {
  using C = GpCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar, tmem_full_bar;
  __shared__ uint32_t tmem_slot;
  B2_STAMP_BEGIN();
#ifdef B2_STEP_TRACE
  __shared__ unsigned long long phase_t[3];  // accumulator visible, pair barrier passed, tail done (thread 64)
#define GP_PHASE(i) do { if (threadIdx.x == 64 && _stamp[0] != 0ull) phase_t[i] = b2_globaltimer(); } while (0)
#else
#define GP_PHASE(i) do { } while (0)
#endif
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t red_base = smem_base + GP_MAX_KB * C::STAGE_BYTES;
  float* red = reinterpret_cast<float*>(smem_al + GP_MAX_KB * C::STAGE_BYTES);
  bf16* cs_rows = reinterpret_cast<bf16*>(smem_al + GP_MAX_KB * C::STAGE_BYTES + C::RED_BYTES);  // [token][cos 64 | sin 64]
  __shared__ int s_slot[C::HALF];
  __shared__ __align__(16) bf16 s_gain[2 * 128];  // QKV mode: QK-norm gains, q then k
  __shared__ __align__(16) bf16 s_ngain[GP_MAX_KB * GP_BK];  // normalise-on-load: the RMSNorm gains of this CTA's K half
  __shared__ float s_rstd[BN];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = gp_cluster_rank();
  const int m_tile = (blockIdx.x >> 1) % m_tiles;
  const int tok0 = ((blockIdx.x >> 1) / m_tiles) * BN;  // first token of this pair's tile (QKV at T > 32 runs two token tiles)
  const int kb_total = K / GP_BK;
  const int nkb = kb_total / 2;          // k-blocks of this CTA: [rank * nkb, rank * nkb + nkb)
  const int kb0 = (int)rank * nkb;

  if (threadIdx.x == 0) {
    // ONE barrier for all operands (all weight bytes + one arrival per X-loader thread).  Two barriers -- MMAs of the
    // first k-blocks issued while the second half of X lands -- measured slower (accumulator ready 2.6 instead of 2.05 us
    // after the wait): X lands within one L2 round trip anyway, and every extra wait / proxy fence costs ~0.1-0.3 us.
    mbar_init(smem_u32(&full_bar), 1 + (blockDim.x - 64));  // all weight bytes + one arrival per X-loader thread
    mbar_init(smem_u32(&tmem_full_bar), 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&tmem_slot), C::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_slot;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_w);
    const uint32_t fb = smem_u32(&full_bar);
    mbar_arrive_expect_tx(fb, (uint32_t)nkb * GP_W_BYTES);
    for (int i = 0; i < nkb; ++i) {
      const uint32_t ws = smem_base + i * C::STAGE_BYTES;
      if (MODE == GP_SWIGLU) {  // 64 gate rows, then the 64 up rows of the same features (two 64-row boxes)
        tma_load_2d(ws, &tmap_w, (kb0 + i) * GP_BK, m_tile * 64, fb);
        tma_load_2d(ws + GP_W_BYTES / 2, &tmap_w, (kb0 + i) * GP_BK, up_row0 + m_tile * 64, fb);
      } else {
        tma_load_2d(ws, &tmap_w, (kb0 + i) * GP_BK, m_tile * GP_BM, fb);
      }
    }
  }
  if (warp == 0) b2_l2_prefetch(nm.next_w, blockIdx.x, gridDim.x, lane);
  // both CTAs of the pair are running (their shared memory is live) before anyone stores into the peer's
  __syncwarp();
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  if (MODE == GP_QKV_ROPE && warp >= 2) {
    // Before the dependency wait: positions and slots were written at the top of the step (several kernels ago), the
    // RoPE table is a constant.  Each token's table row goes to shared memory now, so that the tail below is not a
    // chain of dependent global loads per token.
    const int et = threadIdx.x - 64;
    const int tb = tok0 + (int)rank * C::HALF;
    if (et < 32) reinterpret_cast<uint4*>(s_gain)[et] = reinterpret_cast<const uint4*>(et < 16 ? rp.qw : rp.kw)[et & 15];  // q gains | k gains
    for (int idx = et; idx < C::HALF * 16; idx += (int)blockDim.x - 64) {  // 16 chunks of 16 bytes per 256-byte row
      const int j = idx >> 4, c = idx & 15;
      const int tok = tb + j;
      const int pos = tok < T ? max(rp.positions[tok], 0) : 0;
      reinterpret_cast<uint4*>(cs_rows + j * 128)[c] = reinterpret_cast<const uint4*>(rp.cos_sin + (size_t)pos * 128)[c];
    }
    if (et < C::HALF) s_slot[et] = (tb + et < T) ? rp.slots[tb + et] : -1;
  }
  if (nm.ssq != nullptr && warp >= 2) {  // the gains are constants: staged before the dependency wait
    const int et = threadIdx.x - 64;
    if (et < nkb * (GP_BK / 8)) reinterpret_cast<uint4*>(s_ngain)[et] = reinterpret_cast<const uint4*>(nm.gain + (size_t)kb0 * GP_BK)[et];
  }
  pdl_wait();
  if (threadIdx.x == 0) pdl_launch_dependents();
  B2_STAMP_WAIT();

  if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(GP_BM, BN);
      mbar_wait(smem_u32(&full_bar), 0);
      tc_fence_after();
      for (int i = 0; i < nkb; ++i) {
        const uint32_t ws = smem_base + i * C::STAGE_BYTES;
        const uint64_t da = umma_desc_sw128(ws);
        const uint64_t db = umma_desc_sw128(ws + GP_W_BYTES);
#pragma unroll
        for (int k = 0; k < GP_BK / 16; ++k)
          tc_mma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (i > 0 || k > 0) ? 1u : 0u);
      }
      tc_commit(smem_u32(&tmem_full_bar));
    }
  } else if (warp >= 2) {
    // ---- X rows [0, BN) x this CTA's K half into the 128-byte-swizzled tile layout, 16 bytes per thread and step ----
    const int et = threadIdx.x - 64;
    const bf16* xg = x + (size_t)tok0 * K + (size_t)kb0 * GP_BK;
    constexpr int CHUNKS = BN * 8;                        // 16-byte chunks per k-block tile
    constexpr int CPT = GP_MAX_KB * CHUNKS / GP_LOADERS;  // chunks per thread at most: 16 (BN 64) or 8 (BN 32)
    if (nm.ssq == nullptr) {  // 128 loader threads (a 192-thread block)
#pragma unroll 1
      for (int i = 0; i < nkb; ++i) {
        const uint32_t xs = smem_base + i * C::STAGE_BYTES + GP_W_BYTES;
#pragma unroll
        for (int c0 = 0; c0 < CHUNKS; c0 += 128) {
          const int idx = c0 + et;
          const int r = idx >> 3, c = idx & 7;
          const bool in = tok0 + r < T;
          const bf16* src = xg + (size_t)(in ? r : 0) * K + i * GP_BK + c * 8;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(xs + r * 128 + ((c ^ (r & 7)) << 4)), "l"(src), "r"(in ? 16 : 0)
                       : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    } else {
      // Normalise on load.  Everything this thread needs is requested at once -- its share of the statistic's pieces and
      // its 16-byte chunks of h, through registers -- so the whole stage costs one L2 round trip plus ~25 instructions
      // per chunk.  (A first version copied with cp.async and rescaled the tile in shared memory with 128 threads: 1.4-2.2 us
      // more before the first MMA than the plain copy, which ate what dropping the norm kernel had saved.)
      constexpr int TPN = GP_LOADERS / BN;  // threads per token for the statistic: 4 or 8
      const int j = et / TPN, part = et % TPN;
      const int per = nm.pieces / TPN;      // <= 16
      const bool jin = tok0 + j < T;
      const float* sp = nm.ssq + (size_t)(part * per) * nm.ld + tok0 + (jin ? j : 0);
      float pv[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) pv[c] = c < per ? sp[(size_t)c * nm.ld] : 0.f;
      uint4 xr[CPT];
#pragma unroll
      for (int q = 0; q < CPT; ++q) {
        const int idx = q * GP_LOADERS + et;
        const int i = idx / CHUNKS, r = (idx % CHUNKS) >> 3, c = idx & 7;
        xr[q] = make_uint4(0, 0, 0, 0);  // rows past T are zeros, as the tensor map's out-of-bounds fill would give
        if (i < nkb && tok0 + r < T) xr[q] = *reinterpret_cast<const uint4*>(xg + (size_t)r * K + i * GP_BK + c * 8);
      }
      // rstd: `per` pieces in index order per thread, the TPN partial sums meet in lane order -- a fixed tree, the same in every CTA
      float ssum = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) ssum += pv[c];  // the zeros behind `per` do not change the sum
      ssum += __shfl_xor_sync(0xffffffffu, ssum, 1);
      ssum += __shfl_xor_sync(0xffffffffu, ssum, 2);
      if (TPN == 8) ssum += __shfl_xor_sync(0xffffffffu, ssum, 4);
      if (part == 0) s_rstd[j] = jin ? __fdiv_rn(1.0f, __fsqrt_rn(ssum / (float)K + nm.eps)) : 0.f;
      asm volatile("bar.sync 2, %0;" ::"n"(GP_LOADERS) : "memory");
#pragma unroll
      for (int q = 0; q < CPT; ++q) {
        const int idx = q * GP_LOADERS + et;
        const int i = idx / CHUNKS, r = (idx % CHUNKS) >> 3, c = idx & 7;
        if (i < nkb) {
          const uint4 g4 = *reinterpret_cast<const uint4*>(s_ngain + i * GP_BK + c * 8);
          const float rs = s_rstd[r];
          float v[8];
          unpack8(xr[q], v);
          const uint32_t* g2 = &g4.x;
          uint32_t o2[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t nx = pack_bf16x2(v[2 * e] * rs, v[2 * e + 1] * rs);  // bf16(h * rstd)
            const __nv_bfloat162 y = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&g2[e]), *reinterpret_cast<const __nv_bfloat162*>(&nx));
            o2[e] = *reinterpret_cast<const uint32_t*>(&y);
          }
          *reinterpret_cast<uint4*>(smem_al + i * C::STAGE_BYTES + GP_W_BYTES + r * 128 + ((c ^ (r & 7)) << 4)) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
        }
      }
    }
    fence_proxy_async();
    mbar_arrive(smem_u32(&full_bar));
  }

  // ---- exchange: my accumulator's columns of the OTHER half go to the peer's `red`, mine stay in registers ----
  float own[C::HALF];
  const int q4 = warp & 3;               // TMEM lane quarter this warp may read (epilogue warps 2..5 -> 2,3,0,1)
  const int row = q4 * 32 + lane;        // accumulator row = weight row of the tile
  const bool epi = warp >= 2 && warp < 6;
  if (epi) {
    mbar_wait(smem_u32(&tmem_full_bar), 0);
    GP_PHASE(0);
    tc_fence_after();
    const uint32_t tbase = tmem_d + ((uint32_t)(q4 * 32) << 16);
    const uint32_t peer_red = gp_mapa(red_base, rank ^ 1u);
#pragma unroll
    for (int c0 = 0; c0 < BN; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(tbase + (uint32_t)c0, r);
      tmem_ld_wait();
      const bool mine = (c0 / C::HALF) == (int)rank;  // uniform per chunk: HALF is a multiple of 16 (or the whole tile at BN = 16)
      if (mine) {
#pragma unroll
        for (int j = 0; j < 16; ++j) own[(c0 % C::HALF) + j] = __uint_as_float(r[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          gp_st_cluster_f32(peer_red + (uint32_t)((((c0 % C::HALF) + j) * GP_BM + row) * 4), __uint_as_float(r[j]));
      }
    }
    tc_fence_before();
  }
  // one barrier for the pair: writers release, readers acquire
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  // every epilogue warp finished its tcgen05.ld before it arrived: TMEM goes back now, not at the end of the tail (the CTA
  // then exits as soon as its last store is out; the next kernel's dependency wait returns on the LAST CTA's exit)
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_d, C::TMEM_COLS);
  }

  GP_PHASE(1);
  const int tok_base = tok0 + (int)rank * C::HALF;  // first token this CTA finishes
  if (epi) {
#pragma unroll
    for (int j = 0; j < C::HALF; ++j) own[j] = bf16_round(own[j] + red[j * GP_BM + row]);  // nn.Linear's single rounding
    // the stage ring is free (every MMA has completed): reuse its first bytes as this CTA's staging area
    if (MODE == GP_SWIGLU) {
      // act = bf16(bf16(silu(gate)) * up).  Every thread parks its row's bf16 values in shared memory as
      // [token][128 rows: 64 gate | 64 up]; then the 128 threads share the (token, 8 features) items: two 16-byte reads,
      // eight SiLU products, one 16-byte store each.
      constexpr int RS = GP_BM + 8;  // row stride in elements: 272 bytes, so that the 16-byte reads below spread over the banks
      bf16* sb = reinterpret_cast<bf16*>(smem_al);
#pragma unroll
      for (int j = 0; j < C::HALF; ++j) sb[j * RS + row] = __float2bfloat16_rn(own[j]);
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const int I = N / 2;
      const int et = threadIdx.x - 64;
      for (int item = et; item < C::HALF * 8; item += 128) {
        const int j = item >> 3, fg = item & 7;
        const int tok = tok_base + j;
        if (tok >= T) continue;
        float g[8], u[8], o[8];
        unpack8(*reinterpret_cast<const uint4*>(sb + j * RS + fg * 8), g);
        unpack8(*reinterpret_cast<const uint4*>(sb + j * RS + 64 + fg * 8), u);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = bf16_round(bf16_round(__fdiv_rn(g[e], 1.0f + expf(-g[e]))) * u[e]);
        *reinterpret_cast<uint4*>(out + (size_t)tok * I + m_tile * 64 + fg * 8) = pack8(o);
      }
    } else {
      // one head for HALF tokens: transpose through shared memory (padded rows), then SIXTEEN THREADS PER TOKEN, all
      // tokens at once on all six warps: thread (token, part) owns dims 8 part .. 8 part + 7; its RoPE partners are the same
      // dims +- 64, read from shared memory as well, so the only exchange between threads is the 128-dim sum of squares
      // (four shuffles).  Rounding points are those of qknorm_rope_kv_kernel (elementwise.cu); products of two bf16 values
      // use the packed bf16 multiply, which rounds the exact product once -- bit for bit what rounding the (exact) fp32
      // product gives -- at a third of the instructions: this tail is instruction-bound.
      constexpr int HD = 128;
      constexpr int RS = HD + 8;
      bf16* tq = reinterpret_cast<bf16*>(smem_al);
#pragma unroll
      for (int j = 0; j < C::HALF; ++j) tq[j * RS + row] = __float2bfloat16_rn(own[j]);
    }
  }
  __syncthreads();
  if (MODE == GP_QKV_ROPE) {
    {
      constexpr int HD = 128;
      constexpr int RS = HD + 8;
      const bf16* tq = reinterpret_cast<const bf16*>(smem_al);
      const int head = m_tile;
      const bool is_q = head < rp.nq, is_k = !is_q && head < rp.nq + rp.nkv;
      for (int item = threadIdx.x; item < C::HALF * 16; item += (int)blockDim.x) {  // whole warps: HALF * 16 and the block size are multiples of 32
        const int j = item >> 4, part = item & 15;
        const int tok = tok_base + j;
        const int d0 = part * 8, p0 = d0 ^ 64;
        uint4 raw = *reinterpret_cast<const uint4*>(tq + j * RS + d0);
        if (is_q || is_k) {
          const uint4 praw = *reinterpret_cast<const uint4*>(tq + j * RS + p0);
          float v[8], pv[8];
          unpack8(raw, v);
          unpack8(praw, pv);
          float ss = 0.f;
#pragma unroll
          for (int e = 0; e < 8; ++e) ss += v[e] * v[e];
          ss += __shfl_xor_sync(0xffffffffu, ss, 1);
          ss += __shfl_xor_sync(0xffffffffu, ss, 2);
          ss += __shfl_xor_sync(0xffffffffu, ss, 4);
          ss += __shfl_xor_sync(0xffffffffu, ss, 8);
          const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(ss / (float)HD + rp.eps));
          const bf16* gw = s_gain + (is_q ? 0 : 128);
          const uint4 w4 = *reinterpret_cast<const uint4*>(gw + d0), pw4 = *reinterpret_cast<const uint4*>(gw + p0);
          const uint4 co4 = *reinterpret_cast<const uint4*>(cs_rows + j * HD + (d0 & 63));
          const uint4 sn4 = *reinterpret_cast<const uint4*>(cs_rows + j * HD + 64 + (d0 & 63));
          const uint32_t sign = d0 < 64 ? 0x80008000u : 0u;  // dims 0..63: rot = -partner
          const uint32_t* w2 = &w4.x; const uint32_t* pw2 = &pw4.x; const uint32_t* co2 = &co4.x; const uint32_t* sn2 = &sn4.x;
          uint32_t o2[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t nx = pack_bf16x2(v[2 * e] * rstd, v[2 * e + 1] * rstd);      // bf16(x * rstd)
            const uint32_t pnx = pack_bf16x2(pv[2 * e] * rstd, pv[2 * e + 1] * rstd);
            __nv_bfloat162 n = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&w2[e]), *reinterpret_cast<const __nv_bfloat162*>(&nx));
            __nv_bfloat162 pn = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&pw2[e]), *reinterpret_cast<const __nv_bfloat162*>(&pnx));
            uint32_t rot = *reinterpret_cast<uint32_t*>(&pn) ^ sign;
            __nv_bfloat162 a = __hmul2(n, *reinterpret_cast<const __nv_bfloat162*>(&co2[e]));
            __nv_bfloat162 b = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&rot), *reinterpret_cast<const __nv_bfloat162*>(&sn2[e]));
            const uint32_t au = *reinterpret_cast<uint32_t*>(&a), bu = *reinterpret_cast<uint32_t*>(&b);
            o2[e] = pack_bf16x2(bf16lo(au) + bf16lo(bu), bf16hi(au) + bf16hi(bu));
          }
          raw = make_uint4(o2[0], o2[1], o2[2], o2[3]);
        }
        if (tok < T) {
          if (is_q) {
            *reinterpret_cast<uint4*>(rp.q_out + ((size_t)tok * rp.nq + head) * HD + d0) = raw;
          } else {
            const int slot = s_slot[j];
            if (slot >= 0) {
              const int kvh = is_k ? head - rp.nq : head - rp.nq - rp.nkv;
              const int blk = slot / rp.block_size, off = slot % rp.block_size;
              bf16* base = (is_k ? rp.k_pages : rp.v_pages) + (((size_t)blk * rp.nkv + kvh) * rp.block_size + off) * HD;
              *reinterpret_cast<uint4*>(base + kv_swizzled_chunk(off, part) * 8) = raw;  // 16-byte chunk `part` at its swizzled position
            }
          }
        }
      }
    }
  }
  GP_PHASE(2);
#ifdef B2_STEP_TRACE
  __syncthreads();  // the end stamp is the CTA's last thread's
  if (threadIdx.x == 0 && _stamp[0] != 0ull) {
    unsigned long long* buf = (unsigned long long*)_stamp[0];
    const unsigned long long slot = atomicAdd(buf, 1ull);
    if (slot < buf[1]) {
      unsigned long long* r = buf + 4 + slot * 4;
      r[0] = 101ull | ((unsigned long long)(blockIdx.x & 0xffffffu) << 8) |
             ((unsigned long long)((((unsigned)(N >> 6) & 0xfffu) | (((unsigned)(K >> 6) & 0x7ffu) << 12) | (tag ? (1u << 23) : 0u))) << 40);
      r[1] = phase_t[0]; r[2] = phase_t[1]; r[3] = phase_t[2];
    }
  }
#endif
  B2_STAMP_END_AUX(100, ((unsigned)(N >> 6) & 0xfffu) | (((unsigned)(K >> 6) & 0x7ffu) << 12) | (tag ? (1u << 23) : 0u));
}

template <int BN, int MODE>
int launch_pair(const CUtensorMap& tm_w, const bf16* x, bf16* out, int T, int N, int K, int up_row0, const GpRope& rp, const GpNorm& nm,
                int tag, cudaStream_t s) {
  using C = GpCfg<BN>;
  static B2PerDeviceOnce attr_once, carve_once;
  if (attr_once.need())
    B2_CUDA_OK(cudaFuncSetAttribute(linear_pair_kernel<BN, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
  b2_prefer_max_smem(linear_pair_kernel<BN, MODE>, carve_once);
  const int m_tiles = MODE == GP_SWIGLU ? (N / 2) / 64 : N / GP_BM;
  const int n_tiles = (T + BN - 1) / BN;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * m_tiles * n_tiles);
  cfg.blockDim = dim3(nm.ssq != nullptr ? GP_THREADS : 192);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = b2_pdl_enabled() ? 2 : 1;
  B2_CUDA_OK(cudaLaunchKernelEx(&cfg, linear_pair_kernel<BN, MODE>, tm_w, x, out, T, N, K, up_row0, m_tiles, rp, nm, tag));
  return 0;
}

template <int MODE>
int launch_pair_bn(const CUtensorMap& tm_w, const bf16* x, bf16* out, int T, int N, int K, int up_row0, const GpRope& rp, const GpNorm& nm,
                   int tag, cudaStream_t s) {
  // QKV: 32 heads x 2 K-halves are only 64 CTAs, and its tail (norm + RoPE of 128 dims per token) is instruction-bound:
  // two token tiles of 32 give 128 CTAs with 16 tokens each to finish (each weight tile is then read by two pairs; the
  // second read hits L2; one 64-token tile measured 1.4 % slower per step even with the 320-thread tail).  SwiGLU has 48 tiles
  // x 2 = 96 CTAs already.
  B2_REQUIRE(nm.ssq == nullptr || (nm.gain != nullptr && nm.pieces % 8 == 0 && nm.pieces <= 64 && nm.ld >= T), "pair kernels: bad normalise-on-load arguments");
  if (T <= 32 || MODE == GP_QKV_ROPE) return launch_pair<32, MODE>(tm_w, x, out, T, N, K, up_row0, rp, nm, tag, s);
  return launch_pair<64, MODE>(tm_w, x, out, T, N, K, up_row0, rp, nm, tag, s);
}

}  // namespace

bool pair_shapes_ok(int T, int K) { return T >= 1 && T <= 64 && K % (2 * GP_BK) == 0 && K / (2 * GP_BK) <= GP_MAX_KB; }

// gu [T][2I] is never materialised: act [T][I] = SwiGLU(x . Wgu^T).  tm_gu64: Wgu [2I][K] with a 64-row box.
int launch_gateup_swiglu(const CUtensorMap& tm_gu64, const bf16* x, bf16* act, int T, int I, int K, cudaStream_t s, int tag,
                         const PairNorm& norm) {
  B2_REQUIRE(pair_shapes_ok(T, K) && I % 64 == 0, "gateup_swiglu: T=%d K=%d I=%d not built (T <= 64, K <= 1024 in steps of 128)", T, K, I);
  GpRope none = {};
  const GpNorm nm = {norm.ssq, norm.gain, norm.pieces, norm.ld, norm.eps, norm.next_w};
  return launch_pair_bn<GP_SWIGLU>(tm_gu64, x, act, T, 2 * I, K, I, none, nm, tag, s);
}

// tm_qkv: Wqkv [(nq + 2 nkv) * 128][K] with a 128-row box (one head per tile).
int launch_qkv_rope(const CUtensorMap& tm_qkv, const bf16* x, const bf16* qw, const bf16* kw, const bf16* cos_sin, const int32_t* positions,
                    const int32_t* slots, bf16* q_out, bf16* k_pages, bf16* v_pages, int T, int nq, int nkv, int K, int block_size,
                    float eps, cudaStream_t s, int tag, const PairNorm& norm) {
  B2_REQUIRE(pair_shapes_ok(T, K) && qw != nullptr && kw != nullptr, "qkv_rope: T=%d K=%d not built, or no QK-norm gains", T, K);
  GpRope rp = {qw, kw, cos_sin, positions, slots, q_out, k_pages, v_pages, nq, nkv, block_size, eps};
  const GpNorm nm = {norm.ssq, norm.gain, norm.pieces, norm.ld, norm.eps, norm.next_w};
  return launch_pair_bn<GP_QKV_ROPE>(tm_qkv, x, nullptr, T, (nq + 2 * nkv) * 128, K, 0, rp, nm, tag, s);
}

}  // namespace b2
