// chain.cu -- the decode "chain" between two attention passes as ONE persistent kernel.
//
// Per layer, decode runs eight short, strictly dependent steps besides attention:
//   0 O-proj GEMM  1 residual add + RMSNorm  2 gate-up GEMM  3 SwiGLU gate
//   4 down GEMM    5 residual add + RMSNorm  6 QKV GEMM (next layer)  7 QK-norm + RoPE + KV write
// (HF:models/qwen3/modeling_qwen3.py:305-334 and :252-291).  As separate launches each step costs a
// kernel boundary of a few microseconds although it moves only a few MB, and the boundaries -- not
// HBM -- bounded the step.  Here one CTA per SM stays resident for the whole chain:
//   * GEMM steps use the tcgen05 pipeline of gemm_tc.cu (TMA producer warp, single-thread MMA issuer
//     with a double-buffered TMEM accumulator, four epilogue warps writing split-K partials);
//   * the element-wise steps run on the four epilogue warps and consume the partials in split order
//     (deterministic), with exactly the rounding points of elementwise.cu;
//   * steps are separated by a grid-wide barrier (one atomic + spin per CTA) instead of a launch;
//   * the producer warp pulls the NEXT GEMM step's first weight tiles into shared memory before it
//     joins the barrier -- weights never depend on the barrier -- so the weight stream of step i+2
//     hides behind step i+1.
// All of shared memory goes to the stage ring (8 stages at 64 tokens): a whole weight tile of the next
// GEMM step fits ahead of the barrier.
#include "kernels.h"

namespace b2 {

constexpr int CH_BM = 128;
constexpr int CH_BK = 64;
constexpr int CH_W_BYTES = CH_BM * CH_BK * 2;
constexpr int CH_SMEM_BUDGET = 192 * 1024;
constexpr int CH_THREADS = 192;
constexpr int CH_MAXV = 4;  // 16-byte vectors of a hidden row per epilogue thread (hidden <= 4096)

template <int BN>
struct ChCfg {
  static constexpr int X_BYTES = BN * CH_BK * 2;
  static constexpr int STAGE_BYTES = CH_W_BYTES + X_BYTES;
  static constexpr int STAGES = CH_SMEM_BUDGET / STAGE_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
};

struct ChainMaps {
  CUtensorMap w[4];  // Wo, Wgu, Wd, Wqkv(next layer)   box 64 x 128
  CUtensorMap x[4];  // attn, x, act, x                 box 64 x BN
};

// Optional trace (tools/chain_timeline.py): thread 0 of every CTA stamps %globaltimer when it enters
// each step (slot = step), when its CTA has finished the step's work (16 + step) and when it leaves the
// kernel: trace[blockIdx.x * 32 + slot].
__device__ unsigned long long* g_chain_trace = nullptr;
__device__ __forceinline__ void chain_stamp(int slot) {
  if (g_chain_trace != nullptr && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_chain_trace[blockIdx.x * 32 + slot] = t;
  }
}
// same, by whichever thread calls it (sub-step stamps of the MMA and epilogue warps)
__device__ __forceinline__ void chain_stamp_any(int slot) {
  if (g_chain_trace != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_chain_trace[blockIdx.x * 32 + slot] = t;
  }
}
int chain_set_trace(unsigned long long* dev_buf) {
  B2_CUDA_OK(cudaMemcpyToSymbol(g_chain_trace, &dev_buf, sizeof(dev_buf)));
  return 0;
}

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Grid-wide barrier over the resident CTAs (gridDim.x <= number of SMs, one CTA per SM).
// One 64-bit arrival counter that only ever grows: every launch uses the same grid size G and passes
// whole barriers, so the counter is a multiple of G whenever no barrier is in flight.  A CTA finds the
// base of its launch by rounding what it reads at its start down to a multiple of G -- exact even if
// faster CTAs have already arrived at the first barrier, because none of them can get past it (and so
// the counter cannot reach base + G) before this CTA arrives too.  The k-th barrier completes at
// base + k*G.  Arrival is a fire-and-forget release reduction after a cumulative fence (the CTA barrier
// orders every thread's writes before thread 0's fence); waiting is an acquire-load spin: about one L2
// round trip after the last arrival.
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void chain_grid_sync(unsigned long long* bar, unsigned long long target, int step) {
  __syncthreads();
  chain_stamp(16 + step);
  if (threadIdx.x == 0) {
    __threadfence();
    asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(bar) : "memory");
    while (ld_acquire_u64(bar) < target) {
    }
    fence_proxy_async_global();  // thread 0 is also the TMA producer: order its coming async-proxy reads
  }
  __syncthreads();
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }  // the four epilogue warps

template <int BN>
__global__ void __launch_bounds__(CH_THREADS, 1)
decode_chain_kernel(const __grid_constant__ ChainMaps maps, const ChainParams p) {
  using C = ChCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[C::STAGES];
  __shared__ __align__(8) uint64_t empty_bar[C::STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float red[4];

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = p.T;
  const int n_tiles = (T + BN - 1) / BN;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tmem_full_bar[a]), 1);
      mbar_init(smem_u32(&tmem_empty_bar[a]), 128);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&tmem_slot), C::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  // per-role running counters (the smem ring and the TMEM double buffer run on across steps)
  int it = 0;   // producer: k-blocks issued so far / MMA: k-blocks consumed so far
  int lt = 0;   // MMA and epilogue: tiles so far
  int pre = 0;  // producer: leading k-blocks of the coming GEMM step whose W tile is already in flight

  auto tile_geom = [&](int gi, int tile, int& m_tile, int& n_tile, int& kb0, int& kb1) {
    const int m_tiles = (p.g[gi].N + CH_BM - 1) / CH_BM;
    m_tile = tile % m_tiles;
    n_tile = (tile / m_tiles) % n_tiles;
    const int split = tile / (m_tiles * n_tiles);
    kb0 = split * p.g[gi].kps;
    kb1 = min(p.g[gi].K / CH_BK, kb0 + p.g[gi].kps);
  };
  auto total_tiles = [&](int gi) { return ((p.g[gi].N + CH_BM - 1) / CH_BM) * n_tiles * p.g[gi].nsplit; };
  // producer only: put the W tiles of the first k-blocks of GEMM step gi's first tile in flight
  auto preissue_w = [&](int gi) {
    pre = 0;
    const int tile = blockIdx.x;
    if (tile >= total_tiles(gi)) return;
    int m_tile, n_tile, kb0, kb1;
    tile_geom(gi, tile, m_tile, n_tile, kb0, kb1);
    const int npre = min(kb1 - kb0, C::STAGES);
    for (int i = 0; i < npre; ++i) {
      const int s = (it + i) % C::STAGES;
      mbar_wait(smem_u32(&empty_bar[s]), (((it + i) / C::STAGES) & 1) ^ 1);
      const uint32_t fb = smem_u32(&full_bar[s]);
      mbar_arrive_expect_tx(fb, C::STAGE_BYTES);
      tma_load_2d(smem_base + s * C::STAGE_BYTES, &maps.w[gi], (kb0 + i) * CH_BK, m_tile * CH_BM, fb);
    }
    pre = npre;
  };

  const bool is_producer = (warp == 0 && lane == 0);
  const bool is_mma = (warp == 1 && lane == 0);
  const bool is_epi = warp >= 2;
  const int et = threadIdx.x - 64;  // epilogue thread index 0..127

  // ---- before the dependency wait: weights only ----
  if (is_producer) {
    for (int i = 0; i < 4; ++i) {
      tma_prefetch_desc(&maps.w[i]);
      tma_prefetch_desc(&maps.x[i]);
    }
    for (int ph = p.phase_begin; ph < p.phase_end; ++ph)
      if ((ph & 1) == 0) {
        preissue_w(ph >> 1);
        break;
      }
  }
  chain_stamp(8);
  pdl_wait();
  if (threadIdx.x == 0) pdl_launch_dependents();
  chain_stamp(9);
  unsigned long long bar_target = 0;
  if (threadIdx.x == 0) bar_target = (ld_acquire_u64(p.gbar) / gridDim.x) * gridDim.x;

  for (int ph = p.phase_begin; ph < p.phase_end; ++ph) {
    chain_stamp(ph);
    if ((ph & 1) == 0) {
      // =============================== GEMM step ===============================
      const int gi = ph >> 1;
      const int ntot = total_tiles(gi);
      if (is_producer) {
        bool first = true;
        for (int tile = blockIdx.x; tile < ntot; tile += gridDim.x) {
          int m_tile, n_tile, kb0, kb1;
          tile_geom(gi, tile, m_tile, n_tile, kb0, kb1);
          int kb = kb0;
          if (first) {  // W of the first `pre` k-blocks is already in flight: add the X tiles
            for (int i = 0; i < pre; ++i, ++kb, ++it) {
              const int s = it % C::STAGES;
              tma_load_2d(smem_base + s * C::STAGE_BYTES + CH_W_BYTES, &maps.x[gi], kb * CH_BK, n_tile * BN, smem_u32(&full_bar[s]));
            }
            first = false;
            pre = 0;
          }
          for (; kb < kb1; ++kb, ++it) {
            const int s = it % C::STAGES;
            mbar_wait(smem_u32(&empty_bar[s]), ((it / C::STAGES) & 1) ^ 1);
            const uint32_t fb = smem_u32(&full_bar[s]);
            mbar_arrive_expect_tx(fb, C::STAGE_BYTES);
            const uint32_t ws = smem_base + s * C::STAGE_BYTES;
            tma_load_2d(ws, &maps.w[gi], kb * CH_BK, m_tile * CH_BM, fb);
            tma_load_2d(ws + CH_W_BYTES, &maps.x[gi], kb * CH_BK, n_tile * BN, fb);
          }
        }
        // weights of the next GEMM step do not depend on the barrier: start them now
        for (int nx = ph + 2; nx < p.phase_end; nx += 2) {
          preissue_w(nx >> 1);
          break;
        }
      } else if (is_mma) {
        constexpr uint32_t idesc = umma_idesc_bf16(CH_BM, BN);
        for (int tile = blockIdx.x; tile < ntot; tile += gridDim.x, ++lt) {
          int m_tile, n_tile, kb0, kb1;
          tile_geom(gi, tile, m_tile, n_tile, kb0, kb1);
          const int as = lt & 1;
          mbar_wait(smem_u32(&tmem_empty_bar[as]), ((lt >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
          for (int kb = kb0; kb < kb1; ++kb, ++it) {
            const int s = it % C::STAGES;
            mbar_wait(smem_u32(&full_bar[s]), (it / C::STAGES) & 1);
            tc_fence_after();
            if (ph == 0 && kb == kb0) chain_stamp_any(24);
            const uint32_t ws = smem_base + s * C::STAGE_BYTES;
            const uint64_t da = umma_desc_sw128(ws);
            const uint64_t db = umma_desc_sw128(ws + CH_W_BYTES);
#pragma unroll
            for (int k = 0; k < CH_BK / 16; ++k)
              tc_mma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            tc_commit(smem_u32(&empty_bar[s]));
          }
          tc_commit(smem_u32(&tmem_full_bar[as]));
          if (ph == 0) chain_stamp_any(25);
        }
      } else if (is_epi) {
        const int q = warp & 3;
        const int Nn = p.g[gi].N;
        for (int tile = blockIdx.x; tile < ntot; tile += gridDim.x, ++lt) {
          int m_tile, n_tile, kb0, kb1;
          tile_geom(gi, tile, m_tile, n_tile, kb0, kb1);
          const int split = tile / (((Nn + CH_BM - 1) / CH_BM) * n_tiles);
          const int as = lt & 1;
          mbar_wait(smem_u32(&tmem_full_bar[as]), (lt >> 1) & 1);
          tc_fence_after();
          if (ph == 0 && et == 0) chain_stamp_any(26);
          const int row = m_tile * CH_BM + q * 32 + lane;
          const bool row_ok = row < Nn;
          const int tok0 = n_tile * BN;
          const bool full = row_ok && tok0 + BN <= T;
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
          float* dst = p.partial + ((size_t)split * T + tok0) * Nn + row;
          if (BN == 16) {
            uint32_t r[16];
            tmem_ld16(tbase, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (full || (row_ok && tok0 + j < T)) dst[(size_t)j * Nn] = __uint_as_float(r[j]);
          } else {
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
              uint32_t r[32];
              tmem_ld32(tbase + (uint32_t)c0, r);
              tmem_ld_wait();
              float* d = dst + (size_t)c0 * Nn;
              if (full) {
#pragma unroll
                for (int j = 0; j < 32; ++j) d[(size_t)j * Nn] = __uint_as_float(r[j]);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (row_ok && tok0 + c0 + j < T) d[(size_t)j * Nn] = __uint_as_float(r[j]);
              }
            }
          }
          tc_fence_before();
          mbar_arrive(smem_u32(&tmem_empty_bar[as]));
          if (ph == 0 && et == 0) chain_stamp_any(27);
        }
      }
    } else if (is_epi) {
      // =========================== element-wise step ===========================
      if (ph == 1 || ph == 5) {
        // residual add (HF:...:327,333) + RMSNorm (HF:...:59-64); rounding points as add_rmsnorm_kernel
        const int gi = ph == 1 ? 0 : 2;
        const bool has_delta = (ph == 1) || p.norm_b_has_delta;
        const int ns = p.g[gi].nsplit;
        const bf16* w = ph == 1 ? p.norm_w_a : p.norm_w_b;
        const int H = p.H, nvec = H / 8;
        for (int r = blockIdx.x; r < T; r += gridDim.x) {
          float v[CH_MAXV][8];
          float ss = 0.f;
#pragma unroll
          for (int k = 0; k < CH_MAXV; ++k) {
            const int i = et + k * 128;
            if (i < nvec) {
              unpack8(__ldcg(reinterpret_cast<const uint4*>(p.h + (size_t)r * H) + i), v[k]);
              if (has_delta) {
                float d[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) d[e] = 0.f;
                const float* pp = p.partial + (size_t)r * H + (size_t)i * 8;
                for (int s0 = 0; s0 < ns; s0 += 16) {
                  float4 a[16], b[16];
#pragma unroll
                  for (int j = 0; j < 16; ++j)
                    if (s0 + j < ns) {
                      a[j] = __ldcg(reinterpret_cast<const float4*>(pp + (size_t)(s0 + j) * T * H));
                      b[j] = __ldcg(reinterpret_cast<const float4*>(pp + (size_t)(s0 + j) * T * H) + 1);
                    }
#pragma unroll
                  for (int j = 0; j < 16; ++j)
                    if (s0 + j < ns) {
                      d[0] += a[j].x; d[1] += a[j].y; d[2] += a[j].z; d[3] += a[j].w;
                      d[4] += b[j].x; d[5] += b[j].y; d[6] += b[j].z; d[7] += b[j].w;
                    }
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) v[k][e] = bf16_round(v[k][e] + bf16_round(d[e]));
                reinterpret_cast<uint4*>(p.h + (size_t)r * H)[i] = pack8(v[k]);
              }
#pragma unroll
              for (int e = 0; e < 8; ++e) ss += v[k][e] * v[k][e];
            }
          }
          ss = warp_sum(ss);
          if (ph == 1 && et == 0) chain_stamp_any(28);
          if (lane == 0) red[warp - 2] = ss;
          epi_bar();
          if (ph == 1 && et == 0) chain_stamp_any(29);
          const float tot = ((red[0] + red[1]) + red[2]) + red[3];
          const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(tot / (float)H + p.eps));
#pragma unroll
          for (int k = 0; k < CH_MAXV; ++k) {
            const int i = et + k * 128;
            if (i < nvec) {
              float wv[8], o[8];
              unpack8(reinterpret_cast<const uint4*>(w)[i], wv);
#pragma unroll
              for (int e = 0; e < 8; ++e) o[e] = bf16_round(wv[e] * bf16_round(v[k][e] * rstd));
              reinterpret_cast<uint4*>(p.x + (size_t)r * H)[i] = pack8(o);
            }
          }
          if (ph == 1 && et == 0) chain_stamp_any(30);
          epi_bar();  // red[] is reused by the next row
        }
        fence_proxy_async_global();  // x is read through TMA by the next GEMM step
        if (ph == 1 && et == 0) chain_stamp_any(31);
      } else if (ph == 3) {
        // SwiGLU gate (HF:...:81-83); rounding points as silu_mul_kernel
        const int I = p.I, ns = p.g[1].nsplit;
        const size_t nvec = (size_t)T * I / 8, sstride = (size_t)T * 2 * I;
        for (size_t idx = (size_t)blockIdx.x * 128 + et; idx < nvec; idx += (size_t)gridDim.x * 128) {
          const int t = (int)(idx / (I / 8));
          const int i = (int)(idx % (I / 8)) * 8;
          float g[8], u[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) g[e] = u[e] = 0.f;
          const float* pg = p.partial + (size_t)t * 2 * I + i;
          const float* pu = pg + I;
          for (int s0 = 0; s0 < ns; s0 += 4) {
            float4 a[4], b[4], c[4], d[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (s0 + j < ns) {
                a[j] = __ldcg(reinterpret_cast<const float4*>(pg + (s0 + j) * sstride));
                b[j] = __ldcg(reinterpret_cast<const float4*>(pg + (s0 + j) * sstride) + 1);
                c[j] = __ldcg(reinterpret_cast<const float4*>(pu + (s0 + j) * sstride));
                d[j] = __ldcg(reinterpret_cast<const float4*>(pu + (s0 + j) * sstride) + 1);
              }
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (s0 + j < ns) {
                g[0] += a[j].x; g[1] += a[j].y; g[2] += a[j].z; g[3] += a[j].w; g[4] += b[j].x; g[5] += b[j].y; g[6] += b[j].z; g[7] += b[j].w;
                u[0] += c[j].x; u[1] += c[j].y; u[2] += c[j].z; u[3] += c[j].w; u[4] += d[j].x; u[5] += d[j].y; u[6] += d[j].z; u[7] += d[j].w;
              }
          }
          float o[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float ge = bf16_round(g[e]), ue = bf16_round(u[e]);
            const float sl = bf16_round(__fdiv_rn(ge, 1.0f + expf(-ge)));
            o[e] = bf16_round(sl * ue);
          }
          *reinterpret_cast<uint4*>(p.act + (size_t)t * I + i) = pack8(o);
        }
        fence_proxy_async_global();  // act is read through TMA by the down-projection step
      } else if (ph == 7) {
        // QK-norm + RoPE + KV page write (HF:...:248-249,263-264,151-181,270-271); as qknorm_rope_kv_kernel
        constexpr int HD = 128;
        const int nheads = p.nq + 2 * p.nkv, ns = p.g[3].nsplit;
        const size_t sstride = (size_t)T * nheads * HD;
        for (int item = blockIdx.x * 4 + (warp - 2); item < T * nheads; item += gridDim.x * 4) {
          const int t = item / nheads, head = item % nheads;
          const size_t row = (size_t)t * nheads * HD + (size_t)head * HD + lane * 4;
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int s0 = 0; s0 < ns; s0 += 8) {
            float4 a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (s0 + j < ns) a[j] = __ldcg(reinterpret_cast<const float4*>(p.partial + row + (s0 + j) * sstride));
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (s0 + j < ns) { acc.x += a[j].x; acc.y += a[j].y; acc.z += a[j].z; acc.w += a[j].w; }
          }
          float v[4] = {bf16_round(acc.x), bf16_round(acc.y), bf16_round(acc.z), bf16_round(acc.w)};
          const bool is_q = head < p.nq, is_k = !is_q && head < p.nq + p.nkv;
          if (is_q || is_k) {
            float ss = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
            ss = warp_sum(ss);
            const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(ss / (float)HD + p.eps));
            const uint2 wraw = *reinterpret_cast<const uint2*>((is_q ? p.qw : p.kw) + lane * 4);
            const float wv[4] = {bf16lo(wraw.x), bf16hi(wraw.x), bf16lo(wraw.y), bf16hi(wraw.y)};
            const int pos = max(p.positions[t], 0);
            const bf16* cs = p.cos_sin + (size_t)pos * HD;
            const int ci = (lane * 4) & 63;
            const uint2 craw = *reinterpret_cast<const uint2*>(cs + ci);
            const uint2 sraw = *reinterpret_cast<const uint2*>(cs + 64 + ci);
            const float c[4] = {bf16lo(craw.x), bf16hi(craw.x), bf16lo(craw.y), bf16hi(craw.y)};
            const float sn[4] = {bf16lo(sraw.x), bf16hi(sraw.x), bf16lo(sraw.y), bf16hi(sraw.y)};
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] = bf16_round(wv[e] * bf16_round(v[e] * rstd));
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float partner = __shfl_xor_sync(0xffffffffu, v[e], 16);
              const float rot = (lane < 16) ? -partner : partner;
              v[e] = bf16_round(bf16_round(v[e] * c[e]) + bf16_round(rot * sn[e]));
            }
          }
          uint2 o;
          o.x = pack_bf16x2(v[0], v[1]);
          o.y = pack_bf16x2(v[2], v[3]);
          if (is_q) {
            *reinterpret_cast<uint2*>(p.qrot + ((size_t)t * p.nq + head) * HD + lane * 4) = o;
          } else {
            const int slot = p.slots[t];
            if (slot >= 0) {
              const int kvh = is_k ? head - p.nq : head - p.nq - p.nkv;
              const int blk = slot / p.block_size, off = slot % p.block_size;
              bf16* base = is_k ? p.k_pages : p.v_pages;
              const int chunk = kv_swizzled_chunk(off, lane >> 1);
              *reinterpret_cast<uint2*>(base + (((size_t)blk * p.nkv + kvh) * p.block_size + off) * HD + chunk * 8 + (lane & 1) * 4) = o;
            }
          }
        }
      }
    }
    if (ph + 1 < p.phase_end) {
      bar_target += gridDim.x;
      chain_grid_sync(p.gbar, bar_target, ph);
    }
  }

  tc_fence_before();
  __syncthreads();
  chain_stamp(10);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

template <int BN>
static int launch_chain_bn(const ChainMaps& maps, const ChainParams& p, int grid, cudaStream_t s) {
  using C = ChCfg<BN>;
  static B2PerDeviceOnce attr_once;
  if (attr_once.need())
    B2_CUDA_OK(cudaFuncSetAttribute(decode_chain_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
  B2_CUDA_OK(b2_launch(decode_chain_kernel<BN>, dim3(grid), dim3(CH_THREADS), (size_t)C::SMEM_BYTES, s, maps, p));
  return 0;
}

int launch_decode_chain(const CUtensorMap* w_maps, const CUtensorMap* x_maps, const ChainParams& p, cudaStream_t s) {
  B2_REQUIRE(p.T >= 1 && p.phase_begin >= 0 && p.phase_end <= 8 && p.phase_begin < p.phase_end, "decode_chain: bad phase range");
  B2_REQUIRE(p.H % 8 == 0 && p.H / 8 <= 128 * CH_MAXV, "decode_chain: hidden %d not supported", p.H);
  ChainMaps maps;
  for (int i = 0; i < 4; ++i) {
    maps.w[i] = w_maps[i];
    maps.x[i] = x_maps[i];
  }
  static int sms = device_num_sms();
  switch (p.bn) {
    case 16: return launch_chain_bn<16>(maps, p, sms, s);
    case 32: return launch_chain_bn<32>(maps, p, sms, s);
    case 64: return launch_chain_bn<64>(maps, p, sms, s);
    case 128: return launch_chain_bn<128>(maps, p, sms, s);
    case 256: return launch_chain_bn<256>(maps, p, sms, s);
  }
  b2_set_error("decode_chain: bad token tile %d", p.bn);
  return -22;
}

}  // namespace b2
