// gemm_tc.cu -- tcgen05 / TMEM GEMM for the projections (HF nn.Linear,
// models/qwen3/modeling_qwen3.py:236-247,81-83,451): y[T][N] = x[T][K] . W[N][K]^T.
//
// Layout choice ("swap-AB"): the WEIGHT rows are the MMA M dimension (128 per CTA) and the
// TOKENS are the MMA N dimension (16..256), so a decode batch of 1..64 tokens wastes no
// tensor-core rows and the weight stream -- the only large operand -- is what TMA pipelines.
//   warp 0   : TMA producer  (cp.async.bulk.tensor, 128B-swizzled 64-column k-blocks)
//   warp 1   : TMEM allocator + single-thread tcgen05.mma issuer (accumulator in TMEM)
//   warps 2-5: epilogue -- tcgen05.ld the 128 x BN fp32 accumulator, then
//                TC_EPI_BF16        round to bf16, store y[T][N]
//                TC_EPI_F32_PARTIAL store fp32 split-K partial [split][T][N] (summed in split
//                                   order by the consumer kernel -> deterministic)
//                TC_EPI_ARGMAX      round to bf16, fused arg-max over the weight rows (lowest index on
//                                   ties) via one 64-bit atomicMax per token per tile
//                TC_EPI_ARGMAX_ROWS lm_head: operands trade places (tokens are MMA-M, vocabulary rows
//                                   are MMA-N), so one epilogue thread owns one token and scans its
//                                   logits with a running (value, -index) maximum in registers across
//                                   all of the CTA's tiles: no shuffles, one atomicMax per token per CTA
// Split-K (gridDim.z) exists because N/128 row tiles alone cannot fill 148 SMs for the
// 1024-row projections; HBM bandwidth needs every SM streaming.
#include "kernels.h"

#include <mutex>

namespace b2 {

B2_STEP_TRACE_DEFINE(step_trace_set_gemm)

// Optional phase trace (tools/gemm_timeline.py): when a buffer is registered, every CTA stamps
// %globaltimer at its phase boundaries into trace[blockIdx.x * 8 + slot].
__device__ unsigned long long* g_tc_trace = nullptr;
__device__ __forceinline__ void tc_stamp(int slot) {
  if (g_tc_trace != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_tc_trace[blockIdx.x * 8 + slot] = t;
  }
}

constexpr int TC_BM = 128;
constexpr int TC_BK = 64;
constexpr int TC_W_BYTES = TC_BM * TC_BK * 2;  // 16 KB per stage

// SMALL: a ring short enough (<= 96 KB) for the CTA to be resident on an SM BESIDE a decode-attention CTA (128 KB +
// the 1 KB the system reserves per CTA, of 228 KB), so the O projection -- the kernel that follows attention --
// can start and prefetch its weight tiles while attention's tail is still streaming.
template <int BN, int SMALL = 0>
struct TcCfg {
  static constexpr int X_BYTES = BN * TC_BK * 2;
  static constexpr int STAGE_BYTES = TC_W_BYTES + X_BYTES;
  static constexpr int STAGES = SMALL ? (BN <= 32 ? 4 : 3) : ((BN <= 64) ? 8 : (BN == 128 ? 6 : 4));
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;  // two accumulator stages
};

__device__ __forceinline__ unsigned long long tc_argmax_pack(float v, int idx) {
  uint32_t b = __float_as_uint(v);
  uint32_t key = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  return ((unsigned long long)key << 32) | (unsigned long long)(0xffffffffu - (uint32_t)idx);
}

// x_row0: first row of X this launch reads (a micro-batch of a larger activation buffer; `out` is already offset).
// Persistent: gridDim.x CTAs walk the (m tile, token tile, k split) work list with stride
// gridDim.x.  The smem stage ring runs on across tiles, and the accumulator is double buffered
// in TMEM so the epilogue of tile i overlaps the TMA/MMA main loop of tile i+1.
// SMALL also caps the kernel at 56 registers (launch bound of 6 CTAs per SM): an SM's register file is four banks of
// 16 K, one per warp scheduler, and a decode-attention CTA holds 2 warps x 200 registers in each; this kernel's six
// warps put up to two on a scheduler, so it fits beside that CTA only if 2 x 32 x (200 + R) <= 16384, R <= 56.
// XCP (decode shapes: one tile per CTA, its k-blocks all fit the stage ring): the X tiles do not come through the
// tensor map but through cp.async issued by the four epilogue warps, which are idle until the accumulator is ready.
// Measured in the captured step (tools/step_timeline.py): with TMA the first X tile lands 0.6-0.8 us after the
// dependency wait and every further one 0.35 us later (the tiled copy of a 64-row box is 64 requests of 128 bytes,
// and dozens of CTAs ask for the same rows), so a CTA with 4-6 k-blocks issued its last MMA 1.8-2.4 us after the
// wait; 128 threads x 16-byte LDGSTS have the whole slice in flight at once.
template <int BN, int EPI, int SMALL = 0, int XCP = 0>
__global__ void __launch_bounds__(192, SMALL ? 6 : 1)
linear_tc_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                 void* out, unsigned long long* __restrict__ argmax_packed, int T, int N, int K,
                 int kb_per_split, int m_tiles, int n_tiles, int nsplit, const KvPrefetch pf, int x_row0,
                 const bf16* __restrict__ x_raw, const bf16* resid) {
  using C = TcCfg<BN, SMALL>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[C::STAGES];
  __shared__ __align__(8) uint64_t empty_bar[C::STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_slot;
  __shared__ unsigned long long amax_red[EPI == TC_EPI_ARGMAX ? 4 * BN : 1];
  (void)amax_red;

  if (threadIdx.x == 0) tc_stamp(0);  // CTA started
  B2_STAMP_BEGIN();
#ifdef B2_STEP_TRACE
  __shared__ unsigned long long phase_t[3];  // first operands landed, all MMAs of the first tile issued, accumulator visible
#define TC_PHASE(i) do { if (_stamp[0] != 0ull) phase_t[i] = b2_globaltimer(); } while (0)
#else
#define TC_PHASE(i) do { } while (0)
#endif
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb_total = K / TC_BK;
  const int total_tiles = m_tiles * n_tiles * nsplit;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), XCP ? 1 + 128 : 1);  // XCP: + one arrival per epilogue thread (its share of X)
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
  if (threadIdx.x == 0) tc_stamp(1);  // barriers + TMEM ready

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmap_w);
      tma_prefetch_desc(&tmap_x);
      int it = 0;
      bool waited = false;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m_tile = tile % m_tiles;
        const int n_tile = (tile / m_tiles) % n_tiles;
        const int split = tile / (m_tiles * n_tiles);
        const int kb0 = split * kb_per_split;
        const int kb1 = min(kb_total, kb0 + kb_per_split);
        if (!waited) {
          // The weight stream does not depend on the previous kernel: start it for the first
          // stages of the first tile, THEN wait for the producer of X (programmatic dependent launch).
          // (TC_EPI_ARGMAX_ROWS trades the operands: there the SECOND map holds the weights.)
          constexpr bool kWeightsFirst = (EPI != TC_EPI_ARGMAX_ROWS);
          const int npre = min(kb1 - kb0, C::STAGES);
          if (XCP) {  // one tile, all of its k-blocks resident: the weight tiles are all this thread loads
            // ONE barrier for the whole tile (all weight bytes + the 128 X-loader arrivals): a wait on an mbarrier that
            // has already completed still costs ~0.1 us, and the MMA thread paid it once per k-block
            const uint32_t fb = smem_u32(&full_bar[0]);
            mbar_arrive_expect_tx(fb, (uint32_t)npre * TC_W_BYTES);
            for (int i = 0; i < npre; ++i)
              tma_load_2d(smem_base + i * C::STAGE_BYTES, &tmap_w, (kb0 + i) * TC_BK, m_tile * TC_BM, fb);
            pdl_wait();
            pdl_launch_dependents();
            tc_stamp(2);
            B2_STAMP_WAIT();
            waited = true;
            continue;
          }
          for (int i = 0; i < npre; ++i) {
            const uint32_t fb = smem_u32(&full_bar[i]);
            mbar_arrive_expect_tx(fb, C::STAGE_BYTES);
            if (kWeightsFirst) tma_load_2d(smem_base + i * C::STAGE_BYTES, &tmap_w, (kb0 + i) * TC_BK, m_tile * TC_BM, fb);
            else tma_load_2d(smem_base + i * C::STAGE_BYTES + TC_W_BYTES, &tmap_x, (kb0 + i) * TC_BK, n_tile * BN + x_row0, fb);
          }
          pdl_wait();
          pdl_launch_dependents();
          tc_stamp(2);  // dependency satisfied, the activation tiles go out
          B2_STAMP_WAIT();
          waited = true;
          for (int i = 0; i < npre; ++i) {
            const uint32_t fb = smem_u32(&full_bar[i]);
            if (kWeightsFirst) tma_load_2d(smem_base + i * C::STAGE_BYTES + TC_W_BYTES, &tmap_x, (kb0 + i) * TC_BK, n_tile * BN + x_row0, fb);
            else tma_load_2d(smem_base + i * C::STAGE_BYTES, &tmap_w, (kb0 + i) * TC_BK, m_tile * TC_BM, fb);
          }
          it = npre;
          for (int kb = kb0 + npre; kb < kb1; ++kb, ++it) {
            const int s = it % C::STAGES;
            mbar_wait(smem_u32(&empty_bar[s]), ((it / C::STAGES) & 1) ^ 1);
            const uint32_t fb = smem_u32(&full_bar[s]);
            mbar_arrive_expect_tx(fb, C::STAGE_BYTES);
            const uint32_t ws = smem_base + s * C::STAGE_BYTES;
            tma_load_2d(ws, &tmap_w, kb * TC_BK, m_tile * TC_BM, fb);
            tma_load_2d(ws + TC_W_BYTES, &tmap_x, kb * TC_BK, n_tile * BN + x_row0, fb);
          }
          continue;
        }
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % C::STAGES;
          mbar_wait(smem_u32(&empty_bar[s]), ((it / C::STAGES) & 1) ^ 1);
          const uint32_t fb = smem_u32(&full_bar[s]);
          mbar_arrive_expect_tx(fb, C::STAGE_BYTES);
          const uint32_t ws = smem_base + s * C::STAGE_BYTES;
          tma_load_2d(ws, &tmap_w, kb * TC_BK, m_tile * TC_BM, fb);
          tma_load_2d(ws + TC_W_BYTES, &tmap_x, kb * TC_BK, n_tile * BN + x_row0, fb);
        }
      }
      if (!waited) {
        pdl_wait();
        pdl_launch_dependents();
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(TC_BM, BN);
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
        const int split = tile / (m_tiles * n_tiles);
        const int kb0 = split * kb_per_split;
        const int kb1 = min(kb_total, kb0 + kb_per_split);
        const int as = lt & 1;
        mbar_wait(smem_u32(&tmem_empty_bar[as]), ((lt >> 1) & 1) ^ 1);  // epilogue drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
        if (XCP) {
          // one tile, every k-block has its own stage: wait for all of them, then issue the whole MMA chain back to back
          mbar_wait(smem_u32(&full_bar[0]), 0);
          tc_stamp(3);
          TC_PHASE(0);
          tc_fence_after();
          for (int kb = kb0; kb < kb1; ++kb) {
            const uint32_t ws = smem_base + (kb - kb0) * C::STAGE_BYTES;
            const uint64_t da = umma_desc_sw128(ws);
            const uint64_t db = umma_desc_sw128(ws + TC_W_BYTES);
#pragma unroll
            for (int k = 0; k < TC_BK / 16; ++k)
              tc_mma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tc_commit(smem_u32(&tmem_full_bar[as]));
          tc_stamp(4);
          TC_PHASE(1);
          continue;
        }
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % C::STAGES;
          mbar_wait(smem_u32(&full_bar[s]), (it / C::STAGES) & 1);
          if (it == 0) { tc_stamp(3); TC_PHASE(0); }  // first operands landed
          tc_fence_after();
          const uint32_t ws = smem_base + s * C::STAGE_BYTES;
          const uint64_t da = umma_desc_sw128(ws);
          const uint64_t db = umma_desc_sw128(ws + TC_W_BYTES);
#pragma unroll
          for (int k = 0; k < TC_BK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle row: +2 in the
            // (address >> 4) field of both descriptors
            tc_mma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          if (!XCP) tc_commit(smem_u32(&empty_bar[s]));  // frees the smem stage once these MMAs retire (XCP: no stage is reused)
        }
        tc_commit(smem_u32(&tmem_full_bar[as]));  // accumulator complete
        if (lt == 0) { tc_stamp(4); TC_PHASE(1); }  // all MMAs of the first tile issued
      }
    }
  } else {
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    unsigned long long row_best = 0ull;  // TC_EPI_ARGMAX_ROWS: running arg-max of this thread's token
    int row_best_tok = -1;
    pdl_wait();
    if (XCP) {
      // this CTA's only tile: X rows [tok0, tok0 + BN) x its k-blocks, 16 bytes per cp.async, into the 128-byte-swizzled
      // layout the MMA descriptors expect (chunk c of row r at chunk position c ^ (r & 7); the X tile of a stage is
      // 1024-byte aligned).
      const int tile = blockIdx.x;
      const int n_tile = (tile / m_tiles) % n_tiles;
      const int split = tile / (m_tiles * n_tiles);
      const int kb0 = split * kb_per_split;
      const int nkb = min(kb_total, kb0 + kb_per_split) - kb0;
      const int et = threadIdx.x - 64;
      const bf16* xg = x_raw + (size_t)(n_tile * BN) * K + (size_t)kb0 * TC_BK;
      constexpr int CHUNKS = BN * 8;  // 16-byte chunks per k-block tile
#pragma unroll 1
      for (int i = 0; i < nkb; ++i) {
        const uint32_t xs = smem_base + i * C::STAGE_BYTES + TC_W_BYTES;
#pragma unroll
        for (int c0 = 0; c0 < CHUNKS; c0 += 128) {
          const int idx = c0 + et;
          if (CHUNKS >= 128 || idx < CHUNKS) {
            const int r = idx >> 3, c = idx & 7;
            // rows past T are zero-filled (source size 0), as the tensor map's out-of-bounds fill would
            const bool in = n_tile * BN + r < T;
            const bf16* src = xg + (size_t)(in ? r : 0) * K + i * TC_BK + c * 8;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(xs + r * 128 + ((c ^ (r & 7)) << 4)), "l"(src),
                         "r"(in ? 16 : 0)
                         : "memory");
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      // All copies were issued together and land within one L2 round trip of each other: one wait, one proxy fence
      // (a fence per k-block cost more than the pipelining it allowed), then the arrival on the tile's barrier.
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      fence_proxy_async();
      mbar_arrive(smem_u32(&full_bar[0]));
    }
    if (pf.k_pages != nullptr) {
      // These four warps idle until the first accumulator is ready: spend the time asking the TMA
      // engine to pull this launch's share of the next attention pass's K/V pages into L2.
      const int frac = *pf.frac_q16;
      const long long items = (long long)pf.B * pf.bt_stride * pf.nkv;
      const long long step = (long long)gridDim.x * 128 * pf.nparts;
      for (long long it = ((long long)blockIdx.x * 128 + (threadIdx.x - 64)) * pf.nparts + pf.part; it < items; it += step) {
        const int kvh = (int)(it % pf.nkv);
        const int j = (int)((it / pf.nkv) % pf.bt_stride);
        const int b = (int)(it / ((long long)pf.nkv * pf.bt_stride));
        const int nb = (pf.seq_lens[b] + 15) >> 4;
        if (j < (int)(((long long)nb * frac + 65535) >> 16)) {
          const size_t off = ((size_t)pf.block_table[(size_t)b * pf.bt_stride + j] * pf.nkv + kvh) * (16 * 128);
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pf.k_pages + off), "r"(4096) : "memory");
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pf.v_pages + off), "r"(4096) : "memory");
        }
      }
    }
    int lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const int m_tile = tile % m_tiles;
      const int n_tile = (tile / m_tiles) % n_tiles;
      const int split = tile / (m_tiles * n_tiles);
      const int as = lt & 1;
      mbar_wait(smem_u32(&tmem_full_bar[as]), (lt >> 1) & 1);
      if (lt == 0 && threadIdx.x == 64) { tc_stamp(5); TC_PHASE(2); }  // accumulator visible to the epilogue
      tc_fence_after();
      const int row = m_tile * TC_BM + q * 32 + lane;  // output feature
      const bool row_ok = row < N;
      const int tok0 = n_tile * BN;
      const bool full = row_ok && tok0 + BN <= T;  // whole tile in range: no per-element predicates
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
      if (EPI == TC_EPI_ARGMAX_ROWS) {
        // here the accumulator row is a TOKEN and the columns are vocabulary entries col0 .. col0+BN
        const int tok = row, col0 = n_tile * BN;
        if (tok != row_best_tok) {
          if (row_best_tok >= 0 && row_best_tok < T && row_best != 0ull) atomicMax(argmax_packed + row_best_tok, row_best);
          row_best = 0ull;
          row_best_tok = tok;
        }
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
          uint32_t r[32];
          tmem_ld32(tbase + (uint32_t)c0, r);
          tmem_ld_wait();
          uint32_t packed[16];
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const float a = __uint_as_float(r[j]), b = __uint_as_float(r[j + 1]);
            packed[j >> 1] = pack_bf16x2(a, b);
            const float la = bf16lo(packed[j >> 1]), lb = bf16hi(packed[j >> 1]);
            const int ca = col0 + c0 + j;
            if (ca < N && la == la) { const unsigned long long p = tc_argmax_pack(la, ca); row_best = p > row_best ? p : row_best; }
            if (ca + 1 < N && lb == lb) { const unsigned long long p = tc_argmax_pack(lb, ca + 1); row_best = p > row_best ? p : row_best; }
          }
          if (out != nullptr && tok < T) {
            bf16* d = reinterpret_cast<bf16*>(out) + (size_t)tok * N + col0 + c0;
            if (col0 + c0 + 32 <= N) {
#pragma unroll
              for (int v4 = 0; v4 < 4; ++v4)
                reinterpret_cast<uint4*>(d)[v4] = make_uint4(packed[4 * v4], packed[4 * v4 + 1], packed[4 * v4 + 2], packed[4 * v4 + 3]);
            } else {
              for (int j = 0; j < 32 && col0 + c0 + j < N; ++j)
                d[j] = __ushort_as_bfloat16((unsigned short)((j & 1) ? (packed[j >> 1] >> 16) : (packed[j >> 1] & 0xffffu)));
            }
          }
        }
      } else if (EPI == TC_EPI_F32_PARTIAL) {
        float* dst = reinterpret_cast<float*>(out) + ((size_t)split * T + tok0) * N + row;  // + j * N per token
        if (BN == 16 || SMALL) {  // 16 columns at a time: half the live registers of the 32-column loop below
#pragma unroll 1
          for (int c0 = 0; c0 < BN; c0 += 16) {
            uint32_t r[16];
            tmem_ld16(tbase + (uint32_t)c0, r);
            tmem_ld_wait();
            float* d = dst + (size_t)c0 * N;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (full || (row_ok && tok0 + c0 + j < T)) d[(size_t)j * N] = __uint_as_float(r[j]);
          }
        } else {
#pragma unroll 1
          for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tbase + (uint32_t)c0, r);
            tmem_ld_wait();
            float* d = dst + (size_t)c0 * N;
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; ++j) d[(size_t)j * N] = __uint_as_float(r[j]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (row_ok && tok0 + c0 + j < T) d[(size_t)j * N] = __uint_as_float(r[j]);
            }
          }
        }
      } else if (EPI == TC_EPI_BF16) {
        bf16* dst = reinterpret_cast<bf16*>(out) + (size_t)tok0 * N + row;
        if (BN == 16) {
          uint32_t r[16];
          tmem_ld16(tbase, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (full || (row_ok && tok0 + j < T)) dst[(size_t)j * N] = __float2bfloat16_rn(__uint_as_float(r[j]));
        } else {
#pragma unroll 1
          for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tbase + (uint32_t)c0, r);
            tmem_ld_wait();
            bf16* d = dst + (size_t)c0 * N;
            if (resid != nullptr) {
              // out = bf16(resid + bf16(acc)): nn.Linear's rounding, then the residual add's (HF:...:339-351); resid may be out itself
              // (all 32 loads first: resid and out may be the same buffer, so the compiler will not move a load above a store)
              // and without a predicate per element -- with one the compiler wraps every load in a branch of its own and the
              // 32 L2 round trips run one after the other: 186 instead of 29 us for an 8192 x 1024 x 2048 projection)
              const unsigned short* rs = reinterpret_cast<const unsigned short*>(resid) + (size_t)(tok0 + c0) * N + row;
              if (full) {
                unsigned short rv[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) rv[j] = rs[(size_t)j * N];
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  d[(size_t)j * N] = __float2bfloat16_rn(__uint_as_float((uint32_t)rv[j] << 16) + bf16_round(__uint_as_float(r[j])));
              } else {  // the ragged last token tile / rows past N
#pragma unroll 1
                for (int j = 0; j < 32; ++j)
                  if (row_ok && tok0 + c0 + j < T)
                    d[(size_t)j * N] = __float2bfloat16_rn(__uint_as_float((uint32_t)rs[(size_t)j * N] << 16) + bf16_round(__uint_as_float(r[j])));
              }
            } else if (full) {
#pragma unroll
              for (int j = 0; j < 32; ++j) d[(size_t)j * N] = __float2bfloat16_rn(__uint_as_float(r[j]));
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (row_ok && tok0 + c0 + j < T) d[(size_t)j * N] = __float2bfloat16_rn(__uint_as_float(r[j]));
            }
          }
        }
      } else {
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(tbase + (uint32_t)c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int tok = tok0 + c0 + j;
            const bf16 lb = __float2bfloat16_rn(__uint_as_float(r[j]));
            const float lv = __bfloat162float(lb);
            if (out != nullptr && tok < T && row_ok) reinterpret_cast<bf16*>(out)[(size_t)tok * N + row] = lb;
            unsigned long long p = (row_ok && lv == lv) ? tc_argmax_pack(lv, row) : 0ull;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              unsigned long long other = __shfl_xor_sync(0xffffffffu, p, o);
              p = other > p ? other : p;
            }
            if (lane == 0) amax_red[q * BN + c0 + j] = p;
          }
        }
      }
      // this thread's TMEM reads of the accumulator stage are complete: hand it back to the MMA warp
      tc_fence_before();
      mbar_arrive(smem_u32(&tmem_empty_bar[as]));
      if (EPI == TC_EPI_ARGMAX) {
        asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
        for (int e = threadIdx.x - 64; e < BN; e += 128) {
          const int tok = n_tile * BN + e;
          if (tok < T) {
            unsigned long long p = amax_red[e];
#pragma unroll
            for (int k = 1; k < 4; ++k) p = amax_red[k * BN + e] > p ? amax_red[k * BN + e] : p;
            atomicMax(argmax_packed + tok, p);
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // amax_red is rewritten by the next tile
      }
    }
    if (EPI == TC_EPI_ARGMAX_ROWS && row_best_tok >= 0 && row_best_tok < T && row_best != 0ull)
      atomicMax(argmax_packed + row_best_tok, row_best);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) tc_stamp(6);  // epilogue done
#ifdef B2_STEP_TRACE
  if (threadIdx.x == 0 && _stamp[0] != 0ull) {  // second record: the phases of the first tile
    unsigned long long* buf = (unsigned long long*)_stamp[0];
    const unsigned long long slot = atomicAdd(buf, 1ull);
    if (slot < buf[1]) {
      unsigned long long* r = buf + 4 + slot * 4;
      r[0] = 101ull | ((unsigned long long)(blockIdx.x & 0xffffffu) << 8) |
             ((unsigned long long)((((unsigned)(N >> 6) & 0xfffu) | (((unsigned)(K >> 6) & 0x7ffu) << 12) | (x_row0 != 0 ? (1u << 23) : 0u))) << 40);
      r[1] = phase_t[0]; r[2] = phase_t[1]; r[3] = phase_t[2];
    }
  }
#endif
  B2_STAMP_END_AUX(100 /* any projection: the aux word tells them apart by (N, K) */, ((unsigned)(N >> 6) & 0xfffu) | (((unsigned)(K >> 6) & 0x7ffu) << 12) | (x_row0 != 0 ? (1u << 23) : 0u));
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

}  // namespace b2

int b2_make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return b2_make_tmap_bf16_ex(out, base, rows, cols, ld, 64, box_rows, 1);
}

int b2_make_tmap_bf16_ex(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols, uint32_t box_rows,
                         int swizzle128) {
  auto fn = b2::get_encode_fn();
  B2_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled driver entry point not available");
  B2_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 2) % 16 == 0, "tensor map: base/stride must be 16-byte aligned");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu ld=%llu box=%u)",
             (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows);
  return 0;
}

namespace b2 {

int tc_set_trace(unsigned long long* dev_buf) {
  B2_CUDA_OK(cudaMemcpyToSymbol(g_tc_trace, &dev_buf, sizeof(dev_buf)));
  return 0;
}

int tc_pick_bn(int T) {
  if (T <= 16) return 16;
  if (T <= 32) return 32;
  if (T <= 64) return 64;
  if (T <= 128) return 128;
  return 256;
}

int tc_pick_nsplit(int T, int N, int K, int num_sms) {
  const int bn = tc_pick_bn(T);
  const int tiles = ((N + TC_BM - 1) / TC_BM) * ((T + bn - 1) / bn);
  const int kb_total = K / TC_BK;
  int ns = num_sms / (tiles > 0 ? tiles : 1);
  if (ns < 1) ns = 1;
  if (ns > kb_total) ns = kb_total;
  if (ns > 32) ns = 32;
  // At most 8 splits: past that the consumer kernel's partial-sum loads cost more than the extra streaming CTAs
  // gain (C2 decode on B200, ms/step: cap 32 1.929, 12 1.881, 8 1.877, 6 1.910, 4 1.933, 2 2.077).  B2_SPLIT_MAX overrides.
  static const int cap = [] { const char* e = getenv("B2_SPLIT_MAX"); return e ? atoi(e) : 8; }();
  if (ns > cap && cap >= 1) ns = cap;
  const int kps = (kb_total + ns - 1) / ns;
  return (kb_total + kps - 1) / kps;  // every split non-empty
}

template <int BN, int EPI, int SMALL = 0, int XCP = 0>
static int launch_one(const TcGemmArgs& a, int kb_per_split, cudaStream_t s) {
  using C = TcCfg<BN, SMALL>;
  static B2PerDeviceOnce attr_once, carve_once;
  if (attr_once.need())
    B2_CUDA_OK(cudaFuncSetAttribute(linear_tc_kernel<BN, EPI, SMALL, XCP>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
  b2_prefer_max_smem(linear_tc_kernel<BN, EPI, SMALL, XCP>, carve_once);
  // TC_EPI_ARGMAX_ROWS: tokens are the 128-row M tiles and weight rows the BN-wide N tiles
  const bool rows_mode = (EPI == TC_EPI_ARGMAX_ROWS);
  const int m_tiles = rows_mode ? (a.T + TC_BM - 1) / TC_BM : (a.N + TC_BM - 1) / TC_BM;
  const int n_tiles = rows_mode ? (a.N + BN - 1) / BN : (a.T + BN - 1) / BN;
  const int total = m_tiles * n_tiles * a.nsplit;
  static int sms = device_num_sms();
  const int grid = total < sms ? total : sms;
  if (XCP) {
    B2_REQUIRE(total <= sms && kb_per_split <= C::STAGES && a.x_raw != nullptr, "linear_tc: cp.async X path needs one tile per CTA");
  }
  B2_CUDA_OK(b2_launch(linear_tc_kernel<BN, EPI, SMALL, XCP>, dim3(grid), dim3(192), (size_t)C::SMEM_BYTES, s, a.tmap_w, a.tmap_x, a.out,
                       a.argmax_packed, a.T, a.N, a.K, kb_per_split, m_tiles, n_tiles, a.nsplit, a.prefetch, a.x_row0,
                       reinterpret_cast<const bf16*>(a.x_raw), reinterpret_cast<const bf16*>(a.resid)));
  return 0;
}

// one tile per CTA and every k-block of a split resident in the ring: the shapes of a decode step
template <int BN, int SMALL>
static bool xcp_ok(const TcGemmArgs& a, int kps) {
  static const bool on = [] { const char* e = getenv("B2_XCP"); return !(e && *e == '0'); }();
  static int sms = device_num_sms();
  const int total = ((a.N + TC_BM - 1) / TC_BM) * ((a.T + BN - 1) / BN) * a.nsplit;
  return on && a.x_raw != nullptr && BN <= 64 && total <= sms && kps <= TcCfg<BN, SMALL>::STAGES;
}

template <int BN>
static int launch_bn(const TcGemmArgs& a, int kps, cudaStream_t s) {
  switch (a.epilogue) {
    case TC_EPI_BF16:
      if (BN <= 64 && xcp_ok<(BN <= 64 ? BN : 64), 0>(a, kps)) return launch_one<(BN <= 64 ? BN : 64), TC_EPI_BF16, 0, 1>(a, kps, s);
      return launch_one<BN, TC_EPI_BF16>(a, kps, s);
    case TC_EPI_F32_PARTIAL:
      if (BN <= 64 && a.small_smem) {
        if (xcp_ok<(BN <= 64 ? BN : 64), 1>(a, kps)) return launch_one<(BN <= 64 ? BN : 64), TC_EPI_F32_PARTIAL, 1, 1>(a, kps, s);
        return launch_one<(BN <= 64 ? BN : 64), TC_EPI_F32_PARTIAL, 1>(a, kps, s);
      }
      if (BN <= 64 && xcp_ok<(BN <= 64 ? BN : 64), 0>(a, kps)) return launch_one<(BN <= 64 ? BN : 64), TC_EPI_F32_PARTIAL, 0, 1>(a, kps, s);
      return launch_one<BN, TC_EPI_F32_PARTIAL>(a, kps, s);
    case TC_EPI_ARGMAX: return launch_one<BN, TC_EPI_ARGMAX>(a, kps, s);
    case TC_EPI_ARGMAX_ROWS: return launch_one<BN, TC_EPI_ARGMAX_ROWS>(a, kps, s);
  }
  b2_set_error("linear_tc: bad epilogue %d", a.epilogue);
  return -22;
}

int launch_linear_tc(const TcGemmArgs& a, cudaStream_t s) {
  if (a.T <= 0) return 0;
  B2_REQUIRE(a.K % TC_BK == 0, "linear_tc: K=%d must be a multiple of %d", a.K, TC_BK);
  const int kb_total = a.K / TC_BK;
  B2_REQUIRE(a.nsplit >= 1 && a.nsplit <= kb_total, "linear_tc: nsplit=%d out of range (k-blocks=%d)", a.nsplit, kb_total);
  const int kps = (kb_total + a.nsplit - 1) / a.nsplit;
  B2_REQUIRE((kb_total + kps - 1) / kps == a.nsplit, "linear_tc: nsplit=%d leaves an empty split for %d k-blocks", a.nsplit,
             kb_total);
  B2_REQUIRE(a.nsplit == 1 || a.epilogue == TC_EPI_F32_PARTIAL, "linear_tc: split-K needs the fp32 partial epilogue");
  B2_REQUIRE(a.resid == nullptr || (a.epilogue == TC_EPI_BF16 && a.bn >= 32), "linear_tc: a residual operand needs the bf16 epilogue and a token tile of at least 32");
  B2_REQUIRE(a.x_row0 == 0 || a.epilogue == TC_EPI_BF16 || a.epilogue == TC_EPI_F32_PARTIAL,
             "linear_tc: a row offset into X is only built for the plain epilogues");
  switch (a.bn) {
    case 16: return launch_bn<16>(a, kps, s);
    case 32: return launch_bn<32>(a, kps, s);
    case 64: return launch_bn<64>(a, kps, s);
    case 128: return launch_bn<128>(a, kps, s);
    case 256: return launch_bn<256>(a, kps, s);
  }
  b2_set_error("linear_tc: bad token tile %d", a.bn);
  return -22;
}

}  // namespace b2
