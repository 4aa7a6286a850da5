// gemm_gateup.cu -- prefill-sized gate/up projection with SwiGLU in the epilogue
// (HF models/qwen3/modeling_qwen3.py:81-83: down_proj(act_fn(gate_proj(x)) * up_proj(x)); Phi-3's fused gate_up_proj has the
// same row order, gate rows first).
//
// The stand-alone path writes gu[T][2I] (bf16) and a second kernel reads it back: at 8192 tokens that is 100 MB written
// and 150 MB moved again by silu_mul_kernel (41 us per layer per chunk, a tenth of the C2 prefill).  Here a tile is 128
// gate rows AND the 128 up rows of the same features x 128 tokens: two accumulators in TMEM that share the token operand,
// so the thread that owns TMEM lane r holds gate[r][tok] and up[r][tok] for the same r and the activation needs no
// exchange.  Same roles as gemm_tc.cu (swap-AB: weight rows are MMA-M, tokens MMA-N):
//   warp 0    : TMA producer, per 64-wide k-block: gate rows, up rows, token rows (48 KB, 4 stages)
//   warp 1    : TMEM allocator (512 columns = 2 tiles x (gate 128 | up 128)) + tcgen05.mma issuer
//   warps 2-17: epilogue, four warps per TMEM lane quarter, one 32-token column block each (the activation costs ~60
//               instructions per element: with fewer warps the epilogue, not the MMA, set the tile time):
//               act = bf16( bf16(silu(bf16 gate)) * bf16 up ), the rounding points of silu_mul_kernel -> same bits
// Persistent over (feature tile, token tile) with the feature tile fastest, so the CTAs running together share token rows in L2.
#include "kernels.h"

namespace b2 {

namespace {

constexpr int GU_BM = 128, GU_BN = 128, GU_BK = 64;
constexpr int GU_W_BYTES = GU_BM * GU_BK * 2;          // 16 KB: one operand tile of a k-block
constexpr int GU_STAGE_BYTES = 3 * GU_W_BYTES;         // gate | up | tokens
constexpr int GU_STAGES = 4;
constexpr int GU_OUT_BYTES = 32 * GU_BM * 2;  // staging buffer of one 32-token column block: [32 tokens][128 features] bf16
constexpr int GU_SMEM_BYTES = GU_STAGES * GU_STAGE_BYTES + 4 * GU_OUT_BYTES + 1024;
constexpr int GU_EPI_WARPS = 16;
constexpr int GU_THREADS = 64 + 32 * GU_EPI_WARPS;

__global__ void __launch_bounds__(GU_THREADS, 1)
gateup_swiglu_tc_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                        const __grid_constant__ CUtensorMap tmap_out, int T, int I, int K, int m_tiles, int n_tiles, int x_row0) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[GU_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[GU_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_slot;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb_total = K / GU_BK;
  const int total_tiles = m_tiles * n_tiles;

  if (threadIdx.x == 0) {
    for (int s = 0; s < GU_STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tmem_full_bar[a]), 1);
      mbar_init(smem_u32(&tmem_empty_bar[a]), 32 * GU_EPI_WARPS);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmap_w);
      tma_prefetch_desc(&tmap_x);
      int it = 0;
      bool waited = false;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m_tile = tile % m_tiles, n_tile = tile / m_tiles;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % GU_STAGES;
          mbar_wait(smem_u32(&empty_bar[s]), ((it / GU_STAGES) & 1) ^ 1);
          const uint32_t fb = smem_u32(&full_bar[s]);
          mbar_arrive_expect_tx(fb, GU_STAGE_BYTES);
          const uint32_t ws = smem_base + s * GU_STAGE_BYTES;
          tma_load_2d(ws, &tmap_w, kb * GU_BK, m_tile * GU_BM, fb);                    // gate rows
          tma_load_2d(ws + GU_W_BYTES, &tmap_w, kb * GU_BK, I + m_tile * GU_BM, fb);   // up rows of the same features
          if (!waited && it == GU_STAGES - 1) {
            // the weight tiles of the first ring pass do not depend on the previous kernel; its token rows do
            pdl_wait();
            pdl_launch_dependents();
            waited = true;
            for (int j = 0; j <= it; ++j)
              tma_load_2d(smem_base + (j % GU_STAGES) * GU_STAGE_BYTES + 2 * GU_W_BYTES, &tmap_x, j * GU_BK, n_tile * GU_BN + x_row0,
                          smem_u32(&full_bar[j % GU_STAGES]));
          } else if (waited) {
            tma_load_2d(ws + 2 * GU_W_BYTES, &tmap_x, kb * GU_BK, n_tile * GU_BN + x_row0, fb);
          }
        }
      }
      if (!waited) {
        pdl_wait();
        pdl_launch_dependents();
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(GU_BM, GU_BN);
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
        const int as = lt & 1;
        mbar_wait(smem_u32(&tmem_empty_bar[as]), ((lt >> 1) & 1) ^ 1);  // epilogue drained this accumulator pair
        tc_fence_after();
        const uint32_t d_gate = tmem_base + (uint32_t)(as * 2 * GU_BN), d_up = d_gate + GU_BN;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % GU_STAGES;
          mbar_wait(smem_u32(&full_bar[s]), (it / GU_STAGES) & 1);
          tc_fence_after();
          const uint32_t ws = smem_base + s * GU_STAGE_BYTES;
          const uint64_t dg = umma_desc_sw128(ws), du = umma_desc_sw128(ws + GU_W_BYTES), dx = umma_desc_sw128(ws + 2 * GU_W_BYTES);
#pragma unroll
          for (int k = 0; k < GU_BK / 16; ++k) {
            const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
            tc_mma_bf16(d_gate, dg + (uint64_t)(2 * k), dx + (uint64_t)(2 * k), idesc, acc);
            tc_mma_bf16(d_up, du + (uint64_t)(2 * k), dx + (uint64_t)(2 * k), idesc, acc);
          }
          tc_commit(smem_u32(&empty_bar[s]));
        }
        tc_commit(smem_u32(&tmem_full_bar[as]));
      }
    }
  } else {
    const int q = warp & 3;             // TMEM lane quarter this warp may read
    const int part = (warp - 2) >> 2;   // which of the quarter's four warps: its 32-token column block
    pdl_wait();
    const uint32_t obuf = smem_base + GU_STAGES * GU_STAGE_BYTES;
    uint8_t* obuf_p = smem_raw + (smem_base - smem_u32(smem_raw)) + GU_STAGES * GU_STAGE_BYTES;
    int lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const int m_tile = tile % m_tiles, n_tile = tile / m_tiles;
      const int as = lt & 1;
      mbar_wait(smem_u32(&tmem_full_bar[as]), (lt >> 1) & 1);
      tc_fence_after();
      const int row = m_tile * GU_BM + q * 32 + lane;  // intermediate feature
      const int tok0 = n_tile * GU_BN;
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 2 * GU_BN);
      {
        const int c0 = part * 32;
        uint32_t g[32], u[32];
        tmem_ld32(tbase + (uint32_t)c0, g);
        tmem_ld32(tbase + (uint32_t)(GU_BN + c0), u);
        tmem_ld_wait();
        // both blocks are in registers: the accumulator pair goes back to the MMA warp before the activation math
        tc_fence_before();
        mbar_arrive(smem_u32(&tmem_empty_bar[as]));
        // The block leaves through shared memory and ONE bulk tensor store ([32 tokens][128 features], full 256-byte rows;
        // rows past T are clipped by the tensor map) instead of 32 two-byte stores per thread: see gemm_mc.cu.  The four
        // warps of a column block share staging buffer `part` and named barrier 1 + part.
        unsigned short* ob = reinterpret_cast<unsigned short*>(obuf_p + part * GU_OUT_BYTES) + q * 32 + lane;
        const bool issuer = q == 0 && lane == 0;  // one of the block's 128 threads
        if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the previous tile's store has read the buffer
        asm volatile("bar.sync %0, 128;" ::"r"(1 + part) : "memory");
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float gv = bf16_round(__uint_as_float(g[j])), uv = bf16_round(__uint_as_float(u[j]));
          const float sl = bf16_round(__fdiv_rn(gv, 1.0f + expf(-gv)));
          ob[j * GU_BM] = __bfloat16_as_ushort(__float2bfloat16_rn(sl * uv));
        }
        fence_proxy_async();
        asm volatile("bar.sync %0, 128;" ::"r"(1 + part) : "memory");
        if (issuer) {
          asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmap_out), "r"(m_tile * GU_BM),
                       "r"(tok0 + c0), "r"(obuf + (uint32_t)(part * GU_OUT_BYTES))
                       : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (q == 0 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

bool gateup_prefill_shapes_ok(int T, int I, int K) { return T >= 256 && I % GU_BM == 0 && K % GU_BK == 0 && K / GU_BK >= GU_STAGES; }

// tm_w: the [2I][K] gate|up weight with a 128-row x 64-column box (the projection kernel's map); tm_x128: X with a 128-row box.
int launch_gateup_swiglu_prefill(const CUtensorMap& tm_w, const CUtensorMap& tm_x128, bf16* act, int T, int I, int K, int x_row0,
                                 cudaStream_t s) {
  if (T <= 0) return 0;
  B2_REQUIRE(gateup_prefill_shapes_ok(T, I, K), "gateup_swiglu_prefill: shape T=%d I=%d K=%d not built", T, I, K);
  CUtensorMap tm_out;  // act [T][I], box 128 features x 32 tokens, no swizzle: the staging buffers' layout
  {
    const int rc = b2_make_tmap_bf16_ex(&tm_out, act, (uint64_t)T, (uint64_t)I, (uint64_t)I, GU_BM, 32, /*swizzle128=*/0);
    if (rc != 0) return rc;
  }
  static B2PerDeviceOnce attr_once;
  if (attr_once.need())
    B2_CUDA_OK(cudaFuncSetAttribute(gateup_swiglu_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GU_SMEM_BYTES));
  const int m_tiles = I / GU_BM, n_tiles = (T + GU_BN - 1) / GU_BN;
  const int total = m_tiles * n_tiles;
  static int sms = device_num_sms();
  const int grid = total < sms ? total : sms;
  B2_CUDA_OK(b2_launch(gateup_swiglu_tc_kernel, dim3(grid), dim3(GU_THREADS), (size_t)GU_SMEM_BYTES, s, tm_w, tm_x128, tm_out, T, I, K, m_tiles,
                       n_tiles, x_row0));
  return 0;
}

}  // namespace b2
