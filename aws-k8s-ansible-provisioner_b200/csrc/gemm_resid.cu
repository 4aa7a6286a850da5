// gemm_resid.cu -- the two projections of a decode step that feed the residual stream (attention output projection and
// MLP down projection, HF:models/qwen3/modeling_qwen3.py:247,83) with the residual add and the split-K sum folded in:
//   h[t][n] = bf16( h[t][n] + bf16( sum_k x[t][k] W[n][k] ) )                                   (HF:...:339-351)
// and, as a by-product, the pieces of the NEXT RMSNorm's statistic: ssq[n / 16][t] = sum over 16 features of h[t][n]^2.
// for at most 64 tokens.  Same swap-AB tcgen05 layout as gemm_tc.cu / gemm_pair.cu (weight rows on MMA-M, tokens on MMA-N).
//
// Why: these projections have N = hidden (1024) rows, i.e. 8 row tiles, and need ~64 CTAs to pull their 4-6 MB of weights
// in the ~1 us that matters, so K is split 8 ways.  The stand-alone path writes the 8 fp32 partials to global memory
// (1.6 us of the kernel's 4 us), a second kernel (add_rmsnorm, 1.9 us + 1.3 us hand-off) sums them, adds the residual and
// normalises.  Here the 8 CTAs of a K split are one CLUSTER: each scatters its 128 x BN accumulator into the shared
// memory of the CTA that owns those 16 rows (st.shared::cluster, 16 bytes = four tokens per store), one cluster barrier,
// and every CTA finishes its 16 rows: sum the 8 partials IN RANK ORDER (the order add_rmsnorm used: same bits), round,
// add the residual, round, store h, and write the sum of squares of its 16 features per token.  The consumer
// (gemm_pair.cu, normalise-on-load) adds the hidden/16 pieces in index order and scales the rows as it stages them,
// so no norm kernel runs between this projection and the next one.
#include "kernels.h"

namespace b2 {

B2_STEP_TRACE_DEFINE(step_trace_set_resid)

namespace {

constexpr int GR_BM = 128, GR_BK = 64, GR_KS = 8;  // K splits = cluster size
constexpr int GR_W_BYTES = GR_BM * GR_BK * 2;
constexpr int GR_MAX_KB = 6;                        // k-blocks per CTA (K <= 3072)
constexpr int GR_ROWS = GR_BM / GR_KS;              // rows of the tile a CTA finishes: 16

template <int BN>
struct GrCfg {
  static constexpr int X_BYTES = BN * GR_BK * 2;
  static constexpr int STAGE_BYTES = GR_W_BYTES + X_BYTES;
  // partials for my 16 rows from each of the 8 ranks: [src][row][token] fp32, rows 8..15 shifted by 16 floats so that
  // the two threads of a token (rows 0-7 / 8-15) read different banks
  static constexpr int RED_SRC_FLOATS = GR_ROWS * BN + 16;
  static constexpr int RED_BYTES = GR_KS * RED_SRC_FLOATS * 4;
  static constexpr int SMEM_BYTES = GR_MAX_KB * STAGE_BYTES + RED_BYTES + 1024;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
};

__device__ __forceinline__ uint32_t gr_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t gr_mapa(uint32_t local_smem, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem), "r"(rank));
  return r;
}
__device__ __forceinline__ void gr_st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// grid = 8 * m_tiles CTAs in clusters of 8 (cluster rank = K split); 192 threads: warp 0 TMA producer (weights), warp 1 MMA
// issuer, warps 2-5 fetch X with cp.async, scatter the accumulator, finish 16 rows.
template <int BN>
__global__ void __launch_bounds__(192, 1)
linear_resid_kernel(const __grid_constant__ CUtensorMap tmap_w, const bf16* __restrict__ x, bf16* __restrict__ h, float* __restrict__ ssq,
                    int T, int N, int K, int ssq_ld, const L2Prefetch pf, int tag) {
  using C = GrCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar, tmem_full_bar;
  __shared__ uint32_t tmem_slot;
  B2_STAMP_BEGIN();
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t red_base = smem_base + GR_MAX_KB * C::STAGE_BYTES;
  const float* red = reinterpret_cast<const float*>(smem_al + GR_MAX_KB * C::STAGE_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = gr_cluster_rank();
  const int m_tile = blockIdx.x / GR_KS;
  const int nkb = (K / GR_BK) / GR_KS;   // k-blocks of this CTA: [rank * nkb, rank * nkb + nkb)
  const int kb0 = (int)rank * nkb;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&full_bar), 1 + 128);  // all weight bytes + one arrival per X-loader thread (see gemm_pair.cu)
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
    mbar_arrive_expect_tx(fb, (uint32_t)nkb * GR_W_BYTES);
    for (int i = 0; i < nkb; ++i) tma_load_2d(smem_base + i * C::STAGE_BYTES, &tmap_w, (kb0 + i) * GR_BK, m_tile * GR_BM, fb);
  }
  if (warp == 0) b2_l2_prefetch(pf, blockIdx.x, gridDim.x, lane);  // the next projection's weights, behind this CTA's own
  // every CTA of the cluster is running (its shared memory is live) before anyone stores into it
  __syncwarp();
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  pdl_wait();
  if (threadIdx.x == 0) pdl_launch_dependents();
  B2_STAMP_WAIT();

  // the rows this thread finishes after the exchange: token ft, features n0 + fr0 .. + RPT - 1
  constexpr int RPT = GR_ROWS * BN / 128;  // 8 (BN = 64) or 4 (BN = 32)
  constexpr int TPT = GR_ROWS / RPT;       // threads per token: 2 or 4
  const int et = threadIdx.x - 64;
  const int ft = et / TPT, fr0 = (et % TPT) * RPT;
  const int n0 = m_tile * GR_BM + (int)rank * GR_ROWS;
  uint4 hold = make_uint4(0, 0, 0, 0);

  if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(GR_BM, BN);
      mbar_wait(smem_u32(&full_bar), 0);
      tc_fence_after();
      for (int i = 0; i < nkb; ++i) {
        const uint32_t ws = smem_base + i * C::STAGE_BYTES;
        const uint64_t da = umma_desc_sw128(ws);
        const uint64_t db = umma_desc_sw128(ws + GR_W_BYTES);
#pragma unroll
        for (int k = 0; k < GR_BK / 16; ++k)
          tc_mma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (i > 0 || k > 0) ? 1u : 0u);
      }
      tc_commit(smem_u32(&tmem_full_bar));
    }
  } else if (warp >= 2) {
    // ---- X rows [0, BN) x this CTA's K slice, 16 bytes per cp.async, into the 128-byte-swizzled tile layout ----
    const bf16* xg = x + (size_t)kb0 * GR_BK;
    constexpr int CHUNKS = BN * 8;
#pragma unroll 1
    for (int i = 0; i < nkb; ++i) {
      const uint32_t xs = smem_base + i * C::STAGE_BYTES + GR_W_BYTES;
#pragma unroll
      for (int c0 = 0; c0 < CHUNKS; c0 += 128) {
        const int idx = c0 + et;
        const int r = idx >> 3, c = idx & 7;
        const bool in = r < T;
        const bf16* src = xg + (size_t)(in ? r : 0) * K + i * GR_BK + c * 8;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(xs + r * 128 + ((c ^ (r & 7)) << 4)), "l"(src), "r"(in ? 16 : 0)
                     : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    // the residual rows this thread will add to (written two or three kernels ago, complete before the kernel this one
    // waited for could start): in flight while the operands land
    if (ft < T) {
      if (RPT == 8) hold = *reinterpret_cast<const uint4*>(h + (size_t)ft * N + n0 + fr0);
      else { const uint2 v = *reinterpret_cast<const uint2*>(h + (size_t)ft * N + n0 + fr0); hold.x = v.x; hold.y = v.y; }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    fence_proxy_async();
    mbar_arrive(smem_u32(&full_bar));

    // ---- scatter: accumulator row r belongs to cluster rank r / 16; four tokens per 16-byte store ----
    mbar_wait(smem_u32(&tmem_full_bar), 0);
    tc_fence_after();
    const int q4 = warp & 3;               // TMEM lane quarter this warp may read (warps 2..5 -> 2,3,0,1)
    const int row = q4 * 32 + lane;        // accumulator row = weight row of the tile
    const uint32_t owner = (uint32_t)(row >> 4);
    const int rin = row & 15;
    const uint32_t dst = gr_mapa(red_base, owner) + (uint32_t)(((int)rank * C::RED_SRC_FLOATS + rin * BN + (rin >> 3) * 16) * 4);
    const uint32_t tbase = tmem_d + ((uint32_t)(q4 * 32) << 16);
#pragma unroll
    for (int c0 = 0; c0 < BN; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(tbase + (uint32_t)c0, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        gr_st_cluster_v4(dst + (uint32_t)((c0 + j) * 4), __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                         __uint_as_float(r[j + 3]));
    }
    tc_fence_before();
  }
  // one barrier for the cluster: writers release, readers acquire
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 1) {  // every tcgen05.ld of this CTA is behind the barrier: TMEM goes back before the finishing pass
    tc_fence_after();
    tmem_dealloc(tmem_d, C::TMEM_COLS);
  }

  if (warp >= 2) {
    float acc[RPT];
#pragma unroll
    for (int e = 0; e < RPT; ++e) acc[e] = 0.f;
#pragma unroll
    for (int s = 0; s < GR_KS; ++s) {  // split order, as the stand-alone reduction
      const float* p = red + s * C::RED_SRC_FLOATS + fr0 * BN + (fr0 >> 3) * 16 + ft;
#pragma unroll
      for (int e = 0; e < RPT; ++e) acc[e] += p[e * BN];
    }
    float hv[8];
    unpack8(hold, hv);
    float ss = 0.f;
#pragma unroll
    for (int e = 0; e < RPT; ++e) {
      hv[e] = bf16_round(hv[e] + bf16_round(acc[e]));  // nn.Linear's rounding, then the residual add's
      ss += hv[e] * hv[e];
    }
    ss += __shfl_xor_sync(0xffffffffu, ss, 1);
    if (TPT == 4) ss += __shfl_xor_sync(0xffffffffu, ss, 2);
    if (ft < T) {
      const uint4 o = pack8(hv);
      if (RPT == 8) *reinterpret_cast<uint4*>(h + (size_t)ft * N + n0 + fr0) = o;
      else *reinterpret_cast<uint2*>(h + (size_t)ft * N + n0 + fr0) = make_uint2(o.x, o.y);
      if ((et % TPT) == 0) ssq[(size_t)(n0 / GR_ROWS) * ssq_ld + ft] = ss;
    }
  }
#ifdef B2_STEP_TRACE
  __syncthreads();  // the end stamp is the CTA's last thread's
#endif
  B2_STAMP_END_AUX(100, ((unsigned)(N >> 6) & 0xfffu) | (((unsigned)(K >> 6) & 0x7ffu) << 12) | (tag ? (1u << 23) : 0u));
}

template <int BN>
int launch_resid(const CUtensorMap& tm_w, const bf16* x, bf16* h, float* ssq, int T, int N, int K, int ssq_ld, const L2Prefetch& pf, int tag,
                 cudaStream_t s) {
  using C = GrCfg<BN>;
  static B2PerDeviceOnce attr_once, carve_once;
  if (attr_once.need())
    B2_CUDA_OK(cudaFuncSetAttribute(linear_resid_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
  b2_prefer_max_smem(linear_resid_kernel<BN>, carve_once);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(GR_KS * (N / GR_BM));
  cfg.blockDim = dim3(192);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = GR_KS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = b2_pdl_enabled() ? 2 : 1;
  B2_CUDA_OK(cudaLaunchKernelEx(&cfg, linear_resid_kernel<BN>, tm_w, x, h, ssq, T, N, K, ssq_ld, pf, tag));
  return 0;
}

}  // namespace

bool resid_shapes_ok(int T, int N, int K) {
  return T >= 1 && T <= 64 && N % GR_BM == 0 && K % (GR_KS * GR_BK) == 0 && K / (GR_KS * GR_BK) <= GR_MAX_KB;
}

// tm_w: W [N][K] with a 128-row x 64-column box.  h [T][N] is updated in place; ssq [N / 16][ssq_ld] receives, per token
// column t < T, the sum of squares of the 16 updated features of each piece.
int launch_linear_resid(const CUtensorMap& tm_w, const bf16* x, bf16* h, float* ssq, int T, int N, int K, int ssq_ld, cudaStream_t s, int tag,
                        const L2Prefetch& pf) {
  B2_REQUIRE(resid_shapes_ok(T, N, K) && ssq_ld >= T, "linear_resid: T=%d N=%d K=%d not built (T <= 64, N in steps of 128, K in steps of 512 up to 3072)",
             T, N, K);
  if (T <= 32) return launch_resid<32>(tm_w, x, h, ssq, T, N, K, ssq_ld, pf, tag, s);
  return launch_resid<64>(tm_w, x, h, ssq, T, N, K, ssq_ld, pf, tag, s);
}

}  // namespace b2
