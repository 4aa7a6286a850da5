// gemm_mc.cu -- prefill-sized projection y[T][N] = x[T][K] . W[N][K]^T (bf16 out) with the token operand SHARED between
// the two CTAs of a cluster (HF nn.Linear, models/qwen3/modeling_qwen3.py:236-247).
//
// Why: the persistent 128 x 256 tile kernel of gemm_tc.cu moves 48 KB from L2 into shared memory per 64-wide k-block
// (16 KB of weights + 32 KB of tokens) for 532 cycles of tcgen05.mma: 90 bytes per cycle and SM, more than an SM gets
// from L2 when all 148 ask at once -- ncu shows the tensor pipe 52-68 % active and the share rising with K, i.e. with the
// weight of the main loop (profiles/r02_prefill_full_raw.csv).  Here two CTAs take two ADJACENT weight-row tiles of the same
// 256-token tile: each loads its own 16 KB of weights and only HALF of the token rows (16 KB), and the TMA engine writes that
// half into BOTH CTAs' shared memory (cp.async.bulk.tensor ... .multicast::cluster; the tensor maps cover this GPU's memory
// only).  32 KB per k-block and SM instead of 48.
//   * full barrier of a stage (per CTA): one arrive.expect_tx of 48 KB by the CTA's own producer; bytes arrive from its own
//     weight load, its own token half and the peer's token half
//   * empty barrier of a stage (per CTA): TWO arrivals -- the stage is rewritten by the peer's multicast too, so both CTAs'
//     MMAs must have read it: each MMA thread commits with .multicast::cluster to the barrier of both CTAs
// Everything else is gemm_tc.cu's plain bf16 path: swap-AB (weight rows on MMA-M, tokens on MMA-N), double-buffered TMEM
// accumulator, four epilogue warps, same k order -> the same bits (tests/test_kernels_gpu.py::test_linear_multicast).
#include "kernels.h"

namespace b2 {

namespace {

constexpr int MC_BM = 128, MC_BN = 256, MC_BK = 64;
constexpr int MC_W_BYTES = MC_BM * MC_BK * 2;       // 16 KB
constexpr int MC_XH_BYTES = (MC_BN / 2) * MC_BK * 2;  // 16 KB: one CTA's half of the token tile
constexpr int MC_STAGE_BYTES = MC_W_BYTES + 2 * MC_XH_BYTES;
constexpr int MC_STAGES = 4;
constexpr int MC_OUT_ROWS = 32;                               // tokens per output staging buffer
constexpr int MC_OUT_BYTES = MC_OUT_ROWS * MC_BM * 2;         // 8 KB: [32 tokens][128 features] bf16
constexpr int MC_SMEM_BYTES = MC_STAGES * MC_STAGE_BYTES + 2 * MC_OUT_BYTES + 1024;
constexpr int MC_ROPE_GROUPS = 4;                             // ROPE variant: epilogue warp groups (four warps each)
constexpr int MC_ROPE_ROWS = 16;                              // ... and tokens per staged block (4 KB; the groups' blocks fill the two 8 KB buffers)
constexpr int MC_ROPE_THREADS = 64 + 128 * MC_ROPE_GROUPS;

__device__ __forceinline__ uint32_t mc_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void mc_tma_load_2d_multicast(uint32_t dst_smem, const CUtensorMap* tmap, int c0, int c1, uint32_t bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(
          dst_smem),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void mc_commit_multicast(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask)
               : "memory");
}
__device__ __forceinline__ void mc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ROPE variant (QKV projection of a Qwen3-geometry prefill: one 128-wide head per weight-row tile): the staged 32-token block
// is not stored as it is but finished as qknorm_rope_kv does -- per-head RMS norm of q and k, RoPE, then q through a TMA store
// into q_out and k / v as 16-byte chunks into their cache pages -- by SIXTEEN epilogue warps in four groups that take every
// fourth 16-token block (the tail is ~150 instructions per 8 dims: with eight warps the epilogue was longer than the main loop).  Arithmetic
// and rounding points are those of the decode step's fused QKV kernel (gemm_pair.cu): sixteen threads per token, eight dims each,
// RoPE partner dims +-64 read from the staged block, sum of squares over four shuffles.
struct McRope {
  const bf16 *qw, *kw, *cos_sin;
  const int32_t *positions, *slots;
  bf16 *k_pages, *v_pages;
  int nq, nkv, block_size;
  float eps;
};

// grid = 2 * clusters; cluster c walks the (weight-tile pair, token tile) list with stride `clusters`, the pair index fastest.
template <bool ROPE>
__global__ void __launch_bounds__(ROPE ? MC_ROPE_THREADS : 192, 1)
linear_mc_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_out,
                 int T, int N, int K, int m_pairs, int n_tiles, int x_row0, const McRope rp) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[MC_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MC_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_pos[ROPE ? MC_BN : 1], s_slot[ROPE ? MC_BN : 1];  // positions / cache slots of the tile's tokens
  __shared__ __align__(16) bf16 s_gain[ROPE ? 256 : 8];             // QK-norm gains, q then k
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = mc_cluster_rank();
  const int cluster = blockIdx.x >> 1, clusters = gridDim.x >> 1;
  const int kb_total = K / MC_BK;
  const int total = m_pairs * n_tiles;

  if (threadIdx.x == 0) {
    for (int s = 0; s < MC_STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 2);  // this CTA's MMAs and the peer's
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tmem_full_bar[a]), 1);
      mbar_init(smem_u32(&tmem_empty_bar[a]), ROPE ? MC_ROPE_THREADS - 64 : 128);
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
  mc_cluster_sync();  // the peer's barriers exist before anything of mine can arrive on them
  pdl_wait();
  if (threadIdx.x == 0) pdl_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmap_w);
      tma_prefetch_desc(&tmap_x);
      int it = 0;
      for (int p = cluster; p < total; p += clusters) {
        const int m_tile = 2 * (p % m_pairs) + (int)rank, n_tile = p / m_pairs;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % MC_STAGES;
          mbar_wait(smem_u32(&empty_bar[s]), ((it / MC_STAGES) & 1) ^ 1);  // both CTAs are done with this stage
          const uint32_t fb = smem_u32(&full_bar[s]);
          mbar_arrive_expect_tx(fb, MC_STAGE_BYTES);
          const uint32_t ws = smem_base + s * MC_STAGE_BYTES;
          tma_load_2d(ws, &tmap_w, kb * MC_BK, m_tile * MC_BM, fb);
          // my half of the token rows, into the same place of both CTAs' stage; the bytes are counted on both full barriers
          mc_tma_load_2d_multicast(ws + MC_W_BYTES + rank * MC_XH_BYTES, &tmap_x, kb * MC_BK, n_tile * MC_BN + (int)rank * (MC_BN / 2) + x_row0, fb,
                                   (uint16_t)3);
        }
      }
      // tail: the releases of the last stages (this CTA's MMAs and the PEER's, which arrive through the cluster network) have
      // landed on my barriers before this CTA goes to the final cluster barrier and exits
      for (int j = it > MC_STAGES ? it - MC_STAGES : 0; j < it; ++j) mbar_wait(smem_u32(&empty_bar[j % MC_STAGES]), (j / MC_STAGES) & 1);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(MC_BM, MC_BN);
      int it = 0, lt = 0;
      for (int p = cluster; p < total; p += clusters, ++lt) {
        const int as = lt & 1;
        mbar_wait(smem_u32(&tmem_empty_bar[as]), ((lt >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(as * MC_BN);
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % MC_STAGES;
          mbar_wait(smem_u32(&full_bar[s]), (it / MC_STAGES) & 1);
          tc_fence_after();
          const uint32_t ws = smem_base + s * MC_STAGE_BYTES;
          const uint64_t da = umma_desc_sw128(ws);
          const uint64_t db = umma_desc_sw128(ws + MC_W_BYTES);  // 256 token rows: the two halves are contiguous
#pragma unroll
          for (int k = 0; k < MC_BK / 16; ++k)
            tc_mma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
          mc_commit_multicast(smem_u32(&empty_bar[s]), (uint16_t)3);
        }
        tc_commit(smem_u32(&tmem_full_bar[as]));
      }
    }
  } else {
    if constexpr (!ROPE) {
    // ---- epilogue: the tile leaves through shared memory and TMA stores, 32 tokens x 128 features (8 KB) at a time ----
    // Thread = weight row (TMEM lane), registers = tokens, memory = [token][feature]: written directly, a tile is 32 768
    // two-byte stores.  Measured over the prefill kernels, that cost ~1.7 us per 32 KB of output per tile ON TOP of the main
    // loop (QKV, 64 KB per tile: 10.3 us; gate/up + SwiGLU, same MMA work, 32 KB: 8.6 us; the main loop alone: ~7 us): the
    // scattered stores share the SM's path to L2 with the TMA loads.  Staged as [32 tokens][128 features] (a warp's 32 lanes
    // write 64 contiguous bytes of a token row: no bank conflicts) a chunk goes out as ONE bulk tensor store of full 256-byte
    // rows; rows past T are clipped by the tensor map.
    const int q = warp & 3;
    const int et = threadIdx.x - 64;
    const uint32_t obuf = smem_base + MC_STAGES * MC_STAGE_BYTES;
    uint8_t* obuf_p = smem_raw + (smem_base - smem_u32(smem_raw)) + MC_STAGES * MC_STAGE_BYTES;
    int lt = 0, chunk = 0;
    for (int p = cluster; p < total; p += clusters, ++lt) {
      const int m_tile = 2 * (p % m_pairs) + (int)rank, n_tile = p / m_pairs;
      const int as = lt & 1;
      mbar_wait(smem_u32(&tmem_full_bar[as]), (lt >> 1) & 1);
      tc_fence_after();
      const int tok0 = n_tile * MC_BN;
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * MC_BN);
#pragma unroll 1
      for (int c0 = 0; c0 < MC_BN; c0 += MC_OUT_ROWS, ++chunk) {
        uint32_t r[32];
        tmem_ld32(tbase + (uint32_t)c0, r);
        tmem_ld_wait();
        if (c0 + MC_OUT_ROWS == MC_BN) {  // the whole accumulator is in registers or out: hand it back to the MMA warp
          tc_fence_before();
          mbar_arrive(smem_u32(&tmem_empty_bar[as]));
        }
        const int b = chunk & 1;
        // buffer b was last read by the store issued two chunks ago
        if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
        unsigned short* ob = reinterpret_cast<unsigned short*>(obuf_p + b * MC_OUT_BYTES) + q * 32 + lane;
#pragma unroll
        for (int j = 0; j < 32; ++j) ob[j * MC_BM] = __bfloat16_as_ushort(__float2bfloat16_rn(__uint_as_float(r[j])));
        fence_proxy_async();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmap_out), "r"(m_tile * MC_BM),
                       "r"(tok0 + c0), "r"(obuf + (uint32_t)(b * MC_OUT_BYTES))
                       : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the last stores have left shared memory AND are complete
    } else {
    // ---- ROPE epilogue ----
    constexpr int HD = 128, RB = MC_ROPE_ROWS, NG = MC_ROPE_GROUPS, NE = 128 * NG;
    constexpr int IPT = RB * 16 / 128;            // (token, 8-dim part) items per thread and block: 2
    const int q = warp & 3, g = (warp - 2) >> 2;  // TMEM lane quarter; group 0..3 (blocks g, g + 4, g + 8, ...)
    const int et = threadIdx.x - 64;              // 0..511 over the groups
    const int eg = et & 127;                      // 0..127 within the group
    const int part = eg & 15;                     // this thread's eight dims of a token: 8 part .. 8 part + 7
    const int d0 = part * 8;
    const uint32_t obuf = smem_base + MC_STAGES * MC_STAGE_BYTES + (uint32_t)(g * RB * HD * 2);
    uint8_t* ob_p = smem_raw + (smem_base - smem_u32(smem_raw)) + MC_STAGES * MC_STAGE_BYTES + g * RB * HD * 2;
    if (et < 32) reinterpret_cast<uint4*>(s_gain)[et] = reinterpret_cast<const uint4*>(et < 16 ? rp.qw : rp.kw)[et & 15];
    const uint32_t sign = d0 < 64 ? 0x80008000u : 0u;  // dims 0..63: rot = -partner
    int lt = 0;
    for (int p = cluster; p < total; p += clusters, ++lt) {
      const int head = 2 * (p % m_pairs) + (int)rank, n_tile = p / m_pairs;
      const bool is_q = head < rp.nq, is_k = !is_q && head < rp.nq + rp.nkv;
      const int tok0 = n_tile * MC_BN;
      // every group is done with the previous tile's positions / slots; then this tile's
      asm volatile("bar.sync 5, %0;" ::"n"(NE) : "memory");
      if (et < MC_BN) {
        const int tok = tok0 + et;
        s_pos[et] = tok < T ? max(rp.positions[tok], 0) : 0;
        s_slot[et] = tok < T ? rp.slots[tok] : -1;
      }
      asm volatile("bar.sync 5, %0;" ::"n"(NE) : "memory");
      const int as = lt & 1;
      mbar_wait(smem_u32(&tmem_full_bar[as]), (lt >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * MC_BN);
#pragma unroll 1
      for (int c0 = g * RB; c0 < MC_BN; c0 += NG * RB) {
        // the RoPE table rows of this thread's tokens: requested now, used after the block is staged
        uint4 co[IPT], sn[IPT];
        if (is_q || is_k) {
#pragma unroll
          for (int k = 0; k < IPT; ++k) {
            const bf16* cs = rp.cos_sin + (size_t)s_pos[c0 + (eg >> 4) + 8 * k] * HD + (d0 & 63);
            co[k] = *reinterpret_cast<const uint4*>(cs);
            sn[k] = *reinterpret_cast<const uint4*>(cs + 64);
          }
        }
        uint32_t r[RB];
        tmem_ld16(tbase + (uint32_t)c0, r);
        tmem_ld_wait();
        if (c0 + NG * RB >= MC_BN) {  // this group's last block of the tile: its share of the accumulator is out
          tc_fence_before();
          mbar_arrive(smem_u32(&tmem_empty_bar[as]));
        }
        if (eg == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the group's previous q store has read the buffer
        asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        {
          unsigned short* ob = reinterpret_cast<unsigned short*>(ob_p) + q * 32 + lane;
#pragma unroll
          for (int j = 0; j < RB; ++j) ob[j * MC_BM] = __bfloat16_as_ushort(__float2bfloat16_rn(__uint_as_float(r[j])));
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        uint4 res[IPT];
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
          const int j = (eg >> 4) + 8 * k;  // token of the block; the 16 threads of a token are 16 consecutive lanes
          uint4 raw = *reinterpret_cast<const uint4*>(ob_p + j * 256 + part * 16);
          if (is_q || is_k) {
            const uint4 praw = *reinterpret_cast<const uint4*>(ob_p + j * 256 + (part ^ 8) * 16);
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
            const uint4 w4 = *reinterpret_cast<const uint4*>(gw + d0), pw4 = *reinterpret_cast<const uint4*>(gw + (d0 ^ 64));
            const uint32_t* w2 = &w4.x; const uint32_t* pw2 = &pw4.x; const uint32_t* co2 = &co[k].x; const uint32_t* sn2 = &sn[k].x;
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
          res[k] = raw;
        }
        if (is_q) {
          // back into the staged block (every thread has read its partner's dims: barrier first), then one bulk store
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
#pragma unroll
          for (int k = 0; k < IPT; ++k) *reinterpret_cast<uint4*>(ob_p + ((eg >> 4) + 8 * k) * 256 + part * 16) = res[k];
          fence_proxy_async();
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
          if (eg == 0) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmap_out), "r"(head * HD), "r"(tok0 + c0),
                         "r"(obuf)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        } else {
          const int kvh = is_k ? head - rp.nq : head - rp.nq - rp.nkv;
          bf16* pages = is_k ? rp.k_pages : rp.v_pages;
#pragma unroll
          for (int k = 0; k < IPT; ++k) {
            const int j = (eg >> 4) + 8 * k;
            const int slot = s_slot[c0 + j];
            if (slot >= 0) {
              const int blk = slot / rp.block_size, off = slot % rp.block_size;
              bf16* base = pages + (((size_t)blk * rp.nkv + kvh) * rp.block_size + off) * HD;
              *reinterpret_cast<uint4*>(base + kv_swizzled_chunk(off, part) * 8) = res[k];
            }
          }
        }
      }
    }
    if (eg == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  // no CTA leaves while its peer may still write into its shared memory or arrive on its barriers
  mc_cluster_sync();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

bool linear_mc_shapes_ok(int T, int N, int K) { return T >= 512 && N % (2 * MC_BM) == 0 && K % MC_BK == 0; }

template <bool ROPE>
static int launch_mc(const CUtensorMap& tm_w, const CUtensorMap& tm_x128, const CUtensorMap& tm_out, int T, int N, int K, int x_row0,
                     const McRope& rp, cudaStream_t s) {
  static B2PerDeviceOnce attr_once;
  if (attr_once.need()) B2_CUDA_OK(cudaFuncSetAttribute(linear_mc_kernel<ROPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, MC_SMEM_BYTES));
  const int m_pairs = N / (2 * MC_BM), n_tiles = (T + MC_BN - 1) / MC_BN;
  const int total = m_pairs * n_tiles;
  static int sms = device_num_sms();
  const int clusters = total < sms / 2 ? total : sms / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(ROPE ? MC_ROPE_THREADS : 192);
  cfg.dynamicSmemBytes = MC_SMEM_BYTES;
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
  B2_CUDA_OK(cudaLaunchKernelEx(&cfg, linear_mc_kernel<ROPE>, tm_w, tm_x128, tm_out, T, N, K, m_pairs, n_tiles, x_row0, rp));
  return 0;
}

// tm_w: W [N][K], 128-row x 64-column box; tm_x128: X with a 128-row box (NOT the 256-row one of the single-CTA kernel).
int launch_linear_mc(const CUtensorMap& tm_w, const CUtensorMap& tm_x128, bf16* out, int T, int N, int K, int x_row0, cudaStream_t s) {
  if (T <= 0) return 0;
  B2_REQUIRE(linear_mc_shapes_ok(T, N, K), "linear_mc: shape T=%d N=%d K=%d not built", T, N, K);
  CUtensorMap tm_out;  // out [T][N], box 128 features x 32 tokens, no swizzle: the staging buffer's layout
  {
    const int rc = b2_make_tmap_bf16_ex(&tm_out, out, (uint64_t)T, (uint64_t)N, (uint64_t)N, MC_BM, MC_OUT_ROWS, /*swizzle128=*/0);
    if (rc != 0) return rc;
  }
  return launch_mc<false>(tm_w, tm_x128, tm_out, T, N, K, x_row0, McRope(), s);
}

// Prefill-sized QKV projection + QK-norm + RoPE + K/V page write (Qwen3 geometry: 128-wide heads, norm gains given).
// tm_qkv: Wqkv [(nq + 2 nkv) * 128][K], 128-row box (one head per tile); x_row0: first row of X in tm_x128; positions / slots /
// q_out are already offset to that row.
int launch_qkv_rope_prefill(const CUtensorMap& tm_qkv, const CUtensorMap& tm_x128, const bf16* qw, const bf16* kw, const bf16* cos_sin,
                            const int32_t* positions, const int32_t* slots, bf16* q_out, bf16* k_pages, bf16* v_pages, int T, int nq, int nkv,
                            int K, int block_size, float eps, int x_row0, cudaStream_t s) {
  if (T <= 0) return 0;
  const int N = (nq + 2 * nkv) * 128;
  B2_REQUIRE(linear_mc_shapes_ok(T, N, K) && qw != nullptr && kw != nullptr, "qkv_rope_prefill: T=%d N=%d K=%d not built, or no QK-norm gains", T, N, K);
  CUtensorMap tm_q;  // q_out [T][nq * 128], box one head x 16 tokens
  {
    const int rc = b2_make_tmap_bf16_ex(&tm_q, q_out, (uint64_t)T, (uint64_t)nq * 128, (uint64_t)nq * 128, MC_BM, MC_ROPE_ROWS, /*swizzle128=*/0);
    if (rc != 0) return rc;
  }
  McRope rp = {qw, kw, cos_sin, positions, slots, k_pages, v_pages, nq, nkv, block_size, eps};
  return launch_mc<true>(tm_qkv, tm_x128, tm_q, T, N, K, x_row0, rp, s);
}

}  // namespace b2
