// chain_cluster.cu -- the decode chain between two attention passes in FOUR steps instead of eight.
//
// chain.cu kept the eight steps of the separate kernels (GEMM, then an element-wise kernel that sums
// the split-K partials) and only replaced launches by grid barriers; it lost, because each step is a
// chain of ~1 us L2 round trips whatever separates the steps.  Here every element-wise step is folded
// into the GEMM before or after it, so a layer is O, gate-up, down, QKV(next layer) and nothing else:
//
//   * 128 CTAs = 32 clusters of 4.  A cluster owns one slice of a GEMM's output features (O/down: 32
//     columns of the hidden state, gate-up: 96 gate + 96 up columns, QKV: one 128-wide head); its four
//     CTAs split K four ways.  Tokens sit on the MMA's M side (64 rows used), the weight slice on N.
//   * split-K is reduced INSIDE the cluster through distributed shared memory: every CTA scatters its
//     fp32 accumulator rows to the CTA that owns that quarter of the tokens, one cluster barrier, and the
//     owner adds the four partials in rank order (deterministic) -- no partials in global memory.
//   * the owner then applies what used to be a kernel of its own: residual add + sum of squares (O,
//     down), SwiGLU (gate-up), QK-norm + RoPE + KV page write (QKV).  Rounding points are those of
//     elementwise.cu (HF:models/qwen3/modeling_qwen3.py:59-64,81-83,151-181,248-249,270-271,327,333).
//   * RMSNorm is folded into the NEXT GEMM's operand: the CTAs of gate-up / QKV build their A tile
//     x = bf16(w * bf16(h * rstd)) straight into the swizzled shared-memory layout the MMA reads, from h
//     and the 32 per-slice sums of squares the previous step left behind.
//   * the weight slice of step i+1 is pulled into the other weight buffer while step i runs.
// Built for the bench model's shapes (hidden 1024, 32 internal heads, SwiGLU 3072, at most 64 tokens);
// the engine uses the separate kernels for anything else.
#include "kernels.h"

namespace b2 {

constexpr int FC_CL = 4;          // CTAs per cluster = K splits
constexpr int FC_CLUSTERS = 32;   // output slices per GEMM
constexpr int FC_CTAS = FC_CL * FC_CLUSTERS;
constexpr int FC_WORKERS = 16;    // worker warps: operand builders, accumulator scatter, element-wise tails
constexpr int FC_THREADS = 64 + 32 * FC_WORKERS;  // warp 0 TMA producer, warp 1 MMA issuer, then the workers
constexpr int FC_TOK = 64;        // token rows of the A tile that are real
constexpr int FC_A_STAGES = 4;
constexpr int FC_A_BYTES = FC_TOK * 128;            // one k-block (64 bf16) of 64 token rows
constexpr int FC_WX_BYTES = 96 * 1024;              // gate-up (192 rows x 4 k-blocks) / QKV (128 x 4)
constexpr int FC_WY_BYTES = 48 * 1024;              // O (32 rows x 8 k-blocks) / down (32 x 12)
constexpr int FC_RED_BYTES = 4 * 16 * 192 * 4;      // [source rank][16 tokens][<=192 columns] fp32
constexpr int FC_SMEM_BYTES = FC_A_STAGES * FC_A_BYTES + FC_WX_BYTES + FC_WY_BYTES + FC_RED_BYTES + 1024;
constexpr int FC_TMEM_COLS = 256;

struct FusedMaps {
  CUtensorMap w[4];  // Wo (box 64x32), Wgu (box 64x96), Wd (box 64x32), Wqkv of the next layer (box 64x128)
  CUtensorMap a[2];  // attention output, SwiGLU output (box 64x64)
};

__device__ unsigned long long* g_fc_trace = nullptr;
// `tr` is the trace pointer, read ONCE at kernel start: a load of g_fc_trace at every stamp put a global
// round trip into the producer's critical path.
#define fc_stamp(slot) fc_stamp_impl(tr, (slot), threadIdx.x == 0)
#define fc_stamp_any(slot) fc_stamp_impl(tr, (slot), true)
__device__ __forceinline__ void fc_stamp_impl(unsigned long long* tr, int slot, bool who) {
  if (tr != nullptr && who) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tr[blockIdx.x * 32 + slot] = t;
  }
}
int fused_chain_set_trace(unsigned long long* dev_buf) {
  B2_CUDA_OK(cudaMemcpyToSymbol(g_fc_trace, &dev_buf, sizeof(dev_buf)));
  return 0;
}

__device__ __forceinline__ unsigned long long fc_ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// Grid barrier over the 128 resident CTAs: monotonic 64-bit arrival counter (see chain.cu).
__device__ __forceinline__ void fc_grid_sync(unsigned long long* bar, unsigned long long target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(bar) : "memory");
    while (fc_ld_acquire_u64(bar) < target) {
    }
    asm volatile("fence.proxy.async;" ::: "memory");  // thread 0 issues the coming TMA reads
  }
  __syncthreads();
}
__device__ __forceinline__ void fc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// The release/acquire forms cost every thread a GPU-scope memory barrier (MEMBAR.ALL.GPU + ERRBAR in SASS).
// Only the warps that wrote remote shared memory need the release and only the warps that will read the
// reduce buffer need the acquire; everyone else just takes part in the count.
__device__ __forceinline__ void fc_cluster_sync_roles(bool releases, bool acquires) {
  if (releases) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  else asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  if (acquires) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  else asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t fc_mapa(uint32_t local_smem, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem), "r"(rank));
  return r;
}
__device__ __forceinline__ void fc_st_cluster_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// The element-wise work of a step is a few thousand elements; on four warps it is a latency-bound
// instruction stream (one warp per scheduler, nothing to interleave), which is what made the first
// version of this kernel slower than separate kernels.  Sixteen worker warps give each scheduler four.
__device__ __forceinline__ void fc_epi_bar() { asm volatile("bar.sync 1, 512;" ::: "memory"); }

struct FcPhase {
  int np;    // weight rows (MMA N) per cluster
  int kbs;   // 64-wide k-blocks per CTA
  int wbuf;  // 0 = X buffer, 1 = Y buffer
  int built; // A tile built from h (RMSNorm folded in) instead of loaded by TMA
};
__device__ __forceinline__ FcPhase fc_phase(int ph) {
  switch (ph) {
    case 0: return FcPhase{32, 8, 1, 0};    // O:      K = 2048
    case 1: return FcPhase{192, 4, 0, 1};   // gate-up K = 1024
    case 2: return FcPhase{32, 12, 1, 0};   // down:   K = 3072
    default: return FcPhase{128, 4, 0, 1};  // QKV:    K = 1024
  }
}

__global__ void __launch_bounds__(FC_THREADS, 1)
decode_fused_chain_kernel(const __grid_constant__ FusedMaps maps, const FusedChainParams p) {
  extern __shared__ uint8_t smem_raw[];
  unsigned long long* const tr = g_fc_trace;
  __shared__ __align__(8) uint64_t a_full[FC_A_STAGES];
  __shared__ __align__(8) uint64_t a_empty[FC_A_STAGES];
  __shared__ __align__(8) uint64_t w_full[2];
  __shared__ __align__(8) uint64_t tmem_full;
  __shared__ __align__(8) uint64_t b_loaded;  // builders have their inputs in registers: the weight prefetch may go
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float ss_stage[16];

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_base = smem_base;
  const uint32_t wx_base = a_base + FC_A_STAGES * FC_A_BYTES;
  const uint32_t wy_base = wx_base + FC_WX_BYTES;
  const uint32_t red_base = wy_base + FC_WY_BYTES;
  float* red = reinterpret_cast<float*>(smem_raw + (red_base - smem_u32(smem_raw)));
  uint8_t* a_ptr = smem_raw + (a_base - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = p.T;
  const int cl = blockIdx.x / FC_CL;      // cluster = output slice
  const int rk = blockIdx.x % FC_CL;      // rank in the cluster = K split = owner of tokens [16 rk, 16 rk + 16)

  if (threadIdx.x == 0) {
    for (int s = 0; s < FC_A_STAGES; ++s) {
      mbar_init(smem_u32(&a_full[s]), 1);
      mbar_init(smem_u32(&a_empty[s]), 1);
    }
    mbar_init(smem_u32(&w_full[0]), 1);
    mbar_init(smem_u32(&w_full[1]), 1);
    mbar_init(smem_u32(&tmem_full), 1);
    mbar_init(smem_u32(&b_loaded), 32 * FC_WORKERS);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&tmem_slot), FC_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  const bool is_producer = (warp == 0 && lane == 0);
  const bool is_mma = (warp == 1 && lane == 0);
  const bool is_epi = warp >= 2;
  const int et = threadIdx.x - 64;  // 0..511 on the worker warps
  const int wi = warp - 2;          // worker warp index 0..15

  // weight slice of step ph into its buffer (producer thread only)
  auto issue_w = [&](int ph) {
    const FcPhase f = fc_phase(ph);
    const uint32_t wb = f.wbuf == 0 ? wx_base : wy_base;
    const uint32_t bar = smem_u32(&w_full[f.wbuf]);
    mbar_arrive_expect_tx(bar, (uint32_t)(f.np * f.kbs * 128));
    for (int kb = 0; kb < f.kbs; ++kb) {
      const int col = (rk * f.kbs + kb) * 64;
      const uint32_t dst = wb + (uint32_t)(kb * f.np * 128);
      if (ph == 1) {  // gate rows, then the matching up rows
        tma_load_2d(dst, &maps.w[1], col, cl * 96, bar);
        tma_load_2d(dst + 96 * 128, &maps.w[1], col, p.I + cl * 96, bar);
      } else {
        tma_load_2d(dst, &maps.w[ph], col, cl * f.np, bar);
      }
    }
  };

  if (is_producer) {
    for (int i = 0; i < 4; ++i) tma_prefetch_desc(&maps.w[i]);
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.a[1]);
    issue_w(p.phase_begin);  // weights never depend on the predecessor kernel
  }
  fc_stamp(8);
  // every CTA of the cluster must be running (its shared memory live) before remote stores arrive
  __syncwarp();
  fc_cluster_sync();
  pdl_wait();
  if (threadIdx.x == 0) pdl_launch_dependents();
  fc_stamp(9);
  unsigned long long bar_target = 0;
  if (threadIdx.x == 0) bar_target = (fc_ld_acquire_u64(p.gbar) / gridDim.x) * gridDim.x;

  int ait = 0;                 // A-ring k-blocks so far (same value in every role)
  int wuse[2] = {0, 0};        // completed uses of each weight buffer
  int nph = 0;                 // steps so far (TMEM-full parity)
  int nbuilt = 0;              // built steps so far (b_loaded parity)

  unsigned short h_pre = 0;
  for (int ph = p.phase_begin; ph < p.phase_end; ++ph) {
    const FcPhase f = fc_phase(ph);
    fc_stamp(ph);
    if (is_producer) {
      // The next step's weight slice (48-96 KB per SM) must not queue ahead of this step's latency-critical
      // loads on the SM's return path: it goes out after the A tile's loads (TMA steps) or after the
      // builders have their inputs (built steps).
      if (!f.built) {
        for (int kb = 0; kb < f.kbs; ++kb) {
          const int it = ait + kb, s = it % FC_A_STAGES;
          mbar_wait(smem_u32(&a_empty[s]), ((it / FC_A_STAGES) & 1) ^ 1);
          const uint32_t fb = smem_u32(&a_full[s]);
          mbar_arrive_expect_tx(fb, FC_A_BYTES);
          tma_load_2d(a_base + s * FC_A_BYTES, &maps.a[ph == 0 ? 0 : 1], (rk * f.kbs + kb) * 64, 0, fb);
        }
      } else {
        mbar_wait(smem_u32(&b_loaded), nbuilt & 1);
      }
      if (ph + 1 < p.phase_end) issue_w(ph + 1);  // into the buffer this step does not use
    } else if (is_mma) {
      const uint32_t idesc = umma_idesc_bf16(128, f.np);
      const uint32_t wb = f.wbuf == 0 ? wx_base : wy_base;
      mbar_wait(smem_u32(&w_full[f.wbuf]), wuse[f.wbuf] & 1);
      if (ph == 1) fc_stamp_any(11);
      for (int kb = 0; kb < f.kbs; ++kb) {
        const int it = ait + kb, s = it % FC_A_STAGES;
        mbar_wait(smem_u32(&a_full[s]), (it / FC_A_STAGES) & 1);
        tc_fence_after();
        const uint64_t da = umma_desc_sw128(a_base + s * FC_A_BYTES);
        const uint64_t db = umma_desc_sw128(wb + (uint32_t)(kb * f.np * 128));
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc_mma_bf16(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
        tc_commit(smem_u32(&a_empty[s]));
      }
      tc_commit(smem_u32(&tmem_full));
      if (ph == 1) fc_stamp_any(12);
    } else if (is_epi) {
      if (!f.built) {
        // residual stream value the step's tail will add to (warp = token, lane = column): fetched now
        const int tg0 = rk * 16 + wi;
        h_pre = 0;
        if (tg0 < T) h_pre = __ldcg(reinterpret_cast<const unsigned short*>(p.h + (size_t)tg0 * p.H + cl * 32 + lane));
      } else {
        // ---- A tile = RMSNorm(h) for this CTA's K range (256 columns), written in the MMA's swizzled layout.
        //      warp -> tokens 4*wi..4*wi+3; lane -> 16-byte chunk of the 512-byte row (k-block lane/8, chunk lane%8)
        const float* ssp = ph == 1 ? p.ss_a : p.ss_b;
        const bf16* gain = ph == 1 ? p.norm_w_a : p.norm_w_b;
        const int k0 = rk * 256 + lane * 8;
        uint4 hv[4];
        // lane = slice index: the 32 partial sums of squares of this warp's four tokens
        const float4 p4 = __ldcg(reinterpret_cast<const float4*>(ssp + lane * FC_TOK + 4 * wi));
        const float part[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) hv[q] = __ldcg(reinterpret_cast<const uint4*>(p.h + (size_t)(4 * wi + q) * p.H + k0));
        float g[8];
        unpack8(*reinterpret_cast<const uint4*>(gain + k0), g);
        float rstd[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) rstd[q] = __fdiv_rn(1.0f, __fsqrt_rn(warp_sum(part[q]) / (float)p.H + p.eps));
        {  // inputs are in registers: the producer may start the next step's weight prefetch
          float keep = rstd[3] + bf16lo(hv[3].w);
          asm volatile("" ::"f"(keep));
          mbar_arrive(smem_u32(&b_loaded));
        }
        if (ph == 1 && et == 0) fc_stamp_any(13);
        const int it = ait + (lane >> 3), s = it % FC_A_STAGES, ch = lane & 7;
        mbar_wait(smem_u32(&a_empty[s]), ((it / FC_A_STAGES) & 1) ^ 1);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int t = 4 * wi + q;
          float v[8];
          unpack8(hv[q], v);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = bf16_round(g[e] * bf16_round(v[e] * rstd[q]));
          *reinterpret_cast<uint4*>(a_ptr + s * FC_A_BYTES + t * 128 + ((ch ^ (t & 7)) << 4)) = pack8(v);
        }
        fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
        fc_epi_bar();
        if (et < f.kbs) mbar_arrive(smem_u32(&a_full[(ait + et) % FC_A_STAGES]));
        if (ph == 1 && et == 0) fc_stamp_any(14);
      }
      // ---- accumulator rows -> the CTA that owns their tokens (distributed shared memory) ----
      mbar_wait(smem_u32(&tmem_full), nph & 1);
      tc_fence_after();
      if (ph == 1 && et == 64) fc_stamp_any(15);
      if ((warp & 3) < 2) {  // TMEM lanes 0..63 are the real token rows; a warp can only read its own lane quadrant
        const int tok = (warp & 3) * 32 + lane;
        const uint32_t owner = (uint32_t)(tok >> 4);
        const uint32_t dst0 = fc_mapa(red_base + (uint32_t)(((rk * 16 + (tok & 15)) * f.np) * 4), owner);
        const uint32_t tsrc = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        // the eight eligible warps (two per column group) take the 32-column chunks round robin
        for (int c0 = (wi >> 2) * 32; c0 < f.np; c0 += 128) {
          uint32_t r[32];
          tmem_ld32(tsrc + (uint32_t)c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            fc_st_cluster_f4(dst0 + (uint32_t)((c0 + j) * 4), __uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                             __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        }
        if (ph == 1 && et == 64) fc_stamp_any(20);
      }
    }
    ait += f.kbs;
    wuse[f.wbuf] += 1;
    nph += 1;
    nbuilt += f.built;
    fc_stamp(16 + ph);
    tc_fence_before();
    __syncwarp();
    fc_cluster_sync_roles(is_epi && (warp & 3) < 2, is_epi);  // warp-uniform conditions
    tc_fence_after();
    fc_stamp(24 + ph);

    if (is_epi) {
      // ---- owner: add the four K-split partials in rank order, then the step's element-wise tail.
      //      warp = one of this CTA's 16 tokens, lanes across that token's columns ----
      const int tl = wi;
      const int tg = rk * 16 + tl;            // global token row
      const int slot_stride = 16 * f.np;      // floats per source rank
      const float* rrow = red + tl * f.np;
      if (ph == 0 || ph == 2) {
        // residual add (HF:...:327,333): h = bf16(h + bf16(sum)); and this slice's share of sum(h^2)
        float acc = rrow[lane];
#pragma unroll
        for (int j = 1; j < 4; ++j) acc += rrow[j * slot_stride + lane];
        float v = 0.f;
        if (tg < T) {
          v = bf16_round(__uint_as_float((uint32_t)h_pre << 16) + bf16_round(acc));
          reinterpret_cast<unsigned short*>(p.h)[(size_t)tg * p.H + cl * 32 + lane] = (unsigned short)(__float_as_uint(v) >> 16);
        }
        // the 16 sums of this CTA go out as ONE 64-byte store (two whole sectors): single-lane 4-byte stores
        // to lines that attention has since pushed out of L2 made the readers wait for a DRAM fill
        const float ss = warp_sum(v * v);
        if (lane == 0) ss_stage[tl] = ss;
        fc_epi_bar();
        if (wi == 0 && lane < 16) (ph == 0 ? p.ss_a : p.ss_b)[cl * FC_TOK + rk * 16 + lane] = ss_stage[lane];
      } else if (ph == 1) {
        // SwiGLU (HF:...:81-83): act = bf16(bf16(silu(gate)) * up); outputs lane, lane + 32, lane + 64
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const int j = q * 32 + lane;
          float g = rrow[j], u = rrow[96 + j];
#pragma unroll
          for (int k = 1; k < 4; ++k) {
            g += rrow[k * slot_stride + j];
            u += rrow[k * slot_stride + 96 + j];
          }
          if (tg < T) {
            const float ge = bf16_round(g), ue = bf16_round(u);
            const float sl = bf16_round(__fdiv_rn(ge, 1.0f + expf(-ge)));
            const float o = bf16_round(sl * ue);
            reinterpret_cast<unsigned short*>(p.act)[(size_t)tg * p.I + cl * 96 + j] = (unsigned short)(__float_as_uint(o) >> 16);
          }
        }
        asm volatile("fence.proxy.async;" ::: "memory");  // act is read through TMA by the down step
      } else {
        // QK-norm + RoPE + KV page write for head `cl` (HF:...:248-249,263-264,151-181,270-271); lane owns
        // dims 4*lane..4*lane+3 and its RoPE partner sits in lane ^ 16, exactly as qknorm_rope_kv_kernel
        constexpr int HD = 128;
        const int head = cl;
        const bool is_q = head < p.nq, is_k = !is_q && head < p.nq + p.nkv;
        float4 acc = *reinterpret_cast<const float4*>(rrow + lane * 4);
#pragma unroll
        for (int k = 1; k < 4; ++k) {
          const float4 a = *reinterpret_cast<const float4*>(rrow + k * slot_stride + lane * 4);
          acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
        }
        float v[4] = {bf16_round(acc.x), bf16_round(acc.y), bf16_round(acc.z), bf16_round(acc.w)};
        const int tgc = tg < T ? tg : 0;
        if (is_q || is_k) {
          float ss = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
          ss = warp_sum(ss);
          const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(ss / (float)HD + p.eps));
          const uint2 wraw = *reinterpret_cast<const uint2*>((is_q ? p.qw : p.kw) + lane * 4);
          const float wv[4] = {bf16lo(wraw.x), bf16hi(wraw.x), bf16lo(wraw.y), bf16hi(wraw.y)};
          const int pos = max(p.positions[tgc], 0);
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
        if (tg < T) {
          uint2 o;
          o.x = pack_bf16x2(v[0], v[1]);
          o.y = pack_bf16x2(v[2], v[3]);
          if (is_q) {
            *reinterpret_cast<uint2*>(p.qrot + ((size_t)tg * p.nq + head) * HD + lane * 4) = o;
          } else {
            const int slot = p.slots[tg];
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
    if (ph == 1 && et == 0) fc_stamp_any(21);
    if (ph + 1 < p.phase_end) {
      bar_target += gridDim.x;
      fc_grid_sync(p.gbar, bar_target);
    }
  }
  fc_stamp(10);
  // no CTA may exit while a cluster peer could still address its shared memory
  tc_fence_before();
  __syncthreads();
  fc_cluster_sync();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, FC_TMEM_COLS);
  }
}

int fused_chain_supported(int H, int I, int nq, int nkv, int hd, int T, int block_size) {
  return H == 1024 && I == 3072 && hd == 128 && nq + 2 * nkv == FC_CLUSTERS && nq * hd == 2048 && T >= 1 && T <= FC_TOK &&
         block_size == 16;
}

int launch_decode_fused_chain(const CUtensorMap* w_maps4, const CUtensorMap* a_maps2, const FusedChainParams& p, cudaStream_t s) {
  B2_REQUIRE(fused_chain_supported(p.H, p.I, p.nq, p.nkv, 128, p.T, p.block_size), "fused chain: shapes not built");
  B2_REQUIRE(p.phase_begin >= 0 && p.phase_end <= 4 && p.phase_begin < p.phase_end, "fused chain: bad step range");
  FusedMaps maps;
  for (int i = 0; i < 4; ++i) maps.w[i] = w_maps4[i];
  for (int i = 0; i < 2; ++i) maps.a[i] = a_maps2[i];
  static B2PerDeviceOnce attr_once;
  if (attr_once.need())
    B2_CUDA_OK(cudaFuncSetAttribute(decode_fused_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FC_SMEM_BYTES));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(FC_CTAS);
  cfg.blockDim = dim3(FC_THREADS);
  cfg.dynamicSmemBytes = FC_SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = FC_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = b2_pdl_enabled() ? 2 : 1;
  B2_CUDA_OK(cudaLaunchKernelEx(&cfg, decode_fused_chain_kernel, maps, p));
  return 0;
}

// How many 4-CTA clusters of this kernel the device can hold at once: all 32 must be co-resident
// (the kernel spins on a grid-wide barrier).
int fused_chain_max_clusters() {
  if (cudaFuncSetAttribute(decode_fused_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FC_SMEM_BYTES) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(FC_CTAS);
  cfg.blockDim = dim3(FC_THREADS);
  cfg.dynamicSmemBytes = FC_SMEM_BYTES;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = FC_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, decode_fused_chain_kernel, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // namespace b2
