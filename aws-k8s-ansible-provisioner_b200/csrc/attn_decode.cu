// attn_decode.cu -- paged GQA decode attention, flash-decoding split-K
// (HF:integrations/sdpa_attention.py:40-104 at T == 1; eager form HF:models/qwen3/
// modeling_qwen3.py:196-218).  This is the HBM-dominant kernel of batch-64 decode: per step it
// reads every resident K and V page once (B * S * 114688 B per step).
//
// Work split: one CTA per (kv-split, kv head, sequence); both query heads of the GQA group
// share every K/V page, so each page is fetched once per group.  Inside a CTA each of the 4
// warps runs its own TMA pipeline: lane 0 issues two 4 KB bulk copies (cp.async.bulk, one K
// page-slice and one V page-slice of [16 tokens][128 dims]) per stage onto an mbarrier, three
// stages deep, while the warp computes on the previous stage:
//   scores  s_j = (q . k_j) * scale      fp32 accumulate over bf16 inputs
//   online softmax in fp32; p_j = exp(s_j - m) is ROUNDED TO bf16 before P.V (what torch's CPU
//   flash kernel does for reduced-precision inputs), the row sum uses the unrounded fp32 p_j
//   O += p_j * v_j  in fp32; final O * (1 / l) rounded once to bf16.
// Warp partials merge through shared memory, CTA partials (nsplit > 1) through a global fp32
// scratch that attn_combine_kernel reduces in split order (deterministic, no atomics).
#include <mutex>

#include "../../include/b2_engine.h"
#include "kernels.h"

namespace b2 {

constexpr int AD_WARPS = 4;
constexpr int AD_STAGES = 3;
constexpr int AD_BS = 16;    // tokens per KV page
constexpr int AD_HD = 128;   // head dim
constexpr int AD_TILE = AD_BS * AD_HD * 2;       // 4096 B: one page-slice of K (or V) for one kv head
constexpr int AD_STAGE = 2 * AD_TILE;            // K + V
constexpr int AD_SMEM = AD_WARPS * AD_STAGES * AD_STAGE;  // 96 KB

__global__ void __launch_bounds__(AD_WARPS * 32)
attn_decode_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k_pages, const bf16* __restrict__ v_pages,
                   const int32_t* __restrict__ block_table, int bt_stride, const int32_t* __restrict__ seq_lens,
                   float* __restrict__ scratch, bf16* __restrict__ out, int B, int nq, int nkv, int nsplit, float scale) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[AD_WARPS][AD_STAGES];
  __shared__ float sm_m[AD_WARPS][2], sm_l[AD_WARPS][2];
  __shared__ float sm_o[AD_WARPS][2][AD_HD];

  pdl_launch_dependents();
  pdl_wait();
  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = seq_lens[b];
  const int nblk = (S + AD_BS - 1) / AD_BS;
  const int bps = (nblk + nsplit - 1) / nsplit;
  const int blk0 = split * bps;
  const int blk1 = min(nblk, blk0 + bps);
  const int nb_w = (blk1 - blk0 > warp) ? (blk1 - blk0 - warp + AD_WARPS - 1) / AD_WARPS : 0;

  if (threadIdx.x == 0) {
    for (int w = 0; w < AD_WARPS; ++w)
      for (int s = 0; s < AD_STAGES; ++s) mbar_init(smem_u32(&bars[w][s]), 1);
    mbar_fence_init();
  }
  __syncthreads();

  // thread mapping of the QK phase: 4 lanes per token (part = 32-dim slice), 8 tokens per pass
  const int part = lane & 3, tok8 = lane >> 2;
  const int rot = 2 * (part >> 1) + (tok8 & 1);  // per-thread chunk rotation -> conflict-free LDS.128
  float qreg[2][32];
  {
    const bf16* qb = q + ((size_t)b * nq + (size_t)kvh * 2) * AD_HD;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = 4 * part + ((i + rot) & 3);
#pragma unroll
      for (int h = 0; h < 2; ++h) unpack8(*reinterpret_cast<const uint4*>(qb + h * AD_HD + c * 8), &qreg[h][i * 8]);
    }
  }

  const int32_t* bt = block_table + (size_t)b * bt_stride;
  uint8_t* wsm = smem + (size_t)warp * AD_STAGES * AD_STAGE;
  int cached_group = -1, cached_blk = 0;
  auto get_blk = [&](int i) -> int {  // physical page of this warp's i-th tile (all lanes call)
    const int g = i >> 5;
    if (g != cached_group) {
      const int ii = (g << 5) + lane;
      cached_blk = (ii < nb_w) ? bt[blk0 + warp + AD_WARPS * ii] : 0;
      cached_group = g;
    }
    return __shfl_sync(0xffffffffu, cached_blk, i & 31);
  };
  auto issue = [&](int stage, int phys) {  // lane 0 only
    const uint32_t bar = smem_u32(&bars[warp][stage]);
    const uint32_t dst = smem_u32(wsm + stage * AD_STAGE);
    const size_t off = ((size_t)phys * nkv + kvh) * (AD_BS * AD_HD);
    mbar_arrive_expect_tx(bar, AD_STAGE);
    bulk_g2s(dst, k_pages + off, AD_TILE, bar);
    bulk_g2s(dst + AD_TILE, v_pages + off, AD_TILE, bar);
  };

  for (int i = 0; i < AD_STAGES; ++i) {
    const int phys = get_blk(i);
    if (i < nb_w && lane == 0) issue(i, phys);
  }

  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  float acc[2][8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[0][e] = acc[1][e] = 0.f;
  const int half = lane >> 4, chunk = lane & 15;

  for (int i = 0; i < nb_w; ++i) {
    const int stage = i % AD_STAGES;
    mbar_wait(smem_u32(&bars[warp][stage]), (uint32_t)((i / AD_STAGES) & 1));
    const uint8_t* kbase = wsm + stage * AD_STAGE;
    const uint8_t* vbase = kbase + AD_TILE;
    const int tok_base = (blk0 + warp + AD_WARPS * i) * AD_BS;

    // ---- scores for 16 tokens x 2 query heads ----
    float sc[2][2];
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
      const int tk = pass * 8 + tok8;
      float d0 = 0.f, d1 = 0.f;
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) {
        const int c = 4 * part + ((ii + rot) & 3);
        float kf[8];
        unpack8(*reinterpret_cast<const uint4*>(kbase + tk * (AD_HD * 2) + c * 16), kf);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          d0 = fmaf(qreg[0][ii * 8 + e], kf[e], d0);
          d1 = fmaf(qreg[1][ii * 8 + e], kf[e], d1);
        }
      }
      d0 += __shfl_xor_sync(0xffffffffu, d0, 1);
      d1 += __shfl_xor_sync(0xffffffffu, d1, 1);
      d0 += __shfl_xor_sync(0xffffffffu, d0, 2);
      d1 += __shfl_xor_sync(0xffffffffu, d1, 2);
      const bool valid = tok_base + tk < S;
      sc[pass][0] = valid ? d0 * scale : -INFINITY;
      sc[pass][1] = valid ? d1 * scale : -INFINITY;
    }
    // ---- online softmax (running max / sum per head, warp-uniform max) ----
    float pb[2][2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float mx = fmaxf(sc[0][h], sc[1][h]);
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
      const float m_new = fmaxf(m_run[h], mx);  // finite: every tile holds >= 1 valid token
      const float corr = expf(m_run[h] - m_new);
      const float p0 = expf(sc[0][h] - m_new), p1 = expf(sc[1][h] - m_new);
      l_run[h] = l_run[h] * corr + (p0 + p1);
      m_run[h] = m_new;
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[h][e] *= corr;
      pb[0][h] = bf16_round(p0);
      pb[1][h] = bf16_round(p1);
    }
    // ---- O += P . V : half-warp per token, 16 lanes x 8 dims ----
#pragma unroll
    for (int i2 = 0; i2 < 8; ++i2) {
      const int tk = 2 * i2 + half;
      const int src = 4 * (tk & 7);
      const float p0 = __shfl_sync(0xffffffffu, pb[i2 >> 2][0], src);
      const float p1 = __shfl_sync(0xffffffffu, pb[i2 >> 2][1], src);
      float vf[8];
      unpack8(*reinterpret_cast<const uint4*>(vbase + tk * (AD_HD * 2) + chunk * 16), vf);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        acc[0][e] = fmaf(p0, vf[e], acc[0][e]);
        acc[1][e] = fmaf(p1, vf[e], acc[1][e]);
      }
    }
    __syncwarp();
    const int nxt = i + AD_STAGES;
    const int phys = get_blk(nxt);
    if (nxt < nb_w && lane == 0) issue(stage, phys);
  }

  // ---- merge: lanes -> warp -> CTA ----
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 4);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 8);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 16);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[h][e] += __shfl_xor_sync(0xffffffffu, acc[h][e], 16);
  }
  if (lane == 0) {
    sm_m[warp][0] = m_run[0]; sm_m[warp][1] = m_run[1];
    sm_l[warp][0] = l_run[0]; sm_l[warp][1] = l_run[1];
  }
  if (lane < 16) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int e = 0; e < 8; ++e) sm_o[warp][h][chunk * 8 + e] = acc[h][e];
  }
  __syncthreads();
  const int d = threadIdx.x;  // 128 threads <-> 128 dims
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    float M = sm_m[0][h];
#pragma unroll
    for (int w = 1; w < AD_WARPS; ++w) M = fmaxf(M, sm_m[w][h]);
    float O = 0.f, L = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int w = 0; w < AD_WARPS; ++w) {
        const float f = expf(sm_m[w][h] - M);
        O = fmaf(f, sm_o[w][h][d], O);
        L = fmaf(f, sm_l[w][h], L);
      }
    }
    if (nsplit == 1) {
      const float inv = L > 0.f ? __fdiv_rn(1.0f, L) : 0.f;
      out[((size_t)b * nq + kvh * 2 + h) * AD_HD + d] = __float2bfloat16_rn(O * inv);
    } else {
      const size_t unit = (((size_t)b * nkv + kvh) * nsplit + split) * 2 + h;
      scratch[unit * AD_HD + d] = O;
      if (d == 0) {
        float* ml = scratch + (size_t)B * nkv * nsplit * 2 * AD_HD;
        ml[unit * 2 + 0] = M;
        ml[unit * 2 + 1] = L;
      }
    }
  }
}

// Reduces the split partials of one (sequence, query head) in split order.
__global__ void __launch_bounds__(AD_HD)
attn_combine_kernel(const float* __restrict__ scratch, bf16* __restrict__ out, int B, int nq, int nkv, int nsplit) {
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x / nq, qh = blockIdx.x % nq;
  const int kvh = qh / 2, h = qh % 2;
  const int d = threadIdx.x;
  const float* ml = scratch + (size_t)B * nkv * nsplit * 2 * AD_HD;
  const size_t unit0 = (((size_t)b * nkv + kvh) * nsplit) * 2 + h;
  float M = -INFINITY;
  for (int s = 0; s < nsplit; ++s) M = fmaxf(M, ml[(unit0 + (size_t)s * 2) * 2]);
  float O = 0.f, L = 0.f;
  if (M > -INFINITY) {
    for (int s = 0; s < nsplit; ++s) {
      const size_t u = unit0 + (size_t)s * 2;
      const float f = expf(ml[u * 2] - M);
      O = fmaf(f, scratch[u * AD_HD + d], O);
      L = fmaf(f, ml[u * 2 + 1], L);
    }
  }
  const float inv = L > 0.f ? __fdiv_rn(1.0f, L) : 0.f;
  out[((size_t)b * nq + qh) * AD_HD + d] = __float2bfloat16_rn(O * inv);
}

size_t attn_decode_scratch_floats(int B, int nq, int nkv, int hd, int nsplit) {
  if (nsplit <= 1) return 0;
  return (size_t)B * nkv * nsplit * (nq / nkv) * (hd + 2);
}

int launch_attn_decode(const bf16* q, const bf16* k_pages, const bf16* v_pages, const int32_t* block_table,
                       int bt_stride, const int32_t* seq_lens, float* scratch, bf16* out, int B, int nq, int nkv,
                       int hd, int block_size, int nsplit, cudaStream_t s) {
  if (B <= 0) return 0;
  B2_REQUIRE(hd == AD_HD && block_size == AD_BS, "attn_decode: built for head_dim 128 / page 16 (got %d / %d)", hd, block_size);
  B2_REQUIRE(nq == 2 * nkv, "attn_decode: built for a GQA group of 2 (nq=%d nkv=%d)", nq, nkv);
  B2_REQUIRE(nsplit >= 1 && (nsplit == 1 || scratch != nullptr), "attn_decode: nsplit=%d needs scratch", nsplit);
  static B2PerDeviceOnce attr_once;
  if (attr_once.need()) B2_CUDA_OK(cudaFuncSetAttribute(attn_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AD_SMEM));
  dim3 grid(nsplit, nkv, B);
  const float scale = (float)(1.0 / sqrt((double)hd));  // HF passes head_dim**-0.5 as a double
  B2_CUDA_OK(b2_launch(attn_decode_kernel, grid, dim3(AD_WARPS * 32), (size_t)AD_SMEM, s, q, k_pages, v_pages, block_table,
                       bt_stride, seq_lens, scratch, out, B, nq, nkv, nsplit, scale));
  return 0;
}

int launch_attn_combine(const float* scratch, bf16* out, int B, int nq, int nkv, int hd, int nsplit, cudaStream_t s) {
  if (B <= 0 || nsplit <= 1) return 0;
  B2_REQUIRE(hd == AD_HD, "attn_combine: head_dim %d not built", hd);
  B2_CUDA_OK(b2_launch(attn_combine_kernel, dim3(B * nq), dim3(AD_HD), 0, s, scratch, out, B, nq, nkv, nsplit));
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
// For them the split partials live in the file-scope device array below.  It is part of the
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
                                       int block_size, int nsplit, void* stream) {
  using namespace b2;
  B2_REQUIRE(q && k_pages && v_pages && block_table && seq_lens && out, "b2k_attn_decode_noalloc: null pointer");
  B2_REQUIRE(B >= 1 && nsplit >= 1 && nq >= 1 && nkv >= 1, "b2k_attn_decode_noalloc: bad sizes");
  const size_t need = attn_decode_scratch_floats(B, nq, nkv, hd, nsplit);
  B2_REQUIRE(need <= (size_t)B2K_ATTN_NOALLOC_MAX_FLOATS,
             "b2k_attn_decode_noalloc: batch %d x %d kv heads x %d splits needs %zu floats of scratch, the built-in array holds %d; "
             "use fewer splits or b2k_attn_decode with your own buffer",
             B, nkv, nsplit, need, (int)B2K_ATTN_NOALLOC_MAX_FLOATS);
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // Nothing in the array has to start at zero: every slot the combine pass reads is written by the
  // attention pass of the same call first, so what an earlier call left behind is never observed.
  rc = launch_attn_decode((const bf16*)q, (const bf16*)k_pages, (const bf16*)v_pages, block_table, bt_stride, seq_lens, scratch,
                          (bf16*)out, B, nq, nkv, hd, block_size, nsplit, s);
  if (rc != 0) return rc;
  if (nsplit > 1) rc = launch_attn_combine(scratch, (bf16*)out, B, nq, nkv, hd, nsplit, s);
  return rc;
}
