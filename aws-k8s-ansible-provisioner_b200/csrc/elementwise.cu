// elementwise.cu -- the small HBM/latency-bound kernels between the GEMMs and attention:
// embedding gather, residual-add + RMSNorm, QK-norm + RoPE + KV page write, SwiGLU gate,
// argmax helpers and the device-side decode bookkeeping.  Each kernel states the HF line it
// reproduces and where it rounds to bf16 (the rounding points are the parity contract).
#include "kernels.h"

namespace b2 {

B2_STEP_TRACE_DEFINE(step_trace_set_elementwise)

// ---------------------------------------------------------------------------------------
// a1: nn.Embedding lookup, HF:models/qwen3/modeling_qwen3.py:363,392
// ---------------------------------------------------------------------------------------
__global__ void embed_kernel(const int32_t* __restrict__ ids, const bf16* __restrict__ table, bf16* __restrict__ h,
                             int H, int vocab, unsigned long long* __restrict__ zero_buf, int nzero) {
  B2_STAMP_BEGIN();
  pdl_wait();
  pdl_launch_dependents();
  B2_STAMP_WAIT();
  int t = blockIdx.x;
  // first kernel of every forward: also clears the fused arg-max slots the lm_head will atomicMax into
  if (t == 0 && zero_buf != nullptr)
    for (int i = threadIdx.x; i < nzero; i += blockDim.x) zero_buf[i] = 0ull;
  int id = ids[t];
  if (id < 0 || id >= vocab) id = 0;
  const uint4* src = reinterpret_cast<const uint4*>(table + (size_t)id * H);
  uint4* dst = reinterpret_cast<uint4*>(h + (size_t)t * H);
  for (int i = threadIdx.x; i < H / 8; i += blockDim.x) dst[i] = src[i];
  B2_STAMP_END(B2K_EMBED);
}

int launch_embed(const int32_t* ids, const bf16* table, bf16* h, int T, int H, int vocab, unsigned long long* zero_buf, int nzero,
                 cudaStream_t s) {
  if (T <= 0) return 0;
  static B2PerDeviceOnce c0;
  b2_prefer_max_smem(embed_kernel, c0);
  B2_CUDA_OK(b2_launch(embed_kernel, dim3(T), dim3(128), 0, s, ids, table, h, H, vocab, zero_buf, nzero));
  return 0;
}

__global__ void gather_rows_kernel(const bf16* __restrict__ src, const int32_t* __restrict__ rows, bf16* __restrict__ dst,
                                   int H) {
  pdl_wait();
  pdl_launch_dependents();
  int r = rows[blockIdx.x];
  const uint4* s = reinterpret_cast<const uint4*>(src + (size_t)r * H);
  uint4* d = reinterpret_cast<uint4*>(dst + (size_t)blockIdx.x * H);
  for (int i = threadIdx.x; i < H / 8; i += blockDim.x) d[i] = s[i];
}
int launch_gather_rows(const bf16* src, const int32_t* rows, bf16* dst, int n, int H, cudaStream_t s) {
  if (n <= 0) return 0;
  B2_CUDA_OK(b2_launch(gather_rows_kernel, dim3(n), dim3(128), 0, s, src, rows, dst, H));
  return 0;
}

// ---------------------------------------------------------------------------------------
// a2 + a9: residual add (HF:...:327,333) fused with Qwen3RMSNorm.forward (HF:...:59-64).
//   o      = bf16(sum of split-K partials)            (nn.Linear rounds its output)
//   h      = bf16(h + o)                              (bf16 tensor add)
//   x      = bf16(w * bf16(float(h) * rsqrt(mean(h^2) + eps)))
// One CTA per token, H/8 threads, one 16-byte vector per thread.
// ---------------------------------------------------------------------------------------
template <int MODE>  // 0: no delta, 1: bf16 delta, 2: fp32 partials
__global__ void add_rmsnorm_kernel(bf16* __restrict__ h, const void* __restrict__ delta, int nsplit, size_t split_stride,
                                   const bf16* __restrict__ w, bf16* __restrict__ x, int H, float eps, int early, int tag) {
  __shared__ float red[32];
  B2_STAMP_BEGIN();
  // The kernel that follows a norm is always a projection, and all a projection does before ITS dependency wait is
  // start the TMA stream of its weight tiles (constants).  Releasing it at the very start of this kernel -- before
  // this kernel's own wait -- gives that stream the whole hand-off + norm time instead of the ~1.5 us it had, which
  // was less than a 64-96 KB weight slice takes to arrive (the projection then issued its last MMA ~2 us after its
  // wait; tools/step_timeline.py).  The projection triggers ITS dependents only after its own wait, so the
  // discipline stated in common.cuh still holds for every other kernel.
  if (early & 2) pdl_launch_dependents();
  if (!(early & 1)) pdl_wait();
  const int t = blockIdx.x;
  const int i = threadIdx.x;  // vector index, 8 elements
  // Before the dependency wait: the norm gains (constant) and, when a delta is added, the residual
  // stream h (last written several kernels ago).  With no delta, h comes from the embedding kernel
  // that runs immediately before, so it is read after the wait.
  float wv[8];
  unpack8(reinterpret_cast<const uint4*>(w)[i], wv);
  uint4 hv = make_uint4(0, 0, 0, 0);
  if (MODE != 0) hv = reinterpret_cast<const uint4*>(h + (size_t)t * H)[i];
  pdl_wait();
  if (!(early & 2)) pdl_launch_dependents();
  B2_STAMP_WAIT();
  if (MODE == 0) hv = reinterpret_cast<const uint4*>(h + (size_t)t * H)[i];
  float v[8];
  unpack8(hv, v);
  if (MODE == 1) {
    float d[8];
    unpack8(reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(delta) + (size_t)t * H)[i], d);
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = bf16_round(v[e] + d[e]);
  } else if (MODE == 2) {
    float d[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) d[e] = 0.f;
    const float* p = reinterpret_cast<const float*>(delta) + (size_t)t * H + (size_t)i * 8;
    // all loads of a group of 8 splits (the split-K cap) are issued before any add (the adds stay in split order)
    for (int s0 = 0; s0 < nsplit; s0 += 8) {
      float4 a[8], b[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (s0 + j < nsplit) {
          a[j] = reinterpret_cast<const float4*>(p + (size_t)(s0 + j) * split_stride)[0];
          b[j] = reinterpret_cast<const float4*>(p + (size_t)(s0 + j) * split_stride)[1];
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (s0 + j < nsplit) {
          d[0] += a[j].x; d[1] += a[j].y; d[2] += a[j].z; d[3] += a[j].w;
          d[4] += b[j].x; d[5] += b[j].y; d[6] += b[j].z; d[7] += b[j].w;
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = bf16_round(v[e] + bf16_round(d[e]));
  }
  if (MODE != 0) reinterpret_cast<uint4*>(h + (size_t)t * H)[i] = pack8(v);

  float ss = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) ss += v[e] * v[e];
  ss = warp_sum(ss);
  const int nw = blockDim.x >> 5;
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
  for (int k = 0; k < nw; ++k) tot += red[k];
  const float var = tot / (float)H;
  const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(var + eps));
#pragma unroll
  for (int e = 0; e < 8; ++e) v[e] = bf16_round(wv[e] * bf16_round(v[e] * rstd));
  reinterpret_cast<uint4*>(x + (size_t)t * H)[i] = pack8(v);
  B2_STAMP_END_AUX(B2K_NORM, (unsigned)tag);
}

int launch_add_rmsnorm(bf16* h, ActIn delta, const bf16* w, bf16* x, int T, int H, float eps, cudaStream_t s, int tag) {
  if (T <= 0) return 0;
  B2_REQUIRE(H % 256 == 0 && H <= 8192, "add_rmsnorm: hidden %d must be a multiple of 256 and <= 8192", H);
  const int threads = H / 8;
  static B2PerDeviceOnce c0, c1, c2;
  b2_prefer_max_smem(add_rmsnorm_kernel<0>, c0);
  b2_prefer_max_smem(add_rmsnorm_kernel<1>, c1);
  b2_prefer_max_smem(add_rmsnorm_kernel<2>, c2);
  if (delta.ptr == nullptr)
    B2_CUDA_OK(b2_launch(add_rmsnorm_kernel<0>, dim3(T), dim3(threads), 0, s, h, (const void*)nullptr, 0, (size_t)0, w, x, H, eps, ((b2_pdl_early_mask() & 2) ? 1 : 0) | (b2_pdl_trigger_early() ? 2 : 0), tag));
  else if (!delta.is_f32)
    B2_CUDA_OK(b2_launch(add_rmsnorm_kernel<1>, dim3(T), dim3(threads), 0, s, h, delta.ptr, 1, (size_t)0, w, x, H, eps, ((b2_pdl_early_mask() & 2) ? 1 : 0) | (b2_pdl_trigger_early() ? 2 : 0), tag));
  else
    B2_CUDA_OK(b2_launch(add_rmsnorm_kernel<2>, dim3(T), dim3(threads), 0, s, h, delta.ptr, delta.nsplit, (size_t)T * H, w, x, H,
                         eps, ((b2_pdl_early_mask() & 2) ? 1 : 0) | (b2_pdl_trigger_early() ? 2 : 0), tag));
  return 0;
}

// ---------------------------------------------------------------------------------------
// a4 + a5 + a6: per-head RMSNorm of q and k (HF:...:248-249,263-264), RoPE with the half-split
// rotate (HF:...:151-181; cos/sin were rounded to bf16, HF:...:148) and the KV page write
// (HF:...:270-271).  One warp per (token, head); lane l owns elements 4l..4l+3, so the RoPE
// partner element i +/- 64 lives in lane l ^ 16.
//   n  = bf16(w * bf16(q * rsqrt(mean(q^2)+eps)))
//   q' = bf16(bf16(n*cos) + bf16(rot(n)*sin))
// Variants of the same step (Phi-3, HF:models/phi3/modeling_phi3.py:178-205,237-248): no QK-norm
// (qw == kw == nullptr) and a rotary width other than 128 -- rot_quads = rotary_dim / 8 lanes hold
// the first half, the next rot_quads lanes the second half, later lanes pass through (q_pass);
// heads narrower than 128 are stored zero-padded, and zeros stay zeros through every step here.
// OPT (HF:models/opt/modeling_opt.py:151-157): neither norm nor rotary (rot_quads == 0), but biased
// projections and a pre-scaled query: v = bf16(partials + bias), q = bf16(v * q_mul).
// ---------------------------------------------------------------------------------------
template <bool F32IN>
__global__ void qknorm_rope_kv_kernel(const void* __restrict__ qkv, int nsplit, size_t split_stride,
                                      const bf16* __restrict__ qw, const bf16* __restrict__ kw,
                                      const bf16* __restrict__ cos_sin, const int32_t* __restrict__ positions,
                                      const int32_t* __restrict__ slots, bf16* __restrict__ q_out,
                                      bf16* __restrict__ k_pages, bf16* __restrict__ v_pages, int T, int nq, int nkv,
                                      int block_size, float eps, int early, int rot_quads, const bf16* __restrict__ bias, float q_mul, int tag) {
  constexpr int HD = 128;
  B2_STAMP_BEGIN();
  if (!early) pdl_wait();
  const int nheads = nq + 2 * nkv;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= T * nheads) {
    pdl_wait();
    pdl_launch_dependents();
    return;
  }
  const int t = warp / nheads;
  const int head = warp % nheads;
  const size_t row = (size_t)t * nheads * HD + (size_t)head * HD + lane * 4;
  const bool is_q = head < nq;
  const bool is_k = !is_q && head < nq + nkv;
  // Before the dependency wait: everything that is not the QKV projection's output -- position and
  // slot of the token (written at the top of the step), the RoPE table and the norm gains.
  const int pos = max(positions[t], 0);
  const int slot = slots[t];
  float wv[4] = {1.f, 1.f, 1.f, 1.f}, c[4] = {1.f, 1.f, 1.f, 1.f}, sn[4] = {0.f, 0.f, 0.f, 0.f};
  const bool use_norm = qw != nullptr;
  const bool first_half = lane < rot_quads;
  const bool rotates = lane < 2 * rot_quads;
  const int partner_lane = first_half ? lane + rot_quads : lane - rot_quads;  // == lane ^ 16 for a 128-wide rotary
  if (is_q || is_k) {
    if (use_norm) {
      uint2 wraw = *reinterpret_cast<const uint2*>((is_q ? qw : kw) + lane * 4);
      wv[0] = bf16lo(wraw.x); wv[1] = bf16hi(wraw.x); wv[2] = bf16lo(wraw.y); wv[3] = bf16hi(wraw.y);
    }
    if (rotates) {
      const bf16* cs = cos_sin + (size_t)pos * HD;
      const int ci = (first_half ? lane : lane - rot_quads) * 4;
      uint2 craw = *reinterpret_cast<const uint2*>(cs + ci);
      uint2 sraw = *reinterpret_cast<const uint2*>(cs + 64 + ci);
      c[0] = bf16lo(craw.x); c[1] = bf16hi(craw.x); c[2] = bf16lo(craw.y); c[3] = bf16hi(craw.y);
      sn[0] = bf16lo(sraw.x); sn[1] = bf16hi(sraw.x); sn[2] = bf16lo(sraw.y); sn[3] = bf16hi(sraw.y);
    }
  }
  float bq[4] = {0.f, 0.f, 0.f, 0.f};
  if (F32IN && bias != nullptr) {
    uint2 braw = *reinterpret_cast<const uint2*>(bias + (size_t)head * HD + lane * 4);
    bq[0] = bf16lo(braw.x); bq[1] = bf16hi(braw.x); bq[2] = bf16lo(braw.y); bq[3] = bf16hi(braw.y);
  }
  pdl_wait();
  pdl_launch_dependents();
  B2_STAMP_WAIT();

  float v[4];
  if (F32IN) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* p = reinterpret_cast<const float*>(qkv) + row;
    // all loads of a group of 4 splits go out before any add (the adds stay in split order): one L2 round trip
    // per group instead of one per split
    for (int s0 = 0; s0 < nsplit; s0 += 4) {
      float4 a[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (s0 + j < nsplit) a[j] = *reinterpret_cast<const float4*>(p + (size_t)(s0 + j) * split_stride);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (s0 + j < nsplit) { acc.x += a[j].x; acc.y += a[j].y; acc.z += a[j].z; acc.w += a[j].w; }
    }
    acc.x += bq[0]; acc.y += bq[1]; acc.z += bq[2]; acc.w += bq[3];  // zeros unless the projection has a bias
    v[0] = bf16_round(acc.x); v[1] = bf16_round(acc.y); v[2] = bf16_round(acc.z); v[3] = bf16_round(acc.w);
  } else {
    uint2 a = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(qkv) + row);
    v[0] = bf16lo(a.x); v[1] = bf16hi(a.x); v[2] = bf16lo(a.y); v[3] = bf16hi(a.y);
  }

  if (is_q || is_k) {
    if (use_norm) {
      float ss = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
      ss = warp_sum(ss);
      const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(ss / (float)HD + eps));
#pragma unroll
      for (int e = 0; e < 4; ++e) v[e] = bf16_round(wv[e] * bf16_round(v[e] * rstd));
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float partner = __shfl_sync(0xffffffffu, v[e], partner_lane & 31);
      float rot = first_half ? -partner : partner;
      if (rotates) v[e] = bf16_round(bf16_round(v[e] * c[e]) + bf16_round(rot * sn[e]));
    }
  }
  if (is_q && q_mul != 1.0f) {
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] = bf16_round(v[e] * q_mul);
  }
  uint2 o;
  o.x = pack_bf16x2(v[0], v[1]);
  o.y = pack_bf16x2(v[2], v[3]);
  if (is_q) {
    *reinterpret_cast<uint2*>(q_out + ((size_t)t * nq + head) * HD + lane * 4) = o;
  } else if (slot >= 0) {
    const int kvh = is_k ? head - nq : head - nq - nkv;
    const int blk = slot / block_size, off = slot % block_size;
    bf16* base = is_k ? k_pages : v_pages;
    // lane l holds dims 4l..4l+3 = half of 16-byte chunk l/2; pages are stored chunk-swizzled
    const int chunk = kv_swizzled_chunk(off, lane >> 1);
    *reinterpret_cast<uint2*>(base + (((size_t)blk * nkv + kvh) * block_size + off) * HD + chunk * 8 + (lane & 1) * 4) = o;
  }
  B2_STAMP_END_AUX(B2K_ROPE_KV, (unsigned)tag);
}

// Prefill-sized variant of the kernel above (same arithmetic, same rounding points): with thousands of tokens the
// one-head-per-warp kernel is bound by latency x occupancy (a 256-byte row per warp and a chain of dependent loads:
// measured 82 us for 8192 tokens, 134 MB moved).  Here a warp takes FOUR consecutive heads of one token: position,
// slot and the cos/sin row are fetched once, the four rows are requested together, then processed one by one.
// bf16 input only (a prefill-sized projection is never split along K); needs (nq + 2 nkv) % 4 == 0.
__global__ void qknorm_rope_kv_prefill_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ qw, const bf16* __restrict__ kw,
                                              const bf16* __restrict__ cos_sin, const int32_t* __restrict__ positions,
                                              const int32_t* __restrict__ slots, bf16* __restrict__ q_out,
                                              bf16* __restrict__ k_pages, bf16* __restrict__ v_pages, int T, int nq, int nkv,
                                              int block_size, float eps, int rot_quads) {
  constexpr int HD = 128;
  pdl_wait();
  pdl_launch_dependents();
  const int nheads = nq + 2 * nkv, groups = nheads / 4;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= T * groups) return;
  const int t = warp / groups, head0 = (warp % groups) * 4;
  const bool use_norm = qw != nullptr;
  const bool first_half = lane < rot_quads, rotates = lane < 2 * rot_quads;
  const int partner_lane = first_half ? lane + rot_quads : lane - rot_quads;
  uint2 raw[4];
#pragma unroll
  for (int h = 0; h < 4; ++h)
    raw[h] = *reinterpret_cast<const uint2*>(qkv + (size_t)t * nheads * HD + (size_t)(head0 + h) * HD + lane * 4);
  const int pos = max(positions[t], 0);
  const int slot = slots[t];
  float c[4] = {1.f, 1.f, 1.f, 1.f}, sn[4] = {0.f, 0.f, 0.f, 0.f}, wq4[4] = {1.f, 1.f, 1.f, 1.f}, wk4[4] = {1.f, 1.f, 1.f, 1.f};
  if (rotates) {
    const bf16* cs = cos_sin + (size_t)pos * HD;
    const int ci = (first_half ? lane : lane - rot_quads) * 4;
    const uint2 craw = *reinterpret_cast<const uint2*>(cs + ci);
    const uint2 sraw = *reinterpret_cast<const uint2*>(cs + 64 + ci);
    c[0] = bf16lo(craw.x); c[1] = bf16hi(craw.x); c[2] = bf16lo(craw.y); c[3] = bf16hi(craw.y);
    sn[0] = bf16lo(sraw.x); sn[1] = bf16hi(sraw.x); sn[2] = bf16lo(sraw.y); sn[3] = bf16hi(sraw.y);
  }
  if (use_norm) {
    const uint2 a = *reinterpret_cast<const uint2*>(qw + lane * 4);
    const uint2 b = *reinterpret_cast<const uint2*>(kw + lane * 4);
    wq4[0] = bf16lo(a.x); wq4[1] = bf16hi(a.x); wq4[2] = bf16lo(a.y); wq4[3] = bf16hi(a.y);
    wk4[0] = bf16lo(b.x); wk4[1] = bf16hi(b.x); wk4[2] = bf16lo(b.y); wk4[3] = bf16hi(b.y);
  }
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const int head = head0 + h;
    const bool is_q = head < nq, is_k = !is_q && head < nq + nkv;
    float v[4] = {bf16lo(raw[h].x), bf16hi(raw[h].x), bf16lo(raw[h].y), bf16hi(raw[h].y)};
    if (is_q || is_k) {
      if (use_norm) {
        float ss = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
        ss = warp_sum(ss);
        const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(ss / (float)HD + eps));
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = bf16_round((is_q ? wq4[e] : wk4[e]) * bf16_round(v[e] * rstd));
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float partner = __shfl_sync(0xffffffffu, v[e], partner_lane & 31);
        const float rot = first_half ? -partner : partner;
        if (rotates) v[e] = bf16_round(bf16_round(v[e] * c[e]) + bf16_round(rot * sn[e]));
      }
    }
    uint2 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    if (is_q) {
      *reinterpret_cast<uint2*>(q_out + ((size_t)t * nq + head) * HD + lane * 4) = o;
    } else if (slot >= 0) {
      const int kvh = is_k ? head - nq : head - nq - nkv;
      const int blk = slot / block_size, off = slot % block_size;
      bf16* base = is_k ? k_pages : v_pages;
      const int chunk = kv_swizzled_chunk(off, lane >> 1);
      *reinterpret_cast<uint2*>(base + (((size_t)blk * nkv + kvh) * block_size + off) * HD + chunk * 8 + (lane & 1) * 4) = o;
    }
  }
}

// Same contract, 16 lanes per head (16 bytes each), the bf16 x bf16 products as packed HMUL2 (one rounding, identical to
// rounding the exact fp32 product).  The four-byte-lane kernel above issued ~120 instructions per head per lane and ran
// at a third of the HBM rate (61 us for 8192 tokens x 32 heads); this one is bound by the 134 MB it moves.
// A half-warp walks four heads of one token; rot16 = rotary_dim / 16 lanes hold each half of the rotated span.
__global__ void __launch_bounds__(128)
qknorm_rope_kv_prefill16_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ qw, const bf16* __restrict__ kw,
                                const bf16* __restrict__ cos_sin, const int32_t* __restrict__ positions,
                                const int32_t* __restrict__ slots, bf16* __restrict__ q_out, bf16* __restrict__ k_pages,
                                bf16* __restrict__ v_pages, int T, int nq, int nkv, int block_size, float eps, int rot16) {
  constexpr int HD = 128;
  pdl_wait();
  pdl_launch_dependents();
  const int nheads = nq + 2 * nkv, groups = nheads / 4;
  const long long hw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
  const int lane = threadIdx.x & 31, l16 = lane & 15;
  const bool valid = hw < (long long)T * groups;
  const int t = valid ? (int)(hw / groups) : 0, head0 = valid ? (int)(hw % groups) * 4 : 0;
  const bool use_norm = qw != nullptr;
  const bool first_half = l16 < rot16, rotates = l16 < 2 * rot16;
  const int partner = (lane & 16) | ((first_half ? l16 + rot16 : l16 - rot16) & 15);
  const unsigned hmask = 0xffffu << (lane & 16);  // the two halves of a warp may hold heads of different kinds
  const uint32_t sign = first_half ? 0x80008000u : 0u;  // first half of the rotated span: rot = -partner
  uint4 raw[4];
#pragma unroll
  for (int h = 0; h < 4; ++h)
    raw[h] = *reinterpret_cast<const uint4*>(qkv + ((size_t)t * nheads + head0 + h) * HD + l16 * 8);
  const int pos = max(positions[t], 0);
  const int slot = slots[t];
  uint4 co4 = make_uint4(0, 0, 0, 0), sn4 = co4, wq4 = co4, wk4 = co4;
  if (rotates) {
    const bf16* cs = cos_sin + (size_t)pos * HD + (first_half ? l16 : l16 - rot16) * 8;
    co4 = *reinterpret_cast<const uint4*>(cs);
    sn4 = *reinterpret_cast<const uint4*>(cs + 64);
  }
  if (use_norm) {
    wq4 = *reinterpret_cast<const uint4*>(qw + l16 * 8);
    wk4 = *reinterpret_cast<const uint4*>(kw + l16 * 8);
  }
  const uint32_t* co2 = &co4.x; const uint32_t* sn2 = &sn4.x;
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const int head = head0 + h;
    const bool is_q = head < nq, is_k = !is_q && head < nq + nkv;
    uint32_t x2[4] = {raw[h].x, raw[h].y, raw[h].z, raw[h].w};
    if (is_q || is_k) {  // uniform over the half-warp; the shuffles name only its own lanes
      if (use_norm) {
        float v[8];
        unpack8(raw[h], v);
        // the per-head kernel's summation tree, so that both give the same bits: there lane 2l and 2l+1 hold this
        // lane's two quads and meet in the LAST butterfly step (xor 1), after the steps xor 16, 8, 4, 2
        float sa = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
        float sb = v[4] * v[4] + v[5] * v[5] + v[6] * v[6] + v[7] * v[7];
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
          sa += __shfl_xor_sync(hmask, sa, o);
          sb += __shfl_xor_sync(hmask, sb, o);
        }
        const float ss = sa + sb;
        const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(ss / (float)HD + eps));
        const uint32_t* w2 = is_q ? &wq4.x : &wk4.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const uint32_t nx = pack_bf16x2(v[2 * e] * rstd, v[2 * e + 1] * rstd);  // bf16(x * rstd)
          const __nv_bfloat162 n = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&w2[e]), *reinterpret_cast<const __nv_bfloat162*>(&nx));
          x2[e] = *reinterpret_cast<const uint32_t*>(&n);
        }
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        uint32_t rot = __shfl_sync(hmask, x2[e], partner) ^ sign;
        const __nv_bfloat162 a = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&x2[e]), *reinterpret_cast<const __nv_bfloat162*>(&co2[e]));
        const __nv_bfloat162 b = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&rot), *reinterpret_cast<const __nv_bfloat162*>(&sn2[e]));
        const uint32_t au = *reinterpret_cast<const uint32_t*>(&a), bu = *reinterpret_cast<const uint32_t*>(&b);
        if (rotates) x2[e] = pack_bf16x2(bf16lo(au) + bf16lo(bu), bf16hi(au) + bf16hi(bu));
      }
    }
    const uint4 o = make_uint4(x2[0], x2[1], x2[2], x2[3]);
    if (!valid) continue;
    if (is_q) {
      *reinterpret_cast<uint4*>(q_out + ((size_t)t * nq + head) * HD + l16 * 8) = o;
    } else if (slot >= 0) {
      const int kvh = is_k ? head - nq : head - nq - nkv;
      const int blk = slot / block_size, off = slot % block_size;
      bf16* base = (is_k ? k_pages : v_pages) + (((size_t)blk * nkv + kvh) * block_size + off) * HD;
      *reinterpret_cast<uint4*>(base + kv_swizzled_chunk(off, l16) * 8) = o;
    }
  }
}

int launch_qknorm_rope_kv(ActIn qkv, const bf16* qw, const bf16* kw, const bf16* cos_sin, const int32_t* positions,
                          const int32_t* slots, bf16* q_out, bf16* k_pages, bf16* v_pages, int T, int nq, int nkv,
                          int hd, int block_size, float eps, cudaStream_t s, int rotary_dim, const bf16* bias, float q_mul, int tag) {
  if (T <= 0) return 0;
  B2_REQUIRE(hd == 128, "qknorm_rope_kv: head_dim %d not built (128 only; narrower heads are stored padded)", hd);
  B2_REQUIRE((qw == nullptr) == (kw == nullptr), "qknorm_rope_kv: q and k norm gains must both be given or both be null");
  B2_REQUIRE(rotary_dim >= 0 && rotary_dim <= 128 && rotary_dim % 8 == 0, "qknorm_rope_kv: rotary width %d not supported", rotary_dim);
  B2_REQUIRE(bias == nullptr || qkv.is_f32, "qknorm_rope_kv: a bias needs the projection's fp32 partials (it is added before rounding)");
  const int rot_quads = rotary_dim / 8;
  const int nheads = nq + 2 * nkv;
  if (!qkv.is_f32 && bias == nullptr && q_mul == 1.0f && nheads % 4 == 0 && (long long)T * nheads >= 32768) {
    const long long w4 = (long long)T * (nheads / 4);
    static const bool lanes16 = getenv("B2_ROPE16") == nullptr || atoi(getenv("B2_ROPE16")) != 0;
    if (lanes16 && rotary_dim % 16 == 0) {  // two four-head groups per warp
      B2_CUDA_OK(b2_launch(qknorm_rope_kv_prefill16_kernel, dim3((unsigned)((w4 + 7) / 8)), dim3(128), 0, s, (const bf16*)qkv.ptr, qw, kw,
                           cos_sin, positions, slots, q_out, k_pages, v_pages, T, nq, nkv, block_size, eps, rotary_dim / 16));
      return 0;
    }
    B2_CUDA_OK(b2_launch(qknorm_rope_kv_prefill_kernel, dim3((unsigned)((w4 + 3) / 4)), dim3(128), 0, s, (const bf16*)qkv.ptr, qw, kw,
                         cos_sin, positions, slots, q_out, k_pages, v_pages, T, nq, nkv, block_size, eps, rot_quads));
    return 0;
  }
  const long long warps = (long long)T * nheads;
  const int blocks = (int)((warps + 3) / 4);
  static B2PerDeviceOnce c0, c1;
  b2_prefer_max_smem(qknorm_rope_kv_kernel<true>, c0);
  b2_prefer_max_smem(qknorm_rope_kv_kernel<false>, c1);
  if (qkv.is_f32)
    B2_CUDA_OK(b2_launch(qknorm_rope_kv_kernel<true>, dim3(blocks), dim3(128), 0, s, qkv.ptr, qkv.nsplit,
                         (size_t)T * nheads * hd, qw, kw, cos_sin, positions, slots, q_out, k_pages, v_pages, T, nq, nkv,
                         block_size, eps, b2_pdl_early_mask() & 4, rot_quads, bias, q_mul, tag));
  else
    B2_CUDA_OK(b2_launch(qknorm_rope_kv_kernel<false>, dim3(blocks), dim3(128), 0, s, qkv.ptr, 1, (size_t)0, qw, kw, cos_sin,
                         positions, slots, q_out, k_pages, v_pages, T, nq, nkv, block_size, eps, b2_pdl_early_mask() & 4, rot_quads, bias, q_mul, tag));
  return 0;
}

// ---------------------------------------------------------------------------------------
// a8 (middle): act = bf16( bf16(silu(gate)) * up ), HF:...:81-83.  gate/up are the bf16-rounded
// outputs of their Linear layers; silu is evaluated in fp32 as x / (1 + exp(-x)).
// ---------------------------------------------------------------------------------------
template <bool F32IN>
__global__ void silu_mul_kernel(const void* __restrict__ gu, int nsplit, size_t split_stride, bf16* __restrict__ act,
                                int T, int I, int tag, int trigger_early) {
  B2_STAMP_BEGIN();
  if (trigger_early) pdl_launch_dependents();  // the down projection starts streaming its weights (see add_rmsnorm_kernel)
  pdl_wait();
  if (!trigger_early) pdl_launch_dependents();
  B2_STAMP_WAIT();
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // vector of 8 over [T][I]
  const size_t nvec = (size_t)T * I / 8;
  if (idx >= nvec) return;
  const int t = (int)(idx / (I / 8));
  const int i = (int)(idx % (I / 8)) * 8;
  float g[8], u[8];
  if (F32IN) {
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] = u[e] = 0.f;
    const float* pg = reinterpret_cast<const float*>(gu) + (size_t)t * 2 * I + i;
    const float* pu = pg + I;
    for (int s0 = 0; s0 < nsplit; s0 += 4) {  // loads of up to 4 splits in flight together, adds in split order
      float4 a[4], b[4], c[4], d[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (s0 + j < nsplit) {
          a[j] = reinterpret_cast<const float4*>(pg + (size_t)(s0 + j) * split_stride)[0];
          b[j] = reinterpret_cast<const float4*>(pg + (size_t)(s0 + j) * split_stride)[1];
          c[j] = reinterpret_cast<const float4*>(pu + (size_t)(s0 + j) * split_stride)[0];
          d[j] = reinterpret_cast<const float4*>(pu + (size_t)(s0 + j) * split_stride)[1];
        }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (s0 + j < nsplit) {
          g[0] += a[j].x; g[1] += a[j].y; g[2] += a[j].z; g[3] += a[j].w; g[4] += b[j].x; g[5] += b[j].y; g[6] += b[j].z; g[7] += b[j].w;
          u[0] += c[j].x; u[1] += c[j].y; u[2] += c[j].z; u[3] += c[j].w; u[4] += d[j].x; u[5] += d[j].y; u[6] += d[j].z; u[7] += d[j].w;
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      g[e] = bf16_round(g[e]);
      u[e] = bf16_round(u[e]);
    }
  } else {
    const bf16* p = reinterpret_cast<const bf16*>(gu) + (size_t)t * 2 * I + i;
    unpack8(*reinterpret_cast<const uint4*>(p), g);
    unpack8(*reinterpret_cast<const uint4*>(p + I), u);
  }
  float o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    float sl = bf16_round(__fdiv_rn(g[e], 1.0f + expf(-g[e])));
    o[e] = bf16_round(sl * u[e]);
  }
  *reinterpret_cast<uint4*>(act + (size_t)t * I + i) = pack8(o);
  B2_STAMP_END_AUX(B2K_ACT, (unsigned)tag);
}

int launch_silu_mul(ActIn gu, bf16* act, int T, int I, cudaStream_t s, int tag) {
  if (T <= 0) return 0;
  B2_REQUIRE(I % 8 == 0, "silu_mul: intermediate %d must be a multiple of 8", I);
  const size_t nvec = (size_t)T * I / 8;
  const int blocks = (int)((nvec + 255) / 256);
  static B2PerDeviceOnce c0, c1;
  b2_prefer_max_smem(silu_mul_kernel<true>, c0);
  b2_prefer_max_smem(silu_mul_kernel<false>, c1);
  // two-lane schedule: four warps per CTA (one per warp scheduler, 96 registers) so that a CTA fits into the register
  // banks a decode-attention CTA leaves free (see linear_tc_kernel, SMALL)
  static const bool narrow = [] { const char* mb = getenv("B2_MB"); return mb != nullptr && atoi(mb) > 0; }();
  if (gu.is_f32 && narrow)
    B2_CUDA_OK(b2_launch(silu_mul_kernel<true>, dim3((unsigned)((nvec + 127) / 128)), dim3(128), 0, s, gu.ptr, gu.nsplit,
                         (size_t)T * 2 * I, act, T, I, tag, b2_pdl_trigger_early() ? 1 : 0));
  else if (gu.is_f32)
    B2_CUDA_OK(b2_launch(silu_mul_kernel<true>, dim3(blocks), dim3(256), 0, s, gu.ptr, gu.nsplit, (size_t)T * 2 * I, act, T, I, tag, b2_pdl_trigger_early() ? 1 : 0));
  else
    B2_CUDA_OK(b2_launch(silu_mul_kernel<false>, dim3(blocks), dim3(256), 0, s, gu.ptr, 1, (size_t)0, act, T, I, tag, b2_pdl_trigger_early() ? 1 : 0));
  return 0;
}

// ---------------------------------------------------------------------------------------
// a11: greedy argmax over bf16 logits, lowest index on ties (torch.argmax on CPU).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long argmax_pack(float v, int idx) {
  uint32_t b = __float_as_uint(v);
  uint32_t key = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  return ((unsigned long long)key << 32) | (unsigned long long)(0xffffffffu - (uint32_t)idx);
}

__global__ void argmax_rows_kernel(const bf16* __restrict__ logits, int32_t* __restrict__ out, int V) {
  __shared__ unsigned long long red[32];
  pdl_wait();
  pdl_launch_dependents();
  const bf16* row = logits + (size_t)blockIdx.x * V;
  unsigned long long best = 0;
  for (int i = threadIdx.x; i < V; i += blockDim.x) {
    float v = __bfloat162float(row[i]);
    if (v == v) {
      unsigned long long p = argmax_pack(v, i);
      best = p > best ? p : best;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
    best = other > best ? other : best;
  }
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k) best = red[k] > best ? red[k] : best;
    out[blockIdx.x] = (int32_t)(0xffffffffu - (uint32_t)(best & 0xffffffffull));
  }
}
int launch_argmax_rows(const bf16* logits, int32_t* out, int T, int V, cudaStream_t s) {
  if (T <= 0) return 0;
  B2_CUDA_OK(b2_launch(argmax_rows_kernel, dim3(T), dim3(1024), 0, s, logits, out, V));
  return 0;
}

__global__ void argmax_finalize_kernel(const unsigned long long* __restrict__ packed, int32_t* __restrict__ out, int T) {
  pdl_wait();
  pdl_launch_dependents();
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < T) out[t] = (int32_t)(0xffffffffu - (uint32_t)(packed[t] & 0xffffffffull));
}
int launch_argmax_finalize(const unsigned long long* packed, int32_t* out, int T, cudaStream_t s) {
  if (T <= 0) return 0;
  B2_CUDA_OK(b2_launch(argmax_finalize_kernel, dim3((T + 127) / 128), dim3(128), 0, s, packed, out, T));
  return 0;
}

// ---------------------------------------------------------------------------------------
// decode bookkeeping: one thread per sequence of the running batch.
// ---------------------------------------------------------------------------------------
__global__ void decode_advance_kernel(int32_t* positions, int32_t* seq_lens, int32_t* slots,
                                      const int32_t* __restrict__ block_table, int bt_stride, int block_size, int B,
                                      int32_t* frac_q16, long long l2_budget_bytes, long long bytes_per_page_all_heads) {
  __shared__ int total_pages;
  pdl_wait();
  pdl_launch_dependents();
  if (threadIdx.x == 0) total_pages = 0;
  __syncthreads();
  int my_pages = 0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {  // single CTA
    if (positions[b] < -1) {  // padding row of a graph bucket
      seq_lens[b] = 0;
      slots[b] = -1;
      continue;
    }
    int pos = positions[b] + 1;
    positions[b] = pos;
    seq_lens[b] = pos + 1;
    int blk = block_table[(size_t)b * bt_stride + pos / block_size];
    slots[b] = blk * block_size + pos % block_size;
    my_pages += (pos + block_size) / block_size;
  }
  if (my_pages) atomicAdd(&total_pages, my_pages);
  __syncthreads();
  if (threadIdx.x == 0 && frac_q16 != nullptr) {
    // share of one layer's K/V pages that fits the L2 budget, for the ride-along prefetch
    const long long bytes = (long long)total_pages * bytes_per_page_all_heads;
    long long f = bytes > 0 ? (l2_budget_bytes << 16) / bytes : 65536;
    *frac_q16 = (int32_t)(f > 65536 ? 65536 : f);
  }
}
int launch_decode_advance(int32_t* positions, int32_t* seq_lens, int32_t* slots, const int32_t* block_table,
                          int bt_stride, int block_size, int B, int32_t* frac_q16, long long l2_budget_bytes,
                          long long bytes_per_page_all_heads, cudaStream_t s) {
  if (B <= 0) return 0;
  B2_CUDA_OK(b2_launch(decode_advance_kernel, dim3(1), dim3(256), 0, s, positions, seq_lens, slots, block_table, bt_stride,
                       block_size, B, frac_q16, l2_budget_bytes, bytes_per_page_all_heads));
  return 0;
}

__global__ void logits_to_f32_kernel(const bf16* __restrict__ in, float* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = __bfloat162float(in[i]);
}
int launch_logits_to_f32(const bf16* in, float* out, size_t n, cudaStream_t s) {
  if (n == 0) return 0;
  logits_to_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, out, n);
  B2_CUDA_OK(cudaGetLastError());
  return 0;
}

int device_num_sms() {
  int dev = 0, n = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  return n;
}

}  // namespace b2
