// elementwise_opt.cu -- the small kernels of the OPT variant of the hot path (SURVEY.md 8a row a13,
// HF:models/opt/modeling_opt.py): learned positions, LayerNorm with bias, biased projections and a
// ReLU feed-forward.  The GEMMs and the attention kernels are shared with the other model families;
// these kernels consume the GEMMs' fp32 split-K partials, so every nn.Linear bias is added in fp32
// BEFORE the one rounding to bf16 that nn.Linear performs (the parity contract's rounding point).
#include "kernels.h"

namespace b2 {

// ---------------------------------------------------------------------------------------
// OPTDecoder.forward, HF:...:363-368: h = embed_tokens[id] + embed_positions[position + 2]
// (OPTLearnedPositionalEmbedding, HF:...:45-70: the table is offset by 2), one bf16 add.
// ---------------------------------------------------------------------------------------
__global__ void embed_pos_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ positions,
                                 const bf16* __restrict__ table, const bf16* __restrict__ pos_table, bf16* __restrict__ h, int H,
                                 int vocab, int pos_rows, unsigned long long* __restrict__ zero_buf, int nzero) {
  pdl_wait();
  pdl_launch_dependents();
  const int t = blockIdx.x;
  // first kernel of every forward: also clears the fused arg-max slots the lm_head will atomicMax into
  if (t == 0 && zero_buf != nullptr)
    for (int i = threadIdx.x; i < nzero; i += blockDim.x) zero_buf[i] = 0ull;
  int id = ids[t];
  if (id < 0 || id >= vocab) id = 0;
  int pr = max(positions[t], 0) + 2;  // padding rows of a decode bucket carry position -2
  if (pr >= pos_rows) pr = pos_rows - 1;
  const uint4* a = reinterpret_cast<const uint4*>(table + (size_t)id * H);
  const uint4* b = reinterpret_cast<const uint4*>(pos_table + (size_t)pr * H);
  uint4* dst = reinterpret_cast<uint4*>(h + (size_t)t * H);
  for (int i = threadIdx.x; i < H / 8; i += blockDim.x) {
    float x[8], y[8];
    unpack8(a[i], x);
    unpack8(b[i], y);
#pragma unroll
    for (int e = 0; e < 8; ++e) x[e] = bf16_round(x[e] + y[e]);
    dst[i] = pack8(x);
  }
}

int launch_embed_pos(const int32_t* ids, const int32_t* positions, const bf16* table, const bf16* pos_table, bf16* h, int T, int H,
                     int vocab, int pos_rows, unsigned long long* zero_buf, int nzero, cudaStream_t s) {
  if (T <= 0) return 0;
  B2_CUDA_OK(b2_launch(embed_pos_kernel, dim3(T), dim3(128), 0, s, ids, positions, table, pos_table, h, H, vocab, pos_rows, zero_buf,
                       nzero));
  return 0;
}

// ---------------------------------------------------------------------------------------
// Residual add (HF:...:224,246) of a biased projection, fused with the following nn.LayerNorm
// (HF:...:195-201,214,239; torch's CPU kernel for bf16: moments and the affine map in fp32, one
// rounding).
//   o = bf16(sum of split-K partials + bias)          (nn.Linear rounds its output once)
//   h = bf16(h + o)
//   x = bf16((h - mean) * rsqrt(var + eps) * w + b)    var = mean((h - mean)^2)
// One CTA per token, H/8 threads, one 16-byte vector per thread.
// ---------------------------------------------------------------------------------------
template <bool HAS_DELTA>
__global__ void add_layernorm_kernel(bf16* __restrict__ h, const float* __restrict__ delta, int nsplit, size_t split_stride,
                                     const bf16* __restrict__ delta_bias, const bf16* __restrict__ w, const bf16* __restrict__ b,
                                     bf16* __restrict__ x, int H, float eps) {
  __shared__ float red[2][32];
  const int t = blockIdx.x;
  const int i = threadIdx.x;
  float wv[8], bv[8], db[8];
  unpack8(reinterpret_cast<const uint4*>(w)[i], wv);
  unpack8(reinterpret_cast<const uint4*>(b)[i], bv);
  if (HAS_DELTA) unpack8(reinterpret_cast<const uint4*>(delta_bias)[i], db);
  pdl_wait();
  pdl_launch_dependents();
  float v[8];
  unpack8(reinterpret_cast<const uint4*>(h + (size_t)t * H)[i], v);
  if (HAS_DELTA) {
    float d[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) d[e] = 0.f;
    const float* p = delta + (size_t)t * H + (size_t)i * 8;
    for (int s = 0; s < nsplit; ++s) {  // split order: deterministic
      const float4 a0 = reinterpret_cast<const float4*>(p + (size_t)s * split_stride)[0];
      const float4 a1 = reinterpret_cast<const float4*>(p + (size_t)s * split_stride)[1];
      d[0] += a0.x; d[1] += a0.y; d[2] += a0.z; d[3] += a0.w;
      d[4] += a1.x; d[5] += a1.y; d[6] += a1.z; d[7] += a1.w;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = bf16_round(v[e] + bf16_round(d[e] + db[e]));
    reinterpret_cast<uint4*>(h + (size_t)t * H)[i] = pack8(v);
  }
  const int nw = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float s1 = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) s1 += v[e];
  s1 = warp_sum(s1);
  if (lane == 0) red[0][wid] = s1;
  __syncthreads();
  float tot = 0.f;
  for (int k = 0; k < nw; ++k) tot += red[0][k];
  const float mean = tot / (float)H;
  float s2 = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) s2 += (v[e] - mean) * (v[e] - mean);
  s2 = warp_sum(s2);
  if (lane == 0) red[1][wid] = s2;
  __syncthreads();
  float tot2 = 0.f;
  for (int k = 0; k < nw; ++k) tot2 += red[1][k];
  const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(tot2 / (float)H + eps));
#pragma unroll
  for (int e = 0; e < 8; ++e) v[e] = bf16_round((v[e] - mean) * rstd * wv[e] + bv[e]);
  reinterpret_cast<uint4*>(x + (size_t)t * H)[i] = pack8(v);
}

int launch_add_layernorm(bf16* h, ActIn delta, const bf16* delta_bias, const bf16* w, const bf16* b, bf16* x, int T, int H, float eps,
                         cudaStream_t s) {
  if (T <= 0) return 0;
  B2_REQUIRE(H % 256 == 0 && H <= 8192, "add_layernorm: hidden %d must be a multiple of 256 and <= 8192", H);
  const int threads = H / 8;
  if (delta.ptr == nullptr) {
    B2_CUDA_OK(b2_launch(add_layernorm_kernel<false>, dim3(T), dim3(threads), 0, s, h, (const float*)nullptr, 0, (size_t)0,
                         (const bf16*)nullptr, w, b, x, H, eps));
  } else {
    B2_REQUIRE(delta.is_f32 && delta_bias != nullptr, "add_layernorm: the delta must be fp32 partials plus a bias vector");
    B2_CUDA_OK(b2_launch(add_layernorm_kernel<true>, dim3(T), dim3(threads), 0, s, h, (const float*)delta.ptr, delta.nsplit,
                         (size_t)T * H, delta_bias, w, b, x, H, eps));
  }
  return 0;
}

// ---------------------------------------------------------------------------------------
// fc1 + activation, HF:...:241-242: act = bf16(relu(bf16(sum of partials + bias)))
// ---------------------------------------------------------------------------------------
__global__ void bias_relu_kernel(const float* __restrict__ part, int nsplit, size_t split_stride, const bf16* __restrict__ bias,
                                 bf16* __restrict__ act, int T, int I) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // vector of 8 over [T][I]
  const size_t nvec = (size_t)T * I / 8;
  float bv[8];
  const int i = idx < nvec ? (int)(idx % (I / 8)) * 8 : 0;
  unpack8(*reinterpret_cast<const uint4*>(bias + i), bv);
  pdl_wait();
  pdl_launch_dependents();
  if (idx >= nvec) return;
  const int t = (int)(idx / (I / 8));
  float d[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) d[e] = 0.f;
  const float* p = part + (size_t)t * I + i;
  for (int s = 0; s < nsplit; ++s) {
    const float4 a0 = reinterpret_cast<const float4*>(p + (size_t)s * split_stride)[0];
    const float4 a1 = reinterpret_cast<const float4*>(p + (size_t)s * split_stride)[1];
    d[0] += a0.x; d[1] += a0.y; d[2] += a0.z; d[3] += a0.w;
    d[4] += a1.x; d[5] += a1.y; d[6] += a1.z; d[7] += a1.w;
  }
  float o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] = fmaxf(bf16_round(d[e] + bv[e]), 0.f);
  *reinterpret_cast<uint4*>(act + (size_t)t * I + i) = pack8(o);
}

int launch_bias_relu(ActIn fc1, const bf16* bias, bf16* act, int T, int I, cudaStream_t s) {
  if (T <= 0) return 0;
  B2_REQUIRE(I % 8 == 0 && fc1.is_f32 && bias != nullptr, "bias_relu: needs fp32 partials, a bias and I %% 8 == 0");
  const size_t nvec = (size_t)T * I / 8;
  const int blocks = (int)((nvec + 255) / 256);
  B2_CUDA_OK(b2_launch(bias_relu_kernel, dim3(blocks), dim3(256), 0, s, (const float*)fc1.ptr, fc1.nsplit, (size_t)T * I, bias, act, T, I));
  return 0;
}

}  // namespace b2
