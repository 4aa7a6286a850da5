// gemm_simt.cu -- bandwidth-bound GEMV / skinny GEMM for very small decode batches
// (HF nn.Linear, models/qwen3/modeling_qwen3.py:236-247: y = x W^T, bf16 in, fp32 accumulate,
// one rounding to bf16).  One warp owns one weight row; every lane streams 16-byte pieces of
// the row with L1-bypassing loads (the row is read once) and keeps up to TB token
// accumulators.  For T > TB the weight row is re-read per token chunk -- the tensor-core
// kernel (gemm_tc.cu) is the path for larger batches.
#include "kernels.h"

namespace b2 {

template <int TB, bool OUT_F32>
__global__ void __launch_bounds__(128) linear_simt_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                          void* __restrict__ out, int T, int N, int K) {
  pdl_wait();
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 4 + warp;
  if (n >= N) return;
  const uint4* wrow = reinterpret_cast<const uint4*>(w + (size_t)n * K);
  const int iters = K / 256;  // 32 lanes x 8 elements per iteration
  for (int t0 = 0; t0 < T; t0 += TB) {
    float acc[TB];
#pragma unroll
    for (int j = 0; j < TB; ++j) acc[j] = 0.f;
#pragma unroll 4
    for (int it = 0; it < iters; ++it) {
      const int v = it * 32 + lane;
      float wf[8];
      unpack8(ld_stream(wrow + v), wf);
#pragma unroll
      for (int j = 0; j < TB; ++j) {
        if (t0 + j < T) {
          float xf[8];
          unpack8(reinterpret_cast<const uint4*>(x + (size_t)(t0 + j) * K)[v], xf);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[j] = fmaf(wf[e], xf[e], acc[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < TB; ++j) {
      float r = warp_sum(acc[j]);
      if (lane == 0 && t0 + j < T) {
        if (OUT_F32)
          reinterpret_cast<float*>(out)[(size_t)(t0 + j) * N + n] = r;
        else
          reinterpret_cast<bf16*>(out)[(size_t)(t0 + j) * N + n] = __float2bfloat16_rn(r);
      }
    }
  }
}

int launch_linear_simt(const bf16* x, const bf16* w, void* out, int out_is_f32, int T, int N, int K, cudaStream_t s) {
  if (T <= 0) return 0;
  B2_REQUIRE(K % 256 == 0, "linear_simt: K=%d must be a multiple of 256", K);
  const int blocks = (N + 3) / 4;
#define B2_LAUNCH(TB)                                                                       \
  do {                                                                                      \
    if (out_is_f32)                                                                         \
      B2_CUDA_OK(b2_launch(linear_simt_kernel<TB, true>, dim3(blocks), dim3(128), 0, s, x, w, out, T, N, K));  \
    else                                                                                    \
      B2_CUDA_OK(b2_launch(linear_simt_kernel<TB, false>, dim3(blocks), dim3(128), 0, s, x, w, out, T, N, K)); \
  } while (0)
  if (T == 1)
    B2_LAUNCH(1);
  else if (T == 2)
    B2_LAUNCH(2);
  else if (T <= 4)
    B2_LAUNCH(4);
  else
    B2_LAUNCH(8);
#undef B2_LAUNCH
  return 0;
}

}  // namespace b2
