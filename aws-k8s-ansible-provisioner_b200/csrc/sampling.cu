// sampling.cu -- temperature / top-k / top-p sampling over the bf16 logits of the last forward
// (the API surface of /v1/completions; vLLM:entrypoints/openai/completion/protocol.py:57-96).
// One CTA per sequence.  top-k and top-p are both "largest threshold such that enough
// count / probability mass lies at or above it", found by building the 32-bit orderable key of
// the threshold bit by bit (32 passes over the L2-resident logits row); the draw itself is a
// Gumbel-max over the surviving tokens, so no sort and no prefix scan are needed.
#include "kernels.h"

namespace b2 {

__device__ __forceinline__ uint32_t okey(float v) {
  uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}

template <typename T>
__device__ T block_sum(T v, T* red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  T t = 0;
  for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += red[k];
  return t;
}

__global__ void __launch_bounds__(1024)
sample_kernel(const bf16* __restrict__ logits, int32_t* __restrict__ out, int V, float temperature, int top_k, float top_p,
              uint64_t seed, uint64_t step) {
  __shared__ float redf[32];
  __shared__ int redi[32];
  __shared__ unsigned long long redu[32];
  const bf16* row = logits + (size_t)blockIdx.x * V;
  const float invT = 1.0f / temperature;

  // max (for a stable softmax)
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < V; i += blockDim.x) mx = fmaxf(mx, __bfloat162float(row[i]) * invT);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) redf[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = redf[0];
  for (int k = 1; k < (int)(blockDim.x >> 5); ++k) mx = fmaxf(mx, redf[k]);
  float z = 0.f;
  for (int i = threadIdx.x; i < V; i += blockDim.x) z += __expf(__bfloat162float(row[i]) * invT - mx);
  z = block_sum<float>(z, redf);

  uint32_t thr = 0;  // keys >= thr survive
  if (top_k > 0 && top_k < V) {
    uint32_t cand = 0;
    for (int bit = 31; bit >= 0; --bit) {
      const uint32_t c = cand | (1u << bit);
      int cnt = 0;
      for (int i = threadIdx.x; i < V; i += blockDim.x) cnt += okey(__bfloat162float(row[i])) >= c;
      cnt = block_sum<int>(cnt, redi);
      if (cnt >= top_k) cand = c;
    }
    thr = cand;
  }
  if (top_p > 0.f && top_p < 1.f) {
    uint32_t cand = 0;
    for (int bit = 31; bit >= 0; --bit) {
      const uint32_t c = cand | (1u << bit);
      float mass = 0.f;
      for (int i = threadIdx.x; i < V; i += blockDim.x) {
        const float l = __bfloat162float(row[i]);
        if (okey(l) >= c) mass += __expf(l * invT - mx);
      }
      mass = block_sum<float>(mass, redf);
      if (mass >= top_p * z) cand = c;
    }
    thr = cand > thr ? cand : thr;
  }
  // Gumbel-max draw among survivors
  unsigned long long best = 0;
  for (int i = threadIdx.x; i < V; i += blockDim.x) {
    const float l = __bfloat162float(row[i]);
    if (okey(l) < thr || !(l == l)) continue;
    const uint64_t r = mix64(seed ^ mix64(step * 0x100000001b3ull + blockIdx.x) ^ ((uint64_t)i << 20));
    const float u = ((float)(r >> 40) + 0.5f) * (1.0f / 16777216.0f);
    const float gsc = l * invT - __logf(-__logf(u));
    const unsigned long long p = ((unsigned long long)okey(gsc) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)i);
    best = p > best ? p : best;
  }
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
    best = other > best ? other : best;
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) redu[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k) best = redu[k] > best ? redu[k] : best;
    out[blockIdx.x] = (int32_t)(0xffffffffu - (uint32_t)(best & 0xffffffffull));
  }
}

int launch_sample(const bf16* logits, int32_t* out, int nseq, int V, float temperature, int top_k, float top_p, uint64_t seed,
                  uint64_t step, cudaStream_t s) {
  if (nseq <= 0) return 0;
  sample_kernel<<<nseq, 1024, 0, s>>>(logits, out, V, temperature, top_k, top_p, seed, step);
  B2_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace b2
