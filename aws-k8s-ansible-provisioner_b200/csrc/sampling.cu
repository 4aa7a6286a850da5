// sampling.cu -- temperature / top-k / top-p sampling over the bf16 logits of the last forward
// (the API surface of /v1/completions; vLLM:entrypoints/openai/completion/protocol.py:57-96).
// One CTA per sequence.  top-k and top-p are both "largest threshold such that enough
// count / probability mass lies at or above it", found by building the orderable key of the
// threshold bit by bit (passes over the L2-resident logits row); the draw itself is a
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

// One CTA per listed row, all rows in ONE launch (a launch per row left 147 SMs idle: measured 76 ms per decode step
// for 64 sampled requests).  The logits are bf16, so the orderable 32-bit key has 16 significant bits: 16 threshold
// passes, not 32.  Rows are read as 16-byte vectors (V % 8 == 0 is checked by the launcher).
struct SampleRowDev {
  int row;            // row of the logits matrix
  float temperature;  // <= 0: greedy (lowest index on ties)
  int top_k;
  float top_p;
  unsigned long long seed;
};

__global__ void __launch_bounds__(1024)
sample_rows_kernel(const bf16* __restrict__ logits, const SampleRowDev* __restrict__ rows, int32_t* __restrict__ out, int V,
                   uint64_t step) {
  __shared__ float redf[32];
  __shared__ int redi[32];
  __shared__ unsigned long long redu[32];
  const SampleRowDev rp = rows[blockIdx.x];
  const uint4* row = reinterpret_cast<const uint4*>(logits + (size_t)rp.row * V);
  const int nvec = V / 8;
  const bool greedy = rp.temperature <= 0.f;
  const float invT = greedy ? 1.0f : 1.0f / rp.temperature;
  float mx = -INFINITY, z = 1.f;
  uint32_t thr = 0;  // keys >= thr survive
  if (!greedy) {
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      float v[8];
      unpack8(row[i], v);
#pragma unroll
      for (int e = 0; e < 8; ++e) mx = fmaxf(mx, v[e] * invT);
    }
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) redf[threadIdx.x >> 5] = mx;
    __syncthreads();
    mx = redf[0];
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k) mx = fmaxf(mx, redf[k]);
    const bool want_p = rp.top_p > 0.f && rp.top_p < 1.f;
    if (rp.top_k > 0 && rp.top_k < V) {
      uint32_t cand = 0;
      for (int bit = 31; bit >= 16; --bit) {
        const uint32_t c = cand | (1u << bit);
        int cnt = 0;
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
          float v[8];
          unpack8(row[i], v);
#pragma unroll
          for (int e = 0; e < 8; ++e) cnt += okey(v[e]) >= c;
        }
        cnt = block_sum<int>(cnt, redi);
        if (cnt >= rp.top_k) cand = c;
      }
      thr = cand;
    }
    if (want_p) {
      // top-p is taken over the distribution that is left after top-k (vLLM applies top-k first and renormalises):
      // z is the mass of the top-k survivors, not of the whole vocabulary
      z = 0.f;
      for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
        float v[8];
        unpack8(row[i], v);
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (okey(v[e]) >= thr) z += __expf(v[e] * invT - mx);
      }
      z = block_sum<float>(z, redf);
      uint32_t cand = 0;
      for (int bit = 31; bit >= 16; --bit) {
        const uint32_t c = cand | (1u << bit);
        const uint32_t lo = c > thr ? c : thr;  // only top-k survivors carry mass
        float mass = 0.f;
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
          float v[8];
          unpack8(row[i], v);
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (okey(v[e]) >= lo) mass += __expf(v[e] * invT - mx);
        }
        mass = block_sum<float>(mass, redf);
        if (mass >= rp.top_p * z) cand = c;
      }
      thr = cand > thr ? cand : thr;
    }
  }
  // greedy: arg-max of the logits; otherwise a Gumbel-max draw among the survivors
  unsigned long long best = 0;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    float v[8];
    unpack8(row[i], v);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float l = v[e];
      const int idx = i * 8 + e;
      if (!(l == l) || okey(l) < thr) continue;
      float score = l;
      if (!greedy) {
        // A request that gives a seed gets draws that depend on (seed, token index) only -- the caller advances the
        // seed per generated token -- so its output does not depend on what else is in the batch or on how long the
        // process has run; seed 0 means "no seed given" and mixes in the launch counter and the row.
        const uint64_t salt = rp.seed != 0ull ? 0ull : mix64(step * 0x100000001b3ull + (uint64_t)rp.row);
        const uint64_t r = mix64(rp.seed ^ salt ^ ((uint64_t)idx << 20));
        // u strictly inside (0, 1): 23 random bits + 0.5 is exact in fp32 (24 bits + 0.5 rounds up to 2^24, i.e. u == 1,
        // whose Gumbel score is +inf: that token would win whatever its probability)
        const float u = ((float)(r >> 41) + 0.5f) * (1.0f / 8388608.0f);
        score = l * invT - __logf(-__logf(u));
      }
      const unsigned long long pk = ((unsigned long long)okey(score) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)idx);
      best = pk > best ? pk : best;
    }
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

// rows_host: nrows entries; staged through `stage_dev` (device, >= nrows entries) on the stream.
int launch_sample_rows(const bf16* logits, const SampleRow* rows_host, SampleRow* stage_dev, int32_t* out, int nrows, int V, uint64_t step,
                       cudaStream_t s) {
  if (nrows <= 0) return 0;
  static_assert(sizeof(SampleRow) == sizeof(SampleRowDev), "host and device row records must match");
  B2_REQUIRE(V % 8 == 0, "sampling: vocabulary %d must be a multiple of 8", V);
  B2_CUDA_OK(cudaMemcpyAsync(stage_dev, rows_host, sizeof(SampleRow) * (size_t)nrows, cudaMemcpyHostToDevice, s));
  sample_rows_kernel<<<nrows, 1024, 0, s>>>(logits, reinterpret_cast<const SampleRowDev*>(stage_dev), out, V, step);
  B2_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace b2
