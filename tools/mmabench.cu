// Microbenchmark (GPU): cost of a stream of tcgen05.mma 128 x N x 16 (bf16, both operands in shared memory, 128-byte
// swizzle) issued back to back by one thread, (a) all into ONE TMEM accumulator, (b) round robin over FOUR.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I aws-k8s-ansible-provisioner_b200/csrc tools/mmabench.cu -o /tmp/mmabench && /tmp/mmabench
#include "common.cuh"
void b2_set_error(const char*, ...) {}
bool b2_pdl_enabled() { return false; }

template <int N, int NACC>
__global__ void __launch_bounds__(128, 1) mma_bench(long long* out, int nmma) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_slot), 512); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t da = umma_desc_sw128(base), db = umma_desc_sw128(base + 16384);
    for (int rep = 0; rep < 3; ++rep) {
      long long t0 = clock64();
      for (int i = 0; i < nmma; ++i) {
        const int a = i % NACC;
        tc_mma_bf16(tm + (uint32_t)(a * N), da + (uint64_t)(2 * (i & 3)), db + (uint64_t)(2 * (i & 3)), idesc, i >= NACC ? 1u : 0u);
      }
      long long t1 = clock64();
      tc_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), rep & 1);
      long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

template <int N, int NACC>
void run(long long* d, int nmma) {
  cudaFuncSetAttribute(mma_bench<N, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 32768 + 1024);
  mma_bench<N, NACC><<<1, 128, 16384 + 32768 + 1024>>>(d, nmma);
  long long h[2];
  cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  printf("N=%3d accumulators=%d  %3d MMAs: issue %6lld cyc, issue+complete %6lld cyc = %.1f cyc per MMA (%s)\n", N, NACC, nmma, h[0], h[1],
         (double)h[1] / nmma, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  run<64, 1>(d, 16); run<64, 1>(d, 64); run<64, 2>(d, 64); run<64, 4>(d, 64);
  run<32, 1>(d, 64); run<32, 4>(d, 64);
  run<16, 1>(d, 64); run<16, 4>(d, 64);
  run<128, 1>(d, 64); run<128, 2>(d, 64); run<128, 4>(d, 64);
  run<256, 1>(d, 64); run<256, 2>(d, 64);
  return 0;
}
