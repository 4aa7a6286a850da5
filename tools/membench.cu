// membench.cu -- what HBM read bandwidth can 148 x 8 warp-private TMA bulk-copy pipelines reach, as a
// function of chunk size and address pattern?  (Diagnostic for the decode-attention KV layout.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/membench tools/membench.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int WARPS, int STAGES, int CHUNK>  // CHUNK bytes per bulk copy, 2 copies per stage
__global__ void __launch_bounds__(WARPS * 32, 1) k(const uint8_t* base, size_t stride_a, size_t off_b, int tiles_per_warp, int pattern,
                                                   size_t total_tiles, unsigned* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[WARPS][STAGES];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int w = 0; w < WARPS; ++w) for (int s = 0; s < STAGES; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bars[w][s])), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const size_t gw = (size_t)blockIdx.x * WARPS + warp;
  uint8_t* wsm = smem + (size_t)warp * STAGES * 2 * CHUNK;
  auto addr = [&](int n, int which) -> const uint8_t* {
    size_t t = gw * tiles_per_warp + n;           // tile index on this warp's run
    if (pattern == 1) t = (t * 2654435761ull) % total_tiles;  // scattered pages
    // pattern 0/1: two separate regions (K pool, V pool); pattern 2: K and V adjacent (one 2*CHUNK span)
    if (pattern == 2) return base + t * 2 * CHUNK + which * CHUNK;
    return base + (which ? off_b : 0) + t * stride_a;
  };
  auto issue = [&](int n) {
    const int st = n % STAGES;
    const uint32_t bar = s32(&bars[warp][st]);
    const uint32_t dst = s32(wsm + st * 2 * CHUNK);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(2 * CHUNK) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(addr(n, 0)), "r"(CHUNK), "r"(bar) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst + CHUNK), "l"(addr(n, 1)), "r"(CHUNK), "r"(bar) : "memory");
  };
  if (lane == 0) for (int n = 0; n < STAGES && n < tiles_per_warp; ++n) issue(n);
  unsigned acc = 0;
  for (int n = 0; n < tiles_per_warp; ++n) {
    const int st = n % STAGES;
    const uint32_t bar = s32(&bars[warp][st]);
    uint32_t ok = 0;
    while (!ok) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p; }" : "=r"(ok) : "r"(bar), "r"((n / STAGES) & 1) : "memory");
    acc += *reinterpret_cast<const unsigned*>(wsm + st * 2 * CHUNK + lane * 16);
    __syncwarp();
    if (lane == 0 && n + STAGES < tiles_per_warp) issue(n + STAGES);
  }
  if (acc == 0x12345678u) sink[0] = acc;
}
template <int WARPS, int STAGES, int CHUNK>
void run(const char* name, uint8_t* buf, size_t bytes, int pattern, unsigned* sink) {
  const int grid = 148;
  const size_t total_bytes = bytes ? bytes : (size_t)170 << 20;  // default ~ one layer of C2 K+V
  const size_t total_tiles = total_bytes / (2 * CHUNK);
  const int tpw = (int)(total_tiles / (grid * WARPS));
  const int smem = WARPS * STAGES * 2 * CHUNK;
  CK(cudaFuncSetAttribute(k<WARPS, STAGES, CHUNK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9;
  for (int it = 0; it < 8; ++it) {
    // no dirty-line flush (its write-backs would compete with the reads): slide the window instead so
    // nothing read here is L2-resident (L2 = 126 MB, window step >= 2x the working set)
    const size_t span = total_bytes + ((size_t)64 << 20);
    const size_t nwin = (((size_t)4 << 30) - span) / span;
    const uint8_t* base = buf + (nwin > 0 ? (size_t)(it % nwin) * span : 0);
    cudaEventRecord(e0);
    k<WARPS, STAGES, CHUNK><<<grid, WARPS * 32, smem>>>(base, (size_t)CHUNK, total_bytes / 2, tpw, pattern, total_tiles, sink);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (it > 0 && ms < best) best = ms;
  }
  const double gb = (double)tpw * grid * WARPS * 2 * CHUNK / 1e9;
  printf("%-44s warps %2d stages %d chunk %5d B  in-flight/SM %3d KB : %7.1f us  %7.1f GB/s\n", name, WARPS, STAGES, CHUNK, smem >> 10, best * 1e3, gb / (best * 1e-3));
}
int main() {
  uint8_t* buf; unsigned* sink;
  CK(cudaMalloc(&buf, (size_t)4 << 30)); CK(cudaMalloc(&sink, 64));
  CK(cudaMemset(buf, 1, (size_t)4 << 30));
  run<8, 3, 4096>("seq, K|V far apart, 43 MB", buf, (size_t)43 << 20, 0, sink);
  run<8, 3, 4096>("seq, K|V far apart, 85 MB", buf, (size_t)85 << 20, 0, sink);
  run<8, 3, 4096>("seq, K|V far apart, 340 MB", buf, (size_t)340 << 20, 0, sink);
  run<8, 3, 4096>("seq, K|V far apart, 680 MB", buf, (size_t)680 << 20, 0, sink);
  run<8, 3, 4096>("seq, K|V far apart, 2720 MB", buf, (size_t)2720 << 20, 0, sink);
  run<8, 3, 4096>("scattered, 2720 MB", buf, (size_t)2720 << 20, 1, sink);
  run<8, 3, 4096>("seq adjacent, 2720 MB", buf, (size_t)2720 << 20, 2, sink);
  run<8, 2, 4096>("seq adjacent 2 stages, 2720 MB", buf, (size_t)2720 << 20, 2, sink);
  run<4, 3, 4096>("seq adjacent 4 warps, 2720 MB", buf, (size_t)2720 << 20, 2, sink);
  run<8, 3, 4096>("sequential per warp, K|V far apart", buf, 0, 0, sink);
  run<8, 3, 4096>("scattered 4 KB pages, K|V far apart", buf, 0, 1, sink);
  run<8, 3, 4096>("sequential, K|V adjacent (8 KB spans)", buf, 0, 2, sink);
  run<4, 6, 4096>("sequential, K|V adjacent", buf, 0, 2, sink);
  run<8, 2, 4096>("sequential, K|V adjacent", buf, 0, 2, sink);
  run<8, 1, 8192>("sequential, 16 KB stages", buf, 0, 2, sink);
  run<6, 2, 8192>("sequential, 16 KB stages", buf, 0, 2, sink);
  run<12, 2, 4096>("sequential, K|V adjacent", buf, 0, 2, sink);
  run<16, 3, 2048>("sequential, 2 KB chunks", buf, 0, 2, sink);
  run<13, 2, 4096>("sequential, K|V adjacent", buf, 0, 2, sink);
  run<8, 3, 4096>("scattered, K|V adjacent", buf, 0, 1, sink);
  return 0;
}
