// common.cuh -- shared device helpers for the sm_100a kernels (bf16 packing, mbarrier,
// bulk/TMA copies, tcgen05 wrappers) and host error plumbing.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <string>

// ---------------------------------------------------------------------------------------
// host side error plumbing
// ---------------------------------------------------------------------------------------
void b2_set_error(const char* fmt, ...);
#define B2_CUDA_OK(expr)                                                                      \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      b2_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return -5;                                                                              \
    }                                                                                         \
  } while (0)
#define B2_REQUIRE(cond, ...)    \
  do {                           \
    if (!(cond)) {               \
      b2_set_error(__VA_ARGS__); \
      return -22;                \
    }                            \
  } while (0)

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------------------
// bf16 helpers.  A bf16 is the upper half of an fp32, so widening is a shift/mask.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
  f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 v;
  v.x = pack_bf16x2(f[0], f[1]); v.y = pack_bf16x2(f[2], f[3]);
  v.z = pack_bf16x2(f[4], f[5]); v.w = pack_bf16x2(f[6], f[7]);
  return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// streaming 128-bit load that does not allocate in L1 (weights / KV are read once per step)
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------------------------------
// Programmatic dependent launch (PDL): every kernel is launched with the programmatic stream
// serialization attribute.  Discipline used throughout: a kernel calls pdl_wait() before it touches
// anything its IMMEDIATE predecessor wrote, and only then pdl_launch_dependents().  So when a kernel
// starts, its predecessor has already passed its own wait, i.e. everything older than the immediate
// predecessor is complete and visible: code before pdl_wait() may read constants (weights, tables)
// AND anything produced two or more kernels earlier (block tables, sequence lengths, the residual
// stream, all K/V pages except the slots the previous kernel is writing).
// ---------------------------------------------------------------------------------------
// A kernel's idle lanes ask for the NEXT kernel's weights to be brought into L2 (cp.async.bulk.prefetch.L2, 4 KB per
// request): the next kernel's TMA stream then starts from L2 instead of HBM even when its CTAs could not become
// resident early (see gemm_resid.cu).  part / nparts: this CTA's share.
struct L2Prefetch {
  const void* ptr = nullptr;
  unsigned long long bytes = 0;
};
__device__ __forceinline__ void b2_l2_prefetch(const L2Prefetch& pf, int part, int nparts, int lane) {
  if (pf.ptr == nullptr) return;
  const unsigned long long nchunks = (pf.bytes + 4095ull) >> 12;
  for (unsigned long long c = (unsigned long long)part * 32 + lane; c < nchunks; c += (unsigned long long)nparts * 32) {
    const unsigned long long off = c << 12;
    const unsigned long long left = pf.bytes - off;
    const unsigned sz = (unsigned)(left < 4096ull ? (left & ~15ull) : 4096ull);
    if (sz) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(reinterpret_cast<const char*>(pf.ptr) + off), "r"(sz) : "memory");
  }
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool b2_pdl_enabled();     // consumes a pending b2_pdl_skip_next()
void b2_pdl_skip_next();   // the calling thread's next launch is an ordinary (fully serialised) one
bool b2_pdl_trigger_early();  // B2_PDL_TRIGGER (default 1): norm / SwiGLU kernels release the projection behind them at their start
int b2_pdl_early_mask();  // B2_PDL_EARLY bitmask (diagnostics): 1 attention streams before its wait, 2 norm, 4 rope
// Function attributes (opt-in shared memory size) are per device: remember which devices were set.
struct B2PerDeviceOnce {
  bool done[64] = {false};
  bool need() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
    if (done[d]) return false;
    done[d] = true;
    return true;
  }
};
// Every kernel of the decode path asks for the largest shared-memory carve-out, whatever it uses itself: an SM's
// L1/shared split can only change while the SM is idle, so a CTA that needs 82 KB cannot join a running CTA on an SM
// that was configured for that CTA's 132 KB alone -- it would wait for the SM to drain (two-lane decode, engine.cu).
template <typename... KArgs>
inline void b2_prefer_max_smem(void (*kernel)(KArgs...), B2PerDeviceOnce& once) {
  if (once.need()) {
    // On only with the two-lane schedule (B2_MB > 0) or B2_CARVEOUT=1: the smaller L1 costs a one-lane step 1.5 %.
    static const bool on = [] {
      const char* e = getenv("B2_CARVEOUT");
      if (e) return *e != '0';
      const char* mb = getenv("B2_MB");
      return mb != nullptr && atoi(mb) > 0;
    }();
    if (!on) return;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared) != cudaSuccess)
      cudaGetLastError();  // a preference only
  }
}
template <typename... KArgs, typename... Args>
inline cudaError_t b2_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = b2_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------------------------------
// Step timeline (tools/step_timeline.py): when a trace buffer is registered (b2k_debug_step_trace), thread 0 of
// every CTA of every kernel on the decode path appends one record {class | blockIdx << 8 | smid << 32 | aux << 40, CTA start,
// dependency wait returned, CTA end} (%globaltimer, ns).  buf[0] = record counter (atomic), buf[1] = capacity in
// records, records start at buf[4].  Each .cu file holds its own copy of the pointer (no relocatable device code);
// B2_STEP_TRACE_DEFINE(setter) defines it and its host-side setter.  Compiled in only with -DB2_STEP_TRACE
// (tools/build_variant.sh trace -DB2_STEP_TRACE -> libb2engine_trace.so, which tools/step_timeline.py loads).
// ---------------------------------------------------------------------------------------
#ifdef B2_STEP_TRACE
#define B2_STEP_TRACE_DEFINE(setter)                                              \
  static __constant__ unsigned long long* g_step_trace = nullptr;                 \
  int setter(unsigned long long* dev_buf) {                                       \
    B2_CUDA_OK(cudaMemcpyToSymbol(g_step_trace, &dev_buf, sizeof(dev_buf)));       \
    return 0;                                                                     \
  }
#else
#define B2_STEP_TRACE_DEFINE(setter)                                                                              \
  int setter(unsigned long long* dev_buf) {                                                                       \
    if (dev_buf == nullptr) return 0;                                                                             \
    b2_set_error("step trace: this library was built without -DB2_STEP_TRACE (tools/build_variant.sh trace ...)"); \
    return -38;                                                                                                   \
  }
#endif
// The stamps of a CTA live in shared memory (24 bytes), not in registers: the projection kernel must stay at 72
// registers to fit on an SM next to a decode-attention CTA (engine.cu, two-lane decode).
__device__ __forceinline__ unsigned long long b2_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void b2_stamp_begin(unsigned long long* st, unsigned long long* buf) {
  st[0] = (unsigned long long)buf;
  st[1] = st[2] = 0;
  if (buf != nullptr) st[1] = b2_globaltimer();
}
__device__ __forceinline__ void b2_stamp_wait(unsigned long long* st) {
  if (st[0] != 0ull) st[2] = b2_globaltimer();
}
__device__ __forceinline__ void b2_stamp_end(const unsigned long long* st, int cls, unsigned aux = 0) {
  unsigned long long* buf = (unsigned long long*)st[0];
  if (buf == nullptr) return;
  const unsigned long long t1 = b2_globaltimer();
  const unsigned long long slot = atomicAdd(buf, 1ull);
  if (slot < buf[1]) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    unsigned long long* r = buf + 4 + slot * 4;
    r[0] = (unsigned long long)((unsigned)cls & 0xffu) | ((unsigned long long)(blockIdx.x & 0xffffffu) << 8) |
           ((unsigned long long)(smid & 0xffu) << 32) | ((unsigned long long)(aux & 0xffffffu) << 40);
    r[1] = st[1];
    r[2] = st[2];
    r[3] = t1;
  }
}
// Use inside a kernel (thread 0 only does anything): B2_STAMP_BEGIN at entry, B2_STAMP_WAIT right after
// griddepcontrol.wait, B2_STAMP_END(cls) last.
#ifndef B2_STEP_TRACE  // the stamps cost 1.5 % of a decode step (measured, 258 tiny kernels): only the trace build has them
#define B2_STAMP_BEGIN() do { } while (0)
#define B2_STAMP_WAIT() do { } while (0)
#define B2_STAMP_END(cls) do { } while (0)
#define B2_STAMP_END_AUX(cls, aux) do { } while (0)
#else
#define B2_STAMP_BEGIN() __shared__ unsigned long long _stamp[3]; if (threadIdx.x == 0) b2_stamp_begin(_stamp, g_step_trace)
#define B2_STAMP_WAIT() do { if (threadIdx.x == 0) b2_stamp_wait(_stamp); } while (0)
#define B2_STAMP_END(cls) do { if (threadIdx.x == 0) b2_stamp_end(_stamp, (cls)); } while (0)
#define B2_STAMP_END_AUX(cls, aux) do { if (threadIdx.x == 0) b2_stamp_end(_stamp, (cls), (aux)); } while (0)
#endif

// ---------------------------------------------------------------------------------------
// mbarrier + bulk async copy (TMA engine, SASS: UBLKCP / UTMALDG)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// 1-D bulk copy global -> shared, completion signalled on an mbarrier (bytes % 16 == 0).
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// 2-D tiled TMA load through a tensor map.
__device__ __forceinline__ void tma_load_2d(uint32_t dst_smem, const CUtensorMap* tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst_smem),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------------------
// tcgen05 / TMEM (SASS: UTCHMMA, LDTM, UTCBAR ...)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem], bf16 inputs, fp32 accumulate; issued by ONE thread.
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns -> 16 registers per thread (thread i <-> TMEM lane base+i)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
      "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]),
        "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]),
        "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// the reverse of tmem_ld32: 32 registers per thread -> 32 consecutive fp32 columns of the warp's 32 TMEM lanes
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
      "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]),
      "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
               "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
               "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for a K-major bf16 tile stored as rows of 128 bytes with the
// 128-byte swizzle (what TMA SWIZZLE_128B writes): 8-row groups are 1024 B apart (SBO), LBO is
// unused for swizzled K-major layouts (set to 1 as CUTLASS does), version 1 = Blackwell.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor: D fp32, A/B bf16, both K-major, M x N tile.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// ---------------------------------------------------------------------------------------
// host: tensor-map encoding through the driver entry point (no link-time libcuda dependency)
// ---------------------------------------------------------------------------------------
// 2-D row-major bf16 matrix [rows][cols] with leading dimension ld (elements); box = 64 cols x
// box_rows rows, 128-byte swizzle.
int b2_make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);
// General form: box = box_cols x box_rows elements; swizzle128 = 0 copies rows as they are (for data that is stored swizzled).
int b2_make_tmap_bf16_ex(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols, uint32_t box_rows,
                         int swizzle128);
