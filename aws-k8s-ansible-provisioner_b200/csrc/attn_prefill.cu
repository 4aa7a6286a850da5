// attn_prefill.cu -- causal prefill attention over the paged KV cache
// (HF:integrations/sdpa_attention.py:40-104 with is_causal=True, T > 1).  Ragged batch: the
// host cuts every sequence's new tokens into 64-query tiles (PrefillTile); one CTA handles one
// (tile, query head), 4 warps x 16 query rows, streaming 64-key K/V tiles of its kv head from
// the pages through cp.async into XOR-swizzled shared memory (double buffered).
// Math: S = Q K^T and O += P V on the warp-level tensor-core path (mma.sync m16n8k16, bf16
// in / fp32 accumulate); online softmax in fp32 with P rounded to bf16 for the P.V product --
// the same rounding points as the decode kernel and as torch's CPU flash kernel.
// (Prefill is a one-off per request; the tcgen05 treatment is reserved for the projections,
// which carry ~10x the prefill FLOPs of attention at these sequence lengths.)
#include "kernels.h"

namespace b2 {

constexpr int AP_BM = 64;
constexpr int AP_BN = 64;
constexpr int AP_HD = 128;
constexpr int AP_TILE_BYTES = AP_BN * AP_HD * 2;  // 16 KB
constexpr int AP_SMEM = 4 * AP_TILE_BYTES;        // K,V x 2 buffers = 64 KB

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(128)
attn_prefill_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k_pages, const bf16* __restrict__ v_pages,
                    const int32_t* __restrict__ block_table, int bt_stride, const PrefillTile* __restrict__ tiles,
                    bf16* __restrict__ out, int nq, int nkv, int block_size, float scale_log2, int window) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_wait();
  pdl_launch_dependents();
  const PrefillTile tile = tiles[blockIdx.x];
  const int h = blockIdx.y, kvh = h / (nq / nkv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int kv_len = tile.q_pos0 + tile.nrows;  // keys visible to the last row of the tile
  const int n_kv_tiles = (kv_len + AP_BN - 1) / AP_BN;
  // sliding window (0 = off): a query at position p sees keys p - window + 1 .. p; key tiles wholly below the
  // window of the tile's FIRST row are skipped, the rest is masked per element
  const int j0 = (window > 0 && tile.q_pos0 - window + 1 > 0) ? (tile.q_pos0 - window + 1) / AP_BN : 0;
  const int32_t* bt = block_table + (size_t)tile.bt_row * bt_stride;
  const uint32_t smem0 = smem_u32(smem);

  auto load_tile = [&](int j, int buf) {
    const uint32_t kdst = smem0 + buf * 2 * AP_TILE_BYTES;
    const uint32_t vdst = kdst + AP_TILE_BYTES;
    for (int idx = threadIdx.x; idx < AP_BN * 16; idx += 128) {
      // c = PHYSICAL chunk position.  Pages are stored chunk-swizzled by (row & 7) (kv_swizzled_chunk),
      // and tiles start at multiples of 16 tokens, so (key & 7) == (row in page & 7): the page's own
      // swizzle is exactly the one the ldmatrix reads below expect -- a straight 16-byte copy.
      const int key = idx >> 4, c = idx & 15;
      const int kpos = j * AP_BN + key;
      const uint32_t o = key * (AP_HD * 2) + (c << 4);
      if (kpos < kv_len) {
        const int blk = bt[kpos / block_size];
        const size_t src = (((size_t)blk * nkv + kvh) * block_size + (kpos % block_size)) * AP_HD + c * 8;
        cp_async16(kdst + o, k_pages + src);
        cp_async16(vdst + o, v_pages + src);
      } else {
        *reinterpret_cast<uint4*>(smem + (kdst - smem0) + o) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(smem + (vdst - smem0) + o) = make_uint4(0, 0, 0, 0);
      }
    }
    cp_async_commit();
  };

  load_tile(j0, 0);

  // Q fragments for this warp's 16 rows, straight from global (read once)
  const int row_a = warp * 16 + g, row_b = row_a + 8;
  uint32_t qf[8][4];
  {
    const bf16* qa = q + ((size_t)(tile.q_row0 + row_a) * nq + h) * AP_HD;
    const bf16* qb = q + ((size_t)(tile.q_row0 + row_b) * nq + h) * AP_HD;
    const bool va = row_a < tile.nrows, vb = row_b < tile.nrows;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      qf[ks][0] = va ? *reinterpret_cast<const uint32_t*>(qa + ks * 16 + 2 * t4) : 0u;
      qf[ks][1] = vb ? *reinterpret_cast<const uint32_t*>(qb + ks * 16 + 2 * t4) : 0u;
      qf[ks][2] = va ? *reinterpret_cast<const uint32_t*>(qa + ks * 16 + 8 + 2 * t4) : 0u;
      qf[ks][3] = vb ? *reinterpret_cast<const uint32_t*>(qb + ks * 16 + 8 + 2 * t4) : 0u;
    }
  }

  float oacc[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const int qpos_a = tile.q_pos0 + row_a, qpos_b = tile.q_pos0 + row_b;

  for (int j = j0; j < n_kv_tiles; ++j) {
    const int buf = (j - j0) & 1;
    if (j + 1 < n_kv_tiles) {
      load_tile(j + 1, buf ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const uint32_t ks_base = smem0 + buf * 2 * AP_TILE_BYTES;
    const uint32_t vs_base = ks_base + AP_TILE_BYTES;

    // ---- S = Q K^T (16 rows x 64 keys per warp) ----
    float sacc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) sacc[nt][0] = sacc[nt][1] = sacc[nt][2] = sacc[nt][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 8; ks += 2) {
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int key = nt * 8 + (lane & 7);
        const int c = 2 * ks + (lane >> 3);
        uint32_t r0, r1, r2, r3;
        ldsm_x4(ks_base + key * (AP_HD * 2) + ((c ^ (key & 7)) << 4), r0, r1, r2, r3);
        mma_bf16(sacc[nt], qf[ks], r0, r1);
        mma_bf16(sacc[nt], qf[ks + 1], r2, r3);
      }
    }
    // ---- scale, causal mask, online softmax ----
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int col = j * AP_BN + nt * 8 + 2 * t4 + (e & 1);
        const int qp = (e < 2) ? qpos_a : qpos_b;
        float s = sacc[nt][e] * scale_log2;
        if (col > qp || col >= kv_len || (window > 0 && qp - col >= window)) s = -INFINITY;
        sacc[nt][e] = s;
        mx[e >> 1] = fmaxf(mx[e >> 1], s);
      }
    }
    float corr[2], msub[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_run[r], mx[r]);
      msub[r] = (m_new == -INFINITY) ? 0.f : m_new;
      corr[r] = exp2f(m_run[r] - msub[r]);
      m_run[r] = m_new;
      l_run[r] *= corr[r];
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      oacc[i][0] *= corr[0]; oacc[i][1] *= corr[0];
      oacc[i][2] *= corr[1]; oacc[i][3] *= corr[1];
    }
    uint32_t pa[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float p0 = exp2f(sacc[nt][0] - msub[0]), p1 = exp2f(sacc[nt][1] - msub[0]);
      const float p2 = exp2f(sacc[nt][2] - msub[1]), p3 = exp2f(sacc[nt][3] - msub[1]);
      l_run[0] += p0 + p1;
      l_run[1] += p2 + p3;
      const int kk = nt >> 1;
      if ((nt & 1) == 0) {
        pa[kk][0] = pack_bf16x2(p0, p1);
        pa[kk][1] = pack_bf16x2(p2, p3);
      } else {
        pa[kk][2] = pack_bf16x2(p0, p1);
        pa[kk][3] = pack_bf16x2(p2, p3);
      }
    }
    // ---- O += P V ----
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int dt = 0; dt < 16; dt += 2) {
        const int key = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int c = dt + (lane >> 4);
        uint32_t r0, r1, r2, r3;
        ldsm_x4_t(vs_base + key * (AP_HD * 2) + ((c ^ (key & 7)) << 4), r0, r1, r2, r3);
        mma_bf16(oacc[dt], pa[kk], r0, r1);
        mma_bf16(oacc[dt + 1], pa[kk], r2, r3);
      }
    }
    __syncthreads();  // all warps done with this buffer before it is refilled
  }

#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const float inv_a = l_run[0] > 0.f ? __fdiv_rn(1.0f, l_run[0]) : 0.f;
  const float inv_b = l_run[1] > 0.f ? __fdiv_rn(1.0f, l_run[1]) : 0.f;
  if (row_a < tile.nrows) {
    bf16* o = out + ((size_t)(tile.q_row0 + row_a) * nq + h) * AP_HD;
#pragma unroll
    for (int dt = 0; dt < 16; ++dt)
      *reinterpret_cast<uint32_t*>(o + dt * 8 + 2 * t4) = pack_bf16x2(oacc[dt][0] * inv_a, oacc[dt][1] * inv_a);
  }
  if (row_b < tile.nrows) {
    bf16* o = out + ((size_t)(tile.q_row0 + row_b) * nq + h) * AP_HD;
#pragma unroll
    for (int dt = 0; dt < 16; ++dt)
      *reinterpret_cast<uint32_t*>(o + dt * 8 + 2 * t4) = pack_bf16x2(oacc[dt][2] * inv_b, oacc[dt][3] * inv_b);
  }
}

int launch_attn_prefill(const bf16* q, const bf16* k_pages, const bf16* v_pages, const int32_t* block_table,
                        int bt_stride, const PrefillTile* tiles_dev, int ntiles, bf16* out, int nq, int nkv, int hd,
                        int block_size, cudaStream_t s, float softmax_scale, int window) {
  if (ntiles <= 0) return 0;
  B2_REQUIRE(hd == AP_HD, "attn_prefill: head_dim %d not built (128 only)", hd);
  B2_REQUIRE(nq % nkv == 0, "attn_prefill: nq=%d not a multiple of nkv=%d", nq, nkv);
  static B2PerDeviceOnce attr_once;
  if (attr_once.need()) B2_CUDA_OK(cudaFuncSetAttribute(attn_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AP_SMEM));
  const double scale = softmax_scale > 0.f ? (double)softmax_scale : 1.0 / sqrt((double)hd);
  const float scale_log2 = (float)(scale * 1.4426950408889634);
  dim3 grid(ntiles, nq);
  B2_CUDA_OK(b2_launch(attn_prefill_kernel, grid, dim3(128), (size_t)AP_SMEM, s, q, k_pages, v_pages, block_table, bt_stride,
                       tiles_dev, out, nq, nkv, block_size, scale_log2, window));
  return 0;
}

}  // namespace b2
