// attn_prefill_tc.cu -- causal prefill attention over the paged KV cache on the 5th-generation tensor cores
// (HF:integrations/sdpa_attention.py:40-104 at T > 1; eager form HF:models/qwen3/modeling_qwen3.py:196-218).
// Same contract and the same tile list as attn_prefill.cu (the mma.sync kernel, kept as the cross-check and for
// callers without a tensor map): fp32 scores, fp32 online-softmax statistics, probabilities rounded to bf16 for P.V,
// the row sum from the unrounded fp32 probabilities, fp32 accumulation, one multiply by 1/l, one rounding.
//
// One CTA = one 64-query tile of one sequence x one kv head.  Both query heads of the GQA group sit on the MMA's M side
// (row r = head (r >> 6), query token (r & 63)), so K and V are read once per pair:
//   S[128 x 128 keys]  = Q[128 x 128 dims] . K^T      tcgen05.mma, A and B K-major from shared memory, S in TMEM
//   softmax            : 256 threads, TWO per row (warps w and w + 4 may both read TMEM lanes 32 (w & 3)..+31): each takes
//                        64 of the tile's 128 keys -- tcgen05.ld, scale / mask, row maximum exchanged through shared
//                        memory between the two, running sum kept per thread and added at the end -- and writes its
//                        half of P (bf16) into shared memory in the K-major 128-byte-swizzled layout the next MMA reads
//   PV[128 x 128 dims] = P[128 x 128 keys] . V        V is the B operand in MN-major form: the page layout
//                        [key][128 dims] with its 16-byte chunks XOR-swizzled by (key & 7) IS the canonical MN-major
//                        128-byte-swizzle layout once the two 64-dim halves of every row are stored apart
//   O (registers)      = O * corr + PV                 each of the row's two threads keeps 64 of the 128 output dims
// K and V pages come in through TMA (a 2-D map over the page pool, box 64 dims x 16 keys, no TMA swizzle: the pool is
// stored swizzled): half h of key row k lands at [h][k][128 bytes], which serves as K-major B for Q.K^T and as
// MN-major B for P.V without any re-layout.  Inside a CTA the steps of a key tile run one after the other; the overlap
// comes from two CTAs per SM (96 KB of shared memory, 256 TMEM columns each).
#include "kernels.h"

namespace b2 {

namespace {

constexpr int PT_TQ = 64;                    // query tokens per tile
constexpr int PT_M = 128;                    // MMA rows: 2 heads x 64 tokens
constexpr int PT_BN = 128;                   // keys per key tile
constexpr int PT_HD = 128;
constexpr int PT_HALF_BYTES = PT_BN * 128;   // one 64-dim half of a key tile: 128 keys x 128 bytes = 16 KB
constexpr int PT_TILE_BYTES = 2 * PT_HALF_BYTES;
// Q | K | V, 32 KB each.  P is written over K: once S = Q.K^T is complete the K tile is dead.  96 KB per CTA, so TWO CTAs
// share an SM and one's softmax runs while the other's MMAs and page loads do -- the overlap a deeper in-CTA pipeline
// would otherwise have to provide.
constexpr int PT_SMEM = 3 * PT_TILE_BYTES + 1024;
constexpr int PT_MAX_PAGES = 1024;           // block-table entries staged in shared memory (16 K tokens of context)
constexpr int PT_TMEM_COLS = 256;            // S: columns 0..127, PV: 128..255

// K-major, 128-byte swizzle (rows of 128 bytes, 8-row groups 1024 bytes apart): Q, K and P tiles
__device__ __forceinline__ uint64_t pt_desc_k(uint32_t smem_addr) { return umma_desc_sw128(smem_addr); }
// MN-major, 128-byte swizzle: the V tile.  64 contiguous dims (128 bytes) per key, 8-key groups 1024 bytes apart
// (stride byte offset), the other 64 dims 16 KB further on (leading byte offset).
__device__ __forceinline__ uint64_t pt_desc_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(PT_HALF_BYTES >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ float pt_ex2(float x) {  // 2^x, one MUFU (results below 2^-126 flush to zero: they are probabilities)
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr uint32_t PT_IDESC_QK = umma_idesc_bf16(PT_M, PT_BN);                // both operands K-major
constexpr uint32_t PT_IDESC_PV = umma_idesc_bf16(PT_M, PT_HD) | (1u << 16);  // B (V) MN-major

// Threads: warps 0-7 softmax (two threads per row), warp 8 the MMA issuer (and TMEM allocator), warps 9-12 one lane each
// issuing the TMA loads of a key tile (K pages 0-3 | K pages 4-7 | V pages 0-3 | V pages 4-7).
// Measured with clock64 on the first version of this kernel (one control thread did the MMAs AND all 32 TMA loads of a
// tile, O lived in registers): per key tile the control thread spent ~3 000 cycles just issuing the loads (~100 per
// cp.async.bulk.tensor, serial, in front of the next S = Q.K^T), the softmax threads ~3 000 in the probability pass and
// ~2 200 in the O read-back (both slowed by register spills at the 96-register cap), against ~550 per MMA group.  Hence:
// four loader warps, and O stays in TMEM.
constexpr int PT_THREADS = 13 * 32;
constexpr float PT_RESCALE = 8.0f;  // log2 units: the reference maximum is raised only when the row maximum grew by 2^8

__global__ void __launch_bounds__(PT_THREADS, 2)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v, const bf16* __restrict__ q,
                       const int32_t* __restrict__ block_table, int bt_stride, const PrefillTile* __restrict__ tiles,
                       bf16* __restrict__ out, int nq, int nkv, float scale_log2, int window, int pair_mode) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t k_full, v_full, s_full, p_full, o_full;
  __shared__ uint32_t tmem_slot;
  __shared__ int32_t s_bt[PT_MAX_PAGES];
  __shared__ float s_mx[2][PT_M];  // row maxima of the two key halves of the current tile; at the end: the two row sums
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t q_s = base, k_s = base + PT_TILE_BYTES, v_s = base + 2 * PT_TILE_BYTES, p_s = k_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Blocks are dispatched in index order and a tile late in its sequence has the most key tiles: walking the tile list
  // backwards starts the long CTAs first and leaves the short ones for the kernel's last wave.
  // pair_mode (models with one query head per kv head, run through the GQA-2 layout with an all-zero partner head): the 128
  // MMA rows are TWO consecutive 64-token tiles of one sequence x the real head, instead of one tile x (real, zero) heads --
  // half the CTAs, no MMA rows spent on zeros; the partner head's output rows are not written (the output projection's
  // columns for them are zeros, and the buffer holds finite values from its allocation).  The host pads every sequence to an
  // even number of tiles (a tile with nrows == 0), so tiles 2i and 2i + 1 always belong together.
  const int ti = gridDim.x - 1 - blockIdx.x;
  const PrefillTile tile = tiles[pair_mode ? 2 * ti : ti];                // rows 0..63 (and, without pair_mode, 64..127)
  const PrefillTile tile_b = pair_mode ? tiles[2 * ti + 1] : tile;        // rows 64..127 in pair_mode
  const int kvh = blockIdx.y;
  const int kv_len = (pair_mode && tile_b.nrows > 0) ? tile_b.q_pos0 + tile_b.nrows : tile.q_pos0 + tile.nrows;
  const int n_kt = (kv_len + PT_BN - 1) / PT_BN;
  const int kt0 = (window > 0 && tile.q_pos0 - window + 1 > 0) ? (tile.q_pos0 - window + 1) / PT_BN : 0;
  const int32_t* bt = block_table + (size_t)tile.bt_row * bt_stride;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&k_full), 1);
    mbar_init(smem_u32(&v_full), 1);
    mbar_init(smem_u32(&s_full), 1);
    mbar_init(smem_u32(&p_full), 256);
    mbar_init(smem_u32(&o_full), 1);
    mbar_fence_init();
  }
  if (warp == 8) {
    tmem_alloc(smem_u32(&tmem_slot), PT_TMEM_COLS);
    tmem_relinquish();
  }
  // this sequence's block-table row (written by the host before the forward started) into shared memory: the loader
  // threads would otherwise chain a global load in front of every TMA issue
  for (int i = threadIdx.x; i < n_kt * (PT_BN / 16) && i < PT_MAX_PAGES; i += blockDim.x) s_bt[i] = i < bt_stride ? bt[i] : 0;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  pdl_wait();  // q and the newest K/V rows come from the kernel before this one
  if (threadIdx.x == 0) pdl_launch_dependents();

  if (warp >= 9) {
    // ---- loaders: warp 9 | 10 the K pages 0-3 | 4-7 of a key tile, warp 11 | 12 the V pages; both 64-dim halves each ----
    if (lane == 0) {
      const int j = warp - 9;
      const bool is_v = j >= 2;
      const CUtensorMap* tm = is_v ? &tmap_v : &tmap_k;
      const uint32_t fb = smem_u32(is_v ? &v_full : &k_full);
      const uint32_t dst = (is_v ? v_s : k_s) + (uint32_t)((j & 1) * 4 * 16 * 128);
      tma_prefetch_desc(tm);
      int it = 0;
      for (int kt = kt0; kt < n_kt; ++kt, ++it) {
        // P (in K's place) and V of the previous tile are read by its P.V MMAs: pages may land once those have completed
        if (it > 0) mbar_wait(smem_u32(&o_full), (it - 1) & 1);
        if ((j & 1) == 0) mbar_arrive_expect_tx(fb, PT_TILE_BYTES);  // the other warp of the pair only adds bytes
#pragma unroll
        for (int pg = 0; pg < 4; ++pg) {
          const int row0 = (s_bt[kt * (PT_BN / 16) + (j & 1) * 4 + pg] * nkv + kvh) * 16;  // pages past the sequence are the null page 0
          tma_load_2d(dst + pg * 16 * 128, tm, 0, row0, fb);
          tma_load_2d(dst + PT_HALF_BYTES + pg * 16 * 128, tm, 64, row0, fb);
        }
      }
    }
  } else if (warp == 8) {
    if (lane == 0) {
      mbar_wait(smem_u32(&p_full), 0);  // phase 0 of p_full: the softmax threads have staged Q
      int it = 0;
      for (int kt = kt0; kt < n_kt; ++kt, ++it) {
        mbar_wait(smem_u32(&k_full), it & 1);
        tc_fence_after();
        // S = Q . K^T : K = 128 dims = 2 k-blocks x 4 MMAs of 16
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const uint64_t da = pt_desc_k(q_s + kb * PT_HALF_BYTES), db = pt_desc_k(k_s + kb * PT_HALF_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), PT_IDESC_QK, (kb | k) ? 1u : 0u);
        }
        tc_commit(smem_u32(&s_full));
        // O (+)= P . V : K = 128 keys = 2 k-blocks (of P, which now lies where K was) x 4 MMAs of 16 keys.  O stays in TMEM
        // across key tiles; the softmax threads rescale it in place when a row's reference maximum moves.
        mbar_wait(smem_u32(&p_full), (it + 1) & 1);
        mbar_wait(smem_u32(&v_full), it & 1);
        tc_fence_after();
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const uint64_t da = pt_desc_k(p_s + kb * PT_HALF_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t db = pt_desc_mn(v_s + (uint32_t)((kb * 64 + k * 16) * 128));
            tc_mma_bf16(tmem + 128u, da + (uint64_t)(2 * k), db, PT_IDESC_PV, (it > 0 || (kb | k)) ? 1u : 0u);
          }
        }
        tc_commit(smem_u32(&o_full));
      }
    }
  } else {
    // ---- 256 softmax threads: threads r and r + 128 own MMA row r = (head r >> 6, query token r & 63); `kh` says which
    // 64 keys of a tile (and which 64 output dims) are this thread's ----
    const int r = threadIdx.x & 127, kh = threadIdx.x >> 7;
    const int h = pair_mode ? 0 : r >> 6, t = r & 63;
    const PrefillTile& mt = (r >> 6) ? tile_b : tile;  // the tile this row belongs to
    const bool row_ok = t < mt.nrows;
    const int qpos = mt.q_pos0 + t;
    const int row_kv_len = mt.q_pos0 + mt.nrows;
    // stage Q: rows in the K-major swizzled layout, 16 bytes per thread and step (rows past the tile are zeros)
#pragma unroll
    for (int k = 0; k < (PT_M * 16) / 256; ++k) {  // 8 cp.async per thread, all in flight together
      const int idx = (int)threadIdx.x + k * 256;
      const int row = idx >> 4, c = idx & 15;
      const int rh = row >> 6, rt = row & 63;
      const PrefillTile& st = rh ? tile_b : tile;
      const bool in = rt < st.nrows;
      const bf16* src = q + ((size_t)(st.q_row0 + (in ? rt : 0)) * nq + kvh * 2 + (pair_mode ? 0 : rh)) * PT_HD + c * 8;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(q_s + (c >> 3) * PT_HALF_BYTES + row * 128 + (((c & 7) ^ (row & 7)) << 4)),
                   "l"(src), "r"(in ? 16 : 0)
                   : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    fence_proxy_async();
    mbar_arrive(smem_u32(&p_full));  // phase 0 of p_full = "Q is staged"

    constexpr int HALF = PT_BN / 2;  // keys per thread and tile == output dims per thread
    // m_ref: the value the probabilities of this row are scaled by, 2^(s - m_ref).  It follows the running row maximum
    // lazily -- only when the maximum has grown by more than PT_RESCALE -- so that most key tiles leave O (in TMEM)
    // untouched; probabilities then reach 2^8 at most, and l and O carry the same factor, which cancels in O / l.
    float m_ref = -INFINITY, l_run = 0.f;  // l_run: the sum over THIS thread's keys only
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t orow = trow + 128u + (uint32_t)(kh * HALF);  // this thread's 64 output dims
    uint8_t* prow = sm + PT_TILE_BYTES + kh * PT_HALF_BYTES + r * 128;  // P goes where K was (the S product has consumed K)
    const int pair_bar = 1 + (warp & 3);  // named barrier of the two warps that share a lane quarter
    int it = 0;
    for (int kt = kt0; kt < n_kt; ++kt, ++it) {
      mbar_wait(smem_u32(&s_full), it & 1);
      tc_fence_after();
      const int key0 = kt * PT_BN;
      // Valid keys of this row in this tile are a contiguous range [lo_col, hi_col] (causal bound, sequence end, window).
      // A tile that lies wholly inside the range of EVERY row of the CTA (all but the last one or two tiles of a row
      // block) takes the branch without per-element masks; the test is uniform over the CTA.
      const int hi_col = row_ok ? min(qpos, row_kv_len - 1) : -1;
      const int lo_col = (window > 0 && qpos - window + 1 > 0) ? qpos - window + 1 : 0;
      const bool full_tile = key0 + PT_BN - 1 <= tile.q_pos0 && tile.nrows == PT_TQ && tile_b.nrows == PT_TQ &&
                             (window <= 0 || tile_b.q_pos0 + PT_TQ - 1 - key0 < window);
      // pass 1: the row maximum (the scale is positive: max(s * c) = c * max(s)); four chains instead of one
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
      for (int c0 = kh * HALF; c0 < kh * HALF + HALF; c0 += 16) {
        uint32_t sr[16];
        tmem_ld16(trow + (uint32_t)c0, sr);
        tmem_ld_wait();
        if (full_tile) {
#pragma unroll
          for (int j = 0; j < 16; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(sr[j]));
        } else {
          const int jlo = lo_col - (key0 + c0);
          const unsigned span = (unsigned)(hi_col - lo_col);  // hi < lo (nothing valid) wraps to a huge span: guard below
          const bool any = hi_col >= lo_col;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const bool ok = any && (unsigned)(j - jlo) <= span;
            m4[j & 3] = fmaxf(m4[j & 3], ok ? __uint_as_float(sr[j]) : -INFINITY);
          }
        }
      }
      const float mx_own = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      s_mx[kh][r] = mx_own;
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      const float mx = fmaxf(mx_own, s_mx[kh ^ 1][r]) * scale_log2;  // the same value in both threads of the row
      // raise the reference?  (first tile: O does not exist yet, the first P.V overwrites it)
      const bool raise = mx > m_ref + PT_RESCALE || (m_ref == -INFINITY && mx > -INFINITY);
      if (it > 0 && __any_sync(0xffffffffu, raise)) {
        const float corr = raise ? exp2f(m_ref - mx) : 1.0f;  // m_ref == -inf: 0 (O and l of such a row are zeros so far)
        mbar_wait(smem_u32(&o_full), (it - 1) & 1);  // the previous tile's P.V has landed in O
        tc_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < HALF; c0 += 16) {
          uint32_t ov[16];
          tmem_ld16(orow + (uint32_t)c0, ov);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) ov[j] = __float_as_uint(__uint_as_float(ov[j]) * corr);
          tmem_st16(orow + (uint32_t)c0, ov);
        }
        tmem_st_wait();
        l_run *= corr;
      }
      if (raise) m_ref = mx;
      const float msub = (m_ref == -INFINITY) ? 0.f : m_ref;
      // pass 2: probabilities, their fp32 sum, and P as bf16 into the A-operand tile of the PV product
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
      for (int c0 = kh * HALF; c0 < kh * HALF + HALF; c0 += 16) {  // 16 columns at a time: the kernel lives in 72 registers
        uint32_t sr[16];
        tmem_ld16(trow + (uint32_t)c0, sr);
        tmem_ld_wait();
        float p[16];
        if (full_tile) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            p[j] = pt_ex2(__uint_as_float(sr[j]) * scale_log2 - msub);
            l4[j & 3] += p[j];
          }
        } else {
          const int jlo = lo_col - (key0 + c0);
          const unsigned span = (unsigned)(hi_col - lo_col);
          const bool any = hi_col >= lo_col;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const bool ok = any && (unsigned)(j - jlo) <= span;
            p[j] = ok ? pt_ex2(__uint_as_float(sr[j]) * scale_log2 - msub) : 0.f;
            l4[j & 3] += p[j];
          }
        }
#pragma unroll
        for (int c8 = 0; c8 < 2; ++c8) {  // two 16-byte chunks of 8 keys
          const int chunk = ((c0 >> 3) & 7) + c8;  // 0..7 within this thread's 64-key half of the row
          *reinterpret_cast<uint4*>(prow + ((chunk ^ (r & 7)) << 4)) = pack8(p + c8 * 8);
        }
      }
      l_run += (l4[0] + l4[1]) + (l4[2] + l4[3]);
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(smem_u32(&p_full));
    }
    // ---- O / l, once ----
    s_mx[kh][r] = l_run;  // the pair's last barrier of the loop is behind both threads' reads of s_mx
    asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
    const float l_row = s_mx[0][r] + s_mx[1][r];
    const float inv = l_row > 0.f ? __fdiv_rn(1.0f, l_row) : 0.f;
    mbar_wait(smem_u32(&o_full), (it - 1) & 1);
    tc_fence_after();
    bf16* dst = out + ((size_t)(mt.q_row0 + t) * nq + kvh * 2 + h) * PT_HD + kh * HALF;
#pragma unroll 1
    for (int c0 = 0; c0 < HALF; c0 += 16) {
      uint32_t ov[16];
      tmem_ld16(orow + (uint32_t)c0, ov);
      tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(ov[c * 8 + e]) * inv;
          reinterpret_cast<uint4*>(dst + c0)[c] = pack8(v);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem, PT_TMEM_COLS);
  }
}

}  // namespace

// num_pages: pages in the pool k_pages / v_pages point to (the TMA maps cover exactly the pool).
int launch_attn_prefill_tc(const bf16* q, const bf16* k_pages, const bf16* v_pages, int64_t num_pages, const int32_t* block_table,
                           int bt_stride, const PrefillTile* tiles_dev, int ntiles, bf16* out, int nq, int nkv, int hd, int block_size,
                           cudaStream_t s, float softmax_scale, int window, int pair_mode) {
  if (ntiles <= 0) return 0;
  B2_REQUIRE(hd == PT_HD && block_size == 16 && nq == 2 * nkv, "attn_prefill_tc: built for 128-wide heads in GQA groups of two, 16-token pages");
  B2_REQUIRE(num_pages >= 1, "attn_prefill_tc: the page pool's size is needed for the tensor maps");
  B2_REQUIRE(bt_stride <= PT_MAX_PAGES, "attn_prefill_tc: contexts beyond %d tokens not built (block table row of %d pages)", PT_MAX_PAGES * 16, bt_stride);
  CUtensorMap tk, tv;
  const uint64_t rows = (uint64_t)num_pages * nkv * 16;
  int rc = b2_make_tmap_bf16_ex(&tk, k_pages, rows, PT_HD, PT_HD, 64, 16, /*swizzle128=*/0);
  if (rc != 0) return rc;
  rc = b2_make_tmap_bf16_ex(&tv, v_pages, rows, PT_HD, PT_HD, 64, 16, 0);
  if (rc != 0) return rc;
  static B2PerDeviceOnce attr_once;
  if (attr_once.need()) B2_CUDA_OK(cudaFuncSetAttribute(attn_prefill_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PT_SMEM));
  const double scale = softmax_scale > 0.f ? (double)softmax_scale : 1.0 / sqrt((double)hd);
  const float scale_log2 = (float)(scale * 1.4426950408889634);
  B2_REQUIRE(!pair_mode || ntiles % 2 == 0, "attn_prefill_tc: pair mode needs an even tile list (each sequence padded to an even number of tiles)");
  B2_CUDA_OK(b2_launch(attn_prefill_tc_kernel, dim3(pair_mode ? ntiles / 2 : ntiles, nkv), dim3(PT_THREADS), (size_t)PT_SMEM, s, tk, tv, q, block_table,
                       bt_stride, tiles_dev, out, nq, nkv, scale_log2, window, pair_mode));
  return 0;
}

}  // namespace b2
