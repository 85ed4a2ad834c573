// gemm_tc.cu -- bf16 tensor-core GEMMs for sm_100a: TMA (cp.async.bulk.tensor) feeds a shared-memory
// ring, one elected thread issues tcgen05.mma with the fp32 accumulator in TMEM, epilogue warps read it
// back with tcgen05.ld.  Used by the bf16 path for the hoisted products of the cell stack (reference
// model.py:77 via TF _linear: X_l * Wx for all timesteps at once), the output projection
// (model.py:98-99), every weight / input gradient of model.py:182 and the sampler's per-step products.
//
//   C[M,N] (+)= A[M,K] * B[K,N] (+ bias[N]),   A, B bf16, C fp32 (or bf16, or none: fused LSTM cell)
//
// Either operand may be K-major (K contiguous in memory) or MN-major (M resp. N contiguous); both
// are loaded with 128-byte-swizzled TMA boxes and described to the tensor core with the matching
// canonical shared-memory descriptor, so X^T * dY (both operands MN-major, K = B*T rows) needs no
// transposed copy.
//   gemm_tc_kernel          one tile per CTA, split-K over blockIdx.z with an atomic fp32 epilogue
//                           (weight gradients; 128x128 or 128x256 tiles)
//   gemm_tc_persist_kernel  one CTA per SM walks the tiles, double-buffered accumulator, C by TMA store
//                           (fp32 or bf16) or the sampler's LSTM cell update as the epilogue
#include <stdlib.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 64;
constexpr int STAGES = 3;
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int TC_THREADS = 192;  // warp 0: TMA, warp 1: MMA + TMEM alloc, warps 2-5: epilogue
constexpr int TC_SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TC_SMEM_256 = 4 * (A_BYTES + 256 * BK * 2) + 1024 + 256;

struct TcArgs {
  int M, N, K;
  float *C;
  int ldc;
  const float *bias;
  int accumulate;
  int kb_per_split;
  uint32_t lbo_a, sbo_a, adv_a, lbo_b, sbo_b, adv_b;  // descriptor strides / per-UMMA_K advance (bytes)
  // LSTM-cell epilogue of the persistent kernel (sampler): columns are gate-interleaved (4*unit + gate)
  float *cell_c;                 // [M, H] cell state, updated in place unless the row holds
  __nv_bfloat16 *cell_self;      // this layer's operand rows: h slot at column cell_hoff, row pitch cell_kself
  __nv_bfloat16 *cell_out;       // layer output rows (next layer's operand / head input), row pitch cell_kout
  const float *cell_xin;         // [M, 5] current input point: column 3 > 0 = eoc = hold the state
  int cell_kself, cell_hoff, cell_kout, cell_H;
  int perm_H;                    // split-K kernel: un-interleave the columns of C (see the epilogue)
};

// BN_ = 128: 3 stages, two CTAs per SM.  BN_ = 256: 4 stages, one CTA per SM -- per k-block the tensor core then
// reads 48 KB of operands for twice the math, which takes the main loop off the shared-memory bandwidth limit
// (used for the long-K weight gradients).
template <bool A_KMAJOR, bool B_KMAJOR, int BN_>
__global__ void __launch_bounds__(TC_THREADS, BN_ == 128 ? 2 : 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, TcArgs a) {
  constexpr int BN = BN_;
  constexpr int STAGES = BN_ == 128 ? 3 : 4;
  constexpr int B_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = base + STAGES * STAGE_BYTES;  // full[S], empty[S], tmem_full, tmem holder
  const uint32_t bar_full = bars, bar_empty = bars + 8 * STAGES, bar_tmem = bars + 16 * STAGES;
  const uint32_t tmem_holder = bars + 16 * STAGES + 8;
  uint8_t *gen_base = smem_raw + (base - smem_u32(smem_raw));
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gen_base + STAGES * STAGE_BYTES + 16 * STAGES + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int nkb_total = (a.K + BK - 1) / BK;
  const int kb_beg = blockIdx.z * a.kb_per_split;
  const int kb_end = min(nkb_total, kb_beg + a.kb_per_split);
  const int nkb = kb_end - kb_beg;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    mbar_init(bar_tmem, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_holder), "r"(BN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;

  if (warp == 0) {
    if (elect_one()) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES, ph = (i / STAGES) & 1;
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
        mbar_expect_tx(bar_full + 8 * s, STAGE_BYTES);
        const int k0 = (kb_beg + i) * BK;
        if (A_KMAJOR) {
          tma_load_2d(sa, &tmA, bar_full + 8 * s, k0, m0);
        } else {
          tma_load_2d(sa, &tmA, bar_full + 8 * s, m0, k0);
          tma_load_2d(sa + A_BYTES / 2, &tmA, bar_full + 8 * s, m0 + 64, k0);
        }
        if (B_KMAJOR) {
          tma_load_2d(sb, &tmB, bar_full + 8 * s, k0, n0);
        } else {
#pragma unroll
          for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * 8192, &tmB, bar_full + 8 * s, n0 + 64 * c, k0);
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // instruction descriptor: D fp32, A/B bf16, majors, N>>3, M>>4
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_KMAJOR ? 0u : 1u) << 15) |
                             ((B_KMAJOR ? 0u : 1u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES, ph = (i / STAGES) & 1;
        mbar_wait(bar_full + 8 * s, ph);
        tc_fence_after();
        const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
        const uint64_t da = make_desc(sa, a.lbo_a, a.sbo_a), db = make_desc(sb, a.lbo_b, a.sbo_b);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          tc_mma_bf16(tmem_acc, da + (uint64_t)((a.adv_a * k) >> 4), db + (uint64_t)((a.adv_b * k) >> 4), idesc,
                      (i > 0 || k > 0) ? 1u : 0u);
        }
        tc_commit(bar_empty + 8 * s);  // frees the smem slot once these MMAs have read it
      }
      tc_commit(bar_tmem);  // accumulator complete
    }
  } else {
    // epilogue: warp w may touch TMEM lanes [32*(w%4), +32) = output rows m0 + 32q + lane.  Each
    // 32x32 block of the accumulator is transposed through shared memory (the pipeline stages are
    // idle once the accumulator is complete) so that every global store / atomic instruction covers
    // one contiguous 128-byte row segment.
    const int q = warp & 3;
    if (nkb > 0) {
      mbar_wait(bar_tmem, 0);
      tc_fence_after();
      float *stage = reinterpret_cast<float *>(gen_base) + q * 1024;  // [32][32] floats, rotated columns
      const bool add_bias = a.bias != nullptr && blockIdx.z == 0;
      const int mb = m0 + 32 * q;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tc_ld32(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(c * 32), r);
#pragma unroll
        for (int j = 0; j < 32; ++j) stage[lane * 32 + ((j + lane) & 31)] = __uint_as_float(r[j]);
        __syncwarp();
        const int n = n0 + c * 32 + lane;
        const bool n_ok = n < a.N;
        const float bv = (add_bias && n_ok) ? __ldg(a.bias + n) : 0.f;
        const int rows = min(32, a.M - mb);
        // perm_H > 0: column n of the product is (unit n/4, gate n%4); C is stored gate-major (gate*H + unit)
        const int nd = a.perm_H > 0 ? (n & 3) * a.perm_H + (n >> 2) : n;
        float *dst = a.C + (size_t)mb * a.ldc + nd;
        if (n_ok) {
          if (a.accumulate) {
            for (int rr = 0; rr < rows; ++rr) atomicAdd(dst + (size_t)rr * a.ldc, stage[rr * 32 + ((lane + rr) & 31)] + bv);
          } else {
            for (int rr = 0; rr < rows; ++rr) dst[(size_t)rr * a.ldc] = stage[rr * 32 + ((lane + rr) & 31)] + bv;
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(BN) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// Persistent variant for plain (non-accumulating) products with a 16-byte-aligned C: one CTA per SM walks
// the output tiles, the accumulator is double-buffered in TMEM so the epilogue of tile i overlaps the
// main loop of tile i+1, and C leaves through shared memory as TMA stores (which also clip the M/N edges).
// This is the shape of the hoisted input projections (M = T*B rows, K <= 512): they are bound by the
// write of C, not by the tensor pipe.
// ------------------------------------------------------------------------------------------------
constexpr int P_RING = 2 * BM * 32 * 4;  // two [128 rows][32 cols] fp32 sub-tiles (128B-swizzled): C leaves 32 columns at a time
constexpr int p_stages(int bn) { return bn == 128 ? 5 : 4; }
constexpr int p_smem(int bn) { return p_stages(bn) * (A_BYTES + bn * BK * 2) + P_RING + 1024 /*bias*/ + 1024 /*align*/ + 256; }

// EPI = 0: C (+ bias) leaves through the TMA-store ring.  EPI = 1: the tile's columns are the gate pre-activations
// of 64 hidden units (i, j, f, o interleaved per unit) and the epilogue IS the LSTM cell update of the sampler
// (model.py:82-92 with initial_state = prev_state): nothing but c, h (bf16) and the layer output goes to memory.
// epilogue warps of the persistent kernel (8 were measured slower than 4 for the cell epilogue: 21.3 vs 19.7 ms per
// 8192-sketch call, it is bound by the row-strided c / h accesses, not by issue slots)
constexpr int p_epi_warps(int epi) { return 4; }
constexpr int p_threads(int epi) { return 64 + 32 * p_epi_warps(epi); }

template <bool A_KMAJOR, bool B_KMAJOR, int BN_, int EPI = 0>
__global__ void __launch_bounds__(p_threads(EPI), 1)
gemm_tc_persist_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmC, TcArgs a) {
  constexpr int BN = BN_;
  constexpr int P_STAGES = p_stages(BN_);
  constexpr int B_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gen_base = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t s_ring = base + P_STAGES * STAGE_BYTES;
  const uint32_t s_bias = s_ring + P_RING;
  const uint32_t bars = s_bias + 1024;
  const uint32_t bar_full = bars, bar_empty = bars + 8 * P_STAGES;
  const uint32_t bar_tfull = bars + 16 * P_STAGES, bar_tempty = bar_tfull + 16, tmem_holder = bar_tempty + 16;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gen_base + (tmem_holder - base));
  float *bias_gen = reinterpret_cast<float *>(gen_base + (s_bias - base));
  uint8_t *ring_gen = gen_base + (s_ring - base);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntn = (a.N + BN - 1) / BN, ntm = (a.M + BM - 1) / BM;
  const int ntiles = ntn * ntm;
  const int nkb = (a.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P_STAGES; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar_tfull + 8 * b, 1); mbar_init(bar_tempty + 8 * b, p_epi_warps(EPI)); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmC)) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_holder), "r"(2 * BN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;

  if (warp == 0) {
    if (elect_one()) {
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int m0 = (tile / ntn) * BM, n0 = (tile % ntn) * BN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % P_STAGES, ph = (it / P_STAGES) & 1;
          mbar_wait(bar_empty + 8 * s, ph ^ 1);
          const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
          mbar_expect_tx(bar_full + 8 * s, STAGE_BYTES);
          const int k0 = kb * BK;
          if (A_KMAJOR) {
            tma_load_2d(sa, &tmA, bar_full + 8 * s, k0, m0);
          } else {
            tma_load_2d(sa, &tmA, bar_full + 8 * s, m0, k0);
            tma_load_2d(sa + A_BYTES / 2, &tmA, bar_full + 8 * s, m0 + 64, k0);
          }
          if (B_KMAJOR) {
            tma_load_2d(sb, &tmB, bar_full + 8 * s, k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * 8192, &tmB, bar_full + 8 * s, n0 + 64 * c, k0);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_KMAJOR ? 0u : 1u) << 15) |
                             ((B_KMAJOR ? 0u : 1u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(bar_tempty + 8 * buf, ((lt >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t acc = tmem_acc + (uint32_t)(buf * BN);
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % P_STAGES, ph = (it / P_STAGES) & 1;
          mbar_wait(bar_full + 8 * s, ph);
          tc_fence_after();
          const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
          const uint64_t da = make_desc(sa, a.lbo_a, a.sbo_a), db = make_desc(sb, a.lbo_b, a.sbo_b);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            tc_mma_bf16(acc, da + (uint64_t)((a.adv_a * k) >> 4), db + (uint64_t)((a.adv_b * k) >> 4), idesc,
                        (kb > 0 || k > 0) ? 1u : 0u);
          tc_commit(bar_empty + 8 * s);
        }
        tc_commit(bar_tfull + 8 * buf);
      }
    }
  } else {
    const int q = warp & 3;              // TMEM lane quarter of this warp
    const int et = threadIdx.x - 64;     // 0..127
    const int r = 32 * q + lane;         // accumulator row inside the tile
    const bool issuer = warp == 2 && lane == 0;
    int lt = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
      const int m0 = (tile / ntn) * BM, n0 = (tile % ntn) * BN;
      const int buf = lt & 1;
      // bias slice of this tile (everybody is past the previous tile's last read of it: see the barriers below)
      if constexpr (EPI == 1) {
        for (int j = et; j < BN; j += 32 * p_epi_warps(EPI)) bias_gen[j] = (a.bias != nullptr && n0 + j < a.N) ? __ldg(a.bias + n0 + j) : 0.f;
      } else {
#pragma unroll
        for (int j = 0; j < BN / 128; ++j)
          bias_gen[et + 128 * j] = (a.bias != nullptr && n0 + et + 128 * j < a.N) ? __ldg(a.bias + n0 + et + 128 * j) : 0.f;
      }
      mbar_wait(bar_tfull + 8 * buf, (lt >> 1) & 1);
      tc_fence_after();
      if constexpr (EPI == 1) {
        named_bar_sync(1, 32 * p_epi_warps(EPI));  // bias slice visible
        const int row_g = m0 + r;
        const bool row_ok = row_g < a.M;
        const bool hold = row_ok && __ldg(a.cell_xin + (size_t)row_g * 5 + 3) > 0.f;
        const int Hc = a.cell_H;
#pragma unroll 1
        for (int c = (warp - 2) >> 2; c < BN / 32; c += p_epi_warps(EPI) / 4) {  // warps of a lane quarter alternate chunks
          uint32_t v[32];
          tc_ld32(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * BN + c * 32), v);
          const int ub = (n0 + c * 32) >> 2;  // first of the 8 units of this chunk
          if (row_ok && ub < Hc) {
            float *cp = a.cell_c + (size_t)row_g * Hc + ub;
            const float4 c0 = *reinterpret_cast<const float4 *>(cp), c1 = *reinterpret_cast<const float4 *>(cp + 4);
            const float cin[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
            float cn[8];
            uint32_t hb[4];
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
              float hh[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int u = j + e;
                const float4 bv = *reinterpret_cast<const float4 *>(bias_gen + c * 32 + 4 * u);
                const float ig = sigmoid_approx(__uint_as_float(v[4 * u + 0]) + bv.x);
                const float jg = tanh_approx(__uint_as_float(v[4 * u + 1]) + bv.y);
                const float fg = sigmoid_approx(__uint_as_float(v[4 * u + 2]) + bv.z + 1.0f);
                const float og = sigmoid_approx(__uint_as_float(v[4 * u + 3]) + bv.w);
                cn[u] = fmaf(cin[u], fg, ig * jg);
                hh[e] = tanh_approx(cn[u]) * og;
              }
              const __nv_bfloat162 hv = __floats2bfloat162_rn(hh[0], hh[1]);
              hb[j >> 1] = *reinterpret_cast<const uint32_t *>(&hv);
            }
            const uint4 hq = make_uint4(hb[0], hb[1], hb[2], hb[3]);
            *reinterpret_cast<uint4 *>(a.cell_out + (size_t)row_g * a.cell_kout + ub) = hq;
            if (!hold) {
              *reinterpret_cast<float4 *>(cp) = make_float4(cn[0], cn[1], cn[2], cn[3]);
              *reinterpret_cast<float4 *>(cp + 4) = make_float4(cn[4], cn[5], cn[6], cn[7]);
              *reinterpret_cast<uint4 *>(a.cell_self + (size_t)row_g * a.cell_kself + a.cell_hoff + ub) = hq;
            }
          }
        }
        named_bar_sync(1, 32 * p_epi_warps(EPI));  // everybody is done with the bias slice before the next tile rewrites it
      } else if constexpr (EPI == 2) {
        // bf16 C: 64 columns (one 128-byte swizzled row per accumulator row) per ring slot
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          if (issuer) tma_store_wait_read_n<1>();
          named_bar_sync(1, 128);
          uint8_t *row = ring_gen + (c & 1) * 16384 + r * 128;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t v[32];
            tc_ld32(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * BN + c * 64 + hf * 32), v);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 b0 = *reinterpret_cast<const float4 *>(bias_gen + c * 64 + hf * 32 + 8 * i);
              const float4 b1 = *reinterpret_cast<const float4 *>(bias_gen + c * 64 + hf * 32 + 8 * i + 4);
              const __nv_bfloat162 p0 = __floats2bfloat162_rn(__uint_as_float(v[8 * i + 0]) + b0.x, __uint_as_float(v[8 * i + 1]) + b0.y);
              const __nv_bfloat162 p1 = __floats2bfloat162_rn(__uint_as_float(v[8 * i + 2]) + b0.z, __uint_as_float(v[8 * i + 3]) + b0.w);
              const __nv_bfloat162 p2 = __floats2bfloat162_rn(__uint_as_float(v[8 * i + 4]) + b1.x, __uint_as_float(v[8 * i + 5]) + b1.y);
              const __nv_bfloat162 p3 = __floats2bfloat162_rn(__uint_as_float(v[8 * i + 6]) + b1.z, __uint_as_float(v[8 * i + 7]) + b1.w);
              const uint4 o = make_uint4(*reinterpret_cast<const uint32_t *>(&p0), *reinterpret_cast<const uint32_t *>(&p1),
                                         *reinterpret_cast<const uint32_t *>(&p2), *reinterpret_cast<const uint32_t *>(&p3));
              *reinterpret_cast<uint4 *>(row + (((hf * 4 + i) ^ (r & 7)) << 4)) = o;
            }
          }
          fence_proxy_async_smem();
          named_bar_sync(1, 128);
          if (issuer) {
            if (n0 + c * 64 < a.N) tma_store_2d_nocommit(&tmC, s_ring + (c & 1) * 16384, n0 + c * 64, m0);
            tma_store_commit();
          }
        }
      } else {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        // ring slot c&1: the store issued two chunks ago has finished reading it
        if (issuer) tma_store_wait_read_n<1>();
        named_bar_sync(1, 128);
        uint32_t v[32];
        tc_ld32(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * BN + c * 32), v);
        uint8_t *row = ring_gen + (c & 1) * 16384 + r * 128;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 bv = *reinterpret_cast<const float4 *>(bias_gen + c * 32 + 4 * i);
          float4 o;
          o.x = __uint_as_float(v[4 * i + 0]) + bv.x; o.y = __uint_as_float(v[4 * i + 1]) + bv.y;
          o.z = __uint_as_float(v[4 * i + 2]) + bv.z; o.w = __uint_as_float(v[4 * i + 3]) + bv.w;
          *reinterpret_cast<float4 *>(row + ((i ^ (r & 7)) << 4)) = o;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (issuer) {
          if (n0 + c * 32 < a.N) tma_store_2d_nocommit(&tmC, s_ring + (c & 1) * 16384, n0 + c * 32, m0);
          tma_store_commit();  // one group per chunk, written or not, so the ring accounting stays uniform
        }
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);  // accumulator buffer may be overwritten
    }
    if (issuer) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(2 * BN) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

int make_map_uncached(CUtensorMap *m, const void *ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner, int box_outer) {
  EncodeTiledFn enc = get_encode();
  SK_CHECK(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  SK_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld * 2) % 16 == 0,
           "TMA operand must be 16-byte aligned with a row pitch that is a multiple of 8 bf16 (ld=%lld)", (long long)ld);
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SK_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (inner %lld outer %lld ld %lld)", (int)r,
           (long long)inner, (long long)outer, (long long)ld);
  return 0;
}

// Encoded tensor maps are memoised per thread: step loops (the sampler issues the same three products 300 times)
// would otherwise pay the driver's encode call on every launch.  A map only depends on address, shape and box.
struct MapKey {
  const void *ptr; int64_t d0, d1, d2, d3, s0, s1, s2; int b0, b1, b2, b3, rank, elem, sw;
  bool operator==(const MapKey &o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && d3 == o.d3 && s0 == o.s0 && s1 == o.s1 && s2 == o.s2 &&
           b0 == o.b0 && b1 == o.b1 && b2 == o.b2 && b3 == o.b3 && rank == o.rank && elem == o.elem && sw == o.sw;
  }
};
struct MapSlot { bool used = false; MapKey key; CUtensorMap map; };
constexpr int MAP_SLOTS = 128;
MapSlot *map_slot(const MapKey &k) {
  static thread_local MapSlot slots[MAP_SLOTS];
  uint64_t hsh = reinterpret_cast<uintptr_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
  hsh ^= (uint64_t)k.d0 * 0xC2B2AE3D27D4EB4Full + (uint64_t)k.d1 * 0x165667B19E3779F9ull + (uint64_t)k.b1 * 31 + (uint64_t)k.b0 + (uint64_t)k.s0 * 7;
  return &slots[(hsh >> 32) % MAP_SLOTS];
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows of `ld` elements.
int make_map(CUtensorMap *m, const void *ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner, int box_outer) {
  const MapKey key{ptr, inner, outer, 0, 0, ld, 0, 0, box_inner, box_outer, 0, 0, 2, 2, 1};
  MapSlot *slot = map_slot(key);
  if (slot->used && slot->key == key) { *m = slot->map; return 0; }
  SK_TRY(make_map_uncached(m, ptr, inner, outer, ld, box_inner, box_outer));
  slot->used = true; slot->key = key; slot->map = *m;
  return 0;
}

}  // namespace

int sk_make_tmap(CUtensorMap *m, int elem_bytes, const void *ptr, int rank, const uint64_t *dims,
                 const uint64_t *strides_bytes, const uint32_t *box, bool swizzle128) {
  SK_CHECK(rank >= 2 && rank <= 5, "tensor map rank %d unsupported", rank);
  MapKey key{ptr, (int64_t)dims[0], (int64_t)dims[1], rank > 2 ? (int64_t)dims[2] : 0, rank > 3 ? (int64_t)dims[3] : 0,
             (int64_t)strides_bytes[0], rank > 2 ? (int64_t)strides_bytes[1] : 0, rank > 3 ? (int64_t)strides_bytes[2] : 0,
             (int)box[0], (int)box[1], rank > 2 ? (int)box[2] : 0, rank > 3 ? (int)box[3] : 0, rank, elem_bytes, swizzle128 ? 1 : 0};
  MapSlot *slot = rank <= 4 ? map_slot(key) : nullptr;
  if (slot && slot->used && slot->key == key) { *m = slot->map; return 0; }
  EncodeTiledFn enc = get_encode();
  SK_CHECK(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bx[5], es[5] = {1, 1, 1, 1, 1};
  SK_CHECK(rank >= 2 && rank <= 5, "tensor map rank %d unsupported", rank);
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = enc(m, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank,
                   const_cast<void *>(ptr), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SK_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(rank %d) failed with %d", rank, (int)r);
  if (slot) { slot->used = true; slot->key = key; slot->map = *m; }
  return 0;
}

// Debug knobs for bring-up (descriptor stride sweep from the tests); 0 = use the defaults.
uint32_t g_tc_override[6] = {0, 0, 0, 0, 0, 0};

int sk_gemm_tc(skrnn_handle *h, cudaStream_t st, int M, int N, int K, const void *A, int lda, bool a_kmajor,
               const void *B, int ldb, bool b_kmajor, float *C, int ldc, const float *bias, bool accumulate) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  CUtensorMap tmA, tmB;
  if (a_kmajor) SK_TRY(make_map(&tmA, A, K, M, lda, BK, BM));
  else SK_TRY(make_map(&tmA, A, M, K, lda, 64, BK));
  if (b_kmajor) SK_TRY(make_map(&tmB, B, K, N, ldb, BK, BN));
  else SK_TRY(make_map(&tmB, B, N, K, ldb, 64, BK));
  TcArgs a = {};
  a.M = M; a.N = N; a.K = K; a.C = C; a.ldc = ldc; a.bias = bias; a.accumulate = accumulate ? 1 : 0;
  a.perm_H = accumulate ? h->gemm_perm_H : 0;
  // K-major SW128: rows of 128 B, 8-row groups 1024 B apart; next UMMA_K (16 el) is +32 B.
  // MN-major SW128: 64-element MN chunks 8192 B apart (LBO), 8-k-row groups 1024 B apart (SBO);
  // next UMMA_K is two groups = +2048 B.
  a.lbo_a = a_kmajor ? 0 : 8192; a.sbo_a = 1024; a.adv_a = a_kmajor ? 32 : 2048;
  a.lbo_b = b_kmajor ? 0 : 8192; a.sbo_b = 1024; a.adv_b = b_kmajor ? 32 : 2048;
  if (g_tc_override[0]) a.lbo_a = g_tc_override[0];
  if (g_tc_override[1]) a.sbo_a = g_tc_override[1];
  if (g_tc_override[2]) a.adv_a = g_tc_override[2];
  if (g_tc_override[3]) a.lbo_b = g_tc_override[3];
  if (g_tc_override[4]) a.sbo_b = g_tc_override[4];
  if (g_tc_override[5]) a.adv_b = g_tc_override[5];
  const int nkb = (K + BK - 1) / BK;
  static int wide_env = -1;
  if (wide_env < 0) { const char *e = getenv("SKRNN_GEMM_WIDE"); wide_env = (e && e[0] == '0') ? 0 : 1; }
  static int persist_env = -1;
  if (persist_env < 0) { const char *e = getenv("SKRNN_GEMM_PERSIST"); persist_env = (e && e[0] == '0') ? 0 : 1; }
  const bool persist = persist_env && !accumulate && (ldc % 4) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                       ((N + BN - 1) / BN) * ((M + BM - 1) / BM) >= h->num_sms;
  // 128x256 tiles: long-K accumulating products (weight gradients) and the persistent variant when N allows
  const bool wide = wide_env && N >= 256 && (persist ? ((N + 255) / 256) * ((M + BM - 1) / BM) >= h->num_sms : (accumulate && nkb >= 64));
  if (wide && b_kmajor) SK_TRY(make_map(&tmB, B, K, N, ldb, BK, 256));
  const int bn = wide ? 256 : BN;
  const int gx = (N + bn - 1) / bn, gy = (M + BM - 1) / BM;
  if (persist) {
    CUtensorMap tmC;
    {
      uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
      uint64_t str[1] = {(uint64_t)ldc * 4};
      uint32_t box[2] = {32, (uint32_t)BM};
      SK_TRY(sk_make_tmap(&tmC, 4, C, 2, dims, str, box));
    }
    a.kb_per_split = nkb;
    SkLaunch scope(h, st, SK_CAT_TC_GEMM);
    const int grid = std::min(gx * gy, h->num_sms);
#define SK_TCP_LAUNCH(AK, BKM)                                                                                      \
  do {                                                                                                              \
    if (wide) {                                                                                                     \
      SK_CUDA(cudaFuncSetAttribute(gemm_tc_persist_kernel<AK, BKM, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(256))); \
      gemm_tc_persist_kernel<AK, BKM, 256><<<grid, TC_THREADS, p_smem(256), st>>>(tmA, tmB, tmC, a);                \
    } else {                                                                                                        \
      SK_CUDA(cudaFuncSetAttribute(gemm_tc_persist_kernel<AK, BKM, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(128))); \
      gemm_tc_persist_kernel<AK, BKM, 128><<<grid, TC_THREADS, p_smem(128), st>>>(tmA, tmB, tmC, a);                \
    }                                                                                                               \
  } while (0)
    if (a_kmajor && b_kmajor) SK_TCP_LAUNCH(true, true);
    else if (a_kmajor && !b_kmajor) SK_TCP_LAUNCH(true, false);
    else if (!a_kmajor && b_kmajor) SK_TCP_LAUNCH(false, true);
    else SK_TCP_LAUNCH(false, false);
#undef SK_TCP_LAUNCH
    SK_CUDA(cudaGetLastError());
    return 0;
  }
  int nsplit = 1;
  if (accumulate) {
    int tiles = gx * gy;
    // split K so that the CTAs fill whole waves of the GPU: minimise ceil(tiles*n / slots) / n
    const int slots = (wide ? 1 : 2) * h->num_sms;
    const int nmax = std::max(1, std::min(64, (nkb + 7) / 8));
    double best = 1e30;
    for (int n = 1; n <= nmax; ++n) {
      const double cost = (double)((tiles * n + slots - 1) / slots) / n;
      if (cost < best * 0.97) { best = cost; nsplit = n; }
    }
  }
  a.kb_per_split = (nkb + nsplit - 1) / nsplit;
  nsplit = (nkb + a.kb_per_split - 1) / a.kb_per_split;
  dim3 grid(gx, gy, nsplit);
  SkLaunch scope(h, st, SK_CAT_TC_GEMM);
#define SK_TC_LAUNCH(AK, BKM)                                                                                  \
  do {                                                                                                         \
    if (wide) {                                                                                                \
      SK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<AK, BKM, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_256)); \
      gemm_tc_kernel<AK, BKM, 256><<<grid, TC_THREADS, TC_SMEM_256, st>>>(tmA, tmB, a);                        \
    } else {                                                                                                   \
      SK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<AK, BKM, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM)); \
      gemm_tc_kernel<AK, BKM, 128><<<grid, TC_THREADS, TC_SMEM, st>>>(tmA, tmB, a);                            \
    }                                                                                                          \
  } while (0)
  if (a_kmajor && b_kmajor) SK_TC_LAUNCH(true, true);
  else if (a_kmajor && !b_kmajor) SK_TC_LAUNCH(true, false);
  else if (!a_kmajor && b_kmajor) SK_TC_LAUNCH(false, true);
  else SK_TC_LAUNCH(false, false);
#undef SK_TC_LAUNCH
  SK_CUDA(cudaGetLastError());
  return 0;
}

// gates = A[M,K] * W[K,4H] + b with gate-interleaved columns (4*unit + gate), followed in the same kernel by the
// LSTM cell update of the sampler (see EPI = 1 above).  A: bf16 K-major rows (pitch lda), W: bf16 [K][4H].
int sk_gemm_tc_lstm_cell(skrnn_handle *h, cudaStream_t st, int M, int H, int K, const void *A, int lda, const void *W,
                         const float *bias_il, float *cell_c, __nv_bfloat16 *cell_self, int kself, int hoff,
                         __nv_bfloat16 *cell_out, int kout, const float *xin) {
  if (M <= 0) return 0;
  const int N = 4 * H;
  SK_CHECK(H % 8 == 0 && kself % 8 == 0 && kout % 8 == 0 && hoff % 8 == 0, "lstm cell epilogue: H and row pitches must be multiples of 8");
  CUtensorMap tmA, tmB;
  SK_TRY(make_map(&tmA, A, K, M, lda, BK, BM));
  SK_TRY(make_map(&tmB, W, N, K, N, 64, BK));
  TcArgs a = {};
  a.M = M; a.N = N; a.K = K; a.C = nullptr; a.ldc = 0; a.bias = bias_il; a.accumulate = 0;
  a.lbo_a = 0; a.sbo_a = 1024; a.adv_a = 32;
  a.lbo_b = 8192; a.sbo_b = 1024; a.adv_b = 2048;
  a.kb_per_split = (K + BK - 1) / BK;
  a.cell_c = cell_c; a.cell_self = cell_self; a.cell_out = cell_out; a.cell_xin = xin;
  a.cell_kself = kself; a.cell_hoff = hoff; a.cell_kout = kout; a.cell_H = H;
  const int tiles = ((N + 255) / 256) * ((M + BM - 1) / BM);
  const int grid = std::min(tiles, h->num_sms);
  SkLaunch scope(h, st, SK_CAT_TC_GEMM);
  SK_CUDA(cudaFuncSetAttribute(gemm_tc_persist_kernel<true, false, 256, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(256)));
  gemm_tc_persist_kernel<true, false, 256, 1><<<grid, p_threads(1), p_smem(256), st>>>(tmA, tmB, tmA, a);
  SK_CUDA(cudaGetLastError());
  return 0;
}

// C (bf16) [M,N] = A[M,K] * B[K,N] + bias through the persistent kernel (A bf16 K-major, B bf16 [K][ldb] N-contiguous).
int sk_gemm_tc_bf16out(skrnn_handle *h, cudaStream_t st, int M, int N, int K, const void *A, int lda, const void *B, int ldb,
                       __nv_bfloat16 *C, int ldc, const float *bias) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  SK_CHECK(ldc % 8 == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0, "bf16 GEMM output must be 16-byte aligned (ldc %d)", ldc);
  CUtensorMap tmA, tmB, tmC;
  SK_TRY(make_map(&tmA, A, K, M, lda, BK, BM));
  SK_TRY(make_map(&tmB, B, N, K, ldb, 64, BK));
  {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)ldc * 2};
    uint32_t box[2] = {64, (uint32_t)BM};
    SK_TRY(sk_make_tmap(&tmC, 2, C, 2, dims, str, box));
  }
  TcArgs a = {};
  a.M = M; a.N = N; a.K = K; a.C = nullptr; a.ldc = ldc; a.bias = bias; a.accumulate = 0;
  a.lbo_a = 0; a.sbo_a = 1024; a.adv_a = 32;
  a.lbo_b = 8192; a.sbo_b = 1024; a.adv_b = 2048;
  a.kb_per_split = (K + BK - 1) / BK;
  const bool wide = N >= 256;
  const int tiles = ((N + (wide ? 255 : 127)) / (wide ? 256 : 128)) * ((M + BM - 1) / BM);
  const int grid = std::min(tiles, h->num_sms);
  SkLaunch scope(h, st, SK_CAT_TC_GEMM);
  if (wide) {
    SK_CUDA(cudaFuncSetAttribute(gemm_tc_persist_kernel<true, false, 256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(256)));
    gemm_tc_persist_kernel<true, false, 256, 2><<<grid, p_threads(2), p_smem(256), st>>>(tmA, tmB, tmC, a);
  } else {
    SK_CUDA(cudaFuncSetAttribute(gemm_tc_persist_kernel<true, false, 128, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(128)));
    gemm_tc_persist_kernel<true, false, 128, 2><<<grid, p_threads(2), p_smem(128), st>>>(tmA, tmB, tmC, a);
  }
  SK_CUDA(cudaGetLastError());
  return 0;
}

