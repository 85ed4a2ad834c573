// rnn_tc.cu -- tcgen05 persistent recurrent kernels for the LSTM stack (bf16 tensor-core path):
// the per-timestep h_{t-1} * Wh contraction of reference model.py:77 (TF BasicLSTMCell inside
// custom_rnn_autodecoder, model.py:66-95) and its BPTT counterpart (model.py:182), one launch per
// layer and direction running all T steps.
//
// Partition.  A cluster of CS = H/32 CTAs owns N batch rows (N = 16..128); rank r owns hidden units
// [32r, 32r+32) and keeps the 128 gate columns of those units (row 4u+g <-> unit u, gate g in TF
// order i,j,f,o) resident in shared memory as a bf16 UMMA operand for the whole launch.
//
// Forward step (transposed product, so the MMA M dimension is the 128 gate columns and the batch
// tile is the flexible N dimension):   D^T[128, N] = Wslice[128, K=H] * h_{t-1}^T[K, N]
//   * warp 0 issues TMA loads of the h_{t-1} tile (bf16, written by all ranks in the previous step)
//     and of the hoisted input projection x*Wx+b for this step's gate columns;
//   * warp 1 issues tcgen05.mma (accumulator in TMEM) k-block by k-block as the loads land;
//   * warps 2-5 read the accumulator with tcgen05.ld; lane 4u+g holds gate g of unit u, a 3-round
//     shuffle gives every lane all four gates for its batch columns, then gate nonlinearities, cell
//     update, eoc reset select and output dropout run in registers and h_t goes back to global memory;
//   * ranks meet at one barrier.cluster per step.
// Backward step: every rank multiplies its own d(pre-activation) slice (K = its 128 gate columns,
// written to shared memory in the UMMA layout by its epilogue threads) with Wh[:, those columns],
// giving a partial d(h_{t-1}) for ALL hidden units; partials are exchanged through L2 and summed in a
// fixed rank order after the cluster barrier.
#include <stdlib.h>

#include "common.cuh"
#include "skrnn_math.h"
#include "tc_ptx.cuh"

namespace {

constexpr int TCR_THREADS = 192;

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_id_x() {
  unsigned r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }
__device__ __forceinline__ float tanhf_(float x) {
  float e = __expf(2.0f * fabsf(x));
  float r = 1.0f - 2.0f / (e + 1.0f);
  return copysignf(r, x);
}

__device__ __forceinline__ float drop_scale(const float *mask, int b, int T, int t, int H, int unit, int row0,
                                            uint64_t seed, float keep) {
  if (mask != nullptr) return mask[((int64_t)b * T + t) * H + unit] * (1.0f / keep);
  uint32_t r[4];
  skm_philox4x32((uint32_t)(row0 + b), (uint32_t)t, (uint32_t)(unit >> 2), 0x5eedu, (uint32_t)seed,
                 (uint32_t)(seed >> 32), r);
  return skm_u01(r[unit & 3]) < keep ? (1.0f / keep) : 0.0f;
}

struct TcRnnArgs {
  int B, T, H, G, N, CS, nkb;
  int tmem_cols;
  int row0, dropout;
  float keep;
  uint64_t seed;
  const float *x_bt5;
  const float *state_in;
  float *state_out;
  const float *mask;
  float *acts;            // [T,B,G] fp32: gate activations (forward out) / d(pre) (backward out)
  float *cprev;           // [T,B,H] fp32
  __nv_bfloat16 *out_bf;  // [T,B,H]
  __nv_bfloat16 *hprev_bf;
  const float *dout;      // [T,B,H] fp32
  __nv_bfloat16 *dpre_bf; // [T,B,G]
  float *partial;         // [nclusters][2][CS][H][N]
  const uint8_t *wimg;    // per-rank shared-memory images of the weight operand
  int wimg_bytes;
};

// byte offset of fp32 element (row c, unit u) inside one 128B-swizzled staging plane [N][32]
__device__ __forceinline__ uint32_t plane_off(int c, int u) {
  return (uint32_t)(c * 128 + ((((u >> 2) ^ (c & 7)) & 7) << 4) + ((u & 3) << 2));
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <int NCH>  // N = 16 * NCH batch columns per cluster
__global__ void __launch_bounds__(TCR_THREADS, 1)
lstm_fwd_tc_kernel(const __grid_constant__ CUtensorMap tm_h, const __grid_constant__ CUtensorMap tm_xw, TcRnnArgs a) {
  constexpr int N = 16 * NCH;
  constexpr int NIT = N / 4;  // (unit, column) items per epilogue thread
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const int H = a.H, G = a.G, B = a.B, T = a.T, nkb = a.nkb;
  const uint32_t s_w = base;                                  // nkb x [128 rows][128 B]
  const uint32_t s_h = s_w + (uint32_t)nkb * 16384u;          // nkb x [N rows][128 B]
  const uint32_t s_xw = s_h + (uint32_t)nkb * (N * 128);      // 4 planes x [N][32] fp32
  const uint32_t s_bar = s_xw + 4u * (N * 128);
  const uint32_t bar_h = s_bar;          // [8]
  const uint32_t bar_acc = s_bar + 64;
  const uint32_t bar_xw = s_bar + 72;
  const uint32_t tmem_holder = s_bar + 80;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));
  const uint8_t *xw_gen = gbase + (s_xw - base);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int b0 = cluster_id_x() * N;
  const int u0 = rank * 32;
  const int S = 2 * H;

  // resident weight operand: straight copy of this rank's pre-swizzled image
  {
    const uint4 *src = reinterpret_cast<const uint4 *>(a.wimg + (size_t)rank * a.wimg_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(gbase + (s_w - base));
    for (int i = tid; i < a.wimg_bytes / 16; i += TCR_THREADS) dst[i] = __ldg(src + i);
  }
  if (tid == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(bar_h + 8 * i, 1);
    mbar_init(bar_acc, 1);
    mbar_init(bar_xw, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);

  // epilogue thread <-> accumulator row e = 32*(warp%4) + lane = 4u+g: unit u, gate g; its items are
  // the batch columns c = 4*it + g
  const int e = 32 * (warp & 3) + lane;
  const int eu = (e >> 2) & 31, eg = e & 3;
  float cst[NIT];
  if (warp >= 2) {
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int b = b0 + 4 * it + eg;
      cst[it] = 0.f;
      if (b < B) {
        const float ci = a.state_in[(size_t)b * S + u0 + eu];
        const float hi = a.state_in[(size_t)b * S + H + u0 + eu];
        cst[it] = ci;
        a.cprev[((size_t)0 * B + b) * H + u0 + eu] = ci;
        a.hprev_bf[((size_t)0 * B + b) * H + u0 + eu] = __float2bfloat16_rn(hi);
      }
    }
  }
  __threadfence();
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  cluster_sync_all();

  const uint32_t idesc = make_idesc_bf16(128, N, false, false);

  for (int t = 0; t < T; ++t) {
    if (warp == 0) {
      if (lane == 0) {
        fence_proxy_async();
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_expect_tx(bar_h + 8 * kb, N * 128);
          tma_load_3d(s_h + (uint32_t)kb * (N * 128), &tm_h, bar_h + 8 * kb, kb * 64, b0, t);
        }
        mbar_expect_tx(bar_xw, 4 * N * 128);
        for (int g = 0; g < 4; ++g) tma_load_3d(s_xw + (uint32_t)g * (N * 128), &tm_xw, bar_xw, g * H + u0, b0, t);
      }
    } else if (warp == 1) {
      if (lane == 0) {
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(bar_h + 8 * kb, t & 1);
          tc_fence_after();
          const uint64_t da = make_desc(s_w + (uint32_t)kb * 16384u, 0, 1024);
          const uint64_t db = make_desc(s_h + (uint32_t)kb * (N * 128), 0, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(tmem_acc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
        }
        tc_commit(bar_acc);
      }
    } else {
      const int q = warp & 3;  // TMEM lane quarter this warp may read
      // lane index within the 128 accumulator rows = 32q + lane; must equal e for the gate layout
      mbar_wait(bar_acc, t & 1);
      tc_fence_after();
      mbar_wait(bar_xw, t & 1);
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        uint32_t r[16];
        tc_ld16(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(ch * 16), r);
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const int it = ch * 4 + qq;
          const float x0 = __uint_as_float(r[4 * qq + 0]), x1 = __uint_as_float(r[4 * qq + 1]);
          const float x2 = __uint_as_float(r[4 * qq + 2]), x3 = __uint_as_float(r[4 * qq + 3]);
          // exchange: after round rr this lane holds gate (eg^rr) of column 4*it + eg
          float rc[4];
          rc[0] = eg == 0 ? x0 : (eg == 1 ? x1 : (eg == 2 ? x2 : x3));
#pragma unroll
          for (int rr = 1; rr < 4; ++rr) {
            const int sel = eg ^ rr;
            const float send = sel == 0 ? x0 : (sel == 1 ? x1 : (sel == 2 ? x2 : x3));
            rc[rr] = __shfl_xor_sync(0xffffffffu, send, rr);
          }
          float pre[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int rr = k ^ eg;
            pre[k] = rr == 0 ? rc[0] : (rr == 1 ? rc[1] : (rr == 2 ? rc[2] : rc[3]));
          }
          const int c = 4 * it + eg;
          const int b = b0 + c;
          if (b < B) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              pre[k] += *reinterpret_cast<const float *>(xw_gen + (size_t)k * (N * 128) + plane_off(c, eu));
            const float ig = sigmoidf_(pre[0]), jg = tanhf_(pre[1]), fg = sigmoidf_(pre[2] + 1.0f), og = sigmoidf_(pre[3]);
            const float c2 = cst[it] * fg + ig * jg;
            const float h2 = tanhf_(c2) * og;
            float *ac = a.acts + ((size_t)t * B + b) * G + u0 + eu;
            ac[0] = ig; ac[H] = jg; ac[2 * H] = fg; ac[3 * H] = og;
            const bool eoc = a.x_bt5[((size_t)b * T + t) * 5 + 3] > 0.f;
            float ov = h2;
            if (a.dropout) ov *= drop_scale(a.mask, b, T, t, H, u0 + eu, a.row0, a.seed, a.keep);
            a.out_bf[((size_t)t * B + b) * H + u0 + eu] = __float2bfloat16_rn(ov);
            const float cn = eoc ? a.state_in[(size_t)b * S + u0 + eu] : c2;
            const float hn = eoc ? a.state_in[(size_t)b * S + H + u0 + eu] : h2;
            cst[it] = cn;
            if (t + 1 < T) {
              a.cprev[((size_t)(t + 1) * B + b) * H + u0 + eu] = cn;
              a.hprev_bf[((size_t)(t + 1) * B + b) * H + u0 + eu] = __float2bfloat16_rn(hn);
            } else {
              a.state_out[(size_t)b * S + u0 + eu] = cn;
              a.state_out[(size_t)b * S + H + u0 + eu] = hn;
            }
          }
        }
      }
      __threadfence();
      fence_proxy_async();
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
  }
  __syncthreads();
  if (warp == 1) tc_dealloc(tmem_acc, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
template <int NCH>
__global__ void __launch_bounds__(TCR_THREADS, 1)
lstm_bwd_tc_kernel(const __grid_constant__ CUtensorMap tm_acts, TcRnnArgs a) {
  constexpr int N = 16 * NCH;
  constexpr int NIT = N / 4;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const int H = a.H, G = a.G, B = a.B, T = a.T, CS = a.CS;
  const int nmt = (H + 127) / 128;                 // accumulator m-tiles (all hidden units)
  const int HP = nmt * 128;
  const uint32_t s_w = base;                       // 2 k-blocks x [HP rows][128 B]
  const uint32_t s_b = s_w + 2u * HP * 128u;       // 2 k-blocks x [N rows][128 B]: own d(pre) slice, bf16
  const uint32_t s_ac = s_b + 2u * (N * 128);      // 4 planes x [N][32] fp32: saved gate activations
  const uint32_t s_bar = s_ac + 4u * (N * 128);
  const uint32_t bar_acts = s_bar, bar_ready = s_bar + 8, bar_acc = s_bar + 16, tmem_holder = s_bar + 24;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));
  const uint8_t *ac_gen = gbase + (s_ac - base);
  uint8_t *b_gen = gbase + (s_b - base);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int cid = cluster_id_x();
  const int b0 = cid * N;
  const int u0 = rank * 32;

  {
    const uint4 *src = reinterpret_cast<const uint4 *>(a.wimg + (size_t)rank * a.wimg_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(gbase + (s_w - base));
    for (int i = tid; i < a.wimg_bytes / 16; i += TCR_THREADS) dst[i] = __ldg(src + i);
    uint4 *bz = reinterpret_cast<uint4 *>(b_gen);
    for (int i = tid; i < 2 * N * 128 / 16; i += TCR_THREADS) bz[i] = make_uint4(0, 0, 0, 0);
  }
  if (tid == 0) {
    mbar_init(bar_acts, 1);
    mbar_init(bar_ready, 128);
    mbar_init(bar_acc, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  cluster_sync_all();

  // epilogue thread e: unit u = e/4, column block g = e%4 -> columns [g*NIT, (g+1)*NIT)
  const int e = 32 * (warp & 3) + lane;
  const int eu = (e >> 2) & 31, eg = e & 3;
  float dhc[NIT], dcc[NIT];
#pragma unroll
  for (int it = 0; it < NIT; ++it) { dhc[it] = 0.f; dcc[it] = 0.f; }
  const uint32_t idesc = make_idesc_bf16(128, N, false, false);
  float *part_base = a.partial + (size_t)cid * 2 * CS * H * N;

  for (int t = T - 1, step = 0; t >= 0; --t, ++step) {
    const int par = step & 1;
    if (warp == 0) {
      if (lane == 0) {
        mbar_expect_tx(bar_acts, 4 * N * 128);
        for (int g = 0; g < 4; ++g) tma_load_3d(s_ac + (uint32_t)g * (N * 128), &tm_acts, bar_acts, g * H + u0, b0, t);
      }
    } else if (warp >= 2) {
      // prefetch the per-item global inputs before waiting on the staged activations
      float d_o[NIT], cp[NIT];
      bool eoc[NIT];
#pragma unroll
      for (int it = 0; it < NIT; ++it) {
        const int b = b0 + eg * NIT + it;
        d_o[it] = 0.f; cp[it] = 0.f; eoc[it] = false;
        if (b < B) {
          d_o[it] = a.dout[((size_t)t * B + b) * H + u0 + eu];
          cp[it] = a.cprev[((size_t)t * B + b) * H + u0 + eu];
          eoc[it] = a.x_bt5[((size_t)b * T + t) * 5 + 3] > 0.f;
        }
      }
      mbar_wait(bar_acts, step & 1);
#pragma unroll
      for (int it = 0; it < NIT; ++it) {
        const int c = eg * NIT + it;
        const int b = b0 + c;
        if (b < B) {
          const float ig = *reinterpret_cast<const float *>(ac_gen + 0 * (N * 128) + plane_off(c, eu));
          const float jg = *reinterpret_cast<const float *>(ac_gen + 1 * (N * 128) + plane_off(c, eu));
          const float fg = *reinterpret_cast<const float *>(ac_gen + 2 * (N * 128) + plane_off(c, eu));
          const float og = *reinterpret_cast<const float *>(ac_gen + 3 * (N * 128) + plane_off(c, eu));
          float dd = d_o[it];
          if (a.dropout) dd *= drop_scale(a.mask, b, T, t, H, u0 + eu, a.row0, a.seed, a.keep);
          const float dh2 = dd + (eoc[it] ? 0.f : dhc[it]);
          const float tc = tanhf_(cp[it] * fg + ig * jg);
          const float dc2 = (eoc[it] ? 0.f : dcc[it]) + dh2 * og * (1.0f - tc * tc);
          dcc[it] = dc2 * fg;
          const float d0 = dc2 * jg * ig * (1.0f - ig), d1 = dc2 * ig * (1.0f - jg * jg);
          const float d2 = dc2 * cp[it] * fg * (1.0f - fg), d3 = dh2 * tc * og * (1.0f - og);
          float *ac = a.acts + ((size_t)t * B + b) * G + u0 + eu;
          ac[0] = d0; ac[H] = d1; ac[2 * H] = d2; ac[3 * H] = d3;
          __nv_bfloat16 *db = a.dpre_bf + ((size_t)t * B + b) * G + u0 + eu;
          const __nv_bfloat16 h0 = __float2bfloat16_rn(d0), h1 = __float2bfloat16_rn(d1);
          const __nv_bfloat16 h2 = __float2bfloat16_rn(d2), h3 = __float2bfloat16_rn(d3);
          db[0] = h0; db[H] = h1; db[2 * H] = h2; db[3 * H] = h3;
          // own K slice, local column kk = 4u+g: 8 bytes at (row c, kk = 4u) of the swizzled operand
          const int kk = 4 * eu;
          const uint32_t off = (uint32_t)((kk >> 6) * (N * 128) + c * 128 + (((((kk & 63) >> 3) ^ (c & 7)) & 7) << 4) +
                                          ((kk & 7) << 1));
          uint2 pk;
          pk.x = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          pk.y = (uint32_t)__bfloat16_as_ushort(h2) | ((uint32_t)__bfloat16_as_ushort(h3) << 16);
          *reinterpret_cast<uint2 *>(b_gen + off) = pk;
        }
      }
      fence_proxy_async();
      mbar_arrive(bar_ready);
    }
    if (t > 0) {
      if (warp == 1) {
        if (lane == 0) {
          mbar_wait(bar_ready, step & 1);
          tc_fence_after();
          for (int mt = 0; mt < nmt; ++mt) {
#pragma unroll
            for (int kb = 0; kb < 2; ++kb) {
              const uint64_t da = make_desc(s_w + (uint32_t)kb * (HP * 128) + (uint32_t)mt * 16384u, 0, 1024);
              const uint64_t db = make_desc(s_b + (uint32_t)kb * (N * 128), 0, 1024);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc_mma_bf16(tmem_acc + (uint32_t)(mt * N), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                            (kb > 0 || k > 0) ? 1u : 0u);
            }
          }
          tc_commit(bar_acc);
        }
      } else if (warp >= 2) {
        const int q = warp & 3;
        mbar_wait(bar_acc, step & 1);
        tc_fence_after();
        for (int mt = 0; mt < nmt; ++mt) {
          const int m = mt * 128 + 32 * q + lane;  // hidden unit this lane's accumulator row belongs to
          float *dst = part_base + ((size_t)(par * CS + rank) * H + m) * N;
#pragma unroll
          for (int ch = 0; ch < NCH; ++ch) {
            uint32_t r[16];
            tc_ld16(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(mt * N + ch * 16), r);
            if (m < H) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4 *>(dst + ch * 16 + 4 * j) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
            }
          }
        }
        __threadfence();
      }
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    if (t > 0 && warp >= 2) {
      // d(h_{t-1}) for my units = sum over ranks (fixed order) of their partials
#pragma unroll
      for (int it = 0; it < NIT; ++it) dhc[it] = 0.f;
      for (int r = 0; r < CS; ++r) {
        const float *src = part_base + ((size_t)(par * CS + r) * H + u0 + eu) * N + eg * NIT;
#pragma unroll
        for (int it = 0; it < NIT; it += 4) {
          const float4 v = __ldcg(reinterpret_cast<const float4 *>(src + it));
          dhc[it] += v.x; dhc[it + 1] += v.y; dhc[it + 2] += v.z; dhc[it + 3] += v.w;
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) tc_dealloc(tmem_acc, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// weight images: what each rank copies into shared memory, already in the 128B-swizzled K-major
// UMMA layout.  Wh = rows [I, I+H) of the layer's TF Linear/Matrix, column g*H+v = gate g of unit v.
//   forward : nkb k-blocks x [128 rows m=4u+g][64 k]: value Wh[k][g*H + u0+u]
//   backward: 2 k-blocks x [HP rows m][64 kk], kk = 4u+g: value Wh[m][g*H + u0+u]  (zero for m >= H)
// ------------------------------------------------------------------------------------------------
__global__ void pack_tc_image_kernel(int bwd, const float *__restrict__ mat, int I, int H, int CS, int nkb, int HP,
                                     int img_bytes, uint8_t *__restrict__ dst) {
  const int64_t per_rank = img_bytes / 2;  // bf16 elements
  const int64_t total = per_rank * CS;
  const int G = 4 * H;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int rank = (int)(i / per_rank);
    const int64_t o = i % per_rank;
    const int u0 = rank * 32;
    float v = 0.f;
    int rows = bwd ? HP : 128;
    const int kb = (int)(o / ((int64_t)rows * 64));
    const int64_t w = o % ((int64_t)rows * 64);
    const int m = (int)(w / 64);
    const int within = (int)(w % 64);          // position inside the 128-byte row, in elements
    const int chunk = within >> 3, el = within & 7;
    const int kloc = ((chunk ^ (m & 7)) << 3) + el;  // un-swizzle: logical k within the block
    const int k = kb * 64 + kloc;
    if (!bwd) {
      const int u = m >> 2, g = m & 3;
      if (k < H) v = mat[(size_t)(I + k) * G + g * H + u0 + u];
    } else {
      const int u = k >> 2, g = k & 3;  // k = local gate column 4u+g, 0..127
      if (m < H) v = mat[(size_t)(I + m) * G + g * H + u0 + u];
    }
    reinterpret_cast<__nv_bfloat16 *>(dst)[i] = __float2bfloat16_rn(v);
  }
}

template <int NCH>
int launch_fwd(const CUtensorMap &tm_h, const CUtensorMap &tm_xw, const TcRnnArgs &a, int nclusters, size_t smem,
               cudaStream_t st) {
  auto kern = lstm_fwd_tc_kernel<NCH>;
  SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (a.CS > 8) SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * a.CS);
  cfg.blockDim = dim3(TCR_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = a.CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm_h, tm_xw, a));
  return 0;
}
template <int NCH>
int launch_bwd(const CUtensorMap &tm_acts, const TcRnnArgs &a, int nclusters, size_t smem, cudaStream_t st) {
  auto kern = lstm_bwd_tc_kernel<NCH>;
  SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (a.CS > 8) SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * a.CS);
  cfg.blockDim = dim3(TCR_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = a.CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm_acts, a));
  return 0;
}

size_t smem_need(bool bwd, int H, int N) {
  const int nkb = (H + 63) / 64, HP = (H + 127) / 128 * 128;
  size_t s = bwd ? (size_t)2 * HP * 128 + 2 * (size_t)N * 128 + 4 * (size_t)N * 128
                 : (size_t)nkb * 16384 + (size_t)nkb * N * 128 + 4 * (size_t)N * 128;
  return s + 1024 + 256;
}

}  // namespace

int sk_rnn_tc_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *x_bt5,
                    const float *state_in, float *state_out, const float *mask, uint64_t seed, int row0,
                    bool training) {
  const int H = h->H;
  if (h->cfg.precision != SKRNN_PREC_BF16 || h->cfg.cell != SKRNN_CELL_LSTM) return -1;
  if (!(H == 32 || H == 64 || H == 128 || H == 256 || H == 512)) return -1;
  if (h->tc_rnn_disabled) return -1;
  {
    static int env_off = -1;
    if (env_off < 0) { const char *e = getenv("SKRNN_DISABLE_TC_RNN"); env_off = (e && e[0] == '1') ? 1 : 0; }
    if (env_off) return -1;
  }
  const int CS = H / 32;
  const int nkb = (H + 63) / 64, HP = (H + 127) / 128 * 128;
  // batch tile: the smallest N whose cluster count fits on the GPU at once, within shared memory / TMEM
  int N = 0;
  const int cap_clusters = std::max(1, h->num_sms / CS - (CS >= 8 ? 2 : 0));
  for (int cand : {16, 32, 64, 128}) {
    if (smem_need(bwd, H, cand) > 227 * 1024) break;
    if (bwd && (HP / 128) * cand > 512) break;
    N = cand;
    if ((B + cand - 1) / cand <= cap_clusters) break;
  }
  if (N == 0) return -1;
  const int nclusters = (B + N - 1) / N;
  const size_t smem = smem_need(bwd, H, N);
  const int img_bytes = bwd ? 2 * HP * 128 : nkb * 16384;
  SK_CHECK((int64_t)img_bytes * CS <= h->ws.wpack_floats * 4, "weight image workspace too small");
  {
    SkLaunch scope(h, st, SK_CAT_MISC);
    int64_t total = (int64_t)img_bytes / 2 * CS;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)4 * h->num_sms);
    pack_tc_image_kernel<<<blocks, 256, 0, st>>>(bwd ? 1 : 0, h->params + h->lp[l].w[0], h->lp[l].I, H, CS, nkb, HP,
                                                img_bytes, reinterpret_cast<uint8_t *>(h->ws.wpack));
    SK_CUDA(cudaGetLastError());
  }
  TcRnnArgs a = {};
  a.B = B; a.T = T; a.H = H; a.G = h->G; a.N = N; a.CS = CS; a.nkb = nkb;
  int cols = bwd ? (HP / 128) * N : N;
  int tc = 32;
  while (tc < cols) tc <<= 1;
  a.tmem_cols = tc;
  a.row0 = row0;
  a.dropout = (training && l == h->L - 1 && h->cfg.keep_prob < 1.0f) ? 1 : 0;
  a.keep = h->cfg.keep_prob;
  a.seed = seed;
  a.x_bt5 = x_bt5; a.state_in = state_in; a.state_out = state_out; a.mask = mask;
  a.acts = h->ws.xw[l]; a.cprev = h->ws.cprev[l]; a.out_bf = h->ws.out_bf[l]; a.hprev_bf = h->ws.hprev_bf[l];
  a.dout = h->ws.dout; a.dpre_bf = h->ws.dpre_bf;
  a.wimg = reinterpret_cast<const uint8_t *>(h->ws.wpack);
  a.wimg_bytes = img_bytes;
  CUtensorMap tm_xw;
  {
    uint64_t dims[3] = {(uint64_t)h->G, (uint64_t)B, (uint64_t)T};
    uint64_t str[2] = {(uint64_t)h->G * 4, (uint64_t)B * h->G * 4};
    uint32_t box[3] = {32, (uint32_t)N, 1};
    SK_TRY(sk_make_tmap(&tm_xw, 4, h->ws.xw[l], 3, dims, str, box));
  }
  int rc = 0;
  if (!bwd) {
    CUtensorMap tm_h;
    uint64_t dims[3] = {(uint64_t)H, (uint64_t)B, (uint64_t)T};
    uint64_t str[2] = {(uint64_t)H * 2, (uint64_t)B * H * 2};
    uint32_t box[3] = {64, (uint32_t)N, 1};
    SK_TRY(sk_make_tmap(&tm_h, 2, h->ws.hprev_bf[l], 3, dims, str, box));
    SkLaunch scope(h, st, SK_CAT_RNN_FWD);
    if (N == 16) rc = launch_fwd<1>(tm_h, tm_xw, a, nclusters, smem, st);
    else if (N == 32) rc = launch_fwd<2>(tm_h, tm_xw, a, nclusters, smem, st);
    else if (N == 64) rc = launch_fwd<4>(tm_h, tm_xw, a, nclusters, smem, st);
    else rc = launch_fwd<8>(tm_h, tm_xw, a, nclusters, smem, st);
  } else {
    const int64_t need = (int64_t)nclusters * 2 * CS * H * N;
    if (need > h->ws.partial_floats) {
      SK_CUDA(cudaStreamSynchronize(st));
      cudaFree(h->ws.partial);
      h->ws.partial = nullptr;
      h->ws.partial_floats = 0;
      SK_CUDA(cudaMalloc(&h->ws.partial, need * sizeof(float)));
      h->ws.partial_floats = need;
    }
    a.partial = h->ws.partial;
    SkLaunch scope(h, st, SK_CAT_RNN_BWD);
    if (N == 16) rc = launch_bwd<1>(tm_xw, a, nclusters, smem, st);
    else if (N == 32) rc = launch_bwd<2>(tm_xw, a, nclusters, smem, st);
    else if (N == 64) rc = launch_bwd<4>(tm_xw, a, nclusters, smem, st);
    else rc = launch_bwd<8>(tm_xw, a, nclusters, smem, st);
  }
  if (rc != 0) {
    // e.g. a 16-CTA cluster that this GPU cannot co-schedule: fall back to the CUDA-core recurrent kernels
    fprintf(stderr, "libskrnn: tcgen05 recurrent kernel launch failed (%s); using the fp32 recurrent kernels\n", skrnn_last_error());
    cudaGetLastError();
    h->tc_rnn_disabled = true;
    return -1;
  }
  return 0;
}
