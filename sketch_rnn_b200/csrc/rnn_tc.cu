// rnn_tc.cu -- tcgen05 persistent recurrent kernels for the LSTM stack (bf16 tensor-core path):
// the per-timestep h_{t-1} * Wh contraction of reference model.py:77 (TF BasicLSTMCell inside
// custom_rnn_autodecoder, model.py:66-95) and its BPTT counterpart (model.py:182), one launch per
// layer and direction running all T steps.
//
// Partition.  A cluster of CS = H/32 CTAs owns N batch rows (N = 16..128); rank r owns hidden units
// [32r, 32r+32) and keeps the 128 gate columns of those units (row 4u+g <-> unit u, gate g in TF
// order i,j,f,o) resident for the whole launch: in tensor memory as the A operand of the forward
// product, in shared memory (UMMA K-major, 128B swizzle) for the backward product.
//
// Forward step (transposed product, so the MMA M dimension is the 128 gate columns and the batch
// tile is the flexible N dimension):   D^T[128, N] = Wslice[128, K=H] * h_{t-1}^T[K, N]
//   * warp 0 issues the TMA loads: its share of the h_{t-1} tile (multicast to the whole cluster) and
//     the hoisted projection x*Wx+b of its 32 units (bf16, four gates of a unit adjacent: xw_il);
//   * warp 1 issues tcgen05.mma (accumulator in TMEM);
//   * warps 2-17 read the accumulator with tcgen05.ld; lane 4u+g holds gate g of unit u, a branch-free
//     4-lane exchange gives every lane all four gates for its batch columns, then gate nonlinearities,
//     cell update, eoc reset select and output dropout run in registers; the gate activations go to a
//     bf16 record for BPTT (sign bits of i and f carry the eoc and dropout flags).
// Backward step: every rank multiplies its own d(pre-activation) slice (K = its 128 gate columns,
// written to shared memory in the UMMA layout by its elementwise threads, and sent to the d(pre) array
// from there by TMA) with Wh[:, those columns], giving a partial d(h_{t-1}) for ALL hidden units;
// partials are exchanged as bf16 through L2 and summed in a fixed rank order.
//
// Two generations of kernels live here.  lstm_{fwd,bwd}_tc_kernel: one tile per cluster, one
// barrier.cluster per step (small batches).  lstm_{fwd,bwd}_pp_kernel: two half tiles in flight, no
// cluster barrier in the loop -- a publisher warp signals every rank's mbarrier through the cluster
// shared-memory window once the CTA's h_t (forward, TMA-stored) or partials (backward) are out.
//
// Small tiles (up to 32 rows per cluster) are pure latency chains, and most of a step used to be the exchange through
// global memory.  There the ranks send to each other directly: forward, the publisher warp st.async-es the CTA's rows
// of h_t into every rank's operand tile of the next step (TcRnnArgs::push, lstm_fwd_pp_kernel); backward, every
// elementwise thread st.async-es its partial sums into the destination rank's receive buffer (lstm_bwd_tc_kernel).
// Either way the stores' bytes complete an mbarrier phase on the receiver: no fence, no barrier.cluster, no L2 trip.
#include <stdlib.h>
#include <type_traits>

#include "common.cuh"
#include "skrnn_math.h"
#include "tc_ptx.cuh"

namespace {

constexpr int TCR_EPI_WARPS = 16;                      // 4 TMEM lane quarters x 4 column groups
constexpr int TCR_THREADS = 64 + 32 * TCR_EPI_WARPS;  // warp 0: TMA, warp 1: MMA, warps 2..17: epilogue

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory"); }
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

__device__ __forceinline__ float drop_scale(const float *mask, int b, int T, int t, int H, int unit, int row0,
                                            uint64_t seed, float keep) {
  if (mask != nullptr) return mask[((int64_t)b * T + t) * H + unit] * (1.0f / keep);
  return skm_u01(skm_drop_bits(seed, (uint32_t)(row0 + b), (uint32_t)t, (uint32_t)unit)) < keep ? (1.0f / keep) : 0.0f;
}

struct TcRnnArgs {
  int B, T, H, G, N, CS, nkb;
  int tmem_cols;
  int row0, dropout;
  float keep;
  uint64_t seed;
  const unsigned char *eoc;  // [T,B] end-of-content flag of the input at (t, b)
  const float *state_in;
  float *state_out;
  const float *mask;
  float *acts;            // [T,B,G] fp32: gate activations (forward out) / d(pre) (backward out)
  float *cprev;           // [T,B,H] fp32
  __nv_bfloat16 *out_bf;  // [T,B,H]
  __nv_bfloat16 *hprev_bf;
  __nv_bfloat16 *acts_bf; // [T,B,H,4]: gate activations + flags saved for BPTT
  const float *dout;      // [T,B,H] fp32
  __nv_bfloat16 *dpre_bf; // [T,B,G]
  __nv_bfloat16 *partial; // [nclusters][2][CS][H][N] partial d(h), bf16
  float *dbias;           // [G] bias gradient accumulator (column sums of d(pre)), or null
  const uint8_t *wimg;    // per-rank shared-memory images of the weight operand
  int wimg_bytes;
  int mcast;              // forward: the h tile arrives by TMA multicast (each rank fetches 1/CS of it)
  int sig_async;          // forward, two-half-tile kernel: h stores are announced with st.async instead of release-arrives
  int push;               // forward, two-half-tile kernel, small tiles: every rank sends its units of h_t straight into every
                          // rank's operand tile (st.async + transaction bytes) instead of store -> signal -> multicast load
  int a_tmem;             // forward: weights live in tensor memory (wimg is plain row-major [128][nkb*64] bf16)
  int debug;              // SKRNN_TC_DEBUG=1: one CTA prints its per-phase cycle split
  // time range of this launch (two-half-tile kernels): steps [t0, t1) of the T-step unroll, t0 even.  A layer may be
  // run as several launches so that the GEMMs around it overlap the neighbouring ranges (api.cu)
  int t0, t1;
  float *carry;           // [2][B][H] fp32: d(h), d(c) handed from one BPTT range to the next
};

// Staging planes are [N columns][32 units] fp32 with the TMA 128B swizzle: element (c, u) sits at byte
// c*128 + (((u>>2) ^ (c&7)) << 4) + ((u&3) << 2).
// debug timing: one running 32-bit timestamp per thread, laps accumulate into shared-memory slots
#define TCR_LAP(slot) do { if (dbg) { const uint32_t now_ = (uint32_t)clock(); (slot) += (long long)(now_ - lapt); lapt = now_; } } while (0)

// y_k = x[g ^ k] for g = lo + 2*hi, branch-free (8 selects)
__device__ __forceinline__ void perm4(bool lo, bool hi, float x0, float x1, float x2, float x3, float &y0, float &y1,
                                      float &y2, float &y3) {
  const float a0 = lo ? x1 : x0, a1 = lo ? x0 : x1, a2 = lo ? x3 : x2, a3 = lo ? x2 : x3;
  y0 = hi ? a2 : a0; y1 = hi ? a3 : a1; y2 = hi ? a0 : a2; y3 = hi ? a1 : a3;
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
// N = 16*NCH batch columns per cluster.  Epilogue warp (quarter q = warp%4, column group cg) owns
// accumulator rows 32q..32q+31 (row e = 4u+g: unit u, gate g) and columns [cg*N/4, (cg+1)*N/4).
template <int NCH, bool DBG>
__global__ void __launch_bounds__(TCR_THREADS, 1)
lstm_fwd_tc_kernel(const __grid_constant__ CUtensorMap tm_h, const __grid_constant__ CUtensorMap tm_xw,
                   const __grid_constant__ CUtensorMap tm_act, TcRnnArgs a) {
  constexpr int N = 16 * NCH;
  constexpr int CPW = N / 4;  // columns per epilogue warp
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const int H = a.H, B = a.B, T = a.T, nkb = a.nkb;
  const uint32_t s_w = base;                                  // nkb x [128 rows][128 B] (unless weights sit in TMEM)
  const uint32_t s_h = s_w + (a.a_tmem ? 0u : (uint32_t)nkb * 16384u);  // nkb x [N rows][128 B]
  const uint32_t s_xw = s_h + (uint32_t)nkb * (N * 128);      // [N][32 units][4 gates] bf16: hoisted x*Wx+b of the step
  const uint32_t s_act = s_xw + (uint32_t)N * 256u;               // [N][32 units][4 gates] bf16: activations saved for BPTT
  const uint32_t s_bar = s_act + (uint32_t)N * 256u;
  const uint32_t bar_h = s_bar;          // [8]
  const uint32_t bar_acc = s_bar + 64;
  const uint32_t bar_xw = s_bar + 72;
  const uint32_t bar_out = s_bar + 88;   // epilogue -> warp 0: gate activations are staged, the xw planes are consumed
  const uint32_t tmem_holder = s_bar + 80;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));
  uint8_t *xw_gen = gbase + (s_xw - base);
  uint8_t *act_gen = gbase + (s_act - base);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int b0 = cluster_id_x() * N;
  const int u0 = rank * 32;
  const int S = 2 * H;

  if (!a.a_tmem) {  // resident weight operand: straight copy of this rank's pre-swizzled image
    const uint4 *src = reinterpret_cast<const uint4 *>(a.wimg + (size_t)rank * a.wimg_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(gbase + (s_w - base));
    for (int i = tid; i < a.wimg_bytes / 16; i += TCR_THREADS) dst[i] = __ldg(src + i);
  }
  if (tid == 0) {
    if (a.mcast) {
      for (int kb = 0; kb < nkb; ++kb) mbar_init(bar_h + 8 * kb, 1);  // one per k-block: the product starts as blocks land
    } else {
      mbar_init(bar_h, 32 * TCR_EPI_WARPS);
    }
    mbar_init(bar_acc, 1);
    mbar_init(bar_xw, 1);
    mbar_init(bar_out, 32 * TCR_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);
  {
    uint4 *hz = reinterpret_cast<uint4 *>(gbase + (s_h - base));
    for (int i = tid; i < nkb * N * 8; i += TCR_THREADS) hz[i] = make_uint4(0, 0, 0, 0);
  }

  const int q = warp & 3, cg = (warp - 2) >> 2;
  const int e = 32 * q + lane;
  const int eu = e >> 2, eg = e & 3;
  const int uo = u0 + eu;
  float cst[NCH];
  if (warp >= 2) {
#pragma unroll
    for (int it = 0; it < NCH; ++it) {
      const int c = cg * CPW + 4 * it + eg;
      const int b = b0 + c;
      cst[it] = 0.f;
      if (b < B) {
        const float ci = a.state_in[(size_t)b * S + uo];
        const float hi = a.state_in[(size_t)b * S + H + uo];
        cst[it] = ci;
        a.cprev[(size_t)b * H + uo] = ci;
        a.hprev_bf[(size_t)b * H + uo] = __float2bfloat16_rn(hi);
      }
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  const uint32_t tmem_a = tmem_acc + (uint32_t)((N + 31) / 32 * 32);  // weight operand columns, after the accumulator
  if (a.a_tmem && warp >= 2 && warp < 6) {
    // row e = 32q + lane of the weight slice: nkb*64 bf16 = nkb*32 packed columns, 32 columns per tcgen05.st
    const uint32_t *src = reinterpret_cast<const uint32_t *>(a.wimg + (size_t)rank * a.wimg_bytes) + (size_t)e * (nkb * 32);
    for (int cb = 0; cb < nkb; ++cb) {
      uint32_t v[32];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 x = __ldg(reinterpret_cast<const uint4 *>(src + cb * 32) + j);
        v[4 * j] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
      }
      tc_st32(tmem_a + ((uint32_t)(32 * q) << 16) + (uint32_t)(cb * 32), v);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_sync_all();

  const uint32_t idesc = make_idesc_bf16(128, N, false, false);
  // operand descriptors of k-block 0; the address field counts 16-byte units, so k-block kb adds a constant
  const uint64_t dA0 = make_desc(s_w, 0, 1024), dB0 = make_desc(s_h, 0, 1024);
  __shared__ long long tk[2][6];  // debug cycle counters ([0]: MMA warp, [1]: epilogue thread); shared memory, not registers
  if (tid < 12) (&tk[0][0])[tid] = 0;
  const bool dbg = DBG && rank == 0 && cluster_id_x() == 0 && (tid == 64 || warp == 1);

  // per-thread constants of the elementwise phase: item `it` is column cb + 4*it of unit uo
  const int cb = cg * CPW + eg;
  const bool g_lo = (eg & 1) != 0, g_hi = (eg & 2) != 0;
  uint32_t valid = 0;
#pragma unroll
  for (int it = 0; it < NCH; ++it)
    if (b0 + cb + 4 * it < B) valid |= 1u << it;
  const bool tile_full = b0 + N <= B;
  // staging-plane address of (column, unit): columns 4 apart flip bit 2 of the 128B-swizzle phase
  const uint8_t *xw_p = xw_gen + cb * 256 + eu * 8;                       // (column cb, unit eu): 4 gates, 8 bytes
  uint8_t *act_p = act_gen + cb * 256 + eu * 8;                           // (column cb, unit eu): 4 gates, 8 bytes
  const float *init_p = a.state_in + (size_t)(b0 + cb) * S + uo;          // fed (c, h): the eoc reset source
  const int H4 = 4 * H;
  const size_t BH = (size_t)B * H;
  __nv_bfloat16 *o_p = a.out_bf + (size_t)(b0 + cb) * H + uo;             // out[t]
  __nv_bfloat16 *h_p = a.hprev_bf + ((size_t)B + b0 + cb) * H + uo;       // hprev[t+1]
  float *c_p = a.cprev + ((size_t)B + b0 + cb) * H + uo;                  // cprev[t+1]
  const unsigned char *eoc_p = a.eoc + b0 + cb;                            // eoc[t]

  for (int t = 0; t < T; ++t) {
    uint32_t lapt = 0;
    if (dbg) lapt = (uint32_t)clock();
    if (warp == 0) {
      if (elect_one()) {
        if (a.mcast) {
          // every rank expects the whole tile and fetches one half k-block of it for everybody
          fence_proxy_async();
          for (int kb = 0; kb < nkb; ++kb) mbar_expect_tx(bar_h + 8 * kb, N * 128);
          const int kb = rank >> 1, half = rank & 1;
          tma_load_3d_multicast(s_h + (uint32_t)kb * (N * 128) + (uint32_t)half * (N / 2) * 128, &tm_h, bar_h + 8 * kb, kb * 64,
                                b0 + half * (N / 2), t, (uint16_t)((1u << a.CS) - 1u));
        }
        // last step's activation store has left its staging tile (the epilogue writes it again only after it has
        // seen this step's planes, i.e. after bar_xw)
        tma_store_wait_read();
        mbar_expect_tx(bar_xw, N * 256);
        tma_load_3d(s_xw, &tm_xw, bar_xw, 4 * u0, b0, t);  // the 4 gates of my 32 units for the tile's rows
        // gate activations of this step (bf16, [unit][gate] per row) go to their own array for BPTT
        mbar_wait(bar_out, t & 1);
        tma_store_3d(&tm_act, s_act, 4 * u0, b0, t);
      }
      __syncwarp();
    } else if (warp == 1) {
      if (!a.mcast) mbar_wait(bar_h, t & 1);
      if (lane == 0) TCR_LAP(tk[0][0]);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < 8; ++kb) {
          if (kb < nkb) {
            if (a.mcast) {
              mbar_wait(bar_h + 8 * kb, t & 1);
              tc_fence_after();
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (a.a_tmem)
                tc_mma_bf16_ts(tmem_acc, tmem_a + (uint32_t)(kb * 32 + k * 8), dB0 + (uint64_t)(kb * (N * 8) + 2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
              else
                tc_mma_bf16(tmem_acc, dA0 + (uint64_t)(kb * 1024 + 2 * k), dB0 + (uint64_t)(kb * (N * 8) + 2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
            }
          }
        }
        tc_commit(bar_acc);
      }
      __syncwarp();
      if (lane == 0) TCR_LAP(tk[0][2]);
    } else {
      // h_{t-1} tile of my cluster's rows (all H units, written by all ranks last step): L2 -> swizzled smem
      if (!a.mcast) {
        const int cpr = H >> 3;  // 16-byte chunks per row
        const int et = tid - 64;
        constexpr int GU = NCH >= 2 ? 2 : 1;  // loads in flight per thread (the tile fetch is L2-bandwidth bound)
        constexpr int ET = 32 * TCR_EPI_WARPS;
        const int total = N * cpr;
        const __nv_bfloat16 *hsrc = a.hprev_bf + (size_t)t * B * H;
        for (int i0 = et; i0 < total; i0 += ET * GU) {
          uint4 v[GU];
#pragma unroll
          for (int j = 0; j < GU; ++j) {
            const int i = i0 + j * ET;
            v[j] = make_uint4(0, 0, 0, 0);
            if (i < total) {
              const int n = i / cpr, ch = i - n * cpr;
              if (b0 + n < B) v[j] = __ldcg(reinterpret_cast<const uint4 *>(hsrc + (size_t)(b0 + n) * H) + ch);
            }
          }
#pragma unroll
          for (int j = 0; j < GU; ++j) {
            const int i = i0 + j * ET;
            if (i < total) {
              const int n = i / cpr, ch = i - n * cpr;
              const uint32_t off = (uint32_t)((ch >> 3) * (N * 128) + n * 128 + (((ch & 7) ^ (n & 7)) << 4));
              *reinterpret_cast<uint4 *>(gbase + (s_h - base) + off) = v[j];
            }
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(bar_h);
      }
      // everything that does not depend on the recurrence is fetched before waiting for the MMA
      uint32_t eoc_bits = 0;
      float dsc[NCH];
#pragma unroll
      for (int it = 0; it < NCH; ++it) {
        dsc[it] = 1.0f;
        if ((valid >> it) & 1u) {
          if (__ldg(eoc_p + 4 * it)) eoc_bits |= 1u << it;
          if (a.dropout) dsc[it] = drop_scale(a.mask, b0 + cb + 4 * it, T, t, H, uo, a.row0, a.seed, a.keep);
        }
      }
      const bool more = t + 1 < T;
      TCR_LAP(tk[1][0]);
      mbar_wait(bar_acc, t & 1);
      TCR_LAP(tk[1][1]);
      tc_fence_after();
      uint32_t r[CPW];
      tc_ldn<CPW>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(cg * CPW), r);
      mbar_wait(bar_xw, t & 1);
      TCR_LAP(tk[1][2]);
      constexpr bool DEFER = NCH <= 2;  // small tiles are latency-bound: publish h first, store the rest after arriving
      float sv[DEFER ? NCH : 1][3];
      auto items = [&](auto full_tag) {  // FULL: every column of the tile is a real batch row (no per-item bound checks)
      constexpr bool FULL = decltype(full_tag)::value;
#pragma unroll
      for (int it = 0; it < NCH; ++it) {
        // 4-lane exchange: lane eg sends gate row eg's value of column 4*it + (eg^k) in round k and so collects
        // all four gates of its own column 4*it + eg
        float v0, v1, v2, v3, p0, p1, p2, p3;
        perm4(g_lo, g_hi, __uint_as_float(r[4 * it + 0]), __uint_as_float(r[4 * it + 1]), __uint_as_float(r[4 * it + 2]),
              __uint_as_float(r[4 * it + 3]), v0, v1, v2, v3);
        v1 = __shfl_xor_sync(0xffffffffu, v1, 1);
        v2 = __shfl_xor_sync(0xffffffffu, v2, 2);
        v3 = __shfl_xor_sync(0xffffffffu, v3, 3);
        perm4(g_lo, g_hi, v0, v1, v2, v3, p0, p1, p2, p3);
        if (FULL || ((valid >> it) & 1u)) {
          const uint2 xv = *reinterpret_cast<const uint2 *>(xw_p + it * 1024);  // bf16 i, j, f, o of (column, unit)
          const float ig = sigmoid_approx(p0 + __uint_as_float(xv.x << 16));
          const float jg = tanh_approx(p1 + __uint_as_float(xv.x & 0xffff0000u));
          const float fg = sigmoid_approx(p2 + __uint_as_float(xv.y << 16) + 1.0f);
          const float og = sigmoid_approx(p3 + __uint_as_float(xv.y & 0xffff0000u));
          {  // staged for the TMA store.  i and f are positive: their sign bits carry the step's eoc flag and
             // "this output element was dropped", so BPTT needs no side loads
            const __nv_bfloat162 a01 = __floats2bfloat162_rn(ig, jg), a23 = __floats2bfloat162_rn(fg, og);
            uint32_t w0 = *reinterpret_cast<const uint32_t *>(&a01) & 0xffff7fffu;
            uint32_t w1 = *reinterpret_cast<const uint32_t *>(&a23) & 0xffff7fffu;
            if ((eoc_bits >> it) & 1u) w0 |= 0x8000u;
            if (dsc[it] == 0.0f) w1 |= 0x8000u;
            *reinterpret_cast<uint2 *>(act_p + it * 1024) = make_uint2(w0, w1);
          }
          const float c2 = fmaf(cst[it], fg, ig * jg);
          const float h2 = tanh_approx(c2) * og;
          float cn = c2, hn = h2;
          if ((eoc_bits >> it) & 1u) { cn = __ldg(init_p + (size_t)it * 4 * S); hn = __ldg(init_p + (size_t)it * 4 * S + H); }
          cst[it] = cn;
          if (more) h_p[it * H4] = __float2bfloat16_rn(hn);  // exchange medium first
          if constexpr (DEFER) {
            sv[it][0] = h2 * dsc[it]; sv[it][1] = cn; sv[it][2] = hn;
          } else {
            o_p[it * H4] = __float2bfloat16_rn(h2 * dsc[it]);
            if (more) {
              c_p[it * H4] = cn;
            } else {
              float *so = a.state_out + (size_t)(b0 + cb + 4 * it) * S + uo;
              so[0] = cn; so[H] = hn;
            }
          }
        }
      }
      };
      if (tile_full) items(std::true_type{}); else items(std::false_type{});
      TCR_LAP(tk[1][3]);
      fence_proxy_async_smem();
      mbar_arrive(bar_out);
      if (a.mcast) fence_proxy_async();  // my h stores -> visible to the other ranks' TMA loads
      TCR_LAP(tk[1][5]);
      if constexpr (DEFER) {
        __syncwarp();
        tc_fence_before();
        cluster_arrive();  // release: the h stores above are what the other ranks wait for
#pragma unroll
        for (int it = 0; it < NCH; ++it) {
          if ((valid >> it) & 1u) {
            o_p[it * H4] = __float2bfloat16_rn(sv[it][0]);
            if (more) {
              c_p[it * H4] = sv[it][1];
            } else {
              float *so = a.state_out + (size_t)(b0 + cb + 4 * it) * S + uo;
              so[0] = sv[it][1]; so[H] = sv[it][2];
            }
          }
        }
        o_p += BH; h_p += BH; c_p += BH; eoc_p += B;
        cluster_wait();
        tc_fence_after();
        if (tid == 64) TCR_LAP(tk[1][4]);
        continue;
      }
      o_p += BH; h_p += BH; c_p += BH; eoc_p += B;
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();  // arrive.release publishes the h stores to the other ranks
    tc_fence_after();
    if (tid == 64) TCR_LAP(tk[1][4]);
  }
  if (dbg && tid == 32)
    printf("[tc fwd N=%d] MMA thread cycles/step: barrier->h tile ready %lld, (unused) %lld, issue+commit %lld\n", N,
           tk[0][0] / T, tk[0][1] / T, tk[0][2] / T);
  if (dbg && tid == 64)
    printf("[tc fwd N=%d] epilogue cycles/step: prefetch %lld, wait acc %lld, tmem ld + xw wait %lld, items+stores %lld, fences %lld, cluster barrier %lld\n",
           N, tk[1][0] / T, tk[1][1] / T, tk[1][2] / T, tk[1][3] / T, tk[1][5] / T, tk[1][4] / T);
  if (warp == 0 && elect_one()) tma_store_wait_all();
  __syncthreads();
  if (warp == 1) tc_dealloc(tmem_acc, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// forward, two half tiles in flight ("ping-pong")
// ------------------------------------------------------------------------------------------------
// The kernel above runs its phases one after the other: h tile in, product, elementwise, publish, cluster
// barrier.  Here the cluster's N batch rows are two independent half tiles of N2 = N/2 rows, each with its own
// h tile, accumulator, staging planes and mbarriers.  While the 16 elementwise warps work on one half, the TMA
// loads, the product and the publication round trip of the other half proceed in the background.  There is no
// barrier.cluster in the loop: an extra warp waits for the CTA's elementwise pass of a half (mbarrier), makes
// its h stores visible (fence) and arrives on the `ready` mbarrier of every rank through the cluster's shared
// memory window; a rank's TMA producer issues its share of the next multicast h tile of that half once all
// ranks have arrived.  Requires the multicast path with the weights in tensor memory (N >= 32, CS = 2*nkb).
constexpr int TCR_PP_THREADS = TCR_THREADS + 32;  // + warp 18: publisher

__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}
// 16 bytes into the shared memory of a CTA of the cluster (cluster addresses); the store adds its bytes to the
// transaction count of the mbarrier at `bar` in the same CTA
__device__ __forceinline__ void st_async_v4(uint32_t dst, uint4 v, uint32_t bar) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(bar) : "memory");
}
__device__ __forceinline__ void st_async_v2(uint32_t dst, uint2 v, uint32_t bar) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];"
               ::"r"(dst), "r"(v.x), "r"(v.y), "r"(bar) : "memory");
}
// tcgen05.ld.16x256b: 16 accumulator rows (lanes base .. base+15 of the warp's quarter) x 8*X columns; thread t gets rows
// t/4 and t/4 + 8, columns 2*(t%4), 2*(t%4)+1 (+8 per repeat): registers {row a: c0, c1}, {row b: c0, c1}, then the next 8 columns
__device__ __forceinline__ void tc_ld_16x256b_x2(uint32_t taddr, uint32_t *r) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
}
__device__ __forceinline__ void tc_ld_16x256b_x1(uint32_t taddr, uint32_t *r) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// a partial-sum element group: from L2 (exchange through global memory) or from this CTA's receive buffer (push mode)
template <typename V>
__device__ __forceinline__ V ld_part(const V *p, bool local) { return local ? *p : __ldcg(p); }  // `local` is a compile-time constant at the call sites
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t rank) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar), "r"(rank) : "memory");
}

template <int NCH, bool DBG, int HT = 0>  // NCH items per thread and half tile; HT: rnn_size as a compile-time constant (0: run time)
__global__ void __launch_bounds__(TCR_PP_THREADS, 1)
lstm_fwd_pp_kernel(const __grid_constant__ CUtensorMap tm_h, const __grid_constant__ CUtensorMap tm_xw,
                   const __grid_constant__ CUtensorMap tm_hst, TcRnnArgs a) {
  constexpr int N2 = 16 * NCH, N = 2 * N2;
  constexpr int CPW = N2 / 4;  // columns per elementwise warp and half tile
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  // (with rnn_size known at compile time the per-item offsets of the elementwise stores become immediates)
  const int H = HT ? HT : a.H, B = a.B, T = a.T, nkb = HT ? HT / 64 : a.nkb, CS = HT ? HT / 32 : a.CS;
  const int t0 = a.t0, t1 = a.t1;  // this launch runs steps [t0, t1); t0 is even, so barrier parities follow t
  const uint32_t h_bytes = (uint32_t)nkb * (N2 * 128), xw_bytes = (uint32_t)N2 * 256u, hst_bytes = (uint32_t)N2 * 64u;
  const bool push = a.push != 0;
  const uint32_t s_h = base;                       // [2] x nkb x [N2 rows][128 B]; push mode: two such sets, used in turn by step parity
  const uint32_t s_xw = s_h + (push ? 4 : 2) * h_bytes;  // [2] x [N2][32 units][4 gates] bf16: hoisted x*Wx+b
  const uint32_t s_hst = s_xw + 2 * xw_bytes;      // [2] x [N2][32 units] bf16: my units of h_t on their way out (TMA store); push mode: two sets by step parity
  const uint32_t s_bar = s_hst + (push ? 4 : 2) * hst_bytes;
  // mbarriers, index s = half tile
  const uint32_t bar_h = s_bar, bar_xw = s_bar + 16, bar_acc = s_bar + 32, bar_epi = s_bar + 48, bar_ready = s_bar + 64;
  const uint32_t tmem_holder = s_bar + 80;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int b0 = cluster_id_x() * N;
  const int u0 = rank * 32;
  const int S = 2 * H;

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_h + 8 * s, 1);
      mbar_init(bar_xw + 8 * s, 1);
      mbar_init(bar_acc + 8 * s, 1);
      mbar_init(bar_epi + 8 * s, 32 * TCR_EPI_WARPS);
      mbar_init(bar_ready + 8 * s, a.sig_async ? 1 : CS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);

  const int q = warp & 3, cg = (warp - 2) >> 2;
  const int e = 32 * q + lane;
  const int eu = e >> 2, eg = e & 3;
  const int uo = u0 + eu;
  // The accumulator rows of a lane quarter are ordered [gate][8 units] (the weight image is packed that way), so that
  // two tcgen05.ld.16x256b hand a thread the FOUR GATES of its unit for a pair of columns directly -- no exchange
  // between lanes (3 shuffles + 16 selects per item: it was the largest block of a pass that is bound by issued
  // instructions).  Thread (unit eu, column pair eg) owns columns cb + dco(it), cb = its first one.  Half tiles of 16
  // rows have only 4 columns per warp: there the two warps of a column-group pair read the same 8 columns and take
  // one column of each pair.
  constexpr bool QUAD = true;
  const int cb = NCH == 1 ? 8 * (cg >> 1) + 2 * eg + (cg & 1) : cg * CPW + 2 * eg;
  auto dco = [](int it) { return NCH == 1 ? 0 : 8 * (it >> 1) + (it & 1); };  // column of item `it` minus cb
  float cst[2][NCH];
  if (warp >= 2 && warp < 2 + TCR_EPI_WARPS) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
#pragma unroll
      for (int it = 0; it < NCH; ++it) {
        const int b = b0 + s * N2 + cb + dco(it);
        cst[s][it] = 0.f;
        if (b < B) {
          if (t0 == 0) {
            const float ci = a.state_in[(size_t)b * S + uo];
            const float hi = a.state_in[(size_t)b * S + H + uo];
            cst[s][it] = ci;
            a.cprev[(size_t)b * H + uo] = ci;
            a.hprev_bf[(size_t)b * H + uo] = __float2bfloat16_rn(hi);
          } else {  // a later range of the same unroll: the previous launch left c and h of step t0 in the records
            cst[s][it] = a.cprev[((size_t)t0 * B + b) * H + uo];
          }
        }
      }
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  const uint32_t tmem_a = tmem_acc + (uint32_t)((N + 31) / 32 * 32);  // weight operand columns, after the accumulators
  if (warp >= 2 && warp < 6) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(a.wimg + (size_t)rank * a.wimg_bytes) + (size_t)e * (nkb * 32);
    for (int kb = 0; kb < nkb; ++kb) {
      uint32_t v[32];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 x = __ldg(reinterpret_cast<const uint4 *>(src + kb * 32) + j);
        v[4 * j] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
      }
      tc_st32(tmem_a + ((uint32_t)(32 * q) << 16) + (uint32_t)(kb * 32), v);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_sync_all();  // everybody's barriers are initialised and the step-0 h rows are in global memory

  __shared__ long long tk[24];
  if (tid < 24) tk[tid] = 0;
  __syncthreads();
  const bool dbg0 = DBG && rank == 0 && cluster_id_x() == 0;
  const bool dbg = dbg0 && tid == 64;

  if (warp == 0) {
    // ---------------- TMA producer ----------------
    if (elect_one()) {
      const bool dbg = dbg0;
      uint32_t lapt = 0;
      if (dbg) lapt = (uint32_t)clock();
      const int kb_mine = rank >> 1, half = rank & 1;
      const uint16_t all = (uint16_t)((1u << CS) - 1u);
      auto load_h = [&](int s, int t) {
        // every rank expects the whole half tile and fetches 1/CS of it (half a k-block) for everybody
        mbar_expect_tx(bar_h + 8 * s, h_bytes);
        tma_load_3d_multicast(s_h + s * h_bytes + (uint32_t)kb_mine * (N2 * 128) + (uint32_t)half * (N2 / 2) * 128, &tm_h,
                              bar_h + 8 * s, kb_mine * 64, b0 + s * N2 + half * (N2 / 2), t, all);
      };
      auto load_xw = [&](int s, int t) {
        mbar_expect_tx(bar_xw + 8 * s, xw_bytes);
        tma_load_3d(s_xw + s * xw_bytes, &tm_xw, bar_xw + 8 * s, 4 * u0, b0 + s * N2, t);
      };
      auto retire = [&](int s, int t) {  // the elementwise pass of (s, t) is over: refill its planes
        mbar_wait(bar_epi + 8 * s, t & 1);
        if (t + 1 < t1) load_xw(s, t + 1);
      };
      load_xw(0, t0);
      load_xw(1, t0);
      // sig_async: the ranks announce their h stores with 4-byte st.async stores whose bytes complete bar_ready's
      // phase; this thread states the expected bytes of a phase once the previous one is over
      const bool sig = a.sig_async && !push;
      if (sig && t0 + 1 < t1) { mbar_expect_tx(bar_ready + 0, (uint32_t)CS * 4u); mbar_expect_tx(bar_ready + 8, (uint32_t)CS * 4u); }
      for (int t = t0; t < t1; ++t) {
        // push mode: after the first step the tile is SENT by the ranks' publishers; this thread only states how many
        // bytes the step's barrier phase expects (the previous phase is over: retire() of that half has been passed)
        const bool pushed = push && t > t0;
        if (t > t0 && !push) {
          mbar_wait_cluster(bar_ready + 0, (t - 1) & 1);
          if (sig && t + 1 < t1) mbar_expect_tx(bar_ready + 0, (uint32_t)CS * 4u);
        }
        TCR_LAP(tk[8]);
        if (pushed) mbar_expect_tx(bar_h + 0, h_bytes); else load_h(0, t);
        TCR_LAP(tk[9]);
        if (t > t0) retire(1, t - 1);
        TCR_LAP(tk[10]);
        if (t > t0 && !push) {
          mbar_wait_cluster(bar_ready + 8, (t - 1) & 1);
          if (sig && t + 1 < t1) mbar_expect_tx(bar_ready + 8, (uint32_t)CS * 4u);
        }
        TCR_LAP(tk[11]);
        if (pushed) mbar_expect_tx(bar_h + 8, h_bytes); else load_h(1, t);
        TCR_LAP(tk[12]);
        retire(0, t);
        TCR_LAP(tk[13]);
      }
      if (dbg)
        printf("[tc fwd pp] producer cycles/step: wait ready0 %lld, issue h0 %lld, retire1 %lld, wait ready1 %lld, issue h1 %lld, retire0 %lld\n",
               tk[8] / T, tk[9] / T, tk[10] / T, tk[11] / T, tk[12] / T, tk[13] / T);
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    if (elect_one()) {
      const bool dbg = dbg0;
      uint32_t lapt = 0;
      if (dbg) lapt = (uint32_t)clock();
      const uint32_t idesc = make_idesc_bf16(128, N2, false, false);
      const uint64_t dB0 = make_desc(s_h, 0, 1024);
      for (int t = t0; t < t1; ++t) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (push) {
            mbar_wait_cluster(bar_h + 8 * s, t & 1);
            fence_proxy_async_smem();  // the tile was written by st.async stores: order them before the tensor core's reads
          } else {
            mbar_wait(bar_h + 8 * s, t & 1);
          }
          TCR_LAP(tk[16 + 2 * s]);
          tc_fence_after();
          const uint32_t hset = push ? (uint32_t)(t & 1) * 2u * h_bytes : 0u;
#pragma unroll
          for (int kb = 0; kb < 8; ++kb) {
            if (kb < nkb) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc_mma_bf16_ts(tmem_acc + (uint32_t)(s * N2), tmem_a + (uint32_t)(kb * 32 + k * 8),
                               dB0 + (uint64_t)((hset + s * h_bytes) >> 4) + (uint64_t)(kb * (N2 * 8) + 2 * k), idesc,
                               (kb > 0 || k > 0) ? 1u : 0u);
            }
          }
          tc_commit(bar_acc + 8 * s);
          TCR_LAP(tk[17 + 2 * s]);
        }
      }
      if (dbg)
        printf("[tc fwd pp] MMA thread cycles/step: wait h0 %lld, issue %lld, wait h1 %lld, issue %lld\n", tk[16] / T, tk[17] / T,
               tk[18] / T, tk[19] / T);
    }
    __syncwarp();
  } else if (warp == 2 + TCR_EPI_WARPS) {
    // ---------------- publisher ----------------
    // h_t of a half tile leaves through the async proxy (TMA store from the staging tile), so no generic-to-async
    // proxy fence sits on the critical path; once the store is complete lane r arrives at rank r (the CS remote
    // round trips overlap)
    {
      const bool dbg = dbg0 && lane == 0;
      uint32_t lapt = 0;
      if (dbg) lapt = (uint32_t)clock();
      for (int t = t0; t < t1 && t + 1 < T; ++t) {  // (the last step of a range still stores h for the next launch)
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          mbar_wait(bar_epi + 8 * s, t & 1);  // the staging tile of (s, t) is complete (writers fenced it for the async proxy)
          TCR_LAP(tk[20]);
          if (push) {
            // Small tiles: the round trip store -> completion -> remote arrivals -> multicast load is most of a step.
            // Instead every rank's publisher sends its [N2 rows][32 units] of h_{t+1} straight into all ranks' operand
            // tiles of step t+1 (the set not being read now), 16 bytes per st.async, at the swizzled position the UMMA
            // K-major layout gives them; each store adds its bytes to the receiver's bar_h phase.  The record of h in
            // global memory (hprev, read by the weight-gradient GEMM) still leaves by TMA, off the critical path.
            const uint32_t stg_s = s_hst + ((uint32_t)(t & 1) * 2u + s) * hst_bytes;  // staging tiles alternate by step parity
            if (t + 1 < t1) {
              const uint32_t dst0 = s_h + (uint32_t)((t + 1) & 1) * 2u * h_bytes + s * h_bytes + (uint32_t)(rank >> 1) * (N2 * 128);
              const uint8_t *stg = gbase + (stg_s - base);
              uint4 v[(N2 * 4 + 31) / 32];
              uint32_t off[(N2 * 4 + 31) / 32];
#pragma unroll
              for (int i = 0; i < (N2 * 4 + 31) / 32; ++i) {
                const int c = lane + 32 * i;  // chunk c: row c / 4, 16-byte piece c % 4 of the rank's 64 bytes of that row
                const int n = c >> 2, j = c & 3;
                v[i] = *reinterpret_cast<const uint4 *>(stg + n * 64 + j * 16);
                off[i] = (uint32_t)(n * 128) + ((uint32_t)((4 * (rank & 1) + j) ^ (n & 7)) << 4);
              }
              for (int rr = 0; rr < CS; ++rr) {
                const uint32_t r = (uint32_t)((rank + rr) % CS);
                const uint32_t rd = mapa_u32(dst0, r), rb = mapa_u32(bar_h + 8 * s, r);
#pragma unroll
                for (int i = 0; i < (N2 * 4 + 31) / 32; ++i)
                  if (lane + 32 * i < N2 * 4) st_async_v4(rd + off[i], v[i], rb);
              }
            }
            TCR_LAP(tk[21]);
            // the record in global memory, after the sends.  This staging tile is rewritten two steps from now, and the
            // wait below (one step from now, same half) comes before that
            if (elect_one()) {
              tma_store_wait_read();
              tma_store_3d(&tm_hst, stg_s, u0, b0 + s * N2, t + 1);
            }
            __syncwarp();
            TCR_LAP(tk[22]);
            continue;
          }
          if (elect_one()) {
            tma_store_3d(&tm_hst, s_hst + s * hst_bytes, u0, b0 + s * N2, t + 1);
            tma_store_wait_all();             // h_t[my units] is in global memory before anybody is told
          }
          __syncwarp();
          TCR_LAP(tk[21]);
          if (lane < CS) {
            if (a.sig_async) {
              // the store above is complete (in L2) before this is issued; the announcement itself is a plain 4-byte
              // st.async into a scratch word of rank `lane` that counts on its bar_ready -- no cluster-scope release
              // (which costs a gpu-scope fence, ~1.3 K cycles per half step here)
              if (t + 1 < t1) {
                const uint32_t rs = mapa_u32(s_bar + 96u + (uint32_t)(s * 16 + rank) * 4u, (uint32_t)lane);
                const uint32_t rb = mapa_u32(bar_ready + 8 * s, (uint32_t)lane);
                asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(rs), "r"(1u), "r"(rb) : "memory");
              }
            } else {
              mbar_arrive_remote(bar_ready + 8 * s, (uint32_t)lane);
            }
          }
          __syncwarp();
          TCR_LAP(tk[22]);
        }
      }
      if (push && elect_one()) tma_store_wait_all();
      if (dbg)
        printf("[tc fwd pp] publisher cycles/half step: wait elementwise %lld, h store %lld, remote arrives %lld\n", tk[20] / (2 * T),
               tk[21] / (2 * T), tk[22] / (2 * T));
    }
    __syncwarp();
  } else {
    // ---------------- elementwise warps ----------------
    const bool g_lo = (eg & 1) != 0, g_hi = (eg & 2) != 0;
    uint32_t valid[2] = {0, 0};
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int it = 0; it < NCH; ++it)
        if (b0 + s * N2 + cb + dco(it) < B) valid[s] |= 1u << it;
    uint8_t *xw_gen = gbase + (s_xw - base);
    uint8_t *hst_gen = gbase + (s_hst - base);
    const uint8_t *xw_p = xw_gen + cb * 256 + eu * 8;  // (column cb, unit eu): 4 gates, 8 bytes
    uint8_t *hst_p = hst_gen + cb * 64 + eu * 2;
    const float *init_p = a.state_in + (size_t)(b0 + cb) * S + uo;
    const int H4 = 4 * H;
    const size_t BH = (size_t)B * H, half_off = (size_t)N2 * H;
    __nv_bfloat16 *o_p = a.out_bf + ((size_t)t0 * B + b0 + cb) * H + uo;        // out[t], half 0
    uint2 *act_g = reinterpret_cast<uint2 *>(a.acts_bf) + ((size_t)t0 * B + b0 + cb) * H + uo;  // activation record of step t
    float *c_p = a.cprev + ((size_t)(t0 + 1) * B + b0 + cb) * H + uo;           // cprev[t+1]
    const unsigned char *eoc_p = a.eoc + (size_t)t0 * B + b0 + cb;               // eoc[t]
    const uint32_t drop_unit = skm_drop_unit(a.seed, (uint32_t)uo);
    const float inv_keep = 1.0f / a.keep;
    uint32_t eoc_next = 0;  // flags of the half step about to be processed
#pragma unroll
    for (int it = 0; it < NCH; ++it)
      if (((valid[0] >> it) & 1u) && __ldg(eoc_p + dco(it))) eoc_next |= 1u << it;
    uint32_t lapt = 0;
    if (dbg) lapt = (uint32_t)clock();
    for (int t = t0; t < t1; ++t) {
      const bool more = t + 1 < T;
      const uint32_t hst_par = push ? (uint32_t)(t & 1) * 2u * hst_bytes : 0u;
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint32_t eoc_bits = eoc_next;
        // eoc flags of the following half step: issued now, looked at after this pass
        unsigned char en[NCH];
        {
          const int s2 = s ^ 1;
          const unsigned char *ep = eoc_p + s2 * N2 + (s == 1 ? B : 0);
          const bool in_range = s == 0 || more;
#pragma unroll
          for (int it = 0; it < NCH; ++it) en[it] = (in_range && ((valid[s2] >> it) & 1u)) ? __ldg(ep + dco(it)) : (unsigned char)0;
        }
        float dsc[NCH];
#pragma unroll
        for (int it = 0; it < NCH; ++it) dsc[it] = 1.0f;
        if (a.dropout) {
          if (a.mask != nullptr) {  // injected mask (parity tests)
#pragma unroll
            for (int it = 0; it < NCH; ++it)
              if ((valid[s] >> it) & 1u)
                dsc[it] = a.mask[((int64_t)(b0 + s * N2 + cb + dco(it)) * T + t) * H + uo] * inv_keep;
          } else {
            // rows past the batch get a scale too (never stored): no per-item predicates around the hash
            const uint32_t pre = skm_drop_prefix(a.seed, drop_unit, (uint32_t)t);  // (unit, t) part of the mask hash
            const uint32_t row = (uint32_t)(a.row0 + b0 + s * N2 + cb);
            const float keep = a.keep;
#pragma unroll
            for (int it = 0; it < NCH; ++it)
              dsc[it] = skm_u01(skm_drop_finish(pre, row + (uint32_t)dco(it))) < keep ? inv_keep : 0.0f;
          }
        }
        TCR_LAP(tk[0]);
        mbar_wait(bar_acc + 8 * s, t & 1);
        TCR_LAP(tk[1]);
        tc_fence_after();
        uint32_t sv_a0[NCH], sv_a1[NCH];
        float sv_o[NCH], sv_c[NCH], sv_h[NCH];
        uint32_t r[CPW < 8 ? 8 : CPW];
        if constexpr (QUAD) {
          // rows 32q .. 32q+15 hold gates i, j of the quarter's 8 units, rows 32q+16 .. 32q+31 gates f, o
          const uint32_t ta = tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(s * N2 + (NCH == 1 ? 8 * (cg >> 1) : cg * CPW));
          if constexpr (CPW == 16) {
            tc_ld_16x256b_x2(ta, r);
            tc_ld_16x256b_x2(ta + (16u << 16), r + 8);
          } else {
            tc_ld_16x256b_x1(ta, r);
            tc_ld_16x256b_x1(ta + (16u << 16), r + 4);
          }
          tc_wait_ld();
        } else {
          tc_ldn<CPW>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(s * N2 + cg * CPW), r);
        }
        mbar_wait(bar_xw + 8 * s, t & 1);
        TCR_LAP(tk[2]);
        auto items = [&](auto full_tag) {  // FULL: every column of the half tile is a real batch row
        constexpr bool FULL = decltype(full_tag)::value;
#pragma unroll
        for (int it = 0; it < NCH; ++it) {
          float p0, p1, p2, p3;
          if constexpr (QUAD && NCH == 1) {
            const bool odd = (cg & 1) != 0;  // which column of the pair is mine
            p0 = __uint_as_float(odd ? r[1] : r[0]); p1 = __uint_as_float(odd ? r[3] : r[2]);
            p2 = __uint_as_float(odd ? r[5] : r[4]); p3 = __uint_as_float(odd ? r[7] : r[6]);
          } else if constexpr (QUAD) {
            // per load: {gate a: c0, c1}, {gate b: c0, c1} (then the same for the columns 8 further on)
            constexpr int HALF = CPW / 2;  // registers of one load
            const int j = (it >> 1) * 4 + (it & 1);
            p0 = __uint_as_float(r[j]); p1 = __uint_as_float(r[j + 2]);
            p2 = __uint_as_float(r[HALF + j]); p3 = __uint_as_float(r[HALF + j + 2]);
          } else {
            float v0, v1, v2, v3;
            perm4(g_lo, g_hi, __uint_as_float(r[4 * it + 0]), __uint_as_float(r[4 * it + 1]), __uint_as_float(r[4 * it + 2]),
                  __uint_as_float(r[4 * it + 3]), v0, v1, v2, v3);
            v1 = __shfl_xor_sync(0xffffffffu, v1, 1);
            v2 = __shfl_xor_sync(0xffffffffu, v2, 2);
            v3 = __shfl_xor_sync(0xffffffffu, v3, 3);
            perm4(g_lo, g_hi, v0, v1, v2, v3, p0, p1, p2, p3);
          }
          if (FULL || ((valid[s] >> it) & 1u)) {
            const uint2 xv = *reinterpret_cast<const uint2 *>(xw_p + s * xw_bytes + dco(it) * 256);  // bf16 i, j, f, o
            const float ig = sigmoid_approx(p0 + __uint_as_float(xv.x << 16));
            const float jg = tanh_approx(p1 + __uint_as_float(xv.x & 0xffff0000u));
            const float fg = sigmoid_approx(p2 + __uint_as_float(xv.y << 16) + 1.0f);
            const float og = sigmoid_approx(p3 + __uint_as_float(xv.y & 0xffff0000u));
            {
              const __nv_bfloat162 a01 = __floats2bfloat162_rn(ig, jg), a23 = __floats2bfloat162_rn(fg, og);
              uint32_t w0 = *reinterpret_cast<const uint32_t *>(&a01) & 0xffff7fffu;
              uint32_t w1 = *reinterpret_cast<const uint32_t *>(&a23) & 0xffff7fffu;
              if ((eoc_bits >> it) & 1u) w0 |= 0x8000u;
              if (dsc[it] == 0.0f) w1 |= 0x8000u;
              sv_a0[it] = w0; sv_a1[it] = w1;
            }
            const float c2 = fmaf(cst[s][it], fg, ig * jg);
            const float h2 = tanh_approx(c2) * og;
            float cn = c2, hn = h2;
            if ((eoc_bits >> it) & 1u) {
              const float *ip = init_p + (size_t)(s * N2 + dco(it)) * S;
              cn = __ldg(ip); hn = __ldg(ip + H);
            }
            cst[s][it] = cn;
            *reinterpret_cast<__nv_bfloat16 *>(hst_p + hst_par + s * hst_bytes + dco(it) * 64) = __float2bfloat16_rn(hn);
            sv_o[it] = h2 * dsc[it]; sv_c[it] = cn; sv_h[it] = hn;
          }
        }
        };
        // the records of the step (activations, output, c) are held back until the pass has been announced: the arrive
        // below is a release, which otherwise waits for these twelve global stores per thread to drain
        const bool full_half = b0 + (s + 1) * N2 <= B;
        if (full_half) items(std::true_type{}); else items(std::false_type{});
        fence_proxy_async_smem();  // h staging tile -> TMA store
        tc_fence_before();         // accumulator reads -> next product of this half
        mbar_arrive(bar_epi + 8 * s);
#pragma unroll
        for (int it = 0; it < NCH; ++it) {
          if (full_half || ((valid[s] >> it) & 1u)) {
            const size_t off = s * half_off + (size_t)dco(it) * H;
            act_g[off] = make_uint2(sv_a0[it], sv_a1[it]);
            o_p[off] = __float2bfloat16_rn(sv_o[it]);
            if (more) {
              c_p[off] = sv_c[it];
            } else {
              float *so = a.state_out + (size_t)(b0 + s * N2 + cb + dco(it)) * S + uo;
              so[0] = sv_c[it]; so[H] = sv_h[it];
            }
          }
        }
        eoc_next = 0;
#pragma unroll
        for (int it = 0; it < NCH; ++it)
          if (en[it]) eoc_next |= 1u << it;
        TCR_LAP(tk[3]);
      }
      o_p += BH; act_g += BH; c_p += BH; eoc_p += B;
    }
    if (dbg)
      printf("[tc fwd pp N=2x%d] elementwise thread cycles per half step: prefetch %lld, wait acc %lld, tmem ld + xw wait %lld, items+stores %lld\n",
             N2, tk[0] / (2 * T), tk[1] / (2 * T), tk[2] / (2 * T), tk[3] / (2 * T));
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // nobody leaves while a peer may still multicast into it or arrive on its barriers
  if (warp == 1) tc_dealloc(tmem_acc, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
// Elementwise items: thread (unit u = e/4, sub = e%4, column group cg) owns the NCH columns
// c = cg*N/4 + 4*it + sub of unit u (the 4 lanes of a unit touch consecutive rows of the 128B-swizzled
// staging tiles, i.e. distinct banks).
template <int NCH, bool DBG, int HT = 0>  // HT: rnn_size as a compile-time constant (0: run time)
__global__ void __launch_bounds__(TCR_THREADS, 1)
lstm_bwd_tc_kernel(const __grid_constant__ CUtensorMap tm_acts, const __grid_constant__ CUtensorMap tm_do,
                   const __grid_constant__ CUtensorMap tm_cp, const __grid_constant__ CUtensorMap tm_dp, TcRnnArgs a) {
  constexpr int N = 16 * NCH;
  constexpr int CPW = N / 4;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const int H = HT ? HT : a.H, B = a.B, T = a.T, CS = HT ? HT / 32 : a.CS;
  const int nmt = (H + 127) / 128;                 // accumulator m-tiles (all hidden units)
  const int HP = nmt * 128;
  const uint32_t s_w = base;                       // 2 k-blocks x [HP rows][128 B]
  const uint32_t s_b = s_w + (a.a_tmem ? 0u : 2u * HP * 128u);  // 2 k-blocks x [N rows][128 B]: own d(pre) slice, bf16
  const uint32_t s_ac = s_b + 2u * (N * 128);      // [N][32 units][4 gates] bf16: gate activations saved by the forward
  const uint32_t s_do = s_ac + (uint32_t)N * 256u; // [N][32] fp32, 128B-swizzled: d(out) of the step
  const uint32_t s_cp = s_do + (uint32_t)N * 128u; // [N][32] fp32, 128B-swizzled: c fed to the step
  const uint32_t s_bar = s_cp + (uint32_t)N * 128u;
  const uint32_t bar_acts = s_bar, bar_ready = s_bar + 8, tmem_holder = s_bar + 24;
  const uint32_t bar_acc = s_bar + 32;  // [4]: one per accumulator m-tile, so the read-out starts while later tiles multiply
  // push mode (small tiles): the partial sums go straight into the destination rank's receive buffer (st.async, bytes
  // counted on its bar_rx) instead of through L2 with a barrier.cluster per step
  const bool push = a.push != 0;
  const uint32_t bar_rx = s_bar + 64;
  const uint32_t s_rx = s_bar + 256;    // [parity][src rank][column group][32 units x CPW columns] bf16
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));
  const uint8_t *ac_gen = gbase + (s_ac - base);
  const uint8_t *do_gen = gbase + (s_do - base);
  uint8_t *b_gen = gbase + (s_b - base);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int cid = cluster_id_x();
  const int b0 = cid * N;
  const int u0 = rank * 32;

  {
    const uint4 *src = reinterpret_cast<const uint4 *>(a.wimg + (size_t)rank * a.wimg_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(gbase + (s_w - base));
    if (!a.a_tmem)
      for (int i = tid; i < a.wimg_bytes / 16; i += TCR_THREADS) dst[i] = __ldg(src + i);
    uint4 *bz = reinterpret_cast<uint4 *>(b_gen);
    for (int i = tid; i < 2 * N * 128 / 16; i += TCR_THREADS) bz[i] = make_uint4(0, 0, 0, 0);
  }
  if (tid == 0) {
    mbar_init(bar_acts, 1);
    mbar_init(bar_rx, 1);
    mbar_init(bar_ready, 32 * TCR_EPI_WARPS);
    for (int mt = 0; mt < 4; ++mt) mbar_init(bar_acc + 8 * mt, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  // small tiles leave room in tensor memory for the weight operand (64 columns per m-tile): an MMA then reads its A
  // tile at the 9-cycle rate instead of 39 from shared memory
  const uint32_t tmem_a = tmem_acc + (uint32_t)((nmt * N + 31) / 32 * 32);
  if (a.a_tmem && warp >= 2 && warp < 6) {
    const int qq = warp & 3, ee = 32 * qq + lane;
    for (int mt = 0; mt < nmt; ++mt) {
      const uint32_t *src = reinterpret_cast<const uint32_t *>(a.wimg + (size_t)rank * a.wimg_bytes) + ((size_t)mt * 128 + ee) * 64;
#pragma unroll
      for (int hb = 0; hb < 2; ++hb) {
        uint32_t v[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint4 x = __ldg(reinterpret_cast<const uint4 *>(src + hb * 32) + j);
          v[4 * j] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
        }
        tc_st32(tmem_a + ((uint32_t)(32 * qq) << 16) + (uint32_t)(mt * 64 + hb * 32), v);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  cluster_sync_all();

  const int q = warp & 3, cg = (warp - 2) >> 2;
  const int e = 32 * q + lane;
  const int eu = e >> 2, esub = e & 3;
  const int uo = u0 + eu;
  const int c0 = cg * CPW + esub;
  float dhc[NCH], dcc[NCH];
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};  // bias gradient of my unit over my columns and all steps
  uint64_t dBk[2];
  dBk[0] = make_desc(s_b, 0, 1024);
  dBk[1] = make_desc(s_b + (uint32_t)(N * 128), 0, 1024);
#pragma unroll
  for (int it = 0; it < NCH; ++it) { dhc[it] = 0.f; dcc[it] = 0.f; }
  const uint32_t idesc = make_idesc_bf16(128, N, false, false);
  __nv_bfloat16 *part_base = a.partial + (size_t)cid * 2 * CS * H * N;
  // partial d(h) exchange layout, per parity: [dst rank][src rank][column group][chunk j][unit 32][CK columns]:
  // the writer's lanes (= units of the destination rank) and the reader's lanes both touch whole 128-byte lines
  constexpr int CK = CPW >= 8 ? 8 : CPW;
  constexpr int NJ = CPW / CK;
  const size_t part_blk = (size_t)NJ * 32 * CK;  // one (dst, src, cg) block = 32 units x CPW columns
  const uint32_t rx_bytes = (uint32_t)(CS * 4 * part_blk * 2);  // what a rank receives per step
  const __nv_bfloat16 *rx_gen = reinterpret_cast<const __nv_bfloat16 *>(gbase + (s_rx - base));
  if (push && tid == 0 && T > 1) mbar_expect_tx(bar_rx, rx_bytes);  // step 0's exchange (later ones are armed after each wait)

  __shared__ long long tk[8];  // debug cycle counters (shared memory, not registers)
  if (tid < 8) tk[tid] = 0;
  const bool dbg = DBG && rank == 0 && cid == 0 && tid == 64;
  // per-thread constants of the elementwise phase: item `it` is column c0 + 4*it of unit uo
  uint32_t valid = 0;
#pragma unroll
  for (int it = 0; it < NCH; ++it)
    if (b0 + c0 + 4 * it < B) valid |= 1u << it;
  const bool tile_full = b0 + N <= B;
  const uint8_t *ac0 = ac_gen + c0 * 256 + eu * 8;  // (column c0, unit eu): 4 gates, 8 bytes
  // d(out) / c tiles: columns 4 apart flip bit 2 of the 128B-swizzle phase
  const uint32_t do_sw = ((((uint32_t)(eu >> 2) ^ (uint32_t)c0) & 7u) << 4);
  const uint8_t *do0 = do_gen + c0 * 128 + ((eu & 3) << 2);
  const uint32_t b_sw = (((((uint32_t)(eu & 15) >> 1) ^ (uint32_t)c0) & 7u) << 4);
  uint8_t *bg0 = b_gen + (eu >> 4) * (N * 128) + c0 * 128 + ((eu & 1) << 3);
  const float inv_keep = 1.0f / a.keep;
  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar_acts, N * 512);
    tma_load_3d(s_ac, &tm_acts, bar_acts, 4 * u0, b0, T - 1);
    tma_load_3d(s_do, &tm_do, bar_acts, u0, b0, T - 1);
    tma_load_3d(s_cp, &tm_cp, bar_acts, u0, b0, T - 1);
  }
  __syncwarp();
  for (int t = T - 1, step = 0; t >= 0; --t, ++step) {
    const int par = step & 1;
    uint32_t lapt = 0;
    if (dbg) lapt = (uint32_t)clock();
    if (warp == 0) {
      // the staged activations of step t were requested one step ahead; once the elementwise phase has
      // consumed them (bar_ready), request step t-1's so the load overlaps the product and the exchange
      mbar_wait(bar_ready, step & 1);
      if (elect_one()) {
        // d(pre) of the step leaves as it sits in the operand tile: rows x my 128 gate-interleaved columns
        tma_store_3d(&tm_dp, s_b, 128 * rank, b0, t);
        tma_store_3d(&tm_dp, s_b + (uint32_t)(N * 128), 128 * rank + 64, b0, t);
        if (t > 0) {
          mbar_expect_tx(bar_acts, N * 512);
          tma_load_3d(s_ac, &tm_acts, bar_acts, 4 * u0, b0, t - 1);
          tma_load_3d(s_do, &tm_do, bar_acts, u0, b0, t - 1);
          tma_load_3d(s_cp, &tm_cp, bar_acts, u0, b0, t - 1);
        }
        tma_store_wait_read();  // before the cluster barrier: the next elementwise pass rewrites the operand tile
      }
      __syncwarp();
    } else if (warp >= 2) {
      mbar_wait(bar_acts, step & 1);
      TCR_LAP(tk[0]);
      auto items = [&](auto full_tag) {
      constexpr bool FULL = decltype(full_tag)::value;
#pragma unroll
      for (int it = 0; it < NCH; ++it) {
        if (FULL || ((valid >> it) & 1u)) {
          const uint2 av = *reinterpret_cast<const uint2 *>(ac0 + it * 1024);
          const uint32_t w0 = av.x & 0xffff7fffu, w1 = av.y & 0xffff7fffu;
          const float2 a01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&w0));
          const float2 a23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&w1));
          const float ig = a01.x, jg = a01.y, fg = a23.x, og = a23.y;
          const bool eoc = (av.x & 0x8000u) != 0;  // flags the forward left in the sign bits of i and f
          float dscale = 1.0f;
          if (a.dropout) {
            dscale = (av.y & 0x8000u) ? 0.0f : inv_keep;
            if (a.mask != nullptr) dscale = drop_scale(a.mask, b0 + c0 + 4 * it, T, t, H, uo, a.row0, a.seed, a.keep);
          }
          const uint8_t *dp = do0 + it * 512 + (do_sw ^ (uint32_t)((it & 1) << 6));
          const float d_out = *reinterpret_cast<const float *>(dp) * dscale;
          const float cpv = *reinterpret_cast<const float *>(dp + N * 128);
          const float dh2 = d_out + (eoc ? 0.f : dhc[it]);
          const float tc = tanh_approx(fmaf(cpv, fg, ig * jg));
          const float dc2 = (eoc ? 0.f : dcc[it]) + dh2 * og * (1.0f - tc * tc);
          dcc[it] = dc2 * fg;
          const float d0 = dc2 * jg * ig * (1.0f - ig), d1 = dc2 * ig * (1.0f - jg * jg);
          const float d2 = dc2 * cpv * fg * (1.0f - fg), d3 = dh2 * tc * og * (1.0f - og);
          bsum[0] += d0; bsum[1] += d1; bsum[2] += d2; bsum[3] += d3;
          const __nv_bfloat16 h0 = __float2bfloat16_rn(d0), h1 = __float2bfloat16_rn(d1);
          const __nv_bfloat16 h2 = __float2bfloat16_rn(d2), h3 = __float2bfloat16_rn(d3);
          // own K slice, local column kk = 4u+g: 8 bytes at (row c, kk = 4u) of the swizzled operand (which is also what
          // the TMA store sends to the d(pre) array)
          uint2 pk;
          pk.x = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          pk.y = (uint32_t)__bfloat16_as_ushort(h2) | ((uint32_t)__bfloat16_as_ushort(h3) << 16);
          *reinterpret_cast<uint2 *>(bg0 + it * 512 + (b_sw ^ (uint32_t)((it & 1) << 6))) = pk;
        }
      }
      };
      if (tile_full) items(std::true_type{}); else items(std::false_type{});
      fence_proxy_async_smem();
      mbar_arrive(bar_ready);
      TCR_LAP(tk[1]);
    }
    if (t > 0) {
      if (warp == 1) {
        mbar_wait(bar_ready, step & 1);
        tc_fence_after();
        if (elect_one()) {
          for (int mt = 0; mt < nmt; ++mt) {
#pragma unroll
            for (int kb = 0; kb < 2; ++kb) {
              const uint64_t da = make_desc(s_w + (uint32_t)kb * (HP * 128) + (uint32_t)mt * 16384u, 0, 1024);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (a.a_tmem)
                  tc_mma_bf16_ts(tmem_acc + (uint32_t)(mt * N), tmem_a + (uint32_t)(mt * 64 + kb * 32 + k * 8), dBk[kb] + (uint64_t)(2 * k),
                                 idesc, (kb > 0 || k > 0) ? 1u : 0u);
                else
                  tc_mma_bf16(tmem_acc + (uint32_t)(mt * N), da + (uint64_t)(2 * k), dBk[kb] + (uint64_t)(2 * k), idesc,
                              (kb > 0 || k > 0) ? 1u : 0u);
              }
            }
            tc_commit(bar_acc + 8 * mt);
          }
        }
        __syncwarp();
      } else if (warp >= 2) {
        for (int mt = 0; mt < nmt; ++mt) {
          mbar_wait(bar_acc + 8 * mt, step & 1);
          if (mt == 0) TCR_LAP(tk[2]);
          tc_fence_after();
          const int m = mt * 128 + e;  // hidden unit of this lane's accumulator row
          uint32_t r[CPW];
          tc_ldn<CPW>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(mt * N + cg * CPW), r);
          if (m < H) {
            // rows 32q..32q+31 of this m-tile are the units of destination rank 4*mt + q; lane = unit
            __nv_bfloat16 *dst = part_base + (((size_t)(par * CS + (m >> 5)) * CS + rank) * 4 + cg) * part_blk + lane * CK;
            const uint32_t rdst = push ? mapa_u32(s_rx + (uint32_t)(((((size_t)par * CS + rank) * 4 + cg) * part_blk + lane * CK) * 2), (uint32_t)(m >> 5)) : 0u;
            const uint32_t rbar = push ? mapa_u32(bar_rx, (uint32_t)(m >> 5)) : 0u;
            uint32_t pk[CPW / 2];
#pragma unroll
            for (int j = 0; j < CPW / 2; ++j) {
              // position pos = sub*NCH + it of the block row holds column 4*it + sub (the reader's item order)
              const int ca = 4 * ((2 * j) % NCH) + (2 * j) / NCH, cb2 = 4 * ((2 * j + 1) % NCH) + (2 * j + 1) / NCH;
              const __nv_bfloat162 v = __floats2bfloat162_rn(__uint_as_float(r[ca]), __uint_as_float(r[cb2]));
              pk[j] = *reinterpret_cast<const uint32_t *>(&v);
            }
            if constexpr (CPW >= 8) {
#pragma unroll
              for (int j = 0; j < NJ; ++j) {
                const uint4 v4 = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
                if (push) st_async_v4(rdst + (uint32_t)(j * 32 * CK * 2), v4, rbar);
                else *reinterpret_cast<uint4 *>(dst + j * 32 * CK) = v4;
              }
            } else {
              if (push) st_async_v2(rdst, make_uint2(pk[0], pk[1]), rbar);
              else *reinterpret_cast<uint2 *>(dst) = make_uint2(pk[0], pk[1]);
            }
          }
        }
      }
    }
    TCR_LAP(tk[3]);
    __syncwarp();
    tc_fence_before();
    if (push) {
      __syncthreads();  // (the d(pre) store has read the operand tile, every accumulator has been read out)
      if (t > 0 && warp >= 2) {
        mbar_wait_cluster(bar_rx, step & 1);  // every rank's partials for my units have landed
        // the phase is over, so the next one can be announced; the buffers alternate by step parity, and a rank
        // sends step t-1's partials only after it has consumed step t's, which needs mine: no overwrite before the read
        if (tid == 64 && t > 1) mbar_expect_tx(bar_rx, rx_bytes);
      }
    } else {
      cluster_sync_all();
    }
    tc_fence_after();
    TCR_LAP(tk[4]);
    if (t > 0 && warp >= 2) {
      // d(h_{t-1}) for my units = sum over ranks (fixed order) of their partials
#pragma unroll
      for (int it = 0; it < NCH; ++it) dhc[it] = 0.f;
      // my NCH items sit at positions esub*NCH .. +NCH of the block row: chunk (esub*NCH)/CK, offset (esub*NCH)%CK
      const __nv_bfloat16 *src0 = (push ? rx_gen + (((size_t)par * CS) * 4 + cg) * part_blk
                                        : part_base + (((size_t)(par * CS + rank) * CS) * 4 + cg) * part_blk) +
                                  ((size_t)((esub * NCH) / CK) * 32 + eu) * CK + (esub * NCH) % CK;
      auto sum_parts = [&](auto local_tag) {  // (compile-time flag: a run-time choice per load keeps them from being batched)
      constexpr bool LOCAL = decltype(local_tag)::value;
      constexpr int UR = 8;  // loads in flight (L2 latency bound otherwise); the sum stays in rank order
      for (int r0 = 0; r0 < CS; r0 += UR) {
        if constexpr (NCH == 8) {
          uint4 v[UR];
#pragma unroll
          for (int j = 0; j < UR; ++j)
            v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const uint4 *>(src0 + (size_t)(r0 + j) * 4 * part_blk), LOCAL) : make_uint4(0, 0, 0, 0);
#pragma unroll
          for (int j = 0; j < UR; ++j) {
            const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].x));
            const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].y));
            const float2 f2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].z));
            const float2 f3 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].w));
            dhc[0] += f0.x; dhc[1] += f0.y; dhc[2] += f1.x; dhc[3] += f1.y;
            dhc[4] += f2.x; dhc[5] += f2.y; dhc[6] += f3.x; dhc[7] += f3.y;
          }
        } else if constexpr (NCH == 4) {
          uint2 v[UR];
#pragma unroll
          for (int j = 0; j < UR; ++j)
            v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const uint2 *>(src0 + (size_t)(r0 + j) * 4 * part_blk), LOCAL) : make_uint2(0, 0);
#pragma unroll
          for (int j = 0; j < UR; ++j) {
            const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].x));
            const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].y));
            dhc[0] += f0.x; dhc[1] += f0.y; dhc[2] += f1.x; dhc[3] += f1.y;
          }
        } else if constexpr (NCH == 2) {
          uint32_t v[UR];
#pragma unroll
          for (int j = 0; j < UR; ++j)
            v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const uint32_t *>(src0 + (size_t)(r0 + j) * 4 * part_blk), LOCAL) : 0u;
#pragma unroll
          for (int j = 0; j < UR; ++j) {
            const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j]));
            dhc[0] += f0.x; dhc[1] += f0.y;
          }
        } else {
          unsigned short v[UR];
#pragma unroll
          for (int j = 0; j < UR; ++j)
            v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const unsigned short *>(src0 + (size_t)(r0 + j) * 4 * part_blk), LOCAL) : (unsigned short)0;
#pragma unroll
          for (int j = 0; j < UR; ++j) dhc[0] += __bfloat162float(__ushort_as_bfloat16(v[j]));
        }
      }
      };
      if (push) sum_parts(std::true_type{}); else sum_parts(std::false_type{});
    }
    TCR_LAP(tk[5]);
  }
  if (warp >= 2 && a.dbias != nullptr) {
    // the 4 lanes of a unit (column sub-blocks) first, then one atomic per gate
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float v = bsum[g];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      if (esub == 0) atomicAdd(a.dbias + g * H + uo, v);
    }
  }
  if (dbg)
    printf("[tc bwd N=%d CS=%d] cycles/step: inputs+acts wait %lld, elementwise+stores %lld, MMA wait %lld, tmem ld + partial stores %lld, cluster barrier %lld, partial reads %lld\n",
           N, CS, tk[0] / T, tk[1] / T, tk[2] / T, tk[3] / T, tk[4] / T, tk[5] / T);
  __syncthreads();
  if (push) cluster_sync_all();  // nobody leaves while a peer may still send into it
  if (warp == 1) tc_dealloc(tmem_acc, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// backward, two half tiles in flight
// ------------------------------------------------------------------------------------------------
// Same idea as lstm_fwd_pp_kernel.  Per half tile s and step: E_s (elementwise: d(pre) from the saved
// activations, d(out) and the running d(h), d(c)), the K-split product on the tensor core, R_s (accumulator
// read-out, bf16 partials to L2), publication, P_s (sum of the 16 ranks' partials = d(h) for the next step).
// The 16 elementwise warps run E_0 E_1 R_0 R_1 P_0 P_1 per step: the product of a half runs while they work on
// the other half, and the publication round trip of half 0 (fence + remote mbarrier arrivals, done by a
// publisher warp) hides behind R_1.  No barrier.cluster in the loop.
// V2: the weight operand lives in TENSOR memory (64 columns per m-tile, the A tile of an MMA then costs 32 instead of 48
// cycles at N = 64); to make room the two half tiles share ONE set of accumulator slots (half 1's m-tile mt is multiplied
// once half 0's has been read out of the slot).  The 128 KB of shared memory this frees hold a receive buffer per half: once a
// half's partials are published, the publisher warp fetches this rank's 64 KB slice with ONE bulk copy (async proxy,
// mbarrier transaction bytes), so the elementwise warps sum from shared memory instead of sitting out an L2 round
// trip of 16 dependent-latency loads per thread.
template <int NCH, bool DBG, int HT = 0, bool V2 = false>  // NCH items per thread and half tile; HT: rnn_size at compile time (0: run time)
__global__ void __launch_bounds__(TCR_PP_THREADS, 1)
lstm_bwd_pp_kernel(const __grid_constant__ CUtensorMap tm_acts, const __grid_constant__ CUtensorMap tm_do,
                   const __grid_constant__ CUtensorMap tm_cp, const __grid_constant__ CUtensorMap tm_dp, TcRnnArgs a) {
  constexpr int N2 = 16 * NCH, N = 2 * N2;
  constexpr int CPW = N2 / 4;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const int H = HT ? HT : a.H, B = a.B, T = a.T, CS = HT ? HT / 32 : a.CS;
  const int t0 = a.t0, t1 = a.t1;  // this launch differentiates steps t1-1 .. t0
  const int nmt = (H + 127) / 128;
  const int HP = nmt * 128;
  const uint32_t half_bytes = (uint32_t)N2 * 768u;  // per half: d(pre) operand 256 B/row, activations 256, d(out) 128, c 128
  const uint32_t s_w = base;                        // 2 k-blocks x [HP rows][128 B]; V2: the two receive buffers instead
  const uint32_t rx_bytes = (uint32_t)CS * 256u * CPW;  // one half's partials for my 32 units: [src rank][column group][32 units][CPW] bf16
  const uint32_t s_half = s_w + (V2 ? 2u * rx_bytes : 2u * HP * 128u);
  // offsets inside a half's block
  const uint32_t o_b = 0, o_ac = (uint32_t)N2 * 256u, o_do = (uint32_t)N2 * 512u, o_cp = (uint32_t)N2 * 640u;
  const uint32_t s_bar = s_half + 2 * half_bytes;
  const uint32_t bar_in = s_bar, bar_epi = s_bar + 16, bar_rd = s_bar + 32, bar_ready = s_bar + 48, tmem_holder = s_bar + 64;
  const uint32_t bar_acc = s_bar + 128;  // [2][4]
  const uint32_t bar_rx = s_bar + 80;    // [2]  V2: my slice of a half's partials has landed in shared memory
  const uint32_t bar_free = s_bar + 192; // [4]  V2: half 0's m-tile mt has been read out of its accumulator slot
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));
  uint8_t *half_gen = gbase + (s_half - base);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int cid = cluster_id_x();
  const int b0 = cid * N;
  const int u0 = rank * 32;

  {
    const uint4 *src = reinterpret_cast<const uint4 *>(a.wimg + (size_t)rank * a.wimg_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(gbase + (s_w - base));
    if constexpr (!V2)
      for (int i = tid; i < a.wimg_bytes / 16; i += TCR_PP_THREADS) dst[i] = __ldg(src + i);
    // d(pre) operand tiles start as zeros (rows past the batch stay zero)
    for (int s = 0; s < 2; ++s) {
      uint4 *bz = reinterpret_cast<uint4 *>(half_gen + s * half_bytes + o_b);
      for (int i = tid; i < N2 * 256 / 16; i += TCR_PP_THREADS) bz[i] = make_uint4(0, 0, 0, 0);
    }
  }
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_in + 8 * s, 1);
      mbar_init(bar_epi + 8 * s, 32 * TCR_EPI_WARPS);
      mbar_init(bar_rd + 8 * s, 32 * TCR_EPI_WARPS);
      mbar_init(bar_ready + 8 * s, CS);
      for (int mt = 0; mt < 4; ++mt) mbar_init(bar_acc + 8 * (4 * s + mt), 1);
      mbar_init(bar_rx + 8 * s, 1);
      for (int mt = 0; mt < 4; ++mt) mbar_init(bar_free + 8 * (4 * s + mt), 32 * TCR_EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  const uint32_t tmem_a = tmem_acc + (uint32_t)(4 * N2);  // V2: behind the four accumulator slots
  if constexpr (V2) {
    if (warp >= 2 && warp < 6) {  // image: per m-tile plain row-major [128 rows][128 kk] bf16 (pack mode 3)
      const int qq = warp & 3, ee = 32 * qq + lane;
      for (int mt = 0; mt < nmt; ++mt) {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.wimg + (size_t)rank * a.wimg_bytes) + ((size_t)mt * 128 + ee) * 64;
#pragma unroll
        for (int hb = 0; hb < 2; ++hb) {
          uint32_t v[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint4 x = __ldg(reinterpret_cast<const uint4 *>(src + hb * 32) + j);
            v[4 * j] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
          }
          tc_st32(tmem_a + ((uint32_t)(32 * qq) << 16) + (uint32_t)(mt * 64 + hb * 32), v);
        }
      }
      tc_fence_before();
    }
    __syncthreads();
    tc_fence_after();
  }
  cluster_sync_all();

  __shared__ long long tk[8];
  if (tid < 8) tk[tid] = 0;
  __syncthreads();
  const bool dbg = DBG && rank == 0 && cid == 0 && tid == 64;

  if (warp == 0) {
    // ---------------- TMA producer: the three input tiles of (s, t) ----------------
    if (elect_one()) {
      auto load_in = [&](int s, int t) {
        const uint32_t hb = s_half + s * half_bytes;
        mbar_expect_tx(bar_in + 8 * s, N2 * 512);
        tma_load_3d(hb + o_ac, &tm_acts, bar_in + 8 * s, 4 * u0, b0 + s * N2, t);
        tma_load_3d(hb + o_do, &tm_do, bar_in + 8 * s, u0, b0 + s * N2, t);
        tma_load_3d(hb + o_cp, &tm_cp, bar_in + 8 * s, u0, b0 + s * N2, t);
      };
      load_in(0, t1 - 1);
      load_in(1, t1 - 1);
      for (int t = t1 - 1, step = 0; t >= t0; --t, ++step) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          mbar_wait(bar_epi + 8 * s, step & 1);  // the elementwise pass of (s, t) has consumed its tiles
          // d(pre) of (s, t) leaves as it sits in the operand tile: rows x my 128 gate-interleaved columns
          const uint32_t sb = s_half + s * half_bytes + o_b;
          tma_store_3d(&tm_dp, sb, 128 * rank, b0 + s * N2, t);
          tma_store_3d(&tm_dp, sb + (uint32_t)(N2 * 128), 128 * rank + 64, b0 + s * N2, t);
          if (t > t0) {
            tma_store_wait_read();  // the pass of (s, t-1) rewrites the operand tile once it has its inputs
            load_in(s, t - 1);
          }
        }
        if constexpr (V2) {
          // Between two steps this thread has nothing else to do: as soon as every rank has published a half's
          // partials, fetch my 32 units' slice of all of them (64 KB, contiguous) into that half's receive buffer.
          // Half 0's arrives while the elementwise warps still read out half 1.
          if (t > 0) {
#pragma unroll
            for (int s = 0; s < 2; ++s) {
              mbar_wait_cluster(bar_ready + 8 * s, step & 1);
              asm volatile("fence.proxy.async.global;" ::: "memory");  // ordinary stores -> read through the async proxy
              const __nv_bfloat16 *src = a.partial + (size_t)cid * 4 * ((size_t)CS * CS * 128 * CPW) +
                                         (size_t)(2 * s + (step & 1)) * ((size_t)CS * CS * 128 * CPW) + (size_t)rank * CS * 128 * CPW;
              mbar_expect_tx(bar_rx + 8 * s, rx_bytes);
              for (uint32_t off = 0; off < rx_bytes; off += 16384u)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(s_w + s * rx_bytes + off), "l"(reinterpret_cast<const char *>(src) + off),
                               "r"(rx_bytes - off < 16384u ? rx_bytes - off : 16384u), "r"(bar_rx + 8 * s) : "memory");
            }
          }
        }
      }
      tma_store_wait_all();
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, N2, false, false);
      for (int t = t1 - 1, step = 0; t >= t0 && t > 0; --t, ++step) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          mbar_wait(bar_epi + 8 * s, step & 1);
          tc_fence_after();
          const uint32_t sb = s_half + s * half_bytes + o_b;
          for (int mt = 0; mt < nmt; ++mt) {
            if constexpr (V2) {
              // the four accumulator slots serve both halves in turn: half 1's m-tile mt goes where half 0's was, once
              // that has been read out (half 0's next product needs nothing: the elementwise pass that feeds it comes
              // after half 1's read-out of the step before)
              if (s == 1) {
                mbar_wait(bar_free + 8 * mt, step & 1);
                tc_fence_after();
              }
            }
#pragma unroll
            for (int kb = 0; kb < 2; ++kb) {
              const uint64_t da = make_desc(s_w + (uint32_t)kb * (HP * 128) + (uint32_t)mt * 16384u, 0, 1024);
              const uint64_t db = make_desc(sb + (uint32_t)kb * (N2 * 128), 0, 1024);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if constexpr (V2)
                  tc_mma_bf16_ts(tmem_acc + (uint32_t)(mt * N2), tmem_a + (uint32_t)(mt * 64 + kb * 32 + k * 8),
                                 db + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
                else
                  tc_mma_bf16(tmem_acc + (uint32_t)((s * nmt + mt) * N2), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                              (kb > 0 || k > 0) ? 1u : 0u);
              }
            }
            tc_commit(bar_acc + 8 * (4 * s + mt));
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2 + TCR_EPI_WARPS) {
    // ---------------- publisher ----------------
    for (int t = t1 - 1, step = 0; t >= t0 && t > 0; --t, ++step) {
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        mbar_wait(bar_rd + 8 * s, step & 1);  // this CTA's partial stores of (s, t) happen before this
        // (no gpu-scope fence: the arrivals below are release operations at cluster scope, cumulative over what the wait above observed)
        if (lane < CS) mbar_arrive_remote(bar_ready + 8 * s, (uint32_t)lane);
        __syncwarp();
      }
    }
  } else {
    // ---------------- elementwise warps ----------------
    const int q = warp & 3, cg = (warp - 2) >> 2;
    const int e = 32 * q + lane;
    const int eu = e >> 2, esub = e & 3;
    const int uo = u0 + eu;
    const int c0 = cg * CPW + esub;  // item `it` of half s is column s*N2 + c0 + 4*it of the cluster tile
    float dhc[2][NCH], dcc[2][NCH];
    float bsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int it = 0; it < NCH; ++it) { dhc[s][it] = 0.f; dcc[s][it] = 0.f; }
    uint32_t valid[2] = {0, 0};
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int it = 0; it < NCH; ++it)
        if (b0 + s * N2 + c0 + 4 * it < B) valid[s] |= 1u << it;
    // a range that does not end the sequence continues from the d(h), d(c) its successor in time left behind
    float *carry_h = a.carry + (size_t)(b0 + c0) * H + uo, *carry_c = carry_h + (size_t)B * H;
    if (t1 < T) {
#pragma unroll
      for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int it = 0; it < NCH; ++it)
          if ((valid[s] >> it) & 1u) {
            dhc[s][it] = carry_h[(size_t)(s * N2 + 4 * it) * H];
            dcc[s][it] = carry_c[(size_t)(s * N2 + 4 * it) * H];
          }
    }
    // partial exchange, per half and parity: [dst rank][src rank][column group][chunk j][unit 32][CK columns]
    constexpr int CK = CPW >= 8 ? 8 : CPW;
    constexpr int NJ = CPW / CK;
    const size_t part_blk = (size_t)NJ * 32 * CK;
    const size_t part_par = (size_t)CS * CS * 4 * part_blk;  // one (half, parity) buffer
    __nv_bfloat16 *part_base = a.partial + (size_t)cid * 4 * part_par;
    const uint8_t *ac0 = half_gen + o_ac + c0 * 256 + eu * 8;
    const uint32_t do_sw = ((((uint32_t)(eu >> 2) ^ (uint32_t)c0) & 7u) << 4);
    const uint8_t *do0 = half_gen + o_do + c0 * 128 + ((eu & 3) << 2);
    const uint32_t b_sw = (((((uint32_t)(eu & 15) >> 1) ^ (uint32_t)c0) & 7u) << 4);
    uint8_t *bg0 = half_gen + o_b + (eu >> 4) * (N2 * 128) + c0 * 128 + ((eu & 1) << 3);
    const float inv_keep = 1.0f / a.keep;
    uint32_t lapt = 0;
    if (dbg) lapt = (uint32_t)clock();

    for (int t = t1 - 1, step = 0; t >= t0; --t, ++step) {
      const int par = step & 1;
      // ---- E_s ----
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        mbar_wait(bar_in + 8 * s, step & 1);
        TCR_LAP(tk[0]);
        auto items = [&](auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
#pragma unroll
        for (int it = 0; it < NCH; ++it) {
          if (FULL || ((valid[s] >> it) & 1u)) {
            const uint2 av = *reinterpret_cast<const uint2 *>(ac0 + s * half_bytes + it * 1024);
            const uint32_t w0 = av.x & 0xffff7fffu, w1 = av.y & 0xffff7fffu;
            const float2 a01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&w0));
            const float2 a23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&w1));
            const float ig = a01.x, jg = a01.y, fg = a23.x, og = a23.y;
            const bool eoc = (av.x & 0x8000u) != 0;
            float dscale = 1.0f;
            if (a.dropout) {
              dscale = (av.y & 0x8000u) ? 0.0f : inv_keep;
              if (a.mask != nullptr)
                dscale = drop_scale(a.mask, b0 + s * N2 + c0 + 4 * it, T, t, H, uo, a.row0, a.seed, a.keep);
            }
            const uint8_t *dp = do0 + s * half_bytes + it * 512 + (do_sw ^ (uint32_t)((it & 1) << 6));
            const float d_out = *reinterpret_cast<const float *>(dp) * dscale;
            const float cpv = *reinterpret_cast<const float *>(dp + N2 * 128);
            const float dh2 = d_out + (eoc ? 0.f : dhc[s][it]);
            const float tc = tanh_approx(fmaf(cpv, fg, ig * jg));
            const float dc2 = (eoc ? 0.f : dcc[s][it]) + dh2 * og * (1.0f - tc * tc);
            dcc[s][it] = dc2 * fg;
            const float d0 = dc2 * jg * ig * (1.0f - ig), d1 = dc2 * ig * (1.0f - jg * jg);
            const float d2 = dc2 * cpv * fg * (1.0f - fg), d3 = dh2 * tc * og * (1.0f - og);
            bsum[0] += d0; bsum[1] += d1; bsum[2] += d2; bsum[3] += d3;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(d0), h1 = __float2bfloat16_rn(d1);
            const __nv_bfloat16 h2 = __float2bfloat16_rn(d2), h3 = __float2bfloat16_rn(d3);
            uint2 pk;
            pk.x = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
            pk.y = (uint32_t)__bfloat16_as_ushort(h2) | ((uint32_t)__bfloat16_as_ushort(h3) << 16);
            *reinterpret_cast<uint2 *>(bg0 + s * half_bytes + it * 512 + (b_sw ^ (uint32_t)((it & 1) << 6))) = pk;
          }
        }
        };
        if (b0 + (s + 1) * N2 <= B) items(std::true_type{}); else items(std::false_type{});
        fence_proxy_async_smem();  // d(pre) operand tile -> tcgen05.mma
        mbar_arrive(bar_epi + 8 * s);
        TCR_LAP(tk[1]);
      }
      if (t == 0) break;
      // ---- R_s: accumulator read-out, bf16 partials for every rank ----
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        __nv_bfloat16 *pb = part_base + (size_t)(2 * s + par) * part_par;
        for (int mt = 0; mt < nmt; ++mt) {
          mbar_wait(bar_acc + 8 * (4 * s + mt), step & 1);
          if (mt == 0) TCR_LAP(tk[2]);
          tc_fence_after();
          const int m = mt * 128 + e;
          uint32_t r[CPW];
          if constexpr (V2) {
            tc_ldn<CPW>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(mt * N2 + cg * CPW), r);
            if (s == 0) {  // the slot may take half 1's m-tile now
              tc_fence_before();
              mbar_arrive(bar_free + 8 * mt);
            }
          } else {
            tc_ldn<CPW>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)((s * nmt + mt) * N2 + cg * CPW), r);
          }
          if (m < H) {
            __nv_bfloat16 *dst = pb + (((size_t)(m >> 5) * CS + rank) * 4 + cg) * part_blk + lane * CK;
            uint32_t pk[CPW / 2];
#pragma unroll
            for (int j = 0; j < CPW / 2; ++j) {
              const int ca = 4 * ((2 * j) % NCH) + (2 * j) / NCH, cb2 = 4 * ((2 * j + 1) % NCH) + (2 * j + 1) / NCH;
              const __nv_bfloat162 v = __floats2bfloat162_rn(__uint_as_float(r[ca]), __uint_as_float(r[cb2]));
              pk[j] = *reinterpret_cast<const uint32_t *>(&v);
            }
            if constexpr (CPW >= 8) {
#pragma unroll
              for (int j = 0; j < NJ; ++j)
                *reinterpret_cast<uint4 *>(dst + j * 32 * CK) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
            } else {
              *reinterpret_cast<uint2 *>(dst) = make_uint2(pk[0], pk[1]);
            }
          }
        }
        tc_fence_before();
        mbar_arrive(bar_rd + 8 * s);
        TCR_LAP(tk[3]);
      }
      // ---- P_s: d(h_{t-1}) of my units = sum over ranks (fixed order) of their partials ----
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if constexpr (V2) mbar_wait(bar_rx + 8 * s, step & 1);  // (written by the async proxy; the mbarrier orders it)
        else mbar_wait_cluster(bar_ready + 8 * s, step & 1);
        TCR_LAP(tk[4]);
        const __nv_bfloat16 *src0 = (V2 ? reinterpret_cast<const __nv_bfloat16 *>(gbase + (s_w - base) + s * rx_bytes) + (size_t)cg * part_blk
                                        : part_base + (size_t)(2 * s + par) * part_par + (((size_t)rank * CS) * 4 + cg) * part_blk) +
                                    ((size_t)((esub * NCH) / CK) * 32 + eu) * CK + (esub * NCH) % CK;
#pragma unroll
        for (int it = 0; it < NCH; ++it) dhc[s][it] = 0.f;
        constexpr int UR = 16;  // all ranks' partials in flight at once: one L2 round trip instead of two
        for (int r0 = 0; r0 < CS; r0 += UR) {
          if constexpr (NCH == 4) {
            uint2 v[UR];
#pragma unroll
            for (int j = 0; j < UR; ++j)
              v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const uint2 *>(src0 + (size_t)(r0 + j) * 4 * part_blk), V2) : make_uint2(0, 0);
#pragma unroll
            for (int j = 0; j < UR; ++j) {
              const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].x));
              const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j].y));
              dhc[s][0] += f0.x; dhc[s][1] += f0.y; dhc[s][2] += f1.x; dhc[s][3] += f1.y;
            }
          } else if constexpr (NCH == 1) {
            unsigned short v[UR];
#pragma unroll
            for (int j = 0; j < UR; ++j)
              v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const unsigned short *>(src0 + (size_t)(r0 + j) * 4 * part_blk), V2) : (unsigned short)0;
#pragma unroll
            for (int j = 0; j < UR; ++j) dhc[s][0] += __bfloat162float(__ushort_as_bfloat16(v[j]));
          } else {
            static_assert(NCH == 2, "half tiles of 16, 32 or 64 columns");
            uint32_t v[UR];
#pragma unroll
            for (int j = 0; j < UR; ++j)
              v[j] = (r0 + j < CS) ? ld_part(reinterpret_cast<const uint32_t *>(src0 + (size_t)(r0 + j) * 4 * part_blk), V2) : 0u;
#pragma unroll
            for (int j = 0; j < UR; ++j) {
              const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&v[j]));
              dhc[s][0] += f0.x; dhc[s][1] += f0.y;
            }
          }
        }
        TCR_LAP(tk[5]);
      }
    }
    if (t0 > 0) {  // hand d(h_{t0-1}), d(c_{t0-1}) to the launch that continues below t0
#pragma unroll
      for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int it = 0; it < NCH; ++it)
          if ((valid[s] >> it) & 1u) {
            carry_h[(size_t)(s * N2 + 4 * it) * H] = dhc[s][it];
            carry_c[(size_t)(s * N2 + 4 * it) * H] = dcc[s][it];
          }
    }
    if (a.dbias != nullptr) {
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v = bsum[g];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (esub == 0) atomicAdd(a.dbias + g * H + uo, v);
      }
    }
    if (dbg)
      printf("[tc bwd pp N=2x%d CS=%d] cycles per half step: inputs wait %lld, elementwise+stores %lld, MMA wait %lld, tmem ld + partial stores %lld, wait ready %lld, partial reads %lld\n",
             N2, CS, tk[0] / (2 * T), tk[1] / (2 * T), tk[2] / (2 * T), tk[3] / (2 * T), tk[4] / (2 * T), tk[5] / (2 * T));
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
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
    if (bwd == 4) {  // forward weights for tensor memory, rows of a lane quarter ordered [gate][8 units]: row 32q + 8g + ul
      const int mm = (int)(o / ((int64_t)nkb * 64)), kk2 = (int)(o % ((int64_t)nkb * 64));
      const int u = 8 * (mm >> 5) + (mm & 7), g = (mm >> 3) & 3;
      if (kk2 < H) v = mat[(size_t)(I + kk2) * G + g * H + u0 + u];
    } else if (bwd == 3) {  // backward weights for tensor memory: per m-tile plain row-major [128 rows m][128 kk], kk = 4u+g
      const int mm = (int)(o / 128), kk2 = (int)(o % 128);
      const int u = kk2 >> 2, g = kk2 & 3;
      if (mm < H) v = mat[(size_t)(I + mm) * G + g * H + u0 + u];
    } else if (bwd == 2) {  // forward weights for tensor memory: plain row-major [128 rows m=4u+g][nkb*64]
      const int mm = (int)(o / ((int64_t)nkb * 64)), kk2 = (int)(o % ((int64_t)nkb * 64));
      const int u = mm >> 2, g = mm & 3;
      if (kk2 < H) v = mat[(size_t)(I + kk2) * G + g * H + u0 + u];
    } else if (!bwd) {
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
int launch_fwd(const CUtensorMap &tm_h, const CUtensorMap &tm_xw, const CUtensorMap &tm_act, const TcRnnArgs &a,
               int nclusters, size_t smem, cudaStream_t st) {
  auto kern = a.debug ? lstm_fwd_tc_kernel<NCH, true> : lstm_fwd_tc_kernel<NCH, false>;
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
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm_h, tm_xw, tm_act, a));
  return 0;
}
template <int NCH>
int launch_fwd_pp(const CUtensorMap &tm_h, const CUtensorMap &tm_xw, const CUtensorMap &tm_act, const TcRnnArgs &a,
                  int nclusters, size_t smem, cudaStream_t st) {
  auto kern = a.debug ? lstm_fwd_pp_kernel<NCH, true> : lstm_fwd_pp_kernel<NCH, false>;
  if constexpr (NCH == 4) {
    if (a.H == 512 && !a.debug) kern = lstm_fwd_pp_kernel<NCH, false, 512>;  // the config-4 instance
  }
  if constexpr (NCH == 1) {  // config 1, and 100-200-row shards of config 4
    if (a.H == 512 && !a.debug) kern = lstm_fwd_pp_kernel<NCH, false, 512>;
    if (a.H == 256 && !a.debug) kern = lstm_fwd_pp_kernel<NCH, false, 256>;
  }
  SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (a.CS > 8) SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * a.CS);
  cfg.blockDim = dim3(TCR_PP_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = a.CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm_h, tm_xw, tm_act, a));
  return 0;
}
template <int NCH>
int launch_bwd_pp(const CUtensorMap &tm_acts, const CUtensorMap &tm_do, const CUtensorMap &tm_cp, const CUtensorMap &tm_dp,
                  const TcRnnArgs &a, int nclusters, size_t smem, cudaStream_t st) {
  auto kern = a.debug ? lstm_bwd_pp_kernel<NCH, true> : lstm_bwd_pp_kernel<NCH, false>;
  if constexpr (NCH == 4) {
    if (a.H == 512 && a.debug && a.a_tmem) kern = lstm_bwd_pp_kernel<NCH, true, 512, true>;
    if (a.H == 512 && !a.debug) kern = a.a_tmem ? lstm_bwd_pp_kernel<NCH, false, 512, true> : lstm_bwd_pp_kernel<NCH, false, 512>;  // the config-4 instances
  }
  SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (a.CS > 8) SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * a.CS);
  cfg.blockDim = dim3(TCR_PP_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = a.CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm_acts, tm_do, tm_cp, tm_dp, a));
  return 0;
}
template <int NCH>
int launch_bwd(const CUtensorMap &tm_acts, const CUtensorMap &tm_do, const CUtensorMap &tm_cp, const CUtensorMap &tm_dp,
               const TcRnnArgs &a, int nclusters, size_t smem, cudaStream_t st) {
  auto kern = a.debug ? lstm_bwd_tc_kernel<NCH, true> : lstm_bwd_tc_kernel<NCH, false>;
  if constexpr (NCH <= 2) {  // the small tiles of config 1 and of multi-GPU shards of config 4
    if (a.H == 512 && !a.debug) kern = lstm_bwd_tc_kernel<NCH, false, 512>;
    if (a.H == 256 && !a.debug) kern = lstm_bwd_tc_kernel<NCH, false, 256>;
  }
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
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm_acts, tm_do, tm_cp, tm_dp, a));
  return 0;
}

size_t smem_need(bool bwd, int H, int N, bool a_tmem, bool push = false) {
  const int nkb = (H + 63) / 64, HP = (H + 127) / 128 * 128;
  size_t s = bwd ? (a_tmem ? 0 : (size_t)2 * HP * 128) + 2 * (size_t)N * 128 + 4 * (size_t)N * 128 + (push ? (size_t)2 * (H / 32) * 32 * N * 2 : 0)
                 : (a_tmem ? 0 : (size_t)nkb * 16384) + (size_t)(push ? 2 : 1) * nkb * N * 128 + 2 * (size_t)N * 128 + 2 * (size_t)N * 128;
  return s + 1024 + 256;
}

struct TcPlan { int N; bool a_tmem, mcast_ok, pp; int pp_fwd_env, pp_bwd_env; };

// Smallest batch tile the two-half-tile BPTT kernel takes.  Below it the single-tile kernel wins since it sends its
// partial sums rank to rank (measured at H=512, B=200, N=32: 2.5 ms per layer pass against 3.5 ms).
int bwd_pp_min() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("SKRNN_TC_PP_BWD_MIN"); v = e ? atoi(e) : 64; }
  return v;
}

// batch tile and kernel variant of a layer pass at batch B
TcPlan tc_plan(const skrnn_handle *h, bool bwd, int B) {
  const int H = h->H, CS = H / 32;
  const int nkb = (H + 63) / 64, HP = (H + 127) / 128 * 128;
  TcPlan p = {};
  // batch tile: the smallest N whose cluster count fits on the GPU at once, within shared memory / TMEM
  static int atm_env = -1;
  if (atm_env < 0) { const char *e = getenv("SKRNN_TC_ATMEM"); atm_env = (e && e[0] == '0') ? 0 : 1; }
  const bool a_tmem = !bwd && atm_env;
  int N = 0;
  const int cap_clusters = std::max(1, h->num_sms / CS - (CS >= 8 ? 2 : 0));
  for (int cand : {16, 32, 64, 128}) {
    if (smem_need(bwd, H, cand, a_tmem) > 227 * 1024) break;
    if (a_tmem && (cand + 31) / 32 * 32 + nkb * 32 > 512) break;
    if (bwd && (HP / 128) * cand > 512) break;
    N = cand;
    if ((B + cand - 1) / cand <= cap_clusters) break;
  }
  // The two-half-tile kernels need N >= 32 (two MMA tiles of 16 columns): latency-bound small batches prefer two
  // halves of 16 rows in flight on fewer clusters over one 16-row tile per cluster.
  static int pp_env = -1, ppb_env = -1, mc_env = -1;
  if (pp_env < 0) { const char *e = getenv("SKRNN_TC_PP"); pp_env = (e && e[0] == '0') ? 0 : 1; }
  if (ppb_env < 0) { const char *e = getenv("SKRNN_TC_PP_BWD"); ppb_env = (e && e[0] == '0') ? 0 : 1; }
  if (mc_env < 0) { const char *e = getenv("SKRNN_TC_MCAST"); mc_env = (e && e[0] == '0') ? 0 : 1; }
  const bool mcast_ok = CS >= 2 && CS == 2 * nkb;
  // (measured at H=512, B=100: forward 6.8 K cycles/step with 7 clusters x 16 rows vs 4.6 K with 4 clusters x 2x16;
  // the backward pass is faster the other way round, 5.2 K vs 8.7 K, so only the forward tile is widened)
  if (N == 16 && B > 16 && !bwd && pp_env != 0 && mcast_ok && a_tmem) N = 32;
  p.N = N; p.a_tmem = a_tmem; p.mcast_ok = mcast_ok; p.pp_fwd_env = pp_env; p.pp_bwd_env = ppb_env;
  p.pp = bwd ? (ppb_env && N >= bwd_pp_min()) : (pp_env && mc_env && mcast_ok && a_tmem && N >= 32);
  return p;
}

}  // namespace

// How many time ranges a layer of this handle is run in at batch B (1 = one launch per layer and direction).  Ranges
// let the hoisted GEMMs of a range run next to the neighbouring range's recurrence (api.cu); they need the
// two-half-tile kernels in both directions.  MEASURED at config 4 (lstm H512 L3 B800 T300): 15.4 ms with 1 range,
// 15.8 / 15.9 / 15.6 ms with 2 / 4 / 6 -- the GEMMs do leave the critical path, but a GEMM on the 36 free SMs takes
// L2 bandwidth from the recurrent kernels' exchange and slows them by more than it saves (forward 4.4 -> 5.2 ms, BPTT
// 7.5 -> 8.6 ms).  So the default is 1; SKRNN_TC_RANGES=n turns ranges on (kept, with its parity test, for GPUs or
// shapes where the balance differs).
int sk_rnn_tc_ranges(const skrnn_handle *h, int B, int T) {
  const char *e = getenv("SKRNN_TC_RANGES");  // 1 switches ranges off, n > 1 forces n (read per call: tests toggle it)
  const int env = e ? atoi(e) : 0;
  if (!sk_rnn_tc_eligible(h) || h->ws.bptt_carry == nullptr) return 1;
  if (!tc_plan(h, false, B).pp || !tc_plan(h, true, B).pp) return 1;
  int C = env > 0 ? env : 1;
  while (C > 1 && (T / C < 16 || (T / C) % 2 != 0)) --C;  // ranges start on even steps and are not tiny
  return C;
}

// Static part of the eligibility test: run_forward uses it to decide in which format it hoists x*Wx+b.
bool sk_rnn_tc_eligible(const skrnn_handle *h) {
  const int H = h->H;
  if (h->cfg.precision != SKRNN_PREC_BF16 || h->cfg.cell != SKRNN_CELL_LSTM) return false;
  if (!(H == 32 || H == 64 || H == 128 || H == 256 || H == 512)) return false;
  if (h->tc_rnn_disabled) return false;
  static int env_off = -1;
  if (env_off < 0) { const char *e = getenv("SKRNN_DISABLE_TC_RNN"); env_off = (e && e[0] == '1') ? 1 : 0; }
  return !env_off;
}

int sk_rnn_tc_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *x_bt5,
                    const float *state_in, float *state_out, const float *mask, uint64_t seed, int row0,
                    bool training) {
  const int H = h->H;
  if (!sk_rnn_tc_eligible(h)) return -1;
  if (!bwd && (!h->xw_il_ok[l] || h->ws.xw_il[l] == nullptr)) return -1;  // the projection was hoisted in the fp32 format
  const int CS = H / 32;
  const int nkb = (H + 63) / 64, HP = (H + 127) / 128 * 128;
  const TcPlan plan = tc_plan(h, bwd, B);
  const int N = plan.N;
  const bool a_tmem = plan.a_tmem, mcast_ok = plan.mcast_ok;
  const int pp_env = plan.pp_fwd_env, ppb_env = plan.pp_bwd_env;
  if (N == 0) return -1;
  // a time range [t0, t1) of the unroll (api.cu splits a layer into ranges to overlap the GEMMs around it)
  const int t0 = (h->chunk_t1 > h->chunk_t0) ? h->chunk_t0 : 0, t1 = (h->chunk_t1 > h->chunk_t0) ? h->chunk_t1 : T;
  const bool ranged = t0 != 0 || t1 != T;
  SK_CHECK(!ranged || (plan.pp && t0 % 2 == 0 && t0 >= 0 && t1 <= T), "tcgen05 recurrent kernels: time range [%d,%d) not supported here", t0, t1);
  const int nclusters = (B + N - 1) / N;
  const size_t smem = smem_need(bwd, H, N, a_tmem);
  const int img_bytes = bwd ? 2 * HP * 128 : nkb * 16384;
  // single-tile BPTT kernel with a small tile: the weight operand fits in tensor memory next to the accumulators
  static int batm_env = -1;
  if (batm_env < 0) { const char *e = getenv("SKRNN_TC_ATMEM_BWD"); batm_env = (e && e[0] == '0') ? 0 : 1; }
  const bool bwd_atmem_tc = bwd && !plan.pp && batm_env && ((HP / 128) * N + 31) / 32 * 32 + (HP / 128) * 64 <= 512;
  // two-half-tile BPTT kernel at config 4's shape: weights in tensor memory, accumulator slots shared by the m-tiles,
  // partials fetched into shared memory by one bulk copy per half (lstm_bwd_pp_kernel, V2); SKRNN_TC_BWD_V2=1 selects it
  // (opt-in, read per call: it measured level with the older form -- 7.7 against 7.6 ms for config 4's three layers; the
  // faster products and the shared-memory sums are paid for by the copy sitting behind the publication)
  const char *v2_e = getenv("SKRNN_TC_BWD_V2");
  const bool bwd_v2 = bwd && plan.pp && v2_e && v2_e[0] == '1' && H == 512 && N == 128;
  const bool bwd_atmem = bwd_atmem_tc || bwd_v2;
  SK_CHECK((int64_t)img_bytes * CS <= h->ws.wpack_floats * 4, "weight image workspace too small");
  {
    SkLaunch scope(h, st, SK_CAT_MISC);
    int64_t total = (int64_t)img_bytes / 2 * CS;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)4 * h->num_sms);
    // (forward two-half-tile kernel: rows of a lane quarter ordered [gate][8 units], see lstm_fwd_pp_kernel)
    const bool fwd_quad = !bwd && plan.pp && a_tmem;
    pack_tc_image_kernel<<<blocks, 256, 0, st>>>(bwd ? (bwd_atmem ? 3 : 1) : (a_tmem ? (fwd_quad ? 4 : 2) : 0), h->params + h->lp[l].w[0], h->lp[l].I, H, CS, nkb, HP,
                                                img_bytes, reinterpret_cast<uint8_t *>(h->ws.wpack));
    SK_CUDA(cudaGetLastError());
  }
  TcRnnArgs a = {};
  a.B = B; a.T = T; a.H = H; a.G = h->G; a.N = N; a.CS = CS; a.nkb = nkb;
  int cols = bwd ? (bwd_v2 ? 512 : bwd_atmem ? ((HP / 128) * N + 31) / 32 * 32 + (HP / 128) * 64 : (HP / 128) * N)
                 : (a_tmem ? (N + 31) / 32 * 32 + nkb * 32 : N);
  a.a_tmem = (a_tmem || bwd_atmem) ? 1 : 0;
  int tc = 32;
  while (tc < cols) tc <<= 1;
  a.tmem_cols = tc;
  a.t0 = t0; a.t1 = t1;
  a.carry = h->ws.bptt_carry;
  SK_CHECK(!ranged || !bwd || a.carry != nullptr, "BPTT time ranges need the carry buffer");
  a.row0 = row0;
  a.dropout = (training && l == h->L - 1 && h->cfg.keep_prob < 1.0f) ? 1 : 0;
  a.keep = h->cfg.keep_prob;
  a.seed = seed;
  (void)x_bt5;
  a.eoc = h->ws.eoc; a.state_in = state_in; a.state_out = state_out; a.mask = mask;
  a.acts = h->ws.xw[l]; a.cprev = h->ws.cprev[l]; a.out_bf = h->ws.out_bf[l]; a.hprev_bf = h->ws.hprev_bf[l];
  a.dout = h->ws.dout; a.dpre_bf = h->ws.dpre_bf; a.acts_bf = h->ws.acts_bf[l];
  a.dbias = bwd ? h->grads + h->lp[l].b[0] : nullptr;
  if (bwd) h->bias_grad_done = true;
  a.wimg = reinterpret_cast<const uint8_t *>(h->ws.wpack);
  a.wimg_bytes = img_bytes;
  {
    static int dbg_env = -1;
    if (dbg_env < 0) { const char *e = getenv("SKRNN_TC_DEBUG"); dbg_env = (e && e[0] == '1') ? 1 : 0; }
    a.debug = dbg_env;
  }
  SK_CHECK(h->ws.acts_bf[l] != nullptr, "tcgen05 recurrent kernels: activation workspace missing");
  CUtensorMap tm_act;
  {
    // saved gate activations acts_bf[t][b][4*u + g] (bf16, not swizzled): one box = 32 units x N rows
    uint64_t dims[3] = {(uint64_t)h->G, (uint64_t)B, (uint64_t)T};
    uint64_t str[2] = {(uint64_t)h->G * 2, (uint64_t)B * h->G * 2};
    uint32_t box[3] = {128, (uint32_t)N, 1};
    SK_TRY(sk_make_tmap(&tm_act, 2, h->ws.acts_bf[l], 3, dims, str, box, false));
  }
  int rc = 0;
  if (!bwd) {
    // hoisted projection xw_il[t][b][4*u + g] (bf16, not swizzled, written by run_forward): one box = 32 units x N rows
    if (!h->xw_il_ok[l] || h->ws.xw_il[l] == nullptr) return -1;
    CUtensorMap tm_xw;
    {
      uint64_t dims[3] = {(uint64_t)h->G, (uint64_t)B, (uint64_t)T};
      uint64_t str[2] = {(uint64_t)h->G * 2, (uint64_t)B * h->G * 2};
      uint32_t box[3] = {128, (uint32_t)N, 1};
      SK_TRY(sk_make_tmap(&tm_xw, 2, h->ws.xw_il[l], 3, dims, str, box, false));
    }
    CUtensorMap tm_h;
    uint64_t dims[3] = {(uint64_t)H, (uint64_t)B, (uint64_t)T};
    uint64_t str[2] = {(uint64_t)H * 2, (uint64_t)B * H * 2};
    static int mc_env = -1;
    if (mc_env < 0) { const char *e = getenv("SKRNN_TC_MCAST"); mc_env = (e && e[0] == '0') ? 0 : 1; }
    const bool pp = pp_env && mc_env && mcast_ok && a_tmem && N >= 32;  // two half tiles in flight (lstm_fwd_pp_kernel)
    a.mcast = (mc_env && mcast_ok && (N >= 64 || pp)) ? 1 : 0;
    uint32_t box[3] = {64, (uint32_t)(pp ? N / 4 : (a.mcast ? N / 2 : N)), 1};
    SK_TRY(sk_make_tmap(&tm_h, 2, h->ws.hprev_bf[l], 3, dims, str, box));
    if (pp) {
      // the xw box of a half tile has N/2 rows
      uint64_t xdims[3] = {(uint64_t)h->G, (uint64_t)B, (uint64_t)T};
      uint64_t xstr[2] = {(uint64_t)h->G * 2, (uint64_t)B * h->G * 2};
      uint32_t xbox[3] = {128, (uint32_t)(N / 2), 1};
      SK_TRY(sk_make_tmap(&tm_xw, 2, h->ws.xw_il[l], 3, xdims, xstr, xbox, false));
      // my 32 units of h_t for a half tile's rows leave as one plain (unswizzled) box
      uint32_t sbox[3] = {32, (uint32_t)(N / 2), 1};
      CUtensorMap tm_hst;
      SK_TRY(sk_make_tmap(&tm_hst, 2, h->ws.hprev_bf[l], 3, dims, str, sbox, false));
      // small tiles: h travels rank to rank through distributed shared memory (see the publisher warp); SKRNN_TC_PUSH=<max N>
      // moves the limit (0: never)
      static int push_max = -1;
      if (push_max < 0) { const char *e = getenv("SKRNN_TC_PUSH"); push_max = e ? atoi(e) : 32; }
      a.push = (N <= push_max) ? 1 : 0;
      static int sig_env = -1;
      if (sig_env < 0) { const char *e = getenv("SKRNN_TC_SIG_ASYNC"); sig_env = (e && e[0] == '0') ? 0 : 1; }
      a.sig_async = sig_env;
      const size_t smem_pp = smem_need(false, H, N, a_tmem, a.push != 0);
      SkLaunch scope(h, st, SK_CAT_RNN_FWD);
      if (N == 32) rc = launch_fwd_pp<1>(tm_h, tm_xw, tm_hst, a, nclusters, smem_pp, st);
      else if (N == 64) rc = launch_fwd_pp<2>(tm_h, tm_xw, tm_hst, a, nclusters, smem_pp, st);
      else rc = launch_fwd_pp<4>(tm_h, tm_xw, tm_hst, a, nclusters, smem_pp, st);
      if (rc != 0) {
        fprintf(stderr, "libskrnn: lstm_fwd_pp_kernel launch failed (%s)\n", skrnn_last_error());
        cudaGetLastError();
        return 1;
      }
      return 0;
    }
    SkLaunch scope(h, st, SK_CAT_RNN_FWD);
    if (N == 16) rc = launch_fwd<1>(tm_h, tm_xw, tm_act, a, nclusters, smem, st);
    else if (N == 32) rc = launch_fwd<2>(tm_h, tm_xw, tm_act, a, nclusters, smem, st);
    else if (N == 64) rc = launch_fwd<4>(tm_h, tm_xw, tm_act, a, nclusters, smem, st);
    else rc = launch_fwd<8>(tm_h, tm_xw, tm_act, a, nclusters, smem, st);
  } else {
    const int64_t need = (int64_t)nclusters * 2 * CS * H * N;
    if (need > h->ws.partial_floats) {
      SK_CUDA(cudaStreamSynchronize(st));
      cudaFree(h->ws.partial);
      h->ws.partial = nullptr;
      h->ws.partial_floats = 0;
      SK_CUDA(cudaMalloc(&h->ws.partial, need * sizeof(__nv_bfloat16)));
      h->ws.partial_floats = need;
    }
    a.partial = h->ws.partial;
    const int ppN = (ppb_env && N >= bwd_pp_min()) ? N / 2 : N;  // rows per box (half tiles for the two-half-tile kernel)
    // d(pre) [T,B,4H] bf16 with gate-interleaved columns (4*unit + gate): what the kernel's operand tile holds, stored
    // as two 64-column boxes in the operand's own (128B-swizzled) layout
    CUtensorMap tm_dp;
    {
      uint64_t ddims[3] = {(uint64_t)h->G, (uint64_t)B, (uint64_t)T};
      uint64_t dstr[2] = {(uint64_t)h->G * 2, (uint64_t)B * h->G * 2};
      uint32_t dbox[3] = {64, (uint32_t)ppN, 1};
      SK_TRY(sk_make_tmap(&tm_dp, 2, h->ws.dpre_bf, 3, ddims, dstr, dbox));
    }
    // d(out) and the c fed to each step, [T,B,H] fp32: boxes of 32 units x N rows, 128B-swizzled
    CUtensorMap tm_do, tm_cp;
    {
      uint64_t dims[3] = {(uint64_t)H, (uint64_t)B, (uint64_t)T};
      uint64_t str[2] = {(uint64_t)H * 4, (uint64_t)B * H * 4};
      uint32_t box[3] = {32, (uint32_t)N, 1};
      SK_TRY(sk_make_tmap(&tm_do, 4, h->ws.dout, 3, dims, str, box));
      SK_TRY(sk_make_tmap(&tm_cp, 4, h->ws.cprev[l], 3, dims, str, box));
    }
    if (ppb_env && N >= bwd_pp_min()) {  // two half tiles in flight (lstm_bwd_pp_kernel): boxes of N/2 rows
      uint64_t adims[3] = {(uint64_t)h->G, (uint64_t)B, (uint64_t)T};
      uint64_t astr[2] = {(uint64_t)h->G * 2, (uint64_t)B * h->G * 2};
      uint32_t abox[3] = {128, (uint32_t)(N / 2), 1};
      CUtensorMap tm_act2;
      SK_TRY(sk_make_tmap(&tm_act2, 2, h->ws.acts_bf[l], 3, adims, astr, abox, false));
      uint64_t dims[3] = {(uint64_t)H, (uint64_t)B, (uint64_t)T};
      uint64_t str[2] = {(uint64_t)H * 4, (uint64_t)B * H * 4};
      uint32_t box[3] = {32, (uint32_t)(N / 2), 1};
      SK_TRY(sk_make_tmap(&tm_do, 4, h->ws.dout, 3, dims, str, box));
      SK_TRY(sk_make_tmap(&tm_cp, 4, h->ws.cprev[l], 3, dims, str, box));
      SkLaunch scope(h, st, SK_CAT_RNN_BWD);
      if (N == 32) rc = launch_bwd_pp<1>(tm_act2, tm_do, tm_cp, tm_dp, a, nclusters, smem, st);
      else if (N == 64) rc = launch_bwd_pp<2>(tm_act2, tm_do, tm_cp, tm_dp, a, nclusters, smem, st);
      else rc = launch_bwd_pp<4>(tm_act2, tm_do, tm_cp, tm_dp, a, nclusters, smem, st);
      if (rc != 0) return 1;
      return 0;
    }
    // small tiles: the partial sums travel rank to rank through distributed shared memory; SKRNN_TC_PUSH_BWD=<max N>
    static int pushb_max = -1;
    if (pushb_max < 0) { const char *e = getenv("SKRNN_TC_PUSH_BWD"); pushb_max = e ? atoi(e) : 32; }
    a.push = (N <= pushb_max) ? 1 : 0;
    const size_t smem_b = smem_need(true, H, N, bwd_atmem, a.push != 0);
    if (smem_b > 227 * 1024) { a.push = 0; }
    const size_t smem_b2 = a.push ? smem_b : smem_need(true, H, N, bwd_atmem, false);
    SkLaunch scope(h, st, SK_CAT_RNN_BWD);
    if (N == 16) rc = launch_bwd<1>(tm_act, tm_do, tm_cp, tm_dp, a, nclusters, smem_b2, st);
    else if (N == 32) rc = launch_bwd<2>(tm_act, tm_do, tm_cp, tm_dp, a, nclusters, smem_b2, st);
    else if (N == 64) rc = launch_bwd<4>(tm_act, tm_do, tm_cp, tm_dp, a, nclusters, smem_b2, st);
    else rc = launch_bwd<8>(tm_act, tm_do, tm_cp, tm_dp, a, nclusters, smem_b2, st);
    if (rc != 0) return 1;  // the forward's activations live in this family's format: no fallback possible
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
