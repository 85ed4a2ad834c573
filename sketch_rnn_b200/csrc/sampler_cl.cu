// sampler_cl.cu -- low-latency variant of the reference-precision sampler (reference Model.sample, model.py:187-264)
// for few sketches: the weights of the whole stack are split over a CLUSTER of 16 CTAs and stay in shared memory.
//
// sampler.cu's kernel keeps a group of sketches per CTA and streams every layer's fp32 weights from L2 on every step:
// fine when thousands of sketches amortise it, but one step then costs ~80 us however few sketches there are (a
// dependent chain of ~1000 L2 round trips), and sample.py draws tens of sketches at a time.  Here rank r of the
// cluster owns hidden units [r*UPC, (r+1)*UPC) of every layer (UPC = H/16) and output columns [r*OC, (r+1)*OC), keeps
// exactly those weight columns in its shared memory for the whole launch (kanji model, lstm H = 256 L = 2: 3.3 MB of
// fp32 weights = 204 KB per CTA), and the cluster runs SB = 4 sketches together:
//   per layer:  every rank computes its units' gates for the 4 sketches (one thread per (unit, sketch): all gates of a
//               unit meet in one thread, the cell state lives in its registers) and sends h into EVERY rank's operand
//               buffers with st.async: the store itself reports its bytes on the receiver's mbarrier, so an exchange
//               is "wait until H*SB*4 bytes have landed" -- no barrier.cluster (a gpu-scope fence plus an L1
//               invalidate, ~5 us per step for three of them) in the loop;
//   output:     its OC logits per sketch -> every rank's logits buffer, same mechanism;
//   draw:       every rank draws the next point of all 4 sketches redundantly (two warps per sketch), so the feedback
//               (dx, dy, eos, eoc, cont) needs no further exchange; rank 0 writes the strokes.
// Same arithmetic contract as sampler.cu: every output element is ONE sequential fused-multiply-add chain over
// k = 0..K-1 of [x ; h] * W followed by + bias, and every transcendental comes from skrnn_math.h -- bit-identical to
// oracle/sampler_ref.c given the same uniform draws.
#include <stdlib.h>

#include "common.cuh"
#include "skrnn_math.h"

namespace {

constexpr int SCL_CS = 16;  // CTAs per cluster
constexpr int SCL_SB = 4;   // sketches per cluster
constexpr int SCL_THREADS = 256;  // 8 warps: the draw of a sketch is shared by warps r (mixture) and r + 4 (pen state, normals)

struct SclArgs {
  int n, steps, H, L, M, NOUT, cell;
  int UPC, OC;  // hidden units / output columns per rank
  float temp_mix, temp_pen, data_scale;
  int stop_if_eoc;
  const float *uniforms;  // [n,steps,4] or null
  uint64_t seed;
  int sketch0;
  const float *params;
  int64_t w0[SK_MAX_LAYERS], b0[SK_MAX_LAYERS], w1[SK_MAX_LAYERS], b1[SK_MAX_LAYERS];
  int64_t out_w, out_b;
  float *strokes;  // [n, steps, 5]
  int *lengths;    // [n]
  float *mixp;     // [n, steps, NOUT] or null
  int debug;       // SKRNN_TC_DEBUG=1: block 0 prints its per-phase cycle split
};
#define SCL_LAP(slot) do { if (dbg) { const long long now_ = clock64(); tk[slot] += now_ - lapt; lapt = now_; } } while (0)

__device__ __forceinline__ uint32_t scl_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t scl_cluster_id() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void scl_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// v -> the same shared-memory location in every CTA of the cluster; each store adds its 4 bytes to the transaction
// count of the receiving CTA's mbarrier `bar` (same offset everywhere)
__device__ __forceinline__ void scl_send(const float *local, float v, uint32_t bar) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(local);
#pragma unroll
  for (int r = 0; r < SCL_CS; ++r) {
    uint32_t ra, rb;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(r));
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rb) : "r"(bar), "r"(r));
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(ra), "r"(__float_as_uint(v)), "r"(rb)
                 : "memory");
  }
}
__device__ __forceinline__ void scl_arm(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void scl_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return;
    if (spin > (1u << 24)) {
      printf("libskrnn: cluster sampler exchange timed out (block %d, thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}

// acc[g] = sequential fma chain over k = 0..K-1 of As[k] * W[k][g]  (W: NG adjacent floats per k, `stride` floats from
// one k to the next).  The chain itself costs 4 cycles per k; the operands of the NEXT batch of 8 are requested before
// the current batch is consumed so that the shared-memory latency hides behind it (one warp per scheduler: nothing
// else would).  `As` is 16-byte aligned.
template <int NG>
__device__ __forceinline__ void scl_dot(const float *__restrict__ As, const float *__restrict__ W, int stride, int K, float (&acc)[NG]) {
  constexpr int UB = 8;
  float wc[UB][NG], ac[UB];
  auto load = [&](int k0, float (&w)[UB][NG], float (&av)[UB]) {
    const float4 a0 = *reinterpret_cast<const float4 *>(As + k0), a1 = *reinterpret_cast<const float4 *>(As + k0 + 4);
    av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w; av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
#pragma unroll
    for (int j = 0; j < UB; ++j) {
      const float *wp = W + (size_t)(k0 + j) * stride;
      if constexpr (NG == 4) {
        const float4 v = *reinterpret_cast<const float4 *>(wp);
        w[j][0] = v.x; w[j][1] = v.y; w[j][2] = v.z; w[j][3] = v.w;
      } else if constexpr (NG == 2) {
        const float2 v = *reinterpret_cast<const float2 *>(wp);
        w[j][0] = v.x; w[j][1] = v.y;
      } else {
        w[j][0] = *wp;
      }
    }
  };
  auto fma8 = [&](const float (&w)[UB][NG], const float (&av)[UB]) {
#pragma unroll
    for (int j = 0; j < UB; ++j)
#pragma unroll
      for (int g = 0; g < NG; ++g) acc[g] = SKM_FMA(av[j], w[j][g], acc[g]);
  };
  const int KB = K & ~(2 * UB - 1);  // blocks of 16: two register sets in ping-pong, no copies
  if (KB > 0) {
    float wd[UB][NG], ad[UB];
    load(0, wc, ac);
    int k0 = 0;
    for (; k0 + 2 * UB < KB; k0 += 2 * UB) {
      load(k0 + UB, wd, ad);
      fma8(wc, ac);
      load(k0 + 2 * UB, wc, ac);
      fma8(wd, ad);
    }
    load(k0 + UB, wd, ad);
    fma8(wc, ac);
    fma8(wd, ad);
  }
  for (int k = KB; k < K; ++k) {
    const float av = As[k];
#pragma unroll
    for (int g = 0; g < NG; ++g) acc[g] = SKM_FMA(av, W[(size_t)k * stride + g], acc[g]);
  }
}

// first index with running (fp64) sum >= u, else -1   (get_pi_idx, model.py:189-197)
__device__ __forceinline__ int scl_pick_idx(const float *pdf, int n, float u) {
  double acc = 0.0;
  const double ud = (double)u;
  for (int k = 0; k < n; ++k) {
    acc = SKM_DADD(acc, (double)pdf[k]);
    if (acc >= ud) return k;
  }
  return -1;
}

// shared-memory plan (floats); the same in every CTA
struct SclPlan {
  int w[SK_MAX_LAYERS], b[SK_MAX_LAYERS];    // lstm: [K][UPC][4], [UPC][4]; gru: [K][UPC][2], [UPC][2]; rnn: [K][UPC], [UPC]
  int wc[SK_MAX_LAYERS], bc[SK_MAX_LAYERS];  // gru candidate: [K][UPC], [UPC]
  int wo, bo;                                // [H][OC], [OC]
  int xh[SK_MAX_LAYERS];                     // [SB][Kp]: [x ; h] of the layer (Kp = K rounded up to 4)
  int rh;                                    // gru: [SB][H]  r*h of the layer being computed
  int hlast;                                 // [SB][H]  output of the last layer
  int lg, xin, pdf, flags;                   // [SB][NOUT], [SB][5 -> 8], [SB][M], 2*SB ints
  int dsc;                                   // [SB][8]: pen probabilities, pen index, the two normals of a sketch's draw
  int cst;                                   // lstm cell state: [L][UPC*SB], one slot per cell thread
  int bars;                                  // mbarriers (8 bytes each): per layer A (output), R (gru r*h), Hb (state); then LG
  int total;
};

__host__ __device__ inline int scl_kp(int K) { return (K + 3) & ~3; }

__host__ __device__ inline SclPlan scl_plan(int cell, int H, int L, int M, int NOUT, int UPC, int OC) {
  SclPlan p = {};
  int o = 0;
  const int ng = cell == 0 ? 4 : (cell == 1 ? 2 : 1);
  for (int l = 0; l < L; ++l) {
    const int K = (l == 0 ? 5 : H) + H;
    p.w[l] = o; o += K * UPC * ng;
    p.b[l] = o; o += UPC * ng;
    if (cell == 1) {
      p.wc[l] = o; o += K * UPC;
      p.bc[l] = o; o += UPC;
    }
  }
  p.wo = o; o += H * OC;
  p.bo = o; o += OC;
  o = (o + 3) & ~3;
  for (int l = 0; l < L; ++l) {
    p.xh[l] = o;
    o += SCL_SB * scl_kp((l == 0 ? 5 : H) + H);
  }
  if (cell == 1) { p.rh = o; o += SCL_SB * H; }
  p.hlast = o; o += SCL_SB * H;
  p.lg = o; o += SCL_SB * ((NOUT + 3) & ~3);
  p.xin = o; o += SCL_SB * 8;
  p.pdf = o; o += SCL_SB * ((M + 3) & ~3);
  p.flags = o; o += 2 * SCL_SB;
  p.dsc = o; o += SCL_SB * 8;
  p.cst = o; o += L * UPC * SCL_SB;
  o = (o + 3) & ~3;
  p.bars = o; o += 2 * (3 * SK_MAX_LAYERS + 1);
  p.total = o;
  return p;
}

__global__ void __launch_bounds__(SCL_THREADS, 1)
sample_cluster_kernel(const __grid_constant__ SclArgs a, const __grid_constant__ SclPlan p) {
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int H = a.H, L = a.L, M = a.M, NOUT = a.NOUT, UPC = a.UPC, OC = a.OC, cell = a.cell;
  const int rank = (int)scl_rank();
  const int sk_base = (int)scl_cluster_id() * SCL_SB;
  const int rows_valid = min(SCL_SB, a.n - sk_base);
  const int NOUTP = (NOUT + 3) & ~3, MP = (M + 3) & ~3;
  const int u0 = rank * UPC;
  const int ng = cell == 0 ? 4 : (cell == 1 ? 2 : 1);

  // ---- resident weights: my units' columns of every layer, my columns of the output layer ----
  for (int l = 0; l < L; ++l) {
    const int K = (l == 0 ? 5 : H) + H;
    const float *W = a.params + a.w0[l];
    const float *bias = a.params + a.b0[l];
    for (int i = tid; i < K * UPC * ng; i += SCL_THREADS) {  // (u fastest: runs of UPC contiguous floats in global memory)
      const int u = i % UPC, g = (i / UPC) % ng, k = i / (ng * UPC);
      sm[p.w[l] + (k * UPC + u) * ng + g] = __ldg(W + (size_t)k * (ng * H) + g * H + u0 + u);
    }
    for (int i = tid; i < UPC * ng; i += SCL_THREADS) sm[p.b[l] + i] = bias[(i % ng) * H + u0 + i / ng];
    if (cell == 1) {
      const float *Wc = a.params + a.w1[l];
      for (int i = tid; i < K * UPC; i += SCL_THREADS) sm[p.wc[l] + i] = Wc[(size_t)(i / UPC) * H + u0 + i % UPC];
      for (int i = tid; i < UPC; i += SCL_THREADS) sm[p.bc[l] + i] = a.params[a.b1[l] + u0 + i];
    }
  }
  for (int i = tid; i < H * OC; i += SCL_THREADS) {
    const int j = rank * OC + i % OC;
    sm[p.wo + i] = j < NOUT ? a.params[a.out_w + (size_t)(i / OC) * NOUT + j] : 0.f;
  }
  for (int i = tid; i < OC; i += SCL_THREADS) sm[p.bo + i] = (rank * OC + i < NOUT) ? a.params[a.out_b + rank * OC + i] : 0.f;
  // zero state (model.py:208), first input all zeros (model.py:205)
  for (int i = p.xh[0] + tid; i < p.total; i += SCL_THREADS) sm[i] = 0.f;
  __syncthreads();
  int *flags = reinterpret_cast<int *>(sm + p.flags);  // [SB] done, [SB] length
  if (tid < SCL_SB) { flags[tid] = tid < rows_valid ? 0 : 1; flags[SCL_SB + tid] = 0; }
  const uint32_t bars = (uint32_t)__cvta_generic_to_shared(sm + p.bars);
  auto bar_A = [&](int l) { return bars + 8u * (uint32_t)l; };
  auto bar_R = [&](int l) { return bars + 8u * (uint32_t)(SK_MAX_LAYERS + l); };
  auto bar_H = [&](int l) { return bars + 8u * (uint32_t)(2 * SK_MAX_LAYERS + l); };
  const uint32_t bar_LG = bars + 8u * (uint32_t)(3 * SK_MAX_LAYERS);
  if (tid == 0) {
    for (int i = 0; i < 3 * SK_MAX_LAYERS + 1; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u * (uint32_t)i));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const uint32_t xbytes = (uint32_t)(H * SCL_SB * 4), lgbytes = (uint32_t)(NOUT * SCL_SB * 4);
  float *xin = sm + p.xin;
  float *lg = sm + p.lg;
  scl_cluster_sync();  // nobody sends into a peer before it has zeroed its buffers and initialised its barriers

  // cell threads: one per (unit, sketch); the lstm cell state stays in its registers
  // lstm: the four gate chains of (unit, sketch) are split over two adjacent threads -- (i, j) and (f, o) -- so that the
  // dot products keep all four schedulers busy; the even thread then owns the cell update
  const int pidx = cell == 0 ? (tid >> 1) : tid;
  const int ghalf = cell == 0 ? (tid & 1) : 0;
  const bool dth = pidx < UPC * SCL_SB;  // takes part in the dot products
  const bool cth = dth && ghalf == 0;    // owns (unit, sketch): cell update and sends
  const int cu = pidx % UPC, cs = pidx / UPC;
  float *cst = sm + p.cst + pidx;  // + l * UPC * SB

  int last_step = 0;
  const bool dbg = a.debug && blockIdx.x == 0 && tid == 0;
  long long tk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long lapt = dbg ? clock64() : 0;
  for (int step = 0; step < a.steps; ++step) {
    const uint32_t par = (uint32_t)(step & 1);
    last_step = step;
    // the recurrent state every rank sent at the end of the previous step's layers has landed
    if (step > 0)
      for (int l = 0; l < L; ++l) scl_wait(bar_H(l), par ^ 1u);
    __syncthreads();  // nobody still polls last step's phases when this step's exchanges complete
    if (tid == 0) {   // this step's exchanges: expected bytes per barrier (data may arrive before or after this)
      for (int l = 0; l < L; ++l) {
        scl_arm(bar_A(l), xbytes);
        scl_arm(bar_H(l), xbytes);
        if (cell == 1) scl_arm(bar_R(l), xbytes);
      }
      scl_arm(bar_LG, lgbytes);
    }
    SCL_LAP(0);
    for (int l = 0; l < L; ++l) {
      const int I = (l == 0) ? 5 : H;
      const int K = I + H, Kp = scl_kp(K);
      float *xh = sm + p.xh[l];
      const bool hold = cth && xin[cs * 8 + 3] > 0.f;  // an eoc input holds the state (model.py:82-92,215)
      float hnew = 0.f, hkeep = 0.f;
      if (cell == 0) {
        const float *As = xh + cs * Kp;
        float acc[2] = {0.f, 0.f};
        if (dth) scl_dot<2>(As, sm + p.w[l] + 4 * cu + 2 * ghalf, 4 * UPC, K, acc);
        const float a2 = __shfl_down_sync(0xffffffffu, acc[0], 1), a3 = __shfl_down_sync(0xffffffffu, acc[1], 1);  // f, o from the odd lane
        if (cth) {
          const float a0 = acc[0], a1 = acc[1];
          const float4 bb = reinterpret_cast<const float4 *>(sm + p.b[l])[cu];
          const float c = cst[l * UPC * SCL_SB];
          const float ig = skm_sigmoid(SKM_ADD(a0, bb.x));
          const float jg = skm_tanh(SKM_ADD(a1, bb.y));
          const float fg = skm_sigmoid(SKM_ADD(SKM_ADD(a2, bb.z), 1.0f));
          const float og = skm_sigmoid(SKM_ADD(a3, bb.w));
          const float c2 = SKM_FMA(c, fg, SKM_MUL(ig, jg));
          hnew = SKM_MUL(skm_tanh(c2), og);
          hkeep = hold ? As[I + u0 + cu] : hnew;
          if (!hold) cst[l * UPC * SCL_SB] = c2;
        }
      } else if (cell == 2) {
        if (cth) {
          const float *As = xh + cs * Kp;
          float acc[1] = {0.f};
          scl_dot<1>(As, sm + p.w[l] + cu, UPC, K, acc);
          hnew = skm_tanh(SKM_ADD(acc[0], sm[p.b[l] + cu]));
          hkeep = hold ? As[I + u0 + cu] : hnew;
        }
      } else {
        float az = 0.f, hold_h = 0.f;
        float *rh = sm + p.rh;
        if (cth) {  // gates r, u on [x ; h]
          const float *As = xh + cs * Kp;
          float acc[2] = {0.f, 0.f};
          scl_dot<2>(As, sm + p.w[l] + 2 * cu, 2 * UPC, K, acc);
          float ar = acc[0];
          az = acc[1];
          const float2 bb = reinterpret_cast<const float2 *>(sm + p.b[l])[cu];
          hold_h = As[I + u0 + cu];
          ar = SKM_MUL(skm_sigmoid(SKM_ADD(ar, bb.x)), hold_h);  // r * h
          az = skm_sigmoid(SKM_ADD(az, bb.y));
          scl_send(rh + cs * H + u0 + cu, ar, bar_R(l));
        }
        scl_wait(bar_R(l), par);
        if (cth) {  // candidate on [x ; r*h]
          const float *Wc = sm + p.wc[l] + cu;
          const float *Ax = xh + cs * Kp;
          const float *Ar = rh + cs * H;
          float acc[1] = {0.f};
          if (l == 0) {  // 5 input columns, then r*h (the continuation of the same chain)
            for (int k = 0; k < I; ++k) acc[0] = SKM_FMA(Ax[k], Wc[(size_t)k * UPC], acc[0]);
          } else {
            scl_dot<1>(Ax, Wc, UPC, I, acc);
          }
          scl_dot<1>(Ar, Wc + (size_t)I * UPC, UPC, H, acc);
          const float cand = skm_tanh(SKM_ADD(acc[0], sm[p.bc[l] + cu]));
          hnew = SKM_FMA(az, hold_h, SKM_MUL(SKM_SUB(1.0f, az), cand));
          hkeep = hold ? hold_h : hnew;
        }
      }
      // the layer's output: input of the next layer (nobody reads that buffer now), or of the output projection
      if (cth) {
        if (l + 1 < L) scl_send(sm + p.xh[l + 1] + cs * scl_kp(2 * H) + u0 + cu, hnew, bar_A(l));
        else scl_send(sm + p.hlast + cs * H + u0 + cu, hnew, bar_A(l));
      }
      SCL_LAP(1);
      scl_wait(bar_A(l), par);  // every rank's output is here, so every rank is done reading this layer's [x ; h]
      SCL_LAP(2);
      if (cth) scl_send(xh + cs * Kp + I + u0 + cu, hkeep, bar_H(l));  // waited for at the top of the next step
      SCL_LAP(3);
    }
    // ---- output projection (model.py:98-99): my OC columns for the SB sketches ----
    if (tid < OC * SCL_SB) {
      const int j = tid % OC, s = tid / OC;
      const float *Wo = sm + p.wo + j;
      const float *hl = sm + p.hlast + s * H;
      float acc[1] = {0.f};
      scl_dot<1>(hl, Wo, OC, H, acc);
      if (rank * OC + j < NOUT) scl_send(lg + s * NOUTP + rank * OC + j, SKM_ADD(acc[0], sm[p.bo + j]), bar_LG);
    }
    SCL_LAP(4);
    scl_wait(bar_LG, par);
    SCL_LAP(5);
    // ---- per-sketch draw and feedback (model.py:219-261): warp s draws sketch s; lanes share the transcendentals,
    //      lane 0 does every reduction sequentially, in the order the reference sums ----
    // warp r (< 4): mixture softmax, temperature and the component draw of sketch r; warp r + 4: its pen state and the
    // two standard normals (neither depends on the chosen component), handed over through `dsc`
    const int r = warp & 3;
    const bool dact = r < rows_valid && !flags[r];
    const int sk = sk_base + r;
    const float *z = lg + r * NOUTP;
    float *pd = sm + p.pdf + r * MP;
    float *dsc = sm + p.dsc + r * 8;
    float u4[4] = {0.f, 0.f, 0.f, 0.f};
    int idx = 0;
    if (dact) {
      if (a.uniforms != nullptr) {
        const float *up = a.uniforms + ((size_t)sk * a.steps + step) * 4;
        u4[0] = up[0]; u4[1] = up[1]; u4[2] = up[2]; u4[3] = up[3];
      } else {
        uint32_t rr[4];
        skm_philox4x32((uint32_t)(a.sketch0 + sk), (uint32_t)step, 0u, 0x5a3711eu, (uint32_t)a.seed,
                       (uint32_t)(a.seed >> 32), rr);
        for (int q = 0; q < 4; ++q) u4[q] = skm_u01(rr[q]);
      }
    }
    if (dact && warp < 4) {
      // softmax over the M mixture logits (get_mixture_coef, model.py:152-156)
      float mx = -INFINITY;
      for (int k = lane; k < M; k += 32) mx = fmaxf(mx, z[3 + k]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      for (int k = lane; k < M; k += 32) pd[k] = skm_exp(SKM_SUB(z[3 + k], mx));
      __syncwarp();
      float inv = 0.f;
      if (lane == 0) {
        float sum = 0.f;
        for (int k = 0; k < M; ++k) sum = SKM_ADD(sum, pd[k]);
        inv = SKM_DIV(1.0f, sum);
      }
      inv = __shfl_sync(0xffffffffu, inv, 0);
      for (int k = lane; k < M; k += 32) pd[k] = SKM_MUL(pd[k], inv);
      if (step > 1) {  // temperature only from the third point (model.py:220)
        float m2 = -INFINITY;
        for (int k = lane; k < M; k += 32) {
          const float v = SKM_DIV(skm_log(pd[k]), a.temp_mix);
          pd[k] = v;
          m2 = fmaxf(m2, v);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
        for (int k = lane; k < M; k += 32) pd[k] = skm_exp(SKM_SUB(pd[k], m2));
      }
      __syncwarp();
      float s2 = 0.f;
      if (lane == 0)
        for (int k = 0; k < M; ++k) s2 = SKM_ADD(s2, pd[k]);
      s2 = __shfl_sync(0xffffffffu, s2, 0);
      for (int k = lane; k < M; k += 32) pd[k] = SKM_DIV(pd[k], s2);
      __syncwarp();
      if (lane == 0) {
        idx = scl_pick_idx(pd, M, u4[0]);
        if (idx < 0) idx = M - 1;  // python negative index (model.py:247)
      }
    } else if (dact && lane == 0) {
      // pen state: plain softmax, the pen temperature is never applied (model.py:228-233)
      const float pm = fmaxf(z[0], fmaxf(z[1], z[2]));
      float pe[3];
      pe[0] = skm_exp(SKM_SUB(z[0], pm));
      pe[1] = skm_exp(SKM_SUB(z[1], pm));
      pe[2] = skm_exp(SKM_SUB(z[2], pm));
      const float ps = SKM_ADD(SKM_ADD(pe[0], pe[1]), pe[2]);
      pe[0] = SKM_DIV(pe[0], ps); pe[1] = SKM_DIV(pe[1], ps); pe[2] = SKM_DIV(pe[2], ps);
      const int pen_idx = scl_pick_idx(pe, 3, u4[1]);
      float g1, g2;
      skm_box_muller(u4[2], u4[3], &g1, &g2);
      dsc[0] = pe[0]; dsc[1] = pe[1]; dsc[2] = pe[2]; dsc[3] = (float)pen_idx; dsc[4] = g1; dsc[5] = g2;
    }
    __syncthreads();
    if (dact && warp < 4) {
      if (lane == 0) {
        const float pe[3] = {dsc[0], dsc[1], dsc[2]};
        const int pen_idx = (int)dsc[3];
        const float g1 = dsc[4], g2 = dsc[5];
        float eos = 0.f, eoc = 0.f, cont = 0.f;
        if (pen_idx == 0) eos = 1.f;
        else if (pen_idx == 1) eoc = 1.f;
        else cont = 1.f;
        // bivariate normal of the chosen component (model.py:199-203,247); no temperature
        const float mu1 = z[3 + M + idx], mu2 = z[3 + 2 * M + idx];
        const float sg1 = skm_exp(z[3 + 3 * M + idx]), sg2 = skm_exp(z[3 + 4 * M + idx]);
        const float rho = skm_tanh(z[3 + 5 * M + idx]);
        const float x1 = SKM_FMA(sg1, g1, mu1);
        const float q = SKM_SQRT(SKM_SUB(1.0f, SKM_MUL(rho, rho)));
        const float tt = SKM_FMA(q, g2, SKM_MUL(rho, g1));
        const float x2 = SKM_FMA(sg2, tt, mu2);
        if (rank == 0) {
          float *so = a.strokes + ((size_t)sk * a.steps + step) * 5;
          so[0] = SKM_MUL(x1, a.data_scale);  // model.py:263
          so[1] = SKM_MUL(x2, a.data_scale);
          so[2] = eos; so[3] = eoc; so[4] = cont;
          if (a.mixp != nullptr) {
            float *mo = a.mixp + ((size_t)sk * a.steps + step) * NOUT;
            mo[0] = pe[0]; mo[1] = pe[1]; mo[2] = pe[2];
          }
        }
        // feedback (model.py:259-261): next step's input, also as the x part of layer 0's operand
        xin[r * 8 + 0] = x1; xin[r * 8 + 1] = x2; xin[r * 8 + 2] = eos; xin[r * 8 + 3] = eoc; xin[r * 8 + 4] = cont;
        float *x0 = sm + p.xh[0] + r * scl_kp(5 + H);
        x0[0] = x1; x0[1] = x2; x0[2] = eos; x0[3] = eoc; x0[4] = cont;
        flags[SCL_SB + r] = step + 1;
        if (a.stop_if_eoc && eoc > 0.f) flags[r] = 1;  // model.py:255-257
      }
      if (rank == 0 && a.mixp != nullptr) {
        float *mo = a.mixp + ((size_t)sk * a.steps + step) * NOUT;
        for (int k = lane; k < M; k += 32) {
          // the stray `pi_pdf /= temp_pen` of model.py:230 shows up in the returned params
          mo[3 + k] = (step > 1) ? SKM_DIV(pd[k], a.temp_pen) : pd[k];
          mo[3 + M + k] = z[3 + M + k];
          mo[3 + 2 * M + k] = z[3 + 2 * M + k];
          mo[3 + 3 * M + k] = skm_exp(z[3 + 3 * M + k]);
          mo[3 + 4 * M + k] = skm_exp(z[3 + 4 * M + k]);
          mo[3 + 5 * M + k] = skm_tanh(z[3 + 5 * M + k]);
        }
      }
    }
    SCL_LAP(6);
    __syncthreads();
    SCL_LAP(7);
    int all_done = 1;
    for (int r = 0; r < SCL_SB; ++r) all_done &= flags[r];
    if (all_done) break;  // every rank computed the same flags
  }
  if (rank == 0 && tid < rows_valid) a.lengths[sk_base + tid] = flags[SCL_SB + tid];
  if (dbg)
    printf("[sampler_cl] cycles/step: top wait+arm %lld, layers compute+send %lld, wait outputs %lld, state send %lld, output projection %lld, "
           "wait logits %lld, draw (warp 0) %lld, end barrier %lld\n", tk[0] / (last_step + 1), tk[1] / (last_step + 1), tk[2] / (last_step + 1),
           tk[3] / (last_step + 1), tk[4] / (last_step + 1), tk[5] / (last_step + 1), tk[6] / (last_step + 1), tk[7] / (last_step + 1));
  // nobody leaves while a peer may still send into it: the last step's state exchange has landed here, and once every
  // rank can say so nothing is in flight anywhere in the cluster
  for (int l = 0; l < L; ++l) scl_wait(bar_H(l), (uint32_t)(last_step & 1));
  scl_cluster_sync();
}

}  // namespace

// 0: drawn here; -1: not eligible (the caller falls back to the streaming kernel of sampler.cu); > 0: error
int sk_sample_cluster(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen, int stop_if_eoc,
                      const float *uniforms, uint64_t seed, int sketch0, float *strokes, int *lengths, float *mixp) {
  {
    const char *e = getenv("SKRNN_SAMPLE_CLUSTER");  // read per call: "0" keeps every call on the streaming kernel
    if (e && e[0] == '0') return -1;
  }
  const int H = h->H;
  if (H % SCL_CS != 0 || h->M > 64) return -1;
  const int UPC = H / SCL_CS, OC = (h->NOUT + SCL_CS - 1) / SCL_CS;
  if (2 * UPC * SCL_SB > SCL_THREADS || OC * SCL_SB > SCL_THREADS) return -1;
  const SclPlan p = scl_plan(h->cfg.cell, H, h->L, h->M, h->NOUT, UPC, OC);
  const size_t smem = (size_t)p.total * sizeof(float);
  if (smem > 227 * 1024) return -1;  // the stack does not fit a 16-CTA cluster's shared memory
  SclArgs a = {};
  a.n = n; a.steps = steps; a.H = H; a.L = h->L; a.M = h->M; a.NOUT = h->NOUT; a.cell = h->cfg.cell;
  a.UPC = UPC; a.OC = OC;
  a.temp_mix = temp_mix; a.temp_pen = temp_pen; a.data_scale = h->cfg.data_scale;
  a.stop_if_eoc = stop_if_eoc;
  a.uniforms = uniforms; a.seed = seed; a.sketch0 = sketch0;
  a.params = h->params;
  for (int l = 0; l < h->L; ++l) {
    a.w0[l] = h->lp[l].w[0]; a.b0[l] = h->lp[l].b[0];
    a.w1[l] = h->lp[l].w[1]; a.b1[l] = h->lp[l].b[1];
  }
  a.out_w = h->off_out_w; a.out_b = h->off_out_b;
  a.strokes = strokes; a.lengths = lengths; a.mixp = mixp;
  {
    const char *e = getenv("SKRNN_TC_DEBUG");
    a.debug = (e && e[0] == '1') ? 1 : 0;
  }
  SK_CUDA(cudaFuncSetAttribute(sample_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SK_CUDA(cudaFuncSetAttribute(sample_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t lc = {};
  cudaLaunchAttribute at[1];
  lc.gridDim = dim3((unsigned)(SCL_CS * ((n + SCL_SB - 1) / SCL_SB)));
  lc.blockDim = dim3(SCL_THREADS);
  lc.dynamicSmemBytes = smem;
  lc.stream = st;
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = SCL_CS;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  int nclusters = 0;
  if (cudaOccupancyMaxActiveClusters(&nclusters, sample_cluster_kernel, &lc) != cudaSuccess || nclusters < 1) {
    cudaGetLastError();
    return -1;
  }
  {
    // Worth it up to about four waves of clusters (measured on the kanji model: ~4.1 ms per wave of 300 steps against
    // ~24 ms for the streaming kernel, which takes that long for anything from 1 to ~600 sketches)
    const char *e = getenv("SKRNN_SAMPLE_CLUSTER_MAX");
    const int n_max = e ? atoi(e) : 4 * SCL_SB * nclusters;
    if (n > n_max) return -1;
  }
  SkLaunch scope(h, st, SK_CAT_SAMPLER);
  if (cudaLaunchKernelEx(&lc, sample_cluster_kernel, a, p) != cudaSuccess) {
    cudaGetLastError();  // the cluster could not be placed after all: the streaming kernel takes the call
    return -1;
  }
  return 0;
}
