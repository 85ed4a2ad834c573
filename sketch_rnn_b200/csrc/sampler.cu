// sampler.cu -- persistent batched sampler, reference Model.sample (model.py:187-264).
//
// One launch runs every step of every sketch: a CTA keeps SB sketches in flight, runs the
// single-timestep forward of the cell stack and the MDN head for them, draws the next point
// (temperature, categorical draws, bivariate normal) in-kernel and feeds (dx, dy, eos, eoc, cont)
// back as the next input -- no host round trip per point (the reference pays one sess.run per
// point, model.py:215-217).
//
// This is the reference-precision variant.  Every output element is one sequential
// fused-multiply-add chain over k = 0..K-1 of [x ; h] * W followed by + bias, and all
// transcendental functions come from skrnn_math.h, so the result is bit-identical to the CPU
// restatement in oracle/sampler_ref.c given the same uniform draws.
//
// Thread u of a CTA owns hidden unit u of every layer (blockDim >= H), so all gates of a unit
// meet in one thread and the cell update needs no exchange; the [x ; h] operand of the SB
// sketches is staged k-major in shared memory and broadcast to all threads.  The weights are the
// other operand: every thread copies the columns it owns a few k rows ahead into a private ring in
// shared memory with cp.async (no registers held across the L2 latency, no barrier: a thread only reads
// what it copied itself).  Matrices with several gate columns per unit are read from a gate-interleaved
// copy ([k][unit][gate], packed once per call) so that a k row of a thread is one 8- or 16-byte copy.
#include <stdlib.h>

#include "common.cuh"
#include "skrnn_math.h"

namespace {

struct SampArgs {
  int n, steps, H, L, M, NOUT, cell;
  float temp_mix, temp_pen, data_scale;
  int stop_if_eoc;
  const float *uniforms;  // [n,steps,4] or null
  uint64_t seed;
  int sketch0;
  const float *params;
  const float *wpk[SK_MAX_LAYERS];  // lstm / gru gate matrix of layer l as [k][unit][gate]
  int64_t w0[SK_MAX_LAYERS], b0[SK_MAX_LAYERS], w1[SK_MAX_LAYERS], b1[SK_MAX_LAYERS];
  int64_t out_w, out_b;
  float *state;           // [n, L, S] zero-initialised
  float *strokes;         // [n, steps, 5]
  int *lengths;           // [n]
  float *mixp;            // [n, steps, NOUT] or null
};

// ---- weight stream: per-thread cp.async ring ----
constexpr int SK_KG = 4;     // k rows per copy group
constexpr int SK_NSLOT = 3;  // groups in the ring: two in flight while one is consumed

template <int BYTES>
__device__ __forceinline__ void cp_async_n(uint32_t dst, const float *src) {
  if constexpr (BYTES == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(dst), "l"(src), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Two fused multiply-adds per instruction (FFMA2 on sm_100: c.x = fma(a.x, w, c.x), c.y = fma(a.y, w, c.y), each rounded
// once to nearest like __fmaf_rn; ptxas folds the duplicated w into a scalar operand).  The streaming loop is bound by
// issue slots, and this halves the slots its arithmetic takes.
__device__ __forceinline__ void fma2_rn(float2 &c, const float2 a, const float w) {
  unsigned long long cc = *reinterpret_cast<unsigned long long *>(&c), ww, d;
  const unsigned long long aa = *reinterpret_cast<const unsigned long long *>(&a);
  asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "f"(w));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(aa), "l"(ww), "l"(cc));
  c = *reinterpret_cast<float2 *>(&d);
}

// acc[r][g] = sum over k (in order, one fused multiply-add chain per element) of AT[k][r] * w[k][g], where w[k][0..NG)
// are the NG consecutive floats at wsrc + k * pitch (this thread's columns of row k).  ring: this thread's slice of the
// ring, element (slot, kk) at ring + ((slot * SK_KG + kk) * NT) * NG floats.
template <int SB, int SP, int NG>
__device__ __forceinline__ void stream_dot(const float *__restrict__ wsrc, size_t pitch, int K, const float *__restrict__ AT,
                                           float *__restrict__ ring, int NT, float (&acc)[SB][NG]) {
  const uint32_t row_bytes = (uint32_t)(NT * NG * 4), slot_bytes = SK_KG * row_bytes;
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring), ring_e = ring_s + SK_NSLOT * slot_bytes;
  constexpr int SBH = (SB + 1) / 2;  // sketches in pairs (SP is a multiple of 4, so an odd SB reads a pad lane)
  float2 a2[SBH][NG];
#pragma unroll
  for (int p = 0; p < SBH; ++p)
#pragma unroll
    for (int g = 0; g < NG; ++g) a2[p][g] = make_float2(acc[2 * p][g], 2 * p + 1 < SB ? acc[2 * p + 1][g] : 0.f);
  const int nfull = K / SK_KG, ktail = K - nfull * SK_KG;
  const size_t gstep = (size_t)SK_KG * pitch;
  const float *src = wsrc;  // first row of the next group to copy
  uint32_t si = ring_s;     // its slot
  int gi = 0;
  auto issue = [&]() {
    if (gi < nfull) {
#pragma unroll
      for (int kk = 0; kk < SK_KG; ++kk) cp_async_n<4 * NG>(si + kk * row_bytes, src + (size_t)kk * pitch);
    } else if (gi == nfull) {
      for (int kk = 0; kk < ktail; ++kk) cp_async_n<4 * NG>(si + kk * row_bytes, src + (size_t)kk * pitch);
    }
    cp_async_commit();  // one group per call, empty past the end, so the waits below count uniformly
    ++gi;
    src += gstep;
    si += slot_bytes;
    if (si == ring_e) si = ring_s;
  };
  auto row = [&](const float *rs, const float *at) {
    float w[NG];
    if constexpr (NG == 4) {
      const float4 v = *reinterpret_cast<const float4 *>(rs);
      w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    } else if constexpr (NG == 2) {
      const float2 v = *reinterpret_cast<const float2 *>(rs);
      w[0] = v.x; w[1] = v.y;
    } else {
      w[0] = rs[0];
    }
#pragma unroll
    for (int p = 0; p < SBH; ++p) {
      const float2 av = *reinterpret_cast<const float2 *>(at + 2 * p);
#pragma unroll
      for (int g = 0; g < NG; ++g) fma2_rn(a2[p][g], av, w[g]);
    }
  };
#pragma unroll
  for (int i = 0; i < SK_NSLOT - 1; ++i) issue();
  const uint32_t row_f = row_bytes / 4, slot_f = slot_bytes / 4;
  const float *rs = ring, *const ring_end = ring + SK_NSLOT * slot_f;
  const float *at = AT;
#pragma unroll 1
  for (int grp = 0; grp < nfull; ++grp) {
    issue();
    cp_async_wait<SK_NSLOT - 1>();
#pragma unroll
    for (int kk = 0; kk < SK_KG; ++kk) row(rs + kk * row_f, at + kk * SP);
    at += SK_KG * SP;
    rs += slot_f;
    if (rs == ring_end) rs = ring;
  }
  cp_async_wait<0>();
  for (int kk = 0; kk < ktail; ++kk) row(rs + kk * row_f, at + kk * SP);
#pragma unroll
  for (int r = 0; r < SB; ++r)
#pragma unroll
    for (int g = 0; g < NG; ++g) acc[r][g] = (r & 1) ? a2[r >> 1][g].y : a2[r >> 1][g].x;
}

// gate matrix [K][NG*H] (TF column order gate*H + unit) -> [K][H][NG]
__global__ void sample_pack_kernel(const float *__restrict__ src, float *__restrict__ dst, int K, int H, int NG) {
  const int64_t total = (int64_t)K * H * NG;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % NG), u = (int)((i / NG) % H);
    const int64_t k = i / ((int64_t)NG * H);
    dst[i] = src[k * NG * H + (int64_t)g * H + u];
  }
}

// first index with running (fp64) sum >= u, else -1   (get_pi_idx, model.py:189-197)
__device__ __forceinline__ int pick_idx(const float *pdf, int n, float u) {
  double acc = 0.0;
  const double ud = (double)u;
  for (int k = 0; k < n; ++k) {
    acc = SKM_DADD(acc, (double)pdf[k]);
    if (acc >= ud) return k;
  }
  return -1;
}

template <int SB>
__global__ void __launch_bounds__(512) sample_kernel(SampArgs a) {
  constexpr int SP = (SB + 3) & ~3;  // row pitch of the k-major operand tiles (16-byte rows)
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, NT = blockDim.x;
  const int H = a.H, L = a.L, M = a.M, NOUT = a.NOUT;
  const int S = (a.cell == 0) ? 2 * H : H;
  const int hoff = (a.cell == 0) ? H : 0;
  const int sk_base = blockIdx.x * SB;
  const int rows_valid = min(SB, a.n - sk_base);
  const bool act = tid < H;
  const int u = tid;

  float *AT = smem;                       // [2H rows k][SP]: [x ; h] of the current layer
  float *nxt = AT + (size_t)2 * H * SP;   // [H][SP]: output of the current layer
  float *lg = nxt + (size_t)H * SP;       // [SB][NOUT] logits
  float *xin = lg + (size_t)SB * NOUT;    // [SB][5] current input point
  float *pdf = xin + SB * 5;              // [SB][M] scratch for the per-sketch draw
  int *flags = reinterpret_cast<int *>(pdf + (size_t)SB * M);  // [SB] done, [SB] length
  // weight ring (16-byte aligned; an index into the same array, so that the reads stay shared-memory loads)
  float *ring = smem + ((3 * H * SP + SB * NOUT + SB * 5 + SB * M + 2 * SB + 3) & ~3);

  for (int i = tid; i < SB * 5; i += NT) xin[i] = 0.f;  // first input is all zeros (model.py:205)
  for (int i = tid; i < SB; i += NT) { flags[i] = (i < rows_valid) ? 0 : 1; flags[SB + i] = 0; }
  __syncthreads();

  for (int step = 0; step < a.steps; ++step) {
    for (int l = 0; l < L; ++l) {
      const int I = (l == 0) ? 5 : H;
      const int K = I + H;
      // ---- stage [x ; h] ----
      if (l == 0) {
        for (int i = tid; i < 5 * SB; i += NT) { int k = i / SB, r = i % SB; AT[k * SP + r] = xin[r * 5 + k]; }
      } else {
        for (int i = tid; i < H * SP; i += NT) AT[i] = nxt[i];
      }
      __syncthreads();  // nxt fully copied before anyone overwrites it below
      for (int i = tid; i < H * SB; i += NT) {  // all of a thread's copies in flight at once
        int r = i / H, uu = i % H;
        float *dst = AT + (size_t)(I + uu) * SP + r;
        if (r < rows_valid)
          cp_async_n<4>((uint32_t)__cvta_generic_to_shared(dst), a.state + ((size_t)(sk_base + r) * L + l) * S + hoff + uu);
        else
          *dst = 0.f;
      }
      cp_async_commit();
      cp_async_wait<0>();
      __syncthreads();
      const float *W = a.params + a.w0[l];
      const float *bias = a.params + a.b0[l];
      if (a.cell == 0) {
        if (act) {
          float acc[SB][4];
#pragma unroll
          for (int r = 0; r < SB; ++r) { acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.f; }
          stream_dot<SB, SP, 4>(a.wpk[l] + 4 * u, (size_t)4 * H, K, AT, ring + 4 * tid, NT, acc);
          const float bi = bias[u], bj = bias[H + u], bf = bias[2 * H + u], bo = bias[3 * H + u];
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            if (r >= rows_valid) { nxt[(size_t)u * SP + r] = 0.f; continue; }
            float *st = a.state + ((size_t)(sk_base + r) * L + l) * S;
            float c = st[u];
            float ig = skm_sigmoid(SKM_ADD(acc[r][0], bi));
            float jg = skm_tanh(SKM_ADD(acc[r][1], bj));
            float fg = skm_sigmoid(SKM_ADD(SKM_ADD(acc[r][2], bf), 1.0f));
            float og = skm_sigmoid(SKM_ADD(acc[r][3], bo));
            float c2 = SKM_FMA(c, fg, SKM_MUL(ig, jg));
            float h2 = SKM_MUL(skm_tanh(c2), og);
            nxt[(size_t)u * SP + r] = h2;
            // an eoc input holds the state: initial_state is prev_state (model.py:82-92,215)
            if (!(xin[r * 5 + 3] > 0.f)) { st[u] = c2; st[H + u] = h2; }
          }
        }
      } else if (a.cell == 2) {
        if (act) {
          float acc[SB][1];
#pragma unroll
          for (int r = 0; r < SB; ++r) acc[r][0] = 0.f;
          stream_dot<SB, SP, 1>(W + u, (size_t)H, K, AT, ring + tid, NT, acc);
          const float bb = bias[u];
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            if (r >= rows_valid) { nxt[(size_t)u * SP + r] = 0.f; continue; }
            float h2 = skm_tanh(SKM_ADD(acc[r][0], bb));
            nxt[(size_t)u * SP + r] = h2;
            if (!(xin[r * 5 + 3] > 0.f)) a.state[((size_t)(sk_base + r) * L + l) * S + u] = h2;
          }
        }
      } else {
        float az[SB], hold[SB], ar[SB];
        if (act) {  // gates r, u on [x ; h]
          float g2[SB][2];
#pragma unroll
          for (int r = 0; r < SB; ++r) { g2[r][0] = 0.f; g2[r][1] = 0.f; hold[r] = AT[(size_t)(I + u) * SP + r]; }
          stream_dot<SB, SP, 2>(a.wpk[l] + 2 * u, (size_t)2 * H, K, AT, ring + 2 * tid, NT, g2);
          const float br = bias[u], bz = bias[H + u];
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            ar[r] = SKM_MUL(skm_sigmoid(SKM_ADD(g2[r][0], br)), hold[r]);  // r * h
            az[r] = skm_sigmoid(SKM_ADD(g2[r][1], bz));
          }
        }
        __syncthreads();  // every thread is done reading h from AT
        if (act) {
#pragma unroll
          for (int r = 0; r < SB; ++r) AT[(size_t)(I + u) * SP + r] = ar[r];
        }
        __syncthreads();
        if (act) {  // candidate on [x ; r*h]
          const float *Wc = a.params + a.w1[l];
          const float bc = a.params[a.b1[l] + u];
          float ac[SB][1];
#pragma unroll
          for (int r = 0; r < SB; ++r) ac[r][0] = 0.f;
          stream_dot<SB, SP, 1>(Wc + u, (size_t)H, K, AT, ring + tid, NT, ac);
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            if (r >= rows_valid) { nxt[(size_t)u * SP + r] = 0.f; continue; }
            float cand = skm_tanh(SKM_ADD(ac[r][0], bc));
            float h2 = SKM_FMA(az[r], hold[r], SKM_MUL(SKM_SUB(1.0f, az[r]), cand));
            nxt[(size_t)u * SP + r] = h2;
            if (!(xin[r * 5 + 3] > 0.f)) a.state[((size_t)(sk_base + r) * L + l) * S + u] = h2;
          }
        }
      }
      __syncthreads();
    }
    // ---- output projection (model.py:98-99) ----
    {
      const float *Wo = a.params + a.out_w;
      constexpr int SBA = SB >= 4 ? (((SB / 2) + 1) & ~1) : SB, SBB = SB - SBA;  // sketches of the two thread halves
      if (SBB > 0 && 2 * NOUT <= NT) {
        // NOUT < blockDim / 2: the two halves of the CTA take SBA and SBB of the sketches, so that every warp has work
        const int half = tid >= NT / 2 ? 1 : 0, j = tid - half * (NT / 2);
        if (j < NOUT) {
          const float bb = a.params[a.out_b + j];
          if (half == 0) {
            float acc[SBA][1];
#pragma unroll
            for (int r = 0; r < SBA; ++r) acc[r][0] = 0.f;
            stream_dot<SBA, SP, 1>(Wo + j, (size_t)NOUT, H, nxt, ring + tid, NT, acc);
#pragma unroll
            for (int r = 0; r < SBA; ++r) lg[(size_t)r * NOUT + j] = SKM_ADD(acc[r][0], bb);
          } else {
            constexpr int SBB1 = SBB > 0 ? SBB : 1;
            float acc[SBB1][1];
#pragma unroll
            for (int r = 0; r < SBB1; ++r) acc[r][0] = 0.f;
            stream_dot<SBB1, SP, 1>(Wo + j, (size_t)NOUT, H, nxt + SBA, ring + tid, NT, acc);
#pragma unroll
            for (int r = 0; r < SBB1; ++r) lg[(size_t)(SBA + r) * NOUT + j] = SKM_ADD(acc[r][0], bb);
          }
        }
      } else {
        for (int j = tid; j < NOUT; j += NT) {
          float acc[SB][1];
#pragma unroll
          for (int r = 0; r < SB; ++r) acc[r][0] = 0.f;
          stream_dot<SB, SP, 1>(Wo + j, (size_t)NOUT, H, nxt, ring + tid, NT, acc);
          const float bb = a.params[a.out_b + j];
#pragma unroll
          for (int r = 0; r < SB; ++r) lg[(size_t)r * NOUT + j] = SKM_ADD(acc[r][0], bb);
        }
      }
    }
    __syncthreads();
    // ---- per-sketch draw and feedback (model.py:219-261): one half warp per sketch.  The 16 lanes share the
    //      transcendentals of the mixture weights; lane 0 of the half does every sum sequentially, in the order the
    //      reference sums, and the scalar rest (pen state, the two normals, the chosen component's point) ----
    {
      const int r = tid >> 4, hl = tid & 15;
      const unsigned hmask = 0xffffu << (tid & 16);
      if (r < rows_valid && !flags[r]) {
        const int sk = sk_base + r;
        const float *z = lg + (size_t)r * NOUT;
        float *p = pdf + (size_t)r * M;
        // softmax over the M mixture logits (get_mixture_coef, model.py:152-156)
        float mx = -INFINITY;
        for (int k = hl; k < M; k += 16) mx = fmaxf(mx, z[3 + k]);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(hmask, mx, o, 16));
        for (int k = hl; k < M; k += 16) p[k] = skm_exp(SKM_SUB(z[3 + k], mx));
        __syncwarp(hmask);
        float inv = 0.f;
        if (hl == 0) {
          float sum = 0.f;
          for (int k = 0; k < M; ++k) sum = SKM_ADD(sum, p[k]);
          inv = SKM_DIV(1.0f, sum);
        }
        inv = __shfl_sync(hmask, inv, 0, 16);
        for (int k = hl; k < M; k += 16) p[k] = SKM_MUL(p[k], inv);
        if (step > 1) {  // temperature only from the third point (model.py:220)
          float m2 = -INFINITY;
          for (int k = hl; k < M; k += 16) {
            const float v = SKM_DIV(skm_log(p[k]), a.temp_mix);
            p[k] = v;
            m2 = fmaxf(m2, v);
          }
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) m2 = fmaxf(m2, __shfl_xor_sync(hmask, m2, o, 16));
          for (int k = hl; k < M; k += 16) p[k] = skm_exp(SKM_SUB(p[k], m2));
        }
        __syncwarp(hmask);
        float s2 = 0.f;
        if (hl == 0)
          for (int k = 0; k < M; ++k) s2 = SKM_ADD(s2, p[k]);
        s2 = __shfl_sync(hmask, s2, 0, 16);
        for (int k = hl; k < M; k += 16) p[k] = SKM_DIV(p[k], s2);
        __syncwarp(hmask);
        if (a.mixp != nullptr) {
          float *mo = a.mixp + ((size_t)sk * a.steps + step) * NOUT;
          for (int k = hl; k < M; k += 16) {
            // the stray `pi_pdf /= temp_pen` of model.py:230 shows up in the returned params
            mo[3 + k] = (step > 1) ? SKM_DIV(p[k], a.temp_pen) : p[k];
            mo[3 + M + k] = z[3 + M + k];
            mo[3 + 2 * M + k] = z[3 + 2 * M + k];
            mo[3 + 3 * M + k] = skm_exp(z[3 + 3 * M + k]);
            mo[3 + 4 * M + k] = skm_exp(z[3 + 4 * M + k]);
            mo[3 + 5 * M + k] = skm_tanh(z[3 + 5 * M + k]);
          }
        }
        if (hl == 0) {
          float u4[4];
          if (a.uniforms != nullptr) {
            const float *up = a.uniforms + ((size_t)sk * a.steps + step) * 4;
            u4[0] = up[0]; u4[1] = up[1]; u4[2] = up[2]; u4[3] = up[3];
          } else {
            uint32_t rr[4];
            skm_philox4x32((uint32_t)(a.sketch0 + sk), (uint32_t)step, 0u, 0x5a3711eu, (uint32_t)a.seed,
                           (uint32_t)(a.seed >> 32), rr);
            for (int q = 0; q < 4; ++q) u4[q] = skm_u01(rr[q]);
          }
          int idx = pick_idx(p, M, u4[0]);
          if (idx < 0) idx = M - 1;  // python negative index (model.py:247)
          // pen state: plain softmax, the pen temperature is never applied (model.py:228-233)
          float pm = fmaxf(z[0], fmaxf(z[1], z[2]));
          float pe[3];
          pe[0] = skm_exp(SKM_SUB(z[0], pm));
          pe[1] = skm_exp(SKM_SUB(z[1], pm));
          pe[2] = skm_exp(SKM_SUB(z[2], pm));
          float ps = SKM_ADD(SKM_ADD(pe[0], pe[1]), pe[2]);
          pe[0] = SKM_DIV(pe[0], ps); pe[1] = SKM_DIV(pe[1], ps); pe[2] = SKM_DIV(pe[2], ps);
          int pen_idx = pick_idx(pe, 3, u4[1]);
          float eos = 0.f, eoc = 0.f, cont = 0.f;
          if (pen_idx == 0) eos = 1.f;
          else if (pen_idx == 1) eoc = 1.f;
          else cont = 1.f;
          // bivariate normal of the chosen component (model.py:199-203,247); no temperature
          float mu1 = z[3 + M + idx], mu2 = z[3 + 2 * M + idx];
          float sg1 = skm_exp(z[3 + 3 * M + idx]), sg2 = skm_exp(z[3 + 4 * M + idx]);
          float rho = skm_tanh(z[3 + 5 * M + idx]);
          float g1, g2;
          skm_box_muller(u4[2], u4[3], &g1, &g2);
          float x1 = SKM_FMA(sg1, g1, mu1);
          float q = SKM_SQRT(SKM_SUB(1.0f, SKM_MUL(rho, rho)));
          float tt = SKM_FMA(q, g2, SKM_MUL(rho, g1));
          float x2 = SKM_FMA(sg2, tt, mu2);
          float *so = a.strokes + ((size_t)sk * a.steps + step) * 5;
          so[0] = SKM_MUL(x1, a.data_scale);  // model.py:263
          so[1] = SKM_MUL(x2, a.data_scale);
          so[2] = eos; so[3] = eoc; so[4] = cont;
          if (a.mixp != nullptr) {
            float *mo = a.mixp + ((size_t)sk * a.steps + step) * NOUT;
            mo[0] = pe[0]; mo[1] = pe[1]; mo[2] = pe[2];
          }
          // feedback (model.py:259-261): nobody reads xin between the barrier above and the next one
          xin[r * 5 + 0] = x1; xin[r * 5 + 1] = x2; xin[r * 5 + 2] = eos; xin[r * 5 + 3] = eoc; xin[r * 5 + 4] = cont;
          flags[SB + r] = step + 1;
          if (a.stop_if_eoc && eoc > 0.f) flags[r] = 1;  // model.py:255-257
        }
      }
    }
    __syncthreads();
    int all_done = 1;
    for (int r = 0; r < SB; ++r) all_done &= flags[r];
    if (all_done) break;
  }
  for (int i = tid; i < rows_valid; i += NT) a.lengths[sk_base + i] = flags[SB + i];
}

template <int SB>
int launch_sample(skrnn_handle *h, cudaStream_t st, const SampArgs &a, int threads) {
  constexpr int SP = (SB + 3) & ~3;
  const int ng = a.cell == 0 ? 4 : (a.cell == 1 ? 2 : 1);
  size_t smem = ((size_t)3 * a.H * SP + (size_t)SB * a.NOUT + SB * 5 + (size_t)SB * a.M + 2 * SB) * 4 + 16 +
                (size_t)SK_NSLOT * SK_KG * threads * ng * 4;
  SK_CHECK(smem <= 227 * 1024, "sampler needs %zu B of shared memory (rnn_size too large)", smem);
  SK_CUDA(cudaFuncSetAttribute(sample_kernel<SB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int blocks = (a.n + SB - 1) / SB;
  SkLaunch scope(h, st, SK_CAT_SAMPLER);
  sample_kernel<SB><<<blocks, threads, smem, st>>>(a);
  SK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int sk_sample_f32(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen,
                  int stop_if_eoc, const float *uniforms, uint64_t seed, int sketch0, float *strokes,
                  int *lengths, float *mixp) {
  SK_CHECK(h->H <= 512, "sampler supports rnn_size <= 512 (got %d)", h->H);
  SK_CHECK(n > 0 && steps > 0, "sample: n and num_steps must be positive");
  SK_CUDA(cudaMemsetAsync(strokes, 0, (size_t)n * steps * 5 * sizeof(float), st));
  if (mixp) SK_CUDA(cudaMemsetAsync(mixp, 0, (size_t)n * steps * h->NOUT * sizeof(float), st));
  {
    // few sketches: the cluster-resident kernel (sampler_cl.cu), same arithmetic; it declines (-1) when the stack does
    // not fit a cluster's shared memory or when there are so many sketches that the streaming kernel below wins
    const int r = sk_sample_cluster(h, st, n, steps, temp_mix, temp_pen, stop_if_eoc, uniforms, seed, sketch0, strokes, lengths, mixp);
    if (r >= 0) return r;
  }
  // recurrent state [n, L, S], then the gate-interleaved weight copies (lstm, gru gates)
  const int ng = h->cfg.cell == SKRNN_CELL_LSTM ? 4 : (h->cfg.cell == SKRNN_CELL_GRU ? 2 : 1);
  const int64_t state_n = ((int64_t)n * h->L * h->S + 63) / 64 * 64;
  int64_t need = state_n;
  if (ng > 1)
    for (int l = 0; l < h->L; ++l) need += ((int64_t)(h->lp[l].I + h->H) * ng * h->H + 63) / 64 * 64;
  if (need > h->samp_state_n) {
    if (h->samp_state) cudaFree(h->samp_state);
    h->samp_state = nullptr;
    h->samp_state_n = 0;
    SK_CUDA(cudaMalloc(&h->samp_state, need * sizeof(float)));
    h->samp_state_n = need;
  }
  SK_CUDA(cudaMemsetAsync(h->samp_state, 0, state_n * sizeof(float), st));  // zero_state (model.py:208)
  SampArgs a = {};
  a.n = n; a.steps = steps; a.H = h->H; a.L = h->L; a.M = h->M; a.NOUT = h->NOUT; a.cell = h->cfg.cell;
  a.temp_mix = temp_mix; a.temp_pen = temp_pen; a.data_scale = h->cfg.data_scale;
  a.stop_if_eoc = stop_if_eoc;
  a.uniforms = uniforms; a.seed = seed; a.sketch0 = sketch0;
  a.params = h->params;
  for (int l = 0; l < h->L; ++l) {
    a.w0[l] = h->lp[l].w[0]; a.b0[l] = h->lp[l].b[0];
    a.w1[l] = h->lp[l].w[1]; a.b1[l] = h->lp[l].b[1];
  }
  if (ng > 1) {
    float *dst = h->samp_state + state_n;
    for (int l = 0; l < h->L; ++l) {
      const int K = h->lp[l].I + h->H;
      SkLaunch scope(h, st, SK_CAT_SAMPLER);
      sample_pack_kernel<<<2 * h->num_sms, 256, 0, st>>>(h->params + h->lp[l].w[0], dst, K, h->H, ng);
      SK_CUDA(cudaGetLastError());
      a.wpk[l] = dst;
      dst += ((int64_t)K * ng * h->H + 63) / 64 * 64;
    }
  }
  a.out_w = h->off_out_w; a.out_b = h->off_out_b;
  a.state = h->samp_state; a.strokes = strokes; a.lengths = lengths; a.mixp = mixp;
  int threads = max(256, (h->H + 31) / 32 * 32);
  if (n <= 2) return launch_sample<1>(h, st, a, threads);
  if (n <= 4 * h->num_sms) return launch_sample<4>(h, st, a, threads);
  // Many sketches: 12..16 per CTA, whichever fills the GPU's CTA slots in the fewest (waves x sketches per CTA) --
  // 8192 sketches on 2 x 148 slots are 1.73 waves of 16 but 1.98 waves of 14.
  int sb_env = 0;  // read per call: the parity tests walk through the instances
  { const char *e = getenv("SKRNN_SAMPLE_SB"); sb_env = e ? atoi(e) : 0; }
  int per_sm = 1;
  {
    const int ng = a.cell == 0 ? 4 : (a.cell == 1 ? 2 : 1);
    const size_t smem16 = ((size_t)3 * a.H * 16 + (size_t)16 * a.NOUT + 16 * 5 + (size_t)16 * a.M + 2 * 16) * 4 + 16 +
                          (size_t)SK_NSLOT * SK_KG * threads * ng * 4;
    if (smem16 <= 227 * 1024) {
      SK_CUDA(cudaFuncSetAttribute(sample_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem16));
      SK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sample_kernel<16>, threads, smem16));
      if (per_sm < 1) per_sm = 1;
    }
  }
  const int64_t slots = (int64_t)per_sm * h->num_sms;
  int best = 16;
  int64_t best_cost = INT64_MAX;
  for (int sb = 16; sb >= 12; --sb) {
    const int64_t ctas = (n + sb - 1) / sb, cost = (ctas + slots - 1) / slots * sb;
    if (cost < best_cost) { best_cost = cost; best = sb; }
  }
  if (sb_env >= 12 && sb_env <= 16) best = sb_env;
  switch (best) {
    case 12: return launch_sample<12>(h, st, a, threads);
    case 13: return launch_sample<13>(h, st, a, threads);
    case 14: return launch_sample<14>(h, st, a, threads);
    case 15: return launch_sample<15>(h, st, a, threads);
    default: return launch_sample<16>(h, st, a, threads);
  }
}
