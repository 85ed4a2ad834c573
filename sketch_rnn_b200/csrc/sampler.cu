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
// sketches is staged k-major in shared memory and broadcast to all threads.
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
  int64_t w0[SK_MAX_LAYERS], b0[SK_MAX_LAYERS], w1[SK_MAX_LAYERS], b1[SK_MAX_LAYERS];
  int64_t out_w, out_b;
  float *state;           // [n, L, S] zero-initialised
  float *strokes;         // [n, steps, 5]
  int *lengths;           // [n]
  float *mixp;            // [n, steps, NOUT] or null
};

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
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, NT = blockDim.x;
  const int H = a.H, L = a.L, M = a.M, NOUT = a.NOUT;
  const int S = (a.cell == 0) ? 2 * H : H;
  const int hoff = (a.cell == 0) ? H : 0;
  const int sk_base = blockIdx.x * SB;
  const int rows_valid = min(SB, a.n - sk_base);
  const bool act = tid < H;
  const int u = tid;

  float *AT = smem;                       // [2H rows k][SB]: [x ; h] of the current layer
  float *nxt = AT + (size_t)2 * H * SB;   // [H][SB]: output of the current layer
  float *lg = nxt + (size_t)H * SB;       // [SB][NOUT] logits
  float *xin = lg + (size_t)SB * NOUT;    // [SB][5] current input point
  float *pdf = xin + SB * 5;              // [SB][M] scratch for the per-sketch draw
  int *flags = reinterpret_cast<int *>(pdf + (size_t)SB * M);  // [SB] done, [SB] length

  for (int i = tid; i < SB * 5; i += NT) xin[i] = 0.f;  // first input is all zeros (model.py:205)
  for (int i = tid; i < SB; i += NT) { flags[i] = (i < rows_valid) ? 0 : 1; flags[SB + i] = 0; }
  __syncthreads();

  for (int step = 0; step < a.steps; ++step) {
    for (int l = 0; l < L; ++l) {
      const int I = (l == 0) ? 5 : H;
      const int K = I + H;
      // ---- stage [x ; h] ----
      if (l == 0) {
        for (int i = tid; i < 5 * SB; i += NT) { int k = i / SB, r = i % SB; AT[k * SB + r] = xin[r * 5 + k]; }
      } else {
        for (int i = tid; i < H * SB; i += NT) AT[i] = nxt[i];
      }
      __syncthreads();  // nxt fully copied before anyone overwrites it below
      for (int i = tid; i < H * SB; i += NT) {
        int r = i / H, uu = i % H;
        float v = 0.f;
        if (r < rows_valid) v = a.state[((size_t)(sk_base + r) * L + l) * S + hoff + uu];
        AT[(size_t)(I + uu) * SB + r] = v;
      }
      __syncthreads();
      const float *W = a.params + a.w0[l];
      const float *bias = a.params + a.b0[l];
      if (a.cell == 0) {
        if (act) {
          float acc[SB][4];
#pragma unroll
          for (int r = 0; r < SB; ++r) { acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.f; }
#pragma unroll (SB >= 16 ? 2 : 8)  // weight loads of several k in flight: the chain per accumulator stays sequential
          for (int k = 0; k < K; ++k) {
            const float *wr = W + (size_t)k * (4 * H) + u;
            float w0 = __ldg(wr), w1 = __ldg(wr + H), w2 = __ldg(wr + 2 * H), w3 = __ldg(wr + 3 * H);
#pragma unroll
            for (int r = 0; r < SB; ++r) {
              float av = AT[(size_t)k * SB + r];
              acc[r][0] = SKM_FMA(av, w0, acc[r][0]);
              acc[r][1] = SKM_FMA(av, w1, acc[r][1]);
              acc[r][2] = SKM_FMA(av, w2, acc[r][2]);
              acc[r][3] = SKM_FMA(av, w3, acc[r][3]);
            }
          }
          const float bi = bias[u], bj = bias[H + u], bf = bias[2 * H + u], bo = bias[3 * H + u];
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            if (r >= rows_valid) { nxt[(size_t)u * SB + r] = 0.f; continue; }
            float *st = a.state + ((size_t)(sk_base + r) * L + l) * S;
            float c = st[u];
            float ig = skm_sigmoid(SKM_ADD(acc[r][0], bi));
            float jg = skm_tanh(SKM_ADD(acc[r][1], bj));
            float fg = skm_sigmoid(SKM_ADD(SKM_ADD(acc[r][2], bf), 1.0f));
            float og = skm_sigmoid(SKM_ADD(acc[r][3], bo));
            float c2 = SKM_FMA(c, fg, SKM_MUL(ig, jg));
            float h2 = SKM_MUL(skm_tanh(c2), og);
            nxt[(size_t)u * SB + r] = h2;
            // an eoc input holds the state: initial_state is prev_state (model.py:82-92,215)
            if (!(xin[r * 5 + 3] > 0.f)) { st[u] = c2; st[H + u] = h2; }
          }
        }
      } else if (a.cell == 2) {
        if (act) {
          float acc[SB];
#pragma unroll
          for (int r = 0; r < SB; ++r) acc[r] = 0.f;
#pragma unroll (SB >= 16 ? 2 : 8)  // weight loads of several k in flight: the chain per accumulator stays sequential
          for (int k = 0; k < K; ++k) {
            float w0 = __ldg(W + (size_t)k * H + u);
#pragma unroll
            for (int r = 0; r < SB; ++r) acc[r] = SKM_FMA(AT[(size_t)k * SB + r], w0, acc[r]);
          }
          const float bb = bias[u];
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            if (r >= rows_valid) { nxt[(size_t)u * SB + r] = 0.f; continue; }
            float h2 = skm_tanh(SKM_ADD(acc[r], bb));
            nxt[(size_t)u * SB + r] = h2;
            if (!(xin[r * 5 + 3] > 0.f)) a.state[((size_t)(sk_base + r) * L + l) * S + u] = h2;
          }
        }
      } else {
        float az[SB], hold[SB], ar[SB];
        if (act) {  // gates r, u on [x ; h]
#pragma unroll
          for (int r = 0; r < SB; ++r) { ar[r] = 0.f; az[r] = 0.f; hold[r] = AT[(size_t)(I + u) * SB + r]; }
#pragma unroll (SB >= 16 ? 2 : 8)  // weight loads of several k in flight: the chain per accumulator stays sequential
          for (int k = 0; k < K; ++k) {
            const float *wr = W + (size_t)k * (2 * H) + u;
            float w0 = __ldg(wr), w1 = __ldg(wr + H);
#pragma unroll
            for (int r = 0; r < SB; ++r) {
              float av = AT[(size_t)k * SB + r];
              ar[r] = SKM_FMA(av, w0, ar[r]);
              az[r] = SKM_FMA(av, w1, az[r]);
            }
          }
          const float br = bias[u], bz = bias[H + u];
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            ar[r] = SKM_MUL(skm_sigmoid(SKM_ADD(ar[r], br)), hold[r]);  // r * h
            az[r] = skm_sigmoid(SKM_ADD(az[r], bz));
          }
        }
        __syncthreads();  // every thread is done reading h from AT
        if (act) {
#pragma unroll
          for (int r = 0; r < SB; ++r) AT[(size_t)(I + u) * SB + r] = ar[r];
        }
        __syncthreads();
        if (act) {  // candidate on [x ; r*h]
          const float *Wc = a.params + a.w1[l];
          const float bc = a.params[a.b1[l] + u];
          float ac[SB];
#pragma unroll
          for (int r = 0; r < SB; ++r) ac[r] = 0.f;
#pragma unroll (SB >= 16 ? 2 : 8)  // weight loads of several k in flight: the chain per accumulator stays sequential
          for (int k = 0; k < K; ++k) {
            float w0 = __ldg(Wc + (size_t)k * H + u);
#pragma unroll
            for (int r = 0; r < SB; ++r) ac[r] = SKM_FMA(AT[(size_t)k * SB + r], w0, ac[r]);
          }
#pragma unroll
          for (int r = 0; r < SB; ++r) {
            if (r >= rows_valid) { nxt[(size_t)u * SB + r] = 0.f; continue; }
            float cand = skm_tanh(SKM_ADD(ac[r], bc));
            float h2 = SKM_FMA(az[r], hold[r], SKM_MUL(SKM_SUB(1.0f, az[r]), cand));
            nxt[(size_t)u * SB + r] = h2;
            if (!(xin[r * 5 + 3] > 0.f)) a.state[((size_t)(sk_base + r) * L + l) * S + u] = h2;
          }
        }
      }
      __syncthreads();
    }
    // ---- output projection (model.py:98-99) ----
    {
      const float *Wo = a.params + a.out_w;
      for (int j = tid; j < NOUT; j += NT) {
        float acc[SB];
#pragma unroll
        for (int r = 0; r < SB; ++r) acc[r] = 0.f;
#pragma unroll (SB >= 16 ? 2 : 8)  // weight loads of several k in flight: the chain per accumulator stays sequential
        for (int k = 0; k < H; ++k) {
          float w0 = __ldg(Wo + (size_t)k * NOUT + j);
#pragma unroll
          for (int r = 0; r < SB; ++r) acc[r] = SKM_FMA(nxt[(size_t)k * SB + r], w0, acc[r]);
        }
        const float bb = a.params[a.out_b + j];
#pragma unroll
        for (int r = 0; r < SB; ++r) lg[(size_t)r * NOUT + j] = SKM_ADD(acc[r], bb);
      }
    }
    __syncthreads();
    // ---- per-sketch draw and feedback (model.py:219-261) ----
    if (tid < rows_valid && !flags[tid]) {
      const int r = tid;
      const int sk = sk_base + r;
      const float *z = lg + (size_t)r * NOUT;
      float *p = pdf + (size_t)r * M;
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
      // softmax over the M mixture logits (get_mixture_coef, model.py:152-156)
      float mx = z[3];
      for (int k = 1; k < M; ++k) mx = fmaxf(mx, z[3 + k]);
      float sum = 0.f;
      for (int k = 0; k < M; ++k) { p[k] = skm_exp(SKM_SUB(z[3 + k], mx)); sum = SKM_ADD(sum, p[k]); }
      float inv = SKM_DIV(1.0f, sum);
      for (int k = 0; k < M; ++k) p[k] = SKM_MUL(p[k], inv);
      if (step > 1) {  // temperature only from the third point (model.py:220)
        float m2 = -INFINITY;
        for (int k = 0; k < M; ++k) { p[k] = SKM_DIV(skm_log(p[k]), a.temp_mix); m2 = fmaxf(m2, p[k]); }
        for (int k = 0; k < M; ++k) p[k] = skm_exp(SKM_SUB(p[k], m2));
      }
      float s2 = 0.f;
      for (int k = 0; k < M; ++k) s2 = SKM_ADD(s2, p[k]);
      for (int k = 0; k < M; ++k) p[k] = SKM_DIV(p[k], s2);
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
        for (int k = 0; k < M; ++k) {
          // the stray `pi_pdf /= temp_pen` of model.py:230 shows up in the returned params
          mo[3 + k] = (step > 1) ? SKM_DIV(p[k], a.temp_pen) : p[k];
          mo[3 + M + k] = z[3 + M + k];
          mo[3 + 2 * M + k] = z[3 + 2 * M + k];
          mo[3 + 3 * M + k] = skm_exp(z[3 + 3 * M + k]);
          mo[3 + 4 * M + k] = skm_exp(z[3 + 4 * M + k]);
          mo[3 + 5 * M + k] = skm_tanh(z[3 + 5 * M + k]);
        }
      }
      // feedback (model.py:259-261): nobody reads xin between the barrier above and the next one
      xin[r * 5 + 0] = x1; xin[r * 5 + 1] = x2; xin[r * 5 + 2] = eos; xin[r * 5 + 3] = eoc; xin[r * 5 + 4] = cont;
      flags[SB + r] = step + 1;
      if (a.stop_if_eoc && eoc > 0.f) flags[r] = 1;  // model.py:255-257
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
  size_t smem = ((size_t)3 * a.H * SB + (size_t)SB * a.NOUT + SB * 5 + (size_t)SB * a.M + 2 * SB) * 4;
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
  int64_t need = (int64_t)n * h->L * h->S;
  if (need > h->samp_state_n) {
    if (h->samp_state) cudaFree(h->samp_state);
    h->samp_state = nullptr;
    h->samp_state_n = 0;
    SK_CUDA(cudaMalloc(&h->samp_state, need * sizeof(float)));
    h->samp_state_n = need;
  }
  SK_CUDA(cudaMemsetAsync(h->samp_state, 0, need * sizeof(float), st));  // zero_state (model.py:208)
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
  a.out_w = h->off_out_w; a.out_b = h->off_out_b;
  a.state = h->samp_state; a.strokes = strokes; a.lengths = lengths; a.mixp = mixp;
  int threads = max(256, (h->H + 31) / 32 * 32);
  if (n <= 2) return launch_sample<1>(h, st, a, threads);
  if (n <= 4 * h->num_sms) return launch_sample<4>(h, st, a, threads);
  return launch_sample<16>(h, st, a, threads);
}
