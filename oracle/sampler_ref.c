/*
 * sampler_ref.c -- CPU restatement of Model.sample (reference model.py:187-264) for ONE sketch,
 * in plain C, with the uniform draws injected.
 *
 * TEST INFRASTRUCTURE ONLY: loaded by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
 * leg as the checker for the CUDA sampler; nothing in the shipped package links it.
 *
 * parity unpinned: the reference holds no golden sampler output (its draws are unseeded,
 * model.py:202,226,235) and TensorFlow 0.x cannot run here.  This file is cross-checked against the
 * independent NumPy transcription oracle/sketch_oracle.py::sample to 1e-4 (tests/test_oracle.py).
 *
 * Arithmetic contract shared with the GPU kernel (so outputs can be compared bit for bit): every
 * [x ; h] * W element is one fmaf chain over k ascending starting from 0, then + bias; all
 * transcendental functions and the Box-Muller transform come from skrnn_math.h; sums over mixture
 * components run k ascending in float32; the inverse-CDF scan accumulates in float64 (NumPy 1.10
 * scalar promotion).  Compile with -ffp-contract=off.
 *
 * Declared deviation from the reference: np.random.multivariate_normal (SVD of the covariance, own
 * RNG stream) is replaced by Box-Muller + Cholesky on two injected uniforms -- same distribution.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../sketch_rnn_b200/csrc/skrnn_math.h"

/* y[n] = fma-chain over k of v[k] * W[k*ld + col0 + n], k ascending, then + b[n] */
static void affine(const float *v, int K, const float *W, int ld, int col0, const float *b, int n, float *y) {
  for (int j = 0; j < n; ++j) {
    float acc = 0.0f;
    for (int k = 0; k < K; ++k) acc = SKM_FMA(v[k], W[(size_t)k * ld + col0 + j], acc);
    y[j] = SKM_ADD(acc, b[j]);
  }
}

/* get_pi_idx, model.py:189-197 */
static int pi_idx(double x, const float *pdf, int n) {
  double acc = 0.0;
  for (int i = 0; i < n; ++i) {
    acc += (double)pdf[i];
    if (acc >= x) return i;
  }
  return -1;
}

/* One cell call of layer l on input `in` (width I) and state `st`; writes the cell output to `out`
 * and the new state to `ns`.  TF r0.10-0.12 BasicLSTMCell / GRUCell / BasicRNNCell (model.py:14-27). */
static void cell_step(int cell, int H, int I, const float *P, int64_t w0, int64_t b0, int64_t w1, int64_t b1,
                      const float *in, const float *st, float *out, float *ns, float *scratch) {
  float *v = scratch;             /* [I + H] */
  float *pre = scratch + I + H;   /* [4H]    */
  if (cell == 0) {
    const float *c = st, *h = st + H;
    memcpy(v, in, sizeof(float) * I);
    memcpy(v + I, h, sizeof(float) * H);
    affine(v, I + H, P + w0, 4 * H, 0, P + b0, 4 * H, pre);
    for (int u = 0; u < H; ++u) {
      float ig = skm_sigmoid(pre[u]);
      float jg = skm_tanh(pre[H + u]);
      float fg = skm_sigmoid(SKM_ADD(pre[2 * H + u], 1.0f)); /* forget_bias */
      float og = skm_sigmoid(pre[3 * H + u]);
      float c2 = SKM_FMA(c[u], fg, SKM_MUL(ig, jg));
      float h2 = SKM_MUL(skm_tanh(c2), og);
      ns[u] = c2;
      ns[H + u] = h2;
      out[u] = h2;
    }
  } else if (cell == 1) {
    const float *h = st;
    memcpy(v, in, sizeof(float) * I);
    memcpy(v + I, h, sizeof(float) * H);
    affine(v, I + H, P + w0, 2 * H, 0, P + b0, 2 * H, pre);
    float *z = pre + 2 * H; /* update gate */
    for (int u = 0; u < H; ++u) {
      float r = skm_sigmoid(pre[u]);
      z[u] = skm_sigmoid(pre[H + u]);
      v[I + u] = SKM_MUL(r, h[u]);
    }
    affine(v, I + H, P + w1, H, 0, P + b1, H, pre);
    for (int u = 0; u < H; ++u) {
      float cand = skm_tanh(pre[u]);
      float h2 = SKM_FMA(z[u], h[u], SKM_MUL(SKM_SUB(1.0f, z[u]), cand));
      ns[u] = h2;
      out[u] = h2;
    }
  } else {
    memcpy(v, in, sizeof(float) * I);
    memcpy(v + I, st, sizeof(float) * H);
    affine(v, I + H, P + w0, H, 0, P + b0, H, pre);
    for (int u = 0; u < H; ++u) {
      float h2 = skm_tanh(pre[u]);
      ns[u] = h2;
      out[u] = h2;
    }
  }
}

/* Returns the number of points produced.  strokes [steps,5] (dx,dy scaled by data_scale, model.py:263),
 * mixp [steps, 3+6M] or NULL laid out (pen_pdf[3], pi_pdf[M], mu1, mu2, sigma1, sigma2, corr). */
int skref_sample(int cell, int H, int L, int M, const float *P, const int64_t *w0, const int64_t *b0,
                 const int64_t *w1, const int64_t *b1, int64_t out_w, int64_t out_b, int steps,
                 float temp_mix, float temp_pen, int stop_if_eoc, float data_scale, const float *uniforms,
                 float *strokes, float *mixp) {
  const int S = cell == 0 ? 2 * H : H;
  const int NOUT = 3 + 6 * M;
  float *state = (float *)calloc((size_t)L * S, sizeof(float)); /* zero_state, model.py:208 */
  float *nstate = (float *)calloc((size_t)L * S, sizeof(float));
  float *scratch = (float *)calloc((size_t)8 * H + 16, sizeof(float));
  float *lay = (float *)calloc((size_t)2 * H + 8, sizeof(float));
  float *z = (float *)calloc(NOUT, sizeof(float));
  float *pdf = (float *)calloc(M, sizeof(float));
  float prev_x[5] = {0, 0, 0, 0, 0}; /* model.py:205 */
  int n_out = steps;
  memset(strokes, 0, sizeof(float) * 5 * steps);
  for (int i = 0; i < steps; ++i) {
    /* sess.run([...], {input_data: prev_x, initial_state: prev_state}), model.py:215-217 */
    const float *in = prev_x;
    int I = 5;
    for (int l = 0; l < L; ++l) {
      float *out = lay + (l & 1) * H;
      cell_step(cell, H, I, P, w0[l], b0[l], w1[l], b1[l], in, state + (size_t)l * S, out,
                nstate + (size_t)l * S, scratch);
      in = out;
      I = H;
    }
    /* eoc reset: where(prev_x[3] > 0, initial_state, new_state), model.py:82-92 */
    if (!(prev_x[3] > 0.0f)) memcpy(state, nstate, sizeof(float) * L * S);
    affine(in, H, P + out_w, NOUT, 0, P + out_b, NOUT, z); /* model.py:98-99 */
    /* get_mixture_coef softmax, model.py:152-156 */
    float mx = z[3];
    for (int k = 1; k < M; ++k) mx = fmaxf(mx, z[3 + k]);
    float sum = 0.0f;
    for (int k = 0; k < M; ++k) { pdf[k] = skm_exp(SKM_SUB(z[3 + k], mx)); sum = SKM_ADD(sum, pdf[k]); }
    float inv = SKM_DIV(1.0f, sum);
    for (int k = 0; k < M; ++k) pdf[k] = SKM_MUL(pdf[k], inv);
    if (i > 1) { /* model.py:220-223 */
      float m2 = -INFINITY;
      for (int k = 0; k < M; ++k) { pdf[k] = SKM_DIV(skm_log(pdf[k]), temp_mix); m2 = fmaxf(m2, pdf[k]); }
      for (int k = 0; k < M; ++k) pdf[k] = skm_exp(SKM_SUB(pdf[k], m2));
    }
    float s2 = 0.0f;
    for (int k = 0; k < M; ++k) s2 = SKM_ADD(s2, pdf[k]);
    for (int k = 0; k < M; ++k) pdf[k] = SKM_DIV(pdf[k], s2); /* model.py:224 */
    const float *u = uniforms + (size_t)i * 4;
    int idx = pi_idx((double)u[0], pdf, M); /* model.py:226 */
    if (idx < 0) idx = M - 1;               /* o_mu1[0][-1] */
    float pm = fmaxf(z[0], fmaxf(z[1], z[2])); /* model.py:228-233; `pi_pdf /= temp_pen` never touches pen */
    float pe[3];
    for (int j = 0; j < 3; ++j) pe[j] = skm_exp(SKM_SUB(z[j], pm));
    float ps = SKM_ADD(SKM_ADD(pe[0], pe[1]), pe[2]);
    for (int j = 0; j < 3; ++j) pe[j] = SKM_DIV(pe[j], ps);
    int pen_idx = pi_idx((double)u[1], pe, 3); /* model.py:235 */
    float eos = 0, eoc = 0, cont = 0;
    if (pen_idx == 0) eos = 1; else if (pen_idx == 1) eoc = 1; else cont = 1;
    float mu1 = z[3 + M + idx], mu2 = z[3 + 2 * M + idx];
    float sg1 = skm_exp(z[3 + 3 * M + idx]), sg2 = skm_exp(z[3 + 4 * M + idx]);
    float rho = skm_tanh(z[3 + 5 * M + idx]);
    float g1, g2;
    skm_box_muller(u[2], u[3], &g1, &g2);
    float x1 = SKM_FMA(sg1, g1, mu1);
    float q = SKM_SQRT(SKM_SUB(1.0f, SKM_MUL(rho, rho)));
    float x2 = SKM_FMA(sg2, SKM_FMA(q, g2, SKM_MUL(rho, g1)), mu2);
    float *so = strokes + (size_t)i * 5;
    so[0] = SKM_MUL(x1, data_scale); so[1] = SKM_MUL(x2, data_scale);
    so[2] = eos; so[3] = eoc; so[4] = cont;
    if (mixp) {
      float *mo = mixp + (size_t)i * NOUT;
      mo[0] = pe[0]; mo[1] = pe[1]; mo[2] = pe[2];
      for (int k = 0; k < M; ++k) {
        mo[3 + k] = (i > 1) ? SKM_DIV(pdf[k], temp_pen) : pdf[k]; /* model.py:230,251 */
        mo[3 + M + k] = z[3 + M + k];
        mo[3 + 2 * M + k] = z[3 + 2 * M + k];
        mo[3 + 3 * M + k] = skm_exp(z[3 + 3 * M + k]);
        mo[3 + 4 * M + k] = skm_exp(z[3 + 4 * M + k]);
        mo[3 + 5 * M + k] = skm_tanh(z[3 + 5 * M + k]);
      }
    }
    if (stop_if_eoc && eoc == 1) { n_out = i + 1; break; } /* model.py:255-257 */
    prev_x[0] = x1; prev_x[1] = x2; prev_x[2] = eos; prev_x[3] = eoc; prev_x[4] = cont;
  }
  free(state); free(nstate); free(scratch); free(lay); free(z); free(pdf);
  return n_out;
}

/* element-wise access to the shared math for host-side accuracy tests */
float skref_exp(float x) { return skm_exp(x); }
float skref_log(float x) { return skm_log(x); }
float skref_tanh(float x) { return skm_tanh(x); }
void skref_sincos2pi(float u, float *s, float *c) { skm_sincos2pi(u, s, c); }
void skref_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t *out) {
  skm_philox4x32(c0, c1, c2, c3, k0, k1, out);
}
