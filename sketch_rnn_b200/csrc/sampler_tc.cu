// sampler_tc.cu -- throughput sampler for many concurrent sketches on the bf16 tensor-core path
// (reference Model.sample, model.py:187-264, batched over sketches; BASELINE config 5).
//
// With thousands of sketches in flight one sampling step is three real GEMMs
//   gates_l [n, 4H] = [input_l | h_l] [n, K_l] * W_l [K_l, 4H] + b_l      (l = 0..L-1)
//   logits  [n, NOUT] = h_top [n, H] * output_w + output_b
// The gate GEMMs run the LSTM cell update with the eoc state hold (model.py:82-92 with initial_state = prev_state,
// model.py:215) as their epilogue (gemm_tc_persist_kernel<.,.,256,1>: gate-interleaved weight columns, so the
// pre-activations never leave the SM); the logits GEMM is followed by the per-sketch draw (softmax, temperature from the third
// point on, inverse-CDF categorical draws, Box-Muller bivariate normal, feedback of the sampled point as
// the next input -- model.py:219-261, same quirks as the exact kernel in sampler.cu).  Everything stays on
// the device for all steps; the host only enqueues launches.
//
// This path trades the bit-exact contract of sampler.cu for throughput: operands are rounded to bf16, so
// it is selected only for bf16-precision handles, LSTM cells and n >= 64 sketches; everything else runs the
// exact fp32 persistent kernel.
#include <stdlib.h>

#include "common.cuh"
#include "skrnn_math.h"

namespace {

struct SampBuf {
  __nv_bfloat16 *xh[SK_MAX_LAYERS];  // [n, K_l]: (input | h) of layer l;  K_0 = 8 + H, K_l = 2H
  __nv_bfloat16 *wsm[SK_MAX_LAYERS]; // [K_l, 4H] weights in the same row order
  __nv_bfloat16 *htop;               // [n, H] output of the top layer
  float *cst[SK_MAX_LAYERS];         // [n, H] cell state
  float *bias_il[SK_MAX_LAYERS];     // [4H] bias in the interleaved column order
  float *logits;                     // [n, NOUT]
  float *xin;                        // [n, 5] current input point (fp32)
  int *done;                         // [n]
};

// W rows: layer 0: 5 x rows, 3 zero rows, H h rows; deeper layers: the TF matrix as it is.  Columns are
// gate-interleaved (4*unit + gate instead of TF's gate*H + unit) so that a GEMM tile holds whole units and the cell
// update can be its epilogue; the bias gets the same order.
__global__ void samp_pack_kernel(const float *__restrict__ mat, const float *__restrict__ bias, int I, int H, int Kp,
                                 __nv_bfloat16 *__restrict__ dst, float *__restrict__ bias_il) {
  const int64_t total = (int64_t)Kp * 4 * H;
  const int pad = Kp - (I + H);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i / (4 * H)), ci = (int)(i % (4 * H));
    const int c = (ci & 3) * H + (ci >> 2);  // TF column of interleaved column ci
    float v = 0.f;
    if (k < I) v = mat[(size_t)k * 4 * H + c];
    else if (k >= I + pad) v = mat[(size_t)(k - pad) * 4 * H + c];
    dst[i] = __float2bfloat16_rn(v);
    if (k == 0) bias_il[ci] = bias[c];
  }
}

__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Everything that changes from one skrnn_sample call to the next lives in this device-side record, so the captured
// launch sequence of a call (a CUDA graph of steps*(L+2) kernels) can be replayed by later calls.
struct SampCall {
  const float *uniforms;
  float *strokes;
  int *lengths;
  float *mixp;
  uint64_t seed;
  int sketch0, stop_if_eoc;
  float temp_mix, temp_pen;
};

// One warp per sketch; lanes over mixture components (M <= 32 on this path).
__global__ void samp_draw_kernel(const float *__restrict__ logits, int n, int M, int NOUT, int step, int steps, float data_scale,
                                 const SampCall *__restrict__ call, float *__restrict__ xin, __nv_bfloat16 *__restrict__ xh0,
                                 int k0, int *__restrict__ done) {
  const float *uniforms = call->uniforms;
  float *strokes = call->strokes, *mixp = call->mixp;
  int *lengths = call->lengths;
  const uint64_t seed = call->seed;
  const int sketch0 = call->sketch0, stop_if_eoc = call->stop_if_eoc;
  const float temp_mix = call->temp_mix, temp_pen = call->temp_pen;
  const int lane = threadIdx.x & 31;
  const int sk = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (sk >= n) return;
  if (done[sk]) return;
  const float *z = logits + (size_t)sk * NOUT;
  float u4[4];
  if (uniforms != nullptr) {
    const float *up = uniforms + ((size_t)sk * steps + step) * 4;
    u4[0] = up[0]; u4[1] = up[1]; u4[2] = up[2]; u4[3] = up[3];
  } else {
    uint32_t rr[4];
    skm_philox4x32((uint32_t)(sketch0 + sk), (uint32_t)step, 0u, 0x5a3711eu, (uint32_t)seed, (uint32_t)(seed >> 32), rr);
    for (int q = 0; q < 4; ++q) u4[q] = skm_u01(rr[q]);
  }
  // mixture weights: softmax, then (from the third point on) the temperature re-softmax of model.py:220-224
  const float zl = lane < M ? z[3 + lane] : -INFINITY;
  float mx = warp_max_f(zl);
  float e = lane < M ? __expf(zl - mx) : 0.f;
  float p = e / warp_sum_f(e);
  if (step > 1) {
    const float lp = lane < M ? __logf(fmaxf(p, 1e-37f)) / temp_mix : -INFINITY;
    mx = warp_max_f(lp);
    e = lane < M ? __expf(lp - mx) : 0.f;
    p = e / warp_sum_f(e);
  }
  float cdf = p;  // inclusive scan over lanes
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, cdf, o);
    if (lane >= o) cdf += t;
  }
  const unsigned hit = __ballot_sync(0xffffffffu, lane < M && cdf >= u4[0]);
  const int idx = hit ? (__ffs(hit) - 1) : (M - 1);  // get_pi_idx == -1 -> last component
  // pen state (plain softmax; the reference never applies the pen temperature)
  const float z0 = z[0], z1 = z[1], z2 = z[2];
  const float pm = fmaxf(z0, fmaxf(z1, z2));
  const float e0 = __expf(z0 - pm), e1 = __expf(z1 - pm), e2 = __expf(z2 - pm);
  const float ps = e0 + e1 + e2;
  float eos = 0.f, eoc = 0.f, cont = 0.f;
  if (e0 / ps >= u4[1]) eos = 1.f;
  else if ((e0 + e1) / ps >= u4[1]) eoc = 1.f;
  else cont = 1.f;
  if (mixp != nullptr) {  // same record as the exact kernel (include/skrnn.h: mix_params_dev)
    float *mo = mixp + ((size_t)sk * steps + step) * NOUT;
    if (lane == 0) { mo[0] = e0 / ps; mo[1] = e1 / ps; mo[2] = e2 / ps; }
    if (lane < M) {
      mo[3 + lane] = (step > 1) ? p / temp_pen : p;
      mo[3 + M + lane] = z[3 + M + lane];
      mo[3 + 2 * M + lane] = z[3 + 2 * M + lane];
      mo[3 + 3 * M + lane] = __expf(z[3 + 3 * M + lane]);
      mo[3 + 4 * M + lane] = __expf(z[3 + 4 * M + lane]);
      mo[3 + 5 * M + lane] = tanhf(z[3 + 5 * M + lane]);
    }
  }
  if (lane == 0) {
    const float mu1 = z[3 + M + idx], mu2 = z[3 + 2 * M + idx];
    const float s1 = __expf(z[3 + 3 * M + idx]), s2 = __expf(z[3 + 4 * M + idx]);
    const float rho = tanhf(z[3 + 5 * M + idx]);
    float g1, g2;
    skm_box_muller(u4[2], u4[3], &g1, &g2);
    const float x1 = mu1 + s1 * g1;
    const float x2 = mu2 + s2 * (rho * g1 + sqrtf(fmaxf(0.f, 1.0f - rho * rho)) * g2);
    float *so = strokes + ((size_t)sk * steps + step) * 5;
    so[0] = x1 * data_scale; so[1] = x2 * data_scale; so[2] = eos; so[3] = eoc; so[4] = cont;
    float *xi = xin + (size_t)sk * 5;
    xi[0] = x1; xi[1] = x2; xi[2] = eos; xi[3] = eoc; xi[4] = cont;
    __nv_bfloat16 *xb = xh0 + (size_t)sk * k0;
    xb[0] = __float2bfloat16_rn(x1); xb[1] = __float2bfloat16_rn(x2);
    xb[2] = __float2bfloat16_rn(eos); xb[3] = __float2bfloat16_rn(eoc); xb[4] = __float2bfloat16_rn(cont);
    lengths[sk] = step + 1;
    if (stop_if_eoc && eoc > 0.f) done[sk] = 1;
  }
}

}  // namespace

// Returns 0 when handled, -1 when this path does not apply (caller runs the exact kernel), > 0 on error.
int sk_sample_tc(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen, int stop_if_eoc,
                 const float *uniforms, uint64_t seed, int sketch0, float *strokes, int *lengths, float *mixp) {
  if (h->cfg.precision != SKRNN_PREC_BF16 || h->cfg.cell != SKRNN_CELL_LSTM) return -1;
  if (n < 64 || h->M > 32 || h->params_bf == nullptr) return -1;
  const int H = h->H, L = h->L, G = 4 * H, NOUT = h->NOUT;
  // one allocation for all buffers, kept on the handle
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) / 256 * 256; return o; };
  size_t o_xh[SK_MAX_LAYERS], o_w[SK_MAX_LAYERS], o_c[SK_MAX_LAYERS], o_b[SK_MAX_LAYERS];
  int K[SK_MAX_LAYERS];
  for (int l = 0; l < L; ++l) {
    K[l] = (l == 0) ? 8 + H : 2 * H;
    o_xh[l] = take((size_t)n * K[l] * 2);
    o_w[l] = take((size_t)K[l] * G * 2);
    o_c[l] = take((size_t)n * H * 4);
    o_b[l] = take((size_t)G * 4);
  }
  const size_t o_top = take((size_t)n * H * 2), o_lg = take((size_t)n * NOUT * 4);
  const size_t o_xin = take((size_t)n * 5 * 4), o_done = take((size_t)n * 4), o_call = take(sizeof(SampCall));
  const size_t o_tab = take(sk_sample_tp_scratch_bytes());  // input-term weight images of the persistent kernel (sampler_tp.cu)
  if (off > h->samp_tc_bytes) {
    if (h->samp_graph) { cudaGraphExecDestroy(h->samp_graph); h->samp_graph = nullptr; }
    SK_CUDA(cudaStreamSynchronize(st));
    cudaFree(h->samp_tc);
    h->samp_tc = nullptr;
    h->samp_tc_bytes = 0;
    SK_CUDA(cudaMalloc(&h->samp_tc, off));
    h->samp_tc_bytes = off;
  }
  uint8_t *base = static_cast<uint8_t *>(h->samp_tc);
  SampBuf b;
  for (int l = 0; l < L; ++l) {
    b.xh[l] = reinterpret_cast<__nv_bfloat16 *>(base + o_xh[l]);
    b.wsm[l] = reinterpret_cast<__nv_bfloat16 *>(base + o_w[l]);
    b.cst[l] = reinterpret_cast<float *>(base + o_c[l]);
    b.bias_il[l] = reinterpret_cast<float *>(base + o_b[l]);
  }
  b.htop = reinterpret_cast<__nv_bfloat16 *>(base + o_top);
  b.logits = reinterpret_cast<float *>(base + o_lg);
  b.xin = reinterpret_cast<float *>(base + o_xin);
  b.done = reinterpret_cast<int *>(base + o_done);
  // zero state, zero first input (model.py:205,208); weights in operand order
  SK_CUDA(cudaMemsetAsync(base, 0, off, st));
  SK_CUDA(cudaMemsetAsync(strokes, 0, (size_t)n * steps * 5 * sizeof(float), st));
  SK_CUDA(cudaMemsetAsync(lengths, 0, (size_t)n * sizeof(int), st));
  if (mixp) SK_CUDA(cudaMemsetAsync(mixp, 0, (size_t)n * steps * NOUT * sizeof(float), st));
  SampCall *call_dev = reinterpret_cast<SampCall *>(base + o_call);
  {
    SampCall c;
    c.uniforms = uniforms; c.strokes = strokes; c.lengths = lengths; c.mixp = mixp; c.seed = seed; c.sketch0 = sketch0;
    c.stop_if_eoc = stop_if_eoc; c.temp_mix = temp_mix; c.temp_pen = temp_pen;
    SK_CUDA(cudaMemcpyAsync(call_dev, &c, sizeof(c), cudaMemcpyHostToDevice, st));  // pageable source: staged before return
  }
  SK_TRY(sk_cast_params(h, st));
  for (int l = 0; l < L; ++l) {
    SkLaunch scope(h, st, SK_CAT_SAMPLER);
    samp_pack_kernel<<<4 * h->num_sms, 256, 0, st>>>(h->params + h->lp[l].w[0], h->params + h->lp[l].b[0], h->lp[l].I, H, K[l],
                                                     b.wsm[l], b.bias_il[l]);
    SK_CUDA(cudaGetLastError());
  }
  {
    // rnn_size 256, two layers: every step of every sketch in one persistent kernel (sampler_tp.cu)
    const int r = sk_sample_tp(h, st, n, steps, temp_mix, temp_pen, stop_if_eoc, uniforms, seed, sketch0, strokes, lengths, mixp,
                               b.wsm, b.bias_il, b.cst, reinterpret_cast<float *>(base + o_tab));
    if (r >= 0) return r;
  }
  auto enqueue_steps = [&](cudaStream_t s2) -> int {
    for (int step = 0; step < steps; ++step) {
      for (int l = 0; l < L; ++l) {
        // gates GEMM with the cell update as its epilogue: only c, h and the layer output touch memory
        __nv_bfloat16 *out = (l + 1 < L) ? b.xh[l + 1] : b.htop;
        const int k_out = (l + 1 < L) ? K[l + 1] : H;
        SK_TRY(sk_gemm_tc_lstm_cell(h, s2, n, H, K[l], b.xh[l], K[l], b.wsm[l], b.bias_il[l], b.cst[l], b.xh[l], K[l], K[l] - H,
                                    out, k_out, b.xin));
      }
      SK_TRY(sk_gemm_tc(h, s2, n, NOUT, H, b.htop, H, true, h->params_bf + h->pbf_out_w, h->NOUTP, false, b.logits, NOUT,
                        h->params + h->off_out_b, false));
      SkLaunch scope(h, s2, SK_CAT_SAMPLER);
      samp_draw_kernel<<<(n * 32 + 255) / 256, 256, 0, s2>>>(b.logits, n, h->M, NOUT, step, steps, h->cfg.data_scale, call_dev,
                                                            b.xin, b.xh[0], K[0], b.done);
      SK_CUDA(cudaGetLastError());
    }
    return 0;
  };
  // The step loop is launch-latency bound for small n: it is captured once per (n, steps) into a CUDA graph (on the
  // handle's side stream) and replayed; per-category profiling needs the individual launches.
  static int graph_env = -1;
  if (graph_env < 0) { const char *e = getenv("SKRNN_SAMPLER_GRAPH"); graph_env = (e && e[0] == '0') ? 0 : 1; }
  if (!graph_env || h->prof_on) return enqueue_steps(st);
  if (h->samp_graph == nullptr || h->samp_graph_n != n || h->samp_graph_steps != steps) {
    if (h->samp_graph) { cudaGraphExecDestroy(h->samp_graph); h->samp_graph = nullptr; }
    SK_CUDA(cudaStreamBeginCapture(h->side, cudaStreamCaptureModeThreadLocal));
    const int rc = enqueue_steps(h->side);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(h->side, &graph);
    if (rc != 0 || ce != cudaSuccess || graph == nullptr) {
      if (graph) cudaGraphDestroy(graph);
      cudaGetLastError();
      if (rc != 0) return rc;
      return enqueue_steps(st);  // capture not possible here: plain launches
    }
    const cudaError_t ie = cudaGraphInstantiate(&h->samp_graph, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) {
      h->samp_graph = nullptr;
      cudaGetLastError();
      return enqueue_steps(st);
    }
    h->samp_graph_n = n; h->samp_graph_steps = steps;
  } else {
    h->launches += (int64_t)steps * (L + 2);  // the replayed kernels
  }
  SK_CUDA(cudaGraphLaunch(h->samp_graph, st));
  return 0;
}
