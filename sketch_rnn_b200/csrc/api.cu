// api.cu -- the extern "C" surface of libskrnn.so (include/skrnn.h): handle lifetime, TF-layout
// parameter vector, and the orchestration of one forward / loss+BPTT / optimiser step out of the
// kernels in the other translation units.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[1024] = "";

void sk_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

namespace {

const char *kCellScope[3] = {"BasicLSTMCell", "GRUCell", "BasicRNNCell"};

void add_tensor(skrnn_handle *h, const std::string &name, int rows, int cols, int64_t *off) {
  SkTensor t;
  t.name = name;
  t.offset = h->P;
  t.rows = rows;
  t.cols = cols;
  *off = h->P;
  h->P += (int64_t)rows * cols;
  h->tensors.push_back(t);
}

// tf.trainable_variables() order and TF shapes (model.py:45-47, then MultiRNNCell/Cell{l}).
void build_layout(skrnn_handle *h) {
  const int H = h->H;
  h->P = 0;
  add_tensor(h, "rnn_mdn/output_w", H, h->NOUT, &h->off_out_w);
  add_tensor(h, "rnn_mdn/output_b", 1, h->NOUT, &h->off_out_b);
  for (int l = 0; l < h->L; ++l) {
    SkLayerParams &lp = h->lp[l];
    lp.I = (l == 0) ? 5 : H;
    char base[128];
    snprintf(base, sizeof(base), "rnn_mdn/MultiRNNCell/Cell%d/%s/", l, kCellScope[h->cfg.cell]);
    lp.w[1] = lp.b[1] = 0;
    lp.ncol[1] = 0;
    if (h->cfg.cell == SKRNN_CELL_GRU) {
      lp.ncol[0] = 2 * H;
      lp.ncol[1] = H;
      add_tensor(h, std::string(base) + "Gates/Linear/Matrix", lp.I + H, 2 * H, &lp.w[0]);
      add_tensor(h, std::string(base) + "Gates/Linear/Bias", 1, 2 * H, &lp.b[0]);
      add_tensor(h, std::string(base) + "Candidate/Linear/Matrix", lp.I + H, H, &lp.w[1]);
      add_tensor(h, std::string(base) + "Candidate/Linear/Bias", 1, H, &lp.b[1]);
    } else {
      lp.ncol[0] = h->G;
      add_tensor(h, std::string(base) + "Linear/Matrix", lp.I + H, h->G, &lp.w[0]);
      add_tensor(h, std::string(base) + "Linear/Bias", 1, h->G, &lp.b[0]);
    }
  }
}

void free_workspace(SkWorkspace &w) {
  cudaFree(w.x_tm);
  for (int l = 0; l < SK_MAX_LAYERS; ++l) {
    cudaFree(w.xw[l]);
    cudaFree(w.hprev[l]);
    cudaFree(w.cprev[l]);
    cudaFree(w.out[l]);
    w.xw[l] = w.hprev[l] = w.cprev[l] = w.out[l] = nullptr;
  }
  cudaFree(w.dout);
  cudaFree(w.logits);
  cudaFree(w.eoc);
  w.eoc = nullptr;
  w.x_tm = w.dout = w.logits = nullptr;
  cudaFree(w.x8_bf);
  cudaFree(w.dpre_bf_buf[0]);
  cudaFree(w.dpre_bf_buf[1]);
  cudaFree(w.dlogits_bf);
  w.x8_bf = w.dpre_bf = w.dpre_bf_buf[0] = w.dpre_bf_buf[1] = w.dlogits_bf = nullptr;
  cudaFree(w.xw_tt);
  cudaFree(w.dout_tt);
  w.xw_tt = nullptr;
  w.dout_tt = nullptr;
  w.tt_rows = 0;
  for (int l = 0; l < SK_MAX_LAYERS; ++l) {
    cudaFree(w.out_bf[l]);
    cudaFree(w.hprev_bf[l]);
    cudaFree(w.aux_bf[l]);
    cudaFree(w.acts_bf[l]);
    cudaFree(w.xw_il[l]);
    w.out_bf[l] = w.hprev_bf[l] = w.aux_bf[l] = w.acts_bf[l] = w.xw_il[l] = nullptr;
  }
  w.cap_rows = 0;
  w.f32_acts = false;
}

// bf16 handles whose cell runs on the tcgen05 recurrent kernels keep their own bf16 activation records: the fp32 set
// (xw, hprev, out: 6 of the 10 GB of config 4) is then dead weight and is only allocated if a layer ever falls back
bool needs_f32_acts_upfront(const skrnn_handle *h) {
  return !(h->cfg.precision == SKRNN_PREC_BF16 && (sk_rnn_tc_eligible(h) || sk_rnn1_tc_eligible(h)));
}

int alloc_f32_acts(skrnn_handle *h) {
  SkWorkspace &w = h->ws;
  if (w.f32_acts || w.cap_rows == 0) return 0;
  for (int l = 0; l < h->L; ++l) {
    SK_CUDA(cudaMalloc(&w.xw[l], w.cap_rows * h->G * sizeof(float)));
    SK_CUDA(cudaMalloc(&w.hprev[l], w.cap_rows * h->H * sizeof(float)));
    SK_CUDA(cudaMalloc(&w.out[l], w.cap_rows * h->H * sizeof(float)));
  }
  w.f32_acts = true;
  return 0;
}

int ensure_workspace(skrnn_handle *h, int B, int T) {
  const int64_t rows = (int64_t)B * T;
  const bool tt = h->cfg.precision == SKRNN_PREC_BF16 && h->cfg.cell != SKRNN_CELL_LSTM;
  const int64_t rows_tt = tt ? (int64_t)T * ((B + 15) / 16) * 16 : 0;  // batch padded to the 16-row tiles of rnn1_tc.cu
  SkWorkspace &w = h->ws;
  if (rows <= w.cap_rows && rows_tt <= w.tt_rows && (w.bptt_carry == nullptr || B <= w.carry_B)) return 0;
  SK_CUDA(cudaDeviceSynchronize());
  free_workspace(w);
  SK_CUDA(cudaMalloc(&w.x_tm, rows * 5 * sizeof(float)));
  SK_CUDA(cudaMalloc(&w.eoc, rows));
  for (int l = 0; l < h->L; ++l)
    if (h->cfg.cell != SKRNN_CELL_RNN) SK_CUDA(cudaMalloc(&w.cprev[l], rows * h->H * sizeof(float)));
  SK_CUDA(cudaMalloc(&w.dout, rows * h->H * sizeof(float)));
  SK_CUDA(cudaMalloc(&w.logits, rows * h->NOUT * sizeof(float)));
  if (h->cfg.precision == SKRNN_PREC_BF16) {
    const size_t e = sizeof(__nv_bfloat16);
    SK_CUDA(cudaMalloc(&w.x8_bf, rows * 8 * e));
    SK_CUDA(cudaMemset(w.x8_bf, 0, rows * 8 * e));
    SK_CUDA(cudaMalloc(&w.dpre_bf_buf[0], rows * h->G * e));
    SK_CUDA(cudaMalloc(&w.dpre_bf_buf[1], rows * h->G * e));
    w.dpre_bf = w.dpre_bf_buf[0];
    SK_CUDA(cudaMalloc(&w.dlogits_bf, rows * h->NOUTP * e));
    SK_CUDA(cudaMemset(w.dlogits_bf, 0, rows * h->NOUTP * e));
    for (int l = 0; l < h->L; ++l) {
      SK_CUDA(cudaMalloc(&w.out_bf[l], rows * h->H * e));
      SK_CUDA(cudaMalloc(&w.hprev_bf[l], rows * h->H * e));
      if (h->cfg.cell == SKRNN_CELL_GRU) SK_CUDA(cudaMalloc(&w.aux_bf[l], rows * h->H * e));
      // gate activations saved by the tcgen05 recurrent kernels (gru / rnn: tile-transposed, gru with the fed state as a 4th plane)
      if (!tt) SK_CUDA(cudaMalloc(&w.acts_bf[l], rows * h->G * e));
      else SK_CUDA(cudaMalloc(&w.acts_bf[l], rows_tt * (h->cfg.cell == SKRNN_CELL_GRU ? 4 * h->H : h->H) * e));
      SK_CUDA(cudaMalloc(&w.xw_il[l], rows * h->G * e));                    // hoisted projection in bf16
    }
  }
  if (tt) {
    SK_CUDA(cudaMalloc(&w.xw_tt, rows_tt * h->G * sizeof(__nv_bfloat16)));
    SK_CUDA(cudaMalloc(&w.dout_tt, rows_tt * h->H * sizeof(float)));
    w.tt_rows = rows_tt;
  }
  w.cap_rows = rows;
  if (h->cfg.precision == SKRNN_PREC_BF16 && h->cfg.cell == SKRNN_CELL_LSTM) {
    cudaFree(w.bptt_carry);
    w.bptt_carry = nullptr;
    SK_CUDA(cudaMalloc(&w.bptt_carry, (size_t)2 * B * h->H * sizeof(float)));
    w.carry_B = B;
  }
  if (needs_f32_acts_upfront(h)) SK_TRY(alloc_f32_acts(h));
  return 0;
}

int check_shape(skrnn_handle *h, int B, int T) {
  SK_CHECK(h != nullptr, "null handle");
  SK_CHECK(B > 0 && T > 0, "batch %d and sequence length %d must be positive", B, T);
  return ensure_workspace(h, B, T);
}

// A GEMM operand in both precisions: the fp32 buffer and (bf16 mode) its bf16 mirror.
struct Op {
  const float *f;
  int ldf;
  const __nv_bfloat16 *b;
  int ldb;
};
Op col_offset(Op o, int c) { return Op{o.f + c, o.ldf, o.b ? o.b + c : nullptr, o.ldb}; }
Op layer_input(skrnn_handle *h, int l) {
  if (l == 0) return Op{h->ws.x_tm, 5, h->ws.x8_bf, 8};
  return Op{h->ws.out[l - 1], h->H, h->ws.out_bf[l - 1], h->H};
}
Op weight(skrnn_handle *h, int l, int which) {
  const int nc = h->lp[l].ncol[which];
  return Op{h->params + h->lp[l].w[which], nc, h->params_bf ? h->params_bf + h->pbf_w[l][which] : nullptr, nc};
}
Op out_weight(skrnn_handle *h) {
  return Op{h->params + h->off_out_w, h->NOUT, h->params_bf ? h->params_bf + h->pbf_out_w : nullptr, h->NOUTP};
}
Op top_output(skrnn_handle *h) { return Op{h->ws.out[h->L - 1], h->H, h->ws.out_bf[h->L - 1], h->H}; }
Op hprev_op(skrnn_handle *h, int l) { return Op{h->ws.hprev[l], h->H, h->ws.hprev_bf[l], h->H}; }
Op aux_op(skrnn_handle *h, int l) { return Op{h->ws.cprev[l], h->H, h->ws.aux_bf[l], h->H}; }
Op dpre_op(skrnn_handle *h, int l) { return Op{h->ws.xw[l], h->G, h->ws.dpre_bf, h->G}; }
Op dlogits_op(skrnn_handle *h) { return Op{h->ws.logits, h->NOUT, h->ws.dlogits_bf, h->NOUTP}; }

// C[M,N] (+)= A*B (+bias).  a_kc: A stored [M][ld] (K contiguous) else [K][ld]; b_nc: B stored [K][ld]
// (N contiguous) else [N][ld].  fp32 path: CUDA-core GEMM; bf16 path: tcgen05/TMA GEMM.
int gemm(skrnn_handle *h, cudaStream_t st, int M, int N, int K, Op A, bool a_kc, Op B, bool b_nc, float *C,
         int ldc, const float *bias, bool accumulate) {
  if (h->cfg.precision == SKRNN_PREC_BF16)
    return sk_gemm_tc(h, st, M, N, K, A.b, A.ldb, a_kc, B.b, B.ldb, !b_nc, C, ldc, bias, accumulate);
  return sk_gemm_f32(h, st, M, N, K, A.f, A.ldf, a_kc, B.f, B.ldf, b_nc, C, ldc, bias, accumulate);
}

// d(bias) = column sums of d(pre-activation); the bf16 path reads the bf16 mirror (the tcgen05 BPTT kernel
// does not write an fp32 copy)
int bias_grad(skrnn_handle *h, cudaStream_t st, Op d, int64_t rows, int cols, float *out) {
  if (h->cfg.precision == SKRNN_PREC_BF16) return sk_colsum_bf16(h, st, d.b, rows, cols, d.ldb, out);
  return sk_colsum(h, st, d.f, rows, cols, d.ldf, out);
}

// first step of time range c when T steps are cut into C ranges (even, so that barrier parities follow t)
int range_begin(int T, int C, int c) { return c >= C ? T : (int)((int64_t)T * c / C) & ~1; }

// forward unroll of the whole stack + output projection (model.py:97-99)
int run_forward(skrnn_handle *h, cudaStream_t st, int B, int T, const float *x, const float *state_in,
                float *state_out, const float *mask, uint64_t seed, int row0, bool training, bool want_logits = true) {
  const int H = h->H, G = h->G;
  const int64_t R = (int64_t)B * T;
  SK_TRY(sk_cast_params(h, st));
  SK_TRY(sk_transpose_x(h, st, B, T, x));
  for (int l = 0; l < h->L; ++l) {
    const SkLayerParams &lp = h->lp[l];
    const Op Xo = layer_input(h, l);
    h->xw_il_ok[l] = false;
    if (h->cfg.cell != SKRNN_CELL_LSTM && sk_rnn1_tc_eligible(h) && h->ws.xw_il[l] != nullptr) {
      // tcgen05 gru / rnn kernels: the projection in bf16, TF's column order (half the bytes the recurrent kernel streams)
      const int nmat = h->cfg.cell == SKRNN_CELL_GRU ? 2 : 1;
      for (int wi = 0; wi < nmat; ++wi) {
        const Op W = weight(h, l, wi);
        SK_TRY(sk_gemm_tc_bf16out(h, st, (int)R, lp.ncol[wi], lp.I, Xo.b, Xo.ldb, W.b, W.ldb, h->ws.xw_il[l] + (wi ? 2 * H : 0), G,
                                  h->params + lp.b[wi]));
      }
      h->xw_il_ok[l] = true;
      SK_TRY(sk_rnn_forward_layer(h, st, l, B, T, x, state_in + (size_t)l * B * h->S, state_out + (size_t)l * B * h->S, mask,
                                  seed, row0, training));
      if (h->fwd_family[l] == 3) continue;
      h->xw_il_ok[l] = false;  // declined: redo the layer through the fp32-format projection and the fp32 kernels
      SK_TRY(alloc_f32_acts(h));
    }
    if (h->cfg.cell == SKRNN_CELL_GRU) {
      SK_TRY(gemm(h, st, (int)R, 2 * H, lp.I, Xo, true, weight(h, l, 0), true, h->ws.xw[l], G, h->params + lp.b[0], false));
      SK_TRY(gemm(h, st, (int)R, H, lp.I, Xo, true, weight(h, l, 1), true, h->ws.xw[l] + 2 * H, G, h->params + lp.b[1], false));
    } else {
      const int C = (h->ws.xw_il[l] != nullptr && h->wx_il[l] != nullptr) ? sk_rnn_tc_ranges(h, B, T) : 1;
      if (C > 1) {
        // The layer in C time ranges: the projection of range c runs on the side stream as soon as the layer below
        // has produced that range, i.e. next to the recurrence of the layer below (whose clusters leave a quarter of
        // the SMs free), and this layer's recurrent launch for range c only waits for its own projection.
        const int ip = (lp.I + 7) / 8 * 8;
        h->xw_il_ok[l] = true;
        if (l == 0) SK_CUDA(cudaEventRecord(h->ev_main[0], st));  // x8_bf and the weight copies are ready
        for (int c = 0; c < C; ++c) {
          const int t0 = range_begin(T, C, c), t1 = range_begin(T, C, c + 1);
          SK_CUDA(cudaStreamWaitEvent(h->side, l == 0 ? h->ev_main[0] : h->ev_chunk[1][c], 0));
          SK_TRY(sk_gemm_tc_bf16out(h, h->side, (t1 - t0) * B, G, ip, Xo.b + (size_t)t0 * B * Xo.ldb, Xo.ldb, h->wx_il[l], G,
                                    h->ws.xw_il[l] + (size_t)t0 * B * G, G, h->bias_il[l]));
          SK_CUDA(cudaEventRecord(h->ev_chunk[0][c], h->side));
        }
        for (int c = 0; c < C; ++c) {
          SK_CUDA(cudaStreamWaitEvent(st, h->ev_chunk[0][c], 0));
          h->chunk_t0 = range_begin(T, C, c);
          h->chunk_t1 = range_begin(T, C, c + 1);
          const int rc = sk_rnn_forward_layer(h, st, l, B, T, x, state_in + (size_t)l * B * h->S,
                                              state_out + (size_t)l * B * h->S, mask, seed, row0, training);
          h->chunk_t0 = h->chunk_t1 = 0;
          SK_TRY(rc);
          SK_CHECK(h->fwd_family[l] == 1, "layer %d: the tcgen05 recurrent kernels declined a time range", l);
          SK_CUDA(cudaEventRecord(h->ev_chunk[1][c], st));
        }
        h->fwd_ranges[l] = C;
        continue;
      }
      h->fwd_ranges[l] = 1;
      if (sk_rnn_tc_eligible(h) && h->ws.xw_il[l] != nullptr && h->wx_il[l] != nullptr) {
        // tcgen05 lstm kernels: bf16 projection with the four gates of a unit adjacent (half the bytes of the fp32 one)
        const int ip = (lp.I + 7) / 8 * 8;
        SK_TRY(sk_gemm_tc_bf16out(h, st, (int)R, G, ip, Xo.b, Xo.ldb, h->wx_il[l], G, h->ws.xw_il[l], G, h->bias_il[l]));
        h->xw_il_ok[l] = true;
        SK_TRY(sk_rnn_forward_layer(h, st, l, B, T, x, state_in + (size_t)l * B * h->S,
                                    state_out + (size_t)l * B * h->S, mask, seed, row0, training));
        if (h->fwd_family[l] == 1) continue;
        h->xw_il_ok[l] = false;  // declined (shape, launch failure): redo the layer through the fp32-format projection
        if (!h->warned_tc_fallback) {  // a large performance cliff: say so, once per handle
          h->warned_tc_fallback = true;
          fprintf(stderr, "libskrnn: layer %d (H=%d, B=%d) does not run on the tcgen05 recurrent kernels; this handle "
                          "falls back to the fp32 CUDA-core recurrent kernels between tensor-core GEMMs\n", l, H, B);
        }
        SK_TRY(alloc_f32_acts(h));
      }
      SK_TRY(gemm(h, st, (int)R, G, lp.I, Xo, true, weight(h, l, 0), true, h->ws.xw[l], G, h->params + lp.b[0], false));
    }
    SK_TRY(sk_rnn_forward_layer(h, st, l, B, T, x, state_in + (size_t)l * B * h->S,
                                state_out + (size_t)l * B * h->S, mask, seed, row0, training));
  }
  if (want_logits)
    SK_TRY(gemm(h, st, (int)R, h->NOUT, H, top_output(h), true, out_weight(h), true, h->ws.logits, h->NOUT,
                h->params + h->off_out_b, false));
  return 0;
}

}  // namespace

int sk_ensure_f32_acts(skrnn_handle *h) { return alloc_f32_acts(h); }

extern "C" {

const char *skrnn_last_error(void) { return g_err; }
const char *skrnn_version(void) { return "skrnn 0.1 sm_100a"; }

int skrnn_create(const skrnn_config *cfg, skrnn_handle **out) {
  SK_CHECK(cfg != nullptr && out != nullptr, "skrnn_create: null argument");
  *out = nullptr;
  if (cfg->cell < 0 || cfg->cell > 2) {
    sk_set_error("model type not supported: %d", cfg->cell);  // model.py:21
    return 1;
  }
  SK_CHECK(cfg->rnn_size > 0 && cfg->rnn_size % 8 == 0, "rnn_size %d must be a positive multiple of 8", cfg->rnn_size);
  // every kernel family keeps a unit slice of the recurrent weights resident per CTA of one cluster; the largest
  // plan (cluster of 8 x 64 units fp32, cluster of 16 x 32 units tcgen05, sampler with one thread per unit) ends at 512
  SK_CHECK(cfg->rnn_size <= SKRNN_MAX_RNN_SIZE, "rnn_size %d is not supported: this library runs rnn_size <= %d",
           cfg->rnn_size, SKRNN_MAX_RNN_SIZE);
  SK_CHECK(cfg->num_layers >= 1 && cfg->num_layers <= SK_MAX_LAYERS, "num_layers %d out of range", cfg->num_layers);
  SK_CHECK(cfg->num_mixture >= 1 && cfg->num_mixture <= 128, "num_mixture %d out of range [1,128]", cfg->num_mixture);
  SK_CHECK(cfg->keep_prob > 0.f && cfg->keep_prob <= 1.f, "keep_prob %f out of (0,1]", cfg->keep_prob);
  SK_CHECK(cfg->precision == SKRNN_PREC_F32 || cfg->precision == SKRNN_PREC_BF16, "unknown precision %d", cfg->precision);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    sk_set_error("libskrnn needs an NVIDIA B200 (sm_100) GPU; there is no CPU fallback: %s",
                 e != cudaSuccess ? cudaGetErrorString(e) : "no CUDA device");
    return 2;
  }
  SK_CHECK(cfg->device >= 0 && cfg->device < ndev, "device %d out of range (have %d)", cfg->device, ndev);
  SkDeviceGuard guard(cfg->device);
  cudaDeviceProp prop;
  SK_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  SK_CHECK(prop.major == 10, "libskrnn is built for sm_100a only; device %d is sm_%d%d", cfg->device, prop.major, prop.minor);

  skrnn_handle *h = new skrnn_handle();
  h->cfg = *cfg;
  h->H = cfg->rnn_size;
  h->L = cfg->num_layers;
  h->M = cfg->num_mixture;
  h->NOUT = 3 + 6 * h->M;
  h->NGATE = cfg->cell == SKRNN_CELL_LSTM ? 4 : (cfg->cell == SKRNN_CELL_GRU ? 3 : 1);
  h->G = h->NGATE * h->H;
  h->S = cfg->cell == SKRNN_CELL_LSTM ? 2 * h->H : h->H;
  h->num_sms = prop.multiProcessorCount;
  h->cluster_size = 8;
  build_layout(h);
  auto fail = [&](void) { skrnn_destroy(h); return 1; };
  size_t pb = (size_t)h->P * sizeof(float);
  if (cudaMalloc(&h->params, pb) != cudaSuccess || cudaMalloc(&h->grads, pb) != cudaSuccess ||
      cudaMalloc(&h->adam_m, pb) != cudaSuccess || cudaMalloc(&h->adam_v, pb) != cudaSuccess ||
      cudaMalloc(&h->scal, 8 * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&h->sync_ctr, 4 * sizeof(unsigned int)) != cudaSuccess) {
    sk_set_error("skrnn_create: cudaMalloc of %zu B x4 failed", pb);
    return fail();
  }
  cudaMemset(h->params, 0, pb);
  cudaMemset(h->grads, 0, pb);
  cudaMemset(h->adam_m, 0, pb);
  cudaMemset(h->adam_v, 0, pb);
  cudaMemset(h->scal, 0, 8 * sizeof(double));
  cudaMemset(h->sync_ctr, 0, 4 * sizeof(unsigned int));
  h->NOUTP = (h->NOUT + 7) / 8 * 8;
  if (cfg->precision == SKRNN_PREC_BF16) {
    // bf16 weight mirror: every tensor starts on a 16-byte boundary; output_w rows padded to NOUTP
    std::vector<int64_t> seg;
    int64_t o = 0;
    auto add = [&](int64_t src, int rows, int cols, int ldd, int64_t *dst) {
      *dst = o;
      seg.insert(seg.end(), {src, o, (int64_t)rows * cols, (int64_t)cols, (int64_t)ldd});
      o += ((int64_t)rows * ldd + 7) / 8 * 8;
    };
    add(h->off_out_w, h->H, h->NOUT, h->NOUTP, &h->pbf_out_w);
    for (int l = 0; l < h->L; ++l)
      for (int w = 0; w < (cfg->cell == SKRNN_CELL_GRU ? 2 : 1); ++w)
        add(h->lp[l].w[w], h->lp[l].I + h->H, h->lp[l].ncol[w], h->lp[l].ncol[w], &h->pbf_w[l][w]);
    h->pbf_count = o;
    h->cast_nseg = (int)(seg.size() / 5);
    if (cudaMalloc(&h->params_bf, o * sizeof(__nv_bfloat16)) != cudaSuccess ||
        cudaMalloc(&h->cast_seg_dev, seg.size() * sizeof(int64_t)) != cudaSuccess) {
      sk_set_error("skrnn_create: cudaMalloc of the bf16 weight mirror failed");
      return fail();
    }
    cudaMemset(h->params_bf, 0, o * sizeof(__nv_bfloat16));
    cudaMemcpy(h->cast_seg_dev, seg.data(), seg.size() * sizeof(int64_t), cudaMemcpyHostToDevice);
  }
  if (cfg->precision == SKRNN_PREC_BF16 && cfg->cell == SKRNN_CELL_LSTM) {
    for (int l = 0; l < h->L; ++l) {
      const int ip = (h->lp[l].I + 7) / 8 * 8;
      if (cudaMalloc(&h->wx_il[l], (size_t)ip * h->G * sizeof(__nv_bfloat16)) != cudaSuccess ||
          cudaMalloc(&h->bias_il[l], (size_t)h->G * sizeof(float)) != cudaSuccess) {
        sk_set_error("skrnn_create: cudaMalloc of the interleaved input weights failed");
        return fail();
      }
    }
  }
  if (cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&h->side2, cudaStreamNonBlocking) != cudaSuccess) {
    sk_set_error("skrnn_create: cudaStreamCreate failed");
    return fail();
  }
  for (int i = 0; i < 2; ++i) {
    cudaEventCreateWithFlags(&h->ev_main[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_side[i], cudaEventDisableTiming);
  }
  for (int i = 0; i <= h->L; ++i) cudaEventCreateWithFlags(&h->ev_grad[i], cudaEventDisableTiming);
  for (int k = 0; k < 2; ++k)
    for (int c = 0; c < 8; ++c) cudaEventCreateWithFlags(&h->ev_chunk[k][c], cudaEventDisableTiming);
  h->ws.wpack_floats = (int64_t)h->cluster_size * h->H * 256;
  if (cudaMalloc(&h->ws.wpack, h->ws.wpack_floats * sizeof(float)) != cudaSuccess) {
    sk_set_error("skrnn_create: cudaMalloc of weight pack failed");
    return fail();
  }
  if (cfg->batch_size > 0 && cfg->seq_length > 0 && ensure_workspace(h, cfg->batch_size, cfg->seq_length) != 0)
    return fail();
  if (cudaDeviceSynchronize() != cudaSuccess) {
    sk_set_error("skrnn_create: device error after initialisation");
    return fail();
  }
  *out = h;
  return 0;
}

void skrnn_destroy(skrnn_handle *h) {
  if (!h) return;
  SkDeviceGuard guard(h->cfg.device);
  cudaDeviceSynchronize();
  free_workspace(h->ws);
  cudaFree(h->ws.wpack);
  cudaFree(h->ws.partial);
  h->ws.partial = nullptr;
  h->ws.partial_floats = 0;
  cudaFree(h->params);
  cudaFree(h->grads);
  cudaFree(h->adam_m);
  cudaFree(h->adam_v);
  cudaFree(h->scal);
  cudaFree(h->sync_ctr);
  cudaFree(h->st_x);
  cudaFree(h->st_y);
  cudaFree(h->st_state);
  cudaFree(h->st_loss);
  cudaFree(h->samp_state);
  if (h->samp_graph) cudaGraphExecDestroy(h->samp_graph);
  cudaFree(h->samp_tc);
  cudaFree(h->params_bf);
  cudaFree(h->cast_seg_dev);
  cudaFreeHost(h->pin_x);
  cudaFreeHost(h->pin_y);
  cudaFreeHost(h->pin_state);
  cudaFreeHost(h->pin_loss);
  for (int l = 0; l < SK_MAX_LAYERS; ++l) { cudaFree(h->wx_il[l]); cudaFree(h->bias_il[l]); }
  if (h->side) cudaStreamDestroy(h->side);
  if (h->side2) cudaStreamDestroy(h->side2);
  for (int k = 0; k < 2; ++k)
    for (int c = 0; c < 8; ++c)
      if (h->ev_chunk[k][c]) cudaEventDestroy(h->ev_chunk[k][c]);
  cudaFree(h->ws.bptt_carry);
  for (int i = 0; i < 2; ++i) {
    if (h->ev_main[i]) cudaEventDestroy(h->ev_main[i]);
    if (h->ev_side[i]) cudaEventDestroy(h->ev_side[i]);
  }
  for (int i = 0; i <= SK_MAX_LAYERS; ++i)
    if (h->ev_grad[i]) cudaEventDestroy(h->ev_grad[i]);
  for (auto &r : h->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (auto e : h->prof_pool) cudaEventDestroy(e);
  delete h;
}

int64_t skrnn_param_count(const skrnn_handle *h) { return h ? h->P : -1; }
int32_t skrnn_num_tensors(const skrnn_handle *h) { return h ? (int32_t)h->tensors.size() : -1; }

int skrnn_tensor_info(const skrnn_handle *h, int32_t index, char *name_buf, int32_t name_cap,
                      int64_t *offset, int32_t *rows, int32_t *cols) {
  SK_CHECK(h != nullptr, "null handle");
  SK_CHECK(index >= 0 && index < (int32_t)h->tensors.size(), "tensor index %d out of range", index);
  const SkTensor &t = h->tensors[index];
  if (name_buf && name_cap > 0) {
    strncpy(name_buf, t.name.c_str(), name_cap - 1);
    name_buf[name_cap - 1] = 0;
  }
  if (offset) *offset = t.offset;
  if (rows) *rows = t.rows;
  if (cols) *cols = t.cols;
  return 0;
}

int skrnn_set_params(skrnn_handle *h, const float *host_params, int64_t count) {
  SK_CHECK(h && host_params, "null argument");
  SK_CHECK(count == h->P, "set_params: expected %lld floats, got %lld", (long long)h->P, (long long)count);
  SkDeviceGuard guard(h->cfg.device);
  SK_CUDA(cudaMemcpy(h->params, host_params, count * sizeof(float), cudaMemcpyHostToDevice));
  return 0;
}

int skrnn_get_params(skrnn_handle *h, float *host_params, int64_t count) {
  SK_CHECK(h && host_params, "null argument");
  SK_CHECK(count == h->P, "get_params: expected %lld floats, got %lld", (long long)h->P, (long long)count);
  SkDeviceGuard guard(h->cfg.device);
  SK_CUDA(cudaMemcpy(host_params, h->params, count * sizeof(float), cudaMemcpyDeviceToHost));
  return 0;
}

int skrnn_set_adam_state(skrnn_handle *h, const float *host_m, const float *host_v, int64_t count,
                         double beta1_power, double beta2_power) {
  SK_CHECK(h, "null handle");
  SK_CHECK(count == h->P, "set_adam_state: expected %lld floats, got %lld", (long long)h->P, (long long)count);
  SkDeviceGuard guard(h->cfg.device);
  if (host_m) SK_CUDA(cudaMemcpy(h->adam_m, host_m, count * sizeof(float), cudaMemcpyHostToDevice));
  if (host_v) SK_CUDA(cudaMemcpy(h->adam_v, host_v, count * sizeof(float), cudaMemcpyHostToDevice));
  h->beta1_power = beta1_power;
  h->beta2_power = beta2_power;
  return 0;
}

int skrnn_get_adam_state(skrnn_handle *h, float *host_m, float *host_v, int64_t count,
                         double *beta1_power, double *beta2_power) {
  SK_CHECK(h, "null handle");
  SK_CHECK(count == h->P, "get_adam_state: expected %lld floats, got %lld", (long long)h->P, (long long)count);
  SkDeviceGuard guard(h->cfg.device);
  if (host_m) SK_CUDA(cudaMemcpy(host_m, h->adam_m, count * sizeof(float), cudaMemcpyDeviceToHost));
  if (host_v) SK_CUDA(cudaMemcpy(host_v, h->adam_v, count * sizeof(float), cudaMemcpyDeviceToHost));
  if (beta1_power) *beta1_power = h->beta1_power;
  if (beta2_power) *beta2_power = h->beta2_power;
  return 0;
}

float *skrnn_params_dev(skrnn_handle *h) { return h ? h->params : nullptr; }
float *skrnn_grads_dev(skrnn_handle *h) { return h ? h->grads : nullptr; }
int64_t skrnn_launch_count(const skrnn_handle *h) { return h ? h->launches : -1; }

int skrnn_debug_gemm_bf16(skrnn_handle *h, int32_t M, int32_t N, int32_t K, const void *a_dev, int32_t lda,
                          int32_t a_kmajor, const void *b_dev, int32_t ldb, int32_t b_kmajor, float *c_dev,
                          int32_t ldc, const float *bias_dev, int32_t accumulate, const uint32_t *overrides,
                          void *stream) {
  SK_CHECK(h && a_dev && b_dev && c_dev, "debug_gemm: null argument");
  SkDeviceGuard guard(h->cfg.device);
  for (int i = 0; i < 6; ++i) g_tc_override[i] = overrides ? overrides[i] : 0;
  int rc = sk_gemm_tc(h, (cudaStream_t)stream, M, N, K, a_dev, lda, a_kmajor != 0, b_dev, ldb, b_kmajor != 0,
                      c_dev, ldc, bias_dev, accumulate != 0);
  for (int i = 0; i < 6; ++i) g_tc_override[i] = 0;
  return rc;
}

uint32_t skrnn_crc32c(uint32_t crc, const void *data, size_t n) {
  static uint32_t table[256];
  static bool init = false;
  if (!init) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
      table[i] = c;
    }
    init = true;
  }
  const unsigned char *p = static_cast<const unsigned char *>(data);
  uint32_t c = crc ^ 0xFFFFFFFFu;
  for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

int skrnn_profile_enable(skrnn_handle *h, int32_t on) {
  SK_CHECK(h != nullptr, "null handle");
  h->prof_on = on != 0;
  return 0;
}

int skrnn_profile_read(skrnn_handle *h, double *ms_out, int64_t *count_out) {
  SK_CHECK(h != nullptr && ms_out && count_out, "null argument");
  SkDeviceGuard guard(h->cfg.device);
  SK_CUDA(cudaDeviceSynchronize());
  for (int i = 0; i < SK_NUM_CAT; ++i) { ms_out[i] = 0.0; count_out[i] = 0; }
  for (auto &r : h->prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { ms_out[r.cat] += ms; count_out[r.cat]++; }
    h->prof_pool.push_back(r.a);
    h->prof_pool.push_back(r.b);
  }
  h->prof.clear();
  return 0;
}

int skrnn_forward(skrnn_handle *h, int32_t B, int32_t T, const float *x_dev, const float *state_in_dev,
                  float *mdn_out_dev, float *state_out_dev, void *stream) {
  SK_CHECK(h != nullptr, "null handle");
  SkDeviceGuard guard(h->cfg.device);
  SK_TRY(check_shape(h, B, T));
  SK_CHECK(x_dev && state_in_dev && state_out_dev, "forward: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  SK_TRY(run_forward(h, st, B, T, x_dev, state_in_dev, state_out_dev, nullptr, 0, 0, false));
  if (mdn_out_dev) {
    SK_CUDA(cudaMemsetAsync(h->scal, 0, 3 * sizeof(double), st));
    SK_TRY(sk_head(h, st, B, T, nullptr, mdn_out_dev, false, (int64_t)B * T));
  }
  return 0;
}

int skrnn_loss_grad(skrnn_handle *h, int32_t B, int32_t T, const float *x_dev, const float *y_dev,
                    const float *state_in_dev, const float *dropout_mask_dev, uint64_t seed, int32_t row0,
                    int64_t loss_rows, float *losses_dev, float *state_out_dev, void *stream) {
  SK_CHECK(h != nullptr, "null handle");
  SkDeviceGuard guard(h->cfg.device);
  SK_TRY(check_shape(h, B, T));
  SK_CHECK(x_dev && y_dev && state_in_dev && state_out_dev && losses_dev, "loss_grad: null buffer");
  SK_CHECK(loss_rows >= (int64_t)B * T, "loss_rows %lld smaller than B*T", (long long)loss_rows);
  cudaStream_t st = (cudaStream_t)stream;
  const int H = h->H, G = h->G, NOUT = h->NOUT;
  const int R = B * T;
  SK_CUDA(cudaMemsetAsync(h->grads, 0, (size_t)h->P * sizeof(float), st));
  SK_CUDA(cudaMemsetAsync(h->scal, 0, 3 * sizeof(double), st));
  // bf16 path: the output layer, the head and d(logits) are one kernel and the logits never reach memory
  const bool fused_head = sk_head_tc_eligible(h);
  SK_TRY(run_forward(h, st, B, T, x_dev, state_in_dev, state_out_dev, dropout_mask_dev, seed, row0, true, !fused_head));
  if (fused_head) SK_TRY(sk_head_tc(h, st, B, T, y_dev, loss_rows));
  else SK_TRY(sk_head(h, st, B, T, y_dev, nullptr, true, loss_rows));
  SK_TRY(sk_finalize_losses(h, st, losses_dev));
  // head backward: d output_w, d output_b, d(top layer output)
  SK_TRY(gemm(h, st, H, NOUT, R, top_output(h), false, dlogits_op(h), true, h->grads + h->off_out_w, NOUT, nullptr, true));
  if (fused_head) SK_TRY(sk_colsum_bf16(h, st, h->ws.dlogits_bf, R, NOUT, h->NOUTP, h->grads + h->off_out_b));
  else SK_TRY(sk_colsum(h, st, h->ws.logits, R, NOUT, NOUT, h->grads + h->off_out_b));
  SK_CUDA(cudaEventRecord(h->ev_grad[h->L], st));  // output_w / output_b gradients (and the losses) are final
  SK_TRY(gemm(h, st, R, H, NOUT, dlogits_op(h), true, out_weight(h), false, h->ws.dout, H, nullptr, false));
  // Weight gradients are off the critical path: on the bf16 path layer l's dW products are enqueued on a side
  // stream so that they run next to layer l-1's BPTT kernel (which leaves SMs idle: its clusters do not tile the
  // GPU).  They are handed to the side stream only after that BPTT launch has been enqueued on `st`, so its
  // clusters are dispatched first.  d(pre) is double-buffered for this.
  static int ov_env = -1;
  if (ov_env < 0) { const char *e = getenv("SKRNN_WGRAD_OVERLAP"); ov_env = (e && e[0] == '0') ? 0 : 1; }
  const bool overlap = ov_env && h->cfg.precision == SKRNN_PREC_BF16 && h->cfg.cell != SKRNN_CELL_GRU && h->L > 1;
  struct Deferred { int l = -1; int buf = 0; bool bias_done = false; } def;
  bool side_busy[2] = {false, false};
  auto wgrad = [&](cudaStream_t s2, int l, int buf, bool bias_done) -> int {
    const SkLayerParams &lp = h->lp[l];
    Op dpre = dpre_op(h, l);
    if (h->ws.dpre_bf_buf[buf]) dpre.b = h->ws.dpre_bf_buf[buf];
    float *gW = h->grads + lp.w[0];
    // the tcgen05 lstm kernels leave d(pre) with gate-interleaved columns: the split-K epilogue scatters dW back
    h->gemm_perm_H = (h->fwd_family[l] == 1) ? H : 0;
    int rc = gemm(h, s2, lp.I, G, R, layer_input(h, l), false, dpre, true, gW, G, nullptr, true);
    if (rc == 0) rc = gemm(h, s2, H, G, R, hprev_op(h, l), false, dpre, true, gW + (size_t)lp.I * G, G, nullptr, true);
    h->gemm_perm_H = 0;
    SK_TRY(rc);
    if (!bias_done) SK_TRY(bias_grad(h, s2, dpre, R, G, h->grads + lp.b[0]));
    SK_CUDA(cudaEventRecord(h->ev_grad[l], s2));  // layer l's gradient tensors are final
    return 0;
  };
  auto flush_deferred = [&]() -> int {
    if (def.l < 0) return 0;
    SK_CUDA(cudaStreamWaitEvent(h->side, h->ev_main[def.buf], 0));
    SK_TRY(wgrad(h->side, def.l, def.buf, def.bias_done));
    SK_CUDA(cudaEventRecord(h->ev_side[def.buf], h->side));
    side_busy[def.buf] = true;
    def.l = -1;
    return 0;
  };
  for (int l = h->L - 1; l >= 0; --l) {
    const SkLayerParams &lp = h->lp[l];
    h->bias_grad_done = false;
    const int buf = overlap ? ((h->L - 1 - l) & 1) : 0;
    if (h->ws.dpre_bf_buf[buf]) h->ws.dpre_bf = h->ws.dpre_bf_buf[buf];
    if (side_busy[buf]) {  // the dW products that read this d(pre) copy two layers ago must be done
      SK_CUDA(cudaStreamWaitEvent(st, h->ev_side[buf], 0));
      side_busy[buf] = false;
    }
    const int C = (h->cfg.cell == SKRNN_CELL_LSTM && h->fwd_family[l] == 1) ? h->fwd_ranges[l] : 1;
    const Op X = layer_input(h, l);
    const Op dpre = dpre_op(h, l);
    const int I = lp.I;
    if (C > 1) {
      // BPTT in the forward pass's time ranges, last range first.  d(out) of the layer below for range c is a GEMM on
      // the second side stream that runs next to this layer's BPTT of range c-1; the layer below starts its range c
      // as soon as that GEMM is done.
      for (int c = C - 1; c >= 0; --c) {
        const int t0 = range_begin(T, C, c), t1 = range_begin(T, C, c + 1);
        if (l < h->L - 1) SK_CUDA(cudaStreamWaitEvent(st, h->ev_chunk[0][c], 0));  // d(out) rows of this range
        h->chunk_t0 = t0;
        h->chunk_t1 = t1;
        const int rc = sk_rnn_backward_layer(h, st, l, B, T, x_dev, dropout_mask_dev, seed, row0, true);
        h->chunk_t0 = h->chunk_t1 = 0;
        SK_TRY(rc);
        if (c == C - 1) SK_TRY(flush_deferred());  // the previous layer's dW: enqueued after this layer's first BPTT launch
        if (l > 0) {
          SK_CUDA(cudaEventRecord(h->ev_chunk[1][c], st));
          SK_CUDA(cudaStreamWaitEvent(h->side2, h->ev_chunk[1][c], 0));
          SK_TRY(sk_gemm_tc(h, h->side2, (t1 - t0) * B, I, G, dpre.b + (size_t)t0 * B * G, G, true, h->wx_il[l], G, true,
                            h->ws.dout + (size_t)t0 * B * H, H, nullptr, false));
          SK_CUDA(cudaEventRecord(h->ev_chunk[0][c], h->side2));
        }
      }
      if (overlap && l > 0) {
        SK_CUDA(cudaEventRecord(h->ev_main[buf], st));
        def.l = l; def.buf = buf; def.bias_done = h->bias_grad_done;
      } else {
        SK_TRY(wgrad(st, l, buf, h->bias_grad_done));
      }
      continue;
    }
    SK_TRY(sk_rnn_backward_layer(h, st, l, B, T, x_dev, dropout_mask_dev, seed, row0, true));
    SK_TRY(flush_deferred());  // the previous layer's dW: enqueued after this layer's BPTT launch
    if (h->cfg.cell == SKRNN_CELL_GRU) {
      const Op dg = dpre, dc = col_offset(dpre, 2 * H);
      float *gWg = h->grads + lp.w[0], *gWc = h->grads + lp.w[1];
      SK_TRY(gemm(h, st, I, 2 * H, R, X, false, dg, true, gWg, 2 * H, nullptr, true));
      SK_TRY(gemm(h, st, H, 2 * H, R, hprev_op(h, l), false, dg, true, gWg + (size_t)I * 2 * H, 2 * H, nullptr, true));
      if (!h->bias_grad_done) SK_TRY(bias_grad(h, st, dg, R, 2 * H, h->grads + lp.b[0]));
      SK_TRY(gemm(h, st, I, H, R, X, false, dc, true, gWc, H, nullptr, true));
      SK_TRY(gemm(h, st, H, H, R, aux_op(h, l), false, dc, true, gWc + (size_t)I * H, H, nullptr, true));
      if (!h->bias_grad_done) SK_TRY(bias_grad(h, st, dc, R, H, h->grads + lp.b[1]));
      SK_CUDA(cudaEventRecord(h->ev_grad[l], st));
      if (l > 0) {
        SK_TRY(gemm(h, st, R, I, 2 * H, dg, true, weight(h, l, 0), false, h->ws.dout, H, nullptr, false));
        SK_TRY(gemm(h, st, R, I, H, dc, true, weight(h, l, 1), false, h->ws.dout, H, nullptr, true));
      }
    } else {
      // d(out) of the layer below first: it is what the next BPTT launch waits for
      if (l > 0) {
        if (h->fwd_family[l] == 1)  // interleaved d(pre) against the interleaved copy of Wx (K order matches)
          SK_TRY(sk_gemm_tc(h, st, R, I, G, dpre.b, G, true, h->wx_il[l], G, true, h->ws.dout, H, nullptr, false));
        else
          SK_TRY(gemm(h, st, R, I, G, dpre, true, weight(h, l, 0), false, h->ws.dout, H, nullptr, false));
      }
      if (overlap && l > 0) {
        SK_CUDA(cudaEventRecord(h->ev_main[buf], st));
        def.l = l; def.buf = buf; def.bias_done = h->bias_grad_done;
      } else {
        SK_TRY(wgrad(st, l, buf, h->bias_grad_done));
      }
    }
  }
  SK_TRY(flush_deferred());
  for (int b2 = 0; b2 < 2; ++b2)
    if (side_busy[b2]) SK_CUDA(cudaStreamWaitEvent(st, h->ev_side[b2], 0));  // join
  return 0;
}

int skrnn_grad_segment(const skrnn_handle *h, int32_t index, int64_t *offset, int64_t *count) {
  SK_CHECK(h != nullptr && offset && count, "grad_segment: null argument");
  SK_CHECK(index >= 0 && index <= h->L, "grad_segment: index %d out of range [0,%d]", index, h->L);
  if (index == h->L) {  // output_w, output_b: the head of the flat vector
    *offset = 0;
    *count = h->lp[0].w[0];
  } else {
    *offset = h->lp[index].w[0];
    *count = (index + 1 < h->L ? h->lp[index + 1].w[0] : h->P) - h->lp[index].w[0];
  }
  return 0;
}

int skrnn_stream_wait_grad(skrnn_handle *h, int32_t index, void *stream) {
  SK_CHECK(h != nullptr, "null handle");
  SK_CHECK(index >= 0 && index <= h->L, "stream_wait_grad: index %d out of range [0,%d]", index, h->L);
  SkDeviceGuard guard(h->cfg.device);
  SK_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, h->ev_grad[index], 0));
  return 0;
}

int skrnn_apply_clip_adam(skrnn_handle *h, float lr, float *gnorm_out_dev, void *stream) {
  SK_CHECK(h != nullptr, "null handle");
  SkDeviceGuard guard(h->cfg.device);
  return sk_clip_adam(h, (cudaStream_t)stream, lr, gnorm_out_dev);
}

int skrnn_train_step_host(skrnn_handle *h, int32_t B, int32_t T, const float *x_host, const float *y_host,
                          float *state_host, float lr, uint64_t seed, float *losses_host) {
  SK_CHECK(h != nullptr, "null handle");
  SkDeviceGuard guard(h->cfg.device);
  SK_TRY(check_shape(h, B, T));
  SK_CHECK(x_host && y_host && state_host && losses_host, "train_step_host: null buffer");
  const int64_t rows = (int64_t)B * T;
  const int64_t nstate = (int64_t)h->L * B * h->S;
  if (rows > h->st_rows) {
    cudaFree(h->st_x); cudaFree(h->st_y); cudaFreeHost(h->pin_x); cudaFreeHost(h->pin_y);
    h->st_x = h->st_y = h->pin_x = h->pin_y = nullptr;
    h->st_rows = 0;
    SK_CUDA(cudaMalloc(&h->st_x, rows * 5 * sizeof(float)));
    SK_CUDA(cudaMalloc(&h->st_y, rows * 5 * sizeof(float)));
    SK_CUDA(cudaMallocHost(&h->pin_x, rows * 5 * sizeof(float)));
    SK_CUDA(cudaMallocHost(&h->pin_y, rows * 5 * sizeof(float)));
    h->st_rows = rows;
  }
  if (nstate > h->st_state_n) {
    cudaFree(h->st_state); cudaFreeHost(h->pin_state);
    h->st_state = h->pin_state = nullptr;
    h->st_state_n = 0;
    SK_CUDA(cudaMalloc(&h->st_state, nstate * sizeof(float)));
    SK_CUDA(cudaMallocHost(&h->pin_state, nstate * sizeof(float)));
    h->st_state_n = nstate;
  }
  if (!h->st_loss) {
    SK_CUDA(cudaMalloc(&h->st_loss, 4 * sizeof(float)));
    SK_CUDA(cudaMallocHost(&h->pin_loss, 4 * sizeof(float)));
  }
  cudaStream_t st = 0;
  // page-locked caller buffers are copied from directly; pageable ones go through the handle's pinned staging
  auto pinned = [](const void *p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
  };
  const float *src_x = x_host, *src_y = y_host;
  float *io_state = state_host;
  if (!pinned(x_host)) { memcpy(h->pin_x, x_host, rows * 5 * sizeof(float)); src_x = h->pin_x; }
  if (!pinned(y_host)) { memcpy(h->pin_y, y_host, rows * 5 * sizeof(float)); src_y = h->pin_y; }
  const bool state_pinned = pinned(state_host);
  if (!state_pinned) { memcpy(h->pin_state, state_host, nstate * sizeof(float)); io_state = h->pin_state; }
  SK_CUDA(cudaMemcpyAsync(h->st_x, src_x, rows * 5 * sizeof(float), cudaMemcpyHostToDevice, st));
  SK_CUDA(cudaMemcpyAsync(h->st_y, src_y, rows * 5 * sizeof(float), cudaMemcpyHostToDevice, st));
  SK_CUDA(cudaMemcpyAsync(h->st_state, io_state, nstate * sizeof(float), cudaMemcpyHostToDevice, st));
  SK_TRY(skrnn_loss_grad(h, B, T, h->st_x, h->st_y, h->st_state, nullptr, seed, 0, rows, h->st_loss,
                         h->st_state, st));
  SK_TRY(sk_clip_adam(h, st, lr, h->st_loss + 3));
  SK_CUDA(cudaMemcpyAsync(h->pin_loss, h->st_loss, 4 * sizeof(float), cudaMemcpyDeviceToHost, st));
  SK_CUDA(cudaMemcpyAsync(io_state, h->st_state, nstate * sizeof(float), cudaMemcpyDeviceToHost, st));
  SK_CUDA(cudaStreamSynchronize(st));
  memcpy(losses_host, h->pin_loss, 3 * sizeof(float));
  if (!state_pinned) memcpy(state_host, h->pin_state, nstate * sizeof(float));
  return 0;
}

int skrnn_sample(skrnn_handle *h, int32_t n_sketches, int32_t num_steps, float temp_mixture, float temp_pen,
                 int32_t stop_if_eoc, const float *uniforms_dev, uint64_t seed, int32_t sketch0,
                 float *strokes_dev, int32_t *lengths_dev, float *mix_params_dev, void *stream) {
  SK_CHECK(h != nullptr, "null handle");
  SK_CHECK(strokes_dev && lengths_dev, "sample: null output buffer");
  SK_CHECK(temp_mixture > 0.f && temp_pen > 0.f, "sample: temperatures must be positive");
  SkDeviceGuard guard(h->cfg.device);
  // bf16 handles with many sketches: per-step tensor-core GEMMs (sampler_tc.cu); SKRNN_SAMPLER_EXACT=1
  // keeps the bit-exact fp32 persistent kernel
  const char *ex = getenv("SKRNN_SAMPLER_EXACT");
  if (!(ex && ex[0] == '1')) {
    const int r = sk_sample_tc(h, (cudaStream_t)stream, n_sketches, num_steps, temp_mixture, temp_pen, stop_if_eoc,
                               uniforms_dev, seed, sketch0, strokes_dev, lengths_dev, mix_params_dev);
    if (r >= 0) return r;
  }
  return sk_sample_f32(h, (cudaStream_t)stream, n_sketches, num_steps, temp_mixture, temp_pen, stop_if_eoc,
                       uniforms_dev, seed, sketch0, strokes_dev, lengths_dev, mix_params_dev);
}

}  // extern "C"
