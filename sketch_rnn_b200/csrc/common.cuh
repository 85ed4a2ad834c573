// common.cuh -- internal declarations shared by the libskrnn translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <algorithm>
#include <string>
#include <vector>

#include "../../include/skrnn.h"

#define SK_MAX_LAYERS 8

void sk_set_error(const char *fmt, ...);

#define SK_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      sk_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

#define SK_CHECK(cond, ...)       \
  do {                            \
    if (!(cond)) {                \
      sk_set_error(__VA_ARGS__);  \
      return 1;                   \
    }                             \
  } while (0)

#define SK_TRY(expr)          \
  do {                        \
    int _r = (expr);          \
    if (_r != 0) return _r;   \
  } while (0)

struct SkTensor {
  std::string name;
  int64_t offset;
  int rows, cols;  // bias: rows = 1
};

// Per-layer addressing of the flat parameter / gradient vector (TF layouts, include/skrnn.h).
struct SkLayerParams {
  int I;            // input width (5 for layer 0 else H)
  int64_t w[2];     // Matrix offsets: lstm/rnn {W, -}, gru {Gates, Candidate}
  int64_t b[2];     // Bias offsets
  int ncol[2];      // Matrix column counts
};

struct SkWorkspace {   // time-major activations for one (B, T) capacity
  int64_t cap_rows = 0;  // T*B rows allocated
  bool f32_acts = false; // xw / hprev / out are allocated for cap_rows
  int capB = 0, capT = 0;
  int carry_B = 0;       // batch rows bptt_carry holds
  float *x_tm = nullptr;                    // [T,B,5]
  unsigned char *eoc = nullptr;             // [T,B]: x[b,t,3] > 0
  float *xw[SK_MAX_LAYERS] = {nullptr};     // [T,B,G]: input projection -> gate activations -> d(pre)
  float *hprev[SK_MAX_LAYERS] = {nullptr};  // [T,B,H]: h fed to step t (post eoc reset)
  float *cprev[SK_MAX_LAYERS] = {nullptr};  // [T,B,H]: lstm c fed to step t; gru: r*h of step t
  float *out[SK_MAX_LAYERS] = {nullptr};    // [T,B,H]: layer output (pre reset; top layer: after dropout)
  float *dout = nullptr;                    // [T,B,H]: gradient wrt a layer's output (reused per layer)
  float *logits = nullptr;                  // [T*B,NOUT]: logits -> d(logits)
  float *wpack = nullptr;                   // per-CTA packed recurrent weight slices (fwd or bwd)
  // bf16 copies feeding the tensor-core GEMMs (precision == SKRNN_PREC_BF16 only)
  __nv_bfloat16 *x8_bf = nullptr;                    // [T,B,8]: x padded to 8 columns
  __nv_bfloat16 *out_bf[SK_MAX_LAYERS] = {nullptr};  // [T,B,H]
  __nv_bfloat16 *hprev_bf[SK_MAX_LAYERS] = {nullptr};
  __nv_bfloat16 *aux_bf[SK_MAX_LAYERS] = {nullptr};  // gru r*h
  __nv_bfloat16 *xw_il[SK_MAX_LAYERS] = {nullptr};   // hoisted x*Wx+b in bf16, the input of the tcgen05 recurrent kernels.  lstm:
                                                     // [T,B,H,4], the four gates of a unit adjacent; gru / rnn: [T,B,G] in TF's order
  __nv_bfloat16 *acts_bf[SK_MAX_LAYERS] = {nullptr}; // [T,B,H,4]: lstm gate activations i,j,f,o saved by the tcgen05
                                                     // forward kernel for its BPTT counterpart
  // gru / rnn tensor-core kernels (rnn1_tc.cu): per-step arrays in the tile-transposed layout [T][B/16][2][cols][8];
  // acts_bf[l] of these cells has that layout too (gru: 4H columns, rnn: H)
  __nv_bfloat16 *xw_tt = nullptr;                    // hoisted projection of the layer being run
  float *dout_tt = nullptr;                          // d(out) of the layer being differentiated
  int64_t tt_rows = 0;                               // T * ceil16(B) these were sized for
  __nv_bfloat16 *dpre_bf = nullptr;                  // [T,B,G] of the layer being differentiated (= one of dpre_bf_buf)
  __nv_bfloat16 *dpre_bf_buf[2] = {nullptr, nullptr}; // two copies: layer l's weight gradients overlap layer l-1's BPTT
  __nv_bfloat16 *dlogits_bf = nullptr;               // [T*B,NOUTP]
  int64_t wpack_floats = 0;
  float *bptt_carry = nullptr;              // [2][B][H]: d(h), d(c) between the time ranges of a tcgen05 BPTT pass
  __nv_bfloat16 *partial = nullptr;         // tcgen05 BPTT: per-rank partial d(h) exchange
  int64_t partial_floats = 0;
};

struct skrnn_handle {
  skrnn_config cfg;
  int H, L, M, NOUT, G, S, NGATE;
  int64_t P;
  std::vector<SkTensor> tensors;
  SkLayerParams lp[SK_MAX_LAYERS];
  int64_t off_out_w, off_out_b;
  float *params = nullptr, *grads = nullptr, *adam_m = nullptr, *adam_v = nullptr;
  // bf16 mirror of the weights with 16-byte-aligned tensors: per layer the TF matrices as they are
  // ([I+H, ncol]), and output_w padded to NOUTP columns
  __nv_bfloat16 *params_bf = nullptr;
  int64_t pbf_w[SK_MAX_LAYERS][2];
  int64_t pbf_out_w = 0, pbf_count = 0;
  int cast_nseg = 0;
  int64_t *cast_seg_dev = nullptr;  // [nseg][5] = src offset, dst offset, count, cols, dst ld
  int NOUTP = 0;
  double beta1_power = 0.9, beta2_power = 0.999;
  SkWorkspace ws;
  // small device scratch: [0..2] loss accumulators (double), [3] grad sumsq (double), [4] barrier ctr
  double *scal = nullptr;
  unsigned int *sync_ctr = nullptr;
  // staging buffers for the *_host entry points
  float *st_x = nullptr, *st_y = nullptr, *st_state = nullptr, *st_loss = nullptr;
  float *pin_x = nullptr, *pin_y = nullptr, *pin_state = nullptr, *pin_loss = nullptr;
  int64_t st_rows = 0, st_state_n = 0;
  float *samp_state = nullptr;  // sampler: per-sketch recurrent state [n,L,S]
  int64_t samp_state_n = 0;
  void *samp_tc = nullptr;      // throughput sampler buffers (sampler_tc.cu)
  size_t samp_tc_bytes = 0;
  cudaGraphExec_t samp_graph = nullptr;  // captured step loop of the throughput sampler for (samp_graph_n, samp_graph_steps)
  int samp_graph_n = 0, samp_graph_steps = 0;
  int num_sms = 0;
  int64_t launches = 0;
  int cluster_size = 1;
  bool tc_rnn_disabled = false;
  bool warned_tc_fallback = false;
  bool bias_grad_done = false;  // set by a backward recurrent kernel that accumulated d(bias) itself
  // side stream for work off the critical path of the backward pass (weight gradients), see skrnn_loss_grad
  cudaStream_t side = nullptr, side2 = nullptr;
  cudaEvent_t ev_chunk[2][8] = {{nullptr}};  // [0][c]: a GEMM range is done, [1][c]: a recurrent range is done
  cudaEvent_t ev_main[2] = {nullptr, nullptr}, ev_side[2] = {nullptr, nullptr};
  // recorded by skrnn_loss_grad when a gradient segment is final: [l] = layer l's tensors, [L] = the output layer's
  cudaEvent_t ev_grad[SK_MAX_LAYERS + 1] = {nullptr};
  // lstm, bf16: Wx of every layer with gate-interleaved columns (4*unit + gate), rows padded to 8 for layer 0, and the
  // bias in the same order: the GEMM that produces xw_il
  __nv_bfloat16 *wx_il[SK_MAX_LAYERS] = {nullptr};
  float *bias_il[SK_MAX_LAYERS] = {nullptr};
  int chunk_t0 = 0, chunk_t1 = 0;  // set around a recurrent-layer call that covers only steps [t0, t1) (0, 0: all of them)
  int gemm_perm_H = 0;  // set around a split-K sk_gemm_tc call whose B operand has gate-interleaved columns (d(pre) of
                        // the tcgen05 lstm kernels): the epilogue then scatters C back to TF's gate-major columns
  bool xw_il_ok[SK_MAX_LAYERS] = {false};  // this step's xw_il[l] is what the forward of layer l reads
  int fwd_ranges[SK_MAX_LAYERS] = {0};  // time ranges the forward pass of a layer ran in (its BPTT uses the same)
  int fwd_family[SK_MAX_LAYERS] = {0};  // which kernels ran a layer's forward (0 rnn_f32, 1 rnn_tc, 3 rnn1_tc): they save
                                        // activations in different places, so the backward must be the same family
  // optional per-category kernel timing (skrnn_profile_*): CUDA events around every launch
  bool prof_on = false;
  struct ProfRec { int cat; cudaEvent_t a, b; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> prof_pool;
};

enum { SK_CAT_GEMM = 0, SK_CAT_RNN_FWD = 1, SK_CAT_RNN_BWD = 2, SK_CAT_HEAD = 3, SK_CAT_OPTIM = 4,
       SK_CAT_MISC = 5, SK_CAT_SAMPLER = 6, SK_CAT_TC_GEMM = 7, SK_NUM_CAT = 8 };

// RAII scope of an entry point: makes the handle's device current and puts the caller's device back on exit.
struct SkDeviceGuard {
  int prev = -1;
  bool changed = false;
  explicit SkDeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) changed = cudaSetDevice(dev) == cudaSuccess;
  }
  ~SkDeviceGuard() { if (changed) cudaSetDevice(prev); }
};

// RAII scope: counts the launch and, when profiling is on, brackets it with events on `st`.
struct SkLaunch {
  skrnn_handle *h; cudaStream_t st; cudaEvent_t b = nullptr;
  SkLaunch(skrnn_handle *h_, cudaStream_t st_, int cat) : h(h_), st(st_) {
    h->launches++;
    if (!h->prof_on) return;
    cudaEvent_t a;
    auto get = [&]() { cudaEvent_t e; if (!h->prof_pool.empty()) { e = h->prof_pool.back(); h->prof_pool.pop_back(); } else cudaEventCreate(&e); return e; };
    a = get(); b = get();
    cudaEventRecord(a, st);
    h->prof.push_back({cat, a, b});
  }
  ~SkLaunch() { if (b) cudaEventRecord(b, st); }
};

// ---- kernels.cu launchers (all stream-ordered; return 0 / non-zero with sk_set_error) ----
// C[M,N] (+)= A[M,K] * B[K,N] (+ bias[N]).  a_kcontig: A stored [M][lda] else [K][lda];
// b_ncontig: B stored [K][ldb] else [N][ldb].  accumulate: atomically add into C (split-K ok).
int sk_gemm_f32(skrnn_handle *h, cudaStream_t st, int M, int N, int K, const float *A, int lda,
                bool a_kcontig, const float *B, int ldb, bool b_ncontig, float *C, int ldc,
                const float *bias, bool accumulate);
// bf16 tensor-core GEMM (gemm_tc.cu): operands bf16, K-major (K contiguous) or MN-major.
int sk_gemm_tc(skrnn_handle *h, cudaStream_t st, int M, int N, int K, const void *A, int lda, bool a_kmajor,
               const void *B, int ldb, bool b_kmajor, float *C, int ldc, const float *bias, bool accumulate);
int sk_gemm_tc_lstm_cell(skrnn_handle *h, cudaStream_t st, int M, int H, int K, const void *A, int lda, const void *W,
                         const float *bias_il, float *cell_c, __nv_bfloat16 *cell_self, int kself, int hoff,
                         __nv_bfloat16 *cell_out, int kout, const float *xin);
int sk_gemm_tc_bf16out(skrnn_handle *h, cudaStream_t st, int M, int N, int K, const void *A, int lda, const void *B, int ldb,
                       __nv_bfloat16 *C, int ldc, const float *bias);
extern uint32_t g_tc_override[6];
// Tiled tensor map (rank 2..5), 128B-swizzled unless swizzle128 = false.  elem_bytes: 2 (bf16) or 4 (fp32).
int sk_make_tmap(CUtensorMap *m, int elem_bytes, const void *ptr, int rank, const uint64_t *dims,
                 const uint64_t *strides_bytes, const uint32_t *box, bool swizzle128 = true);
// The fp32 activation set (xw, hprev, out per layer) of the CUDA-core recurrent kernels.  bf16 handles whose layers run
// the tcgen05 recurrent kernels never touch it, so it is allocated on first use (api.cu).
int sk_ensure_f32_acts(skrnn_handle *h);
bool sk_rnn_tc_eligible(const skrnn_handle *h);
int sk_rnn_tc_ranges(const skrnn_handle *h, int B, int T);
// tcgen05 recurrent kernels (rnn_tc.cu): return 0 ok, <0 "not eligible, use the fp32 kernels", >0 error
int sk_rnn_tc_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *x_bt5,
                    const float *state_in, float *state_out, const float *mask, uint64_t seed, int row0,
                    bool training);
// tcgen05 recurrent kernels for the GRU / RNN cells, one CTA per batch tile (rnn1_tc.cu): 0 ok, -1 not eligible, > 0 error
bool sk_rnn1_tc_eligible(const skrnn_handle *h);
int sk_rnn1_tc_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *state_in,
                     float *state_out, const float *mask, uint64_t seed, int row0, bool training);
int sk_colsum_bf16(skrnn_handle *h, cudaStream_t st, const __nv_bfloat16 *A, int64_t rows, int cols, int lda,
                   float *out);
int sk_colsum(skrnn_handle *h, cudaStream_t st, const float *A, int64_t rows, int cols, int lda,
              float *out);  // out[c] += sum_r A[r][c]

int sk_rnn_forward_layer(skrnn_handle *h, cudaStream_t st, int l, int B, int T, const float *x_bt5,
                         const float *state_in, float *state_out, const float *mask, uint64_t seed,
                         int row0, bool training);
int sk_rnn_backward_layer(skrnn_handle *h, cudaStream_t st, int l, int B, int T, const float *x_bt5,
                          const float *mask, uint64_t seed, int row0, bool training);

int sk_head(skrnn_handle *h, cudaStream_t st, int B, int T, const float *y_bt5, float *mdn_out,
            bool want_grad, int64_t loss_rows);
// output layer + MDN head + d(logits) in one tcgen05 kernel (head_tc.cu, bf16 training path)
bool sk_head_tc_eligible(const skrnn_handle *h);
int sk_head_tc(skrnn_handle *h, cudaStream_t st, int B, int T, const float *y_bt5, int64_t loss_rows);
int sk_finalize_losses(skrnn_handle *h, cudaStream_t st, float *losses_dev);
int sk_clip_adam(skrnn_handle *h, cudaStream_t st, float lr, float *gnorm_out);
int sk_cast_params(skrnn_handle *h, cudaStream_t st);
int sk_transpose_x(skrnn_handle *h, cudaStream_t st, int B, int T, const float *x_bt5);
int sk_sample_f32(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen,
                  int stop_if_eoc, const float *uniforms, uint64_t seed, int sketch0, float *strokes,
                  int *lengths, float *mixp);
// bf16 tensor-core throughput sampler (sampler_tc.cu): 0 ok, -1 not applicable, >0 error
int sk_sample_cluster(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen, int stop_if_eoc,
                      const float *uniforms, uint64_t seed, int sketch0, float *strokes, int *lengths, float *mixp);
size_t sk_sample_tp_scratch_bytes();
int sk_sample_tp(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen, int stop_if_eoc,
                 const float *uniforms, uint64_t seed, int sketch0, float *strokes, int *lengths, float *mixp,
                 const __nv_bfloat16 *const *wsm, const float *const *bias_il, float *const *cst, float *tables);
int sk_sample_tc(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen,
                 int stop_if_eoc, const float *uniforms, uint64_t seed, int sketch0, float *strokes,
                 int *lengths, float *mixp);
