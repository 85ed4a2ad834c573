/*
 * skrnn.h -- C ABI of libskrnn.so, the B200 (sm_100a) implementation of sketch-rnn's
 * training step and sampling loop.
 *
 * The reference (ByzanTine/sketch-rnn) has no FFI: its hot path is whatever a
 * tf.Session.run() evaluates.  Each entry point below names the reference call site it
 * replaces (paths relative to the reference repo root).  The Python host code in
 * sketch_rnn_b200/ binds these with ctypes and presents the reference's own
 * Model(args, infer) / train.py / sample.py surface on top.
 *
 * Conventions
 *   - plain C types only; every function returns 0 on success, non-zero on failure, and
 *     never throws.  skrnn_last_error() returns a thread-local message for the last failure.
 *   - all arrays are float32 unless stated.  "dev" pointers are CUDA device pointers on the
 *     handle's device, "host" pointers are ordinary host memory.
 *   - strokes:  [B, T, 5] = (dx, dy, eos, eoc, cont), C-contiguous       (model.py:34-35)
 *   - state:    [L, B, S] = the tuple of L per-layer states; S = 2H for lstm ([c | h]),
 *               H for gru / rnn                                          (model.py:36)
 *   - mdn rows: index b*T + t                                            (model.py:98)
 *   - a handle is bound to one device and is not thread-safe; work is stream-ordered on
 *     the cudaStream_t passed as `void *stream` (NULL = the legacy default stream).
 *   - one process (or thread) per GPU for data-parallel runs; ranks exchange gradients
 *     between skrnn_loss_grad and skrnn_apply_clip_adam.
 */
#ifndef SKRNN_H_
#define SKRNN_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define SKRNN_API __attribute__((visibility("default")))
#else
#define SKRNN_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

enum { SKRNN_CELL_LSTM = 0, SKRNN_CELL_GRU = 1, SKRNN_CELL_RNN = 2 };
/* SKRNN_PREC_F32: fp32 storage and fp32 CUDA-core accumulation everywhere (1e-4 parity path).
 * SKRNN_PREC_BF16: bf16 operands on the tcgen05 tensor cores with fp32 accumulation and fp32
 * state / master weights / optimiser (2e-2 parity path). */
enum { SKRNN_PREC_F32 = 0, SKRNN_PREC_BF16 = 1 };

/* Largest args.rnn_size a handle can be created with (a multiple of 8; the reference takes any size).  Every
 * kernel family keeps a per-CTA slice of the recurrent weights resident on chip and the widest plan ends here;
 * skrnn_create rejects larger values instead of failing at the first step. */
#define SKRNN_MAX_RNN_SIZE 512

/* Hyper-parameters Model.__init__ reads from `args` (model.py:8-47,134,182; train.py:13-44). */
typedef struct skrnn_config {
  int32_t cell;          /* args.model: SKRNN_CELL_*                       (model.py:14-21) */
  int32_t rnn_size;      /* args.rnn_size  H, a multiple of 8, <= SKRNN_MAX_RNN_SIZE       */
  int32_t num_layers;    /* args.num_layers L (reference runs L == 2 only, model.py:82)    */
  int32_t num_mixture;   /* args.num_mixture M; NOUT = 3 + 6M              (model.py:43)   */
  int32_t batch_size;    /* sizing hint: largest B the handle will see                      */
  int32_t seq_length;    /* sizing hint: largest T the handle will see                      */
  float keep_prob;       /* args.keep_prob, output dropout when training   (model.py:29-30) */
  float grad_clip;       /* args.grad_clip                                 (model.py:182)   */
  float stroke_importance_factor; /*                                       (model.py:134)   */
  float data_scale;      /* args.data_scale                                (model.py:263)   */
  int32_t precision;     /* SKRNN_PREC_*                                                    */
  int32_t device;        /* CUDA device ordinal                                             */
} skrnn_config;

typedef struct skrnn_handle skrnn_handle;

SKRNN_API const char *skrnn_last_error(void);
/* Library build tag, e.g. "skrnn 0.1 sm_100a". */
SKRNN_API const char *skrnn_version(void);

/* Replaces graph construction in Model.__init__ (model.py:8-184): allocates weights, Adam
 * slots and activation workspaces on cfg->device.  Weights start at zero; load them with
 * skrnn_set_params.  Fails loudly if no sm_100 device is present. */
SKRNN_API int skrnn_create(const skrnn_config *cfg, skrnn_handle **out);
SKRNN_API void skrnn_destroy(skrnn_handle *h);

/* Flat fp32 parameter vector = tf.trainable_variables() in creation order, each tensor
 * row-major in its TF shape (output_w [H,NOUT], output_b, then per layer Linear/Matrix
 * [I+H, G] with rows [x ; h] and Linear/Bias; model.py:45-47, TF rnn_cell _linear). */
SKRNN_API int64_t skrnn_param_count(const skrnn_handle *h);
SKRNN_API int32_t skrnn_num_tensors(const skrnn_handle *h);
/* name_buf receives the TF variable name (NUL-terminated, truncated to name_cap). */
SKRNN_API int skrnn_tensor_info(const skrnn_handle *h, int32_t index, char *name_buf, int32_t name_cap,
                      int64_t *offset, int32_t *rows, int32_t *cols);

/* Replace tf.train.Saver.restore / save of the trainables (sample.py:59-64, train.py:72-75). */
SKRNN_API int skrnn_set_params(skrnn_handle *h, const float *host_params, int64_t count);
SKRNN_API int skrnn_get_params(skrnn_handle *h, float *host_params, int64_t count);
/* Adam slots `<var>/Adam`, `<var>/Adam_1`, beta1_power, beta2_power (model.py:183). */
SKRNN_API int skrnn_set_adam_state(skrnn_handle *h, const float *host_m, const float *host_v, int64_t count,
                         double beta1_power, double beta2_power);
SKRNN_API int skrnn_get_adam_state(skrnn_handle *h, float *host_m, float *host_v, int64_t count,
                         double *beta1_power, double *beta2_power);
/* Device addresses of the handle-owned flat buffers (for an external all-reduce of the
 * gradient between skrnn_loss_grad and skrnn_apply_clip_adam). */
SKRNN_API float *skrnn_params_dev(skrnn_handle *h);
SKRNN_API float *skrnn_grads_dev(skrnn_handle *h);

/* Replaces sess.run([pi, mu1, mu2, sigma1, sigma2, corr, pen, final_state], feed)
 * (model.py:215-217) and the forward half of train.py:86.  No dropout.
 *   x_dev        [B,T,5]; state_in_dev [L,B,S] (the fed initial_state, also the eoc-reset
 *                source, model.py:82-92)
 *   mdn_out_dev  [B*T, NOUT] or NULL: per row (pen logits[3], pi[M], mu1[M], mu2[M],
 *                sigma1[M], sigma2[M], corr[M]) after get_mixture_coef (model.py:142-163)
 *   state_out_dev [L,B,S] = final_state (model.py:100); may alias state_in_dev. */
SKRNN_API int skrnn_forward(skrnn_handle *h, int32_t B, int32_t T, const float *x_dev,
                  const float *state_in_dev, float *mdn_out_dev, float *state_out_dev,
                  void *stream);

/* Replaces the cost / cost_shape / cost_pen fetches and tf.gradients (model.py:176-182):
 * forward with output dropout, MDN loss, BPTT.  The UNCLIPPED gradient lands in
 * skrnn_grads_dev(h).
 *   dropout_mask_dev  [B,T,H] of 0/1 or NULL.  NULL: draw the mask in-kernel from a counter
 *                     hash keyed by (seed, row0 + b, t, unit) when keep_prob < 1.
 *   row0, loss_rows   data-parallel sharding: this call holds rows [row0, row0+B) of a global
 *                     batch whose losses/gradients are means over loss_rows = B_global*T rows.
 *                     Single GPU: row0 = 0, loss_rows = B*T.
 *   losses_dev        [3] = cost, cost_shape, cost_pen (this shard's share of the mean).
 *   state_out_dev     [L,B,S] final_state; may alias state_in_dev. */
SKRNN_API int skrnn_loss_grad(skrnn_handle *h, int32_t B, int32_t T, const float *x_dev, const float *y_dev,
                    const float *state_in_dev, const float *dropout_mask_dev, uint64_t seed,
                    int32_t row0, int64_t loss_rows, float *losses_dev, float *state_out_dev,
                    void *stream);

/* Data-parallel overlap (no reference counterpart: the reference is single-process; the gradient it would exchange
 * is the output of tf.gradients, model.py:182).  skrnn_loss_grad finishes the flat gradient back to front: the output
 * layer's tensors first, then the cell layers from the top one down.  Segment `index` is L for (output_w, output_b)
 * and l for layer l's tensors; each is one contiguous range [offset, offset + count) of skrnn_grads_dev(h).
 * skrnn_stream_wait_grad makes `stream` wait (cudaStreamWaitEvent, no host blocking) until the most recent
 * skrnn_loss_grad has finished writing that segment, so a caller can start reducing it over the ranks on its own
 * communication stream while the layers below are still being differentiated.  Segment L also covers losses_dev. */
SKRNN_API int skrnn_grad_segment(const skrnn_handle *h, int32_t index, int64_t *offset, int64_t *count);
SKRNN_API int skrnn_stream_wait_grad(skrnn_handle *h, int32_t index, void *stream);

/* Replaces tf.clip_by_global_norm + AdamOptimizer(lr, epsilon=1e-3).apply_gradients
 * (model.py:182-184) on the gradient currently in skrnn_grads_dev(h).
 * gnorm_out_dev: [1] global norm before clipping, or NULL. */
SKRNN_API int skrnn_apply_clip_adam(skrnn_handle *h, float lr, float *gnorm_out_dev, void *stream);

/* One sess.run([cost, cost_shape, cost_pen, final_state, train_op], feed) (train.py:85-86)
 * with HOST buffers: copies x/y/state in, runs loss_grad + clip + Adam, copies the three
 * losses and final_state back, and synchronises.  state_host [L,B,S] is updated in place
 * (the state carry of train.py:85-86).  Page-locked (cudaHostAlloc / cudaHostRegister) buffers are
 * copied from and to directly; pageable ones are staged through pinned memory owned by the handle. */
SKRNN_API int skrnn_train_step_host(skrnn_handle *h, int32_t B, int32_t T, const float *x_host,
                          const float *y_host, float *state_host, float lr, uint64_t seed,
                          float *losses_host);

/* Replaces Model.sample (model.py:187-264) for n_sketches independent sketches at once.
 *   uniforms_dev  [n, num_steps, 4] = (u_pi, u_pen, u_gauss1, u_gauss2) in [0,1) standing in
 *                 for random.random() x2 and np.random.multivariate_normal, or NULL to draw
 *                 them in-kernel from Philox(seed, sketch0 + sketch, step).
 *   strokes_dev   [n, num_steps, 5]; dx,dy already multiplied by data_scale (model.py:263);
 *                 rows past a sketch's length are zero.
 *   lengths_dev   int32 [n]: points produced (num_steps unless stop_if_eoc fired).
 *   mix_params_dev [n, num_steps, NOUT] or NULL: per step (pen_pdf[3], pi_pdf[M] as returned
 *                 in mixture_params incl. the stray /temp_pen, mu1, mu2, sigma1, sigma2, corr). */
SKRNN_API int skrnn_sample(skrnn_handle *h, int32_t n_sketches, int32_t num_steps, float temp_mixture,
                 float temp_pen, int32_t stop_if_eoc, const float *uniforms_dev, uint64_t seed,
                 int32_t sketch0, float *strokes_dev, int32_t *lengths_dev, float *mix_params_dev,
                 void *stream);

/* Number of kernels of this library launched on the handle so far (bench.py's gpu_launches). */
SKRNN_API int64_t skrnn_launch_count(const skrnn_handle *h);

/* Measurement aid (no reference counterpart; train.py:83-92 only prints wall time per batch):
 * when enabled, every kernel launch is bracketed by CUDA events on its stream.
 * skrnn_profile_read synchronises the device, writes the summed milliseconds and launch counts
 * per category into ms_out[8] / count_out[8] and resets.  Categories: 0 fp32 GEMM, 1 recurrent
 * forward, 2 recurrent backward, 3 MDN head, 4 clip+Adam, 5 misc (pack/transpose/colsum),
 * 6 sampler, 7 tensor-core GEMM. */
SKRNN_API int skrnn_profile_enable(skrnn_handle *h, int32_t on);
SKRNN_API int skrnn_profile_read(skrnn_handle *h, double *ms_out, int64_t *count_out);

/* No counterpart in the reference.  The clip + Adam update of skrnn_apply_clip_adam for callers
 * that hold the buffers themselves and cannot let this library allocate device memory: a trainer
 * whose parameters, gradients and Adam slots sit in a framework-managed pool (the pool usually
 * owns the whole GPU already, so a cudaMalloc issued from here would fail or fragment it), or any
 * context in which allocating is not permitted.  Creating a handle is not an option for them --
 * skrnn_create allocates weights, slots and workspaces.
 *   n                 number of float32 parameters, 1 .. SKRNN_NOALLOC_MAX_PARAMS; more is an error
 *   grads_dev         [n] in;  params_dev, m_dev, v_dev  [n] in/out (updated in place)
 *   lr, grad_clip     learning rate and clip norm (model.py:182-183)
 *   beta1_power, beta2_power   0.9^t and 0.999^t for this step t >= 1 (the caller advances them)
 *   gnorm_out_dev     [1] global norm before clipping, or NULL
 * Its only scratch memory is a fixed array of SKRNN_NOALLOC_SCRATCH_BYTES inside the library image
 * (one copy per GPU that loads the library, held until exit).  The call performs no cudaMalloc,
 * cudaFree or allocating library call.  Every call on a device shares that array and it keeps the
 * previous call's contents; what must start at zero is cleared on `stream` at the start of each
 * call.  Calls that overlap on the GPU race on it: keep them on one stream or otherwise serialise
 * them -- host threads included, and the first call included: the first call on a device resolves
 * the array's device address and queries the device (unsynchronised bookkeeping), so issue it from
 * one thread and before beginning any stream capture.  The call runs on the GPU that owns
 * params_dev (the caller's current device is put back before returning); `stream` must belong to
 * that GPU. */
#define SKRNN_NOALLOC_MAX_PARAMS (8192LL * 4096LL)
#define SKRNN_NOALLOC_SCRATCH_BYTES (256 + 8192 * 8)
SKRNN_API int skrnn_clip_adam_noalloc(int64_t n, const float *grads_dev, float *params_dev, float *m_dev,
                            float *v_dev, float lr, float grad_clip, double beta1_power,
                            double beta2_power, float *gnorm_out_dev, void *stream);

/* Host-only helper (no GPU work): CRC-32C (Castagnoli) of `n` bytes continued from `crc` (0 to start), the
 * block checksum of TensorFlow's V1 checkpoint tables that tf.train.Saver reads and writes (train.py:64,74;
 * sample.py:59-64).  Used by sketch_rnn_b200/checkpoint.py. */
SKRNN_API uint32_t skrnn_crc32c(uint32_t crc, const void *data, size_t n);

/* Diagnostic (no reference counterpart): runs the bf16 tcgen05/TMA GEMM building block on caller
 * buffers so tests can check it against a library matmul.  C[M,N] (+)= A*B (+bias), A/B bf16 device
 * matrices, *_kmajor != 0: K contiguous ([M][lda] resp. [N][ldb]); == 0: M resp. N contiguous
 * ([K][lda] resp. [K][ldb]).  overrides: NULL, or six descriptor strides for bring-up sweeps. */
SKRNN_API int skrnn_debug_gemm_bf16(skrnn_handle *h, int32_t M, int32_t N, int32_t K, const void *a_dev,
                          int32_t lda, int32_t a_kmajor, const void *b_dev, int32_t ldb, int32_t b_kmajor,
                          float *c_dev, int32_t ldc, const float *bias_dev, int32_t accumulate,
                          const uint32_t *overrides, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SKRNN_H_ */
