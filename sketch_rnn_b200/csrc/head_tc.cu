// head_tc.cu -- the output linear layer fused with the MDN head, forward and backward, for the bf16 training path:
// reference model.py:98-99 (xw_plus_b on the top layer's outputs), get_mixture_coef (model.py:142-163), tf_2d_normal
// (model.py:112-122), get_lossfunc (model.py:124-139) and the analytic d(cost)/d(logits).
//
// One persistent kernel.  A CTA walks 128-row tiles of the [T*B, H] top-layer output: TMA brings the operand
// k-blocks into a shared-memory ring, tcgen05.mma accumulates the tile's [128, 160] logits in tensor memory (two
// accumulators, so the product of tile i+1 overlaps the head math of tile i), and the epilogue threads -- one per
// row -- read their row out of tensor memory, add the bias, and run the whole head on it: softmax over the mixture
// weights, exp / tanh of the scale and correlation logits, the bivariate density, the clamped mixture NLL, the pen
// cross-entropy with its stroke-importance weight, and the gradient of all of it.  The logits never reach global
// memory: what leaves is d(logits) in bf16 (one TMA store per 64 columns, exactly the operand the two output-layer
// gradient GEMMs read) and the two loss sums.
//
// A row's 147 values do not fit in registers next to the math, so each thread parks its row in shared memory
// (row pitch 161 words: the 32 lanes of a warp hit 32 different banks) and works in place: logits -> (pi, a, b, s1,
// s2, rho) -> gradients -> bf16.
#include "common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int HT_BM = 128, HT_BK = 64, HT_N = 160;   // tile rows, k-block, accumulator columns (NOUT <= 160)
constexpr int HT_STAGES = 2;
constexpr int HT_A_BYTES = HT_BM * HT_BK * 2;          // 16 KB: [128 rows][64 k] K-major
constexpr int HT_B_BYTES = 3 * HT_BK * 64 * 2;         // 24 KB: three boxes of [64 k][64 n] (N-major; the last one half used)
constexpr int HT_STAGE = HT_A_BYTES + HT_B_BYTES;
constexpr int HT_PITCH = 161;                          // words per parked row
constexpr int HT_ROWBUF = HT_BM * HT_PITCH * 4;
constexpr int HT_OUT = 3 * HT_BM * 128;                // three [128 rows][64 bf16] boxes, 128B-swizzled
constexpr int HT_THREADS = 192;                        // warp 0: TMA, warp 1: MMA, warps 2-5: one thread per tile row
constexpr int HT_SMEM = HT_STAGES * HT_STAGE + HT_OUT + HT_ROWBUF + 1024 /*bias*/ + 1024 /*align*/ + 256;

struct HeadTcArgs {
  int R, B, T, H, M, NOUT, NOUTP;
  const float *bias;   // output_b [NOUT]
  const float *y;      // targets [B,T,5] (caller layout)
  float inv_rows, sif;
  double *scal;        // [1] += cost_shape, [2] += cost_pen (already divided by the row count)
};

// The head on one parked row z[0..NOUT): on return z holds d(cost)/d(logits) (zero past NOUT), and the row's two loss
// terms are returned.  Same formulas as mdn_head_kernel (head.cu); the mixture loop runs serially in one thread.
__device__ __forceinline__ float ht_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ht_rsqrt(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// One thread runs 24 mixtures serially with one warp per scheduler, so instruction count is what this kernel costs:
// the transcendental functions are the hardware approximations (ex2 / lg2 / tanh / rcp / rsqrt, relative error
// ~1e-6 -- three orders below the bf16 rounding of the logits this path starts from).
__device__ __forceinline__ void ht_row(float *z, int M, int NOUT, const float *tg, float inv_rows, float sif, float sqrt_sif,
                                       float &row_shape, float &row_pen) {
  const float x1 = tg[0], x2 = tg[1], p_eos = tg[2], p_eoc = tg[3], p_cont = tg[4];
  float *zpi = z + 3, *zm1 = zpi + M, *zm2 = zm1 + M, *zs1 = zm2 + M, *zs2 = zs1 + M, *zr = zs2 + M;
  const float inv_2pi = 0.15915494309189535f;
  float mx = -INFINITY;
  for (int k = 0; k < M; ++k) mx = fmaxf(mx, zpi[k]);
  float se = 0.f;
  for (int k = 0; k < M; ++k) {
    const float e = __expf(zpi[k] - mx);
    zpi[k] = e;
    se += e;
  }
  const float inv_se = ht_rcp(se);
  float S = 0.f;
  for (int k = 0; k < M; ++k) {
    const float pi = zpi[k] * inv_se;
    const float i1 = __expf(-zs1[k]), i2 = __expf(-zs2[k]), rho = tanh_approx(zr[k]);  // 1/sigma1, 1/sigma2
    const float a = (x1 - zm1[k]) * i1, b = (x2 - zm2[k]) * i2;
    const float q = 1.f - rho * rho;
    const float iq = ht_rcp(q);
    const float Z = a * a + b * b - 2.f * rho * a * b;
    const float dens = __expf(-0.5f * Z * iq) * (inv_2pi * i1 * i2 * ht_rsqrt(q));
    S += dens * pi;
    zpi[k] = pi; zm1[k] = a; zm2[k] = b; zs1[k] = i1; zs2[k] = i2; zr[k] = rho;
  }
  row_shape = -__logf(fmaxf(S, 1e-20f));  // model.py:130
  const bool live = S >= 1e-20f;          // the clamp passes no gradient to the mixture
  const float invS = live ? ht_rcp(S) : 0.f;
  for (int k = 0; k < M; ++k) {
    const float pi = zpi[k], a = zm1[k], b = zm2[k], i1 = zs1[k], i2 = zs2[k], rho = zr[k];
    const float q = 1.f - rho * rho;
    const float iq = ht_rcp(q);
    const float Z = a * a + b * b - 2.f * rho * a * b;
    const float dens = __expf(-0.5f * Z * iq) * (inv_2pi * i1 * i2 * ht_rsqrt(q));
    const float gam = pi * dens * invS;
    const float ga = gam * inv_rows;
    zpi[k] = ((live ? pi : 0.f) - gam) * inv_rows;
    zm1[k] = -ga * (a - rho * b) * i1 * iq;
    zm2[k] = -ga * (b - rho * a) * i2 * iq;
    zs1[k] = -ga * (a * (a - rho * b) * iq - 1.f);
    zs2[k] = -ga * (b * (b - rho * a) * iq - 1.f);
    zr[k] = -ga * (a * b + rho * (1.f - Z * iq));
  }
  // pen softmax cross-entropy weighted by the stroke importance (model.py:132-136)
  const float z0 = z[0], z1 = z[1], z2 = z[2];
  const float zm = fmaxf(z0, fmaxf(z1, z2));
  const float e0 = __expf(z0 - zm), e1 = __expf(z1 - zm), e2 = __expf(z2 - zm);
  const float lse = zm + __logf(e0 + e1 + e2);
  const float xent = -(p_eos * (z0 - lse) + p_eoc * (z1 - lse) + p_cont * (z2 - lse));
  const float w = p_cont + sqrt_sif * p_eos + sif * p_eoc;
  row_pen = xent * w;
  const float psum = p_eos + p_eoc + p_cont;
  const float inv_e = ht_rcp(e0 + e1 + e2);
  const float ws = w * inv_rows;
  z[0] = ws * (e0 * inv_e * psum - p_eos);
  z[1] = ws * (e1 * inv_e * psum - p_eoc);
  z[2] = ws * (e2 * inv_e * psum - p_cont);
  for (int c = NOUT; c < HT_N; ++c) z[c] = 0.f;
}

__global__ void __launch_bounds__(HT_THREADS, 1)
head_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmC, HeadTcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t s_out = base + HT_STAGES * HT_STAGE;
  const uint32_t s_row = s_out + HT_OUT;
  const uint32_t s_bias = s_row + HT_ROWBUF;
  const uint32_t bars = s_bias + 1024;
  const uint32_t bar_full = bars, bar_empty = bars + 8 * HT_STAGES;
  const uint32_t bar_tfull = bars + 16 * HT_STAGES, bar_tempty = bar_tfull + 16, tmem_holder = bar_tempty + 16;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gen + (tmem_holder - base));
  float *bias_s = reinterpret_cast<float *>(gen + (s_bias - base));
  float *rowbuf = reinterpret_cast<float *>(gen + (s_row - base));
  uint8_t *out_gen = gen + (s_out - base);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = (a.R + HT_BM - 1) / HT_BM;
  const int nkb = (a.H + HT_BK - 1) / HT_BK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < HT_STAGES; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar_tfull + 8 * b, 1); mbar_init(bar_tempty + 8 * b, 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int j = threadIdx.x; j < HT_N; j += HT_THREADS) bias_s[j] = j < a.NOUT ? __ldg(a.bias + j) : 0.f;
  if (warp == 1) tc_alloc(tmem_holder, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;

  if (warp == 0) {
    // ---------------- TMA producer ----------------
    if (elect_one()) {
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int m0 = tile * HT_BM;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % HT_STAGES, ph = (it / HT_STAGES) & 1;
          mbar_wait(bar_empty + 8 * s, ph ^ 1);
          const uint32_t sa = base + s * HT_STAGE, sb = sa + HT_A_BYTES;
          mbar_expect_tx(bar_full + 8 * s, HT_STAGE);
          tma_load_2d(sa, &tmA, bar_full + 8 * s, kb * HT_BK, m0);
#pragma unroll
          for (int c = 0; c < 3; ++c) tma_load_2d(sb + c * 8192, &tmB, bar_full + 8 * s, 64 * c, kb * HT_BK);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: A K-major, B N-major (64-column chunks 8192 B apart), N = 160 ----------------
    if (elect_one()) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(HT_N >> 3) << 17) |
                             ((uint32_t)(HT_BM >> 4) << 24);
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(bar_tempty + 8 * buf, ((lt >> 1) & 1) ^ 1);  // the epilogue has copied this accumulator out
        tc_fence_after();
        const uint32_t acc = tmem_acc + (uint32_t)(buf * HT_N);
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % HT_STAGES, ph = (it / HT_STAGES) & 1;
          mbar_wait(bar_full + 8 * s, ph);
          tc_fence_after();
          const uint32_t sa = base + s * HT_STAGE, sb = sa + HT_A_BYTES;
          const uint64_t da = make_desc(sa, 0, 1024), db = make_desc(sb, 8192, 1024);
#pragma unroll
          for (int k = 0; k < HT_BK / 16; ++k)
            tc_mma_bf16(acc, da + (uint64_t)(2 * k), db + (uint64_t)(128 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
          tc_commit(bar_empty + 8 * s);
        }
        tc_commit(bar_tfull + 8 * buf);
      }
    }
  } else {
    // ---------------- epilogue: one thread per tile row ----------------
    const int q = warp & 3;
    const int r = 32 * q + lane;
    float *z = rowbuf + r * HT_PITCH;
    const bool issuer = warp == 2 && lane == 0;
    const float sqrt_sif = sqrtf(a.sif);
    double my_shape = 0.0, my_pen = 0.0;
    int lt = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
      const int buf = lt & 1;
      const int row_g = tile * HT_BM + r;       // time-major row t*B + b
      const bool row_ok = row_g < a.R;
      float tg[5] = {0.f, 0.f, 0.f, 0.f, 1.f};
      if (row_ok) {
        const int t = row_g / a.B, b = row_g - t * a.B;
        const float *yp = a.y + ((size_t)b * a.T + t) * 5;
#pragma unroll
        for (int i = 0; i < 5; ++i) tg[i] = __ldg(yp + i);
      }
      mbar_wait(bar_tfull + 8 * buf, (lt >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < HT_N / 32; ++c) {
        uint32_t v[32];
        tc_ld32(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * HT_N + c * 32), v);
#pragma unroll
        for (int i = 0; i < 32; ++i) z[c * 32 + i] = __uint_as_float(v[i]) + bias_s[c * 32 + i];
      }
      tc_fence_before();
      mbar_arrive(bar_tempty + 8 * buf);  // the next product may overwrite the accumulator while the head runs
      if (row_ok) {
        float rs, rp;
        ht_row(z, a.M, a.NOUT, tg, a.inv_rows, a.sif, sqrt_sif, rs, rp);
        my_shape += (double)rs;
        my_pen += (double)rp;
      }
      // the previous tile's store has finished reading the staging boxes
      if (issuer) tma_store_wait_read();
      named_bar_sync(1, 128);
#pragma unroll 1
      for (int c = 0; c < HT_N / 8; ++c) {
        uint32_t p[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const __nv_bfloat162 b2 = __floats2bfloat162_rn(row_ok ? z[c * 8 + 2 * i] : 0.f, row_ok ? z[c * 8 + 2 * i + 1] : 0.f);
          p[i] = *reinterpret_cast<const uint32_t *>(&b2);
        }
        *reinterpret_cast<uint4 *>(out_gen + (c >> 3) * 16384 + r * 128 + (((c & 7) ^ (r & 7)) << 4)) = make_uint4(p[0], p[1], p[2], p[3]);
      }
      fence_proxy_async_smem();
      named_bar_sync(1, 128);
      if (issuer) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          if (64 * c < a.NOUTP) tma_store_2d_nocommit(&tmC, s_out + c * 16384, 64 * c, tile * HT_BM);
        tma_store_commit();
      }
    }
    if (issuer) tma_store_wait_all();
    // loss sums: warp shuffle, then one atomic per warp
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      my_shape += __shfl_xor_sync(0xffffffffu, my_shape, o);
      my_pen += __shfl_xor_sync(0xffffffffu, my_pen, o);
    }
    if (lane == 0) {
      atomicAdd(&a.scal[1], my_shape * (double)a.inv_rows);
      atomicAdd(&a.scal[2], my_pen * (double)a.inv_rows);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tc_dealloc(tmem_acc, 512);
}

}  // namespace

bool sk_head_tc_eligible(const skrnn_handle *h) {
  if (h->cfg.precision != SKRNN_PREC_BF16 || h->NOUT > HT_N || h->H % 8 != 0) return false;
  static int env_off = -1;
  if (env_off < 0) { const char *e = getenv("SKRNN_FUSED_HEAD"); env_off = (e && e[0] == '0') ? 1 : 0; }
  return !env_off;
}

// logits = out_top * output_w + output_b, head, d(logits) -> ws.dlogits_bf; loss sums into h->scal[1..2]
int sk_head_tc(skrnn_handle *h, cudaStream_t st, int B, int T, const float *y_bt5, int64_t loss_rows) {
  const int64_t R = (int64_t)B * T;
  const int H = h->H;
  CUtensorMap tmA, tmB, tmC;
  {
    uint64_t dims[2] = {(uint64_t)H, (uint64_t)R};
    uint64_t str[1] = {(uint64_t)H * 2};
    uint32_t box[2] = {HT_BK, HT_BM};
    SK_TRY(sk_make_tmap(&tmA, 2, h->ws.out_bf[h->L - 1], 2, dims, str, box));
  }
  {
    uint64_t dims[2] = {(uint64_t)h->NOUTP, (uint64_t)H};
    uint64_t str[1] = {(uint64_t)h->NOUTP * 2};
    uint32_t box[2] = {64, HT_BK};
    SK_TRY(sk_make_tmap(&tmB, 2, h->params_bf + h->pbf_out_w, 2, dims, str, box));
  }
  {
    uint64_t dims[2] = {(uint64_t)h->NOUTP, (uint64_t)R};
    uint64_t str[1] = {(uint64_t)h->NOUTP * 2};
    uint32_t box[2] = {64, HT_BM};
    SK_TRY(sk_make_tmap(&tmC, 2, h->ws.dlogits_bf, 2, dims, str, box));
  }
  HeadTcArgs a = {};
  a.R = (int)R; a.B = B; a.T = T; a.H = H; a.M = h->M; a.NOUT = h->NOUT; a.NOUTP = h->NOUTP;
  a.bias = h->params + h->off_out_b;
  a.y = y_bt5;
  a.inv_rows = 1.0f / (float)loss_rows;
  a.sif = h->cfg.stroke_importance_factor;
  a.scal = h->scal;
  const int ntiles = (int)((R + HT_BM - 1) / HT_BM);
  SkLaunch scope(h, st, SK_CAT_HEAD);
  SK_CUDA(cudaFuncSetAttribute(head_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HT_SMEM));
  head_tc_kernel<<<std::min(ntiles, h->num_sms), HT_THREADS, HT_SMEM, st>>>(tmA, tmB, tmC, a);
  SK_CUDA(cudaGetLastError());
  return 0;
}
