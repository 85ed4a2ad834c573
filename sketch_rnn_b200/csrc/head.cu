// head.cu -- MDN head: get_mixture_coef (reference model.py:142-163), tf_2d_normal
// (model.py:112-122), get_lossfunc (model.py:124-139) and the analytic backward of all three
//  in one pass over the logits: one warp per (t,b) row, lanes across the
// mixture components, logits overwritten in place by d(cost)/d(logits).
// Also the fused global-norm clip + Adam update (model.py:182-184).
#include <cooperative_groups.h>

#include "common.cuh"

namespace {

constexpr int KPL = 4;  // mixture components per lane -> num_mixture <= 128

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// rows are time-major: r = t*B + b.  y is the caller's [B,T,5] target.
__global__ void __launch_bounds__(256)
mdn_head_kernel(float *__restrict__ logits, const float *__restrict__ y, float *__restrict__ mdn_out,
                int B, int T, int M, int want_grad, float inv_rows, float sif, double *__restrict__ scal,
                __nv_bfloat16 *__restrict__ dl_bf, int noutp) {
  const int NOUT = 3 + 6 * M;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t R = (int64_t)B * T;
  __shared__ double sh_shape[8], sh_pen[8];
  double my_shape = 0.0, my_pen = 0.0;
  const float sqrt_sif = sqrtf(sif);

  for (int64_t r = (int64_t)blockIdx.x * 8 + wid; r < R; r += (int64_t)gridDim.x * 8) {
    const int b = (int)(r % B), t = (int)(r / B);
    float *z = logits + r * NOUT;
    float x1 = 0.f, x2 = 0.f, p_eos = 0.f, p_eoc = 0.f, p_cont = 1.f;
    if (y != nullptr) {
      const float *tg = y + ((int64_t)b * T + t) * 5;
      x1 = tg[0]; x2 = tg[1]; p_eos = tg[2]; p_eoc = tg[3]; p_cont = tg[4];
    }

    float zpi[KPL], mu1[KPL], mu2[KPL], s1[KPL], s2[KPL], rho[KPL];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < KPL; ++j) {
      int k = lane + 32 * j;
      if (k < M) {
        zpi[j] = z[3 + k];
        mu1[j] = z[3 + M + k];
        mu2[j] = z[3 + 2 * M + k];
        s1[j] = expf(z[3 + 3 * M + k]);
        s2[j] = expf(z[3 + 4 * M + k]);
        rho[j] = tanhf(z[3 + 5 * M + k]);
        mx = fmaxf(mx, zpi[j]);
      }
    }
    mx = warp_max(mx);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < KPL; ++j)
      if (lane + 32 * j < M) {
        zpi[j] = expf(zpi[j] - mx);
        se += zpi[j];
      }
    se = warp_sum(se);
    const float inv_se = 1.0f / se;
    float dens[KPL], pi[KPL];
    float S = 0.f;
#pragma unroll
    for (int j = 0; j < KPL; ++j)
      if (lane + 32 * j < M) {
        pi[j] = zpi[j] * inv_se;
        float n1 = x1 - mu1[j], n2 = x2 - mu2[j];
        float s1s2 = s1[j] * s2[j];
        float a = n1 / s1[j], bb = n2 / s2[j];
        float zz = a * a + bb * bb - 2.f * (rho[j] * (n1 * n2)) / s1s2;
        float neg = 1.f - rho[j] * rho[j];
        dens[j] = expf(-zz / (2.f * neg)) / (6.283185307179586f * (s1s2 * sqrtf(neg)));
        S += dens[j] * pi[j];
      }
    S = warp_sum(S);
    const float row_shape = -logf(fmaxf(S, 1e-20f));

    // pen softmax cross-entropy (all lanes redundantly)
    const float z0 = z[0], z1 = z[1], z2 = z[2];
    const float zm = fmaxf(z0, fmaxf(z1, z2));
    const float e0 = expf(z0 - zm), e1 = expf(z1 - zm), e2 = expf(z2 - zm);
    const float lse = zm + logf(e0 + e1 + e2);
    const float xent = -(p_eos * (z0 - lse) + p_eoc * (z1 - lse) + p_cont * (z2 - lse));
    const float w = p_cont + sqrt_sif * p_eos + sif * p_eoc;
    const float row_pen = xent * w;
    if (lane == 0) {
      my_shape += (double)row_shape;
      my_pen += (double)row_pen;
    }

    if (mdn_out != nullptr) {
      float *o = mdn_out + ((int64_t)b * T + t) * NOUT;
      if (lane == 0) { o[0] = z0; o[1] = z1; o[2] = z2; }
#pragma unroll
      for (int j = 0; j < KPL; ++j) {
        int k = lane + 32 * j;
        if (k < M) {
          o[3 + k] = pi[j];
          o[3 + M + k] = mu1[j];
          o[3 + 2 * M + k] = mu2[j];
          o[3 + 3 * M + k] = s1[j];
          o[3 + 4 * M + k] = s2[j];
          o[3 + 5 * M + k] = rho[j];
        }
      }
    }
    if (want_grad) {
      const bool live = S >= 1e-20f;
      const float invS = live ? 1.0f / S : 0.f;
#pragma unroll
      for (int j = 0; j < KPL; ++j) {
        int k = lane + 32 * j;
        if (k < M) {
          float gam = pi[j] * dens[j] * invS;
          float a = (x1 - mu1[j]) / s1[j], bb = (x2 - mu2[j]) / s2[j];
          float q = 1.f - rho[j] * rho[j];
          float Z = a * a + bb * bb - 2.f * rho[j] * a * bb;
          float ga = gam * inv_rows;
          const float g0 = ((live ? pi[j] : 0.f) - gam) * inv_rows;
          const float g1 = -ga * (a - rho[j] * bb) / (s1[j] * q);
          const float g2 = -ga * (bb - rho[j] * a) / (s2[j] * q);
          const float g3 = -ga * (a * (a - rho[j] * bb) / q - 1.f);
          const float g4 = -ga * (bb * (bb - rho[j] * a) / q - 1.f);
          const float g5 = -ga * (a * bb + rho[j] * (1.f - Z / q));
          z[3 + k] = g0;
          z[3 + M + k] = g1;
          z[3 + 2 * M + k] = g2;
          z[3 + 3 * M + k] = g3;
          z[3 + 4 * M + k] = g4;
          z[3 + 5 * M + k] = g5;
          if (dl_bf != nullptr) {
            __nv_bfloat16 *d = dl_bf + r * noutp + 3 + k;
            d[0] = __float2bfloat16_rn(g0); d[M] = __float2bfloat16_rn(g1); d[2 * M] = __float2bfloat16_rn(g2);
            d[3 * M] = __float2bfloat16_rn(g3); d[4 * M] = __float2bfloat16_rn(g4); d[5 * M] = __float2bfloat16_rn(g5);
          }
        }
      }
      if (lane == 0) {
        const float psum = p_eos + p_eoc + p_cont;
        const float inv_e = 1.0f / (e0 + e1 + e2);
        const float ws = w * inv_rows;
        const float q0 = ws * (e0 * inv_e * psum - p_eos), q1 = ws * (e1 * inv_e * psum - p_eoc);
        const float q2 = ws * (e2 * inv_e * psum - p_cont);
        z[0] = q0;
        z[1] = q1;
        z[2] = q2;
        if (dl_bf != nullptr) {
          dl_bf[r * noutp + 0] = __float2bfloat16_rn(q0);
          dl_bf[r * noutp + 1] = __float2bfloat16_rn(q1);
          dl_bf[r * noutp + 2] = __float2bfloat16_rn(q2);
        }
      }
    }
  }
  if (lane == 0) { sh_shape[wid] = my_shape; sh_pen[wid] = my_pen; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, c = 0.0;
    for (int i = 0; i < 8; ++i) { a += sh_shape[i]; c += sh_pen[i]; }
    atomicAdd(&scal[1], a * (double)inv_rows);
    atomicAdd(&scal[2], c * (double)inv_rows);
  }
}

__global__ void finalize_losses_kernel(const double *__restrict__ scal, float *__restrict__ losses) {
  float s = (float)scal[1], p = (float)scal[2];
  losses[0] = s + p;  // model.py:138
  losses[1] = s;
  losses[2] = p;
}

// Global-norm clip fused into Adam.  Phase 1: every block adds its share of sum(g^2) to a
// double accumulator.  Grid barrier (cooperative launch => all blocks resident).  Phase 2:
// scale = clip*min(1/gn, 1/clip) (TF clip_by_global_norm), then TF ApplyAdam with alpha
// precomputed on the host from the beta powers.  The second read of g hits L2.
__global__ void __launch_bounds__(256)
clip_adam_kernel(int64_t P, const float *__restrict__ g, float *__restrict__ p, float *__restrict__ m,
                 float *__restrict__ v, float alpha, float clip, float one_m_b1, float one_m_b2,
                 float eps, double *__restrict__ scal, unsigned int *__restrict__ ctr,
                 float *__restrict__ gnorm_out) {
  __shared__ double red[8];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  float ss = 0.f;
  double dss = 0.0;
  int cnt = 0;
  for (int64_t i = tid; i < P; i += nth) {
    float x = g[i];
    ss = fmaf(x, x, ss);
    if (++cnt == 64) { dss += (double)ss; ss = 0.f; cnt = 0; }
  }
  dss += (double)ss;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dss += __shfl_xor_sync(0xffffffffu, dss, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dss;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; ++i) t += red[i];
    atomicAdd(&scal[3], t);
    __threadfence();
    atomicAdd(ctr, 1u);
    while (*((volatile unsigned int *)ctr) < gridDim.x) { }
    __threadfence();
  }
  __syncthreads();
  const float gn = (float)sqrt(*((volatile double *)&scal[3]));
  const float scale = clip * fminf(1.0f / gn, 1.0f / clip);
  if (tid == 0 && gnorm_out != nullptr) *gnorm_out = gn;
  for (int64_t i = tid; i < P; i += nth) {
    float gi = g[i] * scale;
    float mi = m[i], vi = v[i];
    mi = mi + (gi - mi) * one_m_b1;
    vi = vi + (gi * gi - vi) * one_m_b2;
    m[i] = mi;
    v[i] = vi;
    p[i] = p[i] - alpha * mi / (sqrtf(vi) + eps);
  }
}

}  // namespace

int sk_head(skrnn_handle *h, cudaStream_t st, int B, int T, const float *y_bt5, float *mdn_out,
            bool want_grad, int64_t loss_rows) {
  SK_CHECK(h->M <= 32 * KPL, "num_mixture %d > %d unsupported", h->M, 32 * KPL);
  int64_t R = (int64_t)B * T;
  int blocks = (int)std::min<int64_t>((R + 7) / 8, (int64_t)8 * h->num_sms);
  SkLaunch scope(h, st, SK_CAT_HEAD);
  mdn_head_kernel<<<blocks, 256, 0, st>>>(h->ws.logits, y_bt5, mdn_out, B, T, h->M, want_grad ? 1 : 0,
                                         (float)(1.0 / (double)loss_rows),
                                         h->cfg.stroke_importance_factor, h->scal,
                                         want_grad ? h->ws.dlogits_bf : nullptr, h->NOUTP);
  SK_CUDA(cudaGetLastError());
  return 0;
}

int sk_finalize_losses(skrnn_handle *h, cudaStream_t st, float *losses_dev) {
  SkLaunch scope(h, st, SK_CAT_MISC);
  finalize_losses_kernel<<<1, 1, 0, st>>>(h->scal, losses_dev);
  SK_CUDA(cudaGetLastError());
  return 0;
}

int sk_clip_adam(skrnn_handle *h, cudaStream_t st, float lr, float *gnorm_out) {
  SK_CUDA(cudaMemsetAsync(&h->scal[3], 0, sizeof(double), st));
  SK_CUDA(cudaMemsetAsync(h->sync_ctr, 0, sizeof(unsigned int), st));
  float alpha = (float)((double)lr * sqrt(1.0 - h->beta2_power) / (1.0 - h->beta1_power));
  int blocks_per_sm = 0;
  SK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, clip_adam_kernel, 256, 0));
  int grid = h->num_sms * max(1, min(blocks_per_sm, 4));
  int64_t need = (h->P + 255) / 256;
  if (need < grid) grid = need < 1 ? 1 : (int)need;
  int64_t P = h->P;
  const float *g = h->grads;
  float *p = h->params, *m = h->adam_m, *v = h->adam_v;
  float clip = h->cfg.grad_clip, omb1 = 1.0f - 0.9f, omb2 = 1.0f - 0.999f, eps = 1e-3f;
  double *scal = h->scal;
  unsigned int *ctr = h->sync_ctr;
  void *args[] = {&P, &g, &p, &m, &v, &alpha, &clip, &omb1, &omb2, &eps, &scal, &ctr, &gnorm_out};
  {
    SkLaunch scope(h, st, SK_CAT_OPTIM);
    SK_CUDA(cudaLaunchCooperativeKernel((void *)clip_adam_kernel, dim3(grid), dim3(256), args, 0, st));
  }
  h->beta1_power *= 0.9;
  h->beta2_power *= 0.999;
  return 0;
}
