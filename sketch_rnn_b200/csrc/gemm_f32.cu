// gemm_f32.cu -- fp32 CUDA-core GEMM used by the reference-precision (1e-4) path for the
// hoisted projections of the cell stack (reference model.py:77, TF _linear on [x,h]: the x half
// of every layer for all timesteps at once), their weight/input gradients (model.py:182) and the
// output projection (model.py:98-99).  128x128x16 tiles, 8x8 register micro-tiles, optional
// split-K with atomic accumulation for the tall-skinny weight-gradient products.
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16;
constexpr int APAD = 4;

// Load 4 consecutive elements along the contiguous axis, guarded.
__device__ __forceinline__ float4 load4_guarded(const float *p, int valid, bool aligned) {
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid >= 4 && aligned) {
    v = *reinterpret_cast<const float4 *>(p);
  } else {
    if (valid > 0) v.x = p[0];
    if (valid > 1) v.y = p[1];
    if (valid > 2) v.z = p[2];
    if (valid > 3) v.w = p[3];
  }
  return v;
}

template <bool A_KCONTIG, bool B_NCONTIG>
__global__ void __launch_bounds__(256)
sgemm_kernel(int M, int N, int K, const float *__restrict__ A, int lda, const float *__restrict__ B,
             int ldb, float *__restrict__ C, int ldc, const float *__restrict__ bias, int accumulate,
             int klen) {
  __shared__ __align__(16) float As[BK][BM + APAD];
  __shared__ __align__(16) float Bs[BK][BN + APAD];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int bm = blockIdx.y * BM, bn = blockIdx.x * BN;
  const int k0 = blockIdx.z * klen;
  const int kend = min(K, k0 + klen);
  const bool a_al = ((reinterpret_cast<uintptr_t>(A) & 15) == 0) && (lda % 4 == 0);
  const bool b_al = ((reinterpret_cast<uintptr_t>(B) & 15) == 0) && (ldb % 4 == 0);

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float4 ra[2], rb[2];
  auto fetch = [&](int kt) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      int f = tid + s * 256;
      if (A_KCONTIG) {
        int row = f >> 2, kq = f & 3;
        int m = bm + row, k = kt + kq * 4;
        int valid = (m < M) ? (kend - k) : 0;
        ra[s] = load4_guarded(A + (size_t)m * lda + k, valid, a_al);
      } else {
        int k = kt + (f >> 5), mq = f & 31;
        int m = bm + mq * 4;
        int valid = (k < kend) ? (M - m) : 0;
        ra[s] = load4_guarded(A + (size_t)k * lda + m, valid, a_al);
      }
      if (B_NCONTIG) {
        int k = kt + (f >> 5), nq = f & 31;
        int n = bn + nq * 4;
        int valid = (k < kend) ? (N - n) : 0;
        rb[s] = load4_guarded(B + (size_t)k * ldb + n, valid, b_al);
      } else {
        int col = f >> 2, kq = f & 3;
        int n = bn + col, k = kt + kq * 4;
        int valid = (n < N) ? (kend - k) : 0;
        rb[s] = load4_guarded(B + (size_t)n * ldb + k, valid, b_al);
      }
    }
  };
  auto stash = [&]() {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      int f = tid + s * 256;
      if (A_KCONTIG) {
        int row = f >> 2, kq = f & 3;
        As[kq * 4 + 0][row] = ra[s].x;
        As[kq * 4 + 1][row] = ra[s].y;
        As[kq * 4 + 2][row] = ra[s].z;
        As[kq * 4 + 3][row] = ra[s].w;
      } else {
        int k = f >> 5, mq = f & 31;
        *reinterpret_cast<float4 *>(&As[k][mq * 4]) = ra[s];
      }
      if (B_NCONTIG) {
        int k = f >> 5, nq = f & 31;
        *reinterpret_cast<float4 *>(&Bs[k][nq * 4]) = rb[s];
      } else {
        int col = f >> 2, kq = f & 3;
        Bs[kq * 4 + 0][col] = rb[s].x;
        Bs[kq * 4 + 1][col] = rb[s].y;
        Bs[kq * 4 + 2][col] = rb[s].z;
        Bs[kq * 4 + 3][col] = rb[s].w;
      }
    }
  };

  if (k0 < kend) fetch(k0);
  for (int kt = k0; kt < kend; kt += BK) {
    __syncthreads();
    stash();
    __syncthreads();
    if (kt + BK < kend) fetch(kt + BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float4 a0 = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
      float4 a1 = *reinterpret_cast<const float4 *>(&As[k][64 + ty * 4]);
      float4 b0 = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
      float4 b1 = *reinterpret_cast<const float4 *>(&Bs[k][64 + tx * 4]);
      float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }

  const bool add_bias = (bias != nullptr) && (blockIdx.z == 0);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int m = bm + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int n = bn + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (n >= N) continue;
      float v = acc[i][j];
      if (add_bias) v += bias[n];
      float *dst = C + (size_t)m * ldc + n;
      if (accumulate) atomicAdd(dst, v);
      else *dst = v;
    }
  }
}

__global__ void colsum_kernel(const float *__restrict__ A, int64_t rows, int cols, int lda,
                              int64_t rows_per_block, float *__restrict__ out) {
  __shared__ float red[8][33];
  int c = blockIdx.x * 32 + threadIdx.x;
  int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
  int64_t r1 = min(rows, r0 + rows_per_block);
  float s = 0.f;
  if (c < cols)
    for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) s += A[r * lda + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    atomicAdd(out + c, t);
  }
}

__global__ void colsum_bf16_kernel(const __nv_bfloat16 *__restrict__ A, int64_t rows, int cols, int lda,
                                   int64_t rows_per_block, float *__restrict__ out) {
  __shared__ float red[8][33];
  int c = blockIdx.x * 32 + threadIdx.x;
  int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
  int64_t r1 = min(rows, r0 + rows_per_block);
  float s = 0.f;
  if (c < cols)
    for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) s += __bfloat162float(A[r * lda + c]);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    atomicAdd(out + c, t);
  }
}

__global__ void transpose_x_kernel(const float *__restrict__ x, float *__restrict__ x_tm,
                                   __nv_bfloat16 *__restrict__ x8, unsigned char *__restrict__ eoc, int B, int T) {
  int64_t n = (int64_t)B * T * 5;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int j = (int)(i % 5);
    int64_t r = i / 5;
    int b = (int)(r % B);
    int t = (int)(r / B);
    float v = x[((int64_t)b * T + t) * 5 + j];
    x_tm[i] = v;
    if (j == 3) eoc[r] = v > 0.f ? 1 : 0;  // end-of-content flag of the step's input, time-major
    if (x8 != nullptr) x8[r * 8 + j] = __float2bfloat16_rn(v);  // columns 5..7 stay zero from allocation
  }
}

// fp32 master weights -> bf16 mirror (16-byte aligned tensors; output_w padded to noutp columns)
__global__ void cast_params_kernel(const float *__restrict__ p, __nv_bfloat16 *__restrict__ pb, int nseg,
                                   const int64_t *__restrict__ seg) {
  // seg[4*s + {0,1,2,3}] = src offset, dst offset, rows*cols (src elements), cols ; dst ld = seg-specific below
  for (int s = 0; s < nseg; ++s) {
    const int64_t so = seg[5 * s], dof = seg[5 * s + 1], cnt = seg[5 * s + 2], cols = seg[5 * s + 3], ldd = seg[5 * s + 4];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < cnt; i += (int64_t)gridDim.x * blockDim.x) {
      int64_t r = i / cols, c = i % cols;
      pb[dof + r * ldd + c] = __float2bfloat16_rn(p[so + i]);
    }
  }
}

// lstm input weights with gate-interleaved columns: dst[k][4u+g] = W[k][g*H+u] for k < I (zero rows up to ipad),
// bias_il[4u+g] = b[g*H+u]
__global__ void interleave_wx_kernel(const float *__restrict__ mat, const float *__restrict__ bias, int I, int ipad, int H,
                                     __nv_bfloat16 *__restrict__ dst, float *__restrict__ bias_il) {
  const int G = 4 * H;
  const int64_t total = (int64_t)ipad * G;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i / G), ci = (int)(i % G);
    const int c = (ci & 3) * H + (ci >> 2);
    dst[i] = __float2bfloat16_rn(k < I ? mat[(size_t)k * G + c] : 0.f);
    if (k == 0) bias_il[ci] = bias[c];
  }
}

}  // namespace

int sk_gemm_f32(skrnn_handle *h, cudaStream_t st, int M, int N, int K, const float *A, int lda,
                bool a_kcontig, const float *B, int ldb, bool b_ncontig, float *C, int ldc,
                const float *bias, bool accumulate) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  int gx = (N + BN - 1) / BN, gy = (M + BM - 1) / BM;
  int nsplit = 1;
  if (accumulate) {
    int tiles = gx * gy;
    int want = (2 * h->num_sms + tiles - 1) / tiles;
    int maxsplit = (K + 8 * BK - 1) / (8 * BK);
    nsplit = max(1, min(want, maxsplit));
  }
  int klen = ((K + nsplit - 1) / nsplit + BK - 1) / BK * BK;
  nsplit = (K + klen - 1) / klen;
  dim3 grid(gx, gy, nsplit);
  int acc = accumulate ? 1 : 0;
  SkLaunch scope(h, st, SK_CAT_GEMM);
  if (a_kcontig && b_ncontig)
    sgemm_kernel<true, true><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, acc, klen);
  else if (a_kcontig && !b_ncontig)
    sgemm_kernel<true, false><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, acc, klen);
  else if (!a_kcontig && b_ncontig)
    sgemm_kernel<false, true><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, acc, klen);
  else
    sgemm_kernel<false, false><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, acc, klen);
  SK_CUDA(cudaGetLastError());
  return 0;
}

int sk_colsum(skrnn_handle *h, cudaStream_t st, const float *A, int64_t rows, int cols, int lda,
              float *out) {
  if (rows <= 0 || cols <= 0) return 0;
  int gx = (cols + 31) / 32;
  int gy = (int)std::min<int64_t>((rows + 255) / 256, (int64_t)std::max(1, 4 * h->num_sms / gx));
  int64_t rpb = (rows + gy - 1) / gy;
  SkLaunch scope(h, st, SK_CAT_MISC);
  colsum_kernel<<<dim3(gx, gy), dim3(32, 8), 0, st>>>(A, rows, cols, lda, rpb, out);
  SK_CUDA(cudaGetLastError());
  return 0;
}

int sk_colsum_bf16(skrnn_handle *h, cudaStream_t st, const __nv_bfloat16 *A, int64_t rows, int cols, int lda,
                   float *out) {
  if (rows <= 0 || cols <= 0) return 0;
  int gx = (cols + 31) / 32;
  int gy = (int)std::min<int64_t>((rows + 255) / 256, (int64_t)std::max(1, 4 * h->num_sms / gx));
  int64_t rpb = (rows + gy - 1) / gy;
  SkLaunch scope(h, st, SK_CAT_MISC);
  colsum_bf16_kernel<<<dim3(gx, gy), dim3(32, 8), 0, st>>>(A, rows, cols, lda, rpb, out);
  SK_CUDA(cudaGetLastError());
  return 0;
}

int sk_transpose_x(skrnn_handle *h, cudaStream_t st, int B, int T, const float *x_bt5) {
  int64_t n = (int64_t)B * T * 5;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)8 * h->num_sms);
  SkLaunch scope(h, st, SK_CAT_MISC);
  transpose_x_kernel<<<blocks, 256, 0, st>>>(x_bt5, h->ws.x_tm, h->ws.x8_bf, h->ws.eoc, B, T);
  SK_CUDA(cudaGetLastError());
  return 0;
}

int sk_cast_params(skrnn_handle *h, cudaStream_t st) {
  if (h->params_bf == nullptr) return 0;
  SkLaunch scope(h, st, SK_CAT_MISC);
  int blocks = (int)std::min<int64_t>((h->P + 255) / 256, (int64_t)4 * h->num_sms);
  cast_params_kernel<<<blocks, 256, 0, st>>>(h->params, h->params_bf, h->cast_nseg, h->cast_seg_dev);
  SK_CUDA(cudaGetLastError());
  for (int l = 0; l < h->L; ++l) {
    if (h->wx_il[l] == nullptr) continue;
    const int I = h->lp[l].I, ip = (I + 7) / 8 * 8;
    const int64_t total = (int64_t)ip * h->G;
    SkLaunch scope2(h, st, SK_CAT_MISC);
    interleave_wx_kernel<<<(int)std::min<int64_t>((total + 255) / 256, (int64_t)2 * h->num_sms), 256, 0, st>>>(
        h->params + h->lp[l].w[0], h->params + h->lp[l].b[0], I, ip, h->H, h->wx_il[l], h->bias_il[l]);
    SK_CUDA(cudaGetLastError());
  }
  return 0;
}
