// rnn_f32.cu -- fp32 persistent recurrent kernels (forward unroll and BPTT) for one layer of the
// cell stack, reference model.py:66-95 (custom_rnn_autodecoder with the end-of-content state
// reset) around TF BasicLSTMCell / GRUCell / BasicRNNCell (model.py:14-27) and their gradients
// (model.py:182).
//
// Decomposition.  The x half of every cell's [x,h]*W is hoisted into one GEMM per layer over
// all timesteps (gemm_f32.cu); what is left per timestep is h_{t-1}[B,H] * Wh[H,G] plus the gate
// math.  One kernel launch runs all T timesteps of one layer:
//   * a thread-block cluster owns RB batch rows; rank r of the cluster owns hidden units
//     [r*HS, (r+1)*HS) and keeps the recurrent weight columns of those units resident in shared
//     memory for the whole launch (packed per rank by pack_weights_kernel);
//   * per step every CTA gathers the full h_{t-1} tile of its rows (written by all ranks of the
//     cluster in the previous step) from L2, multiplies it with its weight slice (K split over
//     the 8 warps, rows x columns register tiles), reduces the K-partials through shared memory
//     and applies the fused gate nonlinearities, cell update, eoc reset select and (top layer)
//     output dropout in the epilogue;
//   * ranks meet at one barrier.cluster (release/acquire) per dependent product.
// Activations are time-major ([T,B,*]) so a cluster's tile at step t is one contiguous block.
#include "common.cuh"
#include "skrnn_math.h"

namespace {

constexpr int NWARP = 8;
constexpr int NTHREADS = 256;

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_nctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_id_x() {
  unsigned r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// 0 or 1/keep for element (global row, t, unit) of the top layer's output (model.py:29-30).
__device__ __forceinline__ float drop_scale(const float *mask, int b, int T, int t, int H, int unit,
                                            int row0, uint64_t seed, float keep) {
  if (mask != nullptr) return mask[((int64_t)b * T + t) * H + unit] * (1.0f / keep);
  return skm_u01(skm_drop_bits(seed, (uint32_t)(row0 + b), (uint32_t)t, (uint32_t)unit)) < keep ? (1.0f / keep) : 0.0f;
}

struct RnnArgs {
  int B, T, H, HS, G;
  int row0;
  int dropout;  // apply output dropout (top layer, training, keep < 1)
  float keep;
  uint64_t seed;
  const float *x_bt5;     // [B,T,5] caller layout: eoc flag = x[b,t,3]
  float *xw;              // [T,B,G]
  float *hprev;           // [T,B,H]
  float *cprev;           // [T,B,H]  (lstm: c_t ; gru: r*h)
  float *out;             // [T,B,H]
  float *dout;            // [T,B,H]  (backward)
  const float *state_in;  // [B,S]
  float *state_out;       // [B,S]
  const float *mask;      // [B,T,H] or null
  const float *wpack;     // packed slices, see pack_weights_kernel
  // optional bf16 mirrors for the tensor-core GEMMs that consume these activations (null in fp32 mode)
  __nv_bfloat16 *out_bf, *hprev_bf, *aux_bf, *dpre_bf;
  int w_in_smem;
  int ncp0, ncp1;         // padded column counts of the (up to) two packed matrices
  int k0, k1;             // K of the two packed matrices
};

// AT holds the gathered A tile transposed and padded: element (k, row) at ((k + k/4)*RB + row).
template <int RB>
__device__ __forceinline__ void gather_tile(const float *__restrict__ src, int lda, int K, int rows_valid,
                                            float *__restrict__ AT, int tid) {
  const int nq = K >> 2;
  for (int idx = tid; idx < RB * nq; idx += NTHREADS) {
    int rr = idx % RB, kq = idx / RB;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rr < rows_valid) v = __ldcg(reinterpret_cast<const float4 *>(src + (size_t)rr * lda + kq * 4));
    float *d = AT + (5 * kq) * RB + rr;
    d[0] = v.x;
    d[RB] = v.y;
    d[2 * RB] = v.z;
    d[3 * RB] = v.w;
  }
}

// part[(warp*RB + row)*NCP + col] = sum over this warp's K range of A[row][k] * W[k][col].
template <int RB, int CPT, bool WSMEM>
__device__ __forceinline__ void gemm_phase(const float *__restrict__ AT, int K, const float *__restrict__ W,
                                           int NCP, float *__restrict__ part, int warp, int lane) {
  float acc[RB][CPT];
#pragma unroll
  for (int r = 0; r < RB; ++r)
#pragma unroll
    for (int j = 0; j < CPT; ++j) acc[r][j] = 0.f;
  const int K4 = K >> 2;
  const int gpw = (K4 + NWARP - 1) / NWARP;
  const int g_beg = warp * gpw, g_end = min(K4, g_beg + gpw);
  for (int g4 = g_beg; g4 < g_end; ++g4) {
    const float4 *ap = reinterpret_cast<const float4 *>(AT + (size_t)(5 * g4) * RB);
    const float *wp = W + (size_t)(4 * g4) * NCP + lane;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float a[RB];
#pragma unroll
      for (int q = 0; q < RB / 4; ++q) {
        float4 t4 = ap[kk * (RB / 4) + q];
        a[4 * q + 0] = t4.x; a[4 * q + 1] = t4.y; a[4 * q + 2] = t4.z; a[4 * q + 3] = t4.w;
      }
      float wv[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) wv[j] = WSMEM ? wp[kk * NCP + 32 * j] : __ldg(wp + kk * NCP + 32 * j);
#pragma unroll
      for (int r = 0; r < RB; ++r)
#pragma unroll
        for (int j = 0; j < CPT; ++j) acc[r][j] = fmaf(a[r], wv[j], acc[r][j]);
    }
  }
#pragma unroll
  for (int r = 0; r < RB; ++r)
#pragma unroll
    for (int j = 0; j < CPT; ++j) part[(size_t)(warp * RB + r) * NCP + lane + 32 * j] = acc[r][j];
}

template <int RB>
__device__ __forceinline__ float sum_part(const float *part, int NCP, int rr, int col) {
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < NWARP; ++w) s += part[(size_t)(w * RB + rr) * NCP + col];
  return s;
}

__device__ __forceinline__ void load_w_smem(float *dst, const float *src, int n, int tid) {
  const float4 *s4 = reinterpret_cast<const float4 *>(src);
  float4 *d4 = reinterpret_cast<float4 *>(dst);
  for (int i = tid; i < (n >> 2); i += NTHREADS) d4[i] = __ldg(s4 + i);
}

constexpr int MAXIT = 4;  // (row,unit) items per thread: RB*HS <= 1024

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
// CELL: 0 lstm, 1 gru, 2 rnn.  CPT0 / CPT1: register columns per thread of the 1st / 2nd product.
template <int CELL, int RB, int CPT0, int CPT1, bool WSMEM>
__global__ void __launch_bounds__(NTHREADS, 1) rnn_fwd_kernel(RnnArgs a) {
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rank = cluster_ctarank();
  const int rowbase = cluster_id_x() * RB;
  const int rows_valid = min(RB, a.B - rowbase);
  const int H = a.H, HS = a.HS, G = a.G, B = a.B, T = a.T;
  const int u0 = rank * HS;
  const int S = (CELL == 0) ? 2 * H : H;
  const int hoff = (CELL == 0) ? H : 0;  // h columns inside the [B,S] state

  // shared memory carve-up
  const int64_t w0n = (int64_t)a.k0 * a.ncp0, w1n = (int64_t)a.k1 * a.ncp1;
  const float *W0 = a.wpack + (size_t)rank * (w0n + w1n);
  const float *W1 = W0 + w0n;
  float *sp = smem;
  if (WSMEM) {
    load_w_smem(sp, W0, (int)(w0n + w1n), tid);
    W0 = sp;
    W1 = sp + w0n;
    sp += w0n + w1n;
  }
  float *AT = sp;   sp += (size_t)(H + H / 4) * RB;
  float *part = sp; sp += (size_t)NWARP * RB * max(a.ncp0, a.ncp1);
  float *cs = sp;   sp += RB * HS;  // lstm: c state ; gru: u gate
  float *hown = sp; sp += RB * HS;  // gru: h of own units

  // t = 0 inputs from the fed initial state
  for (int it = tid; it < RB * HS; it += NTHREADS) {
    int rr = it / HS, u = it % HS, b = rowbase + rr;
    if (b < B) {
      float hi = a.state_in[(size_t)b * S + hoff + u0 + u];
      a.hprev[((size_t)0 * B + b) * H + u0 + u] = hi;
      if (a.hprev_bf) a.hprev_bf[((size_t)0 * B + b) * H + u0 + u] = __float2bfloat16_rn(hi);
      if (CELL == 0) {
        float ci = a.state_in[(size_t)b * S + u0 + u];
        cs[it] = ci;
        a.cprev[((size_t)0 * B + b) * H + u0 + u] = ci;
      }
    }
  }
  cluster_sync_all();

  for (int t = 0; t < T; ++t) {
    const size_t trow = (size_t)t * B + rowbase;
    gather_tile<RB>(a.hprev + trow * H, H, H, rows_valid, AT, tid);
    __syncthreads();
    gemm_phase<RB, CPT0, WSMEM>(AT, H, W0, a.ncp0, part, warp, lane);
    __syncthreads();

    if (CELL == 1) {
      // phase 1 epilogue: r, u gates; publish r*h for the candidate product
      for (int it = tid; it < RB * HS; it += NTHREADS) {
        int rr = it / HS, u = it % HS, b = rowbase + rr;
        if (b >= B) continue;
        float *xw = a.xw + ((size_t)t * B + b) * G;
        float r = sigmoidf_(xw[u0 + u] + sum_part<RB>(part, a.ncp0, rr, u * 2 + 0));
        float z = sigmoidf_(xw[H + u0 + u] + sum_part<RB>(part, a.ncp0, rr, u * 2 + 1));
        int k = u0 + u;
        float hv = AT[(k + (k >> 2)) * RB + rr];
        xw[u0 + u] = r;
        xw[H + u0 + u] = z;
        cs[it] = z;
        hown[it] = hv;
        a.cprev[((size_t)t * B + b) * H + u0 + u] = r * hv;
        if (a.aux_bf) a.aux_bf[((size_t)t * B + b) * H + u0 + u] = __float2bfloat16_rn(r * hv);
      }
      cluster_sync_all();
      gather_tile<RB>(a.cprev + trow * H, H, H, rows_valid, AT, tid);
      __syncthreads();
      gemm_phase<RB, CPT1, WSMEM>(AT, H, W1, a.ncp1, part, warp, lane);
      __syncthreads();
    }

    for (int it = tid; it < RB * HS; it += NTHREADS) {
      int rr = it / HS, u = it % HS, b = rowbase + rr;
      if (b >= B) continue;
      float *xw = a.xw + ((size_t)t * B + b) * G;
      const bool eoc = a.x_bt5[((size_t)b * T + t) * 5 + 3] > 0.f;
      float h2, hnext;
      if (CELL == 0) {
        float i = sigmoidf_(xw[0 * H + u0 + u] + sum_part<RB>(part, a.ncp0, rr, u * 4 + 0));
        float j = tanhf(xw[1 * H + u0 + u] + sum_part<RB>(part, a.ncp0, rr, u * 4 + 1));
        float f = sigmoidf_(xw[2 * H + u0 + u] + sum_part<RB>(part, a.ncp0, rr, u * 4 + 2) + 1.0f);
        float o = sigmoidf_(xw[3 * H + u0 + u] + sum_part<RB>(part, a.ncp0, rr, u * 4 + 3));
        float c = cs[it];
        float c2 = c * f + i * j;
        h2 = tanhf(c2) * o;
        xw[0 * H + u0 + u] = i;
        xw[1 * H + u0 + u] = j;
        xw[2 * H + u0 + u] = f;
        xw[3 * H + u0 + u] = o;
        float cnext = eoc ? a.state_in[(size_t)b * S + u0 + u] : c2;
        cs[it] = cnext;
        if (t + 1 < T) a.cprev[((size_t)(t + 1) * B + b) * H + u0 + u] = cnext;
        else a.state_out[(size_t)b * S + u0 + u] = cnext;
      } else if (CELL == 1) {
        float cand = tanhf(xw[2 * H + u0 + u] + sum_part<RB>(part, a.ncp1, rr, u));
        float z = cs[it], hv = hown[it];
        h2 = z * hv + (1.0f - z) * cand;
        xw[2 * H + u0 + u] = cand;
      } else {
        h2 = tanhf(xw[u0 + u] + sum_part<RB>(part, a.ncp0, rr, u));
        xw[u0 + u] = h2;
      }
      float ov = h2;
      if (a.dropout) ov *= drop_scale(a.mask, b, T, t, H, u0 + u, a.row0, a.seed, a.keep);
      a.out[((size_t)t * B + b) * H + u0 + u] = ov;
      if (a.out_bf) a.out_bf[((size_t)t * B + b) * H + u0 + u] = __float2bfloat16_rn(ov);
      hnext = eoc ? a.state_in[(size_t)b * S + hoff + u0 + u] : h2;
      if (t + 1 < T) {
        a.hprev[((size_t)(t + 1) * B + b) * H + u0 + u] = hnext;
        if (a.hprev_bf) a.hprev_bf[((size_t)(t + 1) * B + b) * H + u0 + u] = __float2bfloat16_rn(hnext);
      }
      else a.state_out[(size_t)b * S + hoff + u0 + u] = hnext;
    }
    cluster_sync_all();
  }
}

// ------------------------------------------------------------------------------------------
// backward (BPTT).  On entry xw holds the saved gate activations; on exit d(pre-activation).
// ------------------------------------------------------------------------------------------
template <int CELL, int RB, int CPT0, int CPT1, bool WSMEM>
__global__ void __launch_bounds__(NTHREADS, 1) rnn_bwd_kernel(RnnArgs a) {
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rank = cluster_ctarank();
  const int rowbase = cluster_id_x() * RB;
  const int rows_valid = min(RB, a.B - rowbase);
  const int H = a.H, HS = a.HS, G = a.G, B = a.B, T = a.T;
  const int u0 = rank * HS;

  const int64_t w0n = (int64_t)a.k0 * a.ncp0, w1n = (int64_t)a.k1 * a.ncp1;
  const float *W0 = a.wpack + (size_t)rank * (w0n + w1n);
  const float *W1 = W0 + w0n;
  float *sp = smem;
  if (WSMEM) {
    load_w_smem(sp, W0, (int)(w0n + w1n), tid);
    W0 = sp;
    W1 = sp + w0n;
    sp += w0n + w1n;
  }
  const int KMAX = max(a.k0, a.k1);
  float *AT = sp;   sp += (size_t)(KMAX + KMAX / 4) * RB;
  float *part = sp; sp += (size_t)NWARP * RB * max(a.ncp0, a.ncp1);
  float *dhc = sp;  sp += RB * HS;  // carried d h_state
  float *dcc = sp;  sp += RB * HS;  // lstm: carried d c_state

  for (int it = tid; it < RB * HS; it += NTHREADS) { dhc[it] = 0.f; dcc[it] = 0.f; }
  __syncthreads();

  for (int t = T - 1; t >= 0; --t) {
    const size_t trow = (size_t)t * B + rowbase;
    // per-thread stash for gru between its phases
    float st_du[MAXIT], st_dh[MAXIT], st_r[MAXIT], st_h[MAXIT], st_z[MAXIT];
    int n_it = 0;
    for (int it = tid; it < RB * HS; it += NTHREADS, ++n_it) {
      int rr = it / HS, u = it % HS, b = rowbase + rr;
      if (b >= B) continue;
      float *xw = a.xw + ((size_t)t * B + b) * G;
      const bool eoc = a.x_bt5[((size_t)b * T + t) * 5 + 3] > 0.f;
      float d_o = a.dout[((size_t)t * B + b) * H + u0 + u];
      if (a.dropout) d_o *= drop_scale(a.mask, b, T, t, H, u0 + u, a.row0, a.seed, a.keep);
      const float dh2 = d_o + (eoc ? 0.f : dhc[it]);
      if (CELL == 0) {
        float i = xw[0 * H + u0 + u], j = xw[1 * H + u0 + u], f = xw[2 * H + u0 + u], o = xw[3 * H + u0 + u];
        float c = a.cprev[((size_t)t * B + b) * H + u0 + u];
        float tc = tanhf(c * f + i * j);
        float dc2 = (eoc ? 0.f : dcc[it]) + dh2 * o * (1.0f - tc * tc);
        float d_oo = dh2 * tc;
        dcc[it] = dc2 * f;
        const float d0 = dc2 * j * i * (1.0f - i), d1 = dc2 * i * (1.0f - j * j);
        const float d2 = dc2 * c * f * (1.0f - f), d3 = d_oo * o * (1.0f - o);
        xw[0 * H + u0 + u] = d0;
        xw[1 * H + u0 + u] = d1;
        xw[2 * H + u0 + u] = d2;
        xw[3 * H + u0 + u] = d3;
        if (a.dpre_bf) {
          __nv_bfloat16 *db = a.dpre_bf + ((size_t)t * B + b) * G + u0 + u;
          db[0] = __float2bfloat16_rn(d0); db[H] = __float2bfloat16_rn(d1);
          db[2 * H] = __float2bfloat16_rn(d2); db[3 * H] = __float2bfloat16_rn(d3);
        }
      } else if (CELL == 1) {
        float r = xw[u0 + u], z = xw[H + u0 + u], cand = xw[2 * H + u0 + u];
        float hv = a.hprev[((size_t)t * B + b) * H + u0 + u];
        float du = dh2 * (hv - cand);
        float dcand = dh2 * (1.0f - z);
        const float dpc = dcand * (1.0f - cand * cand);
        xw[2 * H + u0 + u] = dpc;
        if (a.dpre_bf) a.dpre_bf[((size_t)t * B + b) * G + 2 * H + u0 + u] = __float2bfloat16_rn(dpc);
        if (n_it < MAXIT) { st_du[n_it] = du; st_dh[n_it] = dh2 * z; st_r[n_it] = r; st_h[n_it] = hv; st_z[n_it] = z; }
      } else {
        float h2 = xw[u0 + u];
        const float dp = dh2 * (1.0f - h2 * h2);
        xw[u0 + u] = dp;
        if (a.dpre_bf) a.dpre_bf[((size_t)t * B + b) * G + u0 + u] = __float2bfloat16_rn(dp);
      }
    }
    cluster_sync_all();

    if (CELL == 1) {
      // d(r*h) = dprec * Wc_h^T  (K = H over candidate units)
      gather_tile<RB>(a.xw + trow * G + 2 * H, G, H, rows_valid, AT, tid);
      __syncthreads();
      gemm_phase<RB, CPT0, WSMEM>(AT, H, W0, a.ncp0, part, warp, lane);
      __syncthreads();
      n_it = 0;
      for (int it = tid; it < RB * HS; it += NTHREADS, ++n_it) {
        int rr = it / HS, u = it % HS, b = rowbase + rr;
        if (b >= B) continue;
        float *xw = a.xw + ((size_t)t * B + b) * G;
        float drh = sum_part<RB>(part, a.ncp0, rr, u);
        float r = st_r[n_it], z = st_z[n_it], hv = st_h[n_it];
        float dr = drh * hv;
        st_dh[n_it] += drh * r;
        const float dpr = dr * r * (1.0f - r), dpz = st_du[n_it] * z * (1.0f - z);
        xw[u0 + u] = dpr;
        xw[H + u0 + u] = dpz;
        if (a.dpre_bf) {
          a.dpre_bf[((size_t)t * B + b) * G + u0 + u] = __float2bfloat16_rn(dpr);
          a.dpre_bf[((size_t)t * B + b) * G + H + u0 + u] = __float2bfloat16_rn(dpz);
        }
      }
      cluster_sync_all();
      if (t == 0) break;
      gather_tile<RB>(a.xw + trow * G, G, 2 * H, rows_valid, AT, tid);
      __syncthreads();
      gemm_phase<RB, CPT1, WSMEM>(AT, 2 * H, W1, a.ncp1, part, warp, lane);
      __syncthreads();
      n_it = 0;
      for (int it = tid; it < RB * HS; it += NTHREADS, ++n_it) {
        int rr = it / HS, u = it % HS, b = rowbase + rr;
        if (b >= B) continue;
        dhc[it] = st_dh[n_it] + sum_part<RB>(part, a.ncp1, rr, u);
      }
      __syncthreads();
    } else {
      if (t == 0) break;
      gather_tile<RB>(a.xw + trow * G, G, G, rows_valid, AT, tid);
      __syncthreads();
      gemm_phase<RB, CPT0, WSMEM>(AT, G, W0, a.ncp0, part, warp, lane);
      __syncthreads();
      for (int it = tid; it < RB * HS; it += NTHREADS) {
        int rr = it / HS, u = it % HS, b = rowbase + rr;
        if (b >= B) continue;
        dhc[it] = sum_part<RB>(part, a.ncp0, rr, u);
      }
      __syncthreads();
    }
  }
}

// ------------------------------------------------------------------------------------------
// weight packing: per rank, the K x NCP slices the kernels above multiply with.
//   dir 0 (forward):  lstm  W0[k][u*4+g] = Wh[k][g*H + u0+u]                       K = H
//                     gru   W0[k][u*2+g] = Wg_h[k][g*H + u0+u] ; W1[k][u] = Wc_h[k][u0+u]
//                     rnn   W0[k][u]     = Wh[k][u0+u]
//   dir 1 (backward): lstm  W0[k=g*H+v][u] = Wh[u0+u][g*H+v]                        K = 4H
//                     gru   W0[k=v][u] = Wc_h[u0+u][v] (K=H) ; W1[k][u] = Wg_h[u0+u][k] (K=2H)
//                     rnn   W0[k=v][u] = Wh[u0+u][v]
// Wh = rows [I, I+H) of the layer's TF Linear/Matrix.
// ------------------------------------------------------------------------------------------
__global__ void pack_weights_kernel(int cell, int dir, const float *__restrict__ m0, const float *__restrict__ m1,
                                    int I, int H, int HS, int CS, int k0, int ncp0, int k1, int ncp1,
                                    float *__restrict__ dst) {
  const int64_t per_rank = (int64_t)k0 * ncp0 + (int64_t)k1 * ncp1;
  const int64_t total = per_rank * CS;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int rank = (int)(i / per_rank);
    int64_t o = i % per_rank;
    int which = o >= (int64_t)k0 * ncp0;
    if (which) o -= (int64_t)k0 * ncp0;
    int ncp = which ? ncp1 : ncp0;
    int k = (int)(o / ncp), c = (int)(o % ncp);
    int u0 = rank * HS;
    float v = 0.f;
    if (dir == 0) {
      if (cell == 0) {
        int u = c >> 2, g = c & 3;
        if (u < HS) v = m0[(size_t)(I + k) * (4 * H) + g * H + u0 + u];
      } else if (cell == 1) {
        if (!which) {
          int u = c >> 1, g = c & 1;
          if (u < HS) v = m0[(size_t)(I + k) * (2 * H) + g * H + u0 + u];
        } else if (c < HS) {
          v = m1[(size_t)(I + k) * H + u0 + c];
        }
      } else if (c < HS) {
        v = m0[(size_t)(I + k) * H + u0 + c];
      }
    } else if (c < HS) {
      if (cell == 0) v = m0[(size_t)(I + u0 + c) * (4 * H) + k];
      else if (cell == 1) v = which ? m0[(size_t)(I + u0 + c) * (2 * H) + k] : m1[(size_t)(I + u0 + c) * H + k];
      else v = m0[(size_t)(I + u0 + c) * H + k];
    }
    dst[i] = v;
  }
}

struct Plan {
  int CS, HS, RB;
  int cpt0, cpt1, ncp0, ncp1, k0, k1;
  bool wsmem;
  size_t smem;
};

int make_plan(skrnn_handle *h, int B, int dir, Plan *p) {
  const int H = h->H, cell = h->cfg.cell;
  SK_CHECK(H % 8 == 0, "rnn_size %d must be a multiple of 8", H);
  p->CS = h->cluster_size;
  p->HS = H / p->CS;
  SK_CHECK(p->HS <= 64, "rnn_size %d too large for the fp32 recurrent kernels (max %d)", H, 64 * p->CS);
  p->RB = (B <= 8 * (h->num_sms / p->CS)) ? 8 : 16;
  if (p->HS > 32) p->RB = 8;
  int big = p->HS > 32 ? 2 : 1;
  p->k1 = 0; p->ncp1 = 0; p->cpt1 = 1;
  if (dir == 0) {
    p->k0 = H;
    if (cell == 0) p->cpt0 = 4 * big;
    else if (cell == 1) { p->cpt0 = 2 * big; p->cpt1 = big; p->k1 = H; }
    else p->cpt0 = big;
  } else {
    p->cpt0 = big;
    if (cell == 0) p->k0 = 4 * H;
    else if (cell == 1) { p->k0 = H; p->k1 = 2 * H; p->cpt1 = big; }
    else p->k0 = H;
  }
  p->ncp0 = 32 * p->cpt0;
  p->ncp1 = p->k1 ? 32 * p->cpt1 : 0;
  int kmax = max(p->k0, p->k1);
  size_t wfloats = (size_t)p->k0 * p->ncp0 + (size_t)p->k1 * p->ncp1;
  size_t other = (size_t)(kmax + kmax / 4) * p->RB + (size_t)NWARP * p->RB * max(p->ncp0, p->ncp1) +
                 2 * (size_t)p->RB * p->HS;
  p->wsmem = (wfloats + other) * 4 <= 227 * 1024;
  p->smem = ((p->wsmem ? wfloats : 0) + other) * 4;
  SK_CHECK(p->smem <= 227 * 1024, "recurrent kernel needs %zu B of shared memory", p->smem);
  SK_CHECK((int64_t)wfloats * p->CS <= h->ws.wpack_floats, "wpack workspace too small");
  return 0;
}

template <int CELL, int RB, int C0, int C1, bool WS>
int launch_one(bool bwd, const Plan &p, const RnnArgs &a, int nclusters, cudaStream_t st) {
  auto kern = bwd ? rnn_bwd_kernel<CELL, RB, C0, C1, WS> : rnn_fwd_kernel<CELL, RB, C0, C1, WS>;
  SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * p.CS);
  cfg.blockDim = dim3(NTHREADS);
  cfg.dynamicSmemBytes = p.smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = p.CS;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
  return 0;
}

template <int CELL, int RB, int C0, int C1>
int launch_ws(bool bwd, const Plan &p, const RnnArgs &a, int nc, cudaStream_t st) {
  return p.wsmem ? launch_one<CELL, RB, C0, C1, true>(bwd, p, a, nc, st)
                 : launch_one<CELL, RB, C0, C1, false>(bwd, p, a, nc, st);
}

template <int CELL>
int launch_cell(bool bwd, const Plan &p, const RnnArgs &a, int nc, cudaStream_t st) {
  // (cpt0, cpt1) pairs that make_plan can produce for this cell and direction
  const int c0 = p.cpt0, c1 = p.cpt1;
#define SK_CASE(RBV, A0, A1) \
  if (p.RB == RBV && c0 == A0 && c1 == A1) return launch_ws<CELL, RBV, A0, A1>(bwd, p, a, nc, st);
  if (CELL == 0) {
    SK_CASE(8, 4, 1) SK_CASE(16, 4, 1) SK_CASE(8, 8, 1) SK_CASE(8, 1, 1) SK_CASE(16, 1, 1) SK_CASE(8, 2, 1)
  } else if (CELL == 1) {
    SK_CASE(8, 2, 1) SK_CASE(16, 2, 1) SK_CASE(8, 4, 2) SK_CASE(8, 1, 1) SK_CASE(16, 1, 1) SK_CASE(8, 2, 2)
  } else {
    SK_CASE(8, 1, 1) SK_CASE(16, 1, 1) SK_CASE(8, 2, 1)
  }
#undef SK_CASE
  sk_set_error("no recurrent kernel instance for cell %d RB %d cpt (%d,%d)", CELL, p.RB, c0, c1);
  return 1;
}

int run_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *x_bt5,
              const float *state_in, float *state_out, const float *mask, uint64_t seed, int row0,
              bool training) {
  // The kernel families save a layer's activations in different places, so a layer's BPTT runs in the
  // family that ran its forward pass.
  if (!bwd) {
    h->fwd_family[l] = 0;
    int rc = sk_rnn_tc_layer(h, st, false, l, B, T, x_bt5, state_in, state_out, mask, seed, row0, training);
    if (rc == 0) h->fwd_family[l] = 1;
    if (rc >= 0) return rc;  // handled by the tcgen05 kernels (0) or failed hard (>0)
    rc = sk_rnn1_tc_layer(h, st, false, l, B, T, state_in, state_out, mask, seed, row0, training);
    if (rc == 0) h->fwd_family[l] = 3;
    if (rc >= 0) return rc;
  } else if (h->fwd_family[l] == 3) {
    int rc = sk_rnn1_tc_layer(h, st, true, l, B, T, state_in, state_out, mask, seed, row0, training);
    SK_CHECK(rc == 0, "layer %d: forward ran in the rnn1_tc kernels but their backward failed (rc %d): %s", l, rc,
             skrnn_last_error());
    return 0;
  } else if (h->fwd_family[l] == 1) {
    int rc = sk_rnn_tc_layer(h, st, true, l, B, T, x_bt5, state_in, state_out, mask, seed, row0, training);
    SK_CHECK(rc == 0, "layer %d: forward ran in the rnn_tc kernels but their backward failed (rc %d): %s", l, rc,
             skrnn_last_error());
    return 0;
  }
  Plan p;
  SK_TRY(make_plan(h, B, bwd ? 1 : 0, &p));
  SK_TRY(sk_ensure_f32_acts(h));
  const int cell = h->cfg.cell, H = h->H;
  const SkLayerParams &lp = h->lp[l];
  {
    int64_t total = ((int64_t)p.k0 * p.ncp0 + (int64_t)p.k1 * p.ncp1) * p.CS;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)4 * h->num_sms);
    SkLaunch scope(h, st, SK_CAT_MISC);
    pack_weights_kernel<<<blocks, 256, 0, st>>>(cell, bwd ? 1 : 0, h->params + lp.w[0],
                                               cell == 1 ? h->params + lp.w[1] : nullptr, lp.I, H, p.HS,
                                               p.CS, p.k0, p.ncp0, p.k1, p.ncp1, h->ws.wpack);
    SK_CUDA(cudaGetLastError());
  }
  RnnArgs a = {};
  a.B = B; a.T = T; a.H = H; a.HS = p.HS; a.G = h->G;
  a.row0 = row0;
  a.dropout = (training && l == h->L - 1 && h->cfg.keep_prob < 1.0f) ? 1 : 0;
  a.keep = h->cfg.keep_prob;
  a.seed = seed;
  a.x_bt5 = x_bt5;
  a.xw = h->ws.xw[l];
  a.hprev = h->ws.hprev[l];
  a.cprev = h->ws.cprev[l];
  a.out = h->ws.out[l];
  a.dout = h->ws.dout;
  a.state_in = state_in;
  a.state_out = state_out;
  a.mask = mask;
  a.wpack = h->ws.wpack;
  a.out_bf = h->ws.out_bf[l];
  a.hprev_bf = h->ws.hprev_bf[l];
  a.aux_bf = h->ws.aux_bf[l];
  a.dpre_bf = bwd ? h->ws.dpre_bf : nullptr;
  a.w_in_smem = p.wsmem;
  a.ncp0 = p.ncp0; a.ncp1 = p.ncp1; a.k0 = p.k0; a.k1 = p.k1;
  int nclusters = (B + p.RB - 1) / p.RB;
  int rc;
  SkLaunch scope(h, st, bwd ? SK_CAT_RNN_BWD : SK_CAT_RNN_FWD);
  if (cell == 0) rc = launch_cell<0>(bwd, p, a, nclusters, st);
  else if (cell == 1) rc = launch_cell<1>(bwd, p, a, nclusters, st);
  else rc = launch_cell<2>(bwd, p, a, nclusters, st);
  return rc;
}

}  // namespace

int sk_rnn_forward_layer(skrnn_handle *h, cudaStream_t st, int l, int B, int T, const float *x_bt5,
                         const float *state_in, float *state_out, const float *mask, uint64_t seed,
                         int row0, bool training) {
  return run_layer(h, st, false, l, B, T, x_bt5, state_in, state_out, mask, seed, row0, training);
}

int sk_rnn_backward_layer(skrnn_handle *h, cudaStream_t st, int l, int B, int T, const float *x_bt5,
                          const float *mask, uint64_t seed, int row0, bool training) {
  return run_layer(h, st, true, l, B, T, x_bt5, nullptr, nullptr, mask, seed, row0, training);
}
