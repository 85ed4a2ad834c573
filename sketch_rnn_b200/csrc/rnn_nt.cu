// rnn_nt.cu -- tcgen05 persistent LSTM recurrent kernels, large-batch variant (bf16 tensor-core path).
// Same job as rnn_tc.cu (reference model.py:66-95 unroll around TF BasicLSTMCell, and its BPTT,
// model.py:182) but with the batch on the MMA M dimension:
//
//   forward : D[128 rows, 128 gate cols] = h_{t-1}[128, K=H]      * Wslice^T      (B operand resident)
//   backward: D[128 rows,  32 units   ] = dpre_{t+1}[128, K=4H]  * Wh[units,:]^T (B operand resident)
//
// A cluster of CS = H/32 CTAs owns 128 batch rows; rank r owns hidden units [32r, 32r+32).  The A operand
// (h, resp. d(pre-activation), of ALL units / gate columns, produced by all ranks in the previous step) is
// all-gathered with TMA multicast: the tile is cut into half k-blocks of 64 rows x 64 k (8 KB); rank
// (2j + half) mod CS fetches half k-block (j, half) from L2 once and the hardware delivers it to the same
// ring slot of every CTA of the cluster.  A 4-slot ring of 16 KB k-blocks is recycled inside a step; a slot
// is released cluster-wide by tcgen05.commit multicast to the `free` barrier of every CTA.
//
// Epilogue: 16 warps; warp (q, cg) reads accumulator rows 32q..32q+31 (= batch rows) and the columns of units
// 8cg..8cg+7, so a thread owns 8 consecutive units of one batch row: all four gates of a unit sit in
// consecutive registers (no exchange) and every global access is a 16- or 32-byte vector.
#include <stdlib.h>

#include "common.cuh"
#include "skrnn_math.h"
#include "tc_ptx.cuh"

namespace {

constexpr int NT_EPI_WARPS = 16;
constexpr int NT_THREADS = 64 + 32 * NT_EPI_WARPS;
constexpr int NT_RING = 4;           // ring slots (k-blocks of 128 rows x 64 k = 16 KB)
constexpr int NT_KB_BYTES = 16384;

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_id_x() {
  unsigned r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
// tcgen05.commit that arrives on the mbarrier at the same offset in every CTA of the mask
__device__ __forceinline__ void tc_commit_multicast(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"(mask) : "memory");
}

struct NtArgs {
  int B, T, H, G, CS;
  int nA;        // A k-blocks per step: forward H/64 (>= 1), backward 4H/64
  int row0, dropout;
  float keep;
  uint64_t seed;
  const unsigned char *eoc;  // [T,B]
  const float *state_in;
  float *state_out;
  const float *mask;
  float *acts;               // [T,B,G] fp32: x*Wx+b in, gate activations out (forward); activations in (backward)
  float *cprev;              // [T,B,H]
  __nv_bfloat16 *out_bf, *hprev_bf;
  const float *dout;         // [T,B,H]
  __nv_bfloat16 *dpre_bf;    // [T,B,G]
  const uint8_t *wimg;
  int wimg_bytes;
  int debug;
  int mcast;   // 1: all-gather by TMA multicast; 0: every CTA fetches the whole tile itself (diagnostic)
};

// 8 dropout scales (0 or 1/keep) for units [unit0, unit0+8) of (row, t): 16-bit uniforms from 4 hashes
__device__ __forceinline__ void drop8(const NtArgs &a, int b, int t, int unit0, float *sc) {
  if (a.mask != nullptr) {
    const float4 m0 = *reinterpret_cast<const float4 *>(a.mask + ((size_t)b * a.T + t) * a.H + unit0);
    const float4 m1 = *reinterpret_cast<const float4 *>(a.mask + ((size_t)b * a.T + t) * a.H + unit0 + 4);
    const float ik = 1.0f / a.keep;
    sc[0] = m0.x * ik; sc[1] = m0.y * ik; sc[2] = m0.z * ik; sc[3] = m0.w * ik;
    sc[4] = m1.x * ik; sc[5] = m1.y * ik; sc[6] = m1.z * ik; sc[7] = m1.w * ik;
    return;
  }
  const uint32_t thr = (uint32_t)(a.keep * 65536.0f);
  const float ik = 1.0f / a.keep;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t hsh = skm_drop_bits(a.seed, (uint32_t)(a.row0 + b), (uint32_t)t, (uint32_t)(unit0 + 2 * j));
    sc[2 * j] = (hsh & 0xffffu) < thr ? ik : 0.f;
    sc[2 * j + 1] = (hsh >> 16) < thr ? ik : 0.f;
  }
}

template <bool BWD>
__global__ void __launch_bounds__(NT_THREADS, 1)
lstm_nt_kernel(const __grid_constant__ CUtensorMap tm_a, NtArgs a) {
  constexpr int NB = BWD ? 32 : 128;           // B-operand rows = MMA N
  constexpr int WKB = NB * 128;                // bytes of one resident weight k-block
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const int H = a.H, G = a.G, B = a.B, T = a.T, CS = a.CS, nA = a.nA;
  const uint32_t s_ring = base;  // multicast destination first (low shared-memory offsets)
  const uint32_t s_w = s_ring + NT_RING * NT_KB_BYTES;
  const uint32_t s_bar = s_w + (uint32_t)nA * WKB;
  const uint32_t bar_full = s_bar, bar_free = s_bar + 32, bar_acc = s_bar + 64, tmem_holder = s_bar + 72;
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_holder - base));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = cluster_ctarank();
  const int b0 = cluster_id_x() * 128;
  const int u0 = rank * 32;
  const int S = 2 * H;
  const uint16_t all_mask = (uint16_t)((1u << CS) - 1u);
  const int U = nA >= NT_RING ? nA / NT_RING : 1;  // uses of a ring slot per step

  {  // resident weight operand
    const uint4 *src = reinterpret_cast<const uint4 *>(a.wimg + (size_t)rank * a.wimg_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(gbase + (s_w - base));
    for (int i = tid; i < a.wimg_bytes / 16; i += NT_THREADS) dst[i] = __ldg(src + i);
  }
  if (tid == 0) {
    for (int s = 0; s < NT_RING; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_free + 8 * s, a.mcast ? CS : 1);
    }
    mbar_init(bar_acc, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, BWD ? 32u : 128u);

  // epilogue thread: batch row = 32q + lane, units [8cg, 8cg+8) of this rank
  const int q = warp & 3, cg = (warp - 2) >> 2;
  const int row = 32 * q + lane;
  const int b = b0 + row;
  const bool valid = warp >= 2 && b < B;
  const int un = u0 + 8 * cg;  // first hidden unit of my 8
  float cst[8];                // forward: c state; backward: carried d c
#pragma unroll
  for (int j = 0; j < 8; ++j) cst[j] = 0.f;
  if (!BWD && valid) {
    const float4 c0 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + un);
    const float4 c1 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + un + 4);
    const float4 h0 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + H + un);
    const float4 h1 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + H + un + 4);
    cst[0] = c0.x; cst[1] = c0.y; cst[2] = c0.z; cst[3] = c0.w; cst[4] = c1.x; cst[5] = c1.y; cst[6] = c1.z; cst[7] = c1.w;
    *reinterpret_cast<float4 *>(a.cprev + (size_t)b * H + un) = c0;
    *reinterpret_cast<float4 *>(a.cprev + (size_t)b * H + un + 4) = c1;
    __nv_bfloat162 p[4] = {__floats2bfloat162_rn(h0.x, h0.y), __floats2bfloat162_rn(h0.z, h0.w),
                           __floats2bfloat162_rn(h1.x, h1.y), __floats2bfloat162_rn(h1.z, h1.w)};
    *reinterpret_cast<uint4 *>(a.hprev_bf + (size_t)b * H + un) = *reinterpret_cast<uint4 *>(p);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_holder_ptr;
  if (warp == 1 && elect_one()) {  // arm the first use of every ring slot that will be used
    for (int s = 0; s < NT_RING && s < nA; ++s) mbar_expect_tx(bar_full + 8 * s, NT_KB_BYTES);
  }
  __syncwarp();
  cluster_sync_all();

  const uint32_t idesc = make_idesc_bf16(128, NB, false, false);
  const int nsteps = BWD ? T - 1 : T;  // steps that run the product
  long long tk[6] = {0, 0, 0, 0, 0, 0};
  const bool dbg = a.debug && rank == 0 && cluster_id_x() == 0 && (tid == 64 || tid == 32);

  for (int it = 0; it < T; ++it) {
    const int t = BWD ? T - 1 - it : it;
    // product step index: forward every step; backward steps it >= 1 multiply d(pre)_{t+1}
    const bool do_mm = BWD ? (it >= 1) : true;
    const int gs = BWD ? it - 1 : it;          // running index of product steps
    const int t_src = BWD ? t + 1 : t;         // time slice of the A operand
    long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    if (dbg) c0 = clock64();
    if (do_mm && warp == 0) {
      if (elect_one()) {
        fence_proxy_async();
        const bool mc = a.mcast && CS > 1;
        for (int hb = mc ? rank : 0; hb < 2 * nA; hb += mc ? CS : 1) {
          const int j = hb >> 1, half = hb & 1, s = j % NT_RING;
          const int m = gs * U + j / NT_RING;  // global use index of slot s
          if (m >= 1) mbar_wait_wd(bar_free + 8 * s, (m - 1) & 1, (BWD ? 50000 : 0) + 1000 * gs + 100 + s, 1u << 27);
          const uint32_t dst = s_ring + (uint32_t)s * NT_KB_BYTES + (uint32_t)half * (NT_KB_BYTES / 2);
          if (mc) tma_load_3d_multicast(dst, &tm_a, bar_full + 8 * s, j * 64, b0 + half * 64, t_src, all_mask);
          else tma_load_3d(dst, &tm_a, bar_full + 8 * s, j * 64, b0 + half * 64, t_src);
        }
      }
      __syncwarp();
    } else if (do_mm && warp == 1) {
      if (elect_one()) {
        for (int j = 0; j < nA; ++j) {
          const int s = j % NT_RING;
          const int m = gs * U + j / NT_RING;
          mbar_wait_wd(bar_full + 8 * s, m & 1, (BWD ? 50000 : 0) + 1000 * gs + 200 + 10 * (j / NT_RING) + s, 1u << 23);
          tc_fence_after();
          const uint64_t da = make_desc(s_ring + (uint32_t)s * NT_KB_BYTES, 0, 1024);
          const uint64_t db = make_desc(s_w + (uint32_t)j * WKB, 0, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(tmem_acc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (j > 0 || k > 0) ? 1u : 0u);
          // release the slot in every CTA of the cluster once these MMAs have read it, then re-arm it
          if (a.mcast && CS > 1) tc_commit_multicast(bar_free + 8 * s, all_mask);
          else tc_commit(bar_free + 8 * s);
          mbar_expect_tx(bar_full + 8 * s, NT_KB_BYTES);
        }
        tc_commit(bar_acc);
        if (dbg) tk[5] += clock64() - c0;
      }
      __syncwarp();
    } else if (warp >= 2) {
      if (!BWD) {
        // ---------------- forward epilogue ----------------
        float xw[32];  // x*Wx + b of my 8 units x 4 gates (gate-major in memory: g*H + unit)
        float dsc[8];
        bool eoc = false;
        if (valid) {
          const float *xp = a.acts + ((size_t)t * B + b) * G + un;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const float4 v0 = *reinterpret_cast<const float4 *>(xp + (size_t)g * H);
            const float4 v1 = *reinterpret_cast<const float4 *>(xp + (size_t)g * H + 4);
            xw[8 * g + 0] = v0.x; xw[8 * g + 1] = v0.y; xw[8 * g + 2] = v0.z; xw[8 * g + 3] = v0.w;
            xw[8 * g + 4] = v1.x; xw[8 * g + 5] = v1.y; xw[8 * g + 6] = v1.z; xw[8 * g + 7] = v1.w;
          }
          eoc = __ldg(a.eoc + (size_t)t * B + b) != 0;
          if (a.dropout) drop8(a, b, t, un, dsc);
        }
        if (dbg) c1 = clock64();
        mbar_wait_wd(bar_acc, gs & 1, 300, 1u << 27);
        if (dbg) c2 = clock64();
        tc_fence_after();
        uint32_t r[32];  // accumulator columns 32cg + 4u + g
        tc_ldn<32>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(32 * cg), r);
        if (valid) {
          float hn[8], ov[8];
          float *ga = xw;  // the pre-activation inputs are dead once a unit's gates are computed
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const float ig = sigmoid_approx(__uint_as_float(r[4 * u + 0]) + xw[u]);
            const float jg = tanh_approx(__uint_as_float(r[4 * u + 1]) + xw[8 + u]);
            const float fg = sigmoid_approx(__uint_as_float(r[4 * u + 2]) + xw[16 + u] + 1.0f);
            const float og = sigmoid_approx(__uint_as_float(r[4 * u + 3]) + xw[24 + u]);
            const float c2 = fmaf(cst[u], fg, ig * jg);
            const float h2 = tanh_approx(c2) * og;
            ga[u] = ig; ga[8 + u] = jg; ga[16 + u] = fg; ga[24 + u] = og;
            ov[u] = a.dropout ? h2 * dsc[u] : h2;
            cst[u] = c2;
            hn[u] = h2;
          }
          if (eoc) {  // reset to the fed state (model.py:82-92)
            const float4 c0 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + un);
            const float4 c1 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + un + 4);
            const float4 h0 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + H + un);
            const float4 h1 = *reinterpret_cast<const float4 *>(a.state_in + (size_t)b * S + H + un + 4);
            cst[0] = c0.x; cst[1] = c0.y; cst[2] = c0.z; cst[3] = c0.w; cst[4] = c1.x; cst[5] = c1.y; cst[6] = c1.z; cst[7] = c1.w;
            hn[0] = h0.x; hn[1] = h0.y; hn[2] = h0.z; hn[3] = h0.w; hn[4] = h1.x; hn[5] = h1.y; hn[6] = h1.z; hn[7] = h1.w;
          }
          const size_t ix = ((size_t)t * B + b) * H + un;
          __nv_bfloat162 p[4];
          if (t + 1 < T) {  // what the other ranks wait for goes out first
#pragma unroll
            for (int j = 0; j < 4; ++j) p[j] = __floats2bfloat162_rn(hn[2 * j], hn[2 * j + 1]);
            *reinterpret_cast<uint4 *>(a.hprev_bf + ix + (size_t)B * H) = *reinterpret_cast<uint4 *>(p);
            *reinterpret_cast<float4 *>(a.cprev + ix + (size_t)B * H) = make_float4(cst[0], cst[1], cst[2], cst[3]);
            *reinterpret_cast<float4 *>(a.cprev + ix + (size_t)B * H + 4) = make_float4(cst[4], cst[5], cst[6], cst[7]);
          } else {
            float *so = a.state_out + (size_t)b * S + un;
            *reinterpret_cast<float4 *>(so) = make_float4(cst[0], cst[1], cst[2], cst[3]);
            *reinterpret_cast<float4 *>(so + 4) = make_float4(cst[4], cst[5], cst[6], cst[7]);
            *reinterpret_cast<float4 *>(so + H) = make_float4(hn[0], hn[1], hn[2], hn[3]);
            *reinterpret_cast<float4 *>(so + H + 4) = make_float4(hn[4], hn[5], hn[6], hn[7]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) p[j] = __floats2bfloat162_rn(ov[2 * j], ov[2 * j + 1]);
          *reinterpret_cast<uint4 *>(a.out_bf + ix) = *reinterpret_cast<uint4 *>(p);
          float *ap = a.acts + ((size_t)t * B + b) * G + un;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            *reinterpret_cast<float4 *>(ap + (size_t)g * H) = make_float4(ga[8 * g], ga[8 * g + 1], ga[8 * g + 2], ga[8 * g + 3]);
            *reinterpret_cast<float4 *>(ap + (size_t)g * H + 4) = make_float4(ga[8 * g + 4], ga[8 * g + 5], ga[8 * g + 6], ga[8 * g + 7]);
          }
        }
      } else {
        // ---------------- backward epilogue ----------------
        float ga[32], cp[8], dd[8];
        bool eoc = false;
        if (valid) {
          const float *ap = a.acts + ((size_t)t * B + b) * G + un;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const float4 v0 = *reinterpret_cast<const float4 *>(ap + (size_t)g * H);
            const float4 v1 = *reinterpret_cast<const float4 *>(ap + (size_t)g * H + 4);
            ga[8 * g + 0] = v0.x; ga[8 * g + 1] = v0.y; ga[8 * g + 2] = v0.z; ga[8 * g + 3] = v0.w;
            ga[8 * g + 4] = v1.x; ga[8 * g + 5] = v1.y; ga[8 * g + 6] = v1.z; ga[8 * g + 7] = v1.w;
          }
          const size_t ix = ((size_t)t * B + b) * H + un;
          const float4 c0 = *reinterpret_cast<const float4 *>(a.cprev + ix), c1 = *reinterpret_cast<const float4 *>(a.cprev + ix + 4);
          const float4 d0 = *reinterpret_cast<const float4 *>(a.dout + ix), d1 = *reinterpret_cast<const float4 *>(a.dout + ix + 4);
          cp[0] = c0.x; cp[1] = c0.y; cp[2] = c0.z; cp[3] = c0.w; cp[4] = c1.x; cp[5] = c1.y; cp[6] = c1.z; cp[7] = c1.w;
          dd[0] = d0.x; dd[1] = d0.y; dd[2] = d0.z; dd[3] = d0.w; dd[4] = d1.x; dd[5] = d1.y; dd[6] = d1.z; dd[7] = d1.w;
          eoc = __ldg(a.eoc + (size_t)t * B + b) != 0;
          if (a.dropout) {
            float dsc[8];
            drop8(a, b, t, un, dsc);
#pragma unroll
            for (int u = 0; u < 8; ++u) dd[u] *= dsc[u];
          }
        }
        uint32_t r[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) r[u] = 0u;
        if (do_mm) {
          mbar_wait_wd(bar_acc, gs & 1, 301);
          tc_fence_after();
          tc_ldn<8>(tmem_acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(8 * cg), r);
        }
        if (valid) {
          float dp[32];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const float ig = ga[u], jg = ga[8 + u], fg = ga[16 + u], og = ga[24 + u];
            const float dh2 = dd[u] + (eoc ? 0.f : __uint_as_float(r[u]));
            const float tc = tanh_approx(fmaf(cp[u], fg, ig * jg));
            const float dc2 = (eoc ? 0.f : cst[u]) + dh2 * og * (1.0f - tc * tc);
            cst[u] = dc2 * fg;
            dp[u] = dc2 * jg * ig * (1.0f - ig);
            dp[8 + u] = dc2 * ig * (1.0f - jg * jg);
            dp[16 + u] = dc2 * cp[u] * fg * (1.0f - fg);
            dp[24 + u] = dh2 * tc * og * (1.0f - og);
          }
          __nv_bfloat16 *dpp = a.dpre_bf + ((size_t)t * B + b) * G + un;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            __nv_bfloat162 p[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) p[j] = __floats2bfloat162_rn(dp[8 * g + 2 * j], dp[8 * g + 2 * j + 1]);
            *reinterpret_cast<uint4 *>(dpp + (size_t)g * H) = *reinterpret_cast<uint4 *>(p);
          }
        }
      }
      if (dbg) c3 = clock64();
      fence_proxy_async();  // my global stores -> visible to the other ranks' TMA loads of the next step
      if (dbg && tid == 64) { long long c4 = clock64(); tk[0] += c1 - c0; tk[1] += c2 - c1; tk[2] += c3 - c2; tk[3] += c4 - c3; c0 = c4; }
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    if (dbg && tid == 64) tk[4] += clock64() - c0;
  }
  if (dbg && tid == 64)
    printf("[nt %s] epilogue cycles/step: prefetch %lld, wait acc %lld, ld+math+stores %lld, proxy fence %lld, cluster barrier %lld\n",
           BWD ? "bwd" : "fwd", tk[0] / T, tk[1] / T, tk[2] / T, tk[3] / T, tk[4] / T);
  if (dbg && tid == 32) printf("[nt %s] MMA thread: step start -> all MMAs issued + committed %lld cycles/step\n", BWD ? "bwd" : "fwd", tk[5] / T);
  (void)nsteps;
  __syncthreads();
  if (warp == 1) tc_dealloc(tmem_acc, BWD ? 32u : 128u);
}

// Resident B-operand images, 128B-swizzled K-major, one per rank.
//   forward : nA k-blocks x [128 rows n = 4u+g][64 k]      value Wh[k][g*H + u0+u]
//   backward: nA k-blocks x [ 32 rows u      ][64 k]       value Wh[u0+u][k]          (k = gate column)
// Wh = rows [I, I+H) of the layer's TF Linear/Matrix.
__global__ void pack_nt_image_kernel(int bwd, const float *__restrict__ mat, int I, int H, int CS, int nA,
                                     uint8_t *__restrict__ dst) {
  const int NB = bwd ? 32 : 128;
  const int G = 4 * H;
  const int K = bwd ? G : H;
  const int64_t per_rank = (int64_t)nA * NB * 64;
  const int64_t total = per_rank * CS;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int rank = (int)(i / per_rank);
    const int64_t o = i % per_rank;
    const int kb = (int)(o / (NB * 64));
    const int w = (int)(o % (NB * 64));
    const int m = w / 64, within = w % 64;
    const int kloc = (((within >> 3) ^ (m & 7)) << 3) + (within & 7);
    const int k = kb * 64 + kloc;
    const int u0 = rank * 32;
    float v = 0.f;
    if (k < K) {
      if (!bwd) {
        const int u = m >> 2, g = m & 3;
        v = mat[(size_t)(I + k) * G + g * H + u0 + u];
      } else {
        v = mat[(size_t)(I + u0 + m) * G + k];
      }
    }
    reinterpret_cast<__nv_bfloat16 *>(dst)[i] = __float2bfloat16_rn(v);
  }
}

template <bool BWD>
int launch_nt(const CUtensorMap &tm, const NtArgs &a, int nclusters, size_t smem, cudaStream_t st) {
  auto kern = lstm_nt_kernel<BWD>;
  SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (a.CS > 8) SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * a.CS);
  cfg.blockDim = dim3(NT_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = a.CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tm, a));
  return 0;
}

}  // namespace

// Returns 0 when the layer was run here, -1 when this variant does not apply (caller falls back).
int sk_rnn_nt_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *state_in,
                    float *state_out, const float *mask, uint64_t seed, int row0, bool training) {
  const int H = h->H;
  if (h->cfg.precision != SKRNN_PREC_BF16 || h->cfg.cell != SKRNN_CELL_LSTM) return -1;
  if (!(H == 32 || H == 64 || H == 128 || H == 256 || H == 512)) return -1;
  if (h->nt_rnn_disabled) return -1;
  // Opt-in (SKRNN_NT=1): measured on B200 this variant is not faster than the transposed kernels of
  // rnn_tc.cu yet (profiles/), and its 16-CTA backward all-gather still loses multicast boxes.
  static int env_mode = -2;
  if (env_mode == -2) { const char *e = getenv("SKRNN_NT"); env_mode = e ? atoi(e) : 0; }
  if (env_mode != 1) return -1;
  const int CS = H / 32;
  {
    // the 16-CTA backward all-gather is not reliable yet (multicast boxes go missing): keep the
    // partial-exchange kernel of rnn_tc.cu for H = 512 BPTT unless explicitly enabled
    static int nb_env = -1;
    if (nb_env < 0) { const char *e = getenv("SKRNN_NT_BWD16"); nb_env = (e && e[0] == '1') ? 1 : 0; }
    // (declined for the forward too: a layer's BPTT has to run in the family that saved its activations)
    if (CS > 8 && !nb_env) return -1;
  }
  const int K = bwd ? 4 * H : H;
  const int nA = (K + 63) / 64;
  const int NB = bwd ? 32 : 128;
  const int img_bytes = nA * NB * 128;
  const size_t smem = (size_t)img_bytes + NT_RING * NT_KB_BYTES + 1024 + 256;
  if (smem > 227 * 1024) return -1;
  if (!bwd) {  // same check for the backward configuration of this layer
    const size_t smem_b = (size_t)((4 * H + 63) / 64) * 32 * 128 + NT_RING * NT_KB_BYTES + 1024 + 256;
    if (smem_b > 227 * 1024) return -1;
  }
  SK_CHECK((int64_t)img_bytes * CS <= h->ws.wpack_floats * 4, "weight image workspace too small");
  {
    SkLaunch scope(h, st, SK_CAT_MISC);
    int64_t total = (int64_t)img_bytes / 2 * CS;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)4 * h->num_sms);
    pack_nt_image_kernel<<<blocks, 256, 0, st>>>(bwd ? 1 : 0, h->params + h->lp[l].w[0], h->lp[l].I, H, CS, nA,
                                                reinterpret_cast<uint8_t *>(h->ws.wpack));
    SK_CUDA(cudaGetLastError());
  }
  NtArgs a = {};
  a.B = B; a.T = T; a.H = H; a.G = h->G; a.CS = CS; a.nA = nA;
  a.row0 = row0;
  a.dropout = (training && l == h->L - 1 && h->cfg.keep_prob < 1.0f) ? 1 : 0;
  a.keep = h->cfg.keep_prob;
  a.seed = seed;
  a.eoc = h->ws.eoc; a.state_in = state_in; a.state_out = state_out; a.mask = mask;
  a.acts = h->ws.xw[l]; a.cprev = h->ws.cprev[l]; a.out_bf = h->ws.out_bf[l]; a.hprev_bf = h->ws.hprev_bf[l];
  a.dout = h->ws.dout; a.dpre_bf = h->ws.dpre_bf;
  a.wimg = reinterpret_cast<const uint8_t *>(h->ws.wpack);
  a.wimg_bytes = img_bytes;
  { const char *e = getenv("SKRNN_TC_DEBUG"); a.debug = (e && e[0] == '1') ? 1 : 0; }
  {
    static int mc_env = -1;
    if (mc_env < 0) { const char *e = getenv("SKRNN_NT_MCAST"); mc_env = (e && e[0] == '0') ? 0 : 1; }
    a.mcast = mc_env;
  }
  // A operand: hprev_bf [T,B,H] (forward) or dpre_bf [T,B,G] (backward), boxes of 64 rows x 64 k
  CUtensorMap tm;
  {
    const void *ptr = bwd ? (const void *)h->ws.dpre_bf : (const void *)h->ws.hprev_bf[l];
    uint64_t dims[3] = {(uint64_t)K, (uint64_t)B, (uint64_t)T};
    uint64_t str[2] = {(uint64_t)K * 2, (uint64_t)B * K * 2};
    uint32_t box[3] = {64, 64, 1};
    SK_TRY(sk_make_tmap(&tm, 2, ptr, 3, dims, str, box));
  }
  const int nclusters = (B + 127) / 128;
  int rc;
  {
    SkLaunch scope(h, st, bwd ? SK_CAT_RNN_BWD : SK_CAT_RNN_FWD);
    rc = bwd ? launch_nt<true>(tm, a, nclusters, smem, st) : launch_nt<false>(tm, a, nclusters, smem, st);
  }
  if (rc != 0) {
    fprintf(stderr, "libskrnn: large-batch tcgen05 recurrent kernel launch failed (%s); falling back\n", skrnn_last_error());
    cudaGetLastError();
    h->nt_rnn_disabled = true;
    return -1;
  }
  return 0;
}
