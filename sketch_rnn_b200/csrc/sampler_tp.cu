// sampler_tp.cu -- the throughput sampler as ONE persistent kernel (reference Model.sample, model.py:187-264, many
// sketches in flight; BASELINE config 5: lstm, rnn_size 256, 2 layers).
//
// sampler_tc.cu runs a sampling step as four launches (two gate GEMMs with the cell update as their epilogue, the
// output projection, the draw), 300 x 4 graph-replayed kernels whose ramp-up and drain cost more than their math.
// Sketches never interact, so here a CTA owns 64 sketches for ALL steps and nothing but the weights crosses the SM
// boundary inside the loop:
//   * the operands h0 / h1 of the CTA's 64 sketches stay in shared memory in the tensor core's K-major 128-byte-swizzled
//     layout, double-buffered by step parity (the cell update writes step s+1's operand while step s's products may
//     still read the old one); the cell state c stays in global memory, one row-contiguous slice per thread;
//   * the weights (bf16, gate-interleaved columns, 1.6 MB for the stack) stream from L2 through a five-stage TMA ring
//     (16 KB stages), once per step and CTA: they are the B operand, M = 64 sketches x N = 256 gate columns per accumulator;
//   * warp 0 streams, warp 1 issues tcgen05.mma, eight epilogue warps run the LSTM cell update straight from tensor
//     memory (both groups of four split every chunk: a step is a chain of dependent chunks, so a chunk's latency counts),
//     and after the output projection the per-sketch draw (softmax, temperature, categorical draws, bivariate normal)
//     whose point is the next step's input;
//   * tensor memory holds four M = 64 accumulators of 256 columns (such an accumulator fills lanes 0..15 of every lane
//     quarter and may equally start at lane 16: tools/probe_m64.cu): three rotate over the gate chunks, the fourth belongs
//     to the output projection, so the draw blocks no products and layer 0 opens three chunks before it needs the point;
//   * the input point and the bias ride in one extra K = 16 product per accumulator: operand columns (dx, dy, eos, eoc,
//     cont, 1, 1) against the rows (W_dx, W_dy, W_eos, W_eoc, W_cont, bias rounded to bf16, the rounding's remainder),
//     so that the epilogue reads nothing but tensor memory and the cell state;
//   * an eoc input holds the recurrent state (model.py:82-92 with initial_state = prev_state) while the layer output
//     still feeds the layer above: the new h is always written, and once the products that read it are complete the
//     held rows are copied back from the other buffer.
//   * within a 256-column chunk the gate columns are permuted (tp_src_col) so that tcgen05.ld.16x256b hands each of a
//     warp's 32 threads the four gates of 8 consecutive units for two sketch rows: an M = 64 accumulator only fills 16 of a
//     warp's 32 lanes, and the cell update is bound by the transcendental unit, which a half-empty warp occupies in full.
// Same arithmetic contract as sampler_tc.cu (operands rounded to bf16, fp32 accumulation; not the bit-exact path).
#include <stdlib.h>

#include "common.cuh"
#include "skrnn_math.h"
#include "tc_ptx.cuh"

namespace {

constexpr int TP_M = 64;                 // sketches per CTA (accumulator rows)
constexpr int TP_H = 256, TP_G = 4 * TP_H;
constexpr int TP_BK = 32;                // K rows per ring stage
constexpr int TP_STAGES = 5;
constexpr int TP_CHUNK_BYTES = TP_BK * 128;      // one 64-column chunk of a stage: [32 k rows][128 B]
constexpr int TP_STAGE_BYTES = 4 * TP_CHUNK_BYTES;  // 16 KB
constexpr int TP_KBLK_BYTES = TP_M * 128;        // one [64 rows x 64 k] operand block
constexpr int TP_HBUF_BYTES = 4 * TP_KBLK_BYTES; // 32 KB: h of one layer, one step parity
constexpr int TP_THREADS = 64 + 256;             // warp 0: TMA, warp 1: MMA, warps 2-9: epilogue
constexpr int TP_CHUNKS = 9;                     // accumulators per step: 4 + 4 gate chunks, 1 output projection
constexpr int TP_XROWS = 32;                     // rows of the input-term weight image (7 used; one ring stage)
constexpr int TP_XO_PITCH = 192;                 // columns of the output projection's bias image
constexpr int TP_SMEM = 4 * TP_HBUF_BYTES + TP_STAGES * TP_STAGE_BYTES + TP_KBLK_BYTES + 2 * TP_M * 8 + 256 + 1024 /*align*/;

struct TpArgs {
  int n, steps, M, NOUT, nh;     // nh: columns of the output-projection product (multiple of 64 covering NOUT)
  float data_scale, temp_mix, temp_pen;
  int stop_if_eoc, sketch0;
  uint64_t seed;
  const float *uniforms;
  float *strokes, *mixp;
  int *lengths;
  float *cst0, *cst1;            // [n, 256] cell states, zero on entry
  int debug;                     // SKRNN_TP_DEBUG=1: CTA 0 prints where its MMA warp waited
};

// Column order of the streamed weight images.  Position P = 256 c + 16 j + 8 a + 2 m + b holds gate 2a + b of unit
// 64 c + 16 m + j: tcgen05.ld.16x256b gives thread t (m = t % 4) columns 2m, 2m+1 of every group of 8, so it ends up with
// the gates (i, j | f, o) of units 16m .. 16m+15 of the chunk.  Returns the column of sampler_tc.cu's images (4 unit + gate).
__host__ __device__ inline int tp_src_col(int P) {
  const int c = P >> 8, w = P & 255;
  const int j = w >> 4, a2 = (w >> 3) & 1, m = (w >> 1) & 3, b = w & 1;
  return 4 * (64 * c + 16 * m + j) + 2 * a2 + b;
}

__global__ void tp_wperm_kernel(const __nv_bfloat16 *__restrict__ src, __nv_bfloat16 *__restrict__ dst, int K) {
  const int total = K * TP_G;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i / TP_G, P = i % TP_G;
    dst[i] = src[k * TP_G + tp_src_col(P)];
  }
}

// Input-term weight images [TP_XROWS][1024] (same column order; rows 0..4 of sampler_tc.cu's wsm[0] are the x rows):
// rows 0..4 W_x (layer 0 only), row 5 the bias rounded to bf16, row 6 what the rounding dropped, the rest zero.
__global__ void tp_ximage_kernel(const __nv_bfloat16 *__restrict__ wsm0, const float *__restrict__ bias_il0,
                                 const float *__restrict__ bias_il1, __nv_bfloat16 *__restrict__ wx0,
                                 __nv_bfloat16 *__restrict__ wx1, const float *__restrict__ out_b, int NOUT,
                                 __nv_bfloat16 *__restrict__ wxo) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= TP_G) return;
  const __nv_bfloat16 zero = __float2bfloat16_rn(0.f);
  if (j < TP_XO_PITCH) {  // output projection: bias rows only, natural column order
    for (int k = 0; k < TP_XROWS; ++k) wxo[k * TP_XO_PITCH + j] = zero;
    if (j < NOUT) {
      const float b = out_b[j];
      const __nv_bfloat16 hb = __float2bfloat16_rn(b);
      wxo[5 * TP_XO_PITCH + j] = hb; wxo[6 * TP_XO_PITCH + j] = __float2bfloat16_rn(b - __bfloat162float(hb));
    }
  }
  const int sc = tp_src_col(j);
  for (int k = 0; k < TP_XROWS; ++k) { wx0[k * TP_G + j] = zero; wx1[k * TP_G + j] = zero; }
  for (int k = 0; k < 5; ++k) wx0[k * TP_G + j] = wsm0[k * TP_G + sc];
  const float b0 = bias_il0[sc], b1 = bias_il1[sc];
  const __nv_bfloat16 h0 = __float2bfloat16_rn(b0), h1 = __float2bfloat16_rn(b1);
  wx0[5 * TP_G + j] = h0; wx0[6 * TP_G + j] = __float2bfloat16_rn(b0 - __bfloat162float(h0));
  wx1[5 * TP_G + j] = h1; wx1[6 * TP_G + j] = __float2bfloat16_rn(b1 - __bfloat162float(h1));
}

// tcgen05.ld.16x256b.x8: 16 accumulator rows x 64 columns; thread t gets rows t/4 (registers 4i, 4i+1) and t/4 + 8
// (registers 4i+2, 4i+3), columns 8i + 2(t%4), + 1.  No wait (see tp_ld_wait).
__device__ __forceinline__ void tp_ld16x8_issue(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}

// ---- cluster pieces (sample_tp_kernel<., C > 1>: the gate columns of a tile split over C CTAs) ----
__device__ __forceinline__ uint32_t tp_mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// 16 bytes into the shared memory of a CTA of the cluster; the store adds its bytes to the transaction count of the
// mbarrier at `bar` in that CTA
__device__ __forceinline__ void tp_st_async_v4(uint32_t dst, uint4 v, uint32_t bar) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(bar) : "memory");
}
__device__ __forceinline__ void tp_mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void tp_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// The wait for tp_ld16x8_issue: it also ties the destination registers to itself (the compiler must not consume them
// earlier: it cannot see that the load completes asynchronously).
__device__ __forceinline__ void tp_ld_wait(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                    "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]));
  asm volatile("" : "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                    "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31]));
}

// LSTM cell update of one unit (bias and input term are in the accumulator): returns h', writes c' to cn.
__device__ __forceinline__ float tp_cell(uint32_t vi, uint32_t vj, uint32_t vf, uint32_t vo, float cin, float &cn) {
  const float ig = sigmoid_approx(__uint_as_float(vi));
  const float jg = tanh_approx(__uint_as_float(vj));
  const float fg = sigmoid_approx(__uint_as_float(vf) + 1.0f);
  const float og = sigmoid_approx(__uint_as_float(vo));
  cn = fmaf(cin, fg, ig * jg);
  return tanh_approx(cn) * og;
}

// C: CTAs per 64-sketch tile (cluster size).  C = 1: a CTA computes all four 256-column gate chunks of a layer.  C = 2, 4
// (fewer tiles than a wave of SMs): rank r computes chunks [r * 4 / C, (r + 1) * 4 / C), streams only their weights, and
// sends every new h block into all C operand buffers with st.async (the bytes complete the receivers' mbarriers); output
// projection and draw run redundantly in every CTA (identical operands, identical bits), rank 0 writes the strokes.
template <bool DBG, int C>
__global__ void __launch_bounds__(TP_THREADS, 1)
sample_tp_kernel(const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmWo, const __grid_constant__ CUtensorMap tmX0,
                 const __grid_constant__ CUtensorMap tmX1, const __grid_constant__ CUtensorMap tmXo, const TpArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gen = smem_raw + (base - smem_u32(smem_raw));
  // H0[parity], H1[parity]: operand blocks; ring; xa: input-point operand block (the K = 16 slice `parity` of each row);
  // xs[parity][row] = {unused, pen | done << 8}; barriers
  const uint32_t s_h0 = base, s_h1 = base + 2 * TP_HBUF_BYTES;
  const uint32_t s_ring = base + 4 * TP_HBUF_BYTES;
  const uint32_t s_xa = s_ring + TP_STAGES * TP_STAGE_BYTES;
  const uint32_t s_xs = s_xa + TP_KBLK_BYTES;
  const uint32_t bars = s_xs + 2 * TP_M * 8;
  const uint32_t bar_full = bars, bar_empty = bars + 8 * TP_STAGES;       // ring (5 + 5)
  const uint32_t bar_tfull = bars + 80, bar_tempty = bars + 104;            // the three gate-chunk accumulators
  const uint32_t bar_h0 = bars + 128, bar_h1 = bars + 160;                  // new h k-blocks written (4 each)
  const uint32_t bar_s0 = bars + 192, bar_s1 = bars + 200, bar_x = bars + 208;  // state final for the next step; next input
  const uint32_t bar_tfull_h = bars + 216, bar_tempty_h = bars + 224;       // the output projection's own accumulator
  const uint32_t tmem_holder = bars + 232;
  uint8_t *h0_gen = gen, *h1_gen = gen + 2 * TP_HBUF_BYTES, *xa_gen = gen + (s_xa - base);
  uint2 *xs_gen = reinterpret_cast<uint2 *>(gen + (s_xs - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int CH = 4 / C;      // gate chunks of a layer computed by this CTA
  constexpr int NOPEN = CH < 3 ? CH : 3;  // layer-0 chunks opened before the drawn point is needed
  uint32_t rank = 0;
  if constexpr (C > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int row0 = (blockIdx.x / C) * TP_M;
  const int steps = a.steps;

  for (int i = threadIdx.x; i < 4 * TP_HBUF_BYTES / 16; i += TP_THREADS) reinterpret_cast<uint4 *>(gen)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < TP_KBLK_BYTES / 16; i += TP_THREADS) reinterpret_cast<uint4 *>(xa_gen)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x < 2 * TP_M) xs_gen[threadIdx.x] = make_uint2(0u, 3u);  // first input: zeros (model.py:205)
  __syncthreads();
  if (threadIdx.x < TP_M) {  // the two bias columns of the input operand are constant ones, in both parity slices
    const int r = threadIdx.x;
    const uint4 ones = make_uint4(0u, 0u, 0x3f800000u, 0x00003f80u);  // bf16 columns 5 and 6 = 1.0
#pragma unroll
    for (int par = 0; par < 2; ++par) *reinterpret_cast<uint4 *>(xa_gen + r * 128 + (((2 * par) ^ (r & 7)) << 4)) = ones;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < TP_STAGES; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    for (int b = 0; b < 3; ++b) { mbar_init(bar_tfull + 8 * b, 1); mbar_init(bar_tempty + 8 * b, 8); }
    // new h blocks: written by this CTA's 8 epilogue warps (C = 1), or delivered by st.async from the chunk's owner and
    // counted in bytes (the MMA thread arms them every step)
    for (int k = 0; k < 4; ++k) { mbar_init(bar_h0 + 8 * k, C == 1 ? 8 : 1); mbar_init(bar_h1 + 8 * k, C == 1 ? 8 : 1); }
    mbar_init(bar_s0, 4); mbar_init(bar_s1, 4); mbar_init(bar_x, 4);
    mbar_init(bar_tfull_h, 1); mbar_init(bar_tempty_h, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW0)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW1)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmWo)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmX0)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmX1)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmXo)) : "memory");
  }
  if (warp == 1) tc_alloc(tmem_holder, 512);
  fence_proxy_async_smem();  // the zeroed operand blocks are read by the tensor core
  tc_fence_before();
  __syncthreads();
  if constexpr (C > 1) tp_cluster_sync();  // nobody sends into a CTA whose barriers are not initialised
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t *>(gen + (tmem_holder - base));
  const int nhb = a.nh / 64;  // 64-column boxes of the output projection

  if (warp == 0) {
    // ---- weight stream: the same sequence of stages every step, as far ahead as the ring allows ----
    if (elect_one()) {
      int it = 0;
      // box_bytes: TP_CHUNK_BYTES for the weight maps (32 k rows per box), half of it for the input-term maps (16 rows:
      // their single product reads no more); the 64-column chunks of a stage stay TP_CHUNK_BYTES apart either way
      auto stage = [&](const CUtensorMap *tm, int n0, int k0, int boxes, int box_bytes = TP_CHUNK_BYTES) {
        const int s = it % TP_STAGES, ph = (it / TP_STAGES) & 1;
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        const uint32_t sb = s_ring + s * TP_STAGE_BYTES;
        mbar_expect_tx(bar_full + 8 * s, boxes * box_bytes);
        for (int c = 0; c < boxes; ++c) tma_load_2d(sb + c * TP_CHUNK_BYTES, tm, bar_full + 8 * s, n0 + 64 * c, k0);
        ++it;
      };
      for (int step = 0; step < steps; ++step) {
        // layer 0 (h0 rows of W0 start at row 8; rows 0..7 are the x rows and padding): the h0 stages of the first NOPEN
        // chunks, then their input-term stages, then the remaining chunks -- the order the MMA warp consumes them in.  Every
        // accumulator sums h stages first and the input term last, whichever cluster size computes it: a sketch's
        // points do not depend on how many CTAs share its tile
        for (int i = 0; i < NOPEN; ++i)
          for (int j = 0; j < 8; ++j) stage(&tmW0, 256 * (int)(rank * CH + i), 8 + TP_BK * j, 4);
        for (int i = 0; i < NOPEN; ++i) stage(&tmX0, 256 * (int)(rank * CH + i), 0, 4, TP_CHUNK_BYTES / 2);
        for (int i = NOPEN; i < CH; ++i) {
          for (int j = 0; j < 8; ++j) stage(&tmW0, 256 * (int)(rank * CH + i), 8 + TP_BK * j, 4);
          stage(&tmX0, 256 * (int)(rank * CH + i), 0, 4, TP_CHUNK_BYTES / 2);
        }
        for (int i = 0; i < CH; ++i) {
          const int n0 = 256 * (int)(rank * CH + i);
          for (int j = 0; j < 16; ++j) stage(&tmW1, n0, j < 8 ? TP_H + TP_BK * j : TP_BK * (j - 8), 4);  // h1 rows, then input rows
          stage(&tmX1, n0, 0, 4, TP_CHUNK_BYTES / 2);
        }
        for (int j = 0; j < 8; ++j) stage(&tmWo, 0, TP_BK * j, nhb);
        stage(&tmXo, 0, 0, nhb, TP_CHUNK_BYTES / 2);
      }
    }
  } else if (warp == 1) {
    // ---- products ----
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(TP_M, 256, false, true);
      const uint32_t idesc_o = make_idesc_bf16(TP_M, a.nh, false, true);
      int it = 0;
      uint32_t g = 0;
      long long tw[4] = {0, 0, 0, 0};  // cycles waited for: ring stage, accumulator buffer, new h / state, input point
      const long long t_begin = clock64();
#define TP_TIMED(slot, stmt) do { if constexpr (DBG) { const long long t0_ = clock64(); stmt; tw[slot] += clock64() - t0_; } else { stmt; } } while (0)
      // one ring stage against the operand bytes at a_addr: nk products of K = 16
      auto kstage = [&](uint32_t acc, uint32_t a_addr, uint32_t id, bool first, int nk) {
        const int s = it % TP_STAGES, ph = (it / TP_STAGES) & 1;
        TP_TIMED(0, mbar_wait(bar_full + 8 * s, ph));
        tc_fence_after();
        const uint64_t da = make_desc(a_addr, 0, 1024), db = make_desc(s_ring + s * TP_STAGE_BYTES, TP_CHUNK_BYTES, 1024);
        for (int k = 0; k < nk; ++k)
          tc_mma_bf16(acc, da + (uint64_t)((32 * k) >> 4), db + (uint64_t)((2048 * k) >> 4), id, (!first || k > 0) ? 1u : 0u);
        tc_commit(bar_empty + 8 * s);
        ++it;
      };
      // Tensor memory holds four M = 64 accumulators: an M = 64 accumulator fills lanes 0..15 of every lane quarter and
      // may as well start at lane 16.  Slots 0, 1 (lanes 0.., columns 0 / 256) and 2 (lanes 16.., column 0) rotate over the
      // gate chunks; lanes 16.., columns 256.. belong to the output projection.
      auto slot_addr = [&](uint32_t slot) { return slot == 2 ? tmem + (16u << 16) : tmem + slot * 256; };
      auto acc_begin = [&]() -> uint32_t {
        const uint32_t slot = g % 3;
        TP_TIMED(1, mbar_wait(bar_tempty + 8 * slot, ((g / 3) & 1) ^ 1));
        tc_fence_after();
        return slot_addr(slot);
      };
      auto acc_end = [&]() { tc_commit(bar_tfull + 8 * (g % 3)); ++g; };
      // a new h block is complete: C = 1: the epilogue warps arrived; C > 1: its bytes have landed (st.async from the owner)
      auto wait_h = [&](uint32_t bar, uint32_t par) {
        if constexpr (C == 1) {
          TP_TIMED(2, mbar_wait(bar, par));
        } else {
          TP_TIMED(2, tp_mbar_wait_cluster(bar, par));
          fence_proxy_async_smem();  // written by st.async stores: order them before the tensor core's reads
        }
        tc_fence_after();
      };
      for (int step = 0; step < steps; ++step) {
        const uint32_t p = step & 1, prev = (step - 1) & 1;
        const uint32_t h0s = s_h0 + p * TP_HBUF_BYTES, h0n = s_h0 + (p ^ 1) * TP_HBUF_BYTES;
        const uint32_t h1s = s_h1 + p * TP_HBUF_BYTES, h1n = s_h1 + (p ^ 1) * TP_HBUF_BYTES;
        const uint32_t xa = s_xa + p * 32;  // this step's K = 16 slice of the input operand
        if constexpr (C > 1) {
          // this step's h blocks arrive as bytes: one arrival (this one) plus a block's 8 KB complete a barrier's phase.
          // Every barrier's previous phase is complete: this thread waited for all eight of them during the previous step.
          for (int k = 0; k < 4; ++k) { mbar_expect_tx(bar_h0 + 8 * k, TP_KBLK_BYTES); mbar_expect_tx(bar_h1 + 8 * k, TP_KBLK_BYTES); }
        }
        if (step > 0) { TP_TIMED(2, mbar_wait(bar_s0, prev)); tc_fence_after(); }
        {
          // the first NOPEN chunks: their accumulators take their h0 stages first, then -- once the point drawn at the
          // previous step is known -- their input-term stage (so that the output projection and the draw overlap products)
          uint32_t acc3[NOPEN];
          for (int c = 0; c < NOPEN; ++c) {
            const uint32_t slot = (g + c) % 3;
            TP_TIMED(c == 0 ? 1 : 3, mbar_wait(bar_tempty + 8 * slot, (((g + c) / 3) & 1) ^ 1));
            tc_fence_after();
            acc3[c] = slot_addr(slot);
            for (int j = 0; j < 8; ++j) kstage(acc3[c], h0s + (j >> 1) * TP_KBLK_BYTES + (j & 1) * 64, idesc, j == 0, 2);
          }
          if (step > 0) { TP_TIMED(3, mbar_wait(bar_x, prev)); tc_fence_after(); }
          for (int c = 0; c < NOPEN; ++c) {
            kstage(acc3[c], xa, idesc, false, 1);
            acc_end();
          }
        }
        for (int c = NOPEN; c < CH; ++c) {
          const uint32_t acc = acc_begin();
          for (int j = 0; j < 8; ++j) kstage(acc, h0s + (j >> 1) * TP_KBLK_BYTES + (j & 1) * 64, idesc, j == 0, 2);
          kstage(acc, xa, idesc, false, 1);
          acc_end();
        }
        if (step > 0) { TP_TIMED(2, mbar_wait(bar_s1, prev)); tc_fence_after(); }
        for (int c = 0; c < CH; ++c) {
          const uint32_t acc = acc_begin();
          for (int j = 0; j < 16; ++j) {
            if (j < 8) {
              kstage(acc, h1s + (j >> 1) * TP_KBLK_BYTES + (j & 1) * 64, idesc, j == 0, 2);
            } else {
              const int kb = (j - 8) >> 1;
              if (c == 0 && (j & 1) == 0) wait_h(bar_h0 + 8 * kb, p);
              kstage(acc, h0n + kb * TP_KBLK_BYTES + (j & 1) * 64, idesc, false, 2);
            }
          }
          kstage(acc, xa, idesc, false, 1);
          acc_end();
        }
        {
          // output projection: its own accumulator in lanes 16..31, columns 256 .. 256 + nh, so that the draw does not hold one
          // of the gate chunks' buffers
          TP_TIMED(1, mbar_wait(bar_tempty_h, p ^ 1));
          tc_fence_after();
          const uint32_t acc = tmem + (16u << 16) + 256u;
          for (int j = 0; j < 8; ++j) {
            if ((j & 1) == 0) wait_h(bar_h1 + 8 * (j >> 1), p);
            kstage(acc, h1n + (j >> 1) * TP_KBLK_BYTES + (j & 1) * 64, idesc_o, j == 0, 2);
          }
          kstage(acc, xa, idesc_o, false, 1);  // + output bias
          tc_commit(bar_tfull_h);
        }
      }
#undef TP_TIMED
      if (DBG && blockIdx.x == 0)
        printf("sample_tp: %lld cycles per step; MMA warp waited for ring %lld, accumulator %lld, h/state %lld, logits buffer and input point %lld\n",
               (clock64() - t_begin) / steps, tw[0] / steps, tw[1] / steps, tw[2] / steps, tw[3] / steps);
    }
  } else {
#include "sampler_tp_epi.inl"
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (C > 1) tp_cluster_sync();  // a CTA's shared memory stays up until nobody sends into it any more
  if (warp == 1) tc_dealloc(tmem, 512);
}

}  // namespace

size_t sk_sample_tp_scratch_bytes() { return (((size_t)(8 + TP_H) + 2 * TP_H + 2 * TP_XROWS) * TP_G + (size_t)TP_XROWS * TP_XO_PITCH) * 2; }

// 0: drawn here; -1: not eligible (sampler_tc.cu continues with its step-by-step path); > 0: error.
// wsm / bias_il / cst: the packed operand images and zeroed cell states sampler_tc.cu has just prepared;
// tables: sk_sample_tp_scratch_bytes() of scratch for the column-permuted weight images and the input-term / bias images.
int sk_sample_tp(skrnn_handle *h, cudaStream_t st, int n, int steps, float temp_mix, float temp_pen, int stop_if_eoc,
                 const float *uniforms, uint64_t seed, int sketch0, float *strokes, int *lengths, float *mixp,
                 const __nv_bfloat16 *const *wsm, const float *const *bias_il, float *const *cst, float *tables) {
  {
    const char *e = getenv("SKRNN_SAMPLE_TP");
    if (e && e[0] == '0') return -1;
  }
  if (h->cfg.cell != SKRNN_CELL_LSTM || h->H != TP_H || h->L != 2 || h->M > 32 || h->NOUT > 192 || n < TP_M) return -1;
  CUtensorMap tmW0, tmW1, tmWo, tmX0, tmX1, tmXo;
  __nv_bfloat16 *wp0 = reinterpret_cast<__nv_bfloat16 *>(tables), *wp1 = wp0 + (size_t)(8 + TP_H) * TP_G;
  __nv_bfloat16 *wx0 = wp1 + (size_t)2 * TP_H * TP_G, *wx1 = wx0 + (size_t)TP_XROWS * TP_G;
  __nv_bfloat16 *wxo = wx1 + (size_t)TP_XROWS * TP_G;
  {
    uint64_t dims[2] = {(uint64_t)TP_G, (uint64_t)(8 + TP_H)}, str[1] = {(uint64_t)TP_G * 2};
    uint32_t box[2] = {64, (uint32_t)TP_BK};
    SK_TRY(sk_make_tmap(&tmW0, 2, wp0, 2, dims, str, box));
    dims[1] = 2 * TP_H;
    SK_TRY(sk_make_tmap(&tmW1, 2, wp1, 2, dims, str, box));
    dims[1] = TP_XROWS;
    uint32_t boxx[2] = {64, 16};  // the input-term images feed one K = 16 product per accumulator
    SK_TRY(sk_make_tmap(&tmX0, 2, wx0, 2, dims, str, boxx));
    SK_TRY(sk_make_tmap(&tmX1, 2, wx1, 2, dims, str, boxx));
    uint64_t dimo[2] = {(uint64_t)h->NOUTP, (uint64_t)TP_H}, stro[1] = {(uint64_t)h->NOUTP * 2};
    SK_TRY(sk_make_tmap(&tmWo, 2, h->params_bf + h->pbf_out_w, 2, dimo, stro, box));
    uint64_t dimx[2] = {(uint64_t)TP_XO_PITCH, (uint64_t)TP_XROWS}, strx[1] = {(uint64_t)TP_XO_PITCH * 2};
    SK_TRY(sk_make_tmap(&tmXo, 2, wxo, 2, dimx, strx, boxx));
  }
  TpArgs a = {};
  a.n = n; a.steps = steps; a.M = h->M; a.NOUT = h->NOUT; a.nh = (h->NOUT + 63) / 64 * 64;
  a.data_scale = h->cfg.data_scale; a.temp_mix = temp_mix; a.temp_pen = temp_pen;
  a.stop_if_eoc = stop_if_eoc; a.sketch0 = sketch0; a.seed = seed; a.uniforms = uniforms;
  a.strokes = strokes; a.mixp = mixp; a.lengths = lengths;
  a.cst0 = cst[0]; a.cst1 = cst[1];
  { const char *e = getenv("SKRNN_TP_DEBUG"); a.debug = (e && e[0] == '1') ? 1 : 0; }
  {
    SkLaunch scope(h, st, SK_CAT_SAMPLER);
    tp_wperm_kernel<<<2 * h->num_sms, 256, 0, st>>>(wsm[0], wp0, 8 + TP_H);
    SK_CUDA(cudaGetLastError());
  }
  {
    SkLaunch scope(h, st, SK_CAT_SAMPLER);
    tp_wperm_kernel<<<2 * h->num_sms, 256, 0, st>>>(wsm[1], wp1, 2 * TP_H);
    SK_CUDA(cudaGetLastError());
  }
  {
    SkLaunch scope(h, st, SK_CAT_SAMPLER);
    tp_ximage_kernel<<<TP_G / 256, 256, 0, st>>>(wsm[0], bias_il[0], bias_il[1], wx0, wx1, h->params + h->off_out_b, h->NOUT, wxo);
    SK_CUDA(cudaGetLastError());
  }
  // Fewer tiles than SMs / C: the gate columns of a tile are split over a cluster of C CTAs (see the kernel).
  const int tiles = (n + TP_M - 1) / TP_M;
  auto config = [&](cudaLaunchConfig_t &cfg, cudaLaunchAttribute *at, int c) {
    cfg = cudaLaunchConfig_t{};
    cfg.gridDim = dim3((unsigned)(tiles * c)); cfg.blockDim = dim3(TP_THREADS); cfg.dynamicSmemBytes = TP_SMEM; cfg.stream = st;
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
  };
  // clusters of c CTAs (223 KB of shared memory each) that are resident at once: GPC sizes decide, not SMs / c
  auto resident = [&](auto kern, int c) -> int {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TP_SMEM) != cudaSuccess) { cudaGetLastError(); return 0; }
    cudaLaunchConfig_t cfg; cudaLaunchAttribute at[1];
    config(cfg, at, c);
    int nc = 0;
    if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
    return nc;
  };
  int cl = 1;
  {
    const char *e = getenv("SKRNN_TP_CLUSTER");  // 1, 2, 4: force; default: the largest split whose clusters fit one wave
    const int want = e ? atoi(e) : 0;
    if (want == 1 || want == 2 || want == 4) cl = want;
    else if (4 * tiles <= h->num_sms && tiles <= resident(sample_tp_kernel<false, 4>, 4)) cl = 4;
    else if (2 * tiles <= h->num_sms && tiles <= resident(sample_tp_kernel<false, 2>, 2)) cl = 2;
  }
  SkLaunch scope(h, st, SK_CAT_SAMPLER);
  auto launch = [&](auto kern, int c) -> int {
    SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TP_SMEM));
    cudaLaunchConfig_t cfg; cudaLaunchAttribute at[1];
    config(cfg, at, c);
    SK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmW0, tmW1, tmWo, tmX0, tmX1, tmXo, a));
    return 0;
  };
  if (a.debug && cl == 1) SK_TRY(launch(sample_tp_kernel<true, 1>, 1));
  else if (cl == 4) SK_TRY(launch(sample_tp_kernel<false, 4>, 4));
  else if (cl == 2) SK_TRY(launch(sample_tp_kernel<false, 2>, 2));
  else SK_TRY(launch(sample_tp_kernel<false, 1>, 1));
  SK_CUDA(cudaGetLastError());
  return 0;
}
