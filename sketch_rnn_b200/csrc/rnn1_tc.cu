// rnn1_tc.cu -- tcgen05 persistent recurrent kernels for the GRU and vanilla-RNN cells (bf16 tensor-core path):
// reference model.py:14-25 (GRUCell / BasicRNNCell) unrolled by custom_rnn_autodecoder (model.py:66-95) and their
// BPTT (model.py:182), one launch per layer and direction running all T steps.
//
// Partition.  Unlike the LSTM (rnn_tc.cu), whose recurrent matrix needs a 16-CTA cluster, the recurrent weights of
// these cells fit on one or two SMs at the reference's sizes.  A batch tile of N = 16 rows is owned
//   * by ONE CTA when H <= 128 (all weight tiles in tensor memory), or
//   * by a CLUSTER OF TWO CTAs when 128 < H <= 256: each CTA owns 128 hidden units and keeps their weight rows in
//     tensor memory (gru, H = 256: three 128-column tiles forward, 128 + 256 columns backward).  The contraction runs
//     over all units, so both CTAs need the whole operand tile: each writes the k-blocks of its own units locally and
//     the MMA thread swaps them with the peer by one bulk copy through distributed shared memory per product (the copy
//     reports on the peer's mbarrier: async proxy end to end, no cluster-scope fence in the loop).
//   (SKRNN_R1_CLUSTER=0 keeps H > 128 on one CTA: half of the weight tiles then sit in shared memory, and an MMA that
//   reads its A tile from there runs 39 cycles instead of 9 -- measured 4.9 ms against 3.1 ms per gru training step.)
// The weights stay resident for the whole launch and the recurrent state never goes through L2 between steps: a step is
// MMA -> tcgen05.ld -> gate math in registers -> next operand tile in shared memory (-> peer) -> MMA.
//
// Products are transposed so that the hidden units are the MMA M dimension (128-row tiles; lane = unit) and the
// batch tile is the N dimension:   D^T[units, N] = W^T[units, K] * h^T[K, N].
//   * warp 16 issues tcgen05.mma; the A tiles (weights) are read from tensor memory (TS form) or from shared memory
//     (UMMA K-major, 128B swizzle), whichever the host-side placement gave them.  The same thread sends the operand
//     tiles to global memory with TMA stores: the tile the next product reads IS h_t (forward: the `hprev` record the
//     weight-gradient GEMM reads; GRU: also r*h) resp. d(pre-activation) (backward), already in bf16;
//   * warps 0-15 are elementwise: a warp reads one TMEM lane quarter of one m-tile (lane = unit) for NR = 4 or 8 of
//     the 16 batch columns, applies the gate math in registers (fp32 state in registers), and writes the next B operand
//     (bf16, K-major swizzled) straight into shared memory.  What it reads and writes per step for itself -- the
//     hoisted projection, the saved activations, d(out) -- is kept tile-transposed so that it is one 16-byte (8-byte)
//     access per gate, requested before it waits for the product and prefetched into L2 two steps earlier.
// GRU (TF 0.x semantics): two dependent products per step -- gates r,u on h, then the candidate on r*h.
// BPTT mirrors it: d(r*h) = d(pre_c) * Wc_h^T, then d(h) += [d(pre_r), d(pre_u)] * Wg_h^T.
#include <stdlib.h>
#include <string.h>
#include <type_traits>

#include "common.cuh"
#include "skrnn_math.h"
#include "tc_ptx.cuh"

namespace {

constexpr int R1_N = 16;          // batch rows per CTA (MMA N)
// An elementwise thread handles NR = 4 * MU of the 16 rows for its unit (MU = 128-unit m-tiles of the CTA): always 16
// elementwise warps -- 4 TMEM lane quarters x (16 / NR) row groups x MU m-tiles -- so that a one-m-tile CTA splits the
// per-step chain over twice the threads (these warps are instruction-latency bound, not issue bound).
constexpr int R1_EW = 16;
constexpr int R1_THREADS = 32 * (R1_EW + 1);
constexpr int R1_MAX_TILES = 6;


struct R1Tile {   // one 128-row weight tile (MMA A operand), K = 64 * nkb
  int in_tmem;    // 1: tensor memory at column `off`; 0: shared memory at byte offset `off`
  int off;
  int nkb;
  int img;        // byte offset of the tile's image in the packed buffer (shared-memory tiles pre-swizzled)
};

struct R1Args {
  int B, T, H, G, MU, nkb;
  int ntiles;
  R1Tile tile[R1_MAX_TILES];
  int smem_w_bytes, tmem_cols;
  int row0, dropout;
  float keep;
  uint64_t seed;
  const unsigned char *eoc;  // [T,B]
  const float *state_in;     // [B,H]
  float *state_out;
  const float *mask;         // [B,T,H] or null
  const char *xw_tt;         // tile-transposed bf16 [.., G, NR]: hoisted x*Wx + b
  __nv_bfloat16 *out_bf, *hprev_bf;
  char *acts_tt;             // tile-transposed bf16 [.., AC, NR]: gru r, u, candidate, h fed (AC = 4H); rnn: tanh output (AC = H)
  const char *dout_tt;       // tile-transposed fp32 [.., H, NR]
  float *dbias0, *dbias1;
  const uint8_t *wimg;
  int wimg_rank_bytes;       // cluster of 2: each rank's tiles are a separate image
  int debug;                 // SKRNN_TC_DEBUG=1: CTA 0 prints its per-phase cycle split
  int pf_dist, pf_l1;        // prefetch distance in steps, and whether it targets L1 (else L2)
};

#define R1_WAIT(bar, par, tag) mbar_wait_wd(bar, par, tag, 1u << 22)
// debug timing: laps of one thread accumulate into shared-memory slots
#define R1_LAP(slot) do { if (dbg) { const uint32_t now_ = (uint32_t)clock(); tk[slot] += (long long)(now_ - lapt); lapt = now_; } } while (0)

struct R1Smem {
  uint32_t base, s_w, s_b1, s_b2, s_bar;
  uint8_t *gen;
};
__device__ __forceinline__ R1Smem r1_smem(uint8_t *raw, int w_bytes, int b1_bytes, int b2_bytes) {
  R1Smem s;
  const uint32_t r = smem_u32(raw);
  s.base = (r + 1023u) & ~1023u;
  s.gen = raw + (s.base - r);
  s.s_w = s.base;
  s.s_b1 = s.s_w + (uint32_t)w_bytes;
  s.s_b2 = s.s_b1 + (uint32_t)b1_bytes;
  s.s_bar = s.s_b2 + (uint32_t)b2_bytes;
  return s;
}

// resident weights: shared-memory tiles are straight copies of their pre-swizzled images, tensor-memory tiles are
// written row by row (lane = row) with tcgen05.st
__device__ __forceinline__ void r1_load_smem_tiles(const R1Args &a, const uint8_t *wimg, const R1Smem &s, int tid, int nthreads) {
  for (int j = 0; j < a.ntiles; ++j) {
    if (a.tile[j].in_tmem) continue;
    const uint4 *src = reinterpret_cast<const uint4 *>(wimg + a.tile[j].img);
    uint4 *dst = reinterpret_cast<uint4 *>(s.gen + a.tile[j].off);
    for (int i = tid; i < a.tile[j].nkb * 1024; i += nthreads) dst[i] = __ldg(src + i);
  }
}
__device__ __forceinline__ void r1_load_tmem_tiles(const R1Args &a, const uint8_t *wimg, uint32_t tmem_base, int warp, int lane) {
  if (warp >= 4) return;
  const int e = 32 * warp + lane;
  for (int j = 0; j < a.ntiles; ++j) {
    if (!a.tile[j].in_tmem) continue;
    const int nkb = a.tile[j].nkb;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(wimg + a.tile[j].img) + (size_t)e * (nkb * 32);
    for (int cb = 0; cb < nkb; ++cb) {
      uint32_t v[32];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 x = __ldg(reinterpret_cast<const uint4 *>(src + cb * 32) + i);
        v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
      }
      tc_st32(tmem_base + (uint32_t)a.tile[j].off + ((uint32_t)(32 * warp) << 16) + (uint32_t)(cb * 32), v);
    }
  }
}

// One weight tile as the MMA thread sees it: resolved once before the time loop
struct R1A {
  bool in_tmem;
  uint32_t tm;    // tensor-memory address of the tile (TS form)
  uint64_t desc;  // shared-memory descriptor of its first k-block
};
__device__ __forceinline__ R1A r1_resolve(const R1Tile &tl, uint32_t s_w, uint32_t tmem_base) {
  R1A r;
  r.in_tmem = tl.in_tmem != 0;
  r.tm = tmem_base + (uint32_t)tl.off;
  r.desc = make_desc(s_w + (uint32_t)tl.off, 0, 1024);
  return r;
}
// D[acc] = A(tile) * B over NKBT k-blocks: B is a K-major tile of R1_N rows with descriptor dB0 (k-block stride
// R1_N*128 bytes).  Fully unrolled, all operands precomputed: the issuing thread spends ~2 instructions per MMA, so the
// tensor core's own rate is what counts (measured with tools/mma_rate.cu at N = 16: 9 cycles per MMA with the A tile
// in tensor memory, 39 from shared memory -- the 4 KB operand read -- and no penalty for chaining on one accumulator).
template <int NKBT>
__device__ __forceinline__ void r1_issue(uint32_t acc, const R1A &A, uint64_t dB0, uint32_t idesc) {
  if (A.in_tmem) {
#pragma unroll
    for (int kb = 0; kb < NKBT; ++kb)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        tc_mma_bf16_ts(acc, A.tm + (uint32_t)(kb * 32 + k * 8), dB0 + (uint64_t)(kb * (R1_N * 8) + 2 * k), idesc, (kb | k) ? 1u : 0u);
  } else {
#pragma unroll
    for (int kb = 0; kb < NKBT; ++kb)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        tc_mma_bf16(acc, A.desc + (uint64_t)(kb * 1024 + 2 * k), dB0 + (uint64_t)(kb * (R1_N * 8) + 2 * k), idesc, (kb | k) ? 1u : 0u);
  }
}

// the operand tile at s_b (nkb k-blocks of [R1_N rows][64 bf16], 128B-swizzled) -> columns [col0, col0 + 64*nkb) of rows
// [b0, b0 + R1_N) at step t of a [T,B,ld] bf16 array (rows past B are clipped by the tensor map)
// (k-blocks [kb_lo, kb_hi) of the tile: in a cluster each CTA stores the blocks of its own units)
__device__ __forceinline__ void r1_store_tile(const CUtensorMap *tm, uint32_t s_b, int kb_lo, int kb_hi, int col0, int b0, int t) {
  for (int kb = kb_lo; kb < kb_hi; ++kb) tma_store_3d(tm, s_b + (uint32_t)kb * (R1_N * 128), col0 + 64 * kb, b0, t);
}

// byte offset of element (row n, column k) in such a tile: per-thread part for a fixed k, then the row part
__device__ __forceinline__ uint32_t r1_bcol(int k) { return (uint32_t)((k >> 6) * (R1_N * 128) + ((k & 7) << 1)); }
__device__ __forceinline__ uint32_t r1_bchunk(int k) { return (uint32_t)(((k & 63) >> 3) << 4); }
#define R1_BOFF(colbase, chunk4, n) ((colbase) + (uint32_t)((n) * 128) + ((chunk4) ^ (uint32_t)(((n) & 7) << 4)))
__device__ __forceinline__ void r1_sts_bf16(uint8_t *p, float v) { *reinterpret_cast<__nv_bfloat16 *>(p) = __float2bfloat16_rn(v); }

// Loaded bf16 groups stay raw and are widened only after r1_pin, which the compiler cannot move above the mbarrier
// wait before it (both are volatile asm): otherwise the widening shift is scheduled right behind the load and the
// thread sits out the whole memory latency BEFORE it starts waiting for the product (measured: ~2000 cycles/step).
__device__ __forceinline__ void r1_pin(uint32_t &v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ void r1_pin(float &v) { asm volatile("" : "+f"(v)); }
// Per-step arrays the elementwise threads read or write for THEMSELVES are kept "tile-transposed":
//   X_tt[t][batch tile of 16][row group of NR][column][NR rows]
// so that a thread (one column = hidden unit, NR batch rows) moves its NR values with ONE vector access, coalesced
// across the warp's 32 units -- instead of NR two-byte accesses with NR address computations.  Rows past the batch are
// zero (r1_tt_kernel) or finite (written by the forward kernel).
template <int NR> struct R1G { uint32_t w[NR / 2]; };  // NR bf16 values
template <int NR>
__device__ __forceinline__ R1G<NR> r1_ld_group(const char *p) {
  R1G<NR> g;
  if constexpr (NR == 8) {
    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
    g.w[0] = v.x; g.w[1] = v.y; g.w[2] = v.z; g.w[3] = v.w;
  } else {
    const uint2 v = __ldg(reinterpret_cast<const uint2 *>(p));
    g.w[0] = v.x; g.w[1] = v.y;
  }
  return g;
}
template <int NR>
__device__ __forceinline__ R1G<NR> r1_zero_group() {
  R1G<NR> g;
#pragma unroll
  for (int i = 0; i < NR / 2; ++i) g.w[i] = 0u;
  return g;
}
template <int NR>
__device__ __forceinline__ void r1_pin(R1G<NR> &g) {
#pragma unroll
  for (int i = 0; i < NR / 2; ++i) r1_pin(g.w[i]);
}
template <int NR>
__device__ __forceinline__ float r1_at(const R1G<NR> &g, int n) {  // n is a compile-time constant after unrolling
  const uint32_t w = g.w[n >> 1];
  return __uint_as_float((n & 1) ? (w & 0xffff0000u) : (w << 16));
}
template <int NR>
__device__ __forceinline__ void r1_st_group(char *p, const float (&v)[NR]) {
  uint32_t w[NR / 2];
#pragma unroll
  for (int i = 0; i < NR / 2; ++i) {
    const __nv_bfloat162 b = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<const uint32_t *>(&b);
  }
  if constexpr (NR == 8) *reinterpret_cast<uint4 *>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  else *reinterpret_cast<uint2 *>(p) = make_uint2(w[0], w[1]);
}
template <int NR>
__device__ __forceinline__ void r1_ld_f32(const char *p, float (&d)[NR]) {
#pragma unroll
  for (int i = 0; i < NR / 4; ++i) {
    const float4 v = __ldg(reinterpret_cast<const float4 *>(p) + i);
    d[4 * i] = v.x; d[4 * i + 1] = v.y; d[4 * i + 2] = v.z; d[4 * i + 3] = v.w;
  }
}
__device__ __forceinline__ void r1_prefetch_line(const void *p, int l1) {
  if (l1) asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
  else asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// ---- cluster of 2 (CL = 2): the hidden units are split over two CTAs, each holding the weight rows of its half in
// tensor memory.  Both CTAs need the WHOLE operand tile (the contraction runs over all units): the elementwise warps
// write their units' k-blocks into the local tile exactly as without a cluster, and the MMA thread, once those are
// complete, sends them to the same place in the peer's tile with one bulk copy through distributed shared memory that
// reports its bytes on the peer's receive barrier -- async proxy end to end, no cluster-scope fences.
__device__ __forceinline__ uint32_t r1_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t r1_mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void r1_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void r1_dsmem_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster) : "memory");
}
__device__ __forceinline__ void r1_wait_cluster(uint32_t bar, uint32_t parity, int tag) {
  uint32_t done;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return;
    if (spin > (1u << 22)) {
      printf("libskrnn: receive barrier wait timed out (tag %d, parity %u, block %d)\n", tag, parity, (int)blockIdx.x);
      __trap();
    }
  }
}
// What the MMA thread of a cluster CTA needs to swap k-blocks [kb_lo, kb_hi) of an operand tile with its peer
struct R1Xchg {
  uint32_t peer_tile;   // cluster address of the peer's copy of the tile
  uint32_t peer_bar;    // cluster address of the peer's receive barrier
  uint32_t bar;         // my receive barrier
  uint32_t my_off, my_bytes, peer_bytes;
};
// my blocks of the tile at s_tile (+ `second` bytes further for the second half of a two-part tile) -> peer; then wait
// for the peer's blocks.  `parts` = 1 or 2.
template <int CL>
__device__ __forceinline__ void r1_exchange(const R1Xchg &x, uint32_t s_tile, uint32_t buf_off, int parts, uint32_t second,
                                            uint32_t parity, int tag) {
  if constexpr (CL > 1) {
    mbar_expect_tx(x.bar, x.peer_bytes * (uint32_t)parts);
    for (int p = 0; p < parts; ++p)
      r1_dsmem_copy(x.peer_tile + buf_off + x.my_off + (uint32_t)p * second, s_tile + buf_off + x.my_off + (uint32_t)p * second,
                    x.my_bytes, x.peer_bar);
    r1_wait_cluster(x.bar, parity, tag);
  }
}

// The elementwise warps have written their part of an operand tile: make it visible to the async proxy (tcgen05.mma,
// bulk copies) and tell the MMA thread.  (One CTA-wide named barrier followed by a single fence and arrival was
// measured too: no faster.)
__device__ __forceinline__ void r1_publish(uint32_t bar) {
  fence_proxy_async_smem();
  mbar_arrive(bar);
}

// eoc flag byte of (t, b0 + lane), requested one step ahead; the ballot of it is the step's row mask
__device__ __forceinline__ unsigned char r1_eoc_raw(const unsigned char *eoc, int t, int T, int B, int b0, int lane) {
  const int b = b0 + lane;
  return (t >= 0 && t < T && lane < R1_N && b < B) ? __ldg(eoc + (size_t)t * B + b) : (unsigned char)0;
}

// per-step output-dropout scale of (row b0+n, t, unit): 1/keep or 0 (model.py:29-30); 1 when dropout is off.  Rows past
// the batch get a scale too (their values are never stored).  The thread's rows are [nb, nb + NR) of the tile.
template <int NR>
__device__ __forceinline__ void r1_drop_scales(const R1Args &a, int t, int b0, int nb, int nvalid, int unit, bool uvalid,
                                               uint32_t drop_unit, float (&dsc)[NR]) {
#pragma unroll
  for (int n = 0; n < NR; ++n) dsc[n] = 1.0f;
  if (!a.dropout || !uvalid) return;
  const float inv_keep = 1.0f / a.keep;
  if (a.mask != nullptr) {  // injected mask (parity tests)
    const float *mp = a.mask + ((size_t)(b0 + nb) * a.T + t) * a.H + unit;
    const size_t rs = (size_t)a.T * a.H;
#pragma unroll
    for (int n = 0; n < NR; ++n)
      if (nb + n < nvalid) dsc[n] = __ldg(mp + n * rs) * inv_keep;
    return;
  }
  const uint32_t pre = skm_drop_prefix(a.seed, drop_unit, (uint32_t)t);
  const uint32_t row = (uint32_t)(a.row0 + b0 + nb);
  const float keep = a.keep;
#pragma unroll
  for (int n = 0; n < NR; ++n) dsc[n] = skm_u01(skm_drop_finish(pre, row + (uint32_t)n)) < keep ? inv_keep : 0.0f;
}

// ------------------------------------------------------------------------------------------------
// forward.  CELL: 1 gru, 2 rnn
// ------------------------------------------------------------------------------------------------
// MU: m-tiles (128 hidden units each) of THIS CTA; NKB: k-blocks of the contraction (all H units); CL: CTAs per cluster
template <int CELL, int MU, int NKB, int CL>
__global__ void __launch_bounds__(R1_THREADS, 1)
rnn1_fwd_kernel(const __grid_constant__ CUtensorMap tm_hp, const __grid_constant__ CUtensorMap tm_aux, R1Args a) {
  constexpr int N = R1_N;
  constexpr int NB1 = CELL == 2 ? 2 : 1;  // rnn: the h tile is double-buffered (see the store ordering below)
  extern __shared__ uint8_t smem_raw[];
  constexpr int H = 64 * NKB, G = (CELL == 1 ? 3 : 1) * H;  // rnn_size is a multiple of 64 here: known at compile time
  const int B = a.B, T = a.T;
  constexpr int nkb = NKB;
  constexpr int btile = NKB * N * 128;
  const R1Smem s = r1_smem(smem_raw, a.smem_w_bytes, NB1 * btile, CELL == 1 ? btile : 0);
  const uint32_t bar_op1 = s.s_bar, bar_acc1 = s.s_bar + 8, bar_op2 = s.s_bar + 16, bar_acc2 = s.s_bar + 24;
  const uint32_t tmem_holder = s.s_bar + 32, bar_rx1 = s.s_bar + 40, bar_rx2 = s.s_bar + 48;  // rx: the peer's k-blocks have landed
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(s.gen + (tmem_holder - s.base));
  constexpr int EW = R1_EW, THREADS = R1_THREADS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = CL == 1 ? 0 : (int)r1_cluster_rank();
  const int b0 = (blockIdx.x / CL) * N;
  const uint8_t *wimg = a.wimg + (size_t)rank * a.wimg_rank_bytes;
  __shared__ long long tk[16];
  if (tid < 16) tk[tid] = 0;

  r1_load_smem_tiles(a, wimg, s, tid, THREADS);
  if (tid == 0) {
    mbar_init(bar_op1, 32 * EW);
    mbar_init(bar_acc1, 1);
    mbar_init(bar_op2, 32 * EW);
    mbar_init(bar_acc2, 1);
    mbar_init(bar_rx1, 1);
    mbar_init(bar_rx2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == EW) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);
  {
    uint4 *z = reinterpret_cast<uint4 *>(s.gen + (s.s_b1 - s.base));
    const int n16 = (NB1 + (CELL == 1 ? 1 : 0)) * btile / 16;
    for (int i = tid; i < n16; i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if constexpr (CL > 1) r1_cluster_sync();  // the peer's tiles are zeroed and its barriers initialised before anything is sent there
  const uint32_t tmem_base = *tmem_holder_ptr;
  r1_load_tmem_tiles(a, wimg, tmem_base, warp, lane);

  constexpr int NR = 4 * MU, RG = R1_N / NR;  // rows per thread, row groups
  const int q = warp & 3, hh = (warp >> 2) % RG, mt = warp / (4 * RG);
  const int nb = NR * hh;  // my rows of the tile: [nb, nb + NR)
  const bool active = warp < EW;
  const int unit = (rank * MU + mt) * 128 + 32 * q + lane;
  const bool uvalid = active && unit < H;
  const int nvalid = min(N, B - b0);
  const uint32_t bcol = r1_bcol(unit), bch = r1_bchunk(unit);
  uint8_t *b1_gen = s.gen + (s.s_b1 - s.base);
  uint8_t *b2_gen = s.gen + (s.s_b2 - s.base);
  const int kb_lo = CL == 1 ? 0 : rank * MU * 2, kb_hi = CL == 1 ? nkb : min(nkb, kb_lo + MU * 2);  // the k-blocks of my units
  float hcur[NR];
#pragma unroll
  for (int n = 0; n < NR; ++n) {
    hcur[n] = (uvalid && nb + n < nvalid) ? a.state_in[(size_t)(b0 + nb + n) * H + unit] : 0.f;
    if (uvalid) r1_sts_bf16(b1_gen + R1_BOFF(bcol, bch, nb + n), hcur[n]);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == EW) {
    // ---------------- MMA issuer + TMA stores of the operand tiles ----------------
    // Ordering of a tile's store against its next overwrite: the store is issued right after a tcgen05.commit, the
    // elementwise warps overwrite the tile only after a LATER commit has arrived, and this thread waits for its
    // outstanding stores to have read shared memory before issuing that later commit (one full elementwise phase
    // after the store was issued, so the wait costs nothing).  The rnn has one product per step, hence two h tiles.
    if (elect_one()) {
      const bool dbg = a.debug && blockIdx.x == 0;
      uint32_t lapt = dbg ? (uint32_t)clock() : 0u;
      const uint32_t idesc = make_idesc_bf16(128, N, false, false);
      constexpr int NG = (CELL == 1 ? 2 : 1) * MU;  // tiles of the first product
      R1A A1[NG], A2[MU];
#pragma unroll
      for (int j = 0; j < NG; ++j) A1[j] = r1_resolve(a.tile[j], s.s_w, tmem_base);
#pragma unroll
      for (int j = 0; j < MU; ++j) A2[j] = r1_resolve(a.tile[CELL == 1 ? 2 * MU + j : j], s.s_w, tmem_base);
      const uint64_t dB2 = make_desc(s.s_b2, 0, 1024);
      R1Xchg x1 = {}, x2 = {};
      if constexpr (CL > 1) {
        const int p_lo = (1 - rank) * MU * 2, p_hi = min(nkb, p_lo + MU * 2);
        x1.peer_tile = r1_mapa(s.s_b1, (uint32_t)(1 - rank)); x1.peer_bar = r1_mapa(bar_rx1, (uint32_t)(1 - rank)); x1.bar = bar_rx1;
        x2.peer_tile = r1_mapa(s.s_b2, (uint32_t)(1 - rank)); x2.peer_bar = r1_mapa(bar_rx2, (uint32_t)(1 - rank)); x2.bar = bar_rx2;
        x1.my_off = x2.my_off = (uint32_t)kb_lo * (N * 128);
        x1.my_bytes = x2.my_bytes = (uint32_t)(kb_hi - kb_lo) * (N * 128);
        x1.peer_bytes = x2.peer_bytes = (uint32_t)(p_hi - p_lo) * (N * 128);
      }
      for (int t = 0; t < T; ++t) {
        if (t > 0) {
          R1_WAIT(bar_op1, (t - 1) & 1, 1);
          tc_fence_after();
        }
        const uint32_t b1off = (uint32_t)((NB1 == 2 ? (t & 1) : 0) * btile);
        r1_exchange<CL>(x1, s.s_b1, b1off, 1, 0, t & 1, 21);
        R1_LAP(8);
        const uint32_t sb1 = s.s_b1 + b1off;
        const uint64_t dB1 = make_desc(sb1, 0, 1024);
#pragma unroll
        for (int j = 0; j < NG; ++j) r1_issue<NKB>(tmem_base + (uint32_t)(j * N), A1[j], dB1, idesc);
        tma_store_wait_read();
        tc_commit(bar_acc1);
        r1_store_tile(&tm_hp, sb1, kb_lo, kb_hi, 0, b0, t);  // h fed to step t
        R1_LAP(9);
        if constexpr (CELL == 1) {
          R1_WAIT(bar_op2, t & 1, 2);
          tc_fence_after();
          r1_exchange<CL>(x2, s.s_b2, 0, 1, 0, t & 1, 22);
          R1_LAP(10);
#pragma unroll
          for (int j = 0; j < MU; ++j) r1_issue<NKB>(tmem_base + (uint32_t)((2 * MU + j) * N), A2[j], dB2, idesc);
          tma_store_wait_read();
          tc_commit(bar_acc2);
          r1_store_tile(&tm_aux, s.s_b2, kb_lo, kb_hi, 0, b0, t);  // r*h of step t
          R1_LAP(11);
        }
      }
      tma_store_wait_all();
      if (dbg)
        printf("[rnn1 fwd cell %d] MMA thread cycles/step: wait operand-1 %lld, issue+store %lld, wait operand-2 %lld, issue+store %lld\n",
               CELL, tk[8] / T, tk[9] / T, tk[10] / T, tk[11] / T);
    }
    __syncwarp();
  } else if (active) {
    // ---------------- elementwise warps ----------------
    const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * q) << 16);
    const uint32_t drop_unit = skm_drop_unit(a.seed, (uint32_t)unit);
    // per-thread views of (t, b0, unit) of the step's arrays: base pointer advanced per step, row n at n * row stride
    const int NT = (B + N - 1) / N, tile = blockIdx.x / CL;
    const int AC = CELL == 1 ? 4 * H : H;
    constexpr int GB = NR * 2;  // bytes of one thread's group of bf16 values
    const char *xw_p = a.xw_tt + ((size_t)(tile * RG + hh) * G + unit) * GB;
    char *act_p = a.acts_tt + ((size_t)(tile * RG + hh) * AC + unit) * GB;
    const size_t xw_ts = (size_t)NT * RG * G * GB, act_ts = (size_t)NT * RG * AC * GB;
    const uint32_t gs = (uint32_t)H * GB;  // from one gate's plane to the next
    char *out_p = reinterpret_cast<char *>(a.out_bf + (size_t)(b0 + nb) * H + unit);
    const uint32_t out_rs = (uint32_t)H * 2u;
    const size_t out_ts = (size_t)B * H * 2;
    const bool pf_lane = uvalid && (lane * GB) % 128 == 0;  // one lane per 128-byte line of the warp's groups
    unsigned char eoc_raw = r1_eoc_raw(a.eoc, 0, T, B, b0, lane);
    const bool dbg = a.debug && blockIdx.x == 0 && tid == 0;
    uint32_t lapt = dbg ? (uint32_t)clock() : 0u;

    auto body = [&](auto full_tag, int t) {
      constexpr bool FULL = decltype(full_tag)::value;
      // everything that does not depend on the recurrence is requested before waiting for the product
      R1G<NR> xq0 = r1_zero_group<NR>(), xq1 = xq0, xq2 = xq0;
      if (uvalid) {
        xq0 = r1_ld_group<NR>(xw_p);
        if constexpr (CELL == 1) {
          xq1 = r1_ld_group<NR>(xw_p + gs);
          xq2 = r1_ld_group<NR>(xw_p + 2 * gs);
          r1_st_group<NR>(act_p + 3 * gs, hcur);  // the state fed to this step (BPTT reads it back in the same layout)
        }
      }
      R1_LAP(7);
      if (a.pf_dist >= 0 && pf_lane && t + a.pf_dist < T) {
        const char *pp = xw_p + (size_t)a.pf_dist * xw_ts;
        r1_prefetch_line(pp, a.pf_l1);
        if constexpr (CELL == 1) {
          r1_prefetch_line(pp + gs, a.pf_l1);
          r1_prefetch_line(pp + 2 * gs, a.pf_l1);
        }
      }
      R1_LAP(12);
      const uint32_t eoc_m = __ballot_sync(0xffffffffu, eoc_raw != 0);
      eoc_raw = r1_eoc_raw(a.eoc, t + 1, T, B, b0, lane);
      float dsc[NR];
      r1_drop_scales<NR>(a, t, b0, nb, nvalid, unit, uvalid, drop_unit, dsc);
      float ug[NR];
#pragma unroll
      for (int n = 0; n < NR; ++n) ug[n] = 0.f;
      R1_LAP(0);
      if constexpr (CELL == 1) {
        R1_WAIT(bar_acc1, t & 1, 3);
        R1_LAP(1);
        tc_fence_after();
        r1_pin<NR>(xq0); r1_pin<NR>(xq1);
        uint32_t rr[NR], ru[NR];
        tc_ldn<NR>(lane_addr + (uint32_t)nb + (uint32_t)(mt * N), rr);
        tc_ldn<NR>(lane_addr + (uint32_t)nb + (uint32_t)((MU + mt) * N), ru);
        tc_fence_before();
        float rg[NR];
#pragma unroll
        for (int n = 0; n < NR; ++n) {
          rg[n] = sigmoid_approx(__uint_as_float(rr[n]) + r1_at<NR>(xq0, n));
          ug[n] = sigmoid_approx(__uint_as_float(ru[n]) + r1_at<NR>(xq1, n));
          if (uvalid) r1_sts_bf16(b2_gen + R1_BOFF(bcol, bch, nb + n), rg[n] * hcur[n]);
        }
        r1_publish(bar_op2);
        R1_LAP(2);
        if (uvalid) {
          r1_st_group<NR>(act_p, rg);
          r1_st_group<NR>(act_p + gs, ug);
        }
        R1_LAP(3);
      }
      R1_WAIT(CELL == 1 ? bar_acc2 : bar_acc1, t & 1, 4);
      R1_LAP(4);
      tc_fence_after();
      if constexpr (CELL == 1) r1_pin<NR>(xq2); else r1_pin<NR>(xq0);
      float x2[NR];
      uint32_t rc[NR];
      tc_ldn<NR>(lane_addr + (uint32_t)nb + (uint32_t)((CELL == 1 ? 2 * MU + mt : mt) * N), rc);
      tc_fence_before();
      float h2[NR];
      const uint32_t b1w = NB1 == 2 ? (uint32_t)(((t + 1) & 1) * btile) : 0u;
#pragma unroll
      for (int n = 0; n < NR; ++n) {
        const float cand = tanh_approx(__uint_as_float(rc[n]) + r1_at<NR>(CELL == 1 ? xq2 : xq0, n));
        h2[n] = CELL == 1 ? fmaf(ug[n], hcur[n] - cand, cand) : cand;   // u*h + (1-u)*c
        float hn = h2[n];
        if ((eoc_m >> (nb + n)) & 1u)  // eoc: back to the FED state (model.py:82-92); warp-uniform and rare
          hn = uvalid ? __ldg(a.state_in + (size_t)(b0 + nb + n) * H + unit) : 0.f;
        hcur[n] = hn;
        if (uvalid) r1_sts_bf16(b1_gen + b1w + R1_BOFF(bcol, bch, nb + n), hn);
        x2[n] = CELL == 1 ? cand : h2[n];
      }
      r1_publish(bar_op1);
      R1_LAP(5);
      // off the dependent chain: layer output (pre reset, dropout applied) and the activation BPTT needs
      if (uvalid) {
        r1_st_group<NR>(act_p + (CELL == 1 ? 2 * gs : 0u), x2);  // gru: candidate; rnn: the tanh output
#pragma unroll
        for (int n = 0; n < NR; ++n)
          if (FULL || nb + n < nvalid)
            *reinterpret_cast<__nv_bfloat16 *>(out_p + (uint32_t)n * out_rs) = __float2bfloat16_rn(h2[n] * dsc[n]);
      }
      xw_p += xw_ts; out_p += out_ts; act_p += act_ts;
      R1_LAP(6);
    };
    for (int t = 0; t < T; ++t) {
      if (nvalid == N) body(std::true_type{}, t); else body(std::false_type{}, t);
    }
    if (uvalid) {
#pragma unroll
      for (int n = 0; n < NR; ++n)
        if (nb + n < nvalid) a.state_out[(size_t)(b0 + nb + n) * H + unit] = hcur[n];
    }
    if (dbg)
      printf("[rnn1 fwd cell %d] elementwise thread cycles/step: loads %lld, prefetch %lld, eoc + dropout %lld, wait acc-1 %lld, gates r,u + operand %lld, stores %lld, "
             "wait acc %lld, cell + operand %lld, stores %lld\n", CELL, tk[7] / T, tk[12] / T, tk[0] / T, tk[1] / T, tk[2] / T, tk[3] / T, tk[4] / T, tk[5] / T,
             tk[6] / T);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) r1_cluster_sync();  // nobody leaves while the peer may still write into it or arrive on its barriers
  if (warp == EW) tc_dealloc(tmem_base, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// backward (BPTT).  d(pre-activation) leaves as TMA stores of the operand tiles: no per-thread global stores.
// ------------------------------------------------------------------------------------------------
template <int CELL, int MU, int NKB, int CL>
__global__ void __launch_bounds__(R1_THREADS, 1) rnn1_bwd_kernel(const __grid_constant__ CUtensorMap tm_dp, R1Args a) {
  constexpr int N = R1_N;
  constexpr int NB1 = CELL == 2 ? 2 : 1;
  extern __shared__ uint8_t smem_raw[];
  constexpr int H = 64 * NKB, G = (CELL == 1 ? 3 : 1) * H;  // rnn_size is a multiple of 64 here: known at compile time
  const int B = a.B, T = a.T;
  constexpr int nkb = NKB;
  constexpr int btile = NKB * N * 128;
  // b1: d(pre) of the candidate (gru) / of the cell (rnn, two buffers), K = H;  b2 (gru): d(pre) of the gates, K = 2H
  const R1Smem s = r1_smem(smem_raw, a.smem_w_bytes, NB1 * btile, CELL == 1 ? 2 * btile : 0);
  const uint32_t bar_op1 = s.s_bar, bar_acc1 = s.s_bar + 8, bar_op2 = s.s_bar + 16, bar_acc2 = s.s_bar + 24;
  const uint32_t tmem_holder = s.s_bar + 32, bar_rx1 = s.s_bar + 40, bar_rx2 = s.s_bar + 48;  // rx: the peer's k-blocks have landed
  volatile uint32_t *tmem_holder_ptr = reinterpret_cast<volatile uint32_t *>(s.gen + (tmem_holder - s.base));
  constexpr int EW = R1_EW, THREADS = R1_THREADS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = CL == 1 ? 0 : (int)r1_cluster_rank();
  const int b0 = (blockIdx.x / CL) * N;
  const uint8_t *wimg = a.wimg + (size_t)rank * a.wimg_rank_bytes;
  __shared__ long long tk[16];
  if (tid < 16) tk[tid] = 0;

  r1_load_smem_tiles(a, wimg, s, tid, THREADS);
  if (tid == 0) {
    mbar_init(bar_op1, 32 * EW);
    mbar_init(bar_acc1, 1);
    mbar_init(bar_op2, 32 * EW);
    mbar_init(bar_acc2, 1);
    mbar_init(bar_rx1, 1);
    mbar_init(bar_rx2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == EW) tc_alloc(tmem_holder, (uint32_t)a.tmem_cols);
  {
    uint4 *z = reinterpret_cast<uint4 *>(s.gen + (s.s_b1 - s.base));
    const int n16 = (NB1 + (CELL == 1 ? 2 : 0)) * btile / 16;
    for (int i = tid; i < n16; i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder_ptr;
  r1_load_tmem_tiles(a, wimg, tmem_base, warp, lane);
  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) r1_cluster_sync();  // the peer's tiles are zeroed and its barriers initialised before anything is sent there
  tc_fence_after();

  constexpr int NR = 4 * MU, RG = R1_N / NR;  // rows per thread, row groups
  const int q = warp & 3, hh = (warp >> 2) % RG, mt = warp / (4 * RG);
  const int nb = NR * hh;  // my rows of the tile: [nb, nb + NR)
  const bool active = warp < EW;
  const int unit = (rank * MU + mt) * 128 + 32 * q + lane;
  const bool uvalid = active && unit < H;
  const int nvalid = min(N, B - b0);
  const int kb_lo = CL == 1 ? 0 : rank * MU * 2, kb_hi = CL == 1 ? nkb : min(nkb, kb_lo + MU * 2);  // the k-blocks of my units

  if (warp == EW) {
    // ---------------- MMA issuer + TMA stores of d(pre) (same ordering argument as in the forward kernel) ----------------
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, N, false, false);
      R1A A1[MU], A2[MU];
#pragma unroll
      for (int j = 0; j < MU; ++j) {
        A1[j] = r1_resolve(a.tile[j], s.s_w, tmem_base);
        A2[j] = r1_resolve(a.tile[CELL == 1 ? MU + j : j], s.s_w, tmem_base);
      }
      const uint64_t dB2 = make_desc(s.s_b2, 0, 1024);
      R1Xchg x1 = {}, x2 = {};
      if constexpr (CL > 1) {
        const int p_lo = (1 - rank) * MU * 2, p_hi = min(nkb, p_lo + MU * 2);
        x1.peer_tile = r1_mapa(s.s_b1, (uint32_t)(1 - rank)); x1.peer_bar = r1_mapa(bar_rx1, (uint32_t)(1 - rank)); x1.bar = bar_rx1;
        x2.peer_tile = r1_mapa(s.s_b2, (uint32_t)(1 - rank)); x2.peer_bar = r1_mapa(bar_rx2, (uint32_t)(1 - rank)); x2.bar = bar_rx2;
        x1.my_off = x2.my_off = (uint32_t)kb_lo * (N * 128);
        x1.my_bytes = x2.my_bytes = (uint32_t)(kb_hi - kb_lo) * (N * 128);
        x1.peer_bytes = x2.peer_bytes = (uint32_t)(p_hi - p_lo) * (N * 128);
      }
      for (int t = T - 1, st = 0; t >= 0; --t, ++st) {
        if constexpr (CELL == 1) {
          // d(r*h) = d(pre_c) * Wc_h^T
          R1_WAIT(bar_op1, st & 1, 11);
          tc_fence_after();
          r1_exchange<CL>(x1, s.s_b1, 0, 1, 0, st & 1, 23);
          const uint64_t dB1 = make_desc(s.s_b1, 0, 1024);
#pragma unroll
          for (int j = 0; j < MU; ++j) r1_issue<NKB>(tmem_base + (uint32_t)(j * N), A1[j], dB1, idesc);
          tma_store_wait_read();
          tc_commit(bar_acc1);
          r1_store_tile(&tm_dp, s.s_b1, kb_lo, kb_hi, 2 * H, b0, t);
          R1_WAIT(bar_op2, st & 1, 12);
          tc_fence_after();
          if (CL > 1 && t > 0) r1_exchange<CL>(x2, s.s_b2, 0, 2, (uint32_t)btile, st & 1, 24);
          if (t > 0) {  // d(h_{t-1}) += [d(pre_r), d(pre_u)] * Wg_h^T
#pragma unroll
            for (int j = 0; j < MU; ++j) r1_issue<2 * NKB>(tmem_base + (uint32_t)((MU + j) * N), A2[j], dB2, idesc);
          }
          tma_store_wait_read();
          if (t > 0) tc_commit(bar_acc2);
          r1_store_tile(&tm_dp, s.s_b2, kb_lo, kb_hi, 0, b0, t);              // d(pre_r)
          r1_store_tile(&tm_dp, s.s_b2, nkb + kb_lo, nkb + kb_hi, 0, b0, t);  // d(pre_u)
        } else {
          R1_WAIT(bar_op1, st & 1, 13);
          tc_fence_after();
          if (CL > 1 && t > 0) r1_exchange<CL>(x1, s.s_b1, (uint32_t)((st & 1) * btile), 1, 0, st & 1, 25);
          const uint32_t sb1 = s.s_b1 + (uint32_t)((st & 1) * btile);
          const uint64_t dB1 = make_desc(sb1, 0, 1024);
          if (t > 0) {
#pragma unroll
            for (int j = 0; j < MU; ++j) r1_issue<NKB>(tmem_base + (uint32_t)(j * N), A1[j], dB1, idesc);
          }
          tma_store_wait_read();
          if (t > 0) tc_commit(bar_acc1);
          r1_store_tile(&tm_dp, sb1, kb_lo, kb_hi, 0, b0, t);
        }
      }
      tma_store_wait_all();
    }
    __syncwarp();
  } else if (active) {
    // ---------------- elementwise warps ----------------
    const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * q) << 16);
    const uint32_t drop_unit = skm_drop_unit(a.seed, (uint32_t)unit);
    const uint32_t bcol = r1_bcol(unit), bch = r1_bchunk(unit);
    const uint32_t bcol_u = r1_bcol(H + unit), bch_u = r1_bchunk(H + unit);
    uint8_t *b1_gen = s.gen + (s.s_b1 - s.base);
    uint8_t *b2_gen = s.gen + (s.s_b2 - s.base);
    // per-thread views of (t, tile, unit) of the tile-transposed arrays, walking backwards in t
    const int NT = (B + N - 1) / N, tile = blockIdx.x / CL;
    const int AC = CELL == 1 ? 4 * H : H;
    constexpr int GB = NR * 2, DB = NR * 4;  // bytes of one thread's group of bf16 / fp32 values
    const size_t act_ts = (size_t)NT * RG * AC * GB, do_ts = (size_t)NT * RG * H * DB;
    const char *act_p = a.acts_tt + (size_t)(T - 1) * act_ts + ((size_t)(tile * RG + hh) * AC + unit) * GB;
    const char *do_p = a.dout_tt + (size_t)(T - 1) * do_ts + ((size_t)(tile * RG + hh) * H + unit) * DB;
    const uint32_t gs = (uint32_t)H * GB;
    const bool pf_lane = uvalid && (lane * GB) % 128 == 0, pf_lane_d = uvalid && (lane * DB) % 128 == 0;
    unsigned char eoc_raw = r1_eoc_raw(a.eoc, T - 1, T, B, b0, lane);
    float dhc[NR];
#pragma unroll
    for (int n = 0; n < NR; ++n) dhc[n] = 0.f;
    float bs0 = 0.f, bs1 = 0.f, bs2 = 0.f;  // bias gradients of my unit: rnn: bs0; gru: r, u, c
    const bool dbg = a.debug && blockIdx.x == 0 && tid == 0;
    uint32_t lapt = dbg ? (uint32_t)clock() : 0u;

    auto body = [&](auto full_tag, int t, int st) {
      constexpr bool FULL = decltype(full_tag)::value;
      // saved activations and d(out) of the step: requested before waiting for the previous step's product
      R1G<NR> aq0 = r1_zero_group<NR>(), aq1 = aq0, aq2 = aq0, hq = aq0;
      float d_o[NR];
#pragma unroll
      for (int n = 0; n < NR; ++n) d_o[n] = 0.f;
      if (uvalid) {
        aq0 = r1_ld_group<NR>(act_p);
        r1_ld_f32<NR>(do_p, d_o);
        if constexpr (CELL == 1) {
          aq1 = r1_ld_group<NR>(act_p + gs);
          aq2 = r1_ld_group<NR>(act_p + 2 * gs);
          hq = r1_ld_group<NR>(act_p + 3 * gs);
        }
      }
      R1_LAP(7);
      if (a.pf_dist >= 0 && t - a.pf_dist >= 0) {
        if (pf_lane_d) r1_prefetch_line(do_p - (size_t)a.pf_dist * do_ts, a.pf_l1);
        if (pf_lane) {
          const char *pa = act_p - (size_t)a.pf_dist * act_ts;
          r1_prefetch_line(pa, a.pf_l1);
          if constexpr (CELL == 1) {
            r1_prefetch_line(pa + gs, a.pf_l1);
            r1_prefetch_line(pa + 2 * gs, a.pf_l1);
            r1_prefetch_line(pa + 3 * gs, a.pf_l1);
          }
        }
      }
      R1_LAP(12);
      const uint32_t eoc_m = __ballot_sync(0xffffffffu, eoc_raw != 0);
      eoc_raw = r1_eoc_raw(a.eoc, t - 1, T, B, b0, lane);
      float dsc[NR];
      r1_drop_scales<NR>(a, t, b0, nb, nvalid, unit, uvalid, drop_unit, dsc);
      R1_LAP(0);
      if (st > 0) {  // d(h_t) from the step after: the product issued last iteration
        R1_WAIT(CELL == 1 ? bar_acc2 : bar_acc1, (st - 1) & 1, 14);
        tc_fence_after();
        uint32_t r[NR];
        tc_ldn<NR>(lane_addr + (uint32_t)nb + (uint32_t)((CELL == 1 ? MU + mt : mt) * N), r);
        tc_fence_before();
#pragma unroll
        for (int n = 0; n < NR; ++n) dhc[n] = (CELL == 1 ? dhc[n] : 0.f) + __uint_as_float(r[n]);
      }
      // (see r1_pin: the loads above stay in flight across the wait)
      r1_pin<NR>(aq0);
#pragma unroll
      for (int n = 0; n < NR; ++n) r1_pin(d_o[n]);
      if constexpr (CELL == 1) { r1_pin<NR>(aq1); r1_pin<NR>(aq2); r1_pin<NR>(hq); }
      float a0[NR], a1[NR], a2[NR], hv[NR];
#pragma unroll
      for (int n = 0; n < NR; ++n) {
        a0[n] = r1_at<NR>(aq0, n);
        a1[n] = 0.f; a2[n] = 0.f; hv[n] = 0.f;
        if constexpr (CELL == 1) { a1[n] = r1_at<NR>(aq1, n); a2[n] = r1_at<NR>(aq2, n); hv[n] = r1_at<NR>(hq, n); }
      }
      R1_LAP(1);
      if constexpr (CELL == 2) {
        const uint32_t b1w = (uint32_t)((st & 1) * btile);
#pragma unroll
        for (int n = 0; n < NR; ++n) {
          const float dh2 = d_o[n] * dsc[n] + (((eoc_m >> (nb + n)) & 1u) ? 0.f : dhc[n]);
          const float dp = dh2 * (1.0f - a0[n] * a0[n]);
          bs0 += dp;
          if (uvalid) r1_sts_bf16(b1_gen + b1w + R1_BOFF(bcol, bch, nb + n), dp);
        }
        r1_publish(bar_op1);
        R1_LAP(2);
      } else {
        // a0 = r, a1 = u, a2 = candidate, hv = h fed to the step
        float du[NR];
#pragma unroll
        for (int n = 0; n < NR; ++n) {
          const float dh2 = d_o[n] * dsc[n] + (((eoc_m >> (nb + n)) & 1u) ? 0.f : dhc[n]);
          du[n] = dh2 * (hv[n] - a2[n]);
          const float dpc = dh2 * (1.0f - a1[n]) * (1.0f - a2[n] * a2[n]);
          dhc[n] = dh2 * a1[n];  // direct path h -> h'; the two products add to it below / next iteration
          bs2 += dpc;
          if (uvalid) r1_sts_bf16(b1_gen + R1_BOFF(bcol, bch, nb + n), dpc);
        }
        r1_publish(bar_op1);
        R1_LAP(2);
        R1_WAIT(bar_acc1, st & 1, 15);
        R1_LAP(4);
        tc_fence_after();
        uint32_t r[NR];
        tc_ldn<NR>(lane_addr + (uint32_t)nb + (uint32_t)(mt * N), r);
        tc_fence_before();
#pragma unroll
        for (int n = 0; n < NR; ++n) {
          const float drh = __uint_as_float(r[n]);
          dhc[n] = fmaf(drh, a0[n], dhc[n]);
          const float dpr = drh * hv[n] * a0[n] * (1.0f - a0[n]);
          const float dpu = du[n] * a1[n] * (1.0f - a1[n]);
          bs0 += dpr;
          bs1 += dpu;
          if (uvalid) {
            r1_sts_bf16(b2_gen + R1_BOFF(bcol, bch, nb + n), dpr);
            r1_sts_bf16(b2_gen + R1_BOFF(bcol_u, bch_u, nb + n), dpu);
          }
        }
        r1_publish(bar_op2);
        R1_LAP(5);
      }
      act_p -= act_ts; do_p -= do_ts;
    };
    for (int t = T - 1, st = 0; t >= 0; --t, ++st) {
      if (nvalid == N) body(std::true_type{}, t, st); else body(std::false_type{}, t, st);
    }
    if (uvalid) {
      if constexpr (CELL == 2) {
        atomicAdd(a.dbias0 + unit, bs0);
      } else {
        atomicAdd(a.dbias0 + unit, bs0);
        atomicAdd(a.dbias0 + H + unit, bs1);
        atomicAdd(a.dbias1 + unit, bs2);
      }
    }
    if (dbg)
      printf("[rnn1 bwd cell %d] elementwise thread cycles/step: loads %lld, prefetch %lld, eoc + dropout %lld, wait d(h) product %lld, elementwise-1 + operand %lld, "
             "wait d(r*h) product %lld, elementwise-2 + operand %lld\n", CELL, tk[7] / T, tk[12] / T, tk[0] / T, tk[1] / T, tk[2] / T, tk[4] / T, tk[5] / T);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) r1_cluster_sync();  // nobody leaves while the peer may still write into it or arrive on its barriers
  if (warp == EW) tc_dealloc(tmem_base, (uint32_t)a.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// weight images.  Tile value (m, k): trans ? mat[(I + k) * ld + c0 + m0 + m] : mat[(I + m0 + m) * ld + c0 + k]
//   forward  (trans): A[m][k] = W[k][gate column of unit m0+m]          (h * W: contraction over the input unit k)
//   backward        : A[m][k] = W[m0+m][k]                              (d(pre) * W^T: contraction over the gate column)
// Rows m0+m >= H and columns k >= K are zero.
// ------------------------------------------------------------------------------------------------
struct R1PackTile { const float *mat; int ld, m0, c0, trans, K, in_tmem, img, nkb; };
struct R1PackArgs { R1PackTile t[R1_MAX_TILES]; int ntiles, I, H; };

__global__ void r1_pack_kernel(R1PackArgs p, uint8_t *__restrict__ dst) {
  for (int j = 0; j < p.ntiles; ++j) {
    const R1PackTile &tl = p.t[j];
    const int total = 128 * tl.nkb * 64;
    __nv_bfloat16 *out = reinterpret_cast<__nv_bfloat16 *>(dst + tl.img);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
      int m, k;
      if (tl.in_tmem) {
        m = i / (tl.nkb * 64);
        k = i % (tl.nkb * 64);
      } else {
        const int kb = i / 8192, w = i % 8192;
        m = w / 64;
        const int within = w % 64, chunk = within >> 3, el = within & 7;
        k = kb * 64 + (((chunk ^ (m & 7)) << 3) + el);
      }
      const int u = tl.m0 + m;
      float v = 0.f;
      if (u < p.H && k < tl.K)
        v = tl.trans ? tl.mat[(size_t)(p.I + k) * tl.ld + tl.c0 + u] : tl.mat[(size_t)(p.I + u) * tl.ld + tl.c0 + k];
      out[i] = __float2bfloat16_rn(v);
    }
  }
}

// [T][B][ld >= C] row-major -> tile-transposed [T][NT][16/NR][C][NR] (rows past the batch: zero).  One thread per group.
template <typename E, int NR>
__global__ void r1_tt_kernel(const E *__restrict__ in, int ld, E *__restrict__ out, int B, int C, int NT) {
  constexpr int RG = R1_N / NR;
  const int col = blockIdx.y * blockDim.x + threadIdx.x;
  if (col >= C) return;
  const int rg = blockIdx.x % RG, tile = (blockIdx.x / RG) % NT, t = (blockIdx.x / RG) / NT;
  const int b = tile * R1_N + rg * NR;
  alignas(16) E v[NR];
#pragma unroll
  for (int n = 0; n < NR; ++n) v[n] = (b + n < B) ? in[((size_t)t * B + b + n) * ld + col] : E(0.f);
  E *dst = out + ((size_t)blockIdx.x * C + col) * NR;
  if constexpr (NR * sizeof(E) == 8) {
    *reinterpret_cast<uint2 *>(dst) = *reinterpret_cast<const uint2 *>(v);
  } else {
#pragma unroll
    for (int i = 0; i < (int)(NR * sizeof(E) / 16); ++i) reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(v)[i];
  }
}

}  // namespace

bool sk_rnn1_tc_eligible(const skrnn_handle *h) {
  if (h->cfg.precision != SKRNN_PREC_BF16) return false;
  if (h->cfg.cell != SKRNN_CELL_GRU && h->cfg.cell != SKRNN_CELL_RNN) return false;
  if (h->H % 64 != 0 || h->H > 256) return false;
  static int env_off = -1;
  if (env_off < 0) { const char *e = getenv("SKRNN_DISABLE_TC_RNN"); env_off = (e && e[0] == '1') ? 1 : 0; }
  return !env_off;
}

namespace {

int r1_map(CUtensorMap *m, const void *ptr, int ld, int B, int T) {
  uint64_t dims[3] = {(uint64_t)ld, (uint64_t)B, (uint64_t)T};
  uint64_t str[2] = {(uint64_t)ld * 2, (uint64_t)B * ld * 2};
  uint32_t box[3] = {64, (uint32_t)R1_N, 1};
  return sk_make_tmap(m, 2, ptr, 3, dims, str, box);
}

}  // namespace

// 0: the layer ran here; -1: not eligible (the caller uses the fp32 recurrent kernels); > 0: error
int sk_rnn1_tc_layer(skrnn_handle *h, cudaStream_t st, bool bwd, int l, int B, int T, const float *state_in,
                     float *state_out, const float *mask, uint64_t seed, int row0, bool training) {
  if (!sk_rnn1_tc_eligible(h)) return -1;
  const int H = h->H, G = h->G, cell = h->cfg.cell;
  const bool gru = cell == SKRNN_CELL_GRU;
  if (h->ws.xw_tt == nullptr || h->ws.dout_tt == nullptr || h->ws.tt_rows < (int64_t)T * ((B + R1_N - 1) / R1_N) * R1_N) return -1;
  if (h->ws.acts_bf[l] == nullptr || h->ws.out_bf[l] == nullptr || h->ws.hprev_bf[l] == nullptr || h->ws.dpre_bf == nullptr ||
      h->ws.xw_il[l] == nullptr || (gru && h->ws.aux_bf[l] == nullptr))
    return -1;
  const int MU_all = (H + 127) / 128, nkb = H / 64;
  // H > 128: a cluster of 2 CTAs per batch tile, each holding the weight rows of its 128 units in tensor memory (the
  // whole matrix does not fit in one SM's 512 columns, and a tile read from shared memory runs 4x slower)
  static int env_cl = -1;
  if (env_cl < 0) { const char *e = getenv("SKRNN_R1_CLUSTER"); env_cl = (e && e[0] == '0') ? 0 : 1; }
  const int CL = (MU_all == 2 && env_cl) ? 2 : 1;
  const int MU = MU_all / CL;  // m-tiles per CTA
  const SkLayerParams &lp = h->lp[l];
  const float *W0 = h->params + lp.w[0];                     // rnn: [I+H, H]; gru gates: [I+H, 2H]
  const float *W1 = gru ? h->params + lp.w[1] : nullptr;     // gru candidate: [I+H, H]

  // tiles in the order the kernels index them
  R1PackArgs pk = {};
  pk.I = lp.I; pk.H = H;
  int nt = 0;
  auto add = [&](const float *mat, int ld, int m0, int c0, int trans, int K) {
    R1PackTile &t = pk.t[nt++];
    t.mat = mat; t.ld = ld; t.m0 = m0; t.c0 = c0; t.trans = trans; t.K = K; t.nkb = K / 64;
  };
  for (int r = 0; r < CL; ++r) {  // each rank's tiles, in the same order
    const int u0 = 128 * r * MU;
    if (!bwd) {
      if (gru) {
        for (int g = 0; g < 2; ++g)
          for (int m = 0; m < MU; ++m) add(W0, 2 * H, u0 + 128 * m, g * H, 1, H);
        for (int m = 0; m < MU; ++m) add(W1, H, u0 + 128 * m, 0, 1, H);
      } else {
        for (int m = 0; m < MU; ++m) add(W0, H, u0 + 128 * m, 0, 1, H);
      }
    } else {
      if (gru) {
        for (int m = 0; m < MU; ++m) add(W1, H, u0 + 128 * m, 0, 0, H);
        for (int m = 0; m < MU; ++m) add(W0, 2 * H, u0 + 128 * m, 0, 0, 2 * H);
      } else {
        for (int m = 0; m < MU; ++m) add(W0, H, u0 + 128 * m, 0, 0, H);
      }
    }
  }
  pk.ntiles = nt;
  nt /= CL;  // tiles of one rank: placement and offsets are the same in every CTA
  // placement: the subset of tiles kept in shared memory must fit beside the operand tiles, the rest goes to tensor
  // memory behind the accumulators (512 columns in all).  A tile read from shared memory costs 4 KB of operand
  // traffic per MMA (32 cycles at 128 B/clk against 8 cycles of math at N = 16), so take the feasible split with the
  // MOST tensor memory
  const int btile = nkb * R1_N * 128;
  const int b_bytes = (gru ? (bwd ? 3 : 2) : 2) * btile;
  const int smem_cap = 227 * 1024 - 1024 - 256 - b_bytes;
  const int acc_cols = (!bwd ? (gru ? 3 : 1) : (gru ? 2 : 1)) * MU * R1_N;
  int best = -1, best_cols = -1;
  for (int mask_s = 0; mask_s < (1 << nt); ++mask_s) {
    int sb = 0, cols = acc_cols;
    for (int j = 0; j < nt; ++j) {
      if ((mask_s >> j) & 1) sb += pk.t[j].nkb * 16384;
      else cols += pk.t[j].nkb * 32;
    }
    if (sb <= smem_cap && cols <= 512 && cols > best_cols) { best = mask_s; best_cols = cols; }
  }
  if (best < 0) return -1;
  R1Args a = {};
  int soff = 0, coff = acc_cols, ioff = 0;
  for (int j = 0; j < nt; ++j) {
    R1PackTile &t = pk.t[j];
    t.in_tmem = ((best >> j) & 1) ? 0 : 1;
    t.img = ioff;
    ioff += t.nkb * 16384;
    a.tile[j].in_tmem = t.in_tmem;
    a.tile[j].nkb = t.nkb;
    a.tile[j].img = t.img;
    if (t.in_tmem) { a.tile[j].off = coff; coff += t.nkb * 32; }
    else { a.tile[j].off = soff; soff += t.nkb * 16384; }
  }
  a.wimg_rank_bytes = ioff;
  for (int r = 1; r < CL; ++r)
    for (int j = 0; j < nt; ++j) {
      pk.t[r * nt + j].in_tmem = pk.t[j].in_tmem;
      pk.t[r * nt + j].img = r * a.wimg_rank_bytes + pk.t[j].img;
    }
  ioff *= CL;
  SK_CHECK((int64_t)ioff <= h->ws.wpack_floats * 4, "weight image workspace too small");
  {
    SkLaunch scope(h, st, SK_CAT_MISC);
    r1_pack_kernel<<<std::min(4 * h->num_sms, (ioff / 2 + 255) / 256), 256, 0, st>>>(pk, reinterpret_cast<uint8_t *>(h->ws.wpack));
    SK_CUDA(cudaGetLastError());
  }
  a.B = B; a.T = T; a.H = H; a.G = G; a.MU = MU; a.nkb = nkb; a.ntiles = nt;
  a.smem_w_bytes = soff;
  int tc = 32;
  while (tc < coff) tc <<= 1;
  a.tmem_cols = tc;
  a.row0 = row0;
  a.dropout = (training && l == h->L - 1 && h->cfg.keep_prob < 1.0f) ? 1 : 0;
  a.keep = h->cfg.keep_prob;
  a.seed = seed;
  a.eoc = h->ws.eoc; a.state_in = state_in; a.state_out = state_out; a.mask = mask;
  a.xw_tt = reinterpret_cast<const char *>(h->ws.xw_tt);
  a.out_bf = h->ws.out_bf[l]; a.hprev_bf = h->ws.hprev_bf[l];
  a.acts_tt = reinterpret_cast<char *>(h->ws.acts_bf[l]);  // (allocated with the tile-transposed size for these cells)
  a.dout_tt = reinterpret_cast<const char *>(h->ws.dout_tt);
  a.dbias0 = h->grads + lp.b[0];
  a.dbias1 = gru ? h->grads + lp.b[1] : nullptr;
  a.wimg = reinterpret_cast<const uint8_t *>(h->ws.wpack);
  {
    static int dbg_env = -1;
    if (dbg_env < 0) { const char *e = getenv("SKRNN_TC_DEBUG"); dbg_env = (e && e[0] == '1') ? 1 : 0; }
    a.debug = dbg_env;
    static int pf_d = -1, pf_l = -1;
    if (pf_d < 0) {
      const char *e = getenv("SKRNN_R1_PREFETCH");  // "<distance>[L1]", e.g. "2" (default) or "1L1"
      pf_d = e ? atoi(e) : 2;
      pf_l = (e && strstr(e, "L1")) ? 1 : 0;
    }
    a.pf_dist = pf_d; a.pf_l1 = pf_l;
  }
  const size_t smem = (size_t)soff + b_bytes + 1024 + 256;
  const int grid = CL * ((B + R1_N - 1) / R1_N);
  cudaLaunchConfig_t lc = {};
  cudaLaunchAttribute lattr[1];
  lc.gridDim = dim3((unsigned)grid);
  lc.blockDim = dim3((unsigned)R1_THREADS);
  lc.dynamicSmemBytes = smem;
  lc.stream = st;
  lattr[0].id = cudaLaunchAttributeClusterDimension;
  lattr[0].val.clusterDim.x = (unsigned)CL;
  lattr[0].val.clusterDim.y = 1;
  lattr[0].val.clusterDim.z = 1;
  lc.attrs = lattr;
  lc.numAttrs = 1;
  // kernel instances by (m-tiles per CTA, k-blocks, CTAs per cluster): H = 64, 128, 192, 256
#define R1_DISPATCH(LAUNCH)                                   \
  do {                                                        \
    if (CL == 2 && nkb == 3) { LAUNCH(1, 3, 2); }             \
    else if (CL == 2) { LAUNCH(1, 4, 2); }                    \
    else if (MU == 1 && nkb == 1) { LAUNCH(1, 1, 1); }        \
    else if (MU == 1 && nkb == 2) { LAUNCH(1, 2, 1); }        \
    else if (MU == 2 && nkb == 3) { LAUNCH(2, 3, 1); }        \
    else { LAUNCH(2, 4, 1); }                                 \
  } while (0)
  if (!bwd) {
    if (!h->xw_il_ok[l]) return -1;  // the projection was hoisted in the fp32 format
    CUtensorMap tm_hp, tm_aux;
    SK_TRY(r1_map(&tm_hp, h->ws.hprev_bf[l], H, B, T));
    SK_TRY(r1_map(&tm_aux, gru ? h->ws.aux_bf[l] : h->ws.hprev_bf[l], H, B, T));
    {  // the hoisted projection in the layout the elementwise threads read
      SkLaunch scope(h, st, SK_CAT_MISC);
      const int NT = (B + R1_N - 1) / R1_N;
      if (MU == 1) r1_tt_kernel<__nv_bfloat16, 4><<<dim3((unsigned)(T * NT * 4), (unsigned)((G + 255) / 256)), 256, 0, st>>>(h->ws.xw_il[l], G, h->ws.xw_tt, B, G, NT);
      else r1_tt_kernel<__nv_bfloat16, 8><<<dim3((unsigned)(T * NT * 2), (unsigned)((G + 255) / 256)), 256, 0, st>>>(h->ws.xw_il[l], G, h->ws.xw_tt, B, G, NT);
      SK_CUDA(cudaGetLastError());
    }
    SkLaunch scope(h, st, SK_CAT_RNN_FWD);
#define R1_FWD(MUV, KBV, CLV)                                                                                     \
  do {                                                                                                            \
    auto kern = gru ? rnn1_fwd_kernel<1, MUV, KBV, CLV> : rnn1_fwd_kernel<2, MUV, KBV, CLV>;                      \
    SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                  \
    SK_CUDA(cudaLaunchKernelEx(&lc, kern, tm_hp, tm_aux, a));                                                     \
  } while (0)
    R1_DISPATCH(R1_FWD);
#undef R1_FWD
  } else {
    CUtensorMap tm_dp;
    SK_TRY(r1_map(&tm_dp, h->ws.dpre_bf, G, B, T));
    {  // d(out) of the layer in the layout the elementwise threads read
      SkLaunch scope(h, st, SK_CAT_MISC);
      const int NT = (B + R1_N - 1) / R1_N;
      if (MU == 1) r1_tt_kernel<float, 4><<<dim3((unsigned)(T * NT * 4), (unsigned)((H + 255) / 256)), 256, 0, st>>>(h->ws.dout, H, h->ws.dout_tt, B, H, NT);
      else r1_tt_kernel<float, 8><<<dim3((unsigned)(T * NT * 2), (unsigned)((H + 255) / 256)), 256, 0, st>>>(h->ws.dout, H, h->ws.dout_tt, B, H, NT);
      SK_CUDA(cudaGetLastError());
    }
    SkLaunch scope(h, st, SK_CAT_RNN_BWD);
#define R1_BWD(MUV, KBV, CLV)                                                                                     \
  do {                                                                                                            \
    auto kern = gru ? rnn1_bwd_kernel<1, MUV, KBV, CLV> : rnn1_bwd_kernel<2, MUV, KBV, CLV>;                      \
    SK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                  \
    SK_CUDA(cudaLaunchKernelEx(&lc, kern, tm_dp, a));                                                             \
  } while (0)
    R1_DISPATCH(R1_BWD);
#undef R1_BWD
    h->bias_grad_done = true;
  }
#undef R1_DISPATCH
  SK_CUDA(cudaGetLastError());
  return 0;
}
