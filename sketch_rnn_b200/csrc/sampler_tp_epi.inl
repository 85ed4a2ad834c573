    // ---- epilogue warps of sample_tp_kernel (included into its body) ----
    // Both groups (warps 2-5 / 6-9) work on every accumulator of the sequence [4 chunks of layer 0][4 chunks of layer 1]
    // [output projection] per step: a chunk's 64 units are split between them (the step is a chain of dependent
    // chunks, so a chunk's latency counts, not only the warps' throughput).  An M = 64 accumulator keeps row 16q + i in
    // lane i (< 16) of lane quarter q: the draw and the state hold work thread-per-row on lanes 0..15, the cell update
    // spreads the rows over all 32 lanes (see below).
    const int e = (warp - 2) >> 2;
    const int q = warp & 3;
    const bool active = lane < 16;
    const int r = 16 * q + (lane & 15);
    const int sk = row0 + r;
    const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
    const uint32_t swz = (uint32_t)(r & 7);
    // the cell update's view of the quarter's rows (tcgen05.ld.16x256b): thread (t4 = lane / 4, m4 = lane % 4), rows rA, rB
    const int m4 = lane & 3, rA = 16 * q + (lane >> 2), rB = rA + 8;
    const int skA = row0 + rA, skB = row0 + rB;
    const bool validA = skA < a.n, validB = skB < a.n;
    int pen = 3, done = 0;
    bool hold = false, holdA = false, holdB = false;
    // per step this CTA has CH chunks of layer 0, CH of layer 1 (chunk rank * CH + i of the layer) and the output projection
    constexpr int NACC = 2 * CH + 1;
    const uint32_t nchunks = (uint32_t)steps * NACC;
    for (uint32_t g = 0; g < nchunks; ++g) {
      const int step = (int)(g / NACC), ci = (int)(g % NACC);
      const int cs = ci < 2 * CH ? 0 : 8;   // 8: the output projection
      const int p = step & 1;
      // gate chunks rotate over three accumulators: lanes 0..15 of every quarter, columns 0 and 256, and lanes 16..31,
      // column 0 (an M = 64 accumulator may start at lane 16); the output projection has lanes 16..31, columns 256..
      const uint32_t gg = (uint32_t)(step * 2 * CH + ci);  // gate-chunk count
      const uint32_t buf = gg % 3;
      const uint32_t tacc = cs < 8 ? (buf == 2 ? tq + (16u << 16) : tq + buf * 256) : tq + 256u;
      if (ci == 0) {  // this step's input point (drawn by group 0 after the previous output projection)
        if (step > 0) mbar_wait(bar_x, (uint32_t)((step - 1) & 1));
        const uint2 xv = xs_gen[p * TP_M + r];
        pen = (int)(xv.y & 0xffu);
        done = (int)((xv.y >> 8) & 1u);
        hold = pen == 1;  // an eoc input holds the state
        holdA = (xs_gen[p * TP_M + rA].y & 0xffu) == 1u;
        holdB = (xs_gen[p * TP_M + rB].y & 0xffu) == 1u;
      }
      if (cs < 8) mbar_wait(bar_tfull + 8 * buf, (gg / 3) & 1);
      else mbar_wait(bar_tfull_h, (uint32_t)p);
      tc_fence_after();
      if (cs < 8) {
        const int layer = ci / CH, c = (int)rank * CH + ci % CH;
        if (ci == 2 * CH - 1 && e == 0) {
          // every product of layer 1 (of this CTA) is complete: nobody reads this step's new h0 any more, held rows get the
          // old state back
          if (active && hold) {
            const uint4 *src = reinterpret_cast<const uint4 *>(h0_gen + p * TP_HBUF_BYTES + r * 128);
            uint4 *dst = reinterpret_cast<uint4 *>(h0_gen + (p ^ 1) * TP_HBUF_BYTES + r * 128);
#pragma unroll
            for (int kb = 0; kb < 4; ++kb)
#pragma unroll
              for (int ch = 0; ch < 8; ++ch) dst[kb * (TP_KBLK_BYTES / 16) + ch] = src[kb * (TP_KBLK_BYTES / 16) + ch];
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_s0);
        }
        // tcgen05.ld.16x256b spreads the 16 rows of this warp's lane quarter over all 32 threads: thread (t4, m) gets the
        // four gates of units 16m + 8e .. + 7 of the chunk (tp_src_col) for rows rA and rB = rA + 8, in two passes of 4 units;
        // the second pass's tensor-memory and cell-state loads are in flight during the first
        uint8_t *hblk = (layer == 0 ? h0_gen : h1_gen) + (p ^ 1) * TP_HBUF_BYTES + c * TP_KBLK_BYTES;
        const int uoff = 64 * c + 16 * m4 + 8 * e;
        float *cpA = (layer == 0 ? a.cst0 : a.cst1) + (size_t)(validA ? skA : 0) * TP_H + uoff;
        float *cpB = (layer == 0 ? a.cst0 : a.cst1) + (size_t)(validB ? skB : 0) * TP_H + uoff;
        uint32_t va[32], vb[32];
        const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
        tp_ld16x8_issue(tacc + (uint32_t)(128 * e), va);
        const float4 cA0 = validA ? *reinterpret_cast<const float4 *>(cpA) : zero4;
        const float4 cB0 = validB ? *reinterpret_cast<const float4 *>(cpB) : zero4;
        tp_ld_wait(va);
        tp_ld16x8_issue(tacc + (uint32_t)(128 * e + 64), vb);
        const float4 cA1 = validA ? *reinterpret_cast<const float4 *>(cpA + 4) : zero4;
        const float4 cB1 = validB ? *reinterpret_cast<const float4 *>(cpB + 4) : zero4;
        uint32_t hA[4], hB[4];
        auto pass = [&](int ps, const uint32_t (&v)[32], const float4 &qa, const float4 &qb) {
          const float ca[4] = {qa.x, qa.y, qa.z, qa.w}, cb[4] = {qb.x, qb.y, qb.z, qb.w};
          float na[4], nb[4], ha[4], hb[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            ha[i] = tp_cell(v[8 * i + 0], v[8 * i + 1], v[8 * i + 4], v[8 * i + 5], ca[i], na[i]);
            hb[i] = tp_cell(v[8 * i + 2], v[8 * i + 3], v[8 * i + 6], v[8 * i + 7], cb[i], nb[i]);
          }
#pragma unroll
          for (int i = 0; i < 4; i += 2) {
            const __nv_bfloat162 pa = __floats2bfloat162_rn(ha[i], ha[i + 1]), pb = __floats2bfloat162_rn(hb[i], hb[i + 1]);
            hA[2 * ps + (i >> 1)] = *reinterpret_cast<const uint32_t *>(&pa);
            hB[2 * ps + (i >> 1)] = *reinterpret_cast<const uint32_t *>(&pb);
          }
          if (validA && !holdA) *reinterpret_cast<float4 *>(cpA + 4 * ps) = make_float4(na[0], na[1], na[2], na[3]);
          if (validB && !holdB) *reinterpret_cast<float4 *>(cpB + 4 * ps) = make_float4(nb[0], nb[1], nb[2], nb[3]);
        };
        pass(0, va, cA0, cB0);
        tp_ld_wait(vb);
        pass(1, vb, cA1, cB1);
        {
          const uint32_t chunk = (uint32_t)(2 * m4 + e);
          const uint32_t offA = (uint32_t)(rA * 128) + ((chunk ^ (uint32_t)(rA & 7)) << 4);
          const uint32_t offB = (uint32_t)(rB * 128) + ((chunk ^ (uint32_t)(rB & 7)) << 4);
          if constexpr (C == 1) {
            *reinterpret_cast<uint4 *>(hblk + offA) = make_uint4(hA[0], hA[1], hA[2], hA[3]);
            *reinterpret_cast<uint4 *>(hblk + offB) = make_uint4(hB[0], hB[1], hB[2], hB[3]);
          } else {
            // into every CTA's operand buffer (this one included); the bytes complete the receivers' h-block barriers
            const uint32_t blk = (layer == 0 ? s_h0 : s_h1) + (uint32_t)((p ^ 1) * TP_HBUF_BYTES + c * TP_KBLK_BYTES);
            const uint32_t bar = (layer == 0 ? bar_h0 : bar_h1) + 8u * (uint32_t)c;
#pragma unroll
            for (int rr = 0; rr < C; ++rr) {
              const uint32_t dst = (rank + (uint32_t)rr) % C;
              const uint32_t rb = tp_mapa(bar, dst);
              tp_st_async_v4(tp_mapa(blk + offA, dst), make_uint4(hA[0], hA[1], hA[2], hA[3]), rb);
              tp_st_async_v4(tp_mapa(blk + offB, dst), make_uint4(hB[0], hB[1], hB[2], hB[3]), rb);
            }
          }
        }
        if constexpr (C == 1) fence_proxy_async_smem();  // the new h block is an operand of the tensor core
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (C == 1) mbar_arrive((layer == 0 ? bar_h0 : bar_h1) + 8 * c);
          mbar_arrive(bar_tempty + 8 * buf);
        }
      } else if (e == 1) {
        // output projection complete: nobody reads this step's new h1 any more, held rows get the old state back
        if (active && hold) {
          const uint4 *src = reinterpret_cast<const uint4 *>(h1_gen + p * TP_HBUF_BYTES + r * 128);
          uint4 *dst = reinterpret_cast<uint4 *>(h1_gen + (p ^ 1) * TP_HBUF_BYTES + r * 128);
#pragma unroll
          for (int kb = 0; kb < 4; ++kb)
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) dst[kb * (TP_KBLK_BYTES / 16) + ch] = src[kb * (TP_KBLK_BYTES / 16) + ch];
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_s1);
      } else {
        // ---- the draw (model.py:219-261) from this row's logits (bias included by the product), all in registers; the
        //      logits' accumulator lives in lanes 16..31 of the quarter, so those lanes draw ----
        const bool active = lane >= 16;
        const bool valid = active && sk < a.n;
        const bool writer = rank == 0;  // every CTA of a tile draws the same point; one of them records it
        const int NOUT = a.NOUT, M = a.M;
        uint32_t v0[32], v1[32];
        tc_ld32(tacc, v0);
        tc_ld32(tacc + 32u, v1);
        float pm[32];
        // mixture weights: softmax; from the third point on the reference re-normalises log(pi) / temperature
        // (model.py:220-224), which is the softmax of the logits over the temperature -- one exponential per component
        const float sc = step > 1 ? __frcp_rn(a.temp_mix) : 1.0f;
        float mx = -INFINITY;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          const float zl = __uint_as_float(3 + k < 32 ? v0[(3 + k) & 31] : v1[(3 + k) & 31]);
          pm[k] = k < M ? zl : -INFINITY;
          mx = fmaxf(mx, pm[k]);
        }
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) { pm[k] = __expf((pm[k] - mx) * sc); sum += pm[k]; }
        const float rs = __frcp_rn(sum);
#pragma unroll
        for (int k = 0; k < 32; ++k) pm[k] *= rs;
        float u4[4] = {0.f, 0.f, 0.f, 0.f};
        if (valid) {
          if (a.uniforms != nullptr) {
            const float4 uu = __ldg(reinterpret_cast<const float4 *>(a.uniforms + ((size_t)sk * steps + step) * 4));
            u4[0] = uu.x; u4[1] = uu.y; u4[2] = uu.z; u4[3] = uu.w;
          } else {
            uint32_t rr[4];
            skm_philox4x32((uint32_t)(a.sketch0 + sk), (uint32_t)step, 0u, 0x5a3711eu, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), rr);
#pragma unroll
            for (int k = 0; k < 4; ++k) u4[k] = skm_u01(rr[k]);
          }
        }
        int idx = -1;
        float cdf = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          cdf += pm[k];
          if (k < M && idx < 0 && cdf >= u4[0]) idx = k;
        }
        if (idx < 0) idx = M - 1;  // get_pi_idx == -1 -> last component
        // pen state (plain softmax; the reference never applies the pen temperature)
        const float z0 = __uint_as_float(v0[0]), z1 = __uint_as_float(v0[1]), z2 = __uint_as_float(v0[2]);
        const float pmx = fmaxf(z0, fmaxf(z1, z2));
        const float e0 = __expf(z0 - pmx), e1 = __expf(z1 - pmx), e2 = __expf(z2 - pmx);
        const float ps = e0 + e1 + e2;
        const bool record = a.mixp != nullptr && valid && !done && writer;
        float *mo = a.mixp + (record ? ((size_t)sk * steps + step) * NOUT : 0);
        if (record) {  // same record as the exact kernel (include/skrnn.h: mix_params_dev)
          mo[0] = e0 / ps; mo[1] = e1 / ps; mo[2] = e2 / ps;
#pragma unroll
          for (int k = 0; k < 32; ++k)
            if (k < M) mo[3 + k] = (step > 1) ? pm[k] / a.temp_pen : pm[k];
        }
        // the chosen component's five parameters: column 3 + g M + idx of the logits, picked out of the 32-column pieces
        float val[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
        for (int sub = 0; sub * 32 < NOUT; ++sub) {
          if (sub == 1) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v0[j] = v1[j];
          } else if (sub > 1) {
            tc_ld32(tacc + (uint32_t)(sub * 32), v0);
          }
#pragma unroll
          for (int gq = 1; gq <= 5; ++gq) {
            const int lo = 3 + gq * M;
            if (lo < 32 * sub + 32 && lo + M > 32 * sub) {
              const int want = lo + idx - 32 * sub;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j == want) val[gq] = __uint_as_float(v0[j]);
            }
          }
          if (record) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int col = 32 * sub + j;
              if (col >= 3 + M && col < NOUT) {
                const float zz = __uint_as_float(v0[j]);
                mo[col] = col < 3 + 3 * M ? zz : (col < 3 + 5 * M ? __expf(zz) : tanhf(zz));
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_tempty_h);
        uint32_t xpack = 0u, xflag = 3u | ((uint32_t)done << 8), xpen = 0u, xcont = 0u;
        if (valid && !done) {
          float eos = 0.f, eoc = 0.f, cont = 0.f;
          int pidx = 2;
          if (e0 / ps >= u4[1]) { eos = 1.f; pidx = 0; }
          else if ((e0 + e1) / ps >= u4[1]) { eoc = 1.f; pidx = 1; }
          else cont = 1.f;
          const float mu1 = val[1], mu2 = val[2];
          const float s1 = __expf(val[3]), s2 = __expf(val[4]);
          const float rho = tanhf(val[5]);
          float g1, g2;
          skm_box_muller(u4[2], u4[3], &g1, &g2);
          const float x1 = mu1 + s1 * g1;
          const float x2 = mu2 + s2 * (rho * g1 + sqrtf(fmaxf(0.f, 1.0f - rho * rho)) * g2);
          if (writer) {
            float *so = a.strokes + ((size_t)sk * steps + step) * 5;
            so[0] = x1 * a.data_scale; so[1] = x2 * a.data_scale; so[2] = eos; so[3] = eoc; so[4] = cont;
            a.lengths[sk] = step + 1;
          }
          if (a.stop_if_eoc && eoc > 0.f) done = 1;
          const __nv_bfloat162 xb = __floats2bfloat162_rn(x1, x2);  // the next input, rounded like every operand of this path
          xpack = *reinterpret_cast<const uint32_t *>(&xb);
          xflag = (uint32_t)pidx | ((uint32_t)done << 8);
          xpen = pidx == 0 ? 0x00003f80u : (pidx == 1 ? 0x3f800000u : 0u);  // (eos, eoc) as a bf16 pair
          xcont = pidx == 2 ? 0x00003f80u : 0u;
        }
        if (active) {
          xs_gen[(p ^ 1) * TP_M + r] = make_uint2(xpack, xflag);
          // operand columns (dx, dy, eos, eoc, cont, 1, 1, 0) of the next step's K = 16 slice
          *reinterpret_cast<uint4 *>(xa_gen + r * 128 + (((uint32_t)(2 * (p ^ 1)) ^ swz) << 4)) =
              make_uint4(xpack, xpen, xcont | 0x3f800000u, 0x00003f80u);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_x);
      }
    }
