// Epilogue role of lookup_tc_kernel (included inside the kernel body; warps 2-17, four threads per query = TMEM lane).
        // =============================================================================================================
        // EPT (= 4) threads per query (warps 2 + 4 * part + quad, same TMEM lane quadrant): thread `part` gathers rows
        // 2 * part, 2 * part + 1 of every chunk into the query's window and evaluates the output rows j = part (mod EPT); the
        // threads of a query meet at a named barrier.  (The epilogue is bound by dependent-issue latency -- ncu: stall `wait` --
        // so it wants warps: one per scheduler took 157 us per cfg2 lookup, two 93 us.)
        const int quad = warp & 3;                       // TMEM lane quadrant this warp may read
        const int q = 32 * quad + lane;                  // query index inside the block: row q of operand A = TMEM lane q
        const int ew = warp - 2, et = ew * 32 + lane;    // epilogue warp / thread index
        const int half = ew >> 2;                        // `part`: which of the query's EPT threads
        auto pair_bar = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(2 + quad), "n"(32 * EPT) : "memory"); };
        const int qy = q >> 4, qx = q & 15;              // the block is an 8 x 16 patch of the fmap, row-major
        float* const my = Cs + q;                        // the query's window, one column of [cell][query]: my[cell * NQ]
        const int L = P.L;
        const unsigned HW = (unsigned)P.HW;
        const bool pow2_scale = P.mul != 0.0f;       // 1/sqrt(C) is a power of two: applied exactly to the 81 outputs
        const float out_scale = pow2_scale ? P.mul : 1.0f;

        // bounding box of the block's windows at every level -> red[buf][warp][level] (one int4 per warp and level)
        auto block_origin = [&](long n, int& b, int& qy0, int& qx0) {
            const long blk = (long)blockIdx.x + n * (long)gridDim.x;
            b = (int)(blk / P.qb_per_img);
            const int rem = (int)(blk - (long)b * P.qb_per_img);
            qy0 = (rem / P.qb_x) * QB_H;
            qx0 = (rem % P.qb_x) * QB_W;
        };
        auto load_coords = [&](long n, float& cx, float& cy) {
            int b, qy0, qx0;
            block_origin(n, b, qy0, qx0);
            const float* cp = P.coords + (size_t)b * 2 * HW + (size_t)(qy0 + qy) * P.W + (qx0 + qx);
            cx = __ldg(cp);
            cy = __ldg(cp + HW);
        };
        auto reduce_boxes = [&](float cx, float cy, int buf) {
            for (int l = 0; l < L; ++l) {
                const float inv = __int_as_float((127 - l) << 23);          // 2^-l: the reference's division by 2^l, exactly
                const float lx = __fmul_rn(cx, inv), ly = __fmul_rn(cy, inv);
                const bool dead = !(fabsf(lx) < kDeadCoord) || !(fabsf(ly) < kDeadCoord);
                const int ax = __float2int_rd(lx) - R - 1, ay = __float2int_rd(ly) - R - 1;
                const int wl = P.lw[l], hl = P.lh[l];
                // columns / rows the clamped tap pairs of this query can touch (two zero columns on either side suffice)
                int x0 = dead ? 0x7fffffff : min(max(ax, -2), wl);
                int x1 = dead ? (int)0x80000000 : min(max(ax + PR - 2, -2), wl) + 1;
                int y0 = dead ? 0x7fffffff : min(max(ay, -2), hl);
                int y1 = dead ? (int)0x80000000 : min(max(ay + PR - 2, -2), hl) + 1;
                x0 = __reduce_min_sync(0xffffffffu, x0);
                x1 = __reduce_max_sync(0xffffffffu, x1);
                y0 = __reduce_min_sync(0xffffffffu, y0);
                y1 = __reduce_max_sync(0xffffffffu, y1);
                if (lane == 0 && half == 0) reinterpret_cast<int4*>(red)[(buf * 4 + ew) * CF_MAX_LEVELS + l] = make_int4(x0, x1, y0, y1);
            }
        };
        // epilogue thread 0: combine the four warps' boxes and hand the level table of block n to the producer / MMA warps
        auto publish = [&](long n, int buf) {
            const int slot = (int)(n & 1);
            if (n >= 2) mbar_wait(bar(kInfoEmpty + slot), (uint32_t)(((n >> 1) - 1) & 1));   // both consumers have read block n-2's table
            for (int l = 0; l < L; ++l) {
                int4 bx = reinterpret_cast<const int4*>(red)[(buf * 4 + 0) * CF_MAX_LEVELS + l];
                for (int w2 = 1; w2 < 4; ++w2) {
                    const int4 o = reinterpret_cast<const int4*>(red)[(buf * 4 + w2) * CF_MAX_LEVELS + l];
                    bx.x = min(bx.x, o.x); bx.y = max(bx.y, o.y); bx.z = min(bx.z, o.z); bx.w = max(bx.w, o.w);
                }
                LevelInfo li;
                const bool none = bx.x > bx.y || bx.z > bx.w;               // every query of the block is dead at this level
                li.bx0 = none ? 0 : bx.x;
                li.by0 = none ? 0 : bx.z;
                li.ncx = none ? 0 : (bx.y - bx.x) / CH_W + 1;               // the box is tiled by CH_H x CH_W chunks
                li.ncy = none ? 0 : (bx.w - bx.z) / CH_H + 1;
                info[slot * CF_MAX_LEVELS + l] = li;
            }
            mbar_arrive(bar(kInfoFull + slot));                             // release: the table is visible to the waiters
        };

        float cx = 0.0f, cy = 0.0f, ncx_ = 0.0f, ncy_ = 0.0f;
        if (nblk > 0) {
            load_coords(0, cx, cy);
            reduce_boxes(cx, cy, 0);
            epi_bar();
            if (et == 0) publish(0, 0);
        }
        long t = 0;
        for (long n = 0; n < nblk; ++n) {
            const int slot = (int)(n & 1);
            if (n + 1 < nblk) {
                load_coords(n + 1, ncx_, ncy_);
                reduce_boxes(ncx_, ncy_, (int)((n + 1) & 1));
            }
            epi_bar();      // every epilogue thread has left block n-1; block n's table (written before this barrier) is visible
            if (n + 1 < nblk && et == 0) publish(n + 1, (int)((n + 1) & 1));

            int b, qy0, qx0;
            block_origin(n, b, qy0, qx0);
            float* const ob = P.out + (size_t)b * L * K * HW + (size_t)(qy0 + qy) * P.W + (qx0 + qx);
            for (int l = 0; l < L; ++l) {
                const LevelInfo li = info[slot * CF_MAX_LEVELS + l];
                const int hl = P.lh[l], wl = P.lw[l];
                const float inv = __int_as_float((127 - l) << 23);
                const float lx = __fmul_rn(cx, inv), ly = __fmul_rn(cy, inv);
                const int ax = __float2int_rd(lx) - R - 1, ay = __float2int_rd(ly) - R - 1;   // window anchor: PR x PR cells from here
                // no tap pair of this query touches the image (also NaN / far-out coordinates): zeros, nothing to gather
                const bool dead = !(fabsf(lx) < kDeadCoord) || !(fabsf(ly) < kDeadCoord) || ax + PR - 1 < 0 || ax > wl - 1 ||
                                  ay + PR - 1 < 0 || ay > hl - 1;
                float* const ol = ob + (size_t)l * K * HW;
                // ---- gather: every chunk of the level's bounding box -> the cells of my window it covers ----
                for (int cyy = 0; cyy < li.ncy; ++cyy)
                    for (int cxx = 0; cxx < li.ncx; ++cxx, ++t) {
                        const int acc = (int)(t & 1);
                        mbar_wait(bar(kAccFull + acc), (uint32_t)((t >> 1) & 1));
                        tc_fence_after();
                        // chunk-local (row, col) -> window cell (row + dy, col + dx)
                        const int dx = li.bx0 + cxx * CH_W - ax, dy = li.by0 + cyy * CH_H - ay + half * (CH_H / EPT);
                        // bit c set: chunk column c lies inside my window
                        const int c_lo = max(0, -dx), c_hi = min(CH_W - 1, PR - 1 - dx);
                        const unsigned colmask = (dead || c_lo > c_hi) ? 0u : ((2u << c_hi) - 1u) & ~((1u << c_lo) - 1u);
                        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * NPOS + half * (NPOS / EPT));
                        float* const wbase = my + (dy * PR + dx) * NQ;
#pragma unroll 1
                        for (int h = 0; h < NPOS / (32 * EPT); ++h) {  // two chunk rows (32 positions) per TMEM load
                            float v[32];
                            tmem_ld32(taddr + (uint32_t)(32 * h), v);
                            tmem_wait_ld();
                            const bool r0 = (unsigned)(dy + 2 * h) < (unsigned)PR, r1 = (unsigned)(dy + 2 * h + 1) < (unsigned)PR;
                            // bit k: position k of this load (row k >> 4, column k & 15 of the chunk) lies inside my window
                            const unsigned m32 = (r0 ? colmask : 0u) | (r1 ? (colmask << 16) : 0u);
                            float* const w0 = wbase + (2 * h) * PR * NQ;
                            // the scale test is hoisted out of the 32 stores (left inside, the compiler re-tested it per element:
                            // 7 instructions per store instead of 3 -- a third of the kernel's instructions)
                            if (pow2_scale) {                  // a power-of-two scale is applied to the outputs instead
#pragma unroll
                                for (int k = 0; k < 32; ++k)
                                    if ((m32 >> k) & 1u) w0[((k >> 4) * PR + (k & 15)) * NQ] = v[k];
                            } else {
#pragma unroll
                                for (int k = 0; k < 32; ++k)
                                    if ((m32 >> k) & 1u) w0[((k >> 4) * PR + (k & 15)) * NQ] = __fdiv_rn(v[k], P.div);
                            }
                        }
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar(kAccEmpty + acc));       // TMEM buffer back to the MMA warp
                    }
                pair_bar();            // every thread of the query has written its rows of the window
                // ---- bilinear taps from the window: my output rows j (same arithmetic as the per-query kernels) ----
                if (dead) {
#pragma unroll 1
                    for (int k = half; k < K; k += EPT) __stcs(ol + (size_t)k * HW, 0.0f);
                } else {
                    const float s1x = (float)(wl - 1), s1y = (float)(hl - 1);
                    const float rx = P.rcp_w[l], ry = P.rcp_h[l];
                    int xo[SIDE];
                    float xl[SIDE], xh[SIDE];
#pragma unroll
                    for (int i = 0; i < SIDE; ++i) {
                        int i0;
                        axis_tap(lx, i - R, s1x, rx, i0, xl[i], xh[i]);
                        const int o = min(max(i0 - ax, 0), PR - 2);                   // as the per-query kernels clamp it
                        xo[i] = (min(max(ax + o, -2), wl) - ax) * NQ;                 // pairs left / right of the image: zero columns
                    }
#pragma unroll 1
                    for (int j = half; j < SIDE; j += EPT) {
                        int j0;
                        float wyl, wyh;
                        axis_tap(ly, j - R, s1y, ry, j0, wyl, wyh);
                        const int o = min(max(j0 - ay, 0), PR - 2);
                        const float* row = my + (min(max(ay + o, -2), hl) - ay) * (PR * NQ);
                        float* const oj = ol + (size_t)j * HW;
#pragma unroll
                        for (int i = 0; i < SIDE; ++i) {
                            const float* tp = row + xo[i];
                            float a = __fmul_rn(tp[0], __fmul_rn(xl[i], wyl));           // ATen order: nw, ne, sw, se
                            a = fmaf(tp[NQ], __fmul_rn(xh[i], wyl), a);
                            a = fmaf(tp[PR * NQ], __fmul_rn(xl[i], wyh), a);
                            a = fmaf(tp[(PR + 1) * NQ], __fmul_rn(xh[i], wyh), a);
                            __stcs(oj + (size_t)(i * SIDE) * HW, __fmul_rn(a, out_scale));       // * 1.0f is exact
                        }
                    }
                }
                pair_bar();            // the window may be overwritten by the next level's chunks
            }
            cx = ncx_;
            cy = ncy_;
        }
