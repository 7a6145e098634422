// Loader / epilogue role of corr_bwd_tc_kernel (included inside the kernel body; warps 2 .. 2 + LOADER_WARPS - 1).
        // =============================================================================================================
        // Group g (GROUP_WARPS warps) owns the K blocks kb = g (mod 2).  Its 32 * GROUP_WARPS threads cover one 128 x 64 tile of G:
        //   gemm 0 (rows = p, K = q): thread -> operand rows (tl >> 3) + 32 i, chunk tl & 7: two float4 loads, 8 lanes = 256 contiguous bytes;
        //   gemm 1 (rows = q, K = p): thread -> operand row tl & 127, chunks 4 (tl >> 7) + i: eight scalar loads down a column of G per chunk,
        //   lanes along q (coalesced).
        // Either way the thread ends up with CPT x (8 consecutive K values of one operand row) = CPT 16-byte chunks per precision part.
        const int lw = warp - 2, grp = lw / GROUP_WARPS, tl = (lw % GROUP_WARPS) * 32 + lane;      // tl: 0 .. 32 * GROUP_WARPS - 1
        constexpr int GT = 32 * GROUP_WARPS;             // threads of a group
        const float* const gb = P.g + (size_t)b * HW * HW;

        auto issue = [&](int kb, float* v) {
            const int k0 = kb * BK;
            if (gemm == 0) {
                const int chunk = tl & 7, k = k0 + 8 * chunk;
#pragma unroll
                for (int i = 0; i < CPT; ++i) {
                    const int m = m0 + (tl >> 3) + (GT / 8) * i;
                    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
                    if (m < HW && k < HW) {                  // HW % 8 == 0: a chunk is inside or outside as a whole
                        const float4* src = reinterpret_cast<const float4*>(gb + (size_t)m * HW + k);
                        a = __ldg(src);
                        c = __ldg(src + 1);
                    }
                    v[8 * i + 0] = a.x; v[8 * i + 1] = a.y; v[8 * i + 2] = a.z; v[8 * i + 3] = a.w;
                    v[8 * i + 4] = c.x; v[8 * i + 5] = c.y; v[8 * i + 6] = c.z; v[8 * i + 7] = c.w;
                }
            } else {
                const int m = m0 + (tl & 127), cbase = CPT * (tl >> 7);      // operand row, first chunk of this thread
                const bool row_in = m < HW;
                const float* src = gb + (size_t)(k0 + 8 * cbase) * HW + m;
                const unsigned uHW = (unsigned)HW;               // element offsets inside one pair's G fit 32 bits (HW <= 65535)
#pragma unroll
                for (int i = 0; i < CPT; ++i) {
                    const bool in = row_in && k0 + 8 * (cbase + i) < HW;
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[8 * i + j] = in ? __ldg(src + (unsigned)(8 * i + j) * uHW) : 0.0f;
                }
            }
        };
        auto commit = [&](int kb, const float* v) {
            const int s = kb % NSTAGE;
            mbar_wait(bar(kEmpty + s), (uint32_t)(((kb / NSTAGE) & 1) ^ 1));
            uint8_t* const ahi = gen + (size_t)s * STAGE_BYTES;
#pragma unroll
            for (int i = 0; i < CPT; ++i) {
                const int row = gemm == 0 ? (tl >> 3) + (GT / 8) * i : (tl & 127), chunk = gemm == 0 ? (tl & 7) : CPT * (tl >> 7) + i;
                uint4 hi, lo;
                split8(v + 8 * i, hi, lo);
                const uint32_t off = (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);   // SWIZZLE_128B, K-major
                *reinterpret_cast<uint4*>(ahi + off) = hi;
                *reinterpret_cast<uint4*>(ahi + TILE_BYTES + off) = lo;
            }
            fence_async_smem();          // generic-proxy stores -> visible to the tensor core's async-proxy reads
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kFull + s));
        };

        {
            float va[8 * CPT], vb[8 * CPT];
            int kb = grp;
            if (kb < KB) issue(kb, va);
            while (kb < KB) {
                if (kb + 2 < KB) issue(kb + 2, vb);
                commit(kb, va);
                kb += 2;
                if (kb >= KB) break;
                if (kb + 2 < KB) issue(kb + 2, va);
                commit(kb, vb);
                kb += 2;
            }
        }

        // ===================== epilogue: D[spatial, channel] -> out[b, channel, spatial] / sqrt(C) =====================
        mbar_wait(bar(kAcc), 0u);
        tc_fence_after();
        {
            const int quad = warp & 3;                       // the TMEM lane quadrant this warp may read
            const int m = m0 + quad * 32 + lane;
            float* const o = P.out[gemm] + (size_t)b * P.C * HW + m;
#pragma unroll 1
            constexpr int CW = BN / (LOADER_WARPS / 4);        // columns per warp: the warps of a lane quadrant split the 128 channels
            for (int c0 = (lw >> 2) * CW; c0 < (lw >> 2) * CW + CW; c0 += 32) {
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int c = n0 + c0 + j;
                    if (m < HW && c < P.C) o[(size_t)c * HW] = __fdiv_rn(v[j], P.div);
                }
            }
        }
