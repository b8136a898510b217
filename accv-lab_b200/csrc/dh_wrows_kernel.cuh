// dh_wrows_kernel.cuh — warp-autonomous variant of the rows renderer for maps up to 512 pixels wide (every BASELINE
// config). Included by dh_kernels.cu after dh_rows_kernel.cuh (shares ObjI, load_slot, make_obji, bulk-copy wrappers).
//
// Same output partitioning as rows_kernel (units of <= 64 bands, shared-memory unit list built with warp ballot +
// popc prefix, 64-bit band-occupancy mask, empty bands leave as bulk stores of a constant zero tile), but inside an
// occupied band there is NO block-wide synchronisation: ncu showed rows_kernel stalled on its 4 barriers per band
// (profiles/r01_rows_kernel_cfg5.txt: stall_barrier 4.4 warps/issue). Here
//   * a band is 8 row groups, a row group is G whole rows (<= 2 KB, one contiguous byte range), warp w owns group w;
//   * each warp has two private 2 KB staging buffers; it scans the unit list 32 objects at a time (one per lane,
//     ballot), renders the hits of its own rows (scalar sweep for narrow spans, 128-bit sweep for spans > 32 px),
//     finalises the span union with one expf per touched pixel, fences to the async proxy and issues its OWN
//     cp.async.bulk store; the next group renders into the other buffer while the copy engine drains the first;
//   * a group nobody overlaps is stored straight from the zero tile (no staging at all);
//   * the only block barriers left are the three around the per-unit list build.
#pragma once

namespace dh {

constexpr int kWThreads = 256;
constexpr int kWWarps = kWThreads / 32;
constexpr int kWMaxW = 512;                  // widest map served by this kernel
constexpr int kWGroupBytes = kWMaxW * 4;     // staging buffer of one row group
constexpr int kWZeroBytes = 8192;            // constant zero tile
constexpr int kWCap = 512;                   // unit list capacity
constexpr size_t kWSmemBytes = (size_t)kWZeroBytes + (size_t)kWWarps * 2 * kWGroupBytes + (size_t)kWCap * sizeof(ObjI);

__device__ __forceinline__ unsigned enc_m(float m, bool kneg) {
    const unsigned b = __float_as_uint(m);
    return kneg ? b + 1u : ~b;
}

// One object into the rows [y0, y0+nrows) held in `buf` (row stride W). Warp-collective.
__device__ __forceinline__ void render_obj(unsigned* __restrict__ buf, int y0, int nrows, int W, const ObjI& o, bool kneg,
                                           int lane, int& ux0, int& ux1) {
    const int x0 = max(o.x - o.r, 0), x1 = min(o.x + o.r, W - 1);
    ux0 = min(ux0, x0);
    ux1 = max(ux1, x1);
    for (int rr = 0; rr < nrows; ++rr) {
        const int dy = y0 + rr - o.y;
        if (abs(dy) > o.r) continue;  // warp-uniform
        unsigned* const rowp = buf + rr * W;
        const float fdy = (float)dy;
        const float fi = __fmul_rn(fdy, fdy);
        if (x1 - x0 < 32) {  // narrow span: one pixel per lane
            const int x = x0 + lane;
            if (x <= x1) {
                const float fdx = (float)(x - o.x);
                const unsigned t = enc_m(__fmul_rn(__fadd_rn(fi, __fmul_rn(fdx, fdx)), o.w), kneg);
                rowp[x] = max(rowp[x], t);
            }
        } else {  // wide span: four pixels per lane, 128-bit shared-memory accesses
            for (int xv = (x0 & ~3) + 4 * lane; xv <= x1; xv += 128) {
                uint4 q = *reinterpret_cast<uint4*>(rowp + xv);
                unsigned* qc = reinterpret_cast<unsigned*>(&q);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int x = xv + c;
                    const float fdx = (float)(x - o.x);
                    const unsigned t = enc_m(__fmul_rn(__fadd_rn(fi, __fmul_rn(fdx, fdx)), o.w), kneg);
                    if (x >= x0 && x <= x1) qc[c] = max(qc[c], t);
                }
                *reinterpret_cast<uint4*>(rowp + xv) = q;
            }
        }
    }
}

template <bool ACCUM, int SRC>
__global__ void __launch_bounds__(kWThreads, 4) wrows_kernel(const RenderParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* const zero_tile = reinterpret_cast<float4*>(smem_raw);
    ObjI* const s_unit = reinterpret_cast<ObjI*>(smem_raw + kWZeroBytes + kWWarps * 2 * kWGroupBytes);
    __shared__ int s_cnt;
    __shared__ unsigned long long s_mask;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned* const wbuf0 = reinterpret_cast<unsigned*>(smem_raw + kWZeroBytes + warp * 2 * kWGroupBytes);
    unsigned* const wbuf1 = wbuf0 + kWGroupBytes / 4;
    const bool kneg = p.k < 0.0f;
    const int W = p.W, H = p.H, R = p.band_rows, G = p.band_cols /* rows per group */, w4 = W >> 2;
    const long long plane_elems = (long long)H * W;

    if constexpr (!ACCUM) {
        for (int i = tid; i < kWZeroBytes / 16; i += kWThreads) zero_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        fence_proxy_async();
    }
    int cur = 0;

    for (long long u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const long long plane = u / p.units_per_plane;
        const int ug = (int)(u - plane * p.units_per_plane);
        float* const plane_ptr = p.heat + plane * plane_elems;
        const int b0 = ug * p.bands_per_unit, b1 = min(b0 + p.bands_per_unit, p.bands_y), nb = b1 - b0;
        const int uy0 = b0 * R, uy1 = min(b1 * R, H);

        long long src_base = 0;
        int n = 0, cls = 0;
        bool count_skips = false;
        if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
            const long long sample = plane / p.C;
            cls = (int)(plane - sample * p.C);
            long long ns = p.sizes_i64 ? __ldg(reinterpret_cast<const long long*>(p.sizes) + sample)
                                       : (long long)__ldg(reinterpret_cast<const int*>(p.sizes) + sample);
            ns = ns < 0 ? 0 : (ns > p.T ? p.T : ns);
            n = (int)ns;
            src_base = sample * p.T;
            count_skips = (ug == 0 && cls == 0);
        } else if constexpr (SRC == SRC_CSR) {
            const int o0 = p.csr_off[plane], o1 = p.csr_off[plane + 1];
            src_base = o0;
            n = o1 - o0;
            count_skips = (ug == 0);
        } else {
            n = (int)p.num_objects;
            count_skips = (u == 0);
        }

        // ---- K2: unit list (ballot + popc-prefix compaction) and band-occupancy mask ------------------------
        __syncthreads();  // every warp is done with the previous unit's list (and the zero tile is initialised)
        if (tid == 0) {
            s_cnt = 0;
            s_mask = 0ull;
        }
        __syncthreads();
        for (int base = 0; base < n; base += kWThreads) {
            const int t = base + tid;
            ObjI o;
            bool pred = false;
            if (t < n) {
                int x, y, r;
                if (load_slot<SRC>(p, src_base, t, cls, plane, count_skips, x, y, r))
                    pred = make_obji(x, y, r, p.factor, uy0, uy1, W, o, p.status, count_skips);
            }
            unsigned long long bits = 0ull;
            if (pred) {
                const int lo = max(o.y - o.r, uy0) / R - b0, hi = min(o.y + o.r, uy1 - 1) / R - b0;
                bits = ((2ull << (hi - lo)) - 1ull) << lo;
            }
            const unsigned blo = __reduce_or_sync(0xffffffffu, (unsigned)bits);
            const unsigned bhi = __reduce_or_sync(0xffffffffu, (unsigned)(bits >> 32));
            if (lane == 0 && (blo | bhi)) atomicOr(&s_mask, ((unsigned long long)bhi << 32) | blo);
            const unsigned bal = __ballot_sync(0xffffffffu, pred);
            if (bal) {
                int pos = 0;
                if (lane == 0) pos = atomicAdd(&s_cnt, __popc(bal));
                pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(bal & ((1u << lane) - 1u));
                if (pred && pos < kWCap) s_unit[pos] = o;
            }
        }
        __syncthreads();
        const int ucnt = s_cnt;
        unsigned long long mask = s_mask;
        const bool overflow = ucnt > kWCap;  // rare: warps then scan the source instead of the shared list
        const int lim = overflow ? n : ucnt;

        // ---- FRESH: empty bands leave as bulk stores of the zero tile, spread over the warps ------------------
        if constexpr (!ACCUM) {
            if (lane == 0) {
                for (int b = warp; b < nb; b += kWWarps) {
                    if (!((mask >> b) & 1ull)) {
                        const int by0 = (b0 + b) * R;
                        char* dst = reinterpret_cast<char*>(plane_ptr + (long long)by0 * W);
                        int bytes = min(R, H - by0) * W * 4;
                        while (bytes > 0) {
                            const int c = min(bytes, kWZeroBytes);
                            bulk_store(dst, zero_tile, (unsigned)c);
                            dst += c;
                            bytes -= c;
                        }
                    }
                }
                bulk_commit();
            }
        }

        // ---- occupied bands: warp `warp` owns row group `warp` of every band, no block barrier ------------------
        while (mask) {
            const int b = __ffsll((long long)mask) - 1;
            mask &= mask - 1ull;
            const int y0 = (b0 + b) * R + warp * G;
            if (y0 >= min((b0 + b + 1) * R, H)) continue;
            const int nrows = min(G, H - y0), y1 = y0 + nrows - 1;
            float* const gdst = plane_ptr + (long long)y0 * W;
            unsigned* const buf = cur ? wbuf1 : wbuf0;
            if constexpr (!ACCUM) {
                if (lane == 0) bulk_wait_read<1>();  // the store that last read `buf` is older than my latest group
                __syncwarp();
            }
            bool any = false;
            int ux0 = W, ux1 = -1;
            for (int base = 0; base < lim; base += 32) {
                const int e = base + lane;
                ObjI o;
                o.x = o.y = o.r = 0;
                o.w = 0.f;
                bool hit = false;
                if (e < lim) {
                    if (!overflow) {
                        o = s_unit[e];
                        hit = (o.y + o.r >= y0) && (o.y - o.r <= y1);
                    } else {
                        int x, y, r;
                        if (load_slot<SRC>(p, src_base, e, cls, plane, false, x, y, r))
                            hit = make_obji(x, y, r, p.factor, y0, y1 + 1, W, o, nullptr, false);
                    }
                }
                unsigned bal = __ballot_sync(0xffffffffu, hit);
                if (bal && !any) {  // lazy zero fill of the staging rows
                    any = true;
                    uint4* b4 = reinterpret_cast<uint4*>(buf);
                    for (int i = lane; i < nrows * w4; i += 32) b4[i] = make_uint4(0u, 0u, 0u, 0u);
                    __syncwarp();
                }
                while (bal) {
                    const int src = __ffs(bal) - 1;
                    bal &= bal - 1u;
                    ObjI ob;
                    ob.x = __shfl_sync(0xffffffffu, o.x, src);
                    ob.y = __shfl_sync(0xffffffffu, o.y, src);
                    ob.r = __shfl_sync(0xffffffffu, o.r, src);
                    ob.w = __shfl_sync(0xffffffffu, o.w, src);
                    render_obj(buf, y0, nrows, W, ob, kneg, lane, ux0, ux1);
                    __syncwarp();
                }
            }
            if (any) {
                for (int rr = 0; rr < nrows; ++rr)
                    finalize_row<ACCUM>(buf + rr * W, gdst + (long long)rr * W, ux0, ux1, kneg, p.k, lane);
                if constexpr (!ACCUM) {
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        bulk_store(gdst, buf, (unsigned)(nrows * W * 4));
                        bulk_commit();
                    }
                    cur ^= 1;
                } else {
                    __syncwarp();
                }
            } else if constexpr (!ACCUM) {
                if (lane == 0) {
                    bulk_store(gdst, zero_tile, (unsigned)(nrows * W * 4));
                    bulk_commit();
                }
            }
        }
    }
    if constexpr (!ACCUM) bulk_wait_read<0>();
}

}  // namespace dh
