// dh_tiles_kernel.cuh — the OBJECT-HEAVY renderer: register tiles instead of staged rows.
// Included by dh_kernels.cu after dh_groups_kernel.cuh (shares ObjI, unit_header, scan_slot, Digest, the CLC / mbarrier /
// bulk-store wrappers).
//
// Same result as the reference's per-object max-blend (draw_heatmap_cuda_kernel.cuh:26-49, 76-109), same arithmetic per
// pixel as groups_kernel (m = (dy^2 + dx^2) * (1/var) with explicit IEEE roundings, cuh:45; max_o k*expf(-m_o) ==
// k*expf(-min_o m_o) for k >= 0), for planes on which HUNDREDS of objects overlap (BASELINE cfg4: 1024 objects of
// radius 16-64 on a 512x512 plane, every pixel under ~29 of them). There the launch is bound by instruction issue, not
// by stores (ncu, profiles/r02/ncu_groups_kernel_cfg4.txt: issue-active 84 %, DRAM 0.07 %), and groups_kernel spends
// ~30 warp instructions per 32 object-pixels: every object is swept into shared-memory staging rows with a
// read-modify-write per pixel quad, two rows at a time, and every row of every object pays the row set-up again.
//
// Here a unit is a STRIP of kTRows = 8 whole rows of one plane, cut into tiles of 8 x 64 pixels, one tile per warp at a
// time. A lane owns two columns of the tile (x, x + 32) and keeps the running minimum of m for its 16 pixels in
// REGISTERS (as raw float bits: m >= 0, so unsigned order == float order; 0xFFFFFFFF = untouched). Per object and tile
// what depends on the column (dx, dx^2, |dx| <= r) is computed once, what depends on the row (dy^2, |dy| <= r) once per
// row for both columns: ~10 instructions per row of 64 pixels, ~96 per (object, tile) — about a third of what the
// staged sweep issues. No shared-memory traffic in the sweep except one broadcast 16-byte load per (object, tile).
// The strip's object list is built by the whole CTA exactly as in groups_kernel (scan of the sample's slots, K1 fused,
// ballot + popc compaction), in passes of kTCap slots, so any T is served without an overflow path: the accumulators
// simply persist over the passes. One expf per touched pixel at the end; FRESH writes every pixel of the tile (value or
// zero) with coalesced 128-byte stores, ACCUMULATE read-max-writes the touched ones (cuh:30). A strip without any object
// leaves as bulk stores of the zero tile.
// Selected by the host for T >= 256 and k >= 0 (dh_kernels.cu launch_render); k < 0 keeps the staged renderer, whose
// order-preserving encode covers that sign.
#pragma once

namespace dh {

constexpr int kTRows = 8;     // rows of a strip / tile
constexpr int kTCols = 64;    // columns of a tile (two per lane)
constexpr int kTCap = 1024;   // slots scanned, and list entries held, per pass
constexpr int kTMinB = 4;     // resident CTAs per SM the register budget is cut for

__host__ __device__ constexpr size_t tiles_smem_bytes(bool accum) {
    return (size_t)(accum ? 0 : kGZeroBytes) + (size_t)kTCap * sizeof(ObjI);
}

// All objects of `list[0, cnt)` that overlap the tile's columns [tx0, tx0 + 64) into the lane's 8 x 2 accumulators.
// Every list entry already overlaps the strip's rows (unit cull) and the map's columns. Warp-collective.
__device__ __forceinline__ void sweep_tile(const ObjI* __restrict__ list, int cnt, int tx0, int ty0, int lane,
                                           unsigned (&acc)[kTRows][2]) {
    const int xl = tx0 + lane;  // my first column
    for (int base = 0; base < cnt; base += 32) {
        const int e = base + lane;
        bool hit = false;
        if (e < cnt) {
            const ObjI o = list[e];
            hit = (o.x + o.r >= tx0) && (o.x - o.r < tx0 + kTCols);
        }
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        while (bal) {
            const int src = __ffs(bal) - 1;
            bal &= bal - 1u;
            const ObjI o = list[base + src];  // broadcast
            const float fr = (float)o.r;
            // |x - o.x| and |y - o.y| stay below 2^24 for every object that passed the cull: exact as floats
            const float fdx0 = (float)(xl - o.x), fdx1 = fdx0 + 32.0f;
            const float dx20 = __fmul_rn(fdx0, fdx0), dx21 = __fmul_rn(fdx1, fdx1);
            const bool in0 = fabsf(fdx0) <= fr, in1 = fabsf(fdx1) <= fr;
            float fdy = (float)(ty0 - o.y);
#pragma unroll
            for (int rr = 0; rr < kTRows; ++rr) {
                const float fi = __fmul_rn(fdy, fdy);
                const bool row = fabsf(fdy) <= fr;
                const unsigned m0 = __float_as_uint(__fmul_rn(__fadd_rn(fi, dx20), o.w));
                const unsigned m1 = __float_as_uint(__fmul_rn(__fadd_rn(fi, dx21), o.w));
                if (row && in0) acc[rr][0] = min(acc[rr][0], m0);
                if (row && in1) acc[rr][1] = min(acc[rr][1], m1);
                fdy += 1.0f;
            }
        }
    }
}

template <bool ACCUM, int SRC, bool DIGEST>
__global__ void __launch_bounds__(kGThreads, kTMinB) tiles_kernel(const RenderParams p) {
    static_assert(SRC == SRC_BATCHED || SRC == SRC_BOXES, "tiles_kernel serves the RaggedBatch sources");
    static_assert(!(ACCUM && DIGEST), "the fused digest exists in FRESH mode only");
    constexpr int ZB = ACCUM ? 0 : kGZeroBytes;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* const zero_tile = reinterpret_cast<float4*>(smem_raw);
    ObjI* const s_list = reinterpret_cast<ObjI*>(smem_raw + ZB);
    __shared__ int s_cnt2[2];                          // list size, double-buffered by pass parity (no reset barrier)
    __shared__ __align__(16) uint4 s_clc;              // cluster-launch-control response
    __shared__ __align__(8) unsigned long long s_bar;  // its mbarrier

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = p.W;
    const float k = p.k;
    const long long plane_elems = (long long)p.H * W;
    const int tiles_x = (W + kTCols - 1) / kTCols;
    for (int i = tid; i < ZB / 16; i += kGThreads) zero_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        s_cnt2[0] = s_cnt2[1] = 0;
    }
    fence_proxy_async();
    pdl_wait();  // nothing global before this point (see groups_kernel)
    pdl_launch_dependents();

    long long u = blockIdx.x;
    unsigned par = 0, clc_phase = 0;
    Digest dg;
    for (;;) {
        UnitSrc us;
        int uy0, uy1;
        bool count_skips;
        const int nscan = unit_header<SRC>(p, u, kTRows, us, uy0, uy1, count_skips, nullptr);
        float* const plane_f = p.heat + us.plane * plane_elems;
        const long long index0 = p.digest_base + us.plane * plane_elems;
        bool asked = false;
        for (int t0 = 0; t0 < tiles_x; t0 += kGWarps) {  // one tile per warp and round (one round for W <= 512)
            const int tile = t0 + warp;
            const int tx0 = tile * kTCols;
            unsigned acc[kTRows][2];
#pragma unroll
            for (int rr = 0; rr < kTRows; ++rr) acc[rr][0] = acc[rr][1] = 0xffffffffu;
            bool any = false;  // the strip holds at least one object (same value in every thread)
            for (int s0 = 0; s0 < nscan || s0 == 0; s0 += kTCap, par ^= 1u) {
                __syncthreads();  // everybody is done with the previous pass's list
                if (!asked) {
                    asked = true;
                    if (tid == 0 && p.use_clc) {  // ask for the next unit now; the answer lands while this one is rendered
                        fence_proxy_async();
                        mbar_expect_tx(&s_bar, 16);
                        clc_try_cancel(&s_clc, &s_bar);
                    }
                }
                int& s_cnt = s_cnt2[par];
                const int send = min(nscan, s0 + kTCap);
                for (int base = s0; base < send; base += kGThreads) {
                    ObjI o;
                    const int t = base + tid;
                    const bool pred = (t < send) && scan_slot<SRC>(p, us, t, nscan, uy0, uy1, W, count_skips && t0 == 0, o);
                    const unsigned bal = __ballot_sync(0xffffffffu, pred);
                    if (bal) {
                        int pos = 0;
                        if (lane == 0) pos = atomicAdd(&s_cnt, __popc(bal));
                        pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(bal & ((1u << lane) - 1u));
                        if (pred) s_list[pos] = o;  // pos < kTCap: a pass scans at most kTCap slots
                    }
                }
                __syncthreads();
                const int cnt = s_cnt;
                if (tid == 0) s_cnt2[par ^ 1u] = 0;  // last read before this pass's first barrier
                any |= cnt > 0;
                if (tile < tiles_x && cnt > 0) sweep_tile(s_list, cnt, tx0, uy0, lane, acc);
            }
            if constexpr (!ACCUM) {
                if (!any) {  // nothing on the strip: it leaves from the zero tile, no per-pixel work (and no further round)
                    if (tid == 0) {
                        char* dst = reinterpret_cast<char*>(plane_f + (long long)uy0 * W);
                        long long bytes = (long long)(uy1 - uy0) * W * 4;
                        while (bytes > 0) {
                            const unsigned c = (unsigned)(bytes < kGZeroBytes ? bytes : kGZeroBytes);
                            bulk_store(dst, zero_tile, c);
                            dst += c;
                            bytes -= c;
                        }
                        bulk_commit();
                    }
                    break;
                }
            } else {
                if (!any) break;
            }
            if (tile < tiles_x) {  // K4: one expf per touched pixel (cuh:45), then the blend
                const int xl = tx0 + lane;
#pragma unroll
                for (int rr = 0; rr < kTRows; ++rr) {
                    const int y = uy0 + rr;
                    if (y < uy1) {
                        float* const row = plane_f + (long long)y * W;
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            const int x = xl + 32 * c;
                            if (x < W) {
                                const unsigned t = acc[rr][c];
                                if constexpr (ACCUM) {
                                    if (t != 0xffffffffu) {
                                        const float v = __fmul_rn(expf(-__uint_as_float(t)), k);
                                        const float g = row[x];
                                        if (!(g >= v)) row[x] = v;  // cuh:30: keep old iff old >= v
                                    }
                                } else {
                                    float f = 0.0f;
                                    if (t != 0xffffffffu) {
                                        const float v = __fmul_rn(expf(-__uint_as_float(t)), k);
                                        f = (0.0f >= v) ? 0.0f : v;  // max-blend into the fused zero-init
                                        if constexpr (DIGEST) dg.add(f, index0 + (long long)y * W + x);
                                    }
                                    row[x] = f;
                                }
                            }
                        }
                    }
                }
            }
        }
        // next unit: the one the block scheduler hands over (cluster launch control), else the static stride
        if (p.use_clc) {
            mbar_wait(&s_bar, clc_phase);
            clc_phase ^= 1u;
            unsigned nx;
            if (!clc_response(&s_clc, nx)) break;
            u = nx;
        } else {
            u += gridDim.x;
            if (u >= p.total_units) break;
        }
    }
    if constexpr (!ACCUM) bulk_wait_read<0>();  // the zero tile must outlive every bulk read issued by this thread
    if constexpr (DIGEST) digest_flush(dg, p.digest, tid, lane, warp);
}

}  // namespace dh
