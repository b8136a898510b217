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
//
// Two sweeps. sweep_tile (the first one written, kept as the exact path for unusual objects) predicates every pixel:
// ~11 warp instructions per (object, row of 64 pixels). sweep_tile_fast serves "plain" objects — r <= kTExactR, so
// dy^2 + dx^2 < 2^24 is an exact integer in fp32 however it is formed, and 0 < 1/var <= 1e30, so no in-window m
// overflows — and replaces predication by arithmetic: a pixel outside the object's columns gets dx^2 = +inf, a row
// outside its rows dy = +inf, hence m = +inf, which never beats a real value and doubles as "untouched". What is left
// per (object, row) is FADD2 (dy, dy) + (1, 1), FFMA2 dy * dy + (dx0^2, dx1^2), FMUL2 by (w, w) — the packed fp32
// instructions of sm_100, each component rounded as the scalar instruction would — and, objects taken two at a time,
// one three-input FMNMX3 per pixel pair of objects: 4 instead of 11 issue slots. fma(dy, dy, dx^2) equals the
// reference's (float)(i*i) + (float)(j*j) (cuh:45) bit for bit because every term is an exact integer below 2^24.
// A pass whose list holds an object that is not plain (block-uniform flag set by the scan) switches the tile to the
// predicated sweep for the rest of the unit; the accumulators' "untouched" code is converted once (+inf -> all ones).
#pragma once

namespace dh {

#ifndef DH_T_ROWS
#define DH_T_ROWS 8
#endif
#ifndef DH_T_MINB
#define DH_T_MINB 4
#endif
constexpr int kTRows = DH_T_ROWS;  // rows of a strip / tile
constexpr int kTCols = 64;    // columns of a tile (two per lane)
constexpr int kTCap = 1024;   // slots scanned, and list entries held, per pass
constexpr int kTMinB = DH_T_MINB;  // resident CTAs per SM the register budget is cut for
constexpr int kTExactR = 2896;  // 2 * 2896^2 < 2^24: squared distances of in-window pixels are exact fp32 integers
constexpr unsigned kTNone = 0xffffffffu;      // untouched, predicated sweep (m >= 0: never a real value's bits)
constexpr unsigned kTNoneFast = 0x7f800000u;  // untouched, arithmetic sweep (+inf)

__device__ __forceinline__ bool tile_obj_plain(const ObjI& o) { return o.r <= kTExactR && o.w > 0.0f && o.w <= 1e30f; }

// packed fp32 pairs (sm_100 FADD2 / FMUL2 / FFMA2): both components rounded to nearest even like the scalar forms
__device__ __forceinline__ void fadd2(float& x, float& y, float a, float b) {
    asm("{.reg .b64 p, q, r; mov.b64 p, {%0, %1}; mov.b64 q, {%2, %3}; add.rn.f32x2 r, p, q; mov.b64 {%0, %1}, r;}"
        : "+f"(x), "+f"(y) : "f"(a), "f"(b));
}
__device__ __forceinline__ void fmul2(float& x, float& y, float a, float b) {
    asm("{.reg .b64 p, q, r; mov.b64 p, {%0, %1}; mov.b64 q, {%2, %3}; mul.rn.f32x2 r, p, q; mov.b64 {%0, %1}, r;}"
        : "+f"(x), "+f"(y) : "f"(a), "f"(b));
}
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
    asm("{.reg .b64 p, q, r, s; mov.b64 p, {%2, %3}; mov.b64 q, {%4, %5}; mov.b64 r, {%6, %7}; fma.rn.f32x2 s, p, q, r;"
        " mov.b64 {%0, %1}, s;}" : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// A list entry of the strip: the object in the form both sweeps consume. All four are exact: |x| <= r + W < 2^24 and
// |y0 - y| <= r + kTRows for every object that passed the cull (r <= DH_MAX_RADIUS = 2^22).
struct __align__(16) TileEnt {
    float fx;   // centre column
    float fr;   // radius
    float fdy;  // first row of the strip - centre row
    float w;    // 1 / var
};
static_assert(sizeof(TileEnt) == 16, "one 128-bit shared-memory load per entry");

__host__ __device__ constexpr size_t tiles_queue_bytes() { return (size_t)kGWarps * (kTCap + 2) * sizeof(unsigned short); }
__host__ __device__ constexpr size_t tiles_smem_bytes(bool accum) {
    return (size_t)(accum ? 0 : kGZeroBytes) + (size_t)kTCap * sizeof(TileEnt) + tiles_queue_bytes();
}

// All objects of `list[0, cnt)` that overlap the tile's columns [tx0, tx0 + 64) into the lane's 8 x 2 accumulators
// (raw bits of m, kTNone = untouched): every pixel predicated, any object. Every list entry already overlaps the
// strip's rows (unit cull) and the map's columns. Warp-collective.
__device__ __forceinline__ void sweep_tile(const TileEnt* __restrict__ list, int cnt, int tx0, int lane,
                                           float (&acc)[kTRows][2]) {
    const float xlf = (float)(tx0 + lane);  // my first column
    const float t0f = (float)tx0, t1f = (float)(tx0 + kTCols);
    for (int base = 0; base < cnt; base += 32) {
        const int e = base + lane;
        bool hit = false;
        if (e < cnt) {
            const float2 xr = *reinterpret_cast<const float2*>(&list[e]);
            hit = (xr.x + xr.y >= t0f) && (xr.x - xr.y < t1f);
        }
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        while (bal) {
            const int src = __ffs(bal) - 1;
            bal &= bal - 1u;
            const TileEnt o = list[base + src];  // broadcast
            const float fdx0 = xlf - o.fx, fdx1 = fdx0 + 32.0f;
            const float dx20 = __fmul_rn(fdx0, fdx0), dx21 = __fmul_rn(fdx1, fdx1);
            const bool in0 = fabsf(fdx0) <= o.fr, in1 = fabsf(fdx1) <= o.fr;
            float fdy = o.fdy;
#pragma unroll
            for (int rr = 0; rr < kTRows; ++rr) {
                const float fi = __fmul_rn(fdy, fdy);
                const bool row = fabsf(fdy) <= o.fr;
                const unsigned m0 = __float_as_uint(__fmul_rn(__fadd_rn(fi, dx20), o.w));
                const unsigned m1 = __float_as_uint(__fmul_rn(__fadd_rn(fi, dx21), o.w));
                if (row && in0) acc[rr][0] = __uint_as_float(min(__float_as_uint(acc[rr][0]), m0));
                if (row && in1) acc[rr][1] = __uint_as_float(min(__float_as_uint(acc[rr][1]), m1));
                fdy += 1.0f;
            }
        }
    }
}

// What sweep_pair keeps of an object for one tile: masked dx^2 of the lane's two columns, dy of the tile's first row,
// 1/var, the radius (row mask).
struct TileObj {
    float a0, a1, fdy, w, fr;
    bool allrows;  // every row of the tile lies inside the object's rows (warp-uniform)
};

__device__ __forceinline__ TileObj tile_obj(const TileEnt& o, float xlf) {
    TileObj t;
    const float inf = __uint_as_float(kTNoneFast);
    const float fdx0 = xlf - o.fx, fdx1 = fdx0 + 32.0f;
    t.a0 = fabsf(fdx0) <= o.fr ? __fmul_rn(fdx0, fdx0) : inf;
    t.a1 = fabsf(fdx1) <= o.fr ? __fmul_rn(fdx1, fdx1) : inf;
    t.fdy = o.fdy;
    t.w = o.w;
    t.fr = o.fr;
    t.allrows = (o.fdy >= -o.fr) && (o.fdy + (float)(kTRows - 1) <= o.fr);
    return t;
}

template <bool MASKROWS>
__device__ __forceinline__ void sweep_pair(const TileObj& A, const TileObj& B, float (&acc)[kTRows][2]) {
    const float inf = __uint_as_float(kTNoneFast);
    float fa0 = A.fdy, fa1 = A.fdy, fb0 = B.fdy, fb1 = B.fdy;
#pragma unroll
    for (int rr = 0; rr < kTRows; ++rr) {
        float ya0 = fa0, ya1 = fa1, yb0 = fb0, yb1 = fb1;
        if constexpr (MASKROWS) {  // a row outside the object: dy = +inf, so m = +inf
            const bool ra = fabsf(fa0) <= A.fr, rb = fabsf(fb0) <= B.fr;
            ya0 = ra ? fa0 : inf;
            ya1 = ra ? fa1 : inf;
            yb0 = rb ? fb0 : inf;
            yb1 = rb ? fb1 : inf;
        }
        float sa0, sa1, sb0, sb1;
        ffma2(sa0, sa1, ya0, ya1, ya0, ya1, A.a0, A.a1);  // dy^2 + dx^2: exact (tile_obj_plain)
        ffma2(sb0, sb1, yb0, yb1, yb0, yb1, B.a0, B.a1);
        fmul2(sa0, sa1, A.w, A.w);
        fmul2(sb0, sb1, B.w, B.w);
        acc[rr][0] = fmin3(acc[rr][0], sa0, sb0);
        acc[rr][1] = fmin3(acc[rr][1], sa1, sb1);
        if (rr + 1 < kTRows) {
            fadd2(fa0, fa1, 1.0f, 1.0f);
            fadd2(fb0, fb1, 1.0f, 1.0f);
        }
    }
}

// Same contract as sweep_tile for lists of plain objects only (tile_obj_plain); accumulators hold m as floats, +inf =
// untouched. First the list entries that overlap the tile's columns are compacted into the warp's queue `q` (ballot +
// popc prefix, 16-bit entry indices), then the queue is swept two objects at a time — no bit scanning and no carried
// state between the objects; an odd queue ends with its last object twice (min is idempotent).
__device__ __forceinline__ void sweep_tile_fast(const TileEnt* __restrict__ list, int cnt, int tx0, int lane,
                                                unsigned short* __restrict__ q, float (&acc)[kTRows][2]) {
    const float xlf = (float)(tx0 + lane);
    const float t0f = (float)tx0, t1f = (float)(tx0 + kTCols);
    int n = 0;  // warp-uniform
    for (int base = 0; base < cnt; base += 32) {
        const int e = base + lane;
        bool hit = false;
        if (e < cnt) {
            const float2 xr = *reinterpret_cast<const float2*>(&list[e]);
            hit = (xr.x + xr.y >= t0f) && (xr.x - xr.y < t1f);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (hit) q[n + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)e;
        n += __popc(bal);
    }
    __syncwarp();
    if ((n & 1) && lane == 0) q[n] = q[n - 1];
    __syncwarp();
    const unsigned* const q2 = reinterpret_cast<const unsigned*>(q);
    for (int i = 0; i < n; i += 2) {
        const unsigned pr = q2[i >> 1];  // broadcast
        const TileObj A = tile_obj(list[pr & 0xffffu], xlf), B = tile_obj(list[pr >> 16], xlf);
        if (A.allrows && B.allrows) sweep_pair<false>(A, B, acc);
        else sweep_pair<true>(A, B, acc);
    }
    __syncwarp();  // the queue is rewritten by the next pass
}

template <bool ACCUM, int SRC, bool DIGEST>
__global__ void __launch_bounds__(kGThreads, kTMinB) tiles_kernel(const RenderParams p) {
    static_assert(SRC == SRC_BATCHED || SRC == SRC_BOXES, "tiles_kernel serves the RaggedBatch sources");
    static_assert(!(ACCUM && DIGEST), "the fused digest exists in FRESH mode only");
    constexpr int ZB = ACCUM ? 0 : kGZeroBytes;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* const zero_tile = reinterpret_cast<float4*>(smem_raw);
    TileEnt* const s_list = reinterpret_cast<TileEnt*>(smem_raw + ZB);
    unsigned short* const s_queue = reinterpret_cast<unsigned short*>(smem_raw + ZB + (size_t)kTCap * sizeof(TileEnt));
    __shared__ int s_cnt2[2];                          // list size, double-buffered by pass parity (no reset barrier)
    __shared__ int s_odd2[2];                          // the pass's list holds an object that is not tile_obj_plain
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
        s_odd2[0] = s_odd2[1] = 0;
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
            float acc[kTRows][2];  // min m per pixel: floats while every object was plain, raw bits after that
#pragma unroll
            for (int rr = 0; rr < kTRows; ++rr) acc[rr][0] = acc[rr][1] = __uint_as_float(kTNoneFast);
            bool any = false;   // the strip holds at least one object (same value in every thread)
            bool slow = false;  // a pass held an object that is not plain: predicated sweep from then on (same in every thread)
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
                        if (pred) {  // pos < kTCap: a pass scans at most kTCap slots
                            s_list[pos] = TileEnt{(float)o.x, (float)o.r, (float)(uy0 - o.y), o.w};
                            if (!tile_obj_plain(o)) s_odd2[par] = 1;
                        }
                    }
                }
                __syncthreads();
                const int cnt = s_cnt;
                const bool odd = s_odd2[par] != 0;
                if (tid == 0) s_cnt2[par ^ 1u] = s_odd2[par ^ 1u] = 0;  // last read before this pass's first barrier
                any |= cnt > 0;
                if (odd && !slow) {  // no real value is +inf so far (plain objects): +inf means untouched
                    slow = true;
#pragma unroll
                    for (int rr = 0; rr < kTRows; ++rr) {
                        if (__float_as_uint(acc[rr][0]) == kTNoneFast) acc[rr][0] = __uint_as_float(kTNone);
                        if (__float_as_uint(acc[rr][1]) == kTNoneFast) acc[rr][1] = __uint_as_float(kTNone);
                    }
                }
                if (tile < tiles_x && cnt > 0) {
                    if (slow) sweep_tile(s_list, cnt, tx0, lane, acc);
                    else sweep_tile_fast(s_list, cnt, tx0, lane, s_queue + warp * (kTCap + 2), acc);
                }
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
                const unsigned none = slow ? kTNone : kTNoneFast;
#pragma unroll
                for (int rr = 0; rr < kTRows; ++rr) {
                    const int y = uy0 + rr;
                    if (y < uy1) {
                        float* const row = plane_f + (long long)y * W;
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            const int x = xl + 32 * c;
                            if (x < W) {
                                const unsigned t = __float_as_uint(acc[rr][c]);
                                if constexpr (ACCUM) {
                                    if (t != none) {
                                        const float v = __fmul_rn(expf(-__uint_as_float(t)), k);
                                        const float g = row[x];
                                        if (!(g >= v)) row[x] = v;  // cuh:30: keep old iff old >= v
                                    }
                                } else {
                                    float f = 0.0f;
                                    if (t != none) {
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
