// dh_tiles_kernel.cuh — the OBJECT-HEAVY unit renderer: register tiles instead of staged rows.
// Included by dh_groups_kernel.cuh right before groups_kernel (uses ObjI, Digest); device functions only.
//
// Same result as the reference's per-object max-blend (draw_heatmap_cuda_kernel.cuh:26-49, 76-109), same arithmetic per
// pixel as the staged renderer (m = (dy^2 + dx^2) * (1/var) with explicit IEEE roundings, cuh:45; max_o k*expf(-m_o) ==
// k*expf(-min_o m_o) for k >= 0), for units on which DOZENS TO HUNDREDS of objects land (BASELINE cfg4: 1024 objects of
// radius 16-64 on a 512x512 plane, every pixel under ~29 of them). There a launch is bound by instruction issue, not by
// stores (ncu, profiles/r02/ncu_groups_kernel_cfg4.txt: issue-active 84 %, DRAM 0.07 %), and the staged renderer spends
// ~30 warp instructions per 32 object-pixels: every object is swept into shared-memory staging rows with a
// read-modify-write per pixel quad, two rows at a time, and every row of every object pays the row set-up again.
//
// WHO DECIDES: the device, per unit. The host cannot see how many objects a sample really holds (the sizes live in
// device memory; T is only the padded slot count — CenterPoint-style callers pad ~50 objects to 500 slots), so it only
// selects the groups_kernel instantiation that CONTAINS this renderer (DENSE: T >= 256 and k >= 0); every CTA then looks
// at the list it has just built for its unit and takes this path when the list is long (RenderParams::dense_min <=
// entries <= kGCap), the staged / register paths of groups_kernel otherwise. (The first version was a kernel of its own
// chosen by T alone: cfg4 fine, but the 8-GPU slice of cfg5 padded to T = 256 took 0.73 instead of 0.24 ms.)
//
// A dense unit is cut into STRIPS of kTRows = 8 rows and tiles of 8 x 64 pixels, one tile per warp at a time. A lane owns
// two columns of the tile (x, x + 32) and keeps the running minimum of m for its 16 pixels in REGISTERS. The unit list
// is first rewritten in place into float form (TileEnt), which also tells whether every object is "plain".
//
// Two sweeps. sweep_tile (the exact path for unusual objects) predicates every pixel: ~11 warp instructions per (object,
// row of 64 pixels), accumulators as raw bits of m (m >= 0, so unsigned order == float order; all ones = untouched).
// sweep_tile_fast serves "plain" objects — r <= kTExactR, so dy^2 + dx^2 < 2^24 is an exact integer in fp32 however it
// is formed, and 0 < 1/var <= 1e30, so no in-window m overflows — and replaces predication by arithmetic: a pixel
// outside the object's columns gets dx^2 = +inf, a row outside its rows dy = +inf, hence m = +inf, which never beats a
// real value and doubles as "untouched". What is left per (object, row) is FADD2 (dy, dy) + (1, 1), FFMA2 dy * dy +
// (dx0^2, dx1^2), FMUL2 by (w, w) — the packed fp32 instructions of sm_100, each component rounded as the scalar
// instruction would — and, objects taken two at a time, one three-input FMNMX3 per pixel and pair of objects: 4 instead
// of 11 issue slots. fma(dy, dy, dx^2) equals the reference's (float)(i*i) + (float)(j*j) (cuh:45) bit for bit because
// every term is an exact integer below 2^24. One expf per touched pixel at the end; FRESH writes every pixel of the tile
// (value or zero) with coalesced 128-byte row segments, ACCUMULATE read-max-writes the touched ones (cuh:30).
#pragma once

namespace dh {

constexpr int kTRows = 8;     // rows of a strip / tile (12 and 16 measured: no gain, profiles/r02/tiles/geometry_ab.txt)
constexpr int kTCols = 64;    // columns of a tile (two per lane)
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

// A list entry of a dense unit: the object in the form both sweeps consume. All four are exact: |x| <= r + W < 2^24 and
// |y0 - y| <= r + rows of the unit for every object that passed the cull (r <= DH_MAX_RADIUS = 2^22).
struct __align__(16) TileEnt {
    float fx;   // centre column
    float fr;   // radius
    float fdy;  // first row of the UNIT - centre row
    float w;    // 1 / var
};
static_assert(sizeof(TileEnt) == sizeof(ObjI), "the unit list is rewritten in place");

constexpr int kTQueueBytes = (kGCap + 2) * (int)sizeof(unsigned short);  // per-warp hit queue (lives in the warp's staging buffer)

// entry overlaps the strip whose first row is `dys` rows below the unit's first row, and the tile's columns [t0f, t1f)
__device__ __forceinline__ bool tile_hit(const TileEnt& o, float dys, float t0f, float t1f) {
    const float fdy = o.fdy + dys;
    return (o.fx + o.fr >= t0f) && (o.fx - o.fr < t1f) && (fdy + (float)(kTRows - 1) >= -o.fr) && (fdy <= o.fr);
}

// All objects of `list[0, cnt)` that overlap the tile (rows: the strip `dys` rows into the unit; columns [tx0, tx0 + 64))
// into the lane's 8 x 2 accumulators (raw bits of m, kTNone = untouched): every pixel predicated, any object.
// Warp-collective.
__device__ __forceinline__ void sweep_tile(const TileEnt* __restrict__ list, int cnt, int tx0, float dys, int lane,
                                           float (&acc)[kTRows][2]) {
    const float xlf = (float)(tx0 + lane);  // my first column
    const float t0f = (float)tx0, t1f = (float)(tx0 + kTCols);
    for (int base = 0; base < cnt; base += 32) {
        const int e = base + lane;
        bool hit = false;
        if (e < cnt) hit = tile_hit(list[e], dys, t0f, t1f);
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        while (bal) {
            const int src = __ffs(bal) - 1;
            bal &= bal - 1u;
            const TileEnt o = list[base + src];  // broadcast
            const float fdx0 = xlf - o.fx, fdx1 = fdx0 + 32.0f;
            const float dx20 = __fmul_rn(fdx0, fdx0), dx21 = __fmul_rn(fdx1, fdx1);
            const bool in0 = fabsf(fdx0) <= o.fr, in1 = fabsf(fdx1) <= o.fr;
            float fdy = o.fdy + dys;
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

__device__ __forceinline__ TileObj tile_obj(const TileEnt& o, float xlf, float dys) {
    TileObj t;
    const float inf = __uint_as_float(kTNoneFast);
    const float fdx0 = xlf - o.fx, fdx1 = fdx0 + 32.0f;
    t.a0 = fabsf(fdx0) <= o.fr ? __fmul_rn(fdx0, fdx0) : inf;
    t.a1 = fabsf(fdx1) <= o.fr ? __fmul_rn(fdx1, fdx1) : inf;
    t.fdy = o.fdy + dys;
    t.w = o.w;
    t.fr = o.fr;
    t.allrows = (t.fdy >= -o.fr) && (t.fdy + (float)(kTRows - 1) <= o.fr);
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
// untouched. The tile is rows [strip, strip + nrows) x columns [tx0, min(tx0 + 64, W)).
// 1. Dominance bound (one list entry per lane): among the objects whose window covers the WHOLE tile, the smallest
//    m at the tile point farthest from the centre, M = min_B RN(D2far_B * w_B). Every pixel of the tile ends at m <= M.
// 2. Queue: the entries that overlap the tile AND can still win somewhere on it, RN(D2near_A * w_A) <= M, compacted into
//    the warp's queue `q` (ballot + popc prefix, 16-bit entry indices). An object dropped here has m > M >= the final m
//    on every pixel of the tile, so dropping it changes no bit: D2near / D2far are exact integers (r <= kTExactR), a
//    pixel's m is RN(D2 * w) with D2near <= D2 <= D2far, and rounding is monotone. On cfg4 (every pixel under ~29
//    objects) 45 % of the (object, tile) pairs go this way.
// 3. The queue is swept two objects at a time — no bit scanning and no carried state between the objects; an odd queue
//    ends with its last object twice (min is idempotent).
__device__ __forceinline__ void sweep_tile_fast(const TileEnt* __restrict__ list, int cnt, int tx0, float dys, int nrows,
                                                int W, int lane, unsigned short* __restrict__ q, float (&acc)[kTRows][2]) {
    const float xlf = (float)(tx0 + lane);
    const float x0t = (float)tx0, x1t = (float)min(tx0 + kTCols - 1, W - 1), nr1 = (float)(nrows - 1);
    unsigned M = kTNoneFast;  // bits of +inf; m >= 0, so unsigned order == float order
    for (int base = 0; base < cnt; base += 32) {
        const int e = base + lane;
        unsigned mb = kTNoneFast;
        if (e < cnt) {
            const TileEnt o = list[e];
            const float fdy = o.fdy + dys;  // first row of the tile - centre row
            const float dxf = fmaxf(fabsf(x0t - o.fx), fabsf(x1t - o.fx)), dyf = fmaxf(fabsf(fdy), fabsf(fdy + nr1));
            if (dxf <= o.fr && dyf <= o.fr)  // the window covers the tile
                mb = __float_as_uint(__fmul_rn(__fmaf_rn(dxf, dxf, __fmul_rn(dyf, dyf)), o.w));
        }
        M = min(M, __reduce_min_sync(0xffffffffu, mb));
    }
    int n = 0;  // warp-uniform
    for (int base = 0; base < cnt; base += 32) {
        const int e = base + lane;
        bool hit = false;
        if (e < cnt) {
            const TileEnt o = list[e];
            const float fdy = o.fdy + dys;
            const float dxn = fmaxf(0.0f, fmaxf(x0t - o.fx, o.fx - x1t)), dyn = fmaxf(0.0f, fmaxf(fdy, -(fdy + nr1)));
            hit = dxn <= o.fr && dyn <= o.fr &&  // the window overlaps the tile ...
                  __float_as_uint(__fmul_rn(__fmaf_rn(dxn, dxn, __fmul_rn(dyn, dyn)), o.w)) <= M;  // ... and can still win
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
        const TileObj A = tile_obj(list[pr & 0xffffu], xlf, dys), B = tile_obj(list[pr >> 16], xlf, dys);
        if (A.allrows && B.allrows) sweep_pair<false>(A, B, acc);
        else sweep_pair<true>(A, B, acc);
    }
    __syncwarp();  // the queue is rewritten for the next tile
}

// K4 for one tile: one expf per touched pixel (cuh:45), then the blend. Rows ty0.. (below uy1), columns tx0 + lane (+32).
// 16-bit planes (OUT != OUT_F32, FRESH only): the fp32 value rounded once at the store, as everywhere else.
template <bool ACCUM, bool DIGEST, int OUT>
__device__ __forceinline__ void finish_tile(const float (&acc)[kTRows][2], unsigned none, int tx0, int ty0, int uy1, int W,
                                            float k, char* __restrict__ plane_b, long long index0, int lane, Digest& dg) {
    const int xl = tx0 + lane;
#pragma unroll
    for (int rr = 0; rr < kTRows; ++rr) {
        const int y = ty0 + rr;
        if (y < uy1) {
            float* const row = reinterpret_cast<float*>(plane_b) + (long long)y * W;                    // OUT_F32
            unsigned short* const hrow = reinterpret_cast<unsigned short*>(plane_b) + (long long)y * W;  // 16-bit planes
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
                        if constexpr (OUT == OUT_F32) row[x] = f;
                        else hrow[x] = cvt16<OUT>(f);
                    }
                }
            }
        }
    }
}

// One dense unit: rows [uy0, uy1) of the plane at `plane_b`, strip by strip, every warp of the CTA (block-collective).
// `ucnt` objects landed on the unit. If they fit the unit list (ucnt <= kGCap) the list is rewritten in place into the
// float form (fdy relative to uy0) — the barrier that publishes it also tells whether every object is plain — and every
// tile sweeps it. If not, the sample's slots are scanned again per strip in passes of kGCap slots against the strip's
// rows (a pass cannot overflow the list) and the accumulators of a tile persist over the passes; a pass that holds an
// object that is not plain switches the tile to the predicated sweep for its remaining passes — no real value is +inf
// up to then, so the "untouched" code is converted once. `q`: this warp's hit queue (kTQueueBytes).
template <bool ACCUM, int SRC, bool DIGEST, int OUT>
__device__ __forceinline__ void render_unit_dense(const RenderParams& p, const UnitSrc& us, int nscan, int ucnt, int uy0,
                                                  int uy1, int W, float k, char* __restrict__ plane_b, long long index0,
                                                  ObjI* s_unit, unsigned short* __restrict__ q, int tid, int lane, int warp,
                                                  Digest& dg) {
    __shared__ int s_dcnt[2], s_dodd[2];  // pass list size / "not plain" flag, double-buffered by pass parity
    TileEnt* const s_list = reinterpret_cast<TileEnt*>(s_unit);
    const int tiles_x = (W + kTCols - 1) / kTCols;
    const bool prebuilt = ucnt <= kGCap;  // block-uniform
    bool slow0 = false;
    if (prebuilt) {
        bool odd = false;
        for (int e = tid; e < ucnt; e += kGThreads) {
            const ObjI o = s_unit[e];
            odd = odd || !tile_obj_plain(o);
            s_list[e] = TileEnt{(float)o.x, (float)o.r, (float)(uy0 - o.y), o.w};
        }
        slow0 = __syncthreads_or(odd) != 0;
    } else if (tid == 0) {
        s_dcnt[0] = s_dcnt[1] = s_dodd[0] = s_dodd[1] = 0;  // published by the first pass's first barrier
    }
    const int npass = prebuilt ? 1 : (nscan + kGCap - 1) / kGCap;
    unsigned par = 0;
    for (int ty0 = uy0; ty0 < uy1; ty0 += kTRows) {
        const int ty1 = min(ty0 + kTRows, uy1);
        const float dys = prebuilt ? (float)(ty0 - uy0) : 0.0f;  // pass lists are built relative to the strip
        for (int t0 = 0; t0 < tiles_x; t0 += kGWarps) {
            const int tile = t0 + warp, tx0 = tile * kTCols;
            bool slow = slow0;  // same value in every thread
            float acc[kTRows][2];  // min m per pixel: floats while every object was plain, raw bits after that
#pragma unroll
            for (int rr = 0; rr < kTRows; ++rr) acc[rr][0] = acc[rr][1] = __uint_as_float(slow ? kTNone : kTNoneFast);
            for (int ps = 0; ps < npass; ++ps) {
                int cnt = ucnt;
                if (!prebuilt) {
                    __syncthreads();  // everybody is done with the previous pass's list
                    const int s0 = ps * kGCap, send = min(nscan, s0 + kGCap);
                    for (int base = s0; base < send; base += kGThreads) {
                        ObjI o;
                        const int t = base + tid;
                        const bool pred = (t < send) && scan_slot<SRC>(p, us, t, nscan, ty0, ty1, W, false, o);
                        const unsigned bal = __ballot_sync(0xffffffffu, pred);
                        if (bal) {
                            int pos = 0;
                            if (lane == 0) pos = atomicAdd(&s_dcnt[par], __popc(bal));
                            pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(bal & ((1u << lane) - 1u));
                            if (pred) {  // pos < kGCap: a pass scans at most kGCap slots
                                s_list[pos] = TileEnt{(float)o.x, (float)o.r, (float)(ty0 - o.y), o.w};
                                if (!tile_obj_plain(o)) s_dodd[par] = 1;
                            }
                        }
                    }
                    __syncthreads();
                    cnt = s_dcnt[par];
                    const bool odd = s_dodd[par] != 0;
                    if (tid == 0) s_dcnt[par ^ 1u] = s_dodd[par ^ 1u] = 0;  // last read before this pass's first barrier
                    par ^= 1u;
                    if (odd && !slow) {
                        slow = true;
#pragma unroll
                        for (int rr = 0; rr < kTRows; ++rr) {
                            if (__float_as_uint(acc[rr][0]) == kTNoneFast) acc[rr][0] = __uint_as_float(kTNone);
                            if (__float_as_uint(acc[rr][1]) == kTNoneFast) acc[rr][1] = __uint_as_float(kTNone);
                        }
                    }
                }
                if (tile < tiles_x && cnt > 0) {
                    if (slow) sweep_tile(s_list, cnt, tx0, dys, lane, acc);
                    else sweep_tile_fast(s_list, cnt, tx0, dys, ty1 - ty0, W, lane, q, acc);
                }
            }
            if (tile < tiles_x)
                finish_tile<ACCUM, DIGEST, OUT>(acc, slow ? kTNone : kTNoneFast, tx0, ty0, uy1, W, k, plane_b, index0, lane, dg);
        }
    }
    if (!prebuilt) __syncthreads();  // the list storage goes back to the kernel
}

}  // namespace dh
