// dh_groups_kernel.cuh — the main renderer (maps up to kGMaxW pixels wide: every BASELINE config).
// Included by dh_kernels.cu after dh_rows_kernel.cuh (shares ObjI, load_slot, make_obji, finalize_row, PTX wrappers).
//
// Replaces the reference's one-thread-per-object renderer with its global CAS loop per pixel
// (/root/reference/packages/draw_heatmap/accvlab/draw_heatmap/include/draw_heatmap_cuda_kernel.cuh:26-109) by an
// OUTPUT-partitioned one. What the measurements behind it said (profiles/micro/*.cu, profiles/r01_*):
//   * the same stores run at 7.0-7.4 TB/s when work is handed out in address order by the block scheduler and at
//     6.1 TB/s from a static persistent grid-stride loop  -> ONE CTA PER UNIT, flat grid;
//   * the previous kernel was instruction-issue bound (64 % issue-active, ~300 warp instructions per touched row)
//     -> rows nobody touches cost nothing per row, touched rows take a register-only path when <= 2 objects hit them.
//
// Geometry: a row group = G whole rows of one plane (one contiguous byte range <= kGStageCap); a unit = up to 64
// consecutive groups of one plane (~192 KB). Per unit:
//   K1/K2  the sample's T slots are scanned once (all loads issued up front), radius/centre computed when the source
//          is float boxes, objects culled against the unit rows and compacted with warp ballot + popc prefix into a
//          shared-memory list; every object ORs the groups it overlaps into a 64-bit occupancy mask.
//   K4a    FRESH: runs of untouched groups leave as cp.async.bulk (shared -> global, SASS UBLKCP) from a constant
//          zero tile: no per-row work at all.
//   K3     touched groups are striped over the 8 warps (group g -> warp g % 8). With <= 32 objects in the unit every
//          lane keeps one list entry in registers for the whole unit, so the per-group hit test is a compare + ballot.
//            sparse path (<= 2 hits): per 32-pixel chunk of the hit span each lane evaluates its pixel for both
//              objects in registers (min of m = (i^2+j^2)/var) and applies ONE expf: no shared-memory read-modify-write;
//            dense path: objects are swept into the staging rows as order-preserving uints (max == min m), then the
//              span union is finalised in place with one expf per touched pixel.
//   K4b    FRESH: the staging buffer (G rows, final fp32 values, zeros elsewhere) leaves with one bulk store issued by
//          the warp itself; two buffers per warp alternate. Buffers are all-zero whenever idle: after the copy engine
//          has read one, only the span that was written is cleared.
//          ACCUMULATE (reference semantics): touched pixels only are read-max-written in place with plain ld/st.
// No block barrier after the list build; no atomics on the heatmap anywhere.
#pragma once

namespace dh {

#ifndef DH_G_THREADS
#define DH_G_THREADS 256  // 256 (8 warps, 3 CTAs per SM) or 128 (4 warps, 6 CTAs per SM: A/B builds)
#endif
#ifndef DH_G_MINB
#define DH_G_MINB (768 / DH_G_THREADS)  // resident CTAs per SM the register budget is cut for (24 warps per SM)
#endif
#ifndef DH_G_MINB_ACCUM
// ACCUMULATE (the reference's in-place semantics) reads before it writes and is bound by memory latency, not by stores
// (ncu, profiles/r02/ncu_groups_kernel_accum_b1536.txt: `long_scoreboard` is the top stall, DRAM at 14 %). It needs
// neither the zero tile nor a second staging buffer per warp, so one more CTA fits an SM: 32 resident warps, not 24.
#define DH_G_MINB_ACCUM (1024 / DH_G_THREADS)
#endif
#ifndef DH_G_STAGE_CAP
#define DH_G_STAGE_CAP 4096
#endif
constexpr int kGThreads = DH_G_THREADS;
constexpr int kGWarps = kGThreads / 32;
static_assert(kGWarps == 8 || kGWarps == 4, "groups kernels: 8 or 4 warps per CTA");
constexpr unsigned long long kGStripe = kGWarps == 8 ? 0x0101010101010101ull : 0x1111111111111111ull;  // groups of warp 0
constexpr int kGStageCap = DH_G_STAGE_CAP;  // bytes of one staging buffer (one row group)
constexpr int kGMaxW = kGStageCap / 4;      // widest map served by this kernel
constexpr int kGZeroBytes = kGWarps == 8 ? 8192 : 4096;  // constant zero tile
constexpr int kGCap = kGWarps == 8 ? 512 : 256;           // unit list capacity
constexpr int kGMaxGroups = 64;             // groups per unit (bits of the occupancy mask)
#ifndef DH_SPARSE_MAX
#define DH_SPARSE_MAX 2  // 3 and 4 were measured: registers spill, cfg5 1.57 -> 1.70 / 1.85 ms (profiles/r02/experiments.md section 8)
#endif
constexpr int kSparseMax = DH_SPARSE_MAX;   // objects on a group that are still evaluated in registers (no staging sweep)
constexpr int kGPermCap = 2048;             // samples whose hand-out order is rearranged (heaviest first)
constexpr int kGPermBins = 256;             // counting-sort bins of that rearrangement
constexpr int kGPermSmem = kGPermCap * 2;   // static shared memory it takes (bytes)
constexpr int kGPermScratch = kGWarps * kGPermBins * 2 + 64;  // bytes of the staging area build_perm borrows

// dynamic shared memory: zero tile | 8 warps x 2 staging buffers of `stage` bytes | unit list
// dynamic shared memory: [zero tile] | kGWarps x (2 | 1) staging buffers of `stage` bytes | unit list.
// ACCUMULATE has no zero tile and one staging buffer per warp (scratch of the dense path; nothing leaves from it).
// LEAN (object-heavy FRESH renders, T >= 256): the zero tile stays, but one staging buffer per warp — the render of a
// group takes far longer than the bulk read of the previous one, so there is nothing to double-buffer — which buys twice
// the rows per group (the per-object set-up is paid per group) and the fourth CTA per SM.
__host__ __device__ constexpr size_t groups_smem_bytes(int stage, bool accum = false, bool lean = false) {
    return (size_t)(accum ? 0 : kGZeroBytes) + (size_t)kGWarps * ((accum || lean) ? 1 : 2) * (size_t)stage +
           (size_t)kGCap * sizeof(ObjI);
}

__device__ __forceinline__ unsigned enc_m(float m, bool kneg) {
    const unsigned b = __float_as_uint(m);
    return kneg ? b + 1u : ~b;
}

// 16-bit heatmap planes (FRESH mode only): OUT = 0 fp32 (the reference's only type), 1 fp16, 2 bf16. The value
// is computed in fp32 exactly as before and rounded ONCE, to nearest even, when it is staged for the bulk store.
enum OutType : int { OUT_F32 = 0, OUT_F16 = 1, OUT_BF16 = 2 };
template <int OUT>
__device__ __forceinline__ unsigned short cvt16(float f) {
    if constexpr (OUT == OUT_F16) return __half_as_ushort(__float2half_rn(f));
    else return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}

// Pixel index -> 64-bit digest contribution (== dh_host.cu, oracle/oracle.py:_mix_index): 32-bit IMADs only — the
// renderer is instruction-issue bound, a 64-bit splitmix per touched pixel cost 12 % of the step.
__device__ __forceinline__ unsigned long long mix_index(unsigned long long i) {
    const unsigned lo = (unsigned)i, hi = (unsigned)(i >> 32);
    unsigned h = lo ^ (hi * 0x9E3779B1u);
    h *= 0x85EBCA6Bu;
    h ^= h >> 13;
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return ((unsigned long long)h << 32) | (unsigned long long)(h * 0x27D4EB2Fu + lo);
}

struct Digest {
    unsigned long long bits = 0, hash = 0;
    unsigned cnt = 0;  // per lane; widened when the lanes are combined
    __device__ __forceinline__ void add(float v, long long index) {
        bits += (unsigned long long)__float_as_uint(v);
        if (v > 0.0f) {
            hash += mix_index((unsigned long long)index);
            cnt += 1u;
        }
    }
};

// Dense path, one object into the rows [y0, y0+nrows) held in `buf` (row stride W). Warp-collective.
// (Round 2 tried columns outer / rows inner — the x-only terms computed once per pixel and reused for the rows of the
// group: 27 instead of 40 instructions in the inner body, but more instructions overall (ncu on cfg4: 228 M against
// 201 M warp instructions in this function, 353 against 314 us; profiles/r02/call5/) because with two rows per group the
// second loop nest and its prologue cost more than the shared terms save. Rows outer stays.)
__device__ __forceinline__ void sweep_obj(unsigned* __restrict__ buf, int y0, int nrows, int W, const ObjI& o, bool kneg,
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
            // every x visited lies inside the row (x0 & ~3 >= 0, W % 4 == 0), so the window test is |dx| <= r alone;
            // dx as a float: one conversion per lane, then exact +1 steps (|dx| < 2^24)
            const float fr = (float)o.r;
            for (int xv = (x0 & ~3) + 4 * lane; xv <= x1; xv += 128) {
                uint4 q = *reinterpret_cast<uint4*>(rowp + xv);
                unsigned* qc = reinterpret_cast<unsigned*>(&q);
                const float fdx0 = (float)(xv - o.x);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float fdx = fdx0 + (float)c;
                    const unsigned t = enc_m(__fmul_rn(__fadd_rn(fi, __fmul_rn(fdx, fdx)), o.w), kneg);
                    qc[c] = (fabsf(fdx) <= fr) ? max(qc[c], t) : qc[c];
                }
                *reinterpret_cast<uint4*>(rowp + xv) = q;
            }
        }
    }
}

// Dense path, K4: the span union of one staged row, one expf per touched pixel (cuh:45), in place.
template <bool ACCUM, bool DIGEST, int OUT>
__device__ __forceinline__ void finish_row(unsigned* __restrict__ rowp, float* __restrict__ grow,
                                           unsigned short* __restrict__ hrow, long long gindex, int x0, int x1, bool kneg,
                                           float k, int lane, Digest& dg) {
    for (int x = x0 + lane; x <= x1; x += 32) {
        const unsigned t = rowp[x];
        if (t != 0u) {
            const float m = __uint_as_float(kneg ? t - 1u : ~t);
            const float v = __fmul_rn(expf(-m), k);
            if constexpr (ACCUM) {
                rowp[x] = 0u;  // the staging row is scratch here: back to all-zero
                const float g = grow[x];
                if (!(g >= v)) grow[x] = v;  // cuh:30: keep old iff old >= v
            } else {
                const float f = (0.0f >= v) ? 0.0f : v;  // max-blend into the fused zero-init
                if constexpr (OUT == OUT_F32) {
                    rowp[x] = __float_as_uint(f);
                } else {  // 16-bit planes: the fp32 row is scratch (back to all-zero), the value goes to the 16-bit row
                    rowp[x] = 0u;
                    hrow[x] = cvt16<OUT>(f);
                }
                if constexpr (DIGEST) dg.add(f, gindex + x);
            }
        }
    }
}

// Sparse path: at most two objects (A, and B when hasB) overlap the group's rows. Register-only evaluation, 32 pixels
// at a time over the span the row is hit in; arithmetic of cuh:45 with the sign pulled out of the exponential:
//   k >= 0: max_o k*expf(-m_o) = k*expf(-min_o m_o);  k < 0: the max picks max_o m_o = -min_o(-m_o)
// (expf is monotone: verified exhaustively on the device by tests/test_gpu_properties.py), so `ws` = +-1/var and one
// fminf cover both signs; negation is exact, so every rounding is the reference's.
// Round 2: rows are taken TWO AT A TIME and a group hit by ONE object (the common case) has its own instantiation. What
// depends on x alone (dx, dx^2 — shared by both rows —, the |dx| <= r test) is computed once per 32-pixel chunk, the two
// rows are two independent dependency chains through expf (a heavy unit is bound by the latency of that chain, not by
// issue slots: unit timelines, profiles/r02/experiments.md section 8), and a single object skips the second candidate
// altogether: ~22 instead of ~45 instructions per (row, chunk). The arithmetic per pixel is unchanged (cuh:45).
template <bool ACCUM, bool DIGEST, int OUT, int N>
__device__ __forceinline__ void sparse_rows(unsigned* __restrict__ buf, float* __restrict__ gdst, long long gindex, int y0,
                                            int nrows, int W, const ObjI (&O)[N], bool kneg, float k, int lane, int& ux0,
                                            int& ux1, Digest& dg) {
    int ox0[N], ox1[N];
    float ow[N], fr[N];
#pragma unroll
    for (int j = 0; j < N; ++j) {
        ox0[j] = max(O[j].x - O[j].r, 0);
        ox1[j] = min(O[j].x + O[j].r, W - 1);
        ow[j] = kneg ? -O[j].w : O[j].w;
        fr[j] = (float)O[j].r;
    }
    // one finished FRESH pixel at offset `off` = row * W + x of the group (v = expf(-m) * k already formed)
    auto emit = [&](int off, float v) {
        const float f = (0.0f >= v) ? 0.0f : v;
        if constexpr (OUT == OUT_F32) buf[off] = __float_as_uint(f);
        else reinterpret_cast<unsigned short*>(buf)[off] = cvt16<OUT>(f);  // `buf` = the 16-bit buffer
        if constexpr (DIGEST) dg.add(f, gindex + off);
    };
    for (int rr = 0; rr < nrows; rr += 2) {
        const bool has1 = rr + 1 < nrows;
        bool in0[N], in1[N];  // object j covers row rr / row rr + 1 (warp-uniform)
        float fi0[N], fi1[N];
        int rx0 = W, rx1 = -1;
#pragma unroll
        for (int j = 0; j < N; ++j) {
            const int dy = y0 + rr - O[j].y;
            in0[j] = abs(dy) <= O[j].r;
            in1[j] = has1 && abs(dy + 1) <= O[j].r;
            const float f0 = (float)dy, f1 = (float)(dy + 1);
            fi0[j] = __fmul_rn(f0, f0);
            fi1[j] = __fmul_rn(f1, f1);
            if (in0[j] || in1[j]) {
                rx0 = min(rx0, ox0[j]);
                rx1 = max(rx1, ox1[j]);
            }
        }
        if (rx1 < rx0) continue;
        ux0 = min(ux0, rx0);
        ux1 = max(ux1, rx1);
        for (int xb = rx0; xb <= rx1; xb += 32) {
            const int x = xb + lane;
            float dx2[N];
            bool tx[N];
#pragma unroll
            for (int j = 0; j < N; ++j) {
                const float fdx = (float)(x - O[j].x);  // exact (|dx| < 2^24)
                dx2[j] = __fmul_rn(fdx, fdx);
                tx[j] = fabsf(fdx) <= fr[j] && x <= rx1;
            }
            if constexpr (ACCUM) {
                // The old pixels of BOTH rows are requested first, by the lanes that will blend only, so the two loads
                // are in flight together and the arithmetic runs under them. (Written pixel by pixel — load, blend,
                // store, next load — the second load waits for the first store: the compiler must assume they alias.)
                bool th[2];
                float g[2] = {0.0f, 0.0f};
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    th[h] = false;
#pragma unroll
                    for (int j = 0; j < N; ++j) th[h] = th[h] || ((h ? in1[j] : in0[j]) && tx[j]);
                }
                float* const gp = gdst + (rr * W + x);  // one address per chunk (a group is at most 4 KB: 32-bit offsets)
                if (th[0]) g[0] = gp[0];
                if (th[1]) g[1] = gp[W];
                float v[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float acc = 0.0f;
                    bool touched = false;
#pragma unroll
                    for (int j = 0; j < N; ++j) {
                        const bool t = (h ? in1[j] : in0[j]) && tx[j];
                        const float m = __fmul_rn(__fadd_rn(h ? fi1[j] : fi0[j], dx2[j]), ow[j]);
                        acc = t ? (touched ? fminf(acc, m) : m) : acc;
                        touched = touched || t;
                    }
                    v[h] = __fmul_rn(expf(kneg ? acc : -acc), k);  // = expf(-m) * k
                }
                if (th[0] && !(g[0] >= v[0])) gp[0] = v[0];  // cuh:30
                if (th[1] && !(g[1] >= v[1])) gp[W] = v[1];
            } else {
                const int off = rr * W + x;  // one offset per chunk (a group is at most 4 KB)
#pragma unroll
                for (int h = 0; h < 2; ++h) {  // row rr, row rr + 1
                    float acc = 0.0f;
                    bool touched = false;
#pragma unroll
                    for (int j = 0; j < N; ++j) {
                        const bool t = (h ? in1[j] : in0[j]) && tx[j];
                        const float m = __fmul_rn(__fadd_rn(h ? fi1[j] : fi0[j], dx2[j]), ow[j]);
                        acc = t ? (touched ? fminf(acc, m) : m) : acc;
                        touched = touched || t;
                    }
                    if (touched) emit(h ? off + W : off, __fmul_rn(expf(kneg ? acc : -acc), k));  // = expf(-m) * k
                }
            }
        }
    }
}

// ---- cluster launch control (sm_100): a resident CTA asks the hardware block scheduler for the index of a CTA of
// this grid that has not been launched yet, cancels that launch and does its work itself. The grid still has one CTA
// per unit and units are still handed out dynamically and in address order, but shared-memory set-up is paid once per
// resident CTA and — the point — a CTA starts the next unit's list build while its bulk stores are still draining.
// The request goes out when the current unit is DONE, not at its start (see the end of the unit loop in groups_kernel).
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "DH_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DH_MBAR_DONE;\n"
        "bra DH_MBAR_WAIT;\n"
        "DH_MBAR_DONE:\n"
        "}" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void clc_try_cancel(void* response16, unsigned long long* bar) {
    asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];" ::"r"(
                     smem_u32(response16)),
                 "r"(smem_u32(bar))
                 : "memory");
}
// returns true and the x index of the cancelled CTA, or false when the grid has no CTA left to hand out
__device__ __forceinline__ bool clc_response(const void* response16, unsigned& ctaid_x) {
    unsigned valid, x;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b128 r;\n"
        "ld.shared.b128 r, [%2];\n"
        "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "mov.u32 %1, 0;\n"
        "@p clusterlaunchcontrol.query_cancel.get_first_ctaid::x.b32.b128 %1, r;\n"
        "}"
        : "=r"(valid), "=r"(x)
        : "r"(smem_u32(response16))
        : "memory");
    ctaid_x = x;
    return valid != 0u;
}

__device__ __forceinline__ ObjI shfl_obj(const ObjI& o, int src) {
    ObjI r;
    r.x = __shfl_sync(0xffffffffu, o.x, src);
    r.y = __shfl_sync(0xffffffffu, o.y, src);
    r.r = __shfl_sync(0xffffffffu, o.r, src);
    r.w = __shfl_sync(0xffffffffu, o.w, src);
    return r;
}

// Per-warp staging state: two buffers that are all-zero whenever idle, and the span their last use wrote.
struct WarpStage {
    unsigned* buf0;
    unsigned* buf1;
    bool single = false;  // LEAN layout: one buffer, re-used once the copy engine has read it
    int cur = 0;
    int c0x0 = 1, c0x1 = 0, c1x0 = 1, c1x1 = 0;
};

// Where a unit's objects come from (needed again by the overflow path, which re-scans the source per group).
struct UnitSrc {
    long long plane = 0, src_base = 0;
    int n = 0, cls = 0;
};

// K3 + K4b for ONE touched row group, by one warp. `list`/`ucnt` is the unit list in shared memory, `mine` the
// lane's own entry when the unit has <= 32 objects (`small`). FRESH: the group leaves with one bulk store.
// 16-bit planes (OUT != OUT_F32, FRESH only): the warp's first staging region is the fp32 scratch of the dense path
// (all-zero again after finish_row), the second one holds the two alternating 16-bit output buffers.
template <bool ACCUM, int SRC, bool DIGEST, int OUT>
__device__ __forceinline__ void render_group(const RenderParams& p, WarpStage& ws, char* __restrict__ plane_b,
                                             long long plane_index0, int y0, int nrows, int W, int G, bool kneg, float kk, int lane,
                                             bool small, const ObjI& mine, const ObjI* __restrict__ list, int ucnt,
                                             const UnitSrc& us, Digest& dg) {
    static_assert(OUT == OUT_F32 || !ACCUM, "16-bit planes exist in FRESH mode only");
    constexpr int ES = (OUT == OUT_F32) ? 4 : 2;
    const float k = (SRC == SRC_WEIGHTED) ? kk : p.k;  // per-class peak only in the weighted mode
    const int y1 = y0 + nrows - 1;
    float* const gdst = reinterpret_cast<float*>(plane_b + (long long)y0 * W * ES);  // fp32 view: OUT_F32 only
    const long long gindex = plane_index0 + (long long)y0 * W;
    // `obuf`: what the bulk store reads; `buf`: where the dense path sweeps (the same buffer for fp32 planes)
    unsigned* const obuf = (OUT == OUT_F32) ? (ws.cur ? ws.buf1 : ws.buf0)
                                            : reinterpret_cast<unsigned*>(reinterpret_cast<char*>(ws.buf1) +
                                                                          (ws.cur ? (p.stage_bytes >> 1) : 0));
    unsigned* const buf = (OUT == OUT_F32) ? obuf : ws.buf0;
    if constexpr (!ACCUM) {
        if (lane == 0) {
            if (ws.single) bulk_wait_read<0>();  // one buffer: every store of mine has been read
            else bulk_wait_read<1>();            // the store that last read `obuf` is older than my latest one
        }
        __syncwarp();
        const int cx0 = ws.cur ? ws.c1x0 : ws.c0x0, cx1 = ws.cur ? ws.c1x1 : ws.c0x1;
        if (cx1 >= cx0) {  // back to all-zero: clear only what its last use wrote (4 pixels per store, W % 4 == 0)
            const int q0 = cx0 >> 2, nq = (cx1 >> 2) - q0 + 1, w4 = W >> 2;
            if constexpr (OUT == OUT_F32) {
                uint4* const b4 = reinterpret_cast<uint4*>(obuf) + q0 + lane;
                if (nq <= 32) {  // the usual case (span <= 128 pixels): one store per row
                    if (lane < nq)
                        for (int rr = 0; rr < G; ++rr) b4[rr * w4] = make_uint4(0u, 0u, 0u, 0u);
                } else {
                    for (int rr = 0; rr < G; ++rr)
                        for (int q = lane; q < nq; q += 32) b4[rr * w4 + q - lane] = make_uint4(0u, 0u, 0u, 0u);
                }
            } else {
                uint2* const b2 = reinterpret_cast<uint2*>(obuf) + q0 + lane;
                if (nq <= 32) {
                    if (lane < nq)
                        for (int rr = 0; rr < G; ++rr) b2[rr * w4] = make_uint2(0u, 0u);
                } else {
                    for (int rr = 0; rr < G; ++rr)
                        for (int q = lane; q < nq; q += 32) b2[rr * w4 + q - lane] = make_uint2(0u, 0u);
                }
            }
            __syncwarp();
        }
    }
    int ux0 = W, ux1 = -1;
    bool finish = false;  // dense path used: staged uints still need K4
    if (small) {
        const bool hit = (mine.r >= 0) && (mine.y + mine.r >= y0) && (mine.y - mine.r <= y1);
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        const int nh = __popc(bal);
        if (nh <= kSparseMax) {  // register path: every hit object is evaluated per pixel, nothing goes through staging
            if (nh == 1) {  // the common case: four shuffles, not eight
                const ObjI O1[1] = {shfl_obj(mine, __ffs(bal) - 1)};
                sparse_rows<ACCUM, DIGEST, OUT, 1>(obuf, gdst, gindex, y0, nrows, W, O1, kneg, k, lane, ux0, ux1, dg);
            } else if (nh > 0) {
                ObjI O[kSparseMax];
                unsigned rest = bal;
#pragma unroll
                for (int j = 0; j < kSparseMax; ++j) {
                    const int src = rest ? __ffs(rest) - 1 : 0;
                    rest &= rest - 1u;
                    O[j] = shfl_obj(mine, src);
                }
                if (nh == 2) {
                    const ObjI O2[2] = {O[0], O[1]};
                    sparse_rows<ACCUM, DIGEST, OUT, 2>(obuf, gdst, gindex, y0, nrows, W, O2, kneg, k, lane, ux0, ux1, dg);
                } else {
                    sparse_rows<ACCUM, DIGEST, OUT, kSparseMax>(obuf, gdst, gindex, y0, nrows, W, O, kneg, k, lane, ux0, ux1, dg);
                }
            }
        } else {
            finish = true;
            while (bal) {
                const int src = __ffs(bal) - 1;
                bal &= bal - 1u;
                const ObjI ob = shfl_obj(mine, src);
                sweep_obj(buf, y0, nrows, W, ob, kneg, lane, ux0, ux1);
                __syncwarp();
            }
        }
    } else {
        finish = true;
        const bool overflow = ucnt > kGCap;  // rare: scan the source instead of the (truncated) shared list
        const int lim = overflow ? us.n : ucnt;
        for (int base = 0; base < lim; base += 32) {
            const int e = base + lane;
            ObjI o;
            o.x = o.y = o.r = 0;
            o.w = 0.f;
            bool hit = false;
            if (e < lim) {
                if (!overflow) {
                    o = list[e];
                    hit = (o.y + o.r >= y0) && (o.y - o.r <= y1);
                } else if constexpr (SRC == SRC_BINNED) {
                    o = ld_bin(p.bin_list, us.src_base + e);
                    hit = (o.r >= 0) && (o.y + o.r >= y0) && (o.y - o.r <= y1);
                } else {
                    int x, y, r;
                    if (load_slot<SRC>(p, us.src_base, e, us.cls, us.plane, false, x, y, r))
                        hit = make_obji(x, y, r, p.factor, y0, y1 + 1, W, o, nullptr, false);
                    if constexpr (SRC == SRC_WEIGHTED) {
                        if (hit) o.w = __ldg(p.weights + us.src_base + e);
                    }
                }
            }
            unsigned bal = __ballot_sync(0xffffffffu, hit);
            while (bal) {
                const int src = __ffs(bal) - 1;
                bal &= bal - 1u;
                const ObjI ob = shfl_obj(o, src);
                sweep_obj(buf, y0, nrows, W, ob, kneg, lane, ux0, ux1);
                __syncwarp();
            }
        }
    }
    if (finish && ux1 >= ux0) {
        for (int rr = 0; rr < nrows; ++rr)
            finish_row<ACCUM, DIGEST, OUT>(buf + rr * W, gdst + (long long)rr * W,
                                           reinterpret_cast<unsigned short*>(obuf) + rr * W, gindex + (long long)rr * W, ux0,
                                           ux1, kneg, k, lane, dg);
    }
    if constexpr (!ACCUM) {
        fence_proxy_async();  // my generic-proxy writes to `buf` -> visible to the bulk-copy engine
        __syncwarp();
        if (lane == 0) {
            bulk_store(gdst, obuf, (unsigned)(nrows * W * ES));
            bulk_commit();
        }
        if (ws.cur) {
            ws.c1x0 = ux0;
            ws.c1x1 = ux1;
        } else {
            ws.c0x0 = ux0;
            ws.c0x1 = ux1;
        }
        if (!ws.single) ws.cur ^= 1;
    } else {
        __syncwarp();
    }
}

// Unit -> plane, rows, source range. Returns the number of slots to scan (`nscan`; independent of any load up to 256 slots).
// Heaviest-first hand-out. Units are handed out in index order by the block scheduler, and a unit with many objects
// takes 3-5 x an empty one, so a launch used to end with a long tail of half-idle SMs while the last heavy units
// finished (unit timelines, profiles/r02/call4: the last ~8 us of 37 on cfg3, ~11 us of 211 on the 8-GPU slice of cfg5
// run at half occupancy or less). Every CTA therefore sorts the samples by their object count, descending (counting
// sort over kGPermBins bins in shared memory: `sizes` is all it reads, about a microsecond), and unit index -> sample
// goes through that permutation: the samples with the most objects are rendered first, the launch ends on the cheap ones. Only the last
// kGPermCap samples of the hand-out order are rearranged (what comes before them cannot be in the tail).
// The permutation must be THE SAME in every CTA (each one maps unit indices through its own copy), so the counting sort
// is a stable, atomics-free one: the samples are cut into one contiguous block per warp; every warp histograms its block
// (chunks of 32, __match_any_sync groups equal keys, the lowest lane of a group adds its size), the per-(warp, bin)
// counts are turned into start offsets (bins in order, warps in order inside a bin), and every warp scatters its block
// in the same chunk order. Scratch (keys, 8 x 256 counters) lives in the staging area, which nothing uses yet.
__device__ __forceinline__ void build_perm(const RenderParams& p, unsigned char* __restrict__ scratch,
                                           unsigned short* __restrict__ s_perm, int tid, int lane, int warp) {
    constexpr int PERW = kGPermCap / kGWarps / 32;  // chunks of 32 samples per warp, at most
    const int M = p.perm_n;
    unsigned short* const H = reinterpret_cast<unsigned short*>(scratch);  // [kGWarps][kGPermBins]
    int* const wsum = reinterpret_cast<int*>(scratch + kGWarps * kGPermBins * 2);  // [kGWarps] block-scan scratch
    const int blk = ((M + kGWarps - 1) / kGWarps + 31) & ~31;                 // samples per warp (multiple of 32)
    const int i0 = warp * blk;
    // keys of my (<= PERW) samples: every load issued before anything depends on one
    int key[PERW];
#pragma unroll
    for (int c = 0; c < PERW; ++c) {
        const int i = i0 + c * 32 + lane;
        key[c] = -1;
        if (c * 32 < blk && i < M) {
            const long long s = p.perm_base + i;
            long long ns = p.sizes_i64 ? __ldg(reinterpret_cast<const long long*>(p.sizes) + s)
                                       : (long long)__ldg(reinterpret_cast<const int*>(p.sizes) + s);
            ns = ns < 0 ? 0 : (ns > p.T ? p.T : ns);
            key[c] = (kGPermBins - 1) - (int)(ns * (kGPermBins - 1) / p.T);  // most objects -> bin 0
        }
    }
    for (int i = tid; i < kGWarps * kGPermBins / 2; i += kGThreads) reinterpret_cast<unsigned*>(H)[i] = 0u;
    __syncthreads();
    unsigned short* const myH = H + warp * kGPermBins;
#pragma unroll
    for (int c = 0; c < PERW; ++c) {
        const unsigned act = __ballot_sync(0xffffffffu, key[c] >= 0);
        if (key[c] >= 0) {
            const unsigned peers = __match_any_sync(act, key[c]);
            if (lane == __ffs(peers) - 1) myH[key[c]] = (unsigned short)(myH[key[c]] + __popc(peers));
        }
        __syncwarp();
    }
    __syncthreads();
    {  // counts -> start offsets: thread b owns bin b (kGPermBins <= 2 * kGThreads: two bins per thread when 4 warps)
        constexpr int BPT = (kGPermBins + kGThreads - 1) / kGThreads;
        int tot[BPT], sum = 0;
#pragma unroll
        for (int j = 0; j < BPT; ++j) {
            const int b = tid * BPT + j;
            tot[j] = 0;
            if (b < kGPermBins)
                for (int w = 0; w < kGWarps; ++w) tot[j] += H[w * kGPermBins + b];
            sum += tot[j];
        }
        int incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int run = incl - sum;
        for (int w = 0; w < warp; ++w) run += wsum[w];
#pragma unroll
        for (int j = 0; j < BPT; ++j) {
            const int b = tid * BPT + j;
            if (b < kGPermBins)
                for (int w = 0; w < kGWarps; ++w) {
                    const int c = H[w * kGPermBins + b];
                    H[w * kGPermBins + b] = (unsigned short)run;
                    run += c;
                }
        }
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < PERW; ++c) {
        const unsigned act = __ballot_sync(0xffffffffu, key[c] >= 0);
        if (key[c] >= 0) {
            const unsigned peers = __match_any_sync(act, key[c]);
            const int base = myH[key[c]];
            s_perm[base + __popc(peers & ((1u << lane) - 1u))] = (unsigned short)(i0 + c * 32 + lane);
            __syncwarp(peers);
            if (lane == __ffs(peers) - 1) myH[key[c]] = (unsigned short)(base + __popc(peers));
        }
        __syncwarp();
    }
    __syncthreads();  // the scratch goes back to being staging memory
}

template <int SRC>
__device__ __forceinline__ int unit_header(const RenderParams& p, long long u, int G, UnitSrc& us, int& uy0, int& uy1,
                                           bool& count_skips, const unsigned short* __restrict__ s_perm = nullptr) {
    int ug, bpu;
    if (u < p.tier1_units) {
        us.plane = u / p.units_per_plane;
        ug = (int)(u - us.plane * p.units_per_plane);
        bpu = p.bands_per_unit;
    } else {  // tail tier: smaller units, handed out last
        const long long u2 = u - p.tier1_units;
        const long long pl2 = u2 / p.units_per_plane2;
        us.plane = p.tier1_planes + pl2;
        ug = (int)(u2 - pl2 * p.units_per_plane2);
        bpu = p.bands_per_unit2;
    }
    uy0 = ug * bpu * G;
    uy1 = min(uy0 + bpu * G, p.H);
    if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES || SRC == SRC_WEIGHTED) {
        long long sample = us.plane / p.C;  // so far `plane` is the position in the hand-out order
        us.cls = (int)(us.plane - sample * p.C);
        if (p.perm_n > 0 && sample >= p.perm_base) {
            sample = p.perm_base + s_perm[sample - p.perm_base];
            us.plane = sample * p.C + us.cls;
        }
        long long ns = p.sizes_i64 ? __ldg(reinterpret_cast<const long long*>(p.sizes) + sample)
                                   : (long long)__ldg(reinterpret_cast<const int*>(p.sizes) + sample);
        ns = ns < 0 ? 0 : (ns > p.T ? p.T : ns);
        us.n = (int)ns;
        us.src_base = sample * p.T;
        count_skips = (ug == 0 && us.cls == 0);
        // Scan bound. Up to one step of the block (T <= kGThreads) it is T itself, which does not depend on a load: every
        // slot load is issued at once. Wider slot arrays are mostly padding in practice (500 slots for ~50 objects):
        // there the bound follows the sample's size, rounded up to whole steps — slots from us.n on are never drawn and
        // never counted (cuh:84) — at the price of waiting for the size before the first slot load.
        if (p.T <= kGThreads) return p.T;
        const int steps = (us.n + kGThreads - 1) / kGThreads;
        return min(p.T, steps * kGThreads);
    } else if constexpr (SRC == SRC_CSR) {
        const int o0 = p.csr_off[us.plane], o1 = p.csr_off[us.plane + 1];
        us.src_base = o0;
        us.n = o1 - o0;
        count_skips = (ug == 0);
        return us.n;
    } else {
        us.n = (int)p.num_objects;
        count_skips = (u == 0);
        return us.n;
    }
}

// K1 + cull for slot t of the unit's source: true if the slot is an object of this plane that overlaps rows
// [uy0, uy1). For the batched sources every load of the slot is issued before anything depends on it (one global
// round trip per scan step instead of size -> label -> box); padded slots (t >= n) are valid memory in the RaggedBatch
// layout and are dropped after the fact (cuh:84).
template <int SRC>
__device__ __forceinline__ bool scan_slot(const RenderParams& p, const UnitSrc& us, int t, int nscan, int uy0, int uy1, int W,
                                          bool count_skips, ObjI& o) {
    bool pred = false;
    if constexpr (SRC == SRC_WEIGHTED) {  // integer window + per-object weight prepared by gs_prepare_kernel
        if (t < nscan) {
            const long long i = us.src_base + t;
            int lab = us.cls;
            if (p.labels != nullptr) lab = __ldg(p.labels + i);
            const int2 c = ld_center_i(p, i);
            const int r = __ldg(p.radii + i);
            const float w = __ldg(p.weights + i);
            if (t < us.n && lab == us.cls && r >= 0) {
                const long long X = c.x, Y = c.y, R = r;
                if (!(Y + R < uy0 || Y - R >= uy1 || X + R < 0 || X - R >= W)) {
                    o.x = c.x;
                    o.y = c.y;
                    o.r = r;
                    o.w = w;
                    pred = true;
                }
            }
        }
    } else if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
        if (t < nscan) {
            const long long i = us.src_base + t;
            int lab = us.cls;
            if (p.labels != nullptr) lab = __ldg(p.labels + i);
            int x = 0, y = 0, r = -1;
            if constexpr (SRC == SRC_BOXES) {
                const float2 c = ld_center_f(p, i);
                const float4 b = ld_box_f(p, i);
                if (lab == us.cls || (count_skips && p.status)) radius_center_one(c.x, c.y, b.x, b.y, b.z, b.w, p.osf, x, y, r);
            } else {
                const int2 c = ld_center_i(p, i);
                r = __ldg(p.radii + i);
                x = c.x;
                y = c.y;
            }
            if (t < us.n) {
                if (lab == us.cls)
                    pred = make_obji(x, y, r, p.factor, uy0, uy1, W, o, p.status, count_skips);
                else if (count_skips && p.status) {  // the counting unit sees every slot once: other classes' skips too
                    if (lab < 0 || lab >= p.C) atomicAdd(&p.status->skipped_label, 1u);
                    else if (r > DH_MAX_RADIUS) atomicAdd(&p.status->skipped_radius, 1u);
                }
            }
        }
    } else if (t < us.n) {
        int x, y, r;
        if (load_slot<SRC>(p, us.src_base, t, us.cls, us.plane, count_skips, x, y, r))
            pred = make_obji(x, y, r, p.factor, uy0, uy1, W, o, p.status, count_skips);
    }
    return pred;
}

// groups (relative to the unit) an object overlaps, as a bit range
__device__ __forceinline__ unsigned long long group_bits(const ObjI& o, int uy0, int uy1, int G) {
    const int lo = (max(o.y - o.r, uy0) - uy0) / G, hi = (min(o.y + o.r, uy1 - 1) - uy0) / G;
    return ((2ull << (hi - lo)) - 1ull) << lo;
}

// K4a for the groups whose bit is set in `z` (untouched ones): each maximal run leaves as bulk stores of the zero tile.
// Called by lanes in parallel: lane `g` starts the run that begins at group g (and g + 32).
__device__ __forceinline__ void zero_runs(unsigned long long z, int lane, char* __restrict__ plane_b, int uy0, int uy1,
                                          int W, int G, int es, const float4* zero_tile) {
    bool any = false;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int g = lane + 32 * h;
        const bool start = ((z >> g) & 1ull) && (g == 0 || !((z >> (g - 1)) & 1ull));
        if (start) {
            const unsigned long long inv = ~(z >> g);
            const int run = inv ? __ffsll((long long)inv) - 1 : 64 - g;
            const int r0 = uy0 + g * G, r1 = min(r0 + run * G, uy1);
            char* dst = plane_b + (long long)r0 * W * es;
            int bytes = (r1 - r0) * W * es;
            while (bytes > 0) {
                const int c = min(bytes, kGZeroBytes);
                bulk_store(dst, zero_tile, (unsigned)c);
                dst += c;
                bytes -= c;
            }
            any = true;
        }
    }
    if (any) bulk_commit();
}

// Warp-reduce the per-lane digests, combine the warps in shared memory and issue ONE triple of global atomics per
// CTA (the three words are shared by the whole grid). Contains a block barrier: every thread of the CTA calls it.
__device__ __forceinline__ void digest_flush(Digest& dg, unsigned long long* out3, int tid, int lane, int warp) {
    unsigned long long cnt = dg.cnt;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        dg.bits += __shfl_xor_sync(0xffffffffu, dg.bits, d);
        dg.hash += __shfl_xor_sync(0xffffffffu, dg.hash, d);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    }
    __shared__ unsigned long long s_dg[3][kGWarps];
    if (lane == 0) {
        s_dg[0][warp] = dg.bits;
        s_dg[1][warp] = dg.hash;
        s_dg[2][warp] = cnt;
    }
    __syncthreads();
    if (tid < 3 && out3 != nullptr) {
        unsigned long long t = 0;
#pragma unroll
        for (int w = 0; w < kGWarps; ++w) t += s_dg[tid][w];
        if (t != 0ull) atomicAdd(out3 + tid, t);
    }
}

// The kernel: one CTA per unit (flat grid), resident CTAs take over not-yet-launched ones through cluster launch
// control. Per unit all 8 warps build the list together (three block barriers), warp 7 sends the untouched runs off,
// and the touched groups are rendered striped over the warps (group g -> warp g % 8) with no block barrier.
// Tried and measured slower on B200 (profiles/r01/experiments.md): claiming groups dynamically with shared-memory
// atomics (+23 % instructions, -15 % throughput) and a producer/consumer warp pipeline over two list stages (equal at
// best, 10-25 % slower for units under 128 KB).
#ifdef DH_TRACE_UNITS  // profiling builds only (profiles/scripts/unit_trace.py): per-unit start / end / SM / list size
__device__ unsigned long long* g_trace = nullptr;
__device__ long long g_trace_units = 0;
#endif

}  // namespace dh
#include "dh_tiles_kernel.cuh"  // render_unit_dense: what a DENSE instantiation does with a unit that many objects land on
namespace dh {

// DENSE (selected by the host for T >= 256 padded slots — flat entry: from 16 objects per plane on average — and k >= 0):
// the instantiation
// also contains the register-tile renderer of dh_tiles_kernel.cuh, and each CTA decides PER UNIT, from the list it has
// just built, which renderer the unit gets (see that file's header). Everything else is the plain kernel.
template <bool ACCUM, int SRC, bool DIGEST, int OUT = OUT_F32, bool LEAN = false, bool DENSE = false>
__global__ void __launch_bounds__(kGThreads, (ACCUM || LEAN) ? DH_G_MINB_ACCUM : DH_G_MINB) groups_kernel(const RenderParams p) {
    static_assert(!LEAN || (!ACCUM && OUT == OUT_F32), "the lean layout is a FRESH fp32 variant");
    static_assert(!DENSE || (!LEAN && (SRC == SRC_BATCHED || SRC == SRC_BOXES || SRC == SRC_CSR)),
                  "the per-unit dense renderer serves the RaggedBatch sources and the flat entry's per-plane lists");
    constexpr int ZB = ACCUM ? 0 : kGZeroBytes;    // bytes of the zero tile
    constexpr int NB = (ACCUM || LEAN) ? 1 : 2;    // staging buffers per warp
    constexpr int ES = (OUT == OUT_F32) ? 4 : 2;  // bytes per heatmap element
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* const zero_tile = reinterpret_cast<float4*>(smem_raw);
    const int stage = p.stage_bytes;  // G*W*4 rounded up to 128
    ObjI* const s_unit = reinterpret_cast<ObjI*>(smem_raw + ZB + kGWarps * NB * stage);
    // list size / occupancy mask of the unit, double-buffered by unit parity: thread 0 clears the other slot while
    // this unit is rendered, so no barrier is needed between "cleared" and "scan" (round 1 had one: three per unit)
    __shared__ int s_cnt2[2];
    __shared__ unsigned long long s_mask2[2];
    __shared__ unsigned s_cost2[2];  // DENSE: what the staged paths would roughly spend on the unit's objects (in units of 64)
    __shared__ __align__(16) uint4 s_clc;              // cluster-launch-control response
    __shared__ __align__(8) unsigned long long s_bar;  // its mbarrier
    __shared__ unsigned short s_perm[kGPermCap];       // heaviest-first hand-out order (see build_perm)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool kneg0 = p.k < 0.0f;
    const int W = p.W, G = p.band_cols /* rows per group */;
    const long long plane_elems = (long long)p.H * W;
    {  // Prologue. Only the 8 KB zero tile is cleared by the whole block before the first unit: the untouched runs of
       // the first unit (the bulk of its bytes) can then leave as soon as its list is built. Every warp clears its own
       // two staging buffers later, right before it renders its first touched group (off the path to the first store).
        for (int i = tid; i < ZB / 16; i += kGThreads) zero_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid == 0) {
            mbar_init(&s_bar, 1);
            s_cnt2[0] = s_cnt2[1] = 0;
            s_mask2[0] = s_mask2[1] = 0ull;
            s_cost2[0] = s_cost2[1] = 0u;
        }
        fence_proxy_async();  // zero tile and mbarrier -> visible to the async proxy
    }
    WarpStage ws;
    ws.buf0 = reinterpret_cast<unsigned*>(smem_raw + ZB + warp * NB * stage);
    ws.buf1 = ws.buf0 + (NB - 1) * (stage / 4);  // ACCUMULATE / LEAN: one buffer (ws.cur never flips there)
    ws.single = LEAN;
    bool stage_clean = false;  // my staging buffers are all-zero (established before my first render_group)
    unsigned clc_phase = 0;
    Digest dg;
    // Programmatic dependent launch: everything above overlaps the tail of the previous kernel in the stream when this
    // launch carries cudaLaunchAttributeProgrammaticStreamSerialization; global memory (inputs may be that kernel's
    // outputs, the heatmap may be its input) is touched only after the wait. Without the attribute both are no-ops.
    // The kernel that FOLLOWS may start its own prologue as soon as every CTA of this grid is past this point.
    pdl_wait();
    pdl_launch_dependents();
    if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES || SRC == SRC_WEIGHTED) {
        if (p.perm_n > 0) build_perm(p, smem_raw + ZB, s_perm, tid, lane, warp);
    }

    long long u = blockIdx.x;
    unsigned par = 0;  // parity of the unit in this CTA's sequence
    for (;; par ^= 1u) {
        int& s_cnt = s_cnt2[par];
        unsigned long long& s_mask = s_mask2[par];
        UnitSrc us;
        int uy0, uy1;
        bool count_skips;
        const int nscan = unit_header<SRC>(p, u, G, us, uy0, uy1, count_skips, s_perm);
        char* const plane_b = reinterpret_cast<char*>(p.heat) + us.plane * plane_elems * ES;
        const int ngroups = (uy1 - uy0 + G - 1) / G;
        float kk = p.k;
        bool kneg = kneg0;
        if constexpr (SRC == SRC_WEIGHTED) {
            kk = __ldg(p.k_classes + us.cls);
            kneg = kk < 0.0f;
        }

        __syncthreads();  // every warp is done with the previous unit's list (and shared memory is initialised)
#ifdef DH_TRACE_UNITS
        if (tid == 0 && g_trace && u < g_trace_units) {
            unsigned long long t;
            unsigned sm;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            asm volatile("mov.u32 %0, %smid;" : "=r"(sm));
            g_trace[u * 4 + 0] = t;
            g_trace[u * 4 + 2] = sm;
        }
#endif
        if (tid == 0 && p.use_clc == 2) {  // A/B only: ask for the next unit at the START of this one (round 1 / early round 2)
            fence_proxy_async();
            mbar_expect_tx(&s_bar, 16);
            clc_try_cancel(&s_clc, &s_bar);
        }
        // K1/K2: unit list (ballot + popc-prefix compaction) and group-occupancy mask
        for (int base = 0; base < nscan; base += kGThreads) {
            ObjI o;
            const bool pred = scan_slot<SRC>(p, us, base + tid, nscan, uy0, uy1, W, count_skips, o);
            const unsigned long long bits = pred ? group_bits(o, uy0, uy1, G) : 0ull;
            const unsigned bal = __ballot_sync(0xffffffffu, pred);
            if (bal) {
                const unsigned blo = __reduce_or_sync(0xffffffffu, (unsigned)bits);
                const unsigned bhi = __reduce_or_sync(0xffffffffu, (unsigned)(bits >> 32));
                int pos = 0;
                unsigned cost = 0u;
                if constexpr (DENSE) {  // d (d + 64) / 64, d = window side (clamped): rows x (fixed + per-chunk) work of a
                    // staged object. 32-bit on purpose: a 64-bit shared-memory add is a compare-and-swap loop.
                    const unsigned d = pred ? (unsigned)min(2 * o.r + 1, 1024) : 0u;
                    cost = __reduce_add_sync(0xffffffffu, (d * (d + 64u)) >> 6);
                }
                if (lane == 0) {
                    atomicOr(&s_mask, ((unsigned long long)bhi << 32) | blo);
                    pos = atomicAdd(&s_cnt, __popc(bal));
                    if constexpr (DENSE) atomicAdd(&s_cost2[par], cost);
                }
                pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(bal & ((1u << lane) - 1u));
                if (pred && pos < kGCap) s_unit[pos] = o;
            }
        }
        __syncthreads();
        const int ucnt = s_cnt;
        const unsigned long long mask = s_mask;
        if (tid == 0) {  // the other slot was last read before this unit's first barrier: clear it for the next unit
            s_cnt2[par ^ 1u] = 0;
            s_mask2[par ^ 1u] = 0ull;
            s_cost2[par ^ 1u] = 0u;
        }
        bool dense_done = false;
        if constexpr (DENSE) {
            // Register tiles for this unit (block-uniform test) when the staged paths would spend more on its objects than
            // the tile renderer spends on every pixel of the unit: sum of d (d + 64) over the objects against a multiple
            // of the unit's area (measured breakevens, profiles/r02/experiments.md section 13), and never for a handful
            // of objects.
            const unsigned long long ucost = s_cost2[par];
            if (ucnt >= p.dense_min && ucost * 1024ull >= (unsigned long long)p.dense_cost16 * (unsigned)((uy1 - uy0) * W) &&
                stage >= kTQueueBytes) {
                if constexpr (!ACCUM) {  // the copy engine may still be reading my staging buffer (an earlier unit's group)
                    if (lane == 0) bulk_wait_read<0>();
                    __syncwarp();
                }
                // The queue lives in the staging buffer: all-zero again before the next staged group, which also makes the
                // staging state start over — so it is dead across the tile sweep, whose 16 accumulators need the registers
                // (kept alive, the allocator spilled it inside the per-group path of the sparse units).
                stage_clean = false;
                ws.cur = 0;
                ws.c0x0 = ws.c1x0 = 1;
                ws.c0x1 = ws.c1x1 = 0;
                render_unit_dense<ACCUM, SRC, DIGEST, OUT>(p, us, nscan, ucnt, uy0, uy1, W, p.k, plane_b,
                                                           p.digest_base + us.plane * plane_elems, s_unit,
                                                      reinterpret_cast<unsigned short*>(ws.buf0), tid, lane, warp, dg);
                dense_done = true;
            }
        }
        if (!dense_done) {
        const bool small = ucnt <= 32;  // one list entry per lane, kept in registers for the whole unit
        ObjI mine;
        mine.x = mine.y = 0;
        mine.r = -1;
        mine.w = 0.f;
        if (small && lane < ucnt) mine = s_unit[lane];

#ifndef DH_ACCUM_PREFETCH
#define DH_ACCUM_PREFETCH 1  // 0 off, 1 into L2, 2 into L1 (measured: L2 0.685 ms, L1 0.685 ms, off 0.862 ms on half of cfg5)
#endif
        if constexpr (ACCUM && DH_ACCUM_PREFETCH != 0) {
            // ACCUMULATE reads before it writes, and a warp walks its rows one after the other: every row used to pay a
            // full DRAM round trip (units with 4-15 objects took 21 us against 14 us in FRESH mode, which moves 10 x
            // the bytes). All touched lines of the unit are requested up front instead — objects dealt out to the
            // warps, one row per lane — so the render's loads find them on chip. Only for units of <= 32 objects: with
            // hundreds of overlapping objects (cfg4) the same lines would be requested over and over (+8 %).
            const int np = small ? ucnt : 0;
            for (int e = warp; e < np; e += kGWarps) {
                const ObjI o = s_unit[e];  // broadcast
                const int ry0 = max(o.y - o.r, uy0), ry1 = min(o.y + o.r, uy1 - 1);
                const int x0 = max(o.x - o.r, 0), x1 = min(o.x + o.r, W - 1);
                for (int y = ry0 + lane; y <= ry1; y += 32) {
                    const float* const rp = reinterpret_cast<const float*>(plane_b) + (long long)y * W;
                    for (int x = x0; x <= x1 + 31; x += 32) {  // a line holds 32 pixels; the last step covers x1
                        const float* const a = rp + (x < x1 ? x : x1);
                        if (DH_ACCUM_PREFETCH == 1) prefetch_l2(a);
                        else prefetch_l1(a);
                    }
                }
            }
        }
        if constexpr (!ACCUM) {  // K4a: untouched runs, by warp 7 (it gets the fewest striped groups in a short unit)
            if (warp == kGWarps - 1) {
                const unsigned long long valid = ngroups >= 64 ? ~0ull : ((1ull << ngroups) - 1ull);
                zero_runs(~mask & valid, lane, plane_b, uy0, uy1, W, G, ES, zero_tile);
            }
        }
        // K3/K4b: touched groups, striped over the warps; no block barrier in here
        const long long plane_index0 = p.digest_base + us.plane * plane_elems;
        unsigned long long gm = mask & (kGStripe << warp);
        if (gm != 0ull && !stage_clean) {  // first touched group of this warp: establish the all-zero invariant
            uint4* const b4 = reinterpret_cast<uint4*>(ws.buf0);
            for (int i = lane; i < (NB * stage) / 16; i += 32) b4[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncwarp();
            stage_clean = true;
        }
        while (gm) {
            const int g = __ffsll((long long)gm) - 1;
            gm &= gm - 1ull;
            const int y0 = uy0 + g * G;
            render_group<ACCUM, SRC, DIGEST, OUT>(p, ws, plane_b, plane_index0, y0, min(G, uy1 - y0), W, G, kneg, kk, lane, small,
                                                  mine, s_unit, ucnt, us, dg);
        }
        }  // !dense_done
#ifdef DH_TRACE_UNITS
        if (lane == 0 && g_trace && u < g_trace_units) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            atomicMax(&g_trace[u * 4 + 1], t);
            if (warp == 0) g_trace[u * 4 + 3] = (unsigned long long)ucnt | ((unsigned long long)__popcll(mask) << 32);
        }
#endif
        // next unit: the one the block scheduler hands over (cluster launch control), else the static stride
        if (p.use_clc) {
            // The next unit is claimed only NOW, when this one is done. Claimed at the start of a unit (so that the answer
            // arrives while it is rendered — how this kernel began) the claim of a CTA that then spends 15-20 us on a heavy
            // unit sits on its next unit all that time while other CTAs run dry: unit timelines showed the last 3 us of
            // cfg3 with 2-25 of 444 slots busy, every late unit one that a long-running CTA had reserved. Claiming late
            // makes the hand-out truly dynamic (list scheduling in heaviest-first order); the round trip to the block
            // scheduler hides behind the bulk stores still draining. cfg3 37.1 -> 34.7 us, b384 207.0 -> 205.5 us, cfg5
            // 1.566 -> 1.562 ms (profiles/r02/clc_mode_ab.txt).
            if (tid == 0 && p.use_clc != 2) {
                fence_proxy_async();
                mbar_expect_tx(&s_bar, 16);
                clc_try_cancel(&s_clc, &s_bar);
            }
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
    if constexpr (!ACCUM) bulk_wait_read<0>();  // shared memory must outlive every bulk read issued by this thread
    if constexpr (DIGEST) digest_flush(dg, p.digest, tid, lane, warp);
}

}  // namespace dh
