// dh_kernels.cu — sm_100a kernels + C ABI of the B200-native Gaussian heatmap renderer.
//
// Replaces the reference hot path (citations relative to /root/reference/packages/draw_heatmap):
//   accvlab/draw_heatmap/include/draw_heatmap_cuda_kernel.cuh:26-109  (one thread per object, global CAS atomics)
//   accvlab/draw_heatmap/csrc/draw_heatmap_cuda.cu:29-60              (launchers)
//   tests/_test_helpers.py:20-28                                      (bbox -> integer centre / radius)
//
// Design (not a port): the OUTPUT is partitioned, not the objects.
//   K1  radius_center   : one thread per box, fp32 IEEE, bit-exact vs the reference helper.
//   K2  binning         : every render unit (a group of row bands of one plane) scans the objects of its
//                         sample / plane and compacts the overlapping ones into a shared-memory list with
//                         warp ballot + popc prefix; band lists are compacted again from the unit list.
//                         The flat API additionally counting-sorts objects by plane (match_any aggregated).
//   K3  max-blend       : pixel-owner threads keep max_o(-(i^2+j^2)*vinv_o) in registers (no atomics on
//                         the heatmap, shared or global) and evaluate ONE expf per touched pixel:
//                         max_o k*exp(a_o) == k*exp(max_o a_o) bit-for-bit because expf is monotone
//                         (verified exhaustively on the GPU by
//                         tests/test_gpu_properties.py::test_expf_is_monotone_and_close_to_correctly_rounded).
//   K4  write           : every pixel is stored exactly once with 128-bit coalesced stores; zero-init is
//                         fused (FRESH mode writes zeros where nothing was drawn and never reads).
// No tensor cores: this is HBM-write-bound byte work, not a contraction.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdlib.h>
#include <stdint.h>

#include <atomic>

#include "dh_b200.h"
#include "dh_k1.cuh"

#define DH_STR2(x) #x
#define DH_STR(x) DH_STR2(x)

namespace dh {

constexpr int kThreads = 256;
constexpr int kItemsPerThread = 8;
constexpr int kBandCap = kThreads * kItemsPerThread;  // items (float4 or float) per band
constexpr int kListCap = 1024;                        // objects per shared-memory list

enum Source : int { SRC_BATCHED = 0, SRC_BOXES = 1, SRC_CSR = 2, SRC_FLATSCAN = 3, SRC_WEIGHTED = 4, SRC_BINNED = 5 };

struct __align__(16) Obj {
    float fx, fy, fr, wv;  // centre, radius (exact small ints as floats), signed 1/var
};

struct RenderParams {
    float* heat;
    const int32_t* centers;  // [.,2]   SRC_BATCHED / SRC_CSR / SRC_FLATSCAN
    const int32_t* radii;
    const float* centers_f;  // [.,2]   SRC_BOXES
    const float* bboxes_f;   // [.,4]
    float osf;
    const int32_t* labels;   // may be null (batched sources)
    const void* sizes;       // [B] int32 / int64
    int sizes_i64;
    const int32_t* plane_idx;  // [N] SRC_FLATSCAN
    const int32_t* csr_off;    // [M+1] SRC_CSR
    const int32_t* csr_items;  // [N]
    long long num_objects;     // N (flat sources)
    long long num_planes;
    int C, T, H, W;
    int in_vec;  // centers / centers_f are 8-byte and bboxes_f 16-byte aligned: slots are read with one vector load each
    int wunits;  // W / PX
    int band_rows, band_cols, bands_x, bands_y, bands_per_plane, bands_per_unit, units_per_plane;
    int stage_bytes;  // groups_kernel: bytes of one per-warp staging buffer
    unsigned long long* digest;  // optional fused result digest (FRESH groups_kernel only): bits, hash, count
    long long digest_base;       // element index of heat[0] in the unsharded result
    int use_clc;                 // groups_kernel: fetch further units through cluster launch control
    long long total_units;
    // groups_kernel, tail tier: planes >= tier1_planes are cut into smaller units (handed out last)
    long long tier1_units, tier1_planes;
    int bands_per_unit2, units_per_plane2;
    float factor, k;
    dh_status* status;
    // SRC_WEIGHTED (groups_kernel only; the DALI draw_gaussians fast path): per-object 1/(2 sigma^2) instead of the
    // radius -> sigma rule, per-class peak instead of one k
    const float* weights;    // [B,T]
    const float* k_classes;  // [C]
    // SRC_BINNED (groups2_kernel): output of bin_kernel in the caller's workspace
    // groups_kernel, heaviest-first hand-out: the last perm_n samples (from perm_base on) are rendered in descending order
    // of their object count (0 = hand-out in index order)
    int perm_n;
    long long perm_base;
    // groups_kernel<..., DENSE>: a unit is rendered as register tiles when at least dense_min objects land on it and the
    // sum of d (d + 64) over them (d = window side) reaches dense_cost16 / 16 x the unit's pixel count
    int dense_min, dense_cost16;
    const void* bin_list;  // int4 (x, y, r, bits(1/var)) per object, a plane's objects contiguous
    const void* bin_rec;   // int2 (first entry, count) per plane
};

// ------------------------------------------------------------------------------------------------
// K1 arithmetic: dh_k1.cuh (shared with dh_pack.cu)
__global__ void __launch_bounds__(256) radius_center_kernel(const float* __restrict__ centers,
                                                            const float* __restrict__ bboxes, float osf, long long n,
                                                            int32_t* __restrict__ centers_out,
                                                            int32_t* __restrict__ radii_out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int xi, yi, r;
        radius_center_one(centers[2 * i], centers[2 * i + 1], bboxes[4 * i], bboxes[4 * i + 1], bboxes[4 * i + 2],
                          bboxes[4 * i + 3], osf, xi, yi, r);
        centers_out[2 * i] = xi;
        centers_out[2 * i + 1] = yi;
        radii_out[i] = r;
    }
}

// ------------------------------------------------------------------------------------------------
// Object -> list entry, culled against a pixel rectangle rows [ry0,ry1) x cols [rx0,rx1).
// sigma/var arithmetic restates cuh:61-63 / 90-92 with explicit IEEE intrinsics.
__device__ __forceinline__ bool make_obj(int x, int y, int r, float factor, bool kneg, int ry0, int ry1, int rx0, int rx1,
                                         Obj& o, dh_status* st, bool count) {
    if (r < 0) return false;  // reference: empty loops
    if (r > DH_MAX_RADIUS) {
        if (count && st) atomicAdd(&st->skipped_radius, 1u);
        return false;
    }
    const long long X = x, Y = y, R = r;
    if (Y + R < ry0 || Y - R >= ry1 || X + R < rx0 || X - R >= rx1) return false;
    const int d = 2 * r + 1;
    const float sigma = __fdiv_rn((float)d, factor);
    const float var = __fmul_rn(__fmul_rn(2.0f, sigma), sigma);
    const float vinv = __fdiv_rn(1.0f, var);
    o.fx = (float)x;
    o.fy = (float)y;
    o.fr = (float)r;
    o.wv = kneg ? vinv : -vinv;
    return true;
}

// Warp-ballot + popc-prefix compaction of `pred` lanes into list[] (capacity kListCap); *cnt in shared memory.
__device__ __forceinline__ void compact_push(bool pred, const Obj& o, Obj* list, int* cnt) {
    const unsigned bal = __ballot_sync(0xffffffffu, pred);
    if (bal) {
        const int lane = threadIdx.x & 31;
        int base = 0;
        if (lane == 0) base = atomicAdd(cnt, __popc(bal));
        base = __shfl_sync(0xffffffffu, base, 0);
        const int pos = base + __popc(bal & ((1u << lane) - 1u));
        if (pred && pos < kListCap) list[pos] = o;
    }
}

template <int PX>
struct Vec;
template <>
struct Vec<4> {
    using type = float4;
};
template <>
struct Vec<1> {
    using type = float;
};

template <int PX>
__device__ __forceinline__ void store_px(float* p, const float (&v)[PX]) {
    if constexpr (PX == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
        *p = v[0];
    }
}

// Zero a rectangle rows [y0,y1) x units [cu0,cu1) of one plane (FRESH mode, nothing to draw there).
template <int PX>
__device__ __forceinline__ void zero_rect(float* plane, int W, int wunits, int y0, int y1, int cu0, int cu1) {
    using V = typename Vec<PX>::type;
    V z;
    if constexpr (PX == 4) z = make_float4(0.f, 0.f, 0.f, 0.f); else z = 0.f;
    if (cu0 == 0 && cu1 == wunits) {
        V* p = reinterpret_cast<V*>(plane + (long long)y0 * W);
        const int n = (y1 - y0) * wunits;
#pragma unroll 4
        for (int i = threadIdx.x; i < n; i += kThreads) p[i] = z;
    } else {
        const int cw = cu1 - cu0;
        for (int y = y0; y < y1; ++y) {
            V* p = reinterpret_cast<V*>(plane + (long long)y * W) + cu0;
            for (int i = threadIdx.x; i < cw; i += kThreads) p[i] = z;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2+K3+K4: persistent tile-owner renderer.
template <int PX, bool ACCUM, int SRC>
__global__ void __launch_bounds__(kThreads) render_kernel(const RenderParams p) {
    __shared__ Obj s_unit[kListCap];
    __shared__ Obj s_band[kListCap];
    __shared__ int s_cnt[2];

    const int tid = threadIdx.x;
    const float NEG_INF = __int_as_float(0xff800000);
    const bool kneg = p.k < 0.0f;
    const long long plane_elems = (long long)p.H * p.W;

    for (long long u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const long long plane = u / p.units_per_plane;
        const int ug = (int)(u - plane * p.units_per_plane);
        float* const plane_ptr = p.heat + plane * plane_elems;

        // unit rectangle
        int uy0, uy1, ucu0, ucu1, b0, b1;
        if (p.bands_x == 1) {
            b0 = ug * p.bands_per_unit;
            b1 = min(b0 + p.bands_per_unit, p.bands_y);
            uy0 = b0 * p.band_rows;
            uy1 = min(b1 * p.band_rows, p.H);
            ucu0 = 0;
            ucu1 = p.wunits;
        } else {  // one band per unit, a column segment of one row group
            b0 = ug;
            b1 = ug + 1;
            const int by = ug / p.bands_x, bx = ug - by * p.bands_x;
            uy0 = by * p.band_rows;
            uy1 = min(uy0 + p.band_rows, p.H);
            ucu0 = bx * p.band_cols;
            ucu1 = min(ucu0 + p.band_cols, p.wunits);
        }

        // ---- source range of this unit --------------------------------------------------------
        long long src_base = 0;  // first object slot
        int n = 0;               // slots to scan
        int cls = 0;
        bool count_skips = false;
        if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
            const long long sample = plane / p.C;
            cls = (int)(plane - sample * p.C);
            long long ns = p.sizes_i64 ? reinterpret_cast<const long long*>(p.sizes)[sample]
                                       : (long long)reinterpret_cast<const int*>(p.sizes)[sample];
            ns = ns < 0 ? 0 : (ns > p.T ? p.T : ns);
            n = (int)ns;
            src_base = sample * p.T;
            count_skips = (ug == 0 && cls == 0);
        } else if constexpr (SRC == SRC_CSR) {
            const int o0 = p.csr_off[plane], o1 = p.csr_off[plane + 1];
            src_base = o0;
            n = o1 - o0;
            count_skips = (ug == 0);
        } else {  // SRC_FLATSCAN: every unit looks at all N objects
            n = (int)p.num_objects;
            count_skips = (u == 0);
        }

        // Load slot t (t < n) and cull it against a pixel rectangle.
        auto load_obj = [&](int t, int ry0, int ry1, int rx0, int rx1, bool cnt, Obj& o) -> bool {
            int x, y, r;
            if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
                const long long i = src_base + t;
                if (p.labels != nullptr) {
                    const int lab = p.labels[i];
                    if (lab != cls) {
                        if (cnt && p.status && (lab < 0 || lab >= p.C)) atomicAdd(&p.status->skipped_label, 1u);
                        return false;
                    }
                }
                if constexpr (SRC == SRC_BOXES) {
                    radius_center_one(p.centers_f[2 * i], p.centers_f[2 * i + 1], p.bboxes_f[4 * i], p.bboxes_f[4 * i + 1],
                                      p.bboxes_f[4 * i + 2], p.bboxes_f[4 * i + 3], p.osf, x, y, r);
                } else {
                    x = p.centers[2 * i];
                    y = p.centers[2 * i + 1];
                    r = p.radii[i];
                }
            } else if constexpr (SRC == SRC_CSR) {
                const long long i = p.csr_items[src_base + t];
                x = p.centers[2 * i];
                y = p.centers[2 * i + 1];
                r = p.radii[i];
            } else {
                const int pi = p.plane_idx[t];
                if ((long long)pi != plane) {
                    if (cnt && p.status && (pi < 0 || (long long)pi >= p.num_planes))
                        atomicAdd(&p.status->skipped_plane, 1u);
                    return false;
                }
                x = p.centers[2 * (long long)t];
                y = p.centers[2 * (long long)t + 1];
                r = p.radii[t];
            }
            return make_obj(x, y, r, p.factor, kneg, ry0, ry1, rx0, rx1, o, p.status, cnt);
        };

        // ---- K2a: unit list (ballot + prefix compaction) --------------------------------------
        __syncthreads();  // previous unit is completely done with the shared lists
        if (tid == 0) s_cnt[0] = 0;
        __syncthreads();
        for (int base = 0; base < n; base += kThreads) {
            const int t = base + tid;
            Obj o;
            bool pred = false;
            if (t < n) pred = load_obj(t, uy0, uy1, ucu0 * PX, ucu1 * PX, count_skips, o);
            compact_push(pred, o, s_unit, &s_cnt[0]);
        }
        __syncthreads();
        const int ucnt = s_cnt[0];

        if (ucnt == 0) {
            if constexpr (!ACCUM) zero_rect<PX>(plane_ptr, p.W, p.wunits, uy0, uy1, ucu0, ucu1);
            continue;
        }
        const bool unit_list_ok = ucnt <= kListCap;

        // ---- bands of the unit ----------------------------------------------------------------
        for (int b = b0; b < b1; ++b) {
            int by0, by1, cu0, cu1;
            if (p.bands_x == 1) {
                by0 = b * p.band_rows;
                by1 = min(by0 + p.band_rows, p.H);
                cu0 = 0;
                cu1 = p.wunits;
            } else {
                by0 = uy0; by1 = uy1; cu0 = ucu0; cu1 = ucu1;
            }
            const int cw = cu1 - cu0;
            const int nitems = (by1 - by0) * cw;

            float acc[kItemsPerThread][PX];
            float frow[kItemsPerThread], fcol[kItemsPerThread];
            bool coords_ready = false;

            // Passes: one pass over the unit list, or (overflow) chunks of the global source.
            const int npass = unit_list_ok ? 1 : (n + kListCap - 1) / kListCap;
            bool any = false;
            for (int pass = 0; pass < npass; ++pass) {
                // ---- K2b: band list ------------------------------------------------------------
                __syncthreads();  // previous pass / band finished reading s_band
                if (tid == 0) s_cnt[1] = 0;
                __syncthreads();
                if (unit_list_ok) {
                    const float fy0 = (float)by0, fy1 = (float)by1, fx0 = (float)(cu0 * PX), fx1 = (float)(cu1 * PX);
                    for (int base = 0; base < ucnt; base += kThreads) {
                        const int e = base + tid;
                        Obj o;
                        bool pred = false;
                        if (e < ucnt) {
                            o = s_unit[e];
                            pred = (o.fy + o.fr >= fy0) && (o.fy - o.fr < fy1) && (o.fx + o.fr >= fx0) && (o.fx - o.fr < fx1);
                        }
                        compact_push(pred, o, s_band, &s_cnt[1]);
                    }
                } else {
                    const int c0 = pass * kListCap, c1 = min(n, c0 + kListCap);
                    for (int base = c0; base < c1; base += kThreads) {
                        const int t = base + tid;
                        Obj o;
                        bool pred = false;
                        if (t < c1) pred = load_obj(t, by0, by1, cu0 * PX, cu1 * PX, false, o);
                        compact_push(pred, o, s_band, &s_cnt[1]);
                    }
                }
                __syncthreads();
                const int bcnt = s_cnt[1];
                if (bcnt == 0) continue;
                any = true;

                if (!coords_ready) {
                    coords_ready = true;
                    int q = tid / cw, r = tid - q * cw;
                    const int dq = kThreads / cw, dr = kThreads - dq * cw;
#pragma unroll
                    for (int k = 0; k < kItemsPerThread; ++k) {
                        const bool valid = (k * kThreads + tid) < nitems;
                        frow[k] = valid ? (float)(by0 + q) : 3.0e38f;  // invalid items never pass the row test
                        fcol[k] = (float)((cu0 + r) * PX);
#pragma unroll
                        for (int j = 0; j < PX; ++j) acc[k][j] = NEG_INF;
                        q += dq;
                        r += dr;
                        if (r >= cw) { r -= cw; q += 1; }
                    }
                }

                // ---- K3: max-blend in registers ------------------------------------------------
                for (int e = 0; e < bcnt; ++e) {
                    const Obj o = s_band[e];  // broadcast LDS.128
#pragma unroll
                    for (int k = 0; k < kItemsPerThread; ++k) {
                        const float dy = frow[k] - o.fy;
                        if (fabsf(dy) <= o.fr) {
                            const float dx0 = fcol[k] - o.fx;
                            if (dx0 <= o.fr && dx0 + (float)(PX - 1) >= -o.fr) {
                                const float fi = __fmul_rn(dy, dy);
#pragma unroll
                                for (int j = 0; j < PX; ++j) {
                                    const float dx = dx0 + (float)j;
                                    const float t = __fmul_rn(__fadd_rn(fi, __fmul_rn(dx, dx)), o.wv);
                                    acc[k][j] = (fabsf(dx) <= o.fr) ? fmaxf(acc[k][j], t) : acc[k][j];
                                }
                            }
                        }
                    }
                }
            }

            // ---- K4: write --------------------------------------------------------------------
            if (!any) {
                if constexpr (!ACCUM) zero_rect<PX>(plane_ptr, p.W, p.wunits, by0, by1, cu0, cu1);
                continue;
            }
#pragma unroll
            for (int k = 0; k < kItemsPerThread; ++k) {
                if ((k * kThreads + tid) < nitems) {
                    float* dst = plane_ptr + (long long)((int)frow[k]) * p.W + (int)fcol[k];
                    float v[PX];
                    bool touched[PX];
                    bool any_t = false;
#pragma unroll
                    for (int j = 0; j < PX; ++j) {
                        const float m = acc[k][j];
                        touched[j] = m > NEG_INF;
                        any_t |= touched[j];
                        v[j] = 0.0f;
                        if (touched[j]) v[j] = __fmul_rn(expf(kneg ? -m : m), p.k);
                    }
                    if constexpr (ACCUM) {
                        if (any_t) {
                            float old[PX];
                            if constexpr (PX == 4) {
                                const float4 o4 = *reinterpret_cast<const float4*>(dst);
                                old[0] = o4.x; old[1] = o4.y; old[2] = o4.z; old[3] = o4.w;
                            } else {
                                old[0] = *dst;
                            }
#pragma unroll
                            for (int j = 0; j < PX; ++j)  // cuh:30: keep old iff old >= v
                                v[j] = touched[j] ? ((old[j] >= v[j]) ? old[j] : v[j]) : old[j];
                            store_px<PX>(dst, v);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < PX; ++j) v[j] = touched[j] ? ((0.0f >= v[j]) ? 0.0f : v[j]) : 0.0f;
                        store_px<PX>(dst, v);
                    }
                }
            }
        }
    }
}

}  // namespace dh
#include "dh_rows_kernel.cuh"
#include "dh_groups_kernel.cuh"
#include "dh_groups2_kernel.cuh"
namespace dh {

// ------------------------------------------------------------------------------------------------
// Flat API binning: counting sort of objects by plane (match_any-aggregated counters).
__global__ void __launch_bounds__(256) csr_count_kernel(const int32_t* __restrict__ plane_idx, long long n, long long m,
                                                        int32_t* __restrict__ counts, dh_status* status) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const int idx = (i < n) ? plane_idx[i] : -1;
    const bool valid = (i < n) && idx >= 0 && (long long)idx < m;
    if ((i < n) && !valid && status) atomicAdd(&status->skipped_plane, 1u);
    const unsigned active = __ballot_sync(0xffffffffu, valid);
    if (valid) {
        const unsigned peers = __match_any_sync(active, idx);
        if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&counts[idx], __popc(peers));
    }
}

__global__ void __launch_bounds__(1024) csr_scan_kernel(const int32_t* __restrict__ counts, long long m,
                                                        int32_t* __restrict__ off) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (long long base = 0; base < m; base += 1024) {
        const long long i = base + tid;
        const int c = (i < m) ? counts[i] : 0;
        int incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= d) w += t;
            }
            s_warp[lane] = w;  // inclusive warp sums
        }
        __syncthreads();
        const int carry = s_carry;
        const int wprefix = (warp == 0) ? 0 : s_warp[warp - 1];
        if (i < m) off[i] = carry + wprefix + incl - c;
        __syncthreads();
        if (tid == 1023) s_carry = carry + wprefix + incl;
        __syncthreads();
    }
    if (tid == 0) off[m] = s_carry;
}

__global__ void __launch_bounds__(256) csr_fill_kernel(const int32_t* __restrict__ plane_idx, long long n, long long m,
                                                       const int32_t* __restrict__ off, int32_t* __restrict__ cursor,
                                                       int32_t* __restrict__ items) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const int idx = (i < n) ? plane_idx[i] : -1;
    const bool valid = (i < n) && idx >= 0 && (long long)idx < m;
    const unsigned active = __ballot_sync(0xffffffffu, valid);
    if (valid) {
        const int lane = threadIdx.x & 31;
        const unsigned peers = __match_any_sync(active, idx);
        const int leader = __ffs(peers) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&cursor[idx], __popc(peers));
        base = __shfl_sync(peers, base, leader);
        items[off[idx] + base + __popc(peers & ((1u << lane) - 1u))] = (int32_t)i;
    }
}

// ------------------------------------------------------------------------------------------------
// Write-only ceiling probe: same store instruction and grid shape as the FRESH renderer.
__global__ void __launch_bounds__(kThreads) fill_probe_kernel(float4* __restrict__ dst, long long n4, long long chunk4) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long nchunks = (n4 + chunk4 - 1) / chunk4;
    for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
        float4* p = dst + c * chunk4;
        const int n = (int)min(chunk4, n4 - c * chunk4);
#pragma unroll 4
        for (int i = threadIdx.x; i < n; i += kThreads) p[i] = z;
    }
}

__global__ void __launch_bounds__(256) expf_probe_kernel(unsigned start_bits, long long n, float* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = expf(__uint_as_float(start_bits + (unsigned)i));
}

// ------------------------------------------------------------------------------------------------
// Host side
std::atomic<uint64_t> g_launches{0};
void count_launches(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }  // used by dh_host.cu

struct DeviceInfo {
    int sms = 0;
};

static int device_info(DeviceInfo& di) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    // per-device caches are std::atomic: several host threads may make their first call at once; a duplicated query is
    // harmless (same value), a torn or unsynchronised write would not be
    static std::atomic<int> cached_sms[64];
    if (dev < 0 || dev >= 64) return DH_ERR_NO_DEVICE;
    int sms = cached_sms[dev].load(std::memory_order_relaxed);
    if (sms == 0) {
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return (int)e;
        cached_sms[dev].store(sms, std::memory_order_relaxed);
    }
    di.sms = sms;
    return DH_OK;
}

template <typename K>
static int blocks_per_sm(K kernel, std::atomic<int>& cached, int& out) {
    out = cached.load(std::memory_order_relaxed);
    if (out == 0) {
        int nb = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, kThreads, 0);
        if (e != cudaSuccess) return (int)e;
        out = nb > 0 ? nb : 1;
        cached.store(out, std::memory_order_relaxed);
    }
    return DH_OK;
}

static int check_shape(int64_t H, int64_t W) {
    if (H < 1 || W < 1) return DH_ERR_INVALID_ARGUMENT;
    if (W > (1ll << 23) || H > (1ll << 23) || H * W >= (1ll << 31)) return DH_ERR_UNSUPPORTED_SHAPE;
    return DH_OK;
}

// Partition planes into bands (<= kBandCap items) and units (groups of bands of one plane).
static void make_plan(RenderParams& p, int px, long long max_ctas) {
    p.wunits = p.W / px;
    if (p.wunits <= kBandCap) {
        const int max_rows = kBandCap / p.wunits;
        const long long total_items = p.num_planes * (long long)p.H * p.wunits;
        long long want = total_items / (2 * max_ctas);  // >= 2 bands per resident CTA on small problems
        if (want < kThreads) want = kThreads;
        if (want > kBandCap) want = kBandCap;
        long long rows = want / p.wunits;
        if (rows < 1) rows = 1;
        if (rows > max_rows) rows = max_rows;
        if (rows > p.H) rows = p.H;
        p.bands_x = 1;
        p.bands_y = (int)((p.H + rows - 1) / rows);
        p.band_rows = (int)((p.H + p.bands_y - 1) / p.bands_y);  // balance the last band
        p.bands_y = (p.H + p.band_rows - 1) / p.band_rows;
        p.band_cols = p.wunits;
        p.bands_per_plane = p.bands_y;
        const long long total_bands = p.num_planes * p.bands_y;
        long long bpu = total_bands / (8 * max_ctas);
        if (bpu < 1) bpu = 1;
        if (bpu > p.bands_y) bpu = p.bands_y;
        p.bands_per_unit = (int)bpu;
        p.units_per_plane = (p.bands_y + p.bands_per_unit - 1) / p.bands_per_unit;
    } else {  // a single row does not fit a band: split rows into column segments
        p.band_rows = 1;
        p.band_cols = kBandCap;
        p.bands_x = (p.wunits + kBandCap - 1) / kBandCap;
        p.bands_y = p.H;
        p.bands_per_plane = p.bands_x * p.bands_y;
        p.bands_per_unit = 1;
        p.units_per_plane = p.bands_per_plane;
    }
    p.total_units = p.num_planes * p.units_per_plane;
}

// Plan of the rows kernel: band = R whole rows (<= kTileBytes, a multiple of 8 rows when possible so the 8 warps
// get equal shares), unit = up to 64 bands; enough units for >= 4 per resident CTA when the problem allows it.
static void make_rows_plan(RenderParams& p, long long max_ctas) {
    int R = kTilePx / p.W;
    if (R >= 8) R &= ~7;
    if (R > p.H) R = p.H;
    if (R < 1) R = 1;
    p.band_rows = R;
    p.bands_y = (p.H + R - 1) / R;
    p.bands_x = 1;
    p.bands_per_plane = p.bands_y;
    const long long total_bands = p.num_planes * p.bands_y;
    long long bpu = total_bands / (4 * max_ctas);
    if (bpu < 1) bpu = 1;
    if (bpu > 64) bpu = 64;
    if (bpu > p.bands_y) bpu = p.bands_y;
    p.bands_per_unit = (int)bpu;
    p.units_per_plane = (p.bands_y + p.bands_per_unit - 1) / p.bands_per_unit;
    p.bands_per_unit = (p.bands_y + p.units_per_plane - 1) / p.units_per_plane;  // balance the last unit
    p.units_per_plane = (p.bands_y + p.bands_per_unit - 1) / p.bands_per_unit;
    p.total_units = p.num_planes * p.units_per_plane;
}

template <bool ACCUM, int SRC>
static int launch_rows(cudaStream_t stream, RenderParams& p, int sms) {
    static std::atomic<int> occ_dev[64];  // per instantiation and device (function attributes are per device); also "attribute set"
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DH_ERR_NO_DEVICE;
    int occ = occ_dev[dev].load(std::memory_order_relaxed);
    if (occ == 0) {
        cudaError_t e = cudaFuncSetAttribute(rows_kernel<ACCUM, SRC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kRowsSmemBytes);
        if (e != cudaSuccess) return (int)e;
        int nb = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, rows_kernel<ACCUM, SRC>, kRowsThreads, kRowsSmemBytes);
        if (e != cudaSuccess) return (int)e;
        occ = nb > 0 ? nb : 1;
        occ_dev[dev].store(occ, std::memory_order_relaxed);
    }
    const long long max_ctas = (long long)sms * occ;
    make_rows_plan(p, max_ctas);
    if (p.total_units == 0) return DH_OK;
    const int grid = (int)(p.total_units < max_ctas ? p.total_units : max_ctas);
    rows_kernel<ACCUM, SRC><<<grid, kRowsThreads, kRowsSmemBytes, stream>>>(p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return (int)cudaGetLastError();
}

// Tuning constants. Product builds compile them in; measurement builds (-DDH_TUNING, profiles/scripts/build_variants.sh)
// read the environment once at load time instead, for A/B runs of one binary. The defaults are the measured optima
// (profiles/r01/experiments.md, profiles/r02/experiments.md).
#ifdef DH_TUNING
static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
#define DH_KNOB(name, dflt) env_int(name, dflt)
#else
#define DH_KNOB(name, dflt) (dflt)
#endif
static const int g_unit_bytes = DH_KNOB("DH_G_UNIT_BYTES", 196608);    // target bytes per unit
static const int g_persistent = DH_KNOB("DH_G_PERSISTENT", 0);         // 1: static grid-stride launch (block-scheduler A/B)
static const int g_no_clc = DH_KNOB("DH_G_NOCLC", 0);                  // 1: no cluster launch control
static const int g_single_wave = DH_KNOB("DH_G_SINGLE_WAVE", 1);       // small problems: one wave of larger units
static const int g_heavy_per_slot = DH_KNOB("DH_G_HEAVY_PER_SLOT", 8); // units per resident CTA slot when T >= 256
static const int g_lean = DH_KNOB("DH_G_LEAN", 1);                     // object-heavy FRESH renders: lean shared-memory layout
static const int g_lean_t = DH_KNOB("DH_G_LEAN_T", 256);                // ... from this many padded targets per sample on
static const int g_tiles = DH_KNOB("DH_G_TILES", 1);                   // T >= g_tiles_t, k >= 0: the instantiation with the per-unit register-tile renderer
static const int g_tiles_t = DH_KNOB("DH_G_TILES_T", 256);              // ... from this many padded targets per sample on
static const int g_dense_min = DH_KNOB("DH_G_DENSE_MIN", 8);           // ... which a unit takes when its list holds at least this many objects
static const int g_dense_cost16 = DH_KNOB("DH_G_DENSE_COST16", 48);    // ... whose staged-path cost estimate reaches this / 16 x the unit's pixels (FRESH)
static const int g_dense_cost16_accum = DH_KNOB("DH_G_DENSE_COST16_ACCUM", 112);  // ... the same for ACCUMULATE (measured breakevens: ~3 and ~7)
static const int g_no_bin = DH_KNOB("DH_G_NOBIN", 0);                  // 1: ignore the workspace of the *_ws entries
static const int g_lpt = DH_KNOB("DH_G_LPT", 10);                      // heaviest-first hand-out inside the last g_lpt / 2 waves (0 = off)
static const int g_clc_mode = DH_KNOB("DH_G_CLC_MODE", 1);             // when a CTA claims its next unit: 1 when the current one is done, 2 at its start (A/B)
static const int g_pdl = DH_KNOB("DH_G_PDL", 1);                       // programmatic dependent launch of the groups kernels
static const int g_tail_div = DH_KNOB("DH_G_TAIL_DIV", 4);             // tail-tier units are 1/this of a full unit (1 = off); 2 until units were claimed late, then re-measured
static const int g_tail_cap_div = DH_KNOB("DH_G_TAIL_CAP_DIV", 2);     // ... but at most 1/this of all planes
static const int g_tail_waves_x2 = DH_KNOB("DH_G_TAIL_WAVES_X2", 2);   // tail tier = this/2 waves of full-size units

// Plan of the groups kernels: row group = G whole rows (<= kGStageCap bytes), unit = up to 64 groups of one plane,
// ~192 KB when the problem is large (amortises the per-unit list build; with the tail tier of half-size units the measured
// optimum on cfg5 is 2 units per plane: 1.600 vs 1.631 ms for 3), smaller when it
// is not, so that every resident CTA slot still sees several units.
// `es` = bytes per heatmap element (4, or 2 for the 16-bit planes: rows per group still follow the fp32 scratch).
// Returns false when 16-bit planes cannot be cut into 16-byte aligned groups.
static bool make_groups_plan(RenderParams& p, long long max_ctas, int es = 4, bool accum = false, bool lean = false,
                             bool dense = false) {
    // rows per group: as many as a staging buffer holds while DH_G_MINB(_ACCUM) CTAs still fit the SM's shared memory
    const bool one = accum || lean;  // one staging buffer per warp, four CTAs per SM
    const long long smem_budget = (227ll * 1024) / (one ? DH_G_MINB_ACCUM : DH_G_MINB) - 1024 - kGPermSmem;  // 1 KB: the system's
    const long long stage_budget = (smem_budget - (long long)groups_smem_bytes(0, accum, lean)) / ((one ? 1 : 2) * kGWarps);
    int G = (int)((stage_budget < kGStageCap ? stage_budget : kGStageCap) / (p.W * 4));
    if (G < 1) G = 1;
    if (G > p.H) G = p.H;
    if (es == 2) {  // bulk stores move multiples of 16 bytes from 16-byte aligned addresses
        if (((long long)p.H * p.W) % 8 != 0) return false;
        if (p.W % 8 != 0) G &= ~1;
        if (G < 1) return false;
    }
    p.band_cols = G;  // rows per group
    p.band_rows = G;
    const int groups_y = (p.H + G - 1) / G;
    p.bands_y = groups_y;
    p.bands_x = 1;
    p.bands_per_plane = groups_y;
    const long long group_bytes = (long long)G * p.W * es;
    const long long total_groups = p.num_planes * groups_y;
    // units per resident CTA slot: 4, or 8 when samples can hold many objects (units are then compute-heavy and of
    // similar cost, so the last, partly filled wave is expensive: cfg4 ran 4.1 waves of 1024-object units)
    // (DENSE instantiations render an object-heavy unit with register tiles, whose cost is the same for every strip, and
    // pay one scan of the sample's slots per unit: they take the ordinary plan — cfg4: one wave of 40-row units 0.079 ms,
    // 16-row units in 2.3 waves 0.081, 8-row units in 4.6 waves 0.086.)
    const bool heavy = p.T >= 256 && !dense;
    const long long per_slot = heavy ? g_heavy_per_slot : 4;
    long long gpu = total_groups / (per_slot * max_ctas);
    const long long cap = (g_unit_bytes + group_bytes - 1) / group_bytes;
    if (gpu > cap) gpu = cap;
    // Small problems (fewer planes than resident CTA slots, not object-heavy): ONE wave. Every unit pays ~3 us of pure
    // latency, so a second, partly filled wave of small units costs more than fewer, larger units in a single wave.
    if (g_single_wave && !heavy && p.num_planes <= max_ctas) {
        long long upp = max_ctas / p.num_planes;
        if (upp > groups_y / kGWarps) upp = groups_y / kGWarps;
        if (upp < 1) upp = 1;
        const long long one_wave = (groups_y + upp - 1) / upp;
        if (one_wave <= cap && one_wave > gpu) gpu = one_wave;
    }
    if (dense) {  // units of whole 8-row strips (the tile height of the per-unit dense renderer), at least one
        long long rows = ((gpu * G + kTRows - 1) / kTRows) * kTRows;
        if (rows < kTRows) rows = kTRows;
        gpu = (rows + G - 1) / G;
    } else if (gpu < kGWarps) {
        gpu = kGWarps;  // at least one group per warp, however small the problem
    }
    if (gpu > kGMaxGroups) gpu = kGMaxGroups;
    if (gpu > groups_y) gpu = groups_y;
    p.units_per_plane = (int)((groups_y + gpu - 1) / gpu);
    p.bands_per_unit = (groups_y + p.units_per_plane - 1) / p.units_per_plane;  // balance the last unit
    p.units_per_plane = (groups_y + p.bands_per_unit - 1) / p.bands_per_unit;
    // Tail tier. Units are handed out in address order and a unit with many objects takes 2-3 x the mean, so the
    // launch used to end with a long tail of half-idle SMs (profiles/scripts/unit_trace.py: 12 of 39 us on cfg3).
    // The planes that are handed out last (about one wave of full-size units) are cut into half-size units. Measured
    // gain is small (cfg3 43.7 -> 42.3 us, 8-GPU slice of cfg5 221.0 -> 220.1 us): an EMPTY unit already takes 3-4 us
    // of pure latency (sizes -> slots -> barriers -> next-unit handshake), so finer units buy little; quarter or
    // eighth units and longer tiers were slower (profiles/r01/experiments.md). Since the next unit is claimed when the
    // current one is done (dh_groups_kernel.cuh), quarter-size units are the better tier where they still hold a group
    // per warp (96 / 192 / 384 samples of cfg5: 58.9 / 107.1 / 205.4 us against 59.1 / 107.7 / 205.3 with halves), and
    // cfg3, whose units are too small to be cut in four, is best without a tier (34.3 against 34.7 us):
    // profiles/r02/retune_tail_tier.txt.
    p.tier1_planes = p.num_planes;
    p.bands_per_unit2 = p.bands_per_unit;
    p.units_per_plane2 = p.units_per_plane;
    const int small = g_tail_div > 1 ? p.bands_per_unit / g_tail_div : 0;
    if (small >= kGWarps && p.num_planes * p.units_per_plane > max_ctas) {
        long long planes2 = (g_tail_waves_x2 * max_ctas / 2 + p.units_per_plane - 1) / p.units_per_plane;
        if (planes2 > p.num_planes / g_tail_cap_div) planes2 = p.num_planes / g_tail_cap_div;
        if (planes2 > 0) {
            p.tier1_planes = p.num_planes - planes2;
            p.units_per_plane2 = (groups_y + small - 1) / small;
            p.bands_per_unit2 = (groups_y + p.units_per_plane2 - 1) / p.units_per_plane2;
            p.units_per_plane2 = (groups_y + p.bands_per_unit2 - 1) / p.bands_per_unit2;
        }
    }
    p.tier1_units = p.tier1_planes * p.units_per_plane;
    p.total_units = p.tier1_units + (p.num_planes - p.tier1_planes) * p.units_per_plane2;
    p.stage_bytes = (int)(((long long)G * p.W * 4 + 127) & ~127ll);
    return true;
}

// One launch with the programmatic-dependent-launch attribute: the prologue of the launched kernel (shared-memory
// set-up) may overlap the tail of whatever precedes it in the stream; our kernels wait (griddepcontrol.wait, dh_ptx.cuh
// pdl_wait) before they touch global memory.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

template <bool ACCUM, int SRC, bool DIGEST, int OUT = OUT_F32, bool LEAN = false, bool DENSE = false>
static int launch_groups(cudaStream_t stream, RenderParams& p, int sms) {
    static std::atomic<bool> attr_set_dev[64];  // function attributes are per device: one process may drive several GPUs
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DH_ERR_NO_DEVICE;
    if (!attr_set_dev[dev].load(std::memory_order_acquire)) {
        const int cap = (int)groups_smem_bytes(kGStageCap, ACCUM, LEAN);  // widest map: fewer CTAs per SM then
        cudaError_t e = cudaFuncSetAttribute(groups_kernel<ACCUM, SRC, DIGEST, OUT, LEAN, DENSE>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
        if (e != cudaSuccess) return (int)e;
        attr_set_dev[dev].store(true, std::memory_order_release);
    }
    const long long max_ctas = (long long)sms * ((ACCUM || LEAN) ? DH_G_MINB_ACCUM : DH_G_MINB);
    if (!make_groups_plan(p, max_ctas, OUT == OUT_F32 ? 4 : 2, ACCUM, LEAN, DENSE)) return DH_ERR_UNSUPPORTED_SHAPE;
    p.dense_min = g_dense_min;
    p.dense_cost16 = ACCUM ? g_dense_cost16_accum : g_dense_cost16;
    if (p.total_units == 0) return DH_OK;
    // Flat grid, one CTA per unit: the block scheduler hands units out dynamically and in address order, which
    // balances SMs that see different effective store bandwidth; a static grid-stride loop runs at the pace of the
    // slowest SM (6.1 vs 7.0-7.4 TB/s for the same stores, profiles/micro/fill_bw3.cu). Resident CTAs then take
    // over the not-yet-launched ones through cluster launch control, so only ~sms*DH_G_MINB CTAs ever start.
    long long g = p.total_units;
    p.use_clc = !g_persistent && !g_no_clc && g <= 0x7fffffffll;  // a clamped grid falls back to the static stride
    if (p.use_clc) p.use_clc = g_clc_mode;
    if (g_persistent && g > max_ctas) g = max_ctas;
    if (g > 0x7fffffffll) g = 0x7fffffffll;
    // heaviest-first hand-out (dh_groups_kernel.cuh build_perm): only when the launch has more than one wave — a single
    // wave has no tail to shorten — and the source has per-sample object counts
    p.perm_n = 0;
    p.perm_base = 0;
    if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES || SRC == SRC_WEIGHTED) {
        const long long B = p.num_planes / (p.C > 0 ? p.C : 1);
        if (g_lpt && p.total_units > max_ctas && B >= 2 && p.T >= 1 &&
            (long long)kGWarps * ((ACCUM || LEAN) ? 1 : 2) * p.stage_bytes >= kGPermScratch) {
            // only the samples handed out in the last g_lpt / 2 waves are rearranged: rearranging all of them was measured
            // SLOWER on large batches (cfg5 1.576 -> 1.635 ms): it scatters the set of addresses being written over the whole
            // buffer, and the write rate depends on that set staying compact (profiles/r02/experiments.md)
            long long win = (g_lpt * max_ctas / 2 + (long long)p.C * p.units_per_plane - 1) / ((long long)p.C * p.units_per_plane);
            if (win > kGPermCap) win = kGPermCap;
            if (win > B) win = B;
            p.perm_n = (int)win;
            p.perm_base = B - p.perm_n;
        }
    }
    const cudaError_t e = launch_pdl(groups_kernel<ACCUM, SRC, DIGEST, OUT, LEAN, DENSE>, dim3((unsigned)g), dim3(kGThreads),
                                     groups_smem_bytes(p.stage_bytes, ACCUM, LEAN), stream, p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

// K2 pre-pass + warp-autonomous renderer (dh_groups2_kernel.cuh): two launches, the second one programmatically
// dependent on the first (its prologue overlaps bin_kernel; it waits before it reads the lists).
template <bool ACCUM, bool DIGEST, int OUT = OUT_F32>
static int launch_groups2(cudaStream_t stream, RenderParams& p, int sms, long long B, void* workspace) {
    static std::atomic<bool> attr_set_dev[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DH_ERR_NO_DEVICE;
    if (!attr_set_dev[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(groups2_kernel<ACCUM, DIGEST, OUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)groups_smem_bytes(kGStageCap));
        if (e != cudaSuccess) return (int)e;
        attr_set_dev[dev].store(true, std::memory_order_release);
    }
    const long long max_ctas = (long long)sms * DH_G_MINB;
    if (!make_groups_plan(p, max_ctas, OUT == OUT_F32 ? 4 : 2)) return DH_ERR_UNSUPPORTED_SHAPE;
    if (p.total_units == 0) return DH_OK;
    int2* const rec = reinterpret_cast<int2*>(workspace);
    int4* const list = reinterpret_cast<int4*>(reinterpret_cast<char*>(workspace) + bin_rec_bytes(p.num_planes));
    p.bin_rec = rec;
    p.bin_list = list;
    cudaError_t e;
    if (p.centers_f != nullptr)
        e = launch_pdl(bin_kernel<SRC_BOXES>, dim3((unsigned)B), dim3(kBinThreads), 0, stream, p, list, rec);
    else
        e = launch_pdl(bin_kernel<SRC_BATCHED>, dim3((unsigned)B), dim3(kBinThreads), 0, stream, p, list, rec);
    if (e != cudaSuccess) return (int)e;
    long long g = p.total_units;
    p.use_clc = !g_persistent && !g_no_clc && g <= 0x7fffffffll;
    if (g_persistent && g > max_ctas) g = max_ctas;
    if (g > 0x7fffffffll) g = 0x7fffffffll;
    e = launch_pdl(groups2_kernel<ACCUM, DIGEST, OUT>, dim3((unsigned)g), dim3(kGThreads), groups_smem_bytes(p.stage_bytes),
                   stream, p);
    g_launches.fetch_add(2, std::memory_order_relaxed);
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

// Does the binned two-launch path serve this call? (main-path shape, class counters fit bin_kernel's shared memory,
// list offsets fit int32, workspace present, aligned and large enough)
static bool binned_ok(const RenderParams& p, long long B, const void* workspace, size_t workspace_bytes) {
    if (workspace == nullptr || (reinterpret_cast<uintptr_t>(workspace) & 15u) || g_no_bin) return false;
    if (!((p.W % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.heat) & 15u) == 0) && p.W <= kGMaxW)) return false;
    if (p.C > kBinMaxC || B * (long long)p.T > 0x7fffffffll || B > 0x7fffffffll || p.T < 1) return false;
    return workspace_bytes >= bin_workspace_bytes(B, p.T, p.C);
}

// slot arrays aligned for one vector load per slot? (dh_rows_kernel.cuh ld_center_i / ld_center_f / ld_box_f)
static int inputs_vec_ok(const RenderParams& p) {
    return ((reinterpret_cast<uintptr_t>(p.centers) & 7u) == 0) && ((reinterpret_cast<uintptr_t>(p.centers_f) & 7u) == 0) &&
           ((reinterpret_cast<uintptr_t>(p.bboxes_f) & 15u) == 0);
}

template <int SRC>
static int launch_render(cudaStream_t stream, RenderParams& p, int mode, long long B = 0, void* workspace = nullptr,
                         size_t workspace_bytes = 0) {
    DeviceInfo di;
    int st = device_info(di);
    if (st != DH_OK) return st;
    p.in_vec = inputs_vec_ok(p);
    const bool vec = (p.W % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.heat) & 15u) == 0);
    const bool accum = (mode == DH_MODE_ACCUMULATE);
    if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
        if (binned_ok(p, B, workspace, workspace_bytes)) {  // K2 pre-pass + warp-autonomous renderer
            if (accum) return launch_groups2<true, false>(stream, p, di.sms, B, workspace);
            if (p.digest != nullptr) return launch_groups2<false, true>(stream, p, di.sms, B, workspace);
            return launch_groups2<false, false>(stream, p, di.sms, B, workspace);
        }
    }
    if (vec && p.W <= kGMaxW) {  // main path: row groups striped over warps, per-warp staging + bulk stores
        if constexpr (SRC == SRC_CSR) {
            // Flat entry: the host knows the average number of objects per plane. From 16 on, the instantiation that also
            // contains the register-tile renderer (each unit still decides for itself, dh_tiles_kernel.cuh).
            if (g_tiles && !(p.k < 0.0f) && p.num_objects >= 16 * p.num_planes) {
                if (accum) return launch_groups<true, SRC, false, OUT_F32, false, true>(stream, p, di.sms);
                return launch_groups<false, SRC, false, OUT_F32, false, true>(stream, p, di.sms);
            }
        }
        if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
            // Slot arrays wide enough to hold an object-heavy sample (hundreds of objects per plane: instruction-issue
            // bound, not store bound): the instantiation that also contains the register-tile renderer. T is only the
            // padded slot count — which renderer a unit gets is decided on the device, per unit, from the objects that
            // land on it (dh_tiles_kernel.cuh). k < 0 stays with the staged renderer (its order-preserving encode covers
            // that sign).
            if (g_tiles && p.T >= g_tiles_t && !(p.k < 0.0f)) {
                if (accum) return launch_groups<true, SRC, false, OUT_F32, false, true>(stream, p, di.sms);
                if (p.digest != nullptr) return launch_groups<false, SRC, true, OUT_F32, false, true>(stream, p, di.sms);
                return launch_groups<false, SRC, false, OUT_F32, false, true>(stream, p, di.sms);
            }
        }
        if (accum) return launch_groups<true, SRC, false>(stream, p, di.sms);
        if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES) {
            // object-heavy FRESH renders (hundreds of objects per plane: instruction-issue bound, not store bound): the
            // lean layout — one staging buffer per warp, twice the rows per group, four CTAs per SM
            if (g_lean && p.T >= g_lean_t) {
                if (p.digest != nullptr) return launch_groups<false, SRC, true, OUT_F32, true>(stream, p, di.sms);
                return launch_groups<false, SRC, false, OUT_F32, true>(stream, p, di.sms);
            }
        }
        if (p.digest != nullptr) return launch_groups<false, SRC, true>(stream, p, di.sms);
        return launch_groups<false, SRC, false>(stream, p, di.sms);
    }
    if (p.digest != nullptr) return DH_ERR_UNSUPPORTED_SHAPE;  // fused digest exists on the main path only
    if (vec && p.W <= kTilePx) {  // wide maps: block-cooperative bands + bulk stores
        return accum ? launch_rows<true, SRC>(stream, p, di.sms) : launch_rows<false, SRC>(stream, p, di.sms);
    }
    // generic path (W not a multiple of 4, misaligned base, or rows wider than a tile): pixel-owner threads
    static std::atomic<int> occ[64][2][2];  // per device (occupancy is a per-device property)
    int dev = 0, o = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DH_ERR_NO_DEVICE;
    std::atomic<int>& oc = occ[dev][vec][accum];
    if (vec && accum) st = blocks_per_sm(render_kernel<4, true, SRC>, oc, o);
    else if (vec) st = blocks_per_sm(render_kernel<4, false, SRC>, oc, o);
    else if (accum) st = blocks_per_sm(render_kernel<1, true, SRC>, oc, o);
    else st = blocks_per_sm(render_kernel<1, false, SRC>, oc, o);
    if (st != DH_OK) return st;
    const long long max_ctas = (long long)di.sms * o;
    make_plan(p, vec ? 4 : 1, max_ctas);
    if (p.total_units == 0) return DH_OK;
    const int grid = (int)(p.total_units < max_ctas ? p.total_units : max_ctas);
    if (vec && accum) render_kernel<4, true, SRC><<<grid, kThreads, 0, stream>>>(p);
    else if (vec) render_kernel<4, false, SRC><<<grid, kThreads, 0, stream>>>(p);
    else if (accum) render_kernel<1, true, SRC><<<grid, kThreads, 0, stream>>>(p);
    else render_kernel<1, false, SRC><<<grid, kThreads, 0, stream>>>(p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return (int)cudaGetLastError();
}

// FRESH render into fp16 / bf16 planes: main renderer only.
template <int SRC>
static int launch_render16(cudaStream_t stream, RenderParams& p, int out_dtype, long long B, void* workspace,
                           size_t workspace_bytes) {
    DeviceInfo di;
    int st = device_info(di);
    if (st != DH_OK) return st;
    if (!((p.W % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.heat) & 15u) == 0) && p.W <= kGMaxW))
        return DH_ERR_UNSUPPORTED_SHAPE;
    p.in_vec = inputs_vec_ok(p);
    if (binned_ok(p, B, workspace, workspace_bytes)) {
        if (out_dtype == DH_OUT_F16) return launch_groups2<false, false, OUT_F16>(stream, p, di.sms, B, workspace);
        return launch_groups2<false, false, OUT_BF16>(stream, p, di.sms, B, workspace);
    }
    if (g_tiles && p.T >= g_tiles_t && !(p.k < 0.0f)) {  // as for fp32 planes: the instantiation with the per-unit tile renderer
        if (out_dtype == DH_OUT_F16) return launch_groups<false, SRC, false, OUT_F16, false, true>(stream, p, di.sms);
        return launch_groups<false, SRC, false, OUT_BF16, false, true>(stream, p, di.sms);
    }
    if (out_dtype == DH_OUT_F16) return launch_groups<false, SRC, false, OUT_F16>(stream, p, di.sms);
    return launch_groups<false, SRC, false, OUT_BF16>(stream, p, di.sms);
}

// FRESH render of symmetric integer windows with per-object weights and per-class peaks through the main renderer
// (used by dh_gaussians.cu for the DALI draw_gaussians semantics). DH_ERR_UNSUPPORTED_SHAPE when the map does not
// qualify for groups_kernel (W % 4, alignment, W > kGMaxW): the caller then uses its own kernel.
int launch_weighted_fresh(cudaStream_t stream, float* heat, const int32_t* centers, const int32_t* radii_eff,
                          const float* weights, const int32_t* labels, const void* sizes, int sizes_i64,
                          const float* k_classes, long long B, long long T, long long C, long long H, long long W) {
    DeviceInfo di;
    int st = device_info(di);
    if (st != DH_OK) return st;
    if (!((W % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0) && W <= kGMaxW)) return DH_ERR_UNSUPPORTED_SHAPE;
    RenderParams p{};
    p.heat = heat;
    p.centers = centers;
    p.radii = radii_eff;
    p.weights = weights;
    p.k_classes = k_classes;
    p.labels = labels;
    p.sizes = sizes;
    p.sizes_i64 = sizes_i64;
    p.num_planes = B * C;
    p.C = (int)C;
    p.T = (int)T;
    p.H = (int)H;
    p.W = (int)W;
    p.factor = 1.0f;
    p.k = 1.0f;
    p.in_vec = inputs_vec_ok(p);
    return launch_groups<false, SRC_WEIGHTED, false>(stream, p, di.sms);
}

}  // namespace dh

// ================================================================================================
extern "C" {

int dh_version(void) { return DH_B200_VERSION; }

const char* dh_build_info(void) {
    return "dh_b200 0.1.0 | nvcc " DH_STR(__CUDACC_VER_MAJOR__) "." DH_STR(__CUDACC_VER_MINOR__)
           " | -gencode arch=compute_100a,code=sm_100a | fp32 IEEE (no fast-math) | no CPU fallback";
}

const char* dh_status_string(int status) {
    switch (status) {
        case DH_OK: return "ok";
        case DH_ERR_INVALID_ARGUMENT: return "invalid argument";
        case DH_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
        case DH_ERR_UNSUPPORTED_SHAPE: return "unsupported shape (need 1 <= H,W <= 2^23 and H*W < 2^31)";
        case DH_ERR_NO_DEVICE: return "no usable CUDA device";
        case DH_ERR_STATIC_SCRATCH_EXCEEDED:
            return "problem needs more scratch than the built-in static array holds (DH_NOALLOC_MAX_SCRATCH_BYTES): use the "
                   "*_ws entry with a caller workspace";
        default: break;
    }
    if (status > 0) return cudaGetErrorString((cudaError_t)status);
    return "unknown status";
}

uint64_t dh_launch_count(void) { return dh::g_launches.load(std::memory_order_relaxed); }

int dh_radius_center(void* stream, const float* centers_xy, const float* bboxes_xyxy, float out_size_factor, int64_t n,
                     int32_t* centers_out, int32_t* radii_out) {
    if (n < 0) return DH_ERR_INVALID_ARGUMENT;
    if (n == 0) return DH_OK;
    if (!centers_xy || !bboxes_xyxy || !centers_out || !radii_out) return DH_ERR_INVALID_ARGUMENT;
    dh::DeviceInfo di;
    int st = dh::device_info(di);
    if (st != DH_OK) return st;
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)di.sms * 8;
    if (blocks > cap) blocks = cap;
    dh::radius_center_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(centers_xy, bboxes_xyxy, out_size_factor, n,
                                                                           centers_out, radii_out);
    dh::g_launches.fetch_add(1, std::memory_order_relaxed);
    return (int)cudaGetLastError();
}

static int batched_common(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                          const float* centers_f, const float* bboxes_f, float osf, const void* nums_targets,
                          int nums_dtype, const int32_t* labels, int64_t B, int64_t T, int64_t C, int64_t H, int64_t W,
                          float factor, float k, int mode, dh_status* status, uint64_t* digest = nullptr,
                          int64_t digest_base = 0, void* workspace = nullptr, size_t workspace_bytes = 0) {
    if (B < 0 || T < 0) return DH_ERR_INVALID_ARGUMENT;
    if (mode != DH_MODE_ACCUMULATE && mode != DH_MODE_FRESH) return DH_ERR_INVALID_ARGUMENT;
    if (nums_dtype != DH_SIZES_I32 && nums_dtype != DH_SIZES_I64) return DH_ERR_INVALID_ARGUMENT;
    if (labels == nullptr) C = 1;
    if (C < 0 || C > INT32_MAX) return DH_ERR_INVALID_ARGUMENT;
    if (C == 0) return DH_OK;  // [B,0,H,W]: nothing to draw into
    int st = dh::check_shape(H, W);
    if (st != DH_OK) return st;
    if (B == 0) return DH_OK;
    if (T > INT32_MAX) return DH_ERR_UNSUPPORTED_SHAPE;
    if (!heatmap || !nums_targets) return DH_ERR_INVALID_ARGUMENT;
    if (T > 0 && ((centers_f == nullptr) ? (!centers || !radii) : (!bboxes_f))) return DH_ERR_INVALID_ARGUMENT;
    dh::RenderParams p{};
    p.heat = heatmap;
    p.centers = centers;
    p.radii = radii;
    p.centers_f = centers_f;
    p.bboxes_f = bboxes_f;
    p.osf = osf;
    p.labels = labels;
    p.sizes = nums_targets;
    p.sizes_i64 = (nums_dtype == DH_SIZES_I64);
    p.num_planes = B * C;
    p.C = (int)C;
    p.T = (int)T;
    p.H = (int)H;
    p.W = (int)W;
    p.factor = factor;
    p.k = k;
    p.status = status;
    p.digest = reinterpret_cast<unsigned long long*>(digest);
    p.digest_base = digest_base;
    if (centers_f) return dh::launch_render<dh::SRC_BOXES>((cudaStream_t)stream, p, mode, B, workspace, workspace_bytes);
    return dh::launch_render<dh::SRC_BATCHED>((cudaStream_t)stream, p, mode, B, workspace, workspace_bytes);
}

size_t dh_batched_workspace_bytes(int64_t batch_size, int64_t max_num_targets, int64_t num_classes) {
    if (batch_size < 0 || max_num_targets < 0 || num_classes < 0) return 0;
    if (num_classes < 1) num_classes = 1;
    return dh::bin_workspace_bytes(batch_size, max_num_targets, num_classes) + 16;
}

int dh_draw_heatmap_batched_ws(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                               const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t batch_size,
                               int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                               float diameter_to_sigma_factor, float k_scale, int mode, dh_status* status,
                               void* workspace, size_t workspace_bytes) {
    return batched_common(stream, heatmap, centers, radii, nullptr, nullptr, 1.0f, nums_targets, nums_dtype, labels,
                          batch_size, max_num_targets, num_classes, height, width, diameter_to_sigma_factor, k_scale,
                          mode, status, nullptr, 0, workspace, workspace_bytes);
}

int dh_draw_heatmap_batched(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                            const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t batch_size,
                            int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                            float diameter_to_sigma_factor, float k_scale, int mode, dh_status* status) {
    return dh_draw_heatmap_batched_ws(stream, heatmap, centers, radii, nums_targets, nums_dtype, labels, batch_size,
                                      max_num_targets, num_classes, height, width, diameter_to_sigma_factor, k_scale, mode,
                                      status, nullptr, 0);
}

// ---- allocation-free entry: the bin-list scratch of the two-launch path lives in the library image ---------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Every GPU that loads the library holds
// DH_NOALLOC_MAX_SCRATCH_BYTES of device memory for it until exit. ONE array per device, shared by every call: calls
// that overlap on the GPU (two streams, two threads) race on it. bin_kernel rewrites every (start, count) record and
// every list entry a record can reach before groups2_kernel reads them, so nothing has to start at zero and the last
// call's contents are harmless.
static __device__ __align__(256) unsigned char g_noalloc_scratch[DH_NOALLOC_MAX_SCRATCH_BYTES];

// `g_noalloc_scratch` in HOST code is a host-side stand-in, not a device address: the device address is looked up once
// per device, with that device current, and cached.
static int noalloc_scratch_ptr(void** out) {
    static std::atomic<void*> cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DH_ERR_NO_DEVICE;
    void* ptr = cache[dev].load(std::memory_order_acquire);
    if (ptr == nullptr) {
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess) return (int)e;
        if (ptr == nullptr) return DH_ERR_NO_DEVICE;
        cache[dev].store(ptr, std::memory_order_release);  // racing first calls store the same value
    }
    *out = ptr;
    return DH_OK;
}

size_t dh_noalloc_max_scratch_bytes(void) { return (size_t)DH_NOALLOC_MAX_SCRATCH_BYTES; }

int dh_draw_heatmap_batched_noalloc(void* stream, float* heatmap, const int32_t* centers, const int32_t* radii,
                                    const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t batch_size,
                                    int64_t max_num_targets, int64_t num_classes, int64_t height, int64_t width,
                                    float diameter_to_sigma_factor, float k_scale, int mode, dh_status* status) {
    if (batch_size < 0 || max_num_targets < 0 || num_classes < 0) return DH_ERR_INVALID_ARGUMENT;
    // size check first, without touching a device: beyond the built-in array is an error, never a silent other path
    const int64_t cap = (int64_t)DH_NOALLOC_MAX_SCRATCH_BYTES;
    const int64_t C1 = (labels && num_classes > 1) ? num_classes : 1;
    if (batch_size > 0 && (C1 > cap || max_num_targets > cap || 8 * C1 + 16 * max_num_targets > cap / batch_size))
        return DH_ERR_STATIC_SCRATCH_EXCEEDED;  // (overflow-free bound on B * (8 C + 16 T))
    if (dh_batched_workspace_bytes(batch_size, max_num_targets, C1) > (size_t)cap) return DH_ERR_STATIC_SCRATCH_EXCEEDED;
    void* scratch = nullptr;
    const int st = noalloc_scratch_ptr(&scratch);
    if (st != DH_OK) return st;
    return batched_common(stream, heatmap, centers, radii, nullptr, nullptr, 1.0f, nums_targets, nums_dtype, labels,
                          batch_size, max_num_targets, num_classes, height, width, diameter_to_sigma_factor, k_scale,
                          mode, status, nullptr, 0, scratch, (size_t)DH_NOALLOC_MAX_SCRATCH_BYTES);
}

int dh_draw_heatmap_batched_from_boxes(void* stream, float* heatmap, const float* centers_xy, const float* bboxes_xyxy,
                                       float out_size_factor, const void* nums_targets, int nums_dtype,
                                       const int32_t* labels, int64_t batch_size, int64_t max_num_targets,
                                       int64_t num_classes, int64_t height, int64_t width,
                                       float diameter_to_sigma_factor, float k_scale, int mode, int32_t* centers_out,
                                       int32_t* radii_out, dh_status* status) {
    return dh_draw_heatmap_batched_from_boxes_ws(stream, heatmap, centers_xy, bboxes_xyxy, out_size_factor, nums_targets,
                                                 nums_dtype, labels, batch_size, max_num_targets, num_classes, height, width,
                                                 diameter_to_sigma_factor, k_scale, mode, centers_out, radii_out, status,
                                                 nullptr, 0);
}

int dh_draw_heatmap_batched_from_boxes_ws(void* stream, float* heatmap, const float* centers_xy, const float* bboxes_xyxy,
                                          float out_size_factor, const void* nums_targets, int nums_dtype,
                                          const int32_t* labels, int64_t batch_size, int64_t max_num_targets,
                                          int64_t num_classes, int64_t height, int64_t width,
                                          float diameter_to_sigma_factor, float k_scale, int mode, int32_t* centers_out,
                                          int32_t* radii_out, dh_status* status, void* workspace,
                                          size_t workspace_bytes) {
    if ((centers_out == nullptr) != (radii_out == nullptr)) return DH_ERR_INVALID_ARGUMENT;
    if (batch_size > 0 && max_num_targets > 0 && (!centers_xy || !bboxes_xyxy)) return DH_ERR_INVALID_ARGUMENT;
    if (centers_out != nullptr) {  // integer outputs requested: K1 then the integer-input renderer
        int st = dh_radius_center(stream, centers_xy, bboxes_xyxy, out_size_factor, batch_size * max_num_targets,
                                  centers_out, radii_out);
        if (st != DH_OK) return st;
        return batched_common(stream, heatmap, centers_out, radii_out, nullptr, nullptr, 1.0f, nums_targets, nums_dtype,
                              labels, batch_size, max_num_targets, num_classes, height, width,
                              diameter_to_sigma_factor, k_scale, mode, status, nullptr, 0, workspace, workspace_bytes);
    }
    return batched_common(stream, heatmap, nullptr, nullptr, centers_xy, bboxes_xyxy, out_size_factor, nums_targets,
                          nums_dtype, labels, batch_size, max_num_targets, num_classes, height, width,
                          diameter_to_sigma_factor, k_scale, mode, status, nullptr, 0, workspace, workspace_bytes);
}

int dh_draw_heatmap_batched_from_boxes_digest(void* stream, float* heatmap, const float* centers_xy,
                                              const float* bboxes_xyxy, float out_size_factor,
                                              const void* nums_targets, int nums_dtype, const int32_t* labels,
                                              int64_t batch_size, int64_t max_num_targets, int64_t num_classes,
                                              int64_t height, int64_t width, float diameter_to_sigma_factor,
                                              float k_scale, int64_t index_base, uint64_t* digest_out3,
                                              dh_status* status) {
    return dh_draw_heatmap_batched_from_boxes_digest_ws(stream, heatmap, centers_xy, bboxes_xyxy, out_size_factor,
                                                        nums_targets, nums_dtype, labels, batch_size, max_num_targets,
                                                        num_classes, height, width, diameter_to_sigma_factor, k_scale,
                                                        index_base, digest_out3, status, nullptr, 0);
}

int dh_draw_heatmap_batched_from_boxes_digest_ws(void* stream, float* heatmap, const float* centers_xy,
                                                 const float* bboxes_xyxy, float out_size_factor,
                                                 const void* nums_targets, int nums_dtype, const int32_t* labels,
                                                 int64_t batch_size, int64_t max_num_targets, int64_t num_classes,
                                                 int64_t height, int64_t width, float diameter_to_sigma_factor,
                                                 float k_scale, int64_t index_base, uint64_t* digest_out3,
                                                 dh_status* status, void* workspace, size_t workspace_bytes) {
    if (!digest_out3) return DH_ERR_INVALID_ARGUMENT;
    if (batch_size > 0 && max_num_targets > 0 && (!centers_xy || !bboxes_xyxy)) return DH_ERR_INVALID_ARGUMENT;
    return batched_common(stream, heatmap, nullptr, nullptr, centers_xy, bboxes_xyxy, out_size_factor, nums_targets,
                          nums_dtype, labels, batch_size, max_num_targets, num_classes, height, width,
                          diameter_to_sigma_factor, k_scale, DH_MODE_FRESH, status, digest_out3, index_base, workspace,
                          workspace_bytes);
}

int dh_render_heatmap_batched_16(void* stream, void* heatmap, int out_dtype, const int32_t* centers, const int32_t* radii,
                                 const float* centers_xy, const float* bboxes_xyxy, float out_size_factor,
                                 const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t B, int64_t T,
                                 int64_t C, int64_t H, int64_t W, float factor, float k, dh_status* status) {
    return dh_render_heatmap_batched_16_ws(stream, heatmap, out_dtype, centers, radii, centers_xy, bboxes_xyxy,
                                           out_size_factor, nums_targets, nums_dtype, labels, B, T, C, H, W, factor, k, status,
                                           nullptr, 0);
}

int dh_render_heatmap_batched_16_ws(void* stream, void* heatmap, int out_dtype, const int32_t* centers,
                                    const int32_t* radii, const float* centers_xy, const float* bboxes_xyxy,
                                    float out_size_factor, const void* nums_targets, int nums_dtype, const int32_t* labels,
                                    int64_t B, int64_t T, int64_t C, int64_t H, int64_t W, float factor, float k,
                                    dh_status* status, void* workspace, size_t workspace_bytes) {
    if (out_dtype != DH_OUT_F16 && out_dtype != DH_OUT_BF16) return DH_ERR_INVALID_ARGUMENT;
    if (B < 0 || T < 0) return DH_ERR_INVALID_ARGUMENT;
    if (nums_dtype != DH_SIZES_I32 && nums_dtype != DH_SIZES_I64) return DH_ERR_INVALID_ARGUMENT;
    const bool boxes = (centers_xy != nullptr) || (bboxes_xyxy != nullptr);
    if (boxes && (centers != nullptr || radii != nullptr)) return DH_ERR_INVALID_ARGUMENT;  // one source, not both
    if (labels == nullptr) C = 1;
    if (C < 0 || C > INT32_MAX) return DH_ERR_INVALID_ARGUMENT;
    if (C == 0) return DH_OK;
    int st = dh::check_shape(H, W);
    if (st != DH_OK) return st;
    if (B == 0) return DH_OK;
    if (T > INT32_MAX) return DH_ERR_UNSUPPORTED_SHAPE;
    if (!heatmap || !nums_targets) return DH_ERR_INVALID_ARGUMENT;
    if (T > 0 && (boxes ? (!centers_xy || !bboxes_xyxy) : (!centers || !radii))) return DH_ERR_INVALID_ARGUMENT;
    dh::RenderParams p{};
    p.heat = reinterpret_cast<float*>(heatmap);
    p.centers = centers;
    p.radii = radii;
    p.centers_f = centers_xy;
    p.bboxes_f = bboxes_xyxy;
    p.osf = out_size_factor;
    p.labels = labels;
    p.sizes = nums_targets;
    p.sizes_i64 = (nums_dtype == DH_SIZES_I64);
    p.num_planes = B * C;
    p.C = (int)C;
    p.T = (int)T;
    p.H = (int)H;
    p.W = (int)W;
    p.factor = factor;
    p.k = k;
    p.status = status;
    if (boxes) return dh::launch_render16<dh::SRC_BOXES>((cudaStream_t)stream, p, out_dtype, B, workspace, workspace_bytes);
    return dh::launch_render16<dh::SRC_BATCHED>((cudaStream_t)stream, p, out_dtype, B, workspace, workspace_bytes);
}

// Flat path: direct scan when (units x objects) is small, else counting sort by plane.
static const long long kFlatScanBudget = 1ll << 22;

static size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

size_t dh_draw_heatmap_workspace_bytes(int64_t num_targets, int64_t num_heatmaps) {
    if (num_targets < 0 || num_heatmaps < 0) return 0;
    // counts[M] | cursor[M] | off[M+1] | items[N]
    return align16((size_t)num_heatmaps * 4 * 2) + align16(((size_t)num_heatmaps + 1) * 4) +
           align16((size_t)num_targets * 4) + 16;
}

int dh_draw_heatmap(void* stream_, float* heatmaps, const int32_t* centers, const int32_t* radii,
                    const int32_t* heatmap_idxes, int64_t N, int64_t M, int64_t H, int64_t W, float factor, float k,
                    int mode, void* workspace, size_t workspace_bytes, dh_status* status) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N < 0 || M < 0) return DH_ERR_INVALID_ARGUMENT;
    if (mode != DH_MODE_ACCUMULATE && mode != DH_MODE_FRESH) return DH_ERR_INVALID_ARGUMENT;
    int st = dh::check_shape(H, W);
    if (st != DH_OK) return st;
    if (M == 0) return DH_OK;
    if (N > INT32_MAX || M > INT32_MAX) return DH_ERR_UNSUPPORTED_SHAPE;
    if (!heatmaps) return DH_ERR_INVALID_ARGUMENT;
    if (N > 0 && (!centers || !radii || !heatmap_idxes)) return DH_ERR_INVALID_ARGUMENT;
    if (N == 0 && mode == DH_MODE_ACCUMULATE) return DH_OK;

    dh::RenderParams p{};
    p.heat = heatmaps;
    p.centers = centers;
    p.radii = radii;
    p.plane_idx = heatmap_idxes;
    p.num_objects = N;
    p.num_planes = M;
    p.C = 1;
    p.T = 0;
    p.H = (int)H;
    p.W = (int)W;
    p.factor = factor;
    p.k = k;
    p.status = status;

    // Rough unit count (upper bound: one band per unit) to choose the binning strategy.
    const long long approx_units = M * (((long long)H * W / 4 + dh::kBandCap - 1) / dh::kBandCap);
    if (approx_units * N <= kFlatScanBudget) return dh::launch_render<dh::SRC_FLATSCAN>(stream, p, mode);

    if (workspace == nullptr || workspace_bytes < dh_draw_heatmap_workspace_bytes(N, M)) return DH_ERR_WORKSPACE_TOO_SMALL;
    if (reinterpret_cast<uintptr_t>(workspace) & 15u) return DH_ERR_INVALID_ARGUMENT;
    char* ws = (char*)workspace;
    int32_t* counts = (int32_t*)ws;
    int32_t* cursor = counts + M;
    ws += align16((size_t)M * 8);
    int32_t* off = (int32_t*)ws;
    ws += align16(((size_t)M + 1) * 4);
    int32_t* items = (int32_t*)ws;
    cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)M * 8, stream);
    if (e != cudaSuccess) return (int)e;
    const int nb = (int)((N + 255) / 256);
    dh::csr_count_kernel<<<nb, 256, 0, stream>>>(heatmap_idxes, N, M, counts, status);
    dh::csr_scan_kernel<<<1, 1024, 0, stream>>>(counts, M, off);
    dh::csr_fill_kernel<<<nb, 256, 0, stream>>>(heatmap_idxes, N, M, off, cursor, items);
    dh::g_launches.fetch_add(3, std::memory_order_relaxed);
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    p.csr_off = off;
    p.csr_items = items;
    return dh::launch_render<dh::SRC_CSR>(stream, p, mode);
}

#ifdef DH_TRACE_UNITS
// Profiling builds only (not declared in include/dh_b200.h, not in the shipped library).
int dh_debug_set_trace(void* buf, int64_t units) {
    unsigned long long* b = (unsigned long long*)buf;
    long long n = units;
    cudaError_t e = cudaMemcpyToSymbol(dh::g_trace, &b, sizeof(b));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(dh::g_trace_units, &n, sizeof(n));
    return (int)e;
}
#endif

int dh_fill_probe(void* stream, void* dst, size_t bytes) {
    if (bytes == 0) return DH_OK;
    if (!dst || (bytes & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u)) return DH_ERR_INVALID_ARGUMENT;
    dh::DeviceInfo di;
    int st = dh::device_info(di);
    if (st != DH_OK) return st;
    const long long n4 = (long long)(bytes / 16);
    const long long chunk4 = dh::kBandCap;
    const long long nchunks = (n4 + chunk4 - 1) / chunk4;
    const long long cap = (long long)di.sms * 8;
    dh::fill_probe_kernel<<<(int)(nchunks < cap ? nchunks : cap), dh::kThreads, 0, (cudaStream_t)stream>>>(
        (float4*)dst, n4, chunk4);
    dh::g_launches.fetch_add(1, std::memory_order_relaxed);
    return (int)cudaGetLastError();
}

int dh_expf_probe(void* stream, uint32_t start_bits, int64_t n, float* out) {
    if (n < 0 || (n > 0 && !out)) return DH_ERR_INVALID_ARGUMENT;
    if (n == 0) return DH_OK;
    dh::DeviceInfo di;
    int st = dh::device_info(di);
    if (st != DH_OK) return st;
    long long blocks = (n + 255) / 256;
    if (blocks > (long long)di.sms * 16) blocks = (long long)di.sms * 16;
    dh::expf_probe_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(start_bits, n, out);
    dh::g_launches.fetch_add(1, std::memory_order_relaxed);
    return (int)cudaGetLastError();
}

}  // extern "C"
