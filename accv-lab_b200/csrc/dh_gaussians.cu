// dh_gaussians.cu — N2: the DALI `draw_gaussians` operator semantics on the GPU.
//
// The reference implements this operator for the CPU only
//   /root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc:31-75  (draw_gaussian)
//   .../DrawGaussians.cc:88-157 (RunImpl: out = in; every ACTIVE object -> plane slice_id with k_for_classes[slice_id])
// one thread-pool task per sample, serial loop over objects and window pixels. Differences to the draw_heatmap
// kernel arithmetic: FLOAT radius, sigma = radius * factor, window [floor(c - r), ceil(c + r)]
// (inclusive, can be asymmetric after fp32 rounding), per-class peak k, an `active` flag per object, and
// max(pixel, val) = (pixel < val) ? val : pixel.
//
// Here: output-partitioned like the main renderer. One CTA per unit = a band of whole rows of one (sample, class)
// plane (<= 4096 pixels). Per unit: the sample's T slots are scanned once (active / class / size filter, window
// arithmetic exactly as the reference: floorf/ceilf of one fp32 add, clip, cull against the band), compacted with
// warp ballot + popc prefix into a shared-memory list; every thread then owns 4 consecutive pixels x up to 4 items
// in registers and keeps, per pixel, min_o t_o with t_o = (float)(dy^2+dx^2) * inv_o (max_o for k < 0): ONE expf per
// touched pixel (expf is monotone, tests/test_gpu_properties.py), no atomics; the band is written once with 128-bit
// stores (input copied / zero-init fused for untouched pixels; in-place calls write touched groups only).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "dh_b200.h"
#include "dh_ptx.cuh"

namespace dh {

extern void count_launches(uint64_t n);  // dh_kernels.cu

constexpr int kGsThreads = 256;
constexpr int kGsItems = 4;      // register-resident items (PX pixels each) per thread and pass
constexpr int kGsListCap = 512;  // objects per shared-memory list chunk
constexpr float kGsMaxRadius = 2048.0f;

struct __align__(16) GObj {
    int x0, x1, y0, y1;  // clipped window, inclusive
    float cx, cy, inv;   // centre (exact ints as floats), 1 / (2 sigma^2)
    int pad;
};

struct GsParams {
    const float* in;  // may be null (FRESH) or == out (in place)
    float* out;
    const int32_t* centers;
    const float* radii;
    const int32_t* slice_ids;  // may be null -> class 0
    const uint8_t* active;     // may be null -> all active
    const void* sizes;
    int sizes_i64;
    const float* k_for_classes;
    int num_k;
    long long num_planes, total_units;
    int C, T, H, W, band_rows, bands_per_plane, bands_per_unit, units_per_plane;
    float factor;
    dh_status* status;
    const int* gate;  // optional: the whole grid returns at once when *gate == 0 (fix-up pass with nothing to fix)
    // LOSS mode (N3): draw_heatmap semantics (integer radii, sigma = (2r+1)/factor, one k) + fused Gaussian focal loss
    const int32_t* radii_i;
    float dh_k, alpha, gamma, eps;
    int fast;  // LOSS: MUFU-based log / division (__logf, __fdividef) instead of the IEEE-accurate routines
    double* partial_loss;    // [total_units]
    long long* partial_pos;  // [total_units]
};

// LOSS mode: slot -> list entry with the draw_heatmap arithmetic (reference cuh:61-68,90-97):
// window [x-r, x+r] x [y-r, y+r] clipped, inv = 1/(2 sigma^2) with sigma = (float)(2r+1)/factor. `slice_ids` = labels.
__device__ __forceinline__ bool gs_make_dh(const GsParams& p, long long i, int cls, int uy0, int uy1, GObj& o) {
    if (p.slice_ids != nullptr && p.slice_ids[i] != cls) return false;
    const int r = p.radii_i[i];
    if (r < 0 || r > 2048) return false;
    const int2 cc = *reinterpret_cast<const int2*>(p.centers + 2 * i);
    const long long X = cc.x, Y = cc.y;
    if (Y + r < uy0 || Y - r > uy1 || X + r < 0 || X - r >= p.W || Y + r < 0 || Y - r >= p.H) return false;
    o.x0 = (int)max(X - r, 0ll);
    o.x1 = (int)min(X + r, (long long)p.W - 1);
    o.y0 = (int)max(Y - r, 0ll);
    o.y1 = (int)min(Y + r, (long long)p.H - 1);
    const float sigma = __fdiv_rn((float)(2 * r + 1), p.factor);
    o.inv = __fdiv_rn(1.0f, __fmul_rn(__fmul_rn(2.0f, sigma), sigma));
    o.cx = (float)cc.x;
    o.cy = (float)cc.y;
    o.pad = 0;
    return true;
}

// Slot -> list entry for plane class `cls`, culled against the unit's rows [uy0, uy1]. Window arithmetic of DrawGaussians.cc:34-51.
__device__ __forceinline__ bool gs_make(const GsParams& p, long long i, int cls, int uy0, int uy1, bool count, GObj& o) {
    if (p.active != nullptr && p.active[i] == 0) return false;
    const float r = p.radii[i];
    if (!(fabsf(r) <= kGsMaxRadius)) {  // NaN or huge; counted once per object (by the sample's first unit)
        if (count && p.status) atomicAdd(&p.status->skipped_radius, 1u);
        return false;
    }
    const int c = p.slice_ids != nullptr ? p.slice_ids[i] : 0;
    if (c != cls) {
        if (count && p.status && (c < 0 || c >= p.C || c >= p.num_k)) atomicAdd(&p.status->skipped_label, 1u);
        return false;
    }
    const int2 cc = *reinterpret_cast<const int2*>(p.centers + 2 * i);
    const float fx = (float)cc.x, fy = (float)cc.y;
    // |centre| beyond the map by more than the largest window: nothing to draw (also keeps float -> int casts defined)
    if (fabsf(fx) > 16777216.0f || fabsf(fy) > 16777216.0f) return false;
    const int min_x = (int)floorf(__fsub_rn(fx, r)), max_x = (int)ceilf(__fadd_rn(fx, r));
    const int min_y = (int)floorf(__fsub_rn(fy, r)), max_y = (int)ceilf(__fadd_rn(fy, r));
    if ((max_x < 0 || min_x >= p.W) || (max_y < 0 || min_y >= p.H)) return false;
    o.x0 = min(max(0, min_x), p.W - 1);
    o.x1 = min(max(0, max_x), p.W - 1);
    o.y0 = min(max(0, min_y), p.H - 1);
    o.y1 = min(max(0, max_y), p.H - 1);
    if (o.x0 > o.x1 || o.y0 > o.y1) return false;  // negative radius: empty loops in the reference
    if (o.y1 < uy0 || o.y0 > uy1) return false;
    const float sigma = __fmul_rn(r, p.factor);
    o.inv = __fdiv_rn(1.0f, __fmul_rn(__fmul_rn(2.0f, sigma), sigma));
    o.cx = fx;
    o.cy = fy;
    o.pad = 0;
    return true;
}

constexpr int kGsZeroBytes = 8192;  // constant zero tile of the FRESH mode

// log(a) for the focal loss: a = pred + eps or 1 - pred + eps, i.e. a positive normal float. a = m * 2^e with m in
// [2/3, 4/3), log(m) = f + f^2 P(f) with f = m - 1 and P a degree-7 polynomial fitted to (log1p(f) - f) / f^2 (minimax in
// the relative error of log1p), everything in FMAs: 17 instructions instead of the ~30 of logf (which was a third of the
// loss launch), at the same accuracy — max 1.05 ulp over all normal floats, 0.58 ulp around 1, exact 0 at 1
// (tests/test_focal_loss.py restates it in numpy and checks it against float64 log). Outside the normal range the
// result is patched by selects (no call: the loss body is unrolled 16 times): +0 / -0 and denormals (counted as 0) ->
// -inf, +inf -> +inf, negative or NaN -> NaN.
__device__ __forceinline__ float log_pos(float a) {
    const int ia = __float_as_int(a);
    const int e = (ia - 0x3f2aaaab) & (int)0xff800000;
    const float m = __int_as_float(ia - e);
    const float i = __fmul_rn((float)e, 1.1920928955078125e-07f);  // exponent as a float (e / 2^23)
    const float f = __fadd_rn(m, -1.0f);                           // exact
    const float s = __fmul_rn(f, f);
    float q = __int_as_float(0x3e0fa12d);
    q = fmaf(q, f, __int_as_float(0xbe1d9426));
    q = fmaf(q, f, __int_as_float(0x3e0f512f));
    q = fmaf(q, f, __int_as_float(0xbe281974));
    q = fmaf(q, f, __int_as_float(0x3e4cea14));
    q = fmaf(q, f, __int_as_float(0xbe800a40));
    q = fmaf(q, f, __int_as_float(0x3eaaaa7d));
    q = fmaf(q, f, __int_as_float(0xbeffffeb));
    const float r = fmaf(i, 0.693147182f, fmaf(q, s, f));
    if (a >= 1.17549435e-38f && a <= 3.40282347e38f) return r;
    return (a >= 0.0f) ? (a > 1.0f ? __int_as_float(0x7f800000) : __int_as_float((int)0xff800000)) : __int_as_float(0x7fc00000);
}

// FRESH (no input heatmap, PX == 4): runs of untouched bands leave as bulk stores (SASS UBLKCP) from a zero tile, issued
// by the lanes of warp 0 in parallel — no LSU or register traffic for the bytes nobody draws on.
// LOSS (N3): nothing is drawn to memory. p.in is the prediction, the target value of every pixel is formed in registers
// as in the FRESH renderer and consumed at once by the Gaussian focal loss (DH/docs/intro.rst:7-26); p.out (optional)
// receives d loss / d pred; per-unit partial sums make the reduction order independent of the block scheduler.
// LOSS_MODE: 0 = renderer; 1 = alpha 2 / gamma 4, 1-ulp log (log_pos); 2 = same through the special-function
// unit (__logf); 3 = general exponents (powf); 4, 5, 6 = the same three without d loss / d pred (forward only: the
// gradient's division and extra products were 16 % of the forward launch's instructions). Compile-time so that each
// variant's code stays small: a run-time switch tripled the unrolled loss body and the kernel stalled on instruction
// fetch (ncu: no_instruction 2.8). The gradient's quotient is __fdividef (2 ulp) in every mode: there is no reference
// value to match bit for bit, and the IEEE division cost 14 instructions per pixel.
template <int PX, bool INPLACE, bool FRESH, int LOSS_MODE = 0>
__global__ void __launch_bounds__(kGsThreads, (PX == 4 || LOSS_MODE) ? 3 : 4) gaussians_kernel(const GsParams p) {
    constexpr bool LOSS = LOSS_MODE != 0;
    if (p.gate != nullptr && *p.gate == 0) return;
    constexpr bool GRAD = LOSS_MODE >= 1 && LOSS_MODE <= 3;
    constexpr int LM = LOSS_MODE > 3 ? LOSS_MODE - 3 : LOSS_MODE;
    constexpr bool std_exp = LM == 1 || LM == 2;
    constexpr bool FAST = LM == 2;
    __shared__ GObj s_list[kGsListCap];
    __shared__ int s_cnt;
    __shared__ __align__(128) float4 s_zero[FRESH ? kGsZeroBytes / 16 : 1];
    if constexpr (FRESH) {
        for (int i = threadIdx.x; i < kGsZeroBytes / 16; i += kGsThreads) s_zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        fence_proxy_async();  // visible to the bulk-copy engine (the barriers of the first list build order it)
    }
    const int tid = threadIdx.x, lane = tid & 31;
    const int wunits = p.W / PX;
    const long long plane_elems = (long long)p.H * p.W;
    const float INF = __int_as_float(0x7f800000);

    __shared__ unsigned s_mask;  // bands of the unit some object overlaps
    int drow0[kGsItems], dcol0[kGsItems];  // (row offset, column) of this thread's items in a band
#pragma unroll
    for (int q = 0; q < kGsItems; ++q) {
        const int it = q * kGsThreads + tid;
        drow0[q] = it / wunits;
        dcol0[q] = (it - drow0[q] * wunits) * PX;
    }
    for (long long u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const long long plane = u / p.units_per_plane;
        const int ug = (int)(u - plane * p.units_per_plane);
        const int ub0 = ug * p.bands_per_unit, ub1 = min(ub0 + p.bands_per_unit, p.bands_per_plane);
        const int uy0 = ub0 * p.band_rows, uy1 = min(ub1 * p.band_rows, p.H) - 1;  // inclusive
        const long long sample = plane / p.C;
        const int cls = (int)(plane - sample * p.C);
        long long ns = p.sizes_i64 ? reinterpret_cast<const long long*>(p.sizes)[sample]
                                   : (long long)reinterpret_cast<const int*>(p.sizes)[sample];
        ns = ns < 0 ? 0 : (ns > p.T ? p.T : ns);
        const int n = (int)ns;
        const bool class_ok = LOSS || cls < p.num_k;
        const float k = LOSS ? p.dh_k : (class_ok ? p.k_for_classes[cls] : 0.0f);
        double lsum = 0.0;  // LOSS: this thread's share of the unit
        int npos = 0;
        const bool kneg = k < 0.0f;
        const bool count = (ug == 0 && cls == 0);
        const long long src = sample * p.T;
        const float* const in_plane = p.in ? p.in + plane * plane_elems : nullptr;
        float* const out_plane = p.out ? p.out + plane * plane_elems : nullptr;
        const bool single = n <= kGsListCap;

        auto build = [&](int c0, bool do_count) {  // list of slots [c0, c0 + cap) (block-wide; ends with a barrier)
            __syncthreads();
            if (tid == 0) {
                s_cnt = 0;
                s_mask = 0u;
            }
            __syncthreads();
            const int c1 = min(n, c0 + kGsListCap);
            for (int base = c0; base < c1; base += kGsThreads) {
                const int t = base + tid;
                GObj o;
                bool pred = false;
                if (t < c1 && class_ok) {
                    if constexpr (LOSS) pred = gs_make_dh(p, src + t, cls, uy0, uy1, o);
                    else pred = gs_make(p, src + t, cls, uy0, uy1, do_count, o);
                }
                const unsigned bal = __ballot_sync(0xffffffffu, pred);
                if (bal) {
                    unsigned bits = 0u;
                    if (pred) {
                        const int lo = (max(o.y0, uy0) - uy0) / p.band_rows, hi = (min(o.y1, uy1) - uy0) / p.band_rows;
                        bits = ((2u << (hi - lo)) - 1u) << lo;
                    }
                    bits = __reduce_or_sync(0xffffffffu, bits);
                    int pos = 0;
                    if (lane == 0) {
                        pos = atomicAdd(&s_cnt, __popc(bal));
                        atomicOr(&s_mask, bits);
                    }
                    pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(bal & ((1u << lane) - 1u));
                    if (pred) s_list[pos] = o;
                }
            }
            __syncthreads();
            return s_cnt;
        };

        int cnt = single ? build(0, count) : 0;
        const unsigned mask = (single && !LOSS) ? s_mask : 0xffffffffu;  // LOSS: every pixel has a loss term
        if constexpr (FRESH) {
            if (tid < 32) {  // lane l starts the run of untouched bands that begins at band ub0 + l
                const int nb = ub1 - ub0;
                const unsigned z = ~mask & (nb >= 32 ? 0xffffffffu : ((1u << nb) - 1u));
                if (((z >> tid) & 1u) && (tid == 0 || !((z >> (tid - 1)) & 1u))) {
                    const unsigned inv = ~(z >> tid);
                    const int run = inv ? __ffs(inv) - 1 : 32 - tid;
                    const int r0 = (ub0 + tid) * p.band_rows, r1 = min(r0 + run * p.band_rows, p.H);
                    char* dst = reinterpret_cast<char*>(out_plane + (long long)r0 * p.W);
                    long long bytes = (long long)(r1 - r0) * p.W * 4;
                    while (bytes > 0) {
                        const unsigned c = (unsigned)(bytes < kGsZeroBytes ? bytes : kGsZeroBytes);
                        bulk_store(dst, s_zero, c);
                        dst += c;
                        bytes -= c;
                    }
                    bulk_commit();
                }
            }
        }
        for (int b = ub0; b < ub1; ++b) {
            const int by0 = b * p.band_rows, by1 = min(by0 + p.band_rows, p.H) - 1;  // inclusive
            const int nitems = (by1 - by0 + 1) * wunits;
            if (!((mask >> (b - ub0)) & 1u)) {  // nothing lands in this band
                if constexpr (!INPLACE && !FRESH) {
                    // whole rows: the band is one contiguous range of the plane
                    const long long base = (long long)by0 * wunits;
                    if constexpr (PX == 4) {
                        const float4* const i4 = in_plane ? reinterpret_cast<const float4*>(in_plane) + base : nullptr;
                        float4* const o4 = reinterpret_cast<float4*>(out_plane) + base;
                        // four loads in flight before the first store: written as `o4[it] = i4[it]` the compiler has to
                        // assume that out aliases in and serialises every load behind the previous store
                        for (int it0 = tid; it0 < nitems; it0 += kGsThreads * 4) {
                            float4 a[4];
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const int it = it0 + u * kGsThreads;
                                a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (i4 != nullptr && it < nitems) a[u] = __ldg(i4 + it);
                            }
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const int it = it0 + u * kGsThreads;
                                if (it < nitems) o4[it] = a[u];
                            }
                        }
                    } else {
                        for (int it = tid; it < nitems; it += kGsThreads) out_plane[base + it] = in_plane ? in_plane[base + it] : 0.0f;
                    }
                }
                continue;
            }

            for (int item0 = 0; item0 < nitems; item0 += kGsThreads * kGsItems) {
                float best[kGsItems][PX];
                unsigned touched = 0u;  // bit (q*PX + j)
                int rows[kGsItems], cols[kGsItems];
#pragma unroll
                for (int q = 0; q < kGsItems; ++q) {
                    const int it = item0 + q * kGsThreads + tid;
                    const bool valid = it < nitems;
                    // item -> (row, col): the divisions are hoisted out of the band loop for the first (usually only) pass
                    const int dr = item0 == 0 ? drow0[q] : it / wunits;
                    const int dc = item0 == 0 ? dcol0[q] : (it % wunits) * PX;
                    rows[q] = valid ? by0 + dr : -1 - (1 << 30);  // invalid items fail every row test
                    cols[q] = valid ? dc : 0;
#pragma unroll
                    for (int j = 0; j < PX; ++j) best[q][j] = INF;
                }
                // input pixels of all items up front: the loads are in flight while the objects are evaluated (issued one
                // item at a time next to their stores they serialise on the memory latency: out may alias in)
                float vin[kGsItems][PX];
#pragma unroll
                for (int q = 0; q < kGsItems; ++q) {
#pragma unroll
                    for (int j = 0; j < PX; ++j) vin[q][j] = 0.0f;
                    if (in_plane != nullptr && rows[q] >= 0) {
                        const int idx = rows[q] * p.W + cols[q];
                        if constexpr (PX == 4) {
                            const float4 a = __ldg(reinterpret_cast<const float4*>(in_plane + idx));
                            vin[q][0] = a.x; vin[q][1] = a.y; vin[q][2] = a.z; vin[q][3] = a.w;
                        } else {
                            vin[q][0] = __ldg(in_plane + idx);
                        }
                    }
                }
                for (int c0 = 0; c0 < (single ? 1 : n); c0 += kGsListCap) {
                    if (!single) cnt = build(c0, count && item0 == 0 && b == ub0);
                    for (int e = 0; e < cnt; ++e) {
                        const GObj o = s_list[e];  // broadcast
                        if (o.y1 < by0 || o.y0 > by1) continue;  // block-uniform: the object misses this band of the unit
#pragma unroll
                        for (int q = 0; q < kGsItems; ++q) {
                            if (rows[q] < o.y0 || rows[q] > o.y1) continue;
                            if (cols[q] > o.x1 || cols[q] + (PX - 1) < o.x0) continue;
                            const float fdy = (float)rows[q] - o.cy;
                            const float fy2 = __fmul_rn(fdy, fdy);
#pragma unroll
                            for (int j = 0; j < PX; ++j) {
                                const int x = cols[q] + j;
                                const float fdx = (float)x - o.cx;
                                // (float)(dy^2 + dx^2): both squares are exact (< 2^24), the sum rounds once like the int -> float cast
                                float t = __fmul_rn(__fadd_rn(fy2, __fmul_rn(fdx, fdx)), o.inv);
                                t = kneg ? -t : t;
                                if (x >= o.x0 && x <= o.x1 && t == t) {  // NaN (0 * inf) never wins the reference's max
                                    best[q][j] = fminf(best[q][j], t);
                                    touched |= 1u << (q * PX + j);
                                }
                            }
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < kGsItems; ++q) {
                    if (rows[q] < 0) continue;
                    const unsigned tq = (touched >> (q * PX)) & ((1u << PX) - 1u);
                    if (INPLACE && tq == 0u) continue;
                    const int idx = rows[q] * p.W + cols[q];  // H*W < 2^31
                    float v[PX];
#pragma unroll
                    for (int j = 0; j < PX; ++j) v[j] = vin[q][j];
                    if constexpr (LOSS) {
                        // v[] = prediction. Target of the pixel exactly as the FRESH renderer would have stored it, then
                        //   target == 1: loss = -(1-p)^alpha log(p+eps);  else: loss = -p^alpha log(1-p+eps) (1-target)^gamma
                        // Branch-free (one logarithm per pixel, the four pixels of the item independent of each other: the
                        // two-sided version ran at 54 % issue-active on dependency and instruction-fetch stalls):
                        //   pos = target == 1;  a = pos ? p : 1-p;  b = pos ? 1-p : p;  w = pos ? 1 : (1-target)^gamma
                        //   loss = -log(a+eps) b^alpha w;   d loss / d p = (pos ? -1 : 1) w (b^alpha / (a+eps) - alpha b^(alpha-1) log(a+eps))
                        float lq = 0.0f;
                        float tg[PX];
#pragma unroll
                        for (int j = 0; j < PX; ++j) tg[j] = 0.0f;
                        if (tq != 0u) {  // one expf per pixel of a touched item; untouched pixels hold +inf -> expf(-inf) = 0
#pragma unroll
                            for (int j = 0; j < PX; ++j) {
                                const float val = __fmul_rn(expf(-best[q][j]), k);
                                tg[j] = ((tq & (1u << j)) && !(0.0f >= val)) ? val : 0.0f;
                            }
                        }
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const float pr = v[j], om = 1.0f - pr;
                            const bool pos = tg[j] == 1.0f;
                            const float den = (pos ? pr : om) + p.eps, bs = pos ? om : pr;
                            const float lg = FAST ? __logf(den) : log_pos(den);
                            const float ot = 1.0f - tg[j], ot2 = ot * ot;
                            const float wgt = pos ? 1.0f : (std_exp ? ot2 * ot2 : powf(ot, p.gamma));
                            const float pw = std_exp ? bs * bs : powf(bs, p.alpha);
                            lq += -lg * pw * wgt;
                            npos += pos ? 1 : 0;
                            if constexpr (GRAD) {
                                const float pw1 = std_exp ? bs : powf(bs, p.alpha - 1.0f);
                                const float gg = wgt * (__fdividef(pw, den) - p.alpha * pw1 * lg);
                                v[j] = pos ? -gg : gg;
                            }
                        }
                        lsum += (double)lq;
                        if constexpr (GRAD) {
                            if (p.out != nullptr) {
                                if constexpr (PX == 4) *reinterpret_cast<float4*>(out_plane + idx) = make_float4(v[0], v[1], v[2], v[3]);
                                else out_plane[idx] = v[0];
                            }
                        }
                    } else {
                    if (tq != 0u) {  // one expf per pixel of a touched item, unconditionally: no per-pixel divergence
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const float t = kneg ? -best[q][j] : best[q][j];
                            const float val = __fmul_rn(k, expf(-t));
                            v[j] = ((tq & (1u << j)) && v[j] < val) ? val : v[j];  // std::max(pixel, val)
                        }
                    }
                    if constexpr (PX == 4) *reinterpret_cast<float4*>(out_plane + idx) = make_float4(v[0], v[1], v[2], v[3]);
                    else out_plane[idx] = v[0];
                    }
                }
            }
        }  // bands of the unit
        if constexpr (LOSS) {  // fixed-order reduction of the unit: lanes (xor tree) -> warps (in order) -> partial[u]
            __shared__ double s_l[kGsThreads / 32];
            __shared__ int s_n[kGsThreads / 32];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                lsum += __shfl_xor_sync(0xffffffffu, lsum, d);
                npos += __shfl_xor_sync(0xffffffffu, npos, d);
            }
            __syncthreads();  // previous unit's reader is done
            if (lane == 0) {
                s_l[tid >> 5] = lsum;
                s_n[tid >> 5] = npos;
            }
            __syncthreads();
            if (tid == 0) {
                double a = 0.0;
                long long c = 0;
#pragma unroll
                for (int w = 0; w < kGsThreads / 32; ++w) {
                    a += s_l[w];
                    c += s_n[w];
                }
                p.partial_loss[u] = a;
                p.partial_pos[u] = c;
            }
        }
    }
    if constexpr (FRESH) bulk_wait_read<0>();  // the zero tile must outlive every bulk read this thread issued
}

}  // namespace dh

namespace dh {
// Bands (<= kGsThreads*kGsItems*PX pixels of whole rows) and units (<= 32 bands, ~128 KB) of the pixel-owner kernels.
static void gs_plan(GsParams& p, int sms, bool vec) {
    const long long H = p.H, W = p.W;
    const long long band_px = (long long)kGsThreads * kGsItems * (vec ? 4 : 1);
    long long rows = band_px / W;
    if (rows < 1) rows = 1;
    if (rows > H) rows = H;
    // small problems: more, shorter bands so that every SM gets work
    while (rows > 1 && p.num_planes * ((H + rows - 1) / rows) < 2ll * sms) rows = (rows + 1) / 2;
    p.band_rows = (int)rows;
    p.bands_per_plane = (int)((H + rows - 1) / rows);
    // unit = consecutive bands of one plane sharing one scan of the sample's slots (~128 KB of output, <= 32 bands);
    // fewer bands per unit on small problems so that every SM still sees several units
    long long bpu = (131072 / 4) / (rows * W);
    if (bpu < 1) bpu = 1;
    if (bpu > 32) bpu = 32;
    if (bpu > p.bands_per_plane) bpu = p.bands_per_plane;
    while (bpu > 1 && p.num_planes * ((p.bands_per_plane + bpu - 1) / bpu) < 8ll * sms) bpu = (bpu + 1) / 2;
    p.units_per_plane = (int)((p.bands_per_plane + bpu - 1) / bpu);
    p.bands_per_unit = (int)((p.bands_per_plane + p.units_per_plane - 1) / p.units_per_plane);
    p.units_per_plane = (p.bands_per_plane + p.bands_per_unit - 1) / p.bands_per_unit;
    p.total_units = p.num_planes * p.units_per_plane;
}

// Sum of the per-unit partials in a fixed order (one block): out2[0] = loss sum, out2[1] = number of target == 1 pixels.
__global__ void __launch_bounds__(1024) focal_reduce_kernel(const double* __restrict__ pl, const long long* __restrict__ pp,
                                                            long long n, double* __restrict__ out2) {
    __shared__ double s_a[1024];
    __shared__ long long s_c[1024];
    double a = 0.0;
    long long c = 0;
    for (long long i = threadIdx.x; i < n; i += 1024) {
        a += pl[i];
        c += pp[i];
    }
    s_a[threadIdx.x] = a;
    s_c[threadIdx.x] = c;
    __syncthreads();
    for (int d = 512; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) {
            s_a[threadIdx.x] += s_a[threadIdx.x + d];
            s_c[threadIdx.x] += s_c[threadIdx.x + d];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out2[0] = s_a[0];
        out2[1] = (double)s_c[0];
    }
}
}  // namespace dh

namespace dh {


int launch_weighted_fresh(cudaStream_t stream, float* heat, const int32_t* centers, const int32_t* radii_eff,
                          const float* weights, const int32_t* labels, const void* sizes, int sizes_i64,
                          const float* k_classes, long long B, long long T, long long C, long long H, long long W);  // dh_kernels.cu

// FRESH fast path, pass 0: classify every slot. An object whose reference window is the symmetric integer square
// [c - R, c + R]^2 (always the case unless the fp32 add centre -/+ radius rounds across an integer, the radius is
// negative or sigma^2 under/overflows) is handed to the main renderer (groups_kernel, SRC_WEIGHTED) as (R, 1/(2 sigma^2));
// the rest is flagged for the exact pixel-owner kernel, which then runs in place over the flagged objects only.
__global__ void __launch_bounds__(256) gs_prepare_kernel(const GsParams p, long long B, int32_t* __restrict__ radii_eff,
                                                         float* __restrict__ weights, uint8_t* __restrict__ special,
                                                         int* __restrict__ n_special) {
    const long long total = B * p.T;
    int mine = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long b = i / p.T, t = i - b * p.T;
        const long long n = p.sizes_i64 ? reinterpret_cast<const long long*>(p.sizes)[b] : (long long)reinterpret_cast<const int*>(p.sizes)[b];
        int reff = -1;
        float w = 0.0f;
        bool sp = false;
        const int cls = p.slice_ids ? p.slice_ids[i] : 0;
        if (t < n && (p.active == nullptr || p.active[i] != 0) && cls >= 0 && cls < p.C && cls < p.num_k) {
            const float r = p.radii[i];
            const int2 cc = *reinterpret_cast<const int2*>(p.centers + 2 * i);
            const float fx = (float)cc.x, fy = (float)cc.y;
            if (fabsf(r) <= kGsMaxRadius && fabsf(fx) <= 16777216.0f && fabsf(fy) <= 16777216.0f) {
                const int el = cc.x - (int)floorf(__fsub_rn(fx, r)), er = (int)ceilf(__fadd_rn(fx, r)) - cc.x;
                const int et = cc.y - (int)floorf(__fsub_rn(fy, r)), eb = (int)ceilf(__fadd_rn(fy, r)) - cc.y;
                const float sigma = __fmul_rn(r, p.factor);
                const float inv = __fdiv_rn(1.0f, __fmul_rn(__fmul_rn(2.0f, sigma), sigma));
                if (el == er && et == eb && el == et && el >= 0 && inv > 0.0f && inv < __int_as_float(0x7f800000)) {
                    reff = el;
                    w = inv;
                } else {
                    sp = true;
                }
            }
        }
        radii_eff[i] = reff;
        weights[i] = w;
        special[i] = sp ? 1 : 0;
        mine += sp ? 1 : 0;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(n_special, mine);
}

}  // namespace dh

extern "C" size_t dh_draw_gaussians_workspace_bytes(int64_t B, int64_t T) {
    if (B <= 0 || T <= 0) return 16;
    const size_t n = (size_t)B * (size_t)T;
    return ((n * 4 + 15) & ~(size_t)15) * 2 + ((n + 15) & ~(size_t)15) + 16;
}

extern "C" int dh_draw_gaussians(void* stream, const float* heatmap_in, float* heatmap_out, const int32_t* centers,
                                 const float* radii, const int32_t* slice_ids, const uint8_t* active,
                                 const void* nums_targets, int nums_dtype, const float* k_for_classes, int64_t num_k,
                                 int64_t B, int64_t T, int64_t C, int64_t H, int64_t W, float radius_to_sigma_factor,
                                 dh_status* status, void* workspace, size_t workspace_bytes) {
    if (B < 0 || T < 0 || C < 0 || num_k < 0) return DH_ERR_INVALID_ARGUMENT;
    if (nums_dtype != DH_SIZES_I32 && nums_dtype != DH_SIZES_I64) return DH_ERR_INVALID_ARGUMENT;
    if (H < 1 || W < 1) return DH_ERR_INVALID_ARGUMENT;
    if (W > (1ll << 23) || H > (1ll << 23) || H * W >= (1ll << 31) || T > INT32_MAX || C > INT32_MAX)
        return DH_ERR_UNSUPPORTED_SHAPE;
    if (B == 0 || C == 0) return DH_OK;
    if (!heatmap_out || !nums_targets || (num_k > 0 && !k_for_classes)) return DH_ERR_INVALID_ARGUMENT;
    if (T > 0 && (!centers || !radii)) return DH_ERR_INVALID_ARGUMENT;
    if (reinterpret_cast<uintptr_t>(centers) & 7u) return DH_ERR_INVALID_ARGUMENT;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return DH_ERR_NO_DEVICE;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return DH_ERR_NO_DEVICE;
    const bool inplace = heatmap_in == heatmap_out;
    if (inplace && T == 0) return DH_OK;
    dh::GsParams p{};
    p.in = heatmap_in;
    p.out = heatmap_out;
    p.centers = centers;
    p.radii = radii;
    p.slice_ids = slice_ids;
    p.active = active;
    p.sizes = nums_targets;
    p.sizes_i64 = nums_dtype == DH_SIZES_I64;
    p.k_for_classes = k_for_classes;
    p.num_k = (int)(num_k > INT32_MAX ? INT32_MAX : num_k);
    p.num_planes = B * C;
    p.C = (int)C;
    p.T = (int)T;
    p.H = (int)H;
    p.W = (int)W;
    p.factor = radius_to_sigma_factor;
    p.status = status;
    const bool vec = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(heatmap_out) & 15u) == 0) &&
                     (heatmap_in == nullptr || (reinterpret_cast<uintptr_t>(heatmap_in) & 15u) == 0);
    cudaStream_t s = (cudaStream_t)stream;
    // FRESH fast path (needs the caller's workspace): classify -> main renderer for the symmetric objects (bulk-store /
    // cluster-launch-control path of groups_kernel) -> exact kernel in place over the few flagged objects, if any.
    // Three launches: only worth it when the output is large (measured break-even ~200 MB on B200) — the CALLER decides by
    // passing the workspace or not; the torch binding passes one from 256 MiB of output on.
    if (heatmap_in == nullptr && vec && T > 0 && workspace != nullptr && status == nullptr &&
        workspace_bytes >= dh_draw_gaussians_workspace_bytes(B, T) && (reinterpret_cast<uintptr_t>(workspace) & 15u) == 0) {
        const size_t n = (size_t)B * (size_t)T, a4 = (n * 4 + 15) & ~(size_t)15, a1 = (n + 15) & ~(size_t)15;
        char* ws = (char*)workspace;
        int32_t* radii_eff = (int32_t*)ws;
        float* weights = (float*)(ws + a4);
        uint8_t* special = (uint8_t*)(ws + 2 * a4);
        int* n_special = (int*)(ws + 2 * a4 + a1);
        cudaError_t e = cudaMemsetAsync(n_special, 0, 4, s);
        if (e != cudaSuccess) return (int)e;
        long long pb = ((long long)n + 255) / 256;
        if (pb > (long long)sms * 8) pb = (long long)sms * 8;
        dh::gs_prepare_kernel<<<(int)pb, 256, 0, s>>>(p, B, radii_eff, weights, special, n_special);
        const int st = dh::launch_weighted_fresh(s, heatmap_out, centers, radii_eff, weights, slice_ids, nums_targets,
                                                 p.sizes_i64, k_for_classes, B, T, C, H, W);
        if (st == DH_OK) {
            dh::GsParams q = p;  // exact kernel, in place, flagged objects only; returns at once when there are none
            q.in = heatmap_out;
            q.active = special;
            q.gate = n_special;
            dh::gs_plan(q, sms, vec);
            const long long g2 = q.total_units < 0x7fffffffll ? q.total_units : 0x7fffffffll;
            dh::gaussians_kernel<4, true, false><<<(int)g2, dh::kGsThreads, 0, s>>>(q);
            dh::count_launches(2);  // + the groups_kernel counted by its launcher
            return (int)cudaGetLastError();
        }
        if (st != DH_ERR_UNSUPPORTED_SHAPE) return st;
    }
    dh::gs_plan(p, sms, vec);
    const long long g = p.total_units < 0x7fffffffll ? p.total_units : 0x7fffffffll;
    if (vec) {
        if (inplace) dh::gaussians_kernel<4, true, false><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else if (heatmap_in == nullptr) dh::gaussians_kernel<4, false, true><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else dh::gaussians_kernel<4, false, false><<<(int)g, dh::kGsThreads, 0, s>>>(p);
    } else {
        if (inplace) dh::gaussians_kernel<1, true, false><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else dh::gaussians_kernel<1, false, false><<<(int)g, dh::kGsThreads, 0, s>>>(p);
    }
    dh::count_launches(1);
    return (int)cudaGetLastError();
}

// ================================================================================================================
// Front half of the reference's BoundingBoxToHeatmapConverter (everything between the annotation boxes and
// draw_gaussians), one thread per RaggedBatch slot, one rounding per reference operation. Paths under
// /root/reference/packages/dali_pipeline_framework/accvlab/dali_pipeline_framework:
//   operators_impl/numba_operators/numba_operators.py:811-815    centre from the box
//   operators_impl/python_operator_functions/python_operator_functions.py:140-171  scale, clip, sizes, area fraction
//   processing_steps/bounding_box_to_heatmap_converter.py:352-355  floor -> int32 centre, sub-pixel offset
//   python_operator_functions.py:213-251 + converter :369-371      is_active (class / size / area-fraction / is_valid)
//   numba_operators.py:865-904 + converter :376                    radius (float64 product stored to fp32), clamp
// The reference runs these as ~10 DALI CPU operators (python_function / numba_function) per sample.
namespace dh {

struct ConvParams {
    const float* bboxes;
    const float* centers;    // may be null: centre of the box
    const int32_t* classes;  // may be null: no class check, sizes against row 0 of min_sizes
    const uint8_t* valid;    // may be null
    const void* sizes;
    int sizes_i64;
    const float* scale_xy;    // [B,2]
    const double* min_sizes;  // [num_min_sizes,2] (h,w) or null
    int num_min_sizes, num_classes;
    long long B, T;
    float w_max, h_max;  // heatmap W-1, H-1
    double sf, min_frac;
    float min_r, max_r;
    float* bboxes_clipped;
    int32_t* center_int;
    float* center_offset;
    float* radii;
    uint8_t* is_active;
    float* hw_clipped;
    float* fraction;
};

__global__ void __launch_bounds__(256) converter_kernel(const ConvParams p) {
    const long long total = p.B * p.T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long b = i / p.T;
        const long long t = i - b * p.T;
        long long n = p.sizes_i64 ? reinterpret_cast<const long long*>(p.sizes)[b] : (long long)reinterpret_cast<const int*>(p.sizes)[b];
        float x1 = 0.f, y1 = 0.f, x2 = 0.f, y2 = 0.f, offx = 0.f, offy = 0.f, rad = 0.f, hc = 0.f, wc = 0.f, frac = 0.f;
        int cix = 0, ciy = 0;
        bool act = false;
        if (t < n) {
            const float4 bb = reinterpret_cast<const float4*>(p.bboxes)[i];
            float cx0, cy0;
            if (p.centers) {
                const float2 c = reinterpret_cast<const float2*>(p.centers)[i];
                cx0 = c.x;
                cy0 = c.y;
            } else {
                cx0 = __fdiv_rn(__fadd_rn(bb.x, bb.z), 2.0f);
                cy0 = __fdiv_rn(__fadd_rn(bb.y, bb.w), 2.0f);
            }
            const float sx = p.scale_xy[2 * b], sy = p.scale_xy[2 * b + 1];
            const float xs1 = __fmul_rn(sx, bb.x), xs2 = __fmul_rn(sx, bb.z), ys1 = __fmul_rn(sy, bb.y), ys2 = __fmul_rn(sy, bb.w);
            x1 = fminf(fmaxf(xs1, 0.0f), p.w_max);
            x2 = fminf(fmaxf(xs2, 0.0f), p.w_max);
            y1 = fminf(fmaxf(ys1, 0.0f), p.h_max);
            y2 = fminf(fmaxf(ys2, 0.0f), p.h_max);
            hc = fabsf(__fsub_rn(y2, y1));
            wc = fabsf(__fsub_rn(x2, x1));
            const float ho = fabsf(__fsub_rn(ys2, ys1)), wo = fabsf(__fsub_rn(xs2, xs1));
            frac = __fdiv_rn(__fmul_rn(hc, wc), __fmul_rn(ho, wo));  // 0/0 = NaN: never active
            const float cx = fminf(fmaxf(__fmul_rn(sx, cx0), 0.0f), p.w_max);
            const float cy = fminf(fmaxf(__fmul_rn(sy, cy0), 0.0f), p.h_max);
            const float fx = floorf(cx), fy = floorf(cy);
            cix = (int)fx;
            ciy = (int)fy;
            offx = __fsub_rn(cx, fx);
            offy = __fsub_rn(cy, fy);
            act = true;
            int row = 0;
            if (p.classes) {
                const int c = p.classes[i];
                act = c < p.num_classes;
                row = (act && c >= 0) ? c : 0;
            }
            if (p.min_sizes && p.num_min_sizes > 0) {
                row = row < p.num_min_sizes ? row : p.num_min_sizes - 1;
                act = act && ((double)hc >= p.min_sizes[2 * row]) && ((double)wc >= p.min_sizes[2 * row + 1]);
            }
            act = act && ((double)frac >= p.min_frac);
            if (p.valid) act = act && (p.valid[i] != 0);
            const float l = fminf(x1, x2), r = fmaxf(x1, x2), tp = fminf(y1, y2), bt = fmaxf(y1, y2);
            const float dist = fminf(fminf(__fsub_rn(cx, l), __fsub_rn(cy, tp)), fminf(__fsub_rn(r, cx), __fsub_rn(bt, cy)));
            const double dd = fmax((double)dist, 0.0);
            rad = __double2float_rn(__dmul_rn(dd, p.sf));
            rad = fminf(fmaxf(p.min_r, rad), p.max_r);
        }
        if (p.bboxes_clipped) reinterpret_cast<float4*>(p.bboxes_clipped)[i] = make_float4(x1, y1, x2, y2);
        reinterpret_cast<int2*>(p.center_int)[i] = make_int2(cix, ciy);
        if (p.center_offset) reinterpret_cast<float2*>(p.center_offset)[i] = make_float2(offx, offy);
        p.radii[i] = rad;
        p.is_active[i] = act ? 1 : 0;
        if (p.hw_clipped) reinterpret_cast<float2*>(p.hw_clipped)[i] = make_float2(hc, wc);
        if (p.fraction) p.fraction[i] = frac;
    }
}

}  // namespace dh

extern "C" int dh_boxes_to_gaussian_params(void* stream, const float* bboxes, const float* centers, const int32_t* classes,
                                           const uint8_t* is_valid, const void* nums_targets, int nums_dtype,
                                           const float* scale_xy, const double* min_sizes_hw, int64_t num_min_sizes,
                                           int64_t num_classes, int64_t B, int64_t T, int64_t heatmap_h, int64_t heatmap_w,
                                           double radius_scaling_factor, float min_radius, float max_radius,
                                           double min_fraction_area, float* bboxes_clipped, int32_t* center_int,
                                           float* center_offset, float* radii, uint8_t* is_active, float* hw_clipped,
                                           float* fraction_area) {
    if (B < 0 || T < 0 || heatmap_h < 1 || heatmap_w < 1 || num_min_sizes < 0) return DH_ERR_INVALID_ARGUMENT;
    if (nums_dtype != DH_SIZES_I32 && nums_dtype != DH_SIZES_I64) return DH_ERR_INVALID_ARGUMENT;
    if (heatmap_h > (1ll << 23) || heatmap_w > (1ll << 23) || num_min_sizes > INT32_MAX) return DH_ERR_UNSUPPORTED_SHAPE;
    if (B == 0 || T == 0) return DH_OK;
    if (!bboxes || !nums_targets || !scale_xy || !center_int || !radii || !is_active) return DH_ERR_INVALID_ARGUMENT;
    if (num_min_sizes > 0 && !min_sizes_hw) return DH_ERR_INVALID_ARGUMENT;
    if ((reinterpret_cast<uintptr_t>(bboxes) & 15u) || (reinterpret_cast<uintptr_t>(centers) & 7u) ||
        (reinterpret_cast<uintptr_t>(center_int) & 7u) || (reinterpret_cast<uintptr_t>(bboxes_clipped) & 15u) ||
        (reinterpret_cast<uintptr_t>(center_offset) & 7u) || (reinterpret_cast<uintptr_t>(hw_clipped) & 7u))
        return DH_ERR_INVALID_ARGUMENT;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return DH_ERR_NO_DEVICE;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return DH_ERR_NO_DEVICE;
    dh::ConvParams p{};
    p.bboxes = bboxes;
    p.centers = centers;
    p.classes = classes;
    p.valid = is_valid;
    p.sizes = nums_targets;
    p.sizes_i64 = nums_dtype == DH_SIZES_I64;
    p.scale_xy = scale_xy;
    p.min_sizes = min_sizes_hw;
    p.num_min_sizes = (int)num_min_sizes;
    p.num_classes = (int)(num_classes > INT32_MAX ? INT32_MAX : (num_classes < INT32_MIN ? INT32_MIN : num_classes));
    p.B = B;
    p.T = T;
    p.w_max = (float)(heatmap_w - 1);
    p.h_max = (float)(heatmap_h - 1);
    p.sf = radius_scaling_factor;
    p.min_frac = min_fraction_area;
    p.min_r = min_radius;
    p.max_r = max_radius;
    p.bboxes_clipped = bboxes_clipped;
    p.center_int = center_int;
    p.center_offset = center_offset;
    p.radii = radii;
    p.is_active = is_active;
    p.hw_clipped = hw_clipped;
    p.fraction = fraction_area;
    long long blocks = (B * T + 255) / 256;
    if (blocks > (long long)sms * 8) blocks = (long long)sms * 8;
    dh::converter_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p);
    dh::count_launches(1);
    return (int)cudaGetLastError();
}

// ================================================================================================================
// N3: fused Gaussian-focal-loss consumer. The target heatmap of draw_heatmap_batched is never
// written to HBM: every pixel's target is formed in registers (same arithmetic as the FRESH renderer) and consumed
// by the loss of /root/reference/packages/draw_heatmap/docs/intro.rst:7-26 (the reference states the formula only;
// it has no implementation -> PARITY UNPINNED). Traffic: 4 B/pixel read (+4 B/pixel written when the gradient is
// requested) instead of 4 written + 4 read for the target plus the loss's own passes.
static int focal_device(int& sms) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return DH_ERR_NO_DEVICE;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return DH_ERR_NO_DEVICE;
    return DH_OK;
}

extern "C" size_t dh_focal_loss_workspace_bytes(int64_t B, int64_t C, int64_t H, int64_t W) {
    if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return 16;
    int sms = 0;
    if (focal_device(sms) != DH_OK) sms = 148;
    long long units = 0;
    for (int vec = 0; vec < 2; ++vec) {
        dh::GsParams p{};
        p.num_planes = B * C;
        p.H = (int)H;
        p.W = (int)W;
        dh::gs_plan(p, sms, vec != 0);
        if (p.total_units > units) units = p.total_units;
    }
    return (size_t)units * 16 + 16;
}

extern "C" int dh_gaussian_focal_loss(void* stream, const float* pred, const int32_t* centers, const int32_t* radii,
                                      const void* nums_targets, int nums_dtype, const int32_t* labels, int64_t B,
                                      int64_t T, int64_t C, int64_t H, int64_t W, float diameter_to_sigma_factor,
                                      float k_scale, float alpha, float gamma, float eps, int fast_math,
                                      float* grad_pred, double* loss_out2, void* workspace,
                                      size_t workspace_bytes) {
    if (B < 0 || T < 0 || C < 0) return DH_ERR_INVALID_ARGUMENT;
    if (nums_dtype != DH_SIZES_I32 && nums_dtype != DH_SIZES_I64) return DH_ERR_INVALID_ARGUMENT;
    if (H < 1 || W < 1 || !(k_scale > 0.0f) || !(diameter_to_sigma_factor != 0.0f)) return DH_ERR_INVALID_ARGUMENT;
    if (W > (1ll << 23) || H > (1ll << 23) || H * W >= (1ll << 31) || T > INT32_MAX || C > INT32_MAX)
        return DH_ERR_UNSUPPORTED_SHAPE;
    if (labels == nullptr) C = 1;
    if (!loss_out2) return DH_ERR_INVALID_ARGUMENT;
    cudaStream_t s = (cudaStream_t)stream;
    if (B == 0 || C == 0) return (int)cudaMemsetAsync(loss_out2, 0, 16, s);
    if (!pred || !nums_targets || (T > 0 && (!centers || !radii))) return DH_ERR_INVALID_ARGUMENT;
    if (reinterpret_cast<uintptr_t>(centers) & 7u) return DH_ERR_INVALID_ARGUMENT;
    int sms = 0;
    int st = focal_device(sms);
    if (st != DH_OK) return st;
    dh::GsParams p{};
    p.in = pred;
    p.out = grad_pred;
    p.centers = centers;
    p.radii_i = radii;
    p.slice_ids = labels;
    p.sizes = nums_targets;
    p.sizes_i64 = nums_dtype == DH_SIZES_I64;
    p.num_planes = B * C;
    p.C = (int)C;
    p.T = (int)T;
    p.H = (int)H;
    p.W = (int)W;
    p.factor = diameter_to_sigma_factor;
    p.dh_k = k_scale;
    p.alpha = alpha;
    p.gamma = gamma;
    p.eps = eps;
    p.fast = fast_math != 0;
    const bool vec = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(pred) & 15u) == 0) &&
                     ((reinterpret_cast<uintptr_t>(grad_pred) & 15u) == 0);
    dh::gs_plan(p, sms, vec);
    if (!workspace || workspace_bytes < (size_t)p.total_units * 16 + 16) return DH_ERR_WORKSPACE_TOO_SMALL;
    if (reinterpret_cast<uintptr_t>(workspace) & 15u) return DH_ERR_INVALID_ARGUMENT;
    p.partial_loss = (double*)workspace;
    p.partial_pos = (long long*)((char*)workspace + (((size_t)p.total_units * 8 + 15) & ~(size_t)15));
    const long long g = p.total_units < 0x7fffffffll ? p.total_units : 0x7fffffffll;
    const int mode = (alpha == 2.0f && gamma == 4.0f) ? (fast_math ? 2 : 1) : 3;
    if (vec && grad_pred == nullptr) {  // forward only
        if (mode == 1) dh::gaussians_kernel<4, false, false, 4><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else if (mode == 2) dh::gaussians_kernel<4, false, false, 5><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else dh::gaussians_kernel<4, false, false, 6><<<(int)g, dh::kGsThreads, 0, s>>>(p);
    } else if (vec) {
        if (mode == 1) dh::gaussians_kernel<4, false, false, 1><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else if (mode == 2) dh::gaussians_kernel<4, false, false, 2><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else dh::gaussians_kernel<4, false, false, 3><<<(int)g, dh::kGsThreads, 0, s>>>(p);
    } else {
        if (mode == 1) dh::gaussians_kernel<1, false, false, 1><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else if (mode == 2) dh::gaussians_kernel<1, false, false, 2><<<(int)g, dh::kGsThreads, 0, s>>>(p);
        else dh::gaussians_kernel<1, false, false, 3><<<(int)g, dh::kGsThreads, 0, s>>>(p);
    }
    dh::focal_reduce_kernel<<<1, 1024, 0, s>>>(p.partial_loss, p.partial_pos, p.total_units, loss_out2);
    dh::count_launches(2);
    return (int)cudaGetLastError();
}

// ----------------------------------------------------------------------------------------------------------------
// Backward of the fused loss: grad_out = grad_unscaled * (*scale), the scalar read from DEVICE memory (autograd hands
// the upstream gradient over as a 0-dim CUDA tensor; reading it on the host would synchronise). torch's own
// `tensor * cuda_scalar_tensor` runs at half the copy rate (0.87 ms against 0.43 ms for 1.4 GB, profiles/scripts/
// bwd_probe.py); this is a plain streaming kernel: 8 x 128-bit loads in flight per thread, may run in place.
namespace dh {
constexpr int kScalePerThread = 8;
__global__ void __launch_bounds__(256) scale_kernel(float* __restrict__ dst, const float* __restrict__ src,
                                                    const float* __restrict__ scale, long long n, int vec) {
    const float k = __ldg(scale);
    if (vec) {
        const long long n4 = n >> 2;
        const float4* const s4 = reinterpret_cast<const float4*>(src);
        float4* const d4 = reinterpret_cast<float4*>(dst);
        const long long base = (long long)blockIdx.x * (256 * kScalePerThread) + threadIdx.x;
        float4 v[kScalePerThread];
#pragma unroll
        for (int i = 0; i < kScalePerThread; ++i) {
            const long long j = base + (long long)i * 256;
            if (j < n4) v[i] = s4[j];
        }
#pragma unroll
        for (int i = 0; i < kScalePerThread; ++i) {
            const long long j = base + (long long)i * 256;
            if (j < n4) d4[j] = make_float4(v[i].x * k, v[i].y * k, v[i].z * k, v[i].w * k);
        }
        if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // up to three trailing elements
            const long long j = (n4 << 2) + threadIdx.x;
            dst[j] = src[j] * k;
        }
    } else {
        const long long base = (long long)blockIdx.x * (256 * kScalePerThread) + threadIdx.x;
#pragma unroll
        for (int i = 0; i < kScalePerThread; ++i) {
            const long long j = base + (long long)i * 256;
            if (j < n) dst[j] = src[j] * k;
        }
    }
}
}  // namespace dh

extern "C" int dh_scale_by_device_scalar(void* stream, float* dst, const float* src, const float* scale_device,
                                         int64_t n) {
    if (n < 0) return DH_ERR_INVALID_ARGUMENT;
    if (n == 0) return DH_OK;
    if (!dst || !src || !scale_device) return DH_ERR_INVALID_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return DH_ERR_NO_DEVICE;
    const int vec = ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15u) == 0;
    const long long items = vec ? ((n >> 2) > 0 ? (n >> 2) : 1) : n;
    const long long per_block = 256ll * dh::kScalePerThread;
    const long long blocks = (items + per_block - 1) / per_block;
    if (blocks > 0x7fffffffll) return DH_ERR_UNSUPPORTED_SHAPE;
    dh::scale_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(dst, src, scale_device, n, vec);
    dh::count_launches(1);
    return (int)cudaGetLastError();
}
