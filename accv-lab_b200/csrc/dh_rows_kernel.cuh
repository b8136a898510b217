// dh_rows_kernel.cuh — renderer for WIDE maps (1024 < W <= 4096; everything narrower goes to dh_groups_kernel.cuh):
// row-owner warps + shared-memory tiles + bulk (TMA) stores. Also home of the pieces the other renderers share (ObjI,
// the slot loaders, make_obji).
// Included by dh_kernels.cu after RenderParams / radius_center_one are defined.
//
// Replaces draw_gaussian + atomicMax of the reference
// (/root/reference/packages/draw_heatmap/accvlab/draw_heatmap/include/draw_heatmap_cuda_kernel.cuh:26-49: one thread
// renders a whole object serially with a global CAS loop per pixel) with an OUTPUT-partitioned design:
//   * a band = R whole rows of one plane = ONE contiguous byte range of the heatmap (<= kTileBytes);
//   * a unit = up to 64 consecutive bands of one plane; a persistent CTA walks units grid-stride;
//   * K2: the unit's objects are compacted (warp ballot + popc prefix) into a shared-memory list; each object also ORs
//         the bands it overlaps into a 64-bit occupancy mask; per band the list is compacted once more;
//   * FRESH mode: bands nobody overlaps are written by ONE cp.async.bulk (shared -> global) from a constant zero
//         tile — a single instruction per <= 16 KB of output, no LSU work at all;
//   * K3: in an occupied band every warp owns whole rows of the shared-memory tile: it zero-fills its row, then
//         for every object overlapping the row (warp-uniform test) the lanes sweep ONLY the object's column span and
//         keep min_o (i^2+j^2)*vinv_o as an order-preserving uint (no atomics: the warp owns the row);
//   * K4: the span union of the row is finalised in place with ONE expf per touched pixel
//         (max_o k*exp(-m_o) == k*exp(-min_o m_o), bit for bit, because expf is monotone), then the whole band leaves
//         with one bulk store; two tiles alternate so the store of band i overlaps the rendering of band i+1.
//   * ACCUMULATE mode (reference semantics): only the touched spans are read-max-written with plain coalesced
//         loads/stores; untouched pixels are neither read nor written.
#pragma once
#include "dh_ptx.cuh"

namespace dh {

constexpr int kRowsThreads = 256;
constexpr int kRowsWarps = kRowsThreads / 32;
#ifndef DH_TILE_BYTES
#define DH_TILE_BYTES 16384
#endif
constexpr int kTileBytes = DH_TILE_BYTES;
constexpr int kTilePx = kTileBytes / 4;
constexpr int kRowsCap = 512;  // objects per shared-memory list

struct __align__(16) ObjI {
    int x, y, r;
    float w;  // 1/var (> 0)
};

// entry of a pre-binned object list (bin_kernel, dh_groups2_kernel.cuh)
__device__ __forceinline__ ObjI ld_bin(const void* __restrict__ list, long long i) {
    const int4 q = __ldg(reinterpret_cast<const int4*>(list) + i);
    ObjI o;
    o.x = q.x;
    o.y = q.y;
    o.r = q.z;
    o.w = __int_as_float(q.w);
    return o;
}

// Slot loads of the RaggedBatch arrays: one 8- / 16-byte load when the array base allows it (`p.in_vec`, checked on the
// host), scalar loads otherwise — a contiguous torch tensor may start at any 4-byte offset of its storage, and the
// reference reads these arrays with scalar int loads (cuh:57-58,86-87), so such inputs must keep working.
__device__ __forceinline__ int2 ld_center_i(const RenderParams& p, long long i) {
    if (p.in_vec) return __ldg(reinterpret_cast<const int2*>(p.centers) + i);
    return make_int2(__ldg(p.centers + 2 * i), __ldg(p.centers + 2 * i + 1));
}
__device__ __forceinline__ float2 ld_center_f(const RenderParams& p, long long i) {
    if (p.in_vec) return __ldg(reinterpret_cast<const float2*>(p.centers_f) + i);
    return make_float2(__ldg(p.centers_f + 2 * i), __ldg(p.centers_f + 2 * i + 1));
}
__device__ __forceinline__ float4 ld_box_f(const RenderParams& p, long long i) {
    if (p.in_vec) return __ldg(reinterpret_cast<const float4*>(p.bboxes_f) + i);
    return make_float4(__ldg(p.bboxes_f + 4 * i), __ldg(p.bboxes_f + 4 * i + 1), __ldg(p.bboxes_f + 4 * i + 2),
                       __ldg(p.bboxes_f + 4 * i + 3));
}

// Raw slot -> (x, y, r) for the four sources; false if the slot does not belong to this plane.
template <int SRC>
__device__ __forceinline__ bool load_slot(const RenderParams& p, long long src_base, int t, int cls, long long plane,
                                          bool cnt, int& x, int& y, int& r) {
    if constexpr (SRC == SRC_BATCHED || SRC == SRC_BOXES || SRC == SRC_WEIGHTED) {
        const long long i = src_base + t;
        if (p.labels != nullptr) {
            const int lab = __ldg(p.labels + i);
            if (lab != cls) {
                // the counting unit (class 0, first unit of the sample) sees every slot once: it counts the objects of
                // the OTHER classes that will be skipped too (a bad label, or a valid label with an oversize radius)
                if (cnt && p.status) {
                    if (lab < 0 || lab >= p.C) {
                        atomicAdd(&p.status->skipped_label, 1u);
                    } else {
                        int rr;
                        if constexpr (SRC == SRC_BOXES) {
                            int xx, yy;
                            const float2 c = ld_center_f(p, i);
                            const float4 b = ld_box_f(p, i);
                            radius_center_one(c.x, c.y, b.x, b.y, b.z, b.w, p.osf, xx, yy, rr);
                        } else {
                            rr = __ldg(p.radii + i);
                        }
                        if (rr > DH_MAX_RADIUS) atomicAdd(&p.status->skipped_radius, 1u);
                    }
                }
                return false;
            }
        }
        if constexpr (SRC == SRC_BOXES) {
            const float2 c = ld_center_f(p, i);
            const float4 b = ld_box_f(p, i);
            radius_center_one(c.x, c.y, b.x, b.y, b.z, b.w, p.osf, x, y, r);
        } else {
            const int2 c = ld_center_i(p, i);
            x = c.x;
            y = c.y;
            r = __ldg(p.radii + i);
        }
    } else if constexpr (SRC == SRC_CSR) {
        const long long i = p.csr_items[src_base + t];
        x = p.centers[2 * i];
        y = p.centers[2 * i + 1];
        r = p.radii[i];
    } else {
        const int pi = p.plane_idx[t];
        if ((long long)pi != plane) {
            if (cnt && p.status && (pi < 0 || (long long)pi >= p.num_planes)) atomicAdd(&p.status->skipped_plane, 1u);
            return false;
        }
        x = p.centers[2 * (long long)t];
        y = p.centers[2 * (long long)t + 1];
        r = p.radii[t];
    }
    return true;
}

// Cull against rows [ry0, ry1) of a W-wide map and build the list entry (sigma/var: cuh:61-63 / 90-92).
__device__ __forceinline__ bool make_obji(int x, int y, int r, float factor, int ry0, int ry1, int W, ObjI& o,
                                          dh_status* st, bool count) {
    if (r < 0) return false;  // reference: empty loops
    if (r > DH_MAX_RADIUS) {
        if (count && st) atomicAdd(&st->skipped_radius, 1u);
        return false;
    }
    const long long X = x, Y = y, R = r;
    if (Y + R < ry0 || Y - R >= ry1 || X + R < 0 || X - R >= W) return false;
    const float sigma = __fdiv_rn((float)(2 * r + 1), factor);
    const float var = __fmul_rn(__fmul_rn(2.0f, sigma), sigma);
    o.x = x;
    o.y = y;
    o.r = r;
    o.w = __fdiv_rn(1.0f, var);
    return true;
}

__device__ __forceinline__ void compact_push_i(bool pred, const ObjI& o, ObjI* list, int* cnt) {
    const unsigned bal = __ballot_sync(0xffffffffu, pred);
    if (bal) {
        const int lane = threadIdx.x & 31;
        int base = 0;
        if (lane == 0) base = atomicAdd(cnt, __popc(bal));
        base = __shfl_sync(0xffffffffu, base, 0);
        const int pos = base + __popc(bal & ((1u << lane) - 1u));
        if (pred && pos < kRowsCap) list[pos] = o;
    }
}

constexpr size_t kRowsSmemBytes = 3 * (size_t)kTileBytes + 2 * (size_t)kRowsCap * sizeof(ObjI);

// K3 for one row owned by one warp: sweep the column span of every overlapping object; keep, per pixel, an
// order-preserving uint of m = (i^2 + j^2) * vinv (arithmetic of cuh:45 with the sign pulled out):
//   k >= 0:  t = ~bits(m)    (max t <=> min m <=> max exp(-m));   k < 0:  t = bits(m) + 1  (max t <=> max m)
// t == 0 means "untouched" and is also the bit pattern of the final 0.0f.
__device__ __forceinline__ void render_row(unsigned* __restrict__ rowp, int y, int W, const ObjI* __restrict__ list, int cnt,
                                           bool kneg, int lane, int& ux0, int& ux1) {
    for (int e = 0; e < cnt; ++e) {
        const ObjI o = list[e];  // broadcast LDS.128
        const int dy = y - o.y;
        if (abs(dy) <= o.r) {  // warp-uniform
            const int x0 = max(o.x - o.r, 0), x1 = min(o.x + o.r, W - 1);
            ux0 = min(ux0, x0);
            ux1 = max(ux1, x1);
            const float fdy = (float)dy;
            const float fi = __fmul_rn(fdy, fdy);
            for (int x = x0 + lane; x <= x1; x += 32) {
                const float fdx = (float)(x - o.x);
                const float m = __fmul_rn(__fadd_rn(fi, __fmul_rn(fdx, fdx)), o.w);
                const unsigned b = __float_as_uint(m);
                const unsigned t = kneg ? b + 1u : ~b;
                if (t > rowp[x]) rowp[x] = t;
            }
            __syncwarp();
        }
    }
}

template <bool ACCUM>
__device__ __forceinline__ void finalize_row(unsigned* __restrict__ rowp, float* __restrict__ grow, int x0, int x1, bool kneg,
                                             float k, int lane) {
    for (int x = x0 + lane; x <= x1; x += 32) {
        const unsigned t = rowp[x];
        if (t != 0u) {
            const float m = __uint_as_float(kneg ? t - 1u : ~t);
            const float v = __fmul_rn(expf(-m), k);  // cuh:45
            if constexpr (ACCUM) {
                const float g = grow[x];
                if (!(g >= v)) grow[x] = v;  // cuh:30: keep old iff old >= v
            } else {
                rowp[x] = __float_as_uint((0.0f >= v) ? 0.0f : v);  // max-blend into the fused zero-init
            }
        }
    }
}

template <bool ACCUM, int SRC>
__global__ void __launch_bounds__(kRowsThreads) rows_kernel(const RenderParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned* const tile0 = reinterpret_cast<unsigned*>(smem_raw);
    unsigned* const tile1 = reinterpret_cast<unsigned*>(smem_raw + kTileBytes);
    float4* const zero_tile = reinterpret_cast<float4*>(smem_raw + 2 * kTileBytes);
    ObjI* const s_unit = reinterpret_cast<ObjI*>(smem_raw + 3 * kTileBytes);
    ObjI* const s_band = s_unit + kRowsCap;
    __shared__ int s_cnt[2];
    __shared__ unsigned long long s_mask;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool kneg = p.k < 0.0f;
    const int W = p.W, H = p.H, R = p.band_rows, w4 = W >> 2;
    const long long plane_elems = (long long)H * W;

    if constexpr (!ACCUM) {
        for (int i = tid; i < kTilePx / 4; i += kRowsThreads) zero_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
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

        // ---- K2a: unit list + band-occupancy mask -----------------------------------------------------------
        __syncthreads();  // the previous unit is done with the lists (and the zero tile is initialised)
        if (tid == 0) {
            s_cnt[0] = 0;
            s_mask = 0ull;
        }
        __syncthreads();
        for (int base = 0; base < n; base += kRowsThreads) {
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
            compact_push_i(pred, o, s_unit, &s_cnt[0]);
        }
        __syncthreads();
        const int ucnt = s_cnt[0];
        unsigned long long mask = s_mask;
        const bool single = ucnt <= kRowsCap;

        // ---- FRESH: bands nobody overlaps leave as bulk stores of the constant zero tile ----------------------
        if constexpr (!ACCUM) {
            if (warp == 1) {
                for (int b = lane; b < nb; b += 32) {
                    if (!((mask >> b) & 1ull)) {
                        const int by0 = (b0 + b) * R, rows = min(R, H - by0);
                        bulk_store(plane_ptr + (long long)by0 * W, zero_tile, (unsigned)(rows * W * 4));
                    }
                }
                bulk_commit();
            }
        }

        // ---- occupied bands -----------------------------------------------------------------------------------
        while (mask) {
            const int b = __ffsll((long long)mask) - 1;
            mask &= mask - 1ull;
            const int by0 = (b0 + b) * R, by1 = min(by0 + R, H), nrows = by1 - by0;
            unsigned* const T = ACCUM ? tile0 : (cur ? tile1 : tile0);
            const int npass = single ? 1 : (n + kRowsCap - 1) / kRowsCap;
            for (int pass = 0; pass < npass; ++pass) {
                __syncthreads();  // previous band / pass finished with s_band (and, ACCUM, with the tile)
                if (tid == 0) {
                    s_cnt[1] = 0;
                    if constexpr (!ACCUM) {
                        if (pass == 0) bulk_wait_read<1>();  // the store that last read this tile (2 groups ago) is done
                    }
                }
                __syncthreads();
                // K2b: band list
                if (single) {
                    for (int base = 0; base < ucnt; base += kRowsThreads) {
                        const int e = base + tid;
                        ObjI o;
                        bool pred = false;
                        if (e < ucnt) {
                            o = s_unit[e];
                            pred = (o.y + o.r >= by0) && (o.y - o.r < by1);
                        }
                        compact_push_i(pred, o, s_band, &s_cnt[1]);
                    }
                } else {  // more objects than the list holds: chunks straight from the source, tile persists
                    const int c0 = pass * kRowsCap, c1 = min(n, c0 + kRowsCap);
                    for (int base = c0; base < c1; base += kRowsThreads) {
                        const int t = base + tid;
                        ObjI o;
                        bool pred = false;
                        if (t < c1) {
                            int x, y, r;
                            if (load_slot<SRC>(p, src_base, t, cls, plane, false, x, y, r))
                                pred = make_obji(x, y, r, p.factor, by0, by1, W, o, nullptr, false);
                        }
                        compact_push_i(pred, o, s_band, &s_cnt[1]);
                    }
                }
                __syncthreads();
                const int bcnt = min(s_cnt[1], kRowsCap);
                const bool last = (pass == npass - 1);
                // K3 + K4 (in-tile): every warp owns rows warp, warp+8, ...
                for (int row = warp; row < nrows; row += kRowsWarps) {
                    unsigned* const rowp = T + row * W;
                    if (pass == 0) {
                        uint4* r4 = reinterpret_cast<uint4*>(rowp);
                        for (int i = lane; i < w4; i += 32) r4[i] = make_uint4(0u, 0u, 0u, 0u);
                        __syncwarp();
                    }
                    int ux0 = W, ux1 = -1;
                    render_row(rowp, by0 + row, W, s_band, bcnt, kneg, lane, ux0, ux1);
                    if (last) {
                        if (!single) {
                            ux0 = 0;
                            ux1 = W - 1;
                        }
                        finalize_row<ACCUM>(rowp, plane_ptr + (long long)(by0 + row) * W, ux0, ux1, kneg, p.k, lane);
                    }
                }
            }
            if constexpr (!ACCUM) {
                fence_proxy_async();  // my generic-proxy tile writes -> visible to the bulk-copy engine
                __syncthreads();
                if (tid == 0) {
                    bulk_store(plane_ptr + (long long)by0 * W, T, (unsigned)(nrows * W * 4));
                    bulk_commit();
                }
                cur ^= 1;
            }
        }
    }
    if constexpr (!ACCUM) bulk_wait_read<0>();  // shared memory must outlive every bulk read issued by this thread
}

}  // namespace dh
