// dh_groups2_kernel.cuh — K2 as a PRE-PASS + the warp-autonomous renderer that consumes its lists.
// Included by dh_kernels.cu after dh_groups_kernel.cuh (shares ObjI, render_group, zero_runs, the CLC / mbarrier wrappers).
//
// What round 1's single-launch renderer (groups_kernel) paid per UNIT and this pair pays per OBJECT, once:
//   * every unit re-scanned its sample's T padded slots (label compare on every slot; K1 with 3 IEEE divisions and the
//     sigma / 1/var divisions of cuh:61-63 for every matching one) — cfg3/cfg5: 10 classes x 2-4 units = 20-40 scans of
//     the same 100 slots — between two block barriers, with a third one before the next unit (ncu: `stall barrier`
//     4.9 warps per issue, the top stall; an EMPTY unit cost 2.6-4 us of pure latency);
//   * here bin_kernel reads every slot ONCE: validity (cuh:84), label, K1 when the source is float boxes
//     (tests/_test_helpers.py:20-28), window / sigma arithmetic (cuh:61-68), and counting-sorts the sample's objects by
//     class (shared-memory histogram + block prefix sum + scatter) into a compact per-plane list of 16-byte entries
//     (x, y, r, 1/var) in the caller's workspace, plus one (start, count) record per plane;
//   * groups2_kernel's unit header is ONE 8-byte load; a plane with <= 32 objects (every plane of cfg1-3 and cfg5) is
//     rendered WARP-AUTONOMOUSLY: each warp loads the plane's entries itself (one 16-byte load per lane, L2-resident),
//     derives the unit's group-occupancy mask with a ballot-free __reduce_or, and renders its stripe of groups — no
//     shared-memory list, NO block barrier, warps may run a unit ahead of each other. The next unit still comes from
//     the hardware block scheduler (cluster launch control), handed from the requesting thread to the eight warps
//     through a two-slot full/empty mbarrier ring instead of __syncthreads. Planes with more objects (cfg4) take the
//     block-cooperative list build of groups_kernel, fed from the compact list (no label filter, no divisions).
#pragma once

namespace dh {

constexpr int kBinThreads = 128;
constexpr int kBinMaxC = 2048;  // classes whose counters fit bin_kernel's static shared memory

// Workspace layout of the binned path: [ int2 rec[num_planes] | pad to 16 | int4 list[B*T] ]
__host__ __device__ inline size_t bin_rec_bytes(long long planes) { return ((size_t)planes * 8 + 15) & ~(size_t)15; }
__host__ __device__ inline size_t bin_workspace_bytes(long long B, long long T, long long C) {
    return bin_rec_bytes(B * C) + (size_t)B * (size_t)T * 16;
}

// Slot t of sample `base / T` -> list entry. Invalid objects (r < 0: "empty loops" in the reference; r > DH_MAX_RADIUS;
// window wholly outside the map) become r = -1 entries, which no group test ever hits.
template <int SRC>
__device__ __forceinline__ ObjI bin_make(const RenderParams& p, long long i) {
    int x, y, r;
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
    ObjI o;
    o.x = 0;
    o.y = 0;
    o.r = -1;
    o.w = 0.f;
    if (r > DH_MAX_RADIUS) {
        if (p.status) atomicAdd(&p.status->skipped_radius, 1u);  // every slot is seen exactly once here
        return o;
    }
    make_obji(x, y, r, p.factor, 0, p.H, p.W, o, nullptr, false);  // leaves r = -1 when nothing of it is on the map
    return o;
}

// K2 pre-pass: one CTA per sample.
template <int SRC>
__global__ void __launch_bounds__(kBinThreads) bin_kernel(const RenderParams p, int4* __restrict__ list,
                                                          int2* __restrict__ rec) {
    __shared__ int s_cnt[kBinMaxC];  // per-class count, then the class's write cursor
    __shared__ int s_warp[kBinThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long b = blockIdx.x;
    const long long base = b * p.T;
    // inputs may be the previous kernel's outputs and the workspace its input: nothing global before the wait
    pdl_wait();
    pdl_launch_dependents();
    long long ns = p.sizes_i64 ? __ldg(reinterpret_cast<const long long*>(p.sizes) + b)
                               : (long long)__ldg(reinterpret_cast<const int*>(p.sizes) + b);
    ns = ns < 0 ? 0 : (ns > p.T ? p.T : ns);
    const int n = (int)ns;  // cuh:84: slots t >= n are padding and never drawn
    if (p.labels == nullptr) {  // one plane per sample: slot order is list order
        for (int t = tid; t < n; t += kBinThreads) {
            const ObjI o = bin_make<SRC>(p, base + t);
            list[base + t] = make_int4(o.x, o.y, o.r, __float_as_int(o.w));
        }
        if (tid == 0) rec[b] = make_int2((int)base, n);
        return;
    }
    const int C = p.C;
    for (int c = tid; c < C; c += kBinThreads) s_cnt[c] = 0;
    __syncthreads();
    for (int t = tid; t < n; t += kBinThreads) {
        const int lab = __ldg(p.labels + base + t);
        if (lab >= 0 && lab < C) atomicAdd(&s_cnt[lab], 1);
        else if (p.status) atomicAdd(&p.status->skipped_label, 1u);  // reference: device assert (cuh:102)
    }
    __syncthreads();
    // exclusive prefix sum over the classes: thread `tid` owns classes [tid*per, tid*per + per)
    const int per = (C + kBinThreads - 1) / kBinThreads;
    const int c0 = tid * per, c1 = min(c0 + per, C);
    int mine = 0;
    for (int c = c0; c < c1; ++c) mine += s_cnt[c];
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int run = incl - mine;
    for (int w = 0; w < warp; ++w) run += s_warp[w];
    for (int c = c0; c < c1; ++c) {
        const int cnt = s_cnt[c];
        rec[b * C + c] = make_int2((int)(base + run), cnt);
        s_cnt[c] = run;
        run += cnt;
    }
    __syncthreads();
    for (int t = tid; t < n; t += kBinThreads) {
        const int lab = __ldg(p.labels + base + t);
        if (lab >= 0 && lab < C) {
            const ObjI o = bin_make<SRC>(p, base + t);
            const int pos = atomicAdd(&s_cnt[lab], 1);  // order inside a class is irrelevant: max-blend commutes bit for bit
            list[base + pos] = make_int4(o.x, o.y, o.r, __float_as_int(o.w));
        }
    }
}

__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// unit index -> plane and row range (same two-tier geometry as groups_kernel's unit_header)
__device__ __forceinline__ void unit_rows(const RenderParams& p, long long u, int G, long long& plane, int& uy0, int& uy1) {
    int ug, bpu;
    if (u < p.tier1_units) {
        plane = u / p.units_per_plane;
        ug = (int)(u - plane * p.units_per_plane);
        bpu = p.bands_per_unit;
    } else {
        const long long u2 = u - p.tier1_units;
        const long long pl2 = u2 / p.units_per_plane2;
        plane = p.tier1_planes + pl2;
        ug = (int)(u2 - pl2 * p.units_per_plane2);
        bpu = p.bands_per_unit2;
    }
    uy0 = ug * bpu * G;
    uy1 = min(uy0 + bpu * G, p.H);
}

template <bool ACCUM, bool DIGEST, int OUT = OUT_F32>
__global__ void __launch_bounds__(kGThreads, DH_G_MINB) groups2_kernel(const RenderParams p) {
    constexpr int ES = (OUT == OUT_F32) ? 4 : 2;
    constexpr int SRC = SRC_BINNED;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* const zero_tile = reinterpret_cast<float4*>(smem_raw);
    const int stage = p.stage_bytes;
    ObjI* const s_unit = reinterpret_cast<ObjI*>(smem_raw + kGZeroBytes + kGWarps * 2 * stage);
    __shared__ int s_cnt;
    __shared__ unsigned long long s_mask;
    __shared__ __align__(16) uint4 s_clc[2];               // cluster-launch-control responses (ring of two)
    __shared__ __align__(8) unsigned long long s_full[2];   // response j landed in slot j & 1
    __shared__ __align__(8) unsigned long long s_empty[2];  // all eight warps have read it

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool kneg = p.k < 0.0f;
    const int W = p.W, G = p.band_cols;
    const long long plane_elems = (long long)p.H * W;
    const void* const blist = p.bin_list;
    const int2* const brec = reinterpret_cast<const int2*>(p.bin_rec);
    for (int i = tid; i < kGZeroBytes / 16; i += kGThreads) zero_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid == 0) {
        mbar_init(&s_full[0], 1);
        mbar_init(&s_full[1], 1);
        mbar_init(&s_empty[0], kGWarps);
        mbar_init(&s_empty[1], kGWarps);
    }
    fence_proxy_async();  // zero tile and mbarriers -> visible to the async proxy
    __syncthreads();      // the only unconditional block barrier of the kernel
    WarpStage ws;
    ws.buf0 = reinterpret_cast<unsigned*>(smem_raw + kGZeroBytes + warp * 2 * stage);
    ws.buf1 = ws.buf0 + stage / 4;
    bool stage_clean = false;
    Digest dg;
    pdl_wait();  // the lists are bin_kernel's output; the heatmap may be the previous kernel's input
    pdl_launch_dependents();

    long long u = blockIdx.x;
    unsigned k = 0;  // sequence number of the unit in this CTA; response j (j >= 1) names unit j
    if (p.use_clc && tid == 0) {  // response 1 -> slot 1
        mbar_expect_tx(&s_full[1], 16);
        clc_try_cancel(&s_clc[1], &s_full[1]);
    }
    for (;;) {
        // ONE block barrier per unit keeps the eight warps on the same unit. Letting them drift (no barrier, the ring below
        // alone) was measured SLOWER: 231-237 against 212 us on the 8-GPU slice of cfg5 — with two units in flight per CTA
        // the set of addresses being written doubles (444 x 2 x 186 KB > the 126 MB L2) and the write rate falls, the same
        // effect the write-only microbenchmarks show for larger spans per CTA (profiles/r01/experiments.md section 1,
        // profiles/r02/call4, call6).
        __syncthreads();
        long long plane;
        int uy0, uy1;
        unit_rows(p, u, G, plane, uy0, uy1);
        const int2 rc = __ldg(brec + plane);
        UnitSrc us;
        us.plane = plane;
        us.src_base = rc.x;
        us.n = rc.y;
        char* const plane_b = reinterpret_cast<char*>(p.heat) + plane * plane_elems * ES;
        const int ngroups = (uy1 - uy0 + G - 1) / G;
        const long long plane_index0 = p.digest_base + plane * plane_elems;
        const unsigned long long valid = ngroups >= 64 ? ~0ull : ((1ull << ngroups) - 1ull);
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
        ObjI mine;
        mine.x = mine.y = 0;
        mine.r = -1;
        mine.w = 0.f;
        unsigned long long mask;
        int ucnt;
        bool small;
        if (us.n <= 32) {
            // ---- warp-autonomous: every warp derives the unit's state from the plane's entries by itself ----
            if (lane < us.n) mine = ld_bin(blist, us.src_base + lane);
            const bool hit = (mine.r >= 0) && (mine.y + mine.r >= uy0) && (mine.y - mine.r < uy1);
            const unsigned long long bits = hit ? group_bits(mine, uy0, uy1, G) : 0ull;
            mask = ((unsigned long long)__reduce_or_sync(0xffffffffu, (unsigned)(bits >> 32)) << 32) |
                   (unsigned long long)__reduce_or_sync(0xffffffffu, (unsigned)bits);
            ucnt = us.n;
            small = true;
        } else {
            // ---- block-cooperative list build (planes with many objects), from the compact list ----
            __syncthreads();  // every warp is done with the previous cooperative unit's list
            if (tid == 0) {
                s_cnt = 0;
                s_mask = 0ull;
            }
            __syncthreads();
            for (int base = 0; base < us.n; base += kGThreads) {
                const int t = base + tid;
                ObjI o;
                o.x = o.y = 0;
                o.r = -1;
                o.w = 0.f;
                if (t < us.n) o = ld_bin(blist, us.src_base + t);
                const bool pred = (o.r >= 0) && (o.y + o.r >= uy0) && (o.y - o.r < uy1);
                const unsigned long long bits = pred ? group_bits(o, uy0, uy1, G) : 0ull;
                const unsigned bal = __ballot_sync(0xffffffffu, pred);
                if (bal) {
                    const unsigned blo = __reduce_or_sync(0xffffffffu, (unsigned)bits);
                    const unsigned bhi = __reduce_or_sync(0xffffffffu, (unsigned)(bits >> 32));
                    int pos = 0;
                    if (lane == 0) {
                        atomicOr(&s_mask, ((unsigned long long)bhi << 32) | blo);
                        pos = atomicAdd(&s_cnt, __popc(bal));
                    }
                    pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(bal & ((1u << lane) - 1u));
                    if (pred && pos < kGCap) s_unit[pos] = o;
                }
            }
            __syncthreads();
            ucnt = s_cnt;
            mask = s_mask;
            small = ucnt <= 32;
            if (small && lane < ucnt) mine = s_unit[lane];
        }

        if constexpr (ACCUM && DH_ACCUM_PREFETCH != 0) {  // see groups_kernel: request every touched line up front
            const int np = small ? ucnt : 0;
            for (int e = warp; e < np; e += kGWarps) {
                const ObjI o = shfl_obj(mine, e);
                if (o.r < 0) continue;
                const int ry0 = max(o.y - o.r, uy0), ry1 = min(o.y + o.r, uy1 - 1);
                const int x0 = max(o.x - o.r, 0), x1 = min(o.x + o.r, W - 1);
                for (int y = ry0 + lane; y <= ry1; y += 32) {
                    const float* const rp = reinterpret_cast<const float*>(plane_b) + (long long)y * W;
                    for (int x = x0; x <= x1 + 31; x += 32) {
                        const float* const a = rp + (x < x1 ? x : x1);
                        if (DH_ACCUM_PREFETCH == 1) prefetch_l2(a);
                        else prefetch_l1(a);
                    }
                }
            }
        }
        if constexpr (!ACCUM) {
            // K4a: the untouched runs, by warp 7 (it gets the fewest striped groups in a short unit)
            if (warp == kGWarps - 1) zero_runs(~mask & valid, lane, plane_b, uy0, uy1, W, G, ES, zero_tile);
        }
        unsigned long long gm = mask & (kGStripe << warp);
        if (gm != 0ull && !stage_clean) {  // first touched group of this warp: establish the all-zero staging invariant
            uint4* const b4 = reinterpret_cast<uint4*>(ws.buf0);
            for (int i = lane; i < (2 * stage) / 16; i += 32) b4[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncwarp();
            stage_clean = true;
        }
        while (gm) {
            const int g = __ffsll((long long)gm) - 1;
            gm &= gm - 1ull;
            const int y0 = uy0 + g * G;
            render_group<ACCUM, SRC, DIGEST, OUT>(p, ws, plane_b, plane_index0, y0, min(G, uy1 - y0), W, G, kneg, p.k, lane, small,
                                                  mine, s_unit, ucnt, us, dg);
        }
#ifdef DH_TRACE_UNITS
        if (lane == 0 && g_trace && u < g_trace_units) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            atomicMax(&g_trace[u * 4 + 1], t);
            if (warp == 0) g_trace[u * 4 + 3] = (unsigned long long)ucnt | ((unsigned long long)__popcll(mask) << 32);
        }
#endif
        // ---- next unit ----
        if (!p.use_clc) {
            u += gridDim.x;
            if (u >= p.total_units) break;
            ++k;
            continue;
        }
        const unsigned j = k + 1u, slot = j & 1u;  // response j; its (j-1)/2-th use of the slot
        mbar_wait(&s_full[slot], ((j - 1u) >> 1) & 1u);
        unsigned nx;
        const bool more = clc_response(&s_clc[slot], nx);
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[slot]);  // this warp has read response j
        if (!more) break;  // (no further request may follow a failed one)
        if (tid == 0) {    // request j + 1 -> the other slot, once all warps have read response j - 1 out of it
            const unsigned j2 = j + 1u, s2 = j2 & 1u;
            if (j2 >= 3u) mbar_wait(&s_empty[s2], ((j2 - 3u) >> 1) & 1u);
            fence_proxy_async();
            mbar_expect_tx(&s_full[s2], 16);
            clc_try_cancel(&s_clc[s2], &s_full[s2]);
        }
        u = nx;
        ++k;
    }
    if constexpr (!ACCUM) bulk_wait_read<0>();  // shared memory must outlive every bulk read issued by this thread
    if constexpr (DIGEST) digest_flush(dg, p.digest, tid, lane, warp);
}

}  // namespace dh
