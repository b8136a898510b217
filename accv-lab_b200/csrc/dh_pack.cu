// dh_pack.cu — list-of-tensors -> RaggedBatch packing, optionally fused with K1.
//
// Replaces the step BEFORE the renderer in the reference's callers: a Python loop of one slice-assign kernel per
// sample and per array into a torch.zeros tensor
//   /root/reference/packages/batching_helpers/accvlab/batching_helpers/batched_processing_py.py:411-423 (combine_data)
//   /root/reference/packages/draw_heatmap/examples/draw_heatmap_batched.py:42-50
// by ONE launch per array: a device table of B source pointers + B sizes, every destination element written exactly
// once (payload or the zero padding: the torch.zeros pass is fused in), coalesced.
// dh_pack_radius_center additionally applies K1 (dh_k1.cuh) on the way, so per-sample float centre / box lists become
// the int32 centres + radii RaggedBatch tensors draw_heatmap_batched takes, in one launch.
#include <cuda_runtime.h>
#include <stdint.h>

#include "dh_b200.h"
#include "dh_k1.cuh"

namespace dh {

extern void count_launches(uint64_t n);  // dh_kernels.cu

__device__ __forceinline__ long long clamp_size(const int64_t* __restrict__ sizes, long long b, long long T) {
    long long n = sizes[b];
    return n < 0 ? 0 : (n > T ? T : n);
}

// dst [B, T*R] elements of type E; sample b: the first min(sizes[b],T)*R elements from src[b], zeros after.
// grid.x covers one sample's T*R elements, grid.y strides over samples.
template <typename E>
__global__ void __launch_bounds__(256) pack_kernel(E* __restrict__ dst, const void* const* __restrict__ src,
                                                   const int64_t* __restrict__ sizes, long long B, long long T,
                                                   long long R) {
    const long long per = T * R;
    for (long long b = blockIdx.y; b < B; b += gridDim.y) {
        const long long n = clamp_size(sizes, b, T) * R;
        const E* __restrict__ s = n > 0 ? static_cast<const E*>(src[b]) : nullptr;
        E* __restrict__ d = dst + b * per;
        for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < per; j += (long long)gridDim.x * blockDim.x)
            d[j] = j < n ? s[j] : E{};
    }
}

struct alignas(16) B16 {
    unsigned x = 0, y = 0, z = 0, w = 0;
};

// K1 fused: per-sample float centres [n,2] and boxes [n,4] -> int centres [B,T,2], radii [B,T]; padding = 0.
__global__ void __launch_bounds__(256) pack_k1_kernel(const float* const* __restrict__ csrc, const float* const* __restrict__ bsrc,
                                                      const int64_t* __restrict__ sizes, float osf, long long B, long long T,
                                                      int32_t* __restrict__ centers_out, int32_t* __restrict__ radii_out) {
    for (long long b = blockIdx.y; b < B; b += gridDim.y) {
        const long long n = clamp_size(sizes, b, T);
        const float2* __restrict__ c = n > 0 ? reinterpret_cast<const float2*>(csrc[b]) : nullptr;
        const float4* __restrict__ bx = n > 0 ? reinterpret_cast<const float4*>(bsrc[b]) : nullptr;
        int2* __restrict__ co = reinterpret_cast<int2*>(centers_out) + b * T;
        int32_t* __restrict__ ro = radii_out + b * T;
        for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < T; t += (long long)gridDim.x * blockDim.x) {
            int xi = 0, yi = 0, r = 0;
            if (t < n) {
                const float2 cc = c[t];
                const float4 bb = bx[t];
                radius_center_one(cc.x, cc.y, bb.x, bb.y, bb.z, bb.w, osf, xi, yi, r);
            }
            co[t] = make_int2(xi, yi);
            ro[t] = r;
        }
    }
}

static dim3 pack_grid(long long per, long long B) {
    long long gx = (per + 255) / 256;
    if (gx < 1) gx = 1;
    if (gx > 4096) gx = 4096;
    return dim3((unsigned)gx, (unsigned)(B < 65535 ? B : 65535), 1);
}

}  // namespace dh

extern "C" {

int dh_pack_ragged(void* stream, void* dst, const void* const* src_ptrs, const int64_t* sizes, int64_t B, int64_t T,
                   int64_t row_elems, int elem_bytes) {
    if (B < 0 || T < 0 || row_elems < 0) return DH_ERR_INVALID_ARGUMENT;
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8 && elem_bytes != 16)
        return DH_ERR_INVALID_ARGUMENT;
    if (B == 0 || T == 0 || row_elems == 0) return DH_OK;
    if (!dst || !src_ptrs || !sizes) return DH_ERR_INVALID_ARGUMENT;
    if (reinterpret_cast<uintptr_t>(dst) % (uintptr_t)elem_bytes) return DH_ERR_INVALID_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return DH_ERR_NO_DEVICE;
    const dim3 g = dh::pack_grid(T * row_elems, B);
    cudaStream_t s = (cudaStream_t)stream;
    switch (elem_bytes) {
        case 1: dh::pack_kernel<uint8_t><<<g, 256, 0, s>>>((uint8_t*)dst, src_ptrs, sizes, B, T, row_elems); break;
        case 2: dh::pack_kernel<uint16_t><<<g, 256, 0, s>>>((uint16_t*)dst, src_ptrs, sizes, B, T, row_elems); break;
        case 4: dh::pack_kernel<uint32_t><<<g, 256, 0, s>>>((uint32_t*)dst, src_ptrs, sizes, B, T, row_elems); break;
        case 8: dh::pack_kernel<uint64_t><<<g, 256, 0, s>>>((uint64_t*)dst, src_ptrs, sizes, B, T, row_elems); break;
        default: dh::pack_kernel<dh::B16><<<g, 256, 0, s>>>((dh::B16*)dst, src_ptrs, sizes, B, T, row_elems); break;
    }
    dh::count_launches(1);
    return (int)cudaGetLastError();
}

int dh_pack_radius_center(void* stream, const float* const* centers_ptrs, const float* const* bboxes_ptrs,
                          const int64_t* sizes, float out_size_factor, int64_t B, int64_t T, int32_t* centers_out,
                          int32_t* radii_out) {
    if (B < 0 || T < 0) return DH_ERR_INVALID_ARGUMENT;
    if (B == 0 || T == 0) return DH_OK;
    if (!centers_ptrs || !bboxes_ptrs || !sizes || !centers_out || !radii_out) return DH_ERR_INVALID_ARGUMENT;
    if (reinterpret_cast<uintptr_t>(centers_out) & 7u) return DH_ERR_INVALID_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return DH_ERR_NO_DEVICE;
    dh::pack_k1_kernel<<<dh::pack_grid(T, B), 256, 0, (cudaStream_t)stream>>>(centers_ptrs, bboxes_ptrs, sizes,
                                                                             out_size_factor, B, T, centers_out, radii_out);
    dh::count_launches(1);
    return (int)cudaGetLastError();
}

}  // extern "C"

// ================================================================================================================
// N4: the CenterNet / CenterPoint `gaussian_radius(det_size, min_overlap)` as a K1 mode.
// NOT in the reference (its helper uses the min border distance, dh_k1.cuh) -> PARITY UNPINNED: the
// arithmetic restates the published algorithm (CornerNet `gaussian_radius`, kept by CenterNet utils/image.py and
// CenterPoint det3d core/utils/center_utils.py; mmdet models/utils/gaussian_target.py is the variant with the
// corrected roots), evaluated like upstream in float64 (Python floats), one IEEE operation per upstream operation
// (no FMA contraction: explicit __d*_rn), then radius = max(min_radius, (int)r) as in CenterPoint's target assigner.
namespace dh {

__device__ __forceinline__ double gr_root(double a, double b, double c, bool plus) {
    const double disc = __dsub_rn(__dmul_rn(b, b), __dmul_rn(__dmul_rn(4.0, a), c));
    const double sq = __dsqrt_rn(disc);
    return __ddiv_rn(plus ? __dadd_rn(b, sq) : __dsub_rn(b, sq), __dmul_rn(2.0, a));
}

__global__ void __launch_bounds__(256) gaussian_radius_kernel(const float* __restrict__ det_hw, double mo, int variant,
                                                              int min_radius, long long n, int32_t* __restrict__ out,
                                                              float* __restrict__ out_f) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double h = (double)det_hw[2 * i], w = (double)det_hw[2 * i + 1];
        const double hw = __dadd_rn(h, w), wh = __dmul_rn(w, h);
        const double om = __dsub_rn(1.0, mo);
        double r1, r2, r3;
        if (variant == DH_RADIUS_CORNERNET) {  // a = 1 / 4 / 4*mo, roots (b + sqrt) / 2 (no division by a)
            const double c1 = __ddiv_rn(__dmul_rn(wh, om), __dadd_rn(1.0, mo));
            r1 = __ddiv_rn(__dadd_rn(hw, __dsqrt_rn(__dsub_rn(__dmul_rn(hw, hw), __dmul_rn(4.0, c1)))), 2.0);
            const double b2 = __dmul_rn(2.0, hw), c2 = __dmul_rn(om, wh);
            r2 = __ddiv_rn(__dadd_rn(b2, __dsqrt_rn(__dsub_rn(__dmul_rn(b2, b2), __dmul_rn(16.0, c2)))), 2.0);
            const double a3 = __dmul_rn(4.0, mo), b3 = __dmul_rn(__dmul_rn(-2.0, mo), hw), c3 = __dmul_rn(__dsub_rn(mo, 1.0), wh);
            r3 = __ddiv_rn(__dadd_rn(b3, __dsqrt_rn(__dsub_rn(__dmul_rn(b3, b3), __dmul_rn(__dmul_rn(4.0, a3), c3)))), 2.0);
        } else {  // mmdet: (b - sqrt) / (2a), (b - sqrt) / (2a), (b + sqrt) / (2a)
            r1 = gr_root(1.0, hw, __ddiv_rn(__dmul_rn(wh, om), __dadd_rn(1.0, mo)), false);
            r2 = gr_root(4.0, __dmul_rn(2.0, hw), __dmul_rn(om, wh), false);
            r3 = gr_root(__dmul_rn(4.0, mo), __dmul_rn(__dmul_rn(-2.0, mo), hw), __dmul_rn(__dsub_rn(mo, 1.0), wh), true);
        }
        const double r = fmin(fmin(r1, r2), r3);
        if (out_f) out_f[i] = (float)r;
        int ri = (r >= 2147483647.0) ? 2147483647 : (r > -2147483648.0 ? (int)r : (-2147483647 - 1));  // int(r); NaN -> INT_MIN
        out[i] = ri < min_radius ? min_radius : ri;
    }
}

}  // namespace dh

extern "C" int dh_gaussian_radius(void* stream, const float* det_hw, double min_overlap, int variant, int32_t min_radius,
                                  int64_t n, int32_t* radii_out, float* radii_float_out) {
    if (n < 0 || (variant != DH_RADIUS_CORNERNET && variant != DH_RADIUS_MMDET)) return DH_ERR_INVALID_ARGUMENT;
    if (n == 0) return DH_OK;
    if (!det_hw || !radii_out) return DH_ERR_INVALID_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return DH_ERR_NO_DEVICE;
    long long blocks = (n + 255) / 256;
    if (blocks > 1184) blocks = 1184;
    dh::gaussian_radius_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(det_hw, min_overlap, variant, min_radius, n,
                                                                             radii_out, radii_float_out);
    dh::count_launches(1);
    return (int)cudaGetLastError();
}
