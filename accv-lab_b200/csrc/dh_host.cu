// dh_host.cu — result checksum kernel and the host-buffer entry of the C ABI (include/dh_b200.h).
//
// Neither has a counterpart in the reference (its API is device-tensor in / device-tensor out,
// /root/reference/packages/draw_heatmap/accvlab/draw_heatmap/csrc/draw_heatmap_cuda.cu:91-165, and it has no
// multi-GPU path). They exist for (a) end-to-end measurement from host buffers, (b) non-torch callers,
// (c) verifying that a batch sharded over G GPUs equals the unsharded result without gathering it.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "dh_b200.h"

namespace dh {

extern void count_launches(uint64_t n);  // dh_kernels.cu

// == dh_groups_kernel.cuh mix_index, oracle/oracle.py:_mix_index
__device__ __forceinline__ unsigned long long mix_index(unsigned long long i) {
    const unsigned lo = (unsigned)i, hi = (unsigned)(i >> 32);
    unsigned h = lo ^ (hi * 0x9E3779B1u);
    h *= 0x85EBCA6Bu;
    h ^= h >> 13;
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return ((unsigned long long)h << 32) | (unsigned long long)(h * 0x27D4EB2Fu + lo);
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// Grid-stride over float4 groups (scalar head/tail handled by the same loop at element granularity when the
// pointer is not 16-byte aligned). Wrapping uint64 sums: exactly associative, so the digest is independent of the
// grid shape and of how a batch is split over GPUs.
__global__ void __launch_bounds__(256) checksum_kernel(const float* __restrict__ v, long long n, long long index_base,
                                                       unsigned long long* __restrict__ out3, int vec_ok) {
    unsigned long long s_bits = 0, s_hash = 0, s_cnt = 0;
    const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    auto one = [&](float x, long long i) {
        s_bits += (unsigned long long)__float_as_uint(x);
        if (x > 0.0f) {
            s_hash += mix_index((unsigned long long)(index_base + i));
            s_cnt += 1;
        }
    };
    if (vec_ok) {
        const long long n4 = n >> 2;
        const float4* v4 = reinterpret_cast<const float4*>(v);
        for (long long g = tid; g < n4; g += nthreads) {
            const float4 q = __ldcs(v4 + g);
            one(q.x, 4 * g);
            one(q.y, 4 * g + 1);
            one(q.z, 4 * g + 2);
            one(q.w, 4 * g + 3);
        }
        for (long long i = (n4 << 2) + tid; i < n; i += nthreads) one(v[i], i);
    } else {
        for (long long i = tid; i < n; i += nthreads) one(v[i], i);
    }
    __shared__ unsigned long long sm[3][8];
    s_bits = warp_sum(s_bits);
    s_hash = warp_sum(s_hash);
    s_cnt = warp_sum(s_cnt);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        sm[0][warp] = s_bits;
        sm[1][warp] = s_hash;
        sm[2][warp] = s_cnt;
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        unsigned long long t = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += sm[threadIdx.x][w];
        atomicAdd(&out3[threadIdx.x], t);
    }
}

// ---- helper streams of the host entry (per device, created once) ------------------------------------
// growth factor of the input chunks, smallest input size that is pipelined (KiB); environment-tunable in measurement
// builds only (-DDH_TUNING)
#ifdef DH_TUNING
static int host_env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
#define DH_HOST_KNOB(name, dflt) host_env_int(name, dflt)
#else
#define DH_HOST_KNOB(name, dflt) (dflt)
#endif
static const int g_host_growth = DH_HOST_KNOB("DH_HOST_GROWTH", 4);
static const int g_host_pipeline_kib = DH_HOST_KNOB("DH_HOST_PIPELINE_KIB", 1536);
constexpr int kHostMaxChunks = 16;  // input chunks of the pipelined device-output path
struct HostCtx {
    bool ready = false;
    cudaStream_t s[2] = {nullptr, nullptr};
    cudaStream_t copy = nullptr;  // H2D of input chunk c+1 runs here while chunk c is rendered
    cudaEvent_t fork = nullptr, join[2] = {nullptr, nullptr};
    cudaEvent_t arrived[kHostMaxChunks] = {};
};
// One context (helper streams + events) and one lock PER DEVICE: calls on different GPUs of one process never serialise
// on each other; two calls on the same GPU share its helper streams and are serialised while they enqueue.
static HostCtx g_ctx[64];
static std::mutex g_ctx_mu[64];

static int host_device(int& dev) {
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    return (dev < 0 || dev >= 64) ? DH_ERR_NO_DEVICE : DH_OK;
}

// caller holds g_ctx_mu[dev]
static int host_ctx(int dev, HostCtx*& out) {
    cudaError_t e;
    HostCtx& c = g_ctx[dev];
    if (!c.ready) {
        for (int i = 0; i < 2; ++i) {
            if ((e = cudaStreamCreateWithFlags(&c.s[i], cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
            if ((e = cudaEventCreateWithFlags(&c.join[i], cudaEventDisableTiming)) != cudaSuccess) return (int)e;
        }
        if ((e = cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming)) != cudaSuccess) return (int)e;
        if ((e = cudaStreamCreateWithFlags(&c.copy, cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
        for (int i = 0; i < kHostMaxChunks; ++i)
            if ((e = cudaEventCreateWithFlags(&c.arrived[i], cudaEventDisableTiming)) != cudaSuccess) return (int)e;
        c.ready = true;
    }
    out = &c;
    return DH_OK;
}

// Small inputs in PINNED host memory: one kernel reads the four arrays straight over PCIe (UVA zero-copy) into the
// device scratch — one launch instead of four cudaMemcpyAsync, whose fixed costs (~8 us each for a few hundred KB)
// add up to more than the transfer itself. Pageable memory (not device-accessible) keeps the memcpy path.
struct GatherArgs {
    const void* src[4];
    void* dst[4];
    unsigned long long bytes[4];  // multiples of 4
};
__global__ void __launch_bounds__(256) gather_host_kernel(const GatherArgs a) {
    const long long tid = blockIdx.x * 256ll + threadIdx.x, nthr = (long long)gridDim.x * 256;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const char* const s = (const char*)a.src[k];
        char* const d = (char*)a.dst[k];
        const unsigned long long nb = a.bytes[k];
        if (nb == 0) continue;
        if ((((uintptr_t)s | (uintptr_t)d) & 15u) == 0) {
            const long long n16 = (long long)(nb >> 4);
            for (long long i = tid; i < n16; i += nthr) ((int4*)d)[i] = ((const int4*)s)[i];
            for (long long i = (n16 << 2) + tid; i < (long long)(nb >> 2); i += nthr) ((int*)d)[i] = ((const int*)s)[i];
        } else {
            for (long long i = tid; i < (long long)(nb >> 2); i += nthr) ((int*)d)[i] = ((const int*)s)[i];
        }
    }
}
static const int g_host_gather = DH_HOST_KNOB("DH_HOST_GATHER", 1);
// device-visible alias of a pinned host pointer, or nullptr when the memory is pageable / unknown to the driver
static const void* pinned_alias(const void* h) {
    if (!h) return nullptr;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, h) != cudaSuccess) {
        cudaGetLastError();  // clear
        return nullptr;
    }
    return (at.type == cudaMemoryTypeHost) ? at.devicePointer : nullptr;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct HostLayout {
    size_t off_centers, off_bboxes, off_sizes, off_labels, off_cksum, off_stage[2], total;
    int64_t chunk;
};

static HostLayout host_layout(int64_t B, int64_t T, int64_t C, int64_t H, int64_t W, bool has_labels, int64_t chunk,
                              bool staging) {
    HostLayout L{};
    const int64_t planes_per_sample = has_labels ? C : 1;
    const size_t sample_bytes = (size_t)planes_per_sample * (size_t)H * (size_t)W * 4;
    if (chunk <= 0) {  // default
        chunk = sample_bytes ? (int64_t)((64ull << 20) / sample_bytes) : B;
        if (chunk < 1) chunk = 1;
    }
    if (chunk > B) chunk = B > 0 ? B : 1;
    L.chunk = chunk;
    size_t o = 0;
    L.off_centers = o; o += align256((size_t)B * T * 2 * 4);
    L.off_bboxes = o;  o += align256((size_t)B * T * 4 * 4);
    L.off_sizes = o;   o += align256((size_t)B * 8);
    L.off_labels = o;  o += has_labels ? align256((size_t)B * T * 4) : 0;
    L.off_cksum = o;   o += 256;
    for (int i = 0; i < 2; ++i) {
        L.off_stage[i] = o;
        if (staging) o += align256((size_t)chunk * sample_bytes);
    }
    L.total = o;
    return L;
}

}  // namespace dh

extern "C" {

int dh_checksum(void* stream, const float* values, int64_t n, int64_t index_base, uint64_t* out3) {
    if (n < 0 || !out3) return DH_ERR_INVALID_ARGUMENT;
    if (n == 0) return DH_OK;
    if (!values) return DH_ERR_INVALID_ARGUMENT;
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return (int)e;
    long long blocks = (n / 4 + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > (long long)sms * 8) blocks = (long long)sms * 8;
    const int vec_ok = (reinterpret_cast<uintptr_t>(values) & 15u) == 0;
    dh::checksum_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(values, n, index_base,
                                                                       (unsigned long long*)out3, vec_ok);
    dh::count_launches(1);
    return (int)cudaGetLastError();
}

size_t dh_host_scratch_bytes(int64_t B, int64_t T, int64_t C, int64_t H, int64_t W, int has_labels,
                             int64_t chunk_samples, int staging_buffers) {
    if (B < 0 || T < 0 || C < 0 || H < 0 || W < 0) return 0;
    return dh::host_layout(B, T, C, H, W, has_labels != 0, chunk_samples, staging_buffers != 0).total;
}

int dh_draw_heatmap_batched_from_boxes_host(void* stream_, float* heatmap_host, float* heatmap_device,
                                            uint64_t* checksum_host, const float* centers_xy_host,
                                            const float* bboxes_xyxy_host, float osf, const int64_t* nums_targets_host,
                                            const int32_t* labels_host, int64_t B, int64_t T, int64_t C, int64_t H,
                                            int64_t W, float factor, float k, void* device_scratch,
                                            size_t device_scratch_bytes, int64_t chunk_samples) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (B < 0 || T < 0 || H < 1 || W < 1) return DH_ERR_INVALID_ARGUMENT;
    const bool has_labels = labels_host != nullptr;
    if (!has_labels) C = 1;
    if (C < 0) return DH_ERR_INVALID_ARGUMENT;
    if (!heatmap_host && !heatmap_device) return DH_ERR_INVALID_ARGUMENT;
    if (B == 0 || C == 0) {
        if (checksum_host) checksum_host[0] = checksum_host[1] = checksum_host[2] = 0;
        return DH_OK;
    }
    if (!nums_targets_host || (T > 0 && (!centers_xy_host || !bboxes_xyxy_host))) return DH_ERR_INVALID_ARGUMENT;
    const bool staging = heatmap_host && !heatmap_device;
    const dh::HostLayout L = dh::host_layout(B, T, C, H, W, has_labels, chunk_samples, staging);
    if (!device_scratch || device_scratch_bytes < L.total) return DH_ERR_WORKSPACE_TOO_SMALL;
    if (reinterpret_cast<uintptr_t>(device_scratch) & 15u) return DH_ERR_INVALID_ARGUMENT;

    char* ws = (char*)device_scratch;
    float* d_centers = (float*)(ws + L.off_centers);
    float* d_bboxes = (float*)(ws + L.off_bboxes);
    int64_t* d_sizes = (int64_t*)(ws + L.off_sizes);
    int32_t* d_labels = has_labels ? (int32_t*)(ws + L.off_labels) : nullptr;
    uint64_t* d_ck = (uint64_t*)(ws + L.off_cksum);

    cudaError_t e;
#define DH_TRY(x) do { if ((e = (x)) != cudaSuccess) return (int)e; } while (0)
    const int64_t sample_elems = C * H * W;
    // samples [s0, s0+nb) of the device-resident inputs -> `dst` on stream `cs` (digest fused when the shape allows)
    auto render_range = [&](cudaStream_t cs, float* dst, int64_t s0, int64_t nb) -> int {
        int r = DH_ERR_UNSUPPORTED_SHAPE;
        const int32_t* lab = has_labels ? d_labels + s0 * T : nullptr;
        if (checksum_host)  // digest fused into the render launch: the heatmap is never re-read
            r = dh_draw_heatmap_batched_from_boxes_digest(cs, dst, d_centers + s0 * T * 2, d_bboxes + s0 * T * 4, osf,
                                                          d_sizes + s0, DH_SIZES_I64, lab, nb, T, C, H, W, factor, k,
                                                          s0 * sample_elems, d_ck, nullptr);
        if (r == DH_ERR_UNSUPPORTED_SHAPE) {
            r = dh_draw_heatmap_batched_from_boxes(cs, dst, d_centers + s0 * T * 2, d_bboxes + s0 * T * 4, osf,
                                                   d_sizes + s0, DH_SIZES_I64, lab, nb, T, C, H, W, factor, k,
                                                   DH_MODE_FRESH, nullptr, nullptr, nullptr);
            if (r == DH_OK && checksum_host) r = dh_checksum(cs, dst, nb * sample_elems, s0 * sample_elems, d_ck);
        }
        return r;
    };
    int st;

    // ---- result stays on the device and the inputs are large: pipeline H2D against rendering ----------------
    // Input chunk c+1 crosses PCIe on the copy stream while chunk c is rendered; consecutive chunks alternate
    // between two render streams so the tail of one launch overlaps the head of the next.
    const size_t slot_bytes = 24 + (has_labels ? 4 : 0);
    const size_t in_bytes = (size_t)B * T * slot_bytes;
    if (!heatmap_host && in_bytes >= ((size_t)dh::g_host_pipeline_kib << 10) && B >= 2 && chunk_samples >= 0) {
        // Chunk boundaries. Copying a sample's inputs takes ~1/10 of rendering it, so only the FIRST chunk's copy is
        // exposed: chunks grow geometrically (x4: still hidden at a quarter of the usual PCIe rate) from a first one of
        // ~100 KB of inputs — cfg5: 36 + 144 + 578 + 2314 samples, 4 launches instead of 8 equal ones (each launch
        // costs ~10 us of ramp and tail that the next one only partly hides). A caller-given chunk_samples keeps
        // equal chunks.
        int64_t bounds[dh::kHostMaxChunks + 1];
        int64_t nch;
        if (chunk_samples > 0) {
            nch = (B + chunk_samples - 1) / chunk_samples;
            if (nch > dh::kHostMaxChunks) nch = dh::kHostMaxChunks;
            if (nch > B) nch = B;
            if (nch < 2) nch = 2;
            for (int64_t c = 0; c <= nch; ++c) bounds[c] = B * c / nch;
        } else {
            const int64_t gr = dh::g_host_growth < 2 ? 2 : dh::g_host_growth;
            nch = 2;
            int64_t parts = 1 + gr;  // 1 + gr + ... + gr^(nch-1)
            while (nch < 6 && in_bytes / (size_t)parts > ((size_t)128 << 10)) {
                parts = parts * gr + 1;
                ++nch;
            }
            int64_t acc = 0, w = 1;
            bounds[0] = 0;
            for (int64_t c = 1; c <= nch; ++c, w *= gr) {
                acc += w;
                bounds[c] = (int64_t)((__int128)B * acc / parts);
            }
            if (bounds[1] < 1) bounds[1] = 1;  // B >= 2: the first chunk is never empty
        }
        int dev = 0;
        if ((st = dh::host_device(dev)) != DH_OK) return st;
        std::lock_guard<std::mutex> lock(dh::g_ctx_mu[dev]);
        dh::HostCtx* ctx = nullptr;
        st = dh::host_ctx(dev, ctx);
        if (st != DH_OK) return st;
        DH_TRY(cudaMemcpyAsync(d_sizes, nums_targets_host, (size_t)B * 8, cudaMemcpyHostToDevice, stream));
        if (checksum_host) DH_TRY(cudaMemsetAsync(d_ck, 0, 24, stream));
        DH_TRY(cudaEventRecord(ctx->fork, stream));  // everything that read the scratch inputs before is done
        DH_TRY(cudaStreamWaitEvent(ctx->copy, ctx->fork, 0));
        DH_TRY(cudaStreamWaitEvent(ctx->s[0], ctx->fork, 0));
        DH_TRY(cudaStreamWaitEvent(ctx->s[1], ctx->fork, 0));
        // From here on work may be queued on the helper streams: an error must not return before they are joined to
        // the caller's stream again (a later synchronisation of `stream` has to cover everything this call enqueued).
        st = DH_OK;
#define DH_TRY_J(x) do { if (st == DH_OK && (e = (x)) != cudaSuccess) st = (int)e; } while (0)
        for (int64_t c = 0; c < nch && st == DH_OK; ++c) {
            const int64_t s0 = bounds[c], s1 = bounds[c + 1] > s0 ? bounds[c + 1] : s0, nb = s1 - s0;
            if (nb <= 0) continue;
            DH_TRY_J(cudaMemcpyAsync(d_centers + s0 * T * 2, centers_xy_host + s0 * T * 2, (size_t)nb * T * 8,
                                     cudaMemcpyHostToDevice, ctx->copy));
            DH_TRY_J(cudaMemcpyAsync(d_bboxes + s0 * T * 4, bboxes_xyxy_host + s0 * T * 4, (size_t)nb * T * 16,
                                     cudaMemcpyHostToDevice, ctx->copy));
            if (has_labels)
                DH_TRY_J(cudaMemcpyAsync(d_labels + s0 * T, labels_host + s0 * T, (size_t)nb * T * 4,
                                         cudaMemcpyHostToDevice, ctx->copy));
            DH_TRY_J(cudaEventRecord(ctx->arrived[c], ctx->copy));
            cudaStream_t rs = ctx->s[c & 1];
            DH_TRY_J(cudaStreamWaitEvent(rs, ctx->arrived[c], 0));
            if (st == DH_OK) st = render_range(rs, heatmap_device + s0 * sample_elems, s0, nb);
        }
        for (int j = 0; j < 2; ++j) {  // always: join what was queued
            if (cudaEventRecord(ctx->join[j], ctx->s[j]) == cudaSuccess) cudaStreamWaitEvent(stream, ctx->join[j], 0);
        }
        if (cudaEventRecord(ctx->join[0], ctx->copy) == cudaSuccess) cudaStreamWaitEvent(stream, ctx->join[0], 0);
        if (st != DH_OK) return st;
        if (checksum_host) DH_TRY(cudaMemcpyAsync(checksum_host, d_ck, 24, cudaMemcpyDeviceToHost, stream));
        return DH_OK;
    }

    // ---- inputs: host -> device. Pinned and small: one zero-copy gather launch; else one copy per array ----
    bool gathered = false;
    if (dh::g_host_gather && in_bytes <= ((size_t)8 << 20)) {
        dh::GatherArgs ga{};
        ga.src[0] = T > 0 ? dh::pinned_alias(centers_xy_host) : nullptr;
        ga.src[1] = T > 0 ? dh::pinned_alias(bboxes_xyxy_host) : nullptr;
        ga.src[2] = (T > 0 && has_labels) ? dh::pinned_alias(labels_host) : nullptr;
        ga.src[3] = dh::pinned_alias(nums_targets_host);
        ga.dst[0] = d_centers; ga.dst[1] = d_bboxes; ga.dst[2] = d_labels; ga.dst[3] = d_sizes;
        ga.bytes[0] = T > 0 ? (unsigned long long)B * T * 8 : 0;
        ga.bytes[1] = T > 0 ? (unsigned long long)B * T * 16 : 0;
        ga.bytes[2] = (T > 0 && has_labels) ? (unsigned long long)B * T * 4 : 0;
        ga.bytes[3] = (unsigned long long)B * 8;
        bool ok = true;
        for (int k2 = 0; k2 < 4; ++k2) ok = ok && (ga.bytes[k2] == 0 || ga.src[k2] != nullptr);
        if (ok) {
            const unsigned long long units = (ga.bytes[0] + ga.bytes[1] + ga.bytes[2] + ga.bytes[3]) >> 4;
            long long blocks = (long long)((units + 255) / 256);
            if (blocks < 1) blocks = 1;
            if (blocks > 1184) blocks = 1184;
            dh::gather_host_kernel<<<(int)blocks, 256, 0, stream>>>(ga);
            DH_TRY(cudaGetLastError());
            dh::count_launches(1);
            gathered = true;
        }
    }
    if (!gathered) {
        if (T > 0) {
            DH_TRY(cudaMemcpyAsync(d_centers, centers_xy_host, (size_t)B * T * 2 * 4, cudaMemcpyHostToDevice, stream));
            DH_TRY(cudaMemcpyAsync(d_bboxes, bboxes_xyxy_host, (size_t)B * T * 4 * 4, cudaMemcpyHostToDevice, stream));
            if (has_labels)
                DH_TRY(cudaMemcpyAsync(d_labels, labels_host, (size_t)B * T * 4, cudaMemcpyHostToDevice, stream));
        }
        DH_TRY(cudaMemcpyAsync(d_sizes, nums_targets_host, (size_t)B * 8, cudaMemcpyHostToDevice, stream));
    }
    if (checksum_host) DH_TRY(cudaMemsetAsync(d_ck, 0, 24, stream));

    if (!heatmap_host) {  // result stays on the device, small inputs: one launch
        st = render_range(stream, heatmap_device, 0, B);
        if (st != DH_OK) return st;
        if (checksum_host) DH_TRY(cudaMemcpyAsync(checksum_host, d_ck, 24, cudaMemcpyDeviceToHost, stream));
        return DH_OK;
    }

    // ---- chunked render + copy back, double-buffered over two helper streams ----
    int dev = 0;
    if ((st = dh::host_device(dev)) != DH_OK) return st;
    std::lock_guard<std::mutex> lock(dh::g_ctx_mu[dev]);
    dh::HostCtx* ctx = nullptr;
    st = dh::host_ctx(dev, ctx);
    if (st != DH_OK) return st;
    DH_TRY(cudaEventRecord(ctx->fork, stream));
    DH_TRY(cudaStreamWaitEvent(ctx->s[0], ctx->fork, 0));
    DH_TRY(cudaStreamWaitEvent(ctx->s[1], ctx->fork, 0));
    int64_t i = 0;
    st = DH_OK;
    for (int64_t s0 = 0; s0 < B && st == DH_OK; s0 += L.chunk, ++i) {
        const int64_t nb = (B - s0 < L.chunk) ? (B - s0) : L.chunk;
        cudaStream_t cs = ctx->s[i & 1];
        float* dst = staging ? (float*)(ws + L.off_stage[i & 1]) : heatmap_device + s0 * sample_elems;
        st = render_range(cs, dst, s0, nb);
        DH_TRY_J(cudaMemcpyAsync(heatmap_host + s0 * sample_elems, dst, (size_t)nb * sample_elems * 4,
                                 cudaMemcpyDeviceToHost, cs));
    }
    for (int j = 0; j < 2; ++j) {  // always: join what was queued, also after an error
        if (cudaEventRecord(ctx->join[j], ctx->s[j]) == cudaSuccess) cudaStreamWaitEvent(stream, ctx->join[j], 0);
    }
    if (st != DH_OK) return st;
#undef DH_TRY_J
    if (checksum_host) DH_TRY(cudaMemcpyAsync(checksum_host, d_ck, 24, cudaMemcpyDeviceToHost, stream));
#undef DH_TRY
    return DH_OK;
}

}  // extern "C"
