// fill_bw3.cu — third round: WHY does a flat (non-persistent) store kernel reach 7.4 TB/s when persistent loops
// stall near 6.0-6.3?  Probes: in-order dynamic distribution, throttled bulk stores, plane-per-CTA shapes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fill_bw3 fill_bw3.cu && ./fill_bw3 [bytes]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_store(void* g, const void* s, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ unsigned long long g_counter;

// A: persistent, chunks of `chunk4` float4 handed out in address order by an atomic counter
__global__ void __launch_bounds__(256) st_dyn_kernel(float4* dst, long long n4, int chunk4) {
    __shared__ long long s_c;
    const long long nch = (n4 + chunk4 - 1) / chunk4;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_c = (long long)atomicAdd(&g_counter, 1ull);
        __syncthreads();
        const long long c = s_c;
        if (c >= nch) break;
        float4* p = dst + c * chunk4;
        const int n = (int)min((long long)chunk4, n4 - c * chunk4);
        for (int i = threadIdx.x; i < n; i += 256) p[i] = z;
    }
}
// C: non-persistent, one CTA per `chunk4` float4 (chunk4 = 256: the flat kernel; 23200: a 232x400 plane)
__global__ void __launch_bounds__(256) st_cta_kernel(float4* dst, long long n4, int chunk4) {
    const long long c = blockIdx.x;
    float4* p = dst + c * chunk4;
    const int n = (int)min((long long)chunk4, n4 - c * chunk4);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int i = threadIdx.x; i < n; i += 256) p[i] = z;
}

constexpr int kZero = 16384;
// D/E/F: span-per-CTA with bulk stores of `chunk` bytes from a zero tile, issued by the 8 warp leaders.
// DYN: spans fetched from the atomic counter; else grid-stride (persistent) — or one span per CTA if grid == nspans.
// THR: 0 no throttle; N>0: each leader keeps <= N bulk groups incomplete (cp.async.bulk.wait_group N, full completion)
template <bool DYN, int THR>
__global__ void __launch_bounds__(256) span_bulk_kernel(char* dst, long long nspans, int span_bytes, int chunk) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ long long s_c;
    float4* zero = reinterpret_cast<float4*>(smem);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kZero / 16; i += 256) zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    fence_proxy_async();
    __syncthreads();
    const int per = (span_bytes + chunk - 1) / chunk;
    long long sidx = blockIdx.x;
    for (;;) {
        if constexpr (DYN) {
            __syncthreads();
            if (tid == 0) s_c = (long long)atomicAdd(&g_counter, 1ull);
            __syncthreads();
            sidx = s_c;
        }
        if (sidx >= nspans) break;
        char* base = dst + sidx * (long long)span_bytes;
        if (lane == 0) {
            for (int c = warp; c < per; c += 8) {
                const int off = c * chunk, nb = min(chunk, span_bytes - off);
                bulk_store(base + off, zero, nb);
                bulk_commit();
                if constexpr (THR > 0) bulk_wait_all<THR>();
            }
        }
        if constexpr (!DYN) sidx += gridDim.x;
    }
    bulk_wait_read<0>();
}

struct Arg { char* dst; size_t bytes; int grid, chunk, span; };
static void reset_counter() { unsigned long long z = 0; CK(cudaMemcpyToSymbolAsync(g_counter, &z, 8, 0, cudaMemcpyHostToDevice)); }
static void run_dyn(void* a_) { Arg* a = (Arg*)a_; reset_counter(); st_dyn_kernel<<<a->grid, 256>>>((float4*)a->dst, a->bytes / 16, a->chunk / 16); }
static void run_cta(void* a_) { Arg* a = (Arg*)a_; const long long n4 = a->bytes / 16, c4 = a->chunk / 16; st_cta_kernel<<<(unsigned)((n4 + c4 - 1) / c4), 256>>>((float4*)a->dst, n4, (int)c4); }
template <bool DYN, int THR> static void run_span(void* a_) {
    Arg* a = (Arg*)a_;
    if (DYN) reset_counter();
    CK(cudaFuncSetAttribute(span_bulk_kernel<DYN, THR>, cudaFuncAttributeMaxDynamicSharedMemorySize, kZero));
    const long long nspans = a->bytes / a->span;
    const int grid = a->grid > 0 ? a->grid : (int)nspans;
    span_bulk_kernel<DYN, THR><<<grid, 256, kZero>>>(a->dst, nspans, a->span, a->chunk);
}

static float time_it(void (*fn)(void*), void* arg, int reps) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i) fn(arg);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; ++i) fn(arg);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    return ms / reps;
}

int main(int argc, char** argv) {
    size_t bytes = argc > 1 ? strtoull(argv[1], 0, 10) : 1425408000ull;  // 3840 planes of 232x400 fp32
    char* dst; CK(cudaMalloc(&dst, bytes));
    const int reps = 20;
    auto report = [&](const char* name, float ms) { printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms, bytes / ms / 1e6); fflush(stdout); };
    char n[96];
    Arg a{dst, bytes, 0, 4096, 371200};
    for (int chunk : {4096, 16384, 65536}) for (int k : {4, 8}) {
        a.grid = 148 * k; a.chunk = chunk; snprintf(n, 96, "A st dyn (atomic, in order) c=%d k=%d", chunk, k); report(n, time_it(run_dyn, &a, reps));
    }
    for (int chunk : {4096, 16384, 65536, 371200}) { a.chunk = chunk; snprintf(n, 96, "C st one CTA per chunk c=%d", chunk); report(n, time_it(run_cta, &a, reps)); }
    for (int chunk : {1600, 12800}) {
        a.chunk = chunk; a.span = 371200;
        a.grid = 0;       snprintf(n, 96, "D bulk plane/CTA flat grid c=%d", chunk); report(n, time_it(run_span<false, 0>, &a, reps));
        a.grid = 148 * 4; snprintf(n, 96, "D bulk plane/CTA persistent k=4 c=%d", chunk); report(n, time_it(run_span<false, 0>, &a, reps));
        a.grid = 148 * 4; snprintf(n, 96, "E bulk plane/CTA dyn k=4 c=%d", chunk); report(n, time_it(run_span<true, 0>, &a, reps));
        a.grid = 148 * 1; snprintf(n, 96, "E bulk plane/CTA dyn k=1 c=%d", chunk); report(n, time_it(run_span<true, 0>, &a, reps));
        a.grid = 148 * 4; snprintf(n, 96, "F bulk plane/CTA persistent k=4 thr=1 c=%d", chunk); report(n, time_it(run_span<false, 1>, &a, reps));
        a.grid = 148 * 4; snprintf(n, 96, "F bulk plane/CTA persistent k=4 thr=2 c=%d", chunk); report(n, time_it(run_span<false, 2>, &a, reps));
        a.grid = 148 * 4; snprintf(n, 96, "F bulk plane/CTA persistent k=4 thr=4 c=%d", chunk); report(n, time_it(run_span<false, 4>, &a, reps));
        a.grid = 148 * 4; snprintf(n, 96, "F bulk plane/CTA dyn k=4 thr=2 c=%d", chunk); report(n, time_it(run_span<true, 2>, &a, reps));
        // band-sized spans (8 rows) handed out in address order: the "compact wavefront" renderer shape
        a.span = 12800; a.grid = 148 * 4; snprintf(n, 96, "E bulk band(12.8K)/CTA dyn k=4 c=%d", chunk); report(n, time_it(run_span<true, 0>, &a, reps));
        a.span = 12800; a.grid = 0;       snprintf(n, 96, "D bulk band(12.8K)/CTA flat grid c=%d", chunk); report(n, time_it(run_span<false, 0>, &a, reps));
        a.span = 92800; a.grid = 148 * 4; snprintf(n, 96, "E bulk quarter-plane/CTA dyn k=4 c=%d", chunk); report(n, time_it(run_span<true, 0>, &a, reps));
        a.span = 92800; a.grid = 0;       snprintf(n, 96, "D bulk quarter-plane/CTA flat grid c=%d", chunk); report(n, time_it(run_span<false, 0>, &a, reps));
    }
    return 0;
}
