// fill_bw2.cu — second round of write-only probes: cache hints, 256-bit stores, launch shapes (measurement tool).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fill_bw2 fill_bw2.cu && ./fill_bw2 [bytes] [only-memset]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ unsigned long long policy_evict_first() {
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ unsigned long long policy_evict_last() {
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}

// HINT: 0 plain, 1 .cs, 2 .wt, 3 L2::cache_hint evict_first, 4 v8 (256-bit), 5 v8 + evict_first, 6 .cg, 7 L1::no_allocate
template <int HINT>
__device__ __forceinline__ void store16(float4* p, unsigned long long pol) {
    if constexpr (HINT == 0) asm volatile("st.global.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(0.f) : "memory");
    if constexpr (HINT == 1) asm volatile("st.global.cs.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(0.f) : "memory");
    if constexpr (HINT == 2) asm volatile("st.global.wt.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(0.f) : "memory");
    if constexpr (HINT == 3) asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%1,%1,%1}, %2;" ::"l"(p), "f"(0.f), "l"(pol) : "memory");
    if constexpr (HINT == 6) asm volatile("st.global.cg.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(0.f) : "memory");
    if constexpr (HINT == 7) asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(0.f) : "memory");
}
template <int HINT>
__device__ __forceinline__ void store32(float4* p, unsigned long long pol) {
    if constexpr (HINT == 4) asm volatile("st.global.v8.f32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(p), "f"(0.f) : "memory");
    if constexpr (HINT == 5) asm volatile("st.global.L2::cache_hint.v8.f32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1}, %2;" ::"l"(p), "f"(0.f), "l"(pol) : "memory");
}

template <int HINT>
__global__ void __launch_bounds__(256) st_kernel(float4* dst, long long n4) {
    const unsigned long long pol = policy_evict_first();
    if constexpr (HINT == 4 || HINT == 5) {
        const long long n8 = n4 / 2, stride = (long long)gridDim.x * 256;
#pragma unroll 4
        for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n8; i += stride) store32<HINT>(dst + 2 * i, pol);
    } else {
        const long long stride = (long long)gridDim.x * 256;
#pragma unroll 4
        for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n4; i += stride) store16<HINT>(dst + i, pol);
    }
}

// one float4 per thread, no loop (huge grid)
__global__ void __launch_bounds__(256) st_flat_kernel(float4* dst, long long n4) {
    const long long i = blockIdx.x * 256ll + threadIdx.x;
    if (i < n4) dst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}
// each CTA owns ONE long contiguous span (n4 / grid)
__global__ void __launch_bounds__(256) st_span_kernel(float4* dst, long long n4) {
    const long long per = (n4 + gridDim.x - 1) / gridDim.x, b = blockIdx.x * per, e = min(b + per, n4);
#pragma unroll 4
    for (long long i = b + threadIdx.x; i < e; i += 256) dst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

constexpr int kZero = 16384;
// POL: 0 none, 1 evict_first, 2 evict_last
template <int POL>
__global__ void __launch_bounds__(256) bulk_kernel(char* dst, long long nchunks, int chunk) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* zero = reinterpret_cast<float4*>(smem);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kZero / 16; i += 256) zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    fence_proxy_async();
    __syncthreads();
    const unsigned long long pol = POL == 2 ? policy_evict_last() : policy_evict_first();
    const long long gw = (long long)blockIdx.x * 8 + warp, nw = (long long)gridDim.x * 8;
    if (lane == 0) {
        for (long long c = gw; c < nchunks; c += nw) {
            char* g = dst + c * (long long)chunk;
            if constexpr (POL == 0)
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(zero)), "r"(chunk) : "memory");
            else
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(g), "r"(smem_u32(zero)), "r"(chunk), "l"(pol) : "memory");
            bulk_commit();
        }
    }
    bulk_wait_read<0>();
}

struct Arg { char* dst; size_t bytes; int grid, chunk; };
template <int H> static void run_st(void* a_) { Arg* a = (Arg*)a_; st_kernel<H><<<a->grid, 256>>>((float4*)a->dst, a->bytes / 16); }
static void run_flat(void* a_) { Arg* a = (Arg*)a_; st_flat_kernel<<<(unsigned)((a->bytes / 16 + 255) / 256), 256>>>((float4*)a->dst, a->bytes / 16); }
static void run_span(void* a_) { Arg* a = (Arg*)a_; st_span_kernel<<<a->grid, 256>>>((float4*)a->dst, a->bytes / 16); }
template <int P> static void run_bulk(void* a_) {
    Arg* a = (Arg*)a_;
    CK(cudaFuncSetAttribute(bulk_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, kZero));
    bulk_kernel<P><<<a->grid, 256, kZero>>>(a->dst, (long long)(a->bytes / a->chunk), a->chunk);
}
static void run_memset(void* a_) { Arg* a = (Arg*)a_; CK(cudaMemsetAsync(a->dst, 0, a->bytes)); }
static void run_memset32(void* a_) { Arg* a = (Arg*)a_; CK(cudaMemsetAsync(a->dst, 0x3f, a->bytes)); }

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
    size_t bytes = argc > 1 ? strtoull(argv[1], 0, 10) : 1425408000ull;
    char* dst; CK(cudaMalloc(&dst, bytes));
    const int reps = bytes > (4ull << 30) ? 5 : 20;
    auto report = [&](const char* name, float ms) { printf("%-34s %8.3f ms  %8.1f GB/s\n", name, ms, bytes / ms / 1e6); fflush(stdout); };
    Arg a{dst, bytes, 148 * 8, 1600};
    report("memset 0", time_it(run_memset, &a, reps));
    report("memset 0x3f", time_it(run_memset32, &a, reps));
    if (argc > 2) return 0;
    const char* names[8] = {"plain", ".cs", ".wt", "L2 evict_first", "v8", "v8 evict_first", ".cg", "L1 no_alloc"};
    void (*fns[8])(void*) = {run_st<0>, run_st<1>, run_st<2>, run_st<3>, run_st<4>, run_st<5>, run_st<6>, run_st<7>};
    for (int h = 0; h < 8; ++h) for (int k : {4, 8, 16, 32}) {
        a.grid = 148 * k; char n[64]; snprintf(n, 64, "st %s k=%d", names[h], k); report(n, time_it(fns[h], &a, reps));
    }
    report("st flat (1 float4/thread)", time_it(run_flat, &a, reps));
    for (int k : {1, 4, 8}) { a.grid = 148 * k; char n[64]; snprintf(n, 64, "st span k=%d", k); report(n, time_it(run_span, &a, reps)); }
    for (int chunk : {1600, 12800}) for (int k : {1, 4}) {
        a.grid = 148 * k; a.chunk = chunk; char n[64];
        snprintf(n, 64, "bulk c=%d k=%d", chunk, k); report(n, time_it(run_bulk<0>, &a, reps));
        snprintf(n, 64, "bulk evict_first c=%d k=%d", chunk, k); report(n, time_it(run_bulk<1>, &a, reps));
        snprintf(n, 64, "bulk evict_last c=%d k=%d", chunk, k); report(n, time_it(run_bulk<2>, &a, reps));
    }
    return 0;
}
