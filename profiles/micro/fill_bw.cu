// fill_bw.cu — write-only bandwidth microbenchmarks behind the renderer's design choices (measurement tool, not
// product code):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fill_bw fill_bw.cu && ./fill_bw [bytes]
//   memset            cudaMemsetAsync
//   stv4(k)           st.global.v4 grid-stride, 148*k CTAs x 256 threads
//   bulk_zero(c,k)    cp.async.bulk shared->global of c bytes from a constant zero tile, one elected lane per warp
//   bulk_staged(c,k)  per warp: zero-fill a private c-byte staging buffer (STS.128), fence.proxy.async, bulk store;
//                     two buffers alternate (wait_group.read 1) — the touched-row path of wrows_kernel without rendering
//   bulk_mix(c,k,p)   p% of the chunks take the staged path, the rest the zero-tile path
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
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__global__ void __launch_bounds__(256) stv4_kernel(float4* dst, long long n4) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long stride = (long long)gridDim.x * 256;
#pragma unroll 4
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n4; i += stride) dst[i] = z;
}

// chunk-contiguous variant: a CTA owns a contiguous 16 KB chunk at a time
__global__ void __launch_bounds__(256) stv4_chunk_kernel(float4* dst, long long n4) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long nch = (n4 + 1023) / 1024;
    for (long long c = blockIdx.x; c < nch; c += gridDim.x) {
        float4* p = dst + c * 1024;
        const int n = (int)min(1024ll, n4 - c * 1024);
#pragma unroll
        for (int i = threadIdx.x; i < n; i += 256) p[i] = z;
    }
}

constexpr int kZero = 32768;
// mode 0: all from the zero tile; mode 1: all staged; mode 2: staged iff hash(chunk) % 100 < pct
__global__ void __launch_bounds__(256) bulk_kernel(char* dst, long long nchunks, int chunk, int mode, int pct) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* zero = reinterpret_cast<float4*>(smem);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kZero / 16; i += 256) zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    fence_proxy_async();
    __syncthreads();
    unsigned char* stage = smem + kZero + warp * 2 * chunk;
    int cur = 0;
    const long long gw = (long long)blockIdx.x * 8 + warp, nw = (long long)gridDim.x * 8;
    for (long long c = gw; c < nchunks; c += nw) {
        bool staged = mode == 1;
        if (mode == 2) staged = (int)((unsigned)(c * 2654435761u) >> 8) % 100 < pct;
        char* g = dst + c * (long long)chunk;
        if (!staged) {
            if (lane == 0) { bulk_store(g, zero, chunk); bulk_commit(); }
        } else {
            unsigned char* buf = stage + cur * chunk;
            if (lane == 0) bulk_wait_read<1>();
            __syncwarp();
            uint4* b4 = reinterpret_cast<uint4*>(buf);
            for (int i = lane; i < chunk / 16; i += 32) b4[i] = make_uint4(0u, 0u, 0u, 0u);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) { bulk_store(g, buf, chunk); bulk_commit(); }
            cur ^= 1;
        }
    }
    bulk_wait_read<0>();
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

struct Arg { char* dst; size_t bytes; int k, chunk, mode, pct; };
static void run_memset(void* a_) { Arg* a = (Arg*)a_; CK(cudaMemsetAsync(a->dst, 0, a->bytes)); }
static void run_stv4(void* a_) { Arg* a = (Arg*)a_; stv4_kernel<<<148 * a->k, 256>>>((float4*)a->dst, a->bytes / 16); }
static void run_stv4c(void* a_) { Arg* a = (Arg*)a_; stv4_chunk_kernel<<<148 * a->k, 256>>>((float4*)a->dst, a->bytes / 16); }
static void run_bulk(void* a_) {
    Arg* a = (Arg*)a_;
    const size_t sm = kZero + 16 * (size_t)a->chunk;
    CK(cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    bulk_kernel<<<148 * a->k, 256, sm>>>(a->dst, (long long)(a->bytes / a->chunk), a->chunk, a->mode, a->pct);
}

int main(int argc, char** argv) {
    size_t bytes = argc > 1 ? strtoull(argv[1], 0, 10) : 1425408000ull;  // cfg5 / 8 GPUs
    char* dst; CK(cudaMalloc(&dst, bytes));
    const int reps = bytes > (4ull << 30) ? 5 : 20;
    auto report = [&](const char* name, float ms) { printf("%-34s %8.3f ms  %8.1f GB/s\n", name, ms, bytes / ms / 1e6); fflush(stdout); };
    Arg a{dst, bytes, 4, 1600, 0, 0};
    report("memset", time_it(run_memset, &a, reps));
    for (int k : {2, 4, 8, 16}) { a.k = k; char n[64]; snprintf(n, 64, "stv4 k=%d", k); report(n, time_it(run_stv4, &a, reps)); }
    for (int k : {4, 8}) { a.k = k; char n[64]; snprintf(n, 64, "stv4_chunk16K k=%d", k); report(n, time_it(run_stv4c, &a, reps)); }
    for (int chunk : {1600, 3200, 6400, 12800, 16384, 32768}) for (int k : {1, 2, 4}) {
        if (kZero + 16 * chunk > 227 * 1024 / k) continue;
        a = Arg{dst, bytes, k, chunk, 0, 0}; char n[64]; snprintf(n, 64, "bulk_zero c=%d k=%d", chunk, k); report(n, time_it(run_bulk, &a, reps));
    }
    for (int chunk : {1600, 3200, 6400}) for (int k : {2, 4}) {
        if (kZero + 16 * chunk > 227 * 1024 / k) continue;
        a = Arg{dst, bytes, k, chunk, 1, 0}; char n[64]; snprintf(n, 64, "bulk_staged c=%d k=%d", chunk, k); report(n, time_it(run_bulk, &a, reps));
        for (int pct : {25, 50}) { a.mode = 2; a.pct = pct; snprintf(n, 64, "bulk_mix c=%d k=%d staged=%d%%", chunk, k, pct); report(n, time_it(run_bulk, &a, reps)); }
    }
    return 0;
}
