// hop latency under background gather traffic (cp.async random 8B gathers + plain stores per node)
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define SENT 0x7ff8dead0000beefULL
__device__ __forceinline__ uint64_t ld_u64(const void* p) { uint64_t v; asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_u64(void* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void cpa8(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(s)), "l"(g) : "memory"); }
// W independent chains interleaved: node i depends on node i-W (K words: i-W, i-W-1.. older)
template <int K>
__global__ void k_chain(uint64_t* word, uint32_t n, uint32_t width, const double* bg, uint32_t bgmask, int R, int S, double* sink) {
    __shared__ double buf[8][64];
    if (threadIdx.x & 31) return;
    uint32_t wi = threadIdx.x >> 5;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    uint32_t rng = warp * 2654435761u + 12345u;
    for (uint32_t i = warp; i < n; i += T) {
        for (int r = 0; r < R; ++r) { rng = rng * 1664525u + 1013904223u; cpa8(&buf[wi][r & 63], bg + ((rng >> 4) & bgmask)); }
        asm volatile("cp.async.commit_group;" ::: "memory");
        for (int r = 0; r < S; ++r) { rng = rng * 1664525u + 1013904223u; sink[(rng >> 4) & bgmask] = 1.0; }
        double x = 1.0;
        if (i >= width + K) {
            uint64_t u[K];
            for (;;) {
                bool ready = true;
#pragma unroll
                for (int k = 0; k < K; ++k) { u[k] = ld_u64(word + (i - width - k)); if (u[k] == SENT) ready = false; }
                if (ready) break;
            }
#pragma unroll
            for (int k = 0; k < K; ++k) x += __longlong_as_double((long long)u[k]);
        }
        st_u64(word + i, (uint64_t)__double_as_longlong(x * 1e-3));
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
}
int main() {
    uint64_t* word; CK(cudaMalloc(&word, 1 << 24));
    double *bg, *sink; size_t BG = 1u << 25; CK(cudaMalloc(&bg, BG * 8)); CK(cudaMemset(bg, 0, BG * 8)); CK(cudaMalloc(&sink, BG * 8));
    uint32_t bgmask = BG - 1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (uint32_t width : {1u, 1400u}) for (int R : {0, 8, 16, 32}) for (int S : {0, 6}) {
        uint32_t n = width == 1 ? 50000 : 1000000;
        std::vector<uint64_t> hs(n, SENT); CK(cudaMemcpy(word, hs.data(), n * 8, cudaMemcpyHostToDevice));
        void* args[] = {&word, &n, &width, &bg, &bgmask, &R, &S, &sink};
        cudaEventRecord(e0);
        CK(cudaLaunchCooperativeKernel((void*)k_chain<4>, dim3(3 * 148), dim3(256), args, 0, 0));
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("width=%u R=%d S=%d: %.3f ms total, %.1f ns per level\n", width, R, S, ms, ms * 1e6 / (n / width));
    }
    return 0;
}
