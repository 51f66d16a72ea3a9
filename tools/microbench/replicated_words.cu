// replicated words: producer stores C copies (different lines), consumer keeps C staggered polls in flight
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define SENT 0x7ff8dead0000beefULL
__device__ __forceinline__ uint64_t ld_u64(const void* p) { uint64_t v; asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_u64(void* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
// MODE 0: classic (poll copy 0 of all K words per iteration).  MODE 1: C staggered polls in flight on C copies, all K words per poll
template <int K, int C, int MODE>
__global__ void k_chain(uint64_t* word, uint32_t n, uint32_t width, uint32_t gap, size_t cstride) {
    if (threadIdx.x & 31) return;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n; i += T) {
        double x = 1.0;
        if (i >= width + K) {
            uint64_t u[K];
            if (MODE == 0) {
                for (;;) {
                    bool ready = true;
#pragma unroll
                    for (int k = 0; k < K; ++k) { u[k] = ld_u64(word + (i - width - k)); if (u[k] == SENT) ready = false; }
                    if (ready) break;
                }
            } else {
                uint64_t q[C][K];
                long long next = clock64();
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    while (clock64() < next) {}
                    next += gap;
#pragma unroll
                    for (int k = 0; k < K; ++k) q[c][k] = ld_u64(word + c * cstride + (i - width - k));
                }
                bool done = false;
                while (!done) {
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        if (done) break;
                        bool ready = true;
#pragma unroll
                        for (int k = 0; k < K; ++k) if (q[c][k] == SENT) ready = false;
                        if (ready) {
#pragma unroll
                            for (int k = 0; k < K; ++k) u[k] = q[c][k];
                            done = true;
                        } else {
                            while (clock64() < next) {}
                            next = clock64() + gap;
#pragma unroll
                            for (int k = 0; k < K; ++k) q[c][k] = ld_u64(word + c * cstride + (i - width - k));
                        }
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < K; ++k) x += __longlong_as_double((long long)u[k]);
        }
        uint64_t r = (uint64_t)__double_as_longlong(x * 1e-3);
#pragma unroll
        for (int c = 0; c < C; ++c) st_u64(word + c * cstride + i, r);
    }
}
uint64_t* word; cudaEvent_t e0, e1;
template <int K, int C, int MODE> void run(uint32_t width, uint32_t gap) {
    uint32_t n = width == 1 ? 50000 : 1000000;
    size_t cstride = (1u << 20) + 64;
    std::vector<uint64_t> hs(cstride * 4, SENT); CK(cudaMemcpy(word, hs.data(), hs.size() * 8, cudaMemcpyHostToDevice));
    void* args[] = {&word, &n, &width, &gap, &cstride};
    cudaEventRecord(e0);
    CK(cudaLaunchCooperativeKernel((void*)k_chain<K, C, MODE>, dim3(3 * 148), dim3(256), args, 0, 0));
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("K=%d C=%d mode=%d width=%u gap=%u: %.3f ms, %.1f ns per level\n", K, C, MODE, width, gap, ms, ms * 1e6 / (n / width));
}
int main() {
    CK(cudaMalloc(&word, ((1u << 20) + 64) * 4 * 8)); cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (uint32_t width : {1u, 1400u}) {
        run<1, 1, 0>(width, 0); run<1, 2, 1>(width, 300); run<1, 4, 1>(width, 150); run<1, 4, 1>(width, 80); run<1, 1, 1>(width, 0);
        run<4, 1, 0>(width, 0); run<4, 2, 1>(width, 300); run<4, 4, 1>(width, 150); run<4, 4, 1>(width, 80);
    }
    return 0;
}
