// per-level time of W coupled chains (node i waits for i-W-k, k<K) under different polling strategies
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define SENT 0x7ff8dead0000beefULL
__device__ __forceinline__ uint64_t ld_u64(const void* p) { uint64_t v; asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_u64(void* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
template <int K, int MODE>
__global__ void k_chain(uint64_t* word, uint32_t n, uint32_t width, uint32_t gap) {
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
                unsigned pend = 0;
#pragma unroll
                for (int k = 0; k < K; ++k) { u[k] = ld_u64(word + (i - width - k)); if (u[k] == SENT) pend |= 1u << k; }
                while (pend) {
                    int k = __ffs(pend) - 1;
                    const uint64_t* p = word + (i - width - k);
                    uint64_t v;
                    if (MODE == 1) { while ((v = ld_u64(p)) == SENT) {} }
                    else {
                        // 4 loads in flight, spaced by `gap` cycles; then self-clocked
                        uint64_t q0, q1, q2, q3; long long t0 = clock64();
                        q0 = ld_u64(p); while (clock64() - t0 < gap) {}
                        q1 = ld_u64(p); while (clock64() - t0 < 2 * gap) {}
                        q2 = ld_u64(p); while (clock64() - t0 < 3 * gap) {}
                        q3 = ld_u64(p);
                        for (;;) {
                            if (q0 != SENT) { v = q0; break; } q0 = ld_u64(p);
                            if (q1 != SENT) { v = q1; break; } q1 = ld_u64(p);
                            if (q2 != SENT) { v = q2; break; } q2 = ld_u64(p);
                            if (q3 != SENT) { v = q3; break; } q3 = ld_u64(p);
                        }
                    }
#pragma unroll
                    for (int kk = 0; kk < K; ++kk) if (kk == k) u[kk] = v;
                    pend &= pend - 1;
                }
            }
#pragma unroll
            for (int k = 0; k < K; ++k) x += __longlong_as_double((long long)u[k]);
        }
        st_u64(word + i, (uint64_t)__double_as_longlong(x * 1e-3));
    }
}
uint64_t* word; cudaEvent_t e0, e1;
template <int K, int MODE> void run(uint32_t width, uint32_t gap, int bps) {
    uint32_t n = width == 1 ? 50000 : 1000000;
    std::vector<uint64_t> hs(n, SENT); CK(cudaMemcpy(word, hs.data(), n * 8, cudaMemcpyHostToDevice));
    void* args[] = {&word, &n, &width, &gap};
    cudaEventRecord(e0);
    CK(cudaLaunchCooperativeKernel((void*)k_chain<K, MODE>, dim3(bps * 148), dim3(256), args, 0, 0));
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("K=%d mode=%d width=%u gap=%u bps=%d: %.3f ms, %.1f ns per level\n", K, MODE, width, gap, bps, ms, ms * 1e6 / (n / width));
}
int main() {
    CK(cudaMalloc(&word, 1 << 24)); cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (uint32_t width : {1u, 200u, 1400u}) {
        run<1, 0>(width, 0, 3); run<2, 0>(width, 0, 3); run<4, 0>(width, 0, 3);
        run<4, 1>(width, 0, 3); run<4, 2>(width, 100, 3); run<4, 2>(width, 200, 3);
        run<1, 2>(width, 100, 3);
    }
    run<4, 0>(1400, 0, 1); run<4, 0>(1400, 0, 2); run<4, 0>(1400, 0, 6); run<4, 1>(1400, 0, 6); run<4, 2>(1400, 100, 6);
    return 0;
}
