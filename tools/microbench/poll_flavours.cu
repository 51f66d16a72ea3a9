// hop latency vs polling instruction flavour and number of words polled per iteration
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define SENT 0x7ff8dead0000beefULL
template <int LD> __device__ __forceinline__ uint64_t ldx(const void* p) {
    uint64_t v;
    if (LD == 0) asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else if (LD == 1) asm volatile("ld.volatile.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else if (LD == 2) asm volatile("ld.global.cg.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else if (LD == 3) asm volatile("ld.acquire.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else asm volatile("ld.global.cv.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
template <int ST> __device__ __forceinline__ void stx(void* p, uint64_t v) {
    if (ST == 0) asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else if (ST == 1) asm volatile("st.volatile.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else if (ST == 2) asm volatile("st.global.cg.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else asm volatile("st.global.wt.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// node i waits for words i-1, i-1-s, i-1-2s ... (K words, all long since published except i-1)
template <int LD, int ST, int K>
__global__ void k_chain(uint64_t* word, uint32_t n, uint32_t stride) {
    if (threadIdx.x & 31) return;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n; i += T) {
        double x = 1.0;
        if (i >= 1 + (K - 1) * stride) {
            uint64_t u[K];
            for (;;) {
                bool ready = true;
#pragma unroll
                for (int k = 0; k < K; ++k) { u[k] = ldx<LD>(word + (i - 1 - k * stride)); if (u[k] == SENT) ready = false; }
                if (ready) break;
            }
#pragma unroll
            for (int k = 0; k < K; ++k) x += __longlong_as_double((long long)u[k]);
        }
        stx<ST>(word + i, (uint64_t)__double_as_longlong(x * 1e-3));
    }
}
uint64_t* word; cudaEvent_t e0, e1;
template <int LD, int ST, int K> void run(const char* name, int bps, uint32_t stride) {
    uint32_t n = 100000;
    std::vector<uint64_t> hs(n, SENT); CK(cudaMemcpy(word, hs.data(), n * 8, cudaMemcpyHostToDevice));
    void* args[] = {&word, &n, &stride};
    cudaEventRecord(e0);
    CK(cudaLaunchCooperativeKernel((void*)k_chain<LD, ST, K>, dim3(bps * 148), dim3(256), args, 0, 0));
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%s K=%d bps=%d stride=%u: %.1f ns per hop\n", name, K, bps, stride, ms * 1e6 / n);
}
int main() {
    CK(cudaMalloc(&word, 1 << 23)); cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int bps : {3}) {
        run<0, 0, 1>("ld.relaxed/st.relaxed", bps, 17); run<0, 0, 2>("ld.relaxed/st.relaxed", bps, 17); run<0, 0, 4>("ld.relaxed/st.relaxed", bps, 17); run<0, 0, 8>("ld.relaxed/st.relaxed", bps, 17);
        run<1, 1, 1>("ld.volatile/st.volatile", bps, 17); run<1, 1, 4>("ld.volatile/st.volatile", bps, 17); run<1, 1, 8>("ld.volatile/st.volatile", bps, 17);
        run<2, 2, 1>("ld.cg/st.cg", bps, 17); run<2, 2, 4>("ld.cg/st.cg", bps, 17); run<2, 2, 8>("ld.cg/st.cg", bps, 17);
        run<3, 0, 1>("ld.acquire/st.relaxed", bps, 17); run<3, 0, 4>("ld.acquire/st.relaxed", bps, 17);
        run<4, 3, 1>("ld.cv/st.wt", bps, 17); run<4, 3, 4>("ld.cv/st.wt", bps, 17); run<4, 3, 8>("ld.cv/st.wt", bps, 17);
        run<0, 2, 4>("ld.relaxed/st.cg", bps, 17); run<2, 0, 4>("ld.cg/st.relaxed", bps, 17);
    }
    return 0;
}
