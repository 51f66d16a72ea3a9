// distribution of one hop (store issue -> poll exit on another SM), W independent chains
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define SENT 0x7ff8dead0000beefULL
__device__ __forceinline__ uint64_t ld_u64(const void* p) { uint64_t v; asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_u64(void* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ uint64_t gtime() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
template <int K>
__global__ void k_chain(uint64_t* word, uint32_t n, uint32_t width, uint64_t* t_store, uint64_t* t_det, uint64_t* t_start) {
    if (threadIdx.x & 31) return;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n; i += T) {
        double x = 1.0;
        t_start[i] = gtime();
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
        t_det[i] = gtime();
        st_u64(word + i, (uint64_t)__double_as_longlong(x * 1e-3));
        t_store[i] = gtime();
    }
}
int main() {
    uint64_t *word, *ts, *td, *tb; uint32_t N = 1000000;
    CK(cudaMalloc(&word, N * 8)); CK(cudaMalloc(&ts, N * 8)); CK(cudaMalloc(&td, N * 8)); CK(cudaMalloc(&tb, N * 8));
    for (uint32_t width : {1u, 1400u}) {
        uint32_t n = width == 1 ? 50000 : N;
        std::vector<uint64_t> hs(n, SENT); CK(cudaMemcpy(word, hs.data(), n * 8, cudaMemcpyHostToDevice));
        void* args[] = {&word, &n, &width, &ts, &td, &tb};
        CK(cudaLaunchCooperativeKernel((void*)k_chain<1>, dim3(3 * 148), dim3(256), args, 0, 0));
        CK(cudaDeviceSynchronize());
        std::vector<uint64_t> hts(n), htd(n), htb(n);
        CK(cudaMemcpy(hts.data(), ts, n * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(htd.data(), td, n * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(htb.data(), tb, n * 8, cudaMemcpyDeviceToHost));
        std::vector<long long> lag; long long late = 0;
        for (uint32_t i = width + 1 + n / 10; i < n; ++i) {
            if (htb[i] > hts[i - width]) { late++; continue; }     // worker arrived after the store: not a clean hop
            lag.push_back((long long)htd[i] - (long long)hts[i - width]);
        }
        std::sort(lag.begin(), lag.end());
        size_t m = lag.size();
        printf("width=%u: clean hops %zu (late %lld): p1 %lld p10 %lld p50 %lld p90 %lld p99 %lld p99.9 %lld max %lld ns; span %.1f us\n", width, m, late,
               lag[m / 100], lag[m / 10], lag[m / 2], lag[m * 9 / 10], lag[m * 99 / 100], lag[m * 999 / 1000], lag[m - 1], (hts[n - 1] - htb[0]) / 1e3);
    }
    return 0;
}
