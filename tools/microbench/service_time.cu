// microbenchmarks: hop latency of relaxed ld/st chains, and per-node service time with/without prefetch
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define SENT 0x7ff8dead0000beefULL
__device__ __forceinline__ uint64_t ld_u64(const void* p) { uint64_t v; asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_u64(void* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void touch(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void touch2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// 1. chain: node i waits for word[i - dist] (dist=1: pure chain), handled by worker i % T
__global__ void k_chain(uint64_t* word, uint32_t n, uint32_t dist, uint32_t work) {
    if (threadIdx.x & 31) return;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n; i += T) {
        double x = 1.0;
        if (i >= dist) { uint64_t u; while ((u = ld_u64(word + (i - dist))) == SENT) {} x = __longlong_as_double((long long)u); }
        for (uint32_t k = 0; k < work; ++k) x = x / 1.0000001 + 1e-9;   // dependent f64 divisions
        st_u64(word + i, (uint64_t)__double_as_longlong(x + 1.0));
    }
}
// 2. service: independent nodes; static loads a0..a7[v] then b[c[v]] ; mode 0 none, 1 PF1, 2 PF2(L2), 3 register double buffer of first level
struct Arr { const double* a[8]; const uint32_t* c; const double* b; };
__global__ void k_service(Arr A, const uint32_t* __restrict__ order, uint32_t n, int mode, double* out) {
    if (threadIdx.x & 31) return;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    double acc = 0.0;
    uint32_t pos = warp;
    if (pos >= n) return;
    uint32_t v0 = order[pos], v1 = pos + T < n ? order[pos + T] : 0xffffffffu, v2 = pos + 2 * T < n ? order[pos + 2 * T] : 0xffffffffu;
    for (;;) {
        uint32_t p3 = pos + 3 * T; uint32_t v3 = p3 < n ? order[p3] : 0xffffffffu;
        if (mode == 1) {
            if (v2 != 0xffffffffu) { for (int k = 0; k < 8; ++k) touch(A.a[k] + v2); touch(A.c + v2); }
            if (v1 != 0xffffffffu) { touch(A.b + A.c[v1]); }
        } else if (mode == 2) {
            if (v2 != 0xffffffffu) { for (int k = 0; k < 8; ++k) touch2(A.a[k] + v2); touch2(A.c + v2); }
            if (v1 != 0xffffffffu) { touch2(A.b + A.c[v1]); }
        }
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += A.a[k][v0];
        s += A.b[A.c[v0]];
        acc += s;
        if (v1 == 0xffffffffu) break;
        v0 = v1; v1 = v2; v2 = v3; pos += T;
    }
    out[warp] = acc;
}
// register-pipelined variant: stage 1 (a[k][v], c[v]) one node ahead, stage 2 (b[c]) issued as soon as c is there
__global__ void k_service_pipe(Arr A, const uint32_t* __restrict__ order, uint32_t n, double* out) {
    if (threadIdx.x & 31) return;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    double acc = 0.0;
    uint32_t pos = warp;
    if (pos >= n) return;
    const uint32_t NONE = 0xffffffffu;
    // nodes: i (finish), i+1 (stage 2 issue), i+2 (stage 1 issue), i+3 (order load)
    uint32_t v1 = pos + T < n ? order[pos + T] : NONE, v2 = pos + 2 * T < n ? order[pos + 2 * T] : NONE;
    uint32_t v0 = order[pos];
    double a0[8], a1[8]; uint32_t c0, c1 = 0; double b0;
#pragma unroll
    for (int k = 0; k < 8; ++k) a0[k] = A.a[k][v0];
    c0 = A.c[v0];
    if (v1 != NONE) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a1[k] = A.a[k][v1];
        c1 = A.c[v1];
    }
    b0 = A.b[c0];
    for (;;) {
        uint32_t p3 = pos + 3 * T; uint32_t v3 = p3 < n ? order[p3] : NONE;
        double a2[8]; uint32_t c2 = 0; double b1 = 0.0;
        if (v2 != NONE) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a2[k] = A.a[k][v2];
            c2 = A.c[v2];
        }
        if (v1 != NONE) b1 = A.b[c1];
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += a0[k];
        s += b0;
        acc += s;
        if (v1 == NONE) break;
#pragma unroll
        for (int k = 0; k < 8; ++k) { a0[k] = a1[k]; a1[k] = a2[k]; }
        c0 = c1; c1 = c2; b0 = b1;
        v0 = v1; v1 = v2; v2 = v3; pos += T;
    }
    out[warp] = acc;
}
// cp.async pipeline: the worker lane streams its future nodes' static words into its own shared-memory rings
__device__ __forceinline__ void cpa4(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(s)), "l"(g) : "memory"); }
__device__ __forceinline__ void cpa8(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(s)), "l"(g) : "memory"); }
template <int L>
__global__ void k_service_cpa(Arr A, const uint32_t* __restrict__ order, uint32_t n, double* out) {
    __shared__ uint32_t s_v[8][4 * L];        // ring of node ids
    __shared__ double s_a[8][2 * L][8];       // first level
    __shared__ uint32_t s_c[8][2 * L];
    __shared__ double s_b[8][L];              // second level
    if (threadIdx.x & 31) return;
    const uint32_t wi = threadIdx.x >> 5;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, T = (gridDim.x * blockDim.x) >> 5;
    double acc = 0.0;
    if (warp >= n) return;
    const uint32_t cnt = (n - warp + T - 1) / T;      // nodes of this worker
    // iteration it: issue stage0 for node it+3L-?..; use a uniform schedule: at step t issue s0(t), s1(t-L), s2(t-2L), consume(t-3L)
    for (uint32_t t = 0; t < cnt + 3 * L; ++t) {
        if (t < cnt) cpa4(&s_v[wi][t % (4 * L)], order + (warp + t * T));
        if (t >= L && t - L < cnt) {
            uint32_t i = t - L; uint32_t v = s_v[wi][i % (4 * L)];
#pragma unroll
            for (int k = 0; k < 8; ++k) cpa8(&s_a[wi][i % (2 * L)][k], A.a[k] + v);
            cpa4(&s_c[wi][i % (2 * L)], A.c + v);
        }
        if (t >= 2 * L && t - 2 * L < cnt) {
            uint32_t i = t - 2 * L; uint32_t c = s_c[wi][i % (2 * L)];
            cpa8(&s_b[wi][i % L], A.b + c);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" ::"n"(L - 1) : "memory");
        if (t >= 3 * L) {
            uint32_t i = t - 3 * L;
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < 8; ++k) s += s_a[wi][i % (2 * L)][k];
            s += s_b[wi][i % L];
            acc += s;
        }
    }
    out[warp] = acc;
}
int main() {
    int sms = 148;
    const uint32_t N = 1u << 20;
    uint64_t* word; CK(cudaMalloc(&word, N * 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (uint32_t work : {0u, 4u}) for (uint32_t dist : {1u, 64u}) for (int bps : {1, 2, 4, 8}) {
        uint32_t n = 200000;
        CK(cudaMemset(word, 0, N * 8));
        std::vector<uint64_t> hs(n, SENT); CK(cudaMemcpy(word, hs.data(), n * 8, cudaMemcpyHostToDevice));
        void* args[] = {&word, &n, &dist, &work};
        cudaEventRecord(e0);
        CK(cudaLaunchCooperativeKernel((void*)k_chain, dim3(bps * sms), dim3(256), args, 0, 0));
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("chain work=%u dist=%u bps=%d: %.3f ms, %.1f ns per dependent hop\n", work, dist, bps, ms, ms * 1e6 / (n / dist));
    }
    // service
    Arr A; std::vector<uint32_t> perm(N), cidx(N); for (uint32_t i = 0; i < N; ++i) perm[i] = i;
    std::mt19937 rng(1); std::shuffle(perm.begin(), perm.end(), rng);
    for (uint32_t i = 0; i < N; ++i) cidx[i] = rng() % N;
    for (int k = 0; k < 8; ++k) { double* p; CK(cudaMalloc(&p, N * 8)); CK(cudaMemset(p, 0, N * 8)); A.a[k] = p; }
    { double* p; CK(cudaMalloc(&p, N * 8)); CK(cudaMemset(p, 0, N * 8)); A.b = p; }
    uint32_t *dc, *dorder; CK(cudaMalloc(&dc, N * 4)); CK(cudaMalloc(&dorder, N * 4));
    CK(cudaMemcpy(dc, cidx.data(), N * 4, cudaMemcpyHostToDevice)); A.c = dc;
    double* out; CK(cudaMalloc(&out, 1 << 20));
    for (int locality = 0; locality < 2; ++locality) {
        if (locality) { // order with spatial locality like a level-sorted map: blocks of 16 consecutive
            for (uint32_t i = 0; i < N; ++i) perm[i] = i;
        }
        CK(cudaMemcpy(dorder, perm.data(), N * 4, cudaMemcpyHostToDevice));
        for (int bps : {2, 4}) for (int mode = 0; mode < 6; ++mode) for (int rep = 0; rep < 2; ++rep) {
            uint32_t n = N;
            void* args[] = {&A, &dorder, &n, &mode, &out};
            void* args3[] = {&A, &dorder, &n, &out};
            cudaEventRecord(e0);
            if (mode < 3) CK(cudaLaunchCooperativeKernel((void*)k_service, dim3(bps * sms), dim3(256), args, 0, 0));
            else if (mode == 3) CK(cudaLaunchCooperativeKernel((void*)k_service_pipe, dim3(bps * sms), dim3(256), args3, 0, 0));
            else if (mode == 4) CK(cudaLaunchCooperativeKernel((void*)k_service_cpa<2>, dim3(bps * sms), dim3(256), args3, 0, 0));
            else CK(cudaLaunchCooperativeKernel((void*)k_service_cpa<4>, dim3(bps * sms), dim3(256), args3, 0, 0));
            cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            uint32_t T = bps * sms * 8;
            if (rep) printf("service locality=%d bps=%d mode=%d: %.3f ms, %.0f ns per node per worker\n", locality, bps, mode, ms, ms * 1e6 / (N / T));
        }
    }
    return 0;
}
