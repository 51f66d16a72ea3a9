// cycles per Newton iteration of the warp-per-node variant under different numbers of busy warps per SM
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define FULL 0xffffffffu
__device__ __forceinline__ double widen_fast(float x) {   // normal numbers only (benchmark)
    uint32_t u = __float_as_uint(x);
    uint32_t hi = (u & 0x80000000u) | (((u & 0x7fffffffu) >> 3) + 0x38000000u);
    uint32_t lo = u << 29;
    return __hiloint2double((int)hi, (int)lo);
}
template <int VAR>
__global__ void k(double* out, long long* cyc, const float* rec_h_in, const float* a_in, int iters, double h0, double start, double omega, unsigned pmask) {
    const uint32_t lane = threadIdx.x & 31u;
    const bool part = (pmask >> lane) & 1u;
    float rec_h = rec_h_in[lane & 7], cfs = a_in[lane & 7], cfd = a_in[8 + (lane & 7)];
    float nfl = 1.0f, qfl = 1.5f;
    double new_h_i = start;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        float nh32 = (float)new_h_i;
        float dhv = part ? fmaxf(0.0f, nh32 - rec_h) : 1.0f;
        float sq = (VAR == 4) ? dhv * 0.37f : __fsqrt_rn(dhv);
        float dh_fs = cfs * 1.0f, dh_df = cfd * sq;
        double f = new_h_i - h0, df = 1.0;
        if (VAR == 1) {
            double tf = part ? (double)((dh_fs + dh_df) * dhv) : -0.0, tdf = part ? (double)(nfl * dh_fs + qfl * dh_df) : -0.0;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) { f += __shfl_sync(FULL, tf, kk); df += __shfl_sync(FULL, tdf, kk); }
        } else if (VAR == 2) {
            float tf = part ? (dh_fs + dh_df) * dhv : 1e-30f, tdf = part ? nfl * dh_fs + qfl * dh_df : 1e-30f;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) { f += widen_fast(__shfl_sync(FULL, tf, kk)); df += widen_fast(__shfl_sync(FULL, tdf, kk)); }
        } else if (VAR == 5) {   // 4 slots only
            float tf = part ? (dh_fs + dh_df) * dhv : -0.0f, tdf = part ? nfl * dh_fs + qfl * dh_df : -0.0f;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) { f += (double)__shfl_sync(FULL, tf, kk); df += (double)__shfl_sync(FULL, tdf, kk); }
        } else {
            float tf = part ? (dh_fs + dh_df) * dhv : -0.0f, tdf = part ? nfl * dh_fs + qfl * dh_df : -0.0f;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) { f += (double)__shfl_sync(FULL, tf, kk); df += (double)__shfl_sync(FULL, tdf, kk); }
        }
        double q = (VAR == 3) ? f * df : f / df;
        double hn = new_h_i - q;
        new_h_i = new_h_i * (1.0 - omega) + hn * omega;
        new_h_i = fmin(fmax(new_h_i, 50.0), 60.0);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = new_h_i;
}
template <int VAR> void run(const char* what, int warps_per_sm, double* out, long long* cyc, float* rh, float* a) {
    int iters = 2000;
    k<VAR><<<148, 32 * warps_per_sm>>>(out, cyc, rh, a, iters, 55.0, 54.0, 0.875, 0x17u);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-34s busy warps/SM=%2d: %.0f cycles per iteration\n", what, warps_per_sm, (double)c / iters);
}
int main() {
    double* out; long long* cyc; float *rh, *a; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 64); cudaMalloc(&rh, 64); cudaMalloc(&a, 128);
    float hrh[8] = {50.1f, 49.7f, 51.2f, 48.0f, 52.f, 47.f, 46.f, 53.f}, ha[16];
    for (int i = 0; i < 16; ++i) ha[i] = 0.01f + 0.003f * i;
    cudaMemcpy(rh, hrh, 32, cudaMemcpyHostToDevice); cudaMemcpy(a, ha, 64, cudaMemcpyHostToDevice);
    for (int w : {1, 2, 4, 8, 16}) {
        run<0>("W0 shfl f32 + 16 F2F", w, out, cyc, rh, a);
        run<1>("W1 2 F2F + 64-bit shfl", w, out, cyc, rh, a);
        run<2>("W2 shfl f32 + integer widen", w, out, cyc, rh, a);
        run<3>("W3 = W0 without the division", w, out, cyc, rh, a);
        run<4>("W4 = W0 without the sqrt", w, out, cyc, rh, a);
        run<5>("W5 = W0 with 4 slots", w, out, cyc, rh, a);
    }
    return 0;
}
