#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N 512
template <int OP> __global__ void k(double* out, long long* cyc, double a, double b, float fa, float fb) {
    double x = a; float y = fa;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) {
        if (OP == 0) x = __dadd_rn(x, b);
        else if (OP == 1) x = __fma_rn(x, b, a);
        else if (OP == 2) x = __dmul_rn(x, b);
        else if (OP == 3) x = __ddiv_rn(a, x) ;
        else if (OP == 4) y = __fadd_rn(y, fb);
        else if (OP == 5) y = __fsqrt_rn(y) + fb;
        else if (OP == 6) { y = (float)x; x = (double)y + b; }        // F2F both ways + DADD
        else if (OP == 7) { y = (float)x; x = (double)y; }            // F2F both ways
        else if (OP == 8) { float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y)); y = r; }
        else if (OP == 9) y = fmaxf(y * fb, fa);
        else if (OP == 10) x = fabs(x - b) ;
    }
    long long t1 = clock64();
    out[threadIdx.x] = x + y; cyc[0] = t1 - t0;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 4096); cudaMalloc(&cyc, 64);
    const char* names[] = {"DADD", "DFMA", "DMUL", "DDIV", "FADD", "FSQRT+FADD", "F2F.f32<-f64,F2F.f64<-f32,DADD", "F2F both", "MUFU.RSQ", "FMUL+FMNMX", "DADD+abs"};
    for (int rep = 0; rep < 2; ++rep) {
#define RUN(OP) { k<OP><<<1, 1>>>(out, cyc, 1.0000001, 0.9999999, 1.5f, 1.0000001f); long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("%-34s %.1f cycles per dependent op\n", names[OP], (double)c / N); }
        RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(10)
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0); printf("clock rate attr %d kHz\n", clk);
    return 0;
}
