// ve_stencil.cu — the per-cell (stencil / map) stages of erode() as sm_100a kernels.
// Each kernel cites the reference lines whose arithmetic it reproduces operation by operation
// (device code is built with --fmad=false so a*b+c is never contracted, as in Rust).
#include "ve_internal.cuh"

// ---------------------------------------------------------------------------------------------
// E1: D8 receivers.  util.rs:251-279; ocean predicate of erosion.rs:818 when use_h_le0 != 0.
// E1 and E3b/c read the 3x3 neighbourhood of one f64 field: a CTA stages its 32x8 tile plus a one-cell halo in
// shared memory (coalesced rows, every value fetched from L2/HBM once per CTA instead of up to nine times) and
// the stencil runs out of shared memory.  Cells outside the map hold `pad`, chosen so that the reference's
// in-bounds test needs no branch: +inf is never a lower neighbour, NaN compares false both ways.
#define VE_TX 32
#define VE_TY 32          /* rows per CTA tile; VE_TR rows of threads walk them (several loads in flight per thread) */
#define VE_TR 8
__device__ __forceinline__ void stage_tile(const Geom& g, const double* __restrict__ a, double pad, double (*s)[VE_TX + 2]) {
    const int x0 = (int)blockIdx.x * VE_TX - 1, y0 = (int)blockIdx.y * VE_TY - 1;
#pragma unroll 5
    for (int i = (int)threadIdx.x; i < (VE_TX + 2) * (VE_TY + 2); i += VE_TX * VE_TR) {
        const int ty = i / (VE_TX + 2), tx = i - ty * (VE_TX + 2);
        const int x = x0 + tx, y = y0 + ty;
        s[ty][tx] = g.inb(x, y) ? a[g.idx(x, y)] : pad;
    }
    __syncthreads();
}
__global__ void __launch_bounds__(VE_TX * VE_TR) k_downhill(Geom g, const double* __restrict__ h, const uint8_t* __restrict__ ocean,
                                                            int use_h_le0, int32_t* __restrict__ dh) {
    __shared__ double s[VE_TY + 2][VE_TX + 2];
    stage_tile(g, h, INFINITY, s);
    const int lx = (int)(threadIdx.x % VE_TX);
    const int x = (int)blockIdx.x * VE_TX + lx;
    if (x >= g.nx) return;
#pragma unroll
    for (int ly = (int)(threadIdx.x / VE_TX); ly < VE_TY; ly += VE_TR) {
        const int y = (int)blockIdx.y * VE_TY + ly;
        if (y >= g.ny) break;
        const uint32_t posi = g.idx(x, y);
        double nh = s[ly + 1][lx + 1];
        bool oc = ocean[posi] != 0;
        if (use_h_le0) oc = oc && (nh <= 0.0);
        int best = -1; double besth = nh;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            double nbh = s[ly + 1 + c_ndy[k]][lx + 1 + c_ndx[k]];
            if (nbh < besth) { besth = nbh; best = (int)g.idx(x + c_ndx[k], y + c_ndy[k]); }
        }
        dh[posi] = oc ? -2 : best;
    }
}

static inline dim3 ve_tile_grid(const Geom& g) { return dim3((unsigned)((g.nx + VE_TX - 1) / VE_TX), (unsigned)((g.ny + VE_TY - 1) / VE_TY)); }

int ve_k_downhill(ve_ctx* ctx, const double* h, const uint8_t* ocean, int use_h_le0, int32_t* dh) {
    k_downhill<<<ve_tile_grid(ctx->g), VE_TX * VE_TR, 0, ctx->stream>>>(ctx->g, h, ocean, use_h_le0, dh);
    VE_LAUNCH_CHECK();
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// E3b/E3c: receiver mask, donor count and MFD weights.  erosion.rs:2407-2439, 2444-2479.
__global__ void __launch_bounds__(VE_TX * VE_TR) k_mrec(Geom g, const double* __restrict__ wh, uint8_t* __restrict__ mrec,
                                                        uint8_t* __restrict__ ndon, double* __restrict__ mwrec) {
    __shared__ double s[VE_TY + 2][VE_TX + 2];
    stage_tile(g, wh, __longlong_as_double(0x7ff8000000000000LL), s);
    const int lx = (int)(threadIdx.x % VE_TX);
    const int x = (int)blockIdx.x * VE_TX + lx;
    if (x >= g.nx) return;
    const double SQ = vm_sqrt(32.0 * 32.0 + 32.0 * 32.0);
    for (int ly = (int)(threadIdx.x / VE_TX); ly < VE_TY; ly += VE_TR) {
        const int y = (int)blockIdx.y * VE_TY + ly;
        if (y >= g.ny) break;
        const uint32_t ij = g.idx(x, y);
        double wh_ij = s[ly + 1][lx + 1];
        unsigned mr = 0, nd = 0;
        double wrec[8];
        double sumweight = 0.0; int nrec = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            wrec[k] = 0.0;
            double wh_ijk = s[ly + 1 + c_ndy[k]][lx + 1 + c_ndx[k]];      // NaN outside the map: neither branch is taken
            if (wh_ij > wh_ijk) {
                mr |= 1u << k;
                double lrec = (c_ndx[k] != 0 && c_ndy[k] != 0) ? SQ : 32.0;
                double w = (wh_ij - wh_ijk) / lrec;
                wrec[k] = w; sumweight += w; nrec += 1;
            } else if (wh_ijk > wh_ij) nd += 1;
        }
        mrec[ij] = (uint8_t)mr; ndon[ij] = (uint8_t)nd;
        double slope = sumweight / fmax((double)nrec, 1.0);
        double p_mfd_exp = 0.5 + 0.6 * slope;
        sumweight = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            double w = vm_pow(wrec[k], p_mfd_exp);
            sumweight += w; wrec[k] = w;
        }
        if (sumweight > 0.0) {
#pragma unroll
            for (int k = 0; k < 8; ++k) wrec[k] /= sumweight;
        }
        double2* out = reinterpret_cast<double2*>(mwrec + (size_t)ij * 8);
        out[0] = make_double2(wrec[0], wrec[1]); out[1] = make_double2(wrec[2], wrec[3]);
        out[2] = make_double2(wrec[4], wrec[5]); out[3] = make_double2(wrec[6], wrec[7]);
    }
}

int ve_k_mrec(ve_ctx* ctx) {
    k_mrec<<<ve_tile_grid(ctx->g), VE_TX * VE_TR, 0, ctx->stream>>>(ctx->g, ctx->wh, ctx->mrec, ctx->ndon, ctx->mwrec);
    VE_LAUNCH_CHECK();
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// E5: critical slope from Fbm<Perlin> rock strength.  erosion.rs:550-604; noise 0.9.0
// perlin_3d / Fbm::get restated (crate source is not in the reference tree).
__device__ __forceinline__ double d_grad3(unsigned hsh, double x, double y, double z) {
    switch (hsh & 15u) {
        case 0: return x + y;   case 1: return -x + y;  case 2: return x - y;   case 3: return -x - y;
        case 4: return x + z;   case 5: return -x + z;  case 6: return x - z;   case 7: return -x - z;
        case 8: return y + z;   case 9: return -y + z;  case 10: return y - z;  case 11: return -y - z;
        case 12: return x + y;  case 13: return -x + y; case 14: return -y + z; default: return -y - z;
    }
}
__device__ __forceinline__ double d_quintic(double t) { return t * t * t * (t * (t * 6.0 - 15.0) + 10.0); }

__device__ double d_perlin3(const uint8_t* p, double px, double py, double pz) {
    double fx = floor(px), fy = floor(py), fz = floor(pz);
    long long ix = (long long)fx, iy = (long long)fy, iz = (long long)fz;
    double dx = px - fx, dy = py - fy, dz = pz - fz;
    unsigned x0 = (unsigned)(ix & 0xff), x1 = (unsigned)((ix + 1) & 0xff);
    unsigned y0 = (unsigned)(iy & 0xff), y1 = (unsigned)((iy + 1) & 0xff);
    unsigned z0 = (unsigned)(iz & 0xff), z1 = (unsigned)((iz + 1) & 0xff);
    unsigned a0 = p[x0], a1 = p[x1];
    unsigned b00 = p[a0 ^ y0], b10 = p[a1 ^ y0], b01 = p[a0 ^ y1], b11 = p[a1 ^ y1];
    double g000 = d_grad3(p[b00 ^ z0], dx, dy, dz);
    double g100 = d_grad3(p[b10 ^ z0], dx - 1.0, dy, dz);
    double g010 = d_grad3(p[b01 ^ z0], dx, dy - 1.0, dz);
    double g110 = d_grad3(p[b11 ^ z0], dx - 1.0, dy - 1.0, dz);
    double g001 = d_grad3(p[b00 ^ z1], dx, dy, dz - 1.0);
    double g101 = d_grad3(p[b10 ^ z1], dx - 1.0, dy, dz - 1.0);
    double g011 = d_grad3(p[b01 ^ z1], dx, dy - 1.0, dz - 1.0);
    double g111 = d_grad3(p[b11 ^ z1], dx - 1.0, dy - 1.0, dz - 1.0);
    double u = d_quintic(dx), v = d_quintic(dy), w = d_quintic(dz);
    double k0 = g000;
    double k1 = g100 - g000;
    double k2 = g010 - g000;
    double k3 = g001 - g000;
    double k4 = g000 + g110 - g100 - g010;
    double k5 = g000 + g101 - g100 - g001;
    double k6 = g000 + g011 - g010 - g001;
    double k7 = g100 + g010 + g001 + g111 - g110 - g101 - g011 - g000;
    double r = k0 + k1 * u + k2 * v + k3 * w + k4 * u * v + k5 * u * w + k6 * v * w + k7 * u * v * w;
    r = r * 1.1547005383792515;
    return fmin(1.0, fmax(-1.0, r));
}

// rock strength from the raw noise value rs = rock_strength_nz.get(p) (erosion.rs:589-603)
__device__ __forceinline__ double max_slope_of(const ve_params& P, double rs, double wposz) {
    const double max_angle_range = P.max_max_angle - P.min_max_angle;
    const double logistic_2_base = vm_sqrt(3.0) * 0.63661977236758134308;
    const double center = 0.4, dmin = center - 0.05, dmax = center + 0.05;
    const double logit_center = vm_log(center) - vm_log1p(-center);
    rs = fmin(1.0, fmax(-1.0, rs)) * 0.5 + 0.5;
    double a = fmin(1.0 - 1e-7, fmax(1e-7, rs));
    double hz = fabs(wposz / P.mountain_scale);
    hz = fmin(dmax, fmax(dmin, hz));
    double la = vm_log(a) - vm_log1p(-a);
    double lh = vm_log(hz) - vm_log1p(-hz);
    double arg = 1.0 * la + 1.0 * (lh - logit_center);
    double rock = vm_tanh(arg / logistic_2_base) * 0.5 + 0.5;
    return rock * max_angle_range + P.min_max_angle;
}

__global__ void __launch_bounds__(256) k_max_slope(Geom g, ve_params P, NoiseDev nz, const double* __restrict__ h,
                                                   const double* __restrict__ hscale, double* __restrict__ out,
                                                   double* __restrict__ h_t) {
    extern __shared__ uint8_t s_perm[];
    for (unsigned i = threadIdx.x; i < nz.octaves * 256u; i += blockDim.x) s_perm[i] = nz.perm[i];
    __syncthreads();
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    double z = h[posi];
    h_t[posi] = z; // erosion.rs:854 h.to_vec()
    double px = (double)g.x_of(posi) * 32.0, py = (double)g.y_of(posi) * 32.0;
    double wposz = z / hscale[posi];
    double div_factor = (2.0 * 32.0) / 8.0;
    double pz = wposz * div_factor;
    // Fbm::get
    double result = 0.0, attenuation = nz.persistence;
    px *= nz.frequency; py *= nz.frequency; pz *= nz.frequency;
    for (uint32_t o = 0; o < nz.octaves; ++o) {
        double signal = d_perlin3(s_perm + o * 256u, px, py, pz);
        signal *= attenuation;
        attenuation *= nz.persistence;
        result += signal;
        px *= nz.lacunarity; py *= nz.lacunarity; pz *= nz.lacunarity;
    }
    out[posi] = max_slope_of(P, result * nz.scale_factor, wposz);
}

// host-callback noise (ve_set_noise_cb): sample points out, values back in
__global__ void __launch_bounds__(256) k_noise_points(Geom g, const double* __restrict__ h, const double* __restrict__ hscale,
                                                      double* __restrict__ xyz, double* __restrict__ h_t) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    double z = h[posi];
    h_t[posi] = z; // erosion.rs:854 h.to_vec()
    double wposz = z / hscale[posi];
    double div_factor = (2.0 * 32.0) / 8.0;
    xyz[3 * (size_t)posi] = (double)g.x_of(posi) * 32.0;
    xyz[3 * (size_t)posi + 1] = (double)g.y_of(posi) * 32.0;
    xyz[3 * (size_t)posi + 2] = wposz * div_factor;
}
__global__ void __launch_bounds__(256) k_max_slope_ext(Geom g, ve_params P, const double* __restrict__ rs, const double* __restrict__ h,
                                                       const double* __restrict__ hscale, double* __restrict__ out) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    out[posi] = max_slope_of(P, rs[posi], h[posi] / hscale[posi]);
}

int ve_k_max_slope(ve_ctx* ctx) {
    const size_t n = ctx->g.n;
    if (ctx->noise_cb) {
        double* dxyz;
        VE_TRY(ve_scratch(ctx, 0, n * 24, (void**)&dxyz));
        if (!ctx->noise_host) VE_CUDA(cudaMallocHost((void**)&ctx->noise_host, n * 32));
        double* hxyz = ctx->noise_host; double* hval = ctx->noise_host + 3 * n;
        k_noise_points<<<ve_blocks(n, 256), 256, 0, ctx->stream>>>(ctx->g, ctx->h, ctx->hscale, dxyz, ctx->h_t);
        VE_LAUNCH_CHECK();
        VE_CUDA(cudaMemcpyAsync(hxyz, dxyz, n * 24, cudaMemcpyDeviceToHost, ctx->stream));
        VE_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->noise_cb(ctx->noise_user, hxyz, hval, n);
        double* dval = dxyz;                                   // reuse the scratch for the values
        VE_CUDA(cudaMemcpyAsync(dval, hval, n * 8, cudaMemcpyHostToDevice, ctx->stream));
        k_max_slope_ext<<<ve_blocks(n, 256), 256, 0, ctx->stream>>>(ctx->g, ctx->P, dval, ctx->h, ctx->hscale, ctx->max_slope);
        VE_LAUNCH_CHECK();
        return VE_OK;
    }
    size_t smem = (size_t)ctx->nz.octaves * 256;
    k_max_slope<<<ve_blocks(ctx->g.n, 256), 256, smem, ctx->stream>>>(ctx->g, ctx->P, ctx->nz, ctx->h, ctx->hscale,
                                                                      ctx->max_slope, ctx->h_t);
    VE_LAUNCH_CHECK();
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// the parameter closures of WorldSim::generate (world/src/sim/mod.rs:849-893 scale factors,
// 1102-1337 closures) evaluated per chunk on the device, straight into the context's input arrays.  The noise-rs
// generators they sample are not part of the tree, so the host passes the sampled values (alt_old, uheight,
// rock_strength); everything downstream of the samples is here.  cdf = rank CDF of old_height over the land
// chunks (uniform_noise, util.rs:171-207; mod.rs:1118-1124).
__global__ void __launch_bounds__(256) k_erosion_inputs(Geom g, double grid_scale, double min_epsilon, double max_epsilon, double max_erosion_per_delta_t,
                                                        const float* __restrict__ alt_old, const uint8_t* __restrict__ ocean,
                                                        const float* __restrict__ cdf, const double* __restrict__ uheight_a,
                                                        const double* __restrict__ rock_a, float* __restrict__ uplift, float* __restrict__ nf,
                                                        float* __restrict__ mf, float* __restrict__ gf, float* __restrict__ eps0,
                                                        float* __restrict__ alpha, double* __restrict__ kf, double* __restrict__ kd,
                                                        double* __restrict__ hscale, double* __restrict__ h, double* __restrict__ b,
                                                        double* __restrict__ wh, float* __restrict__ oldh_out) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    const bool oc = ocean[posi] != 0;
    const float n = 1.0f, theta = 0.4f;                                             // n_func 1102-1107, theta_func 1146
    const double height_scale = vm_pow(grid_scale, (double)n);                      // 871
    const double time_scale = vm_pow(grid_scale, (double)n);                        // 873
    const float mountain_scale = 2048.0f;                                           // CONFIG.mountain_scale
    const float old_height = alt_old[posi] * mountain_scale * (float)height_scale;  // 1108-1110
    const float alt = oc ? old_height : (float)((double)old_height / (double)mountain_scale) - 0.5f;   // alt_func 1331-1337
    oldh_out[posi] = alt;
    h[posi] = (double)alt; b[posi] = (double)alt; wh[posi] = 0.0;                   // erosion.rs:2549-2559, 2601
    nf[posi] = n; mf[posi] = theta * n;                                             // erosion.rs:2570
    hscale[posi] = height_scale;                                                    // erosion.rs:848-850
    // uplift_fn 1312-1330 (uplift_uniform[posi].1 = (alt_old_no_ocean.0 - 0.5 + 0.5)^2, 1125-1133), cast at erosion.rs:2607
    double up = 0.0;
    if (!oc) {
        double oheight = (double)cdf[posi] - 0.5;
        double height = (oheight + 0.5) * (oheight + 0.5);
        height = (height - 0.0) / (1.0 - 0.0);
        height = height * (max_epsilon - min_epsilon) + min_epsilon;
        height = (height - min_epsilon) / (max_epsilon - min_epsilon);              // erosion_factor 1142-1144
        up = height * max_erosion_per_delta_t;
    }
    uplift[posi] = (float)up;
    const double kf_scale = vm_pow(grid_scale, -2.0 * (double)(theta * n));         // k_fs_scale 860
    kf[posi] = (oc ? 1.0e-4 : 1.0e-6) * kf_scale;                                   // kf_func 1147-1168
    kd[posi] = (1.0e-2 / 4.0) * (grid_scale * grid_scale / time_scale);             // kd_func 1169-1183, k_d_scale 883
    {   // g_func 1184-1200 with map_edge_factor (util.rs:22-29)
        int ex = g.x_of(posi), ey = g.y_of(posi);
        float fx = (float)(g.nx / 2 - abs(ex - g.nx / 2)) / ((16.0f / 1024.0f) * (float)g.nx);
        float fy = (float)(g.ny / 2 - abs(ey - g.ny / 2)) / ((16.0f / 1024.0f) * (float)g.ny);
        gf[posi] = fminf(fmaxf(fminf(fx, fy), 0.0f), 1.0f) == 0.0f ? 0.0f : 1.0f;
    }
    const float eps_scale = (float)(height_scale / time_scale);                     // epsilon_0_scale 885
    const float alpha_scale = (float)(1.0 / height_scale);                          // alpha_scale 879
    if (oc) {
        eps0[posi] = (2.078e-3f / 4.0f) * eps_scale;                                // 1206-1209
        alpha[posi] = 3.7e-2f * alpha_scale;                                        // 1260-1262
    } else {
        // ustrength 1241-1249 / 1293-1301: logistic_cdf(logit(rock_strength) + logit(uheight) - logit(center))
        const double logistic_2_base = vm_sqrt(3.0) * 0.63661977236758134308;
        const double center = 0.4, dmin = center - 0.05, dmax = center + 0.05;
        const double rs = fmin(1.0 - 1e-7, fmax(1e-7, rock_a[posi]));
        const double uh = fmin(dmax, fmax(dmin, uheight_a[posi]));
        const double arg = 1.0 * (vm_log(rs) - vm_log1p(-rs)) + 1.0 * ((vm_log(uh) - vm_log1p(-uh)) - (vm_log(center) - vm_log1p(-center)));
        const double us = vm_tanh(arg / logistic_2_base) * 0.5 + 0.5;
        eps0[posi] = (float)((1.0 - us) * (2.078e-3 - 5.3e-5) + 5.3e-5) / 4.0f * eps_scale;   // 1255-1256
        alpha[posi] = (float)(us * (4.2e-2 - 1.6e-2) + 1.6e-2) * alpha_scale;                  // 1309-1310
    }
}
__global__ void __launch_bounds__(256) k_land_height(uint32_t n, const float* __restrict__ alt_old, const uint8_t* __restrict__ ocean,
                                                     float height_scale_f, float* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = ocean[i] ? __int_as_float(0x7fc00000) : alt_old[i] * 2048.0f * height_scale_f;   // None on ocean chunks (1118-1124)
}
int ve_k_erosion_inputs(ve_ctx* ctx, double gen_scale, const float* d_alt_old, const uint8_t* d_ocean, const double* d_uheight,
                        const double* d_rock, float* d_cdf, float* d_tmp_f32, float* d_oldh_out, double* max_erosion_out) {
    const Geom& g = ctx->g;
    const uint32_t n = g.n;
    const unsigned nb = ve_blocks(n, 256);
    const double grid_scale = 1.0 / (4.0 / gen_scale);                                          // 849
    const double len_f = (double)n;
    const double min_epsilon = 1.0 / fmax(len_f, 2.220446049250313e-16 * 0.5);                  // 903
    const double max_epsilon = fmin(1.0 - 1.0 / len_f, 1.0 - 2.220446049250313e-16 * 0.5);      // 904
    const double max_erosion_per_delta_t = 64.0 * vm_pow(grid_scale, 1.0);                      // 889-890 (n_approx = 1)
    k_land_height<<<nb, 256, 0, ctx->stream>>>(n, d_alt_old, d_ocean, (float)vm_pow(grid_scale, 1.0), d_tmp_f32);
    VE_LAUNCH_CHECK();
    VE_TRY(ve_g_uniform_noise(ctx, d_tmp_f32, d_cdf, nullptr));
    k_erosion_inputs<<<nb, 256, 0, ctx->stream>>>(g, grid_scale, min_epsilon, max_epsilon, max_erosion_per_delta_t, d_alt_old, d_ocean, d_cdf,
                                                  d_uheight, d_rock, ctx->uplift, ctx->nf, ctx->mf, ctx->gf, ctx->eps0, ctx->alpha, ctx->kf,
                                                  ctx->kd, ctx->hscale, ctx->h, ctx->b, ctx->wh, d_oldh_out);
    VE_LAUNCH_CHECK();
    if (max_erosion_out) *max_erosion_out = max_erosion_per_delta_t;
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// E6: stream-power factors, fused pair.  erosion.rs:870-1006.
// (ncu, round 2: with vm_pow inlined at every call site the kernel stalled on instruction fetch - `no_instruction`
// was its top stall reason - so the receiver loop calls one out-of-line copy)
__device__ __noinline__ double pow_ool(double x, double y) { return vm_pow(x, y); }
__global__ void __launch_bounds__(256) k_kfactors(Geom g, ve_params P, float max_uplift, const int32_t* __restrict__ dh,
                                                  const double* __restrict__ b, const double* __restrict__ h_t,
                                                  const float* __restrict__ nf, const float* __restrict__ mf,
                                                  const double* __restrict__ kf, const float* __restrict__ gf,
                                                  const float* __restrict__ uplift, const double* __restrict__ area,
                                                  const double* __restrict__ max_slope, const uint8_t* __restrict__ mrec,
                                                  const double* __restrict__ mwrec, float* __restrict__ kfs,
                                                  float* __restrict__ kdf) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    float fs[8], df[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { fs[k] = 0.0f; df[k] = 0.0f; }
    if (dh[posi] >= 0) {
        const double dt = (double)max_uplift / 1e-3;
        const double sediment_thickness = 1.0e-4 * dt;
        const double chunk_area = 32.0 * 32.0, p = 1.0, q = 0.2, q_ = 1.5;
        const double SQ = vm_sqrt(32.0 * 32.0 + 32.0 * 32.0);
        double sed = h_t[posi] - b[posi];
        double k_fs = kf[posi] / 1.0;
        double kk_ = ((sed > sediment_thickness) ? (4.0 * k_fs) : k_fs) * dt;
        double n = (double)nf[posi], m = (double)mf[posi];
        double uplift_i = (double)uplift[posi];
        double area_i = area[posi];
        double ms = max_slope[posi];
        float gfi = gf[posi];
        double g_i = (sed > sediment_thickness) ? (double)(1.0f * gfi) : (double)gfi;
        double r0 = ms / P.mid_slope;
        double k_da = P.k_da * (r0 * r0);
        double ms_pow = pow_ool(ms, q_);
        unsigned mr = mrec[posi];
        const double* w = mwrec + (size_t)posi * 8;
        // the neighbour distance takes two values: its powers are evaluated once per cell, not once per receiver
        // (same arguments, same function: same bits)
        const double nd_n_s = pow_ool(32.0, n), nd_n_d = pow_ool(SQ, n), nd_q_s = pow_ool(32.0, q_), nd_q_d = pow_ool(SQ, q_);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            if (!((mr >> kk) & 1u)) continue;
            double wk = w[kk];
            const bool diag = (c_ndx[kk] != 0 && c_ndy[kk] != 0);
            double knew = kk_ * pow_ool(p * chunk_area * (area_i * wk), m) / (diag ? nd_n_d : nd_n_s);
            fs[kk] = (float)knew;
            double k = (wk * (uplift_i + (double)max_uplift * g_i / p)) / (1.0 + k_da * pow_ool(wk * 0.1e6, q)) / ms_pow;
            double knew2 = k * (1.0 + k_da * P.chunk_area_pow_q * pow_ool(area_i * wk, q)) / (diag ? nd_q_d : nd_q_s);
            df[kk] = (float)knew2;
        }
    }
    float4* o1 = reinterpret_cast<float4*>(kfs + (size_t)posi * 8);
    float4* o2 = reinterpret_cast<float4*>(kdf + (size_t)posi * 8);
    o1[0] = make_float4(fs[0], fs[1], fs[2], fs[3]); o1[1] = make_float4(fs[4], fs[5], fs[6], fs[7]);
    o2[0] = make_float4(df[0], df[1], df[2], df[3]); o2[1] = make_float4(df[4], df[5], df[6], df[7]);
}

int ve_k_kfactors(ve_ctx* ctx) {
    k_kfactors<<<ve_blocks(ctx->g.n, 256), 256, 0, ctx->stream>>>(ctx->g, ctx->P, ctx->max_uplift, ctx->dh, ctx->b,
                                                                  ctx->h_t, ctx->nf, ctx->mf, ctx->kf, ctx->gf, ctx->uplift,
                                                                  ctx->area, ctx->max_slope, ctx->mrec, ctx->mwrec,
                                                                  ctx->kfs, ctx->kdf);
    VE_LAUNCH_CHECK();
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// E10: basement + soil production (erosion.rs:1483-1531).  h is the post-solve height.
__global__ void __launch_bounds__(256) k_basement(Geom g, float max_uplift, const double* __restrict__ h,
                                                  double* __restrict__ b, const int32_t* __restrict__ dh,
                                                  const float* __restrict__ uplift, const float* __restrict__ eps0,
                                                  const float* __restrict__ alpha) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    const double dt = (double)max_uplift / 1e-3;
    double h_i = h[posi];
    double new_b_i = b[posi] + (double)uplift[posi];
    int32_t posj = dh[posi];
    double vertical_sed = fmax(h_i - new_b_i, 0.0);
    double h_normal;
    if (posj < 0) h_normal = vertical_sed;
    else {
        double h_j = h[posj];
        double ex = (double)(g.x_of(posi) - g.x_of((uint32_t)posj)) * 32.0;
        double ey = (double)(g.y_of(posi) - g.y_of((uint32_t)posj)) * 32.0;
        double nd2 = ex * ex + ey * ey;
        double dhh = h_i - h_j;
        double h_i_fact = dhh / (nd2 + dhh * dhh);
        double h_i_vertical = 1.0 - h_i_fact * dhh;
        h_normal = vertical_sed * vm_sqrt(h_i_fact * h_i_fact * nd2 + h_i_vertical * h_i_vertical);
    }
    double p_i = (double)eps0[posi] * dt * vm_exp(-(double)alpha[posi] * h_normal);
    new_b_i -= p_i;
    new_b_i = fmin(new_b_i, h_i);
    b[posi] = new_b_i;
}

// E11: lake silt (erosion.rs:1535-1556).  lake_silt / sill are grid-indexed here; lwv keeps the
// reference's mixed index space (slot = the stored sill value).
// Four consecutive cells per thread: the stage is a three-level chain of dependent loads (sill -> lwv[sill] -> h, silt,
// wh) with one cell's worth of bytes in flight per thread; four independent chains per thread keep HBM busy.
#define VE_LS_CELLS 4
__global__ void __launch_bounds__(256) k_lake_silt(Geom g, double* __restrict__ h, const double* __restrict__ wh,
                                                   const int32_t* __restrict__ sill, const double* __restrict__ lake_silt,
                                                   const double* __restrict__ lwv) {
    const uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) * VE_LS_CELLS;
    if (base >= g.n) return;
    int32_t lp[VE_LS_CELLS]; double vol[VE_LS_CELLS];
    if (base + VE_LS_CELLS <= g.n) {
        const int4 q = *reinterpret_cast<const int4*>(sill + base);
        lp[0] = q.x; lp[1] = q.y; lp[2] = q.z; lp[3] = q.w;
    } else {
#pragma unroll
        for (int c = 0; c < VE_LS_CELLS; ++c) lp[c] = base + c < g.n ? sill[base + c] : -1;
    }
#pragma unroll
    for (int c = 0; c < VE_LS_CELLS; ++c) vol[c] = lp[c] >= 0 ? lwv[lp[c]] : 0.0;
#pragma unroll
    for (int c = 0; c < VE_LS_CELLS; ++c) {
        if (lp[c] >= 0 && vol[c] > 0.0) {
            const uint32_t posi = base + c;
            double hv = h[posi];
            hv += fmax(0.0, fmin(lake_silt[posi], vol[c])) / vol[c] * (wh[posi] - hv);
            h[posi] = hv;
        }
    }
}

int ve_k_basement(ve_ctx* ctx) {
    unsigned nb = ve_blocks(ctx->g.n, 256);
    k_basement<<<nb, 256, 0, ctx->stream>>>(ctx->g, ctx->max_uplift, ctx->h, ctx->b, ctx->dh, ctx->uplift, ctx->eps0,
                                            ctx->alpha);
    VE_LAUNCH_CHECK();
    VE_CUDA(cudaMemcpyAsync(ctx->b_spl, ctx->b, (size_t)ctx->g.n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->max_g > 0.0f) {
        k_lake_silt<<<ve_blocks((ctx->g.n + VE_LS_CELLS - 1) / VE_LS_CELLS, 256), 256, 0, ctx->stream>>>(ctx->g, ctx->h, ctx->wh, ctx->sill_cur, ctx->lake_silt, ctx->lwv);
        VE_LAUNCH_CHECK();
    }
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// E12 (per-cell part): effective diffusivity.  erosion.rs:1597-1611, 1627-1673.  wh already
// holds this pass's running max (ve_g_thermal_wh), so wh[posj] is the receiver's new value.
__global__ void __launch_bounds__(256) k_thermal_kd(Geom g, float max_uplift, double kdsed, const double* __restrict__ h,
                                                    const double* __restrict__ b, const double* __restrict__ wh,
                                                    const int32_t* __restrict__ dh, const double* __restrict__ kd,
                                                    const double* __restrict__ max_slope, double* __restrict__ kd_eff) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    const double dt = (double)max_uplift / 1e-3;
    const double sediment_thickness = 1.0e-4 * dt;
    const double max_stable = 32.0 * 32.0 / dt;
    double old_h_i = h[posi];
    double sed = old_h_i - b[posi];
    double ms = max_slope[posi];
    double kv = (sed > sediment_thickness && kdsed > 0.0) ? kdsed : kd[posi];
    int32_t posj = dh[posi];
    if (posj >= 0) {
        double wh_j = wh[posj];
        if (wh_j < old_h_i) {
            double h_k = h[posj];
            double ex = (double)(g.x_of(posi) - g.x_of((uint32_t)posj)) * 32.0;
            double ey = (double)(g.y_of(posi) - g.y_of((uint32_t)posj)) * 32.0;
            double nd = vm_sqrt(ex * ex + ey * ey);
            double dz = fmax(old_h_i - h_k, 0.0);
            double mag_slope = dz / nd;
            if (!(mag_slope >= ms)) {
                double sr0 = mag_slope / ms;
                double slope_ratio = sr0 * sr0;
                double om = 1.0 - slope_ratio;
                double f = slope_ratio * ((3.0 - slope_ratio) / (om * om));
                kv += fmin(kv * f, max_stable);
            }
        }
    }
    kd_eff[posi] = kv;
}

int ve_k_thermal_kd(ve_ctx* ctx) {
    double kdsed = 1.5e-2 / 4.0 * ctx->P.k_d_scale;
    k_thermal_kd<<<ve_blocks(ctx->g.n, 256), 256, 0, ctx->stream>>>(ctx->g, ctx->max_uplift, kdsed, ctx->h, ctx->b, ctx->wh,
                                                                    ctx->dh, ctx->kd, ctx->max_slope, ctx->kd_eff);
    VE_LAUNCH_CHECK();
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// E13: ADI hillslope diffusion (diffusion.rs:41-394) with the Thomas solver (408-472).
// One WARP per line.  The Thomas recurrence itself is serial and keeps the reference's operation
// order (bit-exact): ~130 cycles per element (one f64 division on the chain).  What made a
// thread-per-line version 10x slower than that chain was its nine loads per element with nothing to
// overlap them; here the lanes build the coefficients of 32 consecutive elements at once
// (coalesced along rows), the recurrence then runs out of registers in every lane alike (values by
// shuffle, so the warp stays converged), and lane j keeps element j's results for a coalesced store.
// dir 0: x-sweep over rows 1..ny-2 (reads zp = original h for the explicit y terms, writes z).
// dir 1: y-sweep over columns 1..nx-2 (reads z = x-sweep result, writes zp).
__global__ void __launch_bounds__(128) k_adi_sweep(int nx, int ny, int dir, double dt, const double* __restrict__ kd,
                                                   const double* __restrict__ src, double* __restrict__ dst,
                                                   double* __restrict__ gam, int* __restrict__ fail) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int line = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) + 1;
    int nlines = dir == 0 ? ny : nx;
    if (line >= nlines - 1) return;
    const int n = dir == 0 ? nx : ny;
    const long long sa = dir == 0 ? 1 : nx;       // stride along the line
    const long long sc = dir == 0 ? nx : 1;       // stride across lines
    const long long base = (long long)line * sc;
    const double dx = 32.0, dy = 32.0;            // xl/nx, yl/ny with xl = nx*32 (erosion.rs:1720-1721)
    const double ca = dir == 0 ? (dx * dx) : (dy * dy); // along-line spacing^2
    const double cc = dir == 0 ? (dy * dy) : (dx * dx); // cross-line spacing^2
    // forward elimination.  Element 0 (Dirichlet) goes through the same recurrence as the last one:
    // with diag 1, inf 0 and the initial bet = 1, prev_sup = 0 it yields u = src / 1 exactly.
    double bet = 1.0, uprev = 0.0, prev_sup = 0.0;
    bool bad = false;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        const long long ij = base + (long long)i * sa;
        double inf_i = 0.0, diag_i = 1.0, sup_i = 0.0, f_i = 0.0;
        if (i < n) {
            if (i > 0 && i < n - 1) {
                double kc = kd[ij];
                double fap = (kd[ij + sa] + kc) / 2.0 * (dt / 2.0) / ca;
                double fam = (kd[ij - sa] + kc) / 2.0 * (dt / 2.0) / ca;
                double fcp = (kd[ij + sc] + kc) / 2.0 * (dt / 2.0) / cc;
                double fcm = (kd[ij - sc] + kc) / 2.0 * (dt / 2.0) / cc;
                diag_i = 1.0 + fap + fam;
                sup_i = -fap;
                inf_i = -fam;
                f_i = src[ij] + fcp * src[ij + sc] - (fcp + fcm) * src[ij] + fcm * src[ij - sc];
            } else {
                f_i = src[ij];
            }
        }
        const int cnt = min(32, n - i0);
        double my_gam = 0.0, my_u = 0.0;
#pragma unroll 8
        for (int j = 0; j < cnt; ++j) {
            double inf_j = __shfl_sync(FULL, inf_i, j), diag_j = __shfl_sync(FULL, diag_i, j);
            double f_j = __shfl_sync(FULL, f_i, j), sup_j = __shfl_sync(FULL, sup_i, j);
            double gm = prev_sup / bet;
            bet = diag_j - inf_j * gm;
            if (bet == 0.0) bad = true;
            uprev = (f_j - inf_j * uprev) / bet;
            prev_sup = sup_j;
            if (lane == j) { my_gam = gm; my_u = uprev; }
        }
        if (i < n) { if (i > 0) gam[ij] = my_gam; dst[ij] = my_u; }
    }
    __syncwarp();
    // back substitution, chunks from the end
    double unext = 0.0;
    for (int i0 = ((n - 1) / 32) * 32; i0 >= 0; i0 -= 32) {
        const int i = i0 + lane;
        const long long ij = base + (long long)i * sa;
        double d_i = 0.0, g_i = 0.0;
        if (i < n) { d_i = dst[ij]; if (i < n - 1) g_i = gam[ij + sa]; }
        const int cnt = min(32, n - i0);
        double my_u = 0.0;
#pragma unroll 8
        for (int j = cnt - 1; j >= 0; --j) {
            double d_j = __shfl_sync(FULL, d_i, j), g_j = __shfl_sync(FULL, g_i, j);
            double u = d_j;
            if (i0 + j < n - 1) u -= g_j * unext;
            unext = u;
            if (lane == j) my_u = u;
        }
        if (i < n - 1) dst[ij] = my_u;
    }
    if (bad && lane == 0) *fail = 1;
}

__global__ void __launch_bounds__(256) k_adi_finish(uint32_t n, const double* __restrict__ zp, double* __restrict__ h,
                                                    double* __restrict__ b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double hv = zp[i];
    h[i] = hv;
    b[i] = fmin(hv, b[i]);
}

int ve_k_diffusion(ve_ctx* ctx, double dt, double* h, double* b, const double* kd_eff) {
    const Geom& g = ctx->g;
    size_t nb = (size_t)g.n * 8;
    double *z, *zp, *gam; int* fail;
    VE_TRY(ve_scratch(ctx, 0, nb, (void**)&z));
    VE_TRY(ve_scratch(ctx, 1, nb, (void**)&zp));
    VE_TRY(ve_scratch(ctx, 2, nb, (void**)&gam));
    VE_TRY(ve_scratch(ctx, 3, 256, (void**)&fail));
    VE_CUDA(cudaMemsetAsync(fail, 0, sizeof(int), ctx->stream));
    // zint = h; zintp = zint.clone()  (diffusion.rs:108-120); kdsed = -1 keeps kdint = kd_eff
    VE_CUDA(cudaMemcpyAsync(z, h, nb, cudaMemcpyDeviceToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(zp, h, nb, cudaMemcpyDeviceToDevice, ctx->stream));
    k_adi_sweep<<<ve_blocks((uint64_t)g.ny * 32, 128), 128, 0, ctx->stream>>>(g.nx, g.ny, 0, dt, kd_eff, zp, z, gam, fail);
    VE_LAUNCH_CHECK();
    k_adi_sweep<<<ve_blocks((uint64_t)g.nx * 32, 128), 128, 0, ctx->stream>>>(g.nx, g.ny, 1, dt, kd_eff, z, zp, gam, fail);
    VE_LAUNCH_CHECK();
    k_adi_finish<<<ve_blocks(g.n, 256), 256, 0, ctx->stream>>>(g.n, zp, h, b);
    VE_LAUNCH_CHECK();
    int hfail = 0;
    VE_CUDA(cudaMemcpyAsync(&hfail, fail, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    if (hfail) { ctx->err = "tridag: zero pivot"; return VE_ERR_TRIDAG_SINGULAR; }
    return VE_OK;
}

// ---------------------------------------------------------------------------------------------
// ve_diffusion_static — the ADI diffusion for callers that are not allowed to allocate.
//
// Who needs this: a host that hands this library a GPU whose memory it has already carved up
// completely (an engine or game server running with a fixed device arena), or that calls from a
// section of its frame where the CUDA allocator must not be entered.  Such a caller can own the
// input/output buffers, but it cannot give us the 3 N doubles of work space the solver needs, and
// ve_diffusion() obtains them with cudaMalloc.  Here the work space is a fixed array that is part
// of the library image itself: every GPU that loads libvelerode.so carries these
// VE_STATIC_SCRATCH_BYTES (6 MiB + 256 B) until the process exits, which is why the map size is
// capped at VE_STATIC_MAX_CELLS.
//
// The array's name means something different on each side: device code may use it directly, host
// code only sees a stand-in symbol, so the host asks the runtime for the real device address once
// per device (with that device current) and passes that pointer to the kernels.  That lookup
// happens inside the first call made on a device — make that first call outside any stream
// capture.  All calls on a device share the one array and it keeps whatever the previous call
// left in it; the only word that has to start at zero (the zero-pivot flag) is cleared on every
// call, everything else is overwritten before it is read.  Two calls that overlap on the same GPU
// (different streams, or different host threads) would trample each other's work space: serialise
// them.
#define VE_STATIC_MAX_CELLS (512u * 512u)
#define VE_STATIC_SCRATCH_BYTES (3ull * VE_STATIC_MAX_CELLS * 8ull + 256ull)
static __device__ __align__(256) unsigned char g_ve_static_scratch[VE_STATIC_SCRATCH_BYTES];

#include <mutex>
static std::mutex g_static_mu;
static void* g_static_ptr[64];
static bool g_static_ready[64];

static void static_err(char* buf, size_t cap, const char* msg) {
    if (buf && cap) { snprintf(buf, cap, "%s", msg); }
}

extern "C" int ve_diffusion_static(uint32_t x_lg, uint32_t y_lg, double dt, double* d_h, double* d_b, const double* d_kd_eff,
                                   void* cuda_stream, char* errbuf, size_t errbuf_len) {
    if (!d_h || !d_b || !d_kd_eff || x_lg < 2 || y_lg < 2 || x_lg > 14 || y_lg > 14) {
        static_err(errbuf, errbuf_len, "ve_diffusion_static: null pointer or map size out of range");
        return VE_ERR_INVALID;
    }
    const uint64_t cells = (1ull << x_lg) * (1ull << y_lg);
    if (cells > VE_STATIC_MAX_CELLS) {
        char m[200];
        snprintf(m, sizeof(m), "ve_diffusion_static: %llu cells exceed the built-in work space (max %u cells = 512x512); "
                               "use ve_diffusion, which allocates", (unsigned long long)cells, VE_STATIC_MAX_CELLS);
        static_err(errbuf, errbuf_len, m);
        return VE_ERR_TOO_LARGE;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) { static_err(errbuf, errbuf_len, "ve_diffusion_static: no current CUDA device"); return VE_ERR_CUDA; }
    void* base = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ready[dev]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_ve_static_scratch);   // device address of the array on THIS device
            if (e != cudaSuccess || !p) {
                static_err(errbuf, errbuf_len, cudaGetErrorString(e));
                return VE_ERR_CUDA;
            }
            g_static_ptr[dev] = p;
            g_static_ready[dev] = true;
        }
        base = g_static_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int nx = 1 << x_lg, ny = 1 << y_lg;
    const uint32_t n = (uint32_t)cells;
    const size_t nb = (size_t)n * 8;
    double* z = (double*)base;
    double* zp = z + VE_STATIC_MAX_CELLS;
    double* gam = zp + VE_STATIC_MAX_CELLS;
    int* fail = (int*)((unsigned char*)base + 3ull * VE_STATIC_MAX_CELLS * 8ull);
#define ST_CUDA(call)                                                                     \
    do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { static_err(errbuf, errbuf_len, cudaGetErrorString(e_)); return VE_ERR_CUDA; } } while (0)
    ST_CUDA(cudaMemsetAsync(fail, 0, sizeof(int), st));
    ST_CUDA(cudaMemcpyAsync(z, d_h, nb, cudaMemcpyDeviceToDevice, st));
    ST_CUDA(cudaMemcpyAsync(zp, d_h, nb, cudaMemcpyDeviceToDevice, st));
    k_adi_sweep<<<ve_blocks((uint64_t)ny * 32, 128), 128, 0, st>>>(nx, ny, 0, dt, d_kd_eff, zp, z, gam, fail);
    ST_CUDA(cudaGetLastError());
    k_adi_sweep<<<ve_blocks((uint64_t)nx * 32, 128), 128, 0, st>>>(nx, ny, 1, dt, d_kd_eff, z, zp, gam, fail);
    ST_CUDA(cudaGetLastError());
    k_adi_finish<<<ve_blocks(n, 256), 256, 0, st>>>(n, zp, d_h, d_b);
    ST_CUDA(cudaGetLastError());
    int hfail = 0;
    ST_CUDA(cudaMemcpyAsync(&hfail, fail, sizeof(int), cudaMemcpyDeviceToHost, st));
    ST_CUDA(cudaStreamSynchronize(st));
#undef ST_CUDA
    if (hfail) { static_err(errbuf, errbuf_len, "ve_diffusion_static: tridag zero pivot"); return VE_ERR_TRIDAG_SINGULAR; }
    return VE_OK;
}
