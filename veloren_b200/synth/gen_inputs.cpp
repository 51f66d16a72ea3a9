// gen_inputs.cpp — deterministic synthetic pre-erosion fields for do_erosion.
//
// The reference builds these fields in WorldSim::generate (world/src/sim/mod.rs:849-893,
// 1092-1337) from noise-rs generators seeded by ChaCha; neither crate is on the box, so the
// fields here follow the reference's *recipes* (same formulas, scales, casts and ocean/land
// split) with a counter-hash lattice noise in place of noise-rs.  Only + - * / sqrt floor and
// ve_exact_math.h are used, so the output is bit-identical on every machine.
//
// kind 0: fBm "continent" world (configs C1-C4).  kind 1: depression-heavy plateau (C5).

#include "../../include/ve_exact_math.h"
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

inline uint64_t mix64(uint64_t z) {
    z += 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
inline double hash_unit(uint64_t seed, int64_t x, int64_t y, uint64_t salt) { // [-1, 1)
    uint64_t h = mix64(seed ^ mix64((uint64_t)x * 0x632be59bd9b4e019ULL + salt) ^ mix64((uint64_t)y * 0x8cb92ba72f3d8dd7ULL + 3 * salt + 1));
    return (double)(h >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}
inline double quintic(double t) { return t * t * t * (t * (t * 6.0 - 15.0) + 10.0); }
double value_noise(uint64_t seed, double x, double y, uint64_t salt) {
    double fx = std::floor(x), fy = std::floor(y);
    int64_t ix = (int64_t)fx, iy = (int64_t)fy;
    double u = quintic(x - fx), v = quintic(y - fy);
    double a = hash_unit(seed, ix, iy, salt), b = hash_unit(seed, ix + 1, iy, salt);
    double c = hash_unit(seed, ix, iy + 1, salt), d = hash_unit(seed, ix + 1, iy + 1, salt);
    double ab = a + (b - a) * u, cd = c + (d - c) * u;
    return ab + (cd - ab) * v;
}
double fbm2(uint64_t seed, double x, double y, int octaves, double persistence, uint64_t salt) {
    double amp = 1.0, sum = 0.0, norm = 0.0;
    for (int o = 0; o < octaves; ++o) {
        sum += amp * value_noise(seed, x, y, salt + 17 * (uint64_t)o);
        norm += amp; amp *= persistence; x *= 2.0; y *= 2.0;
    }
    return sum / norm;
}

// util.rs:23-30
float map_edge_factor(int x, int y, int szx, int szy) {
    auto f = [](int e, int sz) { return (float)(sz / 2 - std::abs(e - sz / 2)) / (16.0f / 1024.0f * (float)sz); };
    float m = std::min(f(x, szx), f(y, szy));
    return std::min(1.0f, std::max(0.0f, m));
}

} // namespace

extern "C" {

// scalars out: [0] k_d_scale, [1] k_da, [2] max_erosion_per_delta_t, [3] noise frequency,
// [4] lacunarity, [5] persistence, [6] scale_factor, [7] n_land
int vegen_generate(uint32_t x_lg, uint32_t y_lg, uint64_t seed, int kind, double scale, float* oldh, float* oldb,
                   uint8_t* is_ocean, float* uplift, float* n, float* m, float* g, float* eps0, float* alpha,
                   double* kf, double* kd, double* hscale, uint8_t* perm /*10*256*/, double* scalars) {
    const int nx = 1 << x_lg, ny = 1 << y_lg;
    const size_t N = (size_t)nx * ny;
    const double grid_scale = 1.0 / (4.0 / scale);                    // mod.rs:849
    const double height_scale = grid_scale;                            // n = 1 (871)
    const double time_scale = grid_scale;                              // 873
    const double k_d_scale = grid_scale * grid_scale / time_scale;     // 883
    const double max_erosion_per_delta_t = 64.0 * time_scale;          // 890
    const float mountain_scale = 2048.0f, sea_level = 140.0f;
    const double sea = (double)sea_level / (double)mountain_scale;

    std::vector<float> alt_old(N);
    // alt_old (mod.rs:1078-1088 shape: (base + main) * edge + sea*edge - sea, as f32)
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)N; ++i) {
        int x = (int)(i & (nx - 1)), y = (int)(i >> x_lg);
        double mef = (double)map_edge_factor(x, y, nx, ny);
        double v;
        if (kind == 1) {
            double cx = (x + 0.5 - nx * 0.5) / (nx * 0.5), cy = (y + 0.5 - ny * 0.5) / (ny * 0.5);
            double r2 = cx * cx + cy * cy;
            v = (r2 < 0.64 ? 0.6 : -0.5) + 0.05 * hash_unit(seed, x, y, 99);
            alt_old[i] = (float)v;
            continue;
        }
        // feature size tied to cells (not map fraction) so larger maps hold more continents
        double sx = (double)x / 256.0, sy = (double)y / 256.0;
        double cx = (x + 0.5 - nx * 0.5) / (nx * 0.5), cy = (y + 0.5 - ny * 0.5) / (ny * 0.5);
        double r2 = cx * cx + cy * cy;
        double base = 0.30 * (1.0 - 1.6 * r2);
        double main = 0.75 * fbm2(seed, sx, sy, 9, 0.55, 1);
        double rough = 0.06 * fbm2(seed, sx * 32.0, sy * 32.0, 3, 0.6, 7);
        v = base + main + rough;
        v = std::min(1.067, std::max(-0.946, v));
        alt_old[i] = (float)((v * mef + sea * mef) - sea);
    }
    // util.rs:332-367 get_oceans
    std::memset(is_ocean, 0, N);
    {
        std::vector<size_t> stack;
        auto push = [&](int x, int y) { size_t p = ((size_t)y << x_lg) | (size_t)x; if (alt_old[p] <= 0.0f) stack.push_back(p); };
        for (int x = 0; x < nx; ++x) { push(x, 0); push(x, ny - 1); }
        for (int y = 1; y < ny - 1; ++y) { push(0, y); push(nx - 1, y); }
        const int DX[8] = {-1, 0, 1, -1, 1, -1, 0, 1}, DY[8] = {-1, -1, -1, 0, 0, 1, 1, 1};
        while (!stack.empty()) {
            size_t c = stack.back(); stack.pop_back();
            if (is_ocean[c]) continue;
            is_ocean[c] = 1;
            int x = (int)(c & (nx - 1)), y = (int)(c >> x_lg);
            for (int k = 0; k < 8; ++k) {
                int xx = x + DX[k], yy = y + DY[k];
                if (xx < 0 || yy < 0 || xx >= nx || yy >= ny) continue;
                size_t p = ((size_t)yy << x_lg) | (size_t)xx;
                if (alt_old[p] <= 0.0f) stack.push_back(p);
            }
        }
    }
    // old_height (1108-1110), alt_func (1331-1337)
    std::vector<float> old_height(N);
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)N; ++i) {
        old_height[i] = alt_old[i] * mountain_scale * (float)height_scale;
        float a = is_ocean[i] ? old_height[i] : (float)((double)old_height[i] / (double)mountain_scale) - 0.5f;
        oldh[i] = a; oldb[i] = a;
    }
    // uniform_noise over land (util.rs:171-207; mod.rs:1118-1133): rank CDF, ties by index
    std::vector<uint32_t> land; land.reserve(N);
    for (size_t i = 0; i < N; ++i) if (!is_ocean[i]) land.push_back((uint32_t)i);
    std::sort(land.begin(), land.end(), [&](uint32_t a, uint32_t b) {
        if (old_height[a] != old_height[b]) return old_height[a] < old_height[b];
        return a < b;
    });
    const float total = (float)land.size();
    const double len_f = (double)nx * (double)ny;
    const double min_epsilon = 1.0 / std::max(len_f, 2.220446049250313e-16 * 0.5);
    const double max_epsilon = std::min(1.0 - 1.0 / len_f, 1.0 - 2.220446049250313e-16 * 0.5);
    std::memset(uplift, 0, N * sizeof(float));
    for (size_t r = 0; r < land.size(); ++r) {
        size_t i = land[r];
        float cdf = (float)(1 + r) / total;
        double oheight = (double)cdf - 0.5;
        double height = (oheight + 0.5) * (oheight + 0.5);                     // 1129-1131
        height = height * (max_epsilon - min_epsilon) + min_epsilon;           // 1319
        height = (height - min_epsilon) / (max_epsilon - min_epsilon);         // 1142-1144
        double u = (kind == 1) ? 16.0 : height * max_erosion_per_delta_t;      // 1329 (C5: constant)
        uplift[i] = (float)u;                                                  // erosion.rs:2607
    }
    const float theta = 0.4f;
    const double kfs = vm_pow(grid_scale, -2.0 * (double)(theta * 1.0f));      // 860
    const float eps_scale = (float)(height_scale / time_scale);               // 885
    const float alpha_scale = (float)(1.0 / height_scale);                    // 879
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)N; ++i) {
        int x = (int)(i & (nx - 1)), y = (int)(i >> x_lg);
        n[i] = 1.0f; m[i] = theta * n[i];                                      // erosion.rs:2570
        kf[i] = (is_ocean[i] ? 1.0e-4 : 1.0e-6) * kfs;                         // 1147-1168
        kd[i] = (1.0e-2 / 4.0) * k_d_scale;                                    // 1169-1183
        g[i] = map_edge_factor(x, y, nx, ny) == 0.0f ? 0.0f : 1.0f;            // 1184-1200
        hscale[i] = height_scale;
        if (is_ocean[i]) {
            eps0[i] = (float)(2.078e-3 / 4.0) * eps_scale;                     // 1206-1209
            alpha[i] = 3.7e-2f * alpha_scale;                                  // 1260-1262
        } else {
            double s = 0.5 + 0.5 * fbm2(seed, (double)x / 96.0, (double)y / 96.0, 4, 0.5, 31);
            s = std::min(1.0, std::max(0.0, s));
            eps0[i] = (float)((1.0 - s) * (2.078e-3 - 5.3e-5) + 5.3e-5) / 4.0f * eps_scale; // 1255-1256
            alpha[i] = (float)(s * (4.2e-2 - 1.6e-2) + 1.6e-2) * alpha_scale;               // 1309-1310
        }
    }
    // rock-strength Fbm<Perlin> tables (mod.rs:828-837): one permutation per octave
    for (int o = 0; o < 10; ++o) {
        uint8_t* p = perm + o * 256;
        for (int i = 0; i < 256; ++i) p[i] = (uint8_t)i;
        for (int i = 255; i > 0; --i) {
            uint64_t r = mix64(seed * 0x100000001b3ULL + (uint64_t)o * 1000003ULL + (uint64_t)i);
            int j = (int)(r % (uint64_t)(i + 1));
            std::swap(p[i], p[j]);
        }
    }
    const double continent_scale = scale * 5000.0 / 32.0 * 32.0;               // 776-779
    const double persistence = vm_pow(2.0, -0.75);                             // 833
    double denom = 0.0, pp = 1.0;
    for (int o = 1; o <= 10; ++o) { pp *= persistence; denom += pp; }
    scalars[0] = k_d_scale;
    scalars[1] = 2.5 * vm_pow(grid_scale, -2.0 * 0.2);                          // erosion.rs:750, mod.rs:863
    scalars[2] = max_erosion_per_delta_t;
    scalars[3] = 1.0 * (5000.0 / continent_scale) / (2.0 * 32.0 * 512.0);       // 834-837
    scalars[4] = 2.0;
    scalars[5] = persistence;
    scalars[6] = 1.0 / denom;
    scalars[7] = (double)land.size();
    return 0;
}

} // extern "C"
