/* erosion_oracle.h — C interface of the CPU oracle (TEST INFRASTRUCTURE, not product).
 *
 * The oracle is a line-by-line C++17 restatement of the reference's erosion path
 * (world/src/sim/erosion.rs, diffusion.rs, util.rs; citations beside each function in
 * erosion_oracle.cpp).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  PARITY UNPINNED: the reference has no golden vectors
 * or tests for this path and cannot be compiled here (no Rust toolchain), see DESIGN.md.
 */
#ifndef EROSION_ORACLE_H
#define EROSION_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_ctx orc_ctx;

typedef struct {
    const float* oldh; const float* oldb; const uint8_t* is_ocean;
    const float* uplift; const float* n; const float* m; const float* g;
    const float* epsilon_0; const float* alpha;
    const double* kf; const double* kd; const double* height_scale_cell;
} orc_inputs;

typedef struct {
    double k_d_scale;          /* kdsed = 1.5e-2/4 * k_d_scale (erosion.rs:2635-2636) */
    double k_da;               /* 2.5 * k_da_scale(q) (erosion.rs:750) */
    double chunk_area_pow_q;   /* 1024^0.2 (erosion.rs:926) */
    double min_max_angle, max_max_angle, mid_slope; /* tan 15/60/30 deg (556-557, 863) */
    double mountain_scale;     /* CONFIG.mountain_scale = 2048 (config.rs:62) */
    int use_libm;              /* deviation study, bit mask: 1 = f64 pow/exp/ln/ln_1p/tanh from glibc instead of
                                  ve_exact_math.h; 2 = the f32 powf(dh, 0.5) of the Newton loop from glibc's powf
                                  instead of sqrtf (erosion.rs:1343-1347) */
    int n_threads;             /* 0: omp default */
} orc_params;

typedef struct {
    uint32_t octaves; double frequency, lacunarity, persistence, scale_factor;
    const uint8_t* perm;       /* octaves * 256 */
} orc_noise;

enum orc_array {
    ORC_H = 0, ORC_B, ORC_WH, ORC_DH, ORC_INDIRECTION, ORC_NEWH, ORC_MREC, ORC_MSTACK,
    ORC_MWREC, ORC_AREA, ORC_MAX_SLOPE, ORC_KFS, ORC_KDF, ORC_H_SPL, ORC_LAKE_SILT,
    ORC_LAKE_SILL, ORC_LAKE_WATER_VOLUME, ORC_KD_EFF, ORC_DH0, ORC_WH_MREC, ORC_B_SPL,
    ORC_H_PRE_DIFF, ORC_ARRAY_COUNT
};

void orc_default_params(orc_params* p);
orc_ctx* orc_create(uint32_t x_lg, uint32_t y_lg);
void orc_destroy(orc_ctx*);
int orc_set_inputs(orc_ctx*, const orc_inputs*);
void orc_set_params(orc_ctx*, const orc_params*);
void orc_set_noise(orc_ctx*, const orc_noise*);
/* NoiseFn::get for the fBm described by nz at n points (xyz interleaved): what a host-side noise callback returns */
void orc_noise_eval(const orc_noise* nz, const double* xyz, double* out, size_t n);
/* overwrite the evolving state (h, b, wh) — used to start a step from a given state */
void orc_set_state(orc_ctx*, const double* h, const double* b, const double* wh);
/* one erode() call (erosion.rs:699); returns 0 or a negative code mirroring the panics */
int orc_step(orc_ctx*);
/* do_erosion (erosion.rs:2524): n_steps erode() calls, results copied out */
int orc_run(orc_ctx*, uint32_t n_steps, double* out_h, double* out_b);
/* element size in bytes of array `which`, and copy-out (count = N or 8N) */
size_t orc_array_bytes(orc_ctx*, int which);
int orc_get_array(orc_ctx*, int which, void* dst);
/* scalars of the last step: 0 n_gs, 1 err, 2 maxh, 3 boundary_len, 4 n_lakes (dh==-1 roots),
 * 5 dag_levels (longest multi-receiver path), 6 n_underwater, 7 n_sources */
double orc_get_scalar(orc_ctx*, int which);
/* per-stage seconds of the last step, names via orc_timer_name; returns count */
int orc_get_timers(orc_ctx*, double* dst, int cap);
const char* orc_timer_name(int i);

/* standalone pieces (post-erosion hydrology, mod.rs:1474-1550) */
void orc_downhill(uint32_t x_lg, uint32_t y_lg, const double* h, const uint8_t* ocean, int32_t* dh);
int orc_get_lakes(uint32_t x_lg, uint32_t y_lg, const double* h, int32_t* dh /*in-out*/,
                  int32_t* indirection, uint32_t* newh, double* maxh, uint64_t* boundary_len);
void orc_get_multi_rec(uint32_t x_lg, uint32_t y_lg, const double* h, const int32_t* dh,
                       const uint32_t* newh, double* wh /*in-out*/, uint8_t* mrec,
                       uint32_t* mstack, double* mwrec, int use_libm);
void orc_get_multi_drainage(uint32_t x_lg, uint32_t y_lg, const uint32_t* mstack,
                            const uint8_t* mrec, const double* mwrec, double* area);
void orc_get_drainage(uint32_t x_lg, uint32_t y_lg, const uint32_t* newh, const int32_t* dh,
                      float* flux);
/* literal Planchon-Darboux sweeps (erosion.rs:1738-1811); returns number of sweeps */
int orc_fill_sinks(uint32_t x_lg, uint32_t y_lg, const double* h, const uint8_t* ocean,
                   double* out);
void orc_diffusion(uint32_t nx, uint32_t ny, double xl, double yl, double dt, double* h,
                   double* b, const double* kd, double kdsed);
void orc_get_oceans(uint32_t x_lg, uint32_t y_lg, const double* h, uint8_t* is_ocean);
void orc_water_alt(uint32_t x_lg, uint32_t y_lg, const double* alt, const uint8_t* is_ocean, const int32_t* dh,
                   const int32_t* indirection, double* water_alt_filled, float* water_alt_final);
/* RiverData (erosion.rs:153-185) as a flat record; kind 0 None, 1 Ocean, 2 Lake, 3 River */
typedef struct { float velocity[3]; float spline_derivative[2]; int32_t kind; uint32_t aux[2]; } orc_river;
int orc_get_rivers(uint32_t x_lg, uint32_t y_lg, double continent_scale_hack, const uint32_t* newh, const double* water_alt,
                   const int32_t* dh, const int32_t* indirection, const double* drainage, orc_river* rivers, int use_libm);
/* exact-math probes for tests */
double orc_vm(int fn, double x, double y);

#ifdef __cplusplus
}
#endif
#endif
