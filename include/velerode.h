/* velerode.h — C ABI of libvelerode.so: the B200 (sm_100a) implementation of Veloren's
 * world-map erosion pass.
 *
 * The reference has no FFI boundary for this path; the boundary it offers is the Rust function
 *   pub fn do_erosion(map_size_lg, _max_uplift, n_steps, seed, rock_strength_nz, oldh, oldb,
 *                     is_ocean, uplift, n, theta, kf, kd, g, epsilon_0, alpha, height_scale,
 *                     k_d_scale, k_da_scale, threadpool, report_progress) -> (Box<[f64]>, Box<[f64]>)
 * at world/src/sim/erosion.rs:2524-2547, called from WorldSim::generate
 * (world/src/sim/mod.rs:1370, 1399, 1449).  A Rust shim with that exact signature evaluates
 * each closure once into a flat array (what erosion.rs:2549-2609 already does) and calls the
 * entry points below; INTEGRATION.md shows the shim.  All pointers are HOST pointers unless a
 * comment says otherwise; the library copies what it needs and never keeps caller memory.
 * Every function returns 0 on success or a negative ve_status; ve_last_error() gives text.
 * There is no CPU fallback: without a CUDA device ve_create fails with VE_ERR_CUDA.
 */
#ifndef VELERODE_H
#define VELERODE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ve_ctx ve_ctx;

typedef enum {
    VE_OK = 0,
    VE_ERR_CUDA = -1,              /* any CUDA runtime failure (text in ve_last_error) */
    VE_ERR_INVALID = -2,           /* bad argument / call order */
    VE_ERR_DISCONNECTED_LAKE = -3, /* panic!("Disconnected lake!") erosion.rs:1251, 1617 */
    VE_ERR_TRIDAG_SINGULAR = -4,   /* assert_ne!(bet, 0.0) diffusion.rs:423, 445 */
    VE_ERR_STACK_INCOMPLETE = -5,  /* assert_eq!(stack.len(), nn) erosion.rs:2510; newh 1903/2291 */
    VE_ERR_LAKE_GRAPH = -6,        /* assert!/assert_ne! inside get_lakes (2089, 2174, 2209, 2262) */
    VE_ERR_TOO_LARGE = -7
} ve_status;

/* Per-cell inputs, N = 2^x_lg * 2^y_lg entries each, index = y * nx + x
 * (common/src/terrain/mod.rs:407-418).  Exactly the arrays do_erosion materialises at
 * erosion.rs:2549-2609 plus height_scale(n[i]) (erosion.rs:848-850). */
typedef struct {
    const float* oldh;              /* oldh(posi)                      erosion.rs:2549 */
    const float* oldb;              /* oldb(posi)                      erosion.rs:2555 */
    const uint8_t* is_ocean;        /* is_ocean(posi)                  erosion.rs:818  */
    const float* uplift;            /* uplift(posi) as f32             erosion.rs:2605 */
    const float* n;                 /* n(posi)                         erosion.rs:2561 */
    const float* m;                 /* theta(posi) * n(posi), in f32   erosion.rs:2568 */
    const float* g;                 /*                                 erosion.rs:2586 */
    const float* epsilon_0;         /*                                 erosion.rs:2591 */
    const float* alpha;             /*                                 erosion.rs:2596 */
    const double* kf;               /*                                 erosion.rs:2574 */
    const double* kd;               /*                                 erosion.rs:2580 */
    const double* height_scale_cell;/* height_scale(n(posi))           erosion.rs:848  */
} ve_inputs;

/* Scalars the Rust host evaluates with its own libm so they carry the host's rounding. */
typedef struct {
    double k_d_scale;          /* do_erosion argument; kdsed = 1.5e-2/4*k_d_scale (2635) */
    double k_da;               /* 2.5 * k_da_scale(0.2)                  erosion.rs:745-750 */
    double chunk_area_pow_q;   /* (32*32 as f64).powf(0.2)               erosion.rs:926 */
    double min_max_angle;      /* tan(15 deg)                            erosion.rs:556 */
    double max_max_angle;      /* tan(60 deg)                            erosion.rs:557 */
    double mid_slope;          /* tan(30 deg)                            erosion.rs:863 */
    double mountain_scale;     /* CONFIG.mountain_scale = 2048           world/src/config.rs:62 */
} ve_params;

/* rock_strength_nz as tables: noise 0.9.0 Fbm<Perlin> (world/src/sim/mod.rs:828-837).
 * perm holds `octaves` permutation tables of 256 bytes (Perlin::new(seed + i)). */
typedef struct {
    uint32_t octaves;
    double frequency, lacunarity, persistence, scale_factor;
    const uint8_t* perm;
} ve_noise_fbm_perlin3;

/* Arrays of the last step that ve_get_array can download (test hook). */
typedef enum {
    VE_H = 0, VE_B, VE_WH, VE_DH, VE_INDIRECTION, VE_NEWH, VE_MREC, VE_MSTACK, VE_MWREC, VE_AREA,
    VE_MAX_SLOPE, VE_KFS, VE_KDF, VE_H_SPL, VE_LAKE_SILT, VE_LAKE_SILL, VE_LAKE_WATER_VOLUME,
    VE_KD_EFF, VE_DH0, VE_WH_MREC, VE_B_SPL, VE_H_PRE_DIFF, VE_ARRAY_COUNT
} ve_array;

/* Scalars of the last step (ve_get_scalar). */
typedef enum {
    VE_S_N_GS = 0, VE_S_ERR, VE_S_MAXH, VE_S_BOUNDARY_LEN, VE_S_N_LAKES, VE_S_MAX_UPLIFT, VE_S_MAX_G,
    VE_S_N_UNDERWATER,
    VE_S_UP_MS_TOTAL,        /* device milliseconds spent in up-sweep launches since ve_create (CUDA events) */
    VE_S_UP_LAUNCHES_TOTAL,  /* number of up-sweep launches since ve_create */
    VE_S_LAKE_ROUNDS,        /* fixed-point rounds of the last parallel lake growth (0: the serial path ran) */
    VE_S_WH_UNDECIDED        /* cells the last water-surface sweep (erosion.rs:2383-2400) had to walk in order */
} ve_scalar;

void ve_default_params(ve_params* p);
/* One context per map and per calling thread (batch_generate runs several worlds on
 * independent threads, world/examples/batch_generate.rs:287-352).  device_ids/n_devices:
 * GPUs to use; the context lives on device_ids[0] (NULL -> current device); one map is not split across
 * GPUs, several contexts on several devices may be used from one process. */
int ve_create(ve_ctx** out, uint32_t x_lg, uint32_t y_lg, const int* device_ids, int n_devices);
void ve_destroy(ve_ctx* ctx);
const char* ve_last_error(ve_ctx* ctx);

/* Row-band sharding of ONE map over several GPUs, one process (rank) per GPU (no counterpart in the
 * single-threaded reference).  Every rank creates its own context for the same map on its own device, loads the
 * same inputs, then calls ve_mg_attach with an all-gather supplied by the host (MPI, torch.distributed, ...):
 * cb(user, send, recv, bytes) must deliver every rank's `bytes` bytes, in rank order, into recv on every rank;
 * bytes == 0 is a plain barrier.  The library uses it once to exchange CUDA IPC handles and afterwards as the
 * barrier between sweeps.  From then on ve_step / ve_run* must be called collectively (same calls, same order) on
 * all ranks; every rank ends every step with the complete, identical state.  Sharded today: the Gauss-Seidel
 * sweeps (erosion.rs:1127-1151, 1237-1450): rank r solves the nodes of rows [r*ny/world, (r+1)*ny/world) and reads the
 * values of nodes in other bands directly from their owner's memory over NVLink; every other stage runs
 * replicated.  world must be a power of two <= 8. */
typedef void (*ve_allgather_cb)(void* user, const void* send, void* recv, size_t bytes);
int ve_mg_attach(ve_ctx* ctx, int rank, int world, ve_allgather_cb cb, void* user);

/* Test and tuning knobs (no counterpart in the reference; defaults are what production uses).  Keys:
 * "lanes_up" / "lanes_down" / "lanes_other" (k > 1: thread-per-node sweeps), "df_bps" (resident blocks per SM
 * of the sweeps), "newton" (bit 0: skip the Newton loop, bit 2: force the slow re-solve), "newton_variant",
 * "lwv_small_log" (1: replay, 2: compact the lake_water_volume log before every iteration), "thermal_sweep",
 * "lakes_trace", "lakes_serial" (one-warp heap growth of the lake forest), "poll_sleep_ns", "serial_err" (always
 * sum the Gauss-Seidel error in the reference's serial order), "spin_limit_ms" (give-up time of a stalled sweep),
 * "bps_up" / "bps_down" (resident blocks per SM of the two Gauss-Seidel sweeps), "level_order", "rank_jump",
 * "rank_launches" (one launch per pointer-jumping round instead of one cooperative launch), "quad" (four nodes per
 * warp in the sediment sweep: 1 always, -1 never, 0 from 2048x2048 cells on), "wh_full_sweep" (the water-surface
 * sweep walks every cell), "no_walk_cache" (the stack build never reuses the answer of a pointer walk). */
int ve_set_debug(ve_ctx* ctx, const char* key, long long value);
int ve_set_params(ve_ctx* ctx, const ve_params* p);
int ve_set_noise(ve_ctx* ctx, const ve_noise_fbm_perlin3* nz);
/* Alternative rock-strength source for get_max_slope (erosion.rs:550-604): the host's own
 * `rock_strength_nz: impl NoiseFn<f64, 3>` (erosion.rs:2529), for exact parity with an arbitrary
 * noise-rs function.  Once per erode() the library hands the callback the N sample points
 * (xyz interleaved, the argument of `.get()` at erosion.rs:585-589) and reads N values back; the
 * callback runs on the calling thread.  Costs one 24N-byte download and one 8N-byte upload per step.
 * cb == NULL returns to the table set with ve_set_noise. */
typedef void (*ve_noise_cb)(void* user, const double* xyz, double* out, size_t n);
int ve_set_noise_cb(ve_ctx* ctx, ve_noise_cb cb, void* user);
/* Uploads the inputs and initialises h = oldh as f64, b = oldb as f64, wh = 0
 * (erosion.rs:2549-2559, 2601); computes max_uplift and max_g (2617-2626). */
int ve_set_inputs(ve_ctx* ctx, const ve_inputs* in);
/* SURVEY 8(f) rank 3 — the parameter closures of WorldSim::generate on the device (world/src/sim/mod.rs:849-893 scale
 * factors; 1102-1337 n_func, theta_func, old_height/alt_func, uplift_fn with its uniform_noise rank CDF, kf_func, kd_func,
 * g_func, epsilon_0_func, alpha_func).  Replaces ve_set_inputs: the twelve input arrays are computed in place on the
 * device from what the host samples out of its noise generators (noise-rs is not part of the tree): alt_old[posi].1
 * (mod.rs:1078-1088), uheight = uplift_nz.get(turbed wposf).clamp(-1,1)*0.5+0.5 (1222-1228) and rock_strength =
 * rock_strength_nz.get(wposf3).clamp(-1,1)*0.5+0.5 (1234-1240), one value per chunk.  is_ocean == NULL: get_oceans on
 * alt_old (mod.rs:1092).  gen_scale = GenOpts.scale (grid_scale = scale / 4, mod.rs:849).  `out` (may be NULL) receives
 * copies of the arrays (any member may be NULL) and max_erosion_per_delta_t. */
typedef struct {
    float* oldh; uint8_t* is_ocean; float* uplift; float* n; float* m; float* g; float* epsilon_0; float* alpha;
    double* kf; double* kd; double* height_scale_cell; double max_erosion_per_delta_t;
} ve_inputs_out;
int ve_set_inputs_from_fields(ve_ctx* ctx, double gen_scale, const float* alt_old, const uint8_t* is_ocean, const double* uheight,
                              const double* rock_strength, ve_inputs_out* out);
/* Replace the evolving state (any pointer may be NULL to keep it). */
int ve_set_state(ve_ctx* ctx, const double* h, const double* b, const double* wh);

/* do_erosion's loop (erosion.rs:2645-2681): n_steps erode() calls; `progress` is called on
 * the calling thread with the percentage at the reference's cadence (2648-2651).  out_h/out_b:
 * caller-owned, N doubles each. */
int ve_run(ve_ctx* ctx, uint32_t n_steps, void (*progress)(void* user, double pct), void* user,
           double* out_h, double* out_b);
/* One erode() (erosion.rs:699), state stays on the device. */
int ve_step(ve_ctx* ctx);
/* Device-resident variant used by the benchmark: n_steps erode() calls, no host copies. */
int ve_run_device(ve_ctx* ctx, uint32_t n_steps);
/* Benchmark hook: ve_run_device bracketed by CUDA events recorded on the context's own stream;
 * returns the elapsed device time in milliseconds. */
int ve_run_device_timed(ve_ctx* ctx, uint32_t n_steps, double* elapsed_ms);
int ve_get_array(ve_ctx* ctx, int which, void* dst);
size_t ve_array_bytes(ve_ctx* ctx, int which);
double ve_get_scalar(ve_ctx* ctx, int which);
/* number of kernels launched by this context so far, and seconds spent per stage of the last
 * step (names via ve_timer_name) when ve_set_timing(ctx, 1) is on (adds synchronisations). */
uint64_t ve_launch_count(ve_ctx* ctx);
/* Gauss-Seidel iteration count and lake count of the last `cap` erode() steps (oldest first); returns how
 * many entries were written (at most 256 are kept). */
int ve_get_history(ve_ctx* ctx, int32_t* n_gs, uint32_t* n_lakes, int cap);
int ve_set_timing(ve_ctx* ctx, int on);
int ve_get_timers(ve_ctx* ctx, double* dst, int cap);
const char* ve_timer_name(int i);

/* Post-erosion hydrology calls of WorldSim::generate (world/src/sim/mod.rs:1474-1550): the
 * same kernels, on caller-supplied host arrays.  The context only provides device scratch. */
int ve_downhill(ve_ctx* ctx, const double* h, const uint8_t* is_ocean, int32_t* dh);  /* util.rs:251 */
int ve_get_lakes(ve_ctx* ctx, const double* h, int32_t* dh /*in-out*/, int32_t* indirection,
                 uint32_t* newh, double* maxh, uint64_t* boundary_len);                /* erosion.rs:1825 */
int ve_get_multi_rec(ve_ctx* ctx, const double* h, const int32_t* dh, const uint32_t* newh,
                     double* wh /*in-out*/, uint8_t* mrec, uint32_t* mstack, double* mwrec); /* 2332 */
/* A pure function of its three inputs, like the reference's: the donor lists are rebuilt from mrec and the
 * positions in mstack, the area is gathered along mstack.  An mstack in which a receiver precedes one of its
 * donors (the reference would silently produce nonsense) returns VE_ERR_STACK_INCOMPLETE. */
int ve_get_multi_drainage(ve_ctx* ctx, const uint32_t* mstack, const uint8_t* mrec,
                          const double* mwrec, double* area);                          /* erosion.rs:62 */
int ve_fill_sinks(ve_ctx* ctx, const double* h, const uint8_t* is_ocean, double* out); /* erosion.rs:1738 */
int ve_get_drainage(ve_ctx* ctx, const uint32_t* newh, const int32_t* dh, float* flux); /* erosion.rs:29-58 */
int ve_diffusion(ve_ctx* ctx, double dt, double* h, double* b, const double* kd);       /* diffusion.rs:41 */
/* SURVEY 8(f) rank 1 — the rest of the post-erosion hydrology wiring. */
int ve_get_oceans(ve_ctx* ctx, const double* h, uint8_t* is_ocean);                     /* util.rs:332-367 */
/* water_height_initial (mod.rs:1502-1543) -> fill_sinks (1550) -> final water_alt select (1566-1592).
 * water_alt_filled (f64, may be NULL) is what get_rivers consumes; water_alt_final is the f32 field. */
int ve_water_alt(ve_ctx* ctx, const double* alt, const uint8_t* is_ocean, const int32_t* dh,
                 const int32_t* indirection, double* water_alt_filled, float* water_alt_final);

/* SURVEY 8(f) rank 2 — get_rivers (erosion.rs:219-545).  One record per chunk mirroring RiverData
 * (erosion.rs:153-185): kind 0 None, 1 Ocean, 2 Lake (aux = neighbor_pass_pos as two i32), 3 River
 * (aux = cross_section as two f32).  neighbor_rivers is never filled by the reference and is omitted.
 * newh = get_lakes' order, water_alt = ve_water_alt's filled field, drainage = ve_get_multi_drainage. */
typedef struct { float velocity[3]; float spline_derivative[2]; int32_t kind; uint32_t aux[2]; } ve_river;
int ve_get_rivers(ve_ctx* ctx, double continent_scale_hack, const uint32_t* newh, const double* water_alt,
                  const int32_t* dh, const int32_t* indirection, const double* drainage, ve_river* rivers);

/* SURVEY 8(f) rank 3/4 — pre- and post-processing around the path.
 * uniform_noise (util.rs:171-207): values[i] = the sample, NaN = None; cdf[i] = (1 + rank) / count as f32
 * with ties broken by index, 0 for None. */
int ve_uniform_noise(ve_ctx* ctx, const float* values, float* cdf);
/* map_edge_factor (util.rs:22-29) for every chunk, f32. */
int ve_map_edge_factor(ve_ctx* ctx, float* out);
/* get_horizon_map (util.rs:371-449) for bounds = the whole map, F = f32: raw slope ("angle" before
 * to_angle) and occluder height (before to_height) for the west and the east march. */
int ve_horizon_map(ve_ctx* ctx, float minh, float maxh, const float* h, float* west_angle, float* west_height,
                   float* east_angle, float* east_height);

/* No counterpart in the reference.  Same arithmetic as ve_diffusion, for hosts that may not
 * allocate device memory (a fixed GPU arena, or a call site where the CUDA allocator is off
 * limits): d_h, d_b, d_kd_eff are DEVICE pointers owned by the caller (N doubles each, h and b
 * updated in place) and the solver's work space is an array built into the library
 * (6 MiB per GPU, held until exit), so maps are limited to 512x512 cells; larger ones return
 * VE_ERR_TOO_LARGE with a message in errbuf.  Calls nothing that allocates (the one 4-byte status word it
 * reads back at the end goes through the driver's own staging buffer into a stack variable: a synchronous
 * copy, no allocation).  The first call on a
 * device looks the array's device address up: issue it before any stream capture.  All calls on a
 * device share the array, so calls that overlap on the GPU race; serialise them.  cuda_stream: a
 * cudaStream_t (NULL = default stream); the call returns after the work has finished. */
int ve_diffusion_static(uint32_t x_lg, uint32_t y_lg, double dt, double* d_h, double* d_b,
                        const double* d_kd_eff, void* cuda_stream, char* errbuf, size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif
