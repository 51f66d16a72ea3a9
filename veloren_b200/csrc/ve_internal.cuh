// ve_internal.cuh — context layout and helpers shared by the libvelerode translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <map>
#include <string>
#include <vector>
#include "../../include/velerode.h"
#include "../../include/ve_exact_math.h"

// common/src/terrain/mod.rs:421-430 — slot k <-> 7-k are opposite directions.
__device__ __constant__ static const int c_ndx[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
__device__ __constant__ static const int c_ndy[8] = {-1, -1, -1, 0, 0, 1, 1, 1};

struct Geom {
    uint32_t xlg, ylg;
    int nx, ny;
    uint32_t n;
    __host__ __device__ __forceinline__ int x_of(uint32_t i) const { return (int)(i & ((1u << xlg) - 1u)); }
    __host__ __device__ __forceinline__ int y_of(uint32_t i) const { return (int)(i >> xlg); }
    __host__ __device__ __forceinline__ uint32_t idx(int x, int y) const { return ((uint32_t)y << xlg) | (uint32_t)x; }
    __host__ __device__ __forceinline__ bool inb(int x, int y) const { return x >= 0 && y >= 0 && x < nx && y < ny; }
};

enum VeTimer {
    VT_DOWNHILL, VT_LAKES, VT_WHFILL, VT_MREC, VT_STACK, VT_AREA, VT_SLOPE, VT_KFACT, VT_GS_DOWN, VT_GS_UP, VT_GS_LWV,
    VT_BASEMENT, VT_THERMAL, VT_DIFFUSION, VT_TOTAL, VT_COUNT
};

// Multi-GPU (row bands): cell i belongs to rank i >> shift; p[r] is rank r's copy of the array (its own band is
// the authoritative part).  With one GPU every entry is the local array and shift = 31.
template <class T> struct Remote { T* p[8]; uint32_t shift; };

struct NoiseDev {
    uint32_t octaves;
    double frequency, lacunarity, persistence, scale_factor;
    uint8_t* perm; // device, octaves*256
};

// Every per-cell array lives in HBM for the lifetime of the context (DESIGN.md "data layout").
struct ve_ctx {
    Geom g;
    int device;
    cudaStream_t stream;
    ve_params P;
    NoiseDev nz;
    ve_noise_cb noise_cb; void* noise_user;     // host noise callback (ve_set_noise_cb), NULL = device fBm
    double* noise_host;                          // pinned: 3N points out, N values back (allocated on first use)
    std::string err;
    bool timing;
    uint64_t launches;
    double timers[VT_COUNT];
    int sm_count;
    int df_grid; // resident grid size for the dataflow sweeps

    // inputs
    float *uplift, *nf, *mf, *gf, *eps0, *alpha;
    double *kf, *kd, *hscale;
    uint8_t* ocean;
    float max_uplift, max_g;
    // evolving state
    double *h, *b, *wh;
    // step intermediates (grid-indexed unless noted)
    int32_t *dh, *dh0, *indirection;
    uint32_t *newh, *mstack, *stackpos;
    uint8_t *mrec, *ndon;
    double *mwrec;            // [cell][8]
    double *area, *max_slope, *kd_eff, *h_t, *h_new, *deltah, *elev, *h_p, *lake_silt, *lwv, *wh_prev, *h_spl, *b_spl,
        *wh_mrec, *h_pre_diff;
    float *kfs, *kdf;         // [cell][8]
    int32_t *sill_cur, *sill_prev;
    uint32_t *order_td;       // topological order, sources first (wh desc, stack position asc)
    uint32_t *level;          // longest path to a sink, per node
    uint32_t lanes_down, lanes_up, lanes_other;   // active lanes per warp in the sweeps
    // test / tuning knobs (ve_set_debug); all zero by default
    int dbg_df_bps;           // resident blocks per SM for the sweeps (0: occupancy)
    int bps_up, bps_down;     // resident blocks per SM of the two Gauss-Seidel sweeps (0: occupancy)
    int dbg_newton;           // bit 0: skip the Newton loop, bit 2: force the slow re-solve of every node
    int dbg_lwv_small_log;    // replay the lake_water_volume log before every Gauss-Seidel iteration
    int dbg_thermal_sweep;    // ordered sweep instead of pointer jumping for the thermal water level
    int dbg_lakes_trace;      // print per-phase times of get_lakes to stderr
    int dbg_lakes_serial;     // 1: grow the lake forest with the one-warp heap (round 1) instead of the parallel formulation
    uint32_t lake_rounds;     // fixed-point rounds of the last parallel lake growth (0: serial path was used)
    int dbg_newton_variant;   // Newton loop variant of the up-sweep (0: default)
    int dbg_poll_sleep_ns;    // pause between poll rounds of the warp-per-node sweeps (0: none)
    int dbg_serial_err;       // 1: always recompute the Gauss-Seidel error in the reference's serial order
    uint64_t spin_limit_ns;   // give-up time of a polling worker (0: default scaled with the map size)
    std::map<const void*, std::pair<int, size_t>> occ_cache;   // kernel -> (resident blocks per SM, dynamic smem)
    uint32_t *order_bu;       // nodes sorted by level (longest path to a sink): sinks first
    uint32_t *don_order;      // packed donor slots sorted by stack position
    uint32_t *don_live;       // the same without egress donors (dh < 0), for the sediment sweep
    int dbg_level_order;      // 1: the bottom-up level sweep always walks the wh-sorted stack (the path taken with tie edges)
    int dbg_wh_full_sweep;    // 1: the water-surface sweep walks every cell (no pointer-jumping pre-pass)
    uint32_t n_wh_undecided;  // cells the last water-surface sweep had to walk
    int dbg_rank_launches;    // 1: one kernel launch per pointer-jumping round (else one cooperative launch for all rounds)
    int dbg_no_walk_cache;    // 1: the parent sweep never reuses the answer of a pointer walk
    int dbg_rank_jump;        // 1: list ranking always by pointer jumping over all arcs
    uint32_t *order_q;        // level order padded to quads (four nodes per warp in the down-sweep) + level tables
    uint32_t order_q_cap; uint32_t *order_q_n;   // capacity in entries; padded length on the device (0: unusable)
    int dbg_quad;             // four nodes per warp in the down-sweep: 0 = by map size, 1 = always, -1 = never
    uint32_t *flags;          // dataflow completion epochs
    uint32_t epoch;
    uint32_t* abort_word;     // device word raised by a stalled dataflow sweep
    int df_pending;
    // forest scratch (mstack / lakes)
    uint32_t *f_parent, *f_jump, *f_depth, *f_root;
    // generic scratch
    void* scratch[12];
    size_t scratch_bytes[12];
    void* cub_tmp; size_t cub_tmp_bytes;
    double* d_scalars;        // small device scalar block
    double* h_scalars;        // pinned mirror
    // CUDA events around every up-sweep launch (the dominant kernel), read back at the end of the solve: live
    // per-launch device time without extra synchronisation (bench.py's roofline line)
    cudaEvent_t up_ev[200];
    bool up_ev_ready;
    double up_ms_total; uint64_t up_launches_total;
    // per-step history since ve_create (ring of the last 256 steps): Gauss-Seidel iterations and lake counts
    int hist_n_gs[256]; uint32_t hist_n_lakes[256]; uint64_t hist_count;
    // row-band sharding of the Gauss-Seidel sweeps over several GPUs (ve_mg_attach); world == 1: off
    int mg_rank, mg_world; uint32_t mg_shift;
    ve_allgather_cb mg_cb; void* mg_user;
    std::map<const void*, std::vector<void*>> mg_peers;     // local base pointer -> every rank's base pointer
    uint8_t* mg_uw_flag; double* mg_uw_val; double* mg_errsq; // sweep outputs that live in scratch on one GPU
    uint32_t* mg_order_own;                                   // order_bu restricted to this rank's band (n / world entries)
    // last-step scalars
    int n_gs; double gs_err; double maxh; uint64_t boundary_len; uint32_t n_lakes; uint32_t n_underwater;
};

#define VE_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            char buf_[512];                                                                    \
            snprintf(buf_, sizeof(buf_), "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            ctx->err = buf_;                                                                   \
            return VE_ERR_CUDA;                                                                \
        }                                                                                      \
    } while (0)

#define VE_TRY(call)                   \
    do {                               \
        int rc_ = (call);              \
        if (rc_ != VE_OK) return rc_;  \
    } while (0)

#define VE_LAUNCH_CHECK()                                   \
    do {                                                    \
        ctx->launches++;                                    \
        VE_CUDA(cudaGetLastError());                        \
    } while (0)

static inline unsigned ve_blocks(uint64_t n, unsigned bs) { return (unsigned)((n + bs - 1) / bs); }

// resident blocks per SM of `func` at this block size / dynamic shared memory, cached per context; the
// >48 KB shared-memory opt-in is made per device
// multi-GPU helpers (ve_api.cu): every rank's pointer of a shared array, host barrier, band gather
template <class T> static inline Remote<T> ve_remote(const ve_ctx* ctx, T* local) {
    Remote<T> r;
    for (int i = 0; i < 8; ++i) r.p[i] = local;
    r.shift = 31u;
    if (ctx->mg_world > 1) {
        auto it = ctx->mg_peers.find((const void*)local);
        if (it != ctx->mg_peers.end()) { for (int i = 0; i < ctx->mg_world; ++i) r.p[i] = (T*)it->second[i]; r.shift = ctx->mg_shift; }
    }
    return r;
}
int ve_mg_barrier(ve_ctx* ctx);                                   // stream sync + all ranks arrive
int ve_mg_gather(ve_ctx* ctx, void* local, size_t elem_bytes);    // copy the other ranks' bands into the local replica
int ve_occupancy(ve_ctx* ctx, const void* func, int block, size_t smem, int* bps_out);
uint64_t ve_spin_limit_ns(const ve_ctx* ctx);
// scratch buffer i with at least `bytes` bytes (grown on demand; contents undefined)
int ve_scratch(ve_ctx* ctx, int i, size_t bytes, void** out);
int ve_cub_tmp(ve_ctx* ctx, size_t bytes, void** out);
// synchronise and turn a raised abort word into VE_ERR_STACK_INCOMPLETE
int ve_check_abort(ve_ctx* ctx);
// small device-to-host reads (counts, flags) through the context's pinned block: copy, wait for the stream, hand over
int ve_d2h(ve_ctx* ctx, void* dst, const void* src, size_t bytes);
int ve_d2h2(ve_ctx* ctx, void* dst0, const void* src0, size_t bytes0, void* dst1, const void* src1, size_t bytes1);

// ---- stage entry points (one per translation unit) ------------------------------------------
int ve_k_f32_to_f64(ve_ctx* ctx, const float* a, double* o);      // ve_api.cu
// ve_stencil.cu
int ve_k_downhill(ve_ctx* ctx, const double* h, const uint8_t* ocean, int use_h_le0, int32_t* dh);
int ve_k_mrec(ve_ctx* ctx);                 // mrec, ndon, mwrec from ctx->wh
int ve_k_max_slope(ve_ctx* ctx);
int ve_k_kfactors(ve_ctx* ctx);
int ve_k_basement(ve_ctx* ctx);             // scatter is implicit (grid-indexed), basement + lake silt
int ve_k_thermal_kd(ve_ctx* ctx);
int ve_k_diffusion(ve_ctx* ctx, double dt, double* h, double* b, const double* kd_eff);
int ve_k_erosion_inputs(ve_ctx* ctx, double gen_scale, const float* d_alt_old, const uint8_t* d_ocean, const double* d_uheight,
                        const double* d_rock, float* d_cdf, float* d_tmp_f32, float* d_oldh_out, double* max_erosion_out);
// ve_graph.cu
int ve_g_wh_fill(ve_ctx* ctx);              // erosion.rs:2383-2400 along newh
int ve_g_stack(ve_ctx* ctx);                // exact mstack + stackpos + order_td + don_order
int ve_g_area(ve_ctx* ctx);
int ve_g_area_from_stack(ve_ctx* ctx);      // area from ctx->mstack/mrec/mwrec alone (standalone get_multi_drainage)
int ve_g_spl_solve(ve_ctx* ctx);            // Gauss-Seidel loop (erosion.rs:1076-1467)
int ve_g_thermal_wh(ve_ctx* ctx);           // wh running max along newh (erosion.rs:1620, 1685)
int ve_g_get_drainage(ve_ctx* ctx, const uint32_t* newh, const int32_t* dh, float* flux);
int ve_g_fill_sinks(ve_ctx* ctx, const double* h, const uint8_t* ocean, double* out);
int ve_g_get_oceans(ve_ctx* ctx, const double* h, uint8_t* oc);
int ve_g_uniform_noise(ve_ctx* ctx, const float* values, float* out, uint32_t* count_out);
int ve_g_horizon(ve_ctx* ctx, float minh, float maxh, const float* h, float* a_w, float* h_w, float* a_e, float* h_e);
int ve_g_get_rivers(ve_ctx* ctx, double scale, const uint32_t* newh, const double* water_alt, const int32_t* dh, const int32_t* ind,
                    const double* drainage, void* out_dev);
int ve_g_water_alt(ve_ctx* ctx, const double* alt, const uint8_t* ocean, const int32_t* dh, const int32_t* ind,
                   double* init_scratch, double* filled, float* final_out);
// list ranking over a forest given by successor arcs (shared by stack and lakes)
int ve_list_rank(ve_ctx* ctx, uint32_t n_arcs, uint32_t* succ /*in, destroyed*/, int32_t* w0, int32_t* w1);
// ve_lakes.cu
int ve_l_get_lakes(ve_ctx* ctx, const double* h, int32_t* dh, int32_t* indirection, uint32_t* newh);

// cp.async staging (global -> shared, 4 / 8 / 16 bytes), one commit group per pipeline step
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cpa4(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(s)), "l"(g) : "memory"); }
__device__ __forceinline__ void cpa8(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(s)), "l"(g) : "memory"); }
__device__ __forceinline__ void cpa16(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(s)), "l"(g) : "memory"); }
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cpa_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// neighbour slot -> delta without a table lookup (lane-varying slot): slots 0..7 = the 3x3 ring in row order
__device__ __forceinline__ void slot_delta(int k, int& dx, int& dy) { int kk = k + (k >= 4 ? 1 : 0); dy = kk / 3 - 1; dx = kk - (dy + 1) * 3 - 1; }
