// ve_api.cu — C ABI of libvelerode.so (include/velerode.h) and the erode() step driver.
#include "ve_internal.cuh"
#include <math.h>
#include <string.h>
#include <chrono>
#include <stdlib.h>
#include <mutex>
#include <nvtx3/nvToolsExt.h>

static const char* TIMER_NAMES[VT_COUNT] = {"downhill", "get_lakes", "wh_fill", "mrec", "stack", "area", "max_slope",
                                            "k_factors", "gs_down", "gs_up", "gs_lwv", "basement", "thermal",
                                            "diffusion", "total"};

int ve_scratch(ve_ctx* ctx, int i, size_t bytes, void** out) {
    if (bytes < 256) bytes = 256;
    if (ctx->scratch_bytes[i] < bytes) {
        if (ctx->scratch[i]) { VE_CUDA(cudaStreamSynchronize(ctx->stream)); VE_CUDA(cudaFree(ctx->scratch[i])); ctx->scratch[i] = nullptr; ctx->scratch_bytes[i] = 0; }
        size_t want = bytes + bytes / 8;
        VE_CUDA(cudaMalloc(&ctx->scratch[i], want));
        ctx->scratch_bytes[i] = want;
    }
    *out = ctx->scratch[i];
    return VE_OK;
}
int ve_cub_tmp(ve_ctx* ctx, size_t bytes, void** out) {
    if (bytes < 256) bytes = 256;
    if (ctx->cub_tmp_bytes < bytes) {
        if (ctx->cub_tmp) { VE_CUDA(cudaStreamSynchronize(ctx->stream)); VE_CUDA(cudaFree(ctx->cub_tmp)); ctx->cub_tmp = nullptr; ctx->cub_tmp_bytes = 0; }
        VE_CUDA(cudaMalloc(&ctx->cub_tmp, bytes * 2));
        ctx->cub_tmp_bytes = bytes * 2;
    }
    *out = ctx->cub_tmp;
    return VE_OK;
}
// Function attributes are per device and the occupancy depends on them: the opt-in is made under a mutex
// for the device that is current, the occupancy is cached per context (contexts may live on different
// devices and run on different host threads).
static std::mutex g_attr_mu;
int ve_occupancy(ve_ctx* ctx, const void* func, int block, size_t smem, int* bps_out) {
    auto it = ctx->occ_cache.find(func);
    if (it != ctx->occ_cache.end() && it->second.second == smem) { *bps_out = it->second.first; return VE_OK; }
    if (smem > 48 * 1024) {
        std::lock_guard<std::mutex> lk(g_attr_mu);
        VE_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    int bps = 0;
    VE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, func, block, smem));
    if (bps < 1) { ctx->err = "kernel does not fit an SM"; return VE_ERR_CUDA; }
    ctx->occ_cache[func] = std::make_pair(bps, smem);
    *bps_out = bps;
    return VE_OK;
}
// A polling worker gives up after this long: a legitimately waiting node can wait for most of a sweep
// (~1 s at 8192^2), so the limit scales with the map.
uint64_t ve_spin_limit_ns(const ve_ctx* ctx) {
    if (ctx->spin_limit_ns) return ctx->spin_limit_ns;
    return 5000000000ull + (uint64_t)ctx->g.n * 200ull;
}

int ve_d2h2(ve_ctx* ctx, void* dst0, const void* src0, size_t bytes0, void* dst1, const void* src1, size_t bytes1) {
    // h_scalars[0..15] belong to the Gauss-Seidel loop and ve_check_abort; [16..63] (384 bytes) stage these reads
    char* stage = reinterpret_cast<char*>(ctx->h_scalars + 16);
    if (bytes0 > 192 || bytes1 > 192) { ctx->err = "ve_d2h: read too large for the pinned block"; return VE_ERR_INVALID; }
    VE_CUDA(cudaMemcpyAsync(stage, src0, bytes0, cudaMemcpyDeviceToHost, ctx->stream));
    if (bytes1) VE_CUDA(cudaMemcpyAsync(stage + 192, src1, bytes1, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    memcpy(dst0, stage, bytes0);
    if (bytes1) memcpy(dst1, stage + 192, bytes1);
    return VE_OK;
}
int ve_d2h(ve_ctx* ctx, void* dst, const void* src, size_t bytes) { return ve_d2h2(ctx, dst, src, bytes, nullptr, nullptr, 0); }

int ve_check_abort(ve_ctx* ctx) {
    if (!ctx->df_pending) return VE_OK;                  // no sweep launched since the last look: nothing can have stalled
    volatile uint32_t* ha = reinterpret_cast<volatile uint32_t*>(ctx->h_scalars + 8);    // pinned: a plain async copy
    VE_CUDA(cudaMemcpyAsync((void*)ha, ctx->abort_word, 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint32_t a = *ha;
    ctx->df_pending = 0;
    if (a) {
        cudaMemsetAsync(ctx->abort_word, 0, 4, ctx->stream);
        ctx->err = "dataflow sweep stalled: order is not topological (stack/newh incomplete); the context's step state is undefined, "
                   "reload it with ve_set_inputs / ve_set_state";
        return VE_ERR_STACK_INCOMPLETE;
    }
    return VE_OK;
}

template <class T> static int dev_alloc(ve_ctx* ctx, T** p, size_t count) {
    VE_CUDA(cudaMalloc((void**)p, count * sizeof(T)));
    VE_CUDA(cudaMemset(*p, 0, count * sizeof(T)));
    return VE_OK;
}

__global__ void k_f32_to_f64(uint32_t n, const float* __restrict__ a, double* __restrict__ o) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[i] = (double)a[i];
}

// lake_silt / lake_sill are kept per grid cell on the device; the reference indexes them by
// stack position, so the download permutes them (out[stackpos[i]] = a[i]) for comparison.
template <class T> __global__ void k_to_stack_index(uint32_t n, const T* __restrict__ a, const uint32_t* __restrict__ stackpos, T* __restrict__ o) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[stackpos[i]] = a[i];
}

// ---- row bands over several GPUs ----------------------------------------------------------------------
int ve_mg_barrier(ve_ctx* ctx) {
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->mg_world > 1) ctx->mg_cb(ctx->mg_user, nullptr, nullptr, 0);
    return VE_OK;
}
int ve_mg_gather(ve_ctx* ctx, void* local, size_t elem_bytes) {
    if (ctx->mg_world <= 1) return VE_OK;
    auto it = ctx->mg_peers.find((const void*)local);
    if (it == ctx->mg_peers.end()) { ctx->err = "ve_mg_gather: array is not shared"; return VE_ERR_INVALID; }
    const size_t band = ((size_t)ctx->g.n >> 0) / (size_t)ctx->mg_world * elem_bytes;
    for (int r = 0; r < ctx->mg_world; ++r) {
        if (r == ctx->mg_rank) continue;
        VE_CUDA(cudaMemcpyAsync((char*)local + band * r, (const char*)it->second[r] + band * r, band, cudaMemcpyDefault, ctx->stream));
    }
    return VE_OK;
}
static int mg_share(ve_ctx* ctx, void* local) {
    const int W = ctx->mg_world;
    cudaIpcMemHandle_t mine;
    VE_CUDA(cudaIpcGetMemHandle(&mine, local));
    std::vector<cudaIpcMemHandle_t> all(W);
    ctx->mg_cb(ctx->mg_user, &mine, all.data(), sizeof(cudaIpcMemHandle_t));
    std::vector<void*> ptrs(8, nullptr);
    for (int r = 0; r < W; ++r) {
        if (r == ctx->mg_rank) { ptrs[r] = local; continue; }
        void* p = nullptr;
        VE_CUDA(cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess));
        ptrs[r] = p;
    }
    ctx->mg_peers[(const void*)local] = ptrs;
    return VE_OK;
}
int ve_k_f32_to_f64(ve_ctx* ctx, const float* a, double* o) {
    k_f32_to_f64<<<ve_blocks(ctx->g.n, 256), 256, 0, ctx->stream>>>(ctx->g.n, a, o);
    VE_LAUNCH_CHECK();
    return VE_OK;
}

extern "C" {

void ve_default_params(ve_params* p) {
    const double PI = 3.14159265358979323846;
    p->k_d_scale = 0.5;
    p->k_da = 2.5 * pow(0.5, -2.0 * 0.2);
    p->chunk_area_pow_q = pow(1024.0, 0.2);
    p->min_max_angle = tan(15.0 / 360.0 * 2.0 * PI);
    p->max_max_angle = tan(60.0 / 360.0 * 2.0 * PI);
    p->mid_slope = tan(30.0 / 360.0 * 2.0 * PI);
    p->mountain_scale = 2048.0;
}

static int create_impl(ve_ctx* ctx, uint32_t x_lg, uint32_t y_lg, const int* device_ids, int n_devices) {
    int ndev = 0;
    VE_CUDA(cudaGetDeviceCount(&ndev));
    if (ndev < 1) { ctx->err = "no CUDA device"; return VE_ERR_CUDA; }
    int dev = 0;
    if (device_ids && n_devices > 0) dev = device_ids[0];
    else VE_CUDA(cudaGetDevice(&dev));
    VE_CUDA(cudaSetDevice(dev));
    ctx->device = dev;
    cudaDeviceProp prop;
    VE_CUDA(cudaGetDeviceProperties(&prop, dev));
    ctx->sm_count = prop.multiProcessorCount;
    int coop = 0;
    VE_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    if (!coop) { ctx->err = "device lacks cooperative launch"; return VE_ERR_CUDA; }
    VE_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    Geom& g = ctx->g;
    g.xlg = x_lg; g.ylg = y_lg; g.nx = 1 << x_lg; g.ny = 1 << y_lg; g.n = (uint32_t)g.nx * (uint32_t)g.ny;
    size_t n = g.n;
    VE_TRY(dev_alloc(ctx, &ctx->uplift, n)); VE_TRY(dev_alloc(ctx, &ctx->nf, n)); VE_TRY(dev_alloc(ctx, &ctx->mf, n));
    VE_TRY(dev_alloc(ctx, &ctx->gf, n)); VE_TRY(dev_alloc(ctx, &ctx->eps0, n)); VE_TRY(dev_alloc(ctx, &ctx->alpha, n));
    VE_TRY(dev_alloc(ctx, &ctx->kf, n)); VE_TRY(dev_alloc(ctx, &ctx->kd, n)); VE_TRY(dev_alloc(ctx, &ctx->hscale, n));
    VE_TRY(dev_alloc(ctx, &ctx->ocean, n));
    VE_TRY(dev_alloc(ctx, &ctx->h, n)); VE_TRY(dev_alloc(ctx, &ctx->b, n)); VE_TRY(dev_alloc(ctx, &ctx->wh, n));
    VE_TRY(dev_alloc(ctx, &ctx->dh, n)); VE_TRY(dev_alloc(ctx, &ctx->dh0, n)); VE_TRY(dev_alloc(ctx, &ctx->indirection, n));
    VE_TRY(dev_alloc(ctx, &ctx->newh, n)); VE_TRY(dev_alloc(ctx, &ctx->mstack, n)); VE_TRY(dev_alloc(ctx, &ctx->stackpos, n));
    VE_TRY(dev_alloc(ctx, &ctx->mrec, n)); VE_TRY(dev_alloc(ctx, &ctx->ndon, n));
    VE_TRY(dev_alloc(ctx, &ctx->mwrec, n * 8));
    VE_TRY(dev_alloc(ctx, &ctx->area, n)); VE_TRY(dev_alloc(ctx, &ctx->max_slope, n)); VE_TRY(dev_alloc(ctx, &ctx->kd_eff, n));
    VE_TRY(dev_alloc(ctx, &ctx->h_t, n)); VE_TRY(dev_alloc(ctx, &ctx->h_new, n)); VE_TRY(dev_alloc(ctx, &ctx->deltah, n));
    VE_TRY(dev_alloc(ctx, &ctx->elev, n)); VE_TRY(dev_alloc(ctx, &ctx->h_p, n)); VE_TRY(dev_alloc(ctx, &ctx->lake_silt, n));
    VE_TRY(dev_alloc(ctx, &ctx->lwv, n)); VE_TRY(dev_alloc(ctx, &ctx->wh_prev, n)); VE_TRY(dev_alloc(ctx, &ctx->h_spl, n));
    VE_TRY(dev_alloc(ctx, &ctx->b_spl, n)); VE_TRY(dev_alloc(ctx, &ctx->wh_mrec, n)); VE_TRY(dev_alloc(ctx, &ctx->h_pre_diff, n));
    VE_TRY(dev_alloc(ctx, &ctx->kfs, n * 8)); VE_TRY(dev_alloc(ctx, &ctx->kdf, n * 8));
    VE_TRY(dev_alloc(ctx, &ctx->sill_cur, n)); VE_TRY(dev_alloc(ctx, &ctx->sill_prev, n));
    VE_TRY(dev_alloc(ctx, &ctx->order_td, n)); VE_TRY(dev_alloc(ctx, &ctx->order_bu, n)); VE_TRY(dev_alloc(ctx, &ctx->level, n)); VE_TRY(dev_alloc(ctx, &ctx->don_order, n)); VE_TRY(dev_alloc(ctx, &ctx->don_live, n)); VE_TRY(dev_alloc(ctx, &ctx->flags, n));
    ctx->order_q_cap = (uint32_t)n + 4u * 65536u;
    VE_TRY(dev_alloc(ctx, &ctx->order_q, (size_t)ctx->order_q_cap + 2 * (65536 + 2) + 64));
    ctx->order_q_n = ctx->order_q + ctx->order_q_cap + 2 * (65536 + 2);
    VE_TRY(dev_alloc(ctx, &ctx->f_parent, n)); VE_TRY(dev_alloc(ctx, &ctx->f_jump, n)); VE_TRY(dev_alloc(ctx, &ctx->f_depth, n));
    VE_TRY(dev_alloc(ctx, &ctx->f_root, n));
    VE_TRY(dev_alloc(ctx, &ctx->abort_word, 64));
    VE_TRY(dev_alloc(ctx, &ctx->d_scalars, 64));
    VE_CUDA(cudaMallocHost((void**)&ctx->h_scalars, 64 * sizeof(double)));
    return VE_OK;
}

int ve_create(ve_ctx** out, uint32_t x_lg, uint32_t y_lg, const int* device_ids, int n_devices) {
    if (!out || x_lg < 2 || y_lg < 2 || x_lg + y_lg > 28) return VE_ERR_INVALID;
    ve_ctx* ctx = new ve_ctx();
    memset((void*)&ctx->g, 0, sizeof(Geom));
    ctx->stream = nullptr; ctx->timing = false; ctx->launches = 0; ctx->epoch = 0; ctx->df_pending = 0;
    ctx->max_uplift = 0; ctx->max_g = 0; ctx->n_gs = 0; ctx->gs_err = 0; ctx->maxh = 0; ctx->boundary_len = 0;
    ctx->n_lakes = 0; ctx->n_underwater = 0; ctx->cub_tmp = nullptr; ctx->cub_tmp_bytes = 0;
    memset(ctx->scratch, 0, sizeof(ctx->scratch)); memset(ctx->scratch_bytes, 0, sizeof(ctx->scratch_bytes));
    memset(ctx->timers, 0, sizeof(ctx->timers));
    memset(&ctx->nz, 0, sizeof(ctx->nz));
    ctx->noise_cb = nullptr; ctx->noise_user = nullptr; ctx->noise_host = nullptr;
    ve_default_params(&ctx->P);
    // Sweeps run one warp per node at every map size (measured: 1024^2 .. 8192^2; at 8192^2 the up-sweep
    // takes 0.53 s per step against 1.1 s with eight one-thread nodes per warp).  ve_set_debug("lanes_*", k > 1)
    // selects the thread-per-node kernels instead (kept for comparison and covered by a test).
    ctx->lanes_down = ctx->lanes_up = ctx->lanes_other = 1u;
    ctx->dbg_df_bps = 0; ctx->dbg_newton = 0; ctx->dbg_lwv_small_log = 0; ctx->dbg_thermal_sweep = 0; ctx->dbg_lakes_trace = 0;
    ctx->dbg_lakes_serial = 0; ctx->lake_rounds = 0; ctx->order_q = nullptr; ctx->order_q_n = nullptr; ctx->order_q_cap = 0; ctx->dbg_quad = 0; ctx->dbg_rank_jump = 0; ctx->dbg_no_walk_cache = 0; ctx->dbg_rank_launches = 0; ctx->dbg_wh_full_sweep = 0; ctx->n_wh_undecided = 0; ctx->dbg_level_order = 0; ctx->bps_up = 0; ctx->bps_down = 0;
    ctx->mg_rank = 0; ctx->mg_world = 1; ctx->mg_shift = 31u; ctx->mg_cb = nullptr; ctx->mg_user = nullptr;
    ctx->mg_uw_flag = nullptr; ctx->mg_uw_val = nullptr; ctx->mg_errsq = nullptr; ctx->mg_order_own = nullptr;
    ctx->dbg_newton_variant = 0; ctx->dbg_serial_err = 0; ctx->dbg_poll_sleep_ns = 0; ctx->spin_limit_ns = 0;
    ctx->hist_count = 0;
    ctx->up_ev_ready = false; ctx->up_ms_total = 0.0; ctx->up_launches_total = 0;
    // (bps_up / bps_down = 0: full occupancy.  With lane-0 staging three blocks per SM were 2 % faster; since the
    // staging copies are spread over the lanes the sweeps are no longer issue-bound and full occupancy wins.)
    *out = ctx;
    return create_impl(ctx, x_lg, y_lg, device_ids, n_devices);
}

void ve_destroy(ve_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    void* ptrs[] = {ctx->uplift, ctx->nf, ctx->mf, ctx->gf, ctx->eps0, ctx->alpha, ctx->kf, ctx->kd, ctx->hscale, ctx->ocean,
                    ctx->h, ctx->b, ctx->wh, ctx->dh, ctx->dh0, ctx->indirection, ctx->newh, ctx->mstack, ctx->stackpos,
                    ctx->mrec, ctx->ndon, ctx->mwrec, ctx->area, ctx->max_slope, ctx->kd_eff, ctx->h_t, ctx->h_new,
                    ctx->deltah, ctx->elev, ctx->h_p, ctx->lake_silt, ctx->lwv, ctx->wh_prev, ctx->h_spl, ctx->b_spl,
                    ctx->wh_mrec, ctx->h_pre_diff, ctx->kfs, ctx->kdf, ctx->sill_cur, ctx->sill_prev, ctx->order_td, ctx->order_bu, ctx->level,
                    ctx->don_order, ctx->don_live, ctx->flags, ctx->order_q, ctx->f_parent, ctx->f_jump, ctx->f_depth, ctx->f_root, ctx->abort_word,
                    ctx->d_scalars, ctx->nz.perm, ctx->cub_tmp};
    for (void* p : ptrs) if (p) cudaFree(p);
    for (int i = 0; i < 12; ++i) if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
    for (auto& kv : ctx->mg_peers)
        for (int r = 0; r < ctx->mg_world; ++r) if (r != ctx->mg_rank && kv.second[r]) cudaIpcCloseMemHandle(kv.second[r]);
    if (ctx->mg_uw_flag) cudaFree(ctx->mg_uw_flag);
    if (ctx->mg_uw_val) cudaFree(ctx->mg_uw_val);
    if (ctx->mg_errsq) cudaFree(ctx->mg_errsq);
    if (ctx->mg_order_own) cudaFree(ctx->mg_order_own);
    if (ctx->up_ev_ready) for (int i = 0; i < 200; ++i) cudaEventDestroy(ctx->up_ev[i]);
    if (ctx->h_scalars) cudaFreeHost(ctx->h_scalars);
    if (ctx->noise_host) cudaFreeHost(ctx->noise_host);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* ve_last_error(ve_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int ve_mg_attach(ve_ctx* ctx, int rank, int world, ve_allgather_cb cb, void* user) {
    if (!ctx || !cb || world < 1 || world > 8 || (world & (world - 1)) || rank < 0 || rank >= world) return VE_ERR_INVALID;
    if (ctx->mg_world > 1) { ctx->err = "ve_mg_attach: already attached"; return VE_ERR_INVALID; }
    if (world == 1) return VE_OK;
    if (ctx->g.n / (uint32_t)world < 1024u) { ctx->err = "ve_mg_attach: map too small for this many bands"; return VE_ERR_INVALID; }
    if (ctx->lanes_down != 1 || ctx->lanes_up != 1) { ctx->err = "ve_mg_attach: needs the warp-per-node sweeps"; return VE_ERR_INVALID; }
    VE_CUDA(cudaSetDevice(ctx->device));
    ctx->mg_rank = rank; ctx->mg_world = world; ctx->mg_cb = cb; ctx->mg_user = user;
    uint32_t per = ctx->g.n / (uint32_t)world, sh = 0;
    while ((1u << sh) < per) ++sh;
    ctx->mg_shift = sh;
    const size_t n = ctx->g.n;
    VE_CUDA(cudaMalloc((void**)&ctx->mg_uw_flag, n));
    VE_CUDA(cudaMalloc((void**)&ctx->mg_uw_val, n * 8));
    VE_CUDA(cudaMalloc((void**)&ctx->mg_errsq, n * 8 + 1024 * 8 + 256));
    VE_CUDA(cudaMalloc((void**)&ctx->mg_order_own, n * 4 + 256));
    void* shared[] = {ctx->deltah, ctx->h_new, ctx->wh, ctx->wh_prev, ctx->sill_cur, ctx->sill_prev, ctx->lake_silt,
                      ctx->mg_uw_flag, ctx->mg_uw_val, ctx->mg_errsq};
    for (void* a : shared) VE_TRY(mg_share(ctx, a));
    return ve_mg_barrier(ctx);
}

int ve_set_debug(ve_ctx* ctx, const char* key, long long value) {
    if (!ctx || !key) return VE_ERR_INVALID;
    std::string k(key);
    if (k == "lanes_down") ctx->lanes_down = value < 1 ? 1u : (uint32_t)value;
    else if (k == "lanes_up") ctx->lanes_up = value < 1 ? 1u : (uint32_t)value;
    else if (k == "lanes_other") ctx->lanes_other = value < 1 ? 1u : (uint32_t)value;
    else if (k == "df_bps") ctx->dbg_df_bps = (int)value;
    else if (k == "newton") ctx->dbg_newton = (int)value;
    else if (k == "newton_variant") ctx->dbg_newton_variant = (int)value;
    else if (k == "lwv_small_log") ctx->dbg_lwv_small_log = (int)value;
    else if (k == "thermal_sweep") ctx->dbg_thermal_sweep = (int)value;
    else if (k == "lakes_trace") ctx->dbg_lakes_trace = (int)value;
    else if (k == "lakes_serial") ctx->dbg_lakes_serial = (int)value;
    else if (k == "quad") ctx->dbg_quad = (int)value;
    else if (k == "rank_jump") ctx->dbg_rank_jump = (int)value;
    else if (k == "level_order") ctx->dbg_level_order = (int)value;
    else if (k == "no_walk_cache") ctx->dbg_no_walk_cache = (int)value;
    else if (k == "rank_launches") ctx->dbg_rank_launches = (int)value;
    else if (k == "wh_full_sweep") ctx->dbg_wh_full_sweep = (int)value;
    else if (k == "bps_up") ctx->bps_up = (int)value;
    else if (k == "bps_down") ctx->bps_down = (int)value;
    else if (k == "serial_err") ctx->dbg_serial_err = (int)value;
    else if (k == "poll_sleep_ns") ctx->dbg_poll_sleep_ns = (int)value;
    else if (k == "spin_limit_ms") ctx->spin_limit_ns = (uint64_t)(value < 0 ? 0 : value) * 1000000ull;
    else { ctx->err = "ve_set_debug: unknown key " + k; return VE_ERR_INVALID; }
    return VE_OK;
}

int ve_set_params(ve_ctx* ctx, const ve_params* p) {
    if (!ctx || !p) return VE_ERR_INVALID;
    ctx->P = *p;
    return VE_OK;
}

int ve_set_noise(ve_ctx* ctx, const ve_noise_fbm_perlin3* nz) {
    if (!ctx || !nz || !nz->perm || nz->octaves == 0 || nz->octaves > 32) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    if (ctx->nz.perm) { cudaFree(ctx->nz.perm); ctx->nz.perm = nullptr; }
    VE_CUDA(cudaMalloc((void**)&ctx->nz.perm, (size_t)nz->octaves * 256));
    VE_CUDA(cudaMemcpy(ctx->nz.perm, nz->perm, (size_t)nz->octaves * 256, cudaMemcpyHostToDevice));
    ctx->nz.octaves = nz->octaves; ctx->nz.frequency = nz->frequency; ctx->nz.lacunarity = nz->lacunarity;
    ctx->nz.persistence = nz->persistence; ctx->nz.scale_factor = nz->scale_factor;
    return VE_OK;
}

int ve_set_noise_cb(ve_ctx* ctx, ve_noise_cb cb, void* user) {
    if (!ctx) return VE_ERR_INVALID;
    ctx->noise_cb = cb; ctx->noise_user = user;
    return VE_OK;
}

int ve_set_inputs(ve_ctx* ctx, const ve_inputs* in) {
    if (!ctx || !in) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    cudaStream_t st = ctx->stream;
    float* tmp;
    VE_TRY(ve_scratch(ctx, 0, n * 4, (void**)&tmp));
    // h = oldh as f64, b = oldb as f64, wh = 0 (erosion.rs:2549-2559, 2601)
    VE_CUDA(cudaMemcpyAsync(tmp, in->oldh, n * 4, cudaMemcpyHostToDevice, st));
    k_f32_to_f64<<<ve_blocks(n, 256), 256, 0, st>>>((uint32_t)n, tmp, ctx->h); VE_LAUNCH_CHECK();
    VE_CUDA(cudaStreamSynchronize(st));
    VE_CUDA(cudaMemcpyAsync(tmp, in->oldb, n * 4, cudaMemcpyHostToDevice, st));
    k_f32_to_f64<<<ve_blocks(n, 256), 256, 0, st>>>((uint32_t)n, tmp, ctx->b); VE_LAUNCH_CHECK();
    VE_CUDA(cudaMemsetAsync(ctx->wh, 0, n * 8, st));
    VE_CUDA(cudaMemcpyAsync(ctx->ocean, in->is_ocean, n, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->uplift, in->uplift, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->nf, in->n, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->mf, in->m, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->gf, in->g, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->eps0, in->epsilon_0, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->alpha, in->alpha, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->kf, in->kf, n * 8, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->kd, in->kd, n * 8, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(ctx->hscale, in->height_scale_cell, n * 8, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaStreamSynchronize(st));
    // max_uplift / max_g (erosion.rs:2617-2626): scalar reductions over the host arrays the shim already holds
    float mu = in->uplift[0], mg = in->g[0];
    for (size_t i = 1; i < n; ++i) { if (in->uplift[i] > mu) mu = in->uplift[i]; if (in->g[i] > mg) mg = in->g[i]; }
    ctx->max_uplift = mu; ctx->max_g = mg;
    return VE_OK;
}

int ve_set_inputs_from_fields(ve_ctx* ctx, double gen_scale, const float* alt_old, const uint8_t* is_ocean, const double* uheight,
                              const double* rock_strength, ve_inputs_out* out) {
    if (!ctx || !alt_old || !uheight || !rock_strength || !(gen_scale > 0.0)) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    const size_t n = ctx->g.n;
    cudaStream_t st = ctx->stream;
    // device staging carved from arrays that a step overwrites anyway
    float* d_alt = (float*)ctx->elev; float* d_cdf = (float*)ctx->deltah; float* d_tmp = (float*)ctx->h_p; float* d_oldh = (float*)ctx->h_new;
    double* d_uh = ctx->area; double* d_rock = ctx->max_slope;
    VE_CUDA(cudaMemcpyAsync(d_alt, alt_old, n * 4, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(d_uh, uheight, n * 8, cudaMemcpyHostToDevice, st));
    VE_CUDA(cudaMemcpyAsync(d_rock, rock_strength, n * 8, cudaMemcpyHostToDevice, st));
    if (is_ocean) VE_CUDA(cudaMemcpyAsync(ctx->ocean, is_ocean, n, cudaMemcpyHostToDevice, st));
    else {
        // get_oceans(map_size_lg, |posi| alt_old[posi].1) (mod.rs:1092): flood from the border over alt_old <= 0
        VE_TRY(ve_k_f32_to_f64(ctx, d_alt, ctx->h_t));
        VE_TRY(ve_g_get_oceans(ctx, ctx->h_t, ctx->ocean));
    }
    double max_erosion = 0.0;
    VE_TRY(ve_k_erosion_inputs(ctx, gen_scale, d_alt, ctx->ocean, d_uh, d_rock, d_cdf, d_tmp, d_oldh, &max_erosion));
    // max_uplift / max_g (erosion.rs:2617-2626): reductions over the two f32 arrays, taken on the host
    std::vector<float> hu(n), hg(n);
    VE_CUDA(cudaMemcpyAsync(hu.data(), ctx->uplift, n * 4, cudaMemcpyDeviceToHost, st));
    VE_CUDA(cudaMemcpyAsync(hg.data(), ctx->gf, n * 4, cudaMemcpyDeviceToHost, st));
    if (out) {
        if (out->oldh) VE_CUDA(cudaMemcpyAsync(out->oldh, d_oldh, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->is_ocean) VE_CUDA(cudaMemcpyAsync(out->is_ocean, ctx->ocean, n, cudaMemcpyDeviceToHost, st));
        if (out->uplift) VE_CUDA(cudaMemcpyAsync(out->uplift, ctx->uplift, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->n) VE_CUDA(cudaMemcpyAsync(out->n, ctx->nf, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->m) VE_CUDA(cudaMemcpyAsync(out->m, ctx->mf, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->g) VE_CUDA(cudaMemcpyAsync(out->g, ctx->gf, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->epsilon_0) VE_CUDA(cudaMemcpyAsync(out->epsilon_0, ctx->eps0, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->alpha) VE_CUDA(cudaMemcpyAsync(out->alpha, ctx->alpha, n * 4, cudaMemcpyDeviceToHost, st));
        if (out->kf) VE_CUDA(cudaMemcpyAsync(out->kf, ctx->kf, n * 8, cudaMemcpyDeviceToHost, st));
        if (out->kd) VE_CUDA(cudaMemcpyAsync(out->kd, ctx->kd, n * 8, cudaMemcpyDeviceToHost, st));
        if (out->height_scale_cell) VE_CUDA(cudaMemcpyAsync(out->height_scale_cell, ctx->hscale, n * 8, cudaMemcpyDeviceToHost, st));
        out->max_erosion_per_delta_t = max_erosion;
    }
    VE_CUDA(cudaStreamSynchronize(st));
    float mu = hu[0], mg = hg[0];
    for (size_t i = 1; i < n; ++i) { if (hu[i] > mu) mu = hu[i]; if (hg[i] > mg) mg = hg[i]; }
    ctx->max_uplift = mu; ctx->max_g = mg;
    return VE_OK;
}

int ve_set_state(ve_ctx* ctx, const double* h, const double* b, const double* wh) {
    if (!ctx) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t nb = (size_t)ctx->g.n * 8;
    if (h) VE_CUDA(cudaMemcpyAsync(ctx->h, h, nb, cudaMemcpyHostToDevice, ctx->stream));
    if (b) VE_CUDA(cudaMemcpyAsync(ctx->b, b, nb, cudaMemcpyHostToDevice, ctx->stream));
    if (wh) VE_CUDA(cudaMemcpyAsync(ctx->wh, wh, nb, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

// One NVTX range per stage of erode() (host-side range around the stage's launches; shows up in
// nsys / ncu --nvtx), plus optional synchronising wall-clock timers (ve_set_timing).
struct StageTimer {
    ve_ctx* ctx; int id; std::chrono::steady_clock::time_point t0; bool open;
    ~StageTimer() { if (open) nvtxRangePop(); }                 // error return inside the stage
    StageTimer(ve_ctx* c, int i) : ctx(c), id(i), open(true) {
        nvtxRangePushA(TIMER_NAMES[i]);
        if (ctx->timing) { cudaStreamSynchronize(ctx->stream); t0 = std::chrono::steady_clock::now(); }
    }
    void stop() {
        nvtxRangePop(); open = false;
        if (ctx->timing) {
            cudaStreamSynchronize(ctx->stream);
            ctx->timers[id] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
    }
};

// erode(): erosion.rs:699-1731
int ve_step(ve_ctx* ctx) {
    if (!ctx) return VE_ERR_INVALID;
    if (!ctx->nz.perm && !ctx->noise_cb) { ctx->err = "no rock-strength source: call ve_set_noise or ve_set_noise_cb"; return VE_ERR_INVALID; }
    VE_CUDA(cudaSetDevice(ctx->device));
    const size_t nb8 = (size_t)ctx->g.n * 8;
    memset(ctx->timers, 0, sizeof(ctx->timers));
    auto t_all = std::chrono::steady_clock::now();
    const double dt = (double)ctx->max_uplift / 1e-3;
    { StageTimer t(ctx, VT_DOWNHILL); VE_TRY(ve_k_downhill(ctx, ctx->h, ctx->ocean, 1, ctx->dh)); t.stop(); }          // 815-819
    { StageTimer t(ctx, VT_LAKES); VE_TRY(ve_l_get_lakes(ctx, ctx->h, ctx->dh, ctx->indirection, ctx->newh)); t.stop(); } // 821-822
    { StageTimer t(ctx, VT_WHFILL); VE_TRY(ve_g_wh_fill(ctx)); VE_TRY(ve_check_abort(ctx)); t.stop(); }                 // 2383-2400
    { StageTimer t(ctx, VT_MREC); VE_TRY(ve_k_mrec(ctx));                                                               // 2407-2479
      VE_CUDA(cudaMemcpyAsync(ctx->wh_mrec, ctx->wh, nb8, cudaMemcpyDeviceToDevice, ctx->stream)); t.stop(); }
    { StageTimer t(ctx, VT_STACK); VE_TRY(ve_g_stack(ctx)); VE_TRY(ve_check_abort(ctx)); t.stop(); }                    // 2483-2511
    { StageTimer t(ctx, VT_AREA); VE_TRY(ve_g_area(ctx)); t.stop(); }                                                   // 841
    { StageTimer t(ctx, VT_SLOPE); VE_TRY(ve_k_max_slope(ctx)); t.stop(); }                                             // 845-856
    { StageTimer t(ctx, VT_KFACT); VE_TRY(ve_k_kfactors(ctx)); t.stop(); }                                              // 870-1006
    nvtxRangePushA("gauss_seidel");
    { int rc_gs = ve_g_spl_solve(ctx); nvtxRangePop(); VE_TRY(rc_gs); }                                                  // 1076-1475
    VE_TRY(ve_check_abort(ctx));
    { StageTimer t(ctx, VT_BASEMENT); VE_TRY(ve_k_basement(ctx)); t.stop(); }                                           // 1483-1556
    { StageTimer t(ctx, VT_THERMAL); VE_TRY(ve_g_thermal_wh(ctx)); VE_TRY(ve_k_thermal_kd(ctx));                        // 1597-1696
      VE_CUDA(cudaMemcpyAsync(ctx->h_pre_diff, ctx->h, nb8, cudaMemcpyDeviceToDevice, ctx->stream)); t.stop(); }
    { StageTimer t(ctx, VT_DIFFUSION); VE_TRY(ve_k_diffusion(ctx, dt, ctx->h, ctx->b, ctx->kd_eff)); t.stop(); }        // 1717-1728
    VE_TRY(ve_check_abort(ctx));
    ctx->timers[VT_TOTAL] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_all).count();
    ctx->hist_n_gs[ctx->hist_count & 255u] = ctx->n_gs; ctx->hist_n_lakes[ctx->hist_count & 255u] = ctx->n_lakes;
    ctx->hist_count++;
    return VE_OK;
}

int ve_run_device(ve_ctx* ctx, uint32_t n_steps) {
    for (uint32_t i = 0; i < n_steps; ++i) VE_TRY(ve_step(ctx));
    return VE_OK;
}

int ve_run_device_timed(ve_ctx* ctx, uint32_t n_steps, double* elapsed_ms) {
    if (!ctx || !elapsed_ms) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    VE_CUDA(cudaEventCreate(&e0)); VE_CUDA(cudaEventCreate(&e1));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    VE_CUDA(cudaEventRecord(e0, ctx->stream));
    int rc = ve_run_device(ctx, n_steps);
    if (rc == VE_OK) {
        VE_CUDA(cudaEventRecord(e1, ctx->stream));
        VE_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        VE_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        *elapsed_ms = (double)ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return rc;
}

int ve_run(ve_ctx* ctx, uint32_t n_steps, void (*progress)(void* user, double pct), void* user, double* out_h, double* out_b) {
    if (!ctx) return VE_ERR_INVALID;
    for (uint32_t i = 0; i < n_steps; ++i) {
        uint32_t every = n_steps / 100 > 1 ? n_steps / 100 : 1;       // erosion.rs:2649
        if (progress && i % every == 0) progress(user, ((double)i / (double)n_steps) * 100.0);
        VE_TRY(ve_step(ctx));
    }
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t nb = (size_t)ctx->g.n * 8;
    if (out_h) VE_CUDA(cudaMemcpyAsync(out_h, ctx->h, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_b) VE_CUDA(cudaMemcpyAsync(out_b, ctx->b, nb, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

size_t ve_array_bytes(ve_ctx* ctx, int which) {
    size_t n = ctx->g.n;
    switch (which) {
        case VE_H: case VE_B: case VE_WH: case VE_AREA: case VE_MAX_SLOPE: case VE_H_SPL: case VE_LAKE_SILT:
        case VE_LAKE_WATER_VOLUME: case VE_KD_EFF: case VE_WH_MREC: case VE_B_SPL: case VE_H_PRE_DIFF: return n * 8;
        case VE_MWREC: return n * 64;
        case VE_DH: case VE_DH0: case VE_INDIRECTION: case VE_NEWH: case VE_MSTACK: case VE_LAKE_SILL: return n * 4;
        case VE_MREC: return n;
        case VE_KFS: case VE_KDF: return n * 32;
    }
    return 0;
}

int ve_get_array(ve_ctx* ctx, int which, void* dst) {
    if (!ctx || !dst) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    const void* src = nullptr;
    uint32_t n = ctx->g.n;
    switch (which) {
        case VE_H: src = ctx->h; break;                  case VE_B: src = ctx->b; break;
        case VE_WH: src = ctx->wh; break;                case VE_DH: src = ctx->dh; break;
        case VE_INDIRECTION: src = ctx->indirection; break; case VE_NEWH: src = ctx->newh; break;
        case VE_MREC: src = ctx->mrec; break;            case VE_MSTACK: src = ctx->mstack; break;
        case VE_MWREC: src = ctx->mwrec; break;          case VE_AREA: src = ctx->area; break;
        case VE_MAX_SLOPE: src = ctx->max_slope; break;  case VE_KFS: src = ctx->kfs; break;
        case VE_KDF: src = ctx->kdf; break;              case VE_H_SPL: src = ctx->h_spl; break;
        case VE_LAKE_WATER_VOLUME: src = ctx->lwv; break; case VE_KD_EFF: src = ctx->kd_eff; break;
        case VE_DH0: src = ctx->dh0; break;              case VE_WH_MREC: src = ctx->wh_mrec; break;
        case VE_B_SPL: src = ctx->b_spl; break;          case VE_H_PRE_DIFF: src = ctx->h_pre_diff; break;
        case VE_LAKE_SILT: {
            double* t; VE_TRY(ve_scratch(ctx, 0, (size_t)n * 8, (void**)&t));
            k_to_stack_index<double><<<ve_blocks(n, 256), 256, 0, ctx->stream>>>(n, ctx->lake_silt, ctx->stackpos, t); VE_LAUNCH_CHECK();
            src = t; break;
        }
        case VE_LAKE_SILL: {
            int32_t* t; VE_TRY(ve_scratch(ctx, 0, (size_t)n * 4, (void**)&t));
            k_to_stack_index<int32_t><<<ve_blocks(n, 256), 256, 0, ctx->stream>>>(n, ctx->sill_cur, ctx->stackpos, t); VE_LAUNCH_CHECK();
            src = t; break;
        }
        default: return VE_ERR_INVALID;
    }
    VE_CUDA(cudaMemcpyAsync(dst, src, ve_array_bytes(ctx, which), cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

double ve_get_scalar(ve_ctx* ctx, int which) {
    switch (which) {
        case VE_S_N_GS: return ctx->n_gs;             case VE_S_ERR: return ctx->gs_err;
        case VE_S_MAXH: return ctx->maxh;             case VE_S_BOUNDARY_LEN: return (double)ctx->boundary_len;
        case VE_S_N_LAKES: return (double)ctx->n_lakes; case VE_S_MAX_UPLIFT: return (double)ctx->max_uplift;
        case VE_S_MAX_G: return (double)ctx->max_g;   case VE_S_N_UNDERWATER: return (double)ctx->n_underwater;
        case VE_S_LAKE_ROUNDS: return (double)ctx->lake_rounds;
        case VE_S_WH_UNDECIDED: return (double)ctx->n_wh_undecided;
        case VE_S_UP_MS_TOTAL: return ctx->up_ms_total; case VE_S_UP_LAUNCHES_TOTAL: return (double)ctx->up_launches_total;
    }
    return NAN;
}

uint64_t ve_launch_count(ve_ctx* ctx) { return ctx ? ctx->launches : 0; }
int ve_get_history(ve_ctx* ctx, int32_t* n_gs, uint32_t* n_lakes, int cap) {
    if (!ctx || cap < 0) return 0;
    uint64_t have = ctx->hist_count < 256 ? ctx->hist_count : 256;
    uint64_t k = have < (uint64_t)cap ? have : (uint64_t)cap;
    for (uint64_t i = 0; i < k; ++i) {
        uint64_t idx = (ctx->hist_count - k + i) & 255u;
        if (n_gs) n_gs[i] = ctx->hist_n_gs[idx];
        if (n_lakes) n_lakes[i] = ctx->hist_n_lakes[idx];
    }
    return (int)k;
}
int ve_set_timing(ve_ctx* ctx, int on) { if (!ctx) return VE_ERR_INVALID; ctx->timing = on != 0; return VE_OK; }
int ve_get_timers(ve_ctx* ctx, double* dst, int cap) {
    int k = cap < VT_COUNT ? cap : (int)VT_COUNT;
    for (int i = 0; i < k; ++i) dst[i] = ctx->timers[i];
    return VT_COUNT;
}
const char* ve_timer_name(int i) { return (i >= 0 && i < VT_COUNT) ? TIMER_NAMES[i] : ""; }

// ---- hydrology entry points (world/src/sim/mod.rs:1474-1550) -----------------------------------
int ve_downhill(ve_ctx* ctx, const double* h, const uint8_t* is_ocean, int32_t* dh) {
    if (!ctx || !h || !is_ocean || !dh) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, h, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    uint8_t* oc; VE_TRY(ve_scratch(ctx, 0, n, (void**)&oc));
    VE_CUDA(cudaMemcpyAsync(oc, is_ocean, n, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_k_downhill(ctx, ctx->h_t, oc, 0, ctx->dh));
    VE_CUDA(cudaMemcpyAsync(dh, ctx->dh, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_get_lakes(ve_ctx* ctx, const double* h, int32_t* dh, int32_t* indirection, uint32_t* newh, double* maxh, uint64_t* boundary_len) {
    if (!ctx || !h || !dh || !indirection || !newh) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, h, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->dh, dh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_l_get_lakes(ctx, ctx->h_t, ctx->dh, ctx->indirection, ctx->newh));
    VE_CUDA(cudaMemcpyAsync(dh, ctx->dh, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(indirection, ctx->indirection, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(newh, ctx->newh, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    if (maxh) *maxh = ctx->maxh;
    if (boundary_len) *boundary_len = ctx->boundary_len;
    return VE_OK;
}

int ve_get_multi_rec(ve_ctx* ctx, const double* h, const int32_t* dh, const uint32_t* newh, double* wh, uint8_t* mrec,
                     uint32_t* mstack, double* mwrec) {
    if (!ctx || !h || !dh || !newh || !wh || !mrec || !mstack || !mwrec) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    // uses the context's state slots: h <- h, dh, newh, wh
    VE_CUDA(cudaMemcpyAsync(ctx->h, h, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->dh, dh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->newh, newh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->wh, wh, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_g_wh_fill(ctx)); VE_TRY(ve_check_abort(ctx));
    VE_TRY(ve_k_mrec(ctx));
    VE_TRY(ve_g_stack(ctx)); VE_TRY(ve_check_abort(ctx));
    VE_CUDA(cudaMemcpyAsync(wh, ctx->wh, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(mrec, ctx->mrec, n, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(mstack, ctx->mstack, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(mwrec, ctx->mwrec, n * 64, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_get_multi_drainage(ve_ctx* ctx, const uint32_t* mstack, const uint8_t* mrec, const double* mwrec, double* area) {
    if (!ctx || !mstack || !mrec || !mwrec || !area) return VE_ERR_INVALID;
    // A pure function of its arguments, as in the reference (erosion.rs:62-70): the three arrays are uploaded,
    // the donor lists are rebuilt from them, and the area is gathered along mstack itself (donors first).
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    VE_CUDA(cudaMemcpyAsync(ctx->mstack, mstack, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->mrec, mrec, n, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->mwrec, mwrec, n * 64, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_g_area_from_stack(ctx)); VE_TRY(ve_check_abort(ctx));
    VE_CUDA(cudaMemcpyAsync(area, ctx->area, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_fill_sinks(ve_ctx* ctx, const double* h, const uint8_t* is_ocean, double* out) {
    if (!ctx || !h || !is_ocean || !out) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    uint8_t* oc; VE_TRY(ve_scratch(ctx, 0, n, (void**)&oc));
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, h, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(oc, is_ocean, n, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_g_fill_sinks(ctx, ctx->h_t, oc, ctx->elev));
    VE_CUDA(cudaMemcpyAsync(out, ctx->elev, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_get_oceans(ve_ctx* ctx, const double* h, uint8_t* is_ocean) {
    if (!ctx || !h || !is_ocean) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    uint8_t* oc; VE_TRY(ve_scratch(ctx, 0, n, (void**)&oc));
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, h, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_g_get_oceans(ctx, ctx->h_t, oc));
    VE_CUDA(cudaMemcpyAsync(is_ocean, oc, n, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_water_alt(ve_ctx* ctx, const double* alt, const uint8_t* is_ocean, const int32_t* dh, const int32_t* indirection,
                 double* water_alt_filled, float* water_alt_final) {
    if (!ctx || !alt || !is_ocean || !dh || !indirection || !water_alt_final) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    uint8_t* oc; VE_TRY(ve_scratch(ctx, 0, n, (void**)&oc));
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, alt, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(oc, is_ocean, n, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->dh, dh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->indirection, indirection, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    float* fin = (float*)ctx->h_p;
    VE_TRY(ve_g_water_alt(ctx, ctx->h_t, oc, ctx->dh, ctx->indirection, ctx->deltah, ctx->elev, fin));
    if (water_alt_filled) VE_CUDA(cudaMemcpyAsync(water_alt_filled, ctx->elev, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(water_alt_final, fin, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_get_rivers(ve_ctx* ctx, double continent_scale_hack, const uint32_t* newh, const double* water_alt, const int32_t* dh,
                  const int32_t* indirection, const double* drainage, ve_river* rivers) {
    if (!ctx || !newh || !water_alt || !dh || !indirection || !drainage || !rivers) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    VE_CUDA(cudaMemcpyAsync(ctx->newh, newh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, water_alt, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->dh, dh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->indirection, indirection, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->area, drainage, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    void* out = (void*)ctx->mwrec;   // 64 B per cell available, 32 needed
    VE_TRY(ve_g_get_rivers(ctx, continent_scale_hack, ctx->newh, ctx->h_t, ctx->dh, ctx->indirection, ctx->area, out));
    VE_CUDA(cudaMemcpyAsync(rivers, out, n * sizeof(ve_river), cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_uniform_noise(ve_ctx* ctx, const float* values, float* cdf) {
    if (!ctx || !values || !cdf) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    float* dv = (float*)ctx->elev; float* dout = (float*)ctx->deltah;
    VE_CUDA(cudaMemcpyAsync(dv, values, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_g_uniform_noise(ctx, dv, dout, nullptr));
    VE_CUDA(cudaMemcpyAsync(cdf, dout, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

// util.rs:22-29: distance to the nearest map edge in units of 1/64 of the side, clamped to [0, 1], f32 arithmetic
__global__ void __launch_bounds__(256) k_map_edge_factor(Geom g, float* __restrict__ out) {
    uint32_t posi = blockIdx.x * blockDim.x + threadIdx.x;
    if (posi >= g.n) return;
    int ex = g.x_of(posi), ey = g.y_of(posi);
    int sx = g.nx, sy = g.ny;
    float fx = (float)(sx / 2 - abs(ex - sx / 2)) / ((16.0f / 1024.0f) * (float)sx);
    float fy = (float)(sy / 2 - abs(ey - sy / 2)) / ((16.0f / 1024.0f) * (float)sy);
    out[posi] = fminf(fmaxf(fminf(fx, fy), 0.0f), 1.0f);
}

int ve_map_edge_factor(ve_ctx* ctx, float* out) {
    if (!ctx || !out) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    float* d = (float*)ctx->elev;
    k_map_edge_factor<<<ve_blocks(n, 256), 256, 0, ctx->stream>>>(ctx->g, d);
    VE_LAUNCH_CHECK();
    VE_CUDA(cudaMemcpyAsync(out, d, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_horizon_map(ve_ctx* ctx, float minh, float maxh, const float* h, float* west_angle, float* west_height,
                   float* east_angle, float* east_height) {
    if (!ctx || !h || !west_angle || !west_height || !east_angle || !east_height) return VE_ERR_INVALID;
    if (maxh < minh) { ctx->err = "get_horizon_map: maxh must not be below minh"; return VE_ERR_INVALID; }   // util.rs:382-385
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    float* dh_ = (float*)ctx->elev;
    float* o = (float*)ctx->mwrec;   // 4 * n floats needed, 16 n available
    VE_CUDA(cudaMemcpyAsync(dh_, h, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_g_horizon(ctx, minh, maxh, dh_, o, o + n, o + 2 * n, o + 3 * n));
    VE_CUDA(cudaMemcpyAsync(west_angle, o, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(west_height, o + n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(east_angle, o + 2 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(east_height, o + 3 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_get_drainage(ve_ctx* ctx, const uint32_t* newh, const int32_t* dh, float* flux) {
    if (!ctx || !newh || !dh || !flux) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    VE_CUDA(cudaMemcpyAsync(ctx->newh, newh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->dh, dh, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    float* fl = (float*)ctx->elev;
    VE_TRY(ve_g_get_drainage(ctx, ctx->newh, ctx->dh, fl)); VE_TRY(ve_check_abort(ctx));
    VE_CUDA(cudaMemcpyAsync(flux, fl, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

int ve_diffusion(ve_ctx* ctx, double dt, double* h, double* b, const double* kd) {
    if (!ctx || !h || !b || !kd) return VE_ERR_INVALID;
    VE_CUDA(cudaSetDevice(ctx->device));
    size_t n = ctx->g.n;
    VE_CUDA(cudaMemcpyAsync(ctx->h_t, h, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->deltah, b, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(ctx->elev, kd, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    VE_TRY(ve_k_diffusion(ctx, dt, ctx->h_t, ctx->deltah, ctx->elev));
    VE_CUDA(cudaMemcpyAsync(h, ctx->h_t, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaMemcpyAsync(b, ctx->deltah, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VE_CUDA(cudaStreamSynchronize(ctx->stream));
    return VE_OK;
}

} // extern "C"
