// erosion_oracle.cpp — CPU oracle for the world-map erosion path.  TEST INFRASTRUCTURE.
//
// A restatement (not a copy) of the reference algorithm, following it loop by loop so that
// every order-dependent result (receivers, stack order, lake ids, f64 rounding) is the one the
// reference would produce.  Serial where the reference is serial, `omp parallel for` where it
// uses rayon.  Each function cites the reference lines it follows (paths relative to the
// veloren tree).  PARITY UNPINNED: the reference ships no test or golden vector for this path
// and cannot be built here (no Rust toolchain); see DESIGN.md.
//
// Build: g++ -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
// -ffp-contract=off matters: Rust never fuses a*b+c.

#include "erosion_oracle.h"
#include "../include/ve_exact_math.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <queue>
#include <tuple>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// common/src/terrain/mod.rs:421-430
const int NDX[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
const int NDY[8] = {-1, -1, -1, 0, 0, 1, 1, 1};

struct Grid {
    uint32_t xlg, ylg;
    int nx, ny;
    size_t n;
    // common/src/terrain/mod.rs:407-418
    inline int x_of(size_t i) const { return (int)(i & ((1u << xlg) - 1)); }
    inline int y_of(size_t i) const { return (int)(i >> xlg); }
    inline size_t idx(int x, int y) const { return ((size_t)y << xlg) | (size_t)x; }
    inline bool inb(int x, int y) const { return x >= 0 && y >= 0 && x < nx && y < ny; }
};

Grid make_grid(uint32_t xlg, uint32_t ylg) {
    Grid g; g.xlg = xlg; g.ylg = ylg; g.nx = 1 << xlg; g.ny = 1 << ylg; g.n = (size_t)g.nx * g.ny; return g;
}

struct Math {
    bool libm;    // f64 transcendentals from the host libm instead of ve_exact_math.h
    bool libmf;   // f32 powf(dh, 0.5) of the Newton loop from the host libm's powf instead of sqrtf
    inline double pow_(double x, double y) const { return libm ? std::pow(x, y) : vm_pow(x, y); }
    inline double exp_(double x) const { return libm ? std::exp(x) : vm_exp(x); }
    inline double log_(double x) const { return libm ? std::log(x) : vm_log(x); }
    inline double log1p_(double x) const { return libm ? std::log1p(x) : vm_log1p(x); }
    inline double tanh_(double x) const { return libm ? std::tanh(x) : vm_tanh(x); }
    inline float powf_(float x, float y) const { return libmf ? std::pow(x, y) : vm_powf(x, y); }
};

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------
// util.rs:251-279  downhill
void downhill(const Grid& g, const double* h, const uint8_t* ocean_pred, int32_t* dh) {
#pragma omp parallel for schedule(static)
    for (long long posi = 0; posi < (long long)g.n; ++posi) {
        double nh = h[posi];
        if (ocean_pred[posi]) { dh[posi] = -2; continue; }
        int best = -1; double besth = nh;
        int x = g.x_of(posi), y = g.y_of(posi);
        for (int k = 0; k < 8; ++k) {
            int xx = x + NDX[k], yy = y + NDY[k];
            if (!g.inb(xx, yy)) continue;
            size_t np = g.idx(xx, yy);
            double nbh = h[np];
            if (nbh < besth) { besth = nbh; best = (int)np; }
        }
        dh[posi] = best;
    }
}

// ---------------------------------------------------------------------------------------------
// erosion.rs:1825-2293  get_lakes
struct Pass { int32_t from; uint32_t to; };

int get_lakes(const Grid& g, const double* h, int32_t* downhill_, int32_t* indirection,
              uint32_t* newh_out, double* maxh_out, uint64_t* boundary_len_out, size_t* n_lakes_out) {
    const size_t N = g.n;
    std::vector<size_t> boundary; boundary.reserve(N / 4);
    std::fill(indirection, indirection + N, 0);
    std::vector<uint32_t> newh; newh.reserve(N);
    std::vector<Pass> lakes(N * 8, Pass{-1, 0});
    std::vector<uint32_t> indirection_(N, 0u);
    size_t n_lake_roots = 0;
    // 1865-1902: find all lakes (every dh<0 cell is a root), BFS its uphill tree.
    for (size_t chunk_idx = 0; chunk_idx < N; ++chunk_idx) {
        int32_t dhv = downhill_[chunk_idx];
        if (dhv >= 0) continue;
        if (dhv == -2) boundary.push_back(chunk_idx);
        else if (dhv == -1) n_lake_roots++;
        else return -10;
        size_t start = newh.size();
        uint32_t indirection_idx = (uint32_t)(start * 8);
        newh.push_back((uint32_t)chunk_idx);
        size_t cur = start;
        while (cur < newh.size()) {
            size_t node = newh[cur];
            int x = g.x_of(node), y = g.y_of(node);
            for (int k = 0; k < 8; ++k) { // uphill(): util.rs:240-246
                int xx = x + NDX[k], yy = y + NDY[k];
                if (!g.inb(xx, yy)) continue;
                size_t child = g.idx(xx, yy);
                if (downhill_[child] == (int32_t)node) {
                    indirection[child] = (int32_t)chunk_idx;
                    indirection_[child] = indirection_idx;
                    newh.push_back((uint32_t)child);
                }
            }
            cur++;
        }
        size_t length = (cur - start) * 8;
        indirection[chunk_idx] = -(int32_t)length;
        indirection_[chunk_idx] = indirection_idx;
    }
    if (newh.size() != N) return -11; // assert_eq 1903
    if (n_lakes_out) *n_lakes_out = n_lake_roots;

    double maxh = -std::numeric_limits<double>::infinity();
    // 1915-2019: pass table
    for (size_t t = 0; t < N; ++t) {
        uint32_t chunk_idx_ = newh[t];
        size_t chunk_idx = chunk_idx_;
        uint32_t lake_idx_ = indirection_[chunk_idx];
        size_t lake_idx = lake_idx_;
        double height = h[chunk_idx];
        maxh = std::max(maxh, std::fabs(height));
        int x = g.x_of(chunk_idx), y = g.y_of(chunk_idx);
        for (int k = 0; k < 8; ++k) {
            int xx = x + NDX[k], yy = y + NDY[k];
            if (!g.inb(xx, yy)) continue;
            size_t neighbor_idx = g.idx(xx, yy);
            double neighbor_height = h[neighbor_idx];
            uint32_t neighbor_lake_idx_ = indirection_[neighbor_idx];
            size_t neighbor_lake_idx = neighbor_lake_idx_;
            if (neighbor_lake_idx_ < lake_idx_) {
                Pass* lake = &lakes[lake_idx];
                Pass* neighbor_lake = &lakes[neighbor_lake_idx];
                for (size_t s = 0;; ++s) {
                    Pass& pass = lake[s];
                    if (pass.from == -1) {
                        pass = Pass{(int32_t)chunk_idx_, (uint32_t)neighbor_idx};
                        size_t q = 0;
                        while (neighbor_lake[q].from != -1) ++q;
                        neighbor_lake[q] = Pass{(int32_t)neighbor_idx, chunk_idx_};
                        break;
                    } else if (indirection_[pass.to] == neighbor_lake_idx_) {
                        bool found = false;
                        for (size_t q = 0;; ++q) {
                            Pass& neighbor_pass = neighbor_lake[q];
                            if (neighbor_pass.from == -1) break; // would be the reference's panic
                            if (indirection_[neighbor_pass.to] == lake_idx_) {
                                double pass_height = h[neighbor_pass.to];
                                double neighbor_pass_height = h[pass.to];
                                if (std::max(height, neighbor_height) <
                                    std::max(pass_height, neighbor_pass_height)) {
                                    pass = Pass{(int32_t)chunk_idx_, (uint32_t)neighbor_idx};
                                    neighbor_pass = Pass{(int32_t)neighbor_idx, chunk_idx_};
                                }
                                found = true;
                                break;
                            }
                        }
                        if (!found) return -12;
                        break;
                    }
                }
            }
        }
    }

    // 2022-2065: candidates
    typedef std::tuple<double, uint32_t, uint32_t> Cand;
    std::priority_queue<Cand, std::vector<Cand>, std::greater<Cand>> candidates;
    for (size_t e = 0; e < lakes.size(); ++e) {
        Pass& edge = lakes[e];
        if (edge.from == -1) continue;
        size_t from = (size_t)edge.from;
        int32_t ii = indirection[from];
        size_t lake_idx = ii < 0 ? from : (size_t)ii;
        if (downhill_[lake_idx] == -2) { edge.from = -1; continue; }
        size_t to = edge.to;
        ii = indirection[to];
        lake_idx = ii < 0 ? to : (size_t)ii;
        if (downhill_[lake_idx] == -2) {
            double pass = std::max(h[from], h[to]);
            candidates.push(Cand(pass, (uint32_t)edge.from, edge.to));
        }
    }

    // test/debug aid (ORC_DUMP_LAKES=<file>): the directed pass table as the growth phase sees it, then the
    // connections in the order they are made.  Used by tools/lake_growth_prototype.py.
    FILE* dumpf = nullptr;
    if (const char* dp = getenv("ORC_DUMP_LAKES")) dumpf = fopen(dp, "wb");
    if (dumpf) {
        std::vector<int64_t> rec;
        for (size_t e = 0; e < lakes.size(); ++e) {
            const Pass& edge = lakes[e];
            if (edge.from == -1) continue;
            int32_t ii = indirection[edge.from]; size_t lf = ii < 0 ? (size_t)edge.from : (size_t)ii;
            ii = indirection[edge.to]; size_t lt = ii < 0 ? (size_t)edge.to : (size_t)ii;
            rec.push_back(edge.from); rec.push_back(edge.to); rec.push_back((int64_t)lf); rec.push_back((int64_t)lt);
            rec.push_back(downhill_[lt] == -2 ? 1 : 0);
        }
        int64_t cnt = (int64_t)(rec.size() / 5);
        fwrite(&cnt, 8, 1, dumpf); fwrite(rec.data(), 8, rec.size(), dumpf);
    }
    std::vector<size_t> pass_flows_sorted;
    std::vector<int64_t> dump_conn;
    // 2076-2137
    while (!candidates.empty()) {
        Cand c = candidates.top(); candidates.pop();
        uint32_t chunk_idx = std::get<1>(c), neighbor_idx = std::get<2>(c);
        size_t lake_idx = indirection_[chunk_idx];
        int32_t indirection_idx = indirection[chunk_idx];
        size_t lake_chunk_idx = indirection_idx >= 0 ? (size_t)indirection_idx : (size_t)chunk_idx;
        if (downhill_[lake_chunk_idx] >= 0) continue;
        if (downhill_[lake_chunk_idx] != -1) return -13;
        int32_t neg = indirection_idx < 0 ? indirection_idx : indirection[indirection_idx];
        size_t max_len = (size_t)(-(int64_t)neg);
        downhill_[lake_chunk_idx] = (int32_t)neighbor_idx;
        indirection[lake_chunk_idx] = -(int32_t)chunk_idx;
        pass_flows_sorted.push_back(lake_chunk_idx);
        if (dumpf) { dump_conn.push_back((int64_t)lake_chunk_idx); dump_conn.push_back(chunk_idx); dump_conn.push_back(neighbor_idx); }
        for (size_t s = lake_idx; s < lake_idx + max_len; ++s) {
            Pass& edge_ref = lakes[s];
            if (edge_ref.from == (int32_t)chunk_idx && edge_ref.to == neighbor_idx) continue;
            Pass edge = edge_ref;
            edge_ref = Pass{-1, 0};
            if (edge.from == -1) break;
            int32_t ii = indirection[edge.to];
            size_t neighbor_lake_idx = ii < 0 ? (size_t)edge.to : (size_t)ii;
            if (downhill_[neighbor_lake_idx] != -1) continue;
            double pass = std::max(h[edge.from], h[edge.to]);
            candidates.push(Cand(pass, edge.to, (uint32_t)edge.from));
        }
    }

    if (dumpf) {
        int64_t cnt = (int64_t)(dump_conn.size() / 3);
        fwrite(&cnt, 8, 1, dumpf); fwrite(dump_conn.data(), 8, dump_conn.size(), dumpf);
        fclose(dumpf);
    }
    // 2141-2290: re-route flooded cells of each lake, then final BFS order.
    enum : uint8_t { UnParsed = 0, InQueue = 1, WithRcv = 2 };
    std::vector<uint8_t> tag(N, UnParsed);
    std::vector<size_t> filling_queue;
    std::vector<uint32_t> newh2; newh2.reserve(N);
    auto process = [&](size_t chunk_idx) -> int {
        size_t start = newh2.size();
        int32_t neighbor_pass_idx_s = downhill_[chunk_idx];
        size_t first_idx;
        if (neighbor_pass_idx_s < 0) {
            newh2.push_back((uint32_t)chunk_idx);
            start += 1;
            first_idx = chunk_idx;
        } else {
            size_t neighbor_pass_idx = (size_t)neighbor_pass_idx_s;
            int32_t pass_idx_s = -indirection[chunk_idx];
            if (pass_idx_s < 0) return -14;
            size_t pass_idx = (size_t)pass_idx_s;
            if (tag[pass_idx] != UnParsed) return -15;
            filling_queue.push_back(pass_idx);
            tag[neighbor_pass_idx] = WithRcv;
            tag[pass_idx] = InQueue;
            int ox = g.x_of(neighbor_pass_idx), oy = g.y_of(neighbor_pass_idx);
            double elev = std::max(h[neighbor_pass_idx], h[pass_idx]);
            while (!filling_queue.empty()) {
                size_t node = filling_queue.back(); filling_queue.pop_back();
                int cx = g.x_of(node), cy = g.y_of(node);
                int64_t rcv = -1;
                double rcv_cost = -std::numeric_limits<double>::infinity();
                double odx = (double)(ox - cx), ody = (double)(oy - cy);
                for (int k = 0; k < 8; ++k) {
                    int xx = cx + NDX[k], yy = cy + NDY[k];
                    if (!g.inb(xx, yy)) continue;
                    size_t ineighbor = g.idx(xx, yy);
                    if ((indirection[ineighbor] != (int32_t)chunk_idx && ineighbor != chunk_idx &&
                         ineighbor != neighbor_pass_idx) || h[ineighbor] > elev)
                        continue;
                    double ddx = (double)(xx - cx), ddy = (double)(yy - cy);
                    uint8_t& tg = tag[ineighbor];
                    if (tg == WithRcv) {
                        // vek: v / magnitude, then dot = x*x' + y*y'
                        double mag = std::sqrt(ddx * ddx + ddy * ddy);
                        double ux = ddx / mag, uy = ddy / mag;
                        double cost = odx * ux + ody * uy;
                        if (cost > rcv_cost) { rcv = (int64_t)ineighbor; rcv_cost = cost; }
                    } else if (tg == UnParsed) {
                        filling_queue.push_back(ineighbor);
                        tg = InQueue;
                    }
                }
                if (rcv == -1) return -16; // assert_ne 2262
                downhill_[node] = (int32_t)rcv;
                tag[node] = WithRcv;
            }
            first_idx = neighbor_pass_idx;
        }
        size_t cur = start;
        size_t node = first_idx;
        for (;;) {
            int x = g.x_of(node), y = g.y_of(node);
            for (int k = 0; k < 8; ++k) {
                int xx = x + NDX[k], yy = y + NDY[k];
                if (!g.inb(xx, yy)) continue;
                size_t child = g.idx(xx, yy);
                if (downhill_[child] != (int32_t)node) continue;
                if (indirection[child] == (int32_t)chunk_idx || child == chunk_idx)
                    newh2.push_back((uint32_t)child);
            }
            if (cur == newh2.size()) break;
            node = newh2[cur];
            cur += 1;
        }
        return 0;
    };
    for (size_t c : boundary) { int rc = process(c); if (rc) return rc; }
    for (size_t c : pass_flows_sorted) { int rc = process(c); if (rc) return rc; }
    if (newh2.size() != N) return -17; // assert_eq 2291
    std::memcpy(newh_out, newh2.data(), N * sizeof(uint32_t));
    *maxh_out = maxh;
    *boundary_len_out = boundary.size();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// erosion.rs:2332-2520  get_multi_rec
int get_multi_rec(const Grid& g, const Math& M, const double* h, const int32_t* dh, const uint32_t* newh,
                  double* wh, uint8_t* mrec, uint32_t* mstack, double* mwrec, double dx, double dy,
                  size_t* n_sources) {
    const size_t nn = g.n;
    // 2383-2400
    const double deltah0 = std::numeric_limits<double>::epsilon() + std::numeric_limits<double>::epsilon();
    for (size_t t = 0; t < nn; ++t) {
        size_t ijk = newh[t];
        double h_i = h[ijk];
        int32_t ijr = dh[ijk];
        if (ijr >= 0) {
            double wh_j = wh[ijr];
            if (wh_j >= h_i) {
                double deltah = deltah0 * std::fabs(wh_j);
                wh[ijk] = wh_j + deltah;
            } else wh[ijk] = h_i;
        } else wh[ijk] = h_i;
    }
    std::vector<uint8_t> ndon(nn), vis(nn);
    // 2407-2439
#pragma omp parallel for schedule(static)
    for (long long ij = 0; ij < (long long)nn; ++ij) {
        double wh_ij = wh[ij];
        uint8_t mrec_ij = 0, ndon_ij = 0;
        int x = g.x_of(ij), y = g.y_of(ij);
        for (int k = 0; k < 8; ++k) {
            int xx = x + NDX[k], yy = y + NDY[k];
            if (!g.inb(xx, yy)) continue;
            double wh_ijk = wh[g.idx(xx, yy)];
            if (wh_ij > wh_ijk) mrec_ij |= (uint8_t)(1u << k);
            else if (wh_ijk > wh_ij) ndon_ij += 1;
        }
        mrec[ij] = mrec_ij; ndon[ij] = ndon_ij; vis[ij] = ndon_ij;
    }
    // 2444-2479 (par) — weights
#pragma omp parallel for schedule(static)
    for (long long ij = 0; ij < (long long)nn; ++ij) {
        double sumweight = 0.0;
        double wrec[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int nrec = 0;
        int x = g.x_of(ij), y = g.y_of(ij);
        uint8_t mr = mrec[ij];
        for (int k = 0; k < 8; ++k) {
            if (!((mr >> k) & 1)) continue;
            size_t ijk = g.idx(x + NDX[k], y + NDY[k]);
            double ex = (double)NDX[k] * dx, ey = (double)NDY[k] * dy;
            double lrec = std::sqrt(ex * ex + ey * ey);
            double w = (wh[ij] - wh[ijk]) / lrec;
            wrec[k] = w; sumweight += w; nrec += 1;
        }
        double slope = sumweight / std::max((double)nrec, 1.0);
        double p_mfd_exp = 0.5 + 0.6 * slope;
        sumweight = 0.0;
        for (int k = 0; k < 8; ++k) {
            double w = M.pow_(wrec[k], p_mfd_exp);
            sumweight += w; wrec[k] = w;
        }
        if (sumweight > 0.0) for (int k = 0; k < 8; ++k) wrec[k] /= sumweight;
        for (int k = 0; k < 8; ++k) mwrec[(size_t)ij * 8 + k] = wrec[k];
    }
    // 2483-2511 — stack (serial)
    size_t sp = 0, nsrc = 0;
    std::vector<size_t> parse; parse.reserve(1024);
    for (size_t ij = 0; ij < nn; ++ij) {
        if (ndon[ij] == 0) { parse.push_back(ij); nsrc++; }
        while (!parse.empty()) {
            size_t ijn = parse.back(); parse.pop_back();
            mstack[sp++] = (uint32_t)ijn;
            int x = g.x_of(ijn), y = g.y_of(ijn);
            uint8_t mr = mrec[ijn];
            for (int k = 0; k < 8; ++k) {
                if (!((mr >> k) & 1)) continue;
                size_t ijr = g.idx(x + NDX[k], y + NDY[k]);
                if (vis[ijr] >= 1) {
                    vis[ijr] -= 1;
                    if (vis[ijr] == 0) parse.push_back(ijr);
                }
            }
        }
    }
    if (n_sources) *n_sources = nsrc;
    if (sp != nn) return -20; // assert_eq 2510
    return 0;
}

// erosion.rs:62-97  get_multi_drainage
void get_multi_drainage(const Grid& g, const uint32_t* mstack, const uint8_t* mrec, const double* mwrec, double* area) {
    for (size_t i = 0; i < g.n; ++i) area[i] = 1.0;
    for (size_t t = 0; t < g.n; ++t) {
        size_t ij = mstack[t];
        double area_ij = area[ij];
        int x = g.x_of(ij), y = g.y_of(ij);
        uint8_t mr = mrec[ij];
        for (int k = 0; k < 8; ++k) {
            if (!((mr >> k) & 1)) continue;
            size_t ijr = g.idx(x + NDX[k], y + NDY[k]);
            area[ijr] += area_ij * mwrec[ij * 8 + k];
        }
    }
}

// erosion.rs:29-58 (commented out in the reference)  get_drainage
void get_drainage(const Grid& g, const uint32_t* newh, const int32_t* dh, float* flux) {
    for (size_t i = 0; i < g.n; ++i) flux[i] = 1.0f;
    for (size_t t = g.n; t-- > 0;) {
        size_t c = newh[t];
        int32_t d = dh[c];
        if (d >= 0) flux[d] += flux[c];
    }
}

// ---------------------------------------------------------------------------------------------
// noise 0.9.0 Fbm<Perlin> (NOT in the reference tree; restated from the crate's published
// algorithm; parity with the crate unpinned).  Tables come from the caller.
struct Noise {
    uint32_t octaves = 0; double frequency = 1, lacunarity = 2, persistence = 0.5, scale_factor = 1;
    std::vector<uint8_t> perm;
};

inline double grad3(unsigned hsh, double x, double y, double z) {
    switch (hsh & 15u) {
        case 0: return x + y;   case 1: return -x + y;  case 2: return x - y;   case 3: return -x - y;
        case 4: return x + z;   case 5: return -x + z;  case 6: return x - z;   case 7: return -x - z;
        case 8: return y + z;   case 9: return -y + z;  case 10: return y - z;  case 11: return -y - z;
        case 12: return x + y;  case 13: return -x + y; case 14: return -y + z; default: return -y - z;
    }
}
inline double quintic(double t) { return t * t * t * (t * (t * 6.0 - 15.0) + 10.0); }

double perlin3(const uint8_t* p, double px, double py, double pz) {
    double fx = std::floor(px), fy = std::floor(py), fz = std::floor(pz);
    long long ix = (long long)fx, iy = (long long)fy, iz = (long long)fz;
    double dx = px - fx, dy = py - fy, dz = pz - fz;
    auto hash = [&](long long a, long long b, long long c) -> unsigned {
        unsigned i0 = (unsigned)(a & 0xff);
        unsigned i1 = (unsigned)p[i0] ^ (unsigned)(b & 0xff);
        unsigned i2 = (unsigned)p[i1] ^ (unsigned)(c & 0xff);
        return p[i2];
    };
    double g000 = grad3(hash(ix, iy, iz), dx, dy, dz);
    double g100 = grad3(hash(ix + 1, iy, iz), dx - 1.0, dy, dz);
    double g010 = grad3(hash(ix, iy + 1, iz), dx, dy - 1.0, dz);
    double g110 = grad3(hash(ix + 1, iy + 1, iz), dx - 1.0, dy - 1.0, dz);
    double g001 = grad3(hash(ix, iy, iz + 1), dx, dy, dz - 1.0);
    double g101 = grad3(hash(ix + 1, iy, iz + 1), dx - 1.0, dy, dz - 1.0);
    double g011 = grad3(hash(ix, iy + 1, iz + 1), dx, dy - 1.0, dz - 1.0);
    double g111 = grad3(hash(ix + 1, iy + 1, iz + 1), dx - 1.0, dy - 1.0, dz - 1.0);
    double u = quintic(dx), v = quintic(dy), w = quintic(dz);
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
    return std::min(1.0, std::max(-1.0, r));
}

double fbm3(const Noise& nz, double x, double y, double z) {
    double result = 0.0, attenuation = nz.persistence;
    x *= nz.frequency; y *= nz.frequency; z *= nz.frequency;
    for (uint32_t o = 0; o < nz.octaves; ++o) {
        double signal = perlin3(&nz.perm[(size_t)o * 256], x, y, z);
        signal *= attenuation;
        attenuation *= nz.persistence;
        result += signal;
        x *= nz.lacunarity; y *= nz.lacunarity; z *= nz.lacunarity;
    }
    return result * nz.scale_factor;
}

// erosion.rs:550-604  get_max_slope
void get_max_slope(const Grid& g, const Math& M, const orc_params& P, const Noise& nz, const double* h,
                   const double* hscale, double* out) {
    const double max_angle_range = P.max_max_angle - P.min_max_angle;
    const double logistic_2_base = std::sqrt(3.0) * 0.63661977236758134308; // FRAC_2_PI
    auto logit = [&](double x) { return M.log_(x) - M.log1p_(-x); };
    const double center = 0.4, dmin = center - 0.05, dmax = center + 0.05;
    const double logit_center = logit(center);
#pragma omp parallel for schedule(static)
    for (long long posi = 0; posi < (long long)g.n; ++posi) {
        double z = h[posi];
        double wx = (double)g.x_of(posi) * 32.0, wy = (double)g.y_of(posi) * 32.0;
        double wposz = z / hscale[posi];
        double div_factor = (2.0 * 32.0) / 8.0;
        double rs = fbm3(nz, wx, wy, wposz * div_factor);
        rs = std::min(1.0, std::max(-1.0, rs)) * 0.5 + 0.5;
        double a = std::min(1.0 - 1e-7, std::max(1e-7, rs));
        double hz = std::fabs(wposz / P.mountain_scale);
        hz = std::min(dmax, std::max(dmin, hz));
        double arg = 1.0 * logit(a) + 1.0 * (logit(hz) - logit_center);
        double rock = M.tanh_(arg / logistic_2_base) * 0.5 + 0.5;
        out[posi] = rock * max_angle_range + P.min_max_angle;
    }
}

// ---------------------------------------------------------------------------------------------
// diffusion.rs:408-472  tridag
int tridag(const double* a, const double* b, const double* c, const double* r, double* u, double* gam, size_t n) {
    if (b[0] == 0.0) return -30;
    double bet = b[0];
    u[0] = r[0] / bet;
    for (size_t j = 1; j < n; ++j) {
        gam[j] = c[j - 1] / bet;
        bet = b[j] - a[j] * gam[j];
        if (bet == 0.0) return -31;
        u[j] = (r[j] - a[j] * u[j - 1]) / bet;
    }
    for (size_t j = n - 1; j-- > 0;) u[j] -= gam[j + 1] * u[j + 1];
    return 0;
}

// diffusion.rs:41-394  diffusion
int diffusion(size_t nx, size_t ny, double xl, double yl, double dt, double* h, double* b,
              const double* kd, double kdsed) {
    auto aij = [nx](size_t i, size_t j) { return j * nx + i; };
    const double dx = xl / (double)nx, dy = yl / (double)ny;
    std::vector<double> zint(nx * ny), kdint(nx * ny);
    for (size_t j = 0; j < ny; ++j)
        for (size_t i = 0; i < nx; ++i) {
            size_t ij = aij(i, j);
            zint[ij] = h[ij];
            kdint[ij] = kd[ij];
            if (kdsed > 0.0 && (h[ij] - b[ij]) > 1.0e-6) kdint[ij] = kdsed;
        }
    std::vector<double> zintp = zint;
    {
        std::vector<double> f(nx, 0.0), diag(nx, 0.0), sup(nx, 0.0), inf(nx, 0.0), res(nx, 0.0), gam(nx, 0.0);
        for (size_t j = 1; j + 1 < ny; ++j) {
            for (size_t i = 1; i + 1 < nx; ++i) {
                double factxp = (kdint[aij(i + 1, j)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dx * dx);
                double factxm = (kdint[aij(i - 1, j)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dx * dx);
                double factyp = (kdint[aij(i, j + 1)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dy * dy);
                double factym = (kdint[aij(i, j - 1)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dy * dy);
                diag[i] = 1.0 + factxp + factxm;
                sup[i] = -factxp;
                inf[i] = -factxm;
                f[i] = zintp[aij(i, j)] + factyp * zintp[aij(i, j + 1)] - (factyp + factym) * zintp[aij(i, j)] +
                       factym * zintp[aij(i, j - 1)];
            }
            diag[0] = 1.0; sup[0] = 0.0; f[0] = zintp[aij(0, j)];
            diag[nx - 1] = 1.0; inf[nx - 1] = 0.0; f[nx - 1] = zintp[aij(nx - 1, j)];
            int rc = tridag(inf.data(), diag.data(), sup.data(), f.data(), res.data(), gam.data(), nx);
            if (rc) return rc;
            for (size_t i = 0; i < nx; ++i) zint[aij(i, j)] = res[i];
        }
    }
    {
        std::vector<double> f(ny, 0.0), diag(ny, 0.0), sup(ny, 0.0), inf(ny, 0.0), res(ny, 0.0), gam(ny, 0.0);
        for (size_t i = 1; i + 1 < nx; ++i) {
            for (size_t j = 1; j + 1 < ny; ++j) {
                double factxp = (kdint[aij(i + 1, j)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dx * dx);
                double factxm = (kdint[aij(i - 1, j)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dx * dx);
                double factyp = (kdint[aij(i, j + 1)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dy * dy);
                double factym = (kdint[aij(i, j - 1)] + kdint[aij(i, j)]) / 2.0 * (dt / 2.0) / (dy * dy);
                diag[j] = 1.0 + factyp + factym;
                sup[j] = -factyp;
                inf[j] = -factym;
                f[j] = zint[aij(i, j)] + factxp * zint[aij(i + 1, j)] - (factxp + factxm) * zint[aij(i, j)] +
                       factxm * zint[aij(i - 1, j)];
            }
            diag[0] = 1.0; sup[0] = 0.0; f[0] = zint[aij(i, 0)];
            diag[ny - 1] = 1.0; inf[ny - 1] = 0.0; f[ny - 1] = zint[aij(i, ny - 1)];
            int rc = tridag(inf.data(), diag.data(), sup.data(), f.data(), res.data(), gam.data(), ny);
            if (rc) return rc;
            for (size_t j = 0; j < ny; ++j) zintp[aij(i, j)] = res[j];
        }
    }
    for (size_t ij = 0; ij < nx * ny; ++ij) h[ij] = zintp[ij];
#pragma omp parallel for schedule(static)
    for (long long ij = 0; ij < (long long)(nx * ny); ++ij) b[ij] = std::min(h[ij], b[ij]);
    return 0;
}

// erosion.rs:1738-1811  fill_sinks (literal sweeps, including the stale `nh` in the inner loop)
int fill_sinks(const Grid& g, const double* oldh, const uint8_t* ocean, double* newh) {
    const double inf = std::numeric_limits<double>::infinity();
    for (size_t i = 0; i < g.n; ++i) newh[i] = ocean[i] ? oldh[i] : inf;
    int sweeps = 0;
    for (;;) {
        bool changed = false;
        ++sweeps;
        for (size_t posi = 0; posi < g.n; ++posi) {
            double nh = newh[posi], oh = oldh[posi];
            if (nh == oh) continue;
            int x = g.x_of(posi), y = g.y_of(posi);
            for (int k = 0; k < 8; ++k) {
                int xx = x + NDX[k], yy = y + NDY[k];
                if (!g.inb(xx, yy)) continue;
                double nbh = newh[g.idx(xx, yy)] + 0.0;
                if (oh >= nbh) { newh[posi] = oh; changed = true; break; }
                if (nh > nbh && nbh > oh) { newh[posi] = nbh; changed = true; }
            }
        }
        if (!changed) return sweeps;
    }
}

} // namespace

// =============================================================================================
static const char* TIMER_NAMES[] = {"downhill", "get_lakes", "multi_rec", "drainage", "max_slope", "k_factors",
                                    "gather", "gs_init", "gs_sediment", "gs_elev", "gs_solve", "scatter_basement",
                                    "lake_silt", "thermal", "diffusion", "total"};
enum { T_DOWNHILL, T_LAKES, T_MREC, T_DRAIN, T_SLOPE, T_KFACT, T_GATHER, T_GSINIT, T_GSSED, T_GSELEV, T_GSSOLVE,
       T_SCATTER, T_SILT, T_THERMAL, T_DIFF, T_TOTAL, T_COUNT };

struct orc_ctx {
    Grid g;
    orc_params P;
    Noise nz;
    std::vector<double> h, b, wh, kf, kd, hscale;
    std::vector<float> uplift, n, m, gg, eps0, alpha;
    std::vector<uint8_t> ocean;
    float max_uplift = 0, max_g = 0;
    // last-step intermediates
    std::vector<int32_t> dh, dh0, indirection, lake_sill;
    std::vector<uint32_t> newh, mstack;
    std::vector<uint8_t> mrec;
    std::vector<double> mwrec, area, max_slope, kd_eff, h_spl, b_spl, lake_silt, lwv, wh_mrec, h_pre_diff;
    std::vector<float> kfs, kdf;
    int n_gs = 0; double err = 0, maxh = 0; uint64_t boundary_len = 0; size_t n_lakes = 0, n_sources = 0, n_underwater = 0;
    long dag_levels = -1;
    unsigned long long newton_iters = 0, newton_nodes = 0, newton_max = 0;
    double timers[T_COUNT];
};

extern "C" {

void orc_default_params(orc_params* p) {
    const double PI = 3.14159265358979323846;
    p->k_d_scale = 0.5;
    p->k_da = 2.5 * std::pow(0.5, -2.0 * 0.2);
    p->chunk_area_pow_q = std::pow(1024.0, 0.2);
    p->min_max_angle = std::tan(15.0 / 360.0 * 2.0 * PI);
    p->max_max_angle = std::tan(60.0 / 360.0 * 2.0 * PI);
    p->mid_slope = std::tan(30.0 / 360.0 * 2.0 * PI);
    p->mountain_scale = 2048.0;
    p->use_libm = 0;
    p->n_threads = 0;
}

orc_ctx* orc_create(uint32_t x_lg, uint32_t y_lg) {
    orc_ctx* c = new orc_ctx();
    c->g = make_grid(x_lg, y_lg);
    orc_default_params(&c->P);
    std::memset(c->timers, 0, sizeof(c->timers));
    return c;
}
void orc_destroy(orc_ctx* c) { delete c; }

// erosion.rs:2549-2626 — materialise inputs, reductions
int orc_set_inputs(orc_ctx* c, const orc_inputs* in) {
    size_t N = c->g.n;
    c->h.resize(N); c->b.resize(N); c->wh.assign(N, 0.0);
    for (size_t i = 0; i < N; ++i) { c->h[i] = (double)in->oldh[i]; c->b[i] = (double)in->oldb[i]; }
    c->ocean.assign(in->is_ocean, in->is_ocean + N);
    c->uplift.assign(in->uplift, in->uplift + N);
    c->n.assign(in->n, in->n + N); c->m.assign(in->m, in->m + N); c->gg.assign(in->g, in->g + N);
    c->eps0.assign(in->epsilon_0, in->epsilon_0 + N); c->alpha.assign(in->alpha, in->alpha + N);
    c->kf.assign(in->kf, in->kf + N); c->kd.assign(in->kd, in->kd + N);
    c->hscale.assign(in->height_scale_cell, in->height_scale_cell + N);
    float mu = c->uplift[0], mg = c->gg[0];
    for (size_t i = 1; i < N; ++i) { if (c->uplift[i] > mu) mu = c->uplift[i]; if (c->gg[i] > mg) mg = c->gg[i]; }
    c->max_uplift = mu; c->max_g = mg;
    return 0;
}
void orc_set_params(orc_ctx* c, const orc_params* p) { c->P = *p; }
void orc_set_noise(orc_ctx* c, const orc_noise* nz) {
    c->nz.octaves = nz->octaves; c->nz.frequency = nz->frequency; c->nz.lacunarity = nz->lacunarity;
    c->nz.persistence = nz->persistence; c->nz.scale_factor = nz->scale_factor;
    c->nz.perm.assign(nz->perm, nz->perm + (size_t)nz->octaves * 256);
}
void orc_noise_eval(const orc_noise* nz, const double* xyz, double* out, size_t n) {
    Noise N;
    N.octaves = nz->octaves; N.frequency = nz->frequency; N.lacunarity = nz->lacunarity;
    N.persistence = nz->persistence; N.scale_factor = nz->scale_factor;
    N.perm.assign(nz->perm, nz->perm + (size_t)nz->octaves * 256);
    for (size_t i = 0; i < n; ++i) out[i] = fbm3(N, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
}
void orc_set_state(orc_ctx* c, const double* h, const double* b, const double* wh) {
    size_t N = c->g.n;
    if (h) c->h.assign(h, h + N);
    if (b) c->b.assign(b, b + N);
    if (wh) c->wh.assign(wh, wh + N);
}

// erosion.rs:699-1731  erode
int orc_step(orc_ctx* c) {
    const Grid& g = c->g; const size_t N = g.n; const orc_params& P = c->P;
    Math M; M.libm = (P.use_libm & 1) != 0; M.libmf = (P.use_libm & 2) != 0;
#ifdef _OPENMP
    if (P.n_threads > 0) omp_set_num_threads(P.n_threads);
#endif
    double* h = c->h.data(); double* b = c->b.data(); double* wh = c->wh.data();
    const float* uplift = c->uplift.data(); const float* n_f = c->n.data(); const float* m_f = c->m.data();
    const float* gf = c->gg.data();
    double* T = c->timers; std::memset(T, 0, sizeof(c->timers));
    double t_all = now_s(), t0;

    const double min_erosion_height = 0.0;
    const double dt = (double)c->max_uplift / 1e-3;                 // 735
    const double sediment_thickness = 1.0e-4 * dt;                   // 738
    const double chunk_area = 32.0 * 32.0;
    const double max_stable = 32.0 * 32.0 / dt;                      // 742
    const double q = 0.2, q_ = 1.5;
    const double k_da = P.k_da;
    const int nx = g.nx, ny = g.ny;
    const double dx = 32.0, dy = 32.0;
    const double p = 1.0, k_fs_mult_sed = 4.0, g_fs_mult_sed = 1.0;
    const double kdsed = 1.5e-2 / 4.0 * P.k_d_scale;                 // 2635-2636
    const double SQRT2048 = std::sqrt(32.0 * 32.0 + 32.0 * 32.0);
    auto ndist = [&](int k) { return (NDX[k] != 0 && NDY[k] != 0) ? SQRT2048 : 32.0; };

    c->dh.resize(N); c->dh0.resize(N); c->indirection.resize(N); c->newh.resize(N); c->mrec.resize(N);
    c->mstack.resize(N); c->mwrec.resize(N * 8); c->area.resize(N); c->max_slope.resize(N);
    int32_t* dh = c->dh.data();

    // 815-819
    t0 = now_s();
    {
        std::vector<uint8_t> pred(N);
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)N; ++i) pred[i] = (c->ocean[i] && h[i] <= 0.0) ? 1 : 0;
        downhill(g, h, pred.data(), dh);
    }
    c->dh0 = c->dh;
    T[T_DOWNHILL] = now_s() - t0;
    // 821-822
    t0 = now_s();
    int rc = get_lakes(g, h, dh, c->indirection.data(), c->newh.data(), &c->maxh, &c->boundary_len, &c->n_lakes);
    if (rc) return rc;
    T[T_LAKES] = now_s() - t0;
    // 824-836
    t0 = now_s();
    rc = get_multi_rec(g, M, h, dh, c->newh.data(), wh, c->mrec.data(), c->mstack.data(), c->mwrec.data(), dx, dy,
                       &c->n_sources);
    if (rc) return rc;
    c->wh_mrec.assign(wh, wh + N);
    T[T_MREC] = now_s() - t0;
    // 841
    t0 = now_s();
    get_multi_drainage(g, c->mstack.data(), c->mrec.data(), c->mwrec.data(), c->area.data());
    T[T_DRAIN] = now_s() - t0;
    // 845-856
    t0 = now_s();
    get_max_slope(g, M, P, c->nz, h, c->hscale.data(), c->max_slope.data());
    std::vector<double> max_slopes = c->max_slope; // mutated by the thermal pass
    std::vector<double> h_t(h, h + N);
    T[T_SLOPE] = now_s() - t0;

    const uint8_t* mrec = c->mrec.data(); const uint32_t* mstack = c->mstack.data();
    const double* mwrec = c->mwrec.data(); const double* area = c->area.data();
    const double mid_slope = P.mid_slope;

    // 870-1006: stream power factors
    t0 = now_s();
    c->kfs.assign(N * 8, 0.0f); c->kdf.assign(N * 8, 0.0f);
    float* k_fs_fact = c->kfs.data(); float* k_df_fact = c->kdf.data();
#pragma omp parallel for schedule(static)
    for (long long posi = 0; posi < (long long)N; ++posi) {
        if (dh[posi] < 0) continue;
        double old_b_i = b[posi];
        double sed = h_t[posi] - old_b_i;
        double kd_factor = 1.0;
        double k_fs = c->kf[posi] / kd_factor;
        double k = ((sed > sediment_thickness) ? (k_fs_mult_sed * k_fs) : k_fs) * dt;
        double n = (double)n_f[posi];
        double m = (double)m_f[posi];
        uint8_t mr = mrec[posi];
        for (int kk = 0; kk < 8; ++kk) {
            if (!((mr >> kk) & 1)) continue;
            double neighbor_distance = ndist(kk);
            double knew = k * M.pow_(p * chunk_area * (area[posi] * mwrec[posi * 8 + kk]), m) /
                          M.pow_(neighbor_distance, n);
            k_fs_fact[posi * 8 + kk] = (float)knew;
        }
    }
#pragma omp parallel for schedule(static)
    for (long long posi = 0; posi < (long long)N; ++posi) {
        if (dh[posi] < 0) continue;
        double uplift_i = (double)uplift[posi];
        double area_i = area[posi];
        double max_slope = max_slopes[posi];
        double chunk_area_pow = P.chunk_area_pow_q;
        double old_b_i = b[posi];
        double sed = h_t[posi] - old_b_i;
        double g_i = (sed > sediment_thickness) ? (double)(float)(g_fs_mult_sed * gf[posi]) : (double)gf[posi];
        double r0 = max_slope / mid_slope;
        double kd_factor = r0 * r0;
        double k_da_i = k_da * kd_factor;
        uint8_t mr = mrec[posi];
        for (int kk = 0; kk < 8; ++kk) {
            if (!((mr >> kk) & 1)) continue;
            double mwrec_kk = mwrec[posi * 8 + kk];
            const double chunk_neutral_area = 0.1e6;
            double k = (mwrec_kk * (uplift_i + (double)c->max_uplift * g_i / p)) /
                       (1.0 + k_da_i * M.pow_(mwrec_kk * chunk_neutral_area, q)) / M.pow_(max_slope, q_);
            double neighbor_distance = ndist(kk);
            double knew = k * (1.0 + k_da_i * chunk_area_pow * M.pow_(area_i * mwrec_kk, q)) /
                          M.pow_(neighbor_distance, q_);
            k_df_fact[posi * 8 + kk] = (float)knew;
        }
    }
    T[T_KFACT] = now_s() - t0;

    // 1010-1074
    std::vector<double> lake_water_volume(N, 0.0), elev(N, 0.0), h_p(N, 0.0), deltah(N, 0.0), lake_silt(N, 0.0);
    std::vector<int64_t> lake_sill(N, -1);
    const double tol1 = 1.0e-4 * (c->maxh + 1.0);
    const double tol2 = 1.0e-3 * ((double)c->max_uplift + 1.0);
    const double tol = std::max(tol1, tol2);
    double err = 2.0 * tol;
    int n_gs = 0; const int max_n_gs = 99;
    t0 = now_s();
    std::vector<size_t> mstack_inv(N);
    std::vector<double> h_t_stack(N), h_stack(N), b_stack(N), area_stack(N);
    std::vector<int32_t> dh_stack(N);
    for (size_t stacki = 0; stacki < N; ++stacki) {
        size_t posi = mstack[stacki];
        mstack_inv[posi] = stacki; dh_stack[stacki] = dh[posi]; h_t_stack[stacki] = h_t[posi];
        h_stack[stacki] = h[posi]; b_stack[stacki] = b[posi]; area_stack[stacki] = area[posi];
    }
    T[T_GATHER] = now_s() - t0;
    size_t n_underwater = 0;
    c->newton_iters = c->newton_nodes = c->newton_max = 0;

    while (err > tol && n_gs < max_n_gs) {
        n_gs += 1;
        // 1096-1116
        t0 = now_s();
#pragma omp parallel for schedule(static)
        for (long long s = 0; s < (long long)N; ++s) {
            h_p[s] = h_stack[s];
            size_t posi = mstack[s];
            double uplift_i = (double)uplift[posi];
            deltah[s] = h_t_stack[s] + uplift_i - h_stack[s];
        }
        T[T_GSINIT] += now_s() - t0;
        // 1127-1151: sediment sweep, stack order
        t0 = now_s();
        for (size_t stacki = 0; stacki < N; ++stacki) {
            size_t posi = mstack[stacki];
            int32_t posj = dh_stack[stacki];
            double deltah_i = deltah[stacki];
            if (posj < 0) { lake_silt[stacki] = deltah_i; continue; }
            double uplift_i = (double)uplift[posi];
            double uphill_deltah_i = deltah_i - ((h_t_stack[stacki] + uplift_i) - h_p[stacki]);
            int64_t lposi = lake_sill[stacki];
            if (lposi == (int64_t)stacki) lake_silt[stacki] = (uphill_deltah_i <= 0.0) ? 0.0 : uphill_deltah_i;
            int x = g.x_of(posi), y = g.y_of(posi);
            uint8_t mr = mrec[posi];
            for (int k = 0; k < 8; ++k) {
                if (!((mr >> k) & 1)) continue;
                size_t pj = g.idx(x + NDX[k], y + NDY[k]);
                deltah[mstack_inv[pj]] += deltah_i * mwrec[posi * 8 + k];
            }
        }
        T[T_GSSED] += now_s() - t0;
        // 1178-1227: elev
        t0 = now_s();
#pragma omp parallel for schedule(static)
        for (long long s = 0; s < (long long)N; ++s) {
            size_t posi = mstack[s];
            double uplift_i = (double)uplift[posi];
            if (dh_stack[s] < 0) { elev[s] = h_t_stack[s] + uplift_i; continue; }
            double old_h_after_uplift_i = h_t_stack[s] + uplift_i;
            double area_i = area_stack[s];
            double uphill_silt_i = deltah[s] - (old_h_after_uplift_i - h_p[s]);
            double sed = h_t_stack[s] - b_stack[s];
            double g_i = (sed > sediment_thickness) ? (double)(float)(g_fs_mult_sed * gf[posi]) : (double)gf[posi];
            double g_i_ratio = g_i / (p * area_i);
            elev[s] = old_h_after_uplift_i + uphill_silt_i * g_i_ratio;
        }
        T[T_GSELEV] += now_s() - t0;
        // 1237-1450: implicit solve, reverse stack order
        t0 = now_s();
        double sum_err = 0.0;
        n_underwater = 0;
        for (size_t stacki = N; stacki-- > 0;) {
            size_t posi = mstack[stacki];
            double old_elev_i = elev[stacki];
            double h_t_i = h_t_stack[stacki];
            int32_t posj_s = dh_stack[stacki];
            if (posj_s < 0) {
                if (posj_s == -1) return -40; // panic!("Disconnected lake!")
                double uplift_i = (double)uplift[posi];
                wh[posi] = std::max(min_erosion_height, h_t_i + uplift_i);
                lake_sill[stacki] = (int64_t)posi;
                lake_water_volume[stacki] = 0.0;
            } else {
                size_t posj = (size_t)posj_s;
                double wh_j = wh[posj];
                double old_h_i = h_stack[stacki];
                double new_h_i = old_h_i;
                if (old_elev_i > wh_j) {
                    double dtherm = 0.0;
                    double h0 = old_elev_i + dtherm;
                    double n = (double)n_f[posi];
                    const float* kdf = &k_df_fact[posi * 8];
                    const float* kfs = &k_fs_fact[posi * 8];
                    int x = g.x_of(posi), y = g.y_of(posi);
                    uint8_t mr = mrec[posi];
                    if (std::fabs(n - 1.0) <= 1.0e-3 && std::fabs(q_ - 1.0) <= 1.0e-3) {
                        double f = h0, df = 1.0;
                        for (int kk = 0; kk < 8; ++kk) {
                            if (!((mr >> kk) & 1)) continue;
                            size_t pj = g.idx(x + NDX[kk], y + NDY[kk]);
                            double h_j = h_stack[mstack_inv[pj]];
                            if (old_elev_i > h_j) {
                                double fact = (double)kfs[kk] + (double)kdf[kk];
                                f += fact * h_j; df += fact;
                            }
                        }
                        new_h_i = f / df;
                    } else {
                        double omega1 = 0.875 * n, omega2 = 0.875 / q_;
                        double omega = std::max(omega1, omega2);
                        const double tolp = 1.0e-3;
                        double errp = 2.0 * tolp;
                        float rec_heights[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                        bool mask[8] = {false, false, false, false, false, false, false, false};
                        for (int kk = 0; kk < 8; ++kk) {
                            if (!((mr >> kk) & 1)) continue;
                            size_t pj = g.idx(x + NDX[kk], y + NDY[kk]);
                            double h_j = h_stack[mstack_inv[pj]];
                            if (old_elev_i > h_j) { mask[kk] = true; rec_heights[kk] = (float)h_j; }
                        }
                        c->newton_nodes++; unsigned long long nit = 0;
                        while (errp > tolp) {
                            nit++; c->newton_iters++;
                            double f = new_h_i - h0;
                            double df = 1.0;
                            for (int kk = 0; kk < 8; ++kk) {
                                if (!mask[kk]) continue;
                                float elev_j = rec_heights[kk];
                                float dhv = std::max(0.0f, (float)new_h_i - elev_j);
                                float dh_fs_sample = kfs[kk] * M.powf_(dhv, (float)n - 1.0f);
                                float dh_df_sample = kdf[kk] * M.powf_(dhv, (float)q_ - 1.0f);
                                f += (double)((dh_fs_sample + dh_df_sample) * dhv);
                                df += (double)((float)n * dh_fs_sample + (float)q_ * dh_df_sample);
                            }
                            double hn = new_h_i - f / df;
                            errp = std::fabs(hn - new_h_i);
                            new_h_i = new_h_i * (1.0 - omega) + hn * omega;
                        }
                        if (nit > c->newton_max) c->newton_max = nit;
                    }
                    lake_sill[stacki] = (int64_t)posi;
                    lake_water_volume[stacki] = 0.0;
                    if (new_h_i <= wh_j) new_h_i = wh_j;
                } else {
                    new_h_i = old_elev_i;
                    size_t posj_stack = mstack_inv[posj];
                    int64_t lposj = lake_sill[posj_stack];
                    lake_sill[stacki] = lposj;
                    if (lposj >= 0) lake_water_volume[(size_t)lposj] += (wh_j - old_elev_i);
                    n_underwater++;
                }
                wh[posi] = std::max(wh_j, new_h_i);
                h_stack[stacki] = new_h_i;
            }
            double d = 0.0 + h_stack[stacki] - h_p[stacki];
            sum_err += d * d;
        }
        T[T_GSSOLVE] += now_s() - t0;
        err = std::sqrt(sum_err / (double)N);
        if (c->max_g == 0.0f) err = 0.0;
    }
    c->n_gs = n_gs; c->err = err; c->n_underwater = n_underwater;

    // 1469-1475 scatter, 1483-1531 basement
    t0 = now_s();
#pragma omp parallel for schedule(static)
    for (long long posi = 0; posi < (long long)N; ++posi) h[posi] = h_stack[mstack_inv[posi]];
    c->h_spl.assign(h, h + N);
    {
        std::vector<double> bnew(N);
#pragma omp parallel for schedule(static)
        for (long long posi = 0; posi < (long long)N; ++posi) {
            double h_i = h[posi];
            double old_b_i = b[posi];
            double uplift_i = (double)uplift[posi];
            double new_b_i = old_b_i + uplift_i;
            int32_t posj = dh[posi];
            double vertical_sed = std::max(h_i - new_b_i, 0.0);
            double h_normal;
            if (posj < 0) h_normal = vertical_sed;
            else {
                double h_j = h[posj];
                double ex = (double)(g.x_of(posi) - g.x_of(posj)) * 32.0;
                double ey = (double)(g.y_of(posi) - g.y_of(posj)) * 32.0;
                double nd2 = ex * ex + ey * ey;
                double dhh = h_i - h_j;
                double h_i_fact = dhh / (nd2 + dhh * dhh);
                double h_i_vertical = 1.0 - h_i_fact * dhh;
                h_normal = vertical_sed * std::sqrt(h_i_fact * h_i_fact * nd2 + h_i_vertical * h_i_vertical);
            }
            double p_i = (double)c->eps0[posi] * dt * M.exp_(-(double)c->alpha[posi] * h_normal);
            new_b_i -= p_i;
            new_b_i = std::min(new_b_i, h_i);
            bnew[posi] = new_b_i;
        }
        std::memcpy(b, bnew.data(), N * sizeof(double));
    }
    c->b_spl.assign(b, b + N);
    T[T_SCATTER] = now_s() - t0;
    // 1535-1556 lake silt
    t0 = now_s();
    if (c->max_g > 0.0f) {
#pragma omp parallel for schedule(static)
        for (long long posi = 0; posi < (long long)N; ++posi) {
            size_t stacki = mstack_inv[posi];
            int64_t lposi = lake_sill[stacki];
            if (lposi >= 0) {
                double vol = lake_water_volume[(size_t)lposi];
                if (vol > 0.0) {
                    h[posi] += std::max(0.0, std::min(lake_silt[stacki], vol)) / vol * (wh[posi] - h[posi]);
                }
            }
        }
    }
    T[T_SILT] = now_s() - t0;
    // keep the quirky arrays for inspection (stack-indexed, as in the reference)
    c->lake_silt = lake_silt; c->lwv = lake_water_volume;
    c->lake_sill.resize(N);
    for (size_t i = 0; i < N; ++i) c->lake_sill[i] = (int32_t)lake_sill[i];

    // 1597-1696 thermal pass (dtherm = 0: only kd_eff and wh change)
    t0 = now_s();
    const uint32_t* newh = c->newh.data();
    for (size_t t = 0; t < N; ++t) {
        size_t posi = newh[t];
        double old_h_i = h[posi], old_b_i = b[posi];
        double sed = old_h_i - old_b_i;
        double max_slope = max_slopes[posi];
        max_slopes[posi] = (sed > sediment_thickness && kdsed > 0.0) ? kdsed : c->kd[posi];
        int32_t posj_s = dh[posi];
        if (posj_s < 0) {
            if (posj_s == -1) return -40;
            wh[posi] = std::max(min_erosion_height, old_h_i);
        } else {
            size_t posj = (size_t)posj_s;
            double wh_j = wh[posj];
            double new_h_i = old_h_i;
            if (wh_j < old_h_i) {
                double h_j = h[posj];
                double h_k = h_j;
                double ex = (double)(g.x_of(posi) - g.x_of(posj)) * 32.0;
                double ey = (double)(g.y_of(posi) - g.y_of(posj)) * 32.0;
                double neighbor_distance = std::sqrt(ex * ex + ey * ey);
                double dz = std::max(new_h_i - h_k, 0.0);
                double mag_slope = dz / neighbor_distance;
                if (mag_slope >= max_slope) {
                    // dtherm = 0: nothing changes
                } else {
                    double sr0 = mag_slope / max_slope;
                    double slope_ratio = sr0 * sr0;
                    double om = 1.0 - slope_ratio;
                    double slope_nonlinear_factor = slope_ratio * ((3.0 - slope_ratio) / (om * om));
                    max_slopes[posi] += std::min(max_slopes[posi] * slope_nonlinear_factor, max_stable);
                }
            }
            wh[posi] = std::max(wh_j, new_h_i);
        }
    }
    c->kd_eff = max_slopes;
    c->h_pre_diff.assign(h, h + N);
    T[T_THERMAL] = now_s() - t0;
    // 1717-1728
    t0 = now_s();
    rc = diffusion((size_t)nx, (size_t)ny, (double)nx * dx, (double)ny * dy, dt, h, b, max_slopes.data(), -1.0);
    if (rc) return rc;
    T[T_DIFF] = now_s() - t0;
    T[T_TOTAL] = now_s() - t_all;
    c->dag_levels = -1;
    return 0;
}

int orc_run(orc_ctx* c, uint32_t n_steps, double* out_h, double* out_b) {
    for (uint32_t i = 0; i < n_steps; ++i) { int rc = orc_step(c); if (rc) return rc; }
    size_t N = c->g.n;
    if (out_h) std::memcpy(out_h, c->h.data(), N * 8);
    if (out_b) std::memcpy(out_b, c->b.data(), N * 8);
    return 0;
}

static long compute_dag_levels(orc_ctx* c) {
    // longest path (in nodes) of the multi-receiver DAG, following mstack order
    const Grid& g = c->g; size_t N = g.n;
    if (c->mstack.size() != N) return -1;
    std::vector<int32_t> lvl(N, 0); long mx = 0;
    for (size_t t = 0; t < N; ++t) {
        size_t i = c->mstack[t]; int x = g.x_of(i), y = g.y_of(i); uint8_t mr = c->mrec[i];
        for (int k = 0; k < 8; ++k) if ((mr >> k) & 1) {
            size_t r = g.idx(x + NDX[k], y + NDY[k]);
            if (lvl[r] < lvl[i] + 1) lvl[r] = lvl[i] + 1;
        }
        if (lvl[i] > mx) mx = lvl[i];
    }
    return mx + 1;
}

size_t orc_array_bytes(orc_ctx* c, int which) {
    size_t N = c->g.n;
    switch (which) {
        case ORC_H: case ORC_B: case ORC_WH: case ORC_AREA: case ORC_MAX_SLOPE: case ORC_H_SPL: case ORC_LAKE_SILT:
        case ORC_LAKE_WATER_VOLUME: case ORC_KD_EFF: case ORC_WH_MREC: case ORC_B_SPL: case ORC_H_PRE_DIFF: return N * 8;
        case ORC_MWREC: return N * 64;
        case ORC_DH: case ORC_DH0: case ORC_INDIRECTION: case ORC_NEWH: case ORC_MSTACK: case ORC_LAKE_SILL: return N * 4;
        case ORC_MREC: return N;
        case ORC_KFS: case ORC_KDF: return N * 32;
    }
    return 0;
}

int orc_get_array(orc_ctx* c, int which, void* dst) {
    const void* src = nullptr;
    switch (which) {
        case ORC_H: src = c->h.data(); break;            case ORC_B: src = c->b.data(); break;
        case ORC_WH: src = c->wh.data(); break;          case ORC_DH: src = c->dh.data(); break;
        case ORC_INDIRECTION: src = c->indirection.data(); break; case ORC_NEWH: src = c->newh.data(); break;
        case ORC_MREC: src = c->mrec.data(); break;      case ORC_MSTACK: src = c->mstack.data(); break;
        case ORC_MWREC: src = c->mwrec.data(); break;    case ORC_AREA: src = c->area.data(); break;
        case ORC_MAX_SLOPE: src = c->max_slope.data(); break; case ORC_KFS: src = c->kfs.data(); break;
        case ORC_KDF: src = c->kdf.data(); break;        case ORC_H_SPL: src = c->h_spl.data(); break;
        case ORC_LAKE_SILT: src = c->lake_silt.data(); break; case ORC_LAKE_SILL: src = c->lake_sill.data(); break;
        case ORC_LAKE_WATER_VOLUME: src = c->lwv.data(); break; case ORC_KD_EFF: src = c->kd_eff.data(); break;
        case ORC_DH0: src = c->dh0.data(); break;        case ORC_WH_MREC: src = c->wh_mrec.data(); break;
        case ORC_B_SPL: src = c->b_spl.data(); break;    case ORC_H_PRE_DIFF: src = c->h_pre_diff.data(); break;
        default: return -1;
    }
    size_t nb = orc_array_bytes(c, which);
    if (!src || nb == 0) return -1;
    std::memcpy(dst, src, nb);
    return 0;
}

double orc_get_scalar(orc_ctx* c, int which) {
    switch (which) {
        case 0: return c->n_gs; case 1: return c->err; case 2: return c->maxh; case 3: return (double)c->boundary_len;
        case 4: return (double)c->n_lakes;
        case 5: if (c->dag_levels < 0) c->dag_levels = compute_dag_levels(c); return (double)c->dag_levels;
        case 6: return (double)c->n_underwater; case 7: return (double)c->n_sources;
        case 8: return (double)c->max_uplift; case 9: return (double)c->max_g;
        case 10: return (double)c->newton_iters; case 11: return (double)c->newton_nodes; case 12: return (double)c->newton_max;
    }
    return NAN;
}
int orc_get_timers(orc_ctx* c, double* dst, int cap) {
    int n = std::min(cap, (int)T_COUNT);
    for (int i = 0; i < n; ++i) dst[i] = c->timers[i];
    return T_COUNT;
}
const char* orc_timer_name(int i) { return (i >= 0 && i < T_COUNT) ? TIMER_NAMES[i] : ""; }

void orc_downhill(uint32_t x_lg, uint32_t y_lg, const double* h, const uint8_t* ocean, int32_t* dh) {
    downhill(make_grid(x_lg, y_lg), h, ocean, dh);
}
int orc_get_lakes(uint32_t x_lg, uint32_t y_lg, const double* h, int32_t* dh, int32_t* indirection, uint32_t* newh,
                  double* maxh, uint64_t* boundary_len) {
    return get_lakes(make_grid(x_lg, y_lg), h, dh, indirection, newh, maxh, boundary_len, nullptr);
}
void orc_get_multi_rec(uint32_t x_lg, uint32_t y_lg, const double* h, const int32_t* dh, const uint32_t* newh,
                       double* wh, uint8_t* mrec, uint32_t* mstack, double* mwrec, int use_libm) {
    Math M; M.libm = (use_libm & 1) != 0; M.libmf = (use_libm & 2) != 0;
    get_multi_rec(make_grid(x_lg, y_lg), M, h, dh, newh, wh, mrec, mstack, mwrec, 32.0, 32.0, nullptr);
}
void orc_get_multi_drainage(uint32_t x_lg, uint32_t y_lg, const uint32_t* mstack, const uint8_t* mrec,
                            const double* mwrec, double* area) {
    get_multi_drainage(make_grid(x_lg, y_lg), mstack, mrec, mwrec, area);
}
void orc_get_drainage(uint32_t x_lg, uint32_t y_lg, const uint32_t* newh, const int32_t* dh, float* flux) {
    get_drainage(make_grid(x_lg, y_lg), newh, dh, flux);
}
int orc_fill_sinks(uint32_t x_lg, uint32_t y_lg, const double* h, const uint8_t* ocean, double* out) {
    return fill_sinks(make_grid(x_lg, y_lg), h, ocean, out);
}
void orc_diffusion(uint32_t nx, uint32_t ny, double xl, double yl, double dt, double* h, double* b, const double* kd,
                   double kdsed) {
    diffusion(nx, ny, xl, yl, dt, h, b, kd, kdsed);
}
// util.rs:332-367  get_oceans (stack flood fill from the border over cells with h <= 0)
void orc_get_oceans(uint32_t x_lg, uint32_t y_lg, const double* h, uint8_t* is_ocean) {
    Grid g = make_grid(x_lg, y_lg);
    std::memset(is_ocean, 0, g.n);
    std::vector<size_t> stack;
    auto push = [&](int x, int y) { size_t p = g.idx(x, y); if (h[p] <= 0.0) stack.push_back(p); };
    for (int x = 0; x < g.nx; ++x) { push(x, 0); push(x, g.ny - 1); }
    for (int y = 1; y < g.ny - 1; ++y) { push(0, y); push(g.nx - 1, y); }
    while (!stack.empty()) {
        size_t c = stack.back(); stack.pop_back();
        if (is_ocean[c]) continue;
        is_ocean[c] = 1;
        int x = g.x_of(c), y = g.y_of(c);
        for (int k = 0; k < 8; ++k) {
            int xx = x + NDX[k], yy = y + NDY[k];
            if (!g.inb(xx, yy)) continue;
            size_t p = g.idx(xx, yy);
            if (h[p] <= 0.0) stack.push_back(p);
        }
    }
}
// mod.rs:1502-1543 water_height_initial, 1550 fill_sinks, 1566-1592 final water_alt (f32)
void orc_water_alt(uint32_t x_lg, uint32_t y_lg, const double* alt, const uint8_t* is_ocean, const int32_t* dh,
                   const int32_t* indirection, double* water_alt_filled, float* water_alt_final) {
    Grid g = make_grid(x_lg, y_lg);
    std::vector<double> init(g.n);
    for (size_t i = 0; i < g.n; ++i) {
        int32_t ii = indirection[i];
        size_t lake_idx = ii < 0 ? i : (size_t)ii;
        double chunk_water_alt;
        if (dh[lake_idx] < 0) chunk_water_alt = 0.0;
        else {
            size_t pass_idx = (size_t)(-(int64_t)indirection[lake_idx]);
            double pass_height_i = alt[pass_idx];
            int32_t neighbor_pass_idx = dh[pass_idx];
            double pass_height_j = alt[neighbor_pass_idx];
            chunk_water_alt = std::max(pass_height_i, pass_height_j);
        }
        init[i] = std::max(alt[i], chunk_water_alt);
    }
    fill_sinks(g, init.data(), is_ocean, water_alt_filled);
    for (size_t i = 0; i < g.n; ++i) {
        int32_t ii = indirection[i];
        size_t lake_idx = ii < 0 ? i : (size_t)ii;
        water_alt_final[i] = dh[lake_idx] < 0 ? 0.0f : (float)water_alt_filled[i];
    }
}
// erosion.rs:219-545  get_rivers.  Output record mirrors RiverData (erosion.rs:153-185); kind: 0 None,
// 1 Ocean, 2 Lake (aux = neighbor_pass_pos as i32 bits), 3 River (aux = cross_section).  neighbor_rivers is
// never filled by the reference and is omitted.
int orc_get_rivers(uint32_t x_lg, uint32_t y_lg, double continent_scale_hack, const uint32_t* newh, const double* water_alt,
                   const int32_t* downhill_, const int32_t* indirection, const double* drainage, orc_river* rivers, int use_libm) {
    Grid g = make_grid(x_lg, y_lg);
    Math M; M.libm = (use_libm & 1) != 0; M.libmf = (use_libm & 2) != 0;
    const double rainfall_chunk_rate = (double)(1.0f / (512.0f * 32.0f * 32.0f));   // config.rs:73
    const double river_roughness = (double)0.06125f, river_max_width = (double)2.0f, river_min_height = (double)0.25f,
                 river_width_to_depth = (double)8.0f;                                  // config.rs:74-77
    std::memset(rivers, 0, g.n * sizeof(orc_river));
    const double ncx = 32.0, ncy = 32.0;
    const double mins_per_sec = 1.0 / (continent_scale_hack * continent_scale_hack);
    const double chunk_area_factor = ncx * ncy * mins_per_sec;
    const double derivative_divisor = 1.0;
    for (size_t t = g.n; t-- > 0;) {
        size_t chunk_idx = newh[t];
        int32_t downhill_s = downhill_[chunk_idx];
        if (downhill_s < 0) { rivers[chunk_idx].kind = 1; continue; }
        size_t downhill_idx = (size_t)downhill_s;
        double dx = (double)(g.x_of(downhill_idx) - g.x_of(chunk_idx)), dy = (double)(g.y_of(downhill_idx) - g.y_of(chunk_idx));
        double ndx = ncx * dx, ndy = ncy * dy;
        double chunk_drainage = drainage[chunk_idx];
        double volumetric_flow_rate = chunk_drainage * chunk_area_factor * rainfall_chunk_rate;
        double downhill_drainage = drainage[downhill_idx];
        double incoming_drainage = downhill_drainage - 1.0;
        // get_river_spline_derivative
        double psx = (double)rivers[chunk_idx].spline_derivative[0], psy = (double)rivers[chunk_idx].spline_derivative[1];
        double mdx = ndx - ncx * 2.0 * std::sqrt(2.0), mdy = ndy - ncy * 2.0 * std::sqrt(2.0);
        double e0 = std::fabs(psx / mdx), e1 = std::fabs(psy / mdy);
        double extra_divisor = (e0 > e1) ? e0 : e1;   // reduce_partial_max
        float rsx, rsy;
        if (extra_divisor > 1.0) { rsx = (float)(psx / extra_divisor); rsy = (float)(psy / extra_divisor); }
        else { rsx = (float)psx; rsy = (float)psy; }
        int32_t indirection_idx = indirection[chunk_idx];
        size_t lake_idx;
        if (indirection_idx < 0) {
            size_t pass_idx = (size_t)(-(int64_t)indirection_idx);
            size_t neighbor_pass_idx = (size_t)downhill_[pass_idx];
            rivers[neighbor_pass_idx].kind = 3; rivers[neighbor_pass_idx].aux[0] = 0; rivers[neighbor_pass_idx].aux[1] = 0;
            lake_idx = chunk_idx;
        } else lake_idx = (size_t)indirection_idx;
        size_t pass_idx = downhill_[lake_idx] < 0 ? lake_idx : (size_t)(-(int64_t)indirection[lake_idx]);
        if (pass_idx != downhill_idx) {
            double wx = (ndx * 2.0 - (double)rsx) / derivative_divisor * chunk_drainage / incoming_drainage;
            double wy = (ndy * 2.0 - (double)rsy) / derivative_divisor * chunk_drainage / incoming_drainage;
            rivers[downhill_idx].spline_derivative[0] += (float)wx;
            rivers[downhill_idx].spline_derivative[1] += (float)wy;
        }
        int32_t neighbor_pass_s = downhill_[pass_idx];
        double chunk_water_alt = water_alt[chunk_idx];
        if (neighbor_pass_s >= 0) {
            double lake_water_alt = water_alt[lake_idx];
            if (chunk_water_alt == lake_water_alt) {
                size_t tgt = (pass_idx == chunk_idx) ? downhill_idx : pass_idx;
                orc_river& lake = rivers[chunk_idx];
                lake.spline_derivative[0] = rsx; lake.spline_derivative[1] = rsy;
                lake.kind = 2;
                int32_t px = g.x_of(tgt) * 32, py = g.y_of(tgt) * 32;
                std::memcpy(&lake.aux[0], &px, 4); std::memcpy(&lake.aux[1], &py, 4);
                continue;
            }
        }
        double downhill_water_alt = water_alt[downhill_idx];
        double neighbor_distance = std::sqrt(ndx * ndx + ndy * ndy);
        double dz = downhill_water_alt - chunk_water_alt;
        double slope = std::fabs(dz) / neighbor_distance;
        if (slope == 0.0) return -50; // panic!("Should this happen at all?")
        double slope_sqrt = std::sqrt(slope);
        double almost_velocity = volumetric_flow_rate / slope_sqrt * river_roughness;
        double width = 5.0 * M.pow_(river_width_to_depth * M.pow_(river_width_to_depth + 2.0, 2.0 / 3.0), 3.0 / 8.0) *
                       M.pow_(volumetric_flow_rate, 3.0 / 8.0) * M.pow_(slope, -3.0 / 16.0) * M.pow_(river_roughness, 3.0 / 8.0);
        width = std::max(width, 0.0);
        double height = (width == 0.0) ? river_min_height : M.pow_(almost_velocity / width, 3.0 / 5.0);
        double sgn = dz > 0.0 ? 1.0 : (dz < 0.0 ? -1.0 : (std::signbit(dz) ? -1.0 : 1.0));
        double rdx = ndx, rdy = ndy, rdz = sgn * dz;
        bool is_river = rivers[chunk_idx].kind == 3 || (width >= 0.5 && height >= river_min_height);
        if (is_river) {
            orc_river& dr = rivers[downhill_idx];
            dr.kind = 3; dr.aux[0] = 0; dr.aux[1] = 0;
            double max_width = 32.0 * river_max_width;
            if (width > max_width) { width = max_width; height = M.pow_(almost_velocity / width, 3.0 / 5.0); }
        }
        double velocity_magnitude = 1.0 / river_roughness * M.pow_(height, 2.0 / 3.0) * slope_sqrt;
        float csx = (float)width, csy = (float)height;
        double mag = std::sqrt(rdx * rdx + rdy * rdy + rdz * rdz);
        double vx = rdx / mag * velocity_magnitude, vy = rdy / mag * velocity_magnitude, vz = rdz / mag * velocity_magnitude;
        orc_river& r = rivers[chunk_idx];
        r.velocity[0] = (float)vx; r.velocity[1] = (float)vy; r.velocity[2] = (float)vz;
        r.spline_derivative[0] = rsx; r.spline_derivative[1] = rsy;
        if (is_river) { r.kind = 3; std::memcpy(&r.aux[0], &csx, 4); std::memcpy(&r.aux[1], &csy, 4); }
        else { r.kind = 0; r.aux[0] = 0; r.aux[1] = 0; }
    }
    return 0;
}
double orc_vm(int fn, double x, double y) {
    switch (fn) {
        case 0: return vm_log(x); case 1: return vm_exp(x); case 2: return vm_pow(x, y); case 3: return vm_log1p(x);
        case 4: return vm_tanh(x); case 5: return (double)vm_powf((float)x, (float)y);
    }
    return NAN;
}

} // extern "C"
