// ve_lakes.cu — get_lakes (world/src/sim/erosion.rs:1825-2293) on the GPU.
//
// Reference phases and their GPU formulation:
//  (i)   1865-1902  every dh<0 cell roots a tree; BFS order only matters as a tie-break
//        -> roots by pointer jumping, (depth, pre-order) keys from an Euler tour + list ranking.
//  (ii)  1915-2019  one representative pass per adjacent lake pair: min (pass height, BFS
//        position of the cell in the higher-rooted lake, neighbour slot), strict '<'
//        -> three passes of atomicMin into a hash table keyed by the lake pair.
//  (iii)/(iv) 2028-2137  Prim-style growth from the boundary lakes with a heap keyed
//        (pass height, from, to): order-dependent under ties -> emulated literally by ONE warp
//        with a 32-ary heap (children scanned by the 32 lanes, min by shuffle).
//  (v)   2152-2290  per-lake LIFO flood that re-routes flooded cells; lakes are independent
//        -> one thread per lake; then the final BFS order = stable sort by (block, depth) of the
//        pre-order of the final forest (Euler tour + list ranking again).
#include "ve_internal.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cooperative_groups.h>
#include <algorithm>
#include <string.h>
#include <stdlib.h>
#include <chrono>

__device__ __forceinline__ int l_slot_of(int dx, int dy) { int k = (dy + 1) * 3 + (dx + 1); return k < 4 ? k : k - 1; }
__device__ __forceinline__ uint64_t l_order_bits(double x) {
    x = x + 0.0; // -0.0 and +0.0 compare equal in the reference's f64 comparisons
    uint64_t u = (uint64_t)__double_as_longlong(x);
    return (u & 0x8000000000000000ULL) ? ~u : (u | 0x8000000000000000ULL);
}
__device__ __forceinline__ uint64_t l_hash(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33; return k;
}

// ---- (i) roots -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_root_init(uint32_t n, const int32_t* __restrict__ dh, uint32_t* __restrict__ root) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { int32_t d = dh[i]; root[i] = d < 0 ? i : (uint32_t)d; }
}
__global__ void __launch_bounds__(256) k_root_jump(uint32_t n, uint32_t* root, int* __restrict__ changed) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t r = root[i];
    uint32_t rr = root[r];
    if (rr != r) { root[i] = rr; *changed = 1; }
}
// all rounds of the root search in one cooperative launch: in-place jumping (roots are fixed points, so a pointer read
// mid-update is only ever further along), a grid barrier per round, stop after the first round that moved nothing
__global__ void __launch_bounds__(256) k_root_jump_all(uint32_t n, uint32_t* root, uint32_t* flags, int max_rounds) {
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const uint32_t stride = gridDim.x * blockDim.x;
    for (int r = 0; r < max_rounds; ++r) {
        bool moved = false;
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const uint32_t a = __ldcg(root + i);
            const uint32_t b = __ldcg(root + a);
            if (b != a) { root[i] = b; moved = true; }
        }
        if (__syncthreads_or(moved) && threadIdx.x == 0) flags[r] = 1u;
        grid.sync();
        if (__ldcg(flags + r) == 0u) break;
    }
}
__global__ void __launch_bounds__(256) k_lake_stats(uint32_t n, const int32_t* __restrict__ dh, const uint32_t* __restrict__ root,
                                                    const double* __restrict__ h, uint32_t* __restrict__ size,
                                                    unsigned long long* __restrict__ maxh_bits, uint32_t* __restrict__ counts) {
    // warp-aggregated: a million atomics on the same three words took 0.7 ms
    const unsigned full = 0xffffffffu;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < n;
    if (in) atomicAdd(size + root[i], 1u);
    unsigned long long hb = in ? (unsigned long long)__double_as_longlong(fabs(h[i])) : 0ull; // non-negative: bits are monotone
    uint32_t hi = (uint32_t)(hb >> 32), lo = (uint32_t)hb;
    uint32_t mhi = __reduce_max_sync(full, hi);
    uint32_t mlo = __reduce_max_sync(full, hi == mhi ? lo : 0u);
    int32_t d = in ? dh[i] : 0;
    unsigned c2 = __popc(__ballot_sync(full, in && d == -2)), c1 = __popc(__ballot_sync(full, in && d == -1));
    if ((threadIdx.x & 31u) == 0u) {
        atomicMax(maxh_bits, ((unsigned long long)mhi << 32) | mlo);
        if (c2) atomicAdd(counts + 0, c2);
        if (c1) atomicAdd(counts + 1, c1);
    }
}
// Euler tour of the dh forest (children in ascending slot); every tree ends with END.
__global__ void __launch_bounds__(256) k_tree_arcs(Geom g, const int32_t* __restrict__ dh, uint32_t* __restrict__ succ,
                                                   int32_t* __restrict__ w0, int32_t* __restrict__ w1) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= g.n) return;
    int x = g.x_of(v), y = g.y_of(v);
    uint32_t first = 0xffffffffu;
    for (int k = 0; k < 8; ++k) {
        int xx = x + c_ndx[k], yy = y + c_ndy[k];
        if (!g.inb(xx, yy)) continue;
        uint32_t nb = g.idx(xx, yy);
        if (dh[nb] == (int32_t)v) { first = nb; break; }
    }
    succ[2 * v] = first != 0xffffffffu ? 2 * first : 2 * v + 1;
    w0[2 * v] = 1; w0[2 * v + 1] = 0; w1[2 * v] = 1; w1[2 * v + 1] = -1;
    int32_t p = dh[v];
    uint32_t nxt = 0xffffffffu;
    if (p >= 0) {
        int px = g.x_of((uint32_t)p), py = g.y_of((uint32_t)p);
        int s = l_slot_of(x - px, y - py);
        nxt = 2 * (uint32_t)p + 1;
        for (int k = s + 1; k < 8; ++k) {
            int xx = px + c_ndx[k], yy = py + c_ndy[k];
            if (!g.inb(xx, yy)) continue;
            uint32_t nb = g.idx(xx, yy);
            if (dh[nb] == p) { nxt = 2 * nb; break; }
        }
    }
    succ[2 * v + 1] = nxt;
}
// BFS-position key inside a tree: (depth, pre-order) ascending; 24 + 32 bits.
__global__ void __launch_bounds__(256) k_bfs_key(uint32_t n, const int32_t* __restrict__ w0, const int32_t* __restrict__ w1,
                                                 uint64_t* __restrict__ key, int* __restrict__ bad) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    // a tree's tour sums to 0, so the inclusive suffix of w1 at enter(v) is -depth(v)
    int32_t depth = -w1[2 * v];
    if (depth < 0 || depth >= (1 << 24)) { *bad = 1; depth = 0; }
    uint32_t s = (uint32_t)w0[2 * v];
    key[v] = ((uint64_t)(uint32_t)depth << 32) | (uint64_t)(0xffffffffu - s);
}

// ---- (ii) pass table ---------------------------------------------------------------------------
struct PassTable { uint64_t* key; unsigned long long* hb; unsigned long long* tie; uint64_t* edge; uint32_t mask; };

__device__ __forceinline__ uint32_t pt_find_or_insert(const PassTable& t, uint64_t k) {
    uint32_t s = (uint32_t)l_hash(k) & t.mask;
    for (;;) {
        unsigned long long cur = atomicCAS((unsigned long long*)t.key + s, 0ull, (unsigned long long)k);
        if (cur == 0ull || cur == k) return s;
        s = (s + 1) & t.mask;
    }
}
__device__ __forceinline__ uint32_t pt_find(const PassTable& t, uint64_t k) {
    uint32_t s = (uint32_t)l_hash(k) & t.mask;
    while (t.key[s] != k) s = (s + 1) & t.mask;
    return s;
}
// mode 0: count cross edges; 1: min height; 2: min tie key; 3: write edge
__global__ void __launch_bounds__(256) k_pass_scan(Geom g, int mode, const double* __restrict__ h, const int32_t* __restrict__ dh0,
                                                   const uint32_t* __restrict__ root, const uint64_t* __restrict__ bfskey,
                                                   PassTable t, uint32_t* __restrict__ counter) {
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= g.n) return;
    uint32_t A = root[c];
    bool a_boundary = dh0[A] == -2;
    int x = g.x_of(c), y = g.y_of(c);
    double hc = h[c];
    for (int k = 0; k < 8; ++k) {
        int xx = x + c_ndx[k], yy = y + c_ndy[k];
        if (!g.inb(xx, yy)) continue;
        uint32_t nb = g.idx(xx, yy);
        uint32_t B = root[nb];
        if (!(B < A)) continue;                       // 1934: only the higher-offset lake's cell scans
        if (a_boundary && dh0[B] == -2) continue;     // boundary-boundary passes are deleted in (iii)
        if (mode == 0) { atomicAdd(counter, 1u); continue; }
        uint64_t pk = (((uint64_t)A << 32) | (uint64_t)B) + 1ull;
        unsigned long long hb = l_order_bits(fmax(hc, h[nb]));
        if (mode == 1) {
            uint32_t s = pt_find_or_insert(t, pk);
            atomicMin(t.hb + s, hb);
        } else {
            uint32_t s = pt_find(t, pk);
            if (t.hb[s] != hb) continue;
            unsigned long long tk = (bfskey[c] << 3) | (unsigned long long)k;
            if (mode == 2) atomicMin(t.tie + s, tk);
            else if (t.tie[s] == tk) t.edge[s] = ((uint64_t)c << 32) | (uint64_t)nb;
        }
    }
}

// lake ranks: non-boundary lake roots in ascending index
__global__ void __launch_bounds__(256) k_lake_flags(uint32_t n, const int32_t* __restrict__ dh0, uint32_t* __restrict__ fl) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) fl[i] = dh0[i] == -1 ? 1u : 0u;
}
// adjacency degree / fill over the table; also emits the initial candidates (edges into boundary lakes).
// Each adjacency entry carries everything the serial growth needs: (from,to) cells, the rank of
// the lake on the other side (0xffffffff: boundary lake) and the pass-height key.
__global__ void __launch_bounds__(256) k_adj_build(uint32_t cap, int fill, PassTable t, const int32_t* __restrict__ dh0,
                                                   const uint32_t* __restrict__ lrank, const uint32_t* __restrict__ off,
                                                   uint32_t* __restrict__ deg, uint64_t* __restrict__ adj_ft,
                                                   uint32_t* __restrict__ adj_tgt, uint64_t* __restrict__ adj_k0,
                                                   uint64_t* __restrict__ cand_k0, uint64_t* __restrict__ cand_k1,
                                                   uint4* __restrict__ cand_pl, uint32_t* __restrict__ counter) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= cap || t.key[s] == 0) return;
    uint64_t pk = t.key[s] - 1ull;
    uint32_t A = (uint32_t)(pk >> 32), B = (uint32_t)pk;
    uint64_t e = t.edge[s];
    uint32_t ca = (uint32_t)(e >> 32), cb = (uint32_t)e;
    bool an = dh0[A] == -1, bn = dh0[B] == -1;
    if (an) {
        uint32_t i = atomicAdd(deg + lrank[A], 1u);
        if (fill) { uint32_t o = off[lrank[A]] + i; adj_ft[o] = ((uint64_t)ca << 32) | cb; adj_tgt[o] = bn ? lrank[B] : 0xffffffffu; adj_k0[o] = t.hb[s]; }
    }
    if (bn) {
        uint32_t i = atomicAdd(deg + lrank[B], 1u);
        if (fill) { uint32_t o = off[lrank[B]] + i; adj_ft[o] = ((uint64_t)cb << 32) | ca; adj_tgt[o] = an ? lrank[A] : 0xffffffffu; adj_k0[o] = t.hb[s]; }
    }
    if (fill && (an != bn)) { // exactly one side is a boundary lake: 2057-2064
        uint32_t i = atomicAdd(counter, 1u);
        cand_k0[i] = t.hb[s];
        cand_k1[i] = an ? (((uint64_t)ca << 32) | cb) : (((uint64_t)cb << 32) | ca);
        uint32_t lr = an ? lrank[A] : lrank[B];
        cand_pl[i] = make_uint4(lr, off[lr], off[lr + 1], an ? A : B);
    }
}
__global__ void __launch_bounds__(256) k_lake_roots(uint32_t n, const int32_t* __restrict__ dh0, const uint32_t* __restrict__ lrank,
                                                    uint32_t* __restrict__ lake_root) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && dh0[i] == -1) lake_root[lrank[i]] = i;
}
__global__ void __launch_bounds__(256) k_gather3(uint32_t n, const uint32_t* __restrict__ idx, const uint64_t* __restrict__ a0,
                                                 const uint64_t* __restrict__ a1, const uint4* __restrict__ a2,
                                                 uint64_t* __restrict__ o0, uint64_t* __restrict__ o1, uint4* __restrict__ o2) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t j = idx[i];
    o0[i] = a0[j]; o1[i] = a1[j]; o2[i] = a2[j];
}
__global__ void __launch_bounds__(256) k_gather_u64(uint32_t n, const uint32_t* __restrict__ idx, const uint64_t* __restrict__ a,
                                                    uint64_t* __restrict__ o) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[i] = a[idx[i]];
}
__global__ void __launch_bounds__(256) k_iota32(uint32_t n, uint32_t* p) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}

// ---- (iv) Prim growth, one warp ------------------------------------------------------------------
// 32-ary min-heap on (k0, k1) with payload lr (rank of the lake the candidate would connect).
// The first HEAP_SMEM entries (levels 0..2) live in shared memory, the rest in global memory.
#define HEAP_SMEM_MAX 6400u    /* entries of 32 B: 200 KB of the 227 KB an SM offers */
__device__ __forceinline__ bool key_less(uint64_t a0, uint64_t a1, uint64_t b0, uint64_t b1) {
    return a0 < b0 || (a0 == b0 && a1 < b1);
}
// payload of a heap entry: {rank of the lake it would connect, its adjacency range [b, e), its root cell},
// fetched when the candidate is created so that a pop needs no further dependent loads
struct Heap {
    uint64_t* s0; uint64_t* s1; uint4* s2; uint64_t* g0; uint64_t* g1; uint4* g2; uint32_t ns;
    __device__ __forceinline__ void get(uint32_t i, uint64_t& k0, uint64_t& k1, uint4& pl) const {
        if (i < ns) { k0 = s0[i]; k1 = s1[i]; pl = s2[i]; } else { k0 = g0[i]; k1 = g1[i]; pl = g2[i]; }
    }
    __device__ __forceinline__ void getk(uint32_t i, uint64_t& k0, uint64_t& k1) const {
        if (i < ns) { k0 = s0[i]; k1 = s1[i]; } else { k0 = g0[i]; k1 = g1[i]; }
    }
    __device__ __forceinline__ void set(uint32_t i, uint64_t k0, uint64_t k1, uint4 pl) const {
        if (i < ns) { s0[i] = k0; s1[i] = k1; s2[i] = pl; } else { g0[i] = k0; g1[i] = k1; g2[i] = pl; }
    }
    __device__ __forceinline__ void move(uint32_t dst, uint32_t src) const { uint64_t a, b; uint4 c; get(src, a, b, c); set(dst, a, b, c); }
};
__global__ void __launch_bounds__(32) k_lake_connect(const uint32_t* __restrict__ lake_root, const uint32_t* __restrict__ off,
                                                     const uint64_t* __restrict__ adj_ft, const uint32_t* __restrict__ adj_tgt,
                                                     const uint64_t* __restrict__ adj_k0, uint64_t* hk0, uint64_t* hk1, uint4* hk2,
                                                     uint32_t hn0, int32_t* dh, int32_t* indirection, uint8_t* connected,
                                                     uint64_t* best0, uint64_t* best1,
                                                     uint32_t* __restrict__ flows, uint32_t* __restrict__ n_flows_out,
                                                     uint32_t heap_smem, uint32_t conn_words) {
    extern __shared__ __align__(16) unsigned char heap_raw[];
    uint4* s2 = (uint4*)heap_raw;
    uint64_t* s0 = (uint64_t*)(s2 + heap_smem);
    uint64_t* s1 = s0 + heap_smem;
    uint32_t* cbits = (uint32_t*)(s1 + heap_smem);  // "connected" bitmap (conn_words > 0), else global bytes
    const unsigned lane = threadIdx.x;
    for (uint32_t i = lane; i < conn_words; i += 32) cbits[i] = 0u;
    Heap hp{s0, s1, s2, hk0, hk1, hk2, heap_smem};
    for (uint32_t i = lane; i < heap_smem && i < hn0; i += 32) { s0[i] = hk0[i]; s1[i] = hk1[i]; s2[i] = hk2[i]; }
    __syncwarp();
    uint32_t hn = hn0, n_flows = 0, n_pops = 0, n_push = 0;
    long long t_sift = 0, t_scan = 0, t_push = 0, tc;
    // best[l] = smallest key pushed so far for lake l.  A candidate that is not smaller can never be
    // the one that connects l (the smaller one is popped first and is never stale before l is
    // connected), so it is not pushed at all: same result as the reference's heap, fewer heap ops.
    // (the initial candidates are not entered into best[]: a later, worse candidate for the same lake is
    // then pushed needlessly and popped as stale, which is harmless)
    while (hn > 0) {
        n_pops++;
        tc = clock64();
        uint64_t m0, m1; uint4 mp;
        hp.get(0, m0, m1, mp);
        hn--;
        if (hn > 0) {
            uint64_t l0, l1;
            hp.getk(hn, l0, l1);
            __syncwarp();
            uint32_t i = 0;
            if (hn < heap_smem) {
                // whole heap in shared memory (the usual case): no smem/global tests, and every lane performs
                // the moves (same addresses, same values) - a one-warp kernel pays for every branch region
                const unsigned full = 0xffffffffu;
                for (;;) {
                    const uint32_t cb = 32u * i + 1u;
                    if (cb >= hn) break;
                    const uint32_t c = cb + lane;
                    const bool valid = c < hn;
                    const uint32_t cc = valid ? c : 0u;
                    uint64_t c0 = s0[cc], c1 = s1[cc];
                    if (!valid) { c0 = ~0ull; c1 = ~0ull; }
                    uint32_t w0 = (uint32_t)(c0 >> 32), w1 = (uint32_t)c0, w2 = (uint32_t)(c1 >> 32), w3 = (uint32_t)c1;
                    uint32_t m0r = __reduce_min_sync(full, w0); bool in = valid && (w0 == m0r);
                    uint32_t m1r = __reduce_min_sync(full, in ? w1 : 0xffffffffu); in = in && (w1 == m1r);
                    uint32_t m2r = __reduce_min_sync(full, in ? w2 : 0xffffffffu); in = in && (w2 == m2r);
                    uint32_t m3r = __reduce_min_sync(full, in ? w3 : 0xffffffffu); in = in && (w3 == m3r);
                    unsigned bal = __ballot_sync(full, in);
                    const uint32_t ci = cb + (uint32_t)(bal ? (__ffs(bal) - 1) : 0);
                    c0 = ((uint64_t)m0r << 32) | m1r; c1 = ((uint64_t)m2r << 32) | m3r;
                    if (!key_less(c0, c1, l0, l1)) break;
                    // no barrier needed: every lane makes every store itself, so what a lane reads later
                    // is either its own store or one by lane 0 in the push path, which ends with a barrier
                    uint4 pl = s2[ci];
                    s0[i] = c0; s1[i] = c1; s2[i] = pl;
                    i = ci;
                }
                if (i != hn) { uint4 pl = s2[hn]; s0[i] = l0; s1[i] = l1; s2[i] = pl; }
                __syncwarp();
            } else {
            for (;;) { // sift down: the 32 children of i are scanned by the lanes
                uint32_t c = 32u * i + 1u + lane;
                uint64_t c0 = ~0ull, c1 = ~0ull; uint32_t ci = c;
                if (c < hn) hp.getk(c, c0, c1);
                {
                    // lexicographic min over the lanes with four 32-bit warp reductions (redux.sync) instead of
                    // a shuffle tree on 128-bit keys: in a one-warp kernel every instruction is exposed latency
                    const unsigned full = 0xffffffffu;
                    uint32_t w0 = (uint32_t)(c0 >> 32), w1 = (uint32_t)c0, w2 = (uint32_t)(c1 >> 32), w3 = (uint32_t)c1;
                    uint32_t m0r = __reduce_min_sync(full, w0); bool in = (c < hn) && (w0 == m0r);
                    uint32_t m1r = __reduce_min_sync(full, in ? w1 : 0xffffffffu); in = in && (w1 == m1r);
                    uint32_t m2r = __reduce_min_sync(full, in ? w2 : 0xffffffffu); in = in && (w2 == m2r);
                    uint32_t m3r = __reduce_min_sync(full, in ? w3 : 0xffffffffu); in = in && (w3 == m3r);
                    unsigned bal = __ballot_sync(full, in);
                    int win = bal ? (__ffs(bal) - 1) : 0;
                    c0 = ((uint64_t)m0r << 32) | m1r; c1 = ((uint64_t)m2r << 32) | m3r;
                    ci = 32u * i + 1u + (uint32_t)win;
                }
                if (32u * i + 1u >= hn || !key_less(c0, c1, l0, l1)) break;
                if (lane == 0) hp.move(i, ci);
                __syncwarp();
                i = ci;
            }
            if (lane == 0 && i != hn) hp.move(i, hn);
            __syncwarp();
            }
        }
        (void)m0;
        t_sift += clock64() - tc;
        const uint32_t mlr = mp.x;
        bool already = conn_words ? ((cbits[mlr >> 5] >> (mlr & 31u)) & 1u) : (connected[mlr] != 0);
        if (already) continue;                                        // 2085
        const uint32_t b = mp.y, e = mp.z, lake = mp.w;
        uint32_t from = (uint32_t)(m1 >> 32), to = (uint32_t)m1;
        __syncwarp();
        if (lane == 0) {
            if (conn_words) cbits[mlr >> 5] |= 1u << (mlr & 31u); else connected[mlr] = 1;
            dh[lake] = (int32_t)to;                                   // 2098
            indirection[lake] = -(int32_t)from;                       // 2103
            flows[n_flows] = lake;                                    // 2105
        }
        n_flows++;
        __syncwarp();
        // push the reversed other passes of this lake (2107-2136); lanes evaluate edges in parallel
        for (uint32_t a0 = b; a0 < e; a0 += 32) {
            tc = clock64();
            uint32_t a = a0 + lane;
            uint64_t ft = 0, k0 = 0; bool ok = false; uint4 pl = make_uint4(0, 0, 0, 0);
            if (a < e) {
                ft = adj_ft[a]; k0 = adj_k0[a];
                uint32_t tg = adj_tgt[a];
                ok = (ft != m1) && (tg != 0xffffffffu) && !(conn_words ? ((cbits[tg >> 5] >> (tg & 31u)) & 1u) : (uint32_t)connected[tg]);
                if (ok) {
                    ft = (ft << 32) | (ft >> 32);                     // reversed edge (to, from)
                    uint64_t b0 = best0[tg], b1 = best1[tg];
                    pl = make_uint4(tg, off[tg], off[tg + 1], lake_root[tg]);   // same round trip as best[]
                    ok = key_less(k0, ft, b0, b1);                    // only if it beats that lake's best so far
                }
            }
            unsigned bal = __ballot_sync(0xffffffffu, ok);
            t_scan += clock64() - tc; tc = clock64();
            while (bal) {
                int src = __ffs(bal) - 1; bal &= bal - 1;
                uint64_t p0 = __shfl_sync(0xffffffffu, k0, src), p1 = __shfl_sync(0xffffffffu, ft, src);
                uint4 pp;
                pp.x = __shfl_sync(0xffffffffu, pl.x, src); pp.y = __shfl_sync(0xffffffffu, pl.y, src);
                pp.z = __shfl_sync(0xffffffffu, pl.z, src); pp.w = __shfl_sync(0xffffffffu, pl.w, src);
                if (hn < heap_smem) {
                    // shared-memory heap: sift up in every lane alike (see the sift-down above)
                    best0[pp.x] = p0; best1[pp.x] = p1;   // ok => this candidate beats the lake's best so far
                    uint32_t i = hn;
                    while (i > 0) {
                        uint32_t pi = (i - 1u) / 32u;
                        uint64_t q0 = s0[pi], q1 = s1[pi];
                        if (!key_less(p0, p1, q0, q1)) break;
                        uint4 pl2 = s2[pi];
                        s0[i] = q0; s1[i] = q1; s2[i] = pl2;
                        i = pi;
                    }
                    s0[i] = p0; s1[i] = p1; s2[i] = pp;
                } else if (lane == 0) { // sift up
                    best0[pp.x] = p0; best1[pp.x] = p1;   // ok => this candidate beats the lake's best so far
                    uint32_t i = hn;
                    while (i > 0) {
                        uint32_t pi = (i - 1u) / 32u;
                        uint64_t q0, q1;
                        hp.getk(pi, q0, q1);
                        if (!key_less(p0, p1, q0, q1)) break;
                        hp.move(i, pi);
                        i = pi;
                    }
                    hp.set(i, p0, p1, pp);
                }
                hn++; n_push++;
                __syncwarp();
            }
            t_push += clock64() - tc;
        }
    }
    if (lane == 0) { *n_flows_out = n_flows; n_flows_out[1] = n_pops; n_flows_out[2] = n_push;
                     n_flows_out[3] = (uint32_t)(t_sift >> 10); n_flows_out[4] = (uint32_t)(t_scan >> 10); n_flows_out[5] = (uint32_t)(t_push >> 10); }
}

// ---- (iii)/(iv) parallel formulation ----------------------------------------------------------------
// The reference grows the forest Prim-style from the boundary lakes with a heap keyed
// (pass height, cell on the side being connected, cell on the connected side) (erosion.rs:2060-2137).
// Two facts turn that into parallel work (checked against the oracle's heap loop on every test map):
//  1. If every pass had ONE fixed key, the growth would be Prim's algorithm on a graph with distinct
//     weights: the passes it uses are the minimum spanning tree (all boundary lakes = one vertex S), each
//     lake's pass is its parent edge in the tree rooted at S, and the ORDER of connection is the pre-order
//     of the tree T' in which parent'(v) = the lower end of the nearest ancestor edge heavier than v's own
//     parent edge (S if none), children visited by ascending key: after a heavy edge is taken, everything
//     reachable below it by lighter edges is connected before anything else.
//  2. The key of a pass depends on the direction in which it is crossed, i.e. on which of its two lakes is
//     connected first.  Given a guess of the connection order, orient every pass by it, run 1, and obtain a
//     new order; the reference's order is the unique fixed point of that map (a run of Prim with the keys
//     of a fixed point only ever compares candidates whose key is the one the reference would compute, so
//     it is the reference's run).  Only passes that share their height with another pass can change rank,
//     so the iteration has converged as soon as their orientation repeats: 2-3 rounds in practice.
// Per round: orientation + rank inside the (tiny) equal-height groups, Boruvka rounds in one cooperative
// kernel, rooting by Euler tour + list ranking, nearest-heavier-ancestor by binary lifting, one radix
// sort, a second Euler tour.  No round is sequential in the number of lakes.
namespace cg = cooperative_groups;
#define LG_NONE 0xffffffffu
#define LG_MAX_GROUP 2048u     /* equal-height groups are ranked by brute force; larger ones take the serial path */

__global__ void __launch_bounds__(256) k_lg_edges(uint32_t cap, PassTable t, const int32_t* __restrict__ dh0, const uint32_t* __restrict__ lrank,
                                                  uint32_t L, uint32_t* __restrict__ counter, uint64_t* __restrict__ ehb,
                                                  uint32_t* __restrict__ idx, uint4* __restrict__ erec) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= cap || t.key[s] == 0) return;
    uint64_t pk = t.key[s] - 1ull;
    uint32_t A = (uint32_t)(pk >> 32), B = (uint32_t)pk;
    uint64_t e = t.edge[s];
    uint32_t i = atomicAdd(counter, 1u);
    ehb[i] = t.hb[s];
    idx[i] = i;
    erec[i] = make_uint4(dh0[A] == -1 ? lrank[A] : L, dh0[B] == -1 ? lrank[B] : L, (uint32_t)(e >> 32), (uint32_t)e);
}
__global__ void __launch_bounds__(256) k_lg_gather(uint32_t M, const uint32_t* __restrict__ idx, const uint4* __restrict__ in, uint4* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < M) out[i] = in[idx[i]];
}
// equal-height groups in the height-sorted edge array: [gs, ge) for every position
__global__ void __launch_bounds__(256) k_lg_groups(uint32_t M, const uint64_t* __restrict__ hb, uint32_t* __restrict__ gs, uint32_t* __restrict__ ge,
                                                   int* __restrict__ toolarge) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const uint64_t h = hb[i];
    uint32_t a = i, b = i + 1;
    while (a > 0 && hb[a - 1] == h && i - a < LG_MAX_GROUP) --a;
    while (b < M && hb[b] == h && b - i < LG_MAX_GROUP) ++b;
    if (i - a >= LG_MAX_GROUP || b - i >= LG_MAX_GROUP) *toolarge = 1;
    gs[i] = a; ge[i] = b;
}
// orientation of every pass by the current guess of the connection order; key_lo = (from cell, to cell) as the heap
// would see it; `changed` is raised when a pass that shares its height changes orientation
__global__ void __launch_bounds__(256) k_lg_orient(uint32_t M, const uint4* __restrict__ erec, const uint32_t* __restrict__ pos,
                                                   const uint32_t* __restrict__ gs, const uint32_t* __restrict__ ge, uint8_t* __restrict__ obits,
                                                   uint64_t* __restrict__ key_lo, int first, int* __restrict__ changed) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    uint4 x = erec[i];
    const bool a_later = pos[x.x] > pos[x.y];
    const uint32_t from = a_later ? x.z : x.w, to = a_later ? x.w : x.z;
    key_lo[i] = ((uint64_t)from << 32) | (uint64_t)to;
    const uint8_t bit = a_later ? 1 : 0;
    if (!first && ge[i] - gs[i] > 1u && obits[i] != bit) *changed = 1;
    obits[i] = bit;
}
__global__ void __launch_bounds__(256) k_lg_rank(uint32_t M, const uint32_t* __restrict__ gs, const uint32_t* __restrict__ ge,
                                                 const uint64_t* __restrict__ key_lo, uint32_t* __restrict__ erank, uint32_t* __restrict__ byrank) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const uint32_t a = gs[i], b = ge[i];
    uint32_t r = a;
    const uint64_t k = key_lo[i];
    for (uint32_t j = a; j < b; ++j) r += key_lo[j] < k ? 1u : 0u;
    erank[i] = r;
    byrank[r] = i;
}
// Boruvka: every component takes its lightest outgoing pass (ranks are distinct, so the hooks form a forest plus
// 2-cycles, which are broken towards the smaller id); rounds until no pass joins two components.
__global__ void __launch_bounds__(256) k_lg_boruvka(uint32_t M, uint32_t V, const uint4* __restrict__ erec, const uint32_t* __restrict__ erank,
                                                    const uint32_t* __restrict__ byrank, uint32_t* comp_a, uint32_t* comp_b, uint32_t* best,
                                                    volatile uint32_t* hook, uint8_t* intree, volatile uint32_t* flags, uint32_t* out_sel) {
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (uint32_t v = tid; v < V; v += nth) comp_a[v] = v;
    for (uint32_t e = tid; e < M; e += nth) intree[e] = 0;
    uint32_t* comp = comp_a; uint32_t* comp2 = comp_b;
    grid.sync();
    for (int round = 0; round < 64; ++round) {
        for (uint32_t v = tid; v < V; v += nth) best[v] = LG_NONE;
        grid.sync();
        bool any = false;
        for (uint32_t e = tid; e < M; e += nth) {
            uint4 x = erec[e];
            uint32_t ca = comp[x.x], cb = comp[x.y];
            if (ca != cb) { uint32_t r = erank[e]; atomicMin(best + ca, r); atomicMin(best + cb, r); any = true; }
        }
        if (any) flags[round] = 1u;
        grid.sync();
        if (flags[round] == 0u) break;
        for (uint32_t v = tid; v < V; v += nth) {
            uint32_t h = v;
            if (comp[v] == v) {
                uint32_t b = best[v];
                if (b != LG_NONE) {
                    uint32_t e = byrank[b];
                    uint4 x = erec[e];
                    uint32_t ca = comp[x.x];
                    h = (ca == v) ? comp[x.y] : ca;
                    intree[e] = 1;
                }
            }
            hook[v] = h;
        }
        grid.sync();
        for (uint32_t v = tid; v < V; v += nth) {
            uint32_t w = hook[v];
            if (w != v && hook[w] == v && v < w) hook[v] = v;
        }
        grid.sync();
        for (uint32_t v = tid; v < V; v += nth) {
            uint32_t r = comp[v];
            for (;;) { uint32_t h = hook[r]; if (h == r) break; r = h; }
            comp2[v] = r;
        }
        grid.sync();
        uint32_t* tsw = comp; comp = comp2; comp2 = tsw;
    }
    if (tid == 0) *out_sel = (comp == comp_a) ? 0u : 1u;
}
__global__ void __launch_bounds__(256) k_lg_check(uint32_t V, const uint32_t* __restrict__ comp, int* __restrict__ bad) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < V && comp[v] != comp[V - 1]) *bad = 1;          // some lake cannot reach the boundary
}
__global__ void __launch_bounds__(256) k_lg_deg(uint32_t M, const uint4* __restrict__ erec, const uint8_t* __restrict__ intree, uint32_t* __restrict__ deg) {
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= M || !intree[e]) return;
    uint4 x = erec[e];
    atomicAdd(deg + x.x, 1u); atomicAdd(deg + x.y, 1u);
}
__global__ void __launch_bounds__(256) k_lg_arcs(uint32_t M, const uint4* __restrict__ erec, const uint8_t* __restrict__ intree, const uint32_t* __restrict__ off,
                                                 uint32_t* __restrict__ cur, uint32_t* __restrict__ arc_to, uint32_t* __restrict__ arc_twin,
                                                 uint32_t* __restrict__ arc_e) {
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= M || !intree[e]) return;
    uint4 x = erec[e];
    uint32_t sa = off[x.x] + atomicAdd(cur + x.x, 1u), sb = off[x.y] + atomicAdd(cur + x.y, 1u);
    arc_to[sa] = x.y; arc_twin[sa] = sb; arc_e[sa] = e;
    arc_to[sb] = x.x; arc_twin[sb] = sa; arc_e[sb] = e;
}
// Euler circuit of the doubled tree, cut open at S's first arc
__global__ void __launch_bounds__(256) k_lg_tour(uint32_t A, uint32_t S, const uint32_t* __restrict__ arc_to, const uint32_t* __restrict__ arc_twin,
                                                 const uint32_t* __restrict__ off, uint32_t* __restrict__ succ, int32_t* __restrict__ w0) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A) return;
    uint32_t v = arc_to[i], j = arc_twin[i];
    uint32_t nxt = (j + 1 < off[v + 1]) ? j + 1 : off[v];
    if (nxt == off[S]) nxt = LG_NONE;
    succ[i] = nxt; w0[i] = 1;
}
__global__ void __launch_bounds__(256) k_lg_parent(uint32_t A, uint32_t S, const uint32_t* __restrict__ arc_to, const uint32_t* __restrict__ arc_twin,
                                                   const uint32_t* __restrict__ arc_e, const int32_t* __restrict__ w0, const uint32_t* __restrict__ erank,
                                                   uint32_t* __restrict__ par, uint32_t* __restrict__ pkey, uint32_t* __restrict__ pedge) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { par[S] = S; pkey[S] = LG_NONE; pedge[S] = LG_NONE; }
    if (i >= A) return;
    uint32_t j = arc_twin[i];
    if (w0[i] > w0[j]) {                       // crossed first in this direction: away from S
        uint32_t c = arc_to[i];
        par[c] = arc_to[j]; pkey[c] = erank[arc_e[i]]; pedge[c] = arc_e[i];
    }
}
__global__ void __launch_bounds__(256) k_lg_lift0(uint32_t V, const uint32_t* __restrict__ par, const uint32_t* __restrict__ pkey,
                                                  uint32_t* __restrict__ anc, uint32_t* __restrict__ mx) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    uint32_t p = par[v];
    anc[v] = p; mx[v] = pkey[p];
}
__global__ void __launch_bounds__(256) k_lg_lift(uint32_t V, const uint32_t* __restrict__ anc, const uint32_t* __restrict__ mx,
                                                 uint32_t* __restrict__ anc1, uint32_t* __restrict__ mx1) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    uint32_t a = anc[v];
    anc1[v] = anc[a]; mx1[v] = max(mx[v], mx[a]);
}
// parent in T': the lower end of the nearest ancestor edge heavier than the own parent edge
__global__ void __launch_bounds__(256) k_lg_tp(uint32_t L, uint32_t V, int K, const uint32_t* __restrict__ anc, const uint32_t* __restrict__ mx,
                                               const uint32_t* __restrict__ pkey, uint64_t* __restrict__ skey, uint32_t* __restrict__ sval) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= L) return;
    const uint32_t kv = pkey[v];
    uint32_t cur = v;
    for (int k = K - 1; k >= 0; --k) if (mx[(size_t)k * V + cur] < kv) cur = anc[(size_t)k * V + cur];
    skey[v] = ((uint64_t)anc[cur] << 32) | (uint64_t)kv;
    sval[v] = v;
}
__global__ void __launch_bounds__(256) k_lg_tarcs(uint32_t L, const uint64_t* __restrict__ sk, const uint32_t* __restrict__ sv,
                                                  uint32_t* __restrict__ firstchild, uint32_t* __restrict__ succ) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    uint32_t v = sv[i], p = (uint32_t)(sk[i] >> 32);
    if (i == 0 || (uint32_t)(sk[i - 1] >> 32) != p) firstchild[p] = v;
    succ[2 * v + 1] = (i + 1 < L && (uint32_t)(sk[i + 1] >> 32) == p) ? 2 * sv[i + 1] : 2 * p + 1;
}
__global__ void __launch_bounds__(256) k_lg_tarcs2(uint32_t V, uint32_t S, const uint32_t* __restrict__ firstchild, uint32_t* __restrict__ succ,
                                                   int32_t* __restrict__ w0) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    uint32_t fc = firstchild[v];
    succ[2 * v] = fc != LG_NONE ? 2 * fc : 2 * v + 1;
    w0[2 * v] = v != S ? 1 : 0; w0[2 * v + 1] = 0;
    if (v == S) succ[2 * v + 1] = LG_NONE;
}
__global__ void __launch_bounds__(256) k_lg_pos(uint32_t L, const int32_t* __restrict__ w0, uint32_t* __restrict__ pos, int* __restrict__ bad) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= L) return;
    int64_t t = (int64_t)L - (int64_t)w0[2 * v];
    if (t < 0 || t >= (int64_t)L) { *bad = 1; t = 0; }
    pos[v] = 1u + (uint32_t)t;
}
__global__ void __launch_bounds__(256) k_lg_final(uint32_t L, const uint32_t* __restrict__ pos, const uint32_t* __restrict__ pedge, const uint4* __restrict__ erec,
                                                  const uint32_t* __restrict__ lake_root, uint32_t* __restrict__ flows, int32_t* __restrict__ dh,
                                                  int32_t* __restrict__ indirection) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= L) return;
    uint4 x = erec[pedge[v]];
    uint32_t from = x.x == v ? x.z : x.w, to = x.x == v ? x.w : x.z;
    uint32_t root = lake_root[v];
    dh[root] = (int32_t)to;                    // 2098
    indirection[root] = -(int32_t)from;        // 2103
    flows[pos[v] - 1u] = root;                 // 2105
}
__global__ void __launch_bounds__(256) k_lg_pos_init(uint32_t V, uint32_t* __restrict__ pos) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < V) pos[v] = v == V - 1 ? 0u : v + 1u;
}

// Returns VE_OK with *done = 1 when dh / indirection / flows of all L lakes have been written; *done = 0 asks the
// caller for the serial growth (an equal-height group beyond LG_MAX_GROUP, or no fixed point within the round limit).
static int lake_grow_parallel(ve_ctx* ctx, PassTable t, uint32_t cap, uint32_t n_cross, uint32_t L, const int32_t* dh0, const uint32_t* lrank,
                              const uint32_t* lake_root, int32_t* dh, int32_t* indirection, uint32_t* flows, uint32_t* small, int* done) {
    cudaStream_t st = ctx->stream;
    *done = 0;
    ctx->lake_rounds = 0;
    const uint32_t V = L + 1, S = L;
    int K = 1; while ((1ull << K) < (unsigned long long)V) ++K;
    const size_t Mc = (size_t)n_cross + 1;
    // ---- carve the work arrays
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    size_t eb = up(Mc * 8) * 3 + up(Mc * 4) * 6 + up(Mc * 16) * 2 + up(Mc) * 2;
    uint8_t* E; VE_TRY(ve_scratch(ctx, 8, eb, (void**)&E));
    uint8_t* q = E;
    auto take = [&](size_t bytes) { uint8_t* r = q; q += up(bytes); return r; };
    uint64_t* ehb_in = (uint64_t*)take(Mc * 8); uint64_t* ehb = (uint64_t*)take(Mc * 8); uint64_t* key_lo = (uint64_t*)take(Mc * 8);
    uint32_t* idx_in = (uint32_t*)take(Mc * 4); uint32_t* idx_s = (uint32_t*)take(Mc * 4); uint32_t* gs = (uint32_t*)take(Mc * 4);
    uint32_t* ge = (uint32_t*)take(Mc * 4); uint32_t* erank = (uint32_t*)take(Mc * 4); uint32_t* byrank = (uint32_t*)take(Mc * 4);
    uint4* erec_in = (uint4*)take(Mc * 16); uint4* erec = (uint4*)take(Mc * 16);
    uint8_t* obits = take(Mc); uint8_t* intree = take(Mc);
    const size_t Vc = (size_t)V + 2, Ac = 2 * Vc;
    size_t vb = up(Vc * 4) * (12 + 2 * (size_t)K) + up(Vc * 8) * 2 + up(Ac * 4) * 5;
    uint8_t* W; VE_TRY(ve_scratch(ctx, 9, vb, (void**)&W));
    q = W;
    uint32_t* pos = (uint32_t*)take(Vc * 4); uint32_t* comp_a = (uint32_t*)take(Vc * 4); uint32_t* comp_b = (uint32_t*)take(Vc * 4);
    uint32_t* best = (uint32_t*)take(Vc * 4); uint32_t* hook = (uint32_t*)take(Vc * 4); uint32_t* deg = (uint32_t*)take(Vc * 4);
    uint32_t* off = (uint32_t*)take(Vc * 4); uint32_t* cur = (uint32_t*)take(Vc * 4); uint32_t* par = (uint32_t*)take(Vc * 4);
    uint32_t* pkey = (uint32_t*)take(Vc * 4); uint32_t* pedge = (uint32_t*)take(Vc * 4); uint32_t* firstchild = (uint32_t*)take(Vc * 4);
    uint32_t* anc = (uint32_t*)take(up(Vc * 4) * (size_t)K); uint32_t* mx = (uint32_t*)take(up(Vc * 4) * (size_t)K);
    // anc / mx levels are addressed as k * V + v: pack them densely inside their blocks (K * V <= K * Vc)
    uint64_t* skey = (uint64_t*)take(Vc * 8); uint64_t* skey_s = (uint64_t*)take(Vc * 8);
    uint32_t* arc_to = (uint32_t*)take(Ac * 4); uint32_t* arc_twin = (uint32_t*)take(Ac * 4); uint32_t* arc_e = (uint32_t*)take(Ac * 4);
    uint32_t* succ = (uint32_t*)take(Ac * 4); int32_t* w0 = (int32_t*)take(Ac * 4);
    uint32_t* sval = cur;            // free after the arcs are filled
    uint32_t* sval_s = deg;
    // small: [16] edge counter, [17] toolarge, [18] changed, [19] bad, [20] boruvka comp selector, [32..95] round flags
    VE_CUDA(cudaMemsetAsync(small + 16, 0, 16 * 4, st));
    // ---- undirected passes, sorted by height once
    k_lg_edges<<<ve_blocks(cap, 256), 256, 0, st>>>(cap, t, dh0, lrank, L, small + 16, ehb_in, idx_in, erec_in);
    VE_LAUNCH_CHECK();
    uint32_t M = 0;
    VE_TRY(ve_d2h(ctx, &M, small + 16, 4));
    if (M == 0 || M > n_cross) { ctx->err = "get_lakes: pass table inconsistent"; return VE_ERR_LAKE_GRAPH; }
    const unsigned mb = ve_blocks(M, 256), vbk = ve_blocks(V, 256), lb = ve_blocks(L, 256);
    size_t tmp_bytes = 0; void* tmp;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, ehb_in, ehb, idx_in, idx_s, (int)M, 0, 64, st);
    VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
    VE_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ehb_in, ehb, idx_in, idx_s, (int)M, 0, 64, st));
    ctx->launches += 4;
    k_lg_gather<<<mb, 256, 0, st>>>(M, idx_s, erec_in, erec);
    VE_LAUNCH_CHECK();
    k_lg_groups<<<mb, 256, 0, st>>>(M, ehb, gs, ge, (int*)(small + 17));
    VE_LAUNCH_CHECK();
    k_lg_pos_init<<<vbk, 256, 0, st>>>(V, pos);
    VE_LAUNCH_CHECK();
    int bor_bps = 0;
    VE_TRY(ve_occupancy(ctx, (const void*)k_lg_boruvka, 256, 0, &bor_bps));
    unsigned bor_grid = (unsigned)(bor_bps * ctx->sm_count);
    { unsigned need = ve_blocks(std::max(M, V), 256); if (bor_grid > need) bor_grid = need; if (bor_grid < 1) bor_grid = 1; }
    bool converged = false;
    int round = 0;
    for (; round < 12; ++round) {
        k_lg_orient<<<mb, 256, 0, st>>>(M, erec, pos, gs, ge, obits, key_lo, round == 0 ? 1 : 0, (int*)(small + 18));
        VE_LAUNCH_CHECK();
        uint32_t hflag[4] = {0, 0, 0, 0};   // toolarge, changed, bad, sel
        VE_TRY(ve_d2h(ctx, hflag, small + 17, 16));
        if (hflag[0]) return VE_OK;                                  // a huge equal-height group: serial path
        if (hflag[2]) { ctx->err = "get_lakes: inconsistent lake graph (parallel growth)"; return VE_ERR_LAKE_GRAPH; }
        if (round > 0 && !hflag[1]) { converged = true; break; }   // orientation repeated: the previous round's result stands
        VE_CUDA(cudaMemsetAsync(small + 18, 0, 4, st));
        k_lg_rank<<<mb, 256, 0, st>>>(M, gs, ge, key_lo, erank, byrank);
        VE_LAUNCH_CHECK();
        VE_CUDA(cudaMemsetAsync(small + 32, 0, 64 * 4, st));
        {
            uint32_t Mv = M, Vv = V; const uint4* er = erec; const uint32_t* rk = erank; const uint32_t* br = byrank;
            uint32_t* fl = small + 32; uint32_t* sel = small + 20;
            void* args[] = {(void*)&Mv, (void*)&Vv, (void*)&er, (void*)&rk, (void*)&br, (void*)&comp_a, (void*)&comp_b, (void*)&best,
                            (void*)&hook, (void*)&intree, (void*)&fl, (void*)&sel};
            VE_CUDA(cudaLaunchCooperativeKernel((void*)k_lg_boruvka, dim3(bor_grid), dim3(256), args, 0, st));
            ctx->launches++;
        }
        uint32_t sel = 0;
        VE_TRY(ve_d2h(ctx, &sel, small + 20, 4));
        const uint32_t* comp = sel ? comp_b : comp_a;
        k_lg_check<<<vbk, 256, 0, st>>>(V, comp, (int*)(small + 21));
        VE_LAUNCH_CHECK();
        uint32_t disc = 0;
        VE_TRY(ve_d2h(ctx, &disc, small + 21, 4));
        if (disc) { ctx->err = "Disconnected lake!"; return VE_ERR_DISCONNECTED_LAKE; }
        // rooted spanning tree: parent and parent pass of every lake
        VE_CUDA(cudaMemsetAsync(deg, 0, Vc * 4, st));
        VE_CUDA(cudaMemsetAsync(cur, 0, Vc * 4, st));
        k_lg_deg<<<mb, 256, 0, st>>>(M, erec, intree, deg);
        VE_LAUNCH_CHECK();
        cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, deg, off, (int)(V + 1), st);
        VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
        VE_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, deg, off, (int)(V + 1), st));
        ctx->launches += 2;
        k_lg_arcs<<<mb, 256, 0, st>>>(M, erec, intree, off, cur, arc_to, arc_twin, arc_e);
        VE_LAUNCH_CHECK();
        const uint32_t A = 2 * L;                                    // a spanning tree of V = L + 1 vertices
        k_lg_tour<<<ve_blocks(A, 256), 256, 0, st>>>(A, S, arc_to, arc_twin, off, succ, w0);
        VE_LAUNCH_CHECK();
        VE_TRY(ve_list_rank(ctx, A, succ, w0, nullptr));
        k_lg_parent<<<ve_blocks(A, 256), 256, 0, st>>>(A, S, arc_to, arc_twin, arc_e, w0, erank, par, pkey, pedge);
        VE_LAUNCH_CHECK();
        // T': nearest heavier ancestor edge by binary lifting
        k_lg_lift0<<<vbk, 256, 0, st>>>(V, par, pkey, anc, mx);
        VE_LAUNCH_CHECK();
        for (int k = 0; k + 1 < K; ++k) {
            k_lg_lift<<<vbk, 256, 0, st>>>(V, anc + (size_t)k * V, mx + (size_t)k * V, anc + (size_t)(k + 1) * V, mx + (size_t)(k + 1) * V);
            VE_LAUNCH_CHECK();
        }
        k_lg_tp<<<lb, 256, 0, st>>>(L, V, K, anc, mx, pkey, skey, sval);
        VE_LAUNCH_CHECK();
        cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, skey, skey_s, sval, sval_s, (int)L, 0, 64, st);
        VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
        VE_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, skey, skey_s, sval, sval_s, (int)L, 0, 64, st));
        ctx->launches += 4;
        VE_CUDA(cudaMemsetAsync(firstchild, 0xff, Vc * 4, st));
        k_lg_tarcs<<<lb, 256, 0, st>>>(L, skey_s, sval_s, firstchild, succ);
        VE_LAUNCH_CHECK();
        k_lg_tarcs2<<<vbk, 256, 0, st>>>(V, S, firstchild, succ, w0);
        VE_LAUNCH_CHECK();
        VE_TRY(ve_list_rank(ctx, 2 * V, succ, w0, nullptr));
        k_lg_pos<<<lb, 256, 0, st>>>(L, w0, pos, (int*)(small + 19));
        VE_LAUNCH_CHECK();
    }
    if (!converged) return VE_OK;                                     // no fixed point within the limit: serial path
    ctx->lake_rounds = (uint32_t)round;
    k_lg_final<<<lb, 256, 0, st>>>(L, pos, pedge, erec, lake_root, flows, dh, indirection);
    VE_LAUNCH_CHECK();
    *done = 1;
    return VE_OK;
}

// ---- (v) per-lake flood -----------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_lake_fill(Geom g, uint32_t n_flows, const uint32_t* __restrict__ flows,
                                                   const double* __restrict__ h, const uint32_t* __restrict__ root,
                                                   const uint32_t* __restrict__ start, int32_t* dh, const int32_t* __restrict__ indirection,
                                                   uint8_t* tag, uint32_t* stack, int* __restrict__ bad) {
    uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_flows) return;
    const uint32_t chunk = flows[li];
    const uint32_t nbr_pass = (uint32_t)dh[chunk];
    const int32_t pass_s = -indirection[chunk];
    if (pass_s < 0) { *bad = 1; return; }
    const uint32_t pass = (uint32_t)pass_s;
    uint32_t* st = stack + start[chunk];
    uint32_t sp = 0;
    st[sp++] = pass;
    tag[pass] = 1; // InQueue
    const int ox = g.x_of(nbr_pass), oy = g.y_of(nbr_pass);
    const double elev = fmax(h[nbr_pass], h[pass]);
    while (sp > 0) {
        uint32_t node = st[--sp];
        int cx = g.x_of(node), cy = g.y_of(node);
        int64_t rcv = -1;
        double rcv_cost = -INFINITY;
        double odx = (double)(ox - cx), ody = (double)(oy - cy);
        for (int k = 0; k < 8; ++k) {
            int xx = cx + c_ndx[k], yy = cy + c_ndy[k];
            if (!g.inb(xx, yy)) continue;
            uint32_t nb = g.idx(xx, yy);
            bool mine = root[nb] == chunk;           // indirection[nb] == chunk || nb == chunk
            if ((!mine && nb != nbr_pass) || h[nb] > elev) continue;
            // tags of foreign cells are never written here; the outflow cell counts as WithRcv (2212)
            uint8_t tg = (nb == nbr_pass) ? (uint8_t)2 : tag[nb];
            if (tg == 2) {
                double ddx = (double)c_ndx[k], ddy = (double)c_ndy[k];
                double mag = vm_sqrt(ddx * ddx + ddy * ddy);
                double cost = odx * (ddx / mag) + ody * (ddy / mag);
                if (cost > rcv_cost) { rcv = (int64_t)nb; rcv_cost = cost; }
            } else if (tg == 0) {
                st[sp++] = nb;
                tag[nb] = 1;
            }
        }
        if (rcv < 0) { *bad = 1; return; }            // assert_ne 2262
        dh[node] = (int32_t)rcv;
        tag[node] = 2;
    }
}

// ---- final order ------------------------------------------------------------------------------------
// Blocks: boundary roots (ascending index) then connected lakes in connection order.  Arc ids:
// real node v -> 2v, 2v+1; virtual root of lake with rank r -> 2(n+r), 2(n+r)+1.
__global__ void __launch_bounds__(256) k_final_arcs(Geom g, const int32_t* __restrict__ dh, const int32_t* __restrict__ dh0,
                                                    const uint32_t* __restrict__ root, const uint32_t* __restrict__ lrank,
                                                    const uint32_t* __restrict__ block_of_root, const uint32_t* __restrict__ block_head,
                                                    uint32_t n_blocks, uint32_t* __restrict__ succ, int32_t* __restrict__ w0,
                                                    int32_t* __restrict__ w1) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= g.n) return;
    const uint32_t n = g.n;
    int x = g.x_of(v), y = g.y_of(v);
    uint32_t lake = root[v];
    // first child inside the same lake
    uint32_t first = 0xffffffffu;
    for (int k = 0; k < 8; ++k) {
        int xx = x + c_ndx[k], yy = y + c_ndy[k];
        if (!g.inb(xx, yy)) continue;
        uint32_t nb = g.idx(xx, yy);
        if (dh[nb] == (int32_t)v && root[nb] == lake) { first = nb; break; }
    }
    succ[2 * v] = first != 0xffffffffu ? 2 * first : 2 * v + 1;
    w0[2 * v] = 1; w0[2 * v + 1] = 0; w1[2 * v] = 1; w1[2 * v + 1] = -1;
    int32_t p = dh[v];
    uint32_t nxt;
    if (p < 0) { // ocean root: block root
        uint32_t bi = block_of_root[v];
        nxt = (bi + 1 < n_blocks) ? block_head[bi + 1] : 0xffffffffu;
    } else {
        bool same = root[(uint32_t)p] == lake;
        uint32_t up = same ? (2 * (uint32_t)p + 1) : (2 * (n + lrank[lake]) + 1);
        int px = g.x_of((uint32_t)p), py = g.y_of((uint32_t)p);
        int s = l_slot_of(x - px, y - py);
        nxt = up;
        for (int k = s + 1; k < 8; ++k) {
            int xx = px + c_ndx[k], yy = py + c_ndy[k];
            if (!g.inb(xx, yy)) continue;
            uint32_t nb = g.idx(xx, yy);
            if (dh[nb] == p && root[nb] == lake) { nxt = 2 * nb; break; }
        }
    }
    succ[2 * v + 1] = nxt;
}
__global__ void __launch_bounds__(256) k_final_virtual(Geom g, uint32_t n_flows, const uint32_t* __restrict__ flows,
                                                       const int32_t* __restrict__ dh, const uint32_t* __restrict__ root,
                                                       const uint32_t* __restrict__ lrank, const uint32_t* __restrict__ block_of_root,
                                                       const uint32_t* __restrict__ block_head, uint32_t n_blocks,
                                                       uint32_t* __restrict__ succ, int32_t* __restrict__ w0, int32_t* __restrict__ w1) {
    uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_flows) return;
    uint32_t chunk = flows[li];
    uint32_t a = 2 * (g.n + lrank[chunk]);
    w0[a] = 0; w0[a + 1] = 0; w1[a] = 1; w1[a + 1] = -1;
    uint32_t bi = block_of_root[chunk];
    succ[a + 1] = (bi + 1 < n_blocks) ? block_head[bi + 1] : 0xffffffffu;
}
// first child of each virtual root: neighbours of the outflow cell, ascending slot, in this lake
__global__ void __launch_bounds__(256) k_final_virtual_first(Geom g, uint32_t n_flows, const uint32_t* __restrict__ flows,
                                                             const uint32_t* __restrict__ outflow, const int32_t* __restrict__ dh,
                                                             const uint32_t* __restrict__ root, const uint32_t* __restrict__ lrank,
                                                             uint32_t* __restrict__ succ) {
    uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_flows) return;
    uint32_t chunk = flows[li];
    uint32_t a = 2 * (g.n + lrank[chunk]);
    uint32_t p = outflow[li];
    int px = g.x_of(p), py = g.y_of(p);
    uint32_t first = 0xffffffffu;
    for (int k = 0; k < 8; ++k) {
        int xx = px + c_ndx[k], yy = py + c_ndy[k];
        if (!g.inb(xx, yy)) continue;
        uint32_t nb = g.idx(xx, yy);
        if (dh[nb] == (int32_t)p && root[nb] == chunk) { first = nb; break; }
    }
    succ[a] = first != 0xffffffffu ? 2 * first : a + 1;
}
__global__ void __launch_bounds__(256) k_save_outflow(uint32_t n_flows, const uint32_t* __restrict__ flows, const int32_t* __restrict__ dh,
                                                      uint32_t* __restrict__ outflow) {
    uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li < n_flows) outflow[li] = (uint32_t)dh[flows[li]];
}
// block bookkeeping
__global__ void __launch_bounds__(256) k_boundary_flags(uint32_t n, const int32_t* __restrict__ dh0, uint32_t* __restrict__ fl) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) fl[i] = dh0[i] == -2 ? 1u : 0u;
}
__global__ void __launch_bounds__(256) k_blocks_boundary(uint32_t n, const int32_t* __restrict__ dh0, const uint32_t* __restrict__ brank,
                                                         uint32_t* __restrict__ block_of_root, uint32_t* __restrict__ block_head) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || dh0[i] != -2) return;
    block_of_root[i] = brank[i];
    block_head[brank[i]] = 2 * i;
}
__global__ void __launch_bounds__(256) k_blocks_lakes(uint32_t n, uint32_t n_boundary, uint32_t n_flows, const uint32_t* __restrict__ flows,
                                                      const uint32_t* __restrict__ lrank, uint32_t* __restrict__ block_of_root,
                                                      uint32_t* __restrict__ block_head) {
    uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_flows) return;
    uint32_t chunk = flows[li];
    block_of_root[chunk] = n_boundary + li;
    block_head[n_boundary + li] = 2 * (n + lrank[chunk]);
}
__global__ void __launch_bounds__(256) k_final_keys(uint32_t n, const int32_t* __restrict__ w0, const int32_t* __restrict__ w1,
                                                    const uint32_t* __restrict__ root, const uint32_t* __restrict__ block_of_root,
                                                    uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, int* __restrict__ bad) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int64_t t = (int64_t)n - (int64_t)w0[2 * v];
    int32_t depth = -w1[2 * v];
    if (t < 0 || t >= (int64_t)n || depth < 0) { *bad = 1; return; }
    keys[t] = ((uint64_t)block_of_root[root[v]] << 32) | (uint64_t)(uint32_t)depth;
    vals[t] = v;
}
__global__ void __launch_bounds__(256) k_indirection(uint32_t n, const int32_t* __restrict__ dh0, const uint32_t* __restrict__ root,
                                                     const uint32_t* __restrict__ size, int32_t* __restrict__ ind) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ind[i] = dh0[i] < 0 ? -(int32_t)(8u * size[i]) : (int32_t)root[i];   // 1891, 1900
}

struct PhaseClock {
    bool on; cudaStream_t st; std::chrono::steady_clock::time_point t0; const char* last;
    PhaseClock(cudaStream_t s, bool enable) : on(enable), st(s), last(nullptr) { if (on) { cudaStreamSynchronize(st); t0 = std::chrono::steady_clock::now(); } }
    void mark(const char* name) {
        if (!on) return;
        cudaStreamSynchronize(st);
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[lakes] %-14s %8.3f ms\n", name, std::chrono::duration<double>(t1 - t0).count() * 1e3);
        t0 = t1;
    }
};

int ve_l_get_lakes(ve_ctx* ctx, const double* h, int32_t* dh, int32_t* indirection, uint32_t* newh) {
    const Geom& g = ctx->g;
    PhaseClock pc(ctx->stream, ctx->dbg_lakes_trace != 0);
    const uint32_t n = g.n;
    const unsigned nb = ve_blocks(n, 256);
    cudaStream_t st = ctx->stream;
    // persistent per-call arrays live in the forest scratch of the context
    uint32_t* root = ctx->f_root;       // original lake root of every cell
    uint32_t* size = ctx->f_depth;      // tree size (valid at roots)
    uint32_t* lrank = ctx->f_jump;      // rank of non-boundary lake roots / boundary ranks later
    int32_t* dh0 = ctx->dh0;            // receivers before lake resolution
    VE_CUDA(cudaMemcpyAsync(dh0, dh, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    uint32_t* small;                    // [0] boundary count [1] lake count [2] cross edges [3] cand count [4] n_flows [6..7] maxh bits, [8] changed, [9] bad
    VE_TRY(ve_scratch(ctx, 3, 4096, (void**)&small));
    VE_CUDA(cudaMemsetAsync(small, 0, 4096, st));
    int* changed = (int*)(small + 8);
    int* bad = (int*)(small + 9);
    // (i) roots by pointer jumping
    k_root_init<<<nb, 256, 0, st>>>(n, dh0, root);
    VE_LAUNCH_CHECK();
    if (ctx->dbg_rank_launches) {
        for (int round = 0; round < 64; ++round) {
            VE_CUDA(cudaMemsetAsync(changed, 0, 4, st));
            for (int r = 0; r < 3; ++r) { k_root_jump<<<nb, 256, 0, st>>>(n, root, changed); VE_LAUNCH_CHECK(); }
            int hc = 0;
            VE_TRY(ve_d2h(ctx, &hc, changed, 4));
            if (!hc) break;
        }
    } else {
        uint32_t* flags = small + 128;                                       // 64 round flags (cleared with `small`)
        int max_rounds = 64, bps = 0;
        VE_TRY(ve_occupancy(ctx, (const void*)k_root_jump_all, 256, 0, &bps));
        unsigned grid = (unsigned)(bps * ctx->sm_count);
        if (grid > nb) grid = nb;
        uint32_t nn = n;
        void* args[] = {(void*)&nn, (void*)&root, (void*)&flags, (void*)&max_rounds};
        VE_CUDA(cudaLaunchCooperativeKernel((void*)k_root_jump_all, dim3(grid), dim3(256), args, 0, st));
        ctx->launches++;
    }
    VE_CUDA(cudaMemsetAsync(size, 0, (size_t)n * 4, st));
    k_lake_stats<<<nb, 256, 0, st>>>(n, dh0, root, h, size, (unsigned long long*)(small + 6), small);
    VE_LAUNCH_CHECK();
    k_indirection<<<nb, 256, 0, st>>>(n, dh0, root, size, indirection);
    VE_LAUNCH_CHECK();
    pc.mark("roots");
    // BFS keys of the dh0 forest
    uint32_t* succ; int32_t *w0, *w1; uint64_t* bfskey;
    VE_TRY(ve_scratch(ctx, 0, ((size_t)n * 2 + 64) * 4 * 2, (void**)&succ));   // sized for the final tour too
    VE_TRY(ve_scratch(ctx, 1, ((size_t)n * 2 + 64) * 4 * 2, (void**)&w0));
    VE_TRY(ve_scratch(ctx, 2, ((size_t)n * 2 + 64) * 4 * 2, (void**)&w1));
    VE_TRY(ve_scratch(ctx, 4, (size_t)n * 8, (void**)&bfskey));
    k_tree_arcs<<<nb, 256, 0, st>>>(g, dh0, succ, w0, w1);
    VE_LAUNCH_CHECK();
    VE_TRY(ve_list_rank(ctx, 2 * n, succ, w0, w1));
    k_bfs_key<<<nb, 256, 0, st>>>(n, w0, w1, bfskey, bad);
    VE_LAUNCH_CHECK();
    pc.mark("bfs keys");
    // lake ranks (non-boundary roots, ascending index)
    uint32_t* flagbuf = (uint32_t*)w0; // reuse
    size_t tmp_bytes = 0; void* tmp;
    k_lake_flags<<<nb, 256, 0, st>>>(n, dh0, flagbuf);
    VE_LAUNCH_CHECK();
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flagbuf, lrank, (int)n, st);
    VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
    VE_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flagbuf, lrank, (int)n, st));
    ctx->launches += 2;
    // (ii) count cross edges, size the table
    PassTable t{nullptr, nullptr, nullptr, nullptr, 0};
    k_pass_scan<<<nb, 256, 0, st>>>(g, 0, h, dh0, root, bfskey, t, small + 2);
    VE_LAUNCH_CHECK();
    uint32_t hs[16];
    VE_TRY(ve_d2h(ctx, hs, small, 64));
    const uint32_t n_boundary = hs[0], n_lakes = hs[1], n_cross = hs[2];
    { unsigned long long mb; memcpy(&mb, hs + 6, 8); double mh; memcpy(&mh, &mb, 8); ctx->maxh = mh; }
    ctx->boundary_len = n_boundary; ctx->n_lakes = n_lakes;
    if (hs[9]) { ctx->err = "get_lakes: tree depth overflow"; return VE_ERR_TOO_LARGE; }
    uint32_t n_flows = 0;
    uint32_t *flows = nullptr, *outflow = nullptr;
    VE_TRY(ve_scratch(ctx, 5, ((size_t)n_lakes + 1) * 4 * 3, (void**)&flows));
    outflow = flows + n_lakes + 1;
    if (n_lakes > 0) {
        uint32_t cap = 64;
        while (cap < 2 * n_cross + 2) cap <<= 1;
        uint8_t* tbl;
        VE_TRY(ve_scratch(ctx, 6, (size_t)cap * 32, (void**)&tbl));
        t.key = (uint64_t*)tbl; t.hb = (unsigned long long*)(tbl + (size_t)cap * 8);
        t.tie = (unsigned long long*)(tbl + (size_t)cap * 16); t.edge = (uint64_t*)(tbl + (size_t)cap * 24);
        t.mask = cap - 1;
        auto build_table = [&]() -> int {
            VE_CUDA(cudaMemsetAsync(t.key, 0, (size_t)cap * 8, st));
            VE_CUDA(cudaMemsetAsync(t.hb, 0xff, (size_t)cap * 16, st)); // hb and tie = max
            for (int mode = 1; mode <= 3; ++mode) { k_pass_scan<<<nb, 256, 0, st>>>(g, mode, h, dh0, root, bfskey, t, small + 2); VE_LAUNCH_CHECK(); }
            return VE_OK;
        };
        VE_TRY(build_table());
        pc.mark("pass table");
        // (iii)/(iv): parallel growth; the one-warp heap of round 1 stays as the path for degenerate inputs
        // (a giant equal-height group, no fixed point within the round limit) and behind ve_set_debug("lakes_serial")
        int grown = 0;
        if (!ctx->dbg_lakes_serial) {
            uint32_t* lake_root_p = outflow + n_lakes + 1;
            k_lake_roots<<<nb, 256, 0, st>>>(n, dh0, lrank, lake_root_p);
            VE_LAUNCH_CHECK();
            VE_TRY(lake_grow_parallel(ctx, t, cap, n_cross, n_lakes, dh0, lrank, lake_root_p, dh, indirection, flows, small, &grown));
            if (grown) { n_flows = n_lakes; pc.mark("growth (parallel)"); }
            else {   // the growth's list ranking may have recycled the table's memory
                VE_TRY(ve_scratch(ctx, 6, (size_t)cap * 32, (void**)&tbl));
                t.key = (uint64_t*)tbl; t.hb = (unsigned long long*)(tbl + (size_t)cap * 8);
                t.tie = (unsigned long long*)(tbl + (size_t)cap * 16); t.edge = (uint64_t*)(tbl + (size_t)cap * 24);
                VE_TRY(build_table());
            }
        }
        if (!grown) {
        // adjacency (CSR over non-boundary lakes)
        uint32_t *deg, *off, *lake_root, *adj_tgt, *cidx, *cidx2; uint64_t *adj_ft, *adj_k0, *ck0, *ck1, *ck0s, *ck1s; uint4 *cand_pl, *cand_pl_s;
        size_t lk = (size_t)n_lakes + 1;
        uint8_t* ab;
        size_t adj_cap = 2 * (size_t)n_cross + 2;
        size_t lk16 = ((lk * 4 + 15) / 16) * 16, ac8 = adj_cap * 8, ac4 = ((adj_cap * 4 + 15) / 16) * 16;
        size_t lk8 = ((lk * 8 + 15) / 16) * 16;
        size_t ac16 = adj_cap * 16;
        VE_TRY(ve_scratch(ctx, 7, lk16 * 3 + ac16 * 2 + ac8 * 6 + ac4 * 3 + lk8 * 2 + 256, (void**)&ab));
        deg = (uint32_t*)ab; off = (uint32_t*)(ab + lk16); lake_root = (uint32_t*)(ab + 2 * lk16);
        uint8_t* q = ab + 3 * lk16;
        cand_pl = (uint4*)q; q += ac16; cand_pl_s = (uint4*)q; q += ac16;
        adj_ft = (uint64_t*)q; q += ac8; adj_k0 = (uint64_t*)q; q += ac8; ck0 = (uint64_t*)q; q += ac8; ck1 = (uint64_t*)q; q += ac8;
        ck0s = (uint64_t*)q; q += ac8; ck1s = (uint64_t*)q; q += ac8;
        adj_tgt = (uint32_t*)q; q += ac4;
        cidx = (uint32_t*)q; q += ac4; cidx2 = (uint32_t*)q; q += ac4;
        uint64_t* best0 = (uint64_t*)q; q += lk8; uint64_t* best1 = (uint64_t*)q; q += lk8;
        VE_CUDA(cudaMemsetAsync(best0, 0xff, lk8 * 2, st));
        VE_CUDA(cudaMemsetAsync(deg, 0, lk * 4, st));
        k_lake_roots<<<nb, 256, 0, st>>>(n, dh0, lrank, lake_root);
        VE_LAUNCH_CHECK();
        k_adj_build<<<ve_blocks(cap, 256), 256, 0, st>>>(cap, 0, t, dh0, lrank, off, deg, adj_ft, adj_tgt, adj_k0, ck0, ck1, cand_pl, small + 3);
        VE_LAUNCH_CHECK();
        cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, deg, off, (int)lk, st);
        VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
        VE_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, deg, off, (int)lk, st));
        ctx->launches += 2;
        VE_CUDA(cudaMemsetAsync(deg, 0, lk * 4, st));
        k_adj_build<<<ve_blocks(cap, 256), 256, 0, st>>>(cap, 1, t, dh0, lrank, off, deg, adj_ft, adj_tgt, adj_k0, ck0, ck1, cand_pl, small + 3);
        VE_LAUNCH_CHECK();
        uint32_t n_cand = 0;
        VE_TRY(ve_d2h(ctx, &n_cand, small + 3, 4));
        // initial heap = candidates sorted by (k0, k1) (a sorted array is a valid heap): LSD, k1 first
        uint64_t *h0 = ck0, *h1 = ck1; uint4* h2 = cand_pl;
        if (n_cand > 1) {
            unsigned cb = ve_blocks(n_cand, 256);
            uint64_t* keytmp = ck0s;       // sorted-key output we do not need
            k_iota32<<<cb, 256, 0, st>>>(n_cand, cidx);
            VE_LAUNCH_CHECK();
            cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, ck1, ck1s, cidx, cidx2, (int)n_cand, 0, 64, st);
            VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
            VE_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ck1, ck1s, cidx, cidx2, (int)n_cand, 0, 64, st));
            k_gather_u64<<<cb, 256, 0, st>>>(n_cand, cidx2, ck0, ck1s);                 // k0 in k1-sorted order
            VE_LAUNCH_CHECK();
            VE_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ck1s, keytmp, cidx2, cidx, (int)n_cand, 0, 64, st));
            // cidx = permutation sorted by (k0, k1); gather the three columns into the heap arrays
            uint64_t* g0 = ck0s; uint64_t* g1 = ck1s;
            k_gather3<<<cb, 256, 0, st>>>(n_cand, cidx, ck0, ck1, cand_pl, g0, g1, cand_pl_s);
            VE_LAUNCH_CHECK();
            h0 = g0; h1 = g1; h2 = cand_pl_s;
            ctx->launches += 8;
        }
        pc.mark("adjacency");
        uint8_t* connected;
        VE_TRY(ve_scratch(ctx, 4, std::max((size_t)n * 8, lk), (void**)&connected)); // bfskey no longer needed
        VE_CUDA(cudaMemsetAsync(connected, 0, lk, st));
        {
            int occ_unused = 0;   // per device: opts k_lake_connect in to 200 KB of dynamic shared memory
            VE_TRY(ve_occupancy(ctx, (const void*)k_lake_connect, 32, (size_t)HEAP_SMEM_MAX * 32, &occ_unused));
            // shared memory: connected bitmap first (if it fits), the top of the heap in what is left
            uint32_t conn_words = (uint32_t)((lk + 31) / 32);
            if ((size_t)conn_words * 4 > 96 * 1024) conn_words = 0;
            uint32_t heap_smem = (uint32_t)((HEAP_SMEM_MAX * 32 - (size_t)conn_words * 4) / 32);
            k_lake_connect<<<1, 32, (size_t)heap_smem * 32 + (size_t)conn_words * 4, st>>>(lake_root, off, adj_ft, adj_tgt, adj_k0, h0, h1, h2, n_cand, dh,
                                                                  indirection, connected, best0, best1, flows, small + 12, heap_smem, conn_words);
        }
        VE_LAUNCH_CHECK();
        uint32_t cstat[6] = {0, 0, 0, 0, 0, 0};
        VE_TRY(ve_d2h(ctx, cstat, small + 12, 24));
        n_flows = cstat[0];
        if (pc.on) fprintf(stderr, "[lakes] lakes %u cand %u pops %u pushes %u; kcycles sift %u scan %u push %u\n", n_lakes, n_cand, cstat[1], cstat[2], cstat[3], cstat[4], cstat[5]);
        if (n_flows != n_lakes) { ctx->err = "Disconnected lake!"; return VE_ERR_DISCONNECTED_LAKE; }
        pc.mark("connect");
        }
        // (v) flood: per-lake stacks carved out of one array by an exclusive scan of tree sizes
        k_save_outflow<<<ve_blocks(n_flows, 256), 256, 0, st>>>(n_flows, flows, dh, outflow);
        VE_LAUNCH_CHECK();
        uint32_t* start = (uint32_t*)w1; uint32_t* stack = (uint32_t*)w1 + n; uint8_t* tag = (uint8_t*)w0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, size, start, (int)n, st);
        VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
        VE_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, size, start, (int)n, st));
        ctx->launches += 2;
        VE_CUDA(cudaMemsetAsync(tag, 0, n, st));
        k_lake_fill<<<ve_blocks(n_flows, 128), 128, 0, st>>>(g, n_flows, flows, h, root, start, dh, indirection, tag, stack, bad);
        VE_LAUNCH_CHECK();
    }
    pc.mark("flood");
    // final order: blocks, Euler tour with virtual lake roots, ranking, stable sort by (block, depth)
    uint32_t n_blocks = n_boundary + n_flows;
    uint32_t *brank = (uint32_t*)bfskey /*scratch 4 (>= 8n bytes)*/, *block_of_root = brank + n;
    uint32_t* block_head;
    VE_TRY(ve_scratch(ctx, 6, ((size_t)n_blocks + 1) * 4, (void**)&block_head));
    uint32_t* flagbuf2 = (uint32_t*)w0;
    k_boundary_flags<<<nb, 256, 0, st>>>(n, dh0, flagbuf2);
    VE_LAUNCH_CHECK();
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flagbuf2, brank, (int)n, st);
    VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
    VE_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flagbuf2, brank, (int)n, st));
    ctx->launches += 2;
    k_blocks_boundary<<<nb, 256, 0, st>>>(n, dh0, brank, block_of_root, block_head);
    VE_LAUNCH_CHECK();
    if (n_flows > 0) {
        k_blocks_lakes<<<ve_blocks(n_flows, 256), 256, 0, st>>>(n, n_boundary, n_flows, flows, lrank, block_of_root, block_head);
        VE_LAUNCH_CHECK();
    }
    uint32_t n_arcs = 2 * (n + n_lakes);
    k_final_arcs<<<nb, 256, 0, st>>>(g, dh, dh0, root, lrank, block_of_root, block_head, n_blocks, succ, w0, w1);
    VE_LAUNCH_CHECK();
    if (n_flows > 0) {
        unsigned fb = ve_blocks(n_flows, 256);
        k_final_virtual<<<fb, 256, 0, st>>>(g, n_flows, flows, dh, root, lrank, block_of_root, block_head, n_blocks, succ, w0, w1);
        VE_LAUNCH_CHECK();
        k_final_virtual_first<<<fb, 256, 0, st>>>(g, n_flows, flows, outflow, dh, root, lrank, succ);
        VE_LAUNCH_CHECK();
    }
    VE_TRY(ve_list_rank(ctx, n_arcs, succ, w0, w1));
    uint64_t *keys = (uint64_t*)succ, *keys_out; uint32_t *vals, *vals_out = newh;
    VE_TRY(ve_scratch(ctx, 6, (size_t)n * 8, (void**)&keys_out));
    VE_TRY(ve_scratch(ctx, 7, (size_t)n * 4, (void**)&vals));
    VE_CUDA(cudaMemsetAsync(vals, 0xff, (size_t)n * 4, st));
    k_final_keys<<<nb, 256, 0, st>>>(n, w0, w1, root, block_of_root, keys, vals, bad);
    VE_LAUNCH_CHECK();
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys_out, vals, vals_out, (int)n, 0, 64, st);
    VE_TRY(ve_cub_tmp(ctx, tmp_bytes, &tmp));
    VE_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys_out, vals, vals_out, (int)n, 0, 64, st));
    ctx->launches += 4;
    pc.mark("final order");
    int hbad = 0;
    VE_TRY(ve_d2h(ctx, &hbad, bad, 4));
    if (hbad) { ctx->err = "get_lakes: inconsistent lake graph"; return VE_ERR_LAKE_GRAPH; }
    return VE_OK;
}
