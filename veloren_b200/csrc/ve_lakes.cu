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
    for (int round = 0; round < 64; ++round) {
        VE_CUDA(cudaMemsetAsync(changed, 0, 4, st));
        for (int r = 0; r < 3; ++r) { k_root_jump<<<nb, 256, 0, st>>>(n, root, changed); VE_LAUNCH_CHECK(); }
        int hc = 0;
        VE_CUDA(cudaMemcpyAsync(&hc, changed, 4, cudaMemcpyDeviceToHost, st));
        VE_CUDA(cudaStreamSynchronize(st));
        if (!hc) break;
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
    VE_CUDA(cudaMemcpyAsync(hs, small, 64, cudaMemcpyDeviceToHost, st));
    VE_CUDA(cudaStreamSynchronize(st));
    const uint32_t n_boundary = hs[0], n_lakes = hs[1], n_cross = hs[2];
    { unsigned long long mb; memcpy(&mb, hs + 6, 8); double mh; memcpy(&mh, &mb, 8); ctx->maxh = mh; }
    ctx->boundary_len = n_boundary; ctx->n_lakes = n_lakes;
    if (hs[9]) { ctx->err = "get_lakes: tree depth overflow"; return VE_ERR_TOO_LARGE; }
    uint32_t n_flows = 0;
    uint32_t *flows = nullptr, *outflow = nullptr;
    VE_TRY(ve_scratch(ctx, 5, ((size_t)n_lakes + 1) * 4 * 2, (void**)&flows));
    outflow = flows + n_lakes + 1;
    if (n_lakes > 0) {
        uint32_t cap = 64;
        while (cap < 2 * n_cross + 2) cap <<= 1;
        uint8_t* tbl;
        VE_TRY(ve_scratch(ctx, 6, (size_t)cap * 32, (void**)&tbl));
        t.key = (uint64_t*)tbl; t.hb = (unsigned long long*)(tbl + (size_t)cap * 8);
        t.tie = (unsigned long long*)(tbl + (size_t)cap * 16); t.edge = (uint64_t*)(tbl + (size_t)cap * 24);
        t.mask = cap - 1;
        VE_CUDA(cudaMemsetAsync(t.key, 0, (size_t)cap * 8, st));
        VE_CUDA(cudaMemsetAsync(t.hb, 0xff, (size_t)cap * 16, st)); // hb and tie = max
        for (int mode = 1; mode <= 3; ++mode) { k_pass_scan<<<nb, 256, 0, st>>>(g, mode, h, dh0, root, bfskey, t, small + 2); VE_LAUNCH_CHECK(); }
        pc.mark("pass table");
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
        VE_CUDA(cudaMemcpyAsync(&n_cand, small + 3, 4, cudaMemcpyDeviceToHost, st));
        VE_CUDA(cudaStreamSynchronize(st));
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
        VE_CUDA(cudaMemcpyAsync(cstat, small + 12, 24, cudaMemcpyDeviceToHost, st));
        VE_CUDA(cudaStreamSynchronize(st));
        n_flows = cstat[0];
        if (pc.on) fprintf(stderr, "[lakes] lakes %u cand %u pops %u pushes %u; kcycles sift %u scan %u push %u\n", n_lakes, n_cand, cstat[1], cstat[2], cstat[3], cstat[4], cstat[5]);
        if (n_flows != n_lakes) { ctx->err = "Disconnected lake!"; return VE_ERR_DISCONNECTED_LAKE; }
        pc.mark("connect");
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
    VE_CUDA(cudaMemcpyAsync(&hbad, bad, 4, cudaMemcpyDeviceToHost, st));
    VE_CUDA(cudaStreamSynchronize(st));
    if (hbad) { ctx->err = "get_lakes: inconsistent lake graph"; return VE_ERR_LAKE_GRAPH; }
    return VE_OK;
}
