// euler.cu -- breadth-first ORDER of grid spanning trees without walking them depth by depth (sm_100a).
//
// Same outputs as fedicra_bfs_grid / the reference's bfs_forward (LTF/src/bfs/bfs.cu:46-98 semantics, canonical order:
// depth, then parent position, then vertex id), but every step is data-parallel over the whole batch.  A level-by-level
// walk needs one dependent step per tree depth (1 100 - 2 200 at 256^2); here the depth of the dependency chain is
// logarithmic:
//   1. Euler tour of the tree (each tree edge = two arcs; successor of the arc into u = next tree edge around u)
//      ranked with a splitter walk: one thread per 8 vertices walks its stretch of the tour (~16 arcs), the ~8 192
//      stretches are ranked in shared memory (two short walks between pseudo-random leaders + pointer jumping over the
//      leaders).
//   2. From the ranks: the parent of u is where the tour first enters u; subtree size = half the span between entering
//      and leaving u.
//   3. One prefix sum along the tour of (+1,-1) gives the depth, of (+o,-o) with o(u) = 1 + sizes of u's smaller siblings
//      gives the preorder index that visits children by increasing vertex id.
//   4. BFS position = level_start[depth] + rank of the preorder index among the vertices of that depth (the BFS queue
//      order equals "sort by depth, then preorder"): counted per level.
// Integer-only, deterministic (atomics are used for counters whose order does not reach the result).
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int NT = 256;
constexpr int NIL = -1;
#ifndef SCAN_NT_
#define SCAN_NT_ 256
#endif
#ifndef SCAN_PER_
#define SCAN_PER_ 8
#endif
constexpr int SCAN_NT = SCAN_NT_, SCAN_PER = SCAN_PER_, SCAN_TILE = SCAN_NT * SCAN_PER;   // tiles of the single-pass tour scan
// splitters handed out dynamically inside one CTA of the tour walk (FEDICRA_EULER_SPC overrides; see run time notes below)
// Every walk CTA first packs the whole tree mask into shared memory, so larger images want fewer, longer-lived CTAs:
// measured on B200, chain order of 48 trees: 256^2: 512 -> 313 us, 1024 -> 309, 2048 -> 334; 512^2: 512 -> 1 694 us,
// 1024 -> 1 499, 2048 -> 1 386.
int split_per_cta(int V) {
    const char *v = getenv("FEDICRA_EULER_SPC");
    const int dflt = V > 131072 ? 2048 : 1024;
    const int x = (v && *v) ? atoi(v) : dflt;
    return (x >= 32 && x <= 8192) ? x : dflt;
}

__device__ __forceinline__ int dir_delta(int d, int W) { return d == 0 ? -W : (d == 1 ? -1 : (d == 2 ? 1 : W)); }

// Splitter s sits on a pseudo-randomly chosen vertex of its block of SP = 2^sp_log consecutive vertices (a regular
// placement would put all splitters on a few image columns, and a tour can run for thousands of arcs between two
// columns).  splitter_vertex() and is_splitter_vertex() are inverse to each other.
__device__ __forceinline__ int splitter_vertex(int s, int sp_log, int V) {
    const int v = (s << sp_log) + (int)(((unsigned)s * 2654435761u) >> (32 - sp_log));
    return v < V ? v : (s << sp_log);
}
__device__ __forceinline__ bool is_splitter_vertex(int v, int sp_log, int V) { return splitter_vertex(v >> sp_log, sp_log, V) == v; }

// next tree edge around a vertex with adjacency mask m, cyclically after direction r (r itself when it is the only one)
__device__ __forceinline__ int next_dir(unsigned m, int r) {
    const unsigned rot = ((m | (m << 4)) >> (r + 1)) & 0xFu;
    return (r + __ffs(rot)) & 3;
}

struct Ws {
    unsigned *arc;                  // [B,4V] splitter owning the arc << 18 | offset of the arc inside the splitter's stretch
                                    //        (packed so that the whole table stays L2-resident: its writes are random)
    int *snext, *slen, *sbase;      // [B,NS] next splitter on the tour, stretch length, tour rank of the splitter arc
    int *rin, *rout, *pdir, *size;  // [B,V]  tour rank entering / leaving u, direction of the parent (4 = root), subtree size
    int2 *S;                        // [B,2V] tour-ordered (d preorder = +-o, vertex entered or -1); d depth = sign(d preorder), o >= 1
    int2 *bsum;                     // [B,NBLK] per-block sums of S
    int *depth, *pre;               // [B,V]
    int *cnt, *cursor;              // [B,V+1] vertices per depth; bucket fill cursors   (zeroed per call)
    int2 *tmp;                      // [B,V]  per level, unordered: (preorder, vertex)
    int *pos;                       // [B,V]  BFS position of a vertex
    int *hist;                      // [B,NTILE,32] chain order: vertices per (1024-wide preorder tile, light depth)   (zeroed per call)
    unsigned long long *scan_state; // [B,NSCAN]    chain order: look-back words of the single-pass tour scan             (zeroed per call)
    int *scan_ticket;               // [B]          chain order: tile tickets of that scan                                (zeroed per call)
};

struct Dims {
    int root;  // the vertex the order starts from (0 = the reference's, bfs.cu:69-74)
    int B, V, H, W, SP, SPLOG, NS, T, NBLK, SPC;  // T = 2(V-1) arcs on the tour; NBLK = ceil(T / 1024); SPC splitters per walk CTA
    int NTILE;                                    // ceil(V / 1024)
    int NSCAN;                                    // ceil(T / SCAN_TILE), at least 1: tiles of the single-pass tour scan
};

size_t ws_layout(const Dims &d, char *base, Ws *w) {
    size_t offb = 0;
    auto take = [&](size_t bytes) {
        char *p = base ? base + offb : nullptr;
        offb += align_up(bytes, 256);
        return p;
    };
    const size_t B = d.B, V = d.V;
    Ws t;
    t.cnt = (int *)take(B * (V + 1) * 4);      // cnt and cursor first and adjacent: one memset clears both
    t.cursor = (int *)take(B * (V + 1) * 4);
    t.hist = (int *)take(B * (size_t)d.NTILE * 32 * 4);   // hist .. arc are one range: the chain order clears it at once
    t.scan_state = (unsigned long long *)take(B * (size_t)d.NSCAN * 8);
    t.scan_ticket = (int *)take(B * 4);
    t.arc = (unsigned *)take(B * 4 * V * 4);
    t.snext = (int *)take(B * d.NS * 4);
    t.slen = (int *)take(B * d.NS * 4);
    t.sbase = (int *)take(B * d.NS * 4);
    t.rin = (int *)take(B * V * 4);
    t.rout = (int *)take(B * V * 4);
    t.pdir = (int *)take(B * V * 4);
    t.size = (int *)take(B * V * 4);
    t.S = (int2 *)take(B * 2 * V * 8);
    t.bsum = (int2 *)take(B * (size_t)(d.NBLK > 0 ? d.NBLK : 1) * 8);
    t.depth = (int *)take(B * V * 4);
    t.pre = (int *)take(B * V * 4);
    t.tmp = (int2 *)take(B * V * 8);
    t.pos = (int *)take(B * V * 4);
    if (w) *w = t;
    return offb;
}

Dims make_dims(int B, int H, int W, int root = 0) {
    Dims d;
    d.root = root;
    d.B = B; d.H = H; d.W = W; d.V = H * W;
    {   // splitter spacing: FEDICRA_EULER_SPLOG overrides (2^k vertices per splitter)
        const char *v = getenv("FEDICRA_EULER_SPLOG");
        // measured on B200, 48 trees of 256^2 (whole BFS): 8 vertices per splitter 444 us, 16 -> 460 us, 32 -> 570 us
        const int k = (v && *v) ? atoi(v) : 3;
        d.SPLOG = (k >= 1 && k <= 10) ? k : 3;
        d.SP = 1 << d.SPLOG;
    }
    while ((d.V + d.SP - 1) / d.SP > 16384) { d.SP <<= 1; ++d.SPLOG; }  // at most 16 384 splitters per image (shared-memory ranking: 12 bytes each)
    d.NS = (d.V + d.SP - 1) / d.SP;
    d.T = 2 * (d.V - 1);
    d.NBLK = (d.T + 1023) / 1024;
    d.NTILE = (d.V + 1023) / 1024;
    d.NSCAN = d.T > 0 ? (d.T + SCAN_TILE - 1) / SCAN_TILE : 1;
    d.SPC = split_per_cta(d.V);
    return d;
}

// ---- 1a. every splitter walks its stretch of the tour.  The walk is a chain of dependent mask look-ups (~32 on average,
// a few hundred for the longest stretch), so every CTA first copies the image's masks into shared memory, two per byte.
template <bool SMEM>
__global__ void __launch_bounds__(NT) walk_kernel(const unsigned char *__restrict__ mask, Ws ws, Dims dm) {
    extern __shared__ unsigned char s_mask[];
    const int b = blockIdx.y, V = dm.V, W = dm.W, spl = dm.SPLOG;
    const unsigned char *m = mask + (size_t)b * V;
    if (SMEM) {
        if ((V & 15) == 0 && (((uintptr_t)m) & 15) == 0) {  // 16 masks per load, packed to 8 bytes
            const uint4 *m4 = reinterpret_cast<const uint4 *>(m);
            uint2 *s2 = reinterpret_cast<uint2 *>(s_mask);
            for (int k = threadIdx.x; k < V / 16; k += NT) {
                const uint4 q = m4[k];
                auto pack = [](unsigned x) {  // 4 bytes (one mask each) -> 16 bits (one nibble each)
                    return (x & 0xFu) | ((x >> 4) & 0xF0u) | ((x >> 8) & 0xF00u) | ((x >> 12) & 0xF000u);
                };
                s2[k] = make_uint2(pack(q.x) | (pack(q.y) << 16), pack(q.z) | (pack(q.w) << 16));
            }
        } else {
            for (int k = threadIdx.x; k < (V + 1) / 2; k += NT) {
                const unsigned lo = m[2 * k], hi = (2 * k + 1 < V) ? m[2 * k + 1] : 0u;
                s_mask[k] = (unsigned char)(lo | (hi << 4));
            }
        }
        __syncthreads();
    }
    auto mask_of = [&](int v) -> unsigned {
        if (SMEM) {
            const unsigned byte = s_mask[v >> 1];
            return (v & 1) ? (byte >> 4) : (byte & 15u);
        }
        return m[v];
    };
    // Stretch lengths are geometric (mean ~2 SP, longest a few hundred), so a fixed splitter per thread leaves most lanes
    // idle; instead every CTA owns SPC consecutive splitters and its lanes pull the next one when they are done.
    __shared__ int s_next;
    const int s_begin = blockIdx.x * dm.SPC, s_end = min(s_begin + dm.SPC, dm.NS);
    if (threadIdx.x == 0) s_next = s_begin + 2 * NT;
    __syncthreads();
    unsigned *arc = ws.arc + (size_t)b * 4 * V;
    const int limit = 2 * V;  // a consistent tree mask closes the tour after 2(V-1) arcs; guard against a bad one
    // A lane's next splitter is fetched when it STARTS a stretch, not when it ends one: with ~16 arcs per stretch some lane
    // of a warp finishes on almost every iteration, and the shared-memory atomic's latency would sit on every one of them.
    int s = s_begin + threadIdx.x, sn = s_begin + NT + threadIdx.x, v = 0, d = 0, steps = 0;
    bool walking = false;
    while (true) {
        if (!walking) {
            if (s >= s_end) break;
            v = splitter_vertex(s, spl, V);
            const unsigned mv0 = mask_of(v);
            if (mv0 == 0) {  // single-vertex tree
                ws.snext[(size_t)b * dm.NS + s] = s;
                ws.slen[(size_t)b * dm.NS + s] = 0;
                s = sn;
                sn = atomicAdd(&s_next, 1);
                continue;
            }
            d = __ffs(mv0) - 1;
            steps = 0;
            walking = true;
        }
        arc[4 * v + d] = ((unsigned)s << 18) | ((unsigned)steps & 0x3ffffu);
        ++steps;
        v += dir_delta(d, W);
        const unsigned mv = mask_of(v);
        d = next_dir(mv, 3 - d);
        if ((d == __ffs(mv) - 1 && is_splitter_vertex(v, spl, V)) || steps >= limit) {
            ws.snext[(size_t)b * dm.NS + s] = v >> spl;
            ws.slen[(size_t)b * dm.NS + s] = steps;
            walking = false;
            s = sn;
            sn = atomicAdd(&s_next, 1);
        }
    }
}

// ---- 1b. rank the splitters (tour rank of every stretch's first arc), one CTA per image, in shared memory.  Pointer
// jumping over all NS splitters costs log2(NS) rounds of random shared-memory gathers (27 us for 8 192); instead every
// ~SPACING-th splitter is a LEADER: each leader walks to the next leader summing the stretch lengths on the way (~SPACING
// dependent steps), pointer jumping ranks the <= 1 024 leaders, and a second walk hands the ranks down the segments.
__global__ void __launch_bounds__(1024) rank_splitters_kernel(Ws ws, Dims dm) {
    extern __shared__ int s_mem[];
    const int b = blockIdx.x, tid = threadIdx.x, NS = dm.NS;
    int *len = s_mem;                                               // [NS] stretch length
    short *nxt = reinterpret_cast<short *>(s_mem + NS);            // [NS] next splitter on the tour (NS <= 16 384), NIL at the cut
    // One leader per block of `spacing` (a power of two) consecutive splitters, at a pseudo-random place in the block: the
    // tour mostly steps to a spatial neighbour's splitter, i.e. index +-1 or +- one image row of blocks, so leaders at
    // fixed residues would be met after ~spacing^2 steps (measured: 150 us), random ones after ~spacing.
    int lg = 0;
    while ((1024 << lg) < NS) ++lg;
    const int spacing = 1 << lg, L = (NS + spacing - 1) >> lg;
    auto leader_of = [&](int l) {
        const int s = (l << lg) + ((l == 0 || lg == 0) ? 0 : (int)(((unsigned)l * 2654435761u) >> (32 - lg)));   // splitter 0 (the head) is a leader
        return s < NS ? s : (l << lg);
    };
    auto is_leader = [&](int s) { return leader_of(s >> lg) == s; };
    int *lval[2] = {s_mem + NS + (NS + 1) / 2, s_mem + NS + (NS + 1) / 2 + 1024};       // [L] arcs from the leader to the end
    short *lnxt[2] = {reinterpret_cast<short *>(lval[1] + 1024), reinterpret_cast<short *>(lval[1] + 1024) + 1024};   // next leader (index)
    const int *snext = ws.snext + (size_t)b * NS, *slen = ws.slen + (size_t)b * NS;
    int *sbase = ws.sbase + (size_t)b * NS;
    for (int s = tid; s < NS; s += 1024) {
        const int n = snext[s];
        len[s] = slen[s];
        nxt[s] = (short)((n == 0) ? NIL : n);  // the tour is a cycle through splitter 0 (vertex 0, the root): cut it there
        sbase[s] = 0;                           // (a splitter the walk below never reaches: inconsistent mask)
    }
    __syncthreads();
    int seg = 0, to = NIL;
    if (tid < L) {   // first walk: own stretch, then on to the next leader
        int s = leader_of(tid), steps = 0;
        do {
            seg += len[s];
            s = nxt[s];
        } while (s >= 0 && !is_leader(s) && ++steps < NS);
        to = (s >= 0 && is_leader(s)) ? (s >> lg) : NIL;
        lval[0][tid] = seg;
        lnxt[0][tid] = (short)to;
    }
    __syncthreads();
    int cur = 0;
    for (int span = 1; span < L; span <<= 1) {
        if (tid < L) {
            const int n = lnxt[cur][tid];
            lval[cur ^ 1][tid] = lval[cur][tid] + (n >= 0 ? lval[cur][n] : 0);
            lnxt[cur ^ 1][tid] = (short)((n >= 0) ? lnxt[cur][n] : NIL);
        }
        __syncthreads();
        cur ^= 1;
    }
    const int total = lval[cur][0];   // splitter 0 is a leader and the head of the list
    if (tid < L) {   // second walk: arcs before a stretch = total - (arcs from it to the end)
        int s = leader_of(tid), rest = lval[cur][tid], steps = 0;
        do {
            sbase[s] = total - rest;
            rest -= len[s];
            s = nxt[s];
        } while (s >= 0 && !is_leader(s) && ++steps < NS);
    }
}

__device__ __forceinline__ int arc_rank_raw(const unsigned *arc, const int *sbase, int a) {
    const unsigned r = arc[a];
    return sbase[r >> 18] + (int)(r & 0x3ffffu);
}
// The splitter ranking cuts the (cyclic) tour at vertex 0.  A tour that starts at another root is the same cycle rotated:
// rank' = rank - rank(first arc out of the root) mod T.
__device__ __forceinline__ int root_shift(const unsigned char *mask_b, const unsigned *arc, const int *sbase, int root) {
    if (root == 0) return 0;
    const unsigned m = mask_b[root];
    return m ? arc_rank_raw(arc, sbase, 4 * root + __ffs(m) - 1) : 0;
}
__device__ __forceinline__ int arc_rank(const unsigned *arc, const int *sbase, int a, int shift, int T) {
    const int r = arc_rank_raw(arc, sbase, a) - shift;
    return r < 0 ? r + T : r;
}

// ---- 2. where the tour enters and leaves every vertex
__global__ void __launch_bounds__(NT) enter_leave_kernel(const unsigned char *__restrict__ mask, Ws ws, Dims dm) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= (long long)dm.B * dm.V) return;
    const int b = (int)(idx / dm.V), u = (int)(idx - (long long)b * dm.V), V = dm.V, W = dm.W;
    const unsigned *arc = ws.arc + (size_t)b * 4 * V;
    const int *sbase = ws.sbase + (size_t)b * dm.NS;
    const size_t o = (size_t)b * V + u;
    if (u == dm.root) {
        ws.rin[o] = -1; ws.rout[o] = dm.T; ws.pdir[o] = 4; ws.size[o] = V;
        return;
    }
    const int shift = root_shift(mask + (size_t)b * V, arc, sbase, dm.root), T = dm.T;
    const unsigned m = mask[o];
    int best = 0x7fffffff, pd = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (m & (1u << d)) {
            const int r = arc_rank(arc, sbase, 4 * (u + dir_delta(d, W)) + (3 - d), shift, T);  // the arc neighbour -> u
            if (r < best) { best = r; pd = d; }
        }
    }
    const int out = arc_rank(arc, sbase, 4 * u + pd, shift, T);
    ws.rin[o] = best; ws.rout[o] = out; ws.pdir[o] = pd; ws.size[o] = (out - best + 1) >> 1;
}

// ---- 3a. tour-ordered increments: +1 / -1 for the depth, +o / -o for the preorder with children by increasing id
__global__ void __launch_bounds__(NT) increments_kernel(const unsigned char *__restrict__ mask, Ws ws, Dims dm) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= (long long)dm.B * dm.V) return;
    const int b = (int)(idx / dm.V), u = (int)(idx - (long long)b * dm.V), V = dm.V, W = dm.W;
    if (u == dm.root) return;
    const size_t o = (size_t)b * V;
    const int pd = ws.pdir[o + u], p = u + dir_delta(pd, W), du = 3 - pd;  // u is p's child in direction du
    const unsigned mp = mask[o + p];
    const int ppd = ws.pdir[o + p];
    int offs = 1;
#pragma unroll
    for (int d = 0; d < 3; ++d)
        if (d < du && (mp & (1u << d)) && d != ppd) offs += ws.size[o + p + dir_delta(d, W)];
    int2 *S = ws.S + (size_t)b * 2 * V;
    S[ws.rin[o + u]] = make_int2(offs, u);
    S[ws.rout[o + u]] = make_int2(-offs, -1);
}

// ---- 3b. prefix sums along the tour (block sums, scan of block sums, apply)
__global__ void __launch_bounds__(1024) tour_block_sum_kernel(Ws ws, Dims dm) {
    const int b = blockIdx.y, blk = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int r = blk * 1024 + tid;
    int2 v = make_int2(0, 0);
    if (r < dm.T) {
        const int2 s = ws.S[(size_t)b * 2 * dm.V + r];
        v = make_int2((s.x > 0) - (s.x < 0), s.x);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
    }
    __shared__ int2 s_w[32];
    if (lane == 0) s_w[wid] = v;
    __syncthreads();
    if (wid == 0) {
        int2 t = s_w[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            t.x += __shfl_xor_sync(0xffffffffu, t.x, o);
            t.y += __shfl_xor_sync(0xffffffffu, t.y, o);
        }
        if (lane == 0) ws.bsum[(size_t)b * dm.NBLK + blk] = t;
    }
}

__global__ void __launch_bounds__(1024) tour_scan_sums_kernel(Ws ws, Dims dm) {  // one CTA per image: exclusive scan
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int2 *bs = ws.bsum + (size_t)b * dm.NBLK;
    __shared__ int2 s_w[32];
    __shared__ int2 s_carry;
    if (tid == 0) s_carry = make_int2(0, 0);
    __syncthreads();
    for (int base = 0; base < dm.NBLK; base += 1024) {
        const int i = base + tid;
        const int2 mine = (i < dm.NBLK) ? bs[i] : make_int2(0, 0);
        int2 x = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int tx = __shfl_up_sync(0xffffffffu, x.x, o), ty = __shfl_up_sync(0xffffffffu, x.y, o);
            if (lane >= o) { x.x += tx; x.y += ty; }
        }
        if (lane == 31) s_w[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int2 t = s_w[lane], y = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int tx = __shfl_up_sync(0xffffffffu, y.x, o), ty = __shfl_up_sync(0xffffffffu, y.y, o);
                if (lane >= o) { y.x += tx; y.y += ty; }
            }
            s_w[lane] = make_int2(y.x - t.x, y.y - t.y);
        }
        __syncthreads();
        const int2 carry = s_carry, wo = s_w[wid];
        if (i < dm.NBLK) bs[i] = make_int2(carry.x + wo.x + x.x - mine.x, carry.y + wo.y + x.y - mine.y);
        __syncthreads();
        if (tid == 1023) s_carry = make_int2(carry.x + wo.x + x.x, carry.y + wo.y + x.y);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024) tour_apply_kernel(Ws ws, Dims dm) {
    const int b = blockIdx.y, blk = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, V = dm.V;
    const int r = blk * 1024 + tid;
    int2 sp = make_int2(0, -1);
    if (r < dm.T) sp = ws.S[(size_t)b * 2 * V + r];
    const int3 s = make_int3((sp.x > 0) - (sp.x < 0), sp.x, sp.y);   // (d depth, d preorder, vertex entered or -1)
    int2 x = make_int2(s.x, s.y);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int tx = __shfl_up_sync(0xffffffffu, x.x, o), ty = __shfl_up_sync(0xffffffffu, x.y, o);
        if (lane >= o) { x.x += tx; x.y += ty; }
    }
    __shared__ int2 s_w[32];
    if (lane == 31) s_w[wid] = x;
    __syncthreads();
    if (wid == 0) {
        int2 t = s_w[lane], y = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int tx = __shfl_up_sync(0xffffffffu, y.x, o), ty = __shfl_up_sync(0xffffffffu, y.y, o);
            if (lane >= o) { y.x += tx; y.y += ty; }
        }
        s_w[lane] = make_int2(y.x - t.x, y.y - t.y);
    }
    __syncthreads();
    const size_t o = (size_t)b * V;
    if (blk == 0 && tid == 0) {  // the root is never entered
        ws.depth[o + dm.root] = 0;
        ws.pre[o + dm.root] = 0;
        atomicAdd(ws.cnt + (size_t)b * (V + 1), 1);
    }
    if (s.z >= 0) {  // the tour enters vertex s.z here: inclusive sums = its depth and preorder index
        const int2 base = ws.bsum[(size_t)b * dm.NBLK + blk], wo = s_w[wid];
        const int depth = base.x + wo.x + x.x, pre = base.y + wo.y + x.y;
        ws.depth[o + s.z] = depth;
        ws.pre[o + s.z] = pre;
        atomicAdd(ws.cnt + (size_t)b * (V + 1) + depth, 1);
    }
}

// ---- 4a. level_start = exclusive scan of the per-depth counts; stops as soon as all V vertices are accounted for
__global__ void __launch_bounds__(1024) level_scan_kernel(Ws ws, Dims dm, int *__restrict__ level_start,
                                                          int *__restrict__ n_levels) {
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, V = dm.V;
    const int *cnt = ws.cnt + (size_t)b * (V + 1);
    int *ls = level_start + (size_t)b * (V + 1);
    __shared__ int s_w[32];
    __shared__ int s_carry, s_levels;
    if (tid == 0) { s_carry = 0; s_levels = 0; }
    __syncthreads();
    for (int base = 0; base <= V; base += 1024) {
        const int i = base + tid;
        const int c = (i <= V) ? cnt[i] : 0;
        int x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += t;
        }
        if (lane == 31) s_w[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int t = s_w[lane], y = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int tt = __shfl_up_sync(0xffffffffu, y, o);
                if (lane >= o) y += tt;
            }
            s_w[lane] = y - t;
        }
        __syncthreads();
        const int carry = s_carry, excl = carry + s_w[wid] + x - c;
        if (i <= V && (c > 0 || excl == V)) {  // depths that exist, plus the closing entry ls[n_levels] = V
            ls[i] = excl;
            if (c > 0) atomicMax(&s_levels, i + 1);
        }
        __syncthreads();
        if (tid == 1023) s_carry = excl + c;
        __syncthreads();
        if (s_carry >= V) {  // everything placed; make sure the closing entry lands right after the last depth
            if (tid == 0) {
                ls[s_levels] = V;
                n_levels[b] = s_levels;
            }
            return;
        }
    }
    if (tid == 0) {
        ls[s_levels] = V;
        n_levels[b] = s_levels;
    }
}

// ---- 4b. drop every vertex into its level's bucket (arbitrary order inside the bucket)
__global__ void __launch_bounds__(NT) bucket_kernel(Ws ws, Dims dm, const int *__restrict__ level_start) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= (long long)dm.B * dm.V) return;
    const int b = (int)(idx / dm.V), u = (int)(idx - (long long)b * dm.V), V = dm.V;
    const int d = ws.depth[(size_t)b * V + u];
    const int slot = atomicAdd(ws.cursor + (size_t)b * (V + 1) + d, 1);
    ws.tmp[(size_t)b * V + level_start[(size_t)b * (V + 1) + d] + slot] = make_int2(ws.pre[(size_t)b * V + u], u);
}

// ---- 4c. order every level by preorder index: rank = number of smaller keys (levels are ~45 wide); one warp per level
__global__ void __launch_bounds__(NT) level_rank_kernel(Ws ws, Dims dm, const int *__restrict__ level_start,
                                                        const int *__restrict__ n_levels, int *__restrict__ sorted_index,
                                                        int warps_per_image) {
    const int b = blockIdx.y, lane = threadIdx.x & 31, V = dm.V;
    const int warp = blockIdx.x * (NT / 32) + (threadIdx.x >> 5);
    const int nl = n_levels[b];
    const int *ls = level_start + (size_t)b * (V + 1);
    const int2 *tmp = ws.tmp + (size_t)b * V;
    int *pos = ws.pos + (size_t)b * V, *order = sorted_index + (size_t)b * V;
    for (int L = warp; L < nl; L += warps_per_image) {
        const int lo = ls[L], n = ls[L + 1] - lo;
        if (n <= 128) {   // the usual case: up to four entries per lane, keys exchanged by shuffles (one memory round trip)
            constexpr int R = 4;
            int2 m[R];
            int rk[R];
#pragma unroll
            for (int q = 0; q < R; ++q) {
                m[q] = (q * 32 + lane < n) ? tmp[lo + q * 32 + lane] : make_int2(0x7fffffff, 0);
                rk[q] = 0;
            }
#pragma unroll
            for (int q = 0; q < R; ++q) {
                const int cnt = min(n - q * 32, 32);   // entries held in register q across the warp (warp-uniform)
                for (int j = 0; j < cnt; ++j) {
                    const int kj = __shfl_sync(0xffffffffu, m[q].x, j);
#pragma unroll
                    for (int p = 0; p < R; ++p) rk[p] += kj < m[p].x;
                }
            }
#pragma unroll
            for (int q = 0; q < R; ++q)
                if (q * 32 + lane < n) { order[lo + rk[q]] = m[q].y; pos[m[q].y] = lo + rk[q]; }
            continue;
        }
        for (int i = lane; i < n; i += 32) {
            const int2 me = tmp[lo + i];
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += (tmp[lo + j].x < me.x);
            order[lo + rank] = me.y;
            pos[me.y] = lo + rank;
        }
    }
}

// ---- 5. parent position and children positions of every vertex
__global__ void __launch_bounds__(NT) finalize_kernel(const unsigned char *__restrict__ mask, Ws ws, Dims dm,
                                                      int *__restrict__ sorted_parent, int *__restrict__ sorted_child) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= (long long)dm.B * dm.V) return;
    const int b = (int)(idx / dm.V), u = (int)(idx - (long long)b * dm.V), V = dm.V, W = dm.W;
    const size_t o = (size_t)b * V;
    const int *pos = ws.pos + o;
    const int i = pos[u], pd = ws.pdir[o + u];
    const unsigned m = mask[o + u];
    sorted_parent[o + i] = (u == dm.root) ? 0 : pos[u + dir_delta(pd, W)];
    int c[4] = {0, 0, 0, 0}, k = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d)
        if ((m & (1u << d)) && d != pd) c[k++] = pos[u + dir_delta(d, W)];
    reinterpret_cast<int4 *>(sorted_child)[o + i] = make_int4(c[0], c[1], c[2], c[3]);
}


// =====================================================================================================================
// Chain order: heavy-path decomposition of the rooted tree, laid out for the scan-based filter (csrc/chains.cu).
//
//   heavy child of p  = the child with the largest subtree (smallest vertex id on ties); every other child is light;
//   chain             = maximal path of heavy edges (top = the root or a light child; bottom = a leaf);
//   light depth ld(u) = number of light edges between the root and u  (<= log2 V: a light subtree is at most half
//                       its parent's);
//   position          = vertices sorted by (ld, preorder index of the traversal that visits the heavy child first).
//
// In that order a chain occupies consecutive positions top -> bottom, all chains of one light depth form one contiguous
// GROUP, the root sits at position 0, every parent precedes its children (so sorted_index / sorted_parent /
// sorted_child are valid arrays in the reference's sense, bfs.cu:46-98, with child slot 0 = the heavy child at
// position + 1), and the filter's recurrences become, per group, scans of affine maps along the chains: the depth of
// the dependency chain is the number of groups (8 - 11 for 256^2 .. 512^2 images) instead of the tree depth (~1 500).
// =====================================================================================================================

// direction of the heavy child of p (-1: leaf); sz[d] = subtree size of the child in direction d (0: no child there)
__device__ __forceinline__ int heavy_dir(unsigned m, int pd, const int *__restrict__ size_b, int p, int W, int sz[4]) {
    int best = -1, bsz = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        sz[d] = ((m >> d) & 1u) && d != pd ? size_b[p + dir_delta(d, W)] : 0;
        if (sz[d] > bsz) { bsz = sz[d]; best = d; }
    }
    return best;
}

// Tour entries of the chain order: x = +-o with o(u) = 1 + sizes of the siblings visited before u when the heavy child
// goes first and the light ones follow by increasing id;  y = (u << 1 | light(u)) where the tour enters u,
// -(1 + light(u)) where it leaves.
__device__ __forceinline__ int2 chain_delta(int2 s) {   // (d light depth, d preorder) of one tour entry
    return make_int2(s.y >= 0 ? (s.y & 1) : (s.y + 1), s.x);
}

// ---- chain order, steps 2 + 3a in one pass over the vertices: where the tour enters and leaves u (as enter_leave_kernel),
// then -- straight from the arc ranks around u's parent, without waiting for a size array -- the subtree sizes of u's
// siblings, the parent's heavy child, and u's two tour entries.
__global__ void __launch_bounds__(NT) chain_enter_kernel(const unsigned char *__restrict__ mask, Ws ws, Dims dm) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= (long long)dm.B * dm.V) return;
    const int b = (int)(idx / dm.V), u = (int)(idx - (long long)b * dm.V), V = dm.V, W = dm.W;
    const unsigned *arc = ws.arc + (size_t)b * 4 * V;
    const int *sbase = ws.sbase + (size_t)b * dm.NS;
    const size_t o = (size_t)b * V;
    if (u == dm.root) {
        ws.pdir[o + u] = 4;
        ws.size[o + u] = V;
        return;
    }
    const unsigned m = mask[o + u];
    int rin = 0x7fffffff, pd = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (m & (1u << d)) {
            const int r = arc_rank_raw(arc, sbase, 4 * (u + dir_delta(d, W)) + (3 - d));   // the arc neighbour -> u
            if (r < rin) { rin = r; pd = d; }
        }
    }
    const int rout = arc_rank_raw(arc, sbase, 4 * u + pd);
    ws.pdir[o + u] = pd;
    ws.size[o + u] = (rout - rin + 1) >> 1;
    // around the parent: ranks of the arcs into p (the smallest one comes from p's parent) and out of p
    const int p = u + dir_delta(pd, W), du = 3 - pd;
    const unsigned mp = mask[o + p];
    int pin[4], pout[4], ppd = 4, pmin = 0x7fffffff;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        pin[d] = pout[d] = 0;
        if (mp & (1u << d)) {
            pin[d] = (d == du) ? rout : arc_rank_raw(arc, sbase, 4 * (p + dir_delta(d, W)) + (3 - d));
            pout[d] = (d == du) ? rin : arc_rank_raw(arc, sbase, 4 * p + d);
            if (p != dm.root && pin[d] < pmin) { pmin = pin[d]; ppd = d; }
        }
    }
    int sz[4], hd = -1, hsz = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        sz[d] = ((mp >> d) & 1u) && d != ppd ? (pin[d] - pout[d] + 1) >> 1 : 0;
        if (sz[d] > hsz) { hsz = sz[d]; hd = d; }
    }
    int offs = 1, light = 0;
    if (hd != du) {
        light = 1;
        offs += hsz;
#pragma unroll
        for (int d = 0; d < 3; ++d)
            if (d < du && d != hd) offs += sz[d];
    }
    int2 *S = ws.S + (size_t)b * 2 * V;
    S[rin] = make_int2(offs, (u << 1) | light);
    S[rout] = make_int2(-offs, -(1 + light));
}

// ---- chain order, step 3b: prefix sums along the tour in ONE pass (tiles of 2048 entries handed out by ticket, each
// publishing its aggregate and then its inclusive prefix in one self-validating 64-bit word; a tile looks back over its
// predecessors' words instead of waiting for a separate scan-of-sums launch), keys and histogram written on the way.
__device__ __forceinline__ unsigned long long scan_pack(unsigned flag, int ld, int pre) {
    return ((unsigned long long)flag << 62) | ((unsigned long long)(unsigned)(ld & 0xffff) << 32) | (unsigned)pre;
}
__device__ __forceinline__ int2 scan_unpack(unsigned long long w) {
    return make_int2((int)(short)((w >> 32) & 0xffffu), (int)(unsigned)(w & 0xffffffffull));
}
__global__ void __launch_bounds__(SCAN_NT) chain_scan_kernel(Ws ws, Dims dm) {
    // tree = blockIdx.x: CTAs are handed out x-fastest, so the first tiles of ALL trees start together and the look-back
    // chains of the trees advance side by side (tree = blockIdx.y ran them three waves of ~18 trees after each other)
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, V = dm.V;
    __shared__ int s_blk;
    __shared__ int2 s_w[SCAN_NT / 32], s_excl;
    if (tid == 0) s_blk = atomicAdd(ws.scan_ticket + b, 1);
    __syncthreads();
    const int blk = s_blk;
    const int2 *S = ws.S + (size_t)b * 2 * V;
    const int r0 = blk * SCAN_TILE + tid * SCAN_PER;
    int2 e[SCAN_PER], x[SCAN_PER];
    int2 run = make_int2(0, 0);
#pragma unroll
    for (int k = 0; k < SCAN_PER; k += 2) {
        if (r0 + k + 1 < dm.T) {
            const int4 q = *reinterpret_cast<const int4 *>(S + r0 + k);
            e[k] = make_int2(q.x, q.y);
            e[k + 1] = make_int2(q.z, q.w);
        } else {
            e[k] = (r0 + k < dm.T) ? S[r0 + k] : make_int2(0, -1);
            e[k + 1] = make_int2(0, -1);
        }
    }
#pragma unroll
    for (int k = 0; k < SCAN_PER; ++k) {
        const int2 d = (r0 + k < dm.T) ? chain_delta(e[k]) : make_int2(0, 0);
        run.x += d.x;
        run.y += d.y;
        x[k] = run;                       // inclusive inside the thread
    }
    int2 t = run;                         // inclusive over the warp's threads
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int tx = __shfl_up_sync(0xffffffffu, t.x, o), ty = __shfl_up_sync(0xffffffffu, t.y, o);
        if (lane >= o) { t.x += tx; t.y += ty; }
    }
    if (lane == 31) s_w[wid] = t;
    __syncthreads();
    int2 woff = make_int2(0, 0), agg = make_int2(0, 0);
#pragma unroll
    for (int w2 = 0; w2 < SCAN_NT / 32; ++w2) {
        const int2 v = s_w[w2];
        if (w2 < wid) { woff.x += v.x; woff.y += v.y; }
        agg.x += v.x;
        agg.y += v.y;
    }
    unsigned long long *state = ws.scan_state + (size_t)b * dm.NSCAN;
    if (wid == 0) {   // publish the aggregate, look back for the exclusive prefix, publish the inclusive prefix
        int2 excl = make_int2(0, 0);
        if (blk > 0) {
            if (lane == 0) asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(state + blk), "l"(scan_pack(1u, agg.x, agg.y)));
            int j = blk - 1;
            while (true) {
                const int mine = j - lane;
                unsigned long long wv = 3ull << 62;   // lanes before the first tile: "inclusive, zero"
                if (mine >= 0) {
                    do {
                        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wv) : "l"(state + mine));
                    } while ((wv >> 62) == 0);
                }
                const unsigned fl = (unsigned)(wv >> 62);
                const unsigned done = __ballot_sync(0xffffffffu, fl >= 2u);
                const int stop = done ? __ffs(done) - 1 : 31;          // nearest tile with an inclusive prefix
                int2 v = scan_unpack(wv);
                if (lane > stop || fl == 3u) v = make_int2(0, 0);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
                    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
                }
                excl.x += v.x;
                excl.y += v.y;
                if (done) break;
                j -= 32;
            }
        }
        if (lane == 0) {
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(state + blk), "l"(scan_pack(2u, excl.x + agg.x, excl.y + agg.y)));
            s_excl = excl;
        }
    }
    __syncthreads();
    const int2 excl = s_excl;
    int2 before = make_int2(excl.x + woff.x + t.x - run.x, excl.y + woff.y + t.y - run.y);   // everything ahead of this thread
    int *key = ws.depth + (size_t)b * V;                     // reused: [B,V] keys in preorder
    int *hist = ws.hist + (size_t)b * dm.NTILE * 32;
    if (blk == 0 && tid == 0) {  // the root is never entered
        key[0] = dm.root << 5;
        atomicAdd(hist, 1);
    }
#pragma unroll
    for (int k = 0; k < SCAN_PER; ++k) {
        if (r0 + k < dm.T && e[k].y >= 0) {   // the tour enters vertex e.y >> 1 here
            const int ld = min(before.x + x[k].x, 31), pre = before.y + x[k].y;
            if (pre >= 0 && pre < V) {         // always, for a consistent tree mask
                key[pre] = ((e[k].y >> 1) << 5) | ld;
                atomicAdd(hist + (pre >> 10) * 32 + ld, 1);
            }
        }
    }
}

// one CTA per image, one warp per light depth: exclusive scan of the histogram over the tiles (in place), then the
// group starts = exclusive scan of the group sizes; group_start[g] = V for every g past the last group
__global__ void __launch_bounds__(1024) chain_offsets_kernel(Ws ws, Dims dm, int *__restrict__ group_start) {
    const int b = blockIdx.x, lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    int *hist = ws.hist + (size_t)b * dm.NTILE * 32;
    __shared__ int s_tot[32];
    int carry = 0;
    for (int base = 0; base < dm.NTILE; base += 32) {
        const int t = base + lane;
        const int c = (t < dm.NTILE) ? hist[t * 32 + g] : 0;
        int x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (t < dm.NTILE) hist[t * 32 + g] = carry + x - c;
        carry += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) s_tot[g] = carry;
    __syncthreads();
    if (g == 0) {
        const int c = s_tot[lane];
        int x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        group_start[(size_t)b * 33 + lane] = x - c;
        if (lane == 31) group_start[(size_t)b * 33 + 32] = x;   // = V
    }
}

// position of every vertex: group start + vertices of the same light depth in earlier tiles + in earlier warps of this
// tile + in earlier lanes of this warp
__global__ void __launch_bounds__(1024) chain_place_kernel(Ws ws, Dims dm, const int *__restrict__ group_start,
                                                           int *__restrict__ sorted_index) {
    const int b = blockIdx.y, tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, V = dm.V;
    const int i = tile * 1024 + tid;
    const bool valid = i < V;
    const int key = valid ? ws.depth[(size_t)b * V + i] : 0;
    const int bin = valid ? (key & 31) : 32 + lane;   // idle lanes match nobody
    const unsigned same = __match_any_sync(0xffffffffu, bin);
    const int rank = __popc(same & ((1u << lane) - 1u));
    __shared__ int s_cnt[32][33];
    for (int k = tid; k < 32 * 33; k += 1024) (&s_cnt[0][0])[k] = 0;
    __syncthreads();
    if (valid && rank == 0) s_cnt[wid][bin] = __popc(same);
    __syncthreads();
    {   // exclusive prefix over the warps, one thread per (warp, bin)
        int acc = 0;
        for (int w2 = 0; w2 < wid; ++w2) acc += s_cnt[w2][lane];
        __syncthreads();
        s_cnt[wid][lane] = acc;
    }
    __syncthreads();
    if (!valid) return;
    const int pos = group_start[(size_t)b * 33 + bin] + ws.hist[((size_t)b * dm.NTILE + tile) * 32 + bin] + s_cnt[wid][bin] + rank;
    const int u = key >> 5;
    ws.pos[(size_t)b * V + u] = pos;
    sorted_index[(size_t)b * V + pos] = u;
}

// Embeddings the edge weights of a batch of trees come from: tree t belongs to embedding t / per, image t % per.
constexpr int CHAIN_MAX_EMBEDS = 8;
struct ChainEmbeds {
    const float *ptr[CHAIN_MAX_EMBEDS];
    int D[CHAIN_MAX_EMBEDS];
    float inv_sigma[CHAIN_MAX_EMBEDS];
    int per;              // images per embedding; 0 = no edge weights wanted
};

// parent position; children positions: slot 0 = heavy child (always position + 1), then the light ones by increasing id;
// and, when asked, the filter's edge weight of the position: exp(-|e_u - e_parent|^2 * inv_sigma), the arithmetic of
// edge_weight_fwd_kernel (csrc/refine.cu; tree_filter.py:91-108)
__global__ void __launch_bounds__(NT) chain_finalize_kernel(const unsigned char *__restrict__ mask, Ws ws, Dims dm,
                                                            int *__restrict__ sorted_parent, int *__restrict__ sorted_child,
                                                            ChainEmbeds ce, float *__restrict__ edge_weight) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= (long long)dm.B * dm.V) return;
    const int b = (int)(idx / dm.V), u = (int)(idx - (long long)b * dm.V), V = dm.V, W = dm.W;
    const size_t o = (size_t)b * V;
    const int *pos = ws.pos + o;
    const int i = pos[u], pd = ws.pdir[o + u];
    const int pu = (u == dm.root) ? u : u + dir_delta(pd, W);
    sorted_parent[o + i] = (u == dm.root) ? 0 : pos[pu];
    int sz[4];
    const int hd = heavy_dir(mask[o + u], pd, ws.size + o, u, W, sz);
    int c[4] = {0, 0, 0, 0}, k = 1;
    if (hd >= 0) c[0] = pos[u + dir_delta(hd, W)];
#pragma unroll
    for (int d = 0; d < 4; ++d)
        if (sz[d] > 0 && d != hd) c[k++] = pos[u + dir_delta(d, W)];
    reinterpret_cast<int4 *>(sorted_child)[o + i] = make_int4(c[0], c[1], c[2], c[3]);
    if (ce.per > 0) {
        const int e = b / ce.per, D = ce.D[e];
        const float *x = ce.ptr[e] + (size_t)(b % ce.per) * D * V;
        float d2 = 0.f;
        for (int ch = 0; ch < D; ++ch) {
            const float t = x[(size_t)ch * V + u] - x[(size_t)ch * V + pu];
            d2 = fmaf(t, t, d2);
        }
        edge_weight[o + i] = expf(-d2 * ce.inv_sigma[e]);
    }
}

// Euler tour of every tree ranked, entering / leaving ranks, parent direction and subtree size of every vertex
int run_tour(const unsigned char *tree_mask, const Ws &ws, const Dims &dm, cudaStream_t st, const char *who,
             bool with_enter_leave = true) {
    const int B = dm.B, V = dm.V;
    const unsigned vb = (unsigned)(((long long)B * V + NT - 1) / NT);
    const size_t smem_mask = (size_t)(V + 1) / 2;
    cudaError_t e;
    if (smem_mask <= 200 * 1024) {
        e = cudaFuncSetAttribute(walk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_mask);
        if (e != cudaSuccess) return fedicra_fail(-2, "%s: shared memory: %s", who, cudaGetErrorString(e));
        walk_kernel<true><<<dim3((dm.NS + dm.SPC - 1) / dm.SPC, B), NT, smem_mask, st>>>(tree_mask, ws, dm);
    } else {
        walk_kernel<false><<<dim3((dm.NS + dm.SPC - 1) / dm.SPC, B), NT, 0, st>>>(tree_mask, ws, dm);
    }
    FEDICRA_CHECK_LAUNCH("walk_kernel");
    const size_t smem_rank = ((size_t)dm.NS + (dm.NS + 1) / 2 + 2 * 1024 + 1024) * sizeof(int);   // len, nxt (16 bit), leader tables
    e = cudaFuncSetAttribute(rank_splitters_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rank);
    if (e != cudaSuccess) return fedicra_fail(-2, "%s: shared memory: %s", who, cudaGetErrorString(e));
    rank_splitters_kernel<<<B, 1024, smem_rank, st>>>(ws, dm);
    FEDICRA_CHECK_LAUNCH("rank_splitters_kernel");
    if (with_enter_leave) {
        enter_leave_kernel<<<vb, NT, 0, st>>>(tree_mask, ws, dm);
        FEDICRA_CHECK_LAUNCH("enter_leave_kernel");
    }
    return 0;
}

}  // namespace

extern "C" size_t fedicra_bfs_euler_workspace_bytes(int B, int H, int W) {
    if (B <= 0 || H <= 0 || W <= 0) return 0;
    return ws_layout(make_dims(B, H, W), nullptr, nullptr);
}

extern "C" int fedicra_bfs_euler_rooted(const unsigned char *tree_mask, int B, int H, int W, int root, int *sorted_index,
                                        int *sorted_parent, int *sorted_child, int *level_start, int *n_levels,
                                        void *workspace, size_t workspace_bytes, void *stream);

extern "C" int fedicra_bfs_euler(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index,
                                 int *sorted_parent, int *sorted_child, int *level_start, int *n_levels,
                                 void *workspace, size_t workspace_bytes, void *stream) {
    return fedicra_bfs_euler_rooted(tree_mask, B, H, W, 0, sorted_index, sorted_parent, sorted_child, level_start, n_levels,
                                    workspace, workspace_bytes, stream);
}

extern "C" int fedicra_bfs_euler_rooted(const unsigned char *tree_mask, int B, int H, int W, int root, int *sorted_index,
                                        int *sorted_parent, int *sorted_child, int *level_start, int *n_levels,
                                        void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(tree_mask && sorted_index && sorted_parent && sorted_child && level_start && n_levels && workspace,
                    "bfs_euler: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && H >= 1 && W >= 1 && (long long)H * W <= (1ll << 18),
                    "bfs_euler: bad sizes B=%d H=%d W=%d (at most 262144 pixels per image)", B, H, W);
    FEDICRA_REQUIRE(((uintptr_t)sorted_child & 15) == 0, "bfs_euler: sorted_child must be 16-byte aligned");
    FEDICRA_REQUIRE(root >= 0 && root < H * W, "bfs_euler: root %d outside the %d x %d image", root, H, W);
    const Dims dm = make_dims(B, H, W, root);
    FEDICRA_REQUIRE(workspace_bytes >= ws_layout(dm, nullptr, nullptr), "bfs_euler: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    ws_layout(dm, (char *)workspace, &ws);
    const int V = dm.V;
    // cnt + cursor must start at zero; the arc table is cleared too so that its lines are L2-resident when the walk
    // scatters 4-byte records into them (otherwise every partial-sector write costs a DRAM read-modify-write)
    cudaError_t e = cudaMemsetAsync(ws.cnt, 0, (size_t)((char *)ws.snext - (char *)ws.cnt), st);
    if (e != cudaSuccess) return fedicra_fail(-2, "bfs_euler: memset: %s", cudaGetErrorString(e));
    const unsigned vb = (unsigned)(((long long)B * V + NT - 1) / NT);
    if (int rc = run_tour(tree_mask, ws, dm, st, "bfs_euler")) return rc;
    if (dm.T > 0) {
        increments_kernel<<<vb, NT, 0, st>>>(tree_mask, ws, dm);
        FEDICRA_CHECK_LAUNCH("increments_kernel");
        tour_block_sum_kernel<<<dim3(dm.NBLK, B), 1024, 0, st>>>(ws, dm);
        FEDICRA_CHECK_LAUNCH("tour_block_sum_kernel");
        tour_scan_sums_kernel<<<B, 1024, 0, st>>>(ws, dm);
        FEDICRA_CHECK_LAUNCH("tour_scan_sums_kernel");
    }
    tour_apply_kernel<<<dim3(dm.NBLK > 0 ? dm.NBLK : 1, B), 1024, 0, st>>>(ws, dm);
    FEDICRA_CHECK_LAUNCH("tour_apply_kernel");
    level_scan_kernel<<<B, 1024, 0, st>>>(ws, dm, level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("level_scan_kernel");
    bucket_kernel<<<vb, NT, 0, st>>>(ws, dm, level_start);
    FEDICRA_CHECK_LAUNCH("bucket_kernel");
    const int warps_per_image = 2048;
    level_rank_kernel<<<dim3(warps_per_image / (NT / 32), B), NT, 0, st>>>(ws, dm, level_start, n_levels, sorted_index,
                                                                        warps_per_image);
    FEDICRA_CHECK_LAUNCH("level_rank_kernel");
    finalize_kernel<<<vb, NT, 0, st>>>(tree_mask, ws, dm, sorted_parent, sorted_child);
    FEDICRA_CHECK_LAUNCH("finalize_kernel");
    return 0;
}


namespace {
int chain_order_impl(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index, int *sorted_parent,
                     int *sorted_child, int *group_start, const ChainEmbeds &ce, float *edge_weight, void *workspace,
                     size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(tree_mask && sorted_index && sorted_parent && sorted_child && group_start && workspace,
                    "chain_order: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && H >= 1 && W >= 1 && (long long)H * W <= (1ll << 18),
                    "chain_order: bad sizes B=%d H=%d W=%d (at most 262144 pixels per image)", B, H, W);
    FEDICRA_REQUIRE(((uintptr_t)sorted_child & 15) == 0, "chain_order: sorted_child must be 16-byte aligned");
    const Dims dm = make_dims(B, H, W, 0);
    FEDICRA_REQUIRE(workspace_bytes >= ws_layout(dm, nullptr, nullptr), "chain_order: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    ws_layout(dm, (char *)workspace, &ws);
    const int V = dm.V;
    // the histogram must start at zero; the arc table is cleared with it so that its lines are L2-resident when the
    // walk scatters 4-byte records into them
    cudaError_t e = cudaMemsetAsync(ws.hist, 0, (size_t)((char *)ws.snext - (char *)ws.hist), st);
    if (e != cudaSuccess) return fedicra_fail(-2, "chain_order: memset: %s", cudaGetErrorString(e));
    const unsigned vb = (unsigned)(((long long)B * V + NT - 1) / NT);
    if (int rc = run_tour(tree_mask, ws, dm, st, "chain_order", false)) return rc;
    chain_enter_kernel<<<vb, NT, 0, st>>>(tree_mask, ws, dm);
    FEDICRA_CHECK_LAUNCH("chain_enter_kernel");
    chain_scan_kernel<<<dim3(B, dm.NSCAN), SCAN_NT, 0, st>>>(ws, dm);
    FEDICRA_CHECK_LAUNCH("chain_scan_kernel");
    chain_offsets_kernel<<<B, 1024, 0, st>>>(ws, dm, group_start);
    FEDICRA_CHECK_LAUNCH("chain_offsets_kernel");
    chain_place_kernel<<<dim3(dm.NTILE, B), 1024, 0, st>>>(ws, dm, group_start, sorted_index);
    FEDICRA_CHECK_LAUNCH("chain_place_kernel");
    chain_finalize_kernel<<<vb, NT, 0, st>>>(tree_mask, ws, dm, sorted_parent, sorted_child, ce, edge_weight);
    FEDICRA_CHECK_LAUNCH("chain_finalize_kernel");
    return 0;
}
}  // namespace

extern "C" int fedicra_chain_order(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index,
                                   int *sorted_parent, int *sorted_child, int *group_start, void *workspace,
                                   size_t workspace_bytes, void *stream) {
    ChainEmbeds ce = {};
    return chain_order_impl(tree_mask, B, H, W, sorted_index, sorted_parent, sorted_child, group_start, ce, nullptr,
                            workspace, workspace_bytes, stream);
}

extern "C" int fedicra_chain_order_weights(const unsigned char *tree_mask, const float *const *h_embeds, const int *h_D,
                                           const float *h_inv_sigma, int K, int B, int H, int W, int *sorted_index,
                                           int *sorted_parent, int *sorted_child, int *group_start, float *edge_weight,
                                           void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(h_embeds && h_D && h_inv_sigma && edge_weight, "chain_order_weights: null pointer");
    FEDICRA_REQUIRE(K >= 1 && K <= CHAIN_MAX_EMBEDS && B >= 1, "chain_order_weights: %d embeddings (1..%d)", K, CHAIN_MAX_EMBEDS);
    ChainEmbeds ce = {};
    for (int k = 0; k < K; ++k) {
        FEDICRA_REQUIRE(h_embeds[k] && h_D[k] >= 1, "chain_order_weights: embedding %d is null or has no channels", k);
        ce.ptr[k] = h_embeds[k];
        ce.D[k] = h_D[k];
        ce.inv_sigma[k] = h_inv_sigma[k];
    }
    ce.per = B;
    return chain_order_impl(tree_mask, K * B, H, W, sorted_index, sorted_parent, sorted_child, group_start, ce,
                            edge_weight, workspace, workspace_bytes, stream);
}
