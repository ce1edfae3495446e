// mst.cu -- batched minimum spanning tree on the GPU (sm_100a), Boruvka rounds with 64-bit (weight, edge id) keys.
//
// Replaces mst_forward of the reference (LTF = /root/reference/code/utils/TreeEnergyLoss/kernels/lib_tree_filter):
//   LTF/src/mst/mst.cu:86-117  copies index+weights to the host, runs one std::thread per image of
//   LTF/src/mst/boruvka.cpp:20-112 (sequential Boruvka, union-find) and copies the tree back.
// The reference keeps, per component, the FIRST edge (lowest index) of strictly smallest weight (boruvka.cpp:74-80), so
// its edge set is the unique MST under the total order (weight, edge index).  Here every round each component takes the
// atomicMin of the packed key (orderable fp32 bits << 32 | edge index) over its outgoing edges -- the same total order,
// hence the same edge set, bit for bit, whatever the tie structure.
//
// Three paths, same edge sets:
//   explicit edge list (fedicra_mst_forward): find / hook / compress kernels per round;
//   implicit grid, default: mst_tile_kernel (rounds inside 64x64 tiles out of shared memory, hands over the CONTRACTED
//     graph) + mst_contract_kernel (one CTA per tree finishes on that graph out of shared memory);
//   implicit grid, FEDICRA_MST_FINISH=cluster (round 1): mst_tile_kernel + mst_grid_cluster_kernel (thread-block cluster
//     per tree on vertex lists in L2), or the cluster kernel alone (FEDICRA_MST_TILES=0).
//
// Also fuses MinimumSpanningTree._build_matrix_index / _build_feature_weight (LTF/modules/tree_filter.py:15-34): on the
// implicit 4-connected grid the edge list is never materialised and the weights come straight from the embedding.
#include <cooperative_groups.h>
#include <stdlib.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

typedef unsigned long long u64;
constexpr u64 KEY_INF = ~0ull;

struct Rec;
struct Ws {
    int *comp;       // [B,V] component root of every vertex (fully compressed between rounds)
    int *parent;     // [B,V] hook target of a root (self when not hooked)
    u64 *best;       // [B,V] cheapest outgoing edge key of a root
    float *wgt;      // [B,E] edge weights (grid entry point only)
    unsigned char *intree;  // [B,E]
    int *ncomp;      // [B]
    int *done;       // [B]
    int *act;        // [B,2,V] active-vertex lists (vertices that still see another component), ping-pong
    int *roots;      // [B,2,V] component-root lists, ping-pong
    int *cnt;        // [B,4]   list lengths: act[0], act[1], roots[0], roots[1]
                     //         (contraction: [0] cross-edge records, [2] components left by the tile rounds)
    struct Rec *rec; // [B,E]   contraction: cross edges left by the tile rounds (edge id, the two components)
    unsigned char *tmask;   // [B,V] adjacency mask when the caller wants only the edge list
};

// A cross edge of the contracted graph: `a`, `b` are dense component numbers (per tree, handed out tile by tile);
// b = UNRESOLVED when the far end lies in another tile, whose numbering the emitting tile cannot know.
struct __align__(16) Rec {
    unsigned eid;
    int a, b;
    unsigned ow;   // orderable bits of the edge weight
};
constexpr int UNRESOLVED = -1;

size_t ws_layout(int B, int V, int E, char *base, Ws *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char *p = base ? base + off : nullptr;
        off += align_up(bytes, 256);
        return p;
    };
    Ws t;
    t.comp = (int *)take((size_t)B * V * sizeof(int));
    t.parent = (int *)take((size_t)B * V * sizeof(int));
    t.best = (u64 *)take((size_t)B * V * sizeof(u64));
    t.wgt = (float *)take((size_t)B * E * sizeof(float));
    t.intree = (unsigned char *)take((size_t)B * E);
    t.ncomp = (int *)take((size_t)B * sizeof(int));
    t.done = (int *)take((size_t)B * sizeof(int));
    t.act = (int *)take((size_t)B * 2 * V * sizeof(int));
    t.roots = (int *)take((size_t)B * 2 * V * sizeof(int));
    t.cnt = (int *)take((size_t)B * 4 * sizeof(int));
    t.rec = (Rec *)take((size_t)B * (E > 0 ? E : 1) * sizeof(Rec));
    t.tmask = (unsigned char *)take((size_t)B * V);
    if (w) *w = t;
    return off;
}

__device__ __forceinline__ unsigned orderable(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Grid edge numbering of the reference (tree_filter.py:20-23): vertical edges row-major first, then horizontal ones.
__device__ __forceinline__ void grid_endpoints(int e, int H, int W, int &u, int &v) {
    const int nvert = (H - 1) * W;
    if (e < nvert) {
        u = e;
        v = e + W;
    } else {
        const int r = e - nvert, h = r / (W - 1), w = r - h * (W - 1);
        u = h * W + w;
        v = u + 1;
    }
}

__global__ void mst_init_kernel(Ws ws, int V, int E) {
    const int b = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < V) {
        ws.comp[(size_t)b * V + i] = i;
        ws.parent[(size_t)b * V + i] = i;
        ws.best[(size_t)b * V + i] = KEY_INF;
    }
    if (i < E) ws.intree[(size_t)b * E + i] = 0;
    if (i == 0) {
        ws.ncomp[b] = V;
        ws.done[b] = (V <= 1);
    }
}

template <bool GRID>
__global__ void mst_find_kernel(Ws ws, const int *__restrict__ edge_index, const float *__restrict__ edge_weight, int V,
                                int E, int H, int W) {
    const int b = blockIdx.y;
    if (ws.ncomp[b] <= 1) return;  // this image's tree is complete
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    int u, v;
    if (GRID) {
        grid_endpoints(e, H, W, u, v);
    } else {
        const int2 uv = reinterpret_cast<const int2 *>(edge_index)[(size_t)b * E + e];
        u = uv.x;
        v = uv.y;
    }
    const int *comp = ws.comp + (size_t)b * V;
    const int cu = comp[u], cv = comp[v];
    if (cu == cv) return;
    const u64 key = ((u64)orderable(edge_weight[(size_t)b * E + e]) << 32) | (unsigned)e;
    u64 *best = ws.best + (size_t)b * V;
    if (key < best[cu]) atomicMin(best + cu, key);
    if (key < best[cv]) atomicMin(best + cv, key);
}

template <bool GRID>
__global__ void mst_hook_kernel(Ws ws, const int *__restrict__ edge_index, int V, int E, int H, int W) {
    const int b = blockIdx.y;
    if (ws.ncomp[b] <= 1) return;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= V) return;
    const int *comp = ws.comp + (size_t)b * V;
    if (comp[c] != c) return;  // not a component root
    const u64 *best = ws.best + (size_t)b * V;
    const u64 key = best[c];
    if (key == KEY_INF) return;
    const int e = (int)(unsigned)(key & 0xffffffffull);
    int u, v;
    if (GRID) {
        grid_endpoints(e, H, W, u, v);
    } else {
        const int2 uv = reinterpret_cast<const int2 *>(edge_index)[(size_t)b * E + e];
        u = uv.x;
        v = uv.y;
    }
    const int cu = comp[u], cv = comp[v];
    const int other = (cu == c) ? cv : cu;
    // two components that picked the same edge: the smaller root stays a root, the larger one hooks and records the edge
    if (best[other] == key && c < other) return;
    ws.parent[(size_t)b * V + c] = other;
    ws.intree[(size_t)b * E + e] = 1;
    atomicSub(ws.ncomp + b, 1);
}

__global__ void mst_compress_kernel(Ws ws, int V) {
    const int b = blockIdx.y;
    if (ws.ncomp[b] <= 1 && ws.done[b]) return;  // one last compression after the final hook, then nothing
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    const int *parent = ws.parent + (size_t)b * V;
    int r = ws.comp[(size_t)b * V + i];
    int p = parent[r];
    while (p != r) {
        r = p;
        p = parent[r];
    }
    ws.comp[(size_t)b * V + i] = r;
    ws.best[(size_t)b * V + i] = KEY_INF;
    if (i == 0 && ws.ncomp[b] <= 1) ws.done[b] = 1;
}

// One CTA per image: stable compaction of the selected edges, in increasing edge index.  Unused slots (disconnected
// input) are filled with -1.
template <bool GRID>
__global__ void __launch_bounds__(1024)
mst_compact_kernel(Ws ws, const int *__restrict__ edge_index, int *__restrict__ edge_out, int V, int E, int H, int W,
                   const unsigned char *__restrict__ mask_all) {   // mask_all: take the flags from the adjacency masks (grid)
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned char *flag = ws.intree + (size_t)b * E;
    int2 *out = reinterpret_cast<int2 *>(edge_out) + (size_t)b * (V - 1);
    __shared__ int s_warp[32];
    __shared__ int s_base, s_total;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int e0 = 0; e0 < E; e0 += 1024) {
        const int e = e0 + tid;
        int f = 0;
        if (e < E) {
            if (GRID && mask_all != nullptr) {
                const unsigned char *mk = mask_all + (size_t)b * V;
                const int nvert = (H - 1) * W;
                if (e < nvert) {
                    f = mk[e] & 8;
                } else {
                    const int r = e - nvert, h = r / (W - 1);
                    f = mk[h * W + (r - h * (W - 1))] & 4;
                }
            } else {
                f = flag[e];
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, f);
        if (lane == 0) s_warp[wid] = __popc(m);
        __syncthreads();
        if (wid == 0) {
            const int v = s_warp[lane];
            int x = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += t;
            }
            s_warp[lane] = x - v;  // exclusive prefix over warps
            if (lane == 31) s_total = x;
        }
        __syncthreads();
        const int pos = s_base + s_warp[wid] + __popc(m & ((1u << lane) - 1u));
        if (f && pos < V - 1) {
            int u, v;
            if (GRID) {
                grid_endpoints(e, H, W, u, v);
            } else {
                const int2 uv = reinterpret_cast<const int2 *>(edge_index)[(size_t)b * E + e];
                u = uv.x;
                v = uv.y;
            }
            out[pos] = make_int2(u, v);
        }
        __syncthreads();
        if (tid == 0) s_base += s_total;
        __syncthreads();
    }
    for (int i = s_base + tid; i < V - 1; i += 1024) out[i] = make_int2(-1, -1);
}

// ---------------------------------------------------------------------------------------------------------------
// Persistent variant for the implicit grid: ONE launch for the whole batch.  A thread-block cluster of CL CTAs owns an
// image and runs every Boruvka round itself, separated by cluster barriers instead of kernel boundaries.  Same keys, same
// hooks, hence the same edge set as the multi-launch path (and as the reference).
//
// The rounds work on two shrinking lists instead of on all V vertices:
//   active vertices -- those that still have a neighbour in another component.  A vertex whose four neighbours are all in
//                      its own component can never see an outgoing edge again (components only grow): it is dropped and
//                      its label is set to DEAD, which its neighbours read as "same component as me" (they were, and
//                      stay, in its component).  After a few rounds only the component boundaries are left.
//   roots           -- the current components (one vertex id each); a third of them survive a round.
// find  (active list): vertex-centric, each vertex scans its <= 4 grid edges and issues at most one 64-bit atomicMin on
//                      its component; survivors are appended to the next active list (warp-aggregated atomicAdd).
// hook  (root list):   every root follows its edge; of two roots that picked the same edge the larger one hooks.
// chase (next active list): labels follow the hooks to the new roots; the cheapest-edge slots of the new roots are reset.
// List order depends on atomics, the result does not: min is order-free and hooks depend on the minima only.
constexpr int PNT = 512;
constexpr int DEAD = -1;
constexpr int MAX_EMBEDS = 8;

// The trees of one launch may come from several embeddings of one batch (the image and the decoder heads of the
// multiscale loss): tree t belongs to embedding t / per, image t % per.  No stacked copy is ever made.
struct EmbedSet {
    const float *ptr[MAX_EMBEDS];
    int D[MAX_EMBEDS];
    int per;   // images per embedding
};

__global__ void __launch_bounds__(PNT)
mst_grid_cluster_kernel(Ws ws, EmbedSet es, unsigned char *__restrict__ mask, int H, int W, int max_rounds, int prepared) {
    cg::cluster_group cluster = cg::this_cluster();
    const int CL = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    const int b = blockIdx.x / CL, tid = threadIdx.x, lane = tid & 31;
    const int V = H * W, E = 2 * H * W - H - W, nvert = (H - 1) * W;
    int *comp = ws.comp + (size_t)b * V, *parent = ws.parent + (size_t)b * V;
    u64 *best = ws.best + (size_t)b * V;
    float *wgt = ws.wgt + (size_t)b * E;
    unsigned char *intree = ws.intree + (size_t)b * E;
    int *act = ws.act + (size_t)b * 2 * V, *roots = ws.roots + (size_t)b * 2 * V;
    int *cnt = ws.cnt + (size_t)b * 4;
    const int stride = CL * PNT, first = rank * PNT + tid;

    // ---- init: weights (tree_filter.py:27-34, no fused multiply-add), union state.  `prepared`: mst_tile_kernel has
    //      already run the rounds that stay inside 64x64 tiles and left weights, labels, flags and the two lists
    if (!prepared) {
        const int D = es.D[b / es.per];
        const float *x = es.ptr[b / es.per] + (size_t)(b % es.per) * D * V;
        for (int e = first; e < E; e += stride) {
            int u, v;
            grid_endpoints(e, H, W, u, v);
            float d = 0.f;
            for (int c = 0; c < D; ++c) {
                const float t = __fsub_rn(x[(size_t)c * V + u], x[(size_t)c * V + v]);
                d = __fadd_rn(d, __fmul_rn(t, t));
            }
            wgt[e] = __fadd_rn(d, 1.f);
            intree[e] = 0;
        }
        for (int v = first; v < V; v += stride) {
            comp[v] = v;
            parent[v] = v;
            best[v] = KEY_INF;
        }
        if (rank == 0 && tid < 4) cnt[tid] = 0;
        cluster.sync();
    }

    // Every loop below handles IL list entries per iteration with the loads of all of them issued before the first use:
    // the phases are chains of dependent L2 round trips, and one cluster only has CL*512 threads to hide them with.
    constexpr int IL = 4;
    int cur = 0;            // lists of this round: act + cur*V, roots + cur*V (round 0: the identity, not stored)
    int nact = prepared ? cnt[0] : V, nroots = prepared ? cnt[2] : V;
    for (int round = prepared ? 1 : 0; round < max_rounds + 1 && nroots > 1; ++round) {
        const int *alist = round ? act + (size_t)cur * V : nullptr, *rlist = round ? roots + (size_t)cur * V : nullptr;
        int *anext = act + (size_t)(cur ^ 1) * V, *rnext = roots + (size_t)(cur ^ 1) * V;
        int *cnt_anext = cnt + (cur ^ 1), *cnt_rnext = cnt + 2 + (cur ^ 1);
        // ---- find: cheapest outgoing edge per component
        for (int w0 = first - lane; w0 < nact; w0 += stride * IL) {   // warp-uniform trip count (ballots inside)
            const int i0 = w0 + lane;
            int vv[IL], cv[IL], nb_[IL][4], ee[IL][4];
#pragma unroll
            for (int k = 0; k < IL; ++k) {
                const int i = i0 + k * stride;
                vv[k] = (i < nact) ? (alist ? alist[i] : i) : -1;
            }
#pragma unroll
            for (int k = 0; k < IL; ++k) cv[k] = (vv[k] >= 0) ? comp[vv[k]] : DEAD;
#pragma unroll
            for (int k = 0; k < IL; ++k) {
                const int v = vv[k], h = v / W, w = v - h * W;
                const bool ok = v >= 0;
                ee[k][0] = (ok && h > 0) ? v - W : -1;
                ee[k][1] = (ok && w > 0) ? nvert + h * (W - 1) + w - 1 : -1;
                ee[k][2] = (ok && w < W - 1) ? nvert + h * (W - 1) + w : -1;
                ee[k][3] = (ok && h < H - 1) ? v : -1;
                nb_[k][0] = ee[k][0] >= 0 ? comp[v - W] : DEAD;
                nb_[k][1] = ee[k][1] >= 0 ? comp[v - 1] : DEAD;
                nb_[k][2] = ee[k][2] >= 0 ? comp[v + 1] : DEAD;
                nb_[k][3] = ee[k][3] >= 0 ? comp[v + W] : DEAD;
            }
            float wv[IL][4];
#pragma unroll
            for (int k = 0; k < IL; ++k)
#pragma unroll
                for (int d = 0; d < 4; ++d) {
                    // a DEAD neighbour is in this vertex's component (see above); so is one with the same label
                    const bool out = nb_[k][d] != DEAD && nb_[k][d] != cv[k];
                    wv[k][d] = out ? wgt[ee[k][d]] : 0.f;
                    if (!out) ee[k][d] = -1;
                }
#pragma unroll
            for (int k = 0; k < IL; ++k) {
                u64 key = KEY_INF;
#pragma unroll
                for (int d = 0; d < 4; ++d)
                    if (ee[k][d] >= 0) key = min(key, ((u64)orderable(wv[k][d]) << 32) | (unsigned)ee[k][d]);
                const bool alive = key != KEY_INF;
                if (alive) atomicMin(best + cv[k], key);
                else if (vv[k] >= 0) comp[vv[k]] = DEAD;
                // survivors -> next active list
                const unsigned m = __ballot_sync(0xffffffffu, alive);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    int base = 0;
                    if (lane == leader) base = atomicAdd(cnt_anext, __popc(m));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (alive) anext[base + __popc(m & ((1u << lane) - 1u))] = vv[k];
                }
            }
        }
        cluster.sync();
        // ---- hook: every root follows its edge; of two roots that picked the same edge the larger one hooks
        for (int w0 = first - lane; w0 < nroots; w0 += stride * IL) {
            const int i0 = w0 + lane;
            int cc[IL];
            u64 key[IL];
#pragma unroll
            for (int k = 0; k < IL; ++k) {
                const int i = i0 + k * stride;
                cc[k] = (i < nroots) ? (rlist ? rlist[i] : i) : -1;
            }
#pragma unroll
            for (int k = 0; k < IL; ++k) key[k] = (cc[k] >= 0) ? __ldcg(best + cc[k]) : KEY_INF;   // written by atomics: read at L2
            int other[IL], eid[IL];
#pragma unroll
            for (int k = 0; k < IL; ++k) {
                other[k] = -1;
                eid[k] = 0;
                if (key[k] != KEY_INF) {
                    const int e = (int)(unsigned)(key[k] & 0xffffffffull);
                    int u, v;
                    grid_endpoints(e, H, W, u, v);
                    const int cu = comp[u], cvv = comp[v];   // both endpoints of a cheapest edge are active: labels are fresh
                    other[k] = (cu == cc[k]) ? cvv : cu;
                    eid[k] = e;
                }
            }
#pragma unroll
            for (int k = 0; k < IL; ++k) {
                bool stays = cc[k] >= 0;   // a root without outgoing edge (finished tree) stays too
                if (other[k] >= 0 && !(__ldcg(best + other[k]) == key[k] && cc[k] < other[k])) {
                    parent[cc[k]] = other[k];
                    intree[eid[k]] = 1;
                    stays = false;
                }
                const unsigned m = __ballot_sync(0xffffffffu, stays);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    int base = 0;
                    if (lane == leader) base = atomicAdd(cnt_rnext, __popc(m));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (stays) rnext[base + __popc(m & ((1u << lane) - 1u))] = cc[k];
                }
            }
        }
        cluster.sync();
        // ---- chase: the labels of the surviving vertices follow the hooks to their new roots
        const int nact_next = *reinterpret_cast<volatile int *>(cnt_anext);
        const int nroots_next = *reinterpret_cast<volatile int *>(cnt_rnext);
        for (int i0 = first; i0 < nact_next; i0 += stride * IL) {
            int vv[IL], r[IL], p[IL];
#pragma unroll
            for (int k = 0; k < IL; ++k) vv[k] = (i0 + k * stride < nact_next) ? anext[i0 + k * stride] : -1;
#pragma unroll
            for (int k = 0; k < IL; ++k) r[k] = (vv[k] >= 0) ? comp[vv[k]] : 0;
#pragma unroll
            for (int k = 0; k < IL; ++k) p[k] = parent[r[k]];
            bool moving = true;
            while (moving) {
                moving = false;
#pragma unroll
                for (int k = 0; k < IL; ++k) {
                    if (p[k] != r[k]) {
                        r[k] = p[k];
                        moving = true;
                    }
                }
                if (moving) {
#pragma unroll
                    for (int k = 0; k < IL; ++k) p[k] = parent[r[k]];
                }
            }
#pragma unroll
            for (int k = 0; k < IL; ++k)
                if (vv[k] >= 0) comp[vv[k]] = r[k];
        }
        for (int i = first; i < nroots_next; i += stride) best[rnext[i]] = KEY_INF;
        if (rank == 0 && tid == 0) {   // the lists just consumed become the next round's output lists
            cnt[cur] = 0;
            cnt[2 + cur] = 0;
        }
        cluster.sync();
        cur ^= 1;
        nact = nact_next;
        nroots = nroots_next;
    }

    // ---- 4-bit adjacency mask from the selected-edge flags
    if (mask != nullptr) {
        unsigned char *m = mask + (size_t)b * V;
        for (int v = first; v < V; v += stride) {
            const int h = v / W, w = v - h * W;
            unsigned bits = 0;
            if (h > 0 && intree[v - W]) bits |= 1u;
            if (w > 0 && intree[nvert + h * (W - 1) + w - 1]) bits |= 2u;
            if (w < W - 1 && intree[nvert + h * (W - 1) + w]) bits |= 4u;
            if (h < H - 1 && intree[v]) bits |= 8u;
            m[v] = (unsigned char)bits;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Tile-local Boruvka rounds.  One CTA owns a 64x64 tile of an image and runs, entirely out of shared memory, every
// round that can be decided inside the tile: a component takes the cheapest of ALL its outgoing edges -- those that
// leave the tile included, their weights come from the embedding's halo -- and hooks when that edge lies inside the
// tile; when it leaves the tile the component waits for the global rounds (it may still be hooked INTO by others, after
// which its cheapest edge is looked at again).  A component's cheapest outgoing edge belongs to the MST whatever the
// others do (cut property, keys are distinct), so the edges chosen here are exactly those the untiled rounds would
// choose: same edge set, bit for bit.  The rounds stop after `max_rounds` or when a whole round hooks nothing.  What is
// left is the tile boundaries: a few hundred components per tile instead of 4 096 vertices.
// Output, contract = 1 (default): the adjacency-mask bits of the edges chosen here, dense component numbers on the tile's
// first row and column, and one record per cross edge (see the hand-over below).  contract = 0 (cluster finish): edge
// weights and tree flags of the tile's own edges (down / right of each of its vertices), fresh labels (DEAD for vertices
// whose neighbours all share their component), identity hooks, and the root / active lists.
// FEDICRA_PROF=1: clock64 stamps of tree 0's contraction (start, records, components, loaded, then per round: found,
// located, hooked, chased, compacted, records left) -- read back with fedicra_debug_mst_profile
constexpr int MPROF = 128;   // [0,64) contraction kernel, [64,128) tile kernel (tile 0 of tree 0: start, weights; per round the
                             // phases as (marker, time) pairs -- -10 found, -11 min weight, -12 min code, -13 hooked -- then the
                             // round's end; -2, ranks, tables, emitted, -1)
__device__ long long g_mst_prof[MPROF];
#define MST_STAMP(val)                                                                    \
    do {                                                                                  \
        if (prof && blockIdx.x == 0 && threadIdx.x == 0 && pslot < MPROF) g_mst_prof[pslot++] = (val); \
    } while (0)

constexpr int TS = 64, TV = TS * TS, TNT = 512;
constexpr size_t TILE_SMEM = (size_t)TV * 8 + 2 * (size_t)(TS + 1) * TS * 4 + 2 * (size_t)TV * 2 + (size_t)TV + (size_t)TV * 4;

__global__ void __launch_bounds__(TNT, 2) mst_tile_kernel(Ws ws, EmbedSet es, int H, int W, int tiles_x, int max_rounds,
                                                          unsigned char *__restrict__ mask_out, int contract) {
    // cheapest outgoing edge per component as two 32-bit words (orderable weight, edge code): shared memory has native
    // 32-bit atomicMin.  min weight first, then min code among the edges of that weight.  The code of an edge that
    // touches the tile is type << 14 | row << 7 | column (see s_wv / s_wh); among such edges it orders like the global
    // edge id (vertical edges first, then row-major), so (weight, code) picks the edge (weight, edge id) would.
    extern __shared__ unsigned s_bw[];                                // [TV] orderable weight
    unsigned *s_be = s_bw + TV;                                       // [TV] edge code
    float *s_wv = reinterpret_cast<float *>(s_be + TV);              // [TS+1][TS] edge (row r-1, row r) of column c; 0 = none
    float *s_wh = s_wv + (TS + 1) * TS;                               // [TS][TS+1] edge (col c-1, col c) of row r;    0 = none
    unsigned short *s_lab = reinterpret_cast<unsigned short *>(s_wh + TS * (TS + 1));   // [TV]
    unsigned short *s_par = s_lab + TV;                               // [TV]
    unsigned *s_sel = reinterpret_cast<unsigned *>(s_par + TV);       // [TV] bytes: adjacency mask of the edges chosen here
    unsigned *s_te = s_sel + TV / 4;                                  // [TV] contraction hand-over: third table column
    float *s_x = reinterpret_cast<float *>(s_bw);                     // [(TS+2)][(TS+2)] one channel of the tile + halo (init only)
    constexpr int XS = TS + 2;
    const int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
    const int prof = (contract >> 1) & (blockIdx.y == 0);
    contract &= 1;
    int pslot = 64;
    MST_STAMP(clock64());
    const int ty = blockIdx.x / tiles_x, tx = blockIdx.x - ty * tiles_x, h0 = ty * TS, w0 = tx * TS;
    const int th = min(TS, H - h0), tw = min(TS, W - w0);
    const int V = H * W, E = 2 * H * W - H - W, nvert = (H - 1) * W;
    int *comp = ws.comp + (size_t)b * V, *parent = ws.parent + (size_t)b * V;
    u64 *best = ws.best + (size_t)b * V;
    float *wgt = ws.wgt + (size_t)b * E;
    unsigned char *intree = ws.intree + (size_t)b * E;
    int *act = ws.act + (size_t)b * 2 * V, *roots = ws.roots + (size_t)b * 2 * V;
    int *cnt = ws.cnt + (size_t)b * 4;
    const int D = es.D[b / es.per];
    const float *x = es.ptr[b / es.per] + (size_t)(b % es.per) * D * V;
    auto v_valid = [&](int r, int c) { return c < tw && r <= th && h0 + r >= 1 && h0 + r < H; };      // s_wv[r][c]
    auto h_valid = [&](int r, int c) { return r < th && c <= tw && w0 + c >= 1 && w0 + c < W; };      // s_wh[r][c]
    // ---- weights of every edge that touches the tile: sum_c (delta)^2 + 1, channels in order, no fused multiply-add
    //      (tree_filter.py:27-34); the embedding is staged one channel at a time (tile + one pixel of halo)
    //      The sums live in registers (each thread owns the same nine vertical and nine horizontal edges in every
    //      channel) and the next channel's pixels are already on their way while the current one is being differenced.
    constexpr int NX = (XS * XS + TNT - 1) / TNT, NE = ((TS + 1) * TS + TNT - 1) / TNT;
    float av[NE], ah[NE], xv[NX];
#pragma unroll
    for (int k = 0; k < NE; ++k) av[k] = ah[k] = 0.f;
    auto fetch = [&](int ch) {
#pragma unroll
        for (int k = 0; k < NX; ++k) {
            const int i = tid + k * TNT, r = i / XS, c = i - r * XS, hh = h0 + r - 1, ww = w0 + c - 1;
            xv[k] = (i < XS * XS && hh >= 0 && hh < H && ww >= 0 && ww < W) ? x[(size_t)ch * V + (size_t)hh * W + ww] : 0.f;
        }
    };
    fetch(0);
    for (int ch = 0; ch < D; ++ch) {
        __syncthreads();
#pragma unroll
        for (int k = 0; k < NX; ++k)
            if (tid + k * TNT < XS * XS) s_x[tid + k * TNT] = xv[k];
        __syncthreads();
        if (ch + 1 < D) fetch(ch + 1);
#pragma unroll
        for (int k = 0; k < NE; ++k) {
            const int i = tid + k * TNT;
            if (i < (TS + 1) * TS) {
                const int r = i / TS, c = i - r * TS;   // pixels (r-1, c) and (r, c) of the tile = s_x rows r, r+1, column c+1
                const float t = __fsub_rn(s_x[r * XS + c + 1], s_x[(r + 1) * XS + c + 1]);
                av[k] = __fadd_rn(av[k], __fmul_rn(t, t));
                const int r2 = i / (TS + 1), c2 = i - r2 * (TS + 1);   // pixels (r, c-1) and (r, c) = s_x row r+1, columns c, c+1
                const float t2 = __fsub_rn(s_x[(r2 + 1) * XS + c2], s_x[(r2 + 1) * XS + c2 + 1]);
                ah[k] = __fadd_rn(ah[k], __fmul_rn(t2, t2));
            }
        }
    }
    __syncthreads();   // s_x aliases the cheapest-edge slots that are initialised below
    // finish (+1, or 0 for "no such edge") and publish the edges the tile owns (below / right of its vertices)
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int i = tid + k * TNT;
        if (i >= (TS + 1) * TS) continue;
        {
            const int r = i / TS, c = i - r * TS;
            float wt = 0.f;
            if (v_valid(r, c)) {
                wt = __fadd_rn(av[k], 1.f);
                if (r >= 1 && !contract) {
                    const int e = (h0 + r - 1) * W + w0 + c;
                    wgt[e] = wt;
                    intree[e] = 0;
                }
            }
            s_wv[i] = wt;
        }
        {
            const int r = i / (TS + 1), c = i - r * (TS + 1);
            float wt = 0.f;
            if (h_valid(r, c)) {
                wt = __fadd_rn(ah[k], 1.f);
                if (c >= 1 && !contract) {
                    const int e = nvert + (h0 + r) * (W - 1) + w0 + c - 1;
                    wgt[e] = wt;
                    intree[e] = 0;
                }
            }
            s_wh[i] = wt;
        }
    }
    for (int l = tid; l < TV; l += TNT) {
        s_lab[l] = (unsigned short)l;
        s_par[l] = (unsigned short)l;
        s_bw[l] = 0xffffffffu;
        s_be[l] = 0xffffffffu;
    }
    for (int l = tid; l < TV / 4; l += TNT) s_sel[l] = 0u;
    __syncthreads();
    MST_STAMP(clock64());
    // ---- rounds.  Thread t owns vertices (row (t >> 6) + 8q, column t & 63), q = 0..7.
    constexpr int PER = TV / TNT;
    const int c = tid & (TS - 1), rbase = tid >> 6;
    unsigned interior = 0;   // bit q: all four neighbours share the vertex's component -- for good (components only grow)
#pragma unroll
    for (int q = 0; q < PER; ++q)
        if (rbase + q * (TNT / TS) >= th || c >= tw) interior |= 1u << q;   // outside the image: never looked at
    for (int round = 0; round < max_rounds; ++round) {
        // find: cheapest outgoing edge per component (outgoing = other label, or the other side of the tile boundary)
        unsigned kw[PER], ke[PER];
        unsigned short kc[PER];
        if (round == 0) {
            // every vertex is a component of its own: its cheapest edge is the cheapest of its (at most four) edges -- no
            // labels to compare, no atomics; the slot of a vertex simply takes the code
#pragma unroll
            for (int q = 0; q < PER; ++q) {
                if (interior & (1u << q)) continue;
                const int r = rbase + q * (TNT / TS), l = r * TS + c;
                const float wu = s_wv[l], wd = s_wv[l + TS], wl = s_wh[l + r], wr = s_wh[l + r + 1];
                unsigned bw = 0xffffffffu, be = 0xffffffffu;
                auto take = [&](float wt, unsigned code) {
                    const unsigned o = orderable(wt);
                    if (o < bw) { bw = o; be = code; }
                };
                if (wu > 0.f) take(wu, (unsigned)(r << 7 | c));
                if (wd > 0.f) take(wd, (unsigned)((r + 1) << 7 | c));
                if (wl > 0.f) take(wl, (unsigned)(1 << 14 | r << 7 | c));
                if (wr > 0.f) take(wr, (unsigned)(1 << 14 | r << 7 | (c + 1)));
                s_be[l] = be;
                if (be == 0xffffffffu) interior |= 1u << q;
            }
            __syncthreads();
        } else {
    #pragma unroll
            for (int q = 0; q < PER; ++q) {
                kw[q] = ke[q] = 0xffffffffu;
                kc[q] = 0;
                if (interior & (1u << q)) continue;
                const int r = rbase + q * (TNT / TS), l = r * TS + c;
                const unsigned short cl = s_lab[l];
                kc[q] = cl;
                const float wu = s_wv[l], wd = s_wv[l + TS], wl = s_wh[l + r], wr = s_wh[l + r + 1];
                // the neighbours' labels are read unconditionally (own label where the neighbour is outside the tile), so
                // that all nine loads of a vertex are in flight together instead of behind short-circuit branches
                const bool in_u = r > 0, in_d = r + 1 < th, in_l = c > 0, in_r = c + 1 < tw;
                const unsigned short lu = s_lab[in_u ? l - TS : l], ld = s_lab[in_d ? l + TS : l],
                                     ll = s_lab[in_l ? l - 1 : l], lr = s_lab[in_r ? l + 1 : l];
                // candidates in increasing code order (up, down are vertical: codes r<<7|c and (r+1)<<7|c; then left, right)
                unsigned bw = 0xffffffffu, be = 0xffffffffu;
                auto take = [&](bool out, float wt, unsigned code) {
                    const unsigned o = out ? orderable(wt) : 0xffffffffu;
                    if (o < bw) { bw = o; be = code; }   // strict: on equal weight the earlier (smaller) code stays
                };
                take((wu > 0.f) & (!in_u | (lu != cl)), wu, (unsigned)(r << 7 | c));
                take((wd > 0.f) & (!in_d | (ld != cl)), wd, (unsigned)((r + 1) << 7 | c));
                take((wl > 0.f) & (!in_l | (ll != cl)), wl, (unsigned)(1 << 14 | r << 7 | c));
                take((wr > 0.f) & (!in_r | (lr != cl)), wr, (unsigned)(1 << 14 | r << 7 | (c + 1)));
                kw[q] = bw;
                ke[q] = be;
                if (be == 0xffffffffu) interior |= 1u << q;
            }
            MST_STAMP(-10ll); MST_STAMP(clock64());
            // min weight per component, then min code among the edges of that weight; an atomic is issued only when it
            // can lower the slot.  (Measured: electing one lane per warp when all its candidates share a component -- the
            // usual case once components are large -- costs more in votes and shuffles than it saves in atomics.)
#pragma unroll
            for (int q = 0; q < PER; ++q)
                if (ke[q] != 0xffffffffu && kw[q] < s_bw[kc[q]]) atomicMin(s_bw + kc[q], kw[q]);
            __syncthreads();
            MST_STAMP(-11ll); MST_STAMP(clock64());
#pragma unroll
            for (int q = 0; q < PER; ++q)
                if (ke[q] != 0xffffffffu && s_bw[kc[q]] == kw[q] && ke[q] < s_be[kc[q]]) atomicMin(s_be + kc[q], ke[q]);
            __syncthreads();
            MST_STAMP(-12ll); MST_STAMP(clock64());
        }
        // hook: a root whose cheapest edge lies inside the tile follows it (of a mutual pair the larger root hooks)
        int merged = 0;
        // four vertices at a time, every stage's loads issued together (labels, codes, endpoint labels, the other root's
        // code): the stages are chains of dependent shared-memory reads
#pragma unroll
        for (int h = 0; h < PER; h += 4) {
            int l4[4], la[4], lb[4], other[4];
            unsigned code[4];
            bool act[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int r = rbase + (h + k) * (TNT / TS);
                l4[k] = r * TS + c;
                act[k] = r < th && c < tw && s_lab[l4[k]] == l4[k];
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) code[k] = act[k] ? s_be[l4[k]] : 0xffffffffu;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int horiz = code[k] >> 14, er = (code[k] >> 7) & 127, ec = code[k] & 127;
                act[k] = code[k] != 0xffffffffu && (horiz ? (ec >= 1 && ec < tw) : (er >= 1 && er < th));   // else: leaves the tile
                lb[k] = act[k] ? er * TS + ec : 0;
                la[k] = act[k] ? (horiz ? lb[k] - 1 : lb[k] - TS) : 0;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int ca = s_lab[la[k]], cb = s_lab[lb[k]];
                other[k] = (ca == l4[k]) ? cb : ca;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (!act[k] || (s_be[other[k]] == code[k] && l4[k] < other[k])) continue;   // same edge from both sides: the smaller root stays
                const int horiz = code[k] >> 14;
                s_par[l4[k]] = (unsigned short)other[k];
                if (contract) {   // mask bits of both endpoints: 1 up, 2 left, 4 right, 8 down
                    atomicOr(s_sel + (la[k] >> 2), (horiz ? 4u : 8u) << ((la[k] & 3) * 8));
                    atomicOr(s_sel + (lb[k] >> 2), (horiz ? 2u : 1u) << ((lb[k] & 3) * 8));
                } else {
                    const int er = (code[k] >> 7) & 127, ec = code[k] & 127;
                    intree[horiz ? nvert + (h0 + er) * (W - 1) + w0 + ec - 1 : (h0 + er - 1) * W + w0 + ec] = 1;
                }
                merged = 1;
            }
        }
        merged = __syncthreads_or(merged);
        MST_STAMP(-13ll); MST_STAMP(clock64());
        // chase: labels follow the hooks; cheapest-edge slots are cleared for the next round
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const int l = (rbase + q * (TNT / TS)) * TS + c;
            int r = s_lab[l], p = s_par[r];
            while (p != r) { r = p; p = s_par[r]; }
            s_lab[l] = (unsigned short)r;
            s_bw[l] = 0xffffffffu;
            s_be[l] = 0xffffffffu;
        }
        __syncthreads();
        MST_STAMP(clock64());
        if (!merged) break;
    }
    __shared__ int s_nact, s_nroot, s_abase, s_rbase;
    if (contract) {
        // ---- hand over the CONTRACTED graph: the tile's components get dense numbers (one reservation per tile on the
        //      tree's counter), and every edge the tile owns (below / right of its vertices) whose ends lie in different
        //      components -- or in different tiles -- becomes one record.  The adjacency mask of the edges chosen so far
        //      is final for this tile; the contraction kernel only adds bits.
        if (tid == 0) { s_nact = 0; s_nroot = 0; s_abase = 0; }
        __syncthreads();
        // Of all the edges between the same two components of the tile only the cheapest can ever be chosen, so only that
        // one is handed on (256^2: ~1 400 cross edges per tile, ~300 component pairs): a hash table over the pairs in the
        // cheapest-edge slots, which are free now -- key = the two tile ranks, min weight, then min edge id among the
        // edges of that weight.  An edge that finds no slot (table full around its hash) is handed on as it is.
        // Table of T2 slots in the cheapest-edge arrays; the slot an edge found is remembered (two bytes per edge, in the
        // third array) so that the second and third pass neither look the ranks up nor probe again.
        constexpr int T2 = TV / 2;
        unsigned *tk = s_bw, *tw_ = s_bw + T2, *te = s_be;
        unsigned short *s_slot = reinterpret_cast<unsigned short *>(s_te);   // [2 TV]: edge 2 l + kind -> slot, 0xffff = none
        constexpr unsigned EMPTY = 0xffffffffu;
        for (int l = tid; l < T2; l += TNT) { tk[l] = EMPTY; tw_[l] = EMPTY; te[l] = EMPTY; }
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const int r = rbase + q * (TNT / TS), l = r * TS + c;
            const bool valid = r < th && c < tw;
            const bool is_root = valid && s_lab[l] == l;
            const unsigned m = __ballot_sync(0xffffffffu, is_root);
            if (m) {
                const int leader = __ffs(m) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&s_nroot, __popc(m));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (is_root) s_par[l] = (unsigned short)(base + __popc(m & ((1u << lane) - 1u)));   // dense rank in the tile
            }
        }
        __syncthreads();
        MST_STAMP(-2ll);
        MST_STAMP(clock64());
        // edge k = 2q + kind of this thread: kind 0 = to the vertex below, 1 = to the right.  First which of the 16 join two
        // components inside the tile (in_mask) or leave the tile (out_mask); the passes below only visit those.
        unsigned in_mask = 0, out_mask = 0;
#pragma unroll
        for (int k = 0; k < 2 * PER; ++k) {
            const int r = rbase + (k >> 1) * (TNT / TS), l = r * TS + c;
            const bool valid = r < th && c < tw;
            const bool exists = valid && ((k & 1) ? (w0 + c + 1 < W) : (h0 + r + 1 < H));
            const bool inside = exists && ((k & 1) ? (c + 1 < tw) : (r + 1 < th));
            if (inside && s_lab[l] != s_lab[(k & 1) ? l + 1 : l + TS]) in_mask |= 1u << k;
            if (exists && !inside) out_mask |= 1u << k;
        }
        struct EdgeInfo { unsigned pk, ow, eid; int ra, rb; };
        auto edge = [&](int k, bool inside) {
            EdgeInfo x;
            const int kind = k & 1, r = rbase + (k >> 1) * (TNT / TS), l = r * TS + c;
            x.ra = s_par[s_lab[l]];
            x.rb = inside ? s_par[s_lab[kind ? l + 1 : l + TS]] : 0;
            x.pk = (unsigned)min(x.ra, x.rb) << 12 | (unsigned)max(x.ra, x.rb);
            x.ow = orderable(kind ? s_wh[l + r + 1] : s_wv[l + TS]);
            x.eid = (unsigned)(kind ? nvert + (h0 + r) * (W - 1) + w0 + c : (h0 + r) * W + w0 + c);
            return x;
        };
        auto slot_of = [&](unsigned pk, bool insert) {
            const unsigned h = (pk * 2654435761u) >> 21;
            for (int t = 0; t < 32; ++t) {
                const unsigned sl = (h + t) & (T2 - 1);
                unsigned k = tk[sl];
                if (k == EMPTY && insert) k = atomicCAS(tk + sl, EMPTY, pk), k = (k == EMPTY) ? pk : k;
                if (k == pk) return (int)sl;
                if (k == EMPTY) return -1;
            }
            return -1;
        };
        auto light = [&](int k, int &l, unsigned &ow, unsigned &eid) {   // what the later passes need of edge k: no ranks
            const int kind = k & 1, r = rbase + (k >> 1) * (TNT / TS);
            l = r * TS + c;
            ow = orderable(kind ? s_wh[l + r + 1] : s_wv[l + TS]);
            eid = (unsigned)(kind ? nvert + (h0 + r) * (W - 1) + w0 + c : (h0 + r) * W + w0 + c);
        };
        for (unsigned m = in_mask; m; m &= m - 1) {
            const int k = __ffs(m) - 1;
            const EdgeInfo x = edge(k, true);
            const int sl = slot_of(x.pk, true);
            s_slot[2 * (((rbase + (k >> 1) * (TNT / TS)) * TS) + c) + (k & 1)] = (unsigned short)(sl < 0 ? 0xffff : sl);
            if (sl >= 0 && x.ow < tw_[sl]) atomicMin(tw_ + sl, x.ow);   // (a long common boundary: hundreds of edges on one slot)
        }
        __syncthreads();
        for (unsigned m = in_mask; m; m &= m - 1) {
            int l;
            unsigned ow, eid;
            light(__ffs(m) - 1, l, ow, eid);
            const unsigned sl = s_slot[2 * l + ((__ffs(m) - 1) & 1)];
            if (sl != 0xffffu && tw_[sl] == ow && eid < te[sl]) atomicMin(te + sl, eid);
        }
        __syncthreads();
        unsigned emit_mask = out_mask;
        for (unsigned m = in_mask; m; m &= m - 1) {
            const int k = __ffs(m) - 1;
            int l;
            unsigned ow, eid;
            light(k, l, ow, eid);
            const unsigned sl = s_slot[2 * l + (k & 1)];
            if (sl == 0xffffu || (tw_[sl] == ow && te[sl] == eid)) emit_mask |= 1u << k;
        }
        int nrec = __popc(emit_mask);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) nrec += __shfl_xor_sync(0xffffffffu, nrec, o);
        if (lane == 0 && nrec) atomicAdd(&s_nact, nrec);
        __syncthreads();
        MST_STAMP(clock64());
        if (tid == 0) s_rbase = atomicAdd(cnt + 2, s_nroot);
        if (tid == 32) s_nact = atomicAdd(cnt + 0, s_nact);   // from here on: first record slot of this tile
        __syncthreads();
        Rec *rec = ws.rec + (size_t)b * E + s_nact;
        unsigned char *mask = mask_out + (size_t)b * V;
        const unsigned char *s_sel8 = reinterpret_cast<const unsigned char *>(s_sel);
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const int r = rbase + q * (TNT / TS), l = r * TS + c;
            if (r < th && c < tw) {
                const int gv = (h0 + r) * W + w0 + c;
                mask[gv] = s_sel8[l];
                if (r == 0 || c == 0) comp[gv] = s_rbase + s_par[s_lab[l]];   // the far end of an edge owned by the tile above / to the left
            }
        }
        if (__any_sync(0xffffffffu, emit_mask != 0u)) {
            int mine = __popc(emit_mask), x = mine;   // one reservation per warp, lanes in order
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += t;
            }
            int base = 0;
            if (lane == 31) base = atomicAdd(&s_abase, x);
            base = __shfl_sync(0xffffffffu, base, 31) + x - mine;
            for (unsigned m = emit_mask; m; m &= m - 1) {
                const int k = __ffs(m) - 1;
                const bool inside = (in_mask >> k) & 1u;
                const EdgeInfo e = edge(k, inside);
                Rec y;
                y.eid = e.eid;
                y.a = s_rbase + e.ra;
                y.b = inside ? s_rbase + e.rb : UNRESOLVED;
                y.ow = e.ow;
                rec[base++] = y;
            }
        }
        MST_STAMP(clock64());
        MST_STAMP(-1ll);
        return;
    }
    // ---- hand over to the cluster kernel: global labels, identity hooks, root and active lists.  The two lists are
    //      collected in shared memory (the cheapest-edge slots are free now) and appended with ONE reservation each.
    int *s_act = reinterpret_cast<int *>(s_bw), *s_root = reinterpret_cast<int *>(s_be);
    if (tid == 0) { s_nact = 0; s_nroot = 0; }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < PER; ++q) {
        const int r = rbase + q * (TNT / TS), l = r * TS + c;
        const bool valid = r < th && c < tw;
        bool is_root = false, is_act = false;
        int gv = 0;
        if (valid) {
            gv = (h0 + r) * W + w0 + c;
            const int cl = s_lab[l];
            is_root = cl == l;
            is_act = !(interior & (1u << q));
            comp[gv] = is_act ? (h0 + (cl >> 6)) * W + w0 + (cl & (TS - 1)) : DEAD;
            parent[gv] = gv;
            best[gv] = KEY_INF;
        }
        unsigned m = __ballot_sync(0xffffffffu, is_act);
        if (m) {
            const int leader = __ffs(m) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(&s_nact, __popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (is_act) s_act[base + __popc(m & ((1u << lane) - 1u))] = gv;
        }
        m = __ballot_sync(0xffffffffu, is_root);
        if (m) {
            const int leader = __ffs(m) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(&s_nroot, __popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (is_root) s_root[base + __popc(m & ((1u << lane) - 1u))] = gv;
        }
    }
    __syncthreads();
    if (tid == 0) s_abase = atomicAdd(cnt + 0, s_nact);
    if (tid == 32) s_rbase = atomicAdd(cnt + 2, s_nroot);
    __syncthreads();
    for (int i = tid; i < s_nact; i += TNT) act[s_abase + i] = s_act[i];
    for (int i = tid; i < s_nroot; i += TNT) roots[s_rbase + i] = s_root[i];
}

// ---------------------------------------------------------------------------------------------------------------
// Boruvka rounds on the CONTRACTED graph the tile kernel leaves behind (256^2: ~1 500 components and ~4 500 cross
// edges per tree instead of 65 536 vertices), one CTA per tree, out of shared memory: per round three light passes over
// the edge records (native 32-bit atomicMin of the orderable weight on both components, then of the edge id among the
// edges of that weight -- the same total order as everywhere else, hence the same edge set -- then which record it
// was), the hooks with the mutual-pair rule, a pointer chase over the components, and an in-place ordered compaction of
// the records that still join two components.  A round costs a few microseconds of shared-memory traffic instead of
// ~20 us of dependent L2 round trips in the cluster kernel.
// Storage: records and component arrays in shared memory when they fit (MODE 0); else the records stay where the tile
// kernel put them and are streamed from L2 (MODE 1: component arrays in shared memory, 16-bit numbers; MODE 2: in the
// workspace, any count).
template <int MODE> struct CStore;
template <> struct CStore<0> {
    typedef unsigned short id_t;
    unsigned *e, *w;
    id_t *a, *b;
    unsigned *bw, *be; int *bi; id_t *par, *lab;
    __device__ __forceinline__ void get(int i, unsigned &x, unsigned &ow, int &ca, int &cb) const { x = e[i]; ow = w[i]; ca = a[i]; cb = b[i]; }
    __device__ __forceinline__ void ends(int i, int &ca, int &cb) const { ca = a[i]; cb = b[i]; }
    __device__ __forceinline__ void set(int i, unsigned x, unsigned ow, int ca, int cb) { e[i] = x; w[i] = ow; a[i] = (id_t)ca; b[i] = (id_t)cb; }
};
template <typename ID> struct CStoreGlobal {
    typedef ID id_t;
    Rec *rec;
    unsigned *bw, *be; int *bi; id_t *par, *lab;
    __device__ __forceinline__ void get(int i, unsigned &x, unsigned &ow, int &ca, int &cb) const {
        const int4 t = *reinterpret_cast<const int4 *>(rec + i);
        x = (unsigned)t.x; ca = t.y; cb = t.z; ow = (unsigned)t.w;
    }
    __device__ __forceinline__ void ends(int i, int &ca, int &cb) const { ca = rec[i].a; cb = rec[i].b; }
    __device__ __forceinline__ void set(int i, unsigned x, unsigned ow, int ca, int cb) {
        *reinterpret_cast<int4 *>(rec + i) = make_int4((int)x, ca, cb, (int)ow);
    }
};
template <> struct CStore<1> : CStoreGlobal<unsigned short> {};
template <> struct CStore<2> : CStoreGlobal<int> {};

constexpr int CNT = 1024;   // threads of the contraction CTA

__device__ __forceinline__ void or_mask_bit(unsigned char *mask, int v, unsigned bit) {
    const uintptr_t p = (uintptr_t)(mask + v);
    atomicOr(reinterpret_cast<unsigned *>(p & ~(uintptr_t)3), bit << ((p & 3) * 8));
}

template <int MODE>
__device__ void contract_rounds(CStore<MODE> g, int R, int n, unsigned char *mask, int H, int W, int prof, int pslot) {
    typedef typename CStore<MODE>::id_t id_t;
    __shared__ int s_wcnt[2][32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nvert = (H - 1) * W;
    for (int c = tid; c < R; c += CNT) g.par[c] = (id_t)c;
    while (n > 0) {
        for (int c = tid; c < R; c += CNT) { g.bw[c] = 0xffffffffu; g.be[c] = 0xffffffffu; }
        __syncthreads();
        for (int i = tid; i < n; i += CNT) {   // cheapest weight of every component
            unsigned e, ow; int a, b;
            g.get(i, e, ow, a, b);
            if (ow < g.bw[a]) atomicMin(g.bw + a, ow);
            if (ow < g.bw[b]) atomicMin(g.bw + b, ow);
        }
        __syncthreads();
        for (int i = tid; i < n; i += CNT) {   // smallest edge id among the edges of that weight
            unsigned e, ow; int a, b;
            g.get(i, e, ow, a, b);
            if (ow == g.bw[a] && e < g.be[a]) atomicMin(g.be + a, e);
            if (ow == g.bw[b] && e < g.be[b]) atomicMin(g.be + b, e);
        }
        __syncthreads();
        MST_STAMP(clock64());
        for (int i = tid; i < n; i += CNT) {   // which record it was (edge ids are unique)
            unsigned e, ow; int a, b;
            g.get(i, e, ow, a, b);
            if (g.be[a] == e) g.bi[a] = i;
            if (g.be[b] == e) g.bi[b] = i;
        }
        __syncthreads();
        MST_STAMP(clock64());
        int any = 0;
        for (int c = tid; c < R; c += CNT) {   // hooks: of two roots that picked the same edge the larger one hooks
            if (g.par[c] != c) continue;
            const unsigned e = g.be[c];
            if (e == 0xffffffffu) continue;
            int a, b;
            g.ends(g.bi[c], a, b);
            const int other = (a == c) ? b : a;
            if (g.be[other] == e && c < other) continue;
            g.par[c] = (id_t)other;
            if ((int)e < nvert) {
                or_mask_bit(mask, (int)e, 8u);
                or_mask_bit(mask, (int)e + W, 1u);
            } else {
                const int r = (int)e - nvert, h = r / (W - 1), u = h * W + (r - h * (W - 1));
                or_mask_bit(mask, u, 4u);
                or_mask_bit(mask, u + 1, 2u);
            }
            any = 1;
        }
        if (!__syncthreads_or(any)) break;
        MST_STAMP(clock64());
        for (int c = tid; c < R; c += CNT) {
            int r = c, p = g.par[r];
            while (p != r) { r = p; p = g.par[r]; }
            g.lab[c] = (id_t)r;
        }
        __syncthreads();
        for (int c = tid; c < R; c += CNT) g.par[c] = g.lab[c];
        MST_STAMP(clock64());
        // relabel the records and keep those that still join two components: ordered compaction in place, a chunk of
        // CNT records at a time (a chunk's survivors land at or before the chunk, its reads precede its writes)
        int out = 0, buf = 0;
        for (int i0 = 0; i0 < n; i0 += CNT, buf ^= 1) {
            const int i = i0 + tid;
            unsigned e = 0, ow = 0;
            int a = 0, b = 0;
            if (i < n) {
                g.get(i, e, ow, a, b);
                a = g.lab[a];
                b = g.lab[b];
            }
            const bool keep = a != b;
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) s_wcnt[buf][wid] = __popc(m);
            __syncthreads();
            const int mine = s_wcnt[buf][lane];
            int x = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += t;
            }
            const int total = __shfl_sync(0xffffffffu, x, 31);
            const int before = __shfl_sync(0xffffffffu, x - mine, wid);
            if (keep) g.set(out + before + __popc(m & ((1u << lane) - 1u)), e, ow, a, b);
            out += total;
        }
        __syncthreads();
        n = out;
        MST_STAMP(clock64());
        MST_STAMP((long long)n);
    }
    MST_STAMP(-1ll);
}

__global__ void __launch_bounds__(CNT) mst_contract_kernel(Ws ws, unsigned char *__restrict__ mask_out, int H, int W, int pool_bytes,
                                                           int prof) {
    extern __shared__ __align__(16) unsigned char s_pool[];
    const int b = blockIdx.x, tid = threadIdx.x, V = H * W, E = 2 * H * W - H - W, nvert = (H - 1) * W;
    const int n = ws.cnt[(size_t)b * 4 + 0], R = ws.cnt[(size_t)b * 4 + 2];
    Rec *rec = ws.rec + (size_t)b * E;
    const int *comp = ws.comp + (size_t)b * V;
    unsigned char *mask = mask_out + (size_t)b * V;
    auto far_end = [&](unsigned e) {   // second endpoint (below / right) of a grid edge
        if ((int)e < nvert) return (int)e + W;
        const int r = (int)e - nvert, h = r / (W - 1);
        return h * W + (r - h * (W - 1)) + 1;
    };
    int pslot = 0;
    MST_STAMP(clock64());
    MST_STAMP((long long)n);
    MST_STAMP((long long)R);
    const size_t comps_bytes = (size_t)R * 16 + 32;
    if (R <= 65534 && comps_bytes + (size_t)n * 12 <= (size_t)pool_bytes) {
        CStore<0> g;
        g.bw = reinterpret_cast<unsigned *>(s_pool);
        g.be = g.bw + R;
        g.bi = reinterpret_cast<int *>(g.be + R);
        g.e = reinterpret_cast<unsigned *>(g.bi + R);
        g.w = g.e + n;
        g.par = reinterpret_cast<unsigned short *>(g.w + n);
        g.lab = g.par + R;
        g.a = g.lab + R;
        g.b = g.a + n;
        for (int i = tid; i < n; i += CNT) {
            const int4 t = *reinterpret_cast<const int4 *>(rec + i);
            g.set(i, (unsigned)t.x, (unsigned)t.w, t.y, t.z == UNRESOLVED ? comp[far_end((unsigned)t.x)] : t.z);
        }
        __syncthreads();
        MST_STAMP(clock64());
        contract_rounds<0>(g, R, n, mask, H, W, prof, pslot);
        return;
    }
    for (int i = tid; i < n; i += CNT)
        if (rec[i].b == UNRESOLVED) rec[i].b = comp[far_end(rec[i].eid)];
    if (R <= 65534 && comps_bytes <= (size_t)pool_bytes) {
        CStore<1> g;
        g.rec = rec;
        g.bw = reinterpret_cast<unsigned *>(s_pool);
        g.be = g.bw + R;
        g.bi = reinterpret_cast<int *>(g.be + R);
        g.par = reinterpret_cast<unsigned short *>(g.bi + R);
        g.lab = g.par + R;
        __syncthreads();
        MST_STAMP(clock64());
        contract_rounds<1>(g, R, n, mask, H, W, prof, pslot);
    } else {
        CStore<2> g;
        g.rec = rec;
        g.bw = reinterpret_cast<unsigned *>(ws.best + (size_t)b * V);
        g.be = g.bw + V;
        g.bi = ws.act + (size_t)b * 2 * V;
        g.lab = g.bi + V;
        g.par = ws.parent + (size_t)b * V;
        __syncthreads();
        MST_STAMP(clock64());
        contract_rounds<2>(g, R, n, mask, H, W, prof, pslot);
    }
}

// Tile-local rounds before the finish.  The tile kernel is issue-bound (~10 us per round and CTA, 2.6 waves of CTAs for 48
// trees of 256^2) and its late rounds hook little; measured on B200, 48 trees, total MST time with the contracted
// finish: 3 rounds 245 us, 4 rounds 244 us, 5 rounds 273 us (512^2: 955 / 928 / 1 021); with the cluster finish: 3 rounds
// 381 us, 4 rounds 367 us, 5 rounds 377 us, 6 rounds 393 us, until nothing hooks (7-8 rounds) 420 us, no tile kernel 490 us.
int env_tile_rounds() {
    const char *v = getenv("FEDICRA_MST_TILE_ROUNDS");
    const int x = (v && *v) ? atoi(v) : 4;
    return x >= 1 ? x : 4;
}

int env_tiles() {
    const char *v = getenv("FEDICRA_MST_TILES");
    return !(v && *v == '0');
}

// FEDICRA_MST_FINISH=cluster: the rounds after the tile kernel on the vertex lists, by thread-block clusters (round 1's
// path); default: on the contracted graph
int env_contract() {
    const char *v = getenv("FEDICRA_MST_FINISH");
    return !(v && *v == 'c' && v[1] == 'l');
}

int env_cl(int dflt) {
    const char *v = getenv("FEDICRA_MST_CL");
    const int x = (v && *v) ? atoi(v) : dflt;
    return (x >= 1 && x <= 8) ? x : dflt;
}

int rounds_for(int V) {
    int r = 0;
    while ((1ll << r) < V) ++r;
    return r + 1;
}

template <bool GRID>
int run_mst(const int *edge_index, const float *edge_weight, int B, int E, int V, int H, int W, int *edge_out, Ws ws,
            cudaStream_t st) {
    if (V <= 1 || E <= 0) return 0;  // a single vertex has an empty tree
    const int n = E > V ? E : V;
    const dim3 blk(256);
    mst_init_kernel<<<dim3((n + 255) / 256, B), blk, 0, st>>>(ws, V, E);
    FEDICRA_CHECK_LAUNCH("mst_init_kernel");
    const int rounds = rounds_for(V);
    for (int r = 0; r < rounds; ++r) {
        mst_find_kernel<GRID><<<dim3((E + 255) / 256, B), blk, 0, st>>>(ws, edge_index, edge_weight, V, E, H, W);
        mst_hook_kernel<GRID><<<dim3((V + 255) / 256, B), blk, 0, st>>>(ws, edge_index, V, E, H, W);
        mst_compress_kernel<<<dim3((V + 255) / 256, B), blk, 0, st>>>(ws, V);
    }
    fedicra_count_launches(3 * rounds - 1);
    FEDICRA_CHECK_LAUNCH("mst round kernels");
    if (edge_out != nullptr) {
        mst_compact_kernel<GRID><<<B, 1024, 0, st>>>(ws, edge_index, edge_out, V, E, H, W, nullptr);
        FEDICRA_CHECK_LAUNCH("mst_compact_kernel");
    }
    return 0;
}

}  // namespace

extern "C" size_t fedicra_mst_workspace_bytes(int B, int V, int E) {
    if (B <= 0 || V <= 0 || E < 0) return 0;
    return ws_layout(B, V, E, nullptr, nullptr);
}

extern "C" int fedicra_mst_forward(const int *edge_index, const float *edge_weight, int B, int E, int V, int *edge_out,
                                   void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(edge_index && edge_weight && edge_out && workspace, "mst_forward: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && V >= 1 && E >= V - 1, "mst_forward: bad sizes B=%d E=%d V=%d", B, E, V);
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_mst_workspace_bytes(B, V, E), "mst_forward: workspace too small");
    FEDICRA_REQUIRE(((uintptr_t)edge_index & 7) == 0 && ((uintptr_t)edge_out & 7) == 0,
                    "mst_forward: edge arrays must be 8-byte aligned");
    Ws ws;
    ws_layout(B, V, E, (char *)workspace, &ws);
    return run_mst<false>(edge_index, edge_weight, B, E, V, 0, 0, edge_out, ws, (cudaStream_t)stream);
}

namespace {
int mst_grid_launch(const EmbedSet &es, int B, int H, int W, int *edge_out, unsigned char *tree_mask, void *workspace,
                    size_t workspace_bytes, cudaStream_t st) {
    const int V = H * W, E = 2 * H * W - H - W;
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_mst_workspace_bytes(B, V, E), "mst_grid: workspace too small");
    FEDICRA_REQUIRE(((uintptr_t)edge_out & 7) == 0, "mst_grid: edge_out must be 8-byte aligned");  // (NULL passes)
    Ws ws;
    ws_layout(B, V, E, (char *)workspace, &ws);
    if (E > 0) {
        // one persistent launch: a cluster of CL CTAs per image.
        // the largest cluster that keeps the whole batch in one wave.  63 registers x 512 threads = two CTAs per SM, 296
        // slots, but clusters are placed GPC by GPC and do not pack perfectly: measured on B200 with 48 trees, clusters of
        // 5 (240 CTAs) take 490 us, of 4 570 us, of 6 (288 CTAs, spills into a second wave) 680 us -- so leave 1/8 slack
        int CL = 8;
        while (CL > 1 && (long long)B * CL > 2 * 148 - 37) --CL;
        CL = env_cl(CL);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(B * CL));
        cfg.blockDim = dim3(PNT);
        cfg.dynamicSmemBytes = 0;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)CL;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (env_tiles() && env_contract()) {   // tile rounds, then the contracted graph: one CTA per tree
            constexpr int POOL = 220 * 1024;
            int pool = POOL;   // FEDICRA_MST_POOL_KB: shared memory the contraction may use (tests: forces the streaming modes)
            if (const char *v = getenv("FEDICRA_MST_POOL_KB")) {
                const int x = atoi(v);
                if (x >= 1 && x * 1024 < POOL) pool = x * 1024;
            }
            cudaError_t e0 = cudaFuncSetAttribute(mst_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TILE_SMEM);
            if (e0 == cudaSuccess) e0 = cudaFuncSetAttribute(mst_contract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, POOL);
            if (e0 != cudaSuccess) return fedicra_fail(-2, "mst kernels: %s", cudaGetErrorString(e0));
            unsigned char *mask = tree_mask != nullptr ? tree_mask : ws.tmask;
            static const int prof = [] { const char *v = getenv("FEDICRA_PROF"); return (v && *v && *v != '0') ? 1 : 0; }();
            cudaMemsetAsync(ws.cnt, 0, (size_t)B * 4 * sizeof(int), st);
            const int tiles_x = (W + TS - 1) / TS, tiles_y = (H + TS - 1) / TS;
            mst_tile_kernel<<<dim3(tiles_x * tiles_y, B), TNT, TILE_SMEM, st>>>(ws, es, H, W, tiles_x, env_tile_rounds(), mask, 1 | (prof << 1));
            FEDICRA_CHECK_LAUNCH("mst_tile_kernel");
            mst_contract_kernel<<<B, CNT, POOL, st>>>(ws, mask, H, W, pool, prof);
            FEDICRA_CHECK_LAUNCH("mst_contract_kernel");
            if (edge_out != nullptr) {
                mst_compact_kernel<true><<<B, 1024, 0, st>>>(ws, nullptr, edge_out, V, E, H, W, mask);
                FEDICRA_CHECK_LAUNCH("mst_compact_kernel");
            }
            return 0;
        }
        int prepared = 0;
        if (env_tiles()) {   // rounds that stay inside 64x64 tiles first, out of shared memory
            cudaError_t e0 = cudaFuncSetAttribute(mst_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TILE_SMEM);
            if (e0 != cudaSuccess) return fedicra_fail(-2, "mst_tile_kernel: %s", cudaGetErrorString(e0));
            cudaMemsetAsync(ws.cnt, 0, (size_t)B * 4 * sizeof(int), st);
            const int tiles_x = (W + TS - 1) / TS, tiles_y = (H + TS - 1) / TS;
            mst_tile_kernel<<<dim3(tiles_x * tiles_y, B), TNT, TILE_SMEM, st>>>(ws, es, H, W, tiles_x, env_tile_rounds(), nullptr, 0);
            FEDICRA_CHECK_LAUNCH("mst_tile_kernel");
            prepared = 1;
        }
        cudaError_t e = cudaLaunchKernelEx(&cfg, mst_grid_cluster_kernel, ws, es, tree_mask, H, W, rounds_for(V), prepared);
        fedicra_count_launches(1);
        if (e != cudaSuccess) return fedicra_fail(-2, "mst_grid_cluster_kernel: %s", cudaGetErrorString(e));
        if (edge_out != nullptr) {
            mst_compact_kernel<true><<<B, 1024, 0, st>>>(ws, nullptr, edge_out, V, E, H, W, nullptr);
            FEDICRA_CHECK_LAUNCH("mst_compact_kernel");
        }
        return 0;
    }
    // a single vertex: the empty tree
    if (tree_mask != nullptr) cudaMemsetAsync(tree_mask, 0, (size_t)B * V, st);
    return 0;
}
}  // namespace

extern "C" int fedicra_mst_grid(const float *embed, int B, int D, int H, int W, int *edge_out,
                                unsigned char *tree_mask, void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(embed && workspace && (edge_out || tree_mask || (long long)H * W <= 1),
                    "mst_grid: null pointer (edge_out and tree_mask cannot both be NULL)");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && D > 0 && H >= 1 && W >= 1 && (long long)H * W < (1ll << 30),
                    "mst_grid: bad sizes B=%d D=%d H=%d W=%d", B, D, H, W);
    EmbedSet es = {};
    es.ptr[0] = embed;
    es.D[0] = D;
    es.per = B;
    return mst_grid_launch(es, B, H, W, edge_out, tree_mask, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int fedicra_mst_grid_multi(const float *const *h_embeds, const int *h_D, int K, int B, int H, int W,
                                      int *edge_out, unsigned char *tree_mask, void *workspace, size_t workspace_bytes,
                                      void *stream) {
    FEDICRA_REQUIRE(h_embeds && h_D && workspace && (edge_out || tree_mask || (long long)H * W <= 1),
                    "mst_grid_multi: null pointer (edge_out and tree_mask cannot both be NULL)");
    FEDICRA_REQUIRE(K >= 1 && K <= MAX_EMBEDS, "mst_grid_multi: %d embeddings (1..%d)", K, MAX_EMBEDS);
    FEDICRA_REQUIRE(B > 0 && (long long)B * K <= 65535 && H >= 1 && W >= 1 && (long long)H * W < (1ll << 30),
                    "mst_grid_multi: bad sizes B=%d K=%d H=%d W=%d", B, K, H, W);
    EmbedSet es = {};
    for (int k = 0; k < K; ++k) {
        FEDICRA_REQUIRE(h_embeds[k] && h_D[k] > 0, "mst_grid_multi: embedding %d is null or has no channels", k);
        es.ptr[k] = h_embeds[k];
        es.D[k] = h_D[k];
    }
    es.per = B;
    return mst_grid_launch(es, B * K, H, W, edge_out, tree_mask, workspace, workspace_bytes, (cudaStream_t)stream);
}

/* Debug: clock64 stamps of tree 0's last contraction (FEDICRA_PROF=1); see g_mst_prof.  Synchronises the device. */
extern "C" int fedicra_debug_mst_profile(long long *h_out, int n) {
    if (!h_out || n < MPROF) return fedicra_fail(-1, "debug_mst_profile: need room for %d values", MPROF);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(h_out, g_mst_prof, sizeof(long long) * MPROF);
    if (e != cudaSuccess) return fedicra_fail(-2, "debug_mst_profile: %s", cudaGetErrorString(e));
    return MPROF;
}
