// bfs.cu -- breadth-first ordering of a batch of spanning trees from vertex 0 (sm_100a).
//
// Replaces bfs_forward of the reference (LTF/src/bfs/bfs.cu:100-135: adj_vec_kernel :19-44 and
// breadth_first_sort_kernel :46-98, 64 threads per image, shared atomic tail, spin on __syncthreads).  The three output
// arrays mean the same thing (bfs.cu:84-90):
//   sorted_index[i]    vertex at BFS position i            (position 0 = vertex 0, the root)
//   sorted_parent[i]   BFS position of its parent          (0 for the root)
//   sorted_child[i][k] BFS positions of its children, 0 = none
// but the order is canonical instead of atomic-dependent: by depth, then by parent position, then by vertex id.  It is
// therefore a true level order, children of one vertex are contiguous, and `level_start` (extra output) delimits the
// depths -- which is what lets the tree-filter sweeps run one whole depth per step.
#include "common.cuh"

namespace {

constexpr int MAXADJ = 4;
constexpr int NT = 256;
constexpr int CAP = 2048;  // frontier entries kept in shared memory; wider levels fall back to global memory

struct Ws {
    int *adj;  // [B,V,4]
    int *deg;  // [B,V]
};

size_t ws_layout(int B, int V, char *base, Ws *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char *p = base ? base + off : nullptr;
        off += align_up(bytes, 256);
        return p;
    };
    Ws t;
    t.deg = (int *)take((size_t)B * V * sizeof(int));
    t.adj = (int *)take((size_t)B * V * MAXADJ * sizeof(int));
    if (w) *w = t;
    return off;
}

__global__ void adjacency_kernel(const int *__restrict__ tree, Ws ws, int V) {
    const int b = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= V - 1) return;
    const int2 uv = reinterpret_cast<const int2 *>(tree)[(size_t)b * (V - 1) + e];
    if ((unsigned)uv.x >= (unsigned)V || (unsigned)uv.y >= (unsigned)V) return;  // malformed edge: ignored
    int *deg = ws.deg + (size_t)b * V;
    int *adj = ws.adj + (size_t)b * V * MAXADJ;
    const int su = atomicAdd(deg + uv.x, 1);
    if (su < MAXADJ) adj[uv.x * MAXADJ + su] = uv.y;
    const int sv = atomicAdd(deg + uv.y, 1);
    if (sv < MAXADJ) adj[uv.y * MAXADJ + sv] = uv.x;
}

__device__ __forceinline__ void sort4(int &a, int &b, int &c, int &d) {
#define CSWAP(x, y) { const int lo_ = min(x, y), hi_ = max(x, y); x = lo_; y = hi_; }
    CSWAP(a, b) CSWAP(c, d) CSWAP(a, c) CSWAP(b, d) CSWAP(b, c)
#undef CSWAP
}

__global__ void __launch_bounds__(NT)
bfs_kernel(Ws ws, int V, int *__restrict__ sorted_index, int *__restrict__ sorted_parent,
           int *__restrict__ sorted_child, int *__restrict__ level_start, int *__restrict__ n_levels) {
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int *deg = ws.deg + (size_t)b * V;
    const int4 *adj = reinterpret_cast<const int4 *>(ws.adj) + (size_t)b * V;
    int *order = sorted_index + (size_t)b * V;
    int *ppos = sorted_parent + (size_t)b * V;
    int4 *child = reinterpret_cast<int4 *>(sorted_child) + (size_t)b * V;
    int *lstart = level_start + (size_t)b * (V + 1);

    __shared__ int s_v[2][CAP], s_p[2][CAP];
    __shared__ int s_warp[NT / 32];
    __shared__ int s_chunk_total;

    if (tid == 0) {
        order[0] = 0;
        ppos[0] = 0;
        s_v[0][0] = 0;
        s_p[0][0] = -1;
        lstart[0] = 0;
    }
    __syncthreads();

    int lo = 0, hi = 1, depth = 0, cur = 0;
    while (true) {
        const int n = hi - lo;
        const bool in_smem = n <= CAP;
        int produced = 0;  // children appended so far in this level (uniform across the CTA)
        for (int base = 0; base < n; base += NT) {
            const int j = base + tid;
            int c0 = 0x7fffffff, c1 = 0x7fffffff, c2 = 0x7fffffff, c3 = 0x7fffffff, cnt = 0, v = 0, i = 0;
            if (j < n) {
                i = lo + j;
                int pv;
                if (in_smem) {
                    v = s_v[cur][j];
                    pv = s_p[cur][j];
                } else {
                    v = order[i];
                    pv = (i == 0) ? -1 : order[ppos[i]];
                }
                const int d = min(deg[v], MAXADJ);
                const int4 a = adj[v];
                if (d > 0 && a.x != pv) { c0 = a.x; ++cnt; }
                if (d > 1 && a.y != pv) { c1 = a.y; ++cnt; }
                if (d > 2 && a.z != pv) { c2 = a.z; ++cnt; }
                if (d > 3 && a.w != pv) { c3 = a.w; ++cnt; }
                sort4(c0, c1, c2, c3);  // children in increasing vertex id, absent ones last
            }
            // exclusive scan of cnt over the CTA
            int x = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += t;
            }
            if (lane == 31) s_warp[wid] = x;
            __syncthreads();
            if (wid == 0) {
                int w = (lane < NT / 32) ? s_warp[lane] : 0, y = w;
#pragma unroll
                for (int o = 1; o < NT / 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, y, o);
                    if (lane >= o) y += t;
                }
                if (lane < NT / 32) s_warp[lane] = y - w;
                if (lane == NT / 32 - 1) s_chunk_total = y;
            }
            __syncthreads();
            const int off = produced + s_warp[wid] + (x - cnt);
            if (j < n) {
                const int p0 = hi + off;
                int4 slots = make_int4(0, 0, 0, 0);
                const int cs[4] = {c0, c1, c2, c3};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (k < cnt) {
                        const int pos = p0 + k;
                        order[pos] = cs[k];
                        ppos[pos] = i;
                        if (pos - hi < CAP) {
                            s_v[cur ^ 1][pos - hi] = cs[k];
                            s_p[cur ^ 1][pos - hi] = v;
                        }
                        (&slots.x)[k] = pos;
                    }
                }
                child[i] = slots;
            }
            produced += s_chunk_total;
            __syncthreads();
        }
        if (produced == 0) break;
        lo = hi;
        hi += produced;
        ++depth;
        cur ^= 1;
        if (tid == 0) lstart[depth] = lo;
    }
    if (tid == 0) {
        lstart[depth + 1] = hi;
        n_levels[b] = depth + 1;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Grid specialisation: the tree is given as the 4-bit adjacency mask fedicra_mst_grid writes.  One warp per image walks
// the depths; the masks (packed two per byte) and the frontier live in shared memory, children positions come from
// warp ballots -- no block barrier and no global load on the per-depth dependency chain.  Same outputs, same canonical
// order (children in direction order up, left, right, down = increasing vertex id).
constexpr int GNT = 128;     // threads per CTA: one frontier entry per thread and round
constexpr int SCAP = 4096;   // ring of BFS positions staged in shared memory (power of two)
constexpr int LCAP = 4096;   // level starts staged in shared memory

// A __syncthreads() that follows global stores waits for them to drain (~0.5 us each on B200), so the walk keeps
// everything it produces -- vertex, parent position, children -- in a shared-memory ring indexed by BFS position and
// writes it to global memory in bulk, coalesced, only when the ring is about to wrap (every ~90 depths).
__global__ void __launch_bounds__(GNT)
bfs_grid_kernel(const unsigned char *__restrict__ tree_mask, int V, int W, int *__restrict__ sorted_index,
                int *__restrict__ sorted_parent, int *__restrict__ sorted_child, int *__restrict__ level_start,
                int *__restrict__ n_levels) {
    extern __shared__ unsigned s_raw[];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    unsigned *s_v = s_raw;                                    // [SCAP] vertex | (mask bit pointing back to the parent << 28)
    int *s_p = reinterpret_cast<int *>(s_raw + SCAP);         // [SCAP] parent position
    int *s_c = reinterpret_cast<int *>(s_raw + 2 * SCAP);     // [SCAP] first child position << 3 | child count
    unsigned char *s_mask = reinterpret_cast<unsigned char *>(s_raw + 3 * SCAP);  // [ceil(V/2)] two vertices per byte
    __shared__ int s_wtot[GNT / 32];
    __shared__ int s_ls[LCAP];  // level starts, flushed at the end (deeper trees write the rest directly)
    const unsigned char *gm = tree_mask + (size_t)b * V;
    int *order = sorted_index + (size_t)b * V;
    int *ppos = sorted_parent + (size_t)b * V;
    int4 *child = reinterpret_cast<int4 *>(sorted_child) + (size_t)b * V;
    int *lstart = level_start + (size_t)b * (V + 1);

    for (int k = tid; k < (V + 1) / 2; k += GNT) {
        const unsigned lo = gm[2 * k], hi = (2 * k + 1 < V) ? gm[2 * k + 1] : 0u;
        s_mask[k] = (unsigned char)(lo | (hi << 4));
    }
    if (tid == 0) {
        s_v[0] = 0u | (4u << 28);  // the root has no parent: bit 4 is outside the mask
        s_p[0] = 0;
        s_ls[0] = 0;
    }
    __syncthreads();

    auto flush = [&](int from, int to) {  // ring -> global for positions [from, to), all of them completed
        for (int pos = from + tid; pos < to; pos += GNT) {
            const int r = pos & (SCAP - 1), pk = s_c[r], p0 = pk >> 3, cnt = pk & 7;
            order[pos] = (int)(s_v[r] & 0x0fffffffu);
            ppos[pos] = s_p[r];
            child[pos] = make_int4(cnt > 0 ? p0 : 0, cnt > 1 ? p0 + 1 : 0, cnt > 2 ? p0 + 2 : 0, cnt > 3 ? p0 + 3 : 0);
        }
    };

    const unsigned lt = (1u << lane) - 1u;
    int lo = 0, hi = 1, depth = 0, base = 0;  // ring holds positions [base, ...)
    bool cur_ring = true;                     // entries of the current depth are in the ring (else in global memory)
    while (true) {
        const int n = hi - lo;
        if (!cur_ring) {  // this depth lives in global memory: the ring only holds older, completed positions
            if (base < lo) flush(base, lo);
            base = hi;
            __syncthreads();
        }
        // room for this depth's children (at most 3 per vertex, 4 for the root)?
        bool next_ring = (long long)hi + 3LL * n + 1 - base <= SCAP;
        if (!next_ring && cur_ring) {
            flush(base, lo);
            base = lo;
            next_ring = (long long)hi + 3LL * n + 1 - base <= SCAP;
            __syncthreads();
        }
        int produced = 0;
        for (int b0i = 0; b0i < n; b0i += GNT) {
            const int j = b0i + tid, i = lo + j;
            int v = 0;
            unsigned m = 0;
            if (j < n) {
                unsigned back;
                if (cur_ring) {
                    const unsigned e = s_v[i & (SCAP - 1)];
                    v = (int)(e & 0x0fffffffu);
                    back = e >> 28;
                } else {  // oversized depth, kept in global memory: re-derive the direction of the parent
                    v = order[i];
                    const int d = order[ppos[i]] - v;
                    back = (i == 0) ? 4u : ((d == -W) ? 0u : (d == -1 ? 1u : (d == 1 ? 2u : 3u)));
                }
                const unsigned byte = s_mask[v >> 1];
                m = ((v & 1) ? (byte >> 4) : (byte & 15u)) & ~(1u << back);
            }
            const int cnt = __popc(m);
            const unsigned b0 = __ballot_sync(0xffffffffu, cnt & 1), b1 = __ballot_sync(0xffffffffu, cnt & 2),
                           b2 = __ballot_sync(0xffffffffu, cnt & 4);
            if (lane == 0) s_wtot[wid] = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
            __syncthreads();
            int woff = 0, ctot = 0;
#pragma unroll
            for (int w = 0; w < GNT / 32; ++w) {
                const int t = s_wtot[w];
                woff += (w < wid) ? t : 0;
                ctot += t;
            }
            if (j < n) {
                const int p0 = hi + produced + woff + __popc(b0 & lt) + 2 * __popc(b1 & lt) + 4 * __popc(b2 & lt);
                int k = 0;
#pragma unroll
                for (int d = 0; d < 4; ++d) {  // up, left, right, down = increasing vertex id
                    if (m & (1u << d)) {
                        const int c = v + ((d == 0) ? -W : (d == 1 ? -1 : (d == 2 ? 1 : W)));
                        const int q = p0 + k;
                        if (next_ring) {
                            s_v[q & (SCAP - 1)] = (unsigned)c | ((unsigned)(3 - d) << 28);
                            s_p[q & (SCAP - 1)] = i;
                        } else {
                            order[q] = c;
                            ppos[q] = i;
                        }
                        ++k;
                    }
                }
                if (cur_ring)
                    s_c[i & (SCAP - 1)] = (p0 << 3) | cnt;
                else
                    child[i] = make_int4(cnt > 0 ? p0 : 0, cnt > 1 ? p0 + 1 : 0, cnt > 2 ? p0 + 2 : 0, cnt > 3 ? p0 + 3 : 0);
            }
            produced += ctot;
            __syncthreads();
        }
        if (produced == 0) break;
        lo = hi;
        hi += produced;
        ++depth;
        cur_ring = next_ring;
        if (tid == 0) {
            if (depth < LCAP) s_ls[depth] = lo; else lstart[depth] = lo;
        }
    }
    if (cur_ring) flush(base, hi);
    __syncthreads();
    for (int L = tid; L <= depth && L < LCAP; L += GNT) lstart[L] = s_ls[L];
    if (tid == 0) {
        lstart[depth + 1] = hi;
        n_levels[b] = depth + 1;
    }
}

}  // namespace

extern "C" int fedicra_bfs_grid(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index,
                                int *sorted_parent, int *sorted_child, int *level_start, int *n_levels, void *stream) {
    FEDICRA_REQUIRE(tree_mask && sorted_index && sorted_parent && sorted_child && level_start && n_levels,
                    "bfs_grid: null pointer");
    FEDICRA_REQUIRE(B > 0 && H >= 1 && W >= 1 && (long long)H * W < (1ll << 28), "bfs_grid: bad sizes B=%d H=%d W=%d", B, H, W);
    FEDICRA_REQUIRE(((uintptr_t)sorted_child & 15) == 0, "bfs_grid: sorted_child must be 16-byte aligned");
    const int V = H * W;
    const size_t smem = (size_t)3 * SCAP * sizeof(unsigned) + (size_t)(V + 1) / 2 + 16;  // + 16 KB static
    FEDICRA_REQUIRE(smem + 17 * 1024 <= 227 * 1024, "bfs_grid: %d x %d does not fit the shared-memory walk (use fedicra_bfs_forward)", H, W);
    cudaError_t e = cudaFuncSetAttribute(bfs_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fedicra_fail(-2, "bfs_grid: shared memory: %s", cudaGetErrorString(e));
    bfs_grid_kernel<<<B, GNT, smem, (cudaStream_t)stream>>>(tree_mask, V, W, sorted_index, sorted_parent, sorted_child,
                                                          level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("bfs_grid_kernel");
    return 0;
}

extern "C" size_t fedicra_bfs_workspace_bytes(int B, int V) {
    if (B <= 0 || V <= 0) return 0;
    return ws_layout(B, V, nullptr, nullptr);
}

extern "C" int fedicra_bfs_forward(const int *tree, int B, int V, int *sorted_index, int *sorted_parent,
                                   int *sorted_child, int *level_start, int *n_levels, void *workspace,
                                   size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(sorted_index && sorted_parent && sorted_child && level_start && n_levels && workspace,
                    "bfs_forward: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && V >= 1, "bfs_forward: bad sizes B=%d V=%d", B, V);
    FEDICRA_REQUIRE(V == 1 || tree, "bfs_forward: null tree");
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_bfs_workspace_bytes(B, V), "bfs_forward: workspace too small");
    FEDICRA_REQUIRE(((uintptr_t)tree & 7) == 0 && ((uintptr_t)sorted_child & 15) == 0,
                    "bfs_forward: tree must be 8-byte and sorted_child 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    Ws ws;
    ws_layout(B, V, (char *)workspace, &ws);
    cudaError_t e = cudaMemsetAsync(ws.deg, 0, (size_t)B * V * sizeof(int), st);
    if (e != cudaSuccess) return fedicra_fail(-2, "bfs_forward: memset: %s", cudaGetErrorString(e));
    if (V > 1) {
        adjacency_kernel<<<dim3((V - 1 + 255) / 256, B), 256, 0, st>>>(tree, ws, V);
        FEDICRA_CHECK_LAUNCH("adjacency_kernel");
    }
    bfs_kernel<<<B, NT, 0, st>>>(ws, V, sorted_index, sorted_parent, sorted_child, level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("bfs_kernel");
    return 0;
}
