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

}  // namespace

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
