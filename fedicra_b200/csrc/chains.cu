// chains.cu -- the two-pass tree filter as scans of affine maps along heavy chains (sm_100a).
//
// Replaces refine.cu:19-370 of the reference (leaf_root_aggr / root_leaf_prop / root_leaf_grad: one dependent step per
// tree depth, ~1 500 at 256^2) for trees given in CHAIN ORDER (fedicra_chain_order, csrc/euler.cu): vertices sorted by
// light depth, every heavy chain on consecutive positions top -> bottom.  In that order
//
//   leaf -> root:  x_q = b_q + a_q x_{q+1},   a_q = w_{q+1} if q has a heavy child else 0,
//                                             b_q = f(vertex q) + sum over light children c of w_c x_c   (next group)
//   root -> leaf:  y_q = c_q + a_q y_{q-1},   a_q = w_q unless q is a chain top, then 0,
//                                             c_q = x_q (1 - w_q^2) [+ w_q y_parent for a chain top]   (previous group)
//
// i.e. per group one scan of affine maps x -> a x + b (|a| <= 1: no log domain, no overflow), and the only sequential
// dimension left is the number of groups (8 - 11) instead of the tree depth.
//
// ONE launch per filter call.  Work items = (image, group, 512-position tile), handed out by an atomic ticket in an order
// in which every item only depends on items with smaller tickets (leaf -> root: groups deepest first, tiles by
// decreasing position; then root -> leaf: groups from 0, tiles increasing) -- so a waiting CTA always waits for a CTA
// that is already running, whatever the grid size.
//
// Hand-over between items costs one L2 round trip and no fence: every intermediate word is SELF-VALIDATING.  The
// intermediate arrays start as 0xFFFFFFFF (a NaN no arithmetic produces; the call memsets them), producers store
// results with relaxed gpu-scope stores, and a consumer simply re-reads the words it needs (relaxed gpu-scope loads,
// served by L2) until none of them is the start pattern.  No flags, no release / acquire pairs, no __threadfence.
// A tile also hands its boundary value to the neighbouring tile of the same group as soon as it is known -- which is
// before the tile has heard from its OTHER neighbour unless one chain spans the whole tile -- so the tiles of a group
// do not queue up behind each other.
// All channels of an image ride in one item (C feature channels + the all-ones channel whose result is the
// normaliser), so the topology is read once.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int CT = 128;          // threads per CTA
constexpr int PPT = 4;           // positions per thread
constexpr int TILE = CT * PPT;   // positions per item
constexpr int NW = CT / 32;
#ifndef CHAIN_MINB
#define CHAIN_MINB 5
#endif
constexpr int SYNC_HEAD = 64;    // ints ahead of the hand-over values: [0] ticket - 1, [1] watchdog trips - 1
constexpr unsigned PENDING = 0xffffffffu;

struct ChainArgs {
    const int *vtx, *par;        // [B,V] vertex at position, position of the parent
    const int4 *child;           // [B,V] positions of (heavy child, light children...), 0 = none
    const int *gstart;           // [B,33] first position of every group; V past the last one
    float *w;                    // [B,V] edge weight to the parent, position order (w of the root counts as 0)
    const float *src;            // [B,C,V] vertex order: features (forward) / grad_out (backward)
    float *up, *agg;             // [B,NC,V] position order: leaf->root and root->leaf results (forward: saved; backward: scratch)
    float *out;                  // [B,C,V] vertex order: filter output (forward) / feature gradient (backward)
    const float *up_f, *agg_f;   // backward: what forward saved
    float *gw;                   // backward: [B,V] edge-weight gradient, position order (may be null)
    int *sync;                   // [SYNC_HEAD] counters, all bits set at launch
    float *hand;                 // [2,B,Kmax,8] value a tile hands to its neighbour in the same group (PENDING at launch)
    int B, V, Kmax, vec;         // vec: 16-byte vector access allowed (V % 4 == 0, aligned bases)
    int prof;
};

// relaxed gpu-scope accesses: "strong" operations in the PTX memory model (a relaxed load that polls a location
// written by a relaxed store of another CTA is not a data race), served by L2
__device__ __forceinline__ float ld_strong(const float *p) {
    float v;
    asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ld_strong4(const float *p) {
    float4 v;
    asm volatile("ld.relaxed.gpu.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_strong(float *p, float v) {
    asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(p), "f"(v));
}
__device__ __forceinline__ void st_strong4(float *p, float4 v) {
    asm volatile("st.relaxed.gpu.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
}
__device__ __forceinline__ bool pending(float v) { return __float_as_uint(v) == PENDING; }
// The polling loads are volatile asm statements without a memory clobber, so the compiler may sink an ordinary load
// whose value is first used after a polling loop to below that loop -- and its latency would then sit on the critical
// path.  keep(x) pins the load that defines x ahead of the next volatile asm (no instruction is emitted).
__device__ __forceinline__ void keep(float x) { asm volatile("" ::"f"(x)); }
__device__ __forceinline__ void keep(int x) { asm volatile("" ::"r"(x)); }

// Spin control: every item waits only for items with smaller tickets, which are running, so a wait ends within
// microseconds; the bound (2e10 cycles, about ten seconds -- generous, because wall time also passes while the context is
// time-sliced with other processes or stopped under a debugger) turns a corrupted buffer into a counted error instead of
// a hang.  A feature value with the bit pattern 0xFFFFFFFF (a NaN that arithmetic never produces but passes on) would
// look "pending" for ever and end the same way.
struct Spin {
    long long t0 = 0;
    __device__ __forceinline__ bool again(int *sync) {
        __nanosleep(20);
        if (t0 == 0) { t0 = clock64(); return true; }
        if (clock64() - t0 > 20000000000ll) { atomicAdd(sync + 1, 1); return false; }
        return true;
    }
};

template <typename T>
__device__ __forceinline__ void load4(const T *base, int q, int V, bool vec, T (&v)[PPT], T fill) {
    if (vec) {
        if (q < V) {
            const int4 t = *reinterpret_cast<const int4 *>(base + q);
            v[0] = *reinterpret_cast<const T *>(&t.x); v[1] = *reinterpret_cast<const T *>(&t.y);
            v[2] = *reinterpret_cast<const T *>(&t.z); v[3] = *reinterpret_cast<const T *>(&t.w);
        } else {
            v[0] = v[1] = v[2] = v[3] = fill;
        }
    } else {
#pragma unroll
        for (int j = 0; j < PPT; ++j) v[j] = (q + j < V) ? base[q + j] : fill;
    }
}
// intermediate results: strong stores (other CTAs of this launch poll them)
__device__ __forceinline__ void store4_strong(float *base, int q, bool vec, const float (&v)[PPT], const bool (&ok)[PPT]) {
    if (vec && ok[0] && ok[1] && ok[2] && ok[3]) {
        st_strong4(base + q, make_float4(v[0], v[1], v[2], v[3]));
    } else {
#pragma unroll
        for (int j = 0; j < PPT; ++j)
            if (ok[j]) st_strong(base + q + j, v[j]);
    }
}
__device__ __forceinline__ void store4(float *base, int q, bool vec, const float (&v)[PPT], const bool (&ok)[PPT]) {
    if (vec && ok[0] && ok[1] && ok[2] && ok[3]) {
        *reinterpret_cast<float4 *>(base + q) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int j = 0; j < PPT; ++j)
            if (ok[j]) base[q + j] = v[j];
    }
}

struct Item {
    int b, r, T, id, lo, hi, gs, ge;
};

// FEDICRA_PROF=1: %globaltimer stamps of every item of the first images (claimed, inputs loaded, dependencies met,
// results stored) -- read back with fedicra_debug_chain_profile
constexpr int PROF_B = 2, PROF_K = 768;
__device__ long long g_chain_prof[2 * PROF_B * PROF_K * 4];
__device__ __forceinline__ long long globaltimer() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void stamp(const ChainArgs &a, int phase, const Item &it, int slot) {
    // thread 0 only: coarse, and the divergent branch perturbs the warp it sits in (sub-microsecond figures taken with
    // FEDICRA_PROF=1 are upper bounds); with profiling off the branch is uniform and free
    if (a.prof && threadIdx.x == 0 && it.b < PROF_B && it.id < PROF_K)
        g_chain_prof[((phase * PROF_B + it.b) * PROF_K + it.id) * 4 + slot] = globaltimer();
}

// NC self-validating words p[0], p[stride], ...: re-read until all of them have arrived
template <int NC>
__device__ __forceinline__ void poll_words(const float *p, size_t stride, float (&v)[NC], int *sync) {
    Spin spin;
    while (true) {
        bool all = true;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            v[c] = ld_strong(p + c * stride);
            all = all && !pending(v[c]);
        }
        if (all || !spin.again(sync)) return;
    }
}

template <int MODE, int NC>
__device__ __forceinline__ void up_item(const ChainArgs &a, const Item &it, float (*s_agg)[NC + 1]) {
    constexpr int C = NC - 1;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, V = a.V, b = it.b;
    const bool vec = a.vec != 0;
    const size_t ob = (size_t)b * V;
    const int q0 = it.T * TILE + tid * PPT;
    float *upb = a.up + (size_t)b * NC * V;
    bool ok[PPT];
    int vt[PPT];
#pragma unroll
    for (int j = 0; j < PPT; ++j) ok[j] = q0 + j >= it.lo && q0 + j < it.hi;
    load4(a.vtx + ob, q0, V, vec, vt, 0);
    float av[PPT], bq[PPT][NC];
    int lc[PPT][2];          // first two light children; a third exists only below a root with four neighbours
    bool more = false;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        const int4 ch = ok[j] ? a.child[ob + q0 + j] : make_int4(0, 0, 0, 0);
        lc[j][0] = ch.y; lc[j][1] = ch.z;
        more = more || ch.w > 0;
        av[j] = ok[j] ? ((ch.x != 0) ? a.w[ob + q0 + j + 1] : 0.f) : 1.f;   // outside the item: identity map
        if (!ok[j]) {
#pragma unroll
            for (int c = 0; c < NC; ++c) bq[j][c] = 0.f;
        } else if (MODE == 0) {
#pragma unroll
            for (int c = 0; c < C; ++c) bq[j][c] = a.src[((size_t)b * C + c) * V + vt[j]];
            bq[j][C] = 1.f;
        } else {   // grad_out / Z per channel, and sum_c (grad_out_c / Z) out_c  (refine.cu:347-348, summed over c)
            const float z = a.agg_f[((size_t)b * NC + C) * V + q0 + j], rz = 1.f / z;
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const float gn = a.src[((size_t)b * C + c) * V + vt[j]] * rz;
                bq[j][c] = gn;
                s += gn * (a.agg_f[((size_t)b * NC + c) * V + q0 + j] * rz);
            }
            bq[j][C] = s;
        }
    }
    const bool cross = it.hi == (it.T + 1) * TILE && it.hi < it.ge && a.child[ob + it.hi - 1].x != 0;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        keep(av[j]);
#pragma unroll
        for (int c = 0; c < NC; ++c) keep(bq[j][c]);
    }
    keep((int)cross);
    stamp(a, 0, it, 1);
    // light children (chain tops of the next group): re-read their values until all have arrived, then add w_c x_c in
    // a fixed order: slot by slot, the four positions of a slot together.
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        bool any = false;
#pragma unroll
        for (int j = 0; j < PPT; ++j) any = any || lc[j][s] > 0;
        if (s > 0 && !__any_sync(0xffffffffu, any)) break;
        float wl[PPT], xv[PPT][NC];
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
            wl[j] = lc[j][s] > 0 ? a.w[ob + lc[j][s]] : 0.f;
            keep(wl[j]);
        }
        Spin spin;
        while (true) {
            bool all = true;
#pragma unroll
            for (int j = 0; j < PPT; ++j)
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    xv[j][c] = lc[j][s] > 0 ? ld_strong(upb + (size_t)c * V + lc[j][s]) : 0.f;
                    all = all && !pending(xv[j][c]);
                }
            if (all || !spin.again(a.sync)) break;
        }
#pragma unroll
        for (int j = 0; j < PPT; ++j)
#pragma unroll
            for (int c = 0; c < NC; ++c) bq[j][c] = fmaf(wl[j], xv[j][c], bq[j][c]);
    }
    if (more) {
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
            const int l3 = ok[j] ? a.child[ob + q0 + j].w : 0;
            if (l3 > 0) {
                float xv[NC];
                poll_words<NC>(upb + l3, (size_t)V, xv, a.sync);
                const float wl = a.w[ob + l3];
#pragma unroll
                for (int c = 0; c < NC; ++c) bq[j][c] = fmaf(wl, xv[c], bq[j][c]);
            }
        }
    }
    stamp(a, 0, it, 2);
    // this thread's four positions as one map x_{q0} = X + P x_{q0+4}
    float P = 1.f, X[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) X[c] = 0.f;
#pragma unroll
    for (int j = PPT - 1; j >= 0; --j) {
#pragma unroll
        for (int c = 0; c < NC; ++c) X[c] = fmaf(av[j], X[c], bq[j][c]);
        P *= av[j];
    }
    // suffix scan over the warp, then over the warps
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float Pn = __shfl_down_sync(0xffffffffu, P, o);
        float Xn[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) Xn[c] = __shfl_down_sync(0xffffffffu, X[c], o);
        if (lane + o < 32) {
#pragma unroll
            for (int c = 0; c < NC; ++c) X[c] = fmaf(P, Xn[c], X[c]);
            P *= Pn;
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < NC; ++c) s_agg[wid][c] = X[c];
        s_agg[wid][NC] = P;
    }
    __syncthreads();
    // The tile as one map: value at its first position = Xt + Pt * (value handed in at its end).  Pt is 0 unless one
    // chain spans the whole tile, so the value the PREVIOUS tile of this group waits for is usually known right here.
    float cw[NC], Pt = 1.f, Xt[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) Xt[c] = 0.f;
    for (int w2 = NW - 1; w2 >= 0; --w2) {
        const float Pw = s_agg[w2][NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) Xt[c] = fmaf(Pw, Xt[c], s_agg[w2][c]);
        Pt *= Pw;
    }
    float *head_val = a.hand + ((size_t)b * a.Kmax + it.id) * 8;
    const bool early = Pt == 0.f || !cross;
    if (early && tid == 0) {
#pragma unroll
        for (int c = 0; c < NC; ++c) st_strong(head_val + c, Xt[c]);
    }
    // What flows into this lane's last position is base + coef * (value handed in at the tile's end); coef is 0 for
    // every lane behind the first chain bottom, i.e. for most: those store at once, the others wait for the hand-over.
    float coef = 1.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) cw[c] = 0.f;
    for (int w2 = NW - 1; w2 > wid; --w2) {
        const float Pw = s_agg[w2][NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) cw[c] = fmaf(Pw, cw[c], s_agg[w2][c]);
        coef *= Pw;
    }
    {
        const float Pn = __shfl_down_sync(0xffffffffu, P, 1);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const float Xn = __shfl_down_sync(0xffffffffu, X[c], 1);
            if (lane < 31) cw[c] = fmaf(Pn, cw[c], Xn);
        }
        if (lane < 31) coef *= Pn;
    }
    auto finish = [&](float (&x)[NC]) {
        float res[NC][PPT];
#pragma unroll
        for (int j = PPT - 1; j >= 0; --j) {
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                x[c] = fmaf(av[j], x[c], bq[j][c]);
                res[c][j] = x[c];
            }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) store4_strong(upb + (size_t)c * V, q0, vec, res[c], ok);
    };
    const bool waits = cross && coef != 0.f;
    if (!waits) finish(cw);
    if (waits) {
        float tc[NC];
        poll_words<NC>(head_val + 8, 1, tc, a.sync);      // from the next tile of this group
        if (!early && tid == 0) {
#pragma unroll
            for (int c = 0; c < NC; ++c) st_strong(head_val + c, fmaf(Pt, tc[c], Xt[c]));
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) cw[c] = fmaf(coef, tc[c], cw[c]);
        finish(cw);
    }
    stamp(a, 0, it, 3);
}

template <int MODE, int NC>
__device__ __forceinline__ void down_item(const ChainArgs &a, const Item &it, float (*s_agg)[NC + 1]) {
    constexpr int C = NC - 1;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, V = a.V, b = it.b;
    const bool vec = a.vec != 0;
    const size_t ob = (size_t)b * V;
    const int q0 = it.T * TILE + tid * PPT;
    float *upb = a.up + (size_t)b * NC * V, *aggb = a.agg + (size_t)b * NC * V;
    bool ok[PPT], top[PPT], far[PPT];
    int vt[PPT], pr[PPT];
    float wq[PPT];
#pragma unroll
    for (int j = 0; j < PPT; ++j) ok[j] = q0 + j >= it.lo && q0 + j < it.hi;
    load4(a.vtx + ob, q0, V, vec, vt, 0);
    load4(a.par + ob, q0, V, vec, pr, 0);
    load4(a.w + ob, q0, V, vec, wq, 0.f);
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        if (q0 + j == 0) wq[j] = 0.f;                          // the root has no parent edge (refine.cu:43-46)
        top[j] = q0 + j == 0 || pr[j] != q0 + j - 1;
        far[j] = ok[j] && top[j] && q0 + j != 0;               // its parent sits in the previous group
    }
    if (it.lo == 0 && tid == 0) a.w[ob] = 0.f;                 // and the reference says so in place
    const bool cross = it.lo == it.T * TILE && it.lo > it.gs && a.par[ob + it.lo] == it.lo - 1;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        keep(wq[j]);
        keep(vt[j]);
        keep(pr[j]);
    }
    keep((int)cross);
    stamp(a, 1, it, 1);
    // this tile's leaf->root values (written long ago, as a rule) ...
    float u[NC][PPT], cq[PPT][NC];
    {
        Spin spin;
        while (true) {
            bool all = true;
            if (vec) {
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const float4 t = q0 < V ? ld_strong4(upb + (size_t)c * V + q0) : make_float4(0.f, 0.f, 0.f, 0.f);
                    u[c][0] = t.x; u[c][1] = t.y; u[c][2] = t.z; u[c][3] = t.w;
                }
            } else {
#pragma unroll
                for (int c = 0; c < NC; ++c)
#pragma unroll
                    for (int j = 0; j < PPT; ++j) u[c][j] = ok[j] ? ld_strong(upb + (size_t)c * V + q0 + j) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < PPT; ++j)
#pragma unroll
                for (int c = 0; c < NC; ++c) all = all && !(ok[j] && pending(u[c][j]));
            if (all || !spin.again(a.sync)) break;
        }
    }
    // ... and the root->leaf values of the parents of chain tops (previous group): re-read until all have arrived
    {
        Spin spin;
        while (true) {
            bool all = true;
#pragma unroll
            for (int j = 0; j < PPT; ++j)
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    cq[j][c] = far[j] ? ld_strong(aggb + (size_t)c * V + pr[j]) : 0.f;
                    all = all && !pending(cq[j][c]);
                }
            if (all || !spin.again(a.sync)) break;
        }
    }
    stamp(a, 1, it, 2);
    float av[PPT];
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        const float w = wq[j], k = 1.f - w * w;
        av[j] = ok[j] ? (top[j] ? 0.f : w) : 1.f;
#pragma unroll
        for (int c = 0; c < NC; ++c) cq[j][c] = ok[j] ? fmaf(w, cq[j][c], u[c][j] * k) : 0.f;
    }
    float P = 1.f, X[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) X[c] = 0.f;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
#pragma unroll
        for (int c = 0; c < NC; ++c) X[c] = fmaf(av[j], X[c], cq[j][c]);
        P *= av[j];
    }
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float Pn = __shfl_up_sync(0xffffffffu, P, o);
        float Xn[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) Xn[c] = __shfl_up_sync(0xffffffffu, X[c], o);
        if (lane >= o) {
#pragma unroll
            for (int c = 0; c < NC; ++c) X[c] = fmaf(P, Xn[c], X[c]);
            P *= Pn;
        }
    }
    if (lane == 31) {
#pragma unroll
        for (int c = 0; c < NC; ++c) s_agg[wid][c] = X[c];
        s_agg[wid][NC] = P;
    }
    __syncthreads();
    // as in up_item: the value at this tile's last position = Xt + Pt * (value handed in from the previous tile)
    float cw[NC], Pt = 1.f, Xt[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) Xt[c] = 0.f;
    for (int w2 = 0; w2 < NW; ++w2) {
        const float Pw = s_agg[w2][NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) Xt[c] = fmaf(Pw, Xt[c], s_agg[w2][c]);
        Pt *= Pw;
    }
    float *tail_val = a.hand + ((size_t)(a.B + b) * a.Kmax + it.id) * 8;
    const bool early = Pt == 0.f || !cross;
    if (early && tid == 0) {
#pragma unroll
        for (int c = 0; c < NC; ++c) st_strong(tail_val + c, Xt[c]);
    }
    float coef = 1.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) cw[c] = 0.f;
    for (int w2 = 0; w2 < wid; ++w2) {
        const float Pw = s_agg[w2][NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) cw[c] = fmaf(Pw, cw[c], s_agg[w2][c]);
        coef *= Pw;
    }
    {
        const float Pn = __shfl_up_sync(0xffffffffu, P, 1);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const float Xn = __shfl_up_sync(0xffffffffu, X[c], 1);
            if (lane > 0) cw[c] = fmaf(Pn, cw[c], Xn);
        }
        if (lane > 0) coef *= Pn;
    }
    float res[NC][PPT], cw0[NC];
    auto finish = [&](float (&x)[NC]) {
#pragma unroll
        for (int c = 0; c < NC; ++c) cw0[c] = x[c];
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                x[c] = fmaf(av[j], x[c], cq[j][c]);
                res[c][j] = x[c];
            }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) store4_strong(aggb + (size_t)c * V, q0, vec, res[c], ok);
    };
    const bool waits = cross && coef != 0.f;
    if (!waits) finish(cw);
    if (waits) {
        float tc[NC];
        poll_words<NC>(tail_val - 8, 1, tc, a.sync);       // from the previous tile of this group
        if (!early && tid == NW * 32 - 1) {
#pragma unroll
            for (int c = 0; c < NC; ++c) st_strong(tail_val + c, fmaf(Pt, tc[c], Xt[c]));
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) cw[c] = fmaf(coef, tc[c], cw[c]);
        finish(cw);
    }
    stamp(a, 1, it, 3);
    // nobody inside this launch waits for what follows
    if (MODE == 1 && a.gw != nullptr) {
        // refine.cu:136-199 with both passes folded: U = leaf->root of this call, G = its root->leaf form at the
        // parent, (upf, A) = what forward saved here and at the parent
        float gwv[PPT];
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
            float acc = 0.f;
            if (ok[j] && q0 + j != 0) {
                const float w = wq[j];
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const float gpar = top[j] ? ld_strong(aggb + (size_t)c * V + pr[j]) : (j > 0 ? res[c][j > 0 ? j - 1 : 0] : cw0[c]);
                    const float upf = a.up_f[((size_t)b * NC + c) * V + q0 + j];
                    const float apar = a.agg_f[((size_t)b * NC + c) * V + pr[j]];
                    const float t = u[c][j] * (apar - w * upf) + upf * (gpar - w * u[c][j]);
                    acc += (c < C) ? t : -t;
                }
            }
            gwv[j] = acc;
        }
        store4(a.gw + ob, q0, vec, gwv, ok);
    }
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        if (!ok[j]) continue;
        if (MODE == 0) {
            const float rz = 1.f / res[C][j];
#pragma unroll
            for (int c = 0; c < C; ++c) a.out[((size_t)b * C + c) * V + vt[j]] = res[c][j] * rz;
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) a.out[((size_t)b * C + c) * V + vt[j]] = res[c][j];
        }
    }
}

template <int MODE, int NC>
__global__ void __launch_bounds__(CT, CHAIN_MINB) chain_filter_kernel(ChainArgs a) {
    __shared__ int s_gs[33], s_cum[33], s_ticket;
    __shared__ float s_agg[NW][NC + 1];
    const int tid = threadIdx.x, lane = tid & 31;
    const int per_phase = a.B * a.Kmax;
    while (true) {
        __syncthreads();
        if (tid == 0) s_ticket = atomicAdd(a.sync, 1) + 1;     // the counter starts at -1
        __syncthreads();
        const int t = s_ticket;
        if (t >= 2 * per_phase) break;
        const int phase = t / per_phase, rem = t - phase * per_phase, k = rem / a.B;
        Item it;
        it.b = rem - k * a.B;
        if (tid < 32) {   // tiles per group and their running count: the dense numbering of this image's items
            const int g0 = a.gstart[it.b * 33 + lane], g1 = a.gstart[it.b * 33 + lane + 1];
            const int n = g1 > g0 ? (g1 - 1) / TILE - g0 / TILE + 1 : 0;
            int x = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += y;
            }
            s_gs[lane] = g0;
            s_cum[lane] = x - n;
            if (lane == 31) { s_gs[32] = g1; s_cum[32] = x; }
        }
        __syncthreads();
        const int total = s_cum[32];
        if (k >= total) continue;
        it.id = phase == 0 ? total - 1 - k : k;
        it.r = 0;
        for (int g = 1; g < 32; ++g)
            if (s_cum[g] <= it.id) it.r = g;
        it.gs = s_gs[it.r];
        it.ge = s_gs[it.r + 1];
        it.T = it.gs / TILE + (it.id - s_cum[it.r]);
        it.lo = max(it.gs, it.T * TILE);
        it.hi = min(it.ge, (it.T + 1) * TILE);
        stamp(a, phase, it, 0);
        if (phase == 0) up_item<MODE, NC>(a, it, s_agg);
        else down_item<MODE, NC>(a, it, s_agg);
    }
}

int items_bound(int V) { return (V + TILE - 1) / TILE + 33; }

// counters + hand-over values: everything that must read "all bits set" when the kernel starts
size_t sync_bytes(int B, int V) { return align_up((size_t)(SYNC_HEAD + 2 * (size_t)B * items_bound(V) * 8) * sizeof(int), 256); }

template <int MODE, int NC>
int launch(const ChainArgs &a, cudaStream_t st) {
    static int ctas_per_sm = 0, sms = 0;   // per process; all devices of a box are the same chip
    if (ctas_per_sm == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, chain_filter_kernel<MODE, NC>, CT, 0);
        if (ctas_per_sm < 1) ctas_per_sm = 1;
    }
    const long long items = 2ll * a.B * a.Kmax;
    int per_sm = ctas_per_sm;
    if (const char *v = getenv("FEDICRA_CHAIN_CTAS")) {   // resident CTAs per SM (default: all that fit)
        const int x = atoi(v);
        if (x >= 1 && x < per_sm) per_sm = x;
    }
    const int grid = (int)(items < (long long)sms * per_sm ? items : (long long)sms * per_sm);
    chain_filter_kernel<MODE, NC><<<grid, CT, 0, st>>>(a);
    FEDICRA_CHECK_LAUNCH(MODE == 0 ? "chain_filter_kernel<forward>" : "chain_filter_kernel<backward>");
    return 0;
}

template <int MODE>
int dispatch(const ChainArgs &a, int C, cudaStream_t st) {
    switch (C) {
        case 1: return launch<MODE, 2>(a, st);
        case 2: return launch<MODE, 3>(a, st);
        case 3: return launch<MODE, 4>(a, st);
        case 4: return launch<MODE, 5>(a, st);
    }
    return fedicra_fail(-1, "chain_filter: %d channels (1..4 are built)", C);
}

bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

int prof_enabled() {
    const char *v = getenv("FEDICRA_PROF");
    return (v && *v && *v != '0') ? 1 : 0;
}

}  // namespace

extern "C" int fedicra_chain_filter_max_channels(void) { return 4; }

extern "C" size_t fedicra_chain_filter_saved_offset(int B, int C, int V) {
    if (B <= 0 || C <= 0 || V <= 0) return 0;
    return sync_bytes(B, V);
}

extern "C" size_t fedicra_chain_filter_saved_bytes(int B, int C, int V) {
    if (B <= 0 || C <= 0 || V <= 0) return 0;
    return sync_bytes(B, V) + (size_t)2 * B * (C + 1) * V * sizeof(float);
}

extern "C" size_t fedicra_chain_filter_workspace_bytes(int B, int C, int V) {
    if (B <= 0 || C <= 0 || V <= 0) return 0;
    return sync_bytes(B, V) + (size_t)2 * B * (C + 1) * V * sizeof(float);
}

namespace {
int chain_forward_impl(const float *feature_in, float *edge_weight, const int *sorted_index, const int *sorted_parent,
                       const int *sorted_child, const int *group_start, int B, int C, int V, float *feature_out, void *saved,
                       size_t saved_bytes, void *stream, bool armed) {
    FEDICRA_REQUIRE(feature_in && edge_weight && sorted_index && sorted_parent && sorted_child && group_start &&
                        feature_out && saved, "chain_filter_forward: null pointer");
    FEDICRA_REQUIRE(B > 0 && C >= 1 && C <= 4 && V >= 1, "chain_filter_forward: bad sizes B=%d C=%d V=%d (1..4 channels)", B, C, V);
    FEDICRA_REQUIRE(saved_bytes >= fedicra_chain_filter_saved_bytes(B, C, V), "chain_filter_forward: saved buffer too small");
    FEDICRA_REQUIRE(aligned16(sorted_child) && aligned16(saved), "chain_filter_forward: sorted_child and saved must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    ChainArgs a{};
    a.vtx = sorted_index; a.par = sorted_parent; a.child = reinterpret_cast<const int4 *>(sorted_child);
    a.gstart = group_start; a.w = edge_weight; a.src = feature_in;
    a.sync = (int *)saved; a.hand = (float *)saved + SYNC_HEAD;
    a.up = (float *)((char *)saved + sync_bytes(B, V)); a.agg = a.up + (size_t)B * (C + 1) * V;
    a.out = feature_out;
    a.B = B; a.V = V; a.Kmax = items_bound(V); a.prof = prof_enabled();
    a.vec = (V % 4 == 0) && aligned16(sorted_index) && aligned16(sorted_parent) && aligned16(edge_weight);
    // every intermediate word starts as "pending" (all bits set): the counters and hand-over values at the head of
    // `saved`, and the two result arrays behind them -- one range, one memset
    if (!armed) {
        cudaError_t e = cudaMemsetAsync(saved, 0xff, fedicra_chain_filter_saved_bytes(B, C, V), st);
        if (e != cudaSuccess) return fedicra_fail(-2, "chain_filter_forward: memset: %s", cudaGetErrorString(e));
    }
    return dispatch<0>(a, C, st);
}
}  // namespace

extern "C" int fedicra_chain_filter_forward(const float *feature_in, float *edge_weight, const int *sorted_index,
                                            const int *sorted_parent, const int *sorted_child, const int *group_start,
                                            int B, int C, int V, float *feature_out, void *saved, size_t saved_bytes,
                                            void *stream) {
    return chain_forward_impl(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, group_start, B, C, V,
                              feature_out, saved, saved_bytes, stream, false);
}

extern "C" int fedicra_chain_filter_forward_armed(const float *feature_in, float *edge_weight, const int *sorted_index,
                                                  const int *sorted_parent, const int *sorted_child, const int *group_start,
                                                  int B, int C, int V, float *feature_out, void *saved, size_t saved_bytes,
                                                  void *stream) {
    return chain_forward_impl(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, group_start, B, C, V,
                              feature_out, saved, saved_bytes, stream, true);
}

extern "C" int fedicra_chain_filter_arm(void *buffers, size_t bytes, void *stream) {
    FEDICRA_REQUIRE(buffers != nullptr && bytes > 0, "chain_filter_arm: null pointer or no bytes");
    cudaError_t e = cudaMemsetAsync(buffers, 0xff, bytes, (cudaStream_t)stream);
    if (e != cudaSuccess) return fedicra_fail(-2, "chain_filter_arm: memset: %s", cudaGetErrorString(e));
    return 0;
}

namespace {
int chain_backward_impl(const float *grad_out, float *edge_weight, const int *sorted_index, const int *sorted_parent,
                        const int *sorted_child, const int *group_start, int B, int C, int V, const void *saved,
                        float *grad_feature, float *grad_weight, void *workspace, size_t workspace_bytes, void *stream,
                        bool armed) {
    FEDICRA_REQUIRE(grad_out && edge_weight && sorted_index && sorted_parent && sorted_child && group_start && saved &&
                        grad_feature && workspace, "chain_filter_backward: null pointer");
    FEDICRA_REQUIRE(B > 0 && C >= 1 && C <= 4 && V >= 1, "chain_filter_backward: bad sizes B=%d C=%d V=%d (1..4 channels)", B, C, V);
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_chain_filter_workspace_bytes(B, C, V), "chain_filter_backward: workspace too small");
    FEDICRA_REQUIRE(aligned16(sorted_child), "chain_filter_backward: sorted_child must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    ChainArgs a{};
    a.vtx = sorted_index; a.par = sorted_parent; a.child = reinterpret_cast<const int4 *>(sorted_child);
    a.gstart = group_start; a.w = edge_weight; a.src = grad_out;
    a.up_f = (const float *)((const char *)saved + sync_bytes(B, V)); a.agg_f = a.up_f + (size_t)B * (C + 1) * V;
    a.sync = (int *)workspace; a.hand = (float *)workspace + SYNC_HEAD;
    a.up = (float *)((char *)workspace + sync_bytes(B, V)); a.agg = a.up + (size_t)B * (C + 1) * V;
    a.out = grad_feature; a.gw = grad_weight;
    a.B = B; a.V = V; a.Kmax = items_bound(V); a.prof = prof_enabled();
    a.vec = (V % 4 == 0) && aligned16(sorted_index) && aligned16(sorted_parent) && aligned16(edge_weight) &&
            aligned16(saved) && aligned16(a.up) && (grad_weight == nullptr || aligned16(grad_weight));
    // counters, hand-over values and both scratch arrays are one range: one memset marks them all "pending"
    if (!armed) {
        cudaError_t e = cudaMemsetAsync(workspace, 0xff, fedicra_chain_filter_workspace_bytes(B, C, V), st);
        if (e != cudaSuccess) return fedicra_fail(-2, "chain_filter_backward: memset: %s", cudaGetErrorString(e));
    }
    return dispatch<1>(a, C, st);
}
}  // namespace

extern "C" int fedicra_chain_filter_backward(const float *grad_out, float *edge_weight, const int *sorted_index,
                                             const int *sorted_parent, const int *sorted_child, const int *group_start,
                                             int B, int C, int V, const void *saved, float *grad_feature,
                                             float *grad_weight, void *workspace, size_t workspace_bytes, void *stream) {
    return chain_backward_impl(grad_out, edge_weight, sorted_index, sorted_parent, sorted_child, group_start, B, C, V, saved,
                               grad_feature, grad_weight, workspace, workspace_bytes, stream, false);
}

extern "C" int fedicra_chain_filter_backward_armed(const float *grad_out, float *edge_weight, const int *sorted_index,
                                                   const int *sorted_parent, const int *sorted_child, const int *group_start,
                                                   int B, int C, int V, const void *saved, float *grad_feature,
                                                   float *grad_weight, void *workspace, size_t workspace_bytes, void *stream) {
    return chain_backward_impl(grad_out, edge_weight, sorted_index, sorted_parent, sorted_child, group_start, B, C, V, saved,
                               grad_feature, grad_weight, workspace, workspace_bytes, stream, true);
}

/* Debug: %globaltimer stamps of the last profiled chain-filter launch (FEDICRA_PROF=1), [2 phases][2 images][768 items][4]:
 * claimed, inputs loaded, dependencies met, results stored.  Synchronises the device. */
extern "C" int fedicra_debug_chain_profile(long long *h_out, int n) {
    const int total = 2 * PROF_B * PROF_K * 4;
    if (!h_out || n < total) return fedicra_fail(-1, "debug_chain_profile: need room for %d values", total);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(h_out, g_chain_prof, sizeof(long long) * total);
    if (e != cudaSuccess) return fedicra_fail(-2, "debug_chain_profile: %s", cudaGetErrorString(e));
    return total;
}
