// bands.cu -- depth-banded tree filter: the two-pass filter of refine.cu without its ~1 500-step dependency chain.
//
// What it replaces (LTF = /root/reference/code/utils/TreeEnergyLoss/kernels/lib_tree_filter):
//   refine_forward / refine_backward_feature / refine_backward_weight    LTF/src/refine/refine.cu:201-370
//   (leaf_root_aggr_kernel :70-134, root_leaf_prop_kernel :19-68, root_leaf_grad_kernel :136-199)
// The recurrences are the reference's, in BFS-position space (i = position, p(i) = parent position, w_0 := 0):
//   up[i]  = x[i] + sum_{j child of i} w[j] * up[j]                  (leaf -> root)
//   out[i] = (1 - w[i]^2) * up[i] + w[i] * out[p(i)]                 (root -> leaf)
// A level-synchronous sweep needs one step per tree depth (1 200-1 600 at 256^2, ~3 000 at 512^2).  Here the depths are
// cut into `nb` bands of K consecutive levels and every sweep becomes three short phases:
//   up:    U1  all bands in parallel: local sweep with nothing entering from below  -> partial up at the band's first level
//          U2  one CTA per image, deepest band first: the sum X that enters each band's last level from the band below,
//              X'[j'] = base[j'] + sum_{j in last level of the band below, entry(j) = j'} A[j] * X[j]   (a segmented sum over
//              contiguous runs, done with a warp segmented scan; base = the children's partial sums, written by U1)
//          U3  all bands in parallel: local sweep again with X injected -> exact up everywhere
//   down:  D1  (fused behind U3, same CTA) local top-down pass: out[i] = A[i] * out[entry(i)] + B[i] with entry(i) the
//              parent of i's first-level ancestor, A[i] the product of w from i up to and including that ancestor
//          D2  one CTA per image, shallowest band first: out at each band's last level
//          D3  element-wise: out[i] = A[i] * out[entry(i)] + B[i], fused with the consumer (normalise + scatter, or the
//              edge-weight gradient)
// Dependency chain: 3K + 2nb short steps instead of 2 * depth.  Band state lives in shared memory (a band that does not
// fit falls back to the same code on global memory).  All arithmetic is fp32, deterministic, no atomics.
#include "common.cuh"

namespace {

constexpr int NT = 256;
constexpr int LSCAP = 512;   // level_start entries of a band cached in shared memory

struct TreeView {
    const int *vtx;     // [B,V]   sorted_index
    const int *ppos;    // [B,V]   sorted_parent
    const int4 *child;  // [B,V]   sorted_child (positions, contiguous, 0 = none)
    const int *ls;      // [B,V+1] level_start
    const int *nl;      // [B]
    const float *w;     // [B,V]   edge weight in BFS order
    const int *bands;   // [B,MAXBANDS+2] band plan: [0] = number of bands nb, [1+k] = first level of band k, [1+nb] = n_levels
    int V;
};

constexpr int MAXBANDS = 256;
constexpr int PLAN_STRIDE = MAXBANDS + 2;
constexpr int NBCAP = 36;        // most bands per image the planner produces (the resident kernel's grid is NBCAP + 1 wide)

struct Consts {  // per (tree, edge weights), all [B,V] in position order
    float *A;           // product of w from i up to and including its band's first-level ancestor
    int *ent;           // entry(i): position of the parent of that ancestor (in the last level of the band above)
    int *run_s, *run_e; // for j' in a band's last level: the run of positions j in the NEXT band's last level with entry(j) = j'
    int *pk;            // packed band-local topology: parent << 16 | first child << 3 | #children (offsets from the band's p0)
};

__device__ __forceinline__ bool band_geom(const TreeView &t, int b, int beta, int &L0, int &L1, int &nb) {
    const int *plan = t.bands + (size_t)b * PLAN_STRIDE;
    nb = plan[0];
    if (beta >= nb) return false;
    L0 = plan[1 + beta];
    L1 = plan[2 + beta];
    return true;
}

// Band plan: consecutive levels are grouped greedily so that a band holds at most `pmax` positions (it must fit the
// shared-memory band state) and at most ceil(n_levels / kdiv) levels (the length of the parallel phases).  If that needs
// more than `nbcap` bands the level limit grows by a quarter and the plan is redone; after six such rounds the position
// limit grows too (oversized bands then run on global memory).  One thread per image, level starts cached in shared
// memory.  (Cutting at locally narrow levels was tried: level widths vary too smoothly for it to shrink the stitched
// state.)
constexpr int PLAN_LSCAP = 8192;  // level starts cached in shared memory by the planner

__global__ void __launch_bounds__(256)
band_plan_kernel(const int *__restrict__ level_start, const int *__restrict__ n_levels, int V, int pmax, int kdiv,
                 int nbcap, int *__restrict__ bands) {
    __shared__ int s_ls[PLAN_LSCAP + 1];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int *gls = level_start + (size_t)b * (V + 1);
    int *plan = bands + (size_t)b * PLAN_STRIDE;
    const int nl = n_levels[b];
    const bool cached = nl <= PLAN_LSCAP;
    if (cached)
        for (int L = tid; L <= nl; L += 256) s_ls[L] = gls[L];
    __syncthreads();
    if (tid != 0) return;
    const int *ls = cached ? s_ls : gls;
    int kmax = max(4, (nl + kdiv - 1) / kdiv), escalations = 0;
    while (true) {
        // the cut points are sequential, but each one is a binary search: the band starting at level L0 ends at the
        // largest L1 with L1 - L0 <= kmax and ls[L1] - ls[L0] <= pmax (at least one level)
        int nb = 0, L0 = 0;
        bool overflow = false;
        while (L0 < nl) {
            if (nb >= nbcap) { overflow = true; break; }
            plan[1 + nb] = L0;
            const int p0 = ls[L0];
            int lo = L0 + 1, hi = min(L0 + kmax, nl);
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (ls[mid] - p0 <= pmax) lo = mid; else hi = mid - 1;
            }
            L0 = lo;
            ++nb;
        }
        if (!overflow) {
            plan[0] = nb;
            plan[1 + nb] = nl;
            return;
        }
        kmax += kmax / 4 + 1;
        if (++escalations > 6) pmax += pmax / 4;
    }
}

__device__ __forceinline__ int child_count(const int4 &c) { return (c.x > 0) + (c.y > 0) + (c.z > 0) + (c.w > 0); }

// ------------------------------------------------------------------------------------------------ band constants
template <bool SMEM>
__device__ void consts_body(const TreeView &t, const Consts &k, int b, int beta, int L0, int L1, int p0, int p1, char *smem) {
    const int tid = threadIdx.x, V = t.V, P = p1 - p0;
    const int *ls = t.ls + (size_t)b * (V + 1);
    const int *pp = t.ppos + (size_t)b * V;
    const float *w = t.w + (size_t)b * V;
    float *gA = k.A + (size_t)b * V;
    int *gE = k.ent + (size_t)b * V;
    float *sA = reinterpret_cast<float *>(smem);
    int *sE = reinterpret_cast<int *>(sA + P);
    float *sW = reinterpret_cast<float *>(sE + P);
    int *sP = reinterpret_cast<int *>(sW + P);
    if (SMEM) {  // parents and weights first, in one coalesced pass, so that the level loop never waits on global memory
        for (int i = p0 + tid; i < p1; i += NT) { sW[i - p0] = (i == 0) ? 0.f : w[i]; sP[i - p0] = pp[i]; }
        __syncthreads();
    }
    auto Aat = [&](int i) -> float & { return SMEM ? sA[i - p0] : gA[i]; };
    auto Eat = [&](int i) -> int & { return SMEM ? sE[i - p0] : gE[i]; };
    int lo = ls[L0], hi = ls[L0 + 1];
    for (int L = L0; L < L1; ++L) {
        const int nhi = (L + 1 < L1) ? ls[L + 2] : hi;
        for (int i = lo + tid; i < hi; i += NT) {
            const float wi = SMEM ? sW[i - p0] : ((i == 0) ? 0.f : w[i]);
            const int p = SMEM ? sP[i - p0] : pp[i];
            if (L == L0) {
                Aat(i) = wi; Eat(i) = p;
            } else {
                Aat(i) = wi * Aat(p); Eat(i) = Eat(p);
            }
        }
        __syncthreads();
        lo = hi;
        hi = nhi;
    }
    if (SMEM)
        for (int i = p0 + tid; i < p1; i += NT) { gA[i] = sA[i - p0]; gE[i] = sE[i - p0]; }
    {   // packed topology for the resident kernel (parents above / children below the band are never followed)
        const int4 *child = t.child + (size_t)b * V;
        int *gpk = k.pk + (size_t)b * V;
        for (int i = p0 + tid; i < p1; i += NT) {
            const int4 ch = child[i];
            const int cnt = child_count(ch), par = SMEM ? sP[i - p0] : pp[i];
            const int pl = (par >= p0) ? par - p0 : 0, cl = (cnt > 0 && ch.x < p1) ? ch.x - p0 : 0;
            gpk[i] = (pl << 16) | ((cl & 0x1fff) << 3) | cnt;
        }
    }
    // runs of this band's last-level vertices under each vertex of the previous band's last level (contiguous in the
    // canonical order); written by this CTA for the level just above the band
    if (beta == 0) return;
    int *rs = k.run_s + (size_t)b * V, *re = k.run_e + (size_t)b * V;
    const int q0 = ls[L0 - 1], q1 = ls[L0], e0 = ls[L1 - 1], e1 = ls[L1];
    for (int i = q0 + tid; i < q1; i += NT) { rs[i] = 0; re[i] = 0; }
    __syncthreads();
    for (int j = e0 + tid; j < e1; j += NT) {
        const int en = Eat(j);
        if (j == e0 || Eat(j - 1) != en) rs[en] = j;
        if (j == e1 - 1 || Eat(j + 1) != en) re[en] = j + 1;
    }
}

__global__ void __launch_bounds__(NT) consts_kernel(TreeView t, Consts k, int pcap) {
    extern __shared__ float4 smem4[];
    const int b = blockIdx.y;
    int L0, L1, nb;
    if (!band_geom(t, b, blockIdx.x, L0, L1, nb)) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) const_cast<float *>(t.w)[(size_t)b * t.V] = 0.f;   // refine.cu:43-46
    const int *ls = t.ls + (size_t)b * (t.V + 1);
    const int p0 = ls[L0], p1 = ls[L1];
    if (p1 - p0 <= pcap)
        consts_body<true>(t, k, b, blockIdx.x, L0, L1, p0, p1, reinterpret_cast<char *>(smem4));
    else
        consts_body<false>(t, k, b, blockIdx.x, L0, L1, p0, p1, reinterpret_cast<char *>(smem4));
}

// ------------------------------------------------------------------------------------------------ band up / down
enum { MODE_FWD = 0, MODE_BWD = 1 };

struct SweepArgs {
    // prologue operands
    const float *feat;   // FWD: feature_in [B,C,V] vertex order.  BWD: grad_out [B,C,V] vertex order
    const float *fpos;   // BWD: forward aggregate in position order [B,C+1,V] (channel C = Z)
    int C, NC;           // feature channels; swept channels (FWD: C+1, BWD: C or 2C)
    // per-sweep arrays [B,NC,V], position order
    float *upl;          // (unused slot, kept for the workspace layout)
    float *X;            // sum entering each band's last level from below: U1 writes its base, U2 completes it, U3 injects it
    float *up;           // exact up everywhere (U3 ->)
    float *Bv;           // local down term B (D1 ->)
    float *slab;         // global stand-in for the shared-memory band state (oversized bands only)
};

template <int MODE>
__device__ __forceinline__ float prologue(const SweepArgs &a, int b, int c, int i, int v, int V) {
    if (MODE == MODE_FWD) {
        return (c < a.C) ? a.feat[((size_t)b * a.C + c) * V + v] : 1.f;
    } else {
        const int cc = (c < a.C) ? c : c - a.C;
        const float z = a.fpos[((size_t)b * (a.C + 1) + a.C) * V + i];
        float gn = a.feat[((size_t)b * a.C + cc) * V + v] / z;
        if (c >= a.C) gn *= a.fpos[((size_t)b * (a.C + 1) + cc) * V + i] / z;  // x filtered output (refine.cu:348)
        return gn;
    }
}

constexpr int SWEEP_SLOTS = 64;  // positions of one level handled per pass; a sweep CTA has 64 threads per channel

template <int MODE, bool INJECT, bool SMEM>
__device__ void sweep_body(const TreeView &t, const SweepArgs &a, int b, int beta, int nb, int L0, int L1, int p0, int p1,
                           int pcap, char *smem) {
    const int tid = threadIdx.x, nthr = blockDim.x, V = t.V, NC = a.NC, P = p1 - p0;
    const int *ls = t.ls + (size_t)b * (V + 1);
    const int *vtx = t.vtx + (size_t)b * V, *pp = t.ppos + (size_t)b * V;
    const int4 *child = t.child + (size_t)b * V;
    const float *w = t.w + (size_t)b * V;
    const size_t img = (size_t)b * NC * V;
    // band state: val[c][i], and (shared-memory mode) copies of w / packed parent+children / vertex
    float *sval = reinterpret_cast<float *>(smem);            // [NC][pcap]
    float *sw = sval + (size_t)NC * pcap;                     // [pcap]
    int *spk = reinterpret_cast<int *>(sw + pcap);            // [pcap] (parent - p0) << 15 | (first child - p0) << 3 | #children
    int *svx = spk + pcap;                                    // [pcap] vertex of the position (staging only)
    int *sls = svx + pcap;                                    // [LSCAP+1]
    float *gval = a.slab + img;
    auto VAL = [&](int c, int i) -> float & { return SMEM ? sval[(size_t)c * pcap + (i - p0)] : gval[(size_t)c * V + i]; };
    auto Wat = [&](int i) -> float { return SMEM ? sw[i - p0] : ((i == 0) ? 0.f : w[i]); };
    const bool ls_cached = SMEM && (L1 - L0) <= LSCAP;
    auto LS = [&](int L) -> int { return ls_cached ? sls[L - L0] : ls[L]; };
    // thread -> (channel, slot): blockDim.x = 64 * NC, no division, no idle warp
    const int c = tid / SWEEP_SLOTS, slot = tid % SWEEP_SLOTS;

    // ---- stage the band: topology first (coalesced), then the feature gather with the loads batched four deep
    if (ls_cached)
        for (int L = L0 + tid; L <= L1; L += nthr) sls[L - L0] = ls[L];
    if (SMEM) {
        for (int i = p0 + tid; i < p1; i += nthr) {
            const int4 ch = child[i];
            const int cnt = child_count(ch), par = pp[i];
            // parents above the band / children below it are never followed, so their (out of range) offsets are dropped
            const int pl = (par >= p0) ? par - p0 : 0, cl = (cnt > 0 && ch.x < p1) ? ch.x - p0 : 0;
            sw[i - p0] = (i == 0) ? 0.f : w[i];
            spk[i - p0] = (pl << 15) | (cl << 3) | cnt;
            svx[i - p0] = vtx[i];
        }
        __syncthreads();
    }
    for (int cc = 0; cc < NC; ++cc) {
        int i = p0 + tid;
        for (; i + 3 * nthr < p1; i += 4 * nthr) {
            float v0 = prologue<MODE>(a, b, cc, i, SMEM ? svx[i - p0] : vtx[i], V);
            float v1 = prologue<MODE>(a, b, cc, i + nthr, SMEM ? svx[i + nthr - p0] : vtx[i + nthr], V);
            float v2 = prologue<MODE>(a, b, cc, i + 2 * nthr, SMEM ? svx[i + 2 * nthr - p0] : vtx[i + 2 * nthr], V);
            float v3 = prologue<MODE>(a, b, cc, i + 3 * nthr, SMEM ? svx[i + 3 * nthr - p0] : vtx[i + 3 * nthr], V);
            VAL(cc, i) = v0; VAL(cc, i + nthr) = v1; VAL(cc, i + 2 * nthr) = v2; VAL(cc, i + 3 * nthr) = v3;
        }
        for (; i < p1; i += nthr) VAL(cc, i) = prologue<MODE>(a, b, cc, i, SMEM ? svx[i - p0] : vtx[i], V);
    }
    __syncthreads();

    const bool inject = INJECT && beta < nb - 1;
    if (SMEM) {
        // ================= shared-memory band: tight loops on local indices =================
        float *mv = sval + (size_t)c * pcap;  // this thread's channel row, indexed by position - p0
        // ---- local leaf -> root sweep; the next level's bounds are fetched before the barrier
        int lo = LS(L1 - 1) - p0, hi = P;
        for (int L = L1 - 1; L >= L0; --L) {
            const int nlo = (L > L0) ? LS(L - 1) - p0 : 0;
            if (L == L1 - 1) {
                if (inject)
                    for (int il = lo + slot; il < hi; il += SWEEP_SLOTS) mv[il] += a.X[img + (size_t)c * V + p0 + il];
            } else {
                for (int il = lo + slot; il < hi; il += SWEEP_SLOTS) {
                    const int pk = spk[il], cl = (pk >> 3) & 0xfff, cnt = pk & 7;
                    float acc = mv[il];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (k < cnt) acc = fmaf(mv[cl + k], sw[cl + k], acc);
                    mv[il] = acc;
                }
            }
            __syncthreads();
            hi = lo;
            lo = nlo;
        }
        if (!INJECT) {  // U1: what this band's first level sends up, summed per parent = the base of X for the level above
            if (beta > 0) {
                const int q0 = ls[L0 - 1], nq = p0 - q0;
                for (int s = slot; s < nq; s += SWEEP_SLOTS) {
                    const int4 ch = child[q0 + s];
                    const int cnt = child_count(ch), cl = ch.x - p0;
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (k < cnt) acc = fmaf(mv[cl + k], sw[cl + k], acc);
                    a.X[img + (size_t)c * V + q0 + s] = acc;
                }
            }
            return;
        }
        // ---- U3: exact up leaves; D1: local root -> leaf pass, in place: val <- B
        for (int cc = 0; cc < NC; ++cc)
            for (int il = tid; il < P; il += nthr) a.up[img + (size_t)cc * V + p0 + il] = sval[(size_t)cc * pcap + il];
        __syncthreads();
        lo = 0;
        hi = LS(L0 + 1) - p0;
        for (int L = L0; L < L1; ++L) {
            const int nhi = (L + 1 < L1) ? LS(L + 2) - p0 : hi;
            for (int il = lo + slot; il < hi; il += SWEEP_SLOTS) {
                const float wi = sw[il];
                float bval = mv[il] * (1.f - wi * wi);
                if (L > L0) bval = fmaf(wi, mv[spk[il] >> 15], bval);
                mv[il] = bval;
            }
            __syncthreads();
            lo = hi;
            hi = nhi;
        }
        for (int cc = 0; cc < NC; ++cc)
            for (int il = tid; il < P; il += nthr) a.Bv[img + (size_t)cc * V + p0 + il] = sval[(size_t)cc * pcap + il];
        return;
    }

    // ================= oversized band: same recurrences on global memory =================
    for (int L = L1 - 1; L >= L0; --L) {
        const int lo = LS(L), n = LS(L + 1) - lo;
        const bool last = (L == L1 - 1);
        for (int s = slot; s < n; s += SWEEP_SLOTS) {
            const int i = lo + s;
            float acc = VAL(c, i);
            if (!last) {
                const int4 ch = child[i];
                const int cnt = child_count(ch);
                for (int k = 0; k < cnt; ++k) acc += VAL(c, ch.x + k) * Wat(ch.x + k);
            } else if (inject) {
                acc += a.X[img + (size_t)c * V + i];
            }
            VAL(c, i) = acc;
        }
        __syncthreads();
    }
    if (!INJECT) {
        if (beta > 0) {
            const int q0 = ls[L0 - 1], nq = LS(L0) - q0;
            for (int s = slot; s < nq; s += SWEEP_SLOTS) {
                const int4 ch = child[q0 + s];
                const int cnt = child_count(ch);
                float acc = 0.f;
                for (int k = 0; k < cnt; ++k) acc += VAL(c, ch.x + k) * Wat(ch.x + k);
                a.X[img + (size_t)c * V + q0 + s] = acc;
            }
        }
        return;
    }
    for (int cc = 0; cc < NC; ++cc)
        for (int i = p0 + tid; i < p1; i += nthr) a.up[img + (size_t)cc * V + i] = VAL(cc, i);
    __syncthreads();
    for (int L = L0; L < L1; ++L) {
        const int lo = LS(L), n = LS(L + 1) - lo;
        for (int s = slot; s < n; s += SWEEP_SLOTS) {
            const int i = lo + s;
            const float wi = Wat(i);
            float bval = VAL(c, i) * (1.f - wi * wi);
            if (L > L0) bval += wi * VAL(c, pp[i]);
            VAL(c, i) = bval;
        }
        __syncthreads();
    }
    for (int cc = 0; cc < NC; ++cc)
        for (int i = p0 + tid; i < p1; i += nthr) a.Bv[img + (size_t)cc * V + i] = VAL(cc, i);
}

template <int MODE, bool INJECT>
__global__ void __launch_bounds__(1024) sweep_kernel(TreeView t, SweepArgs a, int pcap) {
    extern __shared__ float4 smem4[];
    const int b = blockIdx.y, beta = blockIdx.x;
    int L0, L1, nb;
    if (!band_geom(t, b, beta, L0, L1, nb)) return;
    const int *ls = t.ls + (size_t)b * (t.V + 1);
    const int p0 = ls[L0], p1 = ls[L1];
    if (p1 - p0 <= pcap)
        sweep_body<MODE, INJECT, true>(t, a, b, beta, nb, L0, L1, p0, p1, pcap, reinterpret_cast<char *>(smem4));
    else
        sweep_body<MODE, INJECT, false>(t, a, b, beta, nb, L0, L1, p0, p1, pcap, reinterpret_cast<char *>(smem4));
}

// ------------------------------------------------------------------------------------------------ D3 epilogues
__device__ __forceinline__ float down_value(const float *A, const int *ent, const float *Bv, const float *og, int i) {
    const float a = A[i];
    float o = Bv[i];
    if (a != 0.f) o += a * og[ent[i]];
    return o;
}

// forward: aggregate in position order (saved), normalised output scattered to vertex order
__global__ void __launch_bounds__(256)
finish_forward_kernel(TreeView t, Consts k, const float *__restrict__ Bv, const float *__restrict__ outE, int C,
                      float *__restrict__ fpos, float *__restrict__ feature_out, float *__restrict__ feature_aggr,
                      float *__restrict__ weight_sum) {
    const int b = blockIdx.y, V = t.V, NC = C + 1;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    const float *A = k.A + (size_t)b * V;
    const int *ent = k.ent + (size_t)b * V;
    const size_t img = (size_t)b * NC * V;
    const int v = t.vtx[(size_t)b * V + i];
    const float z = down_value(A, ent, Bv + img + (size_t)C * V, outE + img + (size_t)C * V, i);
    fpos[img + (size_t)C * V + i] = z;
    if (weight_sum) weight_sum[(size_t)b * V + v] = z;
    for (int c = 0; c < C; ++c) {
        const float o = down_value(A, ent, Bv + img + (size_t)c * V, outE + img + (size_t)c * V, i);
        fpos[img + (size_t)c * V + i] = o;
        feature_out[((size_t)b * C + c) * V + v] = o / z;
        if (feature_aggr) feature_aggr[((size_t)b * C + c) * V + v] = o;
    }
}

// backward: feature gradient scattered to vertex order and, when asked, the edge-weight gradient (refine.cu:136-199)
__global__ void __launch_bounds__(256)
finish_backward_kernel(TreeView t, Consts k, const float *__restrict__ Bv, const float *__restrict__ outE,
                       const float *__restrict__ bup, const float *__restrict__ fup, const float *__restrict__ fpos, int C,
                       int NC, float *__restrict__ grad_feature, float *__restrict__ grad_weight) {
    const int b = blockIdx.y, V = t.V;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    const float *A = k.A + (size_t)b * V;
    const int *ent = k.ent + (size_t)b * V;
    const size_t img = (size_t)b * NC * V, fimg = (size_t)b * (C + 1) * V;
    const int v = t.vtx[(size_t)b * V + i];
    const int p = t.ppos[(size_t)b * V + i];
    const float wi = (i == 0) ? 0.f : t.w[(size_t)b * V + i];
    float gw = 0.f;
    for (int c = 0; c < C; ++c) {
        const float *Bc = Bv + img + (size_t)c * V, *oc = outE + img + (size_t)c * V;
        const float g1 = down_value(A, ent, Bc, oc, i);
        grad_feature[((size_t)b * C + c) * V + v] = g1;
        if (grad_weight != nullptr && i > 0) {
            const float *B2 = Bv + img + (size_t)(C + c) * V, *o2 = outE + img + (size_t)(C + c) * V;
            const float g1p = down_value(A, ent, Bc, oc, p), g2p = down_value(A, ent, B2, o2, p);
            const float u1 = bup[img + (size_t)c * V + i], u2 = bup[img + (size_t)(C + c) * V + i];
            const float fu = fup[fimg + (size_t)c * V + i], zu = fup[fimg + (size_t)C * V + i];
            const float fa_p = fpos[fimg + (size_t)c * V + p], z_p = fpos[fimg + (size_t)C * V + p];
            const float ga = u1 * (fa_p - wi * fu) + fu * (g1p - wi * u1);
            const float gb = u2 * (z_p - wi * zu) + zu * (g2p - wi * u2);
            gw += ga - gb;
        }
    }
    if (grad_weight != nullptr) grad_weight[(size_t)b * V + i] = gw;
}

// ------------------------------------------------------------------------------------------------ host side
// Band limits (host side, functions of the image size only).
#include <stdlib.h>
#include <string.h>
int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
// Plan limits.  Up to 2^17 pixels per image the resident kernel runs the filter: at most 36 bands (37 CTAs per image x 12
// images fill the GPU at three CTAs per SM) of at most 3100 positions (3100 x 24 B + level cache = a third of an SM's
// shared memory).  Larger images would need so many chunks of co-resident images that the multi-launch phases are
// faster: they get that path's plan (bands of 3072 positions, up to 128 of them).
int plan_nbcap(int V) { return env_int("FEDICRA_PLAN_NBCAP", V <= (1 << 17) ? NBCAP : 128); }
int plan_pmax(int V) { return env_int("FEDICRA_PLAN_PMAX", V <= (1 << 17) ? 3100 : 3072); }
#define PLAN_KDIV env_int("FEDICRA_PLAN_KDIV", 30)               // a band holds at most ceil(n_levels / KDIV) levels
int plan_grid(int V) { return plan_nbcap(V); }                    // upper bound of the number of bands
constexpr size_t SMEM_MAX = 227 * 1024;
// the multi-launch kernels keep (NC + 3) words per position; their shared-memory capacity may be below the plan's pmax
int multi_pcap(int V, int NC) {
    const long long fit = ((long long)SMEM_MAX - (LSCAP + 1) * 4 - 64) / ((NC + 3) * 4);
    const int pm = plan_pmax(V);
    const long long cap = fit < 4095 ? fit : 4095;   // 12-bit local offsets in the multi-launch kernels' packed topology
    return (int)(cap < pm ? cap : pm);
}

struct Layout {
    // saved (forward -> backward): A [V]; up, pos [(C+1),V]; ent, run_s, run_e, pk [V]
    static size_t saved_bytes(int B, int C, int V) { return (size_t)B * V * (1 + 2 * (C + 1) + 4) * 4; }
    // work: upl, X, up, Bv, outE, slab  [NCmax,V] each
    // preceded by the resident kernel's sync words, 8 ints per image at a FIXED place (the head of the workspace, sized
    // for the largest batch the API takes) so that no call's scratch ever lands on them: they must be zero before the
    // first use and every launch leaves them at zero
    static constexpr size_t SYNC_BYTES = (size_t)65536 * 8 * sizeof(int);
    static size_t work_bytes(int B, int NCmax, int V) { return SYNC_BYTES + (size_t)B * V * 6 * NCmax * 4; }
};

Consts consts_view(void *saved, int B, int C, int V, float **up, float **pos) {
    float *f = (float *)saved;
    const size_t n = (size_t)B * V;
    Consts k;
    k.A = f;
    *up = f + n;
    *pos = *up + n * (C + 1);
    int *ip = (int *)(*pos + n * (C + 1));
    k.ent = ip;
    k.run_s = ip + n;
    k.run_e = ip + 2 * n;
    k.pk = ip + 3 * n;
    return k;
}

template <typename Kern>
int set_smem(Kern kern, size_t bytes, const char *name) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return fedicra_fail(-2, "%s: cannot reserve %zu bytes of shared memory: %s", name, bytes, cudaGetErrorString(e));
    return 0;
}

#include "stitch2.cuh"

template <int MODE>
int run_sweeps(const TreeView &t, const Consts &k, SweepArgs a, float *outE, int B, cudaStream_t st) {
    const int V = t.V, nb = plan_grid(V), NC = a.NC;
    const int per_pos = (NC + 3) * 4;
    if (NC * SWEEP_SLOTS > 1024) return fedicra_fail(-1, "tree filter: %d channels per sweep exceed %d", NC, 1024 / SWEEP_SLOTS);
    const int nthr_sweep = NC * SWEEP_SLOTS;
    const int pcap = multi_pcap(V, NC);
    const size_t smem = (size_t)pcap * per_pos + (LSCAP + 1) * sizeof(int) + 16;
    if (set_smem(sweep_kernel<MODE, false>, smem, "sweep_kernel") || set_smem(sweep_kernel<MODE, true>, smem, "sweep_kernel")) return -2;
    // band-to-band recurrences: all last levels of a channel in shared memory (stitch2.cuh)
    const size_t smem_s2 = SMEM_MAX;
    const int tcap2 = (int)((smem_s2 - 3 * (MAXBANDS + 2) * sizeof(int) - 64) / 12);
    if (set_smem(stitch2_kernel<true>, smem_s2, "stitch2_kernel") || set_smem(stitch2_kernel<false>, smem_s2, "stitch2_kernel"))
        return -2;
    sweep_kernel<MODE, false><<<dim3(nb, B), nthr_sweep, smem, st>>>(t, a, pcap);
    FEDICRA_CHECK_LAUNCH("sweep_kernel<U1>");
    stitch2_kernel<true><<<dim3(NC, B), ST2_NT, smem_s2, st>>>(t, k, a.X, nullptr, nullptr, NC, tcap2);
    FEDICRA_CHECK_LAUNCH("stitch_up");
    sweep_kernel<MODE, true><<<dim3(nb, B), nthr_sweep, smem, st>>>(t, a, pcap);
    FEDICRA_CHECK_LAUNCH("sweep_kernel<U3D1>");
    stitch2_kernel<false><<<dim3(NC, B), ST2_NT, smem_s2, st>>>(t, k, nullptr, a.Bv, outE, NC, tcap2);
    FEDICRA_CHECK_LAUNCH("stitch_down");
    return 0;
}

#include "resident.cuh"

// Which implementation serves fedicra_tree_filter_*: the resident kernel (default) or the multi-launch phases.
bool use_resident() {
    const char *v = getenv("FEDICRA_FILTER");
    return !(v && strcmp(v, "multi") == 0);
}

// Shared memory of a resident CTA: band state (NC values + weight + packed topology per position) + level cache.
size_t resident_smem(int pcap, int NC) { return (size_t)(pcap + 4) * (NC + 2) * 4 + (LSCAP + 1) * sizeof(int) + 16; }

__device__ long long g_prof[(NBMAX + 1) * PROF_SLOTS];   // phase timestamps of the last profiled launch (FEDICRA_PROF=1)

template <int MODE, int NCT>
int launch_resident(const TreeView &t, const Consts &k, ResidentArgs a, int B, cudaStream_t st) {
    const int V = t.V, NC = a.NC, nthr = NC * RSLOTS, nbcap = plan_nbcap(V);
    const int pcap = plan_pmax(V);
    const size_t smem = resident_smem(pcap, NC);
    if (nthr > 1024 || smem > SMEM_MAX || pcap > 8191 || nbcap > NBMAX) return 1;
    static thread_local int cached_key[2] = {-1, -1}, cached_cap = 0, cached_per_sm = 0;   // (device, nthr, smem) -> resident CTAs
    int dev = 0;
    cudaGetDevice(&dev);
    {   // FEDICRA_WATCHDOG_MS: bound of the inter-CTA waits (0 = none), uploaded once per device
        static thread_local unsigned long long uploaded = 0;
        const char *v = getenv("FEDICRA_WATCHDOG_MS");
        if (v && *v && !(uploaded & (1ull << (dev & 63)))) {
            const unsigned long long ns = (unsigned long long)atoll(v) * 1000000ull;
            cudaMemcpyToSymbolAsync(g_resident_watchdog_ns, &ns, sizeof(ns), 0, cudaMemcpyHostToDevice, st);
            uploaded |= 1ull << (dev & 63);
        }
    }
    const int key0 = nthr * 64 + dev, key1 = (int)smem;
    if (cached_key[0] != key0 || cached_key[1] != key1) {
        if (set_smem(resident_kernel<MODE, NCT>, smem, "resident_kernel")) return -2;
        int per_sm = 0, sms = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resident_kernel<MODE, NCT>, nthr, smem);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return fedicra_fail(-2, "resident_kernel: occupancy query failed: %s", cudaGetErrorString(e));
        cached_key[0] = key0;
        cached_key[1] = key1;
        cached_cap = per_sm * sms;
        cached_per_sm = per_sm;
        // the launches below ask for the whole per-CTA share of the SM (see smem_all)
        const size_t share = per_sm > 0 ? (size_t)233472 / per_sm - 1024 : smem;
        if (set_smem(resident_kernel<MODE, NCT>, share > SMEM_MAX ? SMEM_MAX : (share < smem ? smem : share), "resident_kernel")) return -2;
    }
    const int chunk = cached_cap / (nbcap + 1);
    if (chunk < 1) return 1;
    a.pcap = pcap;
    // the stitcher's capacity comes from the same allocation: take everything the occupancy leaves (it is free)
    size_t smem_all = smem;
    if (cached_per_sm > 0) {
        const size_t share = (size_t)233472 / cached_per_sm - 1024;
        smem_all = share > SMEM_MAX ? SMEM_MAX : share;
        if (smem_all < smem) smem_all = smem;
    }
    a.tcap = (int)((smem_all - 3 * (NBMAX + 2) * sizeof(int) - 64) / (8 + 4 * NC));
    a.prof = nullptr;
    if (env_int("FEDICRA_PROF", 0)) {
        void *ptr = nullptr;
        if (cudaGetSymbolAddress(&ptr, g_prof) == cudaSuccess) a.prof = (long long *)ptr;
    }
    const bool coop = env_int("FEDICRA_COOP", 1) != 0;
    for (int b0 = 0; b0 < B; b0 += chunk) {
        a.b0 = b0;
        const int nbatch = (B - b0 < chunk) ? B - b0 : chunk;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(nbcap + 1, nbatch);
        cfg.blockDim = dim3(nthr);
        cfg.dynamicSmemBytes = smem_all;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = coop ? 1 : 0;
        cudaError_t e = cudaLaunchKernelEx(&cfg, resident_kernel<MODE, NCT>, t, k, a);
        fedicra_count_launches(1);
        if (e != cudaSuccess) return fedicra_fail(-2, "resident_kernel: %s", cudaGetErrorString(e));
        a.prof = nullptr;   // only the first chunk is profiled
    }
    return 0;
}

// Launches the resident kernel over the batch in chunks of as many images as fit on the GPU at once.
// Returns 1 when the configuration cannot run resident (the caller then takes the multi-launch path).
template <int MODE>
int run_resident(const TreeView &t, const Consts &k, const ResidentArgs &a, int B, cudaStream_t st) {
    switch (a.NC) {
        case 1: return launch_resident<MODE, 1>(t, k, a, B, st);
        case 2: return launch_resident<MODE, 2>(t, k, a, B, st);
        case 3: return launch_resident<MODE, 3>(t, k, a, B, st);
        case 4: return launch_resident<MODE, 4>(t, k, a, B, st);
        case 5: return launch_resident<MODE, 5>(t, k, a, B, st);
        case 6: return launch_resident<MODE, 6>(t, k, a, B, st);
        case 8: return launch_resident<MODE, 8>(t, k, a, B, st);
        default: return 1;   // other channel counts take the multi-launch phases
    }
}

}  // namespace

extern "C" int fedicra_debug_filter_profile(long long *h_out, int n) {
    const int total = (NBMAX + 1) * PROF_SLOTS;
    if (h_out == nullptr || n < total) return total;
    cudaError_t e = cudaMemcpyFromSymbol(h_out, g_prof, sizeof(long long) * total);
    if (e != cudaSuccess) return fedicra_fail(-2, "debug_filter_profile: %s", cudaGetErrorString(e));
    return total;
}

extern "C" size_t fedicra_band_plan_ints(int B) { return B > 0 ? (size_t)B * PLAN_STRIDE : 0; }

extern "C" int fedicra_band_plan(const int *level_start, const int *n_levels, int B, int V, int *bands, void *stream) {
    FEDICRA_REQUIRE(level_start && n_levels && bands, "band_plan: null pointer");
    FEDICRA_REQUIRE(B > 0 && V > 0, "band_plan: bad sizes B=%d V=%d", B, V);
    band_plan_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(level_start, n_levels, V, plan_pmax(V), PLAN_KDIV, plan_nbcap(V), bands);
    FEDICRA_CHECK_LAUNCH("band_plan_kernel");
    return 0;
}

extern "C" size_t fedicra_tree_filter_saved_bytes(int B, int C, int V) {
    if (B <= 0 || C <= 0 || V <= 0) return 0;
    return Layout::saved_bytes(B, C, V);
}

extern "C" size_t fedicra_tree_filter_workspace_bytes(int B, int C, int V) {
    if (B <= 0 || C <= 0 || V <= 0) return 0;
    return Layout::work_bytes(B, 2 * C > C + 1 ? 2 * C : C + 1, V);
}

extern "C" int fedicra_tree_filter_forward(const float *feature_in, float *edge_weight, const int *sorted_index,
                                           const int *sorted_parent, const int *sorted_child, const int *level_start,
                                           const int *n_levels, const int *bands, int B, int C, int V,
                                           float *feature_out, float *feature_aggr, float *weight_sum, void *saved,
                                           size_t saved_bytes,
                                           void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(feature_in && edge_weight && sorted_index && sorted_parent && sorted_child && level_start &&
                        n_levels && bands && feature_out && saved && workspace, "tree_filter_forward: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && C > 0 && C <= 64 && V > 0, "tree_filter_forward: bad sizes B=%d C=%d V=%d", B, C, V);
    FEDICRA_REQUIRE(saved_bytes >= fedicra_tree_filter_saved_bytes(B, C, V), "tree_filter_forward: `saved` too small");
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_tree_filter_workspace_bytes(B, C, V), "tree_filter_forward: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = plan_grid(V), NC = C + 1;
    TreeView t{sorted_index, sorted_parent, reinterpret_cast<const int4 *>(sorted_child), level_start, n_levels, edge_weight, bands, V};
    float *fup, *fpos;
    Consts k = consts_view(saved, B, C, V, &fup, &fpos);
    const size_t n = (size_t)B * V * (2 * C > NC ? 2 * C : NC);
    float *wk = (float *)((char *)workspace + Layout::SYNC_BYTES);
    SweepArgs a{feature_in, nullptr, C, NC, wk, wk + n, fup, wk + 2 * n, wk + 4 * n};
    float *outE = wk + 3 * n;

    const int pcap_c = multi_pcap(V, 1);
    const size_t smem_c = (size_t)pcap_c * 16 + 16;
    if (set_smem(consts_kernel, smem_c, "consts_kernel")) return -2;
    consts_kernel<<<dim3(nb, B), NT, smem_c, st>>>(t, k, pcap_c);
    FEDICRA_CHECK_LAUNCH("consts_kernel");
    if (use_resident()) {
        int *sync = reinterpret_cast<int *>(workspace);
        ResidentArgs ra{feature_in, fpos, fup, C, NC, wk + n, nullptr, wk + 2 * n, wk + 3 * n, wk + 4 * n, sync,
                        feature_out, feature_aggr, weight_sum, 0, 0, 0, nullptr};
        const int rc = run_resident<MODE_FWD>(t, k, ra, B, st);
        if (rc <= 0) return rc;   // launched (0) or failed (<0); 1 = configuration not supported, take the phases below
    }
    if (int rc = run_sweeps<MODE_FWD>(t, k, a, outE, B, st)) return rc;
    finish_forward_kernel<<<dim3((V + 255) / 256, B), 256, 0, st>>>(t, k, a.Bv, outE, C, fpos, feature_out, feature_aggr, weight_sum);
    FEDICRA_CHECK_LAUNCH("finish_forward_kernel");
    return 0;
}

extern "C" int fedicra_tree_filter_backward(const float *grad_out, const float *edge_weight, const int *sorted_index,
                                            const int *sorted_parent, const int *sorted_child, const int *level_start,
                                            const int *n_levels, const int *bands, int B, int C, int V,
                                            const void *saved, float *grad_feature, float *grad_weight, void *workspace,
                                            size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(grad_out && edge_weight && sorted_index && sorted_parent && sorted_child && level_start && n_levels &&
                        bands && saved && grad_feature && workspace, "tree_filter_backward: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && C > 0 && C <= 64 && V > 0, "tree_filter_backward: bad sizes B=%d C=%d V=%d", B, C, V);
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_tree_filter_workspace_bytes(B, C, V), "tree_filter_backward: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int NC = grad_weight ? 2 * C : C;
    TreeView t{sorted_index, sorted_parent, reinterpret_cast<const int4 *>(sorted_child), level_start, n_levels, edge_weight, bands, V};
    float *fup, *fpos;
    Consts k = consts_view(const_cast<void *>(saved), B, C, V, &fup, &fpos);
    const size_t n = (size_t)B * V * (2 * C > C + 1 ? 2 * C : C + 1);
    float *wk = (float *)((char *)workspace + Layout::SYNC_BYTES);
    SweepArgs a{grad_out, fpos, C, NC, wk, wk + n, wk + 5 * n, wk + 2 * n, wk + 4 * n};
    float *outE = wk + 3 * n;
    if (use_resident()) {
        int *sync = reinterpret_cast<int *>(workspace);
        ResidentArgs ra{grad_out, fpos, fup, C, NC, wk + n, wk + 5 * n, wk + 2 * n, wk + 3 * n, wk + 4 * n, sync,
                        grad_feature, grad_weight, nullptr, 0, 0, 0, nullptr};
        const int rc = run_resident<MODE_BWD>(t, k, ra, B, st);
        if (rc <= 0) return rc;
    }
    if (int rc = run_sweeps<MODE_BWD>(t, k, a, outE, B, st)) return rc;
    finish_backward_kernel<<<dim3((V + 255) / 256, B), 256, 0, st>>>(t, k, a.Bv, outE, a.up, fup, fpos, C, NC, grad_feature, grad_weight);
    FEDICRA_CHECK_LAUNCH("finish_backward_kernel");
    return 0;
}
