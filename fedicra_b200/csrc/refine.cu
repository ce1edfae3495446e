// refine.cu -- tree filter sweeps on BFS-ordered spanning trees, reference tensor layouts (sm_100a).
//
// Replaces LTF/src/refine/refine.cu of the reference (LTF = /root/reference/code/utils/TreeEnergyLoss/kernels/
// lib_tree_filter): leaf_root_aggr_kernel :70-134, root_leaf_prop_kernel :19-68, root_leaf_grad_kernel :136-199 and the
// three host entry points :201-370, plus TreeFilter2D.build_edge_weight (LTF/modules/tree_filter.py:91-108) and its
// autograd.  The reference gives each (image, channel) 64 threads that poll their parents/children through shared flags
// with a __syncthreads per poll; here the BFS order is a true level order (bfs.cu of this library), so a CTA advances one
// whole depth per barrier and prefetches the next depth's operands while it waits.
//
// Layout reminder: features and outputs are in VERTEX order, edge weights and the *_up
// aggregates are in BFS order.
#include "common.cuh"

namespace {

constexpr int NT = 128;

enum UpMode { UP_PLAIN = 0, UP_ONES = 1, UP_DIV = 2, UP_DIVMUL = 3 };

struct UpArgs {
    const float *in;    // [B,C,V] vertex order (PLAIN / DIV / DIVMUL)
    const float *z;     // [B,V]   vertex order divisor (DIV / DIVMUL)
    const float *mul;   // [B,C,V] vertex order multiplier (DIVMUL)
    float *out;         // [B,C,V] BFS order
    int mode;
};

// up[i] = val(vertex(i)) + sum_children up[c] * w[c], deepest level first  (refine.cu:70-134)
__device__ __forceinline__ void sweep_up(const UpArgs &a, int b, int c, int C, int V, const float *__restrict__ w,
                                         const int *__restrict__ si, const int4 *__restrict__ sc,
                                         const int *__restrict__ ls, int nl) {
    const int tid = threadIdx.x;
    const size_t off = ((size_t)b * C + c) * V;
    const float *in = a.in ? a.in + off : nullptr;
    const float *z = a.z ? a.z + (size_t)b * V : nullptr;
    const float *mul = a.mul ? a.mul + off : nullptr;
    float *out = a.out + off;
    auto value = [&](int v) -> float {
        switch (a.mode) {
            case UP_ONES: return 1.f;
            case UP_DIV: return in[v] / z[v];
            case UP_DIVMUL: return in[v] / z[v] * mul[v];
            default: return in[v];
        }
    };
    int hi = ls[nl], lo = ls[nl - 1];
    // operands of the current level's first item per thread, fetched one level ahead
    float pv = 0.f;
    int4 pc = make_int4(0, 0, 0, 0);
    if (lo + tid < hi) {
        pv = value(si[lo + tid]);
        pc = sc[lo + tid];
    }
    for (int L = nl - 1; L >= 0; --L) {
        const int nlo = (L > 0) ? ls[L - 1] : 0;
        for (int i = lo + tid; i < hi; i += NT) {
            float acc;
            int4 ch;
            if (i == lo + tid) {
                acc = pv;
                ch = pc;
            } else {
                acc = value(si[i]);
                ch = sc[i];
            }
            if (ch.x > 0) acc += out[ch.x] * w[ch.x];
            if (ch.y > 0) acc += out[ch.y] * w[ch.y];
            if (ch.z > 0) acc += out[ch.z] * w[ch.z];
            if (ch.w > 0) acc += out[ch.w] * w[ch.w];
            out[i] = acc;
        }
        if (L > 0 && nlo + tid < lo) {
            pv = value(si[nlo + tid]);
            pc = sc[nlo + tid];
        }
        __syncthreads();
        hi = lo;
        lo = nlo;
    }
}

// out[vertex(i)] = in[i]*(1-w_i^2) + out[vertex(par i)]*w_i, root first, w_0 := 0  (refine.cu:19-68)
__device__ __forceinline__ void sweep_down(const float *in, float *out, const float *__restrict__ w,
                                           const int *__restrict__ si, const int *__restrict__ sp,
                                           const int *__restrict__ ls, int nl) {
    const int tid = threadIdx.x;
    int lo = 0, hi = ls[1];
    float p_in = 0.f, p_w = 0.f;
    int p_v = 0, p_pv = 0;
    if (tid < hi) {
        p_in = in[tid];
        p_w = (tid == 0) ? 0.f : w[tid];
        p_v = si[tid];
        p_pv = si[sp[tid]];
    }
    for (int L = 0; L < nl; ++L) {
        const int nhi = (L + 1 < nl) ? ls[L + 2] : hi;
        for (int i = lo + tid; i < hi; i += NT) {
            float x, wi;
            int v, pv;
            if (i == lo + tid) {
                x = p_in; wi = p_w; v = p_v; pv = p_pv;
            } else {
                x = in[i]; wi = w[i]; v = si[i]; pv = si[sp[i]];
            }
            out[v] = (i == 0) ? x : x * (1.f - wi * wi) + out[pv] * wi;
        }
        if (L + 1 < nl && hi + tid < nhi) {
            const int i = hi + tid;
            p_in = in[i];
            p_w = w[i];
            p_v = si[i];
            p_pv = si[sp[i]];
        }
        __syncthreads();
        lo = hi;
        hi = nhi;
    }
}

__global__ void __launch_bounds__(NT)
up_kernel(UpArgs a, UpArgs ones, int C, int V, float *edge_weight, const int *__restrict__ si,
          const int *__restrict__ sc, const int *__restrict__ ls, const int *__restrict__ nls) {
    // grid (B, C [+1]): channel C, when launched, is the all-ones channel (weight_sum_up)
    const int b = blockIdx.x, c = blockIdx.y;
    const float *w = edge_weight + (size_t)b * V;
    if (c == 0 && threadIdx.x == 0 && ones.out != nullptr) edge_weight[(size_t)b * V] = 0.f;  // refine.cu:43-46
    const int nl = nls[b];
    const int *lsb = ls + (size_t)b * (V + 1);
    const int4 *scb = reinterpret_cast<const int4 *>(sc) + (size_t)b * V;
    if (c < C)
        sweep_up(a, b, c, C, V, w, si + (size_t)b * V, scb, lsb, nl);
    else
        sweep_up(ones, b, 0, 1, V, w, si + (size_t)b * V, scb, lsb, nl);
}

__global__ void __launch_bounds__(NT)
down_kernel(const float *in, float *out, const float *in1, float *out1, int C, int V, const float *__restrict__ edge_weight,
            const int *__restrict__ si, const int *__restrict__ sp, const int *__restrict__ ls,
            const int *__restrict__ nls) {
    // grid (B, C [+1]): channel C, when launched, is the single-channel pair (in1 -> out1)
    const int b = blockIdx.x, c = blockIdx.y;
    const float *w = edge_weight + (size_t)b * V;
    const int *lsb = ls + (size_t)b * (V + 1);
    if (c < C) {
        const size_t off = ((size_t)b * C + c) * V;
        sweep_down(in + off, out + off, w, si + (size_t)b * V, sp + (size_t)b * V, lsb, nls[b]);
    } else {
        sweep_down(in1 + (size_t)b * V, out1 + (size_t)b * V, w, si + (size_t)b * V, sp + (size_t)b * V, lsb, nls[b]);
    }
}

// root_leaf_grad_kernel (refine.cu:136-199).  grid (B, 2C): y <  C  pass A: data = (aggr_up, aggr),     grad chain U1
//                                                           y >= C  pass B: data = (wsum_up, wsum) 1ch, grad chain U2
__global__ void __launch_bounds__(NT)
grad_kernel(const float *__restrict__ aggr_up, const float *__restrict__ aggr, const float *__restrict__ wsum_up,
            const float *__restrict__ wsum, float *u1, float *u2, float *ga, float *gb, int C, int V,
            const float *__restrict__ edge_weight, const int *__restrict__ si, const int *__restrict__ sp,
            const int *__restrict__ ls, const int *__restrict__ nls) {
    const int b = blockIdx.x, tid = threadIdx.x;
    const bool passA = blockIdx.y < C;
    const int c = passA ? blockIdx.y : blockIdx.y - C;
    const size_t off = ((size_t)b * C + c) * V;
    const float *idat = passA ? aggr_up + off : wsum_up + (size_t)b * V;  // BFS order
    const float *odat = passA ? aggr + off : wsum + (size_t)b * V;        // vertex order
    float *g = (passA ? u1 : u2) + off;                                   // BFS order, converted in place
    float *o = (passA ? ga : gb) + off;                                   // BFS order
    const float *w = edge_weight + (size_t)b * V;
    const int *sib = si + (size_t)b * V, *spb = sp + (size_t)b * V, *lsb = ls + (size_t)b * (V + 1);
    const int nl = nls[b];
    if (tid == 0) o[0] = 0.f;
    int lo = lsb[1], hi = (nl > 1) ? lsb[2] : lo;
    for (int L = 1; L < nl; ++L) {
        const int nhi = (L + 1 < nl) ? lsb[L + 2] : hi;
        for (int i = lo + tid; i < hi; i += NT) {
            const int p = spb[i];
            const float wi = w[i], gi = g[i], gp = g[p], di = idat[i];
            const float left = gi * (odat[sib[p]] - wi * di);
            const float right = di * (gp - wi * gi);
            o[i] = left + right;
            g[i] = gi * (1.f - wi * wi) + gp * wi;
        }
        __syncthreads();
        lo = hi;
        hi = nhi;
    }
}

__global__ void div_kernel(const float *__restrict__ num, const float *__restrict__ den, float *__restrict__ out, int C,
                           int V, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const size_t b = i / ((size_t)C * V), v = i % V;
    out[i] = num[i] / den[b * V + v];
}

__global__ void sum_diff_kernel(const float *__restrict__ ga, const float *__restrict__ gb, float *__restrict__ out,
                                int C, int V, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const size_t b = i / V, v = i % V;
    float acc = 0.f;
    for (int c = 0; c < C; ++c) acc += ga[(b * C + c) * V + v] - gb[(b * C + c) * V + v];
    out[i] = acc;
}

__global__ void edge_weight_fwd_kernel(const float *__restrict__ embed, const int *__restrict__ si,
                                       const int *__restrict__ sp, int D, int V, float inv_sigma,
                                       float *__restrict__ ew, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t b = idx / V;
    const int i = (int)(idx % V);
    const int *sib = si + b * V;
    const int v = sib[i], u = sib[sp[b * V + i]];
    const float *e = embed + b * D * V;
    float d = 0.f;
    for (int c = 0; c < D; ++c) {
        const float t = e[(size_t)c * V + v] - e[(size_t)c * V + u];
        d = fmaf(t, t, d);
    }
    ew[idx] = expf(-d * inv_sigma);
}

// grad_embed[v(i)] = k_i (e_v - e_par) - sum_children k_j (e_child - e_v),  k_i = -2/sigma * ew_i * g_i  (k_root = 0)
__global__ void edge_weight_bwd_kernel(const float *__restrict__ embed, const int *__restrict__ si,
                                       const int *__restrict__ sp, const int *__restrict__ sc,
                                       const float *__restrict__ ew, const float *__restrict__ gew, int D, int V,
                                       float inv_sigma, float *__restrict__ grad, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t b = idx / V;
    const int i = (int)(idx % V);
    const int *sib = si + b * V;
    const float *ewb = ew + b * V, *gb = gew + b * V;
    const int v = sib[i], pv = sib[sp[b * V + i]];
    const int4 ch = reinterpret_cast<const int4 *>(sc)[b * V + i];
    const float k = (i > 0) ? -2.f * inv_sigma * ewb[i] * gb[i] : 0.f;
    const int cj[4] = {ch.x, ch.y, ch.z, ch.w};
    float kj[4];
    int vj[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        kj[j] = (cj[j] > 0) ? -2.f * inv_sigma * ewb[cj[j]] * gb[cj[j]] : 0.f;
        vj[j] = (cj[j] > 0) ? sib[cj[j]] : v;
    }
    const float *e = embed + b * D * V;
    float *g = grad + b * D * V;
    for (int c = 0; c < D; ++c) {
        const float ev = e[(size_t)c * V + v];
        float acc = k * (ev - e[(size_t)c * V + pv]);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc -= kj[j] * (e[(size_t)c * V + vj[j]] - ev);
        g[(size_t)c * V + v] = acc;
    }
}

// The same gradient for the trees of several embeddings of one batch in ONE launch (the heads of the multiscale loss:
// their filters run one after the other, their weight gradients are all there when the last filter's backward is done).
constexpr int EWB_MAX = 8;
struct EwbSet {
    const float *embed[EWB_MAX], *ew[EWB_MAX], *gew[EWB_MAX];
    const int *si[EWB_MAX], *sp[EWB_MAX], *sc[EWB_MAX];
    float *grad[EWB_MAX];
    int D[EWB_MAX];
    float inv_sigma[EWB_MAX];
};
__global__ void edge_weight_bwd_multi_kernel(EwbSet es, int V, size_t per) {   // per = B * V positions per embedding
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (idx >= per) return;
    const size_t b = idx / V;
    const int i = (int)(idx % V);
    const int *sib = es.si[k] + b * V;
    const float *ewb = es.ew[k] + b * V, *gb = es.gew[k] + b * V;
    const float inv_sigma = es.inv_sigma[k];
    const int D = es.D[k];
    const int v = sib[i], pv = sib[es.sp[k][b * V + i]];
    const int4 ch = reinterpret_cast<const int4 *>(es.sc[k])[b * V + i];
    const float kk = (i > 0) ? -2.f * inv_sigma * ewb[i] * gb[i] : 0.f;
    const int cj[4] = {ch.x, ch.y, ch.z, ch.w};
    float kj[4];
    int vj[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        kj[j] = (cj[j] > 0) ? -2.f * inv_sigma * ewb[cj[j]] * gb[cj[j]] : 0.f;
        vj[j] = (cj[j] > 0) ? sib[cj[j]] : v;
    }
    const float *e = es.embed[k] + b * D * V;
    float *g = es.grad[k] + b * D * V;
    for (int c = 0; c < D; ++c) {
        const float ev = e[(size_t)c * V + v];
        float acc = kk * (ev - e[(size_t)c * V + pv]);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc -= kj[j] * (e[(size_t)c * V + vj[j]] - ev);
        g[(size_t)c * V + v] = acc;
    }
}

unsigned blocks_for(size_t n, int t) { return (unsigned)((n + t - 1) / t); }

}  // namespace

extern "C" int fedicra_edge_weight_fwd(const float *embed, const int *sorted_index, const int *sorted_parent, int B,
                                       int D, int V, float inv_sigma, float *edge_weight, void *stream) {
    FEDICRA_REQUIRE(embed && sorted_index && sorted_parent && edge_weight, "edge_weight_fwd: null pointer");
    FEDICRA_REQUIRE(B > 0 && D > 0 && V > 0, "edge_weight_fwd: bad sizes B=%d D=%d V=%d", B, D, V);
    const size_t total = (size_t)B * V;
    edge_weight_fwd_kernel<<<blocks_for(total, 256), 256, 0, (cudaStream_t)stream>>>(embed, sorted_index, sorted_parent,
                                                                                     D, V, inv_sigma, edge_weight, total);
    FEDICRA_CHECK_LAUNCH("edge_weight_fwd_kernel");
    return 0;
}

extern "C" int fedicra_edge_weight_bwd(const float *embed, const int *sorted_index, const int *sorted_parent,
                                       const int *sorted_child, const float *edge_weight,
                                       const float *grad_edge_weight, int B, int D, int V, float inv_sigma,
                                       float *grad_embed, void *stream) {
    FEDICRA_REQUIRE(embed && sorted_index && sorted_parent && sorted_child && edge_weight && grad_edge_weight &&
                        grad_embed, "edge_weight_bwd: null pointer");
    FEDICRA_REQUIRE(B > 0 && D > 0 && V > 0, "edge_weight_bwd: bad sizes B=%d D=%d V=%d", B, D, V);
    const size_t total = (size_t)B * V;
    edge_weight_bwd_kernel<<<blocks_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        embed, sorted_index, sorted_parent, sorted_child, edge_weight, grad_edge_weight, D, V, inv_sigma, grad_embed,
        total);
    FEDICRA_CHECK_LAUNCH("edge_weight_bwd_kernel");
    return 0;
}

extern "C" int fedicra_edge_weight_bwd_multi(const float *const *h_embed, const int *const *h_sorted_index,
                                             const int *const *h_sorted_parent, const int *const *h_sorted_child,
                                             const float *const *h_edge_weight, const float *const *h_grad_edge_weight,
                                             const int *h_D, const float *h_inv_sigma, int K, int B, int V,
                                             float *const *h_grad_embed, void *stream) {
    FEDICRA_REQUIRE(h_embed && h_sorted_index && h_sorted_parent && h_sorted_child && h_edge_weight &&
                        h_grad_edge_weight && h_D && h_inv_sigma && h_grad_embed, "edge_weight_bwd_multi: null pointer");
    FEDICRA_REQUIRE(K >= 1 && K <= EWB_MAX && B > 0 && V > 0, "edge_weight_bwd_multi: bad sizes K=%d (1..%d) B=%d V=%d", K,
                    EWB_MAX, B, V);
    EwbSet es = {};
    for (int k = 0; k < K; ++k) {
        FEDICRA_REQUIRE(h_embed[k] && h_sorted_index[k] && h_sorted_parent[k] && h_sorted_child[k] && h_edge_weight[k] &&
                            h_grad_edge_weight[k] && h_grad_embed[k] && h_D[k] > 0,
                        "edge_weight_bwd_multi: entry %d has a null pointer or no channels", k);
        FEDICRA_REQUIRE(((uintptr_t)h_sorted_child[k] & 15) == 0, "edge_weight_bwd_multi: sorted_child must be 16-byte aligned");
        es.embed[k] = h_embed[k]; es.si[k] = h_sorted_index[k]; es.sp[k] = h_sorted_parent[k]; es.sc[k] = h_sorted_child[k];
        es.ew[k] = h_edge_weight[k]; es.gew[k] = h_grad_edge_weight[k]; es.grad[k] = h_grad_embed[k];
        es.D[k] = h_D[k]; es.inv_sigma[k] = h_inv_sigma[k];
    }
    const size_t per = (size_t)B * V;
    edge_weight_bwd_multi_kernel<<<dim3(blocks_for(per, 256), K), 256, 0, (cudaStream_t)stream>>>(es, V, per);
    FEDICRA_CHECK_LAUNCH("edge_weight_bwd_multi_kernel");
    return 0;
}

extern "C" int fedicra_refine_forward(const float *feature_in, float *edge_weight, const int *sorted_index,
                                      const int *sorted_parent, const int *sorted_child, const int *level_start,
                                      const int *n_levels, int B, int C, int V, float *feature_out,
                                      float *feature_aggr, float *feature_aggr_up, float *weight_sum,
                                      float *weight_sum_up, void *stream) {
    FEDICRA_REQUIRE(feature_in && edge_weight && sorted_index && sorted_parent && sorted_child && level_start &&
                        n_levels && feature_out && feature_aggr && feature_aggr_up && weight_sum && weight_sum_up,
                    "refine_forward: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && C > 0 && C < 65535 && V > 0, "refine_forward: bad sizes B=%d C=%d V=%d", B, C, V);
    cudaStream_t st = (cudaStream_t)stream;
    UpArgs a{feature_in, nullptr, nullptr, feature_aggr_up, UP_PLAIN};
    UpArgs ones{nullptr, nullptr, nullptr, weight_sum_up, UP_ONES};
    up_kernel<<<dim3(B, C + 1), NT, 0, st>>>(a, ones, C, V, edge_weight, sorted_index, sorted_child, level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("up_kernel");
    down_kernel<<<dim3(B, C + 1), NT, 0, st>>>(feature_aggr_up, feature_aggr, weight_sum_up, weight_sum, C, V,
                                               edge_weight, sorted_index, sorted_parent, level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("down_kernel");
    const size_t total = (size_t)B * C * V;
    div_kernel<<<blocks_for(total, 256), 256, 0, st>>>(feature_aggr, weight_sum, feature_out, C, V, total);
    FEDICRA_CHECK_LAUNCH("div_kernel");
    return 0;
}

extern "C" size_t fedicra_refine_backward_workspace_bytes(int B, int C, int V) {
    if (B <= 0 || C <= 0 || V <= 0) return 0;
    return 4 * align_up((size_t)B * C * V * sizeof(float), 256);
}

extern "C" int fedicra_refine_backward_feature(const float *edge_weight, const int *sorted_index,
                                               const int *sorted_parent, const int *sorted_child,
                                               const int *level_start, const int *n_levels, const float *weight_sum,
                                               const float *grad_out, int B, int C, int V, float *grad_feature,
                                               void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(edge_weight && sorted_index && sorted_parent && sorted_child && level_start && n_levels &&
                        weight_sum && grad_out && grad_feature && workspace, "refine_backward_feature: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && C > 0 && C < 65535 && V > 0, "refine_backward_feature: bad sizes");
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_refine_backward_workspace_bytes(B, C, V),
                    "refine_backward_feature: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    float *up = (float *)workspace;
    UpArgs a{grad_out, weight_sum, nullptr, up, UP_DIV};
    UpArgs none{nullptr, nullptr, nullptr, nullptr, UP_ONES};
    up_kernel<<<dim3(B, C), NT, 0, st>>>(a, none, C, V, const_cast<float *>(edge_weight), sorted_index, sorted_child,
                                         level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("up_kernel");
    down_kernel<<<dim3(B, C), NT, 0, st>>>(up, grad_feature, nullptr, nullptr, C, V, edge_weight, sorted_index,
                                           sorted_parent, level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("down_kernel");
    return 0;
}

extern "C" int fedicra_refine_backward_weight(const float *edge_weight, const int *sorted_index,
                                              const int *sorted_parent, const int *sorted_child,
                                              const int *level_start, const int *n_levels, const float *feature_out,
                                              const float *feature_aggr, const float *feature_aggr_up,
                                              const float *weight_sum, const float *weight_sum_up,
                                              const float *grad_out, int B, int C, int V, float *grad_weight,
                                              void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(edge_weight && sorted_index && sorted_parent && sorted_child && level_start && n_levels &&
                        feature_out && feature_aggr && feature_aggr_up && weight_sum && weight_sum_up && grad_out &&
                        grad_weight && workspace, "refine_backward_weight: null pointer");
    FEDICRA_REQUIRE(B > 0 && B <= 65535 && C > 0 && 2 * C < 65535 && V > 0, "refine_backward_weight: bad sizes");
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_refine_backward_workspace_bytes(B, C, V),
                    "refine_backward_weight: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t stride = align_up((size_t)B * C * V * sizeof(float), 256) / sizeof(float);
    float *u1 = (float *)workspace, *u2 = u1 + stride, *ga = u2 + stride, *gb = ga + stride;
    UpArgs none{nullptr, nullptr, nullptr, nullptr, UP_ONES};
    UpArgs a1{grad_out, weight_sum, nullptr, u1, UP_DIV};
    UpArgs a2{grad_out, weight_sum, feature_out, u2, UP_DIVMUL};
    float *ew = const_cast<float *>(edge_weight);
    up_kernel<<<dim3(B, C), NT, 0, st>>>(a1, none, C, V, ew, sorted_index, sorted_child, level_start, n_levels);
    up_kernel<<<dim3(B, C), NT, 0, st>>>(a2, none, C, V, ew, sorted_index, sorted_child, level_start, n_levels);
    fedicra_count_launches(1);
    FEDICRA_CHECK_LAUNCH("up_kernel");
    grad_kernel<<<dim3(B, 2 * C), NT, 0, st>>>(feature_aggr_up, feature_aggr, weight_sum_up, weight_sum, u1, u2, ga, gb,
                                               C, V, edge_weight, sorted_index, sorted_parent, level_start, n_levels);
    FEDICRA_CHECK_LAUNCH("grad_kernel");
    const size_t total = (size_t)B * V;
    sum_diff_kernel<<<blocks_for(total, 256), 256, 0, st>>>(ga, gb, grad_weight, C, V, total);
    FEDICRA_CHECK_LAUNCH("sum_diff_kernel");
    return 0;
}
