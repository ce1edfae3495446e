// ala.cu -- element-wise adaptive local aggregation update and the FedAvg pre-scale (sm_100a).
//
// Replaces the per-tensor PyTorch loops of MyModel.set_weights in /root/reference/code/flower_common.py:
//   :553-555  t <- p + (s - p) * w                               (initial blend)
//   :590-597  w <- clamp(w - eta * (g * (s - p)), 0, 1)          (skipped for tensors without a gradient, :593-594)
//   :600-602  t <- p + (s - p) * w
// 3 loops x 42 tensors x ~5 tiny launches per batch in the reference; here one launch over all tensors.  Every product
// and sum is rounded separately (no fused multiply-add), so the result is bit-identical to the torch expressions.
// HBM-bound: 4 arrays read + 2 written = 24 B per element.
#include "common.cuh"

namespace {

__device__ __forceinline__ void ala_elem(float p, float s, float g, bool has_g, float &w, float &t, float eta) {
    const float diff = __fsub_rn(s, p);
    if (has_g) {
        const float step = __fmul_rn(eta, __fmul_rn(g, diff));
        w = fminf(fmaxf(__fsub_rn(w, step), 0.f), 1.f);
    }
    t = __fadd_rn(p, __fmul_rn(diff, w));
}

__device__ __forceinline__ void ala_range(const float *__restrict__ p, const float *__restrict__ s,
                                          const float *__restrict__ g, float *__restrict__ w, float *__restrict__ t,
                                          long long begin, long long end, float eta) {
    const bool has_g = g != nullptr;
    const bool vec = (((uintptr_t)p | (uintptr_t)s | (uintptr_t)g | (uintptr_t)w | (uintptr_t)t) & 15) == 0 && (begin & 3) == 0;
    long long i = begin + (long long)threadIdx.x * 4;
    if (vec) {
        for (; i + 3 < end; i += (long long)blockDim.x * 4) {
            const float4 P = *reinterpret_cast<const float4 *>(p + i), S = *reinterpret_cast<const float4 *>(s + i);
            float4 G = make_float4(0.f, 0.f, 0.f, 0.f);
            if (has_g) G = *reinterpret_cast<const float4 *>(g + i);
            float4 Wv = *reinterpret_cast<float4 *>(w + i), T;
            ala_elem(P.x, S.x, G.x, has_g, Wv.x, T.x, eta);
            ala_elem(P.y, S.y, G.y, has_g, Wv.y, T.y, eta);
            ala_elem(P.z, S.z, G.z, has_g, Wv.z, T.z, eta);
            ala_elem(P.w, S.w, G.w, has_g, Wv.w, T.w, eta);
            if (has_g) *reinterpret_cast<float4 *>(w + i) = Wv;
            *reinterpret_cast<float4 *>(t + i) = T;
        }
        // tail (< 4 elements): the thread whose i straddles `end`
        if (i < end) {
            for (long long k = i; k < end; ++k) {
                float wv = w[k], tv;
                ala_elem(p[k], s[k], has_g ? g[k] : 0.f, has_g, wv, tv, eta);
                if (has_g) w[k] = wv;
                t[k] = tv;
            }
        }
    } else {
        for (long long k = begin + threadIdx.x; k < end; k += blockDim.x) {
            float wv = w[k], tv;
            ala_elem(p[k], s[k], has_g ? g[k] : 0.f, has_g, wv, tv, eta);
            if (has_g) w[k] = wv;
            t[k] = tv;
        }
    }
}

__global__ void __launch_bounds__(256)
ala_flat_kernel(const float *p, const float *s, const float *g, float *w, float *t, long long n, int chunk, float eta) {
    const long long begin = (long long)blockIdx.x * chunk;
    const long long end = begin + chunk < n ? begin + chunk : n;
    ala_range(p, s, g, w, t, begin, end, eta);
}

__global__ void __launch_bounds__(256)
ala_multi_kernel(const float *const *p, const float *const *s, const float *const *g, float *const *w, float *const *t,
                 const long long *sizes, const int *chunk_tensor, const long long *chunk_begin, int chunk, float eta) {
    const int k = chunk_tensor[blockIdx.x];
    const long long begin = chunk_begin[blockIdx.x], n = sizes[k];
    const long long end = begin + chunk < n ? begin + chunk : n;
    ala_range(p[k], s[k], g[k], w[k], t[k], begin, end, eta);
}

__global__ void scale_kernel(float *x, long long n, float alpha) {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if ((((uintptr_t)x) & 15) == 0 && i + 3 < n) {
        float4 v = *reinterpret_cast<float4 *>(x + i);
        v.x *= alpha; v.y *= alpha; v.z *= alpha; v.w *= alpha;
        *reinterpret_cast<float4 *>(x + i) = v;
    } else {
        for (long long k = i; k < n && k < i + 4; ++k) x[k] *= alpha;
    }
}

}  // namespace

extern "C" int fedicra_ala_update_flat(const float *p, const float *s, const float *g, float *w, float *t, int64_t n,
                                       float eta, void *stream) {
    FEDICRA_REQUIRE(p && s && w && t, "ala_update_flat: null pointer");
    FEDICRA_REQUIRE(n >= 0, "ala_update_flat: negative size");
    if (n == 0) return 0;
    const int chunk = 4096;
    const long long blocks = (n + chunk - 1) / chunk;
    FEDICRA_REQUIRE(blocks <= 0x7fffffffLL, "ala_update_flat: array too large");
    ala_flat_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p, s, g, w, t, n, chunk, eta);
    FEDICRA_CHECK_LAUNCH("ala_flat_kernel");
    return 0;
}

extern "C" int fedicra_ala_update_multi(const float *const *p, const float *const *s, const float *const *g,
                                        float *const *w, float *const *t, const int64_t *sizes,
                                        const int *chunk_tensor, const int64_t *chunk_begin, int n_chunks,
                                        int chunk_elems, float eta, void *stream) {
    FEDICRA_REQUIRE(p && s && g && w && t && sizes && chunk_tensor && chunk_begin, "ala_update_multi: null pointer");
    FEDICRA_REQUIRE(n_chunks >= 0 && chunk_elems > 0 && chunk_elems % 4 == 0, "ala_update_multi: bad chunking");
    if (n_chunks == 0) return 0;
    ala_multi_kernel<<<n_chunks, 256, 0, (cudaStream_t)stream>>>(p, s, g, w, t, (const long long *)sizes, chunk_tensor,
                                                                 (const long long *)chunk_begin, chunk_elems, eta);
    FEDICRA_CHECK_LAUNCH("ala_multi_kernel");
    return 0;
}

extern "C" int fedicra_scale_inplace(float *x, int64_t n, float alpha, void *stream) {
    FEDICRA_REQUIRE(x || n == 0, "scale_inplace: null pointer");
    FEDICRA_REQUIRE(n >= 0, "scale_inplace: negative size");
    if (n == 0) return 0;
    const long long blocks = ((n + 3) / 4 + 255) / 256;
    scale_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, n, alpha);
    FEDICRA_CHECK_LAUNCH("scale_kernel");
    return 0;
}
