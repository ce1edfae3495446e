// glue.cu -- the element-wise ends of the tree-energy loss, fused (sm_100a).
//
// What it replaces (reference = /root/reference/code/flower_common.py, class MScaleRecurveTreeEnergyLoss and siblings):
//   :781            prob = torch.softmax(preds, dim=1)                      -> channel_softmax_{fwd,bwd}_kernel
//   :777, :813-816  N = ROI.sum();  loss = (ROI * |prob - AS|).sum();  if N > 0: loss /= N;  weight * loss
//                                                                           -> masked_l1_{fwd,bwd}_kernel
// In PyTorch these are ~20 small kernels per step (sub, abs, mul, sum, clamp, div, sign, neg, ...), each a full HBM round
// trip of a [B,C,H,W] tensor; here each direction is one pass at HBM rate.  The reductions are deterministic: per-CTA
// partial sums in a fixed order, combined by the last CTA to arrive in CTA-index order.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int NT = 256;
constexpr int MAXC = 16;

// prob[b,c,p] = exp(x[b,c,p] - max_c) / sum_c exp(x[b,c,p] - max_c); one thread per pixel, channels strided by P.
__global__ void __launch_bounds__(NT)
channel_softmax_fwd_kernel(const float *__restrict__ x, float *__restrict__ y, int C, long long P, long long BP) {
    const long long i = (long long)blockIdx.x * NT + threadIdx.x;
    if (i >= BP) return;
    const long long b = i / P, p = i - b * P;
    const float *xi = x + b * C * P + p;
    float *yi = y + b * C * P + p;
    float v[MAXC];
    float m = -INFINITY;
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
        if (c < C) {
            v[c] = xi[(long long)c * P];
            m = fmaxf(m, v[c]);
        }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
        if (c < C) {
            v[c] = expf(v[c] - m);
            s += v[c];
        }
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
        if (c < C) yi[(long long)c * P] = v[c] / s;
}

// gx[c] = y[c] * (gy[c] - sum_k gy[k] * y[k])
__global__ void __launch_bounds__(NT)
channel_softmax_bwd_kernel(const float *__restrict__ y, const float *__restrict__ gy, float *__restrict__ gx, int C,
                           long long P, long long BP) {
    const long long i = (long long)blockIdx.x * NT + threadIdx.x;
    if (i >= BP) return;
    const long long b = i / P, p = i - b * P, base = b * C * P + p;
    float yv[MAXC], gv[MAXC];
    float dot = 0.f;
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
        if (c < C) {
            yv[c] = y[base + (long long)c * P];
            gv[c] = gy[base + (long long)c * P];
            dot = fmaf(gv[c], yv[c], dot);
        }
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
        if (c < C) gx[base + (long long)c * P] = yv[c] * (gv[c] - dot);
}

__device__ __forceinline__ float block_sum(float v, float *s_red) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    float t = 0.f;
    if (wid == 0) {
        t = (lane < NT / 32) ? s_red[lane] : 0.f;
        t = warp_sum(t);
    }
    __syncthreads();
    return t;  // valid in warp 0
}

// sums[0] = sum_{b,c,p} roi[b,p] * |prob - filt|, sums[1] = sum_{b,p} roi[b,p];
// loss[0] = weight * sums[0] / max(sums[1], 1)      (flower_common.py:813-818; the `if N > 0` without its host sync)
// scratch: [2*gridDim.x] partials + 1 counter word (zero before the first use, left at zero).
__global__ void __launch_bounds__(NT)
masked_l1_fwd_kernel(const float *__restrict__ prob, const float *__restrict__ filt, const unsigned char *__restrict__ roi,
                     int C, long long P, long long BP, float weight, float *__restrict__ sums, float *__restrict__ loss,
                     float *__restrict__ partial, unsigned *__restrict__ counter) {
    __shared__ float s_red[NT / 32];
    __shared__ bool s_last;
    float acc = 0.f, cnt = 0.f;
    for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < BP; i += (long long)gridDim.x * NT) {
        if (!roi[i]) continue;
        const long long b = i / P, p = i - b * P, base = b * C * P + p;
        cnt += 1.f;
        for (int c = 0; c < C; ++c) acc += fabsf(prob[base + (long long)c * P] - filt[base + (long long)c * P]);
    }
    const float a = block_sum(acc, s_red), n = block_sum(cnt, s_red);
    if (threadIdx.x == 0) {
        partial[2 * blockIdx.x] = a;
        partial[2 * blockIdx.x + 1] = n;
        __threadfence();
        s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    float ta = 0.f, tn = 0.f;   // fixed order: thread t sums partials t, t+NT, ...; then the block tree
    for (int k = threadIdx.x; k < (int)gridDim.x; k += NT) {
        ta += __ldcg(partial + 2 * k);
        tn += __ldcg(partial + 2 * k + 1);
    }
    ta = block_sum(ta, s_red);
    tn = block_sum(tn, s_red);
    if (threadIdx.x == 0) {
        sums[0] = ta;
        sums[1] = tn;
        loss[0] = weight * ta / fmaxf(tn, 1.f);
        *counter = 0u;
    }
}

// g_prob = s * roi * sign(prob - filt), g_filt = -g_prob, with s = gout[0] * weight / max(n[0], 1)  (n == NULL: s = gout[0])
__global__ void __launch_bounds__(NT)
masked_l1_bwd_kernel(const float *__restrict__ prob, const float *__restrict__ filt, const unsigned char *__restrict__ roi,
                     const float *__restrict__ gout, const float *__restrict__ n, float weight, int C, long long P,
                     long long BP, float *__restrict__ g_prob, float *__restrict__ g_filt) {
    const long long i = (long long)blockIdx.x * NT + threadIdx.x;
    if (i >= BP) return;
    const float s = gout[0] * (n ? weight / fmaxf(n[0], 1.f) : 1.f);
    const long long b = i / P, p = i - b * P, base = b * C * P + p;
    const bool on = roi[i] != 0;
    for (int c = 0; c < C; ++c) {
        const long long j = base + (long long)c * P;
        const float d = prob[j] - filt[j];
        const float g = on ? (d > 0.f ? s : (d < 0.f ? -s : 0.f)) : 0.f;   // torch.sign(0) = 0
        if (g_prob) g_prob[j] = g;
        if (g_filt) g_filt[j] = -g;
    }
}

}  // namespace

extern "C" int fedicra_channel_softmax_fwd(const float *x, int B, int C, long long P, float *y, void *stream) {
    FEDICRA_REQUIRE(x && y, "channel_softmax_fwd: null pointer");
    FEDICRA_REQUIRE(B > 0 && C > 0 && C <= MAXC && P > 0, "channel_softmax_fwd: bad sizes B=%d C=%d P=%lld (C <= %d)", B, C, P, MAXC);
    const long long BP = (long long)B * P;
    channel_softmax_fwd_kernel<<<(unsigned)((BP + NT - 1) / NT), NT, 0, (cudaStream_t)stream>>>(x, y, C, P, BP);
    FEDICRA_CHECK_LAUNCH("channel_softmax_fwd_kernel");
    return 0;
}

extern "C" int fedicra_channel_softmax_bwd(const float *y, const float *grad_y, int B, int C, long long P, float *grad_x,
                                           void *stream) {
    FEDICRA_REQUIRE(y && grad_y && grad_x, "channel_softmax_bwd: null pointer");
    FEDICRA_REQUIRE(B > 0 && C > 0 && C <= MAXC && P > 0, "channel_softmax_bwd: bad sizes B=%d C=%d P=%lld (C <= %d)", B, C, P, MAXC);
    const long long BP = (long long)B * P;
    channel_softmax_bwd_kernel<<<(unsigned)((BP + NT - 1) / NT), NT, 0, (cudaStream_t)stream>>>(y, grad_y, grad_x, C, P, BP);
    FEDICRA_CHECK_LAUNCH("channel_softmax_bwd_kernel");
    return 0;
}

constexpr int L1_GRID = 592;   // 4 CTAs per SM

extern "C" size_t fedicra_masked_l1_workspace_bytes(void) { return (2 * L1_GRID + 4) * sizeof(float); }

extern "C" int fedicra_masked_l1_fwd(const float *prob, const float *filtered, const unsigned char *roi, int B, int C,
                                     long long P, float weight, float *sums, float *loss, void *workspace,
                                     size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(prob && filtered && roi && sums && loss && workspace, "masked_l1_fwd: null pointer");
    FEDICRA_REQUIRE(B > 0 && C > 0 && P > 0, "masked_l1_fwd: bad sizes B=%d C=%d P=%lld", B, C, P);
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_masked_l1_workspace_bytes(), "masked_l1_fwd: workspace too small");
    const long long BP = (long long)B * P;
    const long long want = (BP + NT - 1) / NT;
    const int grid = (int)(want < L1_GRID ? want : L1_GRID);
    float *partial = (float *)workspace;
    unsigned *counter = (unsigned *)(partial + 2 * L1_GRID);
    masked_l1_fwd_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(prob, filtered, roi, C, P, BP, weight, sums, loss, partial, counter);
    FEDICRA_CHECK_LAUNCH("masked_l1_fwd_kernel");
    return 0;
}

extern "C" int fedicra_masked_l1_bwd(const float *prob, const float *filtered, const unsigned char *roi,
                                     const float *grad_out, const float *n, float weight, int B, int C, long long P,
                                     float *grad_prob, float *grad_filtered, void *stream) {
    FEDICRA_REQUIRE(prob && filtered && roi && grad_out && (grad_prob || grad_filtered), "masked_l1_bwd: null pointer");
    FEDICRA_REQUIRE(B > 0 && C > 0 && P > 0, "masked_l1_bwd: bad sizes B=%d C=%d P=%lld", B, C, P);
    const long long BP = (long long)B * P;
    masked_l1_bwd_kernel<<<(unsigned)((BP + NT - 1) / NT), NT, 0, (cudaStream_t)stream>>>(prob, filtered, roi, grad_out, n, weight, C,
                                                                                       P, BP, grad_prob, grad_filtered);
    FEDICRA_CHECK_LAUNCH("masked_l1_bwd_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Bilinear resize of the decoder heads to the prediction size, F.interpolate(mode='bilinear', align_corners=False)
// (flower_common.py:789,796,803), and its gradient.  The forward follows torch's CUDA kernel expression for expression
// (source index = max(0, scale * (dst + 0.5) - 0.5), scale = in / out; val = h0 * (w0 * v00 + w1 * v01) + h1 * (w0 * v10
// + w1 * v11)) because the resized head is the embedding the MST is built on and the tree is decided on its exact bits
// (tests compare the outputs bit for bit with torch's).  Several heads share one launch (pointer table in the
// arguments).  The gradient is a gather over the output pixels that read an input pixel -- fixed summation order,
// where torch scatters with atomics.
namespace {

constexpr int RESIZE_MAX = 4;
struct ResizeSet {
    const float *src[RESIZE_MAX];
    float *dst[RESIZE_MAX];
    int h[RESIZE_MAX], w[RESIZE_MAX];
    long long begin[RESIZE_MAX + 1];   // element range of every head in the flat index space of the launch
    int n;
};

// a * x + b * y the way torch's kernel comes out of nvcc: the second product rounded, then one fused multiply-add
// (of the nine ways to round the interpolation expression this is the one whose results equal torch's bit for bit;
// scripts/probe/resize_variants.py tried them all)
__device__ __forceinline__ float combine2(float a, float x, float b, float y) { return __fmaf_rn(a, x, __fmul_rn(b, y)); }

struct Tap {
    int i0, ip;      // first source index, 0/1 step to the second
    float l0, l1;    // weights of the two
};
__device__ __forceinline__ Tap resize_tap(int dst, float scale, int in_size) {
    float r = scale * ((float)dst + 0.5f) - 0.5f;
    r = r < 0.f ? 0.f : r;
    Tap t;
    t.i0 = (int)r;
    t.ip = (t.i0 < in_size - 1) ? 1 : 0;
    t.l1 = r - (float)t.i0;
    t.l0 = 1.f - t.l1;
    return t;
}

__global__ void __launch_bounds__(NT) bilinear_fwd_kernel(ResizeSet rs, int planes, int H, int W) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= rs.begin[rs.n]) return;
    int k = 0;
    while (k + 1 < rs.n && idx >= rs.begin[k + 1]) ++k;
    const long long e = idx - rs.begin[k];
    const int x = (int)(e % W), y = (int)((e / W) % H);
    const long long plane = e / ((long long)W * H);
    const int h = rs.h[k], w = rs.w[k];
    const Tap ty = resize_tap(y, (float)h / (float)H, h), tx = resize_tap(x, (float)w / (float)W, w);
    const float *s = rs.src[k] + plane * h * w;
    const float v00 = s[ty.i0 * w + tx.i0], v01 = s[ty.i0 * w + tx.i0 + tx.ip];
    const float v10 = s[(ty.i0 + ty.ip) * w + tx.i0], v11 = s[(ty.i0 + ty.ip) * w + tx.i0 + tx.ip];
    rs.dst[k][e] = combine2(ty.l0, combine2(tx.l0, v00, tx.l1, v01), ty.l1, combine2(tx.l0, v10, tx.l1, v11));
}

// src = gradient w.r.t. the resized head [planes,H,W]; dst = gradient w.r.t. the head [planes,h,w]
__global__ void __launch_bounds__(NT) bilinear_bwd_kernel(ResizeSet rs, int planes, int H, int W) {
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    if (idx >= rs.begin[rs.n]) return;
    int k = 0;
    while (k + 1 < rs.n && idx >= rs.begin[k + 1]) ++k;
    const long long e = idx - rs.begin[k];
    const int h = rs.h[k], w = rs.w[k];
    const int x = (int)(e % w), y = (int)((e / w) % h);
    const long long plane = e / ((long long)w * h);
    const float sy = (float)h / (float)H, sx = (float)w / (float)W;
    // output rows / columns whose taps can touch (y, x): source index in (y - 1, y + 1)
    const int y_lo = max(0, (int)floorf(((float)y - 1.f + 0.5f) / sy - 0.5f) - 1), y_hi = min(H - 1, (int)ceilf(((float)y + 1.f + 0.5f) / sy - 0.5f) + 1);
    const int x_lo = max(0, (int)floorf(((float)x - 1.f + 0.5f) / sx - 0.5f) - 1), x_hi = min(W - 1, (int)ceilf(((float)x + 1.f + 0.5f) / sx - 0.5f) + 1);
    const float *g = rs.src[k] + plane * H * W;
    float acc = 0.f;
    constexpr int XW = 16;                      // x weights of up to 16 output columns kept in registers (x4 needs 12)
    float wxs[XW];
    const bool cached = x_hi - x_lo < XW;
    if (cached) {
#pragma unroll
        for (int t = 0; t < XW; ++t) {
            const int xx = x_lo + t;
            float wx = 0.f;
            if (xx <= x_hi) {
                const Tap tx = resize_tap(xx, sx, w);
                wx = (tx.i0 == x ? tx.l0 : 0.f) + (tx.i0 + tx.ip == x ? tx.l1 : 0.f);
            }
            wxs[t] = wx;
        }
    }
    for (int yy = y_lo; yy <= y_hi; ++yy) {
        const Tap ty = resize_tap(yy, sy, h);
        const float wy = (ty.i0 == y ? ty.l0 : 0.f) + (ty.i0 + ty.ip == y ? ty.l1 : 0.f);
        if (wy == 0.f) continue;
        float row = 0.f;
        if (cached) {
#pragma unroll
            for (int t = 0; t < XW; ++t)
                if (wxs[t] != 0.f) row = fmaf(wxs[t], g[(long long)yy * W + x_lo + t], row);
        } else {
            for (int xx = x_lo; xx <= x_hi; ++xx) {
                const Tap tx = resize_tap(xx, sx, w);
                const float wx = (tx.i0 == x ? tx.l0 : 0.f) + (tx.i0 + tx.ip == x ? tx.l1 : 0.f);
                if (wx != 0.f) row = fmaf(wx, g[(long long)yy * W + xx], row);
            }
        }
        acc = fmaf(wy, row, acc);
    }
    rs.dst[k][e] = acc;
}

}  // namespace

extern "C" int fedicra_bilinear_resize_fwd(const float *const *h_src, float *const *h_dst, const int *h_h, const int *h_w,
                                           int n, int planes, int H, int W, void *stream) {
    FEDICRA_REQUIRE(h_src && h_dst && h_h && h_w, "bilinear_resize_fwd: null pointer");
    FEDICRA_REQUIRE(n >= 1 && n <= RESIZE_MAX && planes > 0 && H > 0 && W > 0, "bilinear_resize_fwd: bad sizes n=%d (1..%d) planes=%d H=%d W=%d", n, RESIZE_MAX, planes, H, W);
    ResizeSet rs = {};
    rs.n = n;
    for (int k = 0; k < n; ++k) {
        FEDICRA_REQUIRE(h_src[k] && h_dst[k] && h_h[k] > 0 && h_w[k] > 0, "bilinear_resize_fwd: head %d is null or empty", k);
        rs.src[k] = h_src[k]; rs.dst[k] = h_dst[k]; rs.h[k] = h_h[k]; rs.w[k] = h_w[k];
        rs.begin[k + 1] = rs.begin[k] + (long long)planes * H * W;
    }
    bilinear_fwd_kernel<<<(unsigned)((rs.begin[n] + NT - 1) / NT), NT, 0, (cudaStream_t)stream>>>(rs, planes, H, W);
    FEDICRA_CHECK_LAUNCH("bilinear_fwd_kernel");
    return 0;
}

extern "C" int fedicra_bilinear_resize_bwd(const float *const *h_grad_out, float *const *h_grad_in, const int *h_h,
                                           const int *h_w, int n, int planes, int H, int W, void *stream) {
    FEDICRA_REQUIRE(h_grad_out && h_grad_in && h_h && h_w, "bilinear_resize_bwd: null pointer");
    FEDICRA_REQUIRE(n >= 1 && n <= RESIZE_MAX && planes > 0 && H > 0 && W > 0, "bilinear_resize_bwd: bad sizes n=%d (1..%d) planes=%d H=%d W=%d", n, RESIZE_MAX, planes, H, W);
    ResizeSet rs = {};
    rs.n = n;
    for (int k = 0; k < n; ++k) {
        FEDICRA_REQUIRE(h_grad_out[k] && h_grad_in[k] && h_h[k] > 0 && h_w[k] > 0, "bilinear_resize_bwd: head %d is null or empty", k);
        rs.src[k] = h_grad_out[k]; rs.dst[k] = h_grad_in[k]; rs.h[k] = h_h[k]; rs.w[k] = h_w[k];
        rs.begin[k + 1] = rs.begin[k] + (long long)planes * h_h[k] * h_w[k];
    }
    bilinear_bwd_kernel<<<(unsigned)((rs.begin[n] + NT - 1) / NT), NT, 0, (cudaStream_t)stream>>>(rs, planes, H, W);
    FEDICRA_CHECK_LAUNCH("bilinear_bwd_kernel");
    return 0;
}
