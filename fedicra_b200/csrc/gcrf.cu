// gcrf.cu -- Gated CRF loss, fused forward + backward (sm_100a).
//
// Replaces ModelLossSemsegGatedCRF.forward of /root/reference/code/utils/gate_crf_loss.py:20-124 for its live branch
// (no masks, Potts compatibility): the reference materialises F.unfold windows of the features, the kernel values and
// the predictions ([N,*,11,11,H,W] tensors, :89-93, :163-172); here a CTA stages a tile plus halo of the image and of
// the predictions in shared memory and every thread walks the (2R+1)^2 window of 4 neighbouring pixels in registers.
//
//   K_pq   = sum_k w_k * exp(-0.5 * (|xy_p - xy_q|^2 / sxy_k^2 + |I_p - I_q|^2 / srgb_k^2)),  q in window(p) \ {p}
//   agg_p  = sum_q K_pq * y_q            loss = (sum_pq K_pq - sum_p y_p . agg_p) / (N*H*W)
//   d loss / d y_p = -2 agg_p / (N*H*W)  (K is symmetric)
//
// Zero padding quirk kept: F.unfold pads the FEATURES with zeros (:183-188), so a tap q outside the image has
// xy_q = 0, I_q = 0, y_q = 0; it adds exp(-0.5*|f_p|^2) to sum K and nothing to agg.
//
// Roofline: 120 ex2 + ~120*(2Cin+C+3) FP32 ops per pixel against (2C+Cin)*4 bytes -> MUFU/FP32-pipe bound, not HBM.
#include <cuda.h>   // CUtensorMap (types only: the encoder is looked up through the runtime, no -lcuda)
#include <mutex>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace {

constexpr int TW = 32;          // tile width  (pixels)
constexpr int TH = 16;          // tile height (pixels)
constexpr int PX = 4;           // pixels per thread along x
constexpr int BX = TW / PX;     // 8 threads along x
constexpr int NT = BX * TH;     // 128 threads
constexpr int XOFF = 8;         // halo columns kept left of the tile in shared memory: a multiple of 4 floats, because a
                               // tensor-map copy must start on a 16-byte boundary of the image row (the window needs 5)
constexpr float LOG2E = 1.4426950408889634f;

struct FastParams {
    float axy;   // -0.5*log2(e)/sigma_xy^2   (0 when xy is not used)
    float argb;  // -0.5*log2(e)/sigma_rgb^2  (0 when rgb is not used)
    float lw;    // log2(weight)
};

// Workspace: [0] unsigned ticket counter (self-resetting), [64..) one float partial per CTA.
__device__ __forceinline__ void finish_loss(float thread_part, float *partials, unsigned *ticket, float *loss,
                                            double inv_denom) {
    __shared__ float s_warp[32];
    __shared__ double s_dbl[32];
    __shared__ bool s_last;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = (blockDim.x + 31) >> 5;
    float v = warp_sum(thread_part);
    if (lane == 0) s_warp[wid] = v;
    __syncthreads();
    const unsigned nblocks = gridDim.x * gridDim.y * gridDim.z;
    const unsigned bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (tid == 0) {
        float s = 0.f;
        for (int i = 0; i < nw; ++i) s += s_warp[i];
        partials[bid] = s;
        __threadfence();
        unsigned t = atomicAdd(ticket, 1u);
        s_last = (t == nblocks - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // last CTA: fixed-order reduction of all partials in double
    double acc = 0.0;
    for (unsigned i = tid; i < nblocks; i += blockDim.x) acc += (double)__ldcg(partials + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) s_dbl[wid] = acc;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int i = 0; i < nw; ++i) s += s_dbl[i];
        loss[0] = (float)(s * inv_denom);
        *ticket = 0u;  // leave the workspace reusable
    }
}

template <int C, int CIN, int R, bool BORDER>
__device__ __forceinline__ float gcrf_tile_body(const float *s_img, const float *s_y, float *grad, int n, int x0,
                                                int y0, int H, int W, FastParams prm, float gscale) {
    constexpr int HH = TH + 2 * R;
    constexpr int PITCH = TW + 2 * XOFF;
    constexpr int SHIFT = XOFF - R;               // the window row segment starts SHIFT columns into an aligned read
    constexpr int NV4 = (PX + 2 * R + SHIFT) / 4; // whole float4s of that read; one scalar column follows
    constexpr int NV = 4 * NV4 + 1;
    static_assert(XOFF >= R && NV >= PX + 2 * R + SHIFT, "window segment does not fit the aligned read");
    const int tid = threadIdx.x, tx = tid % BX, ty = tid / BX;
    const int px0 = x0 + tx * PX, py = y0 + ty;

    float ip[CIN][PX];
#pragma unroll
    for (int c = 0; c < CIN; ++c)
#pragma unroll
        for (int p = 0; p < PX; ++p) ip[c][p] = s_img[(c * HH + ty + R) * PITCH + tx * PX + XOFF + p];

    float agg[C][PX], ks[PX], oob[PX];
#pragma unroll
    for (int p = 0; p < PX; ++p) {
        ks[p] = 0.f;
        const float fx = (float)(px0 + p), fy = (float)py;
        oob[p] = prm.axy * (fx * fx + fy * fy) + prm.lw;
#pragma unroll
        for (int c = 0; c < C; ++c) agg[c][p] = 0.f;
    }

    // per-|dx| part of the xy exponent
    float cdx[R + 1];
#pragma unroll
    for (int i = 0; i <= R; ++i) cdx[i] = prm.axy * (float)(i * i) + prm.lw;

#pragma unroll 1
    for (int dy = -R; dy <= R; ++dy) {
        const int row = ty + R + dy;
        float si[CIN][NV], sy[C][NV];
#pragma unroll
        for (int c = 0; c < CIN; ++c) {
            const float4 *src = reinterpret_cast<const float4 *>(s_img + (c * HH + row) * PITCH + tx * PX);
#pragma unroll
            for (int v = 0; v < NV4; ++v) {
                float4 t = src[v];
                si[c][4 * v] = t.x; si[c][4 * v + 1] = t.y; si[c][4 * v + 2] = t.z; si[c][4 * v + 3] = t.w;
            }
            si[c][4 * NV4] = s_img[(c * HH + row) * PITCH + tx * PX + 4 * NV4];
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const float4 *src = reinterpret_cast<const float4 *>(s_y + (c * HH + row) * PITCH + tx * PX);
#pragma unroll
            for (int v = 0; v < NV4; ++v) {
                float4 t = src[v];
                sy[c][4 * v] = t.x; sy[c][4 * v + 1] = t.y; sy[c][4 * v + 2] = t.z; sy[c][4 * v + 3] = t.w;
            }
            sy[c][4 * NV4] = s_y[(c * HH + row) * PITCH + tx * PX + 4 * NV4];
        }
        const float rterm = prm.axy * (float)(dy * dy);
        const bool row_in = BORDER ? (py + dy >= 0 && py + dy < H) : true;
#pragma unroll
        for (int dx = -R; dx <= R; ++dx) {
            const float base_in = rterm + cdx[dx < 0 ? -dx : dx];
#pragma unroll
            for (int p = 0; p < PX; ++p) {
                const int q = p + dx + R + SHIFT;
                float d = 0.f;
#pragma unroll
                for (int c = 0; c < CIN; ++c) {
                    const float t = si[c][q] - ip[c][p];
                    d = fmaf(t, t, d);
                }
                float base = base_in;
                if (BORDER) {
                    const int qx = px0 + p + dx;
                    base = (row_in && qx >= 0 && qx < W) ? base_in : oob[p];
                }
                float k = ex2_approx(fmaf(prm.argb, d, base));
                if (dx == 0) k = (dy == 0) ? 0.f : k;  // the centre tap is excluded (gate_crf_loss.py:171)
                ks[p] += k;
#pragma unroll
                for (int c = 0; c < C; ++c) agg[c][p] = fmaf(k, sy[c][q], agg[c][p]);
            }
        }
    }

    // epilogue: gradient store + this thread's share of the loss
    float part = 0.f;
    const bool row_ok = py < H;
#pragma unroll
    for (int p = 0; p < PX; ++p) {
        const bool ok = row_ok && (px0 + p < W);
        float dot = 0.f;
#pragma unroll
        for (int c = 0; c < C; ++c) dot = fmaf(s_y[(c * HH + ty + R) * PITCH + tx * PX + XOFF + p], agg[c][p], dot);
        part += ok ? (ks[p] - dot) : 0.f;
    }
    if (grad != nullptr && row_ok) {
#pragma unroll
        for (int c = 0; c < C; ++c) {
            float *dst = grad + (((size_t)n * C + c) * H + py) * W + px0;
            if ((W & 3) == 0 && px0 + PX <= W) {
                *reinterpret_cast<float4 *>(dst) =
                    make_float4(agg[c][0] * gscale, agg[c][1] * gscale, agg[c][2] * gscale, agg[c][3] * gscale);
            } else {
#pragma unroll
                for (int p = 0; p < PX; ++p)
                    if (px0 + p < W) dst[p] = agg[c][p] * gscale;
            }
        }
    }
    return part;
}

// ---- TMA staging.  The image tile and the prediction tile (each with its halo) are ONE 3-D box each of the tensors
// viewed as [planes, H, W]: box = (PITCH, HH, channels), start = (x0 - XOFF, y0 - R, n * channels).  The tensor-map unit
// zero-fills whatever lies outside the image -- exactly the zero padding of F.unfold (gate_crf_loss.py:183-188) -- and
// lands the box in shared memory in the [channel][row][PITCH] layout the tile body reads.  One thread arms an mbarrier
// with the byte count and issues the two copies; everybody waits on the barrier.  Needs W % 4 == 0 (16-byte row pitch).
__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, int c0, int c1, int c2, unsigned long long *bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_addr(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_addr(bar))
        : "memory");
}

template <int C, int CIN, int R, bool TMA>
__global__ void __launch_bounds__(NT, 3)
gcrf_fast_kernel(const float *__restrict__ y, const float *__restrict__ img, float *__restrict__ grad,
                 float *partials, unsigned *ticket, float *loss, int H, int W, FastParams prm, float gscale,
                 double inv_denom, const __grid_constant__ CUtensorMap map_img, const __grid_constant__ CUtensorMap map_y) {
    constexpr int HH = TH + 2 * R;
    constexpr int PITCH = TW + 2 * XOFF;
    constexpr int IMG_FLOATS = (CIN * HH * PITCH + 31) / 32 * 32;   // the second box starts 128-byte aligned
    extern __shared__ __align__(128) float4 smem4[];
    float *s_img = reinterpret_cast<float *>(smem4);  // [CIN][HH][PITCH]
    float *s_y = s_img + IMG_FLOATS;                   // [C][HH][PITCH]
    const int n = blockIdx.z, x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;

    if (TMA) {
        __shared__ __align__(8) unsigned long long s_bar;
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&s_bar)) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            constexpr unsigned BYTES = (unsigned)((C + CIN) * HH * PITCH * sizeof(float));
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&s_bar)), "r"(BYTES) : "memory");
            tma_load_3d(s_img, &map_img, x0 - XOFF, y0 - R, n * CIN, &s_bar);
            tma_load_3d(s_y, &map_y, x0 - XOFF, y0 - R, n * C, &s_bar);
        }
        unsigned done = 0;
        for (unsigned spins = 0; !done; ++spins) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(smem_addr(&s_bar)) : "memory");
            if (spins > (1u << 26)) __trap();   // a copy that never lands (bad tensor map): fail the launch, do not hang
        }
    } else {
        // stage tile + halo, zero outside the image (coalesced along x)
        for (int idx = threadIdx.x; idx < (CIN + C) * HH * PITCH; idx += NT) {
            const int ch = idx / (HH * PITCH), rem = idx - ch * (HH * PITCH), r = rem / PITCH, c = rem - r * PITCH;
            const int gy = y0 - R + r, gx = x0 - XOFF + c;
            float v = 0.f;
            if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
                v = (ch < CIN) ? __ldg(img + (((size_t)n * CIN + ch) * H + gy) * W + gx)
                               : __ldg(y + (((size_t)n * C + (ch - CIN)) * H + gy) * W + gx);
            }
            (ch < CIN ? s_img : s_y - CIN * HH * PITCH)[idx] = v;
        }
        __syncthreads();
    }

    const bool border = (x0 < R) || (y0 < R) || (x0 + TW + R > W) || (y0 + TH + R > H);
    float part;
    if (border)
        part = gcrf_tile_body<C, CIN, R, true>(s_img, s_y, grad, n, x0, y0, H, W, prm, gscale);
    else
        part = gcrf_tile_body<C, CIN, R, false>(s_img, s_y, grad, n, x0, y0, H, W, prm, gscale);
    finish_loss(part, partials, ticket, loss, inv_denom);
}

// Any radius / channel count / number of descriptors: one pixel per thread straight from global memory.
struct GenericDesc {
    float w, axy, argb;
};
constexpr int MAX_DESC = 8;
struct GenericParams {
    GenericDesc d[MAX_DESC];
    int n;
};

__global__ void __launch_bounds__(128)
gcrf_generic_kernel(const float *__restrict__ y, const float *__restrict__ img, float *__restrict__ grad,
                    float *partials, unsigned *ticket, float *loss, int N, int C, int CIN, int H, int W, int R,
                    GenericParams prm, float gscale, double inv_denom) {
    const size_t npix = (size_t)N * H * W;
    const size_t pix = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    float part = 0.f;
    if (pix < npix) {
        const int n = (int)(pix / ((size_t)H * W));
        const int rem = (int)(pix - (size_t)n * H * W), py = rem / W, px = rem - py * W;
        const float *im = img + (size_t)n * CIN * H * W;
        const float *yy = y + (size_t)n * C * H * W;
        float *g = grad ? grad + (size_t)n * C * H * W : nullptr;
        float ksum = 0.f, dot = 0.f;
        // channel-outer accumulation keeps the state scalar; K is recomputed per class channel (fallback path)
        for (int c = -1; c < C; ++c) {
            float acc = 0.f;
            for (int dy = -R; dy <= R; ++dy)
                for (int dx = -R; dx <= R; ++dx) {
                    if (dy == 0 && dx == 0) continue;
                    const int qy = py + dy, qx = px + dx;
                    const bool in = qy >= 0 && qy < H && qx >= 0 && qx < W;
                    float drgb = 0.f;
                    for (int ch = 0; ch < CIN; ++ch) {
                        const float a = im[((size_t)ch * H + py) * W + px];
                        const float b = in ? im[((size_t)ch * H + qy) * W + qx] : 0.f;
                        drgb = fmaf(a - b, a - b, drgb);
                    }
                    const float dxy = in ? (float)(dx * dx + dy * dy) : (float)px * (float)px + (float)py * (float)py;
                    float k = 0.f;
                    for (int i = 0; i < prm.n; ++i)
                        k = fmaf(prm.d[i].w, ex2_approx(fmaf(prm.d[i].axy, dxy, prm.d[i].argb * drgb)), k);
                    if (c < 0)
                        acc += k;
                    else if (in)
                        acc = fmaf(k, yy[((size_t)c * H + qy) * W + qx], acc);
                }
            if (c < 0) {
                ksum = acc;
            } else {
                dot = fmaf(yy[((size_t)c * H + py) * W + px], acc, dot);
                if (g) g[((size_t)c * H + py) * W + px] = acc * gscale;
            }
        }
        part = ksum - dot;
    }
    finish_loss(part, partials, ticket, loss, inv_denom);
}

// cuTensorMapEncodeTiled, looked up once through the runtime (the library does not link libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn tensor_map_encoder() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *off = getenv("FEDICRA_GCRF_TMA");
        if (off && *off == '0') return;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// [planes, H, W] fp32 tensor, box (pitch, rows, channels), no swizzle, zeros outside
bool make_tile_map(CUtensorMap *m, const float *base, int planes, int H, int W, int pitch, int rows, int channels) {
    EncodeTiledFn enc = tensor_map_encoder();
    if (enc == nullptr || (W & 3) != 0 || ((uintptr_t)base & 15) != 0) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
    const cuuint32_t box[3] = {(cuuint32_t)pitch, (cuuint32_t)rows, (cuuint32_t)channels};
    const cuuint32_t estr[3] = {1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int C, int CIN>
int launch_fast(const float *y, const float *img, float *grad, float *partials, unsigned *ticket, float *loss, int N,
                int H, int W, FastParams prm, float gscale, double inv_denom, cudaStream_t st) {
    constexpr int R = 5;
    constexpr int HH = TH + 2 * R;
    constexpr int PITCH = TW + 2 * XOFF;
    constexpr int IMG_FLOATS = (CIN * HH * PITCH + 31) / 32 * 32;
    const size_t smem = (size_t)(IMG_FLOATS + C * HH * PITCH) * sizeof(float);
    dim3 grid((W + TW - 1) / TW, (H + TH - 1) / TH, N);
    CUtensorMap map_img, map_y;
    memset(&map_img, 0, sizeof(map_img));
    memset(&map_y, 0, sizeof(map_y));
    if (make_tile_map(&map_img, img, N * CIN, H, W, PITCH, HH, CIN) && make_tile_map(&map_y, y, N * C, H, W, PITCH, HH, C))
        gcrf_fast_kernel<C, CIN, R, true><<<grid, NT, smem, st>>>(y, img, grad, partials, ticket, loss, H, W, prm, gscale,
                                                                 inv_denom, map_img, map_y);
    else   // rows that are not a multiple of 16 bytes, or no tensor-map encoder: stage with ordinary loads
        gcrf_fast_kernel<C, CIN, R, false><<<grid, NT, smem, st>>>(y, img, grad, partials, ticket, loss, H, W, prm, gscale,
                                                                  inv_denom, map_img, map_y);
    return 0;
}

size_t n_blocks_upper(int N, int H, int W) {
    const size_t tiles = (size_t)((W + TW - 1) / TW) * ((H + TH - 1) / TH) * N;
    const size_t generic = ((size_t)N * H * W + 127) / 128;
    return tiles > generic ? tiles : generic;
}

}  // namespace

extern "C" size_t fedicra_gcrf_workspace_bytes(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    return 64 + align_up(n_blocks_upper(N, H, W) * sizeof(float), 64);
}

extern "C" int fedicra_gcrf_fwd_bwd(const float *y, const float *image, int N, int C, int Cin, int H, int W,
                                    int radius, const fedicra_gcrf_desc *h_descs, int n_desc, float *loss,
                                    float *grad_y, void *workspace, size_t workspace_bytes, void *stream) {
    FEDICRA_REQUIRE(y && image && loss && workspace && h_descs, "gcrf_fwd_bwd: null pointer");
    FEDICRA_REQUIRE(N > 0 && C > 0 && Cin > 0 && H > 0 && W > 0, "gcrf_fwd_bwd: bad shape N=%d C=%d Cin=%d H=%d W=%d", N,
                    C, Cin, H, W);
    FEDICRA_REQUIRE(radius >= 1 && radius <= 32, "gcrf_fwd_bwd: radius %d outside [1,32]", radius);
    FEDICRA_REQUIRE(n_desc >= 1 && n_desc <= MAX_DESC, "gcrf_fwd_bwd: %d kernel descriptors (1..%d supported)", n_desc,
                    MAX_DESC);
    FEDICRA_REQUIRE(workspace_bytes >= fedicra_gcrf_workspace_bytes(N, H, W), "gcrf_fwd_bwd: workspace too small");
    FEDICRA_REQUIRE(N <= 65535, "gcrf_fwd_bwd: batch %d > 65535", N);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned *ticket = (unsigned *)workspace;
    float *partials = (float *)((char *)workspace + 64);
    const double denom = (double)N * H * W;
    const float gscale = (float)(-2.0 / denom);

    auto coef = [](float sigma) { return sigma > 0.f ? -0.5f * LOG2E / (sigma * sigma) : 0.f; };
    const bool fast_ok = n_desc == 1 && h_descs[0].weight > 0.f && radius == 5 && (C >= 2 && C <= 4) &&
                         (Cin == 1 || Cin == 3);
    if (fast_ok) {
        FastParams prm{coef(h_descs[0].sigma_xy), coef(h_descs[0].sigma_rgb), log2f(h_descs[0].weight)};
#define FEDICRA_GCRF_CASE(c_, ci_)                                                                            \
    if (C == c_ && Cin == ci_) {                                                                               \
        launch_fast<c_, ci_>(y, image, grad_y, partials, ticket, loss, N, H, W, prm, gscale, 1.0 / denom, st); \
        FEDICRA_CHECK_LAUNCH("gcrf_fast_kernel");                                                             \
        return 0;                                                                                              \
    }
        FEDICRA_GCRF_CASE(2, 1)
        FEDICRA_GCRF_CASE(3, 1)
        FEDICRA_GCRF_CASE(4, 1)
        FEDICRA_GCRF_CASE(2, 3)
        FEDICRA_GCRF_CASE(3, 3)
        FEDICRA_GCRF_CASE(4, 3)
#undef FEDICRA_GCRF_CASE
    }
    GenericParams gp;
    gp.n = n_desc;
    for (int i = 0; i < n_desc; ++i) gp.d[i] = GenericDesc{h_descs[i].weight, coef(h_descs[i].sigma_xy), coef(h_descs[i].sigma_rgb)};
    const size_t npix = (size_t)N * H * W;
    const unsigned blocks = (unsigned)((npix + 127) / 128);
    gcrf_generic_kernel<<<blocks, 128, 0, st>>>(y, image, grad_y, partials, ticket, loss, N, C, Cin, H, W, radius, gp,
                                                gscale, 1.0 / denom);
    FEDICRA_CHECK_LAUNCH("gcrf_generic_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of fedicra_gcrf_fwd_bwd.
//
// Who this is for: callers that are not in a position to hand over a scratch buffer because they cannot create one.
// Examples: a host program whose own memory manager has already reserved the GPU's memory and forbids side
// allocations; a service loop with a fixed memory budget where a cudaMalloc (which synchronises the whole device)
// would stall every other stream; a plugin that is handed tensors by its host but has no allocator of its own.
// For them the scratch lives inside the library: one fixed array in device memory, created when the library is loaded
// on a GPU and kept until the process exits.  Every GPU that loads the library pays for it, so it is kept small.
//
// Limit: the array is sized like the workspace fedicra_gcrf_workspace_bytes() asks for -- one float per thread block of
// the kernel that launches the most blocks -- for at most FEDICRA_GCRF_NOALLOC_MAX_CTAS = 32768 blocks, i.e.
//     max(N * ceil(H/16) * ceil(W/32), ceil(N*H*W / 128)) <= 32768.
// For ordinary images that is N*H*W <= 4 194 304 pixels (16 x 512 x 512, 64 x 256 x 256); very thin or very small
// images reach the tile term first (20 000 images of 1 x 33 are 40 000 tiles).  The array is 64 + 4*32768 = 131136
// bytes.  Larger problems are refused with an error; use fedicra_gcrf_fwd_bwd with a caller-owned workspace for those.
//
// The array is shared by all calls on a device and keeps whatever the previous call left in it.  The only part that
// must start from zero is the 4-byte block counter at its front; each call clears it on `stream` before launching.
// Because of the sharing, two calls that may run on the GPU at the same time (different streams, different host
// threads) race on the array and give wrong losses: serialise them.
//
// The first call on each device looks up the array's device address (the name `g_gcrf_noalloc_scratch` below is only a
// host-side handle, not something a kernel can be given).  That lookup is a one-off set-up step: make the first call
// on a device before starting any stream capture on it.  The address is remembered per device ordinal for the life of
// the process: it assumes the primary context and no cudaDeviceReset in between.
namespace {

constexpr int FEDICRA_GCRF_NOALLOC_MAX_CTAS = 32768;
constexpr size_t FEDICRA_GCRF_NOALLOC_BYTES = 64 + (size_t)FEDICRA_GCRF_NOALLOC_MAX_CTAS * sizeof(float);
constexpr int FEDICRA_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_gcrf_noalloc_scratch[FEDICRA_GCRF_NOALLOC_BYTES];

std::mutex g_noalloc_mutex;
void *g_noalloc_ptr[FEDICRA_MAX_DEVICES] = {};  // device address of the array, per device ordinal

// Returns the device address of the scratch array on the CURRENT device (looked up once per device), or nullptr with
// the error text set.
void *gcrf_noalloc_scratch_ptr() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= FEDICRA_MAX_DEVICES) {
        fedicra_fail(-2, "gcrf_fwd_bwd_noalloc: cannot identify the current device (%s)",
                     e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range");
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_gcrf_noalloc_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            fedicra_fail(-2, "gcrf_fwd_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                         cudaGetErrorString(e));
            return nullptr;
        }
        g_noalloc_ptr[dev] = ptr;
    }
    return g_noalloc_ptr[dev];
}

}  // namespace

extern "C" size_t fedicra_gcrf_noalloc_max_bytes(void) { return FEDICRA_GCRF_NOALLOC_BYTES; }

extern "C" int fedicra_gcrf_fwd_bwd_noalloc(const float *y, const float *image, int N, int C, int Cin, int H, int W,
                                            int radius, const fedicra_gcrf_desc *h_descs, int n_desc, float *loss,
                                            float *grad_y, void *stream) {
    FEDICRA_REQUIRE(N > 0 && H > 0 && W > 0, "gcrf_fwd_bwd_noalloc: bad shape N=%d H=%d W=%d", N, H, W);
    const size_t need = fedicra_gcrf_workspace_bytes(N, H, W);
    FEDICRA_REQUIRE(need <= FEDICRA_GCRF_NOALLOC_BYTES,
                    "gcrf_fwd_bwd_noalloc: problem N=%d H=%d W=%d launches up to %zu thread blocks (%zu scratch bytes), the "
                    "built-in array holds %zu bytes = %d thread blocks [max(tiles of 16x32, pixels/128)]; call "
                    "fedicra_gcrf_fwd_bwd with your own workspace instead",
                    N, H, W, (need - 64) / sizeof(float), need, FEDICRA_GCRF_NOALLOC_BYTES, FEDICRA_GCRF_NOALLOC_MAX_CTAS);
    void *scratch = gcrf_noalloc_scratch_ptr();
    if (scratch == nullptr) return -2;
    // the block counter must start at zero whatever an earlier (possibly failed) call left behind
    cudaError_t e = cudaMemsetAsync(scratch, 0, 64, (cudaStream_t)stream);
    if (e != cudaSuccess) return fedicra_fail(-2, "gcrf_fwd_bwd_noalloc: clearing the block counter: %s", cudaGetErrorString(e));
    return fedicra_gcrf_fwd_bwd(y, image, N, C, Cin, H, W, radius, h_descs, n_desc, loss, grad_y, scratch,
                                FEDICRA_GCRF_NOALLOC_BYTES, stream);
}
