// Probe: 3-D TMA tile load with a halo (negative start coordinates, zero fill) into shared memory.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int PITCH, int HH, int CH>
__global__ void probe(const __grid_constant__ CUtensorMap map, float *out, int x0, int y0, int p0, int variant) {
    extern __shared__ __align__(128) float4 smem4[];
    float *s = reinterpret_cast<float *>(smem4);
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned bytes = PITCH * HH * CH * 4;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(smem_addr(s)), "l"(&map), "r"(x0), "r"(y0), "r"(p0), "r"(smem_addr(&bar)) : "memory");
    }
    unsigned done = 0, spins = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_addr(&bar)) : "memory");
        if (++spins > (1u << 22)) { if (threadIdx.x == 0) out[0] = -12345.f; return; }
    }
    for (int i = threadIdx.x; i < PITCH * HH * CH; i += blockDim.x) out[1 + i] = s[i];
    if (threadIdx.x == 0) out[0] = (float)spins;
}
int main() {
    const int N = 2, CH = 3, H = 48, W = 64;
    constexpr int PITCH = 48, HH = 26;
    float *d, *out;
    cudaMalloc(&d, N * CH * H * W * 4);
    cudaMalloc(&out, (1 + PITCH * HH * CH) * 4);
    float *h = (float *)malloc(N * CH * H * W * 4);
    for (int i = 0; i < N * CH * H * W; ++i) h[i] = (float)(i % 1000) + 1.f;
    cudaMemcpy(d, h, N * CH * H * W * 4, cudaMemcpyHostToDevice);
    void *p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry: %d %d %p\n", (int)e, (int)q, p);
    EncodeTiledFn enc = (EncodeTiledFn)p;
    CUtensorMap m;
    cuuint64_t dims[3] = {W, H, N * CH};
    cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
    cuuint32_t box[3] = {PITCH, HH, CH};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    const int starts[3][3] = {{0, 0, 0}, {-8, -5, 3}, {24, 27, 3}};
    for (int t = 0; t < 3; ++t) {
        probe<PITCH, HH, CH><<<1, 128, PITCH * HH * CH * 4>>>(m, out, starts[t][0], starts[t][1], starts[t][2], 0);
        e = cudaDeviceSynchronize();
        printf("start (%d,%d,%d): sync %s\n", starts[t][0], starts[t][1], starts[t][2], cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        float *ho = (float *)malloc((1 + PITCH * HH * CH) * 4);
        cudaMemcpy(ho, out, (1 + PITCH * HH * CH) * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int c = 0; c < CH; ++c) for (int r2 = 0; r2 < HH; ++r2) for (int x = 0; x < PITCH; ++x) {
            int gx = starts[t][0] + x, gy = starts[t][1] + r2, gp = starts[t][2] + c;
            float want = (gx >= 0 && gx < W && gy >= 0 && gy < H && gp < N * CH) ? h[((size_t)gp * H + gy) * W + gx] : 0.f;
            if (ho[1 + (c * HH + r2) * PITCH + x] != want) ++bad;
        }
        printf("  spins %.0f mismatches %d\n", ho[0], bad);
    }
    return 0;
}
