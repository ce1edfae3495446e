// hop_probe.cu -- what one hand-over between two CTAs costs on this GPU: latency of the load flavours that can poll a
// word written by another CTA, and the round trip of a store -> poll -> store -> poll ping-pong between two SMs.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o hop_probe hop_probe.cu && ./hop_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

template <int KIND> __device__ __forceinline__ unsigned ld(const unsigned *p) {
    unsigned v;
    if (KIND == 0) asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(p));
    if (KIND == 1) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
    if (KIND == 2) asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    if (KIND == 3) asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    if (KIND == 4) asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    if (KIND == 5) asm volatile("ld.relaxed.cta.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
template <int KIND> __device__ __forceinline__ void st(unsigned *p, unsigned v) {
    if (KIND == 0) asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v));
    if (KIND == 1) asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v));
    if (KIND == 2) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v));
    if (KIND == 3) asm volatile("st.global.cg.u32 [%0], %1;" ::"l"(p), "r"(v));
}

template <int KIND> __global__ void chase(const unsigned *next, int steps, long long *out) {
    unsigned i = 0;
    for (int k = 0; k < 64; ++k) i = ld<KIND>(next + i);
    const long long t0 = clock64();
    for (int k = 0; k < steps; ++k) i = ld<KIND>(next + i);
    const long long t1 = clock64();
    out[0] = t1 - t0;
    out[1] = i;
}

// CTA 0 and CTA 1 (different SMs): ping-pong through two words `steps` times
template <int LD, int ST, int SLEEP> __global__ void pingpong(unsigned *w, int steps, long long *out) {
    if (threadIdx.x != 0) return;
    unsigned *mine = w + (blockIdx.x == 0 ? 0 : 64), *theirs = w + (blockIdx.x == 0 ? 64 : 0);
    const long long t0 = clock64();
    for (int k = 1; k <= steps; ++k) {
        if (blockIdx.x == 0) {
            st<ST>(mine, k);
            while (ld<LD>(theirs) != (unsigned)k) { if (SLEEP) __nanosleep(SLEEP); }
        } else {
            while (ld<LD>(theirs) != (unsigned)k) { if (SLEEP) __nanosleep(SLEEP); }
            st<ST>(mine, k);
        }
    }
    out[blockIdx.x] = clock64() - t0;
}

// the same with `nwarp` extra warps per CTA and `nbg` extra CTAs polling words that never change (background pollers)
template <int LD> __global__ void pingpong_bg(unsigned *w, int steps, long long *out, volatile unsigned *stop, int words) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (blockIdx.x < 2 && warp == 0) {
        if (lane != 0) return;
        unsigned *mine = w + (blockIdx.x == 0 ? 0 : 64), *theirs = w + (blockIdx.x == 0 ? 64 : 0);
        const long long t0 = clock64();
        for (int k = 1; k <= steps; ++k) {
            if (blockIdx.x == 0) {
                st<0>(mine, k);
                while (ld<LD>(theirs) != (unsigned)k) __nanosleep(20);
            } else {
                while (ld<LD>(theirs) != (unsigned)k) __nanosleep(20);
                st<0>(mine, k);
            }
        }
        out[blockIdx.x] = clock64() - t0;
        if (blockIdx.x == 0) st<0>((unsigned *)stop, 1u);
        return;
    }
    // background: every lane polls `words` scattered words until told to stop
    const unsigned *base = w + 4096 + ((blockIdx.x * blockDim.x + threadIdx.x) * 131u % 1000003u) * 8u;
    unsigned acc = 0;
    while (ld<2>((const unsigned *)stop) == 0u) {
        for (int k = 0; k < words; ++k) acc += ld<LD>(base + k * 65536);
        __nanosleep(20);
    }
    if (acc == 0xdeadbeefu) out[7] = acc;
}

int main() {
    int dev = 0, khz = 0;
    cudaSetDevice(dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int N = 1 << 18;  // 1 MiB of indices: L2-resident, far larger than L1's share of one line per step
    std::vector<unsigned> h(N);
    {   // one random cycle, stride over whole cache lines
        std::vector<unsigned> perm(N / 32);
        for (unsigned i = 0; i < perm.size(); ++i) perm[i] = i;
        srand(1);
        for (size_t i = perm.size() - 1; i > 0; --i) std::swap(perm[i], perm[rand() % (i + 1)]);
        for (size_t i = 0; i < perm.size(); ++i) h[perm[i] * 32] = perm[(i + 1) % perm.size()] * 32;
    }
    unsigned *d_next, *d_w;
    long long *d_out, h_out[8];
    cudaMalloc(&d_next, N * 4);
    cudaMemcpy(d_next, h.data(), N * 4, cudaMemcpyHostToDevice);
    cudaMalloc(&d_w, (size_t)64 << 20);
    cudaMalloc(&d_out, 64);
    const char *names[] = {"ld.global.ca", "ld.global.cg", "ld.relaxed.gpu", "ld.volatile", "ld.acquire.gpu", "ld.relaxed.cta"};
    const int steps = 4000;
    auto report = [&](const char *n) {
        cudaDeviceSynchronize();
        cudaMemcpy(h_out, d_out, 64, cudaMemcpyDeviceToHost);
        printf("chase %-16s %7.1f cycles/step  %6.1f ns\n", n, (double)h_out[0] / steps, (double)h_out[0] / steps * 1e6 / khz);
    };
    chase<0><<<1, 1>>>(d_next, steps, d_out); report(names[0]);
    chase<1><<<1, 1>>>(d_next, steps, d_out); report(names[1]);
    chase<2><<<1, 1>>>(d_next, steps, d_out); report(names[2]);
    chase<3><<<1, 1>>>(d_next, steps, d_out); report(names[3]);
    chase<4><<<1, 1>>>(d_next, steps, d_out); report(names[4]);
    chase<5><<<1, 1>>>(d_next, steps, d_out); report(names[5]);
    auto pp = [&](const char *n) {
        cudaDeviceSynchronize();
        cudaError_t e = cudaGetLastError();
        cudaMemcpy(h_out, d_out, 64, cudaMemcpyDeviceToHost);
        printf("pingpong %-44s %7.1f cycles/round trip  %6.1f ns (one hop = half)  %s\n", n, (double)h_out[0] / steps,
               (double)h_out[0] / steps * 1e6 / khz, e == cudaSuccess ? "" : cudaGetErrorString(e));
    };
    cudaMemset(d_w, 0, 4096); pingpong<2, 0, 0><<<2, 32>>>(d_w, steps, d_out); pp("ld.relaxed.gpu / st.relaxed.gpu, no sleep");
    cudaMemset(d_w, 0, 4096); pingpong<2, 0, 20><<<2, 32>>>(d_w, steps, d_out); pp("ld.relaxed.gpu / st.relaxed.gpu, nanosleep(20)");
    cudaMemset(d_w, 0, 4096); pingpong<2, 0, 100><<<2, 32>>>(d_w, steps, d_out); pp("ld.relaxed.gpu / st.relaxed.gpu, nanosleep(100)");
    cudaMemset(d_w, 0, 4096); pingpong<3, 1, 0><<<2, 32>>>(d_w, steps, d_out); pp("ld.volatile / st.volatile, no sleep");
    cudaMemset(d_w, 0, 4096); pingpong<1, 3, 0><<<2, 32>>>(d_w, steps, d_out); pp("ld.global.cg / st.global.cg, no sleep");
    cudaMemset(d_w, 0, 4096); pingpong<4, 2, 0><<<2, 32>>>(d_w, steps, d_out); pp("ld.acquire.gpu / st.release.gpu, no sleep");
    unsigned *d_stop = d_w + 1024;
    for (int words : {0, 1, 3, 12}) {
        for (int ctas : {2, 148 * 2, 148 * 5}) {
            cudaMemset(d_w, 0, 8192);
            pingpong_bg<2><<<ctas, 128>>>(d_w, steps, d_out, d_stop, words);
            char buf[96];
            snprintf(buf, sizeof buf, "relaxed + %d CTAs x 4 warps polling %d words/lane", ctas, words);
            pp(buf);
        }
    }
    return 0;
}
