// stitch2.cuh -- band-to-band recurrences (U2, D2) of the multi-launch filter, one CTA per (channel, image).
// Included by bands.cu inside its anonymous namespace.
//
// Same recurrences as stitch_up_kernel / stitch_down_kernel (and as the resident kernel's stitcher):
//   U2 (deepest band first):    X[j'] = base[j'] + sum_{j in E(beta), entry(j) = j'} A[j] * X[j]     for j' in E(beta-1)
//   D2 (shallowest band first): out[j] = B[j] + A[j] * out[entry(j)]                                 for j in E(beta)
// with E(beta) the last level of band beta.  The older kernels stream the levels through a cp.async ring, one warp per
// (image, channel), and pay ~1.7 us per band at 512x512 (128 bands).  Here the last levels of ALL bands of one channel
// are loaded into shared memory first (12 bytes per entry: A, entry index, run index, value; ~13 000 entries at 512^2),
// after which a step is a segmented warp scan (U2) or a gather (D2) on shared memory, as in resident.cuh.
// Entries that do not fit (pathological level widths) take the same recurrence on global memory.

constexpr int ST2_NT = 1024;

__device__ __forceinline__ int st2_band_of(const int *s_off, int nb, int x) {   // largest beta with off[beta] <= x
    int lo = 0, hi = nb - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (s_off[mid] <= x) lo = mid; else hi = mid - 1;
    }
    return lo;
}

template <bool UP>
__global__ void __launch_bounds__(ST2_NT)
stitch2_kernel(TreeView t, Consts k, float *__restrict__ X, const float *__restrict__ Bv, float *__restrict__ outE, int NC,
               int tcap) {
    extern __shared__ float4 smem4[];
    char *smem = reinterpret_cast<char *>(smem4);
    int *s_e0 = reinterpret_cast<int *>(smem), *s_n = s_e0 + MAXBANDS + 1, *s_off = s_n + MAXBANDS + 1;
    smem += 3 * (MAXBANDS + 2) * sizeof(int) + 8;
    const int c = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, V = t.V;
    const int *ls = t.ls + (size_t)b * (V + 1);
    const int *plan = t.bands + (size_t)b * PLAN_STRIDE;
    const float *gA = k.A + (size_t)b * V;
    const int *gent = k.ent + (size_t)b * V;
    const size_t row = ((size_t)b * NC + c) * V;
    const int nb = plan[0];
    for (int q = tid; q < nb; q += ST2_NT) {
        const int L1 = plan[2 + q];
        const int e0 = ls[L1 - 1];
        s_e0[q] = e0;
        s_n[q] = ls[L1] - e0;
    }
    __syncthreads();
    if (tid < 32) {   // exclusive scan of the widths by one warp
        int carry = 0;
        for (int q0 = 0; q0 < nb; q0 += 32) {
            const int q = q0 + lane, n = (q < nb) ? s_n[q] : 0;
            int incl = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += up;
            }
            if (q < nb) s_off[q] = carry + incl - n;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) s_off[nb] = carry;
    }
    __syncthreads();
    const int T = s_off[nb];
    float *sA = reinterpret_cast<float *>(smem);                                     // [tcap]
    float *sX = sA + tcap;                                                           // [tcap]
    unsigned short *sEnt = reinterpret_cast<unsigned short *>(sX + tcap);            // [tcap] entry, local to E(beta-1)
    unsigned short *sGd = sEnt + tcap;   // [tcap] bits 0-14: index inside the run of equal entries, bit 15: last of the run
    constexpr int U = 4;
    if (T <= tcap) {
        // ---- A, entry (and for U2 the run structure) of every last-level entry, U loads in flight per thread
        for (int base = tid; base < T; base += U * ST2_NT) {
            int beta[U], en[U], next[U];
            float av[U];
#pragma unroll
            for (int u = 0; u < U; ++u) beta[u] = (base + u * ST2_NT < T) ? st2_band_of(s_off, nb, base + u * ST2_NT) : -1;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                en[u] = next[u] = -1;
                av[u] = 0.f;
                if (beta[u] >= 0) {
                    const int jl = base + u * ST2_NT - s_off[beta[u]], e0 = s_e0[beta[u]], n = s_n[beta[u]];
                    av[u] = gA[e0 + jl];
                    if (beta[u] > 0) {
                        en[u] = gent[e0 + jl];
                        next[u] = (UP && jl < n - 1) ? gent[e0 + jl + 1] : -1;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (beta[u] < 0) continue;
                const int x = base + u * ST2_NT;
                sA[x] = av[u];
                sEnt[x] = (beta[u] > 0) ? (unsigned short)(en[u] - s_e0[beta[u] - 1]) : (unsigned short)0;
                sGd[x] = (next[u] != en[u]) ? (unsigned short)0x8000 : (unsigned short)0;
            }
        }
        // ---- values: U2 starts from the bases written by the first sweep, D2 from the local down terms
        const float *src = UP ? X + row : Bv + row;
        const int Tv = UP ? s_off[nb - 1] : T;   // (the deepest band's last level receives nothing from below)
        for (int base = tid; base < Tv; base += U * ST2_NT) {
            float v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int x = base + u * ST2_NT;
                int pos = 0;
                if (x < Tv) {
                    const int q = st2_band_of(s_off, nb, x);
                    pos = s_e0[q] + (x - s_off[q]);
                }
                v[u] = (x < Tv) ? src[pos] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (base + u * ST2_NT < Tv) sX[base + u * ST2_NT] = v[u];
        }
        __syncthreads();
        if (UP) {   // index inside the run: look back while the entry does not change
            for (int x = tid; x < T; x += ST2_NT) {
                const int first = s_off[st2_band_of(s_off, nb, x)], e = sEnt[x];
                int d = 0;
                while (x - d - 1 >= first && sEnt[x - d - 1] == e && d < 0x7fff) ++d;
                sGd[x] = (unsigned short)(sGd[x] | d);
            }
            __syncthreads();
        }
        if (warp == 0) {
            if (UP) {
                // step beta consumes X of E(beta) and completes X of E(beta-1): one lane per entry of E(beta), products
                // summed per run by a segmented warp scan (two 32-entry chunks at a time), the run's last lane adds the
                // total to its base
                for (int beta = nb - 2; beta >= 1; --beta) {
                    const int off = s_off[beta], qoff = s_off[beta - 1], ne = s_n[beta];
                    float carry = 0.f;
                    for (int jb = 0; jb < ne; jb += 64) {
                        const int j1 = jb + lane, j2 = j1 + 32;
                        const bool ok1 = j1 < ne, ok2 = j2 < ne;
                        const unsigned g1 = ok1 ? sGd[off + j1] : 0u, g2 = ok2 ? sGd[off + j2] : 0u;
                        const int e1 = ok1 ? sEnt[off + j1] : 0, e2 = ok2 ? sEnt[off + j2] : 0;
                        float v1 = ok1 ? sA[off + j1] * sX[off + j1] : 0.f, v2 = ok2 ? sA[off + j2] * sX[off + j2] : 0.f;
                        const float b1 = sX[qoff + e1], b2 = sX[qoff + e2];
                        const int r1 = (int)(g1 & 0x7fffu), r2 = (int)(g2 & 0x7fffu);
                        const int d1 = min(r1, lane), d2 = min(r2, lane);
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const float t1 = __shfl_up_sync(0xffffffffu, v1, o), t2 = __shfl_up_sync(0xffffffffu, v2, o);
                            if (d1 >= o) v1 += t1;
                            if (d2 >= o) v2 += t2;
                        }
                        if (r1 > lane) v1 += carry;
                        const float c1 = __shfl_sync(0xffffffffu, v1, 31);
                        if (r2 > lane) v2 += c1;
                        carry = __shfl_sync(0xffffffffu, v2, 31);
                        if (ok1 && (g1 & 0x8000u)) sX[qoff + e1] = b1 + v1;
                        if (ok2 && (g2 & 0x8000u)) sX[qoff + e2] = b2 + v2;
                    }
                    __syncwarp();
                }
            } else {
                for (int beta = 1; beta < nb; ++beta) {
                    const int off = s_off[beta], qoff = s_off[beta - 1], n = s_n[beta];
                    for (int jl = lane; jl < n; jl += 64) {   // two entries per lane, their loads issued together
                        const int jl2 = jl + 32;
                        const bool two = jl2 < n;
                        const int e1 = sEnt[off + jl], e2 = two ? sEnt[off + jl2] : 0;
                        const float a1 = sA[off + jl], a2 = two ? sA[off + jl2] : 0.f;
                        const float b1 = sX[off + jl], b2 = two ? sX[off + jl2] : 0.f;
                        const float o1 = fmaf(a1, sX[qoff + e1], b1), o2 = fmaf(a2, sX[qoff + e2], b2);
                        sX[off + jl] = o1;
                        if (two) sX[off + jl2] = o2;
                    }
                    __syncwarp();
                }
            }
        }
        __syncthreads();
        float *dst = UP ? X + row : outE + row;
        for (int x = tid; x < Tv; x += ST2_NT) {
            const int q = st2_band_of(s_off, nb, x);
            dst[s_e0[q] + (x - s_off[q])] = sX[x];
        }
        return;
    }
    // ---- oversized: the same recurrences on global memory, one warp
    if (warp != 0) return;
    if (UP) {
        float *Xr = X + row;
        for (int beta = nb - 2; beta >= 1; --beta) {
            const int e0 = s_e0[beta], n = s_n[beta], q0 = s_e0[beta - 1], nq = s_n[beta - 1];
            for (int jl = lane; jl < nq; jl += 32) {
                int lo = 0, hi = n;   // the run under q0 + jl, by bisection on the (monotone) entries
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (gent[e0 + mid] < q0 + jl) lo = mid + 1; else hi = mid; }
                float acc = Xr[q0 + jl];
                for (int jj = lo; jj < n && gent[e0 + jj] == q0 + jl; ++jj) acc = fmaf(gA[e0 + jj], Xr[e0 + jj], acc);
                Xr[q0 + jl] = acc;
            }
            __threadfence_block();
            __syncwarp();
        }
    } else {
        const float *Br = Bv + row;
        float *O = outE + row;
        for (int beta = 0; beta < nb; ++beta) {
            const int e0 = s_e0[beta], n = s_n[beta];
            for (int jl = lane; jl < n; jl += 32) {
                float o = Br[e0 + jl];
                if (beta > 0) o = fmaf(gA[e0 + jl], O[gent[e0 + jl]], o);
                O[e0 + jl] = o;
            }
            __threadfence_block();
            __syncwarp();
        }
    }
}
