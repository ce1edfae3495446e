// resident.cuh -- band-resident tree filter: the phases of bands.cu (U1, U2, U3, D1, D2, D3 + epilogue) in ONE launch.
// Included by bands.cu inside its anonymous namespace (it shares TreeView / Consts / the band plan / prologue()).
//
// What it replaces: the same reference code as bands.cu (LTF/src/refine/refine.cu:19-370), same recurrences, same
// arithmetic.  What changes is where the band state lives between the phases.  In the multi-launch version every phase
// is a kernel: a band is staged into shared memory three times and its partial results travel through global memory.
// Here ONE CTA per (image, band) stays resident for the whole filter and keeps its band (values, weights, packed
// topology) in shared memory from the first sweep to the epilogue; one extra CTA per image (the "stitcher") runs the two
// sequential band-to-band recurrences out of its own shared memory.  Bands and stitcher meet at four per-image
// barriers made of a counter and a flag in global memory:
//
//     bands:     stage, U1 ──cnt_up──▶         ◀──flag_up── U3, D1 ──cnt_down──▶          ◀──flag_down── D3, epilogue
//     stitcher:  preload A/ent ········ U2 (deepest band first) ···················· D2 (shallowest band first)
//
// Waiting on a flag needs every CTA of an image to be on the GPU at the same time: the kernel is launched
// cooperatively with (NBCAP + 1) x Bc CTAs, Bc = as many images as the occupancy calculator says fit at once (the batch
// is cut into such chunks on the host); the planner guarantees nb <= NBCAP bands per image.  A band that does not fit
// the shared-memory budget (possible only for pathological level widths) runs the same code on a global-memory slab.
//
// The sync words start at zero (caller-zeroed workspace, like the Gated CRF scratch) and every launch leaves them at zero.

constexpr int RSLOTS = 64;       // threads per swept channel
#ifndef RES_MINB
#define RES_MINB 3                // resident CTAs per SM the kernel is compiled for (channel counts up to 4)
#endif
constexpr int SYNC_INTS = 8;     // per image: cnt_up, flag_up, cnt_down, flag_down, cnt_fin, (3 spare)

struct ResidentArgs {
    const float *feat;     // FWD: feature_in [B,C,V] vertex order.  BWD: grad_out [B,C,V] vertex order
    float *fpos;           // [B,C+1,V] position order: forward aggregates (FWD writes, BWD reads)
    float *fup;            // [B,C+1,V] position order: forward exact up (FWD writes, BWD reads)
    int C, NC;             // feature channels; swept channels (FWD: C+1, BWD: C or 2C)
    float *X;              // [B,NC,V] values entering the last level of a band from below (only last-level entries used)
    float *up;             // [B,NC,V] BWD: exact up of the swept channels (for the edge-weight gradient)
    float *BvE;            // [B,NC,V] local down term at the last level of every band (stitcher input)
    float *outE;           // [B,NC,V] exact down value at the last level of every band (stitcher output)
    float *slab;           // [B,NC,V] band state of oversized bands
    int *sync;             // [B,SYNC_INTS]
    // epilogue outputs
    float *out0;           // FWD: feature_out [B,C,V] vertex order.   BWD: grad_feature [B,C,V] vertex order
    float *out1;           // FWD: feature_aggr [B,C,V] or NULL.       BWD: grad_weight [B,V] position order or NULL
    float *out2;           // FWD: weight_sum [B,V] or NULL
    int b0;                // first image of this chunk
    int pcap;              // positions per band that fit the shared-memory band state
    int tcap;              // last-level entries (all bands of an image together) that fit the stitcher's shared memory
    long long *prof;       // optional phase timestamps of image b0 (FEDICRA_PROF=1), [gridDim.x][PROF_SLOTS] or NULL
};
constexpr int PROF_SLOTS = 12;
constexpr int NBMAX = 96;        // largest band count a plan may have (stitcher tables)

__device__ __forceinline__ void prof_mark(const ResidentArgs &a, int slot) {
    if (a.prof != nullptr && blockIdx.y == 0 && threadIdx.x == 0) {
        long long tns;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tns));
        a.prof[blockIdx.x * PROF_SLOTS + slot] = tns;
    }
}

__device__ __forceinline__ int ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int *p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_add(int *p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// All threads of the CTA have written their part; publish it and bump the counter.
__device__ __forceinline__ void cta_arrive(int *counter) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        red_release_add(counter, 1);
    }
}
// Longest time a CTA may wait for a counter before the launch is given up (FEDICRA_WATCHDOG_MS, default 10 s, 0 = wait
// for ever).  Wall-clock time also passes while the context is stopped under a debugger, slowed by compute-sanitizer or
// time-sliced with other processes, and a non-cooperative launch (FEDICRA_COOP=0) may legitimately wait for CTAs that
// are not resident yet -- raise or disable the bound there.
__device__ unsigned long long g_resident_watchdog_ns = 10000000000ull;

// Block the CTA until *word == want.  With sound sync words a wait ends within microseconds; one that outlasts the bound
// above means they are corrupted (a workspace that was not zeroed, or shared by overlapping calls): the kernel traps, so
// the launch fails with a (sticky) CUDA error instead of hanging the GPU.
__device__ __forceinline__ void cta_wait(const int *word, int want) {
    if (threadIdx.x == 0) {
        unsigned long long t0 = 0;
        unsigned spins = 0;
        while (ld_acquire(word) != want) {
            __nanosleep(20);
            if ((++spins & 0xfffu) == 0) {   // look at the clock every 4096 polls
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                const unsigned long long limit = g_resident_watchdog_ns;
                if (t0 == 0) t0 = now;
                else if (limit != 0 && now - t0 > limit) __trap();
            }
        }
        __threadfence();
    }
    __syncthreads();
}

// Inputs of the sweeps for one position i (vertex v), all swept channels at once:
//   FWD: x_c = feature_in[c][v] (c < C), x_C = 1                                   (refine.cu:201-249)
//   BWD: x_c = grad_out[c][v] / Z[i],  x_{C+c} = x_c * out[c][i] = x_c * aggr[c][i] / Z[i]      (refine.cu:347-348)
// One reciprocal per position serves every channel (the reference divides channel by channel; fp32 round-off only).
template <int MODE, int NCT>
__device__ __forceinline__ void resident_inputs(const ResidentArgs &a, int b, int i, int v, int V, float *x) {
    const int C = a.C, NC = NCT ? NCT : a.NC;
    if (MODE == MODE_FWD) {
#pragma unroll
        for (int c = 0; c < (NCT ? NCT : 8); ++c)
            if (c < NC) x[c] = (c < C) ? a.feat[((size_t)b * C + c) * V + v] : 1.f;
    } else {
        const size_t fimg = (size_t)b * (C + 1) * V;
        const float rz = 1.f / a.fpos[fimg + (size_t)C * V + i];
#pragma unroll
        for (int c = 0; c < (NCT ? NCT : 8); ++c)
            if (c < C) {
                const float gn = a.feat[((size_t)b * C + c) * V + v] * rz;
                x[c] = gn;
                if (C + c < NC) x[C + c] = gn * (a.fpos[fimg + (size_t)c * V + i] * rz);
            }
    }
}
// The sweeps of different channels never read each other's rows: inside a sweep the two warps of a channel meet at their
// own named barrier (ids 1..NC) instead of stopping the whole CTA at every level.
__device__ __forceinline__ void channel_barrier(int c) {
    asm volatile("bar.sync %0, %1;" ::"r"(c + 1), "n"(RSLOTS) : "memory");
}
__device__ __forceinline__ int sls_or(bool cached, const int *sls, const int *ls, int L, int L0) {
    return cached ? sls[L - L0] : ls[L];
}

// ------------------------------------------------------------------------------------------------ one band, resident
// pk = parent (local) << 16 | first child (local) << 3 | #children; local = position - p0 < 8192 (written by
// consts_kernel).  NCT = swept channels at compile time, so that the memory phases unroll into batches of independent
// loads: a CTA has few threads, its memory phases are one or two round trips deep, and every load that is not in
// flight during a round trip is a microsecond added to the filter.
template <int MODE, bool SMEM, int NCT>
__device__ void resident_band(const TreeView &t, const Consts &k, const ResidentArgs &a, int b, int beta, int nb, int L0,
                              int L1, int p0, int p1, char *smem) {
    const int tid = threadIdx.x, nthr = blockDim.x, V = t.V, NC = NCT, C = a.C, P = p1 - p0, pcap = a.pcap;
    const int *ls = t.ls + (size_t)b * (V + 1);
    const int *vtx = t.vtx + (size_t)b * V + p0, *pp = t.ppos + (size_t)b * V + p0;
    const int4 *child = t.child + (size_t)b * V;
    const float *w = t.w + (size_t)b * V;
    const size_t img = (size_t)b * NC * V;
    // rows of ps = pcap + 4 entries: children are contiguous, so a sweep loads the four slots from its first child on with
    // ONE address and constant offsets and masks the missing ones by their weight; the three slots past the end (zeroed)
    // keep those loads inside the row for the last positions of a band
    const int ps = pcap + 4;
    float *sval = reinterpret_cast<float *>(smem);       // [NC][ps]
    float *sw = sval + (size_t)NC * ps;                  // [ps]
    int *spk = reinterpret_cast<int *>(sw + ps);         // [ps]; after D1: the stitched values of the level above
    int *sls = spk + ps;                                 // [LSCAP+1] level starts (a shared-memory band has <= LSCAP levels)
    float *gval = a.slab + img + p0;                     // oversized band: its state; otherwise: cache of the inputs
    int *sync = a.sync + (size_t)b * SYNC_INTS;
    const bool ls_cached = (L1 - L0) <= LSCAP;
    auto LSl = [&](int L) -> int { return (ls_cached ? sls[L - L0] : ls[L]) - p0; };   // level start, band-local
    auto VAL = [&](int c, int il) -> float & { return SMEM ? sval[c * ps + il] : gval[(size_t)c * V + il]; };
    auto Wl = [&](int il) -> float { return SMEM ? sw[il] : ((p0 + il == 0) ? 0.f : w[p0 + il]); };
    auto kids = [&](int il, int &cl, int &cnt) {
        if (SMEM) {
            const int pk = spk[il];
            cl = (pk >> 3) & 0x1fff;
            cnt = pk & 7;
        } else {
            const int4 ch = child[p0 + il];
            cnt = child_count(ch);
            cl = ch.x - p0;
        }
    };
    auto parent_l = [&](int il) -> int { return SMEM ? (spk[il] >> 16) : (pp[il] - p0); };
    const int c = tid / RSLOTS, slot = tid % RSLOTS;   // blockDim.x = 64 * NC
    constexpr int U = 4;                               // elements per thread whose loads are issued together
    constexpr int UG = (MODE == MODE_FWD && NCT <= 3) ? 8 : 4;   // ... in the first gather (registers permitting)

    // ---- stage the band: level starts, weights, packed topology (two coalesced streams)
    if (ls_cached)
        for (int L = L0 + tid; L <= L1; L += nthr) sls[L - L0] = ls[L];
    if (SMEM) {
        const int *gpk = k.pk + (size_t)b * V + p0;
        const float *gw = w + p0;
#pragma unroll 8
        for (int il = tid; il < P; il += nthr) {
            spk[il] = gpk[il];
            sw[il] = (p0 + il == 0) ? 0.f : gw[il];
        }
        // the four slots after the band's last position (never written otherwise; 0 * garbage could be NaN)
        if (tid < 4 * NC) sval[(tid >> 2) * ps + P + (tid & 3)] = 0.f;
        if (tid < 4) sw[P + tid] = 0.f;
    }
    // inputs: computed once (vertex-order gathers, one reciprocal per position); a shared-memory band also leaves them
    // in the slab, position order, from where the second sweep re-reads them with coalesced loads
    auto gather = [&]() {
        for (int base = tid; base < P; base += UG * nthr) {
            int v[UG];
            float x[UG][NCT];
#pragma unroll
            for (int u = 0; u < UG; ++u) v[u] = (base + u * nthr < P) ? vtx[base + u * nthr] : -1;
#pragma unroll
            for (int u = 0; u < UG; ++u)
                if (v[u] >= 0) resident_inputs<MODE, NCT>(a, b, p0 + base + u * nthr, v[u], V, x[u]);
#pragma unroll
            for (int u = 0; u < UG; ++u)
                if (v[u] >= 0) {
#pragma unroll
                    for (int cc = 0; cc < NCT; ++cc) {
                        VAL(cc, base + u * nthr) = x[u][cc];
                        if (SMEM && !(MODE == MODE_FWD && cc == NCT - 1)) gval[(size_t)cc * V + base + u * nthr] = x[u][cc];
                    }
                }
        }
    };
    auto regather = [&]() {
        if (!SMEM) { gather(); return; }
        for (int base = tid; base < P; base += 2 * U * nthr) {
            float x[2 * U][NCT];
#pragma unroll
            for (int u = 0; u < 2 * U; ++u)
#pragma unroll
                for (int cc = 0; cc < NCT; ++cc)
                    x[u][cc] = (MODE == MODE_FWD && cc == NCT - 1) ? 1.f
                               : ((base + u * nthr < P) ? gval[(size_t)cc * V + base + u * nthr] : 0.f);
#pragma unroll
            for (int u = 0; u < 2 * U; ++u)
                if (base + u * nthr < P) {
#pragma unroll
                    for (int cc = 0; cc < NCT; ++cc) sval[cc * ps + base + u * nthr] = x[u][cc];
                }
        }
    };
    // leaf -> root sweep over the band's levels; `inject`: the last level first receives X from the band below
    auto up_sweep = [&](bool inject) {
        int lo = LSl(L1 - 1), hi = P;
        if (inject)
            for (int il = lo + slot; il < hi; il += RSLOTS) VAL(c, il) += __ldcg(a.X + img + (size_t)c * V + p0 + il);
        if (SMEM) {
            // Software-pipelined and branch-free: a level step is a chain of dependent shared-memory loads (level start
            // -> packed topology -> child weights -> child values) of which only the last link depends on the sweep.
            // The others are fetched one level (level starts: two levels) ahead, the weights of missing children as 0.
            // Between two barriers a thread then computes one address, waits for one batch of five value loads, does
            // four FMAs and a store.  (The loops are issue-bound: three CTAs per SM run them at the same time, so
            // every instruction taken out of the level step shortens it.)
            float *mv = sval + c * ps;         // this thread's channel row
            auto S = [&](int q) -> int { return sls[max(q, 0)] - p0; };   // start of level L0 + q, band-local (S(0) = 0)
            int q = L1 - 2 - L0;               // level being processed = L0 + q = [lo, hi)
            hi = lo;
            lo = S(q);
            int nlo = S(q - 1);
            int il = lo + slot;
            bool act = q >= 0 && il < hi;
            int pk = act ? spk[il] : 0;
            int cl = (pk >> 3) & 0x1fff, cnt = pk & 7;
            const float *pw = sw + cl;
            float w0 = pw[0], w1 = pw[1], w2 = pw[2], w3 = pw[3];
            for (; q >= 0; --q) {
                const int nnlo = S(q - 2);
                const int nil = nlo + slot;
                const bool nact = q >= 1 && nil < lo;
                const int npk = nact ? spk[nil] : 0;    // next level [nlo, lo): its topology does not depend on the values
                channel_barrier(c);                    // the children of this level are final (this channel's 64 threads)
                if (act) {
                    const float *pv = mv + cl;
                    const float v0 = pv[0], v1 = pv[1], v2 = pv[2], v3 = pv[3];
                    float acc = mv[il];
                    acc = fmaf(v0, cnt > 0 ? w0 : 0.f, acc);
                    acc = fmaf(v1, cnt > 1 ? w1 : 0.f, acc);
                    acc = fmaf(v2, cnt > 2 ? w2 : 0.f, acc);
                    acc = fmaf(v3, cnt > 3 ? w3 : 0.f, acc);
                    mv[il] = acc;
                    for (int j = il + RSLOTS; j < hi; j += RSLOTS) {   // levels wider than the slots
                        const int pj = spk[j], cj = (pj >> 3) & 0x1fff, nj = pj & 7;
                        const float *qv = mv + cj, *qw = sw + cj;
                        float aj = mv[j];
                        aj = fmaf(qv[0], nj > 0 ? qw[0] : 0.f, aj);
                        aj = fmaf(qv[1], nj > 1 ? qw[1] : 0.f, aj);
                        aj = fmaf(qv[2], nj > 2 ? qw[2] : 0.f, aj);
                        aj = fmaf(qv[3], nj > 3 ? qw[3] : 0.f, aj);
                        mv[j] = aj;
                    }
                }
                cl = (npk >> 3) & 0x1fff;
                cnt = npk & 7;
                pw = sw + cl;
                w0 = pw[0]; w1 = pw[1]; w2 = pw[2]; w3 = pw[3];
                hi = lo;
                lo = nlo;
                nlo = nnlo;
                il = nil;
                act = nact;
            }
            __syncthreads();
            return;
        }
        for (int L = L1 - 2; L >= L0; --L) {
            hi = lo;
            lo = LSl(L);
            __syncthreads();
            for (int il = lo + slot; il < hi; il += RSLOTS) {
                int cl, cnt;
                kids(il, cl, cnt);
                float acc = VAL(c, il);
                for (int q = 0; q < cnt; ++q) acc = fmaf(VAL(c, cl + q), Wl(cl + q), acc);
                VAL(c, il) = acc;
            }
        }
        __syncthreads();
    };

    prof_mark(a, 0);
    gather();
    __syncthreads();
    prof_mark(a, 1);

    // ---- U1: local leaf -> root sweep with nothing entering from below; what the first level sends up, summed per
    //      parent, is the base of X for the last level of the band above
    up_sweep(false);
    if (beta > 0) {
        const int q0 = ls[L0 - 1], nq = p0 - q0;
        for (int s = slot; s < nq; s += RSLOTS) {
            const int4 ch = child[q0 + s];
            const int cnt = child_count(ch), cl = ch.x - p0;
            float acc = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < cnt) acc = fmaf(VAL(c, cl + q), Wl(cl + q), acc);
            a.X[img + (size_t)c * V + q0 + s] = acc;
        }
    }
    cta_arrive(sync + 0);
    prof_mark(a, 2);
    regather();                // the band's inputs again, while the stitcher works
    prof_mark(a, 3);
    cta_wait(sync + 1, 1);
    prof_mark(a, 4);

    // ---- U3: the same sweep with X injected -> exact up; it leaves for global memory (saved / edge-weight gradient)
    up_sweep(beta < nb - 1);
    float *upout = (MODE == MODE_FWD) ? a.fup : a.up;
    if (MODE == MODE_FWD || a.out1 != nullptr)
        for (int c2 = 0; c2 < NC; ++c2)
            for (int il = tid; il < P; il += nthr) upout[img + (size_t)c2 * V + p0 + il] = VAL(c2, il);
    __syncthreads();
    prof_mark(a, 5);

    // ---- D1: local root -> leaf pass, in place: val <- B
    {
        int lo = 0, hi = LSl(L0 + 1);
        for (int il = lo + slot; il < hi; il += RSLOTS) {   // first level: nothing from above
            const float wi = Wl(il);
            VAL(c, il) *= (1.f - wi * wi);
        }
        if (SMEM) {
            // pipelined like the up sweep: weight and parent of the next level are fetched a level ahead
            float *mv = sval + c * ps;
            const int nlev = L1 - L0;
            auto Eq = [&](int q) -> int { return sls[min(q + 1, nlev)] - p0; };   // end of level L0 + q, band-local
            int q = 1;                                 // level being processed = L0 + q = [lo, hi)
            lo = hi;
            hi = Eq(q);
            int nhi = Eq(q + 1);
            int il = lo + slot;
            bool act = q < nlev && il < hi;
            float wi = act ? sw[il] : 0.f;
            int par = act ? (spk[il] >> 16) : 0;
            for (; q < nlev; ++q) {
                const int nnhi = Eq(q + 2);
                const int nil = hi + slot;
                const bool nact = q + 1 < nlev && nil < nhi;
                const float nwi = nact ? sw[nil] : 0.f;
                const int npar = nact ? (spk[nil] >> 16) : 0;
                channel_barrier(c);                    // the level above is final (this channel's 64 threads)
                if (act) {
                    mv[il] = fmaf(wi, mv[par], mv[il] * (1.f - wi * wi));
                    for (int j = il + RSLOTS; j < hi; j += RSLOTS) {
                        const float wj = sw[j];
                        mv[j] = fmaf(wj, mv[spk[j] >> 16], mv[j] * (1.f - wj * wj));
                    }
                }
                lo = hi;
                hi = nhi;
                nhi = nnhi;
                il = nil;
                act = nact;
                wi = nwi;
                par = npar;
            }
        } else {
            for (int L = L0 + 1; L < L1; ++L) {
                lo = hi;
                hi = LSl(L + 1);
                __syncthreads();
                for (int il = lo + slot; il < hi; il += RSLOTS) {
                    const float wi = Wl(il);
                    VAL(c, il) = fmaf(wi, VAL(c, parent_l(il)), VAL(c, il) * (1.f - wi * wi));
                }
            }
        }
        __syncthreads();
    }
    {
        const int e0l = LSl(L1 - 1);
        for (int cc = 0; cc < NC; ++cc)
            for (int il = e0l + tid; il < P; il += nthr) a.BvE[img + (size_t)cc * V + p0 + il] = VAL(cc, il);
    }
    cta_arrive(sync + 2);
    prof_mark(a, 6);
    cta_wait(sync + 3, 1);
    prof_mark(a, 7);

    // ---- D3: out[i] = B[i] + A[i] * out[entry(i)], in place.  entry(i) lies in the last level of the band above,
    //      positions [q0, p0): its stitched values are staged where the packed topology was (no longer needed)
    const int q0 = (beta > 0) ? ls[L0 - 1] : 0, nq = (beta > 0) ? p0 - q0 : 0;
    float *sE = reinterpret_cast<float *>(spk);          // [NC][nq] when it fits
    const bool e_staged = SMEM && nq * NC <= pcap;
    if (e_staged) {
        for (int x = tid; x < nq * NC; x += nthr) {
            const int cc = x / nq, jl = x - cc * nq;
            sE[x] = __ldcg(a.outE + img + (size_t)cc * V + q0 + jl);
        }
        __syncthreads();
    }
    auto outE_at = [&](int cc, int pos) -> float {      // pos in [q0, p0)
        return e_staged ? sE[cc * nq + (pos - q0)] : __ldcg(a.outE + img + (size_t)cc * V + pos);
    };
    if (beta > 0) {
        const float *A = k.A + (size_t)b * V + p0;
        const int *ent = k.ent + (size_t)b * V + p0;
        for (int base = tid; base < P; base += 2 * U * nthr) {
            float ai[2 * U];
            int en[2 * U];
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) {
                const int il = base + u * nthr;
                ai[u] = (il < P) ? A[il] : 0.f;
                en[u] = (il < P) ? ent[il] : q0;
            }
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) {
                const int il = base + u * nthr;
                if (il < P && ai[u] != 0.f) {
#pragma unroll
                    for (int cc = 0; cc < NCT; ++cc) VAL(cc, il) = fmaf(ai[u], outE_at(cc, en[u]), VAL(cc, il));
                }
            }
        }
    }
    __syncthreads();
    prof_mark(a, 8);

    // ---- epilogue
    if (MODE == MODE_FWD) {
        // aggregates in position order (saved), normalised output scattered to vertex order (refine.cu:245)
        for (int base = tid; base < P; base += 2 * U * nthr) {
            int v[2 * U];
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) v[u] = (base + u * nthr < P) ? vtx[base + u * nthr] : 0;
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) {
                const int il = base + u * nthr, i = p0 + il;
                if (il >= P) continue;
                const float z = VAL(NCT - 1, il);
                a.fpos[img + (size_t)(NCT - 1) * V + i] = z;
                if (a.out2) a.out2[(size_t)b * V + v[u]] = z;
#pragma unroll
                for (int cc = 0; cc < NCT - 1; ++cc) {
                    const float o = VAL(cc, il);
                    a.fpos[img + (size_t)cc * V + i] = o;
                    a.out0[((size_t)b * C + cc) * V + v[u]] = o / z;
                    if (a.out1) a.out1[((size_t)b * C + cc) * V + v[u]] = o;
                }
            }
        }
    } else {
        // feature gradient scattered to vertex order and, when asked, the edge-weight gradient (refine.cu:136-199)
        const size_t fimg = (size_t)b * (C + 1) * V;
        for (int base = tid; base < P; base += U * nthr) {
            int v[U], par[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const bool ok = base + u * nthr < P;
                v[u] = ok ? vtx[base + u * nthr] : 0;
                par[u] = ok ? pp[base + u * nthr] : 0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int il = base + u * nthr, i = p0 + il, p = par[u];
                if (il >= P) continue;
                const float wi = Wl(il);
                float gw = 0.f;
                for (int cc = 0; cc < C; ++cc) {
                    const float g1 = VAL(cc, il);
                    a.out0[((size_t)b * C + cc) * V + v[u]] = g1;
                    if (a.out1 != nullptr && i > 0) {
                        const float g1p = (p >= p0) ? VAL(cc, p - p0) : outE_at(cc, p);
                        const float g2p = (p >= p0) ? VAL(C + cc, p - p0) : outE_at(C + cc, p);
                        const float u1 = a.up[img + (size_t)cc * V + i], u2 = a.up[img + (size_t)(C + cc) * V + i];
                        const float fu = a.fup[fimg + (size_t)cc * V + i], zu = a.fup[fimg + (size_t)C * V + i];
                        const float fa_p = a.fpos[fimg + (size_t)cc * V + p], z_p = a.fpos[fimg + (size_t)C * V + p];
                        const float ga = u1 * (fa_p - wi * fu) + fu * (g1p - wi * u1);
                        const float gb = u2 * (z_p - wi * zu) + zu * (g2p - wi * u2);
                        gw += ga - gb;
                    }
                }
                if (a.out1 != nullptr) a.out1[(size_t)b * V + i] = gw;
            }
        }
    }
    prof_mark(a, 9);
    // the last band through re-arms the flag for the next launch
    if (tid == 0) {
        const int old = atomicAdd(sync + 4, 1);
        if (old == nb - 1) {
            sync[4] = 0;
            sync[3] = 0;
        }
    }
}

// ------------------------------------------------------------------------------------------------ the stitcher
// One CTA per image.  State = the last level E(beta) of every band, all bands side by side ("compact index"
// x = off[beta] + (j - e0[beta])): A, entry (as index into E(beta-1)), the run of E(beta+1) entries below each entry,
// and one value per swept channel.  U2 and D2 are then ~nb short steps on shared memory, one warp per channel.
//   U2 (deepest band first):   X[j'] = base[j'] + sum_{j in E(beta), entry(j) = j'} A[j] * X[j]     for j' in E(beta-1)
//   D2 (shallowest band first): out[j] = B[j] + A[j] * out[entry(j)]                                 for j in E(beta)
// entry(j) is non-decreasing along a level (the canonical BFS order sorts a level by parent position), so the entries
// under one j' are a contiguous run [rs, re) -- summed in index order: deterministic.
// Every transfer between global and shared memory is one flat loop over the compact index (band found by bisection of
// the offsets) in batches of independent loads.
template <int NCT>
__device__ void resident_stitch(const TreeView &t, const Consts &k, const ResidentArgs &a, int b, char *smem) {
    int *s_e0 = reinterpret_cast<int *>(smem), *s_n = s_e0 + NBMAX + 1, *s_off = s_n + NBMAX + 1;   // 3 x (NBMAX+2) ints
    smem += 3 * (NBMAX + 2) * sizeof(int) + 8;
    const int tid = threadIdx.x, nthr = blockDim.x, V = t.V, lane = tid & 31, warp = tid >> 5;
    constexpr int NC = NCT;
    const int *ls = t.ls + (size_t)b * (V + 1);
    const int *plan = t.bands + (size_t)b * PLAN_STRIDE;
    const float *gA = k.A + (size_t)b * V;
    const int *gent = k.ent + (size_t)b * V;
    const size_t img = (size_t)b * NC * V;
    int *sync = a.sync + (size_t)b * SYNC_INTS;
    const int nb = plan[0];
    for (int q = tid; q < nb; q += nthr) {
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
    const int T = s_off[nb], tcap = a.tcap;
    const bool SMEM = T <= tcap;   // shared-memory mode when the last levels of all bands fit together
    float *sA = reinterpret_cast<float *>(smem);                        // [tcap]
    float *sX = sA + tcap;                                              // [NC][tcap]
    unsigned short *sEnt = reinterpret_cast<unsigned short *>(sX + (size_t)NC * tcap);   // [tcap] entry, local to E(beta-1)
    unsigned short *sGd = sEnt + tcap;   // [tcap] bits 0-14: index of the entry inside its run (entries with one entry(j)), bit 15: last of the run
    auto pos_of = [&](int x) -> int {   // global position of compact index x: bisection for the largest beta with off <= x
        int lo = 0, hi = nb - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_off[mid] <= x) lo = mid; else hi = mid - 1;
        }
        return s_e0[lo] + (x - s_off[lo]);
    };
    constexpr int U = 4;
    // sX[c][x] <- src[c][pos(x)] for x < n, all channels, U x NC loads in flight per thread
    auto fetch = [&](const float *src, int n) {
        for (int base = tid; base < n; base += U * nthr) {
            int pos[U];
            float v[U][NCT];
#pragma unroll
            for (int u = 0; u < U; ++u) pos[u] = (base + u * nthr < n) ? pos_of(base + u * nthr) : -1;
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int cc = 0; cc < NCT; ++cc) v[u][cc] = (pos[u] >= 0) ? __ldcg(src + img + (size_t)cc * V + pos[u]) : 0.f;
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (pos[u] >= 0) {
#pragma unroll
                    for (int cc = 0; cc < NCT; ++cc) sX[(size_t)cc * tcap + base + u * nthr] = v[u][cc];
                }
        }
    };
    auto flush = [&](float *dst, int n) {
        for (int x = tid; x < n; x += nthr) {
            const int pos = pos_of(x);
#pragma unroll
            for (int cc = 0; cc < NCT; ++cc) dst[img + (size_t)cc * V + pos] = sX[(size_t)cc * tcap + x];
        }
    };
    prof_mark(a, 0);
    if (SMEM) {
        // ---- while the bands sweep: A, entry and the runs
        for (int base = tid; base < T; base += U * nthr) {
            int beta[U], en[U], next[U];
            float av[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int x = base + u * nthr;
                beta[u] = -1;
                if (x < T) {
                    int lo = 0, hi = nb - 1;
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) >> 1;
                        if (s_off[mid] <= x) lo = mid; else hi = mid - 1;
                    }
                    beta[u] = lo;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                en[u] = next[u] = -1;
                av[u] = 0.f;
                if (beta[u] >= 0) {
                    const int x = base + u * nthr, jl = x - s_off[beta[u]], e0 = s_e0[beta[u]], n = s_n[beta[u]];
                    av[u] = gA[e0 + jl];
                    if (beta[u] > 0) {
                        en[u] = gent[e0 + jl];
                        next[u] = (jl < n - 1) ? gent[e0 + jl + 1] : -1;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (beta[u] < 0) continue;
                const int x = base + u * nthr;
                sA[x] = av[u];
                sEnt[x] = (beta[u] > 0) ? (unsigned short)(en[u] - s_e0[beta[u] - 1]) : (unsigned short)0;
                sGd[x] = (next[u] != en[u]) ? (unsigned short)0x8000 : (unsigned short)0;
            }
        }
        __syncthreads();
        // index inside the run: look back while the entry does not change (runs are a few dozen entries at most)
        for (int x = tid; x < T; x += nthr) {
            int lo = 0, hi = nb - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (s_off[mid] <= x) lo = mid; else hi = mid - 1;
            }
            const int first = s_off[lo], e = sEnt[x];
            int d = 0;
            while (x - d - 1 >= first && sEnt[x - d - 1] == e && d < 0x7fff) ++d;
            sGd[x] = (unsigned short)(sGd[x] | d);
        }
    }
    prof_mark(a, 1);
    // ---- U2
    cta_wait(sync + 0, nb);
    prof_mark(a, 2);
    if (tid == 0) sync[0] = 0;
    const int Tx = s_off[nb - 1];   // entries that receive something from below: all but the deepest band's
    if (SMEM) {
        fetch(a.X, Tx);
        __syncthreads();
        prof_mark(a, 3);
        if (warp < NC) {
            float *X = sX + (size_t)warp * tcap;
            // step beta consumes X of E(beta) and completes X of E(beta-1).  One lane per ENTRY of E(beta) (runs are
            // very uneven: most vertices of a last level are dead ends, a few carry dozens of descendants), products
            // A*X summed per run by a segmented warp scan, the run's last lane adds the total to its base.  Two
            // 32-entry chunks are scanned together so that their shuffles overlap.
            for (int beta = nb - 2; beta >= 1; --beta) {
                const int off = s_off[beta], qoff = s_off[beta - 1], ne = s_n[beta];
                float carry = 0.f;   // scanned value of the previous chunk's last lane
                for (int jb = 0; jb < ne; jb += 64) {
                    const int j1 = jb + lane, j2 = j1 + 32;
                    const bool ok1 = j1 < ne, ok2 = j2 < ne;
                    const unsigned g1 = ok1 ? sGd[off + j1] : 0u, g2 = ok2 ? sGd[off + j2] : 0u;
                    const int e1 = ok1 ? sEnt[off + j1] : 0, e2 = ok2 ? sEnt[off + j2] : 0;
                    float v1 = ok1 ? sA[off + j1] * X[off + j1] : 0.f, v2 = ok2 ? sA[off + j2] * X[off + j2] : 0.f;
                    const float b1 = X[qoff + e1], b2 = X[qoff + e2];
                    const int r1 = (int)(g1 & 0x7fffu), r2 = (int)(g2 & 0x7fffu);
                    const int d1 = min(r1, lane), d2 = min(r2, lane);
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const float t1 = __shfl_up_sync(0xffffffffu, v1, o), t2 = __shfl_up_sync(0xffffffffu, v2, o);
                        if (d1 >= o) v1 += t1;
                        if (d2 >= o) v2 += t2;
                    }
                    if (r1 > lane) v1 += carry;        // the run began in an earlier chunk
                    const float c1 = __shfl_sync(0xffffffffu, v1, 31);
                    if (r2 > lane) v2 += c1;
                    carry = __shfl_sync(0xffffffffu, v2, 31);
                    if (ok1 && (g1 & 0x8000u)) X[qoff + e1] = b1 + v1;
                    if (ok2 && (g2 & 0x8000u)) X[qoff + e2] = b2 + v2;
                }
                __syncwarp();
            }
        }
        __syncthreads();
        prof_mark(a, 4);
        flush(a.X, Tx);
    } else if (warp < NC) {
        // oversized: the same recurrence on global memory, the run found by bisection on the (monotone) entries
        float *X = a.X + img + (size_t)warp * V;
        for (int beta = nb - 2; beta >= 1; --beta) {
            const int e0 = s_e0[beta], n = s_n[beta], q0 = s_e0[beta - 1], nq = s_n[beta - 1];
            for (int jl = lane; jl < nq; jl += 32) {
                int lo = 0, hi = n;
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (gent[e0 + mid] < q0 + jl) lo = mid + 1; else hi = mid; }
                float acc = __ldcg(X + q0 + jl);
                for (int jj = lo; jj < n && gent[e0 + jj] == q0 + jl; ++jj) acc = fmaf(gA[e0 + jj], __ldcg(X + e0 + jj), acc);
                X[q0 + jl] = acc;
            }
            __threadfence_block();
            __syncwarp();
        }
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        st_release(sync + 1, 1);
    }
    prof_mark(a, 5);
    // ---- D2
    cta_wait(sync + 2, nb);
    prof_mark(a, 6);
    if (tid == 0) { sync[2] = 0; sync[1] = 0; }
    if (SMEM) {
        fetch(a.BvE, T);
        __syncthreads();
        prof_mark(a, 7);
        if (warp < NC) {
            float *O = sX + (size_t)warp * tcap;
            for (int beta = 1; beta < nb; ++beta) {
                const int off = s_off[beta], qoff = s_off[beta - 1], n = s_n[beta];
                for (int jl = lane; jl < n; jl += 64) {   // two entries per lane, their loads issued together
                    const int jl2 = jl + 32;
                    const bool two = jl2 < n;
                    const int e1 = sEnt[off + jl], e2 = two ? sEnt[off + jl2] : 0;
                    const float a1 = sA[off + jl], a2 = two ? sA[off + jl2] : 0.f;
                    const float b1 = O[off + jl], b2 = two ? O[off + jl2] : 0.f;
                    const float o1 = fmaf(a1, O[qoff + e1], b1), o2 = fmaf(a2, O[qoff + e2], b2);
                    O[off + jl] = o1;
                    if (two) O[off + jl2] = o2;
                }
                __syncwarp();
            }
        }
        __syncthreads();
        prof_mark(a, 8);
        flush(a.outE, T);
    } else if (warp < NC) {
        const float *Bv = a.BvE + img + (size_t)warp * V;
        float *O = a.outE + img + (size_t)warp * V;
        for (int beta = 0; beta < nb; ++beta) {
            const int e0 = s_e0[beta], n = s_n[beta];
            for (int jl = lane; jl < n; jl += 32) {
                float o = __ldcg(Bv + e0 + jl);
                if (beta > 0) o = fmaf(gA[e0 + jl], __ldcg(O + gent[e0 + jl]), o);
                O[e0 + jl] = o;
            }
            __threadfence_block();
            __syncwarp();
        }
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        st_release(sync + 3, 1);
    }
    prof_mark(a, 9);
}

template <int MODE, int NCT>
__global__ void __launch_bounds__(NCT * RSLOTS, NCT <= 4 ? RES_MINB : (NCT <= 6 ? 2 : 1))
resident_kernel(TreeView t, Consts k, ResidentArgs a) {
    extern __shared__ float4 smem4[];
    char *smem = reinterpret_cast<char *>(smem4);
    const int b = a.b0 + blockIdx.y, beta = blockIdx.x;
    if (beta == (int)gridDim.x - 1) {
        resident_stitch<NCT>(t, k, a, b, smem);
        return;
    }
    int L0, L1, nb;
    if (!band_geom(t, b, beta, L0, L1, nb)) return;
    const int *ls = t.ls + (size_t)b * (t.V + 1);
    const int p0 = ls[L0], p1 = ls[L1];
    if (p1 - p0 <= a.pcap && L1 - L0 <= LSCAP) resident_band<MODE, true, NCT>(t, k, a, b, beta, nb, L0, L1, p0, p1, smem);
    else resident_band<MODE, false, NCT>(t, k, a, b, beta, nb, L0, L1, p0, p1, smem);
}
