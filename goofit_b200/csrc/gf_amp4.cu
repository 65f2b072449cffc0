// gf_amp4.cu — Amp4Body caches and the Monte-Carlo normalisation integral.
//   k_amp4_sf     ONE pass per event set (data or MC), once: get4Vecs, then every spin factor and the pair masses the
//                 lineshapes are evaluated at (parameter-independent kinematics, cached in `kin` planes)
//   k_amp4_cache  ONE pass per event set: every flagged lineshape from the cached pair masses, no four-vectors
//                 (the reference launches one thrust::transform per spin factor / lineshape and
//                 recomputes get4Vecs in each: Amp4Body.cu:432-525, detail/{SF,LS,NormSpin,NormLS}Calculator.cu)
//   k_amp4_amps   A_a = 1/sqrt(nPerm) * sum_perm prod LS * prod SF for every flagged amplitude
//                 (detail/AmpCalc.cu:14-59); also run on the MC sample so that
//   k_amp4_norm   sum_m |sum_a c_a A_a(m)|^2 streams nAmp*16 B per MC event instead of rebuilding the
//                 amplitudes from nLS*16 + nSF*8 B on every call (detail/NormIntegrator.cu:15-68); the
//                 per-event arithmetic is unchanged.
// SoA layout: ls[i*ld + e] (complex), sf[k*ld + e] (real), amps[a*ld + e] (complex).
#include "gf_internal.cuh"
#include "gf_amp4.cuh"
#define GF_GRAM_T 4 // amplitudes per Gram tile side: 16 complex accumulators per thread

struct gf_amp4_set { // one event sample (data or MC) with its caches
    long n = 0, ld = 0;
    const double *cols[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    double *owned_cols = nullptr;
    double *weights = nullptr;
    double *dtime = nullptr, *sigma = nullptr; // per-event decay time and resolution of a WEIGHTED normalisation sample (Amp4Body_TD)
    double2 *ls = nullptr;
    double *sf = nullptr;
    double2 *amps = nullptr;
    double *kin = nullptr; // parameter-independent pair masses of the lineshapes: [n_kin][ld], filled with the spin factors
};
// which kin plane holds the resonant-pair mass (m) and, for three-body pairs, the two-body sub-mass d1 (d) of lineshape i;
// -1: not needed (pairs (12), (34) are data columns; two-body pairs have constant daughters)
struct gf_amp4_kinmap {
    signed char m[64], d[64];
    int n;
};
struct gf_amp4_state {
    int node = -1, slot = -1, nLS = 0, nSF = 0, nAmp = 0, td = 0;
    gf_amp4_kinmap kin{};
    gf_amp4_set data, mc;
    // Gram matrix of the MC amplitudes, tile by tile: [tile][GF_GRAM_T * GF_GRAM_T][re, im]
    double *d_gram = nullptr, *d_gram_partials = nullptr, *h_gram = nullptr;
    unsigned *d_gram_ticket = nullptr;
    int gram_blocks = 0;
    // Gram matrix as the 8x8 MMA tiles of k_amp4_update (gf_amp4_cache_gram)
    double *d_mma = nullptr, *d_mma_partials = nullptr, *h_mma = nullptr;
    unsigned *d_mma_ticket = nullptr;
    int mma_blocks = 0;
    bool mma_valid = false;
    bool gram_valid = false;               // d_gram / h_gram hold every tile of the current MC amplitudes
    unsigned long long gram_pending = 0;   // MC amplitudes rebuilt since the last gf_amp4_gram
};

static void free_set(gf_amp4_set &s) {
    cudaFree(s.owned_cols); cudaFree(s.weights); cudaFree(s.dtime); cudaFree(s.sigma); cudaFree(s.ls); cudaFree(s.sf); cudaFree(s.amps); cudaFree(s.kin);
    s = gf_amp4_set();
}
void gf_amp4_free(gf_ctx *c) {
    for(auto *s : c->amp4) {
        free_set(s->data);
        free_set(s->mc);
        cudaFree(s->d_gram); cudaFree(s->d_gram_partials); cudaFree(s->d_gram_ticket);
        if(s->h_gram) cudaFreeHost(s->h_gram);
        cudaFree(s->d_mma); cudaFree(s->d_mma_partials); cudaFree(s->d_mma_ticket);
        if(s->h_mma) cudaFreeHost(s->h_mma);
        delete s;
    }
    c->amp4.clear();
}
static gf_amp4_state *find_amp4(gf_ctx *c, int node, bool create) {
    for(auto *s : c->amp4)
        if(s->node == node) return s;
    if(!create) return nullptr;
    auto *s = new gf_amp4_state();
    s->node = node;
    c->amp4.push_back(s);
    return s;
}

struct gf_amp4_args {
    const double *cols[5];
    long n, ld;
    double2 *ls;
    double *sf;
    double2 *amps;
    double *kin;
};

// ints of the AMP4 node: {aux slot, nLS, nSF, nAmp, td, ls_node[nLS], sf_node[nSF], start[nAmp], records...}
// record of amplitude a = {nLS_a, nSF_a, nPerm, LS indices..., SF indices...} (Amp4Body.cu:222-231)
// Spin factors are computed once per event sample; their tensor algebra wants many registers, so they get their own
// kernel at one block per SM instead of setting the register budget of the lineshape kernel, which reruns whenever a
// mass or width moves.
#ifndef GF_SF_BLOCKS
#define GF_SF_BLOCKS 2 // measured: first Amp4Body call 75 -> 58 ms at two resident blocks (128 registers) against one (158)
#endif
__global__ void __launch_bounds__(GF_THREADS, GF_SF_BLOCKS) k_amp4_sf(const gf_dev_view V, int node, gf_amp4_args A, const gf_amp4_kinmap K) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const int32_t *in = P.ints + nd.int_off;
    const double *mc = P.consts + nd.con_off; // M, m1, m2, m3, m4
    const int nLS = in[1], nSF = in[2];
    const double M = mc[0], m1 = mc[1], m2 = mc[2], m3 = mc[3], m4 = mc[4];
    const long stride = (long)gridDim.x * blockDim.x;
    for(long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < A.n; e += stride) {
        double m12 = __ldg(A.cols[0] + e), m34 = __ldg(A.cols[1] + e), c12 = __ldg(A.cols[2] + e),
               c34 = __ldg(A.cols[3] + e), phi = __ldg(A.cols[4] + e);
        double vecs[16];
        gf_get4vecs(vecs, m12, m34, c12, c34, phi, M, m1, m2, m3, m4);
        for(int k = 0; k < nSF; ++k) {
            const gf_node &sn = P.nodes[in[5 + nLS + k]];
            int kind = P.ints[sn.int_off];
            A.sf[(size_t)k * A.ld + e] = gf_spinfactor(kind, P.consts + sn.con_off, vecs);
        }
        // The invariant masses a lineshape is evaluated at depend on the event only, never on a fit parameter: keep them,
        // so that the lineshape kernel (which reruns whenever a mass or width moves) needs no four-vectors at all.
        // Each plane is written by the first lineshape that maps to it (same pair => same value).
        for(int i = 0; i < nLS; ++i) {
            if(K.m[i] < 0) continue;
            bool first = true;
            for(int j = 0; j < i; ++j) first = first && K.m[j] != K.m[i];
            if(!first) continue;
            const gf_node &ln = P.nodes[in[5 + i]];
            int pair = (int)(P.consts + ln.con_off)[0];
            double d1, d2;
            double mres = gf_getmass(pair, d1, d2, vecs, m1, m2, m3, m4);
            A.kin[(size_t)K.m[i] * A.ld + e] = mres;
            if(K.d[i] >= 0) A.kin[(size_t)K.d[i] * A.ld + e] = d1;
        }
    }
}

// every flagged lineshape in one pass, from the data columns (pairs (12), (34)) or the cached pair masses: no
// four-vector reconstruction (the reference redoes get4Vecs inside every LSCalculator call, LSCalculator.cu:30-60)
__global__ void __launch_bounds__(GF_THREADS, 3)
    k_amp4_cache(const gf_dev_view V, int node, gf_amp4_args A, unsigned long long ls_mask, const gf_amp4_kinmap K) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const int32_t *in = P.ints + nd.int_off;
    const double *mc = P.consts + nd.con_off; // M, m1, m2, m3, m4
    const int nLS = in[1];
    const double m1 = mc[1], m2 = mc[2], m3 = mc[3], m4 = mc[4];
    const long stride = (long)gridDim.x * blockDim.x;
    // lineshape outer, events inner: everything that depends on the lineshape's parameters only (its q0-type momenta,
    // the k normalisation with two square roots and a division, the pair's constant daughters) is loop-invariant for the
    // event loop and is hoisted out of it by the compiler
    for(int i = 0; i < nLS; ++i) {
        if(!((ls_mask >> i) & 1ull)) continue;
        const gf_node &ln = P.nodes[in[5 + i]];
        const double *lc = P.consts + ln.con_off;
        const double *lp = P.params + ln.par_off;
        const int kind = P.ints[ln.int_off];
        const int pair = (int)lc[0];
        double d1c, d2;
        if(pair < 2) { // LSCalculator.cu:47-53
            d1c = pair == 0 ? m1 : m3;
            d2 = pair == 0 ? m2 : m4;
        } else
            gf_pair_daughters(pair, d1c, d2, m1, m2, m3, m4);
        const double *mcol = pair < 2 ? A.cols[pair] : A.kin + (size_t)K.m[i] * A.ld;
        const double *dcol = (pair >= 2 && K.d[i] >= 0) ? A.kin + (size_t)K.d[i] * A.ld : nullptr;
        double2 *out = A.ls + (size_t)i * A.ld;
        for(long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < A.n; e += stride) {
            const double mres = __ldg(mcol + e);
            const double d1 = dcol ? __ldg(dcol + e) : d1c;
            gf_cplx r = gf_lineshape(kind, lp, lc, mres, d1, d2);
            __stcs(out + e, make_double2(r.re, r.im));
        }
    }
}

// One amplitude with compile-time shape: all of its lineshape / spin-factor loads are issued before the first multiply
// (memory-level parallelism: the generic loop below has one dependent load in flight per thread), then the products are
// formed in exactly the reference's order (AmpCalc.cu:36-57), so the result is bit-identical to the generic path.
// COHERENT: the lineshape planes may have been written earlier in this very kernel (k_amp4_update), so they are read with
// ordinary loads instead of the read-only path.
template <int LS_STEP, int SF_STEP, int NPERM, bool COHERENT = false>
__device__ __forceinline__ double2 gf_amp4_one(const int32_t *rec, const gf_amp4_args &A, long e) {
    constexpr int NLS = LS_STEP * NPERM, NSF = SF_STEP * NPERM;
    double2 L[NLS];
    double S[NSF];
#pragma unroll
    for(int j = 0; j < NLS; ++j) L[j] = COHERENT ? A.ls[(size_t)rec[3 + j] * A.ld + e] : __ldg(A.ls + (size_t)rec[3 + j] * A.ld + e);
#pragma unroll
    for(int j = 0; j < NSF; ++j) S[j] = __ldg(A.sf + (size_t)rec[3 + NLS + j] * A.ld + e);
    gf_cplx total = gf_c(0, 0);
#pragma unroll
    for(int p = 0; p < NPERM; ++p) {
        gf_cplx ret = gf_c(1, 0);
#pragma unroll
        for(int j = 0; j < LS_STEP; ++j) ret = ret * gf_c(L[p * LS_STEP + j].x, L[p * LS_STEP + j].y);
#pragma unroll
        for(int j = 0; j < SF_STEP; ++j) ret = ret * gf_c(S[p * SF_STEP + j], 0.0);
        total = total + ret;
    }
    total = total * (1 / sqrt((double)NPERM));
    return make_double2(total.re, total.im);
}

// amplitude a of event e from the lineshape / spin-factor planes: A_a = 1/sqrt(nPerm) sum_perm prod LS prod SF (AmpCalc.cu:14-59)
template <bool COHERENT>
__device__ __forceinline__ double2 gf_amp4_build(const int32_t *rec, const gf_amp4_args &A, long e) {
    const int numLS = rec[0], numSF = rec[1], nPerm = rec[2];
    // nPerm is 1 or 2 in every decay chain upstream builds: no integer division on the per-event path
    const int LS_step = nPerm == 1 ? numLS : (nPerm == 2 ? numLS >> 1 : numLS / nPerm), SF_step = nPerm == 1 ? numSF : (nPerm == 2 ? numSF >> 1 : numSF / nPerm);
    const int shape = nPerm * 100 + LS_step * 10 + SF_step; // uniform over the grid
    switch(shape) {
    case 221: return gf_amp4_one<2, 1, 2, COHERENT>(rec, A, e); // Bose-symmetrised two-resonance chains (DP4, TDDP4)
    case 121: return gf_amp4_one<2, 1, 1, COHERENT>(rec, A, e);
    case 111: return gf_amp4_one<1, 1, 1, COHERENT>(rec, A, e);
    case 211: return gf_amp4_one<1, 1, 2, COHERENT>(rec, A, e);
    case 131: return gf_amp4_one<3, 1, 1, COHERENT>(rec, A, e);
    case 231: return gf_amp4_one<3, 1, 2, COHERENT>(rec, A, e);
    default: { // any other shape: the reference's loops as they are
        gf_cplx total = gf_c(0, 0);
        for(int p = 0; p < nPerm; ++p) {
            gf_cplx ret = gf_c(1, 0);
            for(int j = p * LS_step; j < (p + 1) * LS_step; ++j) {
                double2 t = A.ls[(size_t)rec[3 + j] * A.ld + e];
                ret = ret * gf_c(t.x, t.y);
            }
            for(int j = p * SF_step; j < (p + 1) * SF_step; ++j) {
                double s = A.sf[(size_t)rec[3 + numLS + j] * A.ld + e];
                ret = ret * gf_c(s, 0.0); // the reference multiplies by complex(sf, 0)
            }
            total = total + ret;
        }
        total = total * (1 / sqrt((double)nPerm));
        return make_double2(total.re, total.im);
    }
    }
}

// The same amplitude from a per-block PLAN in shared memory (k_amp4_update): the plane base pointers are resolved once per block
// instead of once per event (record look-up, 64-bit index * leading dimension), which was a third of the instructions the
// update kernel issued (ncu source view: 450 of 1480 per event). Lineshape planes are read with ordinary loads: some were
// written earlier in this very kernel by the same thread.
#define GF_PLAN_MAXLS 6
#define GF_PLAN_MAXSF 2
struct gf_upd_plan {
    const double2 *ls[GF_PLAN_MAXLS];
    const double *sf[GF_PLAN_MAXSF];
    double2 *out;
    int toff;  // row offset of the amplitude's Re row in the warp's MMA tile, in doubles
    int shape; // nPerm * 100 + LS_step * 10 + SF_step, or 0: not one of the compiled shapes, go through the record
    int amp;
    int same;  // the previous plan of the block has the same shape and the same lineshape planes: their products are reused
};
// prod[p] = the product of the lineshapes of permutation p, kept from the previous amplitude when this one has the very same
// lineshape list (pl.same: the S, P and D waves of one decay chain differ in their spin factors only) - the same intermediate
// value the reference's loop passes through, so nothing changes in the result; 8 of the 16 lineshape loads of a DP4 K* move go
template <int LS_STEP, int SF_STEP, int NPERM>
__device__ __forceinline__ double2 gf_amp4_one_p(const gf_upd_plan &pl, long e, gf_cplx (&prod)[2]) {
    constexpr int NLS = LS_STEP * NPERM, NSF = SF_STEP * NPERM;
    static_assert(NLS <= GF_PLAN_MAXLS && NSF <= GF_PLAN_MAXSF && NPERM <= 2, "plan too small for this shape");
    double S[NSF];
#pragma unroll
    for(int j = 0; j < NSF; ++j) S[j] = __ldg(pl.sf[j] + e);
    if(!pl.same) {
        double2 L[NLS];
#pragma unroll
        for(int j = 0; j < NLS; ++j) L[j] = pl.ls[j][e];
#pragma unroll
        for(int p = 0; p < NPERM; ++p) {
            gf_cplx ret = gf_c(1, 0);
#pragma unroll
            for(int j = 0; j < LS_STEP; ++j) ret = ret * gf_c(L[p * LS_STEP + j].x, L[p * LS_STEP + j].y);
            prod[p] = ret;
        }
    }
    gf_cplx total = gf_c(0, 0);
#pragma unroll
    for(int p = 0; p < NPERM; ++p) {
        gf_cplx ret = prod[p];
#pragma unroll
        for(int j = 0; j < SF_STEP; ++j) ret = ret * gf_c(S[p * SF_STEP + j], 0.0);
        total = total + ret;
    }
    total = total * (1 / sqrt((double)NPERM));
    return make_double2(total.re, total.im);
}
__device__ __forceinline__ int gf_amp4_plan_shape(const int32_t *rec) {
    const int numLS = rec[0], numSF = rec[1], nPerm = rec[2];
    if(nPerm < 1 || nPerm > 2 || numLS % nPerm || numSF % nPerm) return 0;
    const int shape = nPerm * 100 + (numLS / nPerm) * 10 + numSF / nPerm;
    switch(shape) {
    case 221: case 121: case 111: case 211: case 131: case 231: return shape;
    default: return 0;
    }
}

__global__ void __launch_bounds__(GF_THREADS, 3)
    k_amp4_amps(const gf_dev_view V, int node, gf_amp4_args A, unsigned long long amp_mask) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const int32_t *in = P.ints + nd.int_off;
    const int nLS = in[1], nSF = in[2], nAmp = in[3];
    const int32_t *start = in + 5 + nLS + nSF;
    const int32_t *rec0 = start + nAmp;
    const long stride = (long)gridDim.x * blockDim.x;
    for(long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < A.n; e += stride) {
        for(int a = 0; a < nAmp; ++a) {
            if(!((amp_mask >> a) & 1ull)) continue;
            const double2 out = gf_amp4_build<false>(rec0 + start[a], A, e);
            A.amps[(size_t)a * A.ld + e] = out;
        }
    }
}

// sum over MC events of |sum_a A_a(m) c_a|^2 (optionally weighted); deterministic grid reduction
__global__ void __launch_bounds__(GF_THREADS, 4)
    k_amp4_norm(const gf_dev_view V, int node, const double2 *amps, long ld, long n, const double *weights,
                double *partials, unsigned *ticket, double *out, const gf_peer_view PV) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red[32];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const double *coef = P.params + nd.par_off;
    const int nAmp = P.ints[nd.int_off + 3];
    double acc = 0;
    const long stride = (long)gridDim.x * blockDim.x;
    for(long m = (long)blockIdx.x * blockDim.x + threadIdx.x; m < n; m += stride) {
        gf_cplx tot = gf_c(0, 0);
#pragma unroll 4
        for(int a = 0; a < nAmp; ++a) {
            double2 t = __ldg(amps + (size_t)a * ld + m);
            tot = tot + gf_c(t.x, t.y) * gf_c(coef[2 * a], coef[2 * a + 1]);
        }
        double v = tot.re * tot.re + tot.im * tot.im;
        if(weights) v *= __ldg(weights + m);
        acc += v;
    }
    gf_grid_finish(&acc, 1, partials, ticket, out, red, PV);
}

// Amp4Body_TD: four sums over MC events, |A|^2, |B|^2, Re(A B*), Im(A B*) (detail/NormIntegrator_TD.cu:15-99); the
// time integrals are applied to the four sums on the host by the resolution's normalization().
__global__ void __launch_bounds__(GF_THREADS, 4)
    k_amp4_norm_td(const gf_dev_view V, int node, const double2 *amps, long ld, long n, const double *weights,
                   double *partials, unsigned *ticket, double *out, const gf_peer_view PV) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red[32];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const double *p = P.params + nd.par_off;
    const int32_t *in = P.ints + nd.int_off;
    const int nAmp = in[3];
    const int32_t *flags = in + in[4] + 1;
    double acc[4] = {0, 0, 0, 0};
    const long stride = (long)gridDim.x * blockDim.x;
    for(long m = (long)blockIdx.x * blockDim.x + threadIdx.x; m < n; m += stride) {
        double Are, Aim, Bre, Bim;
        gf_amp4td_AB(p, flags, nAmp, [&](int a) { return __ldg(amps + (size_t)a * ld + m); }, Are, Aim, Bre, Bim);
        double w = weights ? __ldg(weights + m) : 1.0;
        acc[0] += w * (Are * Are + Aim * Aim);
        acc[1] += w * (Bre * Bre + Bim * Bim);
        acc[2] += w * (Are * Bre + Aim * Bim);
        acc[3] += w * (Aim * Bre - Are * Bim);
    }
    gf_grid_finish_n(acc, partials, ticket, out, red, PV);
}

// Weighted normalisation events of Amp4Body_TD (detail/NormIntegrator_TD_Weighted.cu:15-122): every MC event carries its own decay
// time, resolution and weight, and the integrand is the time-dependent PDF itself, sum_m w_m R(|A|^2 + |B|^2, |A|^2 - |B|^2,
// A conj(B); tau, t_m, x, y, sigma_m) - no analytic time integral. The amplitudes come from the cached MC planes.
template <bool G3>
__global__ void __launch_bounds__(GF_THREADS, 3)
    k_amp4_norm_tdw(const gf_dev_view V, int node, const double2 *amps, long ld, long n, const double *weights, const double *dtime, const double *sigma,
                    double *partials, unsigned *ticket, double *out, const gf_peer_view PV) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red[32];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const double *p = P.params + nd.par_off;
    const int32_t *in = P.ints + nd.int_off;
    const int nAmp = in[3];
    const int32_t *td = in + in[4];
    const gf_node &rn = P.nodes[td[0]];
    const double *rp = P.params + rn.par_off;
    double acc = 0;
    const long stride = (long)gridDim.x * blockDim.x;
    for(long m = (long)blockIdx.x * blockDim.x + threadIdx.x; m < n; m += stride) {
        const double t = __ldg(dtime + m), sg = __ldg(sigma + m), w = __ldg(weights + m);
        double Are, Aim, Bre, Bim;
        gf_amp4td_AB(p, td + 1, nAmp, [&](int a) { return __ldg(amps + (size_t)a * ld + m); }, Are, Aim, Bre, Bim);
        const double nA = Are * Are + Aim * Aim, nB = Bre * Bre + Bim * Bim;
        const double t3re = Are * Bre + Aim * Bim, t3im = Aim * Bre - Are * Bim; // A conj(B)
        double ret = gf_tres<G3>(rn.op, rp, nA + nB, nA - nB, t3re, t3im, p[0], t, p[1], p[2], sg);
        ret *= w;
        acc += ret;
    }
    gf_grid_finish(&acc, 1, partials, ticket, out, red, PV);
}

// G_ab = sum_m w_m A_a(m) conj(A_b(m)) for one GF_GRAM_T x GF_GRAM_T tile (rows a0.., columns b0..) of the Hermitian
// Gram matrix of the cached MC amplitudes. With G, the normalisation integral of any coefficient set is an
// O(nAmp^2) host sum, sum_ab c_a conj(c_b) G_ab, so the per-call pass over the MC sample (detail/NormIntegrator.cu:15-68,
// NormIntegrator_TD.cu:15-99) is only needed when a lineshape moved. This is the "complex amplitude Gram sum" of the
// north star; plain FP64 FMAs keep it at the HBM roofline (arithmetic intensity 2 x 16 cmul per 128 B), so DMMA would
// add nothing.
__global__ void __launch_bounds__(GF_THREADS, 2)
    k_amp4_gram(const double2 *amps, long ld, long n, const double *weights, int nAmp, int a0, int b0, double *partials,
                unsigned *ticket, double *out, const gf_peer_view PV) {
    __shared__ double red[32];
    double acc[2 * GF_GRAM_T * GF_GRAM_T];
#pragma unroll
    for(int k = 0; k < 2 * GF_GRAM_T * GF_GRAM_T; ++k) acc[k] = 0;
    const long stride = (long)gridDim.x * blockDim.x;
    for(long m = (long)blockIdx.x * blockDim.x + threadIdx.x; m < n; m += stride) {
        double2 A[GF_GRAM_T], B[GF_GRAM_T];
#pragma unroll
        for(int i = 0; i < GF_GRAM_T; ++i) {
            A[i] = a0 + i < nAmp ? __ldg(amps + (size_t)(a0 + i) * ld + m) : make_double2(0.0, 0.0);
            B[i] = b0 + i < nAmp ? __ldg(amps + (size_t)(b0 + i) * ld + m) : make_double2(0.0, 0.0);
        }
        const double w = weights ? __ldg(weights + m) : 1.0;
#pragma unroll
        for(int i = 0; i < GF_GRAM_T; ++i) {
            const double ax = w * A[i].x, ay = w * A[i].y;
#pragma unroll
            for(int j = 0; j < GF_GRAM_T; ++j) { // A_i conj(B_j)
                acc[2 * (i * GF_GRAM_T + j)] += ax * B[j].x + ay * B[j].y;
                acc[2 * (i * GF_GRAM_T + j) + 1] += ay * B[j].x - ax * B[j].y;
            }
        }
    }
    gf_grid_finish_n(acc, partials, ticket, out, red, PV);
}

// ---- fused update: lineshapes -> amplitudes -> Gram matrix in ONE pass over a sample -----------------------------------------
// When a mass or width moves, the reference re-runs one thrust::transform per changed lineshape, one per amplitude, and then
// its NormIntegrator over the whole MC sample (Amp4Body.cu:462-558, detail/NormLSCalculator.cu:11-62, NormIntegrator.cu:15-68).
// k_amp4_update does the whole update of one sample in one pass: per event it evaluates the flagged lineshapes from the cached
// pair masses (stored: a later call may move a different lineshape), rebuilds the amplitudes that contain them (stored), and
// - on the MC sample in Gram mode - accumulates the full Hermitian Gram matrix G_ab = sum_m w_m A_a(m) conj(A_b(m)) with FP64
// tensor-core MMAs (mma.sync m8n8k4): the amplitudes of 32 events are staged per warp in shared memory as the real 16 x 32
// matrix X = [Re A; Im A]; with one 8-amplitude block, Re G = Re.Re^T + Im.Im^T and C2 = Im.Re^T (Im G_ab = C2_ab - C2_ba)
// are three MMAs per four events whose 8x8 accumulators live distributed over the warp (4 registers per thread, instead of the
// 72 a per-thread 8x8 complex accumulator would need - the reason k_amp4_gram tiles 4x4 and re-reads the planes per tile).
// The rebuilt amplitudes never travel through HBM on their way to the sum, and the untouched ones are read once.
#define GF_PF_K 4
// store policy of the planes the update writes (A/B switches; measured)
#ifndef GF_UPD_ST
#define GF_UPD_ST 1
#endif
#if GF_UPD_ST == 1 // amplitudes streamed (evict-first): they are not read again in this launch
#define GF_UPD_STORE_LS(p, v) (*(p) = (v))
#define GF_UPD_STORE_AMP(p, v) __stcs((p), (v))
#elif GF_UPD_ST == 2
#define GF_UPD_STORE_LS(p, v) __stcg((p), (v))
#define GF_UPD_STORE_AMP(p, v) __stcs((p), (v))
#else
#define GF_UPD_STORE_LS(p, v) (*(p) = (v))
#define GF_UPD_STORE_AMP(p, v) (*(p) = (v))
#endif
#ifndef GF_UPD_PF_L2
#define GF_UPD_PF_L2 1
#endif
#ifndef GF_UPD_PF_POS
#define GF_UPD_PF_POS 2
#endif
#ifndef GF_UPD_PF_L1
#define GF_UPD_PF_L1 0 // measured after the L2 request moved late: 1.392 ms with, 1.358 ms without (DP4 K* move)
#endif
#define GF_UPD_TS 36 // tile row stride in doubles: 4 mod 16, so the MMA fragment reads (8 rows x 4 events) of a half-warp hit 16 different 8-byte banks
struct gf_ls_slot { // per-lineshape record in shared memory, filled once per block for the flagged lineshapes
    gf_bw_inv inv;  // parameter-only part (valid when has_inv: constant daughters and a Breit-Wigner-type kind)
    const double *mcol, *dcol, *lp, *lc;
    double2 *out;
    double d1c, d2;
    double r2;       // meson radius squared (Breit-Wigner family)
    int kind, has_inv; // has_inv: 0 no; 1 yes; 2 + L: yes, and a relativistic Breit-Wigner with orbital momentum L <= 2 (gf_ls_bw_pt_t<L>)
    int FF, pad;
};
__device__ __forceinline__ void gf_dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
constexpr int upd_tiles(int nb) { return nb == 1 ? 2 : (nb == 2 ? 6 : 0); } // 8x8 accumulator tiles: (C1, C2) per block pair (0,0) [, (0,1), (1,1)]
// measured (tools/ab_compare.sh, K*(892) move on 10^7 + 10^7 events): the MC copy at three resident blocks (80 registers, small
// spills) against two (128): 1.80 -> 1.66 ms per call (DP4), 2.00 -> 1.77 ms (TDDP4); the 16-amplitude copy keeps two
#ifndef GF_UPD_BLOCKS
#define GF_UPD_BLOCKS 3
#endif
constexpr int upd_blocks_per_sm(int nb) { return nb == 2 ? 2 : GF_UPD_BLOCKS; }
inline size_t gf_upd_smem(size_t prog, int nLS, int nb) {
    return gf_align16(prog) + gf_align16(sizeof(gf_ls_slot) * (size_t)std::max(1, nLS)) + sizeof(double) * (GF_THREADS / 32) * 16 * nb * GF_UPD_TS;
}
// NB = 0: no Gram sum (data sample); 1: up to 8 amplitudes; 2: up to 16
template <int NB>
__global__ void __launch_bounds__(GF_THREADS, upd_blocks_per_sm(NB))
    k_amp4_update(const gf_dev_view V, int node, gf_amp4_args A, unsigned long long ls_mask, unsigned long long amp_mask, const gf_amp4_kinmap K,
                  const double *weights, double *partials, unsigned *ticket, double *out, const gf_peer_view PV) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ unsigned char flagged[64], untouched[64], rebuilt[64];
    __shared__ int nflag, nun, nre;
    __shared__ bool is_last;
    // planes this launch reads, as (base pointer, bytes per event): the 128-byte lines of a warp's NEXT 32 events are requested
    // into L2 while the current ones compute (measured on the DP4 K*(892) move: 1.96 -> 1.77 ms per likelihood call)
    __shared__ const char *pf_base[160];
    __shared__ unsigned char pf_bytes[160];
    __shared__ int npf, npf_l1;
    // resolved per block: where each rebuilt amplitude reads and writes, where each standing one is read, and which lines of the
    // next group every lane requests (GF_PF_K planes per lane: plane (lane >> 2) + 8 k, line lane & 3 of its 32-event span)
    __shared__ gf_upd_plan plans[16];
    __shared__ const double2 *un_ptr[16];
    __shared__ int un_off[16];
    __shared__ const char *pf_lane[GF_PF_K][32];
    __shared__ unsigned char pf_is16[GF_PF_K][32];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const int32_t *in = P.ints + nd.int_off;
    const double *mc = P.consts + nd.con_off; // M, m1, m2, m3, m4
    const int nLS = in[1], nSF = in[2], nAmp = in[3];
    const int32_t *start = in + 5 + nLS + nSF;
    const int32_t *rec0 = start + nAmp;
    gf_ls_slot *slots = reinterpret_cast<gf_ls_slot *>(smem + gf_align16(gf_sprog_bytes(V.n_nodes, V.n_params, V.n_consts, V.n_norms, V.n_ints)));
    double *tiles = reinterpret_cast<double *>(reinterpret_cast<unsigned char *>(slots) + gf_align16(sizeof(gf_ls_slot) * (size_t)(nLS > 1 ? nLS : 1)));
    if((int)threadIdx.x < nLS && ((ls_mask >> threadIdx.x) & 1ull)) {
        const int i = threadIdx.x;
        const gf_node &ln = P.nodes[in[5 + i]];
        gf_ls_slot L;
        L.lc = P.consts + ln.con_off;
        L.lp = P.params + ln.par_off;
        L.kind = P.ints[ln.int_off];
        const int pair = (int)L.lc[0];
        if(pair < 2) { // LSCalculator.cu:47-53
            L.d1c = pair == 0 ? mc[1] : mc[3];
            L.d2 = pair == 0 ? mc[2] : mc[4];
        } else
            gf_pair_daughters(pair, L.d1c, L.d2, mc[1], mc[2], mc[3], mc[4]);
        L.mcol = pair < 2 ? A.cols[pair] : A.kin + (size_t)K.m[i] * A.ld;
        L.dcol = (pair >= 2 && K.d[i] >= 0) ? A.kin + (size_t)K.d[i] * A.ld : nullptr;
        L.out = A.ls + (size_t)i * A.ld;
        L.has_inv = (L.dcol == nullptr && gf_ls_has_inv(L.kind)) ? 1 : 0;
        L.r2 = 0; L.FF = 0; L.pad = 0;
        if(L.has_inv) {
            L.inv = gf_bw_invariants(L.kind == GF_LS_SBW, L.d1c, L.d2, L.lp[0], L.lp[1]);
            const int orbital = (int)L.lc[1];
            L.FF = (int)L.lc[2];
            L.r2 = L.lc[3] * L.lc[3];
            if(L.kind == GF_LS_RBW && orbital >= 0 && orbital <= 2) L.has_inv = 2 + orbital;
        }
        slots[i] = L;
    }
    if(threadIdx.x == 0) {
        int nf = 0, nu = 0;
        for(int i = 0; i < nLS; ++i)
            if((ls_mask >> i) & 1ull) flagged[nf++] = (unsigned char)i;
        for(int a = 0; a < nAmp; ++a)
            if(!((amp_mask >> a) & 1ull)) untouched[nu++] = (unsigned char)a;
        nflag = nf;
        nun = nu;
        int np = 0, nr = 0;
        unsigned long long ls_in = 0, sf_in = 0; // standing lineshapes / spin factors of the amplitudes to rebuild
        for(int a = 0; a < nAmp; ++a) {
            if(!((amp_mask >> a) & 1ull)) continue;
            rebuilt[nr++] = (unsigned char)a;
            const int32_t *rec = rec0 + start[a];
            for(int j = 0; j < rec[0]; ++j) ls_in |= 1ull << rec[3 + j];
            for(int j = 0; j < rec[1]; ++j) sf_in |= 1ull << rec[3 + rec[0] + j];
        }
        nre = nr;
        ls_in &= ~ls_mask;
        for(int i = 0; i < nLS; ++i)
            if((ls_in >> i) & 1ull) { pf_base[np] = reinterpret_cast<const char *>(A.ls + (size_t)i * A.ld); pf_bytes[np++] = 16; }
        for(int k = 0; k < nSF; ++k)
            if((sf_in >> k) & 1ull) { pf_base[np] = reinterpret_cast<const char *>(A.sf + (size_t)k * A.ld); pf_bytes[np++] = 8; }
        npf_l1 = GF_UPD_PF_L1 == 2 ? np : 160; // L1 requests: 2 = only the planes read late in the iteration (by the amplitudes)
        if(NB > 0)
            for(int u = 0; u < nu; ++u) { pf_base[np] = reinterpret_cast<const char *>(A.amps + (size_t)untouched[u] * A.ld); pf_bytes[np++] = 16; }
        for(int f = 0; f < nf && np < 156; ++f) { // pair masses of the flagged lineshapes
            const int i = flagged[f];
            const gf_node &ln = P.nodes[in[5 + i]];
            const int pair = (int)(P.consts + ln.con_off)[0];
            const double *mcol = pair < 2 ? A.cols[pair] : A.kin + (size_t)K.m[i] * A.ld;
            const double *dcol = (pair >= 2 && K.d[i] >= 0) ? A.kin + (size_t)K.d[i] * A.ld : nullptr;
            bool seen = false;
            for(int j = 0; j < np; ++j) seen = seen || pf_base[j] == reinterpret_cast<const char *>(mcol);
            if(!seen) { pf_base[np] = reinterpret_cast<const char *>(mcol); pf_bytes[np++] = 8; }
            if(dcol) { pf_base[np] = reinterpret_cast<const char *>(dcol); pf_bytes[np++] = 8; }
        }
        npf = np;
    }
    __syncthreads();
    if(threadIdx.x < 32 * GF_PF_K) {
        const int k = threadIdx.x >> 5, l = threadIdx.x & 31;
        const int j = (l >> 2) + 8 * k;
        const bool on = j < npf && (l & 3) * 128 < 32 * pf_bytes[j];
        pf_lane[k][l] = on ? pf_base[j] + (l & 3) * 128 : nullptr;
        pf_is16[k][l] = (on && pf_bytes[j] == 16 ? 1 : 0) | (on && j < npf_l1 ? 2 : 0);
    }
    if((int)threadIdx.x < nun) {
        const int a = untouched[threadIdx.x];
        un_ptr[threadIdx.x] = A.amps + (size_t)a * A.ld;
        un_off[threadIdx.x] = ((a >> 3) * 16 + (a & 7)) * GF_UPD_TS;
    }
    if((int)threadIdx.x >= 32 && (int)threadIdx.x - 32 < nre) {
        const int k = threadIdx.x - 32;
        const int a = rebuilt[k];
        const int32_t *rec = rec0 + start[a];
        gf_upd_plan pl;
        pl.shape = gf_amp4_plan_shape(rec);
        pl.amp = a;
        pl.same = 0;
        if(k > 0 && pl.shape) {
            const int32_t *prev = rec0 + start[rebuilt[k - 1]];
            bool same = prev[0] == rec[0] && prev[1] == rec[1] && prev[2] == rec[2];
            for(int j = 0; same && j < rec[0]; ++j) same = prev[3 + j] == rec[3 + j];
            pl.same = same ? 1 : 0;
        }
        pl.out = A.amps + (size_t)a * A.ld;
        pl.toff = ((a >> 3) * 16 + (a & 7)) * GF_UPD_TS;
        for(int j = 0; j < GF_PLAN_MAXLS; ++j) pl.ls[j] = (pl.shape && j < rec[0]) ? A.ls + (size_t)rec[3 + j] * A.ld : nullptr;
        for(int j = 0; j < GF_PLAN_MAXSF; ++j) pl.sf[j] = (pl.shape && j < rec[1]) ? A.sf + (size_t)rec[3 + rec[0] + j] * A.ld : nullptr;
        plans[k] = pl;
    }
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    double *T = tiles + (size_t)wib * 16 * NB * GF_UPD_TS; // rows: block b -> [16 b, 16 b + 8) = Re, [16 b + 8, 16 b + 16) = Im
    if(NB > 0)
        for(int k = threadIdx.x; k < (GF_THREADS / 32) * 16 * NB * GF_UPD_TS; k += blockDim.x) tiles[k] = 0.0; // rows of amplitudes >= nAmp stay zero
    __syncthreads();
    constexpr int NT = upd_tiles(NB);
    double C[NT > 0 ? NT : 1][2];
#pragma unroll
    for(int t = 0; t < (NT > 0 ? NT : 1); ++t) C[t][0] = C[t][1] = 0.0;
    const long warp_stride = (long)gridDim.x * (GF_THREADS / 32) * 32;
    for(long base = ((long)blockIdx.x * (GF_THREADS / 32) + wib) * 32; base < A.n; base += warp_stride) {
        const long e = base + lane;
        const bool live = e < A.n;
        // lane l requests line (l & 3) of planes l >> 2, l >> 2 + 8, ... : the current 32 events into L1 now, the warp's NEXT 32
        // events into L2 at GF_UPD_PF_POS (0: now; 1: after the lineshapes; 2: after the amplitudes). Requested a whole iteration
        // ahead, half of the lines were evicted again before their use (one pass of the grid moves 40-80 MB through L2; ncu:
        // 3.20 GB of DRAM reads against 2.11 GB without prefetch), so the request is made late in the iteration.
        auto request = [&](bool l1, bool l2) {
            const long nb0 = base + warp_stride;
            const size_t off8 = (size_t)nb0 * 8, cur8 = (size_t)base * 8;
            const bool more = nb0 + 32 <= A.n; // whole groups only: the lines of a partial last group may lie past the end of the last plane
#pragma unroll
            for(int k = 0; k < GF_PF_K; ++k) {
                const char *q = pf_lane[k][lane];
                if(q) {
                    const int fl = pf_is16[k][lane];
                    const bool wide = fl & 1;
                    if(l1 && (fl & 2)) asm volatile("prefetch.global.L1 [%0];" ::"l"(q + (wide ? 2 * cur8 : cur8)));
                    if(l2 && more) asm volatile("prefetch.global.L2 [%0];" ::"l"(q + (wide ? 2 * off8 : off8)));
                }
            }
        };
        request(GF_UPD_PF_L1 != 0, GF_UPD_PF_L2 != 0 && GF_UPD_PF_POS == 0);
        if(live) {
            if(NB > 0) { // amplitudes that stand: requested first, four at a time, so that their latency overlaps the lineshape arithmetic
                for(int g = 0; g < nun; g += 4) {
                    double2 v[4];
#pragma unroll
                    for(int j = 0; j < 4; ++j)
                        if(g + j < nun) v[j] = __ldg(un_ptr[g + j] + e);
#pragma unroll
                    for(int j = 0; j < 4; ++j)
                        if(g + j < nun) {
                            double *t = T + un_off[g + j] + lane;
                            t[0] = v[j].x;
                            t[8 * GF_UPD_TS] = v[j].y;
                        }
                }
            }
            for(int f = 0; f < nflag; ++f) {
                const int i = flagged[f];
                const gf_ls_slot &L = slots[i];
                const double mres = __ldg(L.mcol + e);
                gf_cplx r;
                if(L.has_inv == 3)
                    r = gf_ls_bw_pt_t<1>(mres, L.lp[0], L.lp[1], L.FF, L.r2, L.inv);
                else if(L.has_inv == 2)
                    r = gf_ls_bw_pt_t<0>(mres, L.lp[0], L.lp[1], L.FF, L.r2, L.inv);
                else if(L.has_inv == 4)
                    r = gf_ls_bw_pt_t<2>(mres, L.lp[0], L.lp[1], L.FF, L.r2, L.inv);
                else if(L.has_inv)
                    r = gf_lineshape_inv(L.kind, L.lp, L.lc, mres, L.inv);
                else
                    r = gf_lineshape(L.kind, L.lp, L.lc, mres, L.dcol ? __ldg(L.dcol + e) : L.d1c, L.d2);
                GF_UPD_STORE_LS(L.out + e, make_double2(r.re, r.im));
            }
            if(GF_UPD_PF_L2 != 0 && GF_UPD_PF_POS == 1) request(false, true);
            gf_cplx prod[2] = {gf_c(0, 0), gf_c(0, 0)};
            for(int k = 0; k < nre; ++k) {
                const gf_upd_plan &pl = plans[k];
                double2 v;
                switch(pl.shape) { // uniform over the grid
                case 221: v = gf_amp4_one_p<2, 1, 2>(pl, e, prod); break; // Bose-symmetrised two-resonance chains (DP4, TDDP4)
                case 121: v = gf_amp4_one_p<2, 1, 1>(pl, e, prod); break;
                case 111: v = gf_amp4_one_p<1, 1, 1>(pl, e, prod); break;
                case 211: v = gf_amp4_one_p<1, 1, 2>(pl, e, prod); break;
                case 131: v = gf_amp4_one_p<3, 1, 1>(pl, e, prod); break;
                case 231: v = gf_amp4_one_p<3, 1, 2>(pl, e, prod); break;
                default: v = gf_amp4_build<true>(rec0 + start[pl.amp], A, e);
                }
                GF_UPD_STORE_AMP(pl.out + e, v);
                if(NB > 0) {
                    double *t = T + pl.toff + lane;
                    t[0] = v.x;
                    t[8 * GF_UPD_TS] = v.y;
                }
            }
            if(GF_UPD_PF_L2 != 0 && GF_UPD_PF_POS == 2) request(false, true);
        } else if(NB > 0) {
            for(int a = 0; a < nAmp; ++a) {
                T[((a >> 3) * 16 + (a & 7)) * GF_UPD_TS + lane] = 0.0;
                T[((a >> 3) * 16 + 8 + (a & 7)) * GF_UPD_TS + lane] = 0.0;
            }
        }
        if(NB > 0) {
            __syncwarp();
            const double wv = weights ? (live ? __ldg(weights + e) : 0.0) : 1.0;
            const int row = lane >> 2;
#pragma unroll
            for(int r = 0; r < 8; ++r) {
                const int col = 4 * r + (lane & 3);
                const double w = weights ? __shfl_sync(0xffffffffu, wv, col) : 1.0;
                const double re0 = T[row * GF_UPD_TS + col], im0 = T[(8 + row) * GF_UPD_TS + col];
                const double wre0 = weights ? w * re0 : re0, wim0 = weights ? w * im0 : im0;
                gf_dmma(C[0][0], C[0][1], wre0, re0);
                gf_dmma(C[0][0], C[0][1], wim0, im0);
                gf_dmma(C[1][0], C[1][1], wim0, re0);
                if(NB > 1) {
                    const double re1 = T[(16 + row) * GF_UPD_TS + col], im1 = T[(24 + row) * GF_UPD_TS + col];
                    const double wre1 = weights ? w * re1 : re1, wim1 = weights ? w * im1 : im1;
                    gf_dmma(C[2][0], C[2][1], wre0, re1); // block pair (0,1): Re G
                    gf_dmma(C[2][0], C[2][1], wim0, im1);
                    gf_dmma(C[3][0], C[3][1], wim0, re1); // Im G = Im.Re^T - Re.Im^T
                    gf_dmma(C[3][0], C[3][1], -wre0, im1);
                    gf_dmma(C[4][0], C[4][1], wre1, re1); // block pair (1,1)
                    gf_dmma(C[4][0], C[4][1], wim1, im1);
                    gf_dmma(C[5][0], C[5][1], wim1, re1);
                }
            }
            __syncwarp();
        }
    }
    if(NB == 0) return;
    // deterministic reduction of the NT 8x8 tiles: warps in order -> block partial; blocks in order by the last block to arrive
    __syncthreads();
    double *red = tiles; // [warp][NT][64], the amplitude tiles are no longer needed
#pragma unroll
    for(int t = 0; t < NT; ++t) {
        red[((size_t)wib * NT + t) * 64 + (lane >> 2) * 8 + 2 * (lane & 3)] = C[t][0];
        red[((size_t)wib * NT + t) * 64 + (lane >> 2) * 8 + 2 * (lane & 3) + 1] = C[t][1];
    }
    __syncthreads();
    for(int k = threadIdx.x; k < NT * 64; k += blockDim.x) {
        double sum = 0;
        for(int w = 0; w < GF_THREADS / 32; ++w) sum += red[(size_t)w * NT * 64 + k];
        partials[(size_t)k * gridDim.x + blockIdx.x] = sum;
    }
    __threadfence();
    __syncthreads();
    if(threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if(!is_last) return;
    __threadfence();
    double *fin = tiles + (size_t)(GF_THREADS / 32) * NT * 64;
    for(int k = threadIdx.x; k < NT * 64; k += blockDim.x) {
        double sum = 0;
        for(unsigned b = 0; b < gridDim.x; ++b) sum += __ldcg(partials + (size_t)k * gridDim.x + b);
        if(PV.nranks > 1) fin[k] = sum; else out[k] = sum;
    }
    if(threadIdx.x == 0) *ticket = 0;
    if(PV.nranks > 1) {
        __syncthreads();
        gf_peer_allreduce(PV, fin, NT * 64, out);
    }
}

// ---- C ABI ---------------------------------------------------------------------------------------------------
static int amp4_check(gf_ctx *c, int node, gf_amp4_state **out, bool create) {
    if(!c) return gf_fail("null ctx");
    if(node < 0 || node >= (int)c->h_nodes.size() || (c->h_nodes[node].op != GF_OP_AMP4 && c->h_nodes[node].op != GF_OP_AMP4_TD))
        return gf_fail("gf_amp4: node is not an Amp4Body record");
    const gf_node &nd = c->h_nodes[node];
    if(nd.nint < 5 || nd.ncon < 5 || nd.nobs < 6) return gf_fail("gf_amp4: malformed Amp4Body record");
    const int32_t *in = c->h_ints.data() + nd.int_off;
    int nLS = in[1], nSF = in[2], nAmp = in[3];
    if(nLS < 0 || nLS > 64 || nSF < 0 || nSF > 64 || nAmp < 1 || nAmp > 64) return gf_fail("gf_amp4: at most 64 lineshapes / spin factors / amplitudes");
    if(nd.nint < 5 + nLS + nSF + nAmp) return gf_fail("gf_amp4: truncated index block");
    if(nd.npar < 2 * nAmp) return gf_fail("gf_amp4: coefficient block too short");
    int nn = (int)c->h_nodes.size();
    for(int i = 0; i < nLS + nSF; i++)
        if(in[5 + i] < 0 || in[5 + i] >= nn) return gf_fail("gf_amp4: bad lineshape / spin-factor record index");
    for(int a = 0; a < nAmp; a++) {
        int off = 5 + nLS + nSF + nAmp + in[5 + nLS + nSF + a];
        if(off + 3 > nd.nint) return gf_fail("gf_amp4: amplitude record out of range");
        int numLS = in[off], numSF = in[off + 1], nPerm = in[off + 2];
        if(nPerm < 1 || off + 3 + numLS + numSF > nd.nint) return gf_fail("gf_amp4: amplitude record out of range");
        for(int j = 0; j < numLS; j++)
            if(in[off + 3 + j] < 0 || in[off + 3 + j] >= nLS) return gf_fail("gf_amp4: lineshape index out of range");
        for(int j = 0; j < numSF; j++)
            if(in[off + 3 + numLS + j] < 0 || in[off + 3 + numLS + j] >= nSF) return gf_fail("gf_amp4: spin-factor index out of range");
    }
    if(in[0] < 0 || in[0] >= GF_MAX_AUX) return gf_fail("gf_amp4: bad aux slot");
    if(c->h_nodes[node].op == GF_OP_AMP4_TD) {
        int td = in[4];
        if(td < 5 + nLS + nSF + nAmp || td + 1 + nAmp > nd.nint) return gf_fail("gf_amp4: bad time-dependent block offset");
        if(in[td] < 0 || in[td] >= nn || (c->h_nodes[in[td]].op != GF_OP_TRES_TRUTH && c->h_nodes[in[td]].op != GF_OP_TRES_3GAUSS))
            return gf_fail("gf_amp4: not a decay-time resolution record");
        if(c->h_nodes[in[td]].op == GF_OP_TRES_3GAUSS && c->h_nodes[in[td]].npar < 13) return gf_fail("gf_amp4: ThreeGaussResolution needs 13 parameters");
        int ncoef = 0;
        for(int a = 0; a < nAmp; a++) {
            int f = in[td + 1 + a];
            if(f < 0 || f > 2) return gf_fail("gf_amp4: amplitude flag must be 0 (A), 1 (B) or 2 (both)");
            ncoef += f == 2 ? 2 : 1;
        }
        if(nd.npar < 4 + 2 * ncoef) return gf_fail("gf_amp4: time-dependent coefficient block too short");
        if(nd.nobs < 8) return gf_fail("gf_amp4: Amp4Body_TD needs m12, m34, cos12, cos34, phi, event number, dtime, sigma");
    } else if(in[4] != 0)
        return gf_fail("gf_amp4: td offset set on a time-integrated record");
    gf_amp4_state *s = find_amp4(c, node, create);
    if(!s) return gf_fail("gf_amp4: gf_amp4_alloc / gf_amp4_set_mc not called for this node");
    s->slot = in[0]; s->nLS = nLS; s->nSF = nSF; s->nAmp = nAmp; s->td = c->h_nodes[node].op == GF_OP_AMP4_TD ? 1 : 0;
    { // kin planes: one per distinct pair >= 2 (resonant mass) and one more per distinct three-body pair (its d1)
        int slot_m[18], slot_d[18], n = 0;
        for(int p = 0; p < 18; p++) slot_m[p] = slot_d[p] = -1;
        for(int i = 0; i < nLS; i++) {
            int pair = (int)c->h_consts[c->h_nodes[in[5 + i]].con_off];
            s->kin.m[i] = s->kin.d[i] = -1;
            if(pair < 2 || pair > 17) continue;
            if(slot_m[pair] < 0) slot_m[pair] = n++;
            if(pair >= 6 && slot_d[pair] < 0) slot_d[pair] = n++;
            s->kin.m[i] = (signed char)slot_m[pair];
            s->kin.d[i] = (signed char)slot_d[pair];
        }
        s->kin.n = n;
    }
    *out = s;
    return 0;
}

static int alloc_caches(gf_ctx *c, gf_amp4_state *s, gf_amp4_set &S, size_t n) {
    cudaFree(S.ls); cudaFree(S.sf); cudaFree(S.amps);
    S.ls = nullptr; S.sf = nullptr; S.amps = nullptr;
    S.n = (long)n;
    S.ld = (long)((n + 15) & ~size_t(15));
    if(S.ld == 0) S.ld = 16;
    GF_CUDA(cudaMalloc(&S.ls, sizeof(double2) * (size_t)S.ld * std::max(1, s->nLS)));
    GF_CUDA(cudaMalloc(&S.sf, sizeof(double) * (size_t)S.ld * std::max(1, s->nSF)));
    GF_CUDA(cudaMalloc(&S.amps, sizeof(double2) * (size_t)S.ld * s->nAmp));
    cudaFree(S.kin);
    S.kin = nullptr;
    GF_CUDA(cudaMalloc(&S.kin, sizeof(double) * (size_t)S.ld * std::max(1, s->kin.n)));
    GF_CUDA(cudaMemsetAsync(S.ls, 0, sizeof(double2) * (size_t)S.ld * std::max(1, s->nLS), c->stream));
    GF_CUDA(cudaMemsetAsync(S.sf, 0, sizeof(double) * (size_t)S.ld * std::max(1, s->nSF), c->stream));
    GF_CUDA(cudaMemsetAsync(S.amps, 0, sizeof(double2) * (size_t)S.ld * s->nAmp, c->stream));
    return 0;
}

extern "C" int gf_amp4_alloc(gf_ctx *c, int node, size_t n_events) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, true)) return 1;
    if((long)n_events != c->view.n_events) return gf_fail("gf_amp4_alloc: cache size != uploaded events");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    if(alloc_caches(c, s, s->data, n_events)) return 1;
    const gf_node &nd = c->h_nodes[node];
    for(int k = 0; k < 5; k++) s->data.cols[k] = c->view.cols[c->h_ints[nd.obs_off + k]];
    c->view.aux[s->slot] = s->data.amps;
    c->view.aux_ld[s->slot] = s->data.ld;
    return 0;
}

extern "C" int gf_amp4_set_mc(gf_ctx *c, int node, const double *const *mc_cols5, size_t n_mc, const double *weights,
                              const double *dtime, const double *sigma) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, true)) return 1;
    if(n_mc == 0) return gf_fail("gf_amp4_set_mc: no MC events");
    if((dtime == nullptr) != (sigma == nullptr)) return gf_fail("gf_amp4_set_mc: decay times and resolutions come together");
    if(dtime && !weights) return gf_fail("gf_amp4_set_mc: a sample with decay times is a weighted sample");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(s->mc.owned_cols); cudaFree(s->mc.weights); cudaFree(s->mc.dtime); cudaFree(s->mc.sigma);
    s->mc.owned_cols = nullptr; s->mc.weights = nullptr; s->mc.dtime = nullptr; s->mc.sigma = nullptr;
    if(alloc_caches(c, s, s->mc, n_mc)) return 1;
    s->gram_valid = false;
    s->mma_valid = false;
    GF_CUDA(cudaMalloc(&s->mc.owned_cols, sizeof(double) * 5 * (size_t)s->mc.ld));
    for(int k = 0; k < 5; k++) {
        GF_CUDA(cudaMemcpyAsync(s->mc.owned_cols + (size_t)k * s->mc.ld, mc_cols5[k], sizeof(double) * n_mc, cudaMemcpyHostToDevice, c->stream));
        s->mc.cols[k] = s->mc.owned_cols + (size_t)k * s->mc.ld;
    }
    if(weights) {
        GF_CUDA(cudaMalloc(&s->mc.weights, sizeof(double) * (size_t)s->mc.ld)); // padded like the planes: the one-pass update copies 16-byte units
        GF_CUDA(cudaMemsetAsync(s->mc.weights, 0, sizeof(double) * (size_t)s->mc.ld, c->stream));
        GF_CUDA(cudaMemcpyAsync(s->mc.weights, weights, sizeof(double) * n_mc, cudaMemcpyHostToDevice, c->stream));
    }
    if(dtime) {
        GF_CUDA(cudaMalloc(&s->mc.dtime, sizeof(double) * (size_t)s->mc.ld));
        GF_CUDA(cudaMalloc(&s->mc.sigma, sizeof(double) * (size_t)s->mc.ld));
        GF_CUDA(cudaMemcpyAsync(s->mc.dtime, dtime, sizeof(double) * n_mc, cudaMemcpyHostToDevice, c->stream));
        GF_CUDA(cudaMemcpyAsync(s->mc.sigma, sigma, sizeof(double) * n_mc, cudaMemcpyHostToDevice, c->stream));
    }
    GF_CUDA(cudaStreamSynchronize(c->stream)); // host columns may be freed by the caller
    return 0;
}

static unsigned long long mask64(const uint8_t *redo, int n) {
    unsigned long long m = 0;
    for(int i = 0; i < n; i++)
        if(!redo || redo[i]) m |= (1ull << i);
    return m;
}

static int run_set(gf_ctx *c, gf_amp4_state *s, int node, gf_amp4_set &S, int redo_spins, unsigned long long ls_mask,
                   unsigned long long amp_mask) {
    if(S.n == 0 || !S.ls) return 0;
    gf_amp4_args A;
    for(int k = 0; k < 5; k++) A.cols[k] = S.cols[k];
    A.n = S.n; A.ld = S.ld; A.ls = S.ls; A.sf = S.sf; A.amps = S.amps; A.kin = S.kin;
    size_t sm = gf_prog_smem(c);
    int blocks = gf_grid_blocks(c, S.n, 3);
    if(redo_spins) {
        int sf_blocks = gf_grid_blocks(c, S.n, GF_SF_BLOCKS);
        if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_sf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        k_amp4_sf<<<sf_blocks, GF_THREADS, sm, c->stream>>>(c->view, node, A, s->kin);
        c->launches++;
    }
    if(ls_mask) {
        if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_cache, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        k_amp4_cache<<<blocks, GF_THREADS, sm, c->stream>>>(c->view, node, A, ls_mask, s->kin);
        c->launches++;
    }
    if(amp_mask) {
        if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_amps, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        k_amp4_amps<<<gf_grid_blocks(c, S.n, 3), GF_THREADS, sm, c->stream>>>(c->view, node, A, amp_mask);
        c->launches++;
    }
    GF_CUDA(cudaGetLastError());
    return 0;
}

// ls_redo[nLS]: lineshapes whose parameters moved; amp_redo[nAmp]: data amplitudes to rebuild (the caller decides,
// reproducing the reference's flagging, Amp4Body.cu:490-492). MC amplitudes are rebuilt whenever one of their
// lineshapes was flagged, which is what the reference's NormIntegrator sees because it re-reads norm_LS every call.
extern "C" int gf_amp4_cache(gf_ctx *c, int node, int redo_spins, const uint8_t *ls_redo, const uint8_t *amp_redo) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    cudaSetDevice(c->device);
    unsigned long long ls_mask = mask64(ls_redo, s->nLS), amp_mask = mask64(amp_redo, s->nAmp);
    const gf_node &nd = c->h_nodes[node];
    const int32_t *in = c->h_ints.data() + nd.int_off;
    unsigned long long mc_amp_mask = 0;
    for(int a = 0; a < s->nAmp; a++) {
        int off = 5 + s->nLS + s->nSF + s->nAmp + in[5 + s->nLS + s->nSF + a];
        bool need = redo_spins != 0;
        for(int j = 0; j < in[off] && !need; j++) need = (ls_mask >> in[off + 3 + j]) & 1ull;
        if(need) mc_amp_mask |= (1ull << a);
    }
    if(s->data.ls && s->data.n != c->view.n_events) return gf_fail("gf_amp4_cache: cache size != uploaded events");
    if(run_set(c, s, node, s->data, redo_spins, ls_mask, amp_mask)) return 1;
    if(run_set(c, s, node, s->mc, redo_spins, ls_mask, mc_amp_mask)) return 1;
    s->gram_pending |= mc_amp_mask;
    if(mc_amp_mask) s->mma_valid = false;
    return 0;
}

// host view of the Gram matrix from the tiles of k_amp4_update<NB>: tile 2p = Re part, 2p + 1 = Im part of block pair p
static void gram_from_mma_tiles(const double *t, int nb, int n, double *G) {
    auto re = [&](int tile, int i, int j) { return t[(size_t)tile * 64 + i * 8 + j]; };
    for(int a = 0; a < n; a++)
        for(int b = a; b < n; b++) {
            const int ba = a >> 3, bb = b >> 3, i = a & 7, j = b & 7;
            double gr, gi;
            if(ba == bb) { // diagonal block: C1 = Re.Re^T + Im.Im^T, C2 = Im.Re^T, Im G_ab = C2_ab - C2_ba
                const int p = (nb == 1 || ba == 0) ? 0 : 2;
                gr = re(2 * p, i, j);
                gi = re(2 * p + 1, i, j) - re(2 * p + 1, j, i);
            } else { // block pair (0,1): C2 already holds Im.Re^T - Re.Im^T
                gr = re(2, i, j);
                gi = re(3, i, j);
            }
            G[2 * (a * n + b)] = gr;
            G[2 * (a * n + b) + 1] = a == b ? 0.0 : gi;
            if(a != b) {
                G[2 * (b * n + a)] = gr;
                G[2 * (b * n + a) + 1] = -gi;
            }
        }
}
typedef void (*upd_kernel_t)(const gf_dev_view, int, gf_amp4_args, unsigned long long, unsigned long long, const gf_amp4_kinmap, const double *,
                             double *, unsigned *, double *, const gf_peer_view);
static int launch_update(gf_ctx *c, gf_amp4_state *s, int node, gf_amp4_set &S, int nb, unsigned long long ls_mask, unsigned long long amp_mask,
                         double *partials, unsigned *ticket, double *out, const gf_peer_view &PV) {
    gf_amp4_args A;
    for(int k = 0; k < 5; k++) A.cols[k] = S.cols[k];
    A.n = S.n; A.ld = S.ld; A.ls = S.ls; A.sf = S.sf; A.amps = S.amps; A.kin = S.kin;
    upd_kernel_t k = nb == 0 ? k_amp4_update<0> : (nb == 1 ? k_amp4_update<1> : k_amp4_update<2>);
    const size_t sm = gf_upd_smem(gf_prog_smem(c), s->nLS, nb);
    if(sm > 48 * 1024) GF_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    k<<<gf_grid_blocks(c, S.n, upd_blocks_per_sm(nb)), GF_THREADS, sm, c->stream>>>(c->view, node, A, ls_mask, amp_mask, s->kin, S.weights, partials, ticket, out, PV);
    c->launches++;
    GF_CUDA(cudaGetLastError());
    return 0;
}

// gf_amp4_cache followed by gf_amp4_gram, as ONE pass per sample (k_amp4_update): the call the host layer makes in Gram mode
// when a lineshape moved. Falls back to the separate kernels for more than 16 amplitudes or with GOOFIT_B200_NO_FUSED=1.
extern "C" int gf_amp4_cache_gram(gf_ctx *c, int node, int redo_spins, const uint8_t *ls_redo, const uint8_t *amp_redo, double *G_re_im) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    if(!s->mc.amps) return gf_fail("gf_amp4_cache_gram: no MC normalisation sample (gf_amp4_set_mc)");
    if(!G_re_im) return gf_fail("gf_amp4_cache_gram: null output");
    if(!c->fused_on || s->nAmp > 16) {
        if(gf_amp4_cache(c, node, redo_spins, ls_redo, amp_redo)) return 1;
        return gf_amp4_gram(c, node, G_re_im);
    }
    cudaSetDevice(c->device);
    if(s->data.ls && s->data.n != c->view.n_events) return gf_fail("gf_amp4_cache_gram: cache size != uploaded events");
    const unsigned long long ls_mask = mask64(ls_redo, s->nLS), amp_mask = mask64(amp_redo, s->nAmp);
    const gf_node &nd = c->h_nodes[node];
    const int32_t *in = c->h_ints.data() + nd.int_off;
    unsigned long long mc_amp_mask = 0;
    for(int a = 0; a < s->nAmp; a++) {
        int off = 5 + s->nLS + s->nSF + s->nAmp + in[5 + s->nLS + s->nSF + a];
        bool need = redo_spins != 0;
        for(int j = 0; j < in[off] && !need; j++) need = (ls_mask >> in[off + 3 + j]) & 1ull;
        if(need) mc_amp_mask |= (1ull << a);
    }
    const int nb = s->nAmp <= 8 ? 1 : 2, nt = upd_tiles(nb), blocks = gf_grid_blocks(c, s->mc.n, upd_blocks_per_sm(nb));
    const bool mc_work = (ls_mask | mc_amp_mask) != 0 || !s->mma_valid;
    const bool data_work = s->data.n > 0 && s->data.ls && (ls_mask | amp_mask);
    gf_peer_view none{};
    none.nranks = 1;
    if(!s->d_mma || s->mma_blocks < blocks) {
        cudaFree(s->d_mma); cudaFree(s->d_mma_partials); cudaFree(s->d_mma_ticket);
        if(s->h_mma) cudaFreeHost(s->h_mma);
        s->d_mma = s->d_mma_partials = s->h_mma = nullptr;
        s->d_mma_ticket = nullptr;
        GF_CUDA(cudaMalloc(&s->d_mma, sizeof(double) * 6 * 64));
        GF_CUDA(cudaMalloc(&s->d_mma_partials, sizeof(double) * 6 * 64 * (size_t)blocks));
        GF_CUDA(cudaMalloc(&s->d_mma_ticket, sizeof(unsigned)));
        GF_CUDA(cudaMemsetAsync(s->d_mma_ticket, 0, sizeof(unsigned), c->stream));
        GF_CUDA(cudaMallocHost(&s->h_mma, sizeof(double) * 6 * 64));
        s->mma_blocks = blocks;
        s->mma_valid = false;
    }
    gf_amp4_args A;
    size_t sm = gf_prog_smem(c);
    if(redo_spins) {
        if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_sf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        for(gf_amp4_set *S : {&s->mc, &s->data}) {
            if(S->n == 0 || !S->ls) continue;
            for(int k = 0; k < 5; k++) A.cols[k] = S->cols[k];
            A.n = S->n; A.ld = S->ld; A.ls = S->ls; A.sf = S->sf; A.amps = S->amps; A.kin = S->kin;
            k_amp4_sf<<<gf_grid_blocks(c, S->n, GF_SF_BLOCKS), GF_THREADS, sm, c->stream>>>(c->view, node, A, s->kin);
            c->launches++;
        }
        GF_CUDA(cudaGetLastError());
    }
    // MC sample first: its Gram sum is what the host waits for; the data sample's update runs behind it on the stream
    if(mc_work) {
        gf_peer_view PV = gf_next_peer_view(c);
        if(launch_update(c, s, node, s->mc, nb, ls_mask, mc_amp_mask, s->d_mma_partials, s->d_mma_ticket, s->d_mma, PV)) return 1;
        if(c->nccl_comm && !gf_peer_active(c) && gf_allreduce_sum(c, s->d_mma, nt * 64)) return 1;
        GF_CUDA(cudaMemcpyAsync(s->h_mma, s->d_mma, sizeof(double) * nt * 64, cudaMemcpyDeviceToHost, c->stream));
        GF_CUDA(cudaEventRecord(c->ev1, c->stream));
    }
    if(data_work && launch_update(c, s, node, s->data, 0, ls_mask, amp_mask, nullptr, nullptr, nullptr, none)) return 1;
    if(mc_work) {
        GF_CUDA(cudaEventSynchronize(c->ev1)); // the Gram tiles only: the data-sample update keeps running behind the host sum
        if(c->h_err && *c->h_err) return gf_fail("gf_amp4_cache_gram: a rank did not join the peer-memory exchange within the time limit");
        s->mma_valid = true;
    }
    gram_from_mma_tiles(s->h_mma, nb, s->nAmp, G_re_im);
    s->gram_pending |= mc_amp_mask; // the 4x4-tile copy of gf_amp4_gram (if it is ever asked for) is stale for these
    return 0;
}

// the data-sample half of gf_amp4_cache alone: what toy generation needs (Amp4Body_TD::GenerateSig evaluates the PDF with the
// normalisation forced to 1, Amp4Body_TD.cu:764-765,946-950), so no pass over the MC sample is made
extern "C" int gf_amp4_cache_data(gf_ctx *c, int node, int redo_spins, const uint8_t *ls_redo, const uint8_t *amp_redo) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    cudaSetDevice(c->device);
    if(s->data.ls && s->data.n != c->view.n_events) return gf_fail("gf_amp4_cache_data: cache size != uploaded events");
    return run_set(c, s, node, s->data, redo_spins, mask64(ls_redo, s->nLS), mask64(amp_redo, s->nAmp));
}

extern "C" int gf_amp4_norm_async(gf_ctx *c, int node) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    if(!s->mc.amps) return gf_fail("gf_amp4_norm: no MC normalisation sample (gf_amp4_set_mc)");
    cudaSetDevice(c->device);
    size_t sm = gf_prog_smem(c);
    if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_norm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int blocks = gf_grid_blocks(c, s->mc.n, 4);
    if(s->td && s->mc.dtime) { // weighted sample with per-event decay times: ONE sum, the time-dependent PDF itself
        const int32_t *in = c->h_ints.data() + c->h_nodes[node].int_off;
        const bool g3 = c->h_nodes[in[in[4]]].op != GF_OP_TRES_TRUTH;
        auto k = g3 ? k_amp4_norm_tdw<true> : k_amp4_norm_tdw<false>;
        if(sm > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        k<<<gf_grid_blocks(c, s->mc.n, 3), GF_THREADS, sm, c->stream>>>(c->view, node, s->mc.amps, s->mc.ld, s->mc.n, s->mc.weights, s->mc.dtime, s->mc.sigma,
                                                                         c->d_partials + (size_t)c->max_blocks * 2, c->d_ticket + 2, c->d_result + 1, gf_next_peer_view(c));
        c->launches++;
        GF_CUDA(cudaGetLastError());
        if(c->nccl_comm && !gf_peer_active(c) && gf_allreduce_sum(c, c->d_result + 1, 1)) return 1;
        return 0;
    }
    if(s->td) {
        if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_norm_td, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        k_amp4_norm_td<<<blocks, GF_THREADS, sm, c->stream>>>(c->view, node, s->mc.amps, s->mc.ld, s->mc.n, s->mc.weights,
                                                             c->d_partials + (size_t)c->max_blocks * 2, c->d_ticket + 2, c->d_result + 1, gf_next_peer_view(c));
    } else
        k_amp4_norm<<<blocks, GF_THREADS, sm, c->stream>>>(c->view, node, s->mc.amps, s->mc.ld, s->mc.n, s->mc.weights,
                                                          c->d_partials + (size_t)c->max_blocks * 2, c->d_ticket + 2, c->d_result + 1, gf_next_peer_view(c));
    c->launches++;
    GF_CUDA(cudaGetLastError());
    if(c->nccl_comm && !gf_peer_active(c) && gf_allreduce_sum(c, c->d_result + 1, s->td ? 4 : 1)) return 1;
    return 0;
}
extern "C" int gf_amp4_norm(gf_ctx *c, int node, double *out) {
    if(gf_amp4_norm_async(c, node)) return 1;
    gf_amp4_state *s = find_amp4(c, node, false);
    int n = s && s->td && !s->mc.dtime ? 4 : 1;
    GF_CUDA(cudaMemcpyAsync(c->h_result + 1, c->d_result + 1, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    for(int k = 0; k < n; k++) out[k] = c->h_result[1 + k];
    return 0;
}

// Hermitian Gram matrix of the (local shard of the) cached MC amplitudes, all-reduced when a communicator is attached.
// G_re_im[(a * nAmp + b) * 2 + {0,1}] = sum_m w_m A_a(m) conj(A_b(m)). The caller recomputes it after gf_amp4_cache
// touched the MC amplitudes; between such calls the integral needs no device work at all.
extern "C" int gf_amp4_gram(gf_ctx *c, int node, double *G_re_im) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    if(!s->mc.amps) return gf_fail("gf_amp4_gram: no MC normalisation sample (gf_amp4_set_mc)");
    if(!G_re_im) return gf_fail("gf_amp4_gram: null output");
    cudaSetDevice(c->device);
    const int T = GF_GRAM_T, nT = (s->nAmp + T - 1) / T, per = 2 * T * T;
    const int blocks = gf_grid_blocks(c, s->mc.n, 2);
    if(!s->d_gram || s->gram_blocks < blocks) {
        cudaFree(s->d_gram); cudaFree(s->d_gram_partials); cudaFree(s->d_gram_ticket);
        if(s->h_gram) cudaFreeHost(s->h_gram);
        s->d_gram = s->d_gram_partials = s->h_gram = nullptr;
        s->d_gram_ticket = nullptr;
        const int maxTiles = (64 / T) * (64 / T + 1) / 2;
        GF_CUDA(cudaMalloc(&s->d_gram, sizeof(double) * per * maxTiles));
        GF_CUDA(cudaMalloc(&s->d_gram_partials, sizeof(double) * per * (size_t)blocks));
        GF_CUDA(cudaMalloc(&s->d_gram_ticket, sizeof(unsigned)));
        GF_CUDA(cudaMemsetAsync(s->d_gram_ticket, 0, sizeof(unsigned), c->stream));
        GF_CUDA(cudaMallocHost(&s->h_gram, sizeof(double) * per * maxTiles));
        s->gram_blocks = blocks;
        s->gram_valid = false;
    }
    // only the tiles with a rebuilt amplitude among their rows or columns are re-summed; the others keep their
    // (already all-reduced) values
    int t = 0, launched = 0;
    for(int ta = 0; ta < nT; ta++)
        for(int tb = ta; tb < nT; tb++, t++) {
            unsigned long long rows = ((1ull << T) - 1) << (ta * T), cols = ((1ull << T) - 1) << (tb * T);
            if(s->gram_valid && !(s->gram_pending & (rows | cols))) continue;
            double *out = s->d_gram + (size_t)t * per;
            k_amp4_gram<<<blocks, GF_THREADS, 0, c->stream>>>(s->mc.amps, s->mc.ld, s->mc.n, s->mc.weights, s->nAmp, ta * T, tb * T,
                                                             s->d_gram_partials, s->d_gram_ticket, out, gf_next_peer_view(c));
            c->launches++;
            launched++;
            GF_CUDA(cudaGetLastError());
            if(c->nccl_comm && !gf_peer_active(c) && gf_allreduce_sum(c, out, per)) return 1;
            GF_CUDA(cudaMemcpyAsync(s->h_gram + (size_t)t * per, out, sizeof(double) * per, cudaMemcpyDeviceToHost, c->stream));
        }
    if(launched) GF_CUDA(cudaStreamSynchronize(c->stream));
    s->gram_valid = true;
    s->gram_pending = 0;
    t = 0;
    const int n = s->nAmp;
    for(int ta = 0; ta < nT; ta++)
        for(int tb = ta; tb < nT; tb++, t++)
            for(int i = 0; i < T; i++)
                for(int j = 0; j < T; j++) {
                    int a = ta * T + i, b = tb * T + j;
                    if(a >= n || b >= n) continue;
                    double re = s->h_gram[(size_t)t * per + 2 * (i * T + j)], im = s->h_gram[(size_t)t * per + 2 * (i * T + j) + 1];
                    if(ta == tb && b < a) continue; // lower triangle of a diagonal tile: taken from its mirror
                    G_re_im[2 * (a * n + b)] = re;
                    G_re_im[2 * (a * n + b) + 1] = a == b ? 0.0 : im;
                    if(a != b) {
                        G_re_im[2 * (b * n + a)] = re;
                        G_re_im[2 * (b * n + a) + 1] = -im;
                    }
                }
    return 0;
}

extern "C" int gf_amp4_get_amp(gf_ctx *c, int node, int amp, double *out_re_im, size_t n_events) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    if(amp < 0 || amp >= s->nAmp) return gf_fail("gf_amp4_get_amp: bad amplitude index");
    if(!s->data.amps || (long)n_events != s->data.n) return gf_fail("gf_amp4_get_amp: size mismatch");
    cudaSetDevice(c->device);
    GF_CUDA(cudaMemcpyAsync(out_re_im, s->data.amps + (size_t)amp * s->data.ld, sizeof(double2) * n_events, cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
