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
    bool gram_valid = false;               // d_gram / h_gram hold every tile of the current MC amplitudes
    unsigned long long gram_pending = 0;   // MC amplitudes rebuilt since the last gf_amp4_gram
};

static void free_set(gf_amp4_set &s) {
    cudaFree(s.owned_cols); cudaFree(s.weights); cudaFree(s.ls); cudaFree(s.sf); cudaFree(s.amps); cudaFree(s.kin);
    s = gf_amp4_set();
}
void gf_amp4_free(gf_ctx *c) {
    for(auto *s : c->amp4) {
        free_set(s->data);
        free_set(s->mc);
        cudaFree(s->d_gram); cudaFree(s->d_gram_partials); cudaFree(s->d_gram_ticket);
        if(s->h_gram) cudaFreeHost(s->h_gram);
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
__global__ void __launch_bounds__(GF_THREADS, 1) k_amp4_sf(const gf_dev_view V, int node, gf_amp4_args A, const gf_amp4_kinmap K) {
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
template <int LS_STEP, int SF_STEP, int NPERM>
__device__ __forceinline__ double2 gf_amp4_one(const int32_t *rec, const gf_amp4_args &A, long e) {
    constexpr int NLS = LS_STEP * NPERM, NSF = SF_STEP * NPERM;
    double2 L[NLS];
    double S[NSF];
#pragma unroll
    for(int j = 0; j < NLS; ++j) L[j] = __ldg(A.ls + (size_t)rec[3 + j] * A.ld + e);
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
            const int32_t *rec = rec0 + start[a];
            const int numLS = rec[0], numSF = rec[1], nPerm = rec[2];
            const int LS_step = numLS / nPerm, SF_step = numSF / nPerm;
            const int shape = nPerm * 100 + LS_step * 10 + SF_step; // uniform over the grid
            double2 out;
            switch(shape) {
            case 221: out = gf_amp4_one<2, 1, 2>(rec, A, e); break; // Bose-symmetrised two-resonance chains (DP4, TDDP4)
            case 121: out = gf_amp4_one<2, 1, 1>(rec, A, e); break;
            case 111: out = gf_amp4_one<1, 1, 1>(rec, A, e); break;
            case 211: out = gf_amp4_one<1, 1, 2>(rec, A, e); break;
            case 131: out = gf_amp4_one<3, 1, 1>(rec, A, e); break;
            case 231: out = gf_amp4_one<3, 1, 2>(rec, A, e); break;
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
                out = make_double2(total.re, total.im);
            }
            }
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
    (void)dtime; (void)sigma;
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(s->mc.owned_cols); cudaFree(s->mc.weights);
    s->mc.owned_cols = nullptr; s->mc.weights = nullptr;
    if(alloc_caches(c, s, s->mc, n_mc)) return 1;
    s->gram_valid = false;
    GF_CUDA(cudaMalloc(&s->mc.owned_cols, sizeof(double) * 5 * (size_t)s->mc.ld));
    for(int k = 0; k < 5; k++) {
        GF_CUDA(cudaMemcpyAsync(s->mc.owned_cols + (size_t)k * s->mc.ld, mc_cols5[k], sizeof(double) * n_mc, cudaMemcpyHostToDevice, c->stream));
        s->mc.cols[k] = s->mc.owned_cols + (size_t)k * s->mc.ld;
    }
    if(weights) {
        GF_CUDA(cudaMalloc(&s->mc.weights, sizeof(double) * n_mc));
        GF_CUDA(cudaMemcpyAsync(s->mc.weights, weights, sizeof(double) * n_mc, cudaMemcpyHostToDevice, c->stream));
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
        int sf_blocks = gf_grid_blocks(c, S.n, 1);
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
    return 0;
}

extern "C" int gf_amp4_norm_async(gf_ctx *c, int node) {
    gf_amp4_state *s;
    if(amp4_check(c, node, &s, false)) return 1;
    if(!s->mc.amps) return gf_fail("gf_amp4_norm: no MC normalisation sample (gf_amp4_set_mc)");
    cudaSetDevice(c->device);
    size_t sm = gf_prog_smem(c);
    if(sm > 48 * 1024) cudaFuncSetAttribute(k_amp4_norm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int blocks = gf_grid_blocks(c, s->mc.n, 4);
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
    int n = s && s->td ? 4 : 1;
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
    const int T = GF_GRAM_T, nT = (s->nAmp + T - 1) / T, nTiles = nT * (nT + 1) / 2, per = 2 * T * T;
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
