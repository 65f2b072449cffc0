// gf_amp3.cu — Amp3Body caches: per-event resonance waves and the normalisation-grid integrals.
//   k_amp3_waves  all flagged resonances in ONE pass over the events, SoA out   (Amp3Body.cu:288-310,
//                 detail/SpecialResonanceCalculator.cu:12-46 launched once per resonance in the reference)
//   k_amp3_grid   every flagged B_i and the efficiency at each grid bin, once   (detail/SpecialResonanceIntegrator.cu:52-121
//   k_amp3_gram   I_ij = sum_bins B_i conj(B_j) eff, one block per pair          evaluates both amplitudes per pair per bin)
#include "gf_internal.cuh"
#include "gf_res.cuh"

struct gf_amp3_state {
    int node = -1, n_res = 0, slot = -1;
    int td = 0; // GF_OP_AMP3_TD: two cached planes per resonance (A_i at (m12, m13), B_i at (m13, m12)), six sums per ordered pair
    long ld = 0, n_events = 0;
    double2 *d_waves = nullptr;
    // normalisation grid cache
    long n_bins = 0;
    double spec[6] = {0, 0, 0, 0, 0, 0};
    double2 *d_gridB = nullptr; // [n_res][n_bins]
    double *d_gridEff = nullptr;
    double *d_I = nullptr;      // [n_res*n_res*2]
    int *d_pairs = nullptr;     // [n_res*(n_res+1)] pair list of the Gram kernel, uploaded through h_pairs (pinned)
    int *h_pairs = nullptr;
    bool grid_valid = false;
};

static gf_amp3_state *find_amp3(gf_ctx *c, int node, bool create) {
    for(auto *s : c->amp3)
        if(s->node == node) return s;
    if(!create) return nullptr;
    auto *s = new gf_amp3_state();
    s->node = node;
    c->amp3.push_back(s);
    return s;
}

void gf_amp3_free(gf_ctx *c) {
    for(auto *s : c->amp3) {
        cudaFree(s->d_waves); cudaFree(s->d_gridB); cudaFree(s->d_gridEff); cudaFree(s->d_I); cudaFree(s->d_pairs);
        if(s->h_pairs) cudaFreeHost(s->h_pairs);
        delete s;
    }
    c->amp3.clear();
}

__device__ inline gf_decay3 gf_load_decay3(const double *c) {
    gf_decay3 D;
    D.M = c[0]; D.d1 = c[1]; D.d2 = c[2]; D.d3 = c[3]; D.rad = c[4]; D.mrad = c[5];
    D.inv_M = 1.0 / D.M;
    return D;
}

#ifndef GF_WAVES_BLOCKS
#define GF_WAVES_BLOCKS 3 // measured (round 2, after the spills were removed): 3 blocks/SM at 80 registers 1.92 ms, 2 at 128 registers 2.07 ms
#endif
// TD: the time-dependent amplitude caches every resonance twice, at the point and at the point with m12 and m13 exchanged
// (the D0bar amplitude; SpecialWaveCalculator.cu:15-68); observables are dtime, sigma_t, m12, m13, event number, ...
template <bool TD>
__global__ void __launch_bounds__(GF_THREADS, GF_WAVES_BLOCKS)
    k_amp3_waves(const gf_dev_view V, int node, unsigned long long mask, double2 *waves, long ld) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const int32_t *in = P.ints + nd.int_off;
    const int32_t *ob = P.ints + nd.obs_off;
    const int nres = in[1];
    const gf_decay3 D = gf_load_decay3(P.consts + nd.con_off);
    const double *c12 = P.cols[ob[TD ? 2 : 0]], *c13 = P.cols[ob[TD ? 3 : 1]];
    __shared__ gf_amp3_shared S;
    gf_amp3_prepare(S, P, in + 2, nres, mask, D);
    __syncthreads();
    const long stride = (long)gridDim.x * blockDim.x;
    for(long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < V.n_events; e += stride) {
        double m12 = __ldg(c12 + e), m13 = __ldg(c13 + e);
        bool inside = gf_in_dalitz(m12, m13, D.M, D.d1, D.d2, D.d3);
        double m23 = D.M * D.M + D.d1 * D.d1 + D.d2 * D.d2 + D.d3 * D.d3 - m12 - m13;
        // written once, next read by another kernel: streaming stores
        if(!TD)
            gf_amp3_point(S, P, in + 2, D, inside, m12, m13, m23, [&](int i, double2 w) { __stcs(waves + (size_t)i * ld + e, w); });
        else {
            gf_amp3_point(S, P, in + 2, D, inside, m12, m13, m23, [&](int i, double2 w) { __stcs(waves + (size_t)(2 * i) * ld + e, w); });
            gf_amp3_point(S, P, in + 2, D, inside, m13, m12, m23, [&](int i, double2 w) { __stcs(waves + (size_t)(2 * i + 1) * ld + e, w); });
        }
    }
}

// Bin centres as SpecialResonanceIntegrator.cu:60-77, including its use of the m12 bin count for the
// m13 step (slot [2] read twice) — identical for square grids.
// TD (SpecialDalitzIntegrator.cu:52-86): the m13 step uses the m13 bin count, and both A_i and B_i are kept per bin
template <bool TD>
__global__ void __launch_bounds__(GF_THREADS, 1) // <= 57 600 bins: latency-bound, registers are free
    k_amp3_grid(const gf_dev_view V, int node, int eff_first, int eff_root, unsigned long long mask, double lo12,
                double hi12, int n12, double lo13, double hi13, int n13, long n_bins, double2 *gridB, double *gridEff) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &nd = P.nodes[node];
    const int32_t *in = P.ints + nd.int_off;
    const int32_t *ob = P.ints + nd.obs_off;
    const int nres = in[1];
    const gf_decay3 D = gf_load_decay3(P.consts + nd.con_off);
    __shared__ gf_amp3_shared S;
    gf_amp3_prepare(S, P, in + 2, nres, mask, D);
    __syncthreads();
    const long stride = (long)gridDim.x * blockDim.x;
    for(long b = (long)blockIdx.x * blockDim.x + threadIdx.x; b < n_bins; b += stride) {
        int b12 = (int)(b % n12);
        long b13 = b / n12;
        double m12 = hi12 - lo12;
        m12 /= n12;
        m12 *= (b12 + 0.5);
        m12 += lo12;
        double m13 = hi13 - lo13;
        m13 /= TD ? n13 : n12;
        m13 *= (b13 + 0.5);
        m13 += lo13;
        bool inside = gf_in_dalitz(m12, m13, D.M, D.d1, D.d2, D.d3);
        double m23 = D.M * D.M + D.d1 * D.d1 + D.d2 * D.d2 + D.d3 * D.d3 - m12 - m13;
        if(!TD)
            gf_amp3_point(S, P, in + 2, D, inside, m12, m13, m23, [&](int i, double2 w) { gridB[(size_t)i * n_bins + b] = w; });
        else {
            gf_amp3_point(S, P, in + 2, D, inside, m12, m13, m23, [&](int i, double2 w) { gridB[(size_t)(2 * i) * n_bins + b] = w; });
            gf_amp3_point(S, P, in + 2, D, inside, m13, m12, m23, [&](int i, double2 w) { gridB[(size_t)(2 * i + 1) * n_bins + b] = w; });
        }
        gf_evt_vals evt;
#pragma unroll
        for(int k = 0; k < GF_MAX_OBS; k++) evt.v[k] = 0;
        evt.i = b;
        evt.v[ob[TD ? 2 : 0]] = m12;
        evt.v[ob[TD ? 3 : 1]] = m13;
        gridEff[b] = gf_eval_tree(P, V, eff_first, eff_root, evt);
    }
}

// One block per (i,j), i<=j: sum over bins of B_i * conj(B_j) * eff in a fixed order.
__global__ void __launch_bounds__(GF_THREADS)
    k_amp3_gram(const double2 *gridB, const double *gridEff, long n_bins, int nres, const int *pairs, double *I) {
    __shared__ double red[32];
    const int i = pairs[2 * blockIdx.x], j = pairs[2 * blockIdx.x + 1];
    const double2 *Bi = gridB + (size_t)i * n_bins, *Bj = gridB + (size_t)j * n_bins;
    double sr = 0, si = 0;
    for(long b = threadIdx.x; b < n_bins; b += blockDim.x) {
        double2 a = Bi[b], c = Bj[b];
        double e = gridEff[b];
        double pr = a.x * c.x + a.y * c.y; // a * conj(c)
        double pi = a.y * c.x - a.x * c.y;
        sr += pr * e;
        si += pi * e;
    }
    double r = gf_block_sum(sr, red);
    double q = gf_block_sum(si, red);
    if(threadIdx.x == 0) {
        I[2 * (i * nres + j)] = r;
        I[2 * (i * nres + j) + 1] = q;
        if(i != j) { // Hermitian: I_ji = conj(I_ij)
            I[2 * (j * nres + i)] = r;
            I[2 * (j * nres + i) + 1] = -q;
        }
    }
}

// W_ij = sum_e w_i(e) conj(w_j(e)) over the cached waves, for the fit fractions (Amp3Body.cu:376-428: the reference runs one
// transform_reduce of CoefSumFunctor per ordered pair over the whole sample). One block per (pair i <= j, chunk of events);
// the chunk partials are added on the host in chunk order (deterministic), W_ji = conj(W_ij).
#define GF_WG_CHUNKS 8
__global__ void __launch_bounds__(GF_THREADS)
    k_amp3_wave_gram(const double2 *waves, long ld, long n, int nres, double *partials /* [pair][chunk][2] */) {
    __shared__ double red[32];
    int p = blockIdx.x, i = 0;
    while(p >= nres - i) { p -= nres - i; ++i; } // pair index -> (i, j = i + p)
    const int j = i + p;
    const long per = (n + GF_WG_CHUNKS - 1) / GF_WG_CHUNKS, lo = per * blockIdx.y, hi = lo + per < n ? lo + per : n;
    const double2 *A = waves + (size_t)i * ld, *B = waves + (size_t)j * ld;
    double sr = 0, si = 0;
    for(long e = lo + threadIdx.x; e < hi; e += blockDim.x) {
        const double2 a = __ldg(A + e), b = __ldg(B + e);
        sr += a.x * b.x + a.y * b.y; // a conj(b)
        si += a.y * b.x - a.x * b.y;
    }
    const double r = gf_block_sum(sr, red), q = gf_block_sum(si, red);
    if(threadIdx.x == 0) {
        partials[2 * ((size_t)blockIdx.x * GF_WG_CHUNKS + blockIdx.y)] = r;
        partials[2 * ((size_t)blockIdx.x * GF_WG_CHUNKS + blockIdx.y) + 1] = q;
    }
}

// Amp3Body_TD: one block per ORDERED pair (i, j): sums over bins of eff x {A_i A_j*, A_i B_j*, B_i B_j*} in a fixed order
// (SpecialDalitzIntegrator.cu:15-51; A_i B_j* is not Hermitian in (i, j), so all n^2 pairs are summed as upstream does)
__global__ void __launch_bounds__(GF_THREADS)
    k_amp3td_gram(const double2 *grid, const double *gridEff, long n_bins, int nres, const int *pairs, double *I) {
    __shared__ double red[32];
    const int i = pairs[2 * blockIdx.x], j = pairs[2 * blockIdx.x + 1];
    const double2 *Ai = grid + (size_t)(2 * i) * n_bins, *Bi = grid + (size_t)(2 * i + 1) * n_bins;
    const double2 *Aj = grid + (size_t)(2 * j) * n_bins, *Bj = grid + (size_t)(2 * j + 1) * n_bins;
    double s[6] = {0, 0, 0, 0, 0, 0};
    for(long b = threadIdx.x; b < n_bins; b += blockDim.x) {
        const double2 ai = Ai[b], bi = Bi[b], aj = Aj[b], bj = Bj[b];
        const double e = gridEff[b];
        s[0] += (ai.x * aj.x + ai.y * aj.y) * e; // ai conj(aj)
        s[1] += (ai.y * aj.x - ai.x * aj.y) * e;
        s[2] += (ai.x * bj.x + ai.y * bj.y) * e; // ai conj(bj)
        s[3] += (ai.y * bj.x - ai.x * bj.y) * e;
        s[4] += (bi.x * bj.x + bi.y * bj.y) * e; // bi conj(bj)
        s[5] += (bi.y * bj.x - bi.x * bj.y) * e;
    }
#pragma unroll
    for(int k = 0; k < 6; ++k) {
        const double r = gf_block_sum(s[k], red);
        if(threadIdx.x == 0) I[6 * (i * nres + j) + k] = r;
    }
}

static int amp3_check(gf_ctx *c, int node, gf_amp3_state **out, bool create) {
    if(!c) return gf_fail("null ctx");
    if(node < 0 || node >= (int)c->h_nodes.size() || (c->h_nodes[node].op != GF_OP_AMP3 && c->h_nodes[node].op != GF_OP_AMP3_TD))
        return gf_fail("gf_amp3: node is not an Amp3Body record");
    const gf_node &nd = c->h_nodes[node];
    const int td = nd.op == GF_OP_AMP3_TD ? 1 : 0;
    if(nd.nint < 2 || nd.ncon < 6 || nd.nobs < (td ? 5 : 3)) return gf_fail("gf_amp3: malformed Amp3Body record");
    int nres = c->h_ints[nd.int_off + 1];
    if(nres < 1 || nres > 64 || nd.nint < 2 + nres + td) return gf_fail("gf_amp3: 1..64 resonances");
    if(td) {
        if(nd.npar < 5 + 2 * nres) return gf_fail("gf_amp3: Amp3Body_TD needs tau, x, y, deltax, deltay and one coefficient pair per resonance");
        const int rn = c->h_ints[nd.int_off + 2 + nres];
        if(rn < 0 || rn >= (int)c->h_nodes.size() || (c->h_nodes[rn].op != GF_OP_TRES_TRUTH && c->h_nodes[rn].op != GF_OP_TRES_3GAUSS))
            return gf_fail("gf_amp3: not a decay-time resolution record");
    }
    int slot = c->h_ints[nd.int_off];
    if(slot < 0 || slot >= GF_MAX_AUX) return gf_fail("gf_amp3: bad aux slot");
    gf_amp3_state *s = find_amp3(c, node, create);
    if(!s) return gf_fail("gf_amp3: gf_amp3_alloc not called for this node");
    s->n_res = nres;
    s->slot = slot;
    s->td = td;
    *out = s;
    return 0;
}

extern "C" int gf_amp3_alloc(gf_ctx *c, int node, size_t n_events) {
    gf_amp3_state *s;
    if(amp3_check(c, node, &s, true)) return 1;
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(s->d_waves);
    s->d_waves = nullptr;
    s->n_events = (long)n_events;
    s->ld = (long)((n_events + 15) & ~size_t(15));
    if(s->ld == 0) s->ld = 16;
    const size_t planes = (size_t)s->n_res * (1 + s->td);
    GF_CUDA(cudaMalloc(&s->d_waves, sizeof(double2) * (size_t)s->ld * planes));
    GF_CUDA(cudaMemsetAsync(s->d_waves, 0, sizeof(double2) * (size_t)s->ld * planes, c->stream));
    if(!s->d_I) {
        GF_CUDA(cudaMalloc(&s->d_I, sizeof(double) * 6 * s->n_res * s->n_res));
        GF_CUDA(cudaMemsetAsync(s->d_I, 0, sizeof(double) * 6 * s->n_res * s->n_res, c->stream));
    }
    if(!s->d_pairs) {
        GF_CUDA(cudaMalloc(&s->d_pairs, sizeof(int) * 2 * s->n_res * s->n_res));
        GF_CUDA(cudaMallocHost(&s->h_pairs, sizeof(int) * 2 * s->n_res * s->n_res));
    }
    c->view.aux[s->slot] = s->d_waves;
    c->view.aux_ld[s->slot] = s->ld;
    return 0;
}

static unsigned long long mask_of(const uint8_t *redo, int n) {
    unsigned long long m = 0;
    for(int i = 0; i < n; i++)
        if(!redo || redo[i]) m |= (1ull << i);
    return m;
}

extern "C" int gf_amp3_cache_waves(gf_ctx *c, int node, const uint8_t *redo_mask) {
    gf_amp3_state *s;
    if(amp3_check(c, node, &s, false)) return 1;
    if(s->n_events != c->view.n_events) return gf_fail("gf_amp3_cache_waves: cache size != uploaded events");
    unsigned long long mask = mask_of(redo_mask, s->n_res);
    if(!mask || s->n_events == 0) return 0;
    cudaSetDevice(c->device);
    size_t sm = gf_prog_smem(c);
    auto kern = s->td ? k_amp3_waves<true> : k_amp3_waves<false>;
    if(sm > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int blocks = gf_grid_blocks(c, s->n_events, GF_WAVES_BLOCKS);
    kern<<<blocks, GF_THREADS, sm, c->stream>>>(c->view, node, mask, s->d_waves, s->ld);
    c->launches++;
    GF_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int gf_amp3_integrals(gf_ctx *c, int node, const uint8_t *redo_mask, const double *lo_hi_n6, int eff_root,
                                 double *I_re_im) {
    gf_amp3_state *s;
    if(amp3_check(c, node, &s, false)) return 1;
    if(eff_root < 0 || eff_root >= (int)c->h_nodes.size()) return gf_fail("gf_amp3_integrals: bad efficiency root");
    cudaSetDevice(c->device);
    int n12 = (int)std::floor(lo_hi_n6[2] + 0.5), n13 = (int)std::floor(lo_hi_n6[5] + 0.5);
    if(n12 < 1 || n13 < 1) return gf_fail("gf_amp3_integrals: bins < 1");
    long n_bins = (long)n12 * n13;
    bool same = s->grid_valid && s->n_bins == n_bins;
    for(int k = 0; k < 6 && same; k++) same = (s->spec[k] == lo_hi_n6[k]);
    unsigned long long mask = mask_of(redo_mask, s->n_res);
    if(!same) {
        GF_CUDA(cudaStreamSynchronize(c->stream));
        cudaFree(s->d_gridB); cudaFree(s->d_gridEff);
        s->d_gridB = nullptr; s->d_gridEff = nullptr;
        GF_CUDA(cudaMalloc(&s->d_gridB, sizeof(double2) * (size_t)n_bins * s->n_res * (1 + s->td)));
        GF_CUDA(cudaMalloc(&s->d_gridEff, sizeof(double) * (size_t)n_bins));
        s->n_bins = n_bins;
        for(int k = 0; k < 6; k++) s->spec[k] = lo_hi_n6[k];
        mask = mask_of(nullptr, s->n_res); // new grid: everything
        s->grid_valid = true;
    }
    size_t sm = gf_prog_smem(c);
    auto gridk = s->td ? k_amp3_grid<true> : k_amp3_grid<false>;
    if(sm > 48 * 1024) cudaFuncSetAttribute(gridk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int eff_first = eff_root - c->h_nodes[eff_root].size + 1;
    // As in the reference, integrals (and the efficiency they fold in) are refreshed only for flagged
    // resonances; with nothing flagged the cached I_ij are returned (Amp3Body.cu:266-275,313-316).
    if(mask) {
        int blocks = gf_grid_blocks(c, n_bins, 2);
        gridk<<<blocks, GF_THREADS, sm, c->stream>>>(c->view, node, eff_first, eff_root, mask, lo_hi_n6[0], lo_hi_n6[1], n12, lo_hi_n6[3],
                                                     lo_hi_n6[4], n13, n_bins, s->d_gridB, s->d_gridEff);
        c->launches++;
        GF_CUDA(cudaGetLastError());
    }
    // pairs whose integral must be refreshed: either member flagged (Amp3Body.cu:313-316)
    std::vector<int> pairs;
    for(int i = 0; i < s->n_res; i++)
        for(int j = s->td ? 0 : i; j < s->n_res; j++)
            if(((mask >> i) & 1ull) || ((mask >> j) & 1ull)) { pairs.push_back(i); pairs.push_back(j); }
    if(!pairs.empty()) {
        // the pinned list is free again: every earlier call ended with the stream synchronisation below
        std::memcpy(s->h_pairs, pairs.data(), sizeof(int) * pairs.size());
        GF_CUDA(cudaMemcpyAsync(s->d_pairs, s->h_pairs, sizeof(int) * pairs.size(), cudaMemcpyHostToDevice, c->stream));
        if(s->td)
            k_amp3td_gram<<<(int)(pairs.size() / 2), GF_THREADS, 0, c->stream>>>(s->d_gridB, s->d_gridEff, n_bins, s->n_res, s->d_pairs, s->d_I);
        else
            k_amp3_gram<<<(int)(pairs.size() / 2), GF_THREADS, 0, c->stream>>>(s->d_gridB, s->d_gridEff, n_bins, s->n_res,
                                                                             s->d_pairs, s->d_I);
        c->launches++;
        GF_CUDA(cudaGetLastError());
    }
    GF_CUDA(cudaMemcpyAsync(I_re_im, s->d_I, sizeof(double) * (s->td ? 6 : 2) * s->n_res * s->n_res, cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int gf_amp3_get_wave(gf_ctx *c, int node, int res, double *out_re_im, size_t n_events) {
    gf_amp3_state *s;
    if(amp3_check(c, node, &s, false)) return 1;
    if(res < 0 || res >= s->n_res * (1 + s->td)) return gf_fail("gf_amp3_get_wave: bad resonance index");
    if((long)n_events != s->n_events) return gf_fail("gf_amp3_get_wave: size mismatch");
    cudaSetDevice(c->device);
    GF_CUDA(cudaMemcpyAsync(out_re_im, s->d_waves + (size_t)res * s->ld, sizeof(double2) * n_events,
                            cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

// Hermitian matrix of the cached waves, W[(i * n_res + j) * 2 + {re, im}] = sum_e w_i(e) conj(w_j(e)) over the local events
// (summed over ranks when a communicator is attached): what Amp3Body::fit_fractions needs, without moving the waves off the device.
extern "C" int gf_amp3_wave_gram(gf_ctx *c, int node, double *W_re_im) {
    gf_amp3_state *s;
    if(amp3_check(c, node, &s, false)) return 1;
    if(!s->d_waves) return gf_fail("gf_amp3_wave_gram: no cached waves (gf_amp3_cache_waves)");
    if(s->td) return gf_fail("gf_amp3_wave_gram: fit fractions are defined for the time-integrated Amp3Body");
    if(!W_re_im) return gf_fail("gf_amp3_wave_gram: null output");
    cudaSetDevice(c->device);
    const int n = s->n_res, npairs = n * (n + 1) / 2, nv = 2 * npairs * GF_WG_CHUNKS;
    double *d_part = nullptr;
    GF_CUDA(cudaMalloc(&d_part, sizeof(double) * nv));
    k_amp3_wave_gram<<<dim3(npairs, GF_WG_CHUNKS), GF_THREADS, 0, c->stream>>>(s->d_waves, s->ld, s->n_events, n, d_part);
    c->launches++;
    cudaError_t e = cudaGetLastError();
    std::vector<double> part(nv);
    if(e == cudaSuccess) e = cudaMemcpyAsync(part.data(), d_part, sizeof(double) * nv, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    std::vector<double> tri(2 * npairs, 0.0);
    for(int p = 0; p < npairs; p++)
        for(int k = 0; k < GF_WG_CHUNKS; k++) {
            tri[2 * p] += part[2 * (p * GF_WG_CHUNKS + k)];
            tri[2 * p + 1] += part[2 * (p * GF_WG_CHUNKS + k) + 1];
        }
    if(e == cudaSuccess && c->nccl_comm) { // event-sharded: the matrix is a sum over ranks
        e = cudaMemcpyAsync(d_part, tri.data(), sizeof(double) * 2 * npairs, cudaMemcpyHostToDevice, c->stream);
        if(e == cudaSuccess && gf_allreduce_sum(c, d_part, 2 * npairs)) { cudaFree(d_part); return 1; }
        if(e == cudaSuccess) e = cudaMemcpyAsync(tri.data(), d_part, sizeof(double) * 2 * npairs, cudaMemcpyDeviceToHost, c->stream);
        if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    }
    cudaFree(d_part);
    if(e != cudaSuccess) return gf_fail(std::string("gf_amp3_wave_gram: ") + cudaGetErrorString(e));
    int p = 0;
    for(int i = 0; i < n; i++)
        for(int j = i; j < n; j++, p++) {
            W_re_im[2 * (i * n + j)] = tri[2 * p];
            W_re_im[2 * (i * n + j) + 1] = i == j ? 0.0 : tri[2 * p + 1];
            if(i != j) {
                W_re_im[2 * (j * n + i)] = tri[2 * p];
                W_re_im[2 * (j * n + i) + 1] = -tri[2 * p + 1];
            }
        }
    return 0;
}
