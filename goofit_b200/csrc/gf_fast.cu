// gf_fast.cu — shape-specialised copies of the fused likelihood kernel: the same evaluate -> metric -> deterministic grid
// reduction as k_eval_sum (gf_core.cu; reference GooPdf.cu:45-78 + MetricTaker.cu:17-44), but for program shapes that are
// known on the host the per-event interpreter is gone: no node-record decode, no value stack, event-independent subtrees
// (constant efficiency, flat background, AddPdf weights, child normalisations) are evaluated once per thread.
//   shape L   root is ONE exp-type leaf and the metric is the NLL  (examples/exponential: ExpPdf.cu:7-67, GaussianPdf.cu:7-17)
//             -log(exp(a) * norm) is computed as -(a + log norm): no exp, no log per event, two events per 16-byte load
//   shape A   [const eff] Amp4Body | Amp4Body_TD, optionally inside AddPdf(w; amp, const background)
//             (examples/DP4/DP4.cu:214-232, examples/TDDP4/TDDP4.cu:283-300; Amp4Body.cu:53-90, Amp4Body_TD.cu:167-261, AddPdf.cu:16-51)
// Anything else stays on the interpreter. The arithmetic per event is the interpreter's, operation for operation, except
// shape L's log-domain form (pinned at 1e-12 against the reference in tests/test_scale_gpu.py and tests/test_parity_gpu.py).
#include "gf_internal.cuh"

// ---- shape L ---------------------------------------------------------------------------------------------------------
// exponent of an exp-type leaf: leaf value == exp(gf_leaf_arg)
__device__ __forceinline__ double gf_leaf_arg(int op, int npar, const double *p, double x) {
    switch(op) {
    case GF_OP_EXP: return p[0] * x;
    case GF_OP_EXP_OFFSET: x -= p[0]; return p[1] * x;
    case GF_OP_EXP_POLY: {
        double a = 0;
        for(int i = 0; i < npar; ++i) a += gf_pow_i(x, i) * p[i];
        return a;
    }
    case GF_OP_EXP_POLY_OFFSET: {
        x -= p[0];
        double a = 0;
        for(int i = 1; i < npar; ++i) a += gf_pow_i(x, i) * p[i];
        return a;
    }
    case GF_OP_GAUSSIAN: return -0.5 * (x - p[0]) * (x - p[0]) / (p[1] * p[1]);
    default: { // GF_OP_SCALED_GAUSSIAN
        double mean = p[0] + p[2], sigma = p[1] * (1 + p[3]);
        return -0.5 * (x - mean) * (x - mean) / (sigma * sigma);
    }
    }
}
// one event's NLL term. Inside |a| < 600 (and |log norm| < 100, checked by the caller) exp(a) * norm is a positive normal
// number, so the reference's -log(exp(a) * norm) is -(a + log norm) up to rounding; outside, the reference's own form
// (underflow to 0 -> term 0, overflow -> -inf) is evaluated as written.
__device__ __forceinline__ double gf_logleaf_term(double a, double norm, double lognorm, bool shortcut) {
    if(shortcut && fabs(a) < 600.0) return -(a + lognorm);
    double raw = exp(a) * norm;
    return raw > 0.0 ? -log(raw) : 0.0;
}
#define GF_LL_UNROLL 4
__global__ void __launch_bounds__(GF_THREADS, 4)
    k_eval_logleaf(const gf_dev_view V, int node, double *partials, unsigned *ticket, double *out, const gf_peer_view PV) {
    __shared__ double red[32];
    __shared__ double sp[32];
    const gf_node nd = V.nodes[node];
    const int npar = nd.npar < 32 ? nd.npar : 32; // host side refuses longer polynomials for this shape
    if(threadIdx.x < npar) sp[threadIdx.x] = V.params[nd.par_off + threadIdx.x];
    __syncthreads();
    const int op = nd.op;
    const double norm = V.norms[nd.norm_slot];
    const double lognorm = norm > 0 ? log(norm) : 0.0;
    const bool shortcut = norm > 0 && fabs(lognorm) < 100.0;
    const double *col = V.cols[V.ints[nd.obs_off]];
    const double2 *col2 = reinterpret_cast<const double2 *>(col); // columns are 256-B aligned (gf_upload_columns)
    const long npair = V.n_events >> 1;
    const long stride = (long)gridDim.x * blockDim.x;
    double acc0 = 0, acc1 = 0;
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    // GF_LL_UNROLL independent 16-byte loads in flight per thread
    for(; i + (GF_LL_UNROLL - 1) * stride < npair; i += GF_LL_UNROLL * stride) {
        double2 x[GF_LL_UNROLL];
#pragma unroll
        for(int u = 0; u < GF_LL_UNROLL; ++u) x[u] = __ldcs(col2 + i + u * stride);
#pragma unroll
        for(int u = 0; u < GF_LL_UNROLL; ++u) {
            acc0 += gf_logleaf_term(gf_leaf_arg(op, npar, sp, x[u].x), norm, lognorm, shortcut);
            acc1 += gf_logleaf_term(gf_leaf_arg(op, npar, sp, x[u].y), norm, lognorm, shortcut);
        }
    }
    for(; i < npair; i += stride) {
        double2 x = __ldcs(col2 + i);
        acc0 += gf_logleaf_term(gf_leaf_arg(op, npar, sp, x.x), norm, lognorm, shortcut);
        acc1 += gf_logleaf_term(gf_leaf_arg(op, npar, sp, x.y), norm, lognorm, shortcut);
    }
    if((V.n_events & 1) && blockIdx.x == 0 && threadIdx.x == 0)
        acc0 += gf_logleaf_term(gf_leaf_arg(op, npar, sp, col[V.n_events - 1]), norm, lognorm, shortcut);
    double acc = acc0 + acc1;
    gf_grid_finish(&acc, 1, partials, ticket, out, red, PV);
}

// ---- shape A ---------------------------------------------------------------------------------------------------------
struct gf_fast_amp {
    int amp, eff, add, bkg, root; // node indices; add = bkg = -1 without the AddPdf wrapper
};
// TD: 0 = Amp4Body, 1 = Amp4Body_TD with TruthResolution, 2 = Amp4Body_TD with ThreeGaussResolution
// resident blocks per SM: 3 x 256 threads x 128 B of amplitudes requested per thread is several times what HBM latency needs,
// and 80 registers hold a chunk of eight amplitudes without spilling (64 do not); the ThreeGauss copy is erfc-bound
constexpr int fast_amp_blocks_per_sm(int td) { return td == 2 ? 2 : 3; }
template <int TD, bool WRAP>
__global__ void __launch_bounds__(GF_THREADS, fast_amp_blocks_per_sm(TD))
    k_eval_fast_amp4(const gf_dev_view V, const gf_fast_amp F, double *partials, unsigned *ticket, double *out, const gf_peer_view PV) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red[32];
    // cached amplitudes requested together before the first multiply (16 B each); six cover examples/TDDP4 in one batch
    constexpr int GF_FA_CHUNK = TD == 0 ? 8 : 6;
    __shared__ int kk[64]; // Amp4Body_TD: coefficient-pair shift of amplitude i (earlier amplitudes that are in both lists)
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const gf_node &an = P.nodes[F.amp];
    const double *p = P.params + an.par_off;
    const int32_t *in = P.ints + an.int_off;
    const int32_t *ob = P.ints + an.obs_off;
    const int namp = in[3];
    const int32_t *td = in + in[4];
    if(TD != 0) {
        if(threadIdx.x == 0) {
            int k = 0;
            for(int i = 0; i < namp; ++i) {
                kk[i] = k;
                if(td[1 + i] == 2) ++k;
            }
        }
        __syncthreads();
    }
    // event-independent pieces, once per thread, in the interpreter's operation order
    const gf_node &en = P.nodes[F.eff];
    const double eff = P.params[en.par_off + en.npar - 1]; // GF_OP_CONSTANT
    const double norm = P.norms[P.nodes[F.root].norm_slot];
    double w0 = 0, n0 = 0, t2 = 0;
    if(WRAP) { // AddPdf.cu:16-51 with two components: ret = w0 * amp * n0 ; ret += (1 - w0) * bkg * n1
        const gf_node &ad = P.nodes[F.add];
        const int32_t *ai = P.ints + ad.int_off;
        const gf_node &bn = P.nodes[F.bkg];
        w0 = P.params[ad.par_off];
        n0 = P.norms[ai[0]];
        double tw = 0;
        tw += w0;
        t2 = (1 - tw) * P.params[bn.par_off + bn.npar - 1] * P.norms[ai[1]];
    }
    const double2 *w = static_cast<const double2 *>(P.aux[in[0]]);
    const long ld = P.aux_ld[in[0]];
    const int evcol = ob[5];
    const bool ident = (V.identity_mask >> evcol) & 1u;
    const double *evp = P.cols[evcol];
    const double *ctime = TD != 0 ? P.cols[ob[6]] : nullptr, *csig = TD != 0 ? P.cols[ob[7]] : nullptr;
    const gf_node &rn = P.nodes[TD != 0 ? td[0] : 0];
    const double *rp = P.params + rn.par_off;
    double acc = 0;
    const long stride = (long)gridDim.x * blockDim.x;
    for(long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < V.n_events; i += stride) {
        const long e = ident ? i : (long)floor(0.5 + __ldg(evp + i));
        double dtime = 0, sigma = 0;
        if(TD != 0) { dtime = __ldg(ctime + i); sigma = __ldg(csig + i); }
        const double2 *we = w + e;
        double ret;
        if(TD == 0) {
            double re = 0, im = 0;
            for(int a0 = 0; a0 < namp; a0 += GF_FA_CHUNK) {
                double2 c[GF_FA_CHUNK];
#pragma unroll
                for(int j = 0; j < GF_FA_CHUNK; ++j)
                    if(a0 + j < namp) c[j] = __ldg(we + (size_t)(a0 + j) * ld);
#pragma unroll
                for(int j = 0; j < GF_FA_CHUNK; ++j) {
                    if(a0 + j >= namp) break;
                    const double ar = p[2 * (a0 + j)], ai = p[2 * (a0 + j) + 1];
                    re += ar * c[j].x - ai * c[j].y;
                    im += ar * c[j].y + ai * c[j].x;
                }
            }
            ret = (re * re + im * im) * eff;
        } else {
            double Are = 0, Aim = 0, Bre = 0, Bim = 0;
            for(int a0 = 0; a0 < namp; a0 += GF_FA_CHUNK) {
                double2 c[GF_FA_CHUNK];
#pragma unroll
                for(int j = 0; j < GF_FA_CHUNK; ++j)
                    if(a0 + j < namp) c[j] = __ldg(we + (size_t)(a0 + j) * ld);
#pragma unroll
                for(int j = 0; j < GF_FA_CHUNK; ++j) {
                    const int a = a0 + j;
                    if(a >= namp) break;
                    const int flag = td[1 + a], q = 4 + 2 * (a + kk[a]);
                    double cr = p[q], ci = p[q + 1];
                    const double2 me = c[j];
                    const double re = me.x * cr - me.y * ci, im = me.x * ci + me.y * cr;
                    if(flag == 1) {
                        Bre += re; Bim += im;
                    } else {
                        Are += re; Aim += im;
                        if(flag == 2) {
                            cr = p[q + 2]; ci = p[q + 3];
                            Bre += me.x * cr - me.y * ci;
                            Bim += me.x * ci + me.y * cr;
                        }
                    }
                }
            }
            Are *= p[3];
            Aim *= p[3];
            const double nA = Are * Are + Aim * Aim, nB = Bre * Bre + Bim * Bim;
            const double t3re = Are * Bre + Aim * Bim, t3im = Aim * Bre - Are * Bim;
            ret = gf_tres<TD == 2>(rn.op, rp, nA + nB, nA - nB, t3re, t3im, p[0], dtime, p[1], p[2], sigma) * eff;
        }
        if(WRAP) {
            double r = 0;
            r += w0 * ret * n0;
            r += t2;
            ret = r;
        }
        ret *= norm;
        acc += ret > 0.0 ? -log(ret) : 0.0;
    }
    gf_grid_finish(&acc, 1, partials, ticket, out, red, PV);
}

// ---- host side: recognise the shape of nodes [first, root] -------------------------------------------------------------
static bool is_exp_leaf(int op) {
    return op == GF_OP_EXP || op == GF_OP_EXP_OFFSET || op == GF_OP_EXP_POLY || op == GF_OP_EXP_POLY_OFFSET || op == GF_OP_GAUSSIAN
           || op == GF_OP_SCALED_GAUSSIAN;
}
typedef void (*fast_amp_kernel_t)(const gf_dev_view, const gf_fast_amp, double *, unsigned *, double *, const gf_peer_view);
static fast_amp_kernel_t fast_amp_kernel(int td, bool wrap) {
    switch(td * 2 + (wrap ? 1 : 0)) {
    case 0: return k_eval_fast_amp4<0, false>;
    case 1: return k_eval_fast_amp4<0, true>;
    case 2: return k_eval_fast_amp4<1, false>;
    case 3: return k_eval_fast_amp4<1, true>;
    case 4: return k_eval_fast_amp4<2, false>;
    default: return k_eval_fast_amp4<2, true>;
    }
}
int gf_fast_reserve_smem(size_t sm) {
    cudaError_t e = cudaSuccess;
    for(int td = 0; td < 3 && e == cudaSuccess; td++)
        for(int wr = 0; wr < 2 && e == cudaSuccess; wr++)
            e = cudaFuncSetAttribute(fast_amp_kernel(td, wr != 0), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    return e == cudaSuccess ? 0 : 1;
}
// 0: not a known shape (the caller runs the interpreter), 1: launched, -1: error
int gf_fast_launch(gf_ctx *c, int first, int root, int metric, double *out_dev, const gf_peer_view &PV, const char **which) {
    if(!c->fast_on || metric != GF_METRIC_NLL) return 0;
    const std::vector<gf_node> &N = c->h_nodes;
    const int n = root - first + 1;
    if(n == 1 && is_exp_leaf(N[root].op) && N[root].nobs >= 1 && N[root].npar <= 32 && c->cols_owned) {
        // two events per 16-byte load: half as many work items as events
        int b2 = gf_grid_blocks(c, (c->view.n_events + 1) / 2, 4);
        k_eval_logleaf<<<b2, GF_THREADS, 0, c->stream>>>(c->view, root, c->d_partials, c->d_ticket, out_dev, PV);
        if(which) *which = "k_eval_logleaf";
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    gf_fast_amp F{-1, -1, -1, -1, root};
    if(n == 2 && (N[root].op == GF_OP_AMP4 || N[root].op == GF_OP_AMP4_TD)) {
        F.amp = root; F.eff = first;
    } else if(n == 4 && N[root].op == GF_OP_ADD && N[root].nchild == 2 && N[root].npar >= 1 && N[root].nint >= 2
              && (N[first + 1].op == GF_OP_AMP4 || N[first + 1].op == GF_OP_AMP4_TD) && N[first + 1].size == 2) {
        F.amp = first + 1; F.eff = first; F.bkg = first + 2; F.add = root;
        if(N[F.bkg].op != GF_OP_CONSTANT || N[F.bkg].npar < 1) return 0;
    } else
        return 0;
    if(N[F.eff].op != GF_OP_CONSTANT || N[F.eff].npar < 1) return 0;
    const int32_t *in = c->h_ints.data() + N[F.amp].int_off;
    if(in[3] > 64) return 0;
    int td = 0;
    if(N[F.amp].op == GF_OP_AMP4_TD) td = N[in[in[4]]].op == GF_OP_TRES_TRUTH ? 1 : 2;
    const int blocks = gf_grid_blocks(c, c->view.n_events, fast_amp_blocks_per_sm(td));
    fast_amp_kernel(td, F.add >= 0)<<<blocks, GF_THREADS, gf_prog_smem(c), c->stream>>>(c->view, F, c->d_partials, c->d_ticket, out_dev, PV);
    if(which) *which = td == 0 ? "k_eval_fast_amp4" : "k_eval_fast_amp4_td";
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// A/B switches (return the previous setting, -1 for an unknown name):
//   "fast"  shape-specialised likelihood kernels of this file (0: interpreter for every program)
//   "fused" one-pass lineshape -> amplitude -> Gram update of gf_amp4.cu (0: separate kernels + 4x4-tile Gram sums)
extern "C" int gf_set_option(gf_ctx *c, const char *name, int value) {
    if(!c || !name) return -1;
    bool *flag = !std::strcmp(name, "fast") ? &c->fast_on : (!std::strcmp(name, "fused") ? &c->fused_on : nullptr);
    if(!flag) return -1;
    int was = *flag ? 1 : 0;
    *flag = value != 0;
    return was;
}
