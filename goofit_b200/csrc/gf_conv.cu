// gf_conv.cu — ConvolutionPdf workspaces and the per-offset convolution table.
// The reference recomputes sum_i model[i]*resol[i+offset-offsetInBins] for EVERY event and every
// grid point (combine/ConvolutionPdf.cu:27-75: numbins FMAs per evaluation). The sum depends on the
// event only through the integer offsetInBins, so it is computed once per distinct offset (same
// i-order, hence the same rounding) and the per-event op becomes one gather.
#include "gf_internal.cuh"
#include <cmath>

struct gf_conv_state {
    int node = -1, slot = -1, model_root = -1, resol_root = -1;
    double lo = 0, hi = 0, step = 0, obs_lo = 0, obs_hi = 0;
    int n_model = 0, n_resol = 0, offset = 0, kmin = 0, nk = 0;
    double *d_model = nullptr, *d_resol = nullptr, *d_table = nullptr;
};

void gf_conv_free(gf_ctx *c) {
    for(auto *s : c->conv) {
        cudaFree(s->d_model); cudaFree(s->d_resol); cudaFree(s->d_table);
        delete s;
    }
    c->conv.clear();
}

static gf_conv_state *find_conv(gf_ctx *c, int node, bool create) {
    for(auto *s : c->conv)
        if(s->node == node) return s;
    if(!create) return nullptr;
    auto *s = new gf_conv_state();
    s->node = node;
    c->conv.push_back(s);
    return s;
}

// Workspace fill = MetricTaker's binned operator with EvalFunc::Eval on one sub-PDF
// (ConvolutionPdf.cu:331-351, MetricTaker.cu:48-96): bin centre -> slot of the sub-PDF's first observable.
__global__ void __launch_bounds__(GF_THREADS, 4)
    k_conv_fill(const gf_dev_view V, int first, int root, double lo, double hi, int nb, int col, double *out) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const long stride = (long)gridDim.x * blockDim.x;
    for(long b = (long)blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += stride) {
        gf_evt_vals evt;
#pragma unroll
        for(int k = 0; k < GF_MAX_OBS; k++) evt.v[k] = 0;
        evt.i = b;
        double x = hi - lo;
        x /= nb;
        x *= ((int)b + 0.5);
        x += lo;
        evt.v[col] = x;
        out[b] = gf_eval_tree(P, V, first, root, evt);
    }
}

// table[k-kmin] = sum_{i<n_model} model[i] * resol[i + offset - k], i ascending: one thread per k runs the reference's
// sequential FMA chain (same order, same rounding), but the operands come from shared memory: a block of 64 consecutive
// k stages GF_CONV_TILE model points and the GF_CONV_TILE + 63 resolution values they meet, so the chain waits on LDS and
// the FP64 pipe instead of on one global load per term.
#define GF_CONV_TILE 512
__global__ void __launch_bounds__(64)
    k_conv_table(const double *__restrict__ model, const double *__restrict__ resol, int n_model, int n_resol,
                 int offset, int kmin, int nk, double *table) {
    __shared__ double sm_model[GF_CONV_TILE];
    __shared__ double sm_res[GF_CONV_TILE + 64];
    const int tl = threadIdx.x, t = blockIdx.x * 64 + tl;
    const int k0 = kmin + blockIdx.x * 64;
    double ret = 0;
    for(int i0 = 0; i0 < n_model; i0 += GF_CONV_TILE) {
        const int n = min(GF_CONV_TILE, n_model - i0);
        const int jmin = i0 + offset - (k0 + 63); // resolution index met by (i0, the block's last k)
        __syncthreads();
        for(int q = tl; q < GF_CONV_TILE; q += 64) sm_model[q] = q < n ? __ldg(model + i0 + q) : 0.0;
        for(int q = tl; q < GF_CONV_TILE + 64; q += 64) {
            int jj = jmin + q;
            sm_res[q] = (jj >= 0 && jj < n_resol) ? __ldg(resol + jj) : 0.0;
        }
        __syncthreads();
        const double *r = sm_res + (63 - tl);
#pragma unroll 8
        for(int ii = 0; ii < n; ++ii) ret += sm_model[ii] * r[ii];
    }
    if(t < nk) table[t] = ret;
}

static int conv_check(gf_ctx *c, int node, gf_conv_state **out, bool create) {
    if(!c) return gf_fail("null ctx");
    if(node < 0 || node >= (int)c->h_nodes.size() || c->h_nodes[node].op != GF_OP_CONVOLVE)
        return gf_fail("gf_conv: node is not a ConvolutionPdf record");
    const gf_node &nd = c->h_nodes[node];
    if(nd.nint < 3 || nd.ncon < 3 || nd.nobs < 1) return gf_fail("gf_conv: malformed ConvolutionPdf record");
    gf_conv_state *s = find_conv(c, node, create);
    if(!s) return gf_fail("gf_conv: gf_conv_setup not called for this node");
    s->slot = c->h_ints[nd.int_off];
    s->model_root = c->h_ints[nd.int_off + 1];
    s->resol_root = c->h_ints[nd.int_off + 2];
    int nn = (int)c->h_nodes.size();
    if(s->slot < 0 || s->slot >= GF_MAX_AUX || s->model_root < 0 || s->model_root >= nn || s->resol_root < 0
       || s->resol_root >= nn)
        return gf_fail("gf_conv: bad slot / sub-PDF index");
    *out = s;
    return 0;
}

// ConvolutionPdf::setIntegrationConstants (ConvolutionPdf.cu:229-282)
extern "C" int gf_conv_setup(gf_ctx *c, int node, double lo, double hi, double step, double obs_lo, double obs_hi) {
    gf_conv_state *s;
    if(conv_check(c, node, &s, true)) return 1;
    if(!(step > 0) || !(hi > lo)) return gf_fail("gf_conv_setup: bad range");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    s->lo = lo; s->hi = hi; s->step = step; s->obs_lo = obs_lo; s->obs_hi = obs_hi;
    s->n_model = (int)std::floor((hi - lo) / step + 0.5);
    double rlo = lo - obs_hi, rhi = hi - obs_lo;
    s->n_resol = (int)std::floor((rhi - rlo) / step + 0.5);
    s->offset = (int)(obs_hi / step);
    double lbo = lo / step;
    lbo -= std::floor(lbo);
    s->kmin = (int)std::floor(obs_lo / step - lbo) - 1; // one spare entry either side for edge rounding
    int kmax = (int)std::floor(obs_hi / step - lbo) + 1;
    s->nk = kmax - s->kmin + 1;
    if(s->n_model < 1 || s->n_resol < 1 || s->nk < 1) return gf_fail("gf_conv_setup: empty workspace");
    cudaFree(s->d_model); cudaFree(s->d_resol); cudaFree(s->d_table);
    GF_CUDA(cudaMalloc(&s->d_model, sizeof(double) * s->n_model));
    GF_CUDA(cudaMalloc(&s->d_resol, sizeof(double) * s->n_resol));
    GF_CUDA(cudaMalloc(&s->d_table, sizeof(double) * s->nk));
    GF_CUDA(cudaMemset(s->d_model, 0, sizeof(double) * s->n_model));
    GF_CUDA(cudaMemset(s->d_resol, 0, sizeof(double) * s->n_resol));
    GF_CUDA(cudaMemset(s->d_table, 0, sizeof(double) * s->nk));
    c->view.aux[s->slot] = s->d_table;
    c->view.aux_ld[s->slot] = ((long)s->kmin << 32) | (unsigned)(s->nk);
    return 0;
}

extern "C" int gf_conv_fill(gf_ctx *c, int node, int fill_model, int fill_resol) {
    gf_conv_state *s;
    if(conv_check(c, node, &s, false)) return 1;
    if(!fill_model && !fill_resol) return 0;
    cudaSetDevice(c->device);
    size_t sm = gf_prog_smem(c);
    if(sm > 48 * 1024) cudaFuncSetAttribute(k_conv_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if(fill_model) {
        const gf_node &m = c->h_nodes[s->model_root];
        if(m.nobs < 1) return gf_fail("gf_conv_fill: model has no observable");
        int col = c->h_ints[m.obs_off];
        k_conv_fill<<<gf_grid_blocks(c, s->n_model, 4), GF_THREADS, sm, c->stream>>>(
            c->view, s->model_root - m.size + 1, s->model_root, s->lo, s->hi, s->n_model, col, s->d_model);
        c->launches++;
    }
    if(fill_resol) {
        const gf_node &r = c->h_nodes[s->resol_root];
        if(r.nobs < 1) return gf_fail("gf_conv_fill: resolution has no observable");
        int col = c->h_ints[r.obs_off];
        k_conv_fill<<<gf_grid_blocks(c, s->n_resol, 4), GF_THREADS, sm, c->stream>>>(
            c->view, s->resol_root - r.size + 1, s->resol_root, s->lo - s->obs_hi, s->hi - s->obs_lo, s->n_resol, col,
            s->d_resol);
        c->launches++;
    }
    k_conv_table<<<(s->nk + 63) / 64, 64, 0, c->stream>>>(s->d_model, s->d_resol, s->n_model, s->n_resol, s->offset,
                                                         s->kmin, s->nk, s->d_table);
    c->launches++;
    GF_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int gf_conv_get_workspace(gf_ctx *c, int node, int which, double *out, int n) {
    gf_conv_state *s;
    if(conv_check(c, node, &s, false)) return 1;
    const double *src = which == 0 ? s->d_model : (which == 1 ? s->d_resol : s->d_table);
    int have = which == 0 ? s->n_model : (which == 1 ? s->n_resol : s->nk);
    if(n != have) return gf_fail("gf_conv_get_workspace: expected " + std::to_string(have) + " entries");
    cudaSetDevice(c->device);
    GF_CUDA(cudaMemcpyAsync(out, src, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
extern "C" int gf_conv_sizes(gf_ctx *c, int node, int *n_model, int *n_resol, int *n_table) {
    gf_conv_state *s;
    if(conv_check(c, node, &s, false)) return 1;
    if(n_model) *n_model = s->n_model;
    if(n_resol) *n_resol = s->n_resol;
    if(n_table) *n_table = s->nk;
    return 0;
}

// ---- plain tables (SmoothHistogramPdf) ----------------------------------------------------------------------------------------
// One device array per GF_OP_HISTOGRAM record, reachable through the record's aux slot; kept in the same list as the convolution
// workspaces so that gf_reset_aux / gf_destroy release it.
extern "C" int gf_table_set(gf_ctx *c, int node, const double *values, size_t n) {
    if(!c || !values || n == 0) return gf_fail("gf_table_set: null argument or empty table");
    if(node < 0 || node >= (int)c->h_nodes.size() || c->h_nodes[node].op != GF_OP_HISTOGRAM) return gf_fail("gf_table_set: node is not a histogram record");
    const gf_node &nd = c->h_nodes[node];
    if(nd.nint < 1) return gf_fail("gf_table_set: malformed histogram record");
    const int slot = c->h_ints[nd.int_off];
    if(slot < 0 || slot >= GF_MAX_AUX) return gf_fail("gf_table_set: bad aux slot");
    cudaSetDevice(c->device);
    gf_conv_state *s = find_conv(c, node, true);
    s->slot = slot;
    GF_CUDA(cudaStreamSynchronize(c->stream));
    if(!s->d_table || s->nk != (int)n) {
        cudaFree(s->d_table);
        s->d_table = nullptr;
        GF_CUDA(cudaMalloc(&s->d_table, sizeof(double) * n));
        s->nk = (int)n;
    }
    GF_CUDA(cudaMemcpy(s->d_table, values, sizeof(double) * n, cudaMemcpyHostToDevice));
    c->view.aux[slot] = s->d_table;
    c->view.aux_ld[slot] = (long)n;
    return 0;
}
