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

// table[k-kmin] = sum_{i<n_model} model[i] * resol[i + offset - k]. A block computes 64 consecutive k; the operands come from
// shared memory (GF_CONV_TILE model points and the GF_CONV_TILE + 63 resolution values they meet per round), so the sums wait
// on LDS and the FP64 pipe instead of on one global load per term.
// Round 1 ran the reference's sequential FMA chain (one thread per k, i ascending: bit-identical tables) - 64 threads per
// block, two warps per SM, 105 us per 10^4 x 10^4 table, a third of the zachFit call when all three convolutions are refilled.
// Now a block is 256 threads: GF_CONV_SEG groups each sum their share of every tile, a thread holds GF_CONV_KPT consecutive
// entries whose resolution values form a sliding window in registers (one new shared-memory load per KPT FMAs instead of one per
// FMA), and a round's global loads are all issued before the first is stored. The partial sums of an entry are added in a fixed
// order - deterministic and run-to-run identical, but not the reference's order: an entry moves by a few ulp (10^4 positive
// terms), the NLL by < 1e-14 relative against the 1e-12 bar. -DGF_CONV_SEG=1 -DGF_CONV_KPT=1 restores the sequential chain.
#define GF_CONV_TILE 512
#ifndef GF_CONV_SEG
#define GF_CONV_SEG 16 // thread groups per entry (1 with GF_CONV_KPT 1 = the reference's sequential chain)
#endif
#ifndef GF_CONV_KPT
#define GF_CONV_KPT 4 // consecutive k per thread: they share a sliding window of resolution values, one new LDS per KPT FMAs
#endif
#define GF_CONV_KT (64 / GF_CONV_KPT)              // threads along k
#define GF_CONV_THREADS (GF_CONV_KT * GF_CONV_SEG) // 256
// resolution values are stored swizzled (one 8-byte slot skipped every 16) so that the window loads of the GF_CONV_KT threads of a
// group, GF_CONV_KPT doubles apart, fall into different banks
__device__ __forceinline__ int gf_conv_sw(int q) { return q + (q >> 4); }
__global__ void __launch_bounds__(GF_CONV_THREADS)
    k_conv_table(const double *__restrict__ model, const double *__restrict__ resol, int n_model, int n_resol,
                 int offset, int kmin, int nk, double *table) {
    constexpr int NRES = GF_CONV_TILE + 64;
    __shared__ double sm_model[GF_CONV_TILE];
    __shared__ double sm_res[NRES + NRES / 16 + 1];
    __shared__ double sm_part[GF_CONV_SEG][64];
    const int kt = threadIdx.x % GF_CONV_KT, seg = threadIdx.x / GF_CONV_KT;
    const int k0 = kmin + blockIdx.x * 64;
    constexpr int PER = GF_CONV_TILE / GF_CONV_SEG; // terms of a tile per group
    static_assert(GF_CONV_TILE % GF_CONV_SEG == 0 && 64 % GF_CONV_KPT == 0, "tile must split evenly");
    double acc[GF_CONV_KPT];
#pragma unroll
    for(int j = 0; j < GF_CONV_KPT; ++j) acc[j] = 0;
    for(int i0 = 0; i0 < n_model; i0 += GF_CONV_TILE) {
        const int n = min(GF_CONV_TILE, n_model - i0);
        const int jmin = i0 + offset - (k0 + 63); // resolution index met by (i0, the block's last k)
        // this round's values into registers first (every load in flight at once), then into shared memory
        constexpr int NM = (GF_CONV_TILE + GF_CONV_THREADS - 1) / GF_CONV_THREADS, NR = (NRES + GF_CONV_THREADS - 1) / GF_CONV_THREADS;
        double vm[NM], vr[NR];
#pragma unroll
        for(int u = 0; u < NM; ++u) {
            const int q = threadIdx.x + u * GF_CONV_THREADS;
            vm[u] = q < n ? __ldg(model + i0 + q) : 0.0; // zero-padded: the terms past n add exact zeros
        }
#pragma unroll
        for(int u = 0; u < NR; ++u) {
            const int q = threadIdx.x + u * GF_CONV_THREADS, jj = jmin + q;
            vr[u] = (q < NRES && jj >= 0 && jj < n_resol) ? __ldg(resol + jj) : 0.0;
        }
        __syncthreads();
#pragma unroll
        for(int u = 0; u < NM; ++u) {
            const int q = threadIdx.x + u * GF_CONV_THREADS;
            if(q < GF_CONV_TILE) sm_model[q] = vm[u];
        }
#pragma unroll
        for(int u = 0; u < NR; ++u) {
            const int q = threadIdx.x + u * GF_CONV_THREADS;
            if(q < NRES) sm_res[gf_conv_sw(q)] = vr[u];
        }
        __syncthreads();
        // entry k = k0 + KPT kt + j meets model[ii] with sm_res[(63 - KPT kt - j) + ii]: a window of KPT values that slides by one
        // per term. w[j] holds sm_res[base - j + ii].
        const double *m = sm_model + seg * PER;
        const int base = 63 - GF_CONV_KPT * kt + seg * PER;
        double w[GF_CONV_KPT];
#pragma unroll
        for(int j = 0; j < GF_CONV_KPT; ++j) w[j] = sm_res[gf_conv_sw(base - j)];
#pragma unroll 8
        for(int ii = 0; ii < PER; ++ii) {
            const double mv = m[ii];
#pragma unroll
            for(int j = 0; j < GF_CONV_KPT; ++j) acc[j] += mv * w[j];
#pragma unroll
            for(int j = GF_CONV_KPT - 1; j > 0; --j) w[j] = w[j - 1];
            w[0] = sm_res[gf_conv_sw(base + ii + 1)];
        }
    }
#pragma unroll
    for(int j = 0; j < GF_CONV_KPT; ++j) sm_part[seg][GF_CONV_KPT * kt + j] = acc[j];
    __syncthreads();
    if(threadIdx.x < 64 && blockIdx.x * 64 + (int)threadIdx.x < nk) {
        double sum = sm_part[0][threadIdx.x];
#pragma unroll
        for(int g = 1; g < GF_CONV_SEG; ++g) sum += sm_part[g][threadIdx.x];
        table[blockIdx.x * 64 + threadIdx.x] = sum;
    }
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
    k_conv_table<<<(s->nk + 63) / 64, GF_CONV_THREADS, 0, c->stream>>>(s->d_model, s->d_resol, s->n_model, s->n_resol, s->offset,
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
