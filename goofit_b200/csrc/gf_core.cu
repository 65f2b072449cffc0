// gf_core.cu — context, program/param upload, SoA dataset upload, and the three generic kernels:
//   k_eval_sum    fused evaluate -> metric -> deterministic grid reduction   (GooPdf.cu:45-78, MetricTaker.cu:17-44)
//   k_eval_events per-event metric values                                    (GooPdf.cu:112-173)
//   k_grid_sum    mid-point grid integral of a subtree                       (GooPdf.cu:83-108, MetricTaker.cu:48-96)
// No Thrust, no function pointers, no CPU fallback: every entry point fails loudly without a device.
#include "gf_internal.cuh"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <dlfcn.h>

static thread_local std::string g_err;
void gf_set_error(const std::string &s) { g_err = s; }
int gf_fail(const std::string &s) {
    g_err = s;
    return 1;
}
extern "C" const char *gf_last_error(void) { return g_err.c_str(); }
extern "C" int gf_version(void) { return 100; }

// gf_fast.cu: shape-specialised copies of the likelihood kernel (no interpreter); 0 = shape not recognised
int gf_fast_launch(gf_ctx *c, int first, int root, int metric, double *out_dev, const gf_peer_view &PV, const char **which);
int gf_fast_reserve_smem(size_t sm);

size_t gf_prog_smem(const gf_ctx *c) {
    return gf_sprog_bytes((int)c->h_nodes.size(), c->n_params, (int)c->h_consts.size(), c->n_norms, (int)c->h_ints.size());
}
int gf_grid_blocks(const gf_ctx *c, long n_items, int per_sm) {
    long need = (n_items + GF_THREADS - 1) / GF_THREADS;
    long cap = (long)c->sm_count * per_sm;
    long b = std::min(need, cap);
    if(b < 1) b = 1;
    return (int)b;
}
int gf_alloc_aux_slot(gf_ctx *c) {
    for(int i = 0; i < GF_MAX_AUX; i++)
        if(c->view.aux[i] == nullptr && c->view.aux_ld[i] == 0) {
            c->view.aux_ld[i] = -1; // reserved
            return i;
        }
    return -1;
}

// ---- NCCL through dlopen: the library keeps no link-time dependency on it ---------------------------------
struct gf_nccl_uid { char internal[128]; };
namespace {
typedef int (*nccl_get_uid_t)(void *);
typedef int (*nccl_init_rank_t)(void **, int, gf_nccl_uid, int);
typedef int (*nccl_allreduce_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*nccl_allgather_t)(const void *, void *, size_t, int, void *, cudaStream_t);
typedef int (*nccl_destroy_t)(void *);
typedef const char *(*nccl_errstr_t)(int);
}
static void *g_nccl = nullptr;
static nccl_get_uid_t p_uid = nullptr;
static nccl_init_rank_t p_init = nullptr;
static nccl_allreduce_t p_allreduce = nullptr;
static nccl_allgather_t p_allgather = nullptr;
static nccl_destroy_t p_destroy = nullptr;
static nccl_errstr_t p_errstr = nullptr;
static int load_nccl() {
    if(g_nccl) return 0;
    g_nccl = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if(!g_nccl) g_nccl = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if(!g_nccl) return gf_fail(std::string("cannot dlopen libnccl.so.2: ") + dlerror());
    p_uid = (nccl_get_uid_t)dlsym(g_nccl, "ncclGetUniqueId");
    p_init = (nccl_init_rank_t)dlsym(g_nccl, "ncclCommInitRank");
    p_allreduce = (nccl_allreduce_t)dlsym(g_nccl, "ncclAllReduce");
    p_allgather = (nccl_allgather_t)dlsym(g_nccl, "ncclAllGather");
    p_destroy = (nccl_destroy_t)dlsym(g_nccl, "ncclCommDestroy");
    p_errstr = (nccl_errstr_t)dlsym(g_nccl, "ncclGetErrorString");
    if(!p_uid || !p_init || !p_allreduce || !p_destroy) return gf_fail("libnccl: missing symbols");
    return 0;
}
static int nccl_fail(const char *what, int rc) {
    return gf_fail(std::string(what) + ": " + (p_errstr ? p_errstr(rc) : "NCCL error") + " (" + std::to_string(rc) + ")");
}

// ---------------------------------------------------------------------------------------------
// One copy per class of program (HAS = the heavy node kinds compiled in, gf_ops.cuh); eval_sum_launch picks the copy.
// Measured on 10^7 events against the single all-kinds kernel: Amp3Body-only 0.416 ms (was 0.431, no spills left),
// Amp4Body_TD + TruthResolution 0.387 ms (was 0.452). An Amp4Body-only copy came out slower than the all-kinds one
// (0.304 against 0.275 ms, same source, different schedule), so Amp4Body programs keep the all-kinds copy.
constexpr int eval_sum_blocks_per_sm(int) { return 4; } // 3 and 5 measured slower for every copy (tools/ab_compare.sh)
template <int HAS>
__global__ void __launch_bounds__(GF_THREADS, eval_sum_blocks_per_sm(HAS))
    k_eval_sum(const gf_dev_view V, int first, int root, int metric, int n_cols, double *partials, unsigned *ticket,
               double *out, const gf_peer_view PV) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red[32];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const double norm = P.norms[P.nodes[root].norm_slot];
    double acc = 0;
    gf_evt_soa evt{P.cols, V.identity_mask, 0};
    const long stride = (long)gridDim.x * blockDim.x;
    for(long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < V.n_events; i += stride) {
        evt.i = i;
        double raw = gf_eval_tree<HAS>(P, V, first, root, evt);
        acc += gf_metric(metric, raw, norm, evt, V, n_cols);
    }
    gf_grid_finish(&acc, 1, partials, ticket, out, red, PV);
}
// heavy node kinds among nodes [first, root] of the installed program
static int program_class(const gf_ctx *c, int first, int root) {
    int has = 0;
    for(int k = first; k <= root; ++k) {
        switch(c->h_nodes[k].op) {
        case GF_OP_CONVOLVE: has |= GF_HAS_CONV; break;
        case GF_OP_AMP3: has |= GF_HAS_AMP3; break;
        case GF_OP_AMP4: has |= GF_HAS_AMP4; break;
        case GF_OP_AMP3_TD: has |= GF_HAS_AMP3_TD | GF_HAS_3GAUSS; break;
        case GF_OP_COMPOSITE_ARG:
        case GF_OP_COMPOSITE: has |= GF_HAS_COMPOSITE; break;
        case GF_OP_AMP4_TD: {
            has |= GF_HAS_AMP4_TD;
            const int32_t *in = c->h_ints.data() + c->h_nodes[k].int_off; // ints[4] -> {resolution record, flags}
            if(c->h_nodes[in[in[4]]].op != GF_OP_TRES_TRUTH) has |= GF_HAS_3GAUSS;
        } break;
        default: break;
        }
    }
    return has;
}
typedef void (*eval_sum_kernel_t)(const gf_dev_view, int, int, int, int, double *, unsigned *, double *, const gf_peer_view);
static eval_sum_kernel_t eval_sum_kernel(int has) {
    switch(has) {
    case 0: return k_eval_sum<0>;
    case GF_HAS_CONV: return k_eval_sum<GF_HAS_CONV>;
    case GF_HAS_AMP3: return k_eval_sum<GF_HAS_AMP3>;
    case GF_HAS_AMP4_TD: return k_eval_sum<GF_HAS_AMP4_TD>;
    case GF_HAS_AMP4_TD | GF_HAS_3GAUSS: return k_eval_sum<GF_HAS_AMP4_TD | GF_HAS_3GAUSS>;
    default: return k_eval_sum<GF_HAS_ALL>;
    }
}

__global__ void __launch_bounds__(GF_THREADS, 4)
    k_eval_events(const gf_dev_view V, int first, int root, int metric, int n_cols, double *out) {
    extern __shared__ __align__(16) unsigned char smem[];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    const double norm = P.norms[P.nodes[root].norm_slot];
    gf_evt_soa evt{P.cols, V.identity_mask, 0};
    const long stride = (long)gridDim.x * blockDim.x;
    for(long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < V.n_events; i += stride) {
        evt.i = i;
        double raw = gf_eval_tree(P, V, first, root, evt);
        out[i] = gf_metric(metric, raw, norm, evt, V, n_cols);
    }
}

struct gf_grid_spec {
    double lo[GF_MAX_OBS], hi[GF_MAX_OBS];
    int nb[GF_MAX_OBS];
    int col[GF_MAX_OBS];
    int nobs;
    long total;
};

// Bin-centre arithmetic exactly as MetricTaker.cu:75-88: x = (hi-lo)/n * (bin+0.5) + lo, x fastest.
// Reference quirk kept: every dimension's centre is written to the slot of the subtree root's FIRST
// observable (`events[id] = x` with id fixed, MetricTaker.cu:73,85) — exact for 1-D integrals,
// which is all the reference's own configs use.
__global__ void __launch_bounds__(GF_THREADS, 4)
    k_grid_sum(const gf_dev_view V, int first, int root, gf_grid_spec G, double *partials, unsigned *ticket,
               double *out) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red[32];
    gf_sprog P = gf_stage_program(V, smem);
    __syncthreads();
    double acc = 0;
    const long stride = (long)gridDim.x * blockDim.x;
    for(long b = (long)blockIdx.x * blockDim.x + threadIdx.x; b < G.total; b += stride) {
        gf_evt_vals evt;
#pragma unroll
        for(int k = 0; k < GF_MAX_OBS; k++) evt.v[k] = 0;
        evt.i = b;
        long bin = b;
        for(int d = 0; d < G.nobs; ++d) {
            int local = (int)(bin % G.nb[d]);
            double x = G.hi[d] - G.lo[d];
            x /= G.nb[d];
            x *= (local + 0.5);
            x += G.lo[d];
            evt.v[G.col[0]] = x;
            bin /= G.nb[d];
        }
        acc += gf_eval_tree(P, V, first, root, evt);
    }
    gf_peer_view none{};
    none.nranks = 1;
    gf_grid_finish(&acc, 1, partials, ticket, out, red, none); // normalisation grids are replicated on every rank
}

__global__ void k_check_identity(const double *col, long n, int *flag) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if(i < n && col[i] != (double)i) *flag = 0;
}

// ---------------------------------------------------------------------------------------------
extern "C" gf_ctx *gf_create(int device_id) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if(e != cudaSuccess || ndev <= 0) {
        gf_fail(std::string("gf_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path");
        return nullptr;
    }
    if(device_id < 0 || device_id >= ndev) {
        gf_fail("gf_create: bad device id");
        return nullptr;
    }
    if(cudaSetDevice(device_id) != cudaSuccess) {
        gf_fail("gf_create: cudaSetDevice failed");
        return nullptr;
    }
    auto *c = new gf_ctx();
    c->device = device_id;
    {
        const char *e = std::getenv("GOOFIT_B200_NO_FAST"), *f = std::getenv("GOOFIT_B200_NO_FUSED");
        c->fast_on = !(e && e[0] == '1');
        c->fused_on = !(f && f[0] == '1');
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device_id);
    c->sm_count = prop.multiProcessorCount;
    c->max_blocks = c->sm_count * 16;
    // every allocation is checked on its own: a context with a null scratch buffer must never be handed out
    auto ok = [&](cudaError_t e, const char *what) {
        if(e == cudaSuccess) return true;
        gf_fail(std::string("gf_create: ") + what + ": " + cudaGetErrorString(e));
        return false;
    };
    bool good = ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "stream")
                && ok(cudaEventCreate(&c->ev0), "event") && ok(cudaEventCreate(&c->ev1), "event")
                && ok(cudaMalloc(&c->d_partials, sizeof(double) * GF_NPARTIAL * c->max_blocks), "partials")
                && ok(cudaMalloc(&c->d_ticket, sizeof(unsigned) * 8), "ticket")
                && ok(cudaMemset(c->d_ticket, 0, sizeof(unsigned) * 8), "ticket memset")
                && ok(cudaMalloc(&c->d_result, sizeof(double) * GF_NPARTIAL * 64), "result")
                && ok(cudaMemset(c->d_result, 0, sizeof(double) * GF_NPARTIAL * 64), "result memset")
                && ok(cudaHostAlloc(&c->h_result, sizeof(double) * GF_NPARTIAL * 64, cudaHostAllocMapped), "pinned result");
    good = good && ok(cudaHostAlloc(&c->h_done, 64, cudaHostAllocMapped), "pinned completion flag");
    if(good) {
        std::memset(c->h_done, 0, 64);
        if(cudaHostGetDevicePointer((void **)&c->h_done_dev, c->h_done, 0) != cudaSuccess) {
            c->h_done_dev = nullptr;
            cudaGetLastError();
        }
        std::memset(c->h_result, 0, sizeof(double) * GF_NPARTIAL * 64);
        if(cudaHostGetDevicePointer((void **)&c->h_result_dev, c->h_result, 0) != cudaSuccess) {
            c->h_result_dev = nullptr;
            cudaGetLastError();
        }
    }
    if(!good) {
        std::string msg = gf_last_error();
        gf_destroy(c);
        gf_fail(msg);
        return nullptr;
    }
    return c;
}

static void peer_teardown(gf_ctx *c) {
    for(int r = 0; r < GF_MAX_PEERS; r++)
        if(c->peer_mapped[r]) {
            cudaIpcCloseMemHandle(c->peer_mapped[r]);
            c->peer_mapped[r] = nullptr;
        }
    cudaFree(c->peer_buf);
    c->peer_buf = nullptr;
    c->peer_on = false;
    c->peer = gf_peer_view{};
    cudaGetLastError();
}
void gf_amp3_free(gf_ctx *c);
void gf_amp4_free(gf_ctx *c);
void gf_conv_free(gf_ctx *c);

static void free_cols(gf_ctx *c) {
    if(c->d_cols_base && c->cols_owned) cudaFree(c->d_cols_base);
    c->d_cols_base = nullptr;
    for(auto &p : c->view.cols) p = nullptr;
    c->view.n_events = 0;
}

extern "C" void gf_destroy(gf_ctx *c) {
    if(!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    gf_amp3_free(c);
    gf_amp4_free(c);
    gf_conv_free(c);
    free_cols(c);
    cudaFree(c->d_nodes); cudaFree(c->d_params); cudaFree(c->d_consts); cudaFree(c->d_ints); // d_norms lives inside d_params' block
    cudaFree(c->d_partials); cudaFree(c->d_ticket); cudaFree(c->d_result); cudaFree(c->d_events_out);
    cudaFreeHost(c->h_result);
    if(c->h_done) cudaFreeHost(c->h_done);
    for(auto &e : c->prof_ev)
        if(e) cudaEventDestroy(e);
    peer_teardown(c);
    if(c->nccl_comm && p_destroy) p_destroy(c->nccl_comm);
    if(c->h_err) cudaFreeHost(c->h_err);
    for(int i = 0; i < gf_ctx::RING; i++) {
        if(c->ring_buf[i]) cudaFreeHost(c->ring_buf[i]);
        if(c->ring_ev[i]) cudaEventDestroy(c->ring_ev[i]);
    }
    if(c->t0) { cudaEventDestroy(c->t0); cudaEventDestroy(c->t1); }
    if(c->h_pin) cudaFreeHost(c->h_pin);
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1);
    cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int gf_device_info(gf_ctx *c, int *sm_count, size_t *free_bytes, size_t *total_bytes) {
    if(!c) return gf_fail("null ctx");
    cudaSetDevice(c->device);
    if(sm_count) *sm_count = c->sm_count;
    size_t f = 0, t = 0;
    GF_CUDA(cudaMemGetInfo(&f, &t));
    if(free_bytes) *free_bytes = f;
    if(total_bytes) *total_bytes = t;
    return 0;
}

// Pure host-side check of a flattened program (no context, no device): index ranges, then the value stack of every
// evaluable subtree. gf_set_program runs it before anything is committed.
extern "C" int gf_validate_program(const gf_node *nodes, int n_nodes, int n_consts, int n_ints, int n_params, int n_norms) {
    if(!nodes || n_nodes <= 0 || n_params < 0 || n_norms <= 0 || n_consts < 0 || n_ints < 0) return gf_fail("gf_set_program: bad sizes");
    for(int k = 0; k < n_nodes; k++) {
        const gf_node &nd = nodes[k];
        if(nd.par_off < 0 || nd.npar < 0 || nd.par_off + nd.npar > n_params) return gf_fail("gf_set_program: param range, node " + std::to_string(k));
        if(nd.con_off < 0 || nd.ncon < 0 || nd.con_off + nd.ncon > n_consts) return gf_fail("gf_set_program: const range, node " + std::to_string(k));
        if(nd.obs_off < 0 || nd.nobs < 0 || nd.obs_off + nd.nobs > n_ints) return gf_fail("gf_set_program: obs range, node " + std::to_string(k));
        if(nd.int_off < 0 || nd.nint < 0 || nd.int_off + nd.nint > n_ints) return gf_fail("gf_set_program: int range, node " + std::to_string(k));
        if(nd.norm_slot < 0 || nd.norm_slot >= n_norms) return gf_fail("gf_set_program: norm slot, node " + std::to_string(k));
        if(nd.size < 1 || nd.size > k + 1) return gf_fail("gf_set_program: subtree size, node " + std::to_string(k));
        if(nd.nobs > GF_MAX_OBS) return gf_fail("gf_set_program: too many observables");
        if(nd.nchild < 0) return gf_fail("gf_set_program: negative child count, node " + std::to_string(k));
    }
    // Post-order walk of each subtree [k - size + 1, k]: a leaf pushes one value, a combiner pops nchild and pushes one.
    // depth[k] = values on the stack after node k when evaluation starts at the first node of k's subtree. The walk of a
    // larger subtree passes through the same states shifted by what is already pending, so one pass from each root's first
    // node covers it: require pops to be available and the peak to fit GF_MAX_STACK (gf_eval_tree's fixed array).
    {
        auto is_eval = [](int op) { return op < GF_OP_RES_RBW; }; // resonance / lineshape / resolution records never run on the stack
        for(int r = 0; r < n_nodes; r++) {
            if(!is_eval(nodes[r].op)) continue;
            int first = r - nodes[r].size + 1, sp = 0, peak = 0;
            for(int k = first; k <= r; k++) {
                const gf_node &nd = nodes[k];
                if(!is_eval(nd.op)) return gf_fail("gf_set_program: non-evaluable record inside the subtree of node " + std::to_string(r));
                int pops = GF_IS_LEAF(nd.op) ? 0 : (nd.op == GF_OP_CONVOLVE ? 0 : nd.nchild);
                const bool combiner = (nd.op >= GF_OP_ADD && nd.op <= GF_OP_PROD) || (nd.op >= GF_OP_EVTW_ADD && nd.op <= GF_OP_MAPPED);
                if(combiner && (nd.nint < nd.nchild || nd.nchild < 1))
                    return gf_fail("gf_set_program: combiner without child norm slots, node " + std::to_string(k));
                if(nd.op == GF_OP_ADD && nd.npar < nd.nchild - 1) return gf_fail("gf_set_program: AddPdf weights, node " + std::to_string(k));
                if(nd.op == GF_OP_ADD_EXT && nd.npar < nd.nchild) return gf_fail("gf_set_program: AddPdf weights, node " + std::to_string(k));
                if(nd.op == GF_OP_EVTW_ADD && nd.nobs < nd.nchild - 1) return gf_fail("gf_set_program: EventWeightedAddPdf weight columns, node " + std::to_string(k));
                if(nd.op == GF_OP_EVTW_ADD_EXT && nd.nobs < nd.nchild) return gf_fail("gf_set_program: EventWeightedAddPdf weight columns, node " + std::to_string(k));
                if(nd.op == GF_OP_COMPOSITE_ARG && (nd.nchild != 1 || nd.nobs < 1)) return gf_fail("gf_set_program: CompositePdf argument record, node " + std::to_string(k));
                if(nd.op == GF_OP_COMPOSITE && (nd.nchild != 2 || k - 1 - nodes[k - 1].size < 0 || nodes[k - 1 - nodes[k - 1].size].op != GF_OP_COMPOSITE_ARG))
                    return gf_fail("gf_set_program: CompositePdf record without its argument record, node " + std::to_string(k));
                if(nd.op == GF_OP_MAPPED && nd.nchild < 2) return gf_fail("gf_set_program: MappedPdf needs a mapping function and a component, node " + std::to_string(k));
                if(nd.op >= GF_OP_AMP3 && nd.op <= GF_OP_AMP3_TD) pops = 1; // the efficiency
                if(sp < pops) return gf_fail("gf_set_program: value-stack underflow at node " + std::to_string(k));
                sp += 1 - pops;
                if(sp > peak) peak = sp;
            }
            if(sp != 1) return gf_fail("gf_set_program: subtree of node " + std::to_string(r) + " does not reduce to one value");
            if(peak > GF_MAX_STACK)
                return gf_fail("gf_set_program: subtree of node " + std::to_string(r) + " needs a value stack of " + std::to_string(peak)
                               + " > GF_MAX_STACK = " + std::to_string(GF_MAX_STACK) + " (nest the components in smaller AddPdf/ProdPdf groups)");
        }
    }
    if(gf_sprog_bytes(n_nodes, n_params, n_consts, n_norms, n_ints) > 160 * 1024) return gf_fail("gf_set_program: program block exceeds shared memory");
    return 0;
}

extern "C" int gf_set_program(gf_ctx *c, const gf_node *nodes, int n_nodes, const double *consts, int n_consts,
                              const int32_t *ints, int n_ints, int n_params, int n_norms) {
    if(!c) return gf_fail("null ctx");
    if(n_nodes <= 0 || n_params < 0 || n_norms <= 0) return gf_fail("gf_set_program: bad sizes");
    cudaSetDevice(c->device);
    if(gf_validate_program(nodes, n_nodes, n_consts, n_ints, n_params, n_norms)) return 1;
    const size_t sm = gf_sprog_bytes(n_nodes, n_params, n_consts, n_norms, n_ints);
    // allocate into temporaries; host and device state is replaced only when every step succeeded
    gf_node *t_nodes = nullptr;
    double *t_params = nullptr, *t_consts = nullptr, *t_pin = nullptr;
    int32_t *t_ints = nullptr;
    auto bail = [&](const std::string &what) {
        cudaFree(t_nodes); cudaFree(t_params); cudaFree(t_consts); cudaFree(t_ints);
        if(t_pin) cudaFreeHost(t_pin);
        cudaGetLastError();
        return gf_fail("gf_set_program: " + what);
    };
    GF_CUDA(cudaStreamSynchronize(c->stream));
    std::vector<double> ones(n_norms, 1.0);
    const size_t need = sizeof(double) * (size_t)(n_params + n_norms + 16);
    // parameters and normalisations share one allocation so that one copy per likelihood call uploads both
    if(cudaMalloc(&t_nodes, sizeof(gf_node) * n_nodes) != cudaSuccess || cudaMalloc(&t_params, sizeof(double) * (size_t)(n_params + n_norms)) != cudaSuccess
       || cudaMalloc(&t_consts, sizeof(double) * std::max(1, n_consts)) != cudaSuccess || cudaMalloc(&t_ints, sizeof(int32_t) * std::max(1, n_ints)) != cudaSuccess)
        return bail("device allocation failed");
    if(need > c->h_pin_cap && cudaMallocHost(&t_pin, need) != cudaSuccess) return bail("pinned allocation failed");
    if(cudaMemcpy(t_nodes, nodes, sizeof(gf_node) * n_nodes, cudaMemcpyHostToDevice) != cudaSuccess
       || (n_consts && cudaMemcpy(t_consts, consts, sizeof(double) * n_consts, cudaMemcpyHostToDevice) != cudaSuccess)
       || (n_ints && cudaMemcpy(t_ints, ints, sizeof(int32_t) * n_ints, cudaMemcpyHostToDevice) != cudaSuccess)
       || (n_params && cudaMemset(t_params, 0, sizeof(double) * n_params) != cudaSuccess)
       || cudaMemcpy(t_params + n_params, ones.data(), sizeof(double) * n_norms, cudaMemcpyHostToDevice) != cudaSuccess)
        return bail("upload failed");
    if(sm > 48 * 1024) {
        cudaError_t e = cudaSuccess;
        for(int has : {0, (int)GF_HAS_CONV, (int)GF_HAS_AMP3, (int)GF_HAS_AMP4_TD, GF_HAS_AMP4_TD | GF_HAS_3GAUSS, (int)GF_HAS_ALL})
            if(e == cudaSuccess) e = cudaFuncSetAttribute(eval_sum_kernel(has), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if(e == cudaSuccess && gf_fast_reserve_smem(sm)) e = cudaGetLastError();
        if(e == cudaSuccess) e = cudaFuncSetAttribute(k_eval_events, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if(e == cudaSuccess) e = cudaFuncSetAttribute(k_grid_sum, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if(e != cudaSuccess) return bail(std::string("cannot reserve ") + std::to_string(sm) + " B of shared memory for the program block: " + cudaGetErrorString(e));
    }
    // commit
    cudaFree(c->d_nodes); cudaFree(c->d_params); cudaFree(c->d_consts); cudaFree(c->d_ints); // d_norms lives inside d_params' block
    c->d_nodes = t_nodes; c->d_params = t_params; c->d_consts = t_consts; c->d_ints = t_ints;
    c->d_norms = c->d_params + n_params;
    if(t_pin) {
        if(c->h_pin) cudaFreeHost(c->h_pin);
        c->h_pin = t_pin;
        c->h_pin_cap = need;
    }
    c->h_nodes.assign(nodes, nodes + n_nodes);
    c->h_consts.assign(consts, consts + n_consts);
    c->h_ints.assign(ints, ints + n_ints);
    c->n_params = n_params;
    c->n_norms = n_norms;
    gf_dev_view &V = c->view;
    V.nodes = c->d_nodes; V.params = c->d_params; V.consts = c->d_consts; V.norms = c->d_norms; V.ints = c->d_ints;
    V.n_nodes = n_nodes; V.n_params = n_params; V.n_consts = n_consts; V.n_norms = n_norms; V.n_ints = n_ints;
    return 0;
}

extern "C" int gf_comm_unique_id(void *id128) {
    if(load_nccl()) return 1;
    int rc = p_uid(id128);
    return rc ? nccl_fail("ncclGetUniqueId", rc) : 0;
}
// small all-reduce of n <= GF_PEER_SLOT doubles that already sit in device memory (one block)
__global__ void __launch_bounds__(GF_THREADS) k_peer_allreduce(double *vals, int n, const gf_peer_view PV) {
    __shared__ double in[GF_PEER_SLOT];
    for(int k = threadIdx.x; k < n; k += blockDim.x) in[k] = vals[k];
    __syncthreads();
    gf_peer_allreduce(PV, in, n, vals);
}

// Map every rank's exchange buffer into this process (cudaIpc over NVLink / NVSwitch). Collective: every rank calls it with
// the communicator attached. Returns 0 and leaves c->peer_on = false when any rank could not map its peers (all ranks then
// agree to stay on NCCL); fails only on NCCL errors.
static int peer_setup(gf_ctx *c) {
    c->peer_on = false;
    const char *off = std::getenv("GOOFIT_B200_NO_PEER");
    int ok = (off && off[0] == '1') || c->nranks > GF_MAX_PEERS || !p_allgather ? 0 : 1;
    const size_t slots_bytes = sizeof(double) * 2 * GF_MAX_PEERS * GF_PEER_SLOT, flags_bytes = sizeof(unsigned long long) * 2 * GF_MAX_PEERS;
    cudaIpcMemHandle_t mine;
    std::memset(&mine, 0, sizeof mine);
    if(ok) {
        if(!c->h_err) {
            if(cudaHostAlloc(&c->h_err, sizeof(int), cudaHostAllocMapped) != cudaSuccess || cudaHostGetDevicePointer((void **)&c->h_err_dev, c->h_err, 0) != cudaSuccess) ok = 0;
            else *c->h_err = 0;
        }
        if(ok && (cudaMalloc(&c->peer_buf, slots_bytes + flags_bytes) != cudaSuccess || cudaMemset(c->peer_buf, 0, slots_bytes + flags_bytes) != cudaSuccess
                  || cudaDeviceSynchronize() != cudaSuccess || cudaIpcGetMemHandle(&mine, c->peer_buf) != cudaSuccess))
            ok = 0;
        cudaGetLastError();
    }
    // exchange {ok, handle} of every rank through the communicator that was just created
    struct rec { long long ok; cudaIpcMemHandle_t h; };
    rec my{ok, mine};
    rec *d_all = nullptr;
    std::vector<rec> all(c->nranks);
    GF_CUDA(cudaMalloc(&d_all, sizeof(rec) * c->nranks));
    GF_CUDA(cudaMemcpy(d_all + c->rank, &my, sizeof(rec), cudaMemcpyHostToDevice));
    if(p_allgather) {
        int rc = p_allgather(d_all + c->rank, d_all, sizeof(rec), /*ncclChar*/ 0, c->nccl_comm, c->stream);
        if(rc) { cudaFree(d_all); return nccl_fail("ncclAllGather", rc); }
        GF_CUDA(cudaStreamSynchronize(c->stream));
        GF_CUDA(cudaMemcpy(all.data(), d_all, sizeof(rec) * c->nranks, cudaMemcpyDeviceToHost));
    } else
        all[c->rank] = my;
    for(auto &r : all) ok = ok && r.ok;
    if(ok) {
        for(int r = 0; r < c->nranks && ok; r++) {
            void *base = c->peer_buf;
            if(r != c->rank) {
                if(cudaIpcOpenMemHandle(&base, all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cudaGetLastError(); break; }
                c->peer_mapped[r] = base;
            }
            c->peer.slots[r] = static_cast<double *>(base);
            c->peer.flags[r] = reinterpret_cast<unsigned long long *>(static_cast<char *>(base) + slots_bytes);
        }
    }
    // second round: did every rank manage to map every peer?
    double *d_flag = reinterpret_cast<double *>(d_all);
    double v = ok ? 0.0 : 1.0;
    GF_CUDA(cudaMemcpy(d_flag, &v, sizeof v, cudaMemcpyHostToDevice));
    int rc = p_allreduce(d_flag, d_flag, 1, /*ncclDouble*/ 8, /*ncclSum*/ 0, c->nccl_comm, c->stream); // also NCCL's warm-up: channels are set up here, not in the first likelihood call
    if(rc) { cudaFree(d_all); return nccl_fail("ncclAllReduce", rc); }
    GF_CUDA(cudaStreamSynchronize(c->stream));
    GF_CUDA(cudaMemcpy(&v, d_flag, sizeof v, cudaMemcpyDeviceToHost));
    cudaFree(d_all);
    if(v != 0.0) {
        peer_teardown(c);
        return 0;
    }
    c->peer.rank = c->rank;
    c->peer.nranks = c->nranks;
    c->peer.seq = 0;
    c->peer.err = c->h_err_dev;
    c->peer_on = true;
    // one exchange now, so that the first likelihood call finds every mapping touched
    gf_peer_view PV = gf_next_peer_view(c);
    GF_CUDA(cudaMemcpy(c->d_result + GF_NPARTIAL * 8, &v, sizeof v, cudaMemcpyHostToDevice));
    k_peer_allreduce<<<1, GF_THREADS, 0, c->stream>>>(c->d_result + GF_NPARTIAL * 8, 1, PV);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    if(*c->h_err) {
        peer_teardown(c);
        return gf_fail("gf_comm_init: peer-memory warm-up exchange timed out");
    }
    return 0;
}
gf_peer_view gf_next_peer_view(gf_ctx *c) {
    if(!gf_peer_active(c)) {
        gf_peer_view none{};
        none.nranks = 1;
        return none;
    }
    c->peer.seq++;
    c->peer.done = nullptr;
    return c->peer;
}

extern "C" int gf_comm_init(gf_ctx *c, const void *id128, int rank, int nranks) {
    if(!c) return gf_fail("null ctx");
    if(load_nccl()) return 1;
    if(nranks < 1 || rank < 0 || rank >= nranks) return gf_fail("gf_comm_init: bad rank");
    cudaSetDevice(c->device);
    if(c->nccl_comm) gf_comm_destroy(c);
    gf_nccl_uid uid;
    std::memcpy(&uid, id128, sizeof uid);
    void *comm = nullptr;
    int rc = p_init(&comm, nranks, uid, rank);
    if(rc) return nccl_fail("ncclCommInitRank", rc);
    c->nccl_comm = comm;
    c->rank = rank;
    c->nranks = nranks;
    return nranks > 1 ? peer_setup(c) : 0;
}
extern "C" int gf_comm_destroy(gf_ctx *c) {
    if(!c) return gf_fail("null ctx");
    if(c->nccl_comm) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
        peer_teardown(c);
        p_destroy(c->nccl_comm);
        c->nccl_comm = nullptr;
    }
    c->nranks = 1;
    c->rank = 0;
    return 0;
}
extern "C" const char *gf_comm_path(gf_ctx *c) {
    if(!c || !c->nccl_comm || c->nranks <= 1) return "none";
    return c->peer_on ? "peer-memory exchange over NVLink inside the reduction kernel (cudaIpc-mapped slots, rank-ordered sum)"
                      : "in-stream ncclAllReduce";
}
extern "C" int gf_comm_attached(gf_ctx *c) { return c && c->nccl_comm ? c->nranks : 0; }
int gf_allreduce_sum(gf_ctx *c, double *dev, int n) {
    if(!c->nccl_comm) return 0;
    if(gf_peer_active(c) && n <= GF_PEER_SLOT) {
        k_peer_allreduce<<<1, GF_THREADS, 0, c->stream>>>(dev, n, gf_next_peer_view(c));
        c->launches++;
        GF_CUDA(cudaGetLastError());
        return 0;
    }
    int rc = p_allreduce(dev, dev, (size_t)n, /*ncclDouble*/ 8, /*ncclSum*/ 0, c->nccl_comm, c->stream);
    return rc ? nccl_fail("ncclAllReduce", rc) : 0;
}

// Per-event results of a sharded evaluation, gathered on every rank in rank order: the counterpart of the reference's
// MPI_Allgather(v) of evaluate_with_metric (GooPdf.cu:139-172). counts[r] = events of rank r (the caller knows the split).
extern "C" int gf_allgather_events(gf_ctx *c, const double *local, const size_t *counts, double *all) {
    if(!c || !counts || !all) return gf_fail("gf_allgather_events: null argument");
    if(!c->nccl_comm || c->nranks <= 1) {
        if(counts[0] && local) std::memcpy(all, local, sizeof(double) * counts[0]);
        return 0;
    }
    if(!p_allgather) return gf_fail("gf_allgather_events: libnccl has no ncclAllGather");
    cudaSetDevice(c->device);
    size_t mx = 0, total = 0;
    for(int r = 0; r < c->nranks; r++) { mx = std::max(mx, counts[r]); total += counts[r]; }
    if(mx == 0) return 0;
    double *d = nullptr; // [nranks + 1][mx]: the gathered blocks, then this rank's padded block
    GF_CUDA(cudaMalloc(&d, sizeof(double) * mx * (size_t)(c->nranks + 1)));
    double *d_send = d + mx * (size_t)c->nranks;
    cudaError_t e = cudaMemsetAsync(d_send, 0, sizeof(double) * mx, c->stream);
    if(e == cudaSuccess && counts[c->rank]) e = cudaMemcpyAsync(d_send, local, sizeof(double) * counts[c->rank], cudaMemcpyHostToDevice, c->stream);
    int rc = e == cudaSuccess ? p_allgather(d_send, d, mx, /*ncclDouble*/ 8, c->nccl_comm, c->stream) : 0;
    size_t off = 0;
    for(int r = 0; r < c->nranks && e == cudaSuccess && !rc; r++) {
        if(counts[r]) e = cudaMemcpyAsync(all + off, d + mx * (size_t)r, sizeof(double) * counts[r], cudaMemcpyDeviceToHost, c->stream);
        off += counts[r];
    }
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if(rc) return nccl_fail("ncclAllGather", rc);
    if(e != cudaSuccess) return gf_fail(std::string("gf_allgather_events: ") + cudaGetErrorString(e));
    return 0;
}

extern "C" int gf_timer_start(gf_ctx *c) {
    if(!c) return gf_fail("null ctx");
    cudaSetDevice(c->device);
    if(!c->t0) { cudaEventCreate(&c->t0); cudaEventCreate(&c->t1); }
    GF_CUDA(cudaEventRecord(c->t0, c->stream));
    return 0;
}
extern "C" int gf_timer_stop(gf_ctx *c, float *ms) {
    if(!c || !c->t0) return gf_fail("gf_timer_stop: timer not started");
    cudaSetDevice(c->device);
    GF_CUDA(cudaEventRecord(c->t1, c->stream));
    GF_CUDA(cudaEventSynchronize(c->t1));
    GF_CUDA(cudaEventElapsedTime(ms, c->t0, c->t1));
    return 0;
}
// ---- FP64 peak micro-benchmark: the denominator of the FP64-bound kernels' roofline -------------------------------
// 8 independent DFMA chains per thread, 1024 threads per SM-resident block pair; flops = 2 per DFMA.
__global__ void __launch_bounds__(256) k_dfma_peak(double *out, int iters, double a, double b) {
    double x0 = threadIdx.x * 1e-9, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for(int i = 0; i < iters; ++i) {
#pragma unroll
        for(int u = 0; u < 8; ++u) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if(s == 123.456) out[0] = s; // never true; keeps the chains alive
}
extern "C" int gf_fp64_peak(gf_ctx *c, double *tflops) {
    if(!c || !tflops) return gf_fail("gf_fp64_peak: null argument");
    cudaSetDevice(c->device);
    const int blocks = c->sm_count * 8, iters = 4096;
    cudaEvent_t e0, e1;
    GF_CUDA(cudaEventCreate(&e0));
    GF_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for(int rep = 0; rep < 6; ++rep) {
        GF_CUDA(cudaEventRecord(e0, c->stream));
        k_dfma_peak<<<blocks, 256, 0, c->stream>>>(c->d_result + GF_NPARTIAL - 1, iters, 0.999999, 1e-9);
        GF_CUDA(cudaEventRecord(e1, c->stream));
        GF_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        GF_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        double tf = 2.0 * 64.0 * iters * 256.0 * blocks / (ms * 1e-3) / 1e12;
        if(rep > 0 && tf > best) best = tf;
    }
    c->launches += 6;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return 0;
}

extern "C" int gf_set_params_async(gf_ctx *c, const double *params, int n, const double *norms, int n_norms) {
    if(!c) return gf_fail("null ctx");
    if(n != c->n_params) return gf_fail("gf_set_params_async: size mismatch");
    if(norms && n_norms != c->n_norms) return gf_fail("gf_set_params_async: norms size mismatch");
    cudaSetDevice(c->device);
    size_t need = sizeof(double) * (size_t)(c->n_params + c->n_norms + 1);
    if(need > c->ring_cap) {
        GF_CUDA(cudaStreamSynchronize(c->stream));
        for(int i = 0; i < gf_ctx::RING; i++) {
            if(c->ring_buf[i]) cudaFreeHost(c->ring_buf[i]);
            GF_CUDA(cudaMallocHost(&c->ring_buf[i], need));
            if(!c->ring_ev[i]) cudaEventCreateWithFlags(&c->ring_ev[i], cudaEventDisableTiming);
        }
        c->ring_cap = need;
    }
    int k = c->ring_pos;
    c->ring_pos = (k + 1) % gf_ctx::RING;
    cudaEventSynchronize(c->ring_ev[k]); // slot free again? (returns at once for a never-recorded event)
    if(n) std::memcpy(c->ring_buf[k], params, sizeof(double) * n);
    if(norms) std::memcpy(c->ring_buf[k] + n, norms, sizeof(double) * n_norms);
    if(n && norms) // one copy: d_norms directly follows d_params
        GF_CUDA(cudaMemcpyAsync(c->d_params, c->ring_buf[k], sizeof(double) * (n + n_norms), cudaMemcpyHostToDevice, c->stream));
    else if(n)
        GF_CUDA(cudaMemcpyAsync(c->d_params, c->ring_buf[k], sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    else if(norms)
        GF_CUDA(cudaMemcpyAsync(c->d_norms, c->ring_buf[k] + n, sizeof(double) * n_norms, cudaMemcpyHostToDevice, c->stream));
    GF_CUDA(cudaEventRecord(c->ring_ev[k], c->stream));
    return 0;
}

extern "C" int gf_reset_aux(gf_ctx *c) {
    if(!c) return gf_fail("null ctx");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    gf_amp3_free(c);
    gf_amp4_free(c);
    gf_conv_free(c);
    for(int i = 0; i < GF_MAX_AUX; i++) {
        c->view.aux[i]    = nullptr;
        c->view.aux_ld[i] = 0;
    }
    return 0;
}

extern "C" int gf_set_params(gf_ctx *c, const double *params, int n) {
    if(!c) return gf_fail("null ctx");
    if(n != c->n_params) return gf_fail("gf_set_params: expected " + std::to_string(c->n_params) + " got " + std::to_string(n));
    if(n == 0) return 0;
    cudaSetDevice(c->device);
    // the pinned block may still be in flight from the previous call
    GF_CUDA(cudaStreamSynchronize(c->stream));
    std::memcpy(c->h_pin, params, sizeof(double) * n);
    GF_CUDA(cudaMemcpyAsync(c->d_params, c->h_pin, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    return 0;
}

extern "C" int gf_set_norms(gf_ctx *c, const double *norms, int n) {
    if(!c) return gf_fail("null ctx");
    if(n != c->n_norms) return gf_fail("gf_set_norms: size mismatch");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    double *dst = c->h_pin + c->n_params;
    std::memcpy(dst, norms, sizeof(double) * n);
    GF_CUDA(cudaMemcpyAsync(c->d_norms, dst, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    return 0;
}

static int finish_columns(gf_ctx *c, int n_cols, size_t n_events, double total_events) {
    gf_dev_view &V = c->view;
    V.n_events = (long)n_events;
    V.total_events = total_events;
    c->n_cols = n_cols;
    V.identity_mask = 0;
    if(n_events > 0) {
        int *d_flag = nullptr;
        GF_CUDA(cudaMalloc(&d_flag, sizeof(int) * GF_MAX_OBS));
        std::vector<int> ones(GF_MAX_OBS, 1);
        GF_CUDA(cudaMemcpyAsync(d_flag, ones.data(), sizeof(int) * GF_MAX_OBS, cudaMemcpyHostToDevice, c->stream));
        int blocks = (int)((n_events + 255) / 256);
        for(int k = 0; k < n_cols; k++) {
            k_check_identity<<<blocks, 256, 0, c->stream>>>(V.cols[k], (long)n_events, d_flag + k);
            c->launches++;
        }
        std::vector<int> flags(GF_MAX_OBS, 0);
        GF_CUDA(cudaMemcpyAsync(flags.data(), d_flag, sizeof(int) * GF_MAX_OBS, cudaMemcpyDeviceToHost, c->stream));
        GF_CUDA(cudaStreamSynchronize(c->stream));
        cudaFree(d_flag);
        for(int k = 0; k < n_cols; k++)
            if(flags[k]) V.identity_mask |= (1u << k);
    }
    return 0;
}

extern "C" int gf_upload_columns(gf_ctx *c, const double *const *cols, int n_cols, size_t n_events, double total_events) {
    if(!c) return gf_fail("null ctx");
    if(n_cols < 0 || n_cols > GF_MAX_OBS) return gf_fail("gf_upload_columns: 0..16 columns");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    free_cols(c);
    c->cols_owned = true;
    size_t ld = (n_events + 31) & ~size_t(31); // 256-B aligned columns
    if(ld == 0) ld = 32;
    c->view.ld = (long)ld;
    GF_CUDA(cudaMalloc(&c->d_cols_base, sizeof(double) * ld * std::max(1, n_cols)));
    // pinned staging in 32 MiB chunks so pageable user memory streams at full PCIe/C2C rate
    const size_t chunk = size_t(4) << 20; // doubles
    double *stage[2] = {nullptr, nullptr};
    cudaEvent_t done[2];
    if(n_events > 0) {
        for(int b = 0; b < 2; b++) {
            GF_CUDA(cudaMallocHost(&stage[b], sizeof(double) * std::min(chunk, n_events)));
            cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming);
        }
    }
    int buf = 0;
    for(int k = 0; k < n_cols; k++) {
        c->view.cols[k] = c->d_cols_base + (size_t)k * ld;
        for(size_t off = 0; off < n_events; off += chunk) {
            size_t n = std::min(chunk, n_events - off);
            cudaEventSynchronize(done[buf]);
            std::memcpy(stage[buf], cols[k] + off, sizeof(double) * n);
            GF_CUDA(cudaMemcpyAsync(c->d_cols_base + (size_t)k * ld + off, stage[buf], sizeof(double) * n,
                                    cudaMemcpyHostToDevice, c->stream));
            cudaEventRecord(done[buf], c->stream);
            buf ^= 1;
        }
    }
    GF_CUDA(cudaStreamSynchronize(c->stream));
    if(n_events > 0)
        for(int b = 0; b < 2; b++) {
            cudaFreeHost(stage[b]);
            cudaEventDestroy(done[b]);
        }
    return finish_columns(c, n_cols, n_events, total_events);
}

extern "C" int gf_adopt_device_columns(gf_ctx *c, const double *const *dev_cols, int n_cols, size_t n_events,
                                       double total_events) {
    if(!c) return gf_fail("null ctx");
    if(n_cols < 0 || n_cols > GF_MAX_OBS) return gf_fail("gf_adopt_device_columns: 0..16 columns");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    free_cols(c);
    c->cols_owned = false;
    c->view.ld = (long)n_events;
    for(int k = 0; k < n_cols; k++) c->view.cols[k] = dev_cols[k];
    return finish_columns(c, n_cols, n_events, total_events);
}

static int check_root(gf_ctx *c, int root) {
    if(!c) return gf_fail("null ctx");
    if(c->h_nodes.empty()) return gf_fail("no program set");
    if(root < 0 || root >= (int)c->h_nodes.size()) return gf_fail("bad root node");
    return 0;
}

static int eval_sum_launch(gf_ctx *c, int root, int metric, double *out_dev, bool signal_done = false) {
    if(check_root(c, root)) return 1;
    cudaSetDevice(c->device);
    int first = root - c->h_nodes[root].size + 1;
    const int has = program_class(c, first, root);
    int blocks = gf_grid_blocks(c, c->view.n_events, eval_sum_blocks_per_sm(has));
    gf_peer_view PV = gf_next_peer_view(c);
    if(signal_done) {
        PV.done = c->h_done_dev;
        PV.done_seq = ++c->done_seq;
    }
    const bool prof = c->profiling && c->prof_n < gf_ctx::PROF;
    if(prof) cudaEventRecord(c->prof_ev[2 * c->prof_n], c->stream);
    const char *which = "k_eval_sum";
    const int fast = gf_fast_launch(c, first, root, metric, out_dev, PV, &which);
    if(fast < 0) return gf_fail(std::string("likelihood kernel launch failed: ") + cudaGetErrorString(cudaGetLastError()));
    if(fast == 0)
        eval_sum_kernel(has)<<<blocks, GF_THREADS, gf_prog_smem(c), c->stream>>>(
            c->view, first, root, metric, c->n_cols, c->d_partials, c->d_ticket, out_dev, PV);
    c->last_kernel = which;
    if(prof) cudaEventRecord(c->prof_ev[2 * c->prof_n++ + 1], c->stream);
    c->launches++;
    GF_CUDA(cudaGetLastError());
    return 0;
}
extern "C" int gf_eval_sum_async(gf_ctx *c, int root, int metric) {
    if(eval_sum_launch(c, root, metric, c ? c->d_result : nullptr)) return 1;
    if(c->nccl_comm && !gf_peer_active(c) && gf_allreduce_sum(c, c->d_result, 1)) return 1; // peer path: already summed in the kernel
    return 0;
}

// Wait for the completion flag the last block writes behind the result: the host polls one pinned cache line (a few hundred
// ns of latency) instead of a stream synchronisation. The stream is queried now and then so that a faulted kernel cannot
// hang the caller.
static int wait_done(gf_ctx *c) {
    volatile unsigned long long *f = c->h_done;
    const unsigned long long want = c->done_seq;
    for(unsigned spin = 1;; ++spin) {
        if(*f == want) return 0;
        if((spin & 0x3fff) == 0) {
            cudaError_t e = cudaStreamQuery(c->stream);
            if(e == cudaSuccess) { // everything retired: the flag is visible by now (or the kernel never wrote it)
                if(*f == want) return 0;
                return gf_fail("gf_eval_sum: kernel finished without signalling completion");
            }
            if(e != cudaErrorNotReady) return gf_fail(std::string("gf_eval_sum: ") + cudaGetErrorString(e));
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
}

extern "C" int gf_eval_sum(gf_ctx *c, int root, int metric, double *out) {
    if(c && (!c->nccl_comm || gf_peer_active(c)) && c->h_result_dev && c->h_done_dev) {
        // one GPU, or several with the peer-memory exchange: the last block of the reduction writes the (all-rank) sum
        // straight into mapped pinned host memory and then a completion flag, so the call is one launch and a poll of one
        // cache line: no collective kernel, no device-to-host copy, no stream synchronisation
        if(eval_sum_launch(c, root, metric, c->h_result_dev, true)) return 1;
        if(wait_done(c)) return 1;
    } else {
        if(gf_eval_sum_async(c, root, metric)) return 1;
        GF_CUDA(cudaMemcpyAsync(c->h_result, c->d_result, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        GF_CUDA(cudaStreamSynchronize(c->stream));
    }
    if(c->h_err && *c->h_err) return gf_fail("gf_eval_sum: a rank did not join the peer-memory exchange within the time limit");
    *out = c->h_result[0];
    return 0;
}

// opt-in per-kernel timing of the likelihood kernel: CUDA events around each of the next (up to 256) launches
extern "C" int gf_profile_begin(gf_ctx *c) {
    if(!c) return gf_fail("null ctx");
    cudaSetDevice(c->device);
    for(auto &e : c->prof_ev)
        if(!e) GF_CUDA(cudaEventCreate(&e));
    c->prof_n = 0;
    c->profiling = true;
    return 0;
}
// average duration (ms) of the launches recorded since gf_profile_begin; *n_launches = how many
extern "C" int gf_profile_end(gf_ctx *c, float *avg_ms, int *n_launches) {
    if(!c || !avg_ms) return gf_fail("gf_profile_end: null argument");
    cudaSetDevice(c->device);
    c->profiling = false;
    GF_CUDA(cudaStreamSynchronize(c->stream));
    double tot = 0;
    for(int i = 0; i < c->prof_n; i++) {
        float ms = 0;
        GF_CUDA(cudaEventElapsedTime(&ms, c->prof_ev[2 * i], c->prof_ev[2 * i + 1]));
        tot += ms;
    }
    *avg_ms = c->prof_n ? (float)(tot / c->prof_n) : 0.f;
    c->last_ms = *avg_ms;
    if(n_launches) *n_launches = c->prof_n;
    return 0;
}

extern "C" double *gf_partial_device_ptr(gf_ctx *c) { return c ? c->d_result : nullptr; }
extern "C" void *gf_stream(gf_ctx *c) { return c ? (void *)c->stream : nullptr; }
extern "C" int gf_sync(gf_ctx *c) {
    if(!c) return gf_fail("null ctx");
    cudaSetDevice(c->device);
    GF_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
extern "C" float gf_last_kernel_ms(gf_ctx *c) { return c ? c->last_ms : 0.f; }
extern "C" const char *gf_last_kernel_name(gf_ctx *c) { return c ? c->last_kernel : ""; }
extern "C" uint64_t gf_launch_count(gf_ctx *c) { return c ? c->launches : 0; }

extern "C" int gf_eval_events(gf_ctx *c, int root, int metric, double *out_host, size_t n) {
    if(check_root(c, root)) return 1;
    if((long)n != c->view.n_events) return gf_fail("gf_eval_events: n != uploaded events");
    if(n == 0) return 0;
    cudaSetDevice(c->device);
    if(n > c->events_out_cap) {
        cudaFree(c->d_events_out);
        GF_CUDA(cudaMalloc(&c->d_events_out, sizeof(double) * n));
        c->events_out_cap = n;
    }
    int first = root - c->h_nodes[root].size + 1;
    int blocks = gf_grid_blocks(c, c->view.n_events, 4);
    k_eval_events<<<blocks, GF_THREADS, gf_prog_smem(c), c->stream>>>(c->view, first, root, metric, c->n_cols,
                                                                     c->d_events_out);
    c->launches++;
    GF_CUDA(cudaGetLastError());
    GF_CUDA(cudaMemcpyAsync(out_host, c->d_events_out, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int gf_integrate_grid(gf_ctx *c, int root, const double *lo_hi_n, int n_obs, double *out) {
    if(check_root(c, root)) return 1;
    if(n_obs < 1 || n_obs > GF_MAX_OBS) return gf_fail("gf_integrate_grid: bad n_obs");
    const gf_node &nd = c->h_nodes[root];
    if(nd.nobs < 1) return gf_fail("gf_integrate_grid: root has no observables");
    cudaSetDevice(c->device);
    gf_grid_spec G;
    G.nobs = n_obs;
    G.total = 1;
    for(int d = 0; d < GF_MAX_OBS; d++) { G.lo[d] = G.hi[d] = 0; G.nb[d] = 1; G.col[d] = 0; }
    for(int d = 0; d < n_obs; d++) {
        G.lo[d] = lo_hi_n[3 * d];
        G.hi[d] = lo_hi_n[3 * d + 1];
        G.nb[d] = (int)std::floor(lo_hi_n[3 * d + 2] + 0.5);
        if(G.nb[d] < 1) return gf_fail("gf_integrate_grid: bins < 1");
        G.col[d] = c->h_ints[nd.obs_off + std::min(d, nd.nobs - 1)];
        G.total *= G.nb[d];
    }
    int first = root - nd.size + 1;
    int blocks = gf_grid_blocks(c, G.total, 4);
    k_grid_sum<<<blocks, GF_THREADS, gf_prog_smem(c), c->stream>>>(c->view, first, root, G, c->d_partials + (size_t)c->max_blocks * 6,
                                                                  c->d_ticket + 1, c->d_result + GF_NPARTIAL);
    c->launches++;
    GF_CUDA(cudaGetLastError());
    GF_CUDA(cudaMemcpyAsync(c->h_result + GF_NPARTIAL, c->d_result + GF_NPARTIAL, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    GF_CUDA(cudaStreamSynchronize(c->stream));
    *out = c->h_result[GF_NPARTIAL];
    return 0;
}
