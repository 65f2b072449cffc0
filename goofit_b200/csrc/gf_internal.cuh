// gf_internal.cuh — context, shared-memory program block and the per-event stack-machine evaluator.
// Replaces the reference's device_function_table indirect dispatch + ParameterContainer cursor
// (/root/reference/src/PDFs/GooPdf.cu:398-400, include/goofit/PDFs/ParameterContainer.h:9-47).
#pragma once
#include "../../include/goofit_b200.h"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#define GF_THREADS 256
#define GF_MAX_AUX 16

struct gf_amp3_state;
struct gf_amp4_state;
struct gf_conv_state;

// Device-visible view of everything a kernel needs; passed by value as a kernel parameter.
struct gf_dev_view {
    const gf_node *nodes;   // global
    const double *params;
    const double *consts;
    const double *norms;
    const int32_t *ints;
    int n_nodes, n_params, n_consts, n_norms, n_ints;
    const double *cols[GF_MAX_OBS]; // SoA event columns
    unsigned identity_mask;         // bit c set: column c holds 0,1,2,... (event number == row index)
    long n_events;
    long ld;                        // column leading dimension (padded rows)
    double total_events;
    // per-node auxiliary device tables (cached waves / amplitudes / convolution tables)
    const void *aux[GF_MAX_AUX];
    long aux_ld[GF_MAX_AUX];
};

// Program block staged in shared memory by every block.
struct gf_sprog {
    const gf_node *nodes;
    const double *params;
    const double *consts;
    const double *norms;
    const int32_t *ints;
    const double *const *cols; // staged copies of the view's pointer tables (dynamic indexing from shared)
    const void *const *aux;
    const long *aux_ld;
};

__host__ __device__ inline size_t gf_align16(size_t x) { return (x + 15) & ~size_t(15); }

__host__ __device__ inline size_t gf_sprog_bytes(int n_nodes, int n_params, int n_consts, int n_norms, int n_ints) {
    return gf_align16(sizeof(gf_node) * n_nodes) + gf_align16(8 * (size_t)n_params) + gf_align16(8 * (size_t)n_consts)
           + gf_align16(8 * (size_t)n_norms) + gf_align16(4 * (size_t)n_ints) + 8 * (GF_MAX_OBS + 2 * GF_MAX_AUX) + 64;
}

// cooperative copy global -> shared; call from all threads, then __syncthreads()
__device__ inline gf_sprog gf_stage_program(const gf_dev_view &V, unsigned char *smem) {
    gf_sprog P;
    unsigned char *p = smem;
    gf_node *sn = reinterpret_cast<gf_node *>(p);
    p += gf_align16(sizeof(gf_node) * V.n_nodes);
    double *sp = reinterpret_cast<double *>(p);
    p += gf_align16(8 * (size_t)V.n_params);
    double *sc = reinterpret_cast<double *>(p);
    p += gf_align16(8 * (size_t)V.n_consts);
    double *sno = reinterpret_cast<double *>(p);
    p += gf_align16(8 * (size_t)V.n_norms);
    int32_t *si = reinterpret_cast<int32_t *>(p);
    p += gf_align16(4 * (size_t)V.n_ints);
    const double **scol = reinterpret_cast<const double **>(p);
    p += 8 * GF_MAX_OBS;
    const void **saux = reinterpret_cast<const void **>(p);
    p += 8 * GF_MAX_AUX;
    long *sld = reinterpret_cast<long *>(p);
    const int t = threadIdx.x, nt = blockDim.x;
    {
        const int32_t *src = reinterpret_cast<const int32_t *>(V.nodes);
        int32_t *dst = reinterpret_cast<int32_t *>(sn);
        int n = V.n_nodes * (int)(sizeof(gf_node) / 4);
        for(int i = t; i < n; i += nt) dst[i] = src[i];
    }
    for(int i = t; i < V.n_params; i += nt) sp[i] = V.params[i];
    for(int i = t; i < V.n_consts; i += nt) sc[i] = V.consts[i];
    for(int i = t; i < V.n_norms; i += nt) sno[i] = V.norms[i];
    for(int i = t; i < V.n_ints; i += nt) si[i] = V.ints[i];
    if(t < GF_MAX_OBS) scol[t] = V.cols[t];
    if(t < GF_MAX_AUX) { saux[t] = V.aux[t]; sld[t] = V.aux_ld[t]; }
    P.cols = scol; P.aux = saux; P.aux_ld = sld;
    P.nodes = sn; P.params = sp; P.consts = sc; P.norms = sno; P.ints = si;
    return P;
}

// ---- event accessors -------------------------------------------------------------------------
struct gf_evt_soa { // coalesced column reads: consecutive threads -> consecutive rows
    const double *const *cols;
    unsigned identity_mask;
    long i;
    __device__ inline double operator()(int col) const { return __ldg(cols[col] + i); }
    __device__ inline long event_number(int col) const {
        if((identity_mask >> col) & 1u) return i;
        return (long)floor(0.5 + __ldg(cols[col] + i)); // Amp3Body.cu:85-87
    }
};
struct gf_evt_vals { // grid point / single point held in registers
    double v[GF_MAX_OBS];
    long i;
    __device__ inline double operator()(int col) const { return v[col]; }
    __device__ inline long event_number(int col) const { return i; }
};

#include "gf_ops.cuh"

// ---- peer-memory exchange of partial sums (multi-GPU, one process per GPU) ---------------------------------------------
// Replaces MPI_Allreduce of the partial NLL (reference GooPdf.cu:72-75) and the in-stream NCCL all-reduce of round 1. Every
// rank owns an exchange buffer that all peers map over NVLink (cudaIpc). The LAST block of a reduction kernel
//   1. stores this rank's n partial sums into slot [parity][rank] of EVERY rank's buffer (remote stores over NVLink),
//   2. releases a sequence-number flag next to them (fence.sys; st.release.sys),
//   3. waits until all ranks' flags of this sequence number have arrived in its OWN buffer (ld.acquire.sys, local memory),
//   4. adds the n_ranks slots in rank order -> the same bits on every rank, run to run, for a fixed number of ranks.
// One kernel launch, no collective kernel behind it, the result written where the caller wants it (mapped pinned host
// memory on the likelihood path). Two parities suffice: a rank cannot start exchange s+2 before every rank finished s+1,
// which needs every rank to have left exchange s (all exchanges of a context are ordered on its one stream).
#define GF_MAX_PEERS 16
#define GF_PEER_SLOT 512 // doubles per rank and exchange (the largest payload: the six 8x8 tiles of a 16-amplitude Gram matrix, 384)
struct gf_peer_view {
    double *slots[GF_MAX_PEERS];             // rank r's buffer as mapped here: [2][n_ranks][GF_PEER_SLOT]
    unsigned long long *flags[GF_MAX_PEERS]; // rank r's flags: [2][n_ranks]
    int rank, nranks;
    unsigned long long seq; // sequence number of this exchange (host-incremented per launch), starts at 1
    int *err;               // mapped host flag: set when a peer did not show up in time
    // completion flag in mapped pinned host memory (optional): written with `done_seq` after the results, so that the host
    // can wait for a likelihood value by polling one cache line instead of paying a stream synchronisation
    unsigned long long *done;
    unsigned long long done_seq;
};
// results (written by threads 0..n-1 of this block before the call) -> system-visible, then the completion flag
__device__ inline void gf_signal_done(const gf_peer_view &PV) {
    if(PV.done == nullptr) return;
    __threadfence_system();
    __syncthreads();
    if(threadIdx.x == 0) {
        *reinterpret_cast<volatile unsigned long long *>(PV.done) = PV.done_seq;
        __threadfence_system();
    }
}
__device__ inline unsigned long long gf_ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ inline void gf_st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ inline double gf_ld_relaxed_sys(const double *p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
// Called by ALL threads of ONE block. vals[n] must be visible to the whole block (shared or global memory written before a
// __syncthreads). out[k] receives the sum over ranks of vals[k] (written by thread k mod blockDim.x).
__device__ inline void gf_peer_allreduce(const gf_peer_view &PV, const double *vals, int n, double *out) {
    const int t = threadIdx.x, nr = PV.nranks, par = (int)(PV.seq & 1ull);
    __shared__ int timed_out;
    if(t == 0) timed_out = 0;
    for(int i = t; i < n * nr; i += blockDim.x) {
        const int peer = i / n, k = i - peer * n;
        PV.slots[peer][((size_t)par * nr + PV.rank) * GF_PEER_SLOT + k] = vals[k];
    }
    __threadfence_system();
    __syncthreads();
    if(t < nr) gf_st_release_sys(PV.flags[t] + (size_t)par * nr + PV.rank, PV.seq);
    if(t < nr) {
        const unsigned long long *f = PV.flags[PV.rank] + (size_t)par * nr + t;
        const long long t0 = clock64();
        while(gf_ld_acquire_sys(f) < PV.seq) {
            if(clock64() - t0 > 40000000000ll) { // ~20 s at 2 GHz: a rank is missing; fail loudly instead of hanging the GPU
                *PV.err = 1;
                timed_out = 1;
                break;
            }
        }
    }
    __syncthreads();
    for(int k = t; k < n; k += blockDim.x) {
        const double *src = PV.slots[PV.rank] + (size_t)par * nr * GF_PEER_SLOT + k;
        double s = 0;
        for(int r = 0; r < nr; ++r) s += gf_ld_relaxed_sys(src + (size_t)r * GF_PEER_SLOT);
        if(timed_out) s = __longlong_as_double(0x7ff8000000000000ll);
        out[k] = s;
    }
}

// ---- deterministic reduction -----------------------------------------------------------------
// warp shuffle tree -> block tree (shared) -> per-block partial; the last block to arrive sums
// the partials in index order with the same tree. Fixed grid => bit-reproducible run to run.
__device__ inline double gf_warp_sum(double v) {
#pragma unroll
    for(int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ inline double gf_block_sum(double v, double *sh /* >= 32 doubles */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = gf_warp_sum(v);
    __syncthreads();
    if(lane == 0) sh[w] = v;
    __syncthreads();
    double r = 0;
    if(w == 0) {
        r = lane < nw ? sh[lane] : 0.0;
        r = gf_warp_sum(r);
    }
    return r; // valid in thread 0
}
// n_out independent sums laid out partials[k*gridDim.x + block]
// PV.nranks > 1: the per-GPU sums are exchanged over peer memory by the last block and `out` receives the all-rank sums
__device__ inline void gf_grid_finish(double *vals, int n_out, double *partials, unsigned *ticket, double *out,
                                      double *sh, const gf_peer_view &PV) {
    __shared__ bool is_last;
    __shared__ double fin[GF_NPARTIAL];
    for(int k = 0; k < n_out; k++) {
        double r = gf_block_sum(vals[k], sh);
        if(threadIdx.x == 0) partials[(size_t)k * gridDim.x + blockIdx.x] = r;
    }
    __threadfence();
    if(threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if(is_last) {
        __threadfence();
        for(int k = 0; k < n_out; k++) {
            double s = 0;
            for(unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x)
                s += __ldcg(partials + (size_t)k * gridDim.x + b);
            double r = gf_block_sum(s, sh);
            if(threadIdx.x == 0) {
                if(PV.nranks > 1) fin[k] = r; else out[k] = r;
            }
        }
        if(threadIdx.x == 0) *ticket = 0;
        if(PV.nranks > 1) {
            __syncthreads();
            gf_peer_allreduce(PV, fin, n_out, out);
        }
        gf_signal_done(PV);
    }
}

// same, for a compile-time number of sums held in registers: the loops unroll, so `vals` is never indexed dynamically
// (a runtime-indexed array would be demoted to local memory for the whole kernel, including its accumulation loop)
template <int N>
__device__ inline void gf_grid_finish_n(double (&vals)[N], double *partials, unsigned *ticket, double *out, double *sh,
                                        const gf_peer_view &PV) {
    __shared__ bool is_last_n;
    __shared__ double fin_n[N];
#pragma unroll
    for(int k = 0; k < N; k++) {
        double r = gf_block_sum(vals[k], sh);
        if(threadIdx.x == 0) partials[(size_t)k * gridDim.x + blockIdx.x] = r;
    }
    __threadfence();
    if(threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        is_last_n = (t == gridDim.x - 1);
    }
    __syncthreads();
    if(is_last_n) {
        __threadfence();
        for(int k = 0; k < N; k++) {
            double s = 0;
            for(unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x)
                s += __ldcg(partials + (size_t)k * gridDim.x + b);
            double r = gf_block_sum(s, sh);
            if(threadIdx.x == 0) {
                if(PV.nranks > 1) fin_n[k] = r; else out[k] = r;
            }
        }
        if(threadIdx.x == 0) *ticket = 0;
        if(PV.nranks > 1) {
            __syncthreads();
            gf_peer_allreduce(PV, fin_n, N, out);
        }
        gf_signal_done(PV);
    }
}

// ---- host-side context -----------------------------------------------------------------------
struct gf_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // program
    std::vector<gf_node> h_nodes;
    std::vector<double> h_consts;
    std::vector<int32_t> h_ints;
    int n_params = 0, n_norms = 0;
    gf_node *d_nodes = nullptr;
    double *d_params = nullptr, *d_consts = nullptr, *d_norms = nullptr;
    int32_t *d_ints = nullptr;
    double *h_pin = nullptr; // pinned staging for params+norms
    size_t h_pin_cap = 0;
    // data
    double *d_cols_base = nullptr;
    bool cols_owned = true;
    gf_dev_view view{};
    int n_cols = 0;
    // reduction scratch
    double *d_partials = nullptr; // [GF_NPARTIAL * max_blocks]
    unsigned *d_ticket = nullptr;
    double *d_result = nullptr;   // [GF_NPARTIAL]
    double *h_result = nullptr;   // pinned + mapped [GF_NPARTIAL * 64]
    double *h_result_dev = nullptr; // device alias of h_result (zero-copy result of the single-GPU gf_eval_sum)
    double *d_events_out = nullptr;
    size_t events_out_cap = 0;
    int max_blocks = 0;
    float last_ms = 0;
    bool fast_on = true; // shape-specialised likelihood kernels (gf_fast.cu); GOOFIT_B200_NO_FAST=1 or gf_fast_enable(c, 0) turns them off
    bool fused_on = true; // one-pass lineshape -> amplitude -> Gram update (k_amp4_update); GOOFIT_B200_NO_FUSED=1 turns it off
    const char *last_kernel = ""; // which copy of the likelihood kernel the last gf_eval_sum launched
    uint64_t launches = 0;
    // per-kernel CUDA-event timing is opt-in (gf_profile_begin): two event records per launch are not free on a 50 us call
    static const int PROF = 256;
    bool profiling = false;
    int prof_n = 0;
    cudaEvent_t prof_ev[2 * PROF] = {nullptr};
    // completion flag of the synchronous likelihood call (mapped pinned; the last block writes done_seq into it)
    unsigned long long *h_done = nullptr, *h_done_dev = nullptr;
    unsigned long long done_seq = 0;
    // NCCL (dlopen'ed) communicator for the partial-sum all-reduce
    void *nccl_comm = nullptr;
    int nranks = 1, rank = 0;
    // peer-memory exchange (preferred over NCCL for payloads <= GF_PEER_SLOT doubles)
    gf_peer_view peer{};
    bool peer_on = false;
    void *peer_buf = nullptr;                 // this rank's exchange buffer (slots + flags)
    void *peer_mapped[GF_MAX_PEERS] = {nullptr}; // cudaIpcOpenMemHandle mappings of the other ranks' buffers
    int *h_err = nullptr, *h_err_dev = nullptr; // mapped pinned error flag
    // async parameter ring + stopwatch
    static const int RING = 16;
    double *ring_buf[RING] = {nullptr};
    cudaEvent_t ring_ev[RING] = {nullptr};
    size_t ring_cap = 0;
    int ring_pos = 0;
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    std::vector<gf_amp3_state *> amp3;
    std::vector<gf_amp4_state *> amp4;
    std::vector<gf_conv_state *> conv;
};

void gf_set_error(const std::string &s);
int gf_fail(const std::string &s);
#define GF_CUDA(x)                                                                                                     \
    do {                                                                                                               \
        cudaError_t e_ = (x);                                                                                          \
        if(e_ != cudaSuccess)                                                                                          \
            return gf_fail(std::string(#x) + ": " + cudaGetErrorString(e_) + " at " + __FILE__ + ":"                    \
                           + std::to_string(__LINE__));                                                                \
    } while(0)

size_t gf_prog_smem(const gf_ctx *c);
int gf_grid_blocks(const gf_ctx *c, long n_items, int per_sm);
int gf_alloc_aux_slot(gf_ctx *c);
// all-reduce (sum) n doubles in place on the ctx stream when a communicator is attached; no-op otherwise
int gf_allreduce_sum(gf_ctx *c, double *dev, int n);
// the peer view of the NEXT exchange (sequence number advanced) when the peer path is on, else a view with nranks = 1:
// pass it to a reduction kernel; kernels given a view with nranks > 1 leave the all-rank sums in `out`
gf_peer_view gf_next_peer_view(gf_ctx *c);
// after a kernel that was NOT given a peer view: all-reduce through NCCL when a communicator is attached
inline bool gf_peer_active(const gf_ctx *c) { return c->peer_on && c->nranks > 1; }
