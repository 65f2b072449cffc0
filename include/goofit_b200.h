/* goofit_b200.h — C ABI of the B200-native likelihood-evaluation engine.
 *
 * One gf_ctx owns every device allocation of ONE GPU (one process per GPU; multi-GPU runs shard
 * events/MC across processes and all-reduce the partial sums the *_partial calls leave on the device).
 * All entry points return 0 on success, non-zero on failure; gf_last_error() gives the text.
 * No torch / Thrust / C++ types cross this boundary. Not thread-safe per ctx (the reference is
 * single-threaded and global-state too: /root/reference/include/goofit/PDFs/detail/Globals.h:14-49).
 *
 * Reference interfaces replaced (file:line under /root/reference):
 *   flat tables + fn-pointer dispatch   src/PDFs/PdfBase.cu:95-119,268-327  src/PDFs/GooPdf.cu:398-400
 *   GooPdf::reduce_with_metric          src/PDFs/GooPdf.cu:45-78            -> gf_eval_sum
 *   GooPdf::evaluate_with_metric        src/PDFs/GooPdf.cu:112-173          -> gf_eval_events
 *   GooPdf::reduce_with_bins            src/PDFs/GooPdf.cu:83-108           -> gf_integrate_grid
 *   PdfBase::setData upload             src/PDFs/PdfBase.cu:329-530         -> gf_upload_columns
 *   SmartVector sync of params/norms    include/goofit/detail/SmartVectorGPU.h:35-56 -> gf_set_params / gf_set_norms
 *   ConvolutionPdf workspaces           src/PDFs/combine/ConvolutionPdf.cu:313-358 -> gf_conv_*
 *   Amp3Body caches / integrals         src/PDFs/physics/Amp3Body.cu:238-359 -> gf_amp3_*
 *   Amp4Body caches / MC integral       src/PDFs/physics/Amp4Body.cu:398-558 -> gf_amp4_*
 *   Amp4Body_TD / NormEvents_4Body_*    src/PDFs/physics/Amp4Body_TD.cu:744-799 -> gf_amp4_* (td=1)
 */
#ifndef GOOFIT_B200_H
#define GOOFIT_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct gf_ctx gf_ctx;

/* ---- program block ------------------------------------------------------------------------
 * The PDF tree is flattened host-side into an array of fixed-size node records in POST-ORDER
 * (children before parent), so a subtree rooted at node r occupies nodes [r-size+1, r] and is
 * evaluated by a stack machine with no indirect calls. Nodes that are not evaluated per event
 * (ConvolutionPdf's model/resolution, resonances, lineshapes, spin factors) live in the same
 * array outside the main range and are referenced through the node's `ints`. */
typedef struct gf_node {
    int32_t op;        /* gf_op */
    int32_t nchild;    /* children evaluated per event (on the value stack) */
    int32_t size;      /* nodes in the per-event subtree, including this one */
    int32_t par_off;   /* first parameter in params[] */
    int32_t npar;
    int32_t con_off;   /* first constant in consts[] */
    int32_t ncon;
    int32_t obs_off;   /* first observable column index in ints[] */
    int32_t nobs;
    int32_t norm_slot; /* index into norms[] (every node owns one slot) */
    int32_t int_off;   /* node-specific integer block in ints[] */
    int32_t nint;
} gf_node;

enum gf_op {
    GF_OP_EXP = 1,          /* ExpPdf.cu:7    exp(a x) */
    GF_OP_EXP_OFFSET,       /* ExpPdf.cu:19   exp(a (x-x0)) */
    GF_OP_EXP_POLY,         /* ExpPdf.cu:32 */
    GF_OP_EXP_POLY_OFFSET,  /* ExpPdf.cu:50 */
    GF_OP_GAUSSIAN,         /* GaussianPdf.cu:7 */
    GF_OP_POLYNOMIAL,       /* PolynomialPdf.cu:8 */
    GF_OP_OFFSET_POLYNOMIAL,/* PolynomialPdf.cu:27 */
    GF_OP_MULTI_POLYNOMIAL, /* PolynomialPdf.cu:47 */
    GF_OP_ARGUS_UPPER,      /* ArgusPdf.cu:7 */
    GF_OP_ARGUS_LOWER,      /* ArgusPdf.cu:27 */
    GF_OP_KINLIMITBW,       /* KinLimitBWPdf.cu:26 */
    GF_OP_SCALED_GAUSSIAN,  /* ScaledGaussianPdf.cu:9 */
    GF_OP_JOHNSONSU,        /* JohnsonSUPdf.cu:8 */
    GF_OP_CONSTANT,         /* a polynomial with only its x^0 term (PolynomialPdf.cu:8-25 with one weight and lowest degree 0,
                               :47-118 with maxDegree 0): the last parameter of the record; no observable is read */
    GF_OP_BW,               /* BWPdf.cu:6         params: mean, gamma */
    GF_OP_BIFURGAUSS,       /* BifurGaussPdf.cu:6 params: mean, sigmaLeft, sigmaRight */
    GF_OP_CRYSTALBALL,      /* CrystalBallPdf.cu:6 params: mean, sigma, alpha, power */
    GF_OP_EXPGAUS,          /* ExpGausPdf.cu:6    params: mean, sigma, alpha */
    GF_OP_NOVOSIBIRSK,      /* NovosibirskPdf.cu:6 params: mean, sigma, tail */
    GF_OP_STEP,             /* StepPdf.cu:6       params: x0; consts: Up */
    GF_OP_TRIGTHRESH_UPPER, /* TrigThresholdPdf.cu:13 params: thresh, trigConst, linConst */
    GF_OP_TRIGTHRESH_LOWER, /* TrigThresholdPdf.cu:28 */
    GF_OP_TRIGTHRESH2_UPPER,/* TrigThresholdPdf.cu:43 two observables; params: thresh, trigConst, linConst, massConstant */
    GF_OP_TRIGTHRESH2_LOWER,/* TrigThresholdPdf.cu:61 */
    GF_OP_BERNSTEIN,        /* BernsteinPdf.cu:26 params: coefficients; consts: n, lowX, uppX, mindeg */
    GF_OP_CORRGAUSS,        /* CorrGaussianPdf.cu:6 two observables; params: mean1, sigma1, mean2, sigma2, corr */
    GF_OP_BINTRANSFORM,     /* BinTransformPdf.cu:6 consts per observable: lower limit, bin size, number of bins */
    GF_OP_LANDAU,           /* LandauPdf.cu:25    params: mpv, sigma */
    GF_OP_VOIGTIAN,         /* VoigtianPdf.cu:285-320 params in constructor order (m, s, w); as upstream the device function takes the
                               second as the Breit-Wigner width and the third as the Gaussian sigma */
    GF_OP_HISTOGRAM,        /* SmoothHistogramPdf.cu:13-46 ints: aux slot of the (smoothed) bin table, see gf_table_set;
                               consts: nObs, histogram index, nVars, then (bins, lower limit, bin size) per observable */
    GF_OP_VARBINTRANSFORM,  /* utilities/VariableBinTransform1DPdf.cu:8-27 consts: number of limits, the limits; value = how many of them |x| has reached */
    GF_OP_ADD = 32,         /* AddPdf.cu:16  ints: child norm slots */
    GF_OP_ADD_EXT,          /* AddPdf.cu:53 */
    GF_OP_PROD,             /* ProdPdf.cu:9 */
    GF_OP_CONVOLVE,         /* ConvolutionPdf.cu:27  ints: conv id */
    GF_OP_EVTW_ADD,         /* EventWeightedAddPdf.cu:8-44   weights = the first nchild-1 observable columns of the record */
    GF_OP_EVTW_ADD_EXT,     /* EventWeightedAddPdf.cu:46-91  one weight column per component, sum divided by the total weight */
    GF_OP_MAPPED,           /* MappedPdf.cu:6-38  child 0 = mapping function, value rounded to the index of the component to return */
    GF_OP_COMPOSITE_ARG,    /* CompositePdf.cu:6-24  child: the core function g; its value (left on the stack) becomes what the composite's
                             * observable 0 reads as, until the GF_OP_COMPOSITE record that follows the shell */
    GF_OP_COMPOSITE,        /* CompositePdf.cu:26-36 children: the GF_OP_COMPOSITE_ARG record and the shell h; value = h(g(event)), neither normalised */
    /* second range of leaf records (the first is below GF_OP_ADD): GF_IS_LEAF(op) */
    GF_OP_DALITZ_VETO,      /* DalitzVetoPdf.cu:7-33 params: M, m1, m2, m3, then (min, max) per veto; consts: nVetos, pair index per veto */
    GF_OP_SQDALITZ_EFF,     /* SquareDalitzEffPdf.cu:35-74 params: c0..c7 */
    GF_OP_AMP3 = 48,        /* Amp3Body.cu:60  child: efficiency; ints: amp3 id */
    GF_OP_AMP4,             /* Amp4Body.cu:53 */
    GF_OP_AMP4_TD,          /* Amp4Body_TD.cu:167 */
    GF_OP_AMP3_TD,          /* Amp3Body_TD.cu:40-256  child: efficiency; consts as GF_OP_AMP3; params tau, x, y, deltax, deltay, then (re, im) per
                               resonance; obs dtime, sigma_t, m12, m13, event number [, mistag [, charm tag]]; ints {aux slot, n_res,
                               res_node[n_res], resolution record}. Cached planes: A_i at (m12, m13) in plane 2i, B_i at (m13, m12) in 2i+1 */
    /* resonance / lineshape / spin-factor records (never on the per-event stack) */
    GF_OP_RES_RBW = 64,     /* resonances/RBW.cu:9 plainBW<1>; params mass,width; consts spin,cyc,norm */
    GF_OP_RES_RBW_SYM,      /* plainBW<2> */
    GF_OP_RES_GS,           /* resonances/GS.cu:9 */
    GF_OP_RES_NONRES,       /* resonances/NonRes.cu:8 */
    GF_OP_RES_GAUSS,        /* resonances/Gauss.cu:9 */
    GF_OP_RES_LASS,         /* resonances/LASS.cu:9  params mass,width,a,r,R,phiR,B,phiB; consts spin,cyc,norm */
    GF_OP_RES_FLATTE,       /* resonances/FLATTE.cu:18  params mass,g1,g2/g1; consts cyc,symmDP */
    GF_OP_RES_SPLINE,       /* resonances/Spline.cu:10  params re,im per knot; consts cyc,doSwap,nKnots,knots[n],second derivatives[2n] */
    GF_OP_RES_RHOOMEGA,     /* resonances/RhoOmegaMix.cu:9  params omega mass,width, rho mass,width, real, imag, delta; consts spin,cyc,norm */
    GF_OP_RES_KMATRIX,      /* resonances/kMatrix.cu:19  params sA0,sA,s0_prod,s0_scatt, fscat[5], beta(re,im)[5], f_prod(re,im)[5], per pole couplings[5]+mass; consts cyc */
    GF_OP_LS = 96,          /* 4-body lineshape; consts[0] = gf_ls_kind */
    GF_OP_SF,               /* 4-body spin factor; consts[0] = gf_sf_kind */
    /* decay-time resolution records of Amp4Body_TD (never on the per-event stack) */
    GF_OP_TRES_TRUTH = 112, /* physics/TruthResolution.cu:8-28 ; no parameters */
    GF_OP_TRES_3GAUSS       /* physics/ThreeGaussResolution.cu:11-125 ; 13 parameters in constructor order */
};
/* records that push one value without popping any (closed-form shapes) */
#define GF_IS_LEAF(op) ((op) < GF_OP_ADD || ((op) >= GF_OP_DALITZ_VETO && (op) < GF_OP_AMP3))

enum gf_metric { /* MetricPointer.cu:8-58 */
    GF_METRIC_EVAL = 0, GF_METRIC_NLL = 1, GF_METRIC_PROB = 2,
    GF_METRIC_BINAVG = 3, GF_METRIC_BINERR = 4, GF_METRIC_CHISQ = 5
};

const char *gf_last_error(void);
int gf_version(void);

gf_ctx *gf_create(int device_id);
void gf_destroy(gf_ctx *);
int gf_device_info(gf_ctx *, int *sm_count, size_t *free_bytes, size_t *total_bytes);

int gf_set_program(gf_ctx *, const gf_node *nodes, int n_nodes, const double *consts, int n_consts,
                   const int32_t *ints, int n_ints, int n_params, int n_norms);
/* the checks gf_set_program applies, without a context or a device: index ranges of every record, and a simulated
 * post-order walk of every evaluable subtree (pops available at each combiner, peak depth <= GF_MAX_STACK) */
int gf_validate_program(const gf_node *nodes, int n_nodes, int n_consts, int n_ints, int n_params, int n_norms);
/* Drop every per-node device cache (Amp3Body/Amp4Body caches, convolution workspaces) and clear the aux
 * table; called by the host layer before a re-flattened program is installed. */
int gf_reset_aux(gf_ctx *);
int gf_set_params(gf_ctx *, const double *params, int n);
int gf_set_norms(gf_ctx *, const double *norms, int n);

/* Dataset: n_cols host columns of n_events doubles -> 256-B aligned SoA columns in HBM (pinned staging).
 * For binned data the last two columns are (content, error|volume) as PdfBase.cu:485-494.
 * total_events feeds the binned metrics' c_totalEvents. */
int gf_upload_columns(gf_ctx *, const double *const *cols, int n_cols, size_t n_events, double total_events);
/* As above, but the columns are already device pointers (resident data; no host copy). */
int gf_adopt_device_columns(gf_ctx *, const double *const *dev_cols, int n_cols, size_t n_events, double total_events);

/* K1: sum over events of metric(pdf(event), norm[root]).  Deterministic: fixed grid, fixed order. */
int gf_eval_sum(gf_ctx *, int root, int metric, double *out);
/* Same, but leaves the partial sum on the device (for an NCCL all-reduce by the caller) */
int gf_eval_sum_async(gf_ctx *, int root, int metric);
double *gf_partial_device_ptr(gf_ctx *);  /* device double[GF_NPARTIAL]: [0]=NLL sum, [1..4]=MC integral sums */
void *gf_stream(gf_ctx *);                /* cudaStream_t the ctx launches on */
int gf_sync(gf_ctx *);
/* K3: per-event metric values to host (out_host[n_events]) */
int gf_eval_events(gf_ctx *, int root, int metric, double *out_host, size_t n);
/* K2: sum over the mid-point grid of the raw pdf rooted at `root`; lo_hi_n = (lo,hi,nbins) per observable of root */
int gf_integrate_grid(gf_ctx *, int root, const double *lo_hi_n, int n_obs, double *out);

/* ConvolutionPdf (ConvolutionPdf.cu:151-358): ints of the node = {aux_slot, model_root, resol_root}; consts = lo,hi,step */
int gf_conv_setup(gf_ctx *, int node, double lo, double hi, double step, double obs_lo, double obs_hi);
int gf_conv_fill(gf_ctx *, int node, int fill_model, int fill_resol);
int gf_conv_sizes(gf_ctx *, int node, int *n_model, int *n_resol, int *n_table);
int gf_conv_get_workspace(gf_ctx *, int node, int which /*0 model,1 resolution,2 table*/, double *out, int n);

/* Amp3Body: ints of the node = {aux_slot, n_res, res_node[n_res]};  consts = mother,d1,d2,d3,meson_radius,mother_radius */
int gf_amp3_alloc(gf_ctx *, int node, size_t n_events);
int gf_amp3_cache_waves(gf_ctx *, int node, const uint8_t *redo_mask);
int gf_amp3_integrals(gf_ctx *, int node, const uint8_t *redo_mask, const double *lo_hi_n6, int eff_root,
                      double *I_re_im /* n_res*n_res*2, row i col j */);
/* For a GF_OP_AMP3_TD node (Amp3Body_TD.cu:592-760) the same three calls keep TWO planes per resonance (A_i, B_i: SpecialWaveCalculator.cu:15-68),
 * gf_amp3_integrals returns SIX numbers per ordered pair, I[(i*n_res+j)*6 + ..] = sum_bins eff x {Re, Im of A_i A_j*, A_i B_j*, B_i B_j*}
 * (SpecialDalitzIntegrator.cu:15-136), and gf_amp3_get_wave takes a plane index 0 .. 2 n_res - 1. */
int gf_amp3_get_wave(gf_ctx *, int node, int res, double *out_re_im, size_t n_events);
/* Hermitian matrix of the cached waves, W[(i*n_res+j)*2+{re,im}] = sum_e w_i(e) conj(w_j(e)), summed over ranks when a communicator
 * is attached: the device side of Amp3Body::fit_fractions (Amp3Body.cu:376-428, one transform_reduce of CoefSumFunctor per pair) */
int gf_amp3_wave_gram(gf_ctx *, int node, double *W_re_im /* n_res*n_res*2 */);

/* Amp4Body / Amp4Body_TD: ints = {aux slot, nLS, nSF, nAmp, td_off, ls_node[nLS], sf_node[nSF], start[nAmp],
 * records {nLS_a, nSF_a, nPerm, LS idx..., SF idx...}}. Amp4Body: td_off = 0, params = (re, im) per amplitude.
 * Amp4Body_TD: td_off > 0 is the offset (within the node's ints) of {resolution record, flag[nAmp]} with
 * flag 0 = A only, 1 = B only, 2 = both (Amp4Body_TD.cu:373-381); params = tau, x, y, sqrt(WS/RS), then one (re, im)
 * pair per coefficient in the reference's order; obs = m12, m34, cos12, cos34, phi, event number, dtime, sigma_t. */
int gf_amp4_alloc(gf_ctx *, int node, size_t n_events);
int gf_amp4_set_mc(gf_ctx *, int node, const double *const *mc_cols5, size_t n_mc, const double *weights,
                   const double *dtime, const double *sigma);
int gf_amp4_cache(gf_ctx *, int node, int redo_spins, const uint8_t *ls_redo, const uint8_t *amp_redo);
/* sum over the (local shard of the) MC sample, all-reduced when a communicator is attached.
 * Amp4Body: out[0] = sum |sum_a c_a A_a|^2 (detail/NormIntegrator.cu:15-68).
 * Amp4Body_TD: out[0..3] = sum |A|^2, sum |B|^2, sum Re(A B*), sum Im(A B*) (detail/NormIntegrator_TD.cu:15-99).
 * Amp4Body_TD with a WEIGHTED sample (gf_amp4_set_mc given weights, dtime and sigma): out[0] = sum_m w_m pdf(x_m, t_m, sigma_m), the
 * time-dependent PDF at every MC event's own decay time (detail/NormIntegrator_TD_Weighted.cu:15-122). */
int gf_amp4_norm(gf_ctx *, int node, double *out);
int gf_amp4_norm_async(gf_ctx *, int node);
/* Hermitian Gram matrix of the cached MC amplitudes, G[(a*nAmp+b)*2+{re,im}] = sum_m w_m A_a(m) conj(A_b(m)), summed over
 * ranks when a communicator is attached. With it the integral of gf_amp4_norm is the host sum  sum_ab c_a conj(c_b) G_ab
 * (Amp4Body) or the four sums over the A / B coefficient sets (Amp4Body_TD): no per-call pass over the MC sample while
 * the lineshapes stand still. Call it after gf_amp4_cache rebuilt MC amplitudes. */
int gf_amp4_gram(gf_ctx *, int node, double *G_re_im /* nAmp*nAmp*2 */);
/* gf_amp4_cache + gf_amp4_gram as ONE pass per sample: the flagged lineshapes, the amplitudes that contain them and - on the MC
 * sample - the whole Gram matrix, accumulated with FP64 tensor-core MMAs (replaces the re-run of NormLSCalculator + NormIntegrator
 * when a mass or width moves: Amp4Body.cu:462-558, detail/NormLSCalculator.cu:11-62, NormIntegrator.cu:15-68). Up to 16 amplitudes;
 * beyond that (or with the "fused" option off) it is exactly gf_amp4_cache followed by gf_amp4_gram. */
int gf_amp4_cache_gram(gf_ctx *, int node, int redo_spins, const uint8_t *ls_redo, const uint8_t *amp_redo, double *G_re_im);
/* the data-sample half of gf_amp4_cache alone (toy generation evaluates the PDF un-normalised: Amp4Body_TD.cu:946-950) */
int gf_amp4_cache_data(gf_ctx *, int node, int redo_spins, const uint8_t *ls_redo, const uint8_t *amp_redo);
int gf_amp4_get_amp(gf_ctx *, int node, int amp, double *out_re_im, size_t n_events);

/* ---- multi-GPU: one process per GPU, partial sums all-reduced in-stream over NCCL (NVLink/NVSwitch) --------
 * Replaces the reference's MPI_Allreduce of the partial NLL (src/PDFs/GooPdf.cu:72-75). libnccl.so.2 is
 * dlopen'ed on first use. gf_comm_unique_id fills 128 bytes on rank 0; the caller ships them to every rank
 * (any side channel), then each rank calls gf_comm_init. Afterwards gf_eval_sum / gf_amp4_norm return the
 * sum over all ranks (same value on every rank).
 * gf_comm_init also maps every rank's exchange buffer into every process (cudaIpc over NVLink): sums of up to 256
 * doubles are then exchanged by the LAST BLOCK of the reduction kernel itself - remote stores into every peer's slot, a
 * release/acquire flag per rank, a rank-ordered (bit-reproducible) sum - so a likelihood call on N GPUs is still one
 * kernel launch and one synchronisation, with no collective kernel behind it. NCCL stays the fallback (no peer access,
 * GOOFIT_B200_NO_PEER=1) and carries the handles at start-up, where its channel set-up is also paid (not in the fit). */
int gf_comm_unique_id(void *id128);
int gf_comm_init(gf_ctx *, const void *id128, int rank, int nranks);
int gf_comm_destroy(gf_ctx *);
/* how partial sums travel between ranks: "none", the peer-memory exchange, or NCCL (GOOFIT_B200_NO_PEER=1 forces NCCL) */
const char *gf_comm_path(gf_ctx *);
/* number of ranks of the attached communicator, 0 when none is attached (the host layer refuses to evaluate a
 * sharded dataset without one: a partial sum must never be mistaken for the likelihood) */
int gf_comm_attached(gf_ctx *);
/* Per-event results of a sharded evaluation gathered on every rank, in rank order: replaces MPI_Allgather(v) of
 * evaluate_with_metric (src/PDFs/GooPdf.cu:139-172). counts[r] = events of rank r; `all` receives sum(counts) doubles.
 * Without a communicator it copies `local`. */
int gf_allgather_events(gf_ctx *, const double *local, const size_t *counts, double *all);
/* device-side stopwatch on the ctx stream (CUDA events): start, enqueue work, stop -> elapsed ms */
int gf_timer_start(gf_ctx *);
int gf_timer_stop(gf_ctx *, float *ms);
/* as gf_set_params (+ gf_set_norms when norms != NULL), but never blocks on earlier kernels: the values go
 * through a ring of pinned slots, so a caller can queue several evaluations back to back */
int gf_set_params_async(gf_ctx *, const double *params, int n, const double *norms, int n_norms);

/* GENBOD phase space of mother -> 3 or 4 daughters on the device, events [offset, offset + n) of the counter-based stream `seed`
 * (replaces mcbooster::PhaseSpace::Generate + EvaluateArray<Dim5>: Amp4Body.cu:243-289,560-677, Amp4Body_TD.cu:825-1018,
 * detail/Dim5.h:19-111; MCBooster itself is not part of the reference checkout). masses = mother, daughters. Host outputs, each
 * optional: particles[nd][n][4] = (E, px, py, pz), weights[n] relative to the largest possible weight, dim5[5][n] = (m12, m34,
 * cos12, cos34, phi) for four daughters. */
int gf_genbod(gf_ctx *, const double *masses, int nd, size_t n, uint64_t seed, uint64_t offset, double *particles, double *weights, double *dim5);
/* gf_genbod for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). It has no counterpart in the reference. Same
 * arguments, same kernel and bit-identical outputs, but the device staging area is a file-scope __device__ array inside the
 * library (22 doubles per event: 5.5 MiB, held by every GPU that loads the library until the process exits) instead of per-call
 * cudaMalloc / cudaFree: no allocation of any kind is made. n is limited to GF_GENBOD_STATIC_MAX_EVENTS per call (an error names
 * the limit beyond it; walk a longer range with `offset`). All calls on one device share the array: calls that overlap on the
 * GPU (two contexts or two host threads) race on it and must be serialised by the caller. */
#define GF_GENBOD_STATIC_MAX_EVENTS 32768
int gf_genbod_static(gf_ctx *, const double *masses, int nd, size_t n, uint64_t seed, uint64_t offset, double *particles, double *weights,
                     double *dim5);
/* a table of n doubles attached to a GF_OP_HISTOGRAM record (its ints[0] names the aux slot): replaces the device vectors behind
 * dev_smoothed_histograms[] (SmoothHistogramPdf.cu:6-7,165-186). May be called again with new values (the smoothing moved). */
int gf_table_set(gf_ctx *, int node, const double *values, size_t n);
/* measured FP64 FMA peak of this GPU (DFMA micro-benchmark, TFLOP/s, 2 flops per FMA): the roofline denominator of
 * the FP64-bound cache kernels (k_amp3_waves, k_amp4_cache), which MEASURED_PEAKS.json does not carry */
int gf_fp64_peak(gf_ctx *, double *tflops);
/* opt-in timing of the likelihood kernel: gf_profile_begin arms CUDA events around each of the next (up to 256)
 * k_eval_sum launches on the ctx stream; gf_profile_end synchronises and returns their average duration. Off by default:
 * two event records per launch are not free on a 50 us call. gf_last_kernel_ms returns the last such average. */
int gf_profile_begin(gf_ctx *);
int gf_profile_end(gf_ctx *, float *avg_ms, int *n_launches);
float gf_last_kernel_ms(gf_ctx *);
/* which copy of the likelihood kernel the last gf_eval_sum / gf_eval_sum_async launched: "k_eval_sum" (the interpreter) or
 * one of the shape-specialised copies of csrc/gf_fast.cu. GOOFIT_B200_NO_FAST=1 forces the interpreter (A/B runs). */
const char *gf_last_kernel_name(gf_ctx *);
/* A/B switches: "fast" (shape-specialised likelihood kernels, csrc/gf_fast.cu) and "fused" (one-pass lineshape -> amplitude ->
 * Gram update, gf_amp4_cache_gram); value 0 = off. Returns the previous setting, -1 for an unknown name. Both default to on;
 * GOOFIT_B200_NO_FAST=1 / GOOFIT_B200_NO_FUSED=1 in the environment start a context with them off. */
int gf_set_option(gf_ctx *, const char *name, int value);
/* number of kernels this ctx has launched since creation */
uint64_t gf_launch_count(gf_ctx *);

#define GF_NPARTIAL 8
#define GF_MAX_OBS 16
#define GF_MAX_STACK 16

#ifdef __cplusplus
}
#endif
#endif
