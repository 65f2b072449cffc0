// Physics4.h — four-body amplitude PDF: lineshapes, spin factors, amplitudes, Amp4Body
// (reference: include/goofit/PDFs/physics/{Amp4Body.h,Amplitude.h,SpinFactors.h,DalitzPlotHelpers.h},
//  include/goofit/PDFs/physics/lineshapes/{Lineshape,RBW,SBW,LASS,Bugg,Bugg3}.h).
#pragma once
#include <goofit/detail/Physics3.h>

#include <typeinfo>

namespace GooFit {

enum DP4Pair { // DalitzPlotHelpers.h:56-76
    M_12 = 0, M_34, M_13, M_14, M_23, M_24, M_12_3, M_13_2, M_23_1, M_12_4, M_14_2, M_24_1, M_13_4, M_14_3, M_34_1,
    M_23_4, M_24_3, M_34_2
};
enum class FF : unsigned int { One = 0, BL, BL_Prime, BL2 };

class Lineshape : public AmpComponent { // lineshapes/Lineshape.h:22-49
    friend class Amp4BodyBase;
    friend class Amp4Body;
    friend class Amp4Body_TD;

  protected:
    Lineshape(std::string pdf_name, std::string name, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius, int kind);
    unsigned int _L;
    unsigned int _Mpair;
    FF _FormFac;
    fptype _radius;
    int _kind; // gf_ls_kind of the device record

  public:
    virtual bool isEqualByValue(const Lineshape &other) const;
    bool operator==(const Lineshape &L) const { return typeid(*this) == typeid(L) && isEqualByValue(L); }
    int emit(detail::ProgramBuilder &B) override { return emitRecord(B, GF_OP_LS, 0, 1, {_kind}); }
};

namespace Lineshapes {
#define GOOFIT_B200_LS(NAME)                                                                                           \
    class NAME : public Lineshape {                                                                                    \
      public:                                                                                                          \
        NAME(std::string name, Variable mass, Variable width, unsigned int L, unsigned int Mpair,                      \
             FF FormFac = FF::BL_Prime, fptype radius = 1.5);                                                          \
    };
GOOFIT_B200_LS(RBW)   // lineshapes/BW.cu:154-168
GOOFIT_B200_LS(SBW)   // lineshapes/BW.cu:172-186
GOOFIT_B200_LS(LASS)  // lineshapes/BW.cu:190-206
GOOFIT_B200_LS(Bugg)  // lineshapes/Bugg.cu:133-146 (mass and width are accepted and ignored, as upstream)
GOOFIT_B200_LS(Bugg3) // lineshapes/Bugg.cu:148-160
GOOFIT_B200_LS(One)    // lineshapes/One.cu:8-24 (always 1; mass, width ignored)
GOOFIT_B200_LS(NonRes) // lineshapes/NonRes.cu:8-40 (barrier factor only; mass, width ignored)
GOOFIT_B200_LS(Flatte) // lineshapes/Flatte.cu:14-67 (mass floats; width ignored, couplings hard-wired upstream)
#undef GOOFIT_B200_LS
class GLASS final : public Lineshape { // lineshapes/GLASS.cu:64-88 ; AdditionalVars = a, r, phiF, phiR, F
  public:
    GLASS(std::string name, Variable mass, Variable width, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius,
          std::vector<Variable> AdditionalVars);
};
class FOCUS final : public Lineshape { // lineshapes/FOCUS.cu:80-95
  public:
    enum class Mod { Kpi = 0, KEta, I32 };
    FOCUS(std::string name, Mod mod, Variable mass, Variable width, unsigned int L, unsigned int Mpair, FF FormFac = FF::BL_Prime,
          fptype radius = 1.5);
};
/// lineshapes/kMatrix.cu:121-171: the pi pi S-wave K-matrix (five channels, five poles); a_r, a_i are ignored as upstream
class kMatrix final : public Lineshape {
  public:
    kMatrix(std::string name, Variable a_r, Variable a_i, Variable sA0, Variable sA, Variable s0_prod, Variable s0_scatt,
            std::vector<Variable> beta_r, std::vector<Variable> beta_i, std::vector<Variable> f_prod_r, std::vector<Variable> f_prod_i,
            std::vector<Variable> fscat, std::vector<Variable> poles, unsigned int L, unsigned int Mpair, FF FormFac = FF::BL_Prime,
            fptype radius = 1.5);
};
/// lineshapes/GSpline.cu:126-161. spline_t = (min, max, bins) as in lineshapes/GSpline.h
using spline_t = std::tuple<fptype, fptype, unsigned int>;
class GSpline final : public Lineshape {
  public:
    GSpline(std::string name, Variable mass, Variable width, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius,
            std::vector<Variable> AdditionalVars, spline_t SplineInfo);
    static std::vector<fptype> curvatures(const std::vector<Variable> &vars, const spline_t &info);
};
} // namespace Lineshapes

enum class SF_4Body { // SpinFactors.h:18-35
    DtoPP1_PtoSP2_StoP3P4, DtoPP1_PtoVP2_VtoP3P4, DtoV1V2_V1toP1P2_V2toP3P4_S, DtoV1V2_V1toP1P2_V2toP3P4_P,
    DtoV1V2_V1toP1P2_V2toP3P4_D, DtoAP1_AtoVP2_VtoP3P4, DtoAP1_AtoVP2Dwave_VtoP3P4, DtoVS_VtoP1P2_StoP3P4,
    DtoV1P1_V1toV2P2_V2toP3P4, DtoAP1_AtoSP2_StoP3P4, DtoTP1_TtoVP2_VtoP3P4, FF_12_34_L1, FF_12_34_L2, FF_123_4_L1,
    FF_123_4_L2, ONE
};

class SpinFactor final : public AmpComponent { // SpinFactors.h:37-68
    friend class Amp4BodyBase;
    friend class Amp4Body;
    SF_4Body _SF;
    unsigned int _P0, _P1, _P2, _P3;

  public:
    SpinFactor(std::string name, SF_4Body SF, fptype mD0, unsigned int P0, unsigned int P1, unsigned int P2, unsigned int P3);
    SpinFactor(std::string name, SF_4Body SF, unsigned int P0, unsigned int P1, unsigned int P2, unsigned int P3)
        : SpinFactor(name, SF, 1.86484, P0, P1, P2, P3) {}
    bool operator==(const SpinFactor &S) const {
        return S.getName() == getName() && S._SF == _SF && S._P0 == _P0 && S._P1 == _P1 && S._P2 == _P2 && S._P3 == _P3
               && areParamsandConstantsEqualByVal(S);
    }
    int emit(detail::ProgramBuilder &B) override { return emitRecord(B, GF_OP_SF, 0, 1, {static_cast<int32_t>(_SF)}); }
};

class Amplitude : public AmpComponent { // Amplitude.h:18-49
    friend class Amp4BodyBase;
    friend class Amp4Body;
    friend class Amp4Body_TD;

  public:
    Amplitude(std::string uniqueDecayStr, Variable ar, Variable ai, std::vector<Lineshape *> LS, std::vector<SpinFactor *> SF,
              unsigned int nPerm = 1);
    bool operator==(const Amplitude &A) const;
    std::vector<SpinFactor *> getSpinFactors() { return _SF; }
    std::vector<Lineshape *> getLineShapes() { return _LS; }
    bool lineshapeParametersChanged() const;

  private:
    std::string _uniqueDecayStr;
    Variable _ar, _ai;
    std::vector<SpinFactor *> _SF;
    std::vector<Lineshape *> _LS;
    unsigned int _nPerm;
};

struct DecayInfo4 { // DalitzPlotHelpers.h:108-114
    std::vector<fptype> particle_masses;
    fptype meson_radius;
    std::vector<Amplitude *> amplitudes;
    std::vector<Amplitude *> amplitudes_B;
};

/// Deterministic four-body phase-space sample expressed in (m12, m34, cos12, cos34, phi). Replaces the MCBooster
/// PhaseSpace + Dim5 calls of the reference's constructor (Amp4Body.cu:243-289, detail/Dim5.h:19-111); the generator
/// is this build's own (GENBOD sequential two-body decays, counter-based RNG), so toys differ from MCBooster's.
std::vector<std::vector<fptype>> generatePhaseSpace5(const std::vector<fptype> &particle_masses, size_t nEvents,
                                                      unsigned long seed, std::vector<fptype> *weights = nullptr);
/// accept-reject "Unweight" flags for a weighted phase-space sample (counter-based, deterministic in (seed, index))
std::vector<bool> unweightPhaseSpace(const std::vector<fptype> &weights, unsigned long seed);
namespace detail {
/// uniform number in (0, 1) of the counter-based generator: stream `stream` of event `event` of `seed`
double counterUniform(unsigned long seed, size_t event, unsigned long stream);
} // namespace detail
/// (m12, m34, cos12, cos34, phi) of one event given as 4 x (E, px, py, pz) — pinned by tests/simple/VectorsTest.cu
void fourVectorsToDim5(const fptype *p4x4_Epxpypz, fptype *out5);

/// What Amp4Body and Amp4Body_TD share: the unique lineshape / spin-factor lists, the amplitude index records, the MC
/// normalisation sample and the device cache handles (reference: Amp4BodyBase.h + the duplicated bookkeeping of
/// Amp4Body.cu:94-292 and Amp4Body_TD.cu:307-561).
class Amp4BodyBase : public GooPdf {
  public:
    using GooPdf::GooPdf;
    void setForceIntegrals(bool f = true) { forceRedoIntegrals = f; }
    /// How the MC normalisation integral is summed. Gram (default): G_ab = sum_m A_a conj(A_b) is rebuilt only when a
    /// lineshape moved and every call sums c_a conj(c_b) G_ab on the host. PerEvent: the reference's association,
    /// sum_m |sum_a c_a A_a(m)|^2 over the cached MC amplitudes on every call (detail/NormIntegrator.cu:15-68).
    enum class NormMode { Gram, PerEvent };
    void setNormMode(NormMode m) { normMode = m; gramValid = false; }
    NormMode getNormMode() const { return normMode; }
    /// replace the normalisation sample: five columns (m12, m34, cos12, cos34, phi) of equal length
    void setNormEvents(const std::vector<std::vector<fptype>> &cols5);
    const std::vector<std::vector<fptype>> &getNormEvents() {
        materialiseNormEvents();
        return normCols;
    }
    std::vector<std::complex<fptype>> getCachedAmp(size_t a) const;
    size_t numLineshapes() const { return LineShapes.size(); }
    size_t numSpinFactors() const { return SpinFactors.size(); }
    size_t numAmplitudes() const { return nAmps; }
    void emitAux(detail::ProgramBuilder &B) override;
    void afterProgram() override;

  protected:
    /// registers amp's lineshapes and spin factors (unique by value) and appends its record {nLS, nSF, nPerm, idx...}
    void addAmplitudeRecord(Amplitude *amp);
    std::vector<int32_t> headerInts(detail::ProgramBuilder &B, int tdOffset) const;
    void ensureCache();
    /// a normalisation sample still to be generated: GENBOD candidates per seed, accept-reject, compacted or "first nAcc"
    struct NormSpec {
        std::vector<long> seeds;
        unsigned int perBatch = 0;
        bool compact = true;
        std::vector<fptype> masses;
    } normSpec;
    void materialiseNormEvents();
    bool normSampleEmpty{false}; // the generated sample had no accepted event: one candidate stands in, integrals are refused
    void setDataSizeCommon(unsigned int dataSize, unsigned int evtSize);
    /// (re)build the Gram matrix if the MC amplitudes changed; returns G[a * nAmps + b]
    const std::vector<std::complex<fptype>> &gram(bool mcAmpsChanged);
    void refreshCaches(bool any, int spins, const std::vector<uint8_t> &lsRedo, const std::vector<uint8_t> &ampRedo);
    NormMode normMode{NormMode::Gram};
    bool gramValid{false};
    std::vector<std::complex<fptype>> G;
    GooPdf *efficiency{nullptr};
    std::vector<SpinFactor *> SpinFactors;
    std::vector<Lineshape *> LineShapes;
    std::vector<int32_t> ampStarts, ampRecords;
    size_t nAmps{0};
    std::vector<std::vector<fptype>> normCols;
    std::vector<fptype> normDtime, normSigma, normWeights; // weighted normalisation sample (Amp4Body_TD only)
    size_t MCevents{0};
    bool forceRedoIntegrals{true};
    bool cacheAllocated{false};
    int totalEventSize{0};
};

class Amp4Body : public Amp4BodyBase { // Amp4Body.h:33-95
  public:
    Amp4Body(std::string n, std::vector<Observable> observables, DecayInfo4 decay, GooPdf *eff, unsigned int MCeventsNorm = 5e6);
    fptype normalize() override;
    void setDataSize(unsigned int dataSize, unsigned int evtSize = 6) { setDataSizeCommon(dataSize, evtSize); }
    int getMCevents() {
        materialiseNormEvents();
        return static_cast<int>(MCevents);
    }
    int emit(detail::ProgramBuilder &B) override;
    /// Amp4Body.cu:560-677: phase-space events, their (m12, m34, cos12, cos34, phi), weight x |sum c A|^2 x efficiency,
    /// accept/reject flags. As upstream, the generated events REPLACE the PDF's dataset.
    GeneratedSignal GenerateSigColumns(unsigned int numEvents, int seed = 0);
    /// upstream's signature (Amp4Body.h:45-46): the same events as MCBooster host containers, owned by the receiver
    GeneratedSignalTuple<mcbooster::RealVector_h> GenerateSig(unsigned int numEvents, int seed = 0) {
        return toMCBoosterContainers<mcbooster::RealVector_h>(GenerateSigColumns(numEvents, seed));
    }
    void setGenerationOffset(int off) { generation_offset = off; }
    int getGenerationOffset() const { return generation_offset; }

  protected:
    DecayInfo4 decayInfo;
    int generation_offset{0};
    std::unique_ptr<UnbinnedDataSet> generated;
};

// ---- time-dependent four-body amplitude ------------------------------------------------------------------------------
/// MixingTimeResolution.h:18-43. The per-event time functions run on the device (csrc/gf_ops.cuh gf_tres); the time
/// integrals applied to the four MC sums stay on the host, as upstream.
class MixingTimeResolution : public GooPdf {
  public:
    using GooPdf::GooPdf;
    virtual fptype normalization(fptype di1, fptype di2, fptype di3, fptype di4, fptype tau, fptype xmixing, fptype ymixing) const = 0;
    fptype normalize() override { return 1.0; }
};
class TruthResolution : public MixingTimeResolution { // TruthResolution.h / TruthResolution.cu:47-71
  public:
    TruthResolution();
    fptype normalization(fptype di1, fptype di2, fptype di3, fptype di4, fptype tau, fptype xmixing, fptype ymixing) const override;
    int emit(detail::ProgramBuilder &B) override { return emitRecord(B, GF_OP_TRES_TRUTH, 0, 1, {}); }
};
class ThreeGaussResolution : public MixingTimeResolution { // ThreeGaussResolution.h:19-42 / .cu:127-183
  public:
    ThreeGaussResolution(Variable cf, Variable tf, Variable cb, Variable cs, Variable tb, Variable ts, Variable ob, Variable os,
                         Variable sb, Variable common_bias_offset, Variable common_width_scale, Variable common_bias_offset2,
                         Variable common_width_scale2);
    fptype normalization(fptype di1, fptype di2, fptype di3, fptype di4, fptype tau, fptype xmixing, fptype ymixing) const override;
    int emit(detail::ProgramBuilder &B) override { return emitRecord(B, GF_OP_TRES_3GAUSS, 0, 1, {}); }

  private:
    Variable selectionBias;
};

struct DecayInfo4t : public DecayInfo4 { // DalitzPlotHelpers.h:122-133
    Variable _tau, _xmixing, _ymixing, _SqWStoRSrate;
    DecayInfo4t(Variable tau, Variable xmixing, Variable ymixing, Variable SqWStoRSrate)
        : _tau(tau)
        , _xmixing(xmixing)
        , _ymixing(ymixing)
        , _SqWStoRSrate(SqWStoRSrate) {}
};

// ---- normalisation-event containers of Amp4Body_TD (detail/NormEvents_4Body_{Base,DeviceCached,HostCached,Weighted,WeightedDevice}.h)
/// Upstream these own the MC columns and their cached lineshapes / spin factors, in device or host memory, and compute the
/// integral batch by batch. Here they only describe a batch - the caches and the integral live in the engine, where every batch
/// is resident in HBM (and sharded over ranks) - so DeviceCached and HostCached differ in name only.
class NormEvents_4Body_Base {
  public:
    virtual ~NormEvents_4Body_Base() = default;
    bool isWeighted() const { return !weights.empty(); }
    std::vector<std::vector<fptype>> cols;     // m12, m34, cos12, cos34, phi (empty while a generated batch is still pending)
    std::vector<fptype> dtime, sigma, weights; // weighted batches only
    long seed{0};                              // pending generation: GENBOD candidates of this seed, accept-reject, compacted
    unsigned int toGenerate{0};
    std::vector<fptype> masses;
};
class NormEvents_4Body_DeviceCached : public NormEvents_4Body_Base { // NormEvents_4Body_DeviceCached.cu:17-40
  public:
    static std::vector<NormEvents_4Body_Base *> buildBatches(const std::vector<long> &normSeeds, unsigned int numNormEventsToGenPerBatch,
                                                             const std::vector<fptype> &motherAndDaughterMasses);
};
class NormEvents_4Body_HostCached : public NormEvents_4Body_Base { // NormEvents_4Body_HostCached.cu:17-45
  public:
    static std::vector<NormEvents_4Body_Base *> buildBatches(const std::vector<long> &normSeeds, unsigned int numNormEventsToGenPerBatch,
                                                             const std::vector<fptype> &motherAndDaughterMasses);
};
class NormEvents_4Body_Weighted : public NormEvents_4Body_Base { // NormEvents_4Body_Weighted.cu:19-40
  public:
    NormEvents_4Body_Weighted(std::vector<fptype> m12, std::vector<fptype> m34, std::vector<fptype> cos12, std::vector<fptype> cos34,
                              std::vector<fptype> phi, std::vector<fptype> dtime, std::vector<fptype> sigma, std::vector<fptype> weights);
};
class NormEvents_4Body_WeightedDevice : public NormEvents_4Body_Weighted { // NormEvents_4Body_WeightedDevice.cu:19-36
  public:
    using NormEvents_4Body_Weighted::NormEvents_4Body_Weighted;
};

class Amp4Body_TD final : public Amp4BodyBase { // Amp4Body_TD.h:34-137
  public:
    /// normalisation events generated from one seed and kept in HBM (Amp4Body_TD.cu:292-307, NormEvents_4Body_DeviceCached)
    Amp4Body_TD(std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *Tres,
                GooPdf *efficiency, Observable *mistag, long normSeed, unsigned int numNormEventsToGen);
    /// upstream's host-cached batches (Amp4Body_TD.cu:269-290) exist because the sample may not fit a GPU; here every batch
    /// is generated with its own seed and the concatenation is kept in HBM (and sharded over ranks)
    Amp4Body_TD(std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *Tres,
                GooPdf *efficiency, Observable *mistag, const std::vector<long> &normSeeds, unsigned int numNormEventsToGenPerBatch);
    /// caller-supplied batches (Amp4Body_TD.cu:309-325); takes ownership of the pointers as upstream does. Either all batches
    /// are generated ones (DeviceCached / HostCached) or all are weighted.
    Amp4Body_TD(std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *Tres,
                GooPdf *efficiency, Observable *mistag, const std::vector<NormEvents_4Body_Base *> &normEvents);
    fptype normalize() override;
    void setDataSize(unsigned int dataSize, unsigned int evtSize = 8) { setDataSizeCommon(dataSize, evtSize); }
    int getNumAccNormEvents() {
        materialiseNormEvents();
        return normSampleEmpty ? 0 : static_cast<int>(MCevents); // the stand-in candidate of an empty sample does not count
    }
    /// sum of the batches' initial weights: the number of accepted events for generated batches; for weighted batches the sum
    /// upstream forms with an INTEGER accumulator (thrust::reduce(..., 0), NormEvents_4Body_Weighted.cu:36), i.e. sum of floor(w)
    fptype getSumInitNormEventWeights() {
        materialiseNormEvents();
        return weightedNorm ? sumInitWeights : static_cast<fptype>(MCevents);
    }
    bool hasWeightedNormEvents() const { return weightedNorm; }
    /// Amp4Body_TD.cu:825-1018: unweighted phase space, decay times under an exponential envelope, accept/reject flags.
    /// As upstream, the generated events REPLACE the PDF's dataset.
    GeneratedSignal GenerateSigColumns(unsigned int numEvents, int seed = 0);
    /// upstream's signature (Amp4Body_TD.h:94-95): the same events as MCBooster host containers, owned by the receiver
    GeneratedSignalTuple<mcbooster::BoolVector_h> GenerateSig(unsigned int numEvents, int seed = 0) {
        return toMCBoosterContainers<mcbooster::BoolVector_h>(GenerateSigColumns(numEvents, seed));
    }
    void setGenerationOffset(int off) { generation_offset = off; }
    int getGenerationOffset() const { return generation_offset; }
    void setMaxWeight(fptype wmax) { maxWeight = wmax; }
    fptype getMaxWeight() const { return maxWeight; }
    int emit(detail::ProgramBuilder &B) override;
    void emitAux(detail::ProgramBuilder &B) override;

  private:
    void init(std::vector<Observable> &observables, Observable *mistag, const std::vector<long> &seeds, unsigned int perBatch);
    DecayInfo4t decayInfo;
    MixingTimeResolution *resolution;
    std::vector<int32_t> flags;
    bool spinsCalculated{false};
    bool generation_no_norm{false};
    bool weightedNorm{false};
    fptype sumInitWeights{0};
    int generation_offset{0};
    fptype maxWeight{0};
    std::unique_ptr<UnbinnedDataSet> generated;
};

} // namespace GooFit
