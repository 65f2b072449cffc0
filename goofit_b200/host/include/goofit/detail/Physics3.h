// Physics3.h — three-body Dalitz amplitude PDF and its resonance shapes
// (reference: include/goofit/PDFs/physics/{Amp3Body.h,Amp3BodyBase.h,DalitzPlotHelpers.h,DalitzPlotter.h},
//  include/goofit/PDFs/physics/resonances/{Resonance,RBW,GS,NonRes,Gauss}.h).
#pragma once
#include <tuple>

#include <goofit/detail/Pdfs.h>
#include <mcbooster/GContainers.h>
#if GOOFIT_ROOT_FOUND
#include <TH2F.h>
#endif

namespace GooFit {

enum DaughterPair { PAIR_12 = 0, PAIR_13, PAIR_23 };

bool inDalitz(const fptype &m12, const fptype &m13, const fptype &bigM, const fptype &dm1, const fptype &dm2, const fptype &dm3);

/// A component of an amplitude PDF: it is never evaluated per event on its own.
class AmpComponent : public GooPdf {
  public:
    using GooPdf::GooPdf;
    fptype normalize() override { return 1.0; }
};

class ResonancePdf : public AmpComponent {
    friend class Amp3Body;
    friend class Amp3Body_TD;
    friend class Amp3Body_IS;

  public:
    Variable get_amp_real() const { return amp_real; }
    Variable get_amp_img() const { return amp_imag; }

  protected:
    ResonancePdf(std::string pdf_name, std::string name, Variable ar, Variable ai)
        : AmpComponent("Resonances::" + pdf_name, name)
        , amp_real(ar)
        , amp_imag(ai) {}
  public:
    int emit(detail::ProgramBuilder &B) override { return emitRecord(B, opcode_, 0, 1, {}); }

  protected:
    Variable amp_real;
    Variable amp_imag;
};

namespace Resonances {
class RBW : public ResonancePdf { // resonances/RBW.cu:93-119
  public:
    RBW(std::string name, Variable ar, Variable ai, Variable mass, Variable width, unsigned int sp, unsigned int cyc,
        bool norm = true, bool sym = false);
};
class GS : public ResonancePdf { // resonances/GS.cu:81-101
  public:
    GS(std::string name, Variable ar, Variable ai, Variable mass, Variable width, unsigned int sp, unsigned int cyc,
       bool norm = true, bool sym = false);
};
class NonRes : public ResonancePdf { // resonances/NonRes.cu:17-20
  public:
    NonRes(std::string name, Variable ar, Variable ai);
};
class Gauss : public ResonancePdf { // resonances/Gauss.cu:35-46
  public:
    Gauss(std::string name, Variable ar, Variable ai, Variable mean, Variable sigma, unsigned int cyc);
};
class LASS : public ResonancePdf { // resonances/LASS.cu:123-153
  public:
    LASS(std::string name, Variable ar, Variable ai, Variable mass, Variable width, Variable _a, Variable _r, Variable _R,
         Variable _phiR, Variable _B, Variable _phiB, unsigned int sp, unsigned int cyc, bool norm = true);
};
class FLATTE : public ResonancePdf { // resonances/FLATTE.cu:80-97
  public:
    FLATTE(std::string name, Variable ar, Variable ai, Variable mean, Variable g1, Variable rg2og1, unsigned int cyc, bool symmDP);
};
class kMatrix : public ResonancePdf { // resonances/kMatrix.cu:127-177: pi pi S-wave K-matrix, five channels and five poles
  public:
    kMatrix(std::string name, Variable a_r, Variable a_i, Variable sA0, Variable sA, Variable s0_prod, Variable s0_scatt,
            std::vector<Variable> beta_r, std::vector<Variable> beta_i, std::vector<Variable> f_prod_r, std::vector<Variable> f_prod_i,
            std::vector<Variable> fscat, std::vector<Variable> poles, unsigned int L, unsigned int Mpair);
};
class Spline : public ResonancePdf { // resonances/Spline.cu:91-134
  public:
    Spline(std::string name, Variable ar, Variable ai, std::vector<fptype> &HH_bin_limits, std::vector<Variable> &pwa_coefs_reals,
           std::vector<Variable> &pwa_coefs_imags, unsigned int cyc, bool symmDP = false);
    /// second derivatives of the natural-end-slope cubic spline through (x, y), per component (detail/ComplexUtils.h:20-92)
    static std::vector<fptype> secondDerivatives(const std::vector<fptype> &x, const std::vector<fptype> &y);
};
class RhoOmegaMix : public ResonancePdf { // resonances/RhoOmegaMix.cu:139-167 (sym is accepted and ignored, as upstream)
  public:
    RhoOmegaMix(std::string name, Variable ar, Variable ai, Variable omega_mass, Variable omega_width, Variable rho_mass,
                Variable rho_width, Variable real, Variable imag, Variable delta, unsigned int sp, unsigned int cyc, bool norm, bool sym);
};
} // namespace Resonances

/// The events of a GenerateSig call as flat columns (what the Python bindings turn into numpy arrays): particles[d][4*event + {E, px,
/// py, pz}], variables[v][event], the phase-space weight times the un-normalised PDF value, and the accept/reject flag of each event.
struct GeneratedSignal {
    std::vector<std::vector<fptype>> particles;
    std::vector<std::vector<fptype>> variables;
    std::vector<fptype> weights;
    std::vector<bool> flags;
};
/// The same four things as the MCBooster host containers upstream's GenerateSig returns (Amp3Body.h:55-56, Amp4Body.h:45-46,
/// Amp4Body_TD.h:94-95): std::get<0..3> / std::tie on the result work as there, and as there the receiver owns the containers the two
/// sets point to (the examples delete them). FLAGS is RealVector_h (0 or 1) for Amp3Body and Amp4Body, BoolVector_h for Amp4Body_TD.
template <class FLAGS>
using GeneratedSignalTuple = std::tuple<mcbooster::ParticlesSet_h, mcbooster::VariableSet_h, mcbooster::RealVector_h, FLAGS>;
template <class FLAGS>
inline GeneratedSignalTuple<FLAGS> toMCBoosterContainers(const GeneratedSignal &G) {
    const size_t n = G.weights.size();
    mcbooster::ParticlesSet_h particles;
    for(const auto &col : G.particles) {
        auto *p = new mcbooster::Particles_h(n);
        for(size_t i = 0; i < n; i++) (*p)[i].set(col[4 * i], col[4 * i + 1], col[4 * i + 2], col[4 * i + 3]);
        particles.push_back(p);
    }
    mcbooster::VariableSet_h variables;
    for(const auto &col : G.variables) variables.push_back(new mcbooster::RealVector_h(col.begin(), col.end()));
    return GeneratedSignalTuple<FLAGS>(std::move(particles), std::move(variables), G.weights, FLAGS(G.flags.begin(), G.flags.end()));
}
/// GENBOD phase-space events of mother -> 3 or 4 daughters from a counter-based generator: event i of a given seed is
/// the same whatever the batch it is generated in (`offset` = index of the first event). Stand-in for
/// mcbooster::PhaseSpace (not part of the reference checkout); weights are relative to the largest possible weight.
void generatePhaseSpaceParticles(const std::vector<fptype> &mother_and_daughter_masses, size_t nEvents, unsigned long seed,
                                 size_t offset, std::vector<std::vector<fptype>> &particles, std::vector<fptype> &weights);
/// The same stream generated on the GPU (csrc/gf_gen.cu k_genbod, one thread per event): what GenerateSig and the four-body
/// normalisation samples use. `dim5` (four daughters only) receives (m12, m34, cos12, cos34, phi) when not null. Needs a device.
void devicePhaseSpace(const std::vector<fptype> &mother_and_daughter_masses, size_t nEvents, unsigned long seed, size_t offset,
                      std::vector<std::vector<fptype>> *particles, std::vector<fptype> &weights,
                      std::vector<std::vector<fptype>> *dim5 = nullptr);
/// Amp3BodyBase::fillMCFlags (Amp3BodyBase.cu:13-21): accept event i when weight_i / max weight beats a uniform number
std::vector<bool> acceptRejectFlags(const std::vector<fptype> &weights, unsigned long seed, size_t offset);

struct DecayInfo3 { // DalitzPlotHelpers.h:83-92
    fptype motherMass;
    fptype daug1Mass;
    fptype daug2Mass;
    fptype daug3Mass;
    fptype meson_radius;
    fptype mother_meson_radius;
    std::vector<ResonancePdf *> resonances;
};

class Amp3BodyBase : public GooPdf {
  public:
    using GooPdf::GooPdf;
};

class DalitzPlotter;

class Amp3Body : public Amp3BodyBase { // Amp3Body.h:43-111
  public:
    Amp3Body(std::string n, Observable m12, Observable m13, EventNumber eventNumber, DecayInfo3 decay, GooPdf *eff = nullptr);
    fptype normalize() override;
    void setDataSize(unsigned int dataSize, unsigned int evtSize = 3, unsigned int offset = 0);
    void setForceIntegrals(bool f = true) { forceRedoIntegrals = f; }
    std::vector<std::complex<fptype>> getCachedWave(size_t i) const;
    fpcomplex sumCachedWave(size_t i) const;
    DecayInfo3 &getDecayInfo() { return decayInfo; }
    static void resetCacheCounter() {}
    std::vector<std::vector<fptype>> fit_fractions();
    /// Amp3Body.cu:430-547: phase-space events, their (m12^2, m23^2, m13^2), weight x |sum c A|^2 x efficiency, and
    /// accept/reject flags. As upstream, the generated events REPLACE the PDF's dataset. seed 0 = seed 0 (upstream draws one).
    GeneratedSignal GenerateSigColumns(unsigned int numEvents, int seed = 0);
    /// upstream's signature (Amp3Body.h:55-56): the same events as MCBooster host containers, owned by the receiver
    GeneratedSignalTuple<mcbooster::RealVector_h> GenerateSig(unsigned int numEvents, int seed = 0) {
        return toMCBoosterContainers<mcbooster::RealVector_h>(GenerateSigColumns(numEvents, seed));
    }
    void setGenerationOffset(int off) { generation_offset = off; }
    int getGenerationOffset() const { return generation_offset; }
    /// normalisation integrals I_ij of the last normalize() (row i, column j)
    const std::vector<fpcomplex> &getIntegrals() const { return integrals; }
    friend DalitzPlotter;

    void emitAux(detail::ProgramBuilder &B) override;
    int emit(detail::ProgramBuilder &B) override;
    void afterProgram() override;

  protected:
    void ensureCache();

    DecayInfo3 decayInfo;
    Observable _m12, _m13;
    EventNumber _eventNumber;
    GooPdf *efficiency;
    std::vector<fpcomplex> integrals;
    bool forceRedoIntegrals{true};
    bool cacheAllocated{false};
    int totalEventSize{3};
    int generation_offset{0};
    std::unique_ptr<UnbinnedDataSet> generated; // keeps the events of the last GenerateSig alive while they are the dataset
};

struct DecayInfo3t : public DecayInfo3 { // DalitzPlotHelpers.h:99-112
    Variable _tau, _xmixing, _ymixing, _deltax, _deltay;
    DecayInfo3t(Variable tau, Variable xmixing, Variable ymixing, Variable deltax, Variable deltay)
        : _tau(tau)
        , _xmixing(xmixing)
        , _ymixing(ymixing)
        , _deltax(deltax)
        , _deltay(deltay) {}
};
class MixingTimeResolution;
/// Time-dependent Dalitz amplitude with D0 mixing (Amp3Body_TD.h / Amp3Body_TD.cu:40-760): per event
/// R(|A|^2 + |B|^2, |A|^2 - |B|^2, A conj(B); tau, t, x, y, sigma_t) x efficiency with A = sum c_i A_i(m12, m13), B = sum c_i
/// A_i(m13, m12), x and y shifted by +-(deltax, deltay) according to the charm tag, and a mistagged fraction at fixed
/// world-average mixing. Observables: dtime, sigma_t, m12, m13, event number, mistag, charm tag (both mandatory, as upstream
/// dereferences them in its constructor).
class Amp3Body_TD : public Amp3BodyBase {
  public:
    Amp3Body_TD(std::string n, Observable _dtime, Observable _sigmat, Observable m12, Observable m13, EventNumber eventNumber,
                DecayInfo3t decay, MixingTimeResolution *r, GooPdf *eff, Observable *mistag = nullptr, Observable *charmtag = nullptr);
    /// upstream's constructor for several resolution functions binned in the D0 mass: its device code never selects among them
    /// ("Multiple resolution functions not supported yet!", Amp3Body_TD.cu:396), so the first one is used, as there
    Amp3Body_TD(std::string n, Observable _dtime, Observable _sigmat, Observable m12, Observable m13, EventNumber eventNumber,
                DecayInfo3t decay, std::vector<MixingTimeResolution *> &r, GooPdf *eff, Observable md0, Observable *mistag,
                Observable *charmtag);
    fptype normalize() override;
    void setDataSize(unsigned int dataSize, unsigned int evtSize = 5, unsigned int offset = 0);
    void setForceIntegrals(bool f = true) { forceRedoIntegrals = f; }
    void setD0Fraction(fptype d0fraction);
    fptype getD0Fraction() { return _D0Fraction; }
    DecayInfo3t &getDecayInfo() { return decayInfo; }
    static void resetCacheCounter() {}
    /// (A_i, B_i) of resonance i at every event
    std::vector<std::complex<fptype>> getCachedWaveA(size_t i) const;
    std::vector<std::complex<fptype>> getCachedWaveB(size_t i) const;
    void emitAux(detail::ProgramBuilder &B) override;
    int emit(detail::ProgramBuilder &B) override;
    void afterProgram() override;

  protected:
    void ensureCache();
    void init(MixingTimeResolution *r, Observable *mistag, Observable *charmtag);
    DecayInfo3t decayInfo;
    Observable _m12, _m13;
    std::unique_ptr<Observable> _mistag;
    MixingTimeResolution *resolution;
    GooPdf *efficiency;
    std::vector<fptype> integrals; // six numbers per ordered pair (i, j)
    fptype _D0Fraction{0.5};
    bool forceRedoIntegrals{true};
    bool cacheAllocated{false};
    int totalEventSize{7};
};

/// Host helper that evaluates the PDF on the Dalitz grid and draws toy events from it
/// (DalitzPlotter.h:24-141; the ROOT histogram methods are out of scope).
/// physics/DalitzVetoPdf.h: one band of a Dalitz variable to cut out
struct VetoInfo {
    Variable minimum;
    Variable maximum;
    DaughterPair cyclic_index;
    VetoInfo(Variable minimum, Variable maximum, DaughterPair cyclic_index)
        : minimum(minimum)
        , maximum(maximum)
        , cyclic_index(cyclic_index) {}
};
/// 1 inside the Dalitz plot and outside every veto band, 0 elsewhere: a factor of an efficiency function
class DalitzVetoPdf : public GooPdf {
  public:
    DalitzVetoPdf(std::string n, Observable _x, Observable _y, Variable motherM, Variable d1m, Variable d2m, Variable d3m, std::vector<VetoInfo> vetos);
};
/// physics/SquareDalitzEffPdf.h: efficiency polynomial in m23 and theta' over the D -> KS0 h h Dalitz plot
class SquareDalitzEffPdf : public GooPdf {
  public:
    SquareDalitzEffPdf(std::string n, Observable m12, Observable m13, Variable c0, Variable c1, Variable c2, Variable c3, Variable c4, Variable c5,
                       Variable c6, Variable c7);
};

class DalitzPlotter {
  public:
    DalitzPlotter(GooPdf *overallSignal, Amp3Body *signalDalitz);
    void fillDataSetMC(UnbinnedDataSet &dataset, size_t nTotal, unsigned int seed = 0);
    size_t getNumEvents() const { return data.getNumEvents(); }
    size_t getX(size_t event) const { return xbins.at(event); }
    size_t getY(size_t event) const { return ybins.at(event); }
    fptype getXval(size_t event) const { return data.getValue(m12, event); }
    fptype getYval(size_t event) const { return data.getValue(m13, event); }
    fptype getZval(size_t event) const { return mother * mother - getXval(event) * getXval(event) - getYval(event) * getYval(event); }
    fptype getVal(size_t event, size_t num = 0) const { return pdfValues.at(num).at(event); }
    UnbinnedDataSet *getDataSet() { return &data; }
    const Observable &getM12() const { return m12; }
    const Observable &getM13() const { return m13; }
#if GOOFIT_ROOT_FOUND
    /// DalitzPlotter.h:116-137 of the reference: the evaluated grid as a TH2F (bins outside the Dalitz plot stay empty); the caller owns it
    TH2F *make2D(std::string name = "dalitzplot", std::string title = "") {
        auto *dalitzplot = new TH2F(name.c_str(), title.c_str(), (int)m12.getNumBins(), m12.getLowerLimit(), m12.getUpperLimit(),
                                    (int)m13.getNumBins(), m13.getLowerLimit(), m13.getUpperLimit());
        for(size_t j = 0; j < getNumEvents(); ++j) dalitzplot->SetBinContent(1 + (int)getX(j), 1 + (int)getY(j), getVal(j));
        return dalitzplot;
    }
#endif

  private:
    std::vector<size_t> xbins, ybins;
    std::vector<std::vector<fptype>> pdfValues;
    Observable m12, m13;
    EventNumber eventNumber;
    UnbinnedDataSet data;
    fptype mother;
};

} // namespace GooFit
