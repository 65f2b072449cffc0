// amp3.cpp — host half of Amp3Body (reference: src/PDFs/physics/Amp3Body.cu:127-428). The cached
// resonance waves and the grid integrals live in the CUDA library (csrc/gf_amp3.cu); this file keeps the
// reference's cache state machine: what is recomputed on which call (Amp3Body.cu:266-330).
#include <goofit/detail/Physics3.h>

#include <numeric>
#include <random>

namespace GooFit {
using detail::Engine;

bool inDalitz(const fptype &m12, const fptype &m13, const fptype &bigM, const fptype &dm1, const fptype &dm2,
              const fptype &dm3) { // DalitzPlotHelpers.cu:14-59
    fptype dm1pdm2 = dm1 + dm2, bigMmdm3 = bigM - dm3, bigMmdm1 = bigM - dm1;
    fptype m23 = bigM * bigM + dm1 * dm1 + dm2 * dm2 + dm3 * dm3 - m12 - m13;
    if(m23 < 0. || m23 > bigMmdm1 * bigMmdm1) return false;
    bool m12less = (m12 < dm1pdm2 * dm1pdm2) ? false : true;
    bool m12grea = (m12 > bigMmdm3 * bigMmdm3) ? false : true;
    fptype sqrtM12 = sqrt(m12), dm11 = dm1 * dm1, dm22 = dm2 * dm2, dm33 = dm3 * dm3;
    fptype e1star = 0.5 * (m12 - dm22 + dm11) / sqrtM12;
    fptype e3star = 0.5 * (bigM * bigM - m12 - dm33) / sqrtM12;
    fptype r1 = sqrt(e1star * e1star - dm11), r3 = sqrt(e3star * e3star - dm33);
    fptype es = (e1star + e3star) * (e1star + e3star);
    fptype minimum = es - (r1 + r3) * (r1 + r3), maximum = es - (r1 - r3) * (r1 - r3);
    bool m13less = (m13 < minimum) ? false : true, m13grea = (m13 > maximum) ? false : true;
    return m12less && m12grea && m13less && m13grea;
}

namespace Resonances {
RBW::RBW(std::string name, Variable ar, Variable ai, Variable mass, Variable width, unsigned int sp, unsigned int cyc,
         bool norm, bool sym)
    : ResonancePdf("RBW", name, ar, ai) {
    registerParameter(mass);
    registerParameter(width);
    registerConstant(sp);
    registerConstant(cyc);
    registerConstant(norm);
    registerFunction(sym ? "ptr_to_RBW_Sym" : "ptr_to_RBW", sym ? GF_OP_RES_RBW_SYM : GF_OP_RES_RBW);
}
GS::GS(std::string name, Variable ar, Variable ai, Variable mass, Variable width, unsigned int sp, unsigned int cyc,
       bool norm, bool)
    : ResonancePdf("GS", name, ar, ai) {
    registerParameter(mass);
    registerParameter(width);
    registerConstant(sp);
    registerConstant(cyc);
    registerConstant(norm);
    registerFunction("ptr_to_GOUSAK", GF_OP_RES_GS);
}
NonRes::NonRes(std::string name, Variable ar, Variable ai)
    : ResonancePdf("NonRes", name, ar, ai) {
    registerFunction("ptr_to_NONRES", GF_OP_RES_NONRES);
}
Gauss::Gauss(std::string name, Variable ar, Variable ai, Variable mass, Variable width, unsigned int cyc)
    : ResonancePdf("Gauss", name, ar, ai) {
    registerParameter(mass);
    registerParameter(width);
    registerConstant(cyc);
    registerFunction("ptr_to_GAUSSIAN", GF_OP_RES_GAUSS);
}
LASS::LASS(std::string name, Variable ar, Variable ai, Variable mass, Variable width, Variable _a, Variable _r, Variable _R,
           Variable _phiR, Variable _B, Variable _phiB, unsigned int sp, unsigned int cyc, bool norm)
    : ResonancePdf("LASS", name, ar, ai) {
    for(const Variable &v : {mass, width, _a, _r, _R, _phiR, _B, _phiB}) registerParameter(v);
    registerConstant(sp);
    registerConstant(cyc);
    registerConstant(norm);
    registerFunction("ptr_to_LASS", GF_OP_RES_LASS);
}
kMatrix::kMatrix(std::string name, Variable ar, Variable ai, Variable sA0, Variable sA, Variable s0_prod, Variable s0_scatt,
                 std::vector<Variable> beta_r, std::vector<Variable> beta_i, std::vector<Variable> f_prod_r, std::vector<Variable> f_prod_i,
                 std::vector<Variable> fscat, std::vector<Variable> poles, unsigned int L, unsigned int Mpair)
    : ResonancePdf("kMatrix", name, ar, ai) { // resonances/kMatrix.cu:127-177
    (void)L;
    if(fscat.size() != 5) throw GeneralError("You must have {} channels in fscat, not {}", 5, fscat.size());
    if(poles.size() != 5 * 6) throw GeneralError("You must have {}x{} channels in poles, not {}", 5, 6, poles.size());
    registerConstant(Mpair);
    for(const Variable &v : {sA0, sA, s0_prod, s0_scatt}) registerParameter(v);
    for(int i = 0; i < 5; i++) registerParameter(fscat.at(i));
    for(int i = 0; i < 5; i++) {
        registerParameter(beta_r.at(i));
        registerParameter(beta_i.at(i));
    }
    for(int i = 0; i < 5; i++) {
        registerParameter(f_prod_r.at(i));
        registerParameter(f_prod_i.at(i));
    }
    for(int i = 0; i < 30; i++) registerParameter(poles.at(i));
    registerFunction("ptr_to_kMatrix_res", GF_OP_RES_KMATRIX);
}
FLATTE::FLATTE(std::string name, Variable ar, Variable ai, Variable mean, Variable g1, Variable rg2og1, unsigned int cyc, bool symmDP)
    : ResonancePdf("FLATTE", name, ar, ai) {
    registerParameter(mean);
    registerParameter(g1);
    registerParameter(rg2og1);
    registerConstant(cyc);
    registerConstant(symmDP);
    registerFunction("ptr_to_FLATTE", GF_OP_RES_FLATTE);
}
// Tridiagonal solve for the spline's second derivatives with the end slopes fixed to twice the chord slope, as
// complex_derivative() does for the real and the imaginary component independently.
std::vector<fptype> Spline::secondDerivatives(const std::vector<fptype> &x, const std::vector<fptype> &y) {
    if(x.size() != y.size()) throw GeneralError("x and y must have the same dimensions!");
    const int n = static_cast<int>(x.size());
    if(n < 2) throw GeneralError("Spline needs at least two knots");
    std::vector<fptype> u(n), y2(n);
    const fptype slope1 = 2. * (y[1] - y[0]) / (x[1] - x[0]);
    const fptype slopeN = 2. * (y[n - 1] - y[n - 2]) / (x[n - 1] - x[n - 2]);
    if(slope1 > 0.99e30) {
        y2[0] = u[0] = 0.;
    } else {
        y2[0] = -0.5;
        u[0]  = 3.0 / (x[1] - x[0]) * ((y[1] - y[0]) / (x[1] - x[0]) - slope1);
    }
    for(int i = 1; i < n - 1; i++) {
        fptype sig = (x[i] - x[i - 1]) / (x[i + 1] - x[i - 1]);
        fptype p   = sig * y2[i - 1] + 2.0;
        y2[i]      = (sig - 1.0) / p;
        u[i]       = (y[i + 1] - y[i]) / (x[i + 1] - x[i]) - (y[i] - y[i - 1]) / (x[i] - x[i - 1]);
        u[i]       = (6.0 * u[i] / (x[i + 1] - x[i - 1]) - sig * u[i - 1]) / p;
    }
    fptype qn = 0., un = 0.;
    if(!(slopeN > 0.99e30)) {
        qn = 0.5;
        un = (3.0 / (x[n - 1] - x[n - 2])) * (slopeN - (y[n - 1] - y[n - 2]) / (x[n - 1] - x[n - 2]));
    }
    y2[n - 1] = (un - qn * u[n - 2]) / (qn * y2[n - 2] + 1.0);
    for(int k = n - 2; k >= 0; k--) y2[k] = y2[k] * y2[k + 1] + u[k];
    return y2;
}
Spline::Spline(std::string name, Variable ar, Variable ai, std::vector<fptype> &HH_bin_limits, std::vector<Variable> &pwa_coefs_reals,
               std::vector<Variable> &pwa_coefs_imags, unsigned int cyc, bool symmDP)
    : ResonancePdf("Spline", name, ar, ai) {
    const unsigned int nKnobs = HH_bin_limits.size();
    if(pwa_coefs_reals.size() != nKnobs || pwa_coefs_imags.size() != nKnobs)
        throw GeneralError("Spline {}: {} knots need as many real and imaginary coefficients", name, nKnobs);
    registerConstant(cyc);
    registerConstant(symmDP);
    registerConstant(nKnobs);
    std::vector<fptype> re(nKnobs), im(nKnobs);
    for(unsigned int i = 0; i < nKnobs; i++) {
        registerConstant(HH_bin_limits[i]);
        registerParameter(pwa_coefs_reals[i]);
        registerParameter(pwa_coefs_imags[i]);
        re[i] = pwa_coefs_reals[i].getValue();
        im[i] = pwa_coefs_imags[i].getValue();
    }
    // Frozen at construction, like upstream's cDeriatives table (Spline.cu:112 is the only recalculateCache() call); kept per
    // resonance here, where upstream shares one table between all Spline objects.
    auto d2re = secondDerivatives(HH_bin_limits, re), d2im = secondDerivatives(HH_bin_limits, im);
    for(unsigned int i = 0; i < nKnobs; i++) {
        registerConstant(d2re[i]);
        registerConstant(d2im[i]);
    }
    registerFunction("ptr_to_SPLINE", GF_OP_RES_SPLINE);
}
RhoOmegaMix::RhoOmegaMix(std::string name, Variable ar, Variable ai, Variable omega_mass, Variable omega_width, Variable rho_mass,
                         Variable rho_width, Variable real, Variable imag, Variable delta, unsigned int sp, unsigned int cyc, bool norm, bool)
    : ResonancePdf("RHOOMEGAMIX", name, ar, ai) {
    for(auto &v : {omega_mass, omega_width, rho_mass, rho_width, real, imag, delta}) registerParameter(v);
    registerConstant(sp);
    registerConstant(cyc);
    registerConstant(norm);
    registerFunction("ptr_to_RHOOMEGAMIX", GF_OP_RES_RHOOMEGA);
}
} // namespace Resonances

Amp3Body::Amp3Body(std::string n, Observable m12, Observable m13, EventNumber eventNumber, DecayInfo3 decay, GooPdf *eff)
    : Amp3BodyBase("Amp3Body", n, m12, m13, eventNumber)
    , decayInfo(decay)
    , _m12(m12)
    , _m13(m13)
    , _eventNumber(eventNumber)
    , efficiency(eff) {
    if(decayInfo.resonances.empty()) throw GeneralError("Amp3Body {} needs at least one resonance", n);
    if(decayInfo.resonances.size() > 64) throw GeneralError("Amp3Body supports at most 64 resonances");
    if(!efficiency) throw GeneralError("Amp3Body {} needs an efficiency function", n);
    // record constants: decay kinematics (the reference keeps them in __constant__ symbols, Amp3Body.cu:144-149)
    registerConstant(decay.motherMass);
    registerConstant(decay.daug1Mass);
    registerConstant(decay.daug2Mass);
    registerConstant(decay.daug3Mass);
    registerConstant(decay.meson_radius);
    registerConstant(decay.mother_meson_radius);
    for(auto *res : decayInfo.resonances) {
        registerParameter(res->amp_real);
        registerParameter(res->amp_imag);
        components.push_back(res);
    }
    components.push_back(efficiency);
    registerFunction("ptr_to_DalitzPlot", GF_OP_AMP3);
    integrals.assign(decayInfo.resonances.size() * decayInfo.resonances.size(), fpcomplex(0, 0));
    setSeparateNorm();
}
void Amp3Body::emitAux(detail::ProgramBuilder &B) {
    for(auto *res : decayInfo.resonances) res->emit(B);
    efficiency->emitAux(B);
}
int Amp3Body::emit(detail::ProgramBuilder &B) {
    int first = (int)B.nodes.size();
    efficiency->emit(B);
    std::vector<int32_t> ints{B.allocAux(), (int32_t)decayInfo.resonances.size()};
    for(auto *res : decayInfo.resonances) ints.push_back(res->getFunctionIndex());
    return emitRecord(B, GF_OP_AMP3, 1, (int)B.nodes.size() - first + 1, ints);
}
void Amp3Body::afterProgram() {
    cacheAllocated     = false; // device caches belong to the previous program
    forceRedoIntegrals = true;
}
void Amp3Body::setDataSize(unsigned int dataSize, unsigned int evtSize, unsigned int) { // Amp3Body.cu:204-236
    totalEventSize = evtSize;
    if(totalEventSize < 3) throw GeneralError("totalEventSize {} must be 3 or more", totalEventSize);
    numEntries     = dataSize;
    cacheAllocated = false;
    setForceIntegrals();
}
void Amp3Body::ensureCache() {
    if(cacheAllocated) return;
    if(nodeIdx < 0) throw GeneralError("Amp3Body {}: call setData() on the top-level PDF before evaluating", getName());
    Engine &E = Engine::get();
    Engine::check(gf_amp3_alloc(E.ctx(), nodeIdx, E.localEvents), "gf_amp3_alloc");
    cacheAllocated     = true;
    forceRedoIntegrals = true;
}
fptype Amp3Body::normalize() { // Amp3Body.cu:238-359
    recursiveSetNormalization(1.0); // the efficiency is not normalised
    ensureCache();
    Engine &E    = Engine::get();
    size_t nres  = decayInfo.resonances.size();
    std::vector<uint8_t> redo(nres);
    bool any = false;
    for(size_t i = 0; i < nres; ++i) {
        redo[i] = forceRedoIntegrals || decayInfo.resonances[i]->parametersChanged();
        any     = any || redo[i];
    }
    forceRedoIntegrals = false;
    if(any) {
        E.sync();
        Engine::check(gf_amp3_cache_waves(E.ctx(), nodeIdx, redo.data()), "gf_amp3_cache_waves");
        double spec[6] = {_m12.getLowerLimit(), _m12.getUpperLimit(), (double)_m12.getNumBins(),
                          _m13.getLowerLimit(), _m13.getUpperLimit(), (double)_m13.getNumBins()};
        std::vector<double> I(2 * nres * nres);
        Engine::check(gf_amp3_integrals(E.ctx(), nodeIdx, redo.data(), spec, efficiency->getFunctionIndex(), I.data()),
                      "gf_amp3_integrals");
        for(size_t k = 0; k < nres * nres; k++) integrals[k] = fpcomplex(I[2 * k], I[2 * k + 1]);
    }
    fpcomplex sumIntegral(0, 0);
    for(size_t i = 0; i < nres; ++i) {
        fpcomplex amplitude_i(par(2 * i), par(2 * i + 1));
        for(size_t j = 0; j < nres; ++j) {
            fpcomplex amplitude_j(par(2 * j), -par(2 * j + 1));
            sumIntegral += amplitude_i * amplitude_j * integrals[i * nres + j];
        }
    }
    fptype ret           = sumIntegral.real();
    double binSizeFactor = 1;
    binSizeFactor *= _m12.getBinSize();
    binSizeFactor *= _m13.getBinSize();
    ret *= binSizeFactor;
    setNormalization(1.0 / ret);
    return ret;
}
std::vector<std::complex<fptype>> Amp3Body::getCachedWave(size_t i) const {
    if(!cacheAllocated) throw GeneralError("Amp3Body {}: no cached waves yet", getName());
    Engine &E = Engine::get();
    std::vector<std::complex<fptype>> ret(E.localEvents);
    Engine::check(gf_amp3_get_wave(E.ctx(), nodeIdx, (int)i, reinterpret_cast<double *>(ret.data()), ret.size()), "gf_amp3_get_wave");
    return ret;
}
fpcomplex Amp3Body::sumCachedWave(size_t i) const {
    auto w = getCachedWave(i);
    return std::accumulate(w.begin(), w.end(), fpcomplex(0, 0));
}
GeneratedSignal Amp3Body::GenerateSigColumns(unsigned int numEvents, int seed) {
    GeneratedSignal G;
    devicePhaseSpace({decayInfo.motherMass, decayInfo.daug1Mass, decayInfo.daug2Mass, decayInfo.daug3Mass}, numEvents,
                     static_cast<unsigned long>(seed), generation_offset, &G.particles, G.weights);
    auto pairMass2 = [&](int a, int b, size_t i) { // detail/Dim2.h:22-35
        const fptype *p = G.particles[a].data() + 4 * i, *q = G.particles[b].data() + 4 * i;
        fptype e = p[0] + q[0], x = p[1] + q[1], y = p[2] + q[2], z = p[3] + q[3];
        return e * e - x * x - y * y - z * z;
    };
    G.variables.assign(3, std::vector<fptype>(numEvents)); // m12^2, m23^2, m13^2 in upstream's order
    std::vector<fptype> evt(numEvents);
    for(size_t i = 0; i < numEvents; i++) {
        G.variables[0][i] = pairMass2(0, 1, i);
        G.variables[1][i] = pairMass2(1, 2, i);
        G.variables[2][i] = pairMass2(0, 2, i);
        evt[i]            = static_cast<fptype>(i);
    }
    generated = std::make_unique<UnbinnedDataSet>(std::vector<Observable>{_m12, _m13, _eventNumber});
    generated->appendColumns({G.variables[0].data(), G.variables[2].data(), evt.data()}, numEvents);
    setData(generated.get());
    setDataSize(numEvents, 3);
    auto fc = fitControl;
    setFitControl(std::make_shared<EvalFit>()); // the un-normalised value, upstream's ProbFit with the normalisation forced to 1
    normalize();
    std::vector<fptype> values = evaluate_with_metric();
    setFitControl(fc);
    for(size_t i = 0; i < numEvents; i++) G.weights[i] *= values[i];
    G.flags = acceptRejectFlags(G.weights, static_cast<unsigned long>(seed), generation_offset);
    return G;
}

// Amp3Body.cu:376-428. The reference sums Re(c_i conj(c_j) w_i conj(w_j)) over the events with one transform_reduce per ordered
// pair; here the event sums W_ij = sum_e w_i conj(w_j) are taken once on the device (gf_amp3_wave_gram) and the coefficients are
// applied to the n_res^2 results on the host: no wave leaves the GPU.
std::vector<std::vector<fptype>> Amp3Body::fit_fractions() {
    size_t n_res = decayInfo.resonances.size();
    if(!cacheAllocated) throw GeneralError("{}: no cached waves yet (evaluate the PDF first)", getName());
    Engine &E = Engine::get();
    std::vector<std::complex<fptype>> W(n_res * n_res);
    Engine::check(gf_amp3_wave_gram(E.ctx(), nodeIdx, reinterpret_cast<double *>(W.data())), "gf_amp3_wave_gram");
    size_t nEntries = numEntries;
    std::vector<std::vector<fptype>> Amps_int(n_res, std::vector<fptype>(n_res)), ff = Amps_int;
    fptype buffer_all = 0;
    for(size_t i = 0; i < n_res; i++)
        for(size_t j = 0; j < n_res; j++) {
            fpcomplex ci(decayInfo.resonances[i]->amp_real.getValue(), decayInfo.resonances[i]->amp_imag.getValue());
            fpcomplex cj(decayInfo.resonances[j]->amp_real.getValue(), decayInfo.resonances[j]->amp_imag.getValue());
            fptype buffer = (ci * std::conj(cj) * W[i * n_res + j]).real();
            buffer_all += buffer;
            Amps_int[i][j] = buffer / nEntries;
        }
    fptype total_PDF = buffer_all / nEntries;
    for(size_t i = 0; i < n_res; i++)
        for(size_t j = 0; j < n_res; j++) ff[i][j] = Amps_int[i][j] / total_PDF;
    return ff;
}

// ---- DalitzPlotter (include/goofit/PDFs/physics/DalitzPlotter.h:29-95) ---------------------------------
DalitzPlotter::DalitzPlotter(GooPdf *overallSignal, Amp3Body *signalDalitz)
    : m12(signalDalitz->_m12)
    , m13(signalDalitz->_m13)
    , eventNumber(signalDalitz->_eventNumber)
    , data({m12, m13, eventNumber})
    , mother(signalDalitz->decayInfo.motherMass) {
    const DecayInfo3 &d = signalDalitz->decayInfo;
    eventNumber.setValue(0);
    for(size_t i = 0; i < m12.getNumBins(); ++i) {
        m12.setValue(m12.getLowerLimit() + m12.getBinSize() * (i + 0.5));
        for(size_t j = 0; j < m13.getNumBins(); ++j) {
            m13.setValue(m13.getLowerLimit() + m13.getBinSize() * (j + 0.5));
            if(!inDalitz(m12.getValue(), m13.getValue(), d.motherMass, d.daug1Mass, d.daug2Mass, d.daug3Mass)) continue;
            xbins.push_back(i);
            ybins.push_back(j);
            data.addEvent();
            eventNumber.setValue(eventNumber.getValue() + 1);
        }
    }
    auto *old = overallSignal->getData();
    overallSignal->setData(&data);
    signalDalitz->setDataSize(data.getNumEvents());
    pdfValues = overallSignal->getCompProbsAtDataPoints();
    overallSignal->setData(old);
}
// Same draw sequence as the reference (one uniform for the cell, then one per coordinate); the seed is
// an argument instead of std::random_device so toys are reproducible (seed 0 keeps the random behaviour).
void DalitzPlotter::fillDataSetMC(UnbinnedDataSet &dataset, size_t nTotal, unsigned int seed) {
    if(seed == 0) seed = std::random_device{}();
    std::mt19937 gen(seed);
    std::uniform_real_distribution<> unihalf(-.5, .5), uniwhole(0.0, 1.0);
    std::vector<double> integral(pdfValues[0].size());
    std::partial_sum(pdfValues[0].begin(), pdfValues[0].end(), integral.begin());
    for(double &v : integral) v /= integral.back();
    for(size_t i = 0; i < nTotal; i++) {
        double r = uniwhole(gen);
        size_t j = std::lower_bound(integral.begin(), integral.end(), r) - integral.begin();
        double currm12 = data.getValue(m12, j) + m12.getBinSize() * unihalf(gen);
        double currm13 = data.getValue(m13, j) + m13.getBinSize() * unihalf(gen);
        m12.setValue(currm12);
        m13.setValue(currm13);
        eventNumber.setValue(i);
        dataset.addEvent();
    }
}

} // namespace GooFit

// ---- efficiency helpers (physics/DalitzVetoPdf.cu:37-55, physics/SquareDalitzEffPdf.cu:78-95) --------------------------------
namespace GooFit {
DalitzVetoPdf::DalitzVetoPdf(std::string n, Observable _x, Observable _y, Variable motherM, Variable d1m, Variable d2m, Variable d3m,
                             std::vector<VetoInfo> vetos)
    : GooPdf("DalitzVetoPdf", n, _x, _y, motherM, d1m, d2m, d3m) {
    registerConstant(vetos.size());
    for(auto &veto : vetos) {
        registerParameter(veto.minimum);
        registerParameter(veto.maximum);
        registerConstant(veto.cyclic_index);
    }
    registerFunction("ptr_to_DalitzVeto", GF_OP_DALITZ_VETO);
}
SquareDalitzEffPdf::SquareDalitzEffPdf(std::string n, Observable m12, Observable m13, Variable c0, Variable c1, Variable c2, Variable c3, Variable c4,
                                       Variable c5, Variable c6, Variable c7)
    : GooPdf("SquareDalitzEffPdf", n, m12, m13, c0, c1, c2, c3, c4, c5, c6, c7) {
    registerFunction("ptr_to_SquareDalitzEff", GF_OP_SQDALITZ_EFF);
}
} // namespace GooFit
