// amp3td.cpp — Amp3Body_TD: time-dependent Dalitz amplitude with D0 mixing (reference src/PDFs/physics/Amp3Body_TD.cu:262-760,
// detail/SpecialWaveCalculator.cu, detail/SpecialDalitzIntegrator.cu). Host side only: the record it emits is evaluated by
// GF_OP_AMP3_TD (csrc/gf_ops.cuh), its caches and grid integrals by csrc/gf_amp3.cu.
#include <goofit/detail/Physics4.h>

namespace GooFit {
using detail::Engine;

void Amp3Body_TD::init(MixingTimeResolution *r, Observable *mistag, Observable *charmtag) {
    const std::string n = getName();
    if(decayInfo.resonances.empty()) throw GeneralError("Amp3Body_TD {} needs at least one resonance", n);
    if(decayInfo.resonances.size() > 64) throw GeneralError("Amp3Body_TD supports at most 64 resonances");
    if(!efficiency) throw GeneralError("Amp3Body_TD {} needs an efficiency function", n);
    if(!r) throw GeneralError("Amp3Body_TD {} needs a decay-time resolution", n);
    // upstream dereferences both in its initialiser list (Amp3Body_TD.cu:277-278): they are mandatory there too
    if(!mistag || !charmtag) throw GeneralError("Amp3Body_TD {} needs the mistag and charm-tag observables", n);
    resolution = r;
    _mistag    = std::make_unique<Observable>(*mistag);
    registerObservable(*mistag);
    registerObservable(*charmtag);
    totalEventSize = 7;
    registerConstant(decayInfo.motherMass);
    registerConstant(decayInfo.daug1Mass);
    registerConstant(decayInfo.daug2Mass);
    registerConstant(decayInfo.daug3Mass);
    registerConstant(decayInfo.meson_radius);
    registerConstant(decayInfo.mother_meson_radius);
    registerParameter(decayInfo._tau);
    registerParameter(decayInfo._xmixing);
    registerParameter(decayInfo._ymixing);
    registerParameter(decayInfo._deltax);
    registerParameter(decayInfo._deltay);
    setD0Fraction(0.5);
    for(auto *res : decayInfo.resonances) {
        registerParameter(res->amp_real);
        registerParameter(res->amp_imag);
        components.push_back(res);
    }
    components.push_back(resolution);
    components.push_back(efficiency);
    registerFunction("ptr_to_Tddp", GF_OP_AMP3_TD);
    integrals.assign(6 * decayInfo.resonances.size() * decayInfo.resonances.size(), 0.0);
    setSeparateNorm();
}
Amp3Body_TD::Amp3Body_TD(std::string n, Observable _dtime, Observable _sigmat, Observable m12, Observable m13, EventNumber eventNumber,
                         DecayInfo3t decay, MixingTimeResolution *r, GooPdf *eff, Observable *mistag, Observable *charmtag)
    : Amp3BodyBase("Amp3Body_TD", n, _dtime, _sigmat, m12, m13, eventNumber)
    , decayInfo(decay)
    , _m12(m12)
    , _m13(m13)
    , efficiency(eff) {
    init(r, mistag, charmtag);
}
Amp3Body_TD::Amp3Body_TD(std::string n, Observable _dtime, Observable _sigmat, Observable m12, Observable m13, EventNumber eventNumber,
                         DecayInfo3t decay, std::vector<MixingTimeResolution *> &r, GooPdf *eff, Observable, Observable *mistag,
                         Observable *charmtag)
    : Amp3BodyBase("Amp3Body_TD", n, _dtime, _sigmat, m12, m13, eventNumber)
    , decayInfo(decay)
    , _m12(m12)
    , _m13(m13)
    , efficiency(eff) {
    if(r.empty()) throw GeneralError("Amp3Body_TD {} needs a decay-time resolution", n);
    std::cout << "Multiple resolution functions not supported yet!\n"; // as upstream; the first one is used
    init(r[0], mistag, charmtag);
}
void Amp3Body_TD::setD0Fraction(fptype d0fraction) {
    if(d0fraction < 0 || d0fraction > 1) throw GeneralError("D0 fraction {} must be in [0, 1]", d0fraction);
    _D0Fraction = d0fraction;
}
void Amp3Body_TD::emitAux(detail::ProgramBuilder &B) {
    for(auto *res : decayInfo.resonances) res->emit(B);
    resolution->emit(B);
    efficiency->emitAux(B);
}
int Amp3Body_TD::emit(detail::ProgramBuilder &B) {
    int first = (int)B.nodes.size();
    efficiency->emit(B);
    std::vector<int32_t> ints{B.allocAux(), (int32_t)decayInfo.resonances.size()};
    for(auto *res : decayInfo.resonances) ints.push_back(res->getFunctionIndex());
    ints.push_back(resolution->getFunctionIndex());
    return emitRecord(B, GF_OP_AMP3_TD, 1, (int)B.nodes.size() - first + 1, ints);
}
void Amp3Body_TD::afterProgram() {
    cacheAllocated     = false;
    forceRedoIntegrals = true;
}
void Amp3Body_TD::setDataSize(unsigned int dataSize, unsigned int evtSize, unsigned int) { // Amp3Body_TD.cu:530-580
    totalEventSize = evtSize;
    if(totalEventSize < 5) throw GeneralError("totalEventSize {} must be 5 or more", totalEventSize);
    numEntries     = dataSize;
    cacheAllocated = false;
    setForceIntegrals();
}
void Amp3Body_TD::ensureCache() {
    if(cacheAllocated) return;
    if(nodeIdx < 0) throw GeneralError("Amp3Body_TD {}: call setData() on the top-level PDF before evaluating", getName());
    Engine &E = Engine::get();
    Engine::check(gf_amp3_alloc(E.ctx(), nodeIdx, E.localEvents), "gf_amp3_alloc");
    cacheAllocated     = true;
    forceRedoIntegrals = true;
}
static std::vector<std::complex<fptype>> plane(const PdfBase *p, int node, int k, bool ok) {
    if(!ok) throw GeneralError("Amp3Body_TD {}: no cached waves yet", p->getName());
    Engine &E = Engine::get();
    std::vector<std::complex<fptype>> ret(E.localEvents);
    Engine::check(gf_amp3_get_wave(E.ctx(), node, k, reinterpret_cast<double *>(ret.data()), ret.size()), "gf_amp3_get_wave");
    return ret;
}
std::vector<std::complex<fptype>> Amp3Body_TD::getCachedWaveA(size_t i) const { return plane(this, nodeIdx, (int)(2 * i), cacheAllocated); }
std::vector<std::complex<fptype>> Amp3Body_TD::getCachedWaveB(size_t i) const { return plane(this, nodeIdx, (int)(2 * i + 1), cacheAllocated); }

fptype Amp3Body_TD::normalize() { // Amp3Body_TD.cu:592-760
    recursiveSetNormalization(1.0); // the efficiency is not normalised
    ensureCache();
    Engine &E   = Engine::get();
    size_t nres = decayInfo.resonances.size();
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
        Engine::check(gf_amp3_integrals(E.ctx(), nodeIdx, redo.data(), spec, efficiency->getFunctionIndex(), integrals.data()),
                      "gf_amp3_integrals");
    }
    // sum_ij c_i conj(c_j) x {A_i A_j*, A_i B_j*, B_i B_j*}, coefficients from the pushed parameter block (Amp3Body_TD.cu:686-704)
    fpcomplex integralA_2(0, 0), integralB_2(0, 0), integralABs(0, 0);
    for(size_t i = 0; i < nres; ++i) {
        fpcomplex amplitude_i(par(5 + 2 * i), par(6 + 2 * i));
        for(size_t j = 0; j < nres; ++j) {
            fpcomplex amplitude_j(par(5 + 2 * j), -par(6 + 2 * j)); // complex conjugation
            const fptype *I = integrals.data() + 6 * (i * nres + j);
            integralA_2 += amplitude_i * amplitude_j * fpcomplex(I[0], I[1]);
            integralABs += amplitude_i * amplitude_j * fpcomplex(I[2], I[3]);
            integralB_2 += amplitude_i * amplitude_j * fpcomplex(I[4], I[5]);
        }
    }
    double one = integralA_2.real(), two = integralB_2.real(), thr = integralABs.real(), fou = integralABs.imag();
    fptype tau = par(0), x0 = par(1), y0 = par(2), dx = par(3), dy = par(4);
    fptype ret_D0    = resolution->normalization(one, two, thr, fou, tau, x0 + dx, y0 + dy);
    fptype ret_D0bar = resolution->normalization(one, two, thr, fou, tau, x0 - dx, y0 - dy);
    fptype ret       = _D0Fraction * ret_D0 + (1. - _D0Fraction) * ret_D0bar;
    const fptype xfix = 0.0039, yfix = 0.0065, taufix = 0.4101;
    // upstream tests `if(_mistag)` on the Observable object and uses ITS current value, not a per-event one (Amp3Body_TD.cu:735-744)
    const fptype mis = _mistag->getValue();
    if(mis) {
        ret *= (1 - 2 * mis);
        ret += mis * resolution->normalization(one, two, thr, fou, taufix, xfix, yfix);
        ret += mis * resolution->normalization(two, one, thr, -fou, taufix, xfix, yfix);
    }
    double binSizeFactor = 1;
    binSizeFactor *= ((_m12.getUpperLimit() - _m12.getLowerLimit()) / _m12.getNumBins());
    binSizeFactor *= ((_m13.getUpperLimit() - _m13.getLowerLimit()) / _m13.getNumBins());
    ret *= binSizeFactor;
    setNormalization(1.0 / ret);
    return ret;
}

} // namespace GooFit
