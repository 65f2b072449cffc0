// pdfs.cpp — host halves of the basic shapes and combiners (reference: src/PDFs/basic/*.cu,
// src/PDFs/combine/*.cu; line numbers cited per function).
#include <goofit/detail/Pdfs.h>

namespace GooFit {
using detail::Engine;

// ---- ExpPdf (basic/ExpPdf.cu:73-132) -----------------------------------------------------------------
ExpPdf::ExpPdf(std::string n, Observable _x, Variable alpha)
    : GooPdf("ExpPdf", n, _x, alpha) {
    registerFunction("ptr_to_Exp", GF_OP_EXP);
}
ExpPdf::ExpPdf(std::string n, Observable _x, Variable alpha, Variable offset)
    : GooPdf("ExpPdf", n, _x, offset, alpha) {
    registerFunction("ptr_to_ExpOffset", GF_OP_EXP_OFFSET);
}
ExpPdf::ExpPdf(std::string n, Observable _x, std::vector<Variable> &weights)
    : GooPdf("ExpPdf", n, _x) {
    if(weights.empty()) throw GeneralError("Weights are empty!");
    for(Variable &w : weights) registerParameter(w);
    registerFunction("ptr_to_ExpPoly", GF_OP_EXP_POLY);
}
ExpPdf::ExpPdf(std::string n, Observable _x, std::vector<Variable> &weights, Variable offset)
    : GooPdf("ExpPdf", n, _x, offset) {
    if(weights.empty()) throw GeneralError("Weights are empty!");
    for(Variable &w : weights) registerParameter(w);
    registerFunction("ptr_to_ExpPolyOffset", GF_OP_EXP_POLY_OFFSET);
}
fptype ExpPdf::integrate(fptype lo, fptype hi) const {
    fptype alpha = par(0);
    if(0 == alpha) return hi - lo;
    fptype ret = exp(alpha * hi) - exp(alpha * lo);
    ret /= alpha;
    return ret;
}

// ---- GaussianPdf (basic/GaussianPdf.cu:20-35): the analytic integral is over all of R ------------------
GaussianPdf::GaussianPdf(std::string n, Observable _x, Variable mean, Variable sigma)
    : GooPdf("GaussianPdf", n, _x, mean, sigma) {
    registerFunction("ptr_to_Gaussian", GF_OP_GAUSSIAN);
}
fptype GaussianPdf::integrate(fptype, fptype) const {
    static const fptype rootPi = sqrt(atan2(0.0, -1.0));
    static const fptype root2  = sqrt(2.0);
    fptype sigma = par(1);
    sigma *= root2 * rootPi;
    return sigma;
}

// ---- PolynomialPdf (basic/PolynomialPdf.cu:124-215) -----------------------------------------------------
PolynomialPdf::PolynomialPdf(std::string n, Observable _x, std::vector<Variable> weights, unsigned int lowestDegree)
    : GooPdf("PolynomialPdf", n, _x) {
    registerConstant(lowestDegree);
    for(Variable &v : weights) registerParameter(v);
    // w0 * pow(x, 0) is w0 for every x (NaN included): the record needs no observable read on the device
    registerFunction("ptr_to_Polynomial", (weights.size() == 1 && lowestDegree == 0) ? GF_OP_CONSTANT : GF_OP_POLYNOMIAL);
}
PolynomialPdf::PolynomialPdf(std::string n, Observable _x, std::vector<Variable> weights, Variable x0, unsigned int lowestDegree)
    : GooPdf("PolynomialPdf", n, _x)
    , center(new Variable(x0)) {
    registerConstant(lowestDegree);
    for(Variable &v : weights) registerParameter(v);
    registerParameter(x0);
    registerFunction("ptr_to_OffsetPolynomial", GF_OP_OFFSET_POLYNOMIAL);
}
PolynomialPdf::PolynomialPdf(std::string n, std::vector<Observable> obses, std::vector<Variable> coeffs,
                             std::vector<Variable> offsets, unsigned int maxDegree)
    : GooPdf("PolynomialPdf", n) {
    unsigned int numParameters = 1;
    registerConstant(maxDegree);
    for(unsigned int i = 0; i < obses.size(); ++i) {
        registerObservable(obses[i]);
        numParameters *= (maxDegree + 1 + i);
        registerConstant(maxDegree + 1 + i); // padding entries, as the reference
    }
    for(int i = (int)observablesList.size(); i > 1; --i) numParameters /= i;
    while(numParameters > coeffs.size()) {
        std::string varName = getName() + "_extra_coeff_" + std::to_string(coeffs.size());
        coeffs.emplace_back(varName, 0);
        std::cout << "Warning: " << getName() << " created dummy variable " << varName
                  << " (fixed at zero) to account for all terms.\n";
    }
    while(offsets.size() < obses.size())
        offsets.emplace_back(getName() + "_extra_offset_" + std::to_string(offsets.size()), 0);
    for(auto &o : offsets) registerParameter(o);
    for(auto &c : coeffs) registerParameter(c);
    multiDim = true;
    registerFunction("ptr_to_MultiPolynomial", (maxDegree == 0 && coeffs.size() == 1) ? GF_OP_CONSTANT : GF_OP_MULTI_POLYNOMIAL);
    // degree 0 with exactly its one coefficient: the constant leaf (last parameter of the record == p[nobs]); with surplus
    // coefficients (legal: the constructor pads, never truncates) the general opcode keeps reading p[nobs] as upstream
}
fptype PolynomialPdf::integrate(fptype lo, fptype hi) const { // PolynomialPdf.cu:217-243 (marked wrong upstream; unused: no hasAnalyticIntegral)
    fptype lowestDegree = constantsList[0];
    if(center) {
        hi -= center->getValue();
        lo -= center->getValue();
    }
    fptype ret = 0;
    for(size_t i = 0; i < parametersList.size() - (center ? 1 : 0); ++i) {
        fptype powerPlusOne = lowestDegree + i + 1;
        fptype curr = pow(hi, powerPlusOne);
        curr -= pow(lo, powerPlusOne);
        curr /= powerPlusOne;
        ret += parametersList[i].getValue() * curr;
    }
    return ret;
}
fptype PolynomialPdf::getCoefficient(int coef) {
    if(multiDim) throw GeneralError("getCoefficient method of PolynomialPdf not implemented for multi-dimensional polynomials.");
    int lowest = (int)constantsList[0];
    if(coef < lowest) return 0;
    size_t idx = coef - lowest;
    if(idx >= parametersList.size() - (center ? 1 : 0)) return 0;
    return parametersList[idx].getValue();
}

// ---- ArgusPdf (basic/ArgusPdf.cu:52-96) -------------------------------------------------------------------
ArgusPdf::ArgusPdf(std::string n, Observable _x, Variable m0, Variable slope, bool upper)
    : ArgusPdf(n, _x, m0, slope, upper, Variable(n + "powervar", 0.5)) {}
ArgusPdf::ArgusPdf(std::string n, Observable _x, Variable m0, Variable slope, bool upper, Variable power)
    : GooPdf("ArgusPdf", n, _x, m0, slope, power) {
    if(upper)
        registerFunction("ptr_to_Argus_Upper", GF_OP_ARGUS_UPPER);
    else
        registerFunction("ptr_to_Argus_Lower", GF_OP_ARGUS_LOWER);
}
fptype ArgusPdf::integrate(fptype lo, fptype hi) const { // left-edge Riemann sum of the lower form, ArgusPdf.cu:79-96
    double norm = 0;
    fptype m0 = parametersList[0].getValue(), slope = parametersList[1].getValue(), power = parametersList[2].getValue();
    for(int j = 0; j < integrationBins; ++j) {
        double x = hi;
        x -= lo;
        x /= integrationBins;
        x *= j;
        x += lo;
        fptype t = x / m0;
        if(t <= 1) continue;
        t *= t;
        t -= 1;
        norm += x * pow(t, power) * exp(slope * t);
    }
    norm *= ((hi - lo) / integrationBins);
    return norm;
}

// ---- KinLimitBWPdf / ScaledGaussianPdf / JohnsonSUPdf -------------------------------------------------------
KinLimitBWPdf::KinLimitBWPdf(std::string n, Observable _x, Variable mean, Variable width)
    : GooPdf("KinLimitBWPdf", n, _x, mean, width) { // KinLimitBWPdf.cu:58-67
    registerConstant(1.8645);
    registerConstant(0.13957);
    registerFunction("ptr_to_KinLimitBW", GF_OP_KINLIMITBW);
}
void KinLimitBWPdf::setMasses(fptype bigM, fptype smallM) {
    constantsList[0] = bigM;
    constantsList[1] = smallM;
    // the constants live in the flattened program: re-install it so the device sees them (and every cached grid sum,
    // keyed on the program generation, is dropped)
    if(nodeIdx >= 0) {
        Engine &E = Engine::get();
        if(E.root) E.build(E.root);
    }
}
ScaledGaussianPdf::ScaledGaussianPdf(std::string n, Observable _x, Variable mean, Variable sigma, Variable delta, Variable epsilon)
    : GooPdf("ScaledGaussianPdf", n, _x, mean, sigma, delta, epsilon) { // ScaledGaussianPdf.cu:24-31
    registerFunction("ptr_to_ScaledGaussian", GF_OP_SCALED_GAUSSIAN);
}
JohnsonSUPdf::JohnsonSUPdf(std::string n, Observable _x, Variable mean, Variable sigma, Variable gamma, Variable delta)
    : GooPdf("JohnsonSUPdf", n, _x, mean, sigma, gamma, delta) { // JohnsonSUPdf.cu:28-34
    registerFunction("ptr_to_JohnsonSU", GF_OP_JOHNSONSU);
}

// ---- the remaining closed-form shapes of PDFs/basic ---------------------------------------------------------------
BWPdf::BWPdf(std::string n, Observable _x, Variable mean, Variable width)
    : GooPdf("BWPdf", n, _x, mean, width) { // BWPdf.cu:22-27
    registerFunction("ptr_to_BW", GF_OP_BW);
}
BifurGaussPdf::BifurGaussPdf(std::string n, Observable _x, Variable mean, Variable sigmaL, Variable sigmaR)
    : GooPdf("BifurGaussPdf", n, _x, mean, sigmaL, sigmaR) { // BifurGaussPdf.cu:26-31
    registerFunction("ptr_to_BifurGauss", GF_OP_BIFURGAUSS);
}
fptype BifurGaussPdf::integrate(fptype, fptype) const { // BifurGaussPdf.cu:35-45: half of each side's Gaussian normalisation
    fptype sL = par(1), sR = par(2);
    return .5 * (1. / (sqrt(2 * M_PI) * sL)) + .5 * (1. / (sqrt(2 * M_PI) * sR));
}
CrystalBallPdf::CrystalBallPdf(std::string n, Observable _x, Variable mean, Variable sigma, Variable alpha)
    : CrystalBallPdf(n, _x, mean, sigma, alpha, Variable(n + "_n", 2)) {} // CrystalBallPdf.cu:39-40
CrystalBallPdf::CrystalBallPdf(std::string n, Observable _x, Variable mean, Variable sigma, Variable alpha, Variable power)
    : GooPdf("CrystalBallPdf", n, _x, mean, sigma, alpha, power) {
    registerFunction("ptr_to_CrystalBall", GF_OP_CRYSTALBALL);
}
// CrystalBallPdf.cu:50-110 (RooFit's analytic integral): the Gaussian part by erf, the tail by its power-law (or log) primitive
fptype CrystalBallPdf::integrate(fptype lo, fptype hi) const {
    const fptype sqrtPiOver2 = 1.2533141373, sqrt2 = 1.4142135624;
    const fptype mean = par(0), sigma = par(1), alpha = par(2), power = par(3);
    const bool useLog = fabs(power - 1.0) < 1.0e-05;
    fptype tmin = (lo - mean) / sigma, tmax = (hi - mean) / sigma;
    if(alpha < 0) {
        const fptype t = tmin;
        tmin = -tmax;
        tmax = -t;
    }
    const fptype absAlpha = fabs(alpha);
    auto gauss = [&](fptype from, fptype to) { return sigma * sqrtPiOver2 * (erf(to / sqrt2) - erf(from / sqrt2)); };
    if(tmin >= -absAlpha) return gauss(tmin, tmax);
    const fptype a = pow(power / absAlpha, power) * exp(-0.5 * absAlpha * absAlpha), b = power / absAlpha - absAlpha;
    auto tail = [&](fptype to_arg) { // primitive between tmin and the point where (b - t) = to_arg
        if(useLog) return a * sigma * (log(b - tmin) - log(to_arg));
        return a * sigma / (1.0 - power) * (1.0 / (pow(b - tmin, power - 1.0)) - 1.0 / (pow(to_arg, power - 1.0)));
    };
    if(tmax <= -absAlpha) return tail(b - tmax);
    return tail(power / absAlpha) + gauss(-absAlpha, tmax);
}
ExpGausPdf::ExpGausPdf(std::string n, Observable _x, Variable mean, Variable sigma, Variable tau)
    : GooPdf("ExpGausPdf", n, _x, mean, sigma, tau) { // ExpGausPdf.cu:27-32
    registerFunction("ptr_to_ExpGaus", GF_OP_EXPGAUS);
}
NovosibirskPdf::NovosibirskPdf(std::string n, Observable _x, Variable mean, Variable sigma, Variable tail)
    : GooPdf("NovosibirskPdf", n, _x, mean, sigma, tail) { // NovosibirskPdf.cu:42-47
    registerFunction("ptr_to_Novosibirsk", GF_OP_NOVOSIBIRSK);
}
StepPdf::StepPdf(std::string n, Observable _x, Variable x0, int Up)
    : GooPdf("StepPdf", n, _x, x0) { // StepPdf.cu:25-36
    registerConstant(Up);
    registerFunction("ptr_to_Step", GF_OP_STEP);
}
fptype StepPdf::integrate(fptype lo, fptype hi) const { // StepPdf.cu:38-47
    fptype x0 = par(0);
    return (int)constantsList[0] == 1 ? hi - x0 : x0 - lo;
}
TrigThresholdPdf::TrigThresholdPdf(std::string n, Observable _x, Variable thresh, Variable trigConst, Variable linConst, bool upper)
    : GooPdf("TrigThresholdPdf", n, _x, thresh, trigConst, linConst) { // TrigThresholdPdf.cu:84-94
    registerFunction(upper ? "ptr_to_TrigThresholdUpper" : "ptr_to_TrigThresholdLower", upper ? GF_OP_TRIGTHRESH_UPPER : GF_OP_TRIGTHRESH_LOWER);
}
TrigThresholdPdf::TrigThresholdPdf(std::string n, Observable _x, Observable _y, Variable thresh, Variable trigConst, Variable linConst,
                                   Variable massConstant, bool upper)
    : GooPdf("TrigThresholdPdf", n, _x, _y, thresh, trigConst, linConst, massConstant) { // TrigThresholdPdf.cu:96-113
    registerFunction(upper ? "ptr_to_VerySpecialEpisodeTrigThresholdUpper" : "ptr_to_VerySpecialEpisodeTrigThresholdLower",
                     upper ? GF_OP_TRIGTHRESH2_UPPER : GF_OP_TRIGTHRESH2_LOWER);
}
BernsteinPdf::BernsteinPdf(std::string n, Observable obs, std::vector<Variable> coeffs, unsigned int mindeg)
    : GooPdf("BernsteinPdf", n, obs) { // BernsteinPdf.cu:55-72
    registerConstant(int(coeffs.size()));
    registerConstant(obs.getLowerLimit());
    registerConstant(obs.getUpperLimit());
    registerConstant(mindeg);
    for(auto &c : coeffs) registerParameter(c);
    registerFunction("ptr_to_BernsteinPdf", GF_OP_BERNSTEIN);
}
CorrGaussianPdf::CorrGaussianPdf(std::string n, Observable _x, Observable _y, Variable mean1, Variable sigma1, Variable mean2, Variable sigma2,
                                 Variable correlation)
    : GooPdf("CorrGaussianPdf", n, _x, _y, mean1, sigma1, mean2, sigma2, correlation) { // CorrGaussianPdf.cu:31-44
    registerFunction("ptr_to_CorrGaussian", GF_OP_CORRGAUSS);
}
BinTransformPdf::BinTransformPdf(std::string n, std::vector<Observable> obses, std::vector<fptype> limits, std::vector<fptype> binSizes,
                                 std::vector<int> numBins)
    : GooPdf("BinTransformPdf", n) { // BinTransformPdf.cu:31-51
    if(limits.size() < obses.size() || binSizes.size() < obses.size() || numBins.size() < obses.size())
        throw GeneralError("BinTransformPdf {}: one lower limit, bin size and bin count per observable", n);
    for(const auto &o : obses) registerObservable(o);
    observablesList = getObservables();
    for(size_t i = 0; i < obses.size(); ++i) {
        registerConstant(limits[i]);
        registerConstant(binSizes[i]);
        registerConstant(numBins[i]);
    }
    registerFunction("ptr_to_BinTransform", GF_OP_BINTRANSFORM);
}
// ---- SmoothHistogramPdf (basic/SmoothHistogramPdf.cu:107-216) --------------------------------------------------------------
SmoothHistogramPdf::SmoothHistogramPdf(std::string n, BinnedDataSet *hist, Variable smoothing)
    : GooPdf("SmoothHistogramPdf", n, smoothing) {
    if(!hist) throw GeneralError("SmoothHistogramPdf {} needs a histogram", n);
    registerConstant(hist->getObservables().size());
    registerConstant(0); // upstream's index into its table of histograms; the table hangs off the record here
    registerConstant(hist->numVariables());
    for(const Observable &var : hist->getObservables()) {
        registerObservable(var);
        registerConstant(var.getNumBins());
        registerConstant(var.getLowerLimit());
        registerConstant(var.getBinSize());
        nbins.push_back((int)var.getNumBins());
    }
    std::vector<fptype> h(hist->getNumBins());
    for(size_t i = 0; i < h.size(); ++i) {
        h[i] = hist->getBinContent(i);
        totalEvents += h[i];
    }
    if(totalEvents > 0) copyHistogramToDevice(h);
    registerFunction("ptr_to_EvalHistogram", GF_OP_HISTOGRAM);
}
void SmoothHistogramPdf::copyHistogramToDevice(std::vector<fptype> &host_histogram) {
    size_t expected = 1;
    for(int b : nbins) expected *= (size_t)b;
    if(expected != host_histogram.size())
        throw GeneralError("SmoothHistogramPdf {}: histogram has {} bins, expected {}", getName(), host_histogram.size(), expected);
    base     = host_histogram;
    uploaded = false;
}
int SmoothHistogramPdf::emit(detail::ProgramBuilder &B) {
    int slot = B.allocAux();
    return emitRecord(B, GF_OP_HISTOGRAM, 0, 1, {slot});
}
void SmoothHistogramPdf::afterProgram() { uploaded = false; } // a new program: the aux slot is empty again
// the Smoother functor (SmoothHistogramPdf.cu:48-103): every bin with the up to 3^d - 1 neighbours that exist
std::vector<fptype> SmoothHistogramPdf::smoothedTable(fptype smoothing) const {
    const int nv = (int)nbins.size();
    int others = 1;
    for(int v = 0; v < nv; ++v) others *= 3;
    std::vector<fptype> out(base.size());
    for(size_t g = 0; g < base.size(); ++g) {
        fptype central = base[g], total = 0;
        int surrounding = 0;
        for(int i = 0; i < others; ++i) {
            long curr = (long)g, prev = 1, track = (long)g;
            bool off = false;
            int code = i;
            for(int v = 0; v < nv; ++v) {
                const int offset = code % 3 - 1;
                code /= 3;
                curr += offset * prev;
                prev *= nbins[v];
                const long mine = track % nbins[v];
                track /= nbins[v];
                if(mine + offset < 0 || mine + offset >= nbins[v]) off = true;
            }
            if(curr == (long)g || off) continue;
            surrounding++;
            total += base[curr];
        }
        central += total * smoothing;
        central /= (1 + surrounding * smoothing);
        out[g] = central;
    }
    return out;
}
fptype SmoothHistogramPdf::normalize() { // SmoothHistogramPdf.cu:190-214
    if(base.empty()) throw GeneralError("SmoothHistogramPdf {}: no histogram (copyHistogramToDevice)", getName());
    if(nodeIdx < 0) throw GeneralError("SmoothHistogramPdf {}: call setData() on the top-level PDF first", getName());
    const fptype smoothing = par(0);
    std::vector<fptype> table = smoothedTable(smoothing);
    if(!uploaded || uploadedFor != smoothing) {
        Engine &E = Engine::get();
        E.sync();
        Engine::check(gf_table_set(E.ctx(), nodeIdx, table.data(), table.size()), "gf_table_set");
        uploaded    = true;
        uploadedFor = smoothing;
    }
    fptype ret = 0;
    for(fptype v : table) ret += v; // upstream: thrust::reduce over the device vector
    for(size_t v = 0; v < observablesList.size(); ++v) ret *= constantsList[3 + 3 * v + 2];
    setNormalization(1.0 / ret);
    return ret;
}
VoigtianPdf::VoigtianPdf(std::string n, Observable _x, Variable m, Variable s, Variable w)
    : GooPdf("VoigtianPdf", n, _x, m, s, w) { // VoigtianPdf.cu:324-329
    registerFunction("ptr_to_Voigtian", GF_OP_VOIGTIAN);
}
VariableBinTransform1DPdf::VariableBinTransform1DPdf(std::string n, Observable _x, std::vector<fptype> binlimits)
    : GooPdf("VariableBinTransform1DPdf", n, _x) { // VariableBinTransform1DPdf.cu:31-43 (limits exclude the ends of the observable's range)
    registerConstant((fptype)binlimits.size());
    for(fptype l : binlimits) registerConstant(l);
    registerFunction("ptr_to_VarBinTransform1D", GF_OP_VARBINTRANSFORM);
}
LandauPdf::LandauPdf(std::string n, Observable _x, Variable mpv, Variable sigma)
    : GooPdf("LandauPdf", n, _x, mpv, sigma) { // LandauPdf.cu:88-93
    registerFunction("ptr_to_Landau", GF_OP_LANDAU);
}

// ---- AddPdf (combine/AddPdf.cu:85-204) --------------------------------------------------------------------
AddPdf::AddPdf(std::string n, std::vector<Variable> weights, std::vector<PdfBase *> comps)
    : CombinePdf("AddPdf", n)
    , extended(true) {
    if(weights.size() != comps.size() && (weights.size() + 1) != comps.size())
        throw GeneralError("Size of weights {} (+1) != comps {}", weights.size(), comps.size());
    for(PdfBase *p : comps) {
        if(p == nullptr) throw GeneralError("Invalid component");
        components.push_back(p);
    }
    observablesList = getObservables();
    for(auto &w : weights) registerParameter(w);
    if(weights.size() < components.size()) extended = false;
    registerFunction(extended ? "ptr_to_AddPdfsExt" : "ptr_to_AddPdfs", extended ? GF_OP_ADD_EXT : GF_OP_ADD);
}
AddPdf::AddPdf(std::string n, Variable frac1, PdfBase *func1, PdfBase *func2)
    : CombinePdf("AddPdf", n, frac1)
    , extended(false) {
    if(!func1 || !func2) throw GeneralError("Invalid component");
    components.push_back(func1);
    components.push_back(func2);
    observablesList = getObservables();
    registerFunction("ptr_to_AddPdfs", GF_OP_ADD);
}
fptype AddPdf::normalize() { // AddPdf.cu:143-185
    fptype ret = 0, totalWeight = 0;
    for(size_t i = 0; i < components.size() - 1; ++i) {
        fptype weight = parametersList[i].getValue();
        totalWeight += weight;
        fptype curr = components[i]->normalize();
        ret += curr * weight;
    }
    fptype last = components.back()->normalize();
    if(extended) {
        fptype lastWeight = parametersList[components.size() - 1].getValue();
        totalWeight += lastWeight;
        ret += last * lastWeight;
        ret /= totalWeight;
    } else
        ret += (1 - totalWeight) * last;
    setNormalization(1.0);
    if(getCommonNorm())
        for(auto *c : components) c->setNormalization(1.0 / ret);
    return ret;
}
double AddPdf::calculateNLL() { // AddPdf.cu:187-204: extended term  N_exp - N log N_exp
    double ret = GooPdf::calculateNLL();
    if(extended) {
        fptype expEvents = 0;
        for(size_t i = 0; i < components.size(); ++i) expEvents += parametersList[i].getValue();
        ret += (expEvents - numEvents * log(expEvents));
    }
    return ret;
}

// ---- EventWeightedAddPdf (combine/EventWeightedAddPdf.cu:96-160) ---------------------------------------------
EventWeightedAddPdf::EventWeightedAddPdf(std::string n, std::vector<Observable> weights, std::vector<PdfBase *> comps)
    : CombinePdf("EventWeightedAddPdf", n)
    , extended(true) {
    if(weights.size() != comps.size() && (weights.size() + 1) != comps.size())
        throw GeneralError("Size of weights {} (+1) != comps {}", weights.size(), comps.size());
    for(PdfBase *p : comps) {
        if(p == nullptr) throw GeneralError("Invalid component");
        components.push_back(p);
    }
    for(auto &w : weights) registerObservable(w); // before the components' observables: the device op reads the weights first
    if(weights.size() < components.size()) extended = false;
    observablesList = getObservables();
    registerConstant(components.size());
    registerFunction(extended ? "ptr_to_EventWeightedAddPdfsExt" : "ptr_to_EventWeightedAddPdfs", extended ? GF_OP_EVTW_ADD_EXT : GF_OP_EVTW_ADD);
}
fptype EventWeightedAddPdf::normalize() { // per-event weights: only the components can be normalised
    for(PdfBase *comp : components) comp->normalize();
    setNormalization(1.0);
    return 1.0;
}
// ---- MappedPdf (combine/MappedPdf.cu:55-97) ---------------------------------------------------------------------
MappedPdf::MappedPdf(std::string n, GooPdf *m, std::vector<GooPdf *> &t)
    : CombinePdf("MappedPdf", n) {
    if(!m || t.empty()) throw GeneralError("MappedPdf {} needs a mapping function and at least one component", n);
    components.push_back(m);
    for(GooPdf *f : t) {
        if(!f) throw GeneralError("Invalid component");
        components.push_back(f);
    }
    observablesList = getObservables();
    registerConstant(components.size() - 1);
    registerFunction("ptr_to_Mapped", GF_OP_MAPPED);
}
fptype MappedPdf::normalize() {
    fptype ret = 0;
    for(size_t i = 1; i < components.size(); ++i) ret += components[i]->normalize(); // the mapping function is not normalised
    setNormalization(1.0);
    return ret;
}

// ---- CompositePdf (combine/CompositePdf.cu:38-75) -----------------------------------------------------------------
CompositePdf::CompositePdf(std::string n, PdfBase *core, PdfBase *shell)
    : CombinePdf("CompositePdf", n) {
    if(!core || !shell) throw GeneralError("CompositePdf {} needs a core and a shell function", n);
    // one substitution is live at a time on the device: a shell that is itself built on a composite would lose the outer one
    std::vector<PdfBase *> todo{shell};
    while(!todo.empty()) {
        PdfBase *p = todo.back();
        todo.pop_back();
        if(dynamic_cast<CompositePdf *>(p)) throw GeneralError("CompositePdf {}: the shell may not contain another CompositePdf", n);
        for(PdfBase *c : p->getComponents()) todo.push_back(c);
    }
    components.push_back(core);
    components.push_back(shell);
    observablesList = getObservables();
    registerFunction("ptr_to_Composite", GF_OP_COMPOSITE);
}
// core nodes, the argument record (from here on the composite's observable 0 reads as the core's value), shell nodes, the record
int CompositePdf::emit(detail::ProgramBuilder &B) {
    const int first = (int)B.nodes.size();
    const int coreSlot = B.nodes[components[0]->emit(B)].norm_slot;
    emitRecord(B, GF_OP_COMPOSITE_ARG, 1, (int)B.nodes.size() - first + 1, {coreSlot});
    const int shellSlot = B.nodes[components[1]->emit(B)].norm_slot;
    return emitRecord(B, GF_OP_COMPOSITE, 2, (int)B.nodes.size() - first + 1, {coreSlot, shellSlot});
}
fptype CompositePdf::normalize() {
    recursiveSetNormalization(1.0);
    // the components may appear elsewhere in the full function and need their normalisation there (CompositePdf.cu:57-68)
    for(PdfBase *c : components) c->normalize();
    return GooPdf::normalize();
}

// ---- ProdPdf (combine/ProdPdf.cu:30-99) ---------------------------------------------------------------------
ProdPdf::ProdPdf(std::string n, std::vector<PdfBase *> comps)
    : CombinePdf("ProdPdf", n) {
    for(PdfBase *p : comps) {
        if(p == nullptr) throw GeneralError("Invalid component");
        components.push_back(p);
    }
    observablesList = getObservables();
    registerConstant(components.size());
    // the reference compares each component only with its predecessor (ProdPdf.cu:47-63)
    std::vector<Observable> previous;
    for(PdfBase *p : comps) {
        if(varOverlaps) break;
        for(Observable &o : p->getObservables())
            if(std::find(previous.begin(), previous.end(), o) != previous.end()) {
                varOverlaps = true;
                break;
            }
        previous = p->getObservables();
    }
    if(varOverlaps)
        for(PdfBase *p : comps)
            if(p->getSeparateNorm()) varOverlaps = false;
    registerFunction("ptr_to_ProdPdfs", GF_OP_PROD);
}
fptype ProdPdf::normalize() {
    if(varOverlaps) { // shared observable: integrate the product numerically
        recursiveSetNormalization(1.0);
        return GooPdf::normalize();
    }
    for(PdfBase *c : components) c->normalize();
    setNormalization(1.0);
    return 1.0;
}

// ---- ConvolutionPdf (combine/ConvolutionPdf.cu:151-358) -----------------------------------------------------
// The model and resolution sub-PDFs are emitted as auxiliary subtrees; per event the node is a single
// gather from the per-offset table built by gf_conv_fill (see csrc/gf_conv.cu).
ConvolutionPdf::ConvolutionPdf(std::string n, Observable x, GooPdf *m, GooPdf *r)
    : CombinePdf("ConvolutionPdf", n, x)
    , model(m)
    , resolution(r) {
    if(!m || !r) throw GeneralError("Invalid component");
    components.push_back(model);
    components.push_back(resolution);
    registerConstant(-10);
    registerConstant(10);
    registerConstant(0.01);
    registerConstant(0); // lower-bound offset in units of the step (set below): event-independent, kept out of the per-event path
    setLowerBoundOffset();
    registerFunction("ptr_to_ConvolvePdfs", GF_OP_CONVOLVE);
}
ConvolutionPdf::ConvolutionPdf(std::string n, Observable x, GooPdf *m, GooPdf *r, unsigned int others)
    : ConvolutionPdf(std::move(n), x, m, r) {
    // upstream's limit is the size of its shared-memory cache (CONVOLUTION_CACHE_SIZE 512, ConvolutionPdf.cu:218-220)
    if(others > 512) throw GeneralError("numOthers {} must be not be more than the cache size {}", others, 512);
    numOthers = others;
    if(numOthers != 0) properlyInitialised = false;
}
void ConvolutionPdf::registerOthers(std::vector<ConvolutionPdf *> others) {
    const unsigned int numExpectedOthers = numOthers + 1;
    if(numExpectedOthers != others.size()) {
        std::cout << "Problem: " << getName() << " initialized with " << others.size() << " other PDFs, expected " << numExpectedOthers
                  << " (including itself). Returning without initialization.\n";
        return;
    }
    if(std::find(others.begin(), others.end(), this) == others.end()) {
        std::cout << "Problem: " << getName() << " initialized with list that did not include itself. Returning without initialisation.\n";
        return;
    }
    properlyInitialised = true;
}
void ConvolutionPdf::setIntegrationConstants(fptype lo, fptype hi, fptype step) {
    constantsList[0] = lo;
    constantsList[1] = hi;
    constantsList[2] = step;
    setLowerBoundOffset();
    workspacesFilled = false;
    if(nodeIdx >= 0) { // already flattened: refresh the record and the device workspaces
        Engine &E = Engine::get();
        if(E.root) E.build(E.root);
    }
}
void ConvolutionPdf::setLowerBoundOffset() { // ConvolutionPdf.cu:44-46, the part that does not depend on the event
    fptype lbo = constantsList[0] / constantsList[2];
    lbo -= std::floor(lbo);
    constantsList[3] = lbo;
}
void ConvolutionPdf::emitAux(detail::ProgramBuilder &B) {
    model->emitAux(B);
    resolution->emitAux(B);
    model->emit(B);
    resolution->emit(B);
}
int ConvolutionPdf::emit(detail::ProgramBuilder &B) {
    int slot = B.allocAux();
    return emitRecord(B, GF_OP_CONVOLVE, 0, 1, {slot, model->getFunctionIndex(), resolution->getFunctionIndex()});
}
void ConvolutionPdf::afterProgram() {
    Observable dep = observablesList.front();
    Engine &E      = Engine::get();
    Engine::check(gf_conv_setup(E.ctx(), nodeIdx, constantsList[0], constantsList[1], constantsList[2], dep.getLowerLimit(),
                                dep.getUpperLimit()),
                  "gf_conv_setup");
    workspacesFilled = false;
}
fptype ConvolutionPdf::normalize() { // ConvolutionPdf.cu:313-358
    if(nodeIdx < 0) throw GeneralError("ConvolutionPdf {}: call setData() on the top-level PDF first", getName());
    recursiveSetNormalization(1.0);
    bool fm = !workspacesFilled || model->parametersChanged();
    bool fr = !workspacesFilled || resolution->parametersChanged();
    if(fm || fr) {
        Engine &E = Engine::get();
        E.sync();
        Engine::check(gf_conv_fill(E.ctx(), nodeIdx, fm, fr), "gf_conv_fill");
    }
    workspacesFilled = true;
    return GooPdf::normalize();
}

} // namespace GooFit
