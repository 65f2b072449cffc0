// TEST INFRASTRUCTURE (oracle). ROOT-free driver over the UNMODIFIED reference library built by
// oracle/Makefile.ref (libgoofit_ref.so, OpenMP Thrust backend). It is never linked into the product.
//   ref_driver golden <outdir>          write golden fixtures (.npy + manifest.json) for tests/golden
//   ref_driver bench <config> <N> <K> [W] [--data DIR] [--neumaier] [--fit] [--lsmove]
//                                       time K FCN calls on N events (CPU baseline); with --data the events (and the MC
//                                       normalisation sample) are read from DIR/*.npy, so that the product and the
//                                       reference see identical inputs at any size; --neumaier adds a compensated sum of
//                                       the reference's own per-event terms; --fit runs FitManager::fit() and reports the
//                                       wall time, call count and fitted values; --lsmove also times calls on which a
//                                       lineshape mass moves (DP4 / TDDP4)
// Models follow the reference's examples: examples/exponential/exponential.cpp, examples/dalitz/dalitz.cpp:124-322,
// examples/zachFit/zachFit.cpp:143-230, examples/DP4/DP4.cu:51-232, examples/TDDP4/TDDP4.cu:38-156.
#include <goofit/Application.h>
#include <goofit/FitManager.h>
#include <goofit/PDFs/basic/ArgusPdf.h>
#include <goofit/PDFs/basic/ExpPdf.h>
#include <goofit/PDFs/basic/GaussianPdf.h>
#include <goofit/PDFs/basic/JohnsonSUPdf.h>
#include <goofit/PDFs/basic/KinLimitBWPdf.h>
#include <goofit/PDFs/basic/PolynomialPdf.h>
#include <goofit/PDFs/basic/ScaledGaussianPdf.h>
#include <goofit/PDFs/combine/AddPdf.h>
#include <goofit/PDFs/combine/ConvolutionPdf.h>
#include <goofit/PDFs/combine/EventWeightedAddPdf.h>
#include <goofit/PDFs/combine/CompositePdf.h>
#include <goofit/PDFs/basic/BWPdf.h>
#include <goofit/PDFs/basic/BifurGaussPdf.h>
#include <goofit/PDFs/basic/CrystalBallPdf.h>
#include <goofit/PDFs/basic/ExpGausPdf.h>
#include <goofit/PDFs/basic/NovosibirskPdf.h>
#include <goofit/PDFs/basic/StepPdf.h>
#include <goofit/PDFs/basic/TrigThresholdPdf.h>
#include <goofit/PDFs/basic/BernsteinPdf.h>
#include <goofit/PDFs/basic/CorrGaussianPdf.h>
#include <goofit/PDFs/basic/BinTransformPdf.h>
#include <goofit/PDFs/basic/LandauPdf.h>
#include <goofit/PDFs/physics/DalitzVetoPdf.h>
#include <goofit/PDFs/physics/SquareDalitzEffPdf.h>
#include <goofit/PDFs/basic/VoigtianPdf.h>
#include <goofit/PDFs/basic/SmoothHistogramPdf.h>
#include <goofit/PDFs/combine/MappedPdf.h>
#include <goofit/PDFs/combine/ProdPdf.h>
#include <goofit/PDFs/physics/Amp3Body.h>
#include <goofit/PDFs/physics/Amp3Body_TD.h>
#define private public // test infrastructure: read (and, for the at-scale cases, inject) the reference's private MC sample
#define protected public
#include <goofit/PDFs/physics/Amp4Body.h>
#include <goofit/PDFs/physics/detail/NormEvents_4Body_DeviceCached.h>
#include <goofit/PDFs/physics/detail/NormEvents_4Body_Weighted.h>
#include <goofit/PDFs/physics/Amp4Body_TD.h>
#undef private
#undef protected
#include <goofit/PDFs/physics/ThreeGaussResolution.h>
#include <goofit/PDFs/physics/TruthResolution.h>
#include <goofit/PDFs/physics/Amplitude.h>
#include <goofit/PDFs/physics/SpinFactors.h>
#include <goofit/PDFs/physics/lineshapes/FOCUS.h>
#include <goofit/PDFs/physics/lineshapes/Flatte.h>
#include <goofit/PDFs/physics/lineshapes/GLASS.h>
#include <goofit/PDFs/physics/lineshapes/kMatrix.h>
#include <goofit/PDFs/physics/resonances/kMatrix.h>
#include <goofit/PDFs/physics/lineshapes/NonRes.h>
#include <goofit/PDFs/physics/lineshapes/One.h>
#include <goofit/PDFs/physics/lineshapes/Bugg.h>
#include <goofit/PDFs/physics/lineshapes/LASS.h>
#include <goofit/PDFs/physics/lineshapes/RBW.h>
#include <goofit/PDFs/physics/lineshapes/SBW.h>
#include <goofit/PDFs/physics/detail/Dim5.h>
#include <mcbooster/Generate.h>
#include <mcbooster/EvaluateArray.h>
#include <goofit/PDFs/physics/ResonancePdf.h>
#include <goofit/PDFs/physics/resonances/FLATTE.h>
#include <goofit/PDFs/physics/resonances/LASS.h>
#include <goofit/PDFs/physics/resonances/RhoOmegaMix.h>
#include <goofit/PDFs/physics/resonances/Spline.h>
#include <goofit/BinnedDataSet.h>
#include <goofit/FitControl.h>
#include <goofit/UnbinnedDataSet.h>
#include <goofit/Variable.h>

#include <array>
#include <chrono>
#include <climits>
#include <cstdio>
#include <fstream>
#include <functional>
#include <iomanip>
#include <memory>
#include <map>
#include <omp.h>
#include <random>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <vector>

using namespace GooFit;

// ---- tiny .npy writer + manifest -----------------------------------------------------------------
static std::string g_out;
static std::ostringstream g_manifest;
static bool g_first = true;

static void save_npy(const std::string &name, const double *v, size_t rows, size_t cols = 0) {
    std::string shape = cols ? "(" + std::to_string(rows) + ", " + std::to_string(cols) + ")" : "(" + std::to_string(rows) + ",)";
    std::string hdr = "{'descr': '<f8', 'fortran_order': False, 'shape': " + shape + ", }";
    size_t total = 10 + hdr.size() + 1;
    size_t pad = (64 - total % 64) % 64;
    hdr += std::string(pad, ' ') + "\n";
    std::ofstream f(g_out + "/" + name + ".npy", std::ios::binary);
    const char magic[] = "\x93NUMPY\x01\x00";
    f.write(magic, 8);
    unsigned short hl = (unsigned short)hdr.size();
    f.write(reinterpret_cast<const char *>(&hl), 2);
    f.write(hdr.data(), hdr.size());
    f.write(reinterpret_cast<const char *>(v), sizeof(double) * rows * (cols ? cols : 1));
}
static void save_npy(const std::string &name, const std::vector<double> &v) { save_npy(name, v.data(), v.size()); }
static void put(const std::string &key, double v) {
    char buf[64];
    snprintf(buf, sizeof buf, "%.17g", v);
    g_manifest << (g_first ? "" : ",\n") << "  \"" << key << "\": " << buf;
    g_first = false;
}
static void put_vec(const std::string &key, const std::vector<double> &v) {
    g_manifest << (g_first ? "" : ",\n") << "  \"" << key << "\": [";
    for(size_t i = 0; i < v.size(); i++) {
        char buf[64];
        snprintf(buf, sizeof buf, "%.17g", v[i]);
        g_manifest << (i ? ", " : "") << buf;
    }
    g_manifest << "]";
    g_first = false;
}

static std::vector<double> column(UnbinnedDataSet &d, const Observable &o) {
    std::vector<double> r(d.getNumEvents());
    for(size_t i = 0; i < r.size(); i++) r[i] = d.getValue(o, i);
    return r;
}

// per-event Prob values of the root (getCompProbsAtDataPoints()[0]) and the reference's own NLL
static void dump_pdf(const std::string &tag, GooPdf &pdf, UnbinnedDataSet &data, bool with_grid, Observable *gridvar,
                     bool comps = true) {
    pdf.setData(&data);
    double nll = pdf.calculateNLL();
    put(tag + ".nll", nll);
    if(comps) {
        auto probs = pdf.getCompProbsAtDataPoints();
        save_npy(tag + ".prob", probs[0]);
        for(size_t k = 1; k < probs.size(); k++) save_npy(tag + ".comp" + std::to_string(k - 1), probs[k]);
    }
    pdf.setFitControl(std::make_shared<UnbinnedNllFit>());
    pdf.setData(&data);
    put(tag + ".nll_again", pdf.calculateNLL());
    if(with_grid && gridvar) {
        auto g = pdf.evaluateAtPoints(*gridvar);
        save_npy(tag + ".grid", g);
        pdf.setFitControl(std::make_shared<UnbinnedNllFit>());
        pdf.setData(&data);
    }
}

// ---- case: 1-D basic shapes and combiners ------------------------------------------------------
static void fill_1d(UnbinnedDataSet &d, Observable &x, size_t n, unsigned seed, int kind, double a = 1.5) {
    std::mt19937 gen(seed);
    std::uniform_real_distribution<double> uni(x.getLowerLimit(), x.getUpperLimit());
    std::exponential_distribution<double> ex(a);
    std::normal_distribution<double> ga(0.5 * (x.getLowerLimit() + x.getUpperLimit()),
                                        0.15 * (x.getUpperLimit() - x.getLowerLimit()));
    while(d.getNumEvents() < n) {
        double v = kind == 0 ? uni(gen) : (kind == 1 ? x.getLowerLimit() + ex(gen) : ga(gen));
        if(v <= x.getLowerLimit() || v >= x.getUpperLimit()) continue;
        x.setValue(v);
        d.addEvent();
    }
}

static void case_basic() {
    { // exponential example (examples/exponential/exponential.cpp) + NormalizeTest.cpp:42-44
        Observable x{"x", 0, 10};
        UnbinnedDataSet d{x};
        fill_1d(d, x, 2000, 137, 1, 1.0);
        save_npy("exp.x", column(d, x));
        Variable alpha{"alpha", -2, 0.1, -10, 10};
        ExpPdf pdf{"exppdf", x, alpha};
        pdf.setData(&d);
        put("exp.integrate_0_10", pdf.integrate(0, 10));
        put("exp.normalize", pdf.normalize());
        dump_pdf("exp", pdf, d, true, &x);
        FitManager fit{&pdf};
        fit.setVerbosity(0);
        auto fm = fit.fit();
        put("exp.fit_alpha", alpha.getValue());
        put("exp.fit_alpha_err", alpha.getError());
        put("exp.fit_nll", pdf.calculateNLL());
        put("exp.fit_valid", (double)(int)fm.IsValid());
    }
    { // Gaussian (NormalizeTest.cpp:84-99,189-218)
        Observable x{"x", -10, 10};
        UnbinnedDataSet d{x};
        fill_1d(d, x, 1500, 11, 2);
        save_npy("gauss.x", column(d, x));
        Variable m{"m", 1.5, 0.1, -10, 10}, s{"s", 2.5, 0.1, 0.5, 10};
        GaussianPdf pdf{"gauss", x, m, s};
        pdf.setData(&d);
        put("gauss.normalize", pdf.normalize());
        dump_pdf("gauss", pdf, d, true, &x);
    }
    { // offset exponential, exp-poly, polynomial
        Observable x{"x", 0.5, 4};
        UnbinnedDataSet d{x};
        fill_1d(d, x, 1000, 5, 0);
        save_npy("shapes.x", column(d, x));
        Variable a{"a", -0.7, 0.1, -5, 5}, x0{"x0", 0.25, 0.1, -1, 1};
        ExpPdf eo{"expoff", x, a, x0};
        dump_pdf("expoff", eo, d, true, &x);
        std::vector<Variable> w{Variable{"w0", 0.3, 0.1, -5, 5}, Variable{"w1", -0.8, 0.1, -5, 5}, Variable{"w2", 0.05, 0.1, -5, 5}};
        ExpPdf ep{"exppoly", x, w};
        dump_pdf("exppoly", ep, d, true, &x);
        ExpPdf epo{"exppolyoff", x, w, x0};
        dump_pdf("exppolyoff", epo, d, true, &x);
        std::vector<Variable> c{Variable{"c0", 1.0, 0.1, -5, 5}, Variable{"c1", 0.4, 0.1, -5, 5}, Variable{"c2", 0.15, 0.1, -5, 5}};
        PolynomialPdf po{"poly", x, c};
        dump_pdf("poly", po, d, true, &x);
        PolynomialPdf po1{"poly_low1", x, c, 1};
        dump_pdf("poly_low1", po1, d, true, &x);
        Variable jm{"jm", 2.0, 0.1, 0, 4}, js{"js", 0.6, 0.1, 0.1, 4}, jg{"jg", -0.4, 0.1, -3, 3}, jd{"jd", 1.3, 0.1, 0.1, 5};
        JohnsonSUPdf jo{"johnson", x, jm, js, jg, jd};
        dump_pdf("johnson", jo, d, true, &x);
        Variable sm{"sm", 2.1, 0.1, 0, 4}, ss{"ss", 0.7, 0.1, 0.1, 4}, sd{"sd", 0.1, 0.1, -1, 1}, se{"se", 0.2, 0.1, -1, 1};
        ScaledGaussianPdf sg{"scaledgauss", x, sm, ss, sd, se};
        dump_pdf("scaledgauss", sg, d, true, &x);
        Variable am0{"am0", 4.2, 0.1, 4, 5}, asl{"asl", -1.5, 0.1, -10, 10}, apw{"apw", 0.5, 0.1, 0.1, 3};
        ArgusPdf au{"argus_upper", x, am0, asl, true, apw};
        dump_pdf("argus_upper", au, d, true, &x);
        Variable bm0{"bm0", 0.3, 0.1, 0, 0.5};
        ArgusPdf al{"argus_lower", x, bm0, asl, false, apw};
        // the lower Argus grows fast; fine for value parity
        Variable asl2{"asl2", -0.02, 0.01, -1, 1};
        ArgusPdf al2{"argus_lower", x, bm0, asl2, false, apw};
        dump_pdf("argus_lower", al2, d, true, &x);
    }
    { // zachFit-like pieces: KinLimitBW and Argus on the delta-m range (zachFit.cpp:143-230)
        Observable dm{"dm", 0.1395, 0.1665};
        dm.setNumBins(270);
        UnbinnedDataSet d{dm};
        fill_1d(d, dm, 1200, 99, 0);
        save_npy("dm.x", column(d, dm));
        Variable mean{"kl_mean", 0.1454, 0.0001, 0.1450, 0.1458}, width{"kl_width", 0.0004, 0.0001, 0.0001, 0.001};
        KinLimitBWPdf kl{"klbw", dm, mean, width};
        kl.setMasses(1.8645, 0.13957);
        dump_pdf("klbw", kl, d, true, &dm);
        Variable m0{"arg_m0", 0.13957, 0.0001, 0.1, 0.2}, slope{"arg_slope", -1.0, 0.1, -100, 10}, power{"arg_power", 0.5, 0.01, 0.1, 2};
        ArgusPdf bkg{"argus_dm", dm, m0, slope, false, power};
        dump_pdf("argus_dm", bkg, d, true, &dm);
        // convolution: KinLimitBW (x) ScaledGaussian, as zachFit.cpp:213-221 with a coarser step
        Variable dummyzero{"dummyzero", 0}, rs{"res_sigma", 0.00025, 0.00001, 0.0001, 0.001};
        Variable delta{"delta", 0.00002, 0.00001, -0.001, 0.001}, eps{"epsilon", 0.05, 0.01, -0.5, 0.5};
        ScaledGaussianPdf res{"resolution", dm, dummyzero, rs, delta, eps};
        ConvolutionPdf conv{"conv", dm, &kl, &res};
        conv.setIntegrationConstants(0.1395, 0.1665, 0.00003);
        dump_pdf("conv", conv, d, true, &dm, false);
        Variable frac{"bkg_frac", 0.3, 0.01, 0, 1};
        AddPdf total{"total", frac, &bkg, &conv};
        dump_pdf("convadd", total, d, true, &dm);
        // second parameter point: only the resolution moves, then only the model
        total.setData(&d);
        static_cast<GooPdf &>(total).calculateNLL();
        rs.setValue(0.0003); rs.setChanged(true); mean.setChanged(false); width.setChanged(false);
        total.updateVariable(rs, 0.0003); total.updateParameters();
        put("convadd.nll_res_moved", static_cast<GooPdf &>(total).calculateNLL());
        rs.setChanged(false); width.setChanged(true);
        total.updateVariable(width, 0.0005); total.updateParameters();
        put("convadd.nll_model_moved", static_cast<GooPdf &>(total).calculateNLL());
    }
    { // AddNormTest.cpp:16-44 non-extended sum, extended sum, and a 2-D product (NormalizeTest.cpp:147-169)
        Observable x{"x", 0, 10};
        UnbinnedDataSet d{x};
        fill_1d(d, x, 1500, 7, 1, 0.6);
        save_npy("add.x", column(d, x));
        Variable m{"am", 4.0, 0.1, 0, 10}, s{"as", 1.2, 0.1, 0.1, 5}, al{"aalpha", -0.5, 0.1, -5, 5}, f{"afrac", 0.35, 0.1, 0, 1};
        GaussianPdf g{"ag", x, m, s};
        ExpPdf e{"ae", x, al};
        AddPdf add{"add", f, &g, &e};
        dump_pdf("add", add, d, true, &x);
        Variable n1{"n1", 500, 10, 0, 5000}, n2{"n2", 1100, 10, 0, 5000};
        AddPdf ext{"addext", std::vector<Variable>{n1, n2}, std::vector<PdfBase *>{&g, &e}};
        dump_pdf("addext", ext, d, true, &x);
        Observable y{"y", -5, 5};
        UnbinnedDataSet d2{x, y};
        std::mt19937 gen(21);
        std::uniform_real_distribution<double> ux(0, 10), uy(-5, 5);
        for(int i = 0; i < 800; i++) { x.setValue(ux(gen)); y.setValue(uy(gen)); d2.addEvent(); }
        save_npy("prod.x", column(d2, x));
        save_npy("prod.y", column(d2, y));
        Variable my{"my", 0.5, 0.1, -5, 5}, sy{"sy", 1.7, 0.1, 0.1, 5};
        GaussianPdf gy{"gy", y, my, sy};
        ProdPdf prod{"prod", {&g, &gy}};
        dump_pdf("prod", prod, d2, false, nullptr);
    }
}

// ---- dalitz model (examples/dalitz/dalitz.cpp:124-322) -------------------------------------------
// ---- binned fits (tests/simple/BinningTest.cpp:20-188: binned exp / gauss fits; MetricPointer.cu:23-58 metrics) -------
static void case_binned() {
    { // exponential, 100 bins
        Observable x{"x", 0, 10};
        x.setNumBins(100);
        UnbinnedDataSet u{x};
        fill_1d(u, x, 20000, 137, 1, 1.5);
        save_npy("binned.exp_x", column(u, x));
        BinnedDataSet b{x};
        for(size_t i = 0; i < u.getNumEvents(); i++) {
            x.setValue(u.getValue(x, i));
            b.addEvent();
        }
        Variable alpha{"alpha", -2, 0.1, -10, 10};
        ExpPdf pdf{"exppdf", x, alpha};
        pdf.setFitControl(std::make_shared<BinnedNllFit>());
        pdf.setData(&b);
        put("binned.exp_nll", pdf.calculateNLL());
        pdf.setFitControl(std::make_shared<BinnedChisqFit>());
        pdf.setData(&b);
        put("binned.exp_chisq", pdf.calculateNLL());
        pdf.setFitControl(std::make_shared<BinnedErrorFit>());
        for(size_t i = 0; i < b.getNumBins(); i++) b.setBinError(i, std::sqrt(std::max(1.0, (double)b.getBinContent(i))));
        pdf.setData(&b);
        put("binned.exp_binerr", pdf.calculateNLL());
        pdf.setFitControl(std::make_shared<BinnedNllFit>());
        pdf.setData(&b);
        FitManager fit{&pdf};
        fit.setVerbosity(0);
        auto fm = fit.fit();
        put("binned.exp_fit_alpha", alpha.getValue());
        put("binned.exp_fit_alpha_err", alpha.getError());
        put("binned.exp_fit_valid", (double)(int)fm.IsValid());
    }
    { // Gaussian, 200 bins, chi-square fit
        Observable x{"x", -10, 10};
        x.setNumBins(200);
        UnbinnedDataSet u{x};
        fill_1d(u, x, 30000, 11, 2);
        save_npy("binned.gauss_x", column(u, x));
        BinnedDataSet b{x};
        for(size_t i = 0; i < u.getNumEvents(); i++) {
            x.setValue(u.getValue(x, i));
            b.addEvent();
        }
        Variable m{"m", 1.0, 0.1, -10, 10}, sg{"s", 2.0, 0.1, 0.5, 10};
        GaussianPdf pdf{"gauss", x, m, sg};
        pdf.setFitControl(std::make_shared<BinnedNllFit>());
        pdf.setData(&b);
        put("binned.gauss_nll", pdf.calculateNLL());
        pdf.setFitControl(std::make_shared<BinnedChisqFit>());
        pdf.setData(&b);
        put("binned.gauss_chisq", pdf.calculateNLL());
        FitManager fit{&pdf};
        fit.setVerbosity(0);
        auto fm = fit.fit();
        put_vec("binned.gauss_fit", {m.getValue(), sg.getValue()});
        put_vec("binned.gauss_fit_err", {m.getError(), sg.getError()});
        put("binned.gauss_fit_valid", (double)(int)fm.IsValid());
    }
}

static const fptype _mD0 = 1.86484, _mD02 = _mD0 * _mD0, _mD02inv = 1. / _mD02;
static const fptype piPlusMass = 0.13957018, piZeroMass = 0.1349766;

struct DalitzModel {
    Observable m12{"m12", 0, 3}, m13{"m13", 0, 3};
    EventNumber eventNumber{"eventNumber"};
    std::vector<Variable> masses, widths; // one per resonance (NonRes: dummies)
    DecayInfo3 info;
    GooPdf *eff = nullptr;
    Amp3Body *signal = nullptr;
    std::vector<Variable> coefs; // (re, im) interleaved
};

static DalitzModel *make_dalitz(bool float_masses, int grid, int eff_kind = 0) {
    auto *M = new DalitzModel();
    M->m12.setNumBins(grid);
    M->m13.setNumBins(grid);
    DecayInfo3 &d = M->info;
    d.motherMass = _mD0; d.daug1Mass = piZeroMass; d.daug2Mass = piPlusMass; d.daug3Mass = piPlusMass;
    d.meson_radius = 1.5; d.mother_meson_radius = 5.0;
    Variable rhoM("rho_mass", 0.7758, 0.01, 0.7, 0.8), rhoW("rho_width", 0.1503, 0.01, 0.1, 0.2);
    Variable m1450("rhop_1450_mass", 1.465, 0.01, 1.0, 2.0), w1450("rhop_1450_width", 0.400, 0.01, 0.01, 5.0);
    Variable m1700("rhop_1700_mass", 1.720, 0.01, 1.6, 1.9), w1700("rhop_1700_width", 0.250, 0.01, 0.1, 1.0);
    struct R { const char *n; double re, im; Variable m, w; int spin; int cyc; };
    std::vector<R> rs = {
        {"rhop", 1, 0, rhoM, rhoW, 1, PAIR_12},
        {"rho0", 0.565, 0.164, rhoM, rhoW, 1, PAIR_23},
        {"rhom", 0.714, -0.025, rhoM, rhoW, 1, PAIR_13},
        {"rhop_1450", -0.174, -0.117, m1450, w1450, 1, PAIR_12},
        {"rho0_1450", 0.325, 0.057, m1450, w1450, 1, PAIR_23},
        {"rhom_1450", 0.788, 0.226, m1450, w1450, 1, PAIR_13},
        {"rhop_1700", 2.151, -0.658, m1700, w1700, 1, PAIR_12},
        {"rho0_1700", 2.400, -0.734, m1700, w1700, 1, PAIR_23},
        {"rhom_1700", 1.286, -1.532, m1700, w1700, 1, PAIR_13},
        {"f0_980", 0.008 * (-_mD02), -0.013 * (-_mD02), Variable("f0_980_mass", 0.980, 0.01, 0.8, 1.2), Variable("f0_980_width", 0.044, 0.001, 0.001, 0.08), 0, PAIR_23},
        {"f0_1370", -0.058 * (-_mD02), 0.026 * (-_mD02), Variable("f0_1370_mass", 1.434, 0.01, 1.2, 1.6), Variable("f0_1370_width", 0.173, 0.01, 0.01, 0.4), 0, PAIR_23},
        {"f0_1500", 0.057 * (-_mD02), 0.012 * (-_mD02), Variable("f0_1500_mass", 1.507, 0.01, 1.3, 1.7), Variable("f0_1500_width", 0.109, 0.01, 0.01, 0.3), 0, PAIR_23},
        {"f0_1710", 0.070 * (-_mD02), 0.087 * (-_mD02), Variable("f0_1710_mass", 1.714, 0.01, 1.5, 2.9), Variable("f0_1710_width", 0.140, 0.01, 0.01, 0.5), 0, PAIR_23},
        {"f2_1270", -1.027 * (-_mD02inv), -0.162 * (-_mD02inv), Variable("f2_1270_mass", 1.2754, 0.01, 1.0, 1.5), Variable("f2_1270_width", 0.1851, 0.01, 0.01, 0.4), 2, PAIR_23},
        {"f0_600", 0.068 * (-_mD02), 0.010 * (-_mD02), Variable("f0_600_mass", 0.500, 0.01, 0.3, 0.7), Variable("f0_600_width", 0.400, 0.01, 0.2, 0.6), 0, PAIR_23},
    };
    // order as dalitz.cpp:283-298: nonr first
    Variable nre("nonr_amp_real", -0.5595, 0.001, 0, 0), nim("nonr_amp_imag", 0.108761, 0.1, 0, 0);
    M->coefs.push_back(nre); M->coefs.push_back(nim);
    d.resonances.push_back(new Resonances::NonRes("nonr", nre, nim));
    for(size_t k = 0; k < rs.size(); k++) {
        R &r = rs[k];
        Variable re = k == 0 ? Variable(std::string(r.n) + "_amp_real", r.re) : Variable(std::string(r.n) + "_amp_real", r.re, 0.001, 0, 0);
        Variable im = k == 0 ? Variable(std::string(r.n) + "_amp_imag", r.im) : Variable(std::string(r.n) + "_amp_imag", r.im, 0.1, 0, 0);
        M->coefs.push_back(re); M->coefs.push_back(im);
        M->masses.push_back(r.m); M->widths.push_back(r.w);
        d.resonances.push_back(new Resonances::RBW(r.n, re, im, r.m, r.w, r.spin, r.cyc));
    }
    if(!float_masses)
        for(auto *res : d.resonances) res->setParameterConstantness(true);
    Variable one("constantOne", 1), zero("constantZero", 0);
    if(eff_kind == 1) { // veto bands on two of the Dalitz variables
        Variable vM("vetoM", _mD0), v1("veto1", piZeroMass), v2("veto2", piPlusMass), v3("veto3", piPlusMass);
        Variable lo1("vetoLo1", 0.9), hi1("vetoHi1", 1.1), lo2("vetoLo2", 0.2), hi2("vetoHi2", 0.3);
        std::vector<VetoInfo> vetos{VetoInfo(lo1, hi1, PAIR_12), VetoInfo(lo2, hi2, PAIR_23)};
        auto *veto = new DalitzVetoPdf("veto", M->m12, M->m13, vM, v1, v2, v3, vetos);
        M->eff = veto; // one record: upstream takes the LAST function of the table as the efficiency (Amp3Body.cu:196-203) and
                       // skips exactly one record for it outside the Dalitz plot (:72-82), so a composed efficiency is not usable there
    } else if(eff_kind == 2) { // the D -> KS0 h h efficiency polynomial (its own masses are built in)
        Variable c0("sq0", 0.1), c1("sq1", -0.2), c2("sq2", 0.3), c3("sq3", 0.15), c4("sq4", 0.05), c5("sq5", 1.0), c6("sq6", 0.01), c7("sq7", -0.02);
        M->eff = new SquareDalitzEffPdf("sqeff", M->m12, M->m13, c0, c1, c2, c3, c4, c5, c6, c7);
    } else
    M->eff = new PolynomialPdf("constantEff", std::vector<Observable>{M->m12, M->m13}, std::vector<Variable>{one},
                               std::vector<Variable>{zero, zero}, 0);
    M->signal = new Amp3Body("signalPDF", M->m12, M->m13, M->eventNumber, d, M->eff);
    return M;
}

static bool cpu_in_dalitz(double m12, double m13) {
    return inDalitz(m12, m13, _mD0, piZeroMass, piPlusMass, piPlusMass);
}

// uniform events inside the Dalitz boundary (seeded); evtNum = row
static void fill_dalitz_uniform(DalitzModel &M, UnbinnedDataSet &d, size_t n, unsigned long seed) {
    std::mt19937_64 gen(seed);
    std::uniform_real_distribution<double> u(0.0, 3.0);
    while(d.getNumEvents() < n) {
        double a = u(gen), b = u(gen);
        if(!cpu_in_dalitz(a, b)) continue;
        M.m12.setValue(a); M.m13.setValue(b); M.eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
}

static void case_dalitz() {
    DalitzModel *M = make_dalitz(true, 120);
    UnbinnedDataSet d({M->m12, M->m13, M->eventNumber});
    fill_dalitz_uniform(*M, d, 600, 2024);
    // a few points outside the boundary exercise the zero branch (Amp3Body.cu:74-85)
    for(double a : {0.05, 2.9, 1.0}) {
        M->m12.setValue(a); M->m13.setValue(a > 2 ? 2.9 : 0.01); M->eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
    save_npy("dalitz.m12", column(d, M->m12));
    save_npy("dalitz.m13", column(d, M->m13));
    Amp3Body *s = M->signal;
    s->setData(&d);
    s->setDataSize(d.getNumEvents());
    double nll = s->calculateNLL();
    put("dalitz.nll", nll);
    put("dalitz.integral", 1.0 / s->getNormalization());
    std::vector<double> waves;
    for(size_t i = 0; i < M->info.resonances.size(); i++) {
        auto w = s->getCachedWave(i);
        for(auto &c : w) { waves.push_back(c.real()); waves.push_back(c.imag()); }
    }
    save_npy("dalitz.waves", waves.data(), M->info.resonances.size(), 2 * d.getNumEvents());
    { // per-event probabilities through a ProdPdf wrapper, as DalitzPlotter.h:60-63 does
        ProdPdf wrap{"wrap", {s}};
        wrap.setData(&d);
        s->setDataSize(d.getNumEvents());
        auto probs = wrap.getCompProbsAtDataPoints();
        save_npy("dalitz.prob", probs[0]);
    }
    s->setFitControl(std::make_shared<UnbinnedNllFit>());
    s->setData(&d);
    s->setDataSize(d.getNumEvents());
    put("dalitz.nll_again", s->calculateNLL());
    // second point: coefficients move, masses stay -> caches must be reused (Amp3Body.cu:266-275)
    for(auto &m : M->masses) m.setChanged(false);
    for(auto &w : M->widths) w.setChanged(false);
    s->updateVariable(M->coefs[4], 0.6); s->updateVariable(M->coefs[7], -0.1); s->updateParameters();
    put("dalitz.nll_coef_moved", s->calculateNLL());
    put("dalitz.integral_coef_moved", 1.0 / s->getNormalization());
    // third point: one mass and one width move -> waves 0-2 resp. 9 and their integrals redone
    M->masses[0].setChanged(true); M->widths[9].setChanged(true);
    s->updateVariable(M->masses[0], 0.77); s->updateVariable(M->widths[9], 0.05); s->updateParameters();
    put("dalitz.nll_mass_moved", s->calculateNLL());
    put("dalitz.integral_mass_moved", 1.0 / s->getNormalization());
    { // fit fractions from the cached waves at this point (Amp3Body.cu:376-428)
        auto ff = s->fit_fractions();
        std::vector<double> flat;
        for(auto &row : ff) flat.insert(flat.end(), row.begin(), row.end());
        save_npy("dalitz.fitfrac", flat.data(), ff.size(), ff.size());
    }
}

// ---- the other three-body resonance shapes: LASS, FLATTE (with and without the m12<->m13 swap), GS, Gauss, symmetrised RBW ----
// (resonances/{LASS,FLATTE,GS,Gauss,RBW}.cu), on a K K pi-like Dalitz plot with every shape parameter free
static void case_zoo3() {
    Observable m12{"m12", 0, 3}, m13{"m13", 0, 3};
    EventNumber eventNumber{"eventNumber"};
    m12.setNumBins(90);
    m13.setNumBins(90);
    DecayInfo3 info;
    info.motherMass = _mD0; info.daug1Mass = piZeroMass; info.daug2Mass = piPlusMass; info.daug3Mass = piPlusMass;
    info.meson_radius = 1.5; info.mother_meson_radius = 5.0;
    std::vector<Variable> pars;
    auto F = [&](const char *n, double v, double lo, double hi) {
        Variable x(n, v, 0.001, lo, hi);
        pars.push_back(x);
        return x;
    };
    Variable one("one", 1), zero("zero", 0);
    // one Spline only: upstream keeps the second derivatives of every Spline in the same __device__ table
    std::vector<fptype> knots = {0.1, 0.45, 0.8, 1.2, 1.6, 2.0, 2.5, 2.95};
    std::vector<Variable> sp_re, sp_im;
    for(size_t i = 0; i < knots.size(); i++) {
        sp_re.push_back(F(("sp_re" + std::to_string(i)).c_str(), 0.4 * std::cos(1.3 * i) + 0.1 * i, -5, 5));
        sp_im.push_back(F(("sp_im" + std::to_string(i)).c_str(), 0.3 * std::sin(0.9 * i) - 0.05 * i, -5, 5));
    }
    info.resonances = {
        new Resonances::Spline("spline", Variable("sp_r", 0.6), Variable("sp_i", -0.2), knots, sp_re, sp_im, PAIR_13, true),
        new Resonances::LASS("lass", Variable("lass_r", 0.8), Variable("lass_i", -0.3), F("lass_m", 1.425, 1.3, 1.5), F("lass_w", 0.27, 0.1, 0.4),
                             F("lass_a", 2.07, 1, 3), F("lass_r_", 3.32, 2, 4), F("lass_R", 1.0, 0.5, 1.5), F("lass_phiR", -1.9, -3, 3),
                             F("lass_B", 0.96, 0.5, 1.5), F("lass_phiB", 0.0017, -1, 1), 0, PAIR_12, true),
        new Resonances::LASS("lass1", Variable("lass1_r", 0.2), Variable("lass1_i", 0.1), F("lass1_m", 0.896, 0.8, 1.0), F("lass1_w", 0.05, 0.01, 0.1),
                             F("lass1_a", 1.8, 1, 3), F("lass1_r_", 1.76, 1, 4), F("lass1_R", 1.0, 0.5, 1.5), F("lass1_phiR", 0.3, -3, 3),
                             F("lass1_B", 0.6, 0.1, 1.5), F("lass1_phiB", -0.2, -1, 1), 1, PAIR_13, false),
        new Resonances::FLATTE("f0", Variable("f0_r", 0.5), Variable("f0_i", 0.4), F("f0_m", 0.965, 0.9, 1.0), F("f0_g1", 0.165, 0.1, 0.3),
                               F("f0_rg", 4.21, 3, 5), PAIR_23, false),
        new Resonances::FLATTE("f0s", Variable("f0s_r", -0.3), Variable("f0s_i", 0.2), F("f0s_m", 0.98, 0.9, 1.05), F("f0s_g1", 0.2, 0.1, 0.3),
                               F("f0s_rg", 3.5, 3, 5), PAIR_12, true),
        new Resonances::GS("rho_gs", Variable("gs_r", 1.0), Variable("gs_i", 0.0), F("gs_m", 0.7758, 0.7, 0.85), F("gs_w", 0.1503, 0.1, 0.2), 1, PAIR_23),
        new Resonances::Gauss("gauss", Variable("ga_r", 0.1), Variable("ga_i", 0.05), F("ga_m", 1.2, 1.0, 1.4), F("ga_s", 0.1, 0.05, 0.3), PAIR_13),
        new Resonances::RBW("rho_sym", Variable("rs_r", 0.3), Variable("rs_i", -0.2), F("rs_m", 0.7758, 0.7, 0.85), F("rs_w", 0.15, 0.1, 0.2), 1, PAIR_12,
                            true, true),
        new Resonances::RBW("f2_plain", Variable("f2_r", 0.2), Variable("f2_i", 0.3), F("f2_m", 1.2754, 1.2, 1.35), F("f2_w", 0.1851, 0.1, 0.3), 2, PAIR_23,
                            false),
        new Resonances::RhoOmegaMix("rho_omega", Variable("ro_r", 0.7), Variable("ro_i", 0.1), F("om_m", 0.78265, 0.75, 0.8),
                                    F("om_w", 0.00849, 0.005, 0.02), F("ro_m", 0.7690, 0.7, 0.85), F("ro_w", 0.1509, 0.1, 0.2),
                                    F("mix_re", 0.00185, -0.1, 0.1), F("mix_im", -0.0007, -0.1, 0.1), F("mix_delta", 1.0, 0.5, 1.5), 1, PAIR_23,
                                    true, false),
        new Resonances::RhoOmegaMix("rho_omega_u", Variable("rou_r", -0.2), Variable("rou_i", 0.3), F("omu_m", 0.78265, 0.75, 0.8),
                                    F("omu_w", 0.00849, 0.005, 0.02), F("rou_m", 0.7690, 0.7, 0.85), F("rou_w", 0.1509, 0.1, 0.2),
                                    F("mixu_re", 0.003, -1, 1), F("mixu_im", -0.002, -1, 1), F("mixu_delta", 0.8, 0.5, 1.5), 1, PAIR_12,
                                    false, false),
    };
    std::vector<Observable> obs = {m12, m13};
    auto *eff = new PolynomialPdf("constantEff", obs, std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
    auto *s = new Amp3Body("zoo3", m12, m13, eventNumber, info, eff);
    UnbinnedDataSet d({m12, m13, eventNumber});
    std::mt19937_64 gen(77);
    std::uniform_real_distribution<double> u(0, 3);
    while(d.getNumEvents() < 500) {
        double a = u(gen), b = u(gen);
        if(!inDalitz(a, b, _mD0, piZeroMass, piPlusMass, piPlusMass)) continue;
        m12.setValue(a); m13.setValue(b); eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
    save_npy("zoo3.m12", column(d, m12));
    save_npy("zoo3.m13", column(d, m13));
    s->setData(&d);
    s->setDataSize(d.getNumEvents());
    put("zoo3.nll", s->calculateNLL());
    put("zoo3.integral", 1.0 / s->getNormalization());
    std::vector<double> waves;
    for(size_t i = 0; i < info.resonances.size(); i++) {
        auto w = s->getCachedWave(i);
        for(auto &c : w) { waves.push_back(c.real()); waves.push_back(c.imag()); }
    }
    save_npy("zoo3.waves", waves.data(), info.resonances.size(), 2 * d.getNumEvents());
    // the LASS effective range, the FLATTE g1 coupling and the GS mass move
    for(auto &v : pars) v.setChanged(false);
    // (looked up by name: the order in which F() ran inside each constructor call is the compiler's choice)
    auto P = [&](const char *n) -> Variable & {
        for(auto &v : pars)
            if(v.getName() == n) return v;
        throw std::runtime_error(n);
    };
    P("lass_r_").setChanged(true); P("f0_g1").setChanged(true); P("gs_m").setChanged(true); P("mix_im").setChanged(true);
    s->updateVariable(P("lass_r_"), 3.1); s->updateVariable(P("f0_g1"), 0.18); s->updateVariable(P("gs_m"), 0.79);
    s->updateVariable(P("mix_im"), 0.004);
    P("sp_re3").setChanged(true); // the spline's second derivatives stay at their construction values upstream
    s->updateVariable(P("sp_re3"), 0.9);
    s->updateParameters();
    put("zoo3.nll_moved", s->calculateNLL());
    put("zoo3.integral_moved", 1.0 / s->getNormalization());
}

// accept-reject toy from the model itself, then a full Minuit2 fit with the reference's FitManager
static void case_dalitz_fit() {
    DalitzModel *M = make_dalitz(false, 240);
    UnbinnedDataSet cand({M->m12, M->m13, M->eventNumber});
    fill_dalitz_uniform(*M, cand, 60000, 77);
    Amp3Body *s = M->signal;
    std::vector<double> probs;
    {
        ProdPdf wrap{"wrap", {s}};
        wrap.setData(&cand);
        s->setDataSize(cand.getNumEvents());
        probs = wrap.getCompProbsAtDataPoints()[0];
    }
    double pmax = 0;
    for(double p : probs) pmax = std::max(pmax, p);
    std::mt19937_64 gen(78);
    std::uniform_real_distribution<double> u(0, 1);
    UnbinnedDataSet d({M->m12, M->m13, M->eventNumber});
    for(size_t i = 0; i < probs.size(); i++) {
        if(u(gen) * pmax > probs[i]) continue;
        M->m12.setValue(cand.getValue(M->m12, i)); M->m13.setValue(cand.getValue(M->m13, i));
        M->eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
    save_npy("dalitzfit.m12", column(d, M->m12));
    save_npy("dalitzfit.m13", column(d, M->m13));
    s->setFitControl(std::make_shared<UnbinnedNllFit>());
    s->setData(&d);
    s->setDataSize(d.getNumEvents());
    put("dalitzfit.nll_start", s->calculateNLL());
    FitManager fit(s);
    fit.setVerbosity(0);
    auto t0 = std::chrono::steady_clock::now();
    auto fm = fit.fit();
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::vector<double> val, err;
    for(auto &c : M->coefs) { val.push_back(c.getValue()); err.push_back(c.getError()); }
    put_vec("dalitzfit.values", val);
    put_vec("dalitzfit.errors", err);
    put("dalitzfit.fval", fm.Fval());
    put("dalitzfit.nfcn", fm.NFcn());
    put("dalitzfit.valid", (double)(int)fm.IsValid());
    put("dalitzfit.seconds", secs);
}

// ---- DP4 model (examples/DP4/DP4.cu:51-232) ------------------------------------------------------------
struct DP4Model {
    Observable m12{"m12", 0, 3}, m34{"m34", 0, 3}, cos12{"cos12", -1, 1}, cos34{"m34", -1, 1}, phi{"phi", -3.5, 3.5};
    EventNumber eventNumber{"eventNumber", 0, INT_MAX};
    std::vector<Variable> lsvars; // RhoMass, RhoWidth, KstarM, KstarW, ...
    std::vector<Variable> coefs;
    DecayInfo4 info;
    GooPdf *eff = nullptr;
    Amp4Body *dp = nullptr;
    PolynomialPdf *backgr = nullptr;
    AddPdf *signal = nullptr;
};
static DP4Model *make_dp4(unsigned int mcEvents, bool float_ls) {
    auto *M = new DP4Model();
    const fptype mD0 = 1.8645, piP = 0.13957018, Km = 0.493677;
    M->info.meson_radius    = 1.5;
    M->info.particle_masses = {mD0, piP, piP, Km, piP};
    auto V = [&](const char *n, double v, double lo, double hi) {
        Variable x = float_ls ? Variable(n, v, 0.001, lo, hi) : Variable(n, v);
        M->lsvars.push_back(x);
        return x;
    };
    Variable RhoMass = V("rho_mass", 0.77526, 0.6, 0.9), RhoWidth = V("rho_width", 0.1478, 0.05, 0.3);
    Variable KstarM = V("KstarM", 0.89581, 0.8, 1.0), KstarW = V("KstarW", 0.0474, 0.01, 0.1);
    Variable f600M = V("f600M", 0.519, 0.3, 0.7), f600W = V("f600W", 0.454, 0.2, 0.7);
    Variable a1M = V("a1M", 1.23, 1.0, 1.5), a1W = V("a1W", 0.42, 0.2, 0.7);
    Variable K1M = V("K1M", 1.272, 1.1, 1.4), K1W = V("K1W", 0.09, 0.03, 0.2);
    Variable K1430M = V("K1430M", 1.414, 1.2, 1.6), K1430W = V("K1430W", 0.29, 0.1, 0.5);
    auto SF2 = [&](SF_4Body k, std::array<int, 4> a, std::array<int, 4> b) {
        return std::vector<SpinFactor *>{new SpinFactor("SF", k, mD0, a[0], a[1], a[2], a[3]),
                                         new SpinFactor("SF", k, mD0, b[0], b[1], b[2], b[3])};
    };
    auto SFKRS = SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_S, {0, 1, 2, 3}, {3, 1, 2, 0});
    auto SFKRP = SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_P, {0, 1, 2, 3}, {3, 1, 2, 0});
    auto SFKRD = SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_D, {0, 1, 2, 3}, {3, 1, 2, 0});
    auto SFKF  = SF2(SF_4Body::DtoVS_VtoP1P2_StoP3P4, {2, 3, 0, 1}, {2, 0, 3, 1});
    auto SFKK  = SF2(SF_4Body::DtoAP1_AtoSP2_StoP3P4, {0, 1, 3, 2}, {3, 1, 0, 2});
    auto SFK1R = SF2(SF_4Body::DtoAP1_AtoVP2_VtoP3P4, {3, 2, 0, 1}, {0, 2, 3, 1});
    auto SFA1R = SF2(SF_4Body::DtoAP1_AtoVP2_VtoP3P4, {2, 3, 0, 1}, {2, 0, 3, 1});
    auto SFA1RD = SF2(SF_4Body::DtoAP1_AtoVP2Dwave_VtoP3P4, {2, 3, 0, 1}, {2, 0, 3, 1});
    auto rho = [&](unsigned p) { return (Lineshape *)new Lineshapes::RBW("rho(770)", RhoMass, RhoWidth, 1, p); };
    auto kst = [&](unsigned p) { return (Lineshape *)new Lineshapes::RBW("K*(892)bar", KstarM, KstarW, 1, p); };
    std::vector<Lineshape *> LSKRS = {rho(M_12), kst(M_34), rho(M_24), kst(M_13)};
    std::vector<Lineshape *> LSKRP = {rho(M_12), kst(M_34), rho(M_24), kst(M_13)};
    std::vector<Lineshape *> LSKRD = {rho(M_12), kst(M_34), rho(M_24), kst(M_13)};
    std::vector<Lineshape *> LSKF  = {kst(M_34), new Lineshapes::Bugg("f600", f600M, f600W, 0, M_12), kst(M_13),
                                      new Lineshapes::Bugg("f600", f600M, f600W, 0, M_24)};
    std::vector<Lineshape *> LSKK  = {new Lineshapes::SBW("K(1)(1270)bar", K1M, K1W, 1, M_34_2),
                                      new Lineshapes::LASS("K(0)*(1430)bar", K1430M, K1430W, 0, M_34),
                                      new Lineshapes::SBW("K(1)(1270)bar2", K1M, K1W, 1, M_13_2),
                                      new Lineshapes::LASS("K(0)*(1430)bar2", K1430M, K1430W, 0, M_13)};
    std::vector<Lineshape *> LSK1R = {new Lineshapes::SBW("K(1)(1270)bar", K1M, K1W, 0, M_12_3), rho(M_12),
                                      new Lineshapes::SBW("K(1)(1270)bar", K1M, K1W, 0, M_24_3), rho(M_24)};
    std::vector<Lineshape *> LSA1R = {new Lineshapes::SBW("a(1)(1260)+", a1M, a1W, 0, M_12_4), rho(M_12),
                                      new Lineshapes::SBW("a(1)(1260)+", a1M, a1W, 0, M_24_1), rho(M_24)};
    std::vector<Lineshape *> LSA1RD = {new Lineshapes::SBW("a(1)(1260)+", a1M, a1W, 2, M_12_4), rho(M_12),
                                       new Lineshapes::SBW("a(1)(1260)+", a1M, a1W, 2, M_24_1), rho(M_24)};
    auto C = [&](const char *n, double v, bool fl) {
        Variable x = fl ? Variable(n, v, 0.001, 0, 0) : Variable(n, v);
        M->coefs.push_back(x);
        return x;
    };
    Variable AS_re = C("amp_real1", -0.1, true), AS_im = C("amp_imag1", 0.1, true);
    auto *AS = new Amplitude("K*(892)rho(770)_S", AS_re, AS_im, LSKRS, SFKRS, 2);
    Variable AP_re = C("amp_real2", -0.02, true), AP_im = C("amp_imag2", -0.07, true);
    auto *AP = new Amplitude("K*(892)rho(770)_P", AP_re, AP_im, LSKRP, SFKRP, 2);
    Variable AD_re = C("amp_real3", -0.4, true), AD_im = C("amp_imag3", 0.4, true);
    auto *AD = new Amplitude("K*(892)rho(770)_D", AD_re, AD_im, LSKRD, SFKRD, 2);
    Variable KF_re = C("amp_real4", 0.0120787, false), KF_im = C("amp_imag4", -0.0332525, false);
    auto *KF = new Amplitude("KF", KF_re, KF_im, LSKF, SFKF, 2);
    Variable KK_re = C("amp_real5", 0.0109033, false), KK_im = C("amp_imag5", -0.00186219, false);
    auto *KK = new Amplitude("LSKK", KK_re, KK_im, LSKK, SFKK, 2);
    Variable K1R_re = C("amp_real6", -0.10728, false), K1R_im = C("amp_imag6", -0.130213, false);
    auto *K1R = new Amplitude("LSK1R", K1R_re, K1R_im, LSK1R, SFK1R, 2);
    Variable A1R_re = C("amp_real7", 1.0, false), A1R_im = C("amp_imag7", 0.0, false);
    auto *A1R = new Amplitude("LSA1R", A1R_re, A1R_im, LSA1R, SFA1R, 2);
    Variable A1RD_re = C("amp_real8", -0.94921, false), A1RD_im = C("amp_imag8", -1.73407, false);
    auto *A1RD = new Amplitude("LSA1RD", A1RD_re, A1RD_im, LSA1RD, SFA1RD, 2);
    M->info.amplitudes = {KF, AS, AP, AD, KK, K1R, A1R, A1RD};
    Variable one("constantOne", 1), zero("constantZero", 0);
    std::vector<Observable> obs = {M->m12, M->m34, M->cos12, M->cos34, M->phi, M->eventNumber};
    M->eff = new PolynomialPdf("constantEff", obs, std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
    M->dp  = new Amp4Body("test", obs, M->info, M->eff, mcEvents);
    Variable constant("constant", 0.1), constant2("constant2", 1.0);
    M->backgr = new PolynomialPdf("backgr", M->m12, {constant});
    M->signal = new AddPdf("signal", constant2, M->dp, M->backgr);
    return M;
}
// phase-space data events through the shim generator + the reference's own Dim5 map
static void fill_dp4_phsp(DP4Model &M, UnbinnedDataSet &d, size_t n, long seed) {
    std::vector<mcbooster::GReal_t> masses(M.info.particle_masses.begin() + 1, M.info.particle_masses.end());
    mcbooster::PhaseSpace phsp(M.info.particle_masses[0], masses, n, seed);
    phsp.Generate(mcbooster::Vector4R(M.info.particle_masses[0], 0.0, 0.0, 0.0));
    auto d1 = phsp.GetDaughters(0), d2 = phsp.GetDaughters(1), d3 = phsp.GetDaughters(2), d4 = phsp.GetDaughters(3);
    mcbooster::ParticlesSet_d pset = {&d1, &d2, &d3, &d4};
    mcbooster::RealVector_d v0(n), v1(n), v2(n), v3(n), v4(n);
    mcbooster::VariableSet_d vs = {&v0, &v1, &v2, &v3, &v4};
    Dim5 eval;
    mcbooster::EvaluateArray<Dim5>(eval, pset, vs);
    for(size_t i = 0; i < n; i++) {
        M.m12.setValue(v0[i]); M.m34.setValue(v1[i]); M.cos12.setValue(v2[i]); M.cos34.setValue(v3[i]); M.phi.setValue(v4[i]);
        M.eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
}
static std::vector<double> host_copy(const mcbooster::RealVector_d &v) { return std::vector<double>(v.begin(), v.end()); }

static void case_dp4() {
    DP4Model *M = make_dp4(12000, true);
    UnbinnedDataSet d({M->m12, M->m34, M->cos12, M->cos34, M->phi, M->eventNumber});
    fill_dp4_phsp(*M, d, 500, 4242);
    save_npy("dp4.m12", column(d, M->m12)); save_npy("dp4.m34", column(d, M->m34)); save_npy("dp4.cos12", column(d, M->cos12));
    save_npy("dp4.cos34", column(d, M->cos34)); save_npy("dp4.phi", column(d, M->phi));
    Amp4Body *dp = M->dp;
    save_npy("dp4.mc_m12", host_copy(dp->norm_M12)); save_npy("dp4.mc_m34", host_copy(dp->norm_M34));
    save_npy("dp4.mc_cos12", host_copy(dp->norm_CosTheta12)); save_npy("dp4.mc_cos34", host_copy(dp->norm_CosTheta34));
    save_npy("dp4.mc_phi", host_copy(dp->norm_phi));
    put("dp4.mc_events", dp->getMCevents());
    GooPdf *sig = M->signal;
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 6);
    put("dp4.nll", sig->calculateNLL());
    put("dp4.integral", 1.0 / dp->getNormalization());
    { // cached data amplitudes, [evt * nAmp + a]
        std::vector<double> amps;
        thrust::host_vector<fpcomplex> h = *(dp->cachedAMPs);
        for(size_t i = 0; i < h.size(); i++) { amps.push_back(h[i].real()); amps.push_back(h[i].imag()); }
        save_npy("dp4.amps", amps.data(), d.getNumEvents(), 16);
    }
    auto probs = sig->getCompProbsAtDataPoints();
    save_npy("dp4.prob", probs[0]);
    sig->setFitControl(std::make_shared<UnbinnedNllFit>());
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 6);
    put("dp4.nll_again", sig->calculateNLL());
    // coefficients move, lineshapes stay
    for(auto &v : M->lsvars) v.setChanged(false);
    sig->updateVariable(M->coefs[0], -0.12); sig->updateVariable(M->coefs[5], 0.45); sig->updateParameters();
    put("dp4.nll_coef_moved", sig->calculateNLL());
    put("dp4.integral_coef_moved", 1.0 / dp->getNormalization());
    // the K*(892) mass moves: lineshape index 0 (first unique lineshape) is flagged
    M->lsvars[2].setChanged(true);
    sig->updateVariable(M->lsvars[2], 0.8975); sig->updateParameters();
    put("dp4.nll_kstar_moved", sig->calculateNLL());
    put("dp4.integral_kstar_moved", 1.0 / dp->getNormalization());
}

// ---- the remaining four-body lineshapes (One, NonRes, Flatte, GLASS, FOCUS x3) and the nine spin factors DP4 does not use ----
static void case_zoo4() {
    Observable m12{"m12", 0, 3}, m34{"m34", 0, 3}, cos12{"cos12", -1, 1}, cos34{"m34", -1, 1}, phi{"phi", -3.5, 3.5};
    EventNumber eventNumber{"eventNumber", 0, INT_MAX};
    const fptype mD0 = 1.8645, piP = 0.13957018, Km = 0.493677;
    DecayInfo4 info;
    info.meson_radius    = 1.5;
    info.particle_masses = {mD0, piP, piP, Km, piP};
    std::vector<Variable> pars;
    auto F = [&](const char *n, double v, double lo, double hi) {
        Variable x(n, v, 0.001, lo, hi);
        pars.push_back(x);
        return x;
    };
    auto P = [&](const char *n) -> Variable & {
        for(auto &v : pars)
            if(v.getName() == n) return v;
        throw std::runtime_error(n);
    };
    Variable dm = F("dummy_m", 1.0, 0.5, 1.5), dw = F("dummy_w", 0.1, 0.05, 0.5);
    Variable k0m = F("K0s1430_m", 1.425, 1.3, 1.5), k0w = F("K0s1430_w", 0.27, 0.1, 0.4);
    Variable ga = F("glass_a", 2.07, 1, 3), gr = F("glass_r", 3.32, 2, 4), gphiF = F("glass_phiF", 0.2, -1, 1);
    Variable gphiR = F("glass_phiR", -0.3, -1, 1), gF = F("glass_F", 0.9, 0.5, 1.5);
    Variable f0m = F("f0_m", 0.965, 0.9, 1.0);
    Variable rhoM = F("rho_m", 0.77526, 0.6, 0.9), rhoW = F("rho_w", 0.1478, 0.05, 0.3);
    Variable ksM = F("Kstar_m", 0.89581, 0.8, 1.0), ksW = F("Kstar_w", 0.0474, 0.01, 0.1);
    auto SF = [&](SF_4Body k, int a, int b, int c, int d) { return new SpinFactor("SF", k, mD0, a, b, c, d); };
    using LSv = std::vector<Lineshape *>;
    using SFv = std::vector<SpinFactor *>;
    LSv L1 = {new Lineshapes::FOCUS("focus_kpi", Lineshapes::FOCUS::Mod::Kpi, dm, dw, 0, M_34),
              new Lineshapes::NonRes("nonres_p", dm, dw, 1, M_12)};
    SFv S1 = {SF(SF_4Body::DtoPP1_PtoSP2_StoP3P4, 0, 1, 2, 3), SF(SF_4Body::FF_12_34_L1, 0, 1, 2, 3)};
    LSv L2 = {new Lineshapes::FOCUS("focus_keta", Lineshapes::FOCUS::Mod::KEta, dm, dw, 0, M_13),
              new Lineshapes::FOCUS("focus_i32", Lineshapes::FOCUS::Mod::I32, dm, dw, 0, M_24)};
    SFv S2 = {SF(SF_4Body::DtoPP1_PtoVP2_VtoP3P4, 3, 2, 0, 1), SF(SF_4Body::FF_12_34_L2, 0, 1, 2, 3)};
    LSv L3 = {new Lineshapes::GLASS("glass", k0m, k0w, 0, M_34, FF::BL2, 1.5, {ga, gr, gphiF, gphiR, gF}),
              new Lineshapes::One("one", dm, dw, 0, M_12)};
    SFv S3 = {SF(SF_4Body::DtoV1P1_V1toV2P2_V2toP3P4, 0, 1, 2, 3), SF(SF_4Body::FF_123_4_L1, 0, 1, 2, 3)};
    LSv L4 = {new Lineshapes::Flatte("f0_980", f0m, dw, 0, M_12), new Lineshapes::NonRes("nonres_d", dm, dw, 2, M_34_2)};
    SFv S4 = {SF(SF_4Body::DtoTP1_TtoVP2_VtoP3P4, 2, 3, 0, 1), SF(SF_4Body::FF_123_4_L2, 0, 1, 2, 3)};
    LSv L5 = {new Lineshapes::RBW("rho", rhoM, rhoW, 1, M_12), new Lineshapes::RBW("Kstar", ksM, ksW, 1, M_34)};
    SFv S5 = {SF(SF_4Body::ONE, 0, 1, 2, 3)};
    info.amplitudes = {new Amplitude("A1", Variable("a1r", 1.0), Variable("a1i", 0.0), L1, S1, 1),
                       new Amplitude("A2", Variable("a2r", 0.4), Variable("a2i", -0.7), L2, S2, 1),
                       new Amplitude("A3", Variable("a3r", -0.3), Variable("a3i", 0.5), L3, S3, 1),
                       new Amplitude("A4", Variable("a4r", 0.8), Variable("a4i", 0.2), L4, S4, 1),
                       new Amplitude("A5", Variable("a5r", -0.05), Variable("a5i", 0.02), L5, S5, 1)};
    Variable one("constantOne", 1), zero("constantZero", 0);
    std::vector<Observable> obs = {m12, m34, cos12, cos34, phi, eventNumber};
    auto *eff = new PolynomialPdf("constantEff", obs, std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
    auto *dp  = new Amp4Body("zoo4", obs, info, eff, 40000);
    UnbinnedDataSet d(obs);
    { // phase-space data events (shim generator + the reference's Dim5 map), as fill_dp4_phsp
        size_t n = 400;
        std::vector<mcbooster::GReal_t> masses(info.particle_masses.begin() + 1, info.particle_masses.end());
        mcbooster::PhaseSpace phsp(mD0, masses, n, 991);
        phsp.Generate(mcbooster::Vector4R(mD0, 0.0, 0.0, 0.0));
        auto d1 = phsp.GetDaughters(0), d2 = phsp.GetDaughters(1), d3 = phsp.GetDaughters(2), d4 = phsp.GetDaughters(3);
        mcbooster::ParticlesSet_d pset = {&d1, &d2, &d3, &d4};
        mcbooster::RealVector_d v0(n), v1(n), v2(n), v3(n), v4(n);
        mcbooster::VariableSet_d vs = {&v0, &v1, &v2, &v3, &v4};
        Dim5 eval;
        mcbooster::EvaluateArray<Dim5>(eval, pset, vs);
        for(size_t i = 0; i < n; i++) {
            m12.setValue(v0[i]); m34.setValue(v1[i]); cos12.setValue(v2[i]); cos34.setValue(v3[i]); phi.setValue(v4[i]);
            eventNumber.setValue(d.getNumEvents());
            d.addEvent();
        }
    }
    save_npy("zoo4.m12", column(d, m12)); save_npy("zoo4.m34", column(d, m34)); save_npy("zoo4.cos12", column(d, cos12));
    save_npy("zoo4.cos34", column(d, cos34)); save_npy("zoo4.phi", column(d, phi));
    save_npy("zoo4.mc_m12", host_copy(dp->norm_M12)); save_npy("zoo4.mc_m34", host_copy(dp->norm_M34));
    save_npy("zoo4.mc_cos12", host_copy(dp->norm_CosTheta12)); save_npy("zoo4.mc_cos34", host_copy(dp->norm_CosTheta34));
    save_npy("zoo4.mc_phi", host_copy(dp->norm_phi));
    put("zoo4.mc_events", dp->getMCevents());
    dp->setData(&d);
    dp->setDataSize(d.getNumEvents(), 6);
    put("zoo4.nll", dp->calculateNLL());
    put("zoo4.integral", 1.0 / dp->getNormalization());
    {
        std::vector<double> amps;
        thrust::host_vector<fpcomplex> h = *(dp->cachedAMPs);
        for(size_t i = 0; i < h.size(); i++) { amps.push_back(h[i].real()); amps.push_back(h[i].imag()); }
        save_npy("zoo4.amps", amps.data(), d.getNumEvents(), 10);
    }
    for(auto &v : pars) v.setChanged(false);
    P("glass_phiF").setChanged(true); P("f0_m").setChanged(true);
    dp->updateVariable(P("glass_phiF"), 0.35); dp->updateVariable(P("f0_m"), 0.98);
    dp->updateParameters();
    put("zoo4.nll_moved", dp->calculateNLL());
    put("zoo4.integral_moved", 1.0 / dp->getNormalization());
}

// ---- pi pi S-wave K-matrix: three-body resonance (resonances/kMatrix.cu) and four-body lineshape (lineshapes/kMatrix.cu) ----
// Anisovich-Sarantsev pole masses and couplings (the values AmpGen / the BaBar D0 -> KS pi pi analyses use); production vector
// and direct terms arbitrary. Inputs are those of the zoo3 / zoo4 cases (same seeds), so no new event files are written.
struct KMatrixPars {
    Variable sA0{"sA0", -0.15}, sA{"sA", 1.0}, s0_prod{"s0_prod", -0.07, 0.001, -5, 5}, s0_scatt{"s0_scatt", -3.92637};
    std::vector<Variable> beta_r, beta_i, fprod_r, fprod_i, fscat, poles;
    KMatrixPars() {
        const double br[5] = {8.5, 12.2, 29.2, 10.8, 3.1}, bi[5] = {-4.2, 9.6, -2.1, -5.3, 7.7};
        const double fr[5] = {-4.7, -23.2, -1.9, -14.7, 5.0}, fi[5] = {6.4, -10.9, 15.3, 2.0, -3.3};
        const double fs[5] = {0.23399, 0.15044, -0.20545, 0.32825, 0.35412};
        const double g[5][5] = {{0.22889, -0.55377, 0.00000, -0.39899, -0.34639},
                                {0.94128, 0.55095, 0.00000, 0.39065, 0.31503},
                                {0.36856, 0.23888, 0.55639, 0.18340, 0.18681},
                                {0.33650, 0.40907, 0.85679, 0.19906, -0.00984},
                                {0.18171, -0.17558, -0.79658, -0.00355, 0.22358}};
        const double pm[5] = {0.65100, 1.20360, 1.55817, 1.21000, 1.82206};
        for(int i = 0; i < 5; i++) {
            std::string k = std::to_string(i);
            beta_r.emplace_back("beta_r" + k, br[i], 0.01, -100, 100);
            beta_i.emplace_back("beta_i" + k, bi[i], 0.01, -100, 100);
            fprod_r.emplace_back("fprod_r" + k, fr[i]);
            fprod_i.emplace_back("fprod_i" + k, fi[i]);
            fscat.emplace_back("fscat" + k, fs[i]);
            for(int j = 0; j < 5; j++) poles.emplace_back("g" + k + std::to_string(j), g[i][j]);
            poles.emplace_back("pmass" + k, pm[i]);
        }
    }
};
static void case_kmatrix() {
    { // three-body: D0 -> pi0 pi+ pi- with the pi+ pi- S-wave as a K-matrix and a rho+
        Observable m12{"m12", 0, 3}, m13{"m13", 0, 3};
        EventNumber eventNumber{"eventNumber"};
        m12.setNumBins(90);
        m13.setNumBins(90);
        DecayInfo3 info;
        info.motherMass = _mD0; info.daug1Mass = piZeroMass; info.daug2Mass = piPlusMass; info.daug3Mass = piPlusMass;
        info.meson_radius = 1.5; info.mother_meson_radius = 5.0;
        auto *K = new KMatrixPars();
        info.resonances = {
            new Resonances::kMatrix("kmat", Variable("km_r", 0.7), Variable("km_i", -0.4), K->sA0, K->sA, K->s0_prod, K->s0_scatt, K->beta_r, K->beta_i,
                                    K->fprod_r, K->fprod_i, K->fscat, K->poles, 0, PAIR_23),
            new Resonances::RBW("rhop", Variable("rhop_r", 1.0), Variable("rhop_i", 0.0), Variable("rhop_m", 0.7758), Variable("rhop_w", 0.1503), 1, PAIR_12),
        };
        Variable one("one", 1), zero("zero", 0);
        std::vector<Observable> obs = {m12, m13};
        auto *eff = new PolynomialPdf("constantEff", obs, std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
        auto *s = new Amp3Body("kmatrix3", m12, m13, eventNumber, info, eff);
        UnbinnedDataSet d({m12, m13, eventNumber});
        std::mt19937_64 gen(77); // the events of case_zoo3
        std::uniform_real_distribution<double> u(0, 3);
        while(d.getNumEvents() < 500) {
            double a = u(gen), b = u(gen);
            if(!inDalitz(a, b, _mD0, piZeroMass, piPlusMass, piPlusMass)) continue;
            m12.setValue(a); m13.setValue(b); eventNumber.setValue(d.getNumEvents());
            d.addEvent();
        }
        s->setData(&d);
        s->setDataSize(d.getNumEvents());
        put("kmatrix3.nll", s->calculateNLL());
        put("kmatrix3.integral", 1.0 / s->getNormalization());
        std::vector<double> waves;
        auto w = s->getCachedWave(0);
        for(auto &c : w) { waves.push_back(c.real()); waves.push_back(c.imag()); }
        save_npy("kmatrix3.wave", waves.data(), 1, 2 * d.getNumEvents());
        K->beta_r[1].setChanged(true); K->s0_prod.setChanged(true);
        s->updateVariable(K->beta_r[1], 11.0); s->updateVariable(K->s0_prod, -0.2); s->updateParameters();
        put("kmatrix3.nll_moved", s->calculateNLL());
        put("kmatrix3.integral_moved", 1.0 / s->getNormalization());
    }
    { // four-body: D0 -> K- pi+ pi+ pi- with a (pi pi)_S K-matrix x K*(892) amplitude and a rho x K* one
        Observable m12{"m12", 0, 3}, m34{"m34", 0, 3}, cos12{"cos12", -1, 1}, cos34{"m34", -1, 1}, phi{"phi", -3.5, 3.5};
        EventNumber eventNumber{"eventNumber", 0, INT_MAX};
        const fptype mD0 = 1.8645, piP = 0.13957018, Km = 0.493677;
        DecayInfo4 info;
        info.meson_radius    = 1.5;
        info.particle_masses = {mD0, piP, piP, Km, piP};
        auto *K = new KMatrixPars();
        Variable rhoM("rho_m", 0.77526), rhoW("rho_w", 0.1478), ksM("Kstar_m", 0.89581), ksW("Kstar_w", 0.0474);
        std::vector<Lineshape *> L1 = {new Lineshapes::kMatrix("kmat", Variable("km_r", 1), Variable("km_i", 0), K->sA0, K->sA, K->s0_prod, K->s0_scatt,
                                                                K->beta_r, K->beta_i, K->fprod_r, K->fprod_i, K->fscat, K->poles, 0, M_12, FF::BL_Prime, 1.5),
                                       new Lineshapes::RBW("Kstar", ksM, ksW, 1, M_34)};
        std::vector<SpinFactor *> S1 = {new SpinFactor("SF", SF_4Body::DtoVS_VtoP1P2_StoP3P4, mD0, 2, 3, 0, 1)};
        std::vector<Lineshape *> L2 = {new Lineshapes::RBW("rho", rhoM, rhoW, 1, M_12), new Lineshapes::RBW("Kstar", ksM, ksW, 1, M_34)};
        std::vector<SpinFactor *> S2 = {new SpinFactor("SF", SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_S, mD0, 0, 1, 2, 3)};
        info.amplitudes = {new Amplitude("KS", Variable("a1r", 0.02), Variable("a1i", 0.01), L1, S1, 1),
                           new Amplitude("RK", Variable("a2r", 1.0), Variable("a2i", 0.0), L2, S2, 1)};
        Variable one("constantOne", 1), zero("constantZero", 0);
        std::vector<Observable> obs = {m12, m34, cos12, cos34, phi, eventNumber};
        auto *eff = new PolynomialPdf("constantEff", obs, std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
        auto *dp  = new Amp4Body("kmatrix4", obs, info, eff, 40000); // the MC sample of case_zoo4
        UnbinnedDataSet d(obs);
        { // the data events of case_zoo4
            size_t n = 400;
            std::vector<mcbooster::GReal_t> masses(info.particle_masses.begin() + 1, info.particle_masses.end());
            mcbooster::PhaseSpace phsp(mD0, masses, n, 991);
            phsp.Generate(mcbooster::Vector4R(mD0, 0.0, 0.0, 0.0));
            auto d1 = phsp.GetDaughters(0), d2 = phsp.GetDaughters(1), d3 = phsp.GetDaughters(2), d4 = phsp.GetDaughters(3);
            mcbooster::ParticlesSet_d pset = {&d1, &d2, &d3, &d4};
            mcbooster::RealVector_d v0(n), v1(n), v2(n), v3(n), v4(n);
            mcbooster::VariableSet_d vs = {&v0, &v1, &v2, &v3, &v4};
            Dim5 eval;
            mcbooster::EvaluateArray<Dim5>(eval, pset, vs);
            for(size_t i = 0; i < n; i++) {
                m12.setValue(v0[i]); m34.setValue(v1[i]); cos12.setValue(v2[i]); cos34.setValue(v3[i]); phi.setValue(v4[i]);
                eventNumber.setValue(d.getNumEvents());
                d.addEvent();
            }
        }
        put("kmatrix4.mc_events", dp->getMCevents());
        dp->setData(&d);
        dp->setDataSize(d.getNumEvents(), 6);
        put("kmatrix4.nll", dp->calculateNLL());
        put("kmatrix4.integral", 1.0 / dp->getNormalization());
        {
            std::vector<double> amps;
            thrust::host_vector<fpcomplex> h = *(dp->cachedAMPs);
            for(size_t i = 0; i < h.size(); i++) { amps.push_back(h[i].real()); amps.push_back(h[i].imag()); }
            save_npy("kmatrix4.amps", amps.data(), d.getNumEvents(), 4);
        }
        K->beta_r[1].setChanged(true); K->s0_prod.setChanged(true);
        dp->updateVariable(K->beta_r[1], 11.0); dp->updateVariable(K->s0_prod, -0.2); dp->updateParameters();
        put("kmatrix4.nll_moved", dp->calculateNLL());
        put("kmatrix4.integral_moved", 1.0 / dp->getNormalization());
    }
}

// ---- Amp3Body_TD: time-dependent Dalitz amplitude with mixing (Amp3Body_TD.cu:40-760) -----------------------------------------
// D0 -> pi0 pi+ pi- with the three rho(770) charge states, the f0(980) and a non-resonant term; Dalitz points of case_zoo3 (same
// seed), decay time ~ exp(-t / tau), sigma_t uniform, charm tag +1 / -1 / 0 in turn, one mistag value for all events (upstream's
// normalisation reads the Observable object's own value: Amp3Body_TD.cu:735-744).
static void td3_one(const std::string &tag, int three_gauss, double mistag_value) {
    Observable dtime{"dtime", 0, 10}, sigmat{"sigmat", -3, 3}, m12{"m12", 0, 3}, m13{"m13", 0, 3};
    Observable mistag{"mistag", 0, 1}, charmtag{"charmtag", -2, 2};
    EventNumber eventNumber{"eventNumber"};
    m12.setNumBins(120);
    m13.setNumBins(120);
    Variable tau("tau", 0.4101, 0.001, 0.3, 0.5), xm("xmixing", 0.0049, 0.001, -0.1, 0.1), ym("ymixing", 0.0061, 0.001, -0.1, 0.1);
    Variable dx("deltax", 0.0011, 0.001, -0.1, 0.1), dy("deltay", -0.0007, 0.001, -0.1, 0.1);
    DecayInfo3t info{tau, xm, ym, dx, dy};
    info.motherMass = _mD0; info.daug1Mass = piZeroMass; info.daug2Mass = piPlusMass; info.daug3Mass = piPlusMass;
    info.meson_radius = 1.5; info.mother_meson_radius = 5.0;
    Variable rhoM("rho_mass", 0.7758, 0.01, 0.7, 0.8), rhoW("rho_width", 0.1503, 0.01, 0.1, 0.2);
    Variable f0M("f0_980_mass", 0.980, 0.01, 0.8, 1.2), f0W("f0_980_width", 0.044, 0.001, 0.001, 0.08);
    Variable c4r("rhom_amp_real", 0.714, 0.001, 0, 0);
    info.resonances = {
        new Resonances::NonRes("nonr", Variable("nonr_amp_real", -0.5595), Variable("nonr_amp_imag", 0.108761)),
        new Resonances::RBW("rhop", Variable("rhop_amp_real", 1.0), Variable("rhop_amp_imag", 0.0), rhoM, rhoW, 1, PAIR_12),
        new Resonances::RBW("rho0", Variable("rho0_amp_real", 0.565), Variable("rho0_amp_imag", 0.164), rhoM, rhoW, 1, PAIR_23),
        new Resonances::RBW("rhom", c4r, Variable("rhom_amp_imag", -0.025), rhoM, rhoW, 1, PAIR_13),
        new Resonances::RBW("f0_980", Variable("f0_amp_real", 0.008 * (-_mD02)), Variable("f0_amp_imag", -0.013 * (-_mD02)), f0M, f0W, 0, PAIR_23),
    };
    MixingTimeResolution *res;
    if(three_gauss) {
        auto R = [&](const char *n, double v) { return Variable(n, v); };
        res = new ThreeGaussResolution(R("coreFrac", 0.8), R("tailFrac", 0.15), R("coreBias", 0.05), R("coreScale", 1.0), R("tailBias", 0.1),
                                       R("tailScale", 2.0), R("outlBias", -0.05), R("outlScale", 5.0), R("selBias", 0.02), R("cbo", 0.01),
                                       R("cws", 1.1), R("cbo2", -0.02), R("cws2", 0.9));
    } else
        res = new TruthResolution();
    Variable one("constantOne", 1), zero("constantZero", 0);
    auto *eff = new PolynomialPdf("constantEff", std::vector<Observable>{m12, m13}, std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
    auto *dp  = new Amp3Body_TD("tddalitz", dtime, sigmat, m12, m13, eventNumber, info, res, eff, &mistag, &charmtag);
    UnbinnedDataSet d({dtime, sigmat, m12, m13, eventNumber, mistag, charmtag});
    std::mt19937_64 gen(77), gent(4711);
    std::uniform_real_distribution<double> u(0, 3), u1(0, 1);
    std::vector<double> T, S, C;
    while(d.getNumEvents() < 500) {
        double a = u(gen), b = u(gen);
        if(!inDalitz(a, b, _mD0, piZeroMass, piPlusMass, piPlusMass)) continue;
        double t;
        do t = -0.4101 * std::log(1.0 - u1(gent)); while(t >= 10.0);
        double sg = 0.02 + 0.06 * u1(gent);
        int n = d.getNumEvents();
        double ct = n % 3 == 0 ? 1.0 : (n % 3 == 1 ? -1.0 : 0.0);
        m12.setValue(a); m13.setValue(b); eventNumber.setValue(n);
        dtime.setValue(t); sigmat.setValue(sg); mistag.setValue(mistag_value); charmtag.setValue(ct);
        T.push_back(t); S.push_back(sg); C.push_back(ct);
        d.addEvent();
    }
    if(tag == "td3") { // one copy of the inputs (the Dalitz points are those of zoo3)
        save_npy("td3.dtime", T); save_npy("td3.sigmat", S); save_npy("td3.charmtag", C);
    }
    dp->setData(&d);
    dp->setDataSize(d.getNumEvents(), 7);
    put(tag + ".nll", dp->calculateNLL());
    put(tag + ".integral", 1.0 / dp->getNormalization());
    {
        std::vector<double> waves;
        for(size_t i = 0; i < info.resonances.size(); i++) {
            thrust::host_vector<WaveHolder_s> h = dp->getCachedWaveNoCopy(i);
            for(size_t e = 0; e < h.size(); e++) {
                waves.push_back(h[e].ai_real); waves.push_back(h[e].ai_imag); waves.push_back(h[e].bi_real); waves.push_back(h[e].bi_imag);
            }
        }
        save_npy(tag + ".waves", waves.data(), info.resonances.size(), 4 * d.getNumEvents());
    }
    { // per-event probabilities through a ProdPdf wrapper, as for the time-integrated amplitude (DalitzPlotter.h:60-63)
        ProdPdf wrap{"wrap", {dp}};
        wrap.setData(&d);
        dp->setDataSize(d.getNumEvents(), 7);
        auto probs = wrap.getCompProbsAtDataPoints();
        save_npy(tag + ".prob", probs[0]);
    }
    dp->setFitControl(std::make_shared<UnbinnedNllFit>());
    dp->setData(&d);
    dp->setDataSize(d.getNumEvents(), 7);
    put(tag + ".nll_again", dp->calculateNLL());
    // mixing parameters and one coefficient move: caches and integrals are reused
    rhoM.setChanged(false); rhoW.setChanged(false); f0M.setChanged(false); f0W.setChanged(false);
    dp->updateVariable(xm, 0.0123); dp->updateVariable(ym, -0.0071); dp->updateVariable(dx, 0.002); dp->updateVariable(tau, 0.4153);
    dp->updateVariable(c4r, 0.75);
    dp->updateParameters();
    put(tag + ".nll_mix_moved", dp->calculateNLL());
    put(tag + ".integral_mix_moved", 1.0 / dp->getNormalization());
    // the rho mass moves: three of the five waves and their integrals are redone
    rhoM.setChanged(true);
    dp->updateVariable(rhoM, 0.77); dp->updateParameters();
    put(tag + ".nll_mass_moved", dp->calculateNLL());
    put(tag + ".integral_mass_moved", 1.0 / dp->getNormalization());
}
static void case_td3() {
    td3_one("td3", 0, 0.0);
    td3_one("td3mis", 0, 0.05);
    td3_one("td3g3", 1, 0.05);
}

// ---- EventWeightedAddPdf (per-event weights, plain and extended) and MappedPdf (combine/*.cu) ---------------------------------
static void case_combine2() {
    Observable x{"x", -5, 5}, w1{"w1", 0, 1}, w2{"w2", 0, 5}, sel{"sel", 0, 2};
    std::mt19937_64 gen(9001);
    std::normal_distribution<double> g(0.3, 1.2);
    std::uniform_real_distribution<double> u(0, 1);
    std::vector<double> X, W1, W2, S;
    while(X.size() < 700) {
        double v = g(gen);
        if(v <= -5 || v >= 5) continue;
        X.push_back(v); W1.push_back(0.2 + 0.6 * u(gen)); W2.push_back(0.5 + 3.0 * u(gen)); S.push_back((double)(X.size() % 3));
    }
    save_npy("combine2.x", X); save_npy("combine2.w1", W1); save_npy("combine2.w2", W2); save_npy("combine2.sel", S);
    Variable m1("m1", 0.2, 0.1, -3, 3), s1("s1", 1.1, 0.1, 0.3, 4), m2("m2", -0.8, 0.1, -3, 3), s2("s2", 2.0, 0.1, 0.3, 4), al("alpha", -0.3, 0.1, -2, 2);
    { // plain: one weight column, last component takes 1 - w
        GaussianPdf ga("ga", x, m1, s1);
        ExpPdf ex("ex", x, al);
        EventWeightedAddPdf ew("ew", {w1}, {&ga, &ex});
        UnbinnedDataSet d({w1, x});
        for(size_t i = 0; i < X.size(); i++) { x.setValue(X[i]); w1.setValue(W1[i]); d.addEvent(); }
        ew.setData(&d);
        put("combine2.ew_nll", ew.calculateNLL());
        auto probs = ew.getCompProbsAtDataPoints();
        save_npy("combine2.ew_prob", probs[0]);
    }
    { // extended: one weight column per component, divided by the total weight
        GaussianPdf ga("ga", x, m1, s1), gb("gb", x, m2, s2);
        EventWeightedAddPdf ew("ewx", {w1, w2}, {&ga, &gb});
        UnbinnedDataSet d({w1, w2, x});
        for(size_t i = 0; i < X.size(); i++) { x.setValue(X[i]); w1.setValue(W1[i]); w2.setValue(W2[i]); d.addEvent(); }
        ew.setData(&d);
        put("combine2.ewx_nll", ew.calculateNLL());
        auto probs = ew.getCompProbsAtDataPoints();
        save_npy("combine2.ewx_prob", probs[0]);
    }
    { // mapped: sel = 0, 1, 2 picks one of three shapes; the mapping function is the degree-1 polynomial 0 + 1 * sel
        Variable zero("zero", 0), one("one", 1);
        PolynomialPdf map("map", sel, std::vector<Variable>{zero, one});
        GaussianPdf ga("ga", x, m1, s1), gb("gb", x, m2, s2);
        ExpPdf ex("ex", x, al);
        std::vector<GooPdf *> t = {&ga, &gb, &ex};
        MappedPdf mp("mapped", &map, t);
        UnbinnedDataSet d({sel, x});
        for(size_t i = 0; i < X.size(); i++) { x.setValue(X[i]); sel.setValue(S[i]); d.addEvent(); }
        mp.setData(&d);
        put("combine2.map_nll", mp.calculateNLL());
        auto probs = mp.getCompProbsAtDataPoints();
        save_npy("combine2.map_prob", probs[0]);
    }
}

// ---- the remaining closed-form shapes of PDFs/basic: BW, BifurGauss, CrystalBall (both tails), ExpGaus, Novosibirsk (with and
// without tail), Step (both directions), TrigThreshold (four forms), Bernstein, CorrGaussian, Landau (every range), BinTransform
static void case_basic2() {
    Observable x{"x", -4, 6};
    UnbinnedDataSet d{x};
    fill_1d(d, x, 900, 21, 2);
    save_npy("basic2.x", column(d, x));
    Variable m{"b2m", 0.8, 0.1, -4, 6}, w{"b2w", 1.3, 0.1, 0.1, 5};
    { BWPdf p{"bw", x, m, w}; dump_pdf("bw", p, d, true, &x); }
    Variable sl{"b2sl", 0.9, 0.1, 0.1, 5}, sr{"b2sr", 1.7, 0.1, 0.1, 5};
    { BifurGaussPdf p{"bifur", x, m, sl, sr}; dump_pdf("bifur", p, d, true, &x); }
    Variable ap{"b2ap", 1.1, 0.1, 0.1, 5}, an{"b2an", -0.9, 0.1, -5, -0.1}, pw{"b2pw", 2.5, 0.1, 0.5, 10};
    { CrystalBallPdf p{"cbpos", x, m, w, ap, pw}; dump_pdf("cbpos", p, d, true, &x); }
    { CrystalBallPdf p{"cbneg", x, m, w, an}; dump_pdf("cbneg", p, d, true, &x); }
    Variable tau{"b2tau", 0.7, 0.1, 0.05, 5};
    { ExpGausPdf p{"expgaus", x, m, sl, tau}; dump_pdf("expgaus", p, d, true, &x); }
    Variable tail{"b2tail", 0.35, 0.05, -2, 2}, notail{"b2notail", 0.0};
    { NovosibirskPdf p{"novo", x, m, w, tail}; dump_pdf("novo", p, d, true, &x); }
    { NovosibirskPdf p{"novo0", x, m, w, notail}; dump_pdf("novo0", p, d, true, &x); }
    Variable x0{"b2x0", 0.4, 0.1, -4, 6};
    { // a bare step is 0 for half of the events: as a factor of a product with a Gaussian, and summed with one
        GaussianPdf g{"b2g", x, m, sr};
        Variable f{"b2f", 0.3, 0.1, 0, 1};
        StepPdf up{"stepup", x, x0, 1}, dn{"stepdn", x, x0, 0};
        put("stepup.integral", up.integrate(-4, 6));
        put("stepdn.integral", dn.integrate(-4, 6));
        AddPdf a1{"stepup_add", f, &up, &g}, a2{"stepdn_add", f, &dn, &g};
        dump_pdf("stepup_add", a1, d, true, &x);
        dump_pdf("stepdn_add", a2, d, true, &x);
    }
    Variable th{"b2th", 1.0, 0.1, -4, 6}, tc{"b2tc", 0.12, 0.01, 0.01, 2}, lc{"b2lc", 0.6, 0.01, 0, 1};
    { TrigThresholdPdf p{"trigup", x, th, tc, lc, true}; dump_pdf("trigup", p, d, true, &x); }
    { TrigThresholdPdf p{"trigdn", x, th, tc, lc, false}; dump_pdf("trigdn", p, d, true, &x); }
    std::vector<Variable> bc{Variable{"bc0", 1.0, 0.1, 0, 5}, Variable{"bc1", 0.3, 0.1, 0, 5}, Variable{"bc2", 1.6, 0.1, 0, 5}, Variable{"bc3", 0.7, 0.1, 0, 5}};
    { BernsteinPdf p{"bern", x, bc}; dump_pdf("bern", p, d, true, &x); }
    { BernsteinPdf p{"bern1", x, bc, 1}; dump_pdf("bern1", p, d, true, &x); }
    { // Landau: narrow and wide, so that the events reach every range of (x - mpv) / sigma the approximation distinguishes
        Variable mpv{"b2mpv", -0.5, 0.1, -4, 6}, s1{"b2ls1", 0.02, 0.01, 0.001, 5}, s2{"b2ls2", 0.6, 0.1, 0.001, 5};
        LandauPdf p1{"landau_narrow", x, mpv, s1};
        dump_pdf("landau_narrow", p1, d, true, &x);
        LandauPdf p2{"landau_wide", x, mpv, s2};
        dump_pdf("landau_wide", p2, d, true, &x);
    }
    { // two observables: CorrGaussian, the two-observable threshold forms, and the bin number as the mapping function of a MappedPdf
        Observable y{"y", -3, 3};
        y.setNumBins(60);
        x.setNumBins(80);
        UnbinnedDataSet d2({x, y});
        std::mt19937_64 gen(77);
        std::normal_distribution<double> gy(0.2, 1.1);
        for(size_t i = 0; i < d.getNumEvents();) {
            double v = gy(gen);
            if(v <= -3 || v >= 3) continue;
            x.setValue(d.getValue(x, i)); y.setValue(v); d2.addEvent(); ++i;
        }
        save_npy("basic2.y", column(d2, y));
        Variable m2{"b2m2", 0.1, 0.1, -3, 3}, s2{"b2s2", 0.8, 0.1, 0.1, 5}, corr{"b2corr", 0.25, 0.05, -1, 1};
        CorrGaussianPdf cg{"corrgauss", x, y, m, w, m2, s2, corr};
        dump_pdf("corrgauss", cg, d2, false, nullptr);
        Variable mc{"b2mc", 3.0, 0.1, 0, 10};
        TrigThresholdPdf t2u{"trig2up", x, y, th, tc, lc, mc, true}, t2d{"trig2dn", x, y, th, tc, lc, mc, false};
        dump_pdf("trig2up", t2u, d2, false, nullptr);
        dump_pdf("trig2dn", t2d, d2, false, nullptr);
        BinTransformPdf bt{"bintr", {x, y}, {-4, -3}, {5, 3}, {2, 2}}; // 2 x 2 bins -> 0..3
        GaussianPdf g0{"bt0", x, m, sl}, g1{"bt1", x, m, sr}, g2{"bt2", y, m2, s2};
        ExpPdf g3{"bt3", y, corr};
        std::vector<GooPdf *> t{&g0, &g1, &g2, &g3};
        MappedPdf mp{"bintr_mapped", &bt, t};
        dump_pdf("bintr_mapped", mp, d2, false, nullptr);
        x.setNumBins(100);
    }
}

// ---- SmoothHistogramPdf: a 1-D and a 2-D histogram, each bin averaged with its neighbours (basic/SmoothHistogramPdf.cu)
static void case_smoothhist() {
    Observable x{"x", -4, 6}, y{"y", -3, 3};
    x.setNumBins(50);
    y.setNumBins(12);
    std::mt19937_64 gen(515);
    std::normal_distribution<double> gx(0.9, 1.6), gy(-0.3, 1.2);
    BinnedDataSet h1(x);
    Observable x2{"x", -4, 6};
    x2.setNumBins(20);
    BinnedDataSet h2({x2, y});
    std::vector<double> c1, c2;
    for(int i = 0; i < 6000; i++) {
        double vx = gx(gen), vy = gy(gen);
        if(vx <= -4 || vx >= 6 || vy <= -3 || vy >= 3) continue;
        x.setValue(vx); h1.addEvent();
        x2.setValue(vx); y.setValue(vy); h2.addEvent();
    }
    for(size_t i = 0; i < h1.getNumBins(); i++) c1.push_back(h1.getBinContent(i));
    for(size_t i = 0; i < h2.getNumBins(); i++) c2.push_back(h2.getBinContent(i));
    save_npy("smoothhist.h1", c1); save_npy("smoothhist.h2", c2);
    UnbinnedDataSet d{x}, d2({x2, y});
    std::vector<double> X, Y;
    while(X.size() < 700) {
        double vx = gx(gen), vy = gy(gen);
        if(vx <= -4 || vx >= 6 || vy <= -3 || vy >= 3) continue;
        X.push_back(vx); Y.push_back(vy);
        x.setValue(vx); d.addEvent();
        x2.setValue(vx); y.setValue(vy); d2.addEvent();
    }
    save_npy("smoothhist.x", X); save_npy("smoothhist.y", Y);
    Variable sm1{"sm1", 0.4, 0.05, 0, 2}, sm2{"sm2", 0.25, 0.05, 0, 2}, sm0{"sm0", 0.0};
    { SmoothHistogramPdf p{"smooth1", &h1, sm1}; dump_pdf("smooth1", p, d, true, &x); put("smooth1.normalize", p.normalize()); }
    { SmoothHistogramPdf p{"smooth0", &h1, sm0}; dump_pdf("smooth0", p, d, true, &x); }
    { SmoothHistogramPdf p{"smooth2", &h2, sm2}; dump_pdf("smooth2", p, d2, false, nullptr); put("smooth2.normalize", p.normalize()); }
}

// ---- VoigtianPdf: wide, narrow (the h = 0.53 node grid near the real axis) and very narrow with events on the peak (the Pade form
// for |z|^2 < 1e-7), and its two degenerate forms (basic/VoigtianPdf.cu). The constructor's (m, s, w) reach the device as (m, w, s).
static void case_voigt() {
    Observable x{"x", -4, 6};
    UnbinnedDataSet d{x};
    fill_1d(d, x, 800, 31, 2);
    const double peak = 0.8;
    for(double dx : {0.0, 1e-4, -1e-4, 2.5e-5}) { x.setValue(peak + dx); d.addEvent(); }
    save_npy("voigt.x", column(d, x));
    Variable m{"vm", peak, 0.1, -4, 6};
    Variable a1{"va1", 0.9, 0.1, 0, 5}, b1{"vb1", 0.7, 0.1, 0, 5};
    { VoigtianPdf p{"voigt_wide", x, m, a1, b1}; dump_pdf("voigt_wide", p, d, true, &x); }
    Variable a2{"va2", 0.001, 0.0001, 0, 5}, b2{"vb2", 0.5, 0.1, 0, 5};
    { VoigtianPdf p{"voigt_narrow", x, m, a2, b2}; dump_pdf("voigt_narrow", p, d, true, &x); }
    Variable a3{"va3", 0.0002, 0.0001, 0, 5};
    { VoigtianPdf p{"voigt_tiny", x, m, a3, b2}; dump_pdf("voigt_tiny", p, d, true, &x); }
    Variable zero{"vzero", 0.0};
    { VoigtianPdf p{"voigt_bw", x, m, a1, zero}; dump_pdf("voigt_bw", p, d, true, &x); }
    { VoigtianPdf p{"voigt_gauss", x, m, zero, b1}; dump_pdf("voigt_gauss", p, d, true, &x); }
}

// ---- DalitzVetoPdf and SquareDalitzEffPdf on their own, and a vetoed, sloped efficiency inside Amp3Body (physics/DalitzVetoPdf.cu,
// physics/SquareDalitzEffPdf.cu, Amp3Body.cu:60-122)
static void case_dalitzeff() {
    DalitzModel *M = make_dalitz(false, 120, 1);
    UnbinnedDataSet d({M->m12, M->m13, M->eventNumber});
    fill_dalitz_uniform(*M, d, 500, 777);
    for(double a : {0.05, 2.9}) { // outside the boundary
        M->m12.setValue(a); M->m13.setValue(a > 2 ? 2.9 : 0.01); M->eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
    save_npy("dalitzeff.m12", column(d, M->m12));
    save_npy("dalitzeff.m13", column(d, M->m13));
    Amp3Body *sig = M->signal;
    sig->setData(&d);
    sig->setDataSize(d.getNumEvents());
    put("dalitzeff.nll", sig->calculateNLL());
    ProdPdf wrap{"wrap", {sig}};
    wrap.setData(&d);
    sig->setDataSize(d.getNumEvents());
    save_npy("dalitzeff.prob", wrap.getCompProbsAtDataPoints()[0]);
    // (on their own neither helper can be normalised: MetricTaker.cu:73-88 writes every dimension's bin centre into the first
    // observable's slot, m13 stays 0 and no grid point lies inside the Dalitz plot - so both are pinned as the efficiency of an Amp3Body)
    DalitzModel *M2 = make_dalitz(false, 120, 2);
    UnbinnedDataSet e({M2->m12, M2->m13, M2->eventNumber});
    for(size_t i = 0; i < d.getNumEvents(); i++) {
        M2->m12.setValue(d.getValue(M->m12, i)); M2->m13.setValue(d.getValue(M->m13, i)); M2->eventNumber.setValue(i);
        e.addEvent();
    }
    Amp3Body *sig2 = M2->signal;
    sig2->setData(&e);
    sig2->setDataSize(e.getNumEvents());
    put("dalitzsq.nll", sig2->calculateNLL());
    ProdPdf wrap2{"wrap2", {sig2}};
    wrap2.setData(&e);
    sig2->setDataSize(e.getNumEvents());
    save_npy("dalitzsq.prob", wrap2.getCompProbsAtDataPoints()[0]);
}

// ---- CompositePdf: f(x) = h(g(x)), core g = 0.5 + 0.8 x, shell h = Gaussian; a second one with a two-observable core (combine/CompositePdf.cu)
static void case_composite() {
    Observable x{"x", -5, 5}, y{"y", -2, 2};
    std::mt19937_64 gen(4711);
    std::normal_distribution<double> g(-0.2, 1.4);
    std::uniform_real_distribution<double> u(-2, 2);
    std::vector<double> X, Y;
    while(X.size() < 600) {
        double v = g(gen);
        if(v <= -5 || v >= 5) continue;
        X.push_back(v); Y.push_back(u(gen));
    }
    save_npy("composite.x", X); save_npy("composite.y", Y);
    Variable c0("c0", 0.5), c1("c1", 0.8), mean("cmean", 0.3, 0.1, -3, 3), sigma("csigma", 1.5, 0.1, 0.3, 4);
    {
        PolynomialPdf core("core", x, std::vector<Variable>{c0, c1});
        GaussianPdf shell("shell", x, mean, sigma);
        CompositePdf comp("comp", &core, &shell);
        UnbinnedDataSet d({x});
        for(size_t i = 0; i < X.size(); i++) { x.setValue(X[i]); d.addEvent(); }
        comp.setData(&d);
        put("composite.nll", comp.calculateNLL());
        save_npy("composite.prob", comp.getCompProbsAtDataPoints()[0]);
        // as a component of a sum, next to a plain shape of the same observable
        Variable al("calpha", -0.2, 0.1, -2, 2), frac("cfrac", 0.6, 0.1, 0, 1);
        ExpPdf ex("cex", x, al);
        AddPdf sum("csum", frac, &comp, &ex);
        sum.setData(&d);
        put("composite.sum_nll", static_cast<GooPdf &>(sum).calculateNLL());
        save_npy("composite.sum_prob", sum.getCompProbsAtDataPoints()[0]);
    }
    { // core of two observables (x first: its column is the one the shell reads the core value from)
        Variable k0("k0", 0.25), zero("kzero", 0);
        PolynomialPdf core("core2", std::vector<Observable>{x, y}, std::vector<Variable>{k0, c1, c1, k0}, std::vector<Variable>{zero, zero}, 1);
        GaussianPdf shell("shell2", x, mean, sigma);
        CompositePdf comp("comp2", &core, &shell);
        UnbinnedDataSet d({x, y});
        for(size_t i = 0; i < X.size(); i++) { x.setValue(X[i]); y.setValue(Y[i]); d.addEvent(); }
        comp.setData(&d);
        put("composite.two_nll", comp.calculateNLL());
        save_npy("composite.two_prob", comp.getCompProbsAtDataPoints()[0]);
    }
}

static void case_dp4_fit() {
    DP4Model *M = make_dp4(60000, false);
    UnbinnedDataSet cand({M->m12, M->m34, M->cos12, M->cos34, M->phi, M->eventNumber});
    fill_dp4_phsp(*M, cand, 40000, 99);
    GooPdf *sig = M->signal;
    Amp4Body *dp = M->dp;
    // toy at the "correct fixed values" of DP4.cu:140-145
    double truth[6] = {-0.115177, 0.153976, -0.0298697, -0.0722874, -0.452212, 0.426521};
    for(int k = 0; k < 6; k++) { M->coefs[k].setValue(truth[k]); }
    sig->setData(&cand);
    dp->setDataSize(cand.getNumEvents(), 6);
    auto probs = sig->getCompProbsAtDataPoints()[0];
    double pmax = 0;
    for(double p : probs) pmax = std::max(pmax, p);
    std::mt19937_64 gen(100);
    std::uniform_real_distribution<double> u(0, 1);
    UnbinnedDataSet d({M->m12, M->m34, M->cos12, M->cos34, M->phi, M->eventNumber});
    for(size_t i = 0; i < probs.size(); i++) {
        if(u(gen) * pmax > probs[i]) continue;
        M->m12.setValue(cand.getValue(M->m12, i)); M->m34.setValue(cand.getValue(M->m34, i));
        M->cos12.setValue(cand.getValue(M->cos12, i)); M->cos34.setValue(cand.getValue(M->cos34, i));
        M->phi.setValue(cand.getValue(M->phi, i)); M->eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
    save_npy("dp4fit.m12", column(d, M->m12)); save_npy("dp4fit.m34", column(d, M->m34)); save_npy("dp4fit.cos12", column(d, M->cos12));
    save_npy("dp4fit.cos34", column(d, M->cos34)); save_npy("dp4fit.phi", column(d, M->phi));
    save_npy("dp4fit.mc_m12", host_copy(dp->norm_M12)); save_npy("dp4fit.mc_m34", host_copy(dp->norm_M34));
    save_npy("dp4fit.mc_cos12", host_copy(dp->norm_CosTheta12)); save_npy("dp4fit.mc_cos34", host_copy(dp->norm_CosTheta34));
    save_npy("dp4fit.mc_phi", host_copy(dp->norm_phi));
    double start[6] = {-0.1, 0.1, -0.02, -0.07, -0.4, 0.4};
    for(int k = 0; k < 6; k++) M->coefs[k].setValue(start[k]);
    sig->setFitControl(std::make_shared<UnbinnedNllFit>());
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 6);
    FitManager fit(sig);
    fit.setVerbosity(0);
    put("dp4fit.nll_start", (*fit.getFCN())());
    auto fm = fit.fit();
    std::vector<double> val, err;
    for(int k = 0; k < 6; k++) { val.push_back(M->coefs[k].getValue()); err.push_back(M->coefs[k].getError()); }
    put_vec("dp4fit.values", val);
    put_vec("dp4fit.errors", err);
    put("dp4fit.fval", fm.Fval());
    put("dp4fit.nfcn", fm.NFcn());
    put("dp4fit.valid", (double)(int)fm.IsValid());
}

// ---- TDDP4 model (examples/TDDP4/TDDP4.cu:38-156; fit set-up after examples/TDDP4WeightedMC/ToyModel.cu:114-151) ----
struct TDModel {
    Observable m12{"m12", 0, 3}, m34{"m34", 0, 3}, cos12{"cos12", -1, 1}, cos34{"cos34", -1, 1}, phi{"phi", -3.5, 3.5};
    EventNumber eventNumber{"eventNumber", 0, INT_MAX};
    Observable dtime{"dtime", 0, 10}, sigmat{"sigmat", -3, 3};
    std::vector<Variable> mix;    // tau, x, y
    std::vector<Variable> lsvars; // RhoMass, RhoWidth, KstarM, KstarW
    std::vector<Variable> resvars;
    std::vector<double> pm;
    MixingTimeResolution *res = nullptr;
    GooPdf *eff = nullptr;
    Amp4Body_TD *dp = nullptr;
    PolynomialPdf *backgr = nullptr;
    AddPdf *signal = nullptr;
    std::vector<Observable> obs() { return {m12, m34, cos12, cos34, phi, eventNumber, dtime, sigmat}; }
};
// shared: 0 = A and B lists hold distinct Amplitude objects (flags 0/1, as TDDP4.cu); 1 = the P-wave amplitude object is in
// both lists (flag 2: one cached amplitude, two coefficients, Amp4Body_TD.cu:373-381)
static TDModel *make_tddp4(int three_gauss, unsigned int mcToGen, bool float_ls, bool float_mix, int shared = 0,
                           const std::vector<NormEvents_4Body_Base *> *normEvents = nullptr) {
    auto *M = new TDModel();
    const fptype mD0 = 1.8645, piP = 0.13957018, Km = 0.493677;
    Variable tau = float_mix ? Variable("tau", 0.4101, 0.001, 0.100, 0.900) : Variable("tau", 0.4101);
    Variable xm = float_mix ? Variable("xmixing", 0.005, 0.001, -0.15, 0.15) : Variable("xmixing", 0.005);
    Variable ym = float_mix ? Variable("ymixing", 0.01, 0.001, -0.15, 0.15) : Variable("ymixing", 0.01);
    M->mix = {tau, xm, ym};
    DecayInfo4t DI{tau, xm, ym, Variable("SqWStoRSrate", 1.0 / sqrt(300.0))};
    DI.meson_radius    = 1.5;
    DI.particle_masses = {mD0, piP, piP, Km, piP};
    M->pm              = DI.particle_masses;
    auto V = [&](const char *n, double v, double lo, double hi) {
        Variable x = float_ls ? Variable(n, v, 0.001, lo, hi) : Variable(n, v);
        M->lsvars.push_back(x);
        return x;
    };
    Variable RhoMass = V("rho_mass", 0.77526, 0.7, 0.85), RhoWidth = V("rho_width", 0.1478, 0.1, 0.2);
    Variable KstarM = V("KstarM", 0.89581, 0.85, 0.95), KstarW = V("KstarW", 0.0474, 0.03, 0.07);
    auto SF2 = [&](SF_4Body k) {
        return std::vector<SpinFactor *>{new SpinFactor("SF", k, mD0, 0, 1, 2, 3), new SpinFactor("SF", k, mD0, 3, 1, 2, 0)};
    };
    auto LS4 = [&]() {
        return std::vector<Lineshape *>{new Lineshapes::RBW("rho(770)", RhoMass, RhoWidth, 1, M_12, FF::BL2),
                                        new Lineshapes::RBW("K*(892)bar", KstarM, KstarW, 1, M_34, FF::BL2),
                                        new Lineshapes::RBW("rho(770)", RhoMass, RhoWidth, 1, M_24, FF::BL2),
                                        new Lineshapes::RBW("K*(892)bar", KstarM, KstarW, 1, M_13, FF::BL2)};
    };
    auto *S  = new Amplitude("K*(892)rho(770)_S", Variable("amp_real1", 1.0), Variable("amp_imag1", 0.0), LS4(),
                            SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_S), 2);
    auto *P  = new Amplitude("K*(892)rho(770)_P", Variable("amp_real2", 0.526), Variable("amp_imag2", -0.626), LS4(),
                            SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_P), 2);
    auto *D  = new Amplitude("K*(892)rho(770)_D", Variable("amp_real3", 26.537), Variable("amp_imag3", 12.284), LS4(),
                            SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_D), 2);
    auto *SB = new Amplitude("B_K*(892)rho(770)_S", Variable("Bamp_real1", 1.0), Variable("Bamp_imag1", 0), LS4(),
                             SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_S), 2);
    auto *PB = new Amplitude("B_K*(892)rho(770)_P", Variable("Bamp_real2", -0.145), Variable("Bamp_imag2", 0.86), LS4(),
                             SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_P), 2);
    auto *DB = new Amplitude("B_K*(892)rho(770)_D", Variable("Bamp_real3", 24.343), Variable("Bamp_imag3", 5.329), LS4(),
                             SF2(SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_D), 2);
    DI.amplitudes_B = {S, P, D};
    DI.amplitudes   = shared ? std::vector<Amplitude *>{SB, P, DB} : std::vector<Amplitude *>{SB, PB, DB};
    if(three_gauss) {
        auto R = [&](const char *n, double v) {
            Variable x(n, v);
            M->resvars.push_back(x);
            return x;
        };
        M->res = new ThreeGaussResolution(R("coreFrac", 0.8), R("tailFrac", 0.15), R("coreBias", 0.05), R("coreScale", 1.0),
                                          R("tailBias", 0.1), R("tailScale", 2.0), R("outlBias", -0.05), R("outlScale", 5.0),
                                          R("selBias", 0.02), R("cbo", 0.01), R("cws", 1.1), R("cbo2", -0.02), R("cws2", 0.9));
    } else
        M->res = new TruthResolution();
    Variable one("constantOne", 1), zero("constantZero", 0);
    M->eff = new PolynomialPdf("constantEff", M->obs(), std::vector<Variable>{one}, std::vector<Variable>{zero, zero}, 0);
    M->dp  = normEvents ? new Amp4Body_TD("test", M->obs(), DI, M->res, M->eff, nullptr, *normEvents)
                        : new Amp4Body_TD("test", M->obs(), DI, M->res, M->eff, nullptr, (long)7, mcToGen);
    Variable constant("constant", 0.1), constant2("constant2", 1.0);
    M->backgr = new PolynomialPdf("backgr", M->m12, {constant});
    M->signal = new AddPdf("signal", constant2, M->dp, M->backgr);
    return M;
}
// phase-space kinematics + exponential decay time + uniform per-event resolution
static void fill_td_phsp(TDModel &M, UnbinnedDataSet &d, size_t n, long seed) {
    std::vector<mcbooster::GReal_t> masses(M.pm.begin() + 1, M.pm.end());
    mcbooster::PhaseSpace phsp(M.pm[0], masses, n, seed);
    phsp.Generate(mcbooster::Vector4R(M.pm[0], 0.0, 0.0, 0.0));
    auto d1 = phsp.GetDaughters(0), d2 = phsp.GetDaughters(1), d3 = phsp.GetDaughters(2), d4 = phsp.GetDaughters(3);
    mcbooster::ParticlesSet_d pset = {&d1, &d2, &d3, &d4};
    mcbooster::RealVector_d v0(n), v1(n), v2(n), v3(n), v4(n);
    mcbooster::VariableSet_d vs = {&v0, &v1, &v2, &v3, &v4};
    Dim5 eval;
    mcbooster::EvaluateArray<Dim5>(eval, pset, vs);
    std::mt19937_64 gen(seed * 7919 + 1);
    std::uniform_real_distribution<double> u(0, 1);
    for(size_t i = 0; i < n; i++) {
        M.m12.setValue(v0[i]); M.m34.setValue(v1[i]); M.cos12.setValue(v2[i]); M.cos34.setValue(v3[i]); M.phi.setValue(v4[i]);
        double t;
        do t = -0.4101 * std::log(1.0 - u(gen)); while(t >= 10.0);
        M.dtime.setValue(t);
        M.sigmat.setValue(0.02 + 0.06 * u(gen));
        M.eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
}
static void save_td_data(const std::string &tag, TDModel &M, UnbinnedDataSet &d) {
    save_npy(tag + ".m12", column(d, M.m12)); save_npy(tag + ".m34", column(d, M.m34)); save_npy(tag + ".cos12", column(d, M.cos12));
    save_npy(tag + ".cos34", column(d, M.cos34)); save_npy(tag + ".phi", column(d, M.phi));
    save_npy(tag + ".dtime", column(d, M.dtime)); save_npy(tag + ".sigmat", column(d, M.sigmat));
}
static void save_td_mc(const std::string &tag, TDModel &M) {
    auto *ne = dynamic_cast<NormEvents_4Body_DeviceCached *>(M.dp->_normEvents[0].get());
    save_npy(tag + ".mc_m12", host_copy(ne->_norm_M12_d)); save_npy(tag + ".mc_m34", host_copy(ne->_norm_M34_d));
    save_npy(tag + ".mc_cos12", host_copy(ne->_norm_CosTheta12_d)); save_npy(tag + ".mc_cos34", host_copy(ne->_norm_CosTheta34_d));
    save_npy(tag + ".mc_phi", host_copy(ne->_norm_phi_d));
    put(tag + ".mc_events", M.dp->getNumAccNormEvents());
}
static void td_one(const std::string &tag, int three_gauss, int shared) {
    TDModel *M = make_tddp4(three_gauss, 16000, true, true, shared);
    UnbinnedDataSet d(M->obs());
    fill_td_phsp(*M, d, 400, 31337);
    if(tag == "tddp4") { // one copy of the inputs: every td case uses the same data and MC sample
        save_td_data(tag, *M, d);
        save_td_mc(tag, *M);
    }
    GooPdf *sig = M->signal;
    Amp4Body_TD *dp = M->dp;
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 8);
    put(tag + ".nll", sig->calculateNLL());
    put(tag + ".integral", 1.0 / dp->getNormalization());
    {
        std::vector<double> amps;
        thrust::host_vector<fpcomplex> h = *(dp->_cachedAMPs);
        size_t nAmp = h.size() / d.getNumEvents();
        for(size_t i = 0; i < h.size(); i++) { amps.push_back(h[i].real()); amps.push_back(h[i].imag()); }
        save_npy(tag + ".amps", amps.data(), d.getNumEvents(), 2 * nAmp);
    }
    auto probs = sig->getCompProbsAtDataPoints();
    save_npy(tag + ".prob", probs[0]);
    sig->setFitControl(std::make_shared<UnbinnedNllFit>());
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 8);
    put(tag + ".nll_again", sig->calculateNLL());
    // nothing moves, but the parameter block is re-pushed from the Variables (PdfBase.cu:131-140)
    sig->updateParameters();
    put(tag + ".nll_after_update", sig->calculateNLL());
    put(tag + ".integral_after_update", 1.0 / dp->getNormalization());
    // mixing parameters move, lineshapes stay: cached amplitudes are reused, only the MC integral's time factors change
    for(auto &v : M->lsvars) v.setChanged(false);
    sig->updateVariable(M->mix[0], 0.4153); sig->updateVariable(M->mix[1], 0.0123); sig->updateVariable(M->mix[2], -0.0071);
    sig->updateParameters();
    put(tag + ".nll_mix_moved", sig->calculateNLL());
    put(tag + ".integral_mix_moved", 1.0 / dp->getNormalization());
    // the K*(892) mass moves
    M->lsvars[2].setChanged(true);
    sig->updateVariable(M->lsvars[2], 0.8975); sig->updateParameters();
    put(tag + ".nll_kstar_moved", sig->calculateNLL());
    put(tag + ".integral_kstar_moved", 1.0 / dp->getNormalization());
}
// Weighted normalisation events (detail/NormEvents_4Body_Weighted.cu, NormIntegrator_TD_Weighted.cu): the MC sample of the other
// td cases, each event with its own decay time, resolution and weight; the integral is sum_m w_m pdf(x_m, t_m) / sum of weights.
static void td_weighted(const std::string &tag, int three_gauss) {
    TDModel *M0 = make_tddp4(0, 16000, true, true, 0);
    auto *ne0 = dynamic_cast<NormEvents_4Body_DeviceCached *>(M0->dp->_normEvents[0].get());
    auto c0 = host_copy(ne0->_norm_M12_d), c1 = host_copy(ne0->_norm_M34_d), c2 = host_copy(ne0->_norm_CosTheta12_d),
         c3 = host_copy(ne0->_norm_CosTheta34_d), c4 = host_copy(ne0->_norm_phi_d);
    size_t n = c0.size();
    std::mt19937_64 gen(20261105);
    std::uniform_real_distribution<double> u(0, 1);
    std::vector<double> t(n), sg(n), w(n);
    for(size_t i = 0; i < n; i++) {
        do t[i] = -0.4101 * std::log(1.0 - u(gen)); while(t[i] >= 10.0);
        sg[i] = 0.02 + 0.06 * u(gen);
        w[i]  = 1.0 + 2.0 * u(gen); // the reference sums these into an INT accumulator (thrust::reduce(..., 0)): see sum_init_weights
    }
    if(tag == "tddp4w") {
        save_npy("tddp4w.mc_dtime", t); save_npy("tddp4w.mc_sigma", sg); save_npy("tddp4w.mc_weight", w);
    }
    auto H = [](const std::vector<double> &v) { return mcbooster::RealVector_h(v.begin(), v.end()); };
    std::vector<NormEvents_4Body_Base *> ne = {new NormEvents_4Body_Weighted(H(c0), H(c1), H(c2), H(c3), H(c4), H(t), H(sg), H(w))};
    TDModel *M = make_tddp4(three_gauss, 0, true, true, 0, &ne);
    UnbinnedDataSet d(M->obs());
    fill_td_phsp(*M, d, 400, 31337); // the data of the other td cases
    GooPdf *sig = M->signal;
    Amp4Body_TD *dp = M->dp;
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 8);
    put(tag + ".sum_init_weights", dp->getSumInitNormEventWeights());
    put(tag + ".nll", sig->calculateNLL());
    put(tag + ".integral", 1.0 / dp->getNormalization());
    for(auto &v : M->lsvars) v.setChanged(false);
    sig->updateVariable(M->mix[0], 0.4153); sig->updateVariable(M->mix[1], 0.0123); sig->updateVariable(M->mix[2], -0.0071);
    sig->updateParameters();
    put(tag + ".nll_mix_moved", sig->calculateNLL());
    put(tag + ".integral_mix_moved", 1.0 / dp->getNormalization());
    M->lsvars[2].setChanged(true);
    sig->updateVariable(M->lsvars[2], 0.8975); sig->updateParameters();
    put(tag + ".nll_kstar_moved", sig->calculateNLL());
    put(tag + ".integral_kstar_moved", 1.0 / dp->getNormalization());
}
static void case_tddp4w() {
    td_weighted("tddp4w", 0);
    td_weighted("tddp4wg3", 1);
}
static void case_tddp4() {
    td_one("tddp4", 0, 0);
    td_one("tddp4g3", 1, 0);
    td_one("tddp4sh", 0, 1);
}

static void case_tddp4_fit() {
    TDModel *M = make_tddp4(0, 200000, false, true);
    UnbinnedDataSet cand(M->obs());
    fill_td_phsp(*M, cand, 400000, 2718);
    GooPdf *sig = M->signal;
    Amp4Body_TD *dp = M->dp;
    // toy: accept-reject of the candidates' kinematics against the model evaluated at each candidate's own decay time,
    // divided by the exponential the times were drawn from (so the accepted sample follows the full time-dependent model)
    sig->setData(&cand);
    dp->setDataSize(cand.getNumEvents(), 8);
    auto probs = sig->getCompProbsAtDataPoints()[0];
    std::vector<double> w(probs.size());
    double wmax = 0;
    for(size_t i = 0; i < probs.size(); i++) {
        w[i] = probs[i] / std::exp(-cand.getValue(M->dtime, i) / 0.4101);
        wmax = std::max(wmax, w[i]);
    }
    std::mt19937_64 gen(101);
    std::uniform_real_distribution<double> u(0, 1);
    UnbinnedDataSet d(M->obs());
    for(size_t i = 0; i < probs.size(); i++) {
        if(u(gen) * wmax > w[i]) continue;
        M->m12.setValue(cand.getValue(M->m12, i)); M->m34.setValue(cand.getValue(M->m34, i));
        M->cos12.setValue(cand.getValue(M->cos12, i)); M->cos34.setValue(cand.getValue(M->cos34, i));
        M->phi.setValue(cand.getValue(M->phi, i)); M->dtime.setValue(cand.getValue(M->dtime, i));
        M->sigmat.setValue(cand.getValue(M->sigmat, i)); M->eventNumber.setValue(d.getNumEvents());
        d.addEvent();
    }
    save_td_data("tddp4fit", *M, d);
    save_td_mc("tddp4fit", *M);
    M->mix[0].setValue(0.40); M->mix[1].setValue(0.01); M->mix[2].setValue(0.0);
    sig->setFitControl(std::make_shared<UnbinnedNllFit>());
    sig->setData(&d);
    dp->setDataSize(d.getNumEvents(), 8);
    FitManager fit(sig);
    fit.setVerbosity(0);
    put("tddp4fit.nll_start", (*fit.getFCN())());
    auto fm = fit.fit();
    std::vector<double> val, err;
    for(int k = 0; k < 3; k++) { val.push_back(M->mix[k].getValue()); err.push_back(M->mix[k].getError()); }
    put_vec("tddp4fit.values", val);
    put_vec("tddp4fit.errors", err);
    put("tddp4fit.fval", fm.Fval());
    put("tddp4fit.nfcn", fm.NFcn());
    put("tddp4fit.valid", (double)(int)fm.IsValid());
}

// ---- .npy reader (float64, C order, 1-D) --------------------------------------------------------------------------------
static std::vector<double> load_npy(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if(!f) throw std::runtime_error("cannot open " + path);
    char magic[8];
    f.read(magic, 8);
    if(std::string(magic, 6) != "\x93NUMPY") throw std::runtime_error(path + ": not a .npy file");
    size_t hl = 0;
    if(magic[6] == 1) {
        unsigned short h;
        f.read(reinterpret_cast<char *>(&h), 2);
        hl = h;
    } else {
        unsigned int h;
        f.read(reinterpret_cast<char *>(&h), 4);
        hl = h;
    }
    std::string hdr(hl, ' ');
    f.read(&hdr[0], hl);
    if(hdr.find("'<f8'") == std::string::npos || hdr.find("'fortran_order': False") == std::string::npos)
        throw std::runtime_error(path + ": need little-endian float64 in C order");
    size_t p0 = hdr.find("'shape': (") + 10, p1 = hdr.find(')', p0), total = 1;
    std::stringstream ss(hdr.substr(p0, p1 - p0));
    std::string tok;
    bool any = false;
    while(std::getline(ss, tok, ',')) {
        if(tok.find_first_of("0123456789") == std::string::npos) continue;
        total *= std::stoul(tok);
        any = true;
    }
    if(!any) total = 1;
    std::vector<double> v(total);
    f.read(reinterpret_cast<char *>(v.data()), sizeof(double) * total);
    if(!f) throw std::runtime_error(path + ": truncated");
    return v;
}
static bool file_exists(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0;
}

// compensated sum (MathUtils.cpp:11-30 is the reference's own Neumaier summation; restated here for a host vector)
static double neumaier(const std::vector<double> &v) {
    double sum = 0, c = 0;
    for(double x : v) {
        double t = sum + x;
        c += std::fabs(sum) >= std::fabs(x) ? (sum - t) + x : (x - t) + sum;
        sum = t;
    }
    return sum + c;
}
static double neumaier_nll_of_probs(const std::vector<double> &probs) {
    std::vector<double> t(probs.size());
    for(size_t i = 0; i < probs.size(); i++) t[i] = probs[i] > 0 ? -std::log(probs[i]) : 0.0; // MetricPointer.cu:13-21
    return neumaier(t);
}

// ---- the full zachFit tree (examples/zachFit/zachFit.cpp:143-230), MC-stage variables fixed at their start values --------
struct ZachModel {
    Observable dm{"dm", 0.1395, 0.1665};
    std::vector<Variable> free; // width_bw, delta, epsilon, slope, bkg_frac
    std::vector<GooPdf *> keep;
    AddPdf *total = nullptr;
};
static ZachModel *make_zach(bool johnson) {
    auto *Z = new ZachModel();
    Observable &dm = Z->dm;
    dm.setNumBins(2700);
    Variable mean1("kpi_mc_mean1", 0.145402), mean2("kpi_mc_mean2", 0.145465), mean3("kpi_mc_mean3", 0.145404);
    Variable sigma1("kpi_mc_sigma1", 0.00010), sigma2("kpi_mc_sigma2", 0.00075), sigma3("kpi_mc_sigma3", 0.00020);
    Variable pimass("kpi_mc_pimass", 0.13957), aslope("kpi_mc_aslope", -20.0), apower("kpi_mc_apower", 1.3);
    Variable gfrac1("kpi_mc_gfrac1", 0.65), gfrac2("kpi_mc_gfrac2", 0.02), afrac("kpi_mc_afrac", 0.005);
    auto *argus = new ArgusPdf("argus", dm, pimass, aslope, false, apower);
    Variable dummyzero("kpi_rd_dummyzero", 0);
    Variable delta("kpi_rd_delta", 0.000002, -0.00005, 0.00005), epsilon("kpi_rd_epsilon", 0.05, -0.1, 0.2);
    auto *r1 = new ScaledGaussianPdf("resolution1", dm, dummyzero, sigma1, delta, epsilon);
    auto *r2 = new ScaledGaussianPdf("resolution2", dm, dummyzero, sigma2, delta, epsilon);
    auto *r3 = new ScaledGaussianPdf("resolution3", dm, dummyzero, sigma3, delta, epsilon);
    Variable width_bw("kpi_rd_width_bw", 0.0001, 0.00001, 0.0005);
    GooPdf *m1, *m2, *m3;
    if(johnson) { // BASELINE.json words the config as "JohnsonSU (x) resolution": same tree, JohnsonSU as the convolved model
        Variable jg("kpi_rd_jsu_gamma", -0.3), jd("kpi_rd_jsu_delta", 1.1);
        m1 = new JohnsonSUPdf("jsu1", dm, mean1, width_bw, jg, jd);
        m2 = new JohnsonSUPdf("jsu2", dm, mean2, width_bw, jg, jd);
        m3 = new JohnsonSUPdf("jsu3", dm, mean3, width_bw, jg, jd);
    } else {
        m1 = new KinLimitBWPdf("rbw1", dm, mean1, width_bw);
        m2 = new KinLimitBWPdf("rbw2", dm, mean2, width_bw);
        m3 = new KinLimitBWPdf("rbw3", dm, mean3, width_bw);
    }
    auto *s1 = new ConvolutionPdf("signal1", dm, m1, r1);
    auto *s2 = new ConvolutionPdf("signal2", dm, m2, r2);
    auto *s3 = new ConvolutionPdf("signal3", dm, m3, r3);
    s1->setIntegrationConstants(0.1395, 0.1665, 0.0000027);
    s2->setIntegrationConstants(0.1395, 0.1665, 0.0000027);
    s3->setIntegrationConstants(0.1395, 0.1665, 0.0000027);
    auto *signal = new AddPdf("signal", {gfrac1, gfrac2, afrac}, {s1, s2, argus, s3});
    Variable slope("kpi_rd_slope", -1.0, 0.1, -35.0, 25.0);
    auto *bkg = new ArgusPdf("bkg", dm, pimass, slope, false);
    Variable bkg_frac("kpi_rd_bkg_frac", 0.03, 0.0, 0.3);
    Z->total = new AddPdf("total", {bkg_frac}, {bkg, signal});
    Z->free  = {width_bw, delta, epsilon, slope, bkg_frac};
    return Z;
}

// ---- CPU baseline timing / at-scale parity values ------------------------------------------------------------------------
// Each step = one Minuit-style call through the reference's own FCN::operator(): Params::from_minuit_vector semantics
// (Params.cpp:47-57) then calculateNLL(). The movers follow the schedule bench.py uses on the product side:
// value_j(k) = start_j * (1 + 1e-3 * ((k + j) % 3 - 1)).
struct ScaleOpts {
    std::string data;
    bool neumaier = false, fit = false, lsmove = false;
};
static int bench(const std::string &config, size_t N, int K, int W, const ScaleOpts &opt) {
    int threads = omp_get_max_threads();
    double t_setup = 0, t_first = 0;
    std::vector<double> ms;
    double last = 0, first_nll = 0, neum = 0, ms_lsmove = 0, last_lsmove = 0;
    size_t n_mc = 0;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto secs = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
    auto col = [&](const std::string &name) {
        std::vector<double> v = load_npy(opt.data + "/" + name + ".npy");
        if(N && v.size() > N) v.resize(N);
        return v;
    };
    GooPdf *top = nullptr;
    std::vector<Variable> movers, report;          // Minuit-vector entries that move per call; values reported after a fit
    std::function<void()> after_setdata = [] {};   // setDataSize of the amplitude PDFs
    std::function<std::vector<double>()> probs;    // per-event normalised probabilities (for the compensated sum)
    Variable *ls_var = nullptr;                    // lineshape mass moved by --lsmove
    std::unique_ptr<UnbinnedDataSet> d;
    if(config == "dalitz") {
        DalitzModel *M = make_dalitz(false, 240);
        d.reset(new UnbinnedDataSet({M->m12, M->m13, M->eventNumber}));
        if(opt.data.empty())
            fill_dalitz_uniform(*M, *d, N, 12345);
        else {
            auto a = col("m12"), b = col("m13");
            for(size_t i = 0; i < a.size(); i++) {
                M->m12.setValue(a[i]); M->m13.setValue(b[i]); M->eventNumber.setValue(i);
                d->addEvent();
            }
        }
        top = M->signal;
        after_setdata = [M, &d] { M->signal->setDataSize(d->getNumEvents()); };
        for(size_t c = 2; c < M->coefs.size(); c++) movers.push_back(M->coefs[c]);
        report = M->coefs;
        probs = [M, &d] {
            ProdPdf wrap{"wrap", {M->signal}};
            wrap.setData(d.get());
            M->signal->setDataSize(d->getNumEvents());
            return wrap.getCompProbsAtDataPoints()[0];
        };
    } else if(config == "exponential") {
        auto *x = new Observable("x", 0, 10);
        d.reset(new UnbinnedDataSet(*x));
        if(opt.data.empty())
            fill_1d(*d, *x, N, 137, 1, 1.0);
        else
            for(double v : col("x")) { x->setValue(v); d->addEvent(); }
        auto *alpha = new Variable("alpha", -2, 0.1, -10, 10);
        auto *pdf = new ExpPdf("exppdf", *x, *alpha);
        top = pdf;
        movers = {*alpha};
        report = movers;
        probs = [pdf] { return pdf->getCompProbsAtDataPoints()[0]; };
    } else if(config == "zachfit" || config == "zachfit_jsu") {
        ZachModel *Z = make_zach(config == "zachfit_jsu");
        d.reset(new UnbinnedDataSet(Z->dm));
        if(opt.data.empty()) throw std::runtime_error("zachfit needs --data DIR with dm.npy");
        for(double v : col("dm")) { Z->dm.setValue(v); d->addEvent(); }
        top = Z->total;
        movers = Z->free;
        report = movers;
        probs = [Z] { return Z->total->getCompProbsAtDataPoints()[0]; };
    } else if(config == "dp4") {
        DP4Model *M = make_dp4(opt.data.empty() ? (unsigned)std::max<size_t>(N, 1000) : 1000u, opt.lsmove);
        d.reset(new UnbinnedDataSet({M->m12, M->m34, M->cos12, M->cos34, M->phi, M->eventNumber}));
        if(opt.data.empty())
            fill_dp4_phsp(*M, *d, N, 4242);
        else {
            auto a = col("m12"), b = col("m34"), c = col("cos12"), e = col("cos34"), f = col("phi");
            for(size_t i = 0; i < a.size(); i++) {
                M->m12.setValue(a[i]); M->m34.setValue(b[i]); M->cos12.setValue(c[i]); M->cos34.setValue(e[i]); M->phi.setValue(f[i]);
                M->eventNumber.setValue(i);
                d->addEvent();
            }
            // inject the product's MC normalisation sample into the reference's private containers (Amp4Body.cu:270-289)
            Amp4Body *dp = M->dp;
            size_t keepN = N;
            N = 0; // MC columns are read whole
            auto m0 = col("mc_m12"), m1 = col("mc_m34"), m2 = col("mc_cos12"), m3 = col("mc_cos34"), m4 = col("mc_phi");
            N = keepN;
            dp->norm_M12 = mcbooster::RealVector_d(m0.begin(), m0.end());
            dp->norm_M34 = mcbooster::RealVector_d(m1.begin(), m1.end());
            dp->norm_CosTheta12 = mcbooster::RealVector_d(m2.begin(), m2.end());
            dp->norm_CosTheta34 = mcbooster::RealVector_d(m3.begin(), m3.end());
            dp->norm_phi = mcbooster::RealVector_d(m4.begin(), m4.end());
            dp->MCevents = m0.size();
            dp->norm_SF = mcbooster::RealVector_d(m0.size() * dp->SpinFactors.size());
            dp->norm_LS = mcbooster::mc_device_vector<fpcomplex>(m0.size() * dp->LineShapes.size());
        }
        n_mc = M->dp->getMCevents();
        top = M->signal;
        after_setdata = [M, &d] { M->dp->setDataSize(d->getNumEvents(), 6); };
        for(int c = 0; c < 6; c++) movers.push_back(M->coefs[c]);
        report = movers;
        if(opt.lsmove) ls_var = &M->lsvars[2]; // K*(892) mass; a fit floats the six coefficients only (DP4.cu:140-177)
        probs = [M] { return M->signal->getCompProbsAtDataPoints()[0]; };
    } else if(config == "tddp4") {
        TDModel *M = make_tddp4(0, opt.data.empty() ? (unsigned)std::max<size_t>(N, 1000) : 1000u, opt.lsmove, true);
        d.reset(new UnbinnedDataSet(M->obs()));
        if(opt.data.empty())
            fill_td_phsp(*M, *d, N, 31337);
        else {
            auto a = col("m12"), b = col("m34"), c = col("cos12"), e = col("cos34"), f = col("phi"), t = col("dtime"), g = col("sigmat");
            for(size_t i = 0; i < a.size(); i++) {
                M->m12.setValue(a[i]); M->m34.setValue(b[i]); M->cos12.setValue(c[i]); M->cos34.setValue(e[i]); M->phi.setValue(f[i]);
                M->dtime.setValue(t[i]); M->sigmat.setValue(g[i]); M->eventNumber.setValue(i);
                d->addEvent();
            }
            auto *ne = dynamic_cast<NormEvents_4Body_DeviceCached *>(M->dp->_normEvents[0].get());
            size_t keepN = N;
            N = 0;
            auto m0 = col("mc_m12"), m1 = col("mc_m34"), m2 = col("mc_cos12"), m3 = col("mc_cos34"), m4 = col("mc_phi");
            N = keepN;
            ne->_norm_M12_d = mcbooster::RealVector_d(m0.begin(), m0.end());
            ne->_norm_M34_d = mcbooster::RealVector_d(m1.begin(), m1.end());
            ne->_norm_CosTheta12_d = mcbooster::RealVector_d(m2.begin(), m2.end());
            ne->_norm_CosTheta34_d = mcbooster::RealVector_d(m3.begin(), m3.end());
            ne->_norm_phi_d = mcbooster::RealVector_d(m4.begin(), m4.end());
            ne->_totNumAccNormEvents = m0.size();
            ne->_sumInitWeights = m0.size();
            ne->_norm_SF_d.resize(0);
            ne->_norm_LS_d.resize(0);
        }
        n_mc = M->dp->getNumAccNormEvents();
        top = M->signal;
        after_setdata = [M, &d] { M->dp->setDataSize(d->getNumEvents(), 8); };
        movers = M->mix;
        report = movers;
        if(opt.lsmove) // C5 as configured floats the rho / K* masses and widths too (TDDP4.cu:50-53): report them after a fit
            for(auto &v : M->lsvars) report.push_back(v);
        if(opt.lsmove) ls_var = &M->lsvars[2];
        probs = [M] { return M->signal->getCompProbsAtDataPoints()[0]; };
    } else {
        fprintf(stderr, "unknown bench config %s\n", config.c_str());
        return 2;
    }
    const size_t n_events = d->getNumEvents();
    auto t0 = now();
    top->setData(d.get());
    after_setdata();
    auto t1 = now();
    first_nll = top->calculateNLL();
    auto t2 = now();
    t_setup = secs(t0, t1); t_first = secs(t1, t2);
    std::string fit_json;
    {
        FitManager fm(top);
        FCN *fcn = fm.getFCN();
        std::vector<double> v0 = fm.getParams()->make_minuit_vector(), v = v0;
        std::vector<int> idx;
        for(auto &m : movers) idx.push_back(m.getFitterIndex());
        for(int k = 0; k < W + K; k++) {
            for(size_t c = 0; c < idx.size(); c++) v[idx[c]] = v0[idx[c]] * (1.0 + 1e-3 * ((k + (int)c) % 3 - 1));
            auto a = now();
            last = (*fcn)(v);
            auto b = now();
            if(k >= W) ms.push_back(1e3 * secs(a, b));
        }
        if(opt.lsmove && ls_var && !ls_var->IsFixed()) { // a lineshape mass moves on every call: caches + integral rebuilt
            int li = ls_var->getFitterIndex();
            std::vector<double> t;
            for(int k = 0; k < 4; k++) {
                v[li] = v0[li] * (1.0 + 1e-4 * (k + 1));
                auto a = now();
                last_lsmove = (*fcn)(v);
                auto b = now();
                if(k >= 1) t.push_back(1e3 * secs(a, b));
            }
            for(double x : t) ms_lsmove += x / t.size();
            v[li] = v0[li];
        }
        // back to the last steady-state point, so that last_nll is what the caches hold from here on
        last = (*fcn)(v);
        if(opt.neumaier) {
            neum = neumaier_nll_of_probs(probs());
            top->setFitControl(std::make_shared<UnbinnedNllFit>());
            top->setData(d.get());
            after_setdata();
        }
        if(opt.fit) (*fcn)(v0); // the schedule moved the Variables: a fit starts from the constructor values again
    }
    if(opt.fit) {
        FitManager fit(top);
        fit.setVerbosity(0);
        auto a = now();
        auto fmin = fit.fit();
        double wall = secs(a, now());
        std::ostringstream o;
        o << std::setprecision(17) << ", \"fit\": {\"wall_s\": " << wall << ", \"nfcn\": " << fmin.NFcn() << ", \"valid\": " << (bool(fit) ? "true" : "false")
          << ", \"fval\": " << fmin.Fval() << ", \"edm\": " << fmin.Edm() << ", \"names\": [";
        for(size_t i = 0; i < report.size(); i++) o << (i ? ", " : "") << "\"" << report[i].getName() << "\"";
        o << "], \"values\": [";
        for(size_t i = 0; i < report.size(); i++) o << (i ? ", " : "") << report[i].getValue();
        o << "], \"errors\": [";
        for(size_t i = 0; i < report.size(); i++) o << (i ? ", " : "") << report[i].getError();
        o << "]}";
        fit_json = o.str();
    }
    double tot = 0;
    for(double m : ms) tot += m;
    double mean = ms.empty() ? 0 : tot / ms.size();
    printf("{\"config\": \"%s\", \"n_events\": %zu, \"n_mc\": %zu, \"steps\": %d, \"warmup\": %d, \"threads\": %d, \"ms_per_step\": %.6f, "
           "\"events_x_evals_per_s\": %.6e, \"setdata_s\": %.4f, \"first_call_s\": %.4f, \"first_nll\": %.17g, \"last_nll\": %.17g",
           config.c_str(), n_events, n_mc, K, W, threads, mean, mean > 0 ? (double)n_events / (mean * 1e-3) : 0.0, t_setup, t_first, first_nll, last);
    if(opt.neumaier) printf(", \"neumaier_last_nll\": %.17g", neum);
    if(opt.lsmove && ls_var) printf(", \"lsmove_ms_per_step\": %.6f, \"lsmove_last_nll\": %.17g", ms_lsmove, last_lsmove);
    printf("%s}\n", fit_json.c_str());
    return 0;
}

int main(int argc, char **argv) {
    if(argc < 2) {
        fprintf(stderr, "usage: ref_driver golden <outdir> | bench <config> <N> <K> [W]\n");
        return 2;
    }
    std::string mode = argv[1];
    try {
        if(mode == "golden") {
            g_out = argc > 2 ? argv[2] : ".";
            mkdir(g_out.c_str(), 0755);
            std::string only = argc > 3 ? argv[3] : "all";
            if(only == "all" || only == "basic") case_basic();
            if(only == "all" || only == "binned") case_binned();
            if(only == "all" || only == "dalitz") case_dalitz();
            if(only == "all" || only == "zoo3") case_zoo3();
            if(only == "all" || only == "zoo4") case_zoo4();
            if(only == "all" || only == "kmatrix") case_kmatrix();
            if(only == "all" || only == "td3") case_td3();
            if(only == "all" || only == "combine2") case_combine2();
            if(only == "all" || only == "composite") case_composite();
            if(only == "all" || only == "basic2") case_basic2();
            if(only == "all" || only == "smoothhist") case_smoothhist();
            if(only == "all" || only == "voigt") case_voigt();
            if(only == "all" || only == "dalitzeff") case_dalitzeff();
            if(only == "all" || only == "dalitzfit") case_dalitz_fit();
            if(only == "all" || only == "dp4") case_dp4();
            if(only == "all" || only == "dp4fit") case_dp4_fit();
            if(only == "all" || only == "tddp4") case_tddp4();
            if(only == "all" || only == "tddp4w") case_tddp4w();
            if(only == "all" || only == "tddp4fit") case_tddp4_fit();
            std::ofstream f(g_out + "/manifest" + (only == "all" ? "" : "_" + only) + ".json");
            f << "{\n" << g_manifest.str() << "\n}\n";
            return 0;
        }
        if(mode == "bench") {
            if(argc < 5) return 2;
            ScaleOpts opt;
            int W = 1;
            for(int i = 5; i < argc; i++) {
                std::string a = argv[i];
                if(a == "--data" && i + 1 < argc) opt.data = argv[++i];
                else if(a == "--neumaier") opt.neumaier = true;
                else if(a == "--fit") opt.fit = true;
                else if(a == "--lsmove") opt.lsmove = true;
                else W = std::stoi(a);
            }
            return bench(argv[2], std::stoul(argv[3]), std::stoi(argv[4]), W, opt);
        }
    } catch(const std::exception &e) {
        fprintf(stderr, "ref_driver: %s\n", e.what());
        return 1;
    }
    return 2;
}
