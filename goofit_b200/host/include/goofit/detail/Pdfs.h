// Pdfs.h — the basic shapes and combiners named by the north star, same constructors as the
// reference headers (/root/reference/include/goofit/PDFs/basic/*.h, PDFs/combine/*.h). Each class
// only records an opcode + its parameters/constants; the per-event math lives in csrc/gf_ops.cuh.
#pragma once
#include <goofit/detail/Core.h>

namespace GooFit {

class ExpPdf : public GooPdf { // basic/ExpPdf.h
  public:
    ExpPdf(std::string n, Observable _x, Variable alpha);
    ExpPdf(std::string n, Observable _x, std::vector<Variable> &weights);
    ExpPdf(std::string n, Observable _x, Variable alpha, Variable offset);
    ExpPdf(std::string n, Observable _x, std::vector<Variable> &weights, Variable offset);
    fptype integrate(fptype lo, fptype hi) const override;
    bool hasAnalyticIntegral() const override { return opcode_ == GF_OP_EXP; }
};

class GaussianPdf : public GooPdf { // basic/GaussianPdf.h
  public:
    GaussianPdf(std::string n, Observable _x, Variable m, Variable s);
    fptype integrate(fptype lo, fptype hi) const override;
    bool hasAnalyticIntegral() const override { return true; }
};

class PolynomialPdf : public GooPdf { // basic/PolynomialPdf.h
  public:
    PolynomialPdf(std::string n, Observable _x, std::vector<Variable> weights, unsigned int lowestDegree = 0);
    PolynomialPdf(std::string n, Observable _x, std::vector<Variable> weights, Variable x0, unsigned int lowestDegree = 0);
    PolynomialPdf(std::string n, std::vector<Observable> obses, std::vector<Variable> coeffs, std::vector<Variable> offsets,
                  unsigned int maxDegree = 0);
    fptype integrate(fptype lo, fptype hi) const override;
    fptype getCoefficient(int coef);

  private:
    std::unique_ptr<Variable> center;
    bool multiDim = false;
};

class ArgusPdf : public GooPdf { // basic/ArgusPdf.h
  public:
    ArgusPdf(std::string n, Observable _x, Variable m, Variable s, bool upper);
    ArgusPdf(std::string n, Observable _x, Variable m, Variable s, bool upper, Variable power);
    bool hasAnalyticIntegral() const override { return false; }
    fptype integrate(fptype lo, fptype hi) const override;
};

class KinLimitBWPdf : public GooPdf { // basic/KinLimitBWPdf.h
  public:
    KinLimitBWPdf(std::string n, Observable _x, Variable m, Variable s);
    bool hasAnalyticIntegral() const override { return false; }
    void setMasses(fptype bigM, fptype smallM);
};

class ScaledGaussianPdf : public GooPdf { // basic/ScaledGaussianPdf.h
  public:
    ScaledGaussianPdf(std::string n, Observable _x, Variable m, Variable s, Variable d, Variable e);
    bool hasAnalyticIntegral() const override { return false; }
};

class JohnsonSUPdf : public GooPdf { // basic/JohnsonSUPdf.h
  public:
    JohnsonSUPdf(std::string n, Observable _x, Variable m, Variable s, Variable g, Variable d);
    fptype integrate(fptype lo, fptype hi) const override { return 1.0; }
    bool hasAnalyticIntegral() const override { return true; }
};

// ---- the remaining closed-form shapes of PDFs/basic ----
class BWPdf : public GooPdf { // basic/BWPdf.h: non-relativistic Breit-Wigner
  public:
    BWPdf(std::string n, Observable _x, Variable m, Variable s);
};
class BifurGaussPdf : public GooPdf { // basic/BifurGaussPdf.h: Gaussian with different widths left and right of the mean
  public:
    BifurGaussPdf(std::string n, Observable _x, Variable m, Variable sL, Variable sR);
    fptype integrate(fptype lo, fptype hi) const override; // upstream defines it and leaves hasAnalyticIntegral false
};
class CrystalBallPdf : public GooPdf { // basic/CrystalBallPdf.h: Gaussian core with a power-law tail
  public:
    CrystalBallPdf(std::string n, Observable _x, Variable m, Variable s, Variable a);
    CrystalBallPdf(std::string n, Observable _x, Variable m, Variable s, Variable a, Variable power);
    fptype integrate(fptype lo, fptype hi) const override;
};
class ExpGausPdf : public GooPdf { // basic/ExpGausPdf.h: exponential convolved with a Gaussian
  public:
    ExpGausPdf(std::string n, Observable _x, Variable m, Variable s, Variable t);
};
class NovosibirskPdf : public GooPdf { // basic/NovosibirskPdf.h
  public:
    NovosibirskPdf(std::string n, Observable _x, Variable m, Variable s, Variable t);
};
class StepPdf : public GooPdf { // basic/StepPdf.h: 0 / 1 around x0
  public:
    StepPdf(std::string n, Observable _x, Variable x0, int Up = 1);
    fptype integrate(fptype lo, fptype hi) const override;
    bool hasAnalyticIntegral() const override { return true; }
};
class TrigThresholdPdf : public GooPdf { // basic/TrigThresholdPdf.h: sine-shaped efficiency edge
  public:
    TrigThresholdPdf(std::string n, Observable _x, Variable thresh, Variable trigConst, Variable linConst, bool upper = true);
    TrigThresholdPdf(std::string n, Observable _x, Observable _y, Variable thresh, Variable trigConst, Variable linConst, Variable massConstant,
                     bool upper);
};
class BernsteinPdf : public GooPdf { // basic/BernsteinPdf.h: Bernstein polynomial on the observable's range
  public:
    BernsteinPdf(std::string n, Observable vars, std::vector<Variable> coeffs, unsigned int mindeg = 0);
    bool hasAnalyticIntegral() const override { return false; }
};
class CorrGaussianPdf : public GooPdf { // basic/CorrGaussianPdf.h: 2-D Gaussian whose y width grows with the x distance
  public:
    CorrGaussianPdf(std::string n, Observable _x, Observable _y, Variable mean1, Variable sigma1, Variable mean2, Variable sigma2,
                    Variable correlation);
};
class BinTransformPdf : public GooPdf { // basic/BinTransformPdf.h: N-D coordinates -> one global bin number
  public:
    BinTransformPdf(std::string n, std::vector<Observable> obses, std::vector<fptype> limits, std::vector<fptype> binSizes, std::vector<int> numBins);
};
/// basic/SmoothHistogramPdf.h: the content of the event's bin in a histogram, each bin averaged with its neighbours:
/// (own + s sum(neighbours)) / (1 + s n_neighbours) for the smoothing parameter s. The smoothed table lives on the device.
class SmoothHistogramPdf : public GooPdf {
  public:
    SmoothHistogramPdf(std::string n, BinnedDataSet *x, Variable smoothing);
    fptype normalize() override;
    bool hasAnalyticIntegral() const override { return false; }
    void extractHistogram(std::vector<fptype> &host_hist) const { host_hist = base; }
    void copyHistogramToDevice(std::vector<fptype> &host_histogram);
    int emit(detail::ProgramBuilder &B) override;
    void afterProgram() override;

  private:
    std::vector<fptype> smoothedTable(fptype smoothing) const;
    std::vector<fptype> base;
    std::vector<int> nbins;
    fptype totalEvents = 0, uploadedFor = 0;
    bool uploaded = false;
};
class VoigtianPdf : public GooPdf { // basic/VoigtianPdf.h: Breit-Wigner convolved with a Gaussian
  public:
    VoigtianPdf(std::string n, Observable _x, Variable m, Variable s, Variable w);
};
/// utilities/VariableBinTransform1DPdf.h: the number of the variable-width bin |x| falls in (a mapping function for MappedPdf)
class VariableBinTransform1DPdf : public GooPdf {
  public:
    VariableBinTransform1DPdf(std::string n, Observable _x, std::vector<fptype> binlimits);
};
class LandauPdf : public GooPdf { // basic/LandauPdf.h
  public:
    LandauPdf(std::string n, Observable _x, Variable mpv, Variable sigma);
};

class CombinePdf : public GooPdf { // PDFs/CombinePdf.h
  public:
    using GooPdf::GooPdf;
};

class AddPdf : public CombinePdf { // combine/AddPdf.h
  public:
    AddPdf(std::string n, std::vector<Variable> weights, std::vector<PdfBase *> comps);
    AddPdf(std::string n, Variable frac1, PdfBase *func1, PdfBase *func2);
    fptype normalize() override;
    bool hasAnalyticIntegral() const override { return false; }
    double calculateNLL() override;

  private:
    bool extended;
};

/// combine/EventWeightedAddPdf.h: AddPdf whose weights are per-event observables (the first observables of the record)
class EventWeightedAddPdf : public CombinePdf {
  public:
    EventWeightedAddPdf(std::string n, std::vector<Observable> weights, std::vector<PdfBase *> comps);
    fptype normalize() override;
    bool hasAnalyticIntegral() const override { return false; }

  private:
    bool extended;
};
/// combine/MappedPdf.h: the value of the mapping function m (rounded) selects which of the functions t is evaluated
class MappedPdf : public CombinePdf {
  public:
    MappedPdf(std::string n, GooPdf *m, std::vector<GooPdf *> &t);
    fptype normalize() override;
    bool hasAnalyticIntegral() const override { return false; }
};

/// combine/CompositePdf.h: f(x) = h(g(x)) for a core function g of any number of observables and a shell h of one. Neither is
/// normalised inside the composite; the composite itself is, by the usual grid integral (CompositePdf.cu:38-75).
class CompositePdf : public CombinePdf {
  public:
    CompositePdf(std::string n, PdfBase *core, PdfBase *shell);
    fptype normalize() override;
    bool hasAnalyticIntegral() const override { return false; }
    int emit(detail::ProgramBuilder &B) override;
};

class ProdPdf : public CombinePdf { // combine/ProdPdf.h
  public:
    ProdPdf(std::string n, std::vector<PdfBase *> comps);
    fptype normalize() override;
    bool hasAnalyticIntegral() const override { return false; }

  private:
    bool varOverlaps{false};
};

class ConvolutionPdf : public CombinePdf { // combine/ConvolutionPdf.h
  public:
    ConvolutionPdf(std::string n, Observable _x, GooPdf *model, GooPdf *resolution);
    /// ConvolutionPdf.cu:180-226 of the reference: the form for several convolutions that are evaluated side by side (the targets of a
    /// MappedPdf), which upstream lets load their model workspaces into one shared-memory cache. Here every convolution reads its
    /// own per-offset table (csrc/gf_conv.cu), so there is nothing to share: the values are those of the four-argument form. The
    /// bookkeeping is kept: the PDF counts as not properly initialised until registerOthers has seen a list of numOthers + 1
    /// convolutions that includes this one.
    ConvolutionPdf(std::string n, Observable _x, GooPdf *model, GooPdf *resolution, unsigned int numOthers);
    /// ConvolutionPdf.cu:284-311: same checks and messages as upstream; returns without effect when the list is wrong
    void registerOthers(std::vector<ConvolutionPdf *> others);
    bool isProperlyInitialised() const { return properlyInitialised; }
    fptype normalize() override;
    void setIntegrationConstants(fptype lo, fptype hi, fptype step);

    void emitAux(detail::ProgramBuilder &B) override;
    int emit(detail::ProgramBuilder &B) override;
    void afterProgram() override;

  protected:

  private:
    void setLowerBoundOffset();
    GooPdf *model;
    GooPdf *resolution;
    bool workspacesFilled{false};
    unsigned int numOthers{0};
};

} // namespace GooFit
