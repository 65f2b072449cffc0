// Fit.h — Minuit2 glue, unchanged in behaviour from the reference (include/goofit/fitting/{Params,FCN,
// FitManagerMinuit2}.h). Minuit2 itself stays on the host; it is the reference's bundled third-party
// library (extern/Minuit2), compiled as-is into goofit_b200/lib/libMinuit2.so.
#pragma once
#include <Minuit2/FCNBase.h>
#include <Minuit2/FunctionMinimum.h>
#include <Minuit2/MnScan.h>
#include <Minuit2/MnUserParameterState.h>
#include <Minuit2/MnUserParameters.h>

#include <goofit/detail/Core.h>

namespace GooFit {
namespace Minuit2 = ROOT::Minuit2;
class FCN;

class Params : public Minuit2::MnUserParameters {
    friend FCN;

  protected:
    std::vector<Variable> vars_;
    PdfBase *pdf_;
    bool do_record_{false};
    std::vector<std::vector<double>> recorded_;
    // updateVariable() matches parameters by NAME over the whole tree (PdfBase.cu:121-129): resolved once per tree
    // structure instead of once per Minuit call (30 parameters x ~100 string compares is most of a 50 us call otherwise)
    std::vector<std::vector<Variable *>> targets_;
    uint64_t targets_epoch_{~uint64_t(0)};
    void resolveTargets();

  public:
    using MnUserParameters::MnUserParameters;
    Params(PdfBase &pdf);
    std::vector<Variable> GetGooFitParams() const { return vars_; }
    void SetGooFitParams(const Minuit2::MnUserParameterState &input);
    size_t size() const { return vars_.size(); }
    std::vector<double> make_minuit_vector() const;
    void from_minuit_vector(const std::vector<double> &values, bool force_changed = false);
    void set_record(bool do_record = true) { do_record_ = do_record; }
    std::vector<std::vector<double>> get_recorded() const { return recorded_; }
};

class FCN : public Minuit2::FCNBase {
    friend class FitManagerMinuit2;

  protected:
    Params *params_;

  public:
    FCN(Params &params);
    double operator()(const std::vector<double> &pars) const override;
    double operator()() const;
    double Up() const override { return 0.5; }
    Params *GetParams() { return params_; }
    /// number of likelihood evaluations since the last fit() started (the reference's host_callnumber)
    static int callNumber();
};

enum class FitErrors : int { Valid = 0, NotRun = 50, InValid };

class FitManagerMinuit2 {
  public:
    FitManagerMinuit2(PdfBase *dat);
    Minuit2::FunctionMinimum fit();
    void setMaxCalls(unsigned int max_calls = 0) { maxfcn_ = max_calls; }
    Params *getParams() { return &upar_; }
    FCN *getFCN() { return &fcn_; }
    Minuit2::MnScan getMnScan();
    operator bool() const { return retval_ == FitErrors::Valid; }
    operator int() const { return static_cast<int>(retval_); }
    void setVerbosity(int value) { verbosity = value; }
    int getVerbosity() const { return verbosity; }
    std::vector<std::pair<double, double>> getMinosErrors() const { return minos_errors; }
    void setMinos(bool minos_flag = true) { minos = minos_flag; }
    void setStrategy(int strat) { m_strategy = strat; }
    /// wall time of the last minimisation and the number of FCN calls ("The minimization took", FitManagerMinuit2.cpp:43-87)
    double lastFitSeconds() const { return seconds_; }
    unsigned lastFitCalls() const { return ncalls_; }

  private:
    Params upar_;
    FCN fcn_;
    unsigned int maxfcn_{0};
    FitErrors retval_{FitErrors::NotRun};
    int verbosity{3};
    std::vector<std::pair<double, double>> minos_errors;
    bool minos{false};
    int m_strategy = 1;
    double seconds_{0};
    unsigned ncalls_{0};
};
using FitManager = FitManagerMinuit2;

} // namespace GooFit
