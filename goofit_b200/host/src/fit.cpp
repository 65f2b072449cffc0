// fit.cpp — Params / FCN / FitManagerMinuit2 (reference: src/goofit/{Params,FCN,FitManagerMinuit2}.cpp).
// The fit-parity-relevant settings are the reference's: MnStrategy(1), Up()=0.5, Minuit's default call
// limit, fixed parameters added value-only, unlimited parameters when lower==upper, blind offsets
// subtracted before evaluation, validity = IsValid && HasCovariance && !IsAboveMaxEdm && !HasReachedCallLimit.
#include <goofit/detail/Fit.h>

#include <Minuit2/MnMigrad.h>
#include <Minuit2/MnMinos.h>
#include <Minuit2/MnPrint.h>
#include <Minuit2/MnStrategy.h>

#include <chrono>

namespace GooFit {

static int g_callnumber = 0;
int FCN::callNumber() { return g_callnumber; }

Params::Params(PdfBase &pdf)
    : pdf_(&pdf) {
    vars_ = pdf_->getParameters();
    for(Variable &var : vars_) {
        bool added;
        if(var.IsFixed())
            added = Add(var.getName(), var.getValue());
        else if(var.getLowerLimit() == var.getUpperLimit())
            added = Add(var.getName(), var.getValue(), var.getError());
        else
            added = Add(var.getName(), var.getValue(), var.getError(), var.getLowerLimit(), var.getUpperLimit());
        if(!added) throw std::runtime_error("The name " + var.getName() + " appears more than once!");
        var.setFitterIndex(Index(var.getName()));
    }
}
void Params::SetGooFitParams(const Minuit2::MnUserParameterState &input) {
    for(Variable &var : vars_) {
        int k = var.getFitterIndex();
        var.setValue(input.Value(k));
        var.setError(input.Error(k));
        SetValue(k, var.getValue());
        SetError(k, var.getError());
    }
}
std::vector<double> Params::make_minuit_vector() const {
    std::vector<double> v(max_fitter_index(vars_) + 1);
    for(const Variable &var : vars_) v.at(var.getFitterIndex()) = var.getValue();
    return v;
}
void Params::resolveTargets() {
    targets_.assign(vars_.size(), {});
    for(size_t i = 0; i < vars_.size(); i++) pdf_->collectParametersByName(vars_[i].getName(), targets_[i]);
    targets_epoch_ = PdfBase::structureEpoch();
}
void Params::from_minuit_vector(const std::vector<double> &values, bool force_changed) {
    if(targets_epoch_ != PdfBase::structureEpoch()) resolveTargets();
    for(size_t i = 0; i < vars_.size(); i++) {
        Variable &var = vars_[i];
        double nv = values.at(var.getFitterIndex());
        var.setChanged(force_changed ? true : var.getValue() != nv);
        const double v = nv - var.getBlind(Variable::Key());
        for(Variable *t : targets_[i]) t->setValue(v); // == pdf_->updateVariable(var, v)
    }
    if(do_record_) recorded_.push_back(values);
    pdf_->updateParameters();
}

FCN::FCN(Params &params)
    : params_(&params) {
    g_callnumber = 0;
    for(Variable &var : params_->vars_) var.setChanged(true);
}
double FCN::operator()(const std::vector<double> &pars) const {
    params_->from_minuit_vector(pars);
    double nll = params_->pdf_->calculateNLL();
    g_callnumber++;
    return nll;
}
double FCN::operator()() const {
    std::vector<double> pars = params_->make_minuit_vector();
    params_->from_minuit_vector(pars, true);
    double nll = params_->pdf_->calculateNLL();
    g_callnumber++;
    return nll;
}

FitManagerMinuit2::FitManagerMinuit2(PdfBase *dat)
    : upar_(*dat)
    , fcn_(upar_) {}

Minuit2::FunctionMinimum FitManagerMinuit2::fit() {
    auto old_level = Minuit2::MnPrint::GlobalLevel();
    Minuit2::MnPrint::SetGlobalLevel(verbosity);
    g_callnumber = 0;
    auto t0 = std::chrono::steady_clock::now();
    Minuit2::MnStrategy strat(m_strategy);
    Minuit2::MnMigrad migrad{fcn_, upar_, strat};
    Minuit2::FunctionMinimum min = migrad(maxfcn_);
    if(minos) {
        Minuit2::MnMinos mn{fcn_, min};
        minos_errors.clear();
        for(Variable &var : upar_.GetGooFitParams())
            if(!var.IsFixed()) minos_errors.push_back(mn(var.getFitterIndex()));
        int k = 0;
        std::cout << "1-sigma minos errors: " << std::endl;
        for(Variable &var : upar_.GetGooFitParams())
            if(!var.IsFixed()) {
                std::cout << var.getName() << ": " << minos_errors[k].first << " " << minos_errors[k].second << std::endl;
                k++;
            }
    }
    seconds_ = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    ncalls_  = min.NFcn();
    if(verbosity > 0) {
        std::cout << min;
        std::cout << "The minimization took: " << seconds_ << " s\n";
        std::cout << "Average time per call: " << (ncalls_ ? 1e3 * seconds_ / ncalls_ : 0.0) << " ms" << std::endl;
    }
    if(min.IsValid() && min.HasCovariance() && !min.IsAboveMaxEdm() && !min.HasReachedCallLimit())
        retval_ = FitErrors::Valid;
    else {
        if(verbosity > 0) {
            std::cout << "HesseFailed: " << min.HesseFailed() << "\nHasCovariance: " << min.HasCovariance()
                      << "\nHasValidCovariance: " << min.HasValidCovariance()
                      << "\nHasValidParameters: " << min.HasValidParameters() << "\nIsAboveMaxEdm: " << min.IsAboveMaxEdm()
                      << "\nHasReachedCallLimit: " << min.HasReachedCallLimit()
                      << "\nHasAccurateCovar: " << min.HasAccurateCovar() << "\nHasPosDefCovar : " << min.HasPosDefCovar()
                      << "\nHasMadePosDefCovar : " << min.HasMadePosDefCovar() << std::endl;
        }
        retval_ = FitErrors::InValid;
    }
    upar_.SetGooFitParams(min.UserState());
    Minuit2::MnPrint::SetGlobalLevel(old_level);
    return min;
}
Minuit2::MnScan FitManagerMinuit2::getMnScan() { return Minuit2::MnScan{fcn_, upar_}; }

// PdfBase::fitTo / fit (src/goofit/PdfBase.cpp:140-152)
ROOT::Minuit2::FunctionMinimum PdfBase::fitTo(DataSet *data, int verbosity) {
    auto *old = getData();
    setData(data);
    auto funmin = fit(verbosity);
    setData(old);
    return funmin;
}
ROOT::Minuit2::FunctionMinimum PdfBase::fit(int verbosity) {
    FitManager fitter{this};
    fitter.setVerbosity(verbosity);
    return fitter.fit();
}

} // namespace GooFit
