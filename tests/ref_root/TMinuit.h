// tests/ref_root stand-in (see TH1.h): the part of TMinuit's interface goofit/fitting/FitManagerMinuit1.h drives, carried by
// Minuit2 (the C++ rewrite of the same algorithms, bundled with the reference) — there is no ROOT on either machine. TEST
// INFRASTRUCTURE: it lets the reference's ROOT-gated tests and examples that fit through FitManagerMinuit1 compile unmodified and
// run; a user with ROOT gets the real TMinuit. Parameters with lower == upper are unlimited, as in TMinuit; UP (the error
// definition) is TMinuit's default 1; crossed limits are swapped as TMinuit does; the status is 0 after a valid minimum and 4 otherwise, as TMinuit::Migrad returns it.
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include <Minuit2/FCNBase.h>
#include <Minuit2/FunctionMinimum.h>
#include <Minuit2/MnHesse.h>
#include <Minuit2/MnMigrad.h>
#include <Minuit2/MnPrint.h>
#include <Minuit2/MnUserParameterState.h>
#include <Minuit2/MnUserParameters.h>

#include "Rtypes.h"
#include "TString.h"

class TMinuit {
  public:
    explicit TMinuit(Int_t maxpar = 25)
        : maxpar_(maxpar) {}
    virtual ~TMinuit() = default;

    /// TMinuit's callback: value of the function at `par` (all parameters, by number). Overridden by a subclass, or forwarded to the
    /// function given to SetFCN
    virtual Int_t Eval(Int_t npar, Double_t *grad, Double_t &fval, Double_t *par, Int_t flag) {
        if(fcn_) fcn_(npar, grad, fval, par, flag);
        return 0;
    }
    void SetFCN(void (*fcn)(Int_t &, Double_t *, Double_t &, Double_t *, Int_t)) { fcn_ = fcn; }

    Int_t DefineParameter(Int_t parNo, const char *name, Double_t initVal, Double_t initErr, Double_t lowerLimit, Double_t upperLimit) {
        if(parNo != (Int_t)names_.size() || parNo >= maxpar_ || !(initErr >= 0)) return 4;
        if(lowerLimit > upperLimit) std::swap(lowerLimit, upperLimit); // TMinuit: "PARAMETER LIMITS HAVE BEEN REVERSED" (GooFit's constants come with crossed limits)
        std::string n = name ? name : "";
        if(lowerLimit == upperLimit)
            upar_.Add(n, initVal, initErr);
        else
            upar_.Add(n, initVal, initErr, lowerLimit, upperLimit);
        names_.push_back(n);
        return 0;
    }
    Int_t GetNumPars() const { return (Int_t)names_.size(); }
    Int_t FixParameter(Int_t parNo) {
        upar_.Fix(parNo);
        return 0;
    }
    void SetPrintLevel(Int_t level = 0) { printLevel_ = level; }
    void SetErrorDef(Double_t up) { up_ = up; }
    Int_t GetStatus() const { return status_; }

    Int_t Migrad() { return run(0); }
    /// "MIGRAD" (plist[0] = call limit), "HESSE"; "MINOS" and "IMPROVE" are accepted and skipped
    void mnexcm(const char *command, Double_t *plist, Int_t llist, Int_t &ierflg) {
        const std::string c = command ? command : "";
        ierflg              = 0;
        if(c == "MIGRAD" || c == "MIGrad")
            ierflg = run(llist > 0 && plist ? (unsigned int)plist[0] : 0);
        else if(c == "HESSE" || c == "HESse")
            hesse(llist > 0 && plist ? (unsigned int)plist[0] : 0);
    }
    Int_t GetParameter(Int_t parNo, Double_t &currentValue, Double_t &currentError) const {
        if(parNo < 0 || parNo >= GetNumPars()) return -1;
        const ROOT::Minuit2::MnUserParameters &p = state_ ? state_->Parameters() : upar_;
        currentValue = p.Value(parNo);
        currentError = p.Error(parNo);
        return 0;
    }
    void mnstat(Double_t &fmin, Double_t &fedm, Double_t &errdef, Int_t &npari, Int_t &nparx, Int_t &istat) const {
        fmin   = state_ ? state_->Fval() : 0.0;
        fedm   = state_ ? state_->Edm() : 0.0;
        errdef = up_;
        npari  = (Int_t)upar_.VariableParameters();
        nparx  = GetNumPars();
        istat  = state_ && state_->HasCovariance() ? (status_ == 0 ? 3 : 1) : 0;
    }

  private:
    struct Fcn : ROOT::Minuit2::FCNBase {
        TMinuit *self;
        explicit Fcn(TMinuit *s)
            : self(s) {}
        double operator()(const std::vector<double> &x) const override {
            std::vector<double> p(x);
            double f = 0;
            self->Eval((Int_t)p.size(), nullptr, f, p.data(), 4);
            return f;
        }
        double Up() const override { return self->up_; }
    };
    const ROOT::Minuit2::MnUserParameterState &start() {
        if(!state_) state_ = std::make_unique<ROOT::Minuit2::MnUserParameterState>(upar_);
        return *state_;
    }
    Int_t run(unsigned int maxcalls) {
        ROOT::Minuit2::MnPrint::SetGlobalLevel(printLevel_ < 0 ? 0 : printLevel_);
        Fcn fcn(this);
        ROOT::Minuit2::MnMigrad migrad(fcn, start(), ROOT::Minuit2::MnStrategy(1));
        ROOT::Minuit2::FunctionMinimum min = migrad(maxcalls);
        state_  = std::make_unique<ROOT::Minuit2::MnUserParameterState>(min.UserState());
        status_ = min.IsValid() ? 0 : 4;
        return status_;
    }
    void hesse(unsigned int maxcalls) {
        Fcn fcn(this);
        ROOT::Minuit2::MnHesse h(ROOT::Minuit2::MnStrategy(1));
        state_ = std::make_unique<ROOT::Minuit2::MnUserParameterState>(h(fcn, start(), maxcalls));
    }

    void (*fcn_)(Int_t &, Double_t *, Double_t &, Double_t *, Int_t){nullptr};
    Int_t maxpar_;
    Int_t printLevel_{0};
    Int_t status_{0};
    Double_t up_{1.0};
    std::vector<std::string> names_;
    ROOT::Minuit2::MnUserParameters upar_;
    std::unique_ptr<ROOT::Minuit2::MnUserParameterState> state_;
};
