// fitting/FitManagerMinuit1.h — the reference's TMinuit-based fit manager (include/goofit/fitting/FitManagerMinuit1.h,
// src/goofit/FitManagerMinuit1.cpp), for builds that have ROOT (GOOFIT_ROOT_FOUND=1, as upstream: without ROOT its header does not
// compile either). Header-only, because the libraries of this repository never link ROOT: TMinuit drives, Minuit1::Eval copies its
// parameter vector into the GooFit variables and returns the likelihood of the CUDA path, exactly as FCN::operator() does for Minuit2.
#pragma once

#include <goofit/Version.h>
#if !GOOFIT_ROOT_FOUND
#error "goofit/fitting/FitManagerMinuit1.h needs ROOT's TMinuit: compile with -DGOOFIT_ROOT_FOUND=1 and ROOT's include path, or use FitManagerMinuit2"
#endif

#include <TMinuit.h>

#include <cmath>
#include <iostream>
#include <vector>

#include <goofit/Color.h>
#include <goofit/Log.h>
#include <goofit/detail/Core.h>

namespace GooFit {

class Minuit1 : public TMinuit {
    PdfBase *pdfPointer;
    std::vector<Variable> vars;

  public:
    /// one TMinuit parameter per GooFit parameter, in the PDF's own order; fixed ones are fixed in TMinuit too
    explicit Minuit1(PdfBase *pdf)
        : TMinuit(static_cast<Int_t>(pdf->getParameters().size()) + 1)
        , pdfPointer(pdf)
        , vars(pdf->getParameters()) {
        int index = 0;
        for(Variable &var : vars) {
            var.setFitterIndex(index);
            const Int_t err = DefineParameter(index, var.getName().c_str(), var.getValue(), var.getError(), var.getLowerLimit(), var.getUpperLimit());
            if(GetNumPars() != index + 1)
                throw GeneralError("Error when implementing param {} (possibly invalid error/lowerlimit/upperlimit values)!", var.getName());
            if(err != 0) throw GeneralError("Was not able to implement param {} (error {})", var.getName(), err);
            if(var.IsFixed()) FixParameter(index);
            ++index;
        }
        pdfPointer->copyParams();
    }

    /// TMinuit's callback: `par` holds every parameter (fixed ones included) by fitter index; no gradient is provided
    Int_t Eval(Int_t, Double_t *, Double_t &fval, Double_t *par, Int_t) override {
        for(Variable &var : vars) {
            const double v = par[var.getFitterIndex()];
            if(std::isnan(v)) GOOFIT_WARN("Variable {} at {} is NaN", var.getName(), var.getIndex());
            var.setChanged(var.getValue() != v);
            var.setValue(v);
            pdfPointer->updateVariable(var, v - var.getBlind(Variable::Key()));
        }
        pdfPointer->updateParameters();
        fval = pdfPointer->calculateNLL();
        return 0;
    }

    /// (spelling as upstream)
    std::vector<Variable> getVaraibles() const { return vars; }
};

class FitManagerMinuit1 {
  public:
    explicit FitManagerMinuit1(PdfBase *dat)
        : minuit_(dat) {}

    void setMaxCalls(double mxc) { overrideCallLimit = mxc; }
    void useHesseBefore(bool use = true) { _useHesseBefore = use; }
    void useHesse(bool use = true) { _useHesse = use; }
    void useMinos(bool use = true) { _useMinos = use; }
    void useImprove(bool use = true) { _useImprove = use; }
    void setVerbosity(int v) { minuit_.SetPrintLevel(v - 1); }

    operator bool() const { return minuit_.GetStatus() == 0; }
    operator int() const { return minuit_.GetStatus(); }

    /// plain MIGRAD, or with a call limit the sequence HESSE / MIGRAD / HESSE / MINOS / IMPROVE that the switches select; the fitted
    /// values and errors end up in the Variables
    void fit() {
        for(Variable &var : minuit_.getVaraibles()) var.setChanged(true);
        std::cout << GooFit::gray << GooFit::bold;
        if(overrideCallLimit > 0) {
            std::cout << "Calling MIGRAD with call limit " << overrideCallLimit << std::endl;
            double plist[1] = {overrideCallLimit};
            int err         = 0;
            if(_useHesseBefore) minuit_.mnexcm("HESSE", plist, 1, err);
            minuit_.mnexcm("MIGRAD", plist, 1, err);
            if(_useHesse) minuit_.mnexcm("HESSE", plist, 1, err);
            if(_useMinos) minuit_.mnexcm("MINOS", plist, 1, err);
            if(_useImprove) minuit_.mnexcm("IMPROVE", plist, 1, err);
        } else {
            minuit_.Migrad();
        }
        std::cout << GooFit::reset;
        for(Variable &var : minuit_.getVaraibles()) {
            double value = 0, error = 0;
            minuit_.GetParameter(var.getFitterIndex(), value, error);
            var.setValue(value);
            var.setError(error);
        }
    }

    Minuit1 *getMinuitObject() { return &minuit_; }

    void getMinuitStatus(double &fmin, double &fedm, double &errdef, int &npari, int &nparx, int &istat) {
        minuit_.mnstat(fmin, fedm, errdef, npari, nparx, istat);
        std::cout << "mnstat(fmin = " << fmin << ", fedm = " << fedm << ", errdef = " << errdef << ", npari = " << npari << ", nparx = " << nparx
                  << ", istat = " << istat << ")" << std::endl;
    }

  private:
    double overrideCallLimit{-1};
    bool _useHesseBefore{true};
    bool _useHesse{true};
    bool _useMinos{false};
    bool _useImprove{false};
    Minuit1 minuit_;
};

} // namespace GooFit
