// tests/ref_root — TEST INFRASTRUCTURE, not part of the product. A headless stand-in for the handful of ROOT classes the reference's
// examples touch (examples/dalitz/dalitz.cpp, addition.cpp, simpleFit.cpp, convolution.cpp, 2d_plot.cpp, zachFit.cpp), so that those
// sources compile UNMODIFIED against this repo's <goofit/...> headers with GOOFIT_ROOT_FOUND=1 on a machine without ROOT (there is no
// ROOT and no graphics stack on the build or GPU box). Histograms hold their contents in memory with ROOT's bin numbering (0 =
// underflow, 1..n, n+1 = overflow); canvases, styles, legends and files accept every call and draw or write nothing. A user with a
// real ROOT puts its include path first and gets the real classes.
#pragma once
#include <cmath>
#include <string>
#include <vector>
#include "TString.h"

enum EColorStandIn { kWhite = 0, kBlack = 1, kRed = 632, kGreen = 416, kBlue = 600, kYellow = 400, kMagenta = 616, kCyan = 432, kViridis = 112, kRainBow = 55 };

enum ELineStyleStandIn { kSolid = 1, kDashed = 2, kDotted = 3, kDashDotted = 4 };

class TAxis {
  public:
    TAxis() = default;
    TAxis(int n, double lo, double hi)
        : n_(n)
        , lo_(lo)
        , hi_(hi) {}
    int GetNbins() const { return n_; }
    double GetXmin() const { return lo_; }
    double GetXmax() const { return hi_; }
    double GetBinWidth(int) const { return n_ > 0 ? (hi_ - lo_) / n_ : 0.0; }
    double GetBinCenter(int b) const { return lo_ + (b - 0.5) * GetBinWidth(b); }
    void SetTitle(const char *) {}
    void SetRangeUser(double, double) {}
    void SetLabelSize(double) {}
    void SetTitleSize(double) {}
    void SetTitleOffset(double) {}

  private:
    int n_{0};
    double lo_{0}, hi_{0};
};

class TObject {
  public:
    virtual ~TObject() = default;
    virtual void Draw(const char * = "") {}
    virtual void Write(const char * = nullptr) {}
};

class TAttStandIn { // line, marker and fill attributes: accepted and forgotten
  public:
    void SetLineColor(int) {}
    void SetLineWidth(int) {}
    void SetLineStyle(int) {}
    void SetMarkerStyle(int) {}
    void SetMarkerSize(double) {}
    void SetMarkerColor(int) {}
    void SetFillColor(int) {}
    void SetFillStyle(int) {}
};

class TH1 : public TObject, public TAttStandIn {
  public:
    TH1(const char *name, const char *title, int nx, double xlo, double xhi)
        : name_(name ? name : "")
        , title_(title ? title : "")
        , nx_(nx)
        , xlo_(xlo)
        , xhi_(xhi)
        , c_(nx + 2, 0.0)
        , e_(nx + 2, -1.0)
        , ax_(nx, xlo, xhi) {}
    int FindBin(double x) const { return x < xlo_ ? 0 : (x >= xhi_ ? nx_ + 1 : 1 + (int)std::floor((x - xlo_) / (xhi_ - xlo_) * nx_)); }
    int Fill(double x, double w = 1.0) {
        int b = FindBin(x);
        c_[b] += w;
        entries_ += 1;
        return b;
    }
    void SetBinContent(int b, double v) {
        if(b >= 0 && b < (int)c_.size()) c_[b] = v;
    }
    double GetBinContent(int b) const { return b >= 0 && b < (int)c_.size() ? c_[b] : 0.0; }
    void SetBinError(int b, double v) {
        if(b >= 0 && b < (int)e_.size()) e_[b] = v;
    }
    double GetBinError(int b) const { return b >= 0 && b < (int)e_.size() ? (e_[b] >= 0 ? e_[b] : std::sqrt(std::fabs(c_[b]))) : 0.0; }
    double GetBinCenter(int b) const { return xlo_ + (b - 0.5) * (xhi_ - xlo_) / nx_; }
    double GetBinWidth(int) const { return (xhi_ - xlo_) / nx_; }
    int GetNbinsX() const { return nx_; }
    double GetEntries() const { return entries_; }
    double Integral() const {
        double s = 0;
        for(int b = 1; b <= nx_; b++) s += c_[b];
        return s;
    }
    void Scale(double f) {
        for(auto &v : c_) v *= f;
    }
    void SetStats(bool) {}
    void SetTitle(const char *t) { title_ = t ? t : ""; }
    void SetMinimum(double) {}
    void SetMaximum(double) {}
    const char *GetName() const { return name_.c_str(); }
    TAxis *GetXaxis() { return &ax_; }
    TAxis *GetYaxis() { return &ay_; }
    TAxis *GetZaxis() { return &az_; }

  protected:
    std::string name_, title_;
    int nx_;
    double xlo_, xhi_;
    std::vector<double> c_, e_;
    double entries_{0};
    TAxis ax_, ay_, az_;
};
