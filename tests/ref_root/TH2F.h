// tests/ref_root stand-in (see TH1.h): a two-dimensional histogram, global bin = bx + (nx + 2) * by as in ROOT
#pragma once
#include "TH1.h"
class TH2F : public TObject, public TAttStandIn {
  public:
    TH2F(const char *name, const char *title, int nx, double xlo, double xhi, int ny, double ylo, double yhi)
        : name_(name ? name : "")
        , title_(title ? title : "")
        , nx_(nx)
        , ny_(ny)
        , xlo_(xlo)
        , xhi_(xhi)
        , ylo_(ylo)
        , yhi_(yhi)
        , c_((size_t)(nx + 2) * (ny + 2), 0.0)
        , ax_(nx, xlo, xhi)
        , ay_(ny, ylo, yhi) {}
    int FindBinX(double x) const { return x < xlo_ ? 0 : (x >= xhi_ ? nx_ + 1 : 1 + (int)std::floor((x - xlo_) / (xhi_ - xlo_) * nx_)); }
    int FindBinY(double y) const { return y < ylo_ ? 0 : (y >= yhi_ ? ny_ + 1 : 1 + (int)std::floor((y - ylo_) / (yhi_ - ylo_) * ny_)); }
    int Fill(double x, double y, double w = 1.0) {
        int b = FindBinX(x) + (nx_ + 2) * FindBinY(y);
        c_[b] += w;
        entries_ += 1;
        return b;
    }
    void SetBinContent(int bx, int by, double v) {
        if(bx >= 0 && bx <= nx_ + 1 && by >= 0 && by <= ny_ + 1) c_[bx + (size_t)(nx_ + 2) * by] = v;
    }
    double GetBinContent(int bx, int by) const { return bx >= 0 && bx <= nx_ + 1 && by >= 0 && by <= ny_ + 1 ? c_[bx + (size_t)(nx_ + 2) * by] : 0.0; }
    int GetNbinsX() const { return nx_; }
    int GetNbinsY() const { return ny_; }
    double GetEntries() const { return entries_; }
    double Integral() const {
        double s = 0;
        for(int by = 1; by <= ny_; by++)
            for(int bx = 1; bx <= nx_; bx++) s += c_[bx + (size_t)(nx_ + 2) * by];
        return s;
    }
    void SetStats(bool) {}
    void SetTitle(const char *t) { title_ = t ? t : ""; }
    void SetMinimum(double) {}
    void SetMaximum(double) {}
    const char *GetName() const { return name_.c_str(); }
    TAxis *GetXaxis() { return &ax_; }
    TAxis *GetYaxis() { return &ay_; }
    TAxis *GetZaxis() { return &az_; }

  private:
    std::string name_, title_;
    int nx_, ny_;
    double xlo_, xhi_, ylo_, yhi_;
    std::vector<double> c_;
    double entries_{0};
    TAxis ax_, ay_, az_;
};
