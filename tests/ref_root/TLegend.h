// tests/ref_root stand-in (see TH1.h)
#pragma once
#include "TH1.h"
class TLegend : public TObject, public TAttStandIn {
  public:
    TLegend(double, double, double, double, const char * = "") {}
    void AddEntry(const TObject *, const char *, const char * = "") {}
    void SetBorderSize(int) {}
    void SetTextSize(double) {}
};
