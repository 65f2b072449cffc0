// tests/ref_root stand-in (see TH1.h)
#pragma once
#include "TH1.h"
class TLatex : public TObject {
  public:
    TLatex() = default;
    TLatex(double, double, const char *) {}
    void DrawLatex(double, double, const char *) {}
    void SetTextSize(double) {}
    void SetNDC(bool = true) {}
};
