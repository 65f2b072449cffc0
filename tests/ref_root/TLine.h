// tests/ref_root stand-in (see TH1.h)
#pragma once
#include "TH1.h"
class TLine : public TObject, public TAttStandIn {
  public:
    TLine() = default;
    TLine(double, double, double, double) {}
    void DrawLine(double, double, double, double) {}
};
