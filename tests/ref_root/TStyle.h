// tests/ref_root stand-in (see TH1.h): style settings are accepted and forgotten
#pragma once
#include "TH1.h"
class TStyle {
  public:
    void SetCanvasBorderMode(int) {}
    void SetCanvasColor(int) {}
    void SetFrameBorderMode(int) {}
    void SetFrameFillColor(int) {}
    void SetPadBorderMode(int) {}
    void SetPadColor(int) {}
    void SetStatColor(int) {}
    void SetTitleColor(int) {}
    void SetTitleFillColor(int) {}
    void SetFillColor(int) {}
    void SetFuncWidth(int) {}
    void SetLineWidth(int) {}
    void SetLineColor(int) {}
    void SetPalette(int, int * = nullptr) {}
    void SetNumberContours(int) {}
    void SetOptStat(int) {}
    void SetOptFit(int) {}
    void SetPadRightMargin(double) {}
    void SetPadLeftMargin(double) {}
};
inline TStyle *gStyleStandIn() {
    static TStyle s;
    return &s;
}
#define gStyle (gStyleStandIn())
