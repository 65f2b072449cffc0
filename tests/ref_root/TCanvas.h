// tests/ref_root stand-in (see TH1.h): a canvas that draws nothing; SaveAs only records the name it was asked for
#pragma once
#include <string>
#include <vector>
#include "TH1.h"
class TVirtualPad {
  public:
    void SetLogy(bool = true) {}
    void SetLogx(bool = true) {}
    void SetLogz(bool = true) {}
    TVirtualPad *cd(int = 0) { return this; }
    void Divide(int, int) {}
    void Clear() {}
    void Update() {}
    void SaveAs(const char *name) { saved().emplace_back(name ? name : ""); }
    void Print(const char *name) { SaveAs(name); }
    /// every file name a canvas of this process was asked to write (what a test can look at)
    static std::vector<std::string> &saved() {
        static std::vector<std::string> s;
        return s;
    }
};
class TCanvas : public TVirtualPad {
  public:
    TCanvas() = default;
    TCanvas(const char *, const char * = "", int = 0, int = 0, int = 0, int = 0) {}
};
