// tests/ref_root stand-in (see TH1.h): a file that stores nothing
#pragma once
#include "TH1.h"
class TFile {
  public:
    TFile(const char *, const char * = "") {}
    static TFile *Open(const char *n, const char *o = "") { return new TFile(n, o); }
    bool IsZombie() const { return false; }
    bool cd() { return true; }
    void Write() {}
    void Close() {}
    template <class T>
    T *Get(const char *) { return nullptr; }
};
