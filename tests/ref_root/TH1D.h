// tests/ref_root stand-in (see TH1.h)
#pragma once
#include "TH1.h"
class TH1D : public TH1 {
  public:
    using TH1::TH1;
};
