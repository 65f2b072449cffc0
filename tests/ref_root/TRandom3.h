// tests/ref_root stand-in (see TRandom.h)
#pragma once
#include "TRandom.h"
class TRandom3 : public TRandom {
  public:
    explicit TRandom3(unsigned long seed = 4357)
        : TRandom(seed) {}
};
