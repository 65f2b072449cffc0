// TEST INFRASTRUCTURE (oracle). Stand-in for GooFit/MCBooster@c8bcee8f (submodule absent from
// /root/reference/extern/MCBooster). Only the typedef surface the reference's call sites use
// (SURVEY.md §8c). Written from the call sites, not from MCBooster sources. Parity of the
// phase-space GENERATOR against real MCBooster: unpinned (irrelevant to NLL parity: oracle
// and product consume identical events).
#pragma once
#include <cstddef>
#include <cstdint>
namespace mcbooster {
typedef double GReal_t;
typedef long GLong_t;
typedef unsigned long GULong_t;
typedef int GInt_t;
typedef unsigned int GUInt_t;
typedef bool GBool_t;
} // namespace mcbooster
