// Version.h - what the reference generates from include/goofit/Version.h.in at configure time: the version it was studied at, and
// the feature switches of this build (no ROOT, kMatrix present, the CUDA backend).
#pragma once

#define GOOFIT_VERSION_MAJOR 2
#define GOOFIT_VERSION_MINOR 3
#define GOOFIT_VERSION_PATCH 0
#define GOOFIT_VERSION_TWEAK 0
#define GOOFIT_VERSION "2.3.0-b200"

#ifndef GOOFIT_SOURCE_DIR
#define GOOFIT_SOURCE_DIR "."
#endif

#define CMAKE_CXX_FLAGS ""
#define CMAKE_BUILD_TYPE "Release"
#define GOOFIT_DEVICE "CUDA"
#define GOOFIT_HOST "CPP"
#define GOOFIT_ARCH_FLAGS "sm_100a"

// ROOT is not part of this build; a user who has it compiles with -DGOOFIT_ROOT_FOUND=1 (upstream's configure-time switch, Version.h.in:18;
// its CMake also defines ROOT_FOUND) and gets the header-only ROOT conveniences: GooPdf::plotToROOT, DalitzPlotter::make2D, setROOTStyle
#ifndef GOOFIT_ROOT_FOUND
#ifdef ROOT_FOUND
#define GOOFIT_ROOT_FOUND 1
#else
#define GOOFIT_ROOT_FOUND 0
#endif
#endif
#if GOOFIT_ROOT_FOUND && !defined(ROOT_FOUND)
#define ROOT_FOUND 1
#endif
#define GOOFIT_SPLASH 1
#define GOOFIT_KMATRIX 1
