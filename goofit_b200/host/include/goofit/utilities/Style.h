// utilities/Style.h — include/goofit/utilities/Style.h of the reference: the default ROOT style of GooFit's plots, for builds that
// have ROOT (GOOFIT_ROOT_FOUND=1). `inline` here: upstream's definition in a header breaks the one-definition rule when two
// translation units include it.
#pragma once

#include <goofit/Version.h>

#if GOOFIT_ROOT_FOUND
#include <RVersion.h>
#include <TCanvas.h>
#include <TStyle.h>
#endif

namespace GooFit {

#if GOOFIT_ROOT_FOUND
inline void setROOTStyle() {
    gStyle->SetNumberContours(512);
    gStyle->SetOptStat(0);
    gStyle->SetPadRightMargin(.13);
#if ROOT_VERSION_CODE < ROOT_VERSION(6, 6, 0)
    gStyle->SetPalette(55, 0); // kRainBow, which ROOT < 6.06 does not name
#else
    gStyle->SetPalette(kViridis, 0);
#endif
}
#endif

} // namespace GooFit
