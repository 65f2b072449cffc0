// The reference's examples/exponential/exponential.cpp flow, compiled unchanged in spirit against the B200 host API:
// 10^6 generated events, ExpPdf unbinned fit with Minuit2, alpha must come out in (-1.01, -0.99).
//   g++ -std=c++17 -Iinclude -Igoofit_b200/host/include -I$REF/extern/Minuit2/inc examples/exponential.cpp \
//       -Lgoofit_b200/lib -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,goofit_b200/lib -o build/exponential
#include <goofit/Application.h>
#include <goofit/FitManager.h>
#include <goofit/PDFs/basic/ExpPdf.h>
#include <goofit/UnbinnedDataSet.h>
#include <goofit/Variable.h>

#include <random>

using namespace GooFit;

int main(int argc, char **argv) {
    GooFit::Application app("Exponential example", argc, argv);
    GOOFIT_PARSE(app);

    Observable xvar{"xvar", 0, std::log(1. + RAND_MAX / 2.)};
    UnbinnedDataSet data{xvar};
    std::mt19937 gen(137);
    std::exponential_distribution<double> d(1.0);
    while(data.getNumEvents() < 1000000) {
        double v = d(gen);
        if(v < xvar.getUpperLimit()) {
            xvar.setValue(v);
            data.addEvent();
        }
    }
    Variable alpha{"alpha", -2, 0.1, -10, 10};
    ExpPdf exppdf{"exppdf", xvar, alpha};
    exppdf.setData(&data);
    FitManager fitter{&exppdf};
    fitter.fit();
    if(alpha.getValue() > -0.99 || alpha.getValue() < -1.01) return 1;
    return fitter;
}
