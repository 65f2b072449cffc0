#include <goofit/MathUtils.h>
int main() {
    std::vector<GooFit::fptype> v{1.0, 1e100, 1.0, -1e100};
    return GooFit::MathUtils::doNeumaierSummation(v) == 2.0 && GooFit::MathUtils::doNeumaierSummation({0.25}) == 0.25 ? 0 : 1;
}
