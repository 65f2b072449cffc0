#include <goofit/FunctorWriter.h>
#include <goofit/PDFs/basic/GaussianPdf.h>
#include <goofit/PDFs/basic/ExpPdf.h>
#include <goofit/PDFs/combine/AddPdf.h>
#include <cstdio>
using namespace GooFit;
int main() {
    Observable x{"x", -5, 5};
    Variable m{"m", 0.123456789012345, 0.1, -5, 5}, s{"s", 1.5, 0.1, 0.1, 5}, a{"a", -0.3, 0.1, -2, 2}, f{"f", 0.4, 0.1, 0, 1};
    GaussianPdf g{"g", x, m, s};
    ExpPdf e{"e", x, a};
    AddPdf sum{"sum", f, &g, &e};
    writeToFile(&sum, "fw_params.txt");
    m.setValue(9); s.setValue(9); a.setValue(9); f.setValue(9); m.setError(7);
    readFromFile(&sum, "fw_params.txt");
    if(m.getValue() != 0.123456789012345 || s.getValue() != 1.5 || a.getValue() != -0.3 || f.getValue() != 0.4 || m.getError() != 0.1) return 1;
    std::vector<fptype> v{1.5, -2.25, 1e-300}, w;
    writeListOfNumbers(v, "fw_list.txt");
    readListOfNumbers(w, "fw_list.txt");
    return v == w ? 0 : 2;
}
