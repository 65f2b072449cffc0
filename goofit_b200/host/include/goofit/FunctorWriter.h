// FunctorWriter.h - parameter values of a PDF tree to and from a text file, and plain lists of numbers (reference
// goofit/FunctorWriter.h:9-14, src/goofit/FunctorWriter.cpp:10-85). One line per parameter: name value error lower upper.
// Host-only; the lists are std::vector here (thrust::host_vector upstream).
#pragma once
#include <fstream>
#include <iomanip>
#include <limits>
#include <map>
#include <string>
#include <vector>

#include <goofit/detail/Core.h>

namespace GooFit {
inline void writeToFile(PdfBase *pdf, const char *fname) {
    std::ofstream out(fname);
    out << std::setprecision(std::numeric_limits<fptype>::max_digits10); // upstream writes 6 digits; a round trip here is exact
    for(const Variable &p : pdf->getParameters())
        out << p.getName() << " " << p.getValue() << " " << p.getError() << " " << p.getLowerLimit() << " " << p.getUpperLimit() << std::endl;
}
/// lines whose first word is not a parameter of the tree are skipped; reading stops once every parameter was seen
inline void readFromFile(PdfBase *pdf, const char *fname) {
    std::vector<Variable> params = pdf->getParameters();
    std::map<std::string, Variable *> byName;
    for(Variable &p : params) byName[p.getName()] = &p;
    std::ifstream in(fname);
    std::string word, rest;
    size_t seen = 0;
    while(in >> word) {
        auto it = byName.find(word);
        if(it == byName.end()) {
            std::getline(in, rest);
            continue;
        }
        fptype value, error, lo, hi;
        if(!(in >> value >> error >> lo >> hi)) break;
        it->second->setValue(value);
        it->second->setError(error);
        it->second->setLowerLimit(lo);
        it->second->setUpperLimit(hi);
        if(++seen == byName.size()) break;
    }
}
inline void writeListOfNumbers(std::vector<fptype> &target, const char *fname) {
    std::ofstream out(fname);
    out << std::setprecision(std::numeric_limits<fptype>::max_digits10);
    for(fptype t : target) out << t << " ";
}
inline void readListOfNumbers(std::vector<fptype> &target, const char *fname) {
    std::ifstream in(fname);
    fptype v;
    while(in >> v) target.push_back(v);
}
} // namespace GooFit
