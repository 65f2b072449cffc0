// tests/ref_root stand-in (see TH1.h): ROOT's string as a std::string with Data() and the implicit conversion to const char *
#pragma once
#include <string>
class TString : public std::string {
  public:
    TString() = default;
    TString(const char *s)
        : std::string(s ? s : "") {}
    TString(const std::string &s)
        : std::string(s) {}
    const char *Data() const { return c_str(); }
    int Length() const { return (int)size(); }
    operator const char *() const { return c_str(); }
};
