// tests/ref_root stand-in (see TH1.h): a tree that counts what it is asked to store. Branch() remembers the address and type letter of
// each leaf ("name/D", "name/I", "name/F"); Fill() reads the current values into columns in memory, so a test can look at them.
#pragma once
#include <map>
#include <string>
#include <vector>
#include "TH1.h"
class TTree : public TObject {
  public:
    TTree(const char *name = "", const char *title = "")
        : name_(name ? name : "")
        , title_(title ? title : "") {}
    template <class T>
    int Branch(const char *name, T *address, const char *leaflist = "") {
        std::string l = leaflist ? leaflist : "";
        char type = l.size() >= 2 && l[l.size() - 2] == '/' ? l.back() : 'D';
        leaves_.push_back({name ? name : "", (const void *)address, type});
        columns_.emplace_back();
        return (int)leaves_.size();
    }
    int Fill() {
        for(size_t i = 0; i < leaves_.size(); i++) {
            const Leaf &L = leaves_[i];
            double v = L.type == 'I' ? (double)*(const int *)L.address : (L.type == 'F' ? (double)*(const float *)L.address : *(const double *)L.address);
            columns_[i].push_back(v);
        }
        return (int)++entries_;
    }
    long long GetEntries() const { return entries_; }
    const std::vector<double> &column(const std::string &name) const {
        static const std::vector<double> none;
        for(size_t i = 0; i < leaves_.size(); i++)
            if(leaves_[i].name == name) return columns_[i];
        return none;
    }
    const char *GetName() const { return name_.c_str(); }

  private:
    struct Leaf {
        std::string name;
        const void *address;
        char type;
    };
    std::string name_, title_;
    std::vector<Leaf> leaves_;
    std::vector<std::vector<double>> columns_;
    long long entries_{0};
};
