// Color.h - the reference's terminal colours (goofit/Color.h:7-21, there rang::fg / rang::style values) as plain ANSI sequences:
// `std::cout << GooFit::red << "text" << GooFit::reset`. Nothing is written when the stream is not a terminal.
#pragma once
#include <cstdio>
#include <iostream>
#include <unistd.h>

namespace GooFit {
struct TermCode {
    int code;
};
inline std::ostream &operator<<(std::ostream &os, TermCode c) {
    const bool tty = (&os == &std::cout && isatty(fileno(stdout))) || (&os == &std::cerr && isatty(fileno(stderr)));
    if(tty) os << "\033[" << c.code << "m";
    return os;
}
constexpr TermCode black{30}, red{31}, green{32}, yellow{33}, blue{34}, magenta{35}, cyan{36}, gray{37};
constexpr TermCode reset{0}, bold{1}, dim{2}, italic{3};
} // namespace GooFit
