// Test infrastructure: lets the reference's OWN Catch2 test sources (tests/convert, tests/simple, tests/PDFs of the reference checkout)
// compile unmodified and run against this repo. The reference carries Catch2 as an empty submodule (extern/catch2) and no copy of it
// exists in this image, so this header provides the small part of its interface those tests use - TEST_CASE, SECTION, CHECK, REQUIRE,
// CHECK_THROWS_AS, CHECK_THAT with ContainsSubstring, INFO, CAPTURE, Approx with margin / epsilon, the _a literal and
// Catch::Session::run - written from that interface's documentation. A failed CHECK is reported with file, line and the expression
// text and counted; REQUIRE also leaves the test case. A test case with SECTIONs is run once per top-level section.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <exception>
#include <functional>
#include <iostream>
#include <limits>
#include <sstream>
#include <string>
#include <vector>

namespace Catch {

class Approx {
  public:
    explicit Approx(double v)
        : value_(v) {}
    Approx &margin(double m) {
        margin_ = m;
        return *this;
    }
    Approx &epsilon(double e) {
        epsilon_ = e;
        return *this;
    }
    Approx &scale(double s) {
        scale_ = s;
        return *this;
    }
    Approx operator-() const {
        Approx r(*this);
        r.value_ = -value_;
        return r;
    }
    bool matches(double other) const {
        // Catch2's rule: within the absolute margin, or within epsilon relative to scale + |value|
        if(std::fabs(other - value_) <= margin_) return true;
        return std::fabs(other - value_) <= epsilon_ * (scale_ + std::fabs(std::isinf(value_) ? 0.0 : value_));
    }
    double value() const { return value_; }

  private:
    double value_, margin_ = 0.0, epsilon_ = std::numeric_limits<float>::epsilon() * 100, scale_ = 0.0;
};
template <class T> bool operator==(const T &l, const Approx &r) { return r.matches(static_cast<double>(l)); }
template <class T> bool operator==(const Approx &l, const T &r) { return l.matches(static_cast<double>(r)); }
template <class T> bool operator!=(const T &l, const Approx &r) { return !r.matches(static_cast<double>(l)); }
template <class T> bool operator!=(const Approx &l, const T &r) { return !l.matches(static_cast<double>(r)); }
template <class T> bool operator<=(const T &l, const Approx &r) { return static_cast<double>(l) < r.value() || r.matches(static_cast<double>(l)); }
template <class T> bool operator>=(const T &l, const Approx &r) { return static_cast<double>(l) > r.value() || r.matches(static_cast<double>(l)); }
inline std::ostream &operator<<(std::ostream &o, const Approx &a) { return o << "Approx(" << a.value() << ")"; }

namespace literals {
inline Approx operator"" _a(long double v) { return Approx(static_cast<double>(v)); }
inline Approx operator"" _a(unsigned long long v) { return Approx(static_cast<double>(v)); }
} // namespace literals

namespace Matchers {
struct StringContains {
    std::string needle;
    bool match(const std::string &s) const { return s.find(needle) != std::string::npos; }
    std::string describe() const { return "contains: \"" + needle + "\""; }
};
inline StringContains ContainsSubstring(const std::string &s) { return StringContains{s}; }
} // namespace Matchers

namespace shim {
struct TestCase {
    const char *name, *tags;
    void (*fn)();
};
struct LeaveTestCase {};
struct State {
    std::vector<TestCase> cases;
    int checks = 0, failed = 0;
    int section_target = 0, sections_seen = 0, section_depth = 0;
    std::vector<std::string> info;
};
inline State &state() {
    static State s;
    return s;
}
struct Registrar {
    Registrar(const char *name, const char *tags, void (*fn)()) { state().cases.push_back({name, tags, fn}); }
};
inline void report(bool ok, const char *macro, const char *expr, const char *file, int line, const std::string &extra = "") {
    State &s = state();
    s.checks++;
    if(ok) return;
    s.failed++;
    std::cerr << file << ":" << line << ": FAILED: " << macro << "( " << expr << " )" << (extra.empty() ? "" : "  " + extra) << "\n";
    for(const std::string &i : s.info) std::cerr << "    with: " << i << "\n";
}
// one top-level SECTION per run of the test case; nested sections run with their parent
struct SectionGuard {
    bool run;
    explicit SectionGuard(const char *) {
        State &s = state();
        if(s.section_depth == 0) {
            run = s.sections_seen == s.section_target;
            s.sections_seen++;
        } else
            run = true;
        if(run) s.section_depth++;
    }
    ~SectionGuard() {
        if(run) state().section_depth--;
    }
    explicit operator bool() const { return run; }
};
struct InfoGuard {
    explicit InfoGuard(const std::string &m) { state().info.push_back(m); }
    ~InfoGuard() { state().info.pop_back(); }
};
} // namespace shim

class Session {
  public:
    // arguments: test-case names (or substrings) to run; none = all. Returns the number of failed assertions (capped at 255).
    int run(int argc, char **argv) {
        shim::State &s = shim::state();
        int ran = 0;
        for(const shim::TestCase &t : s.cases) {
            bool wanted = argc <= 1;
            for(int i = 1; i < argc; ++i)
                if(argv[i][0] != '-' && (std::strstr(t.name, argv[i]) || std::strstr(t.tags, argv[i]))) wanted = true;
            if(!wanted) continue;
            ran++;
            const int before = s.failed;
            s.section_target = 0;
            do {
                s.sections_seen = 0;
                s.section_depth = 0;
                try {
                    t.fn();
                } catch(const shim::LeaveTestCase &) {
                } catch(const std::exception &e) {
                    s.failed++;
                    std::cerr << t.name << ": unexpected exception: " << e.what() << "\n";
                } catch(...) {
                    s.failed++;
                    std::cerr << t.name << ": unexpected exception\n";
                }
                s.section_target++;
            } while(s.section_target < s.sections_seen);
            std::cerr << (s.failed == before ? "[ passed ] " : "[ FAILED ] ") << t.name << "\n";
        }
        std::cerr << ran << " test cases, " << s.checks << " assertions, " << s.failed << " failed\n";
        return s.failed > 255 ? 255 : s.failed;
    }
};
} // namespace Catch

#define CATCH_SHIM_CAT2(a, b) a##b
#define CATCH_SHIM_CAT(a, b) CATCH_SHIM_CAT2(a, b)
#define CATCH_SHIM_TEST(fn, ...)                                                                                                                \
    static void fn();                                                                                                                          \
    static ::Catch::shim::Registrar CATCH_SHIM_CAT(fn, _reg)(CATCH_SHIM_FIRST(__VA_ARGS__, ""), CATCH_SHIM_SECOND(__VA_ARGS__, "", ""), &fn);  \
    static void fn()
#define CATCH_SHIM_FIRST(a, ...) a
#define CATCH_SHIM_SECOND(a, b, ...) b
#define TEST_CASE(...) CATCH_SHIM_TEST(CATCH_SHIM_CAT(catch_shim_test_, __COUNTER__), __VA_ARGS__)
// TEST_CASE_METHOD(Fixture, name, tags): the body is a member function of a class derived from the fixture
#define CATCH_SHIM_METHOD(cls, fixture, ...)                                                                                                    \
    namespace {                                                                                                                                \
    struct cls : fixture {                                                                                                                     \
        void catch_shim_body();                                                                                                                \
        static void catch_shim_run() {                                                                                                         \
            cls obj;                                                                                                                           \
            obj.catch_shim_body();                                                                                                             \
        }                                                                                                                                      \
    };                                                                                                                                         \
    static ::Catch::shim::Registrar CATCH_SHIM_CAT(cls, _reg)(CATCH_SHIM_FIRST(__VA_ARGS__, ""), CATCH_SHIM_SECOND(__VA_ARGS__, "", ""),        \
                                                              &cls::catch_shim_run);                                                           \
    }                                                                                                                                          \
    void cls::catch_shim_body()
#define TEST_CASE_METHOD(fixture, ...) CATCH_SHIM_METHOD(CATCH_SHIM_CAT(catch_shim_fixture_, __COUNTER__), fixture, __VA_ARGS__)
#define SECTION(...) if(::Catch::shim::SectionGuard CATCH_SHIM_CAT(catch_shim_section_, __COUNTER__){CATCH_SHIM_FIRST(__VA_ARGS__, "")})
#define CHECK(...) ::Catch::shim::report(static_cast<bool>(__VA_ARGS__), "CHECK", #__VA_ARGS__, __FILE__, __LINE__)
#define CHECK_FALSE(...) ::Catch::shim::report(!static_cast<bool>(__VA_ARGS__), "CHECK_FALSE", #__VA_ARGS__, __FILE__, __LINE__)
#define REQUIRE(...)                                                                                                                            \
    do {                                                                                                                                       \
        const int catch_shim_before = ::Catch::shim::state().failed;                                                                            \
        ::Catch::shim::report(static_cast<bool>(__VA_ARGS__), "REQUIRE", #__VA_ARGS__, __FILE__, __LINE__);                                     \
        if(::Catch::shim::state().failed != catch_shim_before) throw ::Catch::shim::LeaveTestCase();                                            \
    } while(0)
#define CHECK_THROWS_AS(expr, type)                                                                                                             \
    do {                                                                                                                                       \
        bool catch_shim_ok = false;                                                                                                            \
        try {                                                                                                                                  \
            static_cast<void>(expr);                                                                                                           \
        } catch(const type &) {                                                                                                                \
            catch_shim_ok = true;                                                                                                              \
        } catch(...) {                                                                                                                         \
        }                                                                                                                                      \
        ::Catch::shim::report(catch_shim_ok, "CHECK_THROWS_AS", #expr ", " #type, __FILE__, __LINE__);                                          \
    } while(0)
#define CHECK_THAT(value, matcher)                                                                                                              \
    do {                                                                                                                                       \
        auto catch_shim_m = (matcher);                                                                                                         \
        ::Catch::shim::report(catch_shim_m.match(value), "CHECK_THAT", #value ", " #matcher, __FILE__, __LINE__, catch_shim_m.describe());      \
    } while(0)
#define INFO(msg)                                                                                                                               \
    std::ostringstream CATCH_SHIM_CAT(catch_shim_os_, __LINE__);                                                                                \
    CATCH_SHIM_CAT(catch_shim_os_, __LINE__) << msg;                                                                                            \
    ::Catch::shim::InfoGuard CATCH_SHIM_CAT(catch_shim_info_, __LINE__)(CATCH_SHIM_CAT(catch_shim_os_, __LINE__).str())
#define CAPTURE(...) INFO(#__VA_ARGS__ " := " << (__VA_ARGS__))
