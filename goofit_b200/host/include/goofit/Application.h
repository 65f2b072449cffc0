// Application.h — minimal stand-in for the reference's CLI11-based GooFit::Application (src/goofit/Application.cpp):
// parses --gpu-dev N, -q/--quiet and --nosplash, selects the CUDA device before the first PDF is evaluated and, when
// launched with RANK/WORLD_SIZE/LOCAL_RANK in the environment, takes its event shard from them AND attaches the NCCL
// communicator the partial sums are reduced over (Engine::attachCommFromEnv; the reference's MPI rank -> GPU mapping
// plus its MPI_Init, Application.cpp:163-207): sharding without the reduction never happens.
// Where CLI11's headers are on the include path (the reference bundles them under extern/CLI11/include) Application IS a CLI::App,
// as upstream's (include/goofit/Application.h:48-93): add_option / add_flag / add_subcommand / footer / --config and the GooFit
// option group (--gpu-dev, --info-only, -q, --nosplash) behave as there, so examples/dalitz/dalitz.cpp, zachFit.cpp, simpleFit.cpp
// compile as they are. Without those headers (GOOFIT_B200_NO_CLI11, or no <CLI/CLI.hpp>) the minimal stand-in below is used.
#pragma once
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>

#if defined(__has_include) && !defined(GOOFIT_B200_NO_CLI11)
#if __has_include(<CLI/CLI.hpp>)
#define GOOFIT_B200_HAVE_CLI11 1
#include <CLI/CLI.hpp>
#endif
#endif

#include <goofit/Color.h>
#include <goofit/Version.h>
#include <goofit/detail/Core.h>

namespace GooFit {
/// Application.h:37-46 of the reference: what the start-up banner and `--info` print
inline std::string goofit_info_version() { return std::string("GooFit (B200-native likelihood engine) ") + GOOFIT_VERSION; }
inline std::string goofit_info_device(int gpuDev_) {
    int sm = 0;
    size_t free_b = 0, total_b = 0;
    detail::Engine &E = detail::Engine::get();
    if(gpuDev_ >= 0) E.setDevice(gpuDev_);
    if(gf_device_info(E.ctx(), &sm, &free_b, &total_b) != 0) return "no CUDA device";
    return "CUDA device " + std::to_string(gpuDev_ < 0 ? 0 : gpuDev_) + ": " + std::to_string(sm) + " SMs, " + std::to_string(total_b >> 30) + " GiB ("
           + std::to_string(free_b >> 30) + " GiB free), sm_100a kernels";
}
inline void print_splash() { std::cout << goofit_info_version() << std::endl; }
inline void print_goofit_info(int gpuDev_ = 0) { std::cout << goofit_info_version() << "\n" << goofit_info_device(gpuDev_) << std::endl; }
namespace detail {
/// RANK / WORLD_SIZE / LOCAL_RANK of a torchrun- or SLURM-style launch: the device of this process and, for more than one rank,
/// the event shard together with the communicator its partial sums are reduced over
inline void applyLaunchEnvironment(int &gpuDev) {
    const char *lr = std::getenv("LOCAL_RANK"), *r = std::getenv("RANK"), *w = std::getenv("WORLD_SIZE");
    if(gpuDev < 0 && lr) gpuDev = std::atoi(lr);
    if(gpuDev >= 0) Engine::get().setDevice(gpuDev);
    if(r && w && std::atoi(w) > 1) Engine::get().attachCommFromEnv(std::atoi(r), std::atoi(w));
}
/// Application.cpp:295-333: the working directory, the directory of the executable, then `base` under GOOFIT_SOURCE_DIR (an
/// environment variable or compile-time macro here)
inline std::string findInputFile(int argc, char **argv, const std::string &input_str, const std::string &base) {
    auto exists = [](const std::string &p) {
        std::ifstream f(p);
        return f.good();
    };
    if(exists(input_str)) return input_str;
    std::string prog = argc > 0 && argv && argv[0] ? argv[0] : "";
    size_t loc = prog.rfind('/');
    if(loc != std::string::npos && exists(prog.substr(0, loc) + "/" + input_str)) return prog.substr(0, loc) + "/" + input_str;
    const char *env = std::getenv("GOOFIT_SOURCE_DIR");
#ifdef GOOFIT_SOURCE_DIR
    std::string src = env ? env : GOOFIT_SOURCE_DIR;
#else
    std::string src = env ? env : ".";
#endif
    if(!base.empty() && exists(src + "/" + base + "/" + input_str)) return src + "/" + base + "/" + input_str;
    if(exists(src + "/../" + input_str)) return src + "/../" + input_str;
    throw GeneralError("Could not find file: {}", input_str);
}
} // namespace detail

#ifdef GOOFIT_B200_HAVE_CLI11
// the names upstream imports from CLI11 (Application.h:24-33)
using CLI::App;
using CLI::ExistingDirectory;
using CLI::ExistingFile;
using CLI::ExitCodes;
using CLI::FileError;
using CLI::NonexistentPath;
using CLI::Option;
using CLI::ParseError;
using CLI::Success;

class Application : public CLI::App {
  protected:
    int gpuDev_{-1}; // -1: not given; LOCAL_RANK of the launcher if there is one, else device 0
    bool quiet_{false};
    bool splash_{false}; // upstream's sense (Application.cpp:226,246): set by --nosplash, the splash is printed while it is false
    int argc_;
    char **argv_;

  public:
    Application(std::string description, int argc, char **argv)
        : App(std::move(description))
        , argc_(argc)
        , argv_(argv) {
        fallthrough(); // Application.cpp:207: options of the application are visible to its subcommands
        add_option("--gpu-dev", gpuDev_, "GPU device to use")->group("GooFit");
        add_flag_function(
            "--info-only",
            [](std::int64_t) {
                print_splash();
                print_goofit_info(-1);
                throw CLI::Success();
            },
            "Show the available GPU devices and exit")
            ->group("GooFit");
        auto *quiet = add_flag("-q,--quiet", quiet_, "Reduce the verbosity of the Application")->group("GooFit");
        add_flag("--nosplash", splash_, "Do not print a splash")->group("GooFit")->excludes(quiet);
        set_config("--config", "config.ini", "An ini file with command line options in it")->group("GooFit");
        std::atexit([]() { std::cout << GooFit::reset; }); // Application.cpp:231: colour off at exit
    }
    Application(int argc, char **argv)
        : Application("", argc, argv) {}

    int get_device() const { return gpuDev_ < 0 ? 0 : gpuDev_; }
    /// argc and argv are stored: parse them (throws CLI::ParseError, which GOOFIT_PARSE turns into the exit code)
    void run() { parse(argc_, argv_); }
    /// called by CLI11 once the options are read, before any callback of a subcommand (Application.cpp:241-250)
    void pre_callback() override {
        detail::applyLaunchEnvironment(gpuDev_);
        if(!quiet_) {
            if(!splash_) print_splash();
            print_goofit_info(get_device());
        }
    }
    int exit(const CLI::Error &e) {
        std::cout << (e.get_exit_code() == 0 ? blue : red);
        int rval = CLI::App::exit(e);
        std::cout << GooFit::reset;
        return rval;
    }
    int exit(const GeneralError &e) {
        std::cerr << e.what() << std::endl;
        return e.get_exit_code();
    }
    bool quiet() const { return quiet_; }
    /// make the selected GPU current again (Application.cpp: cudaSetDevice; here the engine owns the context of its device)
    void set_device() const {
        if(gpuDev_ >= 0) detail::Engine::get().setDevice(gpuDev_);
    }
    /// upstream enables SIGFPE traps on the host; kept as a no-op: the device code reports through its status, never through signals
    static void set_floating_exceptions() {}
    std::string get_filename(const std::string &input_str, std::string base = "") const { return detail::findInputFile(argc_, argv_, input_str, base); }
};
} // namespace GooFit
#define GOOFIT_PARSE(app, ...)                                                                                         \
    try {                                                                                                              \
        app.run();                                                                                                     \
    } catch(const GooFit::ParseError &e) {                                                                             \
        return app.exit(e);                                                                                            \
    } catch(const GooFit::GeneralError &e) {                                                                           \
        return app.exit(e);                                                                                            \
    }
#else
class Application {
  public:
    Application(std::string description, int argc, char **argv)
        : description_(std::move(description))
        , argc_(argc)
        , argv_(argv) {}
    Application(int argc, char **argv)
        : Application("", argc, argv) {}
    /// parse and apply; throws GeneralError on a malformed option
    void run() {
        for(int i = 1; i < argc_; i++) {
            std::string a = argv_[i];
            if(a == "--gpu-dev" && i + 1 < argc_)
                gpuDev_ = std::atoi(argv_[++i]);
            else if(a == "-q" || a == "--quiet")
                quiet_ = true;
            else if(a == "--nosplash")
                splash_ = false;
        }
        detail::applyLaunchEnvironment(gpuDev_);
        if(splash_ && !quiet_) std::cout << "GooFit (B200-native likelihood engine) " << description_ << std::endl;
    }
    int exit(const GeneralError &e) {
        std::cerr << e.what() << std::endl;
        return e.get_exit_code();
    }
    bool quiet() const { return quiet_; }
    int get_device() const { return gpuDev_ < 0 ? 0 : gpuDev_; }
    /// make the selected GPU current again (Application.cpp: cudaSetDevice; here the engine owns the context of its device)
    void set_device() const {
        if(gpuDev_ >= 0) detail::Engine::get().setDevice(gpuDev_);
    }
    /// upstream enables SIGFPE traps on the host; kept as a no-op: the device code reports through its status, never through signals
    static void set_floating_exceptions() {}
    /// where an input file of an example lives (Application.cpp:295-333)
    std::string get_filename(const std::string &input_str, std::string base = "") const { return detail::findInputFile(argc_, argv_, input_str, base); }

  private:
    std::string description_;
    int argc_;
    char **argv_;
    int gpuDev_{-1};
    bool quiet_{false}, splash_{true};
};
} // namespace GooFit
#define GOOFIT_PARSE(app, ...)                                                                                             \
    try {                                                                                                              \
        app.run();                                                                                                     \
    } catch(const GooFit::GeneralError &e) {                                                                           \
        return app.exit(e);                                                                                            \
    }
#endif
