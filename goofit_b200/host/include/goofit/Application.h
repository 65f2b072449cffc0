// Application.h — minimal stand-in for the reference's CLI11-based GooFit::Application (src/goofit/Application.cpp):
// parses --gpu-dev N, -q/--quiet and --nosplash, selects the CUDA device before the first PDF is evaluated and, when
// launched with RANK/WORLD_SIZE/LOCAL_RANK in the environment, takes its event shard from them AND attaches the NCCL
// communicator the partial sums are reduced over (Engine::attachCommFromEnv; the reference's MPI rank -> GPU mapping
// plus its MPI_Init, Application.cpp:163-207): sharding without the reduction never happens. The CLI11 option machinery itself is out of scope.
#pragma once
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>

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
        const char *lr = std::getenv("LOCAL_RANK"), *r = std::getenv("RANK"), *w = std::getenv("WORLD_SIZE");
        if(gpuDev_ < 0 && lr) gpuDev_ = std::atoi(lr);
        if(gpuDev_ >= 0) detail::Engine::get().setDevice(gpuDev_);
        if(r && w && std::atoi(w) > 1) detail::Engine::get().attachCommFromEnv(std::atoi(r), std::atoi(w));
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
    /// where an input file of an example lives (Application.cpp:295-333): the working directory, the directory of the
    /// executable, then `base` under GOOFIT_SOURCE_DIR (an environment variable or compile-time macro here)
    std::string get_filename(const std::string &input_str, std::string base = "") const {
        auto exists = [](const std::string &p) {
            std::ifstream f(p);
            return f.good();
        };
        if(exists(input_str)) return input_str;
        std::string prog = argc_ > 0 && argv_ && argv_[0] ? argv_[0] : "";
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
