// gpu-feature-discovery: B200-native drop-in for the reference daemon's entry point.
//
// Replaces cmd/gpu-feature-discovery/main.go:22-210 (main, start, run) and watchers.go:26-31:
// same flags / environment / config file, same output file life-cycle, same exit codes, same log lines.
#include <cerrno>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <system_error>
#include <thread>
#include <exception>

#include <unistd.h>

#include "config.hpp"
#include "go_compat.hpp"
#include "labeler.hpp"

using gocompat::logf;

namespace {

double now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6;
}

const char *signal_name(int s) {  // syscall.Signal.String()
    switch (s) {
    case SIGHUP: return "hangup";
    case SIGINT: return "interrupt";
    case SIGTERM: return "terminated";
    case SIGQUIT: return "quit";
    default: return "signal";
    }
}

// run (main.go:148-210): returns restart; throws gfdlm::PassError on error
bool run(gfdlm::Manager &mgr, const gfdcfg::Config &config, const sigset_t &sigs) {
    struct Cleanup {  // the deferred removeOutputFile (main.go:149-156)
        const gfdcfg::Config &c;
        ~Cleanup() {
            if (!*c.oneshot && c.passes <= 0 && !c.output_file->empty()) {
                try {
                    gfdlm::remove_output_file(*c.output_file);
                } catch (const gfdlm::PassError &e) {
                    logf("Warning: Error removing output file: %s", e.what());
                }
            }
        }
    } cleanup{config};

    // lm.NewTimestampLabeler: evaluated ONCE per run(), so every rerun carries the same timestamp (main_test.go:266-267)
    gfdlm::Labels timestamp;
    if (!*config.no_timestamp) timestamp["nvidia.com/gfd.timestamp"] = std::to_string((long long)time(nullptr));
    // NEW opt-in probes: also once per run(); they allocate GPU memory and take tens of ms per GPU
    gfdlm::Labels probes = gfdlm::probe_labels(config);

    int passes_done = 0;
    while (true) {
        double t0 = now_ms();
        gfdlm::PhaseTimes ph;
        // The vGPU scan (sysfs reads, no NVML) has no input from the NVML bracket, so it runs beside it on a thread of its own;
        // its result -- or its error -- is looked at only where the reference would get to it (the lazy part of
        // labelers.Labels(), after the NVML labeler has been built: main.go:160-172), so an NVML failure still wins.
        gfdlm::Labels vgpu;
        std::exception_ptr vgpu_error;
        auto scan = [&] {
            try {
                vgpu = gfdlm::vgpu_labels(config);
            } catch (...) {
                vgpu_error = std::current_exception();
            }
        };
        std::thread vgpu_thread;
        try {
            vgpu_thread = std::thread(scan);
        } catch (const std::system_error &) {
            // no thread to be had (a pids limit on the pod): the scan runs after the NVML bracket, where the reference has it
        }
        const bool scan_in_thread = vgpu_thread.joinable();
        struct Join {
            std::thread &t;
            ~Join() { if (t.joinable()) t.join(); }
        } join{vgpu_thread};
        gfdlm::Labels nvml = gfdlm::nvml_labels(mgr, config, &ph);  // lm.NewLabelers: the eager NVML part
        double t1 = now_ms();
        gfdlm::Labels labels = timestamp;                      // lm.Merge(timestampLabeler, loopLabelers): later wins
        for (auto &kv : nvml) labels[kv.first] = kv.second;
        if (scan_in_thread) vgpu_thread.join();
        else scan();
        try {
            if (vgpu_error) std::rethrow_exception(vgpu_error);
        } catch (const gfdlm::PassError &e) {
            throw gfdlm::PassError(std::string("error generating labels: error generating labels: ") + e.what());
        }
        for (auto &kv : vgpu) labels[kv.first] = kv.second;
        for (auto &kv : probes) labels[kv.first] = kv.second;
        double t2 = now_ms();

        if (labels.size() <= 1) logf("Warning: no labels generated from any source");
        logf("Writing labels to output file");
        try {
            gfdlm::write_labels(labels, *config.output_file);
        } catch (const gfdlm::PassError &e) {
            throw gfdlm::PassError("error writing file '" + *config.output_file + "': " + e.what());
        }
        double t3 = now_ms();
        if (config.timing)
            logf("timing: pass %.3f ms (nvml %.3f, vgpu %.3f, write %.3f) [init %.3f, attach %.3f, labels %.3f, shutdown %.3f]", t3 - t0, t1 - t0,
                 t2 - t1, t3 - t2, ph.init_ms, ph.attach_ms, ph.labels_ms, ph.shutdown_ms);

        if (*config.oneshot) return false;
        if (config.passes > 0 && passes_done + 1 >= config.passes) return false;
        ++passes_done;

        logf("Sleeping for %s", gocompat::duration_string(*config.sleep_interval_ns).c_str());
        const int64_t ns = *config.sleep_interval_ns < 0 ? 0 : *config.sleep_interval_ns;
        timespec start;
        clock_gettime(CLOCK_MONOTONIC, &start);
        int64_t left = ns;
        while (true) {
            timespec to;
            to.tv_sec = (time_t)(left / 1000000000ll);
            to.tv_nsec = (long)(left % 1000000000ll);
            int s = sigtimedwait(&sigs, nullptr, &to);
            if (s == SIGHUP) { logf("Received SIGHUP, restarting."); return true; }
            if (s > 0) { logf("Received signal \"%s\", shutting down.", signal_name(s)); return false; }
            if (errno != EINTR) break;  // EAGAIN: rerunTimeout fired
            timespec now;
            clock_gettime(CLOCK_MONOTONIC, &now);
            left = ns - ((now.tv_sec - start.tv_sec) * 1000000000ll + (now.tv_nsec - start.tv_nsec));
            if (left <= 0) break;
        }
    }
}

// start (main.go:109-146)
int start(const gfdcfg::ParsedFlags &flags) {
    struct Bye { ~Bye() { logf("Exiting"); } } bye;
    logf("Starting OS watcher.");
    sigset_t sigs;  // newOSWatcher(SIGHUP, SIGINT, SIGTERM, SIGQUIT): queue them instead of dying
    sigemptyset(&sigs);
    sigaddset(&sigs, SIGHUP);
    sigaddset(&sigs, SIGINT);
    sigaddset(&sigs, SIGTERM);
    sigaddset(&sigs, SIGQUIT);
    sigprocmask(SIG_BLOCK, &sigs, nullptr);

    while (true) {
        logf("Loading configuration.");
        gfdcfg::Config config;
        try {
            config = gfdcfg::load_config(flags);
        } catch (const std::runtime_error &e) {
            logf("Error: %s", e.what());
            return 1;
        }
        logf("\nRunning with config:\n%s", config.to_json().c_str());
        bool restart;
        try {
            std::unique_ptr<gfdlm::Manager> mgr = gfdlm::new_manager(config);
            logf("Start running");
            restart = run(*mgr, config, sigs);
        } catch (const gfdlm::PassError &e) {
            logf("Error: %s", e.what());
            return 1;
        }
        if (!restart) return 0;
    }
}

}  // namespace

int main(int argc, char **argv) {
    gfdcfg::CliResult res;
    auto flags = gfdcfg::parse_cli(argc, argv, &res);
    switch (res.action) {
    case gfdcfg::CliAction::Help:
        fputs(gfdcfg::help_text().c_str(), stdout);
        return 0;
    case gfdcfg::CliAction::Version:
        fputs(gfdcfg::version_text().c_str(), stdout);
        return 0;
    case gfdcfg::CliAction::Error:
        if (!res.usage_error.empty()) {
            fprintf(stdout, "%s\n\n", res.usage_error.c_str());
            fputs(gfdcfg::help_text().c_str(), stdout);
        }
        logf("Error: %s", res.error.c_str());
        return 1;
    case gfdcfg::CliAction::Run:
        break;
    }
    return start(*flags);
}
