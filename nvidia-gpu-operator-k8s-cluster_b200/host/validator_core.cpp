// validator_core.cpp — see validator_core.h.  Plain C++17 + POSIX; no CUDA headers here: the GPU is
// reached only through the Api table (the C ABI of include/b200validate.h).
#include "validator_core.h"

#include <algorithm>
#include <arpa/inet.h>
#include <cerrno>
#include <cinttypes>
#include <cmath>
#include <csignal>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <dirent.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

namespace b200v_host {

namespace {

const char *kOpNames[B200V_OP_COUNT] = {"copy", "scale", "add", "triad", "vectoradd", "triad_unfused"};

std::string env_or(const char *name, const std::string &dflt) {
    const char *v = std::getenv(name);
    return (v && *v) ? std::string(v) : dflt;
}

bool parse_size(const std::string &s, size_t &out) {
    if (s.empty()) return false;
    char *end = nullptr;
    errno = 0;
    unsigned long long v = std::strtoull(s.c_str(), &end, 10);
    if (errno || !end || *end) return false;
    out = (size_t)v;
    return true;
}

bool parse_ll(const std::string &s, long long &out) {
    if (s.empty()) return false;
    char *end = nullptr;
    errno = 0;
    long long v = std::strtoll(s.c_str(), &end, 10);
    if (errno || !end || *end) return false;
    out = v;
    return true;
}

int env_int(const char *name, int dflt) {
    long long v;
    return parse_ll(env_or(name, ""), v) ? (int)v : dflt;
}

bool ends_with(const std::string &s, const std::string &suffix) {
    return s.size() >= suffix.size() && s.compare(s.size() - suffix.size(), suffix.size(), suffix) == 0;
}

}  // namespace

bool parse_bool(const std::string &s, bool dflt) {
    if (s.empty()) return dflt;
    if (s == "1" || s == "true" || s == "TRUE" || s == "True" || s == "t" || s == "T") return true;
    if (s == "0" || s == "false" || s == "FALSE" || s == "False" || s == "f" || s == "F") return false;
    return dflt;
}

bool parse_config(int argc, char **argv, Config &cfg) {
    cfg.component = env_or("COMPONENT", "");
    cfg.with_wait = parse_bool(env_or("WITH_WAIT", ""), false);
    cfg.with_workload = parse_bool(env_or("WITH_WORKLOAD", ""), true);
    cfg.output_dir = env_or("OUTPUT_DIR", "/run/nvidia/validations");
    cfg.node_name = env_or("NODE_NAME", "");
    cfg.operator_namespace = env_or("OPERATOR_NAMESPACE", "");
    cfg.visible_devices = env_or("NVIDIA_VISIBLE_DEVICES", "");
    cfg.mode = env_or("B200V_MODE", "validate");
    size_t elems;
    if (parse_size(env_or("B200V_ELEMS", ""), elems)) cfg.elems = elems;
    cfg.gpus = env_int("B200V_GPUS", 0);
    cfg.sleep_seconds = env_int("B200V_SLEEP_SECONDS", 5);
    cfg.max_retries = env_int("B200V_MAX_RETRIES", 0);
    long long inj;
    if (parse_ll(env_or("B200V_INJECT_FAULT", ""), inj)) cfg.inject_fault = inj;
    cfg.child_device = env_int("B200V_CHILD_DEVICE", -1);
    cfg.allocatable_gpus = env_or("B200V_ALLOCATABLE_GPUS", "");
    cfg.allocatable_file = env_or("B200V_ALLOCATABLE_FILE", "");
    cfg.plugin_retries = env_int("B200V_PLUGIN_RETRIES", 30);
    cfg.metrics_port = env_int("METRICS_PORT", 8000);
    cfg.write_prom = parse_bool(env_or("B200V_WRITE_PROM", ""), false);
    cfg.nvidia_smi = env_or("B200V_NVIDIA_SMI", "nvidia-smi");
    cfg.driver_root = env_or("DRIVER_ROOT", "/run/nvidia/driver");
    char self[4096];
    ssize_t len = readlink("/proc/self/exe", self, sizeof(self) - 1);
    cfg.self_exe = len > 0 ? std::string(self, (size_t)len) : (argc > 0 ? argv[0] : "");

    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i], val;
        bool has_val = false;
        size_t eq = a.find('=');
        if (a.rfind("--", 0) == 0 && eq != std::string::npos) {
            val = a.substr(eq + 1);
            a = a.substr(0, eq);
            has_val = true;
        }
        auto need = [&](std::string &dst) -> bool {
            if (has_val) { dst = val; return true; }
            if (i + 1 >= argc) { cfg.error = "flag needs an argument: " + a; return false; }
            dst = argv[++i];
            return true;
        };
        auto flag_bool = [&](bool &dst) {
            // urfave/cli style: a bare bool flag means true; --flag=false is accepted
            dst = has_val ? parse_bool(val, true) : true;
        };
        std::string tmp;
        if (a == "--component" || a == "-c") { if (!need(cfg.component)) return false; }
        else if (a == "--with-wait" || a == "-w") flag_bool(cfg.with_wait);
        else if (a == "--with-workload" || a == "-l") flag_bool(cfg.with_workload);
        else if (a == "--output-dir" || a == "-o") { if (!need(cfg.output_dir)) return false; }
        else if (a == "--node-name") { if (!need(cfg.node_name)) return false; }
        else if (a == "--namespace" || a == "-n") { if (!need(cfg.operator_namespace)) return false; }
        else if (a == "--mode") { if (!need(cfg.mode)) return false; }
        else if (a == "--metrics-port" || a == "-p") {
            if (!need(tmp)) return false;
            long long port;
            if (!parse_ll(tmp, port) || port < 0 || port > 65535) { cfg.error = "bad --metrics-port: " + tmp; return false; }
            cfg.metrics_port = (int)port;
        }
        else if (a == "--elems") {
            if (!need(tmp)) return false;
            if (!parse_size(tmp, cfg.elems)) { cfg.error = "bad --elems: " + tmp; return false; }
        } else if (a == "--gpus") {
            if (!need(tmp)) return false;
            long long g;
            if (!parse_ll(tmp, g) || g < 0) { cfg.error = "bad --gpus: " + tmp; return false; }
            cfg.gpus = (int)g;
        } else {
            cfg.error = "flag provided but not defined: " + a;
            return false;
        }
    }
    if (cfg.mode != "validate" && cfg.mode != "stream" && cfg.mode != "nccl" && cfg.mode != "nvls") {
        cfg.error = "B200V_MODE must be validate, stream, nccl or nvls, not " + cfg.mode;
        return false;
    }
    return true;
}

// ---- status files ---------------------------------------------------------------------------------

std::string status_path(const Config &cfg, const std::string &name) {
    return cfg.output_dir + (ends_with(cfg.output_dir, "/") ? "" : "/") + name;
}

bool create_status_file(const Config &cfg, const std::string &name) {
    const std::string p = status_path(cfg, name);
    int fd = ::open(p.c_str(), O_CREAT | O_WRONLY | O_TRUNC, 0644);
    if (fd < 0) return false;
    ::close(fd);
    return true;
}

bool delete_status_file(const Config &cfg, const std::string &name) {
    const std::string p = status_path(cfg, name);
    if (::unlink(p.c_str()) == 0) return true;
    return errno == ENOENT;
}

namespace {
// Removes cuda-gpu<i>-ready left by an earlier run (same effect as upstream's preStop rm -f).
bool delete_per_gpu_files(const Config &cfg) {
    DIR *d = ::opendir(cfg.output_dir.c_str());
    if (!d) return errno == ENOENT;
    bool ok = true;
    while (dirent *e = ::readdir(d)) {
        std::string name = e->d_name;
        if (name.rfind("cuda-gpu", 0) == 0 && ends_with(name, "-ready")) ok = delete_status_file(cfg, name) && ok;
    }
    ::closedir(d);
    return ok;
}
}  // namespace

// ---- the sample -------------------------------------------------------------------------------------

long long verify_result(const float *A, const float *B, const float *C, size_t n) {
    for (size_t i = 0; i < n; ++i)
        if (std::fabs(A[i] + B[i] - C[i]) > 1e-5) return (long long)i;
    return -1;
}

int run_vectoradd_sample(const Api &api, int dev, size_t n, long long inject_fault, FILE *out, FILE *err,
                         b200v_timing *timing_out, const std::function<void()> *before_device) {
    std::fprintf(out, "[Vector addition of %zu elements]\n", n);
    float *A = (float *)std::malloc(n ? n * sizeof(float) : 1);
    float *B = (float *)std::malloc(n ? n * sizeof(float) : 1);
    float *C = (float *)std::malloc(n ? n * sizeof(float) : 1);
    struct Free {
        float *a, *b, *c;
        ~Free() { std::free(a); std::free(b); std::free(c); }
    } guard{A, B, C};
    if (!A || !B || !C) {
        std::fprintf(err, "Failed to allocate host vectors!\n");
        return 1;
    }
    // Host init as in the sample: glibc rand() at its default seed, A and B drawn alternately.
    std::srand(1);
    for (size_t i = 0; i < n; ++i) {
        A[i] = std::rand() / (float)RAND_MAX;
        B[i] = std::rand() / (float)RAND_MAX;
    }
    if (before_device) (*before_device)();   // fan-out children meet here: host init done, GPU work next
    std::fprintf(out, "Copy input data from the host memory to the CUDA device\n");
    b200v_timing t;
    std::memset(&t, 0, sizeof t);
    int rc = api.vectoradd(dev, A, B, C, n, &t);
    if (timing_out) *timing_out = t;
    if (rc != B200V_OK) {
        std::fprintf(err, "Failed to launch vectorAdd kernel (error code %s)!\n", api.strerror(rc));
        return 1;
    }
    std::fprintf(out, "CUDA kernel launch with %d blocks of %d threads\n", t.grid, t.block);
    std::fprintf(out, "Copy output data from the CUDA device to the host memory\n");
    if (inject_fault >= 0 && (size_t)inject_fault < n) C[inject_fault] += 1.0f;

    long long bad = verify_result(A, B, C, n);
    if (bad >= 0) {
        std::fprintf(err, "Result verification failed at element %lld!\n", bad);
        return 1;
    }
    // Stricter than the sample: IEEE fp32 add is deterministic, so the bytes must match too.
    for (size_t i = 0; i < n; ++i) {
        const float want = A[i] + B[i] + 0.0f;
        if (std::memcmp(&want, &C[i], sizeof(float)) != 0) {
            std::fprintf(err, "Result verification failed at element %zu! (bitwise)\n", i);
            return 1;
        }
    }
    std::fprintf(out, "Test PASSED\n");
    std::fprintf(out, "Done\n");
    std::fflush(out);
    return 0;
}

// ---- process helpers ---------------------------------------------------------------------------------

int run_command(const std::string &cmd) {
    int st = std::system(cmd.c_str());
    if (st == -1) return -1;
    if (WIFEXITED(st)) return WEXITSTATUS(st);
    return 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
}

// Fan-out plumbing.  Three kinds of pipe connect the parent to its per-device children:
//   ready   every child writes one byte once its GPU context exists and its host inputs are filled
//   go      the parent closes the write end when all children are ready (or after
//           B200V_FANOUT_BARRIER_TIMEOUT_S); children block on the read end until then, so the GPU
//           phases of all devices start together (SURVEY.md §8d cfg3 "started behind a barrier")
//   result  one per child: its JSON result row, read by the parent after the child exits
// The child learns its three descriptors from B200V_FANOUT_FDS="ready_w,go_r,result_w".
std::vector<int> fan_out(const Config &cfg, int ndev, const std::vector<std::string> &argv,
                         std::vector<std::string> *rows) {
    std::vector<pid_t> pids((size_t)ndev, -1);
    std::vector<int> results((size_t)ndev, 1);
    if (rows) rows->assign((size_t)ndev, std::string());
    int ready[2] = {-1, -1}, go[2] = {-1, -1};
    std::vector<int> res_r((size_t)ndev, -1);
    const bool piped = ::pipe2(ready, O_CLOEXEC) == 0 && ::pipe2(go, O_CLOEXEC) == 0;
    std::fflush(nullptr);
    for (int i = 0; i < ndev; ++i) {
        int res[2] = {-1, -1};
        const bool have_res = piped && ::pipe2(res, O_CLOEXEC) == 0;
        pid_t pid = ::fork();
        if (pid == 0) {
            ::setenv("B200V_CHILD_DEVICE", std::to_string(i).c_str(), 1);
            if (have_res) {
                // only these three survive the exec; every other pipe end is close-on-exec
                for (int fd : {ready[1], go[0], res[1]}) ::fcntl(fd, F_SETFD, 0);
                const std::string fds = std::to_string(ready[1]) + "," + std::to_string(go[0]) + "," +
                                        std::to_string(res[1]);
                ::setenv("B200V_FANOUT_FDS", fds.c_str(), 1);
            } else {
                ::unsetenv("B200V_FANOUT_FDS");
            }
            std::vector<char *> av;
            av.push_back(const_cast<char *>(cfg.self_exe.c_str()));
            for (const std::string &a : argv) av.push_back(const_cast<char *>(a.c_str()));
            av.push_back(nullptr);
            ::execv(cfg.self_exe.c_str(), av.data());
            ::_exit(127);
        }
        pids[(size_t)i] = pid;
        if (have_res) {
            ::close(res[1]);
            res_r[(size_t)i] = res[0];
        }
    }
    if (piped) {
        // The barrier: wait for one byte per child.  A child that dies early closes its end, and
        // when every writer is gone read() returns 0, so a failed child cannot wedge the others
        // beyond the timeout.
        ::close(ready[1]);
        ::close(go[0]);
        int arrived = 0;
        const int limit_ms = env_int("B200V_FANOUT_BARRIER_TIMEOUT_S", 300) * 1000;
        struct timespec t0;
        ::clock_gettime(CLOCK_MONOTONIC, &t0);
        while (arrived < ndev) {
            struct timespec now;
            ::clock_gettime(CLOCK_MONOTONIC, &now);
            const long waited = (now.tv_sec - t0.tv_sec) * 1000L + (now.tv_nsec - t0.tv_nsec) / 1000000L;
            if (waited >= limit_ms) break;
            struct pollfd pfd = {ready[0], POLLIN, 0};
            const int pr = ::poll(&pfd, 1, (int)(limit_ms - waited));
            if (pr < 0 && errno == EINTR) continue;
            if (pr <= 0) break;
            char buf[64];
            const ssize_t got = ::read(ready[0], buf, sizeof buf);
            if (got <= 0) break;
            arrived += (int)got;
        }
        ::close(go[1]);          // releases every child blocked on the go pipe
        ::close(ready[0]);
    } else {
        for (int fd : {ready[0], ready[1], go[0], go[1]}) if (fd >= 0) ::close(fd);
    }
    for (int i = 0; i < ndev; ++i) {
        if (pids[(size_t)i] >= 0) {
            int st = 0;
            if (::waitpid(pids[(size_t)i], &st, 0) >= 0)
                results[(size_t)i] = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
        }
        if (res_r[(size_t)i] >= 0) {
            // a row is far smaller than a pipe buffer, so the child never blocked writing it
            std::string row;
            char buf[4096];
            ssize_t got;
            while ((got = ::read(res_r[(size_t)i], buf, sizeof buf)) > 0) row.append(buf, (size_t)got);
            ::close(res_r[(size_t)i]);
            if (rows) (*rows)[(size_t)i] = row;
        }
    }
    return results;
}

FanoutChannel::FanoutChannel() {
    const std::string v = env_or("B200V_FANOUT_FDS", "");
    int a = -1, b = -1, c = -1;
    if (std::sscanf(v.c_str(), "%d,%d,%d", &a, &b, &c) == 3 && a >= 0 && b >= 0 && c >= 0) {
        ready_fd = a;
        go_fd = b;
        result_fd = c;
    }
}

void FanoutChannel::arrive_and_wait() {
    if (ready_fd >= 0) {
        const char one = 1;
        ssize_t ignored = ::write(ready_fd, &one, 1);
        (void)ignored;
        ::close(ready_fd);
        ready_fd = -1;
    }
    if (go_fd >= 0) {
        char buf;
        while (::read(go_fd, &buf, 1) < 0 && errno == EINTR) {}   // returns 0 when the parent closes it
        ::close(go_fd);
        go_fd = -1;
    }
}

void FanoutChannel::report(const std::string &row) {
    if (result_fd < 0) return;
    size_t off = 0;
    while (off < row.size()) {
        const ssize_t w = ::write(result_fd, row.data() + off, row.size() - off);
        if (w <= 0) break;
        off += (size_t)w;
    }
    ::close(result_fd);
    result_fd = -1;
}

FanoutChannel::~FanoutChannel() {
    // a child that leaves without arriving must not hold the others up: closing the write end of
    // `ready` is what lets the parent see that nobody else is coming
    for (int *fd : {&ready_fd, &go_fd, &result_fd})
        if (*fd >= 0) ::close(*fd);
}

namespace {
// Value of "key":<number> in a flat JSON row written by this file (no nesting, no escapes).
double row_number(const std::string &row, const std::string &key, double dflt) {
    const std::string k = "\"" + key + "\":";
    const size_t at = row.find(k);
    if (at == std::string::npos) return dflt;
    char *end = nullptr;
    const double v = std::strtod(row.c_str() + at + k.size(), &end);
    return end && end != row.c_str() + at + k.size() ? v : dflt;
}
}  // namespace

// One fan-out child: validates exactly one device, writes no status file (the parent does), and
// hands the parent one JSON row with what it measured.
int fanout_child_main(const Api &api, const Config &cfg, FILE *out, FILE *err) {
    FanoutChannel ch;
    const int dev = cfg.child_device;
    const long long inj = (env_int("B200V_INJECT_DEVICE", -1) < 0 || env_int("B200V_INJECT_DEVICE", -1) == dev)
                              ? cfg.inject_fault : -1;
    // Create the GPU context (and the library's streams) before the barrier, so that what starts
    // together after it is GPU work and not context creation.
    {
        float a[16] = {0}, b[16] = {0}, c[16];
        (void)api.vectoradd(dev, a, b, c, 16, nullptr);
    }
    const std::function<void()> meet = [&ch] { ch.arrive_and_wait(); };
    b200v_timing t;
    std::memset(&t, 0, sizeof t);
    const int rc = run_vectoradd_sample(api, dev, cfg.elems, inj, out, err, &t, &meet);
    const double bytes = 12.0 * (double)cfg.elems;
    char row[1024];
    int len = std::snprintf(row, sizeof row,
                            "{\"mode\":\"fanout-child\",\"gpu\":%d,\"elems\":%zu,\"passed\":%d,\"h2d_ms\":%.4f,"
                            "\"kernel_ms\":%.4f,\"d2h_ms\":%.4f,\"total_ms\":%.4f,\"kernel_launches\":%d,"
                            "\"kernel_gbs\":%.1f,\"e2e_gbs\":%.2f",
                            dev, cfg.elems, rc == 0 ? 1 : 0, t.h2d_ms, t.kernel_ms, t.d2h_ms, t.total_ms,
                            t.kernel_launches, t.kernel_ms > 0 ? bytes / (t.kernel_ms * 1e-3) / 1e9 : 0.0,
                            t.total_ms > 0 ? bytes / (t.total_ms * 1e-3) / 1e9 : 0.0);
    // Device-resident HBM figure (cfg3's "aggregate HBM GB/s"): the same vectorAdd kernel on arrays
    // generated on the GPU, one launch per timed sample.  On for validation sizes that are HBM-bound
    // anyway (B200V_ELEMS >= 2^22), or at the size B200V_FANOUT_HBM_LOG2 names; 0 turns it off.
    int hbm_log2 = env_int("B200V_FANOUT_HBM_LOG2", -1);
    if (hbm_log2 < 0) {
        hbm_log2 = 0;
        if (cfg.elems >= ((size_t)1 << 22))
            while (((size_t)2 << hbm_log2) <= cfg.elems) ++hbm_log2;
    }
    int rc2 = 0;
    if (rc == 0 && hbm_log2 > 0 && hbm_log2 <= 32) {
        b200v_stream_result r;
        std::memset(&r, 0, sizeof r);
        const int src = api.stream(dev, B200V_OP_VECTORADD, (size_t)1 << hbm_log2, 0.0f,
                                   0xB200ull + 0x1000ull * (uint64_t)dev, 5, 20, &r);
        if (src != B200V_OK || r.mismatches) {
            std::fprintf(err, "device-resident vectorAdd on gpu %d failed: %s\n", dev,
                         api.strerror(src != B200V_OK ? src : B200V_E_MISMATCH));
            rc2 = 1;
        }
        len += std::snprintf(row + len, sizeof row - (size_t)len,
                             ",\"hbm_log2_n\":%d,\"hbm_median_ms\":%.5f,\"hbm_median_gbs\":%.1f,\"hbm_best_gbs\":%.1f,"
                             "\"hbm_mismatches\":%" PRIu64 ",\"hbm_digest\":\"%016" PRIx64 "\"",
                             hbm_log2, r.median_ms, r.median_gbs, r.best_gbs, r.mismatches, r.digest);
    }
    std::snprintf(row + len, sizeof row - (size_t)len, "}\n");
    ch.report(row);
    return (rc == 0 && rc2 == 0) ? 0 : 1;
}

// ---- COMPONENT handlers --------------------------------------------------------------------------------

namespace {
void write_prom(const Config &cfg, const std::string &file, const std::string &text);

// Prints every child's row, then the whole-node figures: per-GPU and summed
// device-resident HBM GB/s, the job figure "all GPUs' bytes over the slowest GPU's time", and
// the same for the host-buffer call.  Leaves fanout-bandwidth.prom for the metrics mode.
void report_fanout(const Config &cfg, int ndev, const std::vector<std::string> &rows, FILE *out) {
    std::string per_hbm, per_e2e;
    double sum_hbm = 0, slowest_hbm_ms = 0, slowest_total_ms = 0, hbm_bytes = 0;
    int with_hbm = 0, reported = 0;
    std::string prom = "# HELP b200v_fanout_hbm_bandwidth_gbs device-resident vectorAdd GB/s of each GPU, all GPUs running at once\n"
                       "# TYPE b200v_fanout_hbm_bandwidth_gbs gauge\n";
    for (int i = 0; i < ndev; ++i) {
        const std::string &row = rows[(size_t)i];
        if (row.empty()) continue;
        ++reported;
        std::fputs(row.c_str(), out);
        const double hbm = row_number(row, "hbm_median_gbs", -1.0), hbm_ms = row_number(row, "hbm_median_ms", 0.0);
        const double e2e = row_number(row, "e2e_gbs", 0.0), total_ms = row_number(row, "total_ms", 0.0);
        char buf[128];
        if (hbm >= 0) {
            ++with_hbm;
            sum_hbm += hbm;
            slowest_hbm_ms = std::max(slowest_hbm_ms, hbm_ms);
            hbm_bytes = 12.0 * std::ldexp(1.0, (int)row_number(row, "hbm_log2_n", 0.0));
            std::snprintf(buf, sizeof buf, "%s%.1f", per_hbm.empty() ? "" : ",", hbm);
            per_hbm += buf;
            std::snprintf(buf, sizeof buf, "b200v_fanout_hbm_bandwidth_gbs{gpu=\"%d\"} %.1f\n", i, hbm);
            prom += buf;
        }
        slowest_total_ms = std::max(slowest_total_ms, total_ms);
        std::snprintf(buf, sizeof buf, "%s%.2f", per_e2e.empty() ? "" : ",", e2e);
        per_e2e += buf;
    }
    if (!reported) return;
    const double job_hbm = slowest_hbm_ms > 0 ? with_hbm * hbm_bytes / (slowest_hbm_ms * 1e-3) / 1e9 : 0.0;
    const double job_e2e = slowest_total_ms > 0 ? reported * 12.0 * (double)cfg.elems / (slowest_total_ms * 1e-3) / 1e9 : 0.0;
    std::fprintf(out,
                 "{\"mode\":\"fanout\",\"gpus\":%d,\"reported\":%d,\"elems\":%zu,\"per_gpu_hbm_gbs\":[%s],"
                 "\"sum_hbm_gbs\":%.1f,\"aggregate_hbm_gbs\":%.1f,\"slowest_hbm_ms\":%.5f,\"per_gpu_e2e_gbs\":[%s],"
                 "\"aggregate_e2e_gbs\":%.2f,\"slowest_total_ms\":%.4f}\n",
                 ndev, reported, cfg.elems, per_hbm.c_str(), sum_hbm, job_hbm, slowest_hbm_ms, per_e2e.c_str(), job_e2e,
                 slowest_total_ms);
    std::fflush(out);
    if (with_hbm) {
        char buf[512];
        std::snprintf(buf, sizeof buf,
                      "# HELP b200v_fanout_hbm_aggregate_gbs all GPUs' bytes over the slowest GPU's time\n"
                      "# TYPE b200v_fanout_hbm_aggregate_gbs gauge\nb200v_fanout_hbm_aggregate_gbs %.1f\n", job_hbm);
        prom += buf;
        write_prom(cfg, "fanout-bandwidth.prom", prom);
    }
}
}  // namespace

int validate_cuda(const Api &api, const Config &cfg, FILE *out, FILE *err) {
    if (!delete_status_file(cfg, "cuda-ready") || !delete_per_gpu_files(cfg)) {
        std::fprintf(err, "unable to remove stale status files under %s: %s\n", cfg.output_dir.c_str(),
                     std::strerror(errno));
        return 1;
    }
    if (!cfg.with_workload) {
        // Upstream's behaviour with WITH_WORKLOAD=false: no workload is run, the status file is
        // simply (re)created so the next init container is not blocked.
        if (!create_status_file(cfg, "cuda-ready")) {
            std::fprintf(err, "unable to create status file %s: %s\n", status_path(cfg, "cuda-ready").c_str(),
                         std::strerror(errno));
            return 1;
        }
        std::fprintf(out, "cuda workload validation is disabled (WITH_WORKLOAD=false)\n");
        return 0;
    }
    int ndev = 0;
    int rc = api.device_count(&ndev);
    if (rc != B200V_OK || ndev < 1) {
        std::fprintf(err, "cuda validation failed: %s\n", api.strerror(rc != B200V_OK ? rc : B200V_E_NODEVICE));
        return 1;
    }
    if (cfg.gpus > ndev) {
        std::fprintf(err, "cuda validation failed: B200V_GPUS=%d but only %d device(s) visible\n", cfg.gpus, ndev);
        return 1;
    }
    const int use = cfg.gpus > 0 ? cfg.gpus : ndev;
    std::vector<int> results;
    // One device: the sample runs in this process, as upstream's does.  B200V_FANOUT_FORCE=1 sends
    // even a single device through the child-process path (start barrier, result row, aggregate).
    if (use == 1 && !parse_bool(env_or("B200V_FANOUT_FORCE", ""), false)) {
        results.push_back(run_vectoradd_sample(api, 0, cfg.elems, cfg.inject_fault, out, err));
    } else {
        // One validator instance per device: independent shards, no exchange.
        std::vector<std::string> args = {"--component", "cuda", "--output-dir", cfg.output_dir, "--elems",
                                         std::to_string(cfg.elems)};
        std::vector<std::string> rows;
        results = fan_out(cfg, use, args, &rows);
        report_fanout(cfg, use, rows, out);
    }
    bool all = true;
    for (int i = 0; i < use; ++i) {
        if (results[(size_t)i] == 0) {
            if (!create_status_file(cfg, "cuda-gpu" + std::to_string(i) + "-ready")) {
                std::fprintf(err, "unable to create status file for gpu %d: %s\n", i, std::strerror(errno));
                all = false;
            }
        } else {
            std::fprintf(err, "cuda validation failed on gpu %d (exit code %d)\n", i, results[(size_t)i]);
            all = false;
        }
    }
    if (!all) return 1;
    if (!create_status_file(cfg, "cuda-ready")) {
        std::fprintf(err, "unable to create status file %s: %s\n", status_path(cfg, "cuda-ready").c_str(),
                     std::strerror(errno));
        return 1;
    }
    std::fprintf(out, "cuda workload validation is successful on %d device(s)\n", use);
    return 0;
}

namespace {
int validate_with_smi(const Config &cfg, const std::string &file, const std::string &cmd, const char *what,
                      FILE *out, FILE *err) {
    if (!delete_status_file(cfg, file)) {
        std::fprintf(err, "unable to remove stale %s: %s\n", file.c_str(), std::strerror(errno));
        return 1;
    }
    std::fflush(nullptr);
    int rc = run_command(cmd);
    if (rc != 0) {
        std::fprintf(err, "%s validation failed: `%s` exited with %d\n", what, cmd.c_str(), rc);
        return 1;
    }
    if (!create_status_file(cfg, file)) {
        std::fprintf(err, "unable to create status file %s: %s\n", file.c_str(), std::strerror(errno));
        return 1;
    }
    std::fprintf(out, "%s validation is successful\n", what);
    return 0;
}
}  // namespace

int validate_driver(const Config &cfg, FILE *out, FILE *err) {
    // Containerised driver: nvidia-smi lives under the driver root (what the runbook execs into at
    // reference README.md:152); host-installed driver: plain nvidia-smi.
    struct stat st;
    const std::string in_root = cfg.driver_root + "/usr/bin/nvidia-smi";
    std::string cmd = cfg.nvidia_smi;
    if (cfg.nvidia_smi == "nvidia-smi" && ::stat(in_root.c_str(), &st) == 0)
        cmd = "chroot " + cfg.driver_root + " nvidia-smi";
    return validate_with_smi(cfg, "driver-ready", cmd, "driver", out, err);
}

int validate_toolkit(const Config &cfg, FILE *out, FILE *err) {
    return validate_with_smi(cfg, "toolkit-ready", cfg.nvidia_smi, "toolkit", out, err);
}

// ---- COMPONENT=plugin --------------------------------------------------------------------------------------

namespace {
// First integer that follows `key` in `text` (skipping separators and quotes), or -1.
long long number_after(const std::string &text, const std::string &key) {
    size_t at = text.find(key);
    if (at == std::string::npos) return -1;
    at += key.size();
    while (at < text.size() && (text[at] == ':' || text[at] == '"' || text[at] == '\'' || text[at] == ' ' ||
                                text[at] == '\t' || text[at] == '='))
        ++at;
    size_t end = at;
    while (end < text.size() && text[end] >= '0' && text[end] <= '9') ++end;
    long long v;
    return (end > at && parse_ll(text.substr(at, end - at), v)) ? v : -1;
}

std::string read_file(const std::string &path, bool &ok) {
    std::string s;
    FILE *f = std::fopen(path.c_str(), "rb");
    ok = f != nullptr;
    if (!f) return s;
    char buf[4096];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, f)) > 0) s.append(buf, got);
    std::fclose(f);
    return s;
}
}  // namespace

long long read_allocatable_gpus(const Config &cfg) {
    long long v;
    if (!cfg.allocatable_gpus.empty()) return parse_ll(cfg.allocatable_gpus, v) && v >= 0 ? v : 0;
    if (cfg.allocatable_file.empty()) return -1;
    bool ok = false;
    std::string text = read_file(cfg.allocatable_file, ok);
    if (!ok) return 0;                                   // not written yet: nothing advertised so far
    while (!text.empty() && (text.back() == '\n' || text.back() == ' ' || text.back() == '\r')) text.pop_back();
    if (parse_ll(text, v)) return v >= 0 ? v : 0;        // a bare number
    v = number_after(text, "nvidia.com/gpu");            // `nvidia.com/gpu: 8`, JSON or describe output
    return v >= 0 ? v : 0;
}

int validate_plugin(const Api &api, const Config &cfg, FILE *out, FILE *err) {
    if (!delete_status_file(cfg, "plugin-ready")) {
        std::fprintf(err, "unable to remove stale plugin-ready: %s\n", std::strerror(errno));
        return 1;
    }
    long long gpus = read_allocatable_gpus(cfg);
    if (gpus == -1) {
        std::fprintf(err, "plugin validation needs the node's allocatable nvidia.com/gpu count: set "
                          "B200V_ALLOCATABLE_GPUS or B200V_ALLOCATABLE_FILE (no Kubernetes client in this build)\n");
        return 1;
    }
    for (int attempt = 1; gpus <= 0 && attempt < cfg.plugin_retries; ++attempt) {
        std::fprintf(err, "GPU resources are not yet discovered by the node, retry: %d\n", attempt);
        std::fflush(nullptr);
        ::sleep((unsigned)cfg.sleep_seconds);
        gpus = read_allocatable_gpus(cfg);
    }
    if (gpus <= 0) {
        std::fprintf(err, "GPU resources are not discovered by the node\n");
        return 1;
    }
    std::fprintf(out, "node advertises nvidia.com/gpu: %lld\n", gpus);
    if (cfg.with_workload) {
        // Upstream runs the workload in a pod limited to nvidia.com/gpu: 1; here the container's
        // own first device (what the device plugin granted it) runs it in-process.
        if (run_vectoradd_sample(api, 0, cfg.elems, cfg.inject_fault, out, err) != 0) return 1;
    }
    if (!create_status_file(cfg, "plugin-ready")) {
        std::fprintf(err, "unable to create status file plugin-ready: %s\n", std::strerror(errno));
        return 1;
    }
    // Upstream's file is empty and only its presence is read; the count the device plugin
    // advertised is kept in it so the metrics mode can export it without a Kubernetes client.
    if (FILE *f = std::fopen(status_path(cfg, "plugin-ready").c_str(), "wb")) {
        std::fprintf(f, "nvidia.com/gpu: %lld\n", gpus);
        std::fclose(f);
    }
    std::fprintf(out, "plugin validation is successful\n");
    return 0;
}

// ---- COMPONENT=metrics ---------------------------------------------------------------------------------------

std::string render_metrics(const Config &cfg) {
    std::string m;
    struct stat st;
    const char *components[] = {"driver", "toolkit", "cuda", "plugin"};
    for (const char *c : components) {
        const std::string name = std::string("gpu_operator_node_") + c + "_ready";
        const bool ready = ::stat(status_path(cfg, std::string(c) + "-ready").c_str(), &st) == 0;
        m += "# HELP " + name + " 1 when " + c + " validation has left its status file on this node\n";
        m += "# TYPE " + name + " gauge\n";
        m += name + (ready ? " 1\n" : " 0\n");
        // when it last succeeded = when its status file was (re)created
        const std::string ts = std::string("gpu_operator_node_") + c + "_validation_last_success_ts_seconds";
        m += "# HELP " + ts + " unix time of the last successful " + c + " validation (0 = never)\n";
        m += "# TYPE " + ts + " gauge\n";
        m += ts + " " + std::to_string(ready ? (long long)st.st_mtime : 0LL) + "\n";
    }
    // per-GPU fan-out files and bandwidth results
    std::vector<std::string> gpu_lines, prom_files;
    if (DIR *d = ::opendir(cfg.output_dir.c_str())) {
        while (dirent *e = ::readdir(d)) {
            const std::string f = e->d_name;
            if (f.rfind("cuda-gpu", 0) == 0 && ends_with(f, "-ready")) {
                const std::string idx = f.substr(8, f.size() - 8 - 6);
                gpu_lines.push_back("b200v_node_cuda_gpu_ready{gpu=\"" + idx + "\"} 1\n");
            } else if (ends_with(f, ".prom")) {
                prom_files.push_back(f);
            }
        }
        ::closedir(d);
    }
    std::sort(gpu_lines.begin(), gpu_lines.end());
    std::sort(prom_files.begin(), prom_files.end());
    m += "# HELP b200v_node_cuda_gpu_ready 1 for every GPU whose own vectorAdd validation passed\n";
    m += "# TYPE b200v_node_cuda_gpu_ready gauge\n";
    for (const std::string &l : gpu_lines) m += l;
    // Upstream's meaning: the GPU count the device plugin advertises, i.e. what plugin validation
    // read from the node's allocatable nvidia.com/gpu and left in plugin-ready (0 until it has run).
    long long advertised = 0;
    {
        bool ok = false;
        const std::string text = read_file(status_path(cfg, "plugin-ready"), ok);
        if (ok) advertised = std::max(0LL, number_after(text, "nvidia.com/gpu"));
    }
    m += "# HELP gpu_operator_node_device_plugin_devices_total GPUs the device plugin advertises on this node (allocatable nvidia.com/gpu seen by plugin validation)\n";
    m += "# TYPE gpu_operator_node_device_plugin_devices_total gauge\n";
    m += "gpu_operator_node_device_plugin_devices_total " + std::to_string(advertised) + "\n";
    // Several result files may describe the same metric family (one file per GPU): its HELP and
    // TYPE lines go out once.
    std::vector<std::string> seen_meta;
    for (const std::string &f : prom_files) {
        bool ok = false;
        const std::string text = read_file(status_path(cfg, f), ok);
        if (!ok) continue;
        size_t at = 0;
        while (at < text.size()) {
            size_t end = text.find('\n', at);
            if (end == std::string::npos) end = text.size();
            const std::string line = text.substr(at, end - at);
            at = end + 1;
            if (line.empty()) continue;
            if (line.rfind("# HELP ", 0) == 0 || line.rfind("# TYPE ", 0) == 0) {
                if (std::find(seen_meta.begin(), seen_meta.end(), line) != seen_meta.end()) continue;
                seen_meta.push_back(line);
            }
            m += line + "\n";
        }
    }
    return m;
}

int serve_metrics(const Config &cfg, FILE *out, FILE *err) {
    if (parse_bool(env_or("B200V_METRICS_ONESHOT", ""), false)) {
        const std::string m = render_metrics(cfg);
        std::fwrite(m.data(), 1, m.size(), out);
        std::fflush(out);
        return 0;
    }
    const int max_requests = env_int("B200V_METRICS_MAX_REQUESTS", 0);   // 0 = serve forever
    int srv = ::socket(AF_INET, SOCK_STREAM, 0);
    if (srv < 0) {
        std::fprintf(err, "metrics: socket: %s\n", std::strerror(errno));
        return 1;
    }
    int one = 1;
    ::setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr;
    std::memset(&addr, 0, sizeof addr);
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons((uint16_t)cfg.metrics_port);
    if (::bind(srv, (sockaddr *)&addr, sizeof addr) != 0 || ::listen(srv, 16) != 0) {
        std::fprintf(err, "metrics: cannot listen on port %d: %s\n", cfg.metrics_port, std::strerror(errno));
        ::close(srv);
        return 1;
    }
    socklen_t alen = sizeof addr;
    ::getsockname(srv, (sockaddr *)&addr, &alen);
    std::fprintf(out, "metrics: listening on :%d/metrics\n", (int)ntohs(addr.sin_port));
    std::fflush(out);
    for (int served = 0; max_requests == 0 || served < max_requests; ++served) {
        int c = ::accept(srv, nullptr, nullptr);
        if (c < 0) {
            if (errno == EINTR) { --served; continue; }
            break;
        }
        char req[2048];
        ssize_t got = ::recv(c, req, sizeof req - 1, 0);
        req[got > 0 ? got : 0] = 0;
        const bool is_metrics = std::strncmp(req, "GET /metrics", 12) == 0 || std::strncmp(req, "GET / ", 6) == 0;
        const std::string body = is_metrics ? render_metrics(cfg) : std::string("not found\n");
        std::string resp = std::string(is_metrics ? "HTTP/1.1 200 OK\r\n" : "HTTP/1.1 404 Not Found\r\n") +
                           "Content-Type: text/plain; version=0.0.4\r\nConnection: close\r\nContent-Length: " +
                           std::to_string(body.size()) + "\r\n\r\n" + body;
        size_t off = 0;
        while (off < resp.size()) {
            ssize_t w = ::send(c, resp.data() + off, resp.size() - off, MSG_NOSIGNAL);
            if (w <= 0) break;
            off += (size_t)w;
        }
        ::close(c);
    }
    ::close(srv);
    return 0;
}

namespace {
// Textfile-collector style result files the metrics mode re-exports.
void write_prom(const Config &cfg, const std::string &file, const std::string &text) {
    if (!cfg.write_prom) return;
    const std::string tmp = status_path(cfg, file + ".tmp"), dst = status_path(cfg, file);
    if (FILE *f = std::fopen(tmp.c_str(), "wb")) {
        std::fwrite(text.data(), 1, text.size(), f);
        std::fclose(f);
        std::rename(tmp.c_str(), dst.c_str());
    }
}
}  // namespace

// ---- sweeps -------------------------------------------------------------------------------------------

int run_stream_sweep(const Api &api, const Config &cfg, FILE *out, FILE *err) {
    int ndev = 0;
    int rc = api.device_count(&ndev);
    if (rc != B200V_OK || ndev < 1) {
        std::fprintf(err, "stream sweep failed: %s\n", api.strerror(rc != B200V_OK ? rc : B200V_E_NODEVICE));
        return 1;
    }
    // B200V_GPUS > 1: one sweep per device, each in its own process, all at once — the per-GPU
    // fan-out of the validation applied to the bandwidth sweep.  Each child prints
    // its rows with its own "gpu" index; the sweep passes when every child does.
    if (cfg.child_device < 0 && cfg.gpus > 1) {
        if (cfg.gpus > ndev) {
            std::fprintf(err, "stream sweep failed: B200V_GPUS=%d but only %d device(s) visible\n", cfg.gpus, ndev);
            return 1;
        }
        std::vector<int> results = fan_out(cfg, cfg.gpus, {"--mode", "stream"});
        int bad = 0;
        for (int i = 0; i < cfg.gpus; ++i)
            if (results[(size_t)i] != 0) {
                std::fprintf(err, "stream sweep failed on gpu %d (exit code %d)\n", i, results[(size_t)i]);
                ++bad;
            }
        return bad ? 1 : 0;
    }
    const int dev = cfg.child_device >= 0 ? cfg.child_device : 0;
    if (cfg.child_device >= 0) {
        // per-GPU sweeps start together: create the context, then meet the other children
        b200v_info warm;
        api.device_info(dev, &warm);
        FanoutChannel ch;
        ch.arrive_and_wait();
    }
    const int lo = env_int("B200V_MIN_LOG2", 18), hi = env_int("B200V_MAX_LOG2", 32);
    const int ops[] = {B200V_OP_COPY, B200V_OP_SCALE, B200V_OP_ADD, B200V_OP_TRIAD};
    b200v_info info;
    std::memset(&info, 0, sizeof info);
    api.device_info(dev, &info);
    const double nominal = 2.0 * info.mem_clock_khz * 1e3 * info.mem_bus_bits / 8.0 / 1e9;
    int failures = 0;
    std::string prom = "# HELP b200v_hbm_bandwidth_gbs median GB/s of the STREAM op on HBM-bound sizes\n"
                       "# TYPE b200v_hbm_bandwidth_gbs gauge\n";
    for (int lg = lo; lg <= hi; ++lg) {
        const size_t n = (size_t)1 << lg;
        for (int op : ops) {
            b200v_stream_result r;
            rc = api.stream(dev, op, n, 3.0f, 0xB200, 5, 20, &r);
            if (rc != B200V_OK && rc != B200V_E_MISMATCH) {
                std::fprintf(err, "stream %s n=2^%d failed: %s\n", kOpNames[op], lg, api.strerror(rc));
                ++failures;
                continue;
            }
            if (r.mismatches) ++failures;
            std::fprintf(out,
                         "{\"mode\":\"stream\",\"gpu\":%d,\"op\":\"%s\",\"log2_n\":%d,\"bytes_per_array\":%zu,"
                         "\"best_ms\":%.5f,\"median_ms\":%.5f,\"best_gbs\":%.1f,\"median_gbs\":%.1f,"
                         "\"l2_resident\":%d,\"frac_of_nominal\":%.4f,\"nominal_gbs\":%.0f,\"grid\":%d,\"block\":%d,"
                         "\"mismatches\":%" PRIu64 ",\"digest\":\"%016" PRIx64 "\"}\n",
                         dev, kOpNames[op], lg, n * 4, r.best_ms, r.median_ms, r.best_gbs, r.median_gbs,
                         r.l2_resident, nominal > 0 ? r.median_gbs / nominal : 0.0, nominal, r.grid, r.block,
                         r.mismatches, r.digest);
            std::fflush(out);
            if (!r.l2_resident && !r.mismatches) {
                char line[256];
                std::snprintf(line, sizeof line, "b200v_hbm_bandwidth_gbs{gpu=\"%d\",op=\"%s\",log2_n=\"%d\"} %.1f\n",
                              dev, kOpNames[op], lg, r.median_gbs);
                prom += line;
            }
        }
    }
    write_prom(cfg, cfg.child_device >= 0 ? "hbm-bandwidth-gpu" + std::to_string(dev) + ".prom" : "hbm-bandwidth.prom",
               prom);
    return failures ? 1 : 0;
}

int run_nccl_probe(const Api &api, const Config &cfg, FILE *out, FILE *err) {
    // B200V_MODE=nvls runs the same sweep with the hand-written NVLink-multicast kernel.
    const bool nvls = cfg.mode == "nvls";
    const char *mode = nvls ? "nvls" : "nccl";
    int ndev = 0;
    int rc = api.device_count(&ndev);
    if (rc != B200V_OK || ndev < 1) {
        std::fprintf(err, "nccl probe failed: %s\n", api.strerror(rc != B200V_OK ? rc : B200V_E_NODEVICE));
        return 1;
    }
    const int use = cfg.gpus > 0 ? cfg.gpus : ndev;
    if (use > ndev) {
        std::fprintf(err, "nccl probe failed: B200V_GPUS=%d but only %d device(s) visible\n", use, ndev);
        return 1;
    }
    std::vector<int> devs((size_t)use);
    for (int i = 0; i < use; ++i) devs[(size_t)i] = i;
    std::vector<b200v_ar_row> rows(64);
    int nrows = 0;
    const size_t lo = (size_t)1 << env_int("B200V_MIN_LOG2", 10), hi = (size_t)1 << env_int("B200V_MAX_LOG2", 30);
    // NCCL prints its version banner on stdout at communicator creation; keep stdout for the JSON
    // rows by pointing fd 1 at stderr for the duration of the call.
    std::fflush(nullptr);
    const int saved_fd = ::dup(1);
    if (saved_fd >= 0) ::dup2(2, 1);
    if (nvls && !api.nvls_allreduce_probe) rc = B200V_E_BADARG;
    else if (nvls) rc = api.nvls_allreduce_probe(use, devs.data(), std::max<size_t>(lo, 16), hi, 5, 20, rows.data(),
                                                 (int)rows.size(), &nrows);
    else rc = api.allreduce_probe(use, devs.data(), lo, hi, 5, 20, rows.data(), (int)rows.size(), &nrows);
    std::fflush(nullptr);
    if (saved_fd >= 0) {
        ::dup2(saved_fd, 1);
        ::close(saved_fd);
    }
    int failures = rc != B200V_OK ? 1 : 0;
    if (rc != B200V_OK) std::fprintf(err, "%s probe failed: %s\n", mode, api.strerror(rc));
    const double min_busbw = std::atof(env_or("B200V_MIN_BUSBW_GBS", "0").c_str());
    std::string prom = "# HELP b200v_nvlink_allreduce_busbw_gbs all-reduce bus bandwidth over NVLink, GB/s\n"
                       "# TYPE b200v_nvlink_allreduce_busbw_gbs gauge\n";
    for (int i = 0; i < nrows; ++i) {
        const b200v_ar_row &r = rows[(size_t)i];
        if (r.mismatches) ++failures;
        char line[256];
        std::snprintf(line, sizeof line,
                      "b200v_nvlink_allreduce_busbw_gbs{impl=\"%s\",ranks=\"%d\",bytes=\"%" PRIu64 "\"} %.2f\n", mode, use,
                      r.bytes, r.busbw_gbs);
        prom += line;
        std::fprintf(out,
                     "{\"mode\":\"%s\",\"ranks\":%d,\"bytes\":%" PRIu64 ",\"best_ms\":%.5f,\"median_ms\":%.5f,"
                     "\"algbw_gbs\":%.2f,\"busbw_gbs\":%.2f,\"frac_of_900\":%.4f,\"mismatches\":%" PRIu64 "}\n",
                     mode, use, r.bytes, r.best_ms, r.median_ms, r.algbw_gbs, r.busbw_gbs, r.busbw_gbs / 900.0,
                     r.mismatches);
    }
    if (nrows > 0 && min_busbw > 0 && rows[(size_t)nrows - 1].busbw_gbs < min_busbw) {
        std::fprintf(err, "%s probe failed: busBW %.1f GB/s below the required %.1f GB/s\n", mode,
                     rows[(size_t)nrows - 1].busbw_gbs, min_busbw);
        ++failures;
    }
    std::fflush(out);
    if (!failures) write_prom(cfg, nvls ? "nvlink-busbw-nvls.prom" : "nvlink-busbw.prom", prom);
    return failures ? 1 : 0;
}

// ---- executables -----------------------------------------------------------------------------------------

int validator_main(const Api &api, int argc, char **argv) {
    Config cfg;
    if (!parse_config(argc, argv, cfg)) {
        std::fprintf(stderr, "Incorrect Usage. %s\n", cfg.error.c_str());
        return 1;
    }
    if (cfg.mode != "validate") {
        // A health probe must not hang the node's validation: a collective that never completes
        // (a wedged link, a bad NCCL_* override) ends the process with exit code 1 after
        // B200V_PROBE_TIMEOUT_S seconds (default 600).
        const int limit = env_int("B200V_PROBE_TIMEOUT_S", 600);
        if (limit > 0) {
            ::signal(SIGALRM, [](int) {
                static const char msg[] = "probe timed out (B200V_PROBE_TIMEOUT_S)\n";
                ssize_t ignored = ::write(2, msg, sizeof msg - 1);
                (void)ignored;
                ::_exit(1);
            });
            ::alarm((unsigned)limit);
        }
    }
    if (cfg.mode == "stream") return run_stream_sweep(api, cfg, stdout, stderr);
    if (cfg.mode == "nccl" || cfg.mode == "nvls") return run_nccl_probe(api, cfg, stdout, stderr);
    if (cfg.component.empty()) {
        std::fprintf(stderr, "invalid -c <component-name> flag: must not be empty string\n");
        return 1;
    }
    if (cfg.component == "metrics") return serve_metrics(cfg, stdout, stderr);
    if (cfg.child_device >= 0) return fanout_child_main(api, cfg, stdout, stderr);
    auto once = [&]() -> int {
        if (cfg.component == "cuda") return validate_cuda(api, cfg, stdout, stderr);
        if (cfg.component == "driver") return validate_driver(cfg, stdout, stderr);
        if (cfg.component == "toolkit") return validate_toolkit(cfg, stdout, stderr);
        if (cfg.component == "plugin") return validate_plugin(api, cfg, stdout, stderr);
        return -1;
    };
    int tries = 0;
    for (;;) {
        int rc = once();
        if (rc == -1) {
            // mofed, vfio-pci, vgpu-manager, vgpu-devices, cc-manager and nvidia-fs belong to other
            // parts of the operator; this image covers the GPU workload path only.
            std::fprintf(stderr, "invalid component specified for validation: %s (this build supports cuda, "
                                 "driver, toolkit, plugin, metrics)\n", cfg.component.c_str());
            return 1;
        }
        if (rc == 0) return 0;
        ++tries;
        if (!cfg.with_wait || (cfg.max_retries > 0 && tries >= cfg.max_retries)) return 1;
        std::fprintf(stderr, "%s is not ready, retrying after %d seconds\n", cfg.component.c_str(), cfg.sleep_seconds);
        std::fflush(nullptr);
        ::sleep((unsigned)cfg.sleep_seconds);
    }
}

int vectoradd_main(const Api &api, int argc, char **argv) {
    // The sample takes no arguments; B200V_ELEMS / --elems and B200V_GPUS are the only additions.
    Config cfg;
    if (!parse_config(argc, argv, cfg)) {
        std::fprintf(stderr, "Incorrect Usage. %s\n", cfg.error.c_str());
        return 1;
    }
    const int dev = cfg.child_device >= 0 ? cfg.child_device : 0;
    int rc = run_vectoradd_sample(api, dev, cfg.elems, cfg.inject_fault, stdout, stderr);
    api.shutdown();
    return rc == 0 ? 0 : 1;
}

}  // namespace b200v_host
