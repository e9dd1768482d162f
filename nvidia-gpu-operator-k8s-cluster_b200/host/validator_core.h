// validator_core.h — host side of the validation workload, above the C ABI.
//
// Mirrors the operator-validator contract (upstream NVIDIA/gpu-operator
// validator/main.go and cuda-samples vectorAdd.cu, neither on disk here): the executables
// `vectorAdd` and `nvidia-validator`, the COMPONENT / WITH_WAIT / WITH_WORKLOAD / OUTPUT_DIR
// environment, exit code 0/1, and the <OUTPUT_DIR>/*-ready status files that the next init
// container and the node-status-exporter (reference README.md:107) read.
//
// The reference's host language is Go; Go is not installed in this image, so this C++ twin is
// what is built and tested and go/validator.go carries the same logic as cgo source.
// Everything GPU-side goes through the `Api` table, which the executables fill with the symbols
// of libb200validate.so and the host-logic tests fill with scripted fakes (no GPU needed there).
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

#include "b200validate.h"

namespace b200v_host {

struct Api {
    int (*device_count)(int *);
    int (*device_info)(int, b200v_info *);
    int (*vectoradd)(int, const float *, const float *, float *, size_t, b200v_timing *);
    int (*stream)(int, int, size_t, float, uint64_t, int, int, b200v_stream_result *);
    int (*allreduce_probe)(int, const int *, size_t, size_t, int, int, b200v_ar_row *, int, int *);
    const char *(*strerror)(int);
    int (*shutdown)(void);
    int (*nvls_allreduce_probe)(int, const int *, size_t, size_t, int, int, b200v_ar_row *, int, int *);
};

// The table bound to libb200validate.so (real_api.cpp; linked only into the product binaries).
Api real_api();

struct Config {
    // upstream names (env, each mirrored by a flag)
    std::string component;                       // COMPONENT / --component
    bool with_wait = false;                      // WITH_WAIT / --with-wait
    bool with_workload = true;                   // WITH_WORKLOAD / --with-workload
    std::string output_dir = "/run/nvidia/validations";   // OUTPUT_DIR / --output-dir
    std::string node_name;                       // NODE_NAME / --node-name
    std::string operator_namespace;              // OPERATOR_NAMESPACE / --namespace
    std::string visible_devices;                 // NVIDIA_VISIBLE_DEVICES (informational)
    // new, optional; an unmodified DaemonSet never sets them
    std::string mode = "validate";               // B200V_MODE: validate | stream | nccl
    size_t elems = 50000;                        // B200V_ELEMS (the sample's numElements)
    int gpus = 0;                                // B200V_GPUS: 0 = every visible device
    int sleep_seconds = 5;                       // B200V_SLEEP_SECONDS between WITH_WAIT retries
    int max_retries = 0;                         // B200V_MAX_RETRIES: 0 = forever (upstream)
    long long inject_fault = -1;                 // B200V_INJECT_FAULT: corrupt this output element
    int child_device = -1;                       // B200V_CHILD_DEVICE: set by the fan-out parent
    std::string allocatable_gpus;                // B200V_ALLOCATABLE_GPUS: node allocatable nvidia.com/gpu
    std::string allocatable_file;                // B200V_ALLOCATABLE_FILE: file holding that count
    int plugin_retries = 30;                     // B200V_PLUGIN_RETRIES (upstream: 30 x 5 s)
    int metrics_port = 8000;                     // METRICS_PORT / --metrics-port (upstream default)
    bool write_prom = false;                     // B200V_WRITE_PROM: sweeps leave *.prom files
    std::string nvidia_smi = "nvidia-smi";       // B200V_NVIDIA_SMI: command for driver/toolkit
    std::string driver_root = "/run/nvidia/driver";
    std::string self_exe;                        // path re-executed for fan-out children
    std::string error;                           // set by parse_config on a bad flag
};

// Reads the environment, then argv (flags win).  Returns false and sets cfg.error on a bad flag.
bool parse_config(int argc, char **argv, Config &cfg);
bool parse_bool(const std::string &s, bool dflt);

// Status files (SURVEY.md §8b "Outputs").
std::string status_path(const Config &cfg, const std::string &name);
bool create_status_file(const Config &cfg, const std::string &name);
bool delete_status_file(const Config &cfg, const std::string &name);   // true when absent after

// The sample: host init with rand(), device add through the C ABI, host verification loop.
// Prints the sample's lines to `out` / `err`.  Returns 0 on "Test PASSED", 1 otherwise.
// timing_out (optional) receives the library's timing of the device call; before_device (optional)
// runs after the host init and before the device call (the fan-out barrier).
int run_vectoradd_sample(const Api &api, int dev, size_t n, long long inject_fault, FILE *out,
                         FILE *err, b200v_timing *timing_out = nullptr,
                         const std::function<void()> *before_device = nullptr);

// The sample's host verification loop: first index with fabs(A[i]+B[i]-C[i]) > 1e-5, or -1.
long long verify_result(const float *A, const float *B, const float *C, size_t n);

// COMPONENT handlers.  Each returns 0 (status file written) or 1 (no status file left behind).
int validate_cuda(const Api &api, const Config &cfg, FILE *out, FILE *err);
int validate_driver(const Config &cfg, FILE *out, FILE *err);
int validate_toolkit(const Config &cfg, FILE *out, FILE *err);

// COMPONENT=plugin: wait until the node advertises nvidia.com/gpu (the field the
// runbook greps at reference README.md:122), then run the workload and create plugin-ready.  This
// build has no Kubernetes client, so the allocatable count comes from B200V_ALLOCATABLE_GPUS or
// from B200V_ALLOCATABLE_FILE (a bare number, or text containing `nvidia.com/gpu: N` as printed
// by `kubectl describe node` / a downward-API-fed sidecar).  -1 = no source configured.
long long read_allocatable_gpus(const Config &cfg);
int validate_plugin(const Api &api, const Config &cfg, FILE *out, FILE *err);

// COMPONENT=metrics: node-status-exporter compatible gauges from the status
// files (nodeStatusExporter.enabled=true at reference README.md:107), plus every *.prom file the
// sweeps left in OUTPUT_DIR, so HBM and NVLink bandwidth become node health signals.
std::string render_metrics(const Config &cfg);
int serve_metrics(const Config &cfg, FILE *out, FILE *err);

// B200V_MODE=stream / nccl: the bandwidth sweeps, one JSON line per row on `out`.
int run_stream_sweep(const Api &api, const Config &cfg, FILE *out, FILE *err);
int run_nccl_probe(const Api &api, const Config &cfg, FILE *out, FILE *err);

// Entry points of the two executables.
int validator_main(const Api &api, int argc, char **argv);
int vectoradd_main(const Api &api, int argc, char **argv);

// Runs a shell command, returns its exit status (-1 if it could not be started).
int run_command(const std::string &cmd);

// Spawns `self_exe` once per device with B200V_CHILD_DEVICE=i, holds the children at a start
// barrier until all are ready, and waits; results[i] = exit code, (*rows)[i] = the JSON row child i
// reported ("" if none).
std::vector<int> fan_out(const Config &cfg, int ndev, const std::vector<std::string> &argv,
                         std::vector<std::string> *rows = nullptr);

// The child's end of the fan-out pipes (B200V_FANOUT_FDS); every method is a no-op when the
// process was not started by fan_out.
struct FanoutChannel {
    int ready_fd = -1, go_fd = -1, result_fd = -1;
    FanoutChannel();                        // reads B200V_FANOUT_FDS
    FanoutChannel(const FanoutChannel &) = delete;
    FanoutChannel &operator=(const FanoutChannel &) = delete;
    void arrive_and_wait();                 // "I am ready", then block until every child is
    void report(const std::string &row);    // hand the parent this child's result row
    ~FanoutChannel();
};

// A fan-out child of COMPONENT=cuda: one device, no status files, one result row.
int fanout_child_main(const Api &api, const Config &cfg, FILE *out, FILE *err);

}  // namespace b200v_host
