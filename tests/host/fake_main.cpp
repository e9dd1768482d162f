// fake_main.cpp — TEST ONLY.  Links the product's host logic (host/validator_core.cpp) against a
// scripted stand-in for the C ABI so the validator contract (exit codes, status files, fan-out,
// retries, output lines) can be exercised on a machine with no GPU.  Never shipped: the product
// executables link real_api.cpp and libb200validate.so instead.
//
// Script (environment):
//   FAKE_ENTRY       validator (default) | vectoradd
//   FAKE_NDEV        visible device count (default 1; 0 = no device)
//   FAKE_MODE        ok (default) | cudaerr | wrong
//   FAKE_FAIL_DEVICE device the failure mode applies to (-1 = all, default)
#include <cstdlib>
#include <cstring>
#include <unistd.h>

#include "validator_core.h"

namespace {

int env_int(const char *name, int dflt) {
    const char *v = std::getenv(name);
    return (v && *v) ? std::atoi(v) : dflt;
}

bool failing(int dev) {
    const int f = env_int("FAKE_FAIL_DEVICE", -1);
    return f < 0 || f == dev;
}

const char *mode() {
    const char *m = std::getenv("FAKE_MODE");
    return (m && *m) ? m : "ok";
}

int fake_device_count(int *n) {
    *n = env_int("FAKE_NDEV", 1);
    return *n > 0 ? B200V_OK : B200V_E_NODEVICE;
}

int fake_device_info(int, b200v_info *out) {
    std::memset(out, 0, sizeof *out);
    std::strcpy(out->name, "FAKE B200");
    out->sm_count = 148;
    out->mem_clock_khz = 3996000;
    out->mem_bus_bits = 8192;
    return B200V_OK;
}

int fake_vectoradd(int dev, const float *a, const float *b, float *c, size_t n, b200v_timing *t) {
    if (dev < 0 || dev >= env_int("FAKE_NDEV", 1)) return B200V_E_BADARG;
    if (!std::strcmp(mode(), "cudaerr") && failing(dev)) return 700;   // cudaErrorIllegalAddress
    for (size_t i = 0; i < n; ++i) c[i] = a[i] + b[i] + 0.0f;
    if (!std::strcmp(mode(), "wrong") && failing(dev) && n > 17) c[17] = -1.0f;
    if (t) {
        std::memset(t, 0, sizeof *t);
        t->grid = (int)((n + 2047) / 2048);
        t->block = 256;
        t->chunks = 1;
        t->kernel_launches = 1;
    }
    return B200V_OK;
}

int fake_stream(int, int op, size_t n, float, uint64_t, int, int iters, b200v_stream_result *r) {
    std::memset(r, 0, sizeof *r);
    r->best_ms = r->median_ms = r->mean_ms = 1.0f;
    r->bytes_per_launch = n * (op >= B200V_OP_ADD ? 12 : 8);
    r->best_gbs = r->median_gbs = (double)r->bytes_per_launch / 1e6;
    r->digest = 0x1234;
    r->kernel_launches = iters;
    if (!std::strcmp(mode(), "wrong")) { r->mismatches = 1; return B200V_E_MISMATCH; }
    return B200V_OK;
}

int fake_allreduce(int ndev, const int *, size_t lo, size_t hi, int, int, b200v_ar_row *rows, int max_rows,
                   int *nrows) {
    if (!std::strcmp(mode(), "cudaerr")) return B200V_E_NCCL;
    if (!std::strcmp(mode(), "hang")) for (;;) ::pause();     // a collective that never completes
    int k = 0;
    for (size_t b = lo; b <= hi && k < max_rows; b *= 2) {
        std::memset(&rows[k], 0, sizeof rows[k]);
        rows[k].bytes = b;
        rows[k].best_ms = rows[k].median_ms = 1.0f;
        rows[k].algbw_gbs = 400.0;
        rows[k].busbw_gbs = 400.0 * 2 * (ndev - 1) / ndev;
        if (!std::strcmp(mode(), "wrong")) rows[k].mismatches = 3;
        ++k;
    }
    *nrows = k;
    return B200V_OK;
}

const char *fake_strerror(int code) {
    switch (code) {
    case 0: return "ok";
    case B200V_E_NODEVICE: return "no CUDA device";
    case B200V_E_NCCL: return "NCCL error";
    case 700: return "an illegal memory access was encountered";
    default: return "fake error";
    }
}

int fake_shutdown(void) { return 0; }

}  // namespace

int main(int argc, char **argv) {
    b200v_host::Api api{fake_device_count, fake_device_info, fake_vectoradd, fake_stream,
                        fake_allreduce,    fake_strerror,    fake_shutdown,    fake_allreduce};
    const char *entry = std::getenv("FAKE_ENTRY");
    if (entry && !std::strcmp(entry, "vectoradd")) return b200v_host::vectoradd_main(api, argc, argv);
    return b200v_host::validator_main(api, argc, argv);
}
