// real_api.cpp — binds the Api table to libb200validate.so.  Linked into the product executables
// only; the host-logic tests link their own scripted table instead.
#include "validator_core.h"

namespace b200v_host {

Api real_api() {
    Api a;
    a.device_count = b200v_device_count;
    a.device_info = b200v_device_info;
    a.vectoradd = b200v_vectoradd;
    a.stream = b200v_stream;
    a.allreduce_probe = b200v_allreduce_probe;
    a.strerror = b200v_strerror;
    a.shutdown = b200v_shutdown;
    a.nvls_allreduce_probe = b200v_nvls_allreduce_probe;
    return a;
}

}  // namespace b200v_host
