// vectorAdd — drop-in for the cuda-sample binary the workload pod runs (`args: ["vectorAdd"]`):
// same stdout lines, same exit code, sm_100a kernel underneath.
#include "validator_core.h"

int main(int argc, char **argv) {
    return b200v_host::vectoradd_main(b200v_host::real_api(), argc, argv);
}
