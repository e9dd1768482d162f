// nvidia-validator — drop-in for the operator-validator entrypoint on the GPU workload path
// (`command: ["sh","-c"], args: ["nvidia-validator"]` with COMPONENT=cuda|driver|toolkit).
#include "validator_core.h"

int main(int argc, char **argv) {
    return b200v_host::validator_main(b200v_host::real_api(), argc, argv);
}
