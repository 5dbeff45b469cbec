// K3 + K4 kernels of the two-launch tick (blfik_split.cuh)
#define BLFIK_INSTANTIATE
#include "../../include/blfik.h"
#include "blfik_inst.h"
namespace blfik {
BLFIK_SOLVE_INSTANCES(BLFIK_DEF_SOLVE)
}
