// factored-form kernels for wide models (blfik_eq.cuh)
#define BLFIK_INSTANTIATE
#include "../../include/blfik.h"
#include "blfik_inst.h"
namespace blfik {
BLFIK_EQ_INSTANCES(BLFIK_DEF_EQ)
}
