// fused register-resident tick kernels, up to 32 velocity variables (iCub-26)
#define BLFIK_INSTANTIATE
#include "../../include/blfik.h"
#include "blfik_inst.h"
namespace blfik {
BLFIK_FAST_INSTANCES_32(BLFIK_DEF_FAST)
}
