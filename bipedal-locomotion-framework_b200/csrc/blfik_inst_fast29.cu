// fused register-resident tick kernels, ergoCub-sized (29 velocity variables)
#define BLFIK_INSTANTIATE
#include "../../include/blfik.h"
#include "blfik_inst.h"
namespace blfik {
BLFIK_FAST_INSTANCES_29(BLFIK_DEF_FAST)
}
