// blfik_inst.h -- the tick-kernel instantiations are compiled in translation units of their own (blfik_inst_*.cu), so that
// the library builds in parallel; blfik_capi.cu only takes their addresses.
#pragma once
#include "blfik_fast.cuh"
#include "blfik_split.cuh"
#include "blfik_eq.cuh"

#define BLFIK_FAST_INSTANCES_29(X) X(29, 16, false) X(29, 18, false) X(29, 30, false) X(29, 32, false) X(29, 32, true)
#define BLFIK_FAST_INSTANCES_32(X) X(32, 16, false) X(32, 20, false) X(32, 32, false) X(32, 32, true)
#define BLFIK_EQ_INSTANCES(X) X(0) X(16) X(24)
#define BLFIK_SOLVE_INSTANCES(X) X(29, 16) X(29, 18) X(29, 30) X(29, 32) X(32, 16) X(32, 20) X(32, 32)

namespace blfik {
#define BLFIK_DECL_FAST(NV, KP, AM) extern template __global__ void blfik_tick_fast_kernel<NV, KP, AM>(const FastParams);
#define BLFIK_DECL_SOLVE(NV, KP) extern template __global__ void blfik_solve_kernel<NV, KP>(const SolveParams);
#define BLFIK_DECL_EQ(KP) extern template __global__ void blfik_tick_eq_kernel<KP>(const TickParams);
#ifndef BLFIK_INSTANTIATE
BLFIK_EQ_INSTANCES(BLFIK_DECL_EQ)
BLFIK_FAST_INSTANCES_29(BLFIK_DECL_FAST)
BLFIK_FAST_INSTANCES_32(BLFIK_DECL_FAST)
BLFIK_SOLVE_INSTANCES(BLFIK_DECL_SOLVE)
#endif
#define BLFIK_DEF_FAST(NV, KP, AM) template __global__ void blfik_tick_fast_kernel<NV, KP, AM>(const FastParams);
#define BLFIK_DEF_EQ(KP) template __global__ void blfik_tick_eq_kernel<KP>(const TickParams);
#define BLFIK_DEF_SOLVE(NV, KP) template __global__ void blfik_solve_kernel<NV, KP>(const SolveParams);
}  // namespace blfik
