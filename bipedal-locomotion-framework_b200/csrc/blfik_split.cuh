// blfik_split.cuh -- the tick as TWO launches, each with the mapping its half of the work wants.
//
// The fused register-resident kernel (blfik_fast.cuh) runs K1 / K2 one warp per robot: 24 of 32 lanes walk the tree, one lane
// per task evaluates a control law, and every lane pays the type dispatch of the task loop -- 40 % of its instructions, at
// the occupancy the factorisation dictates (ncu, profiles/r1_v6_*).  Here:
//
//   blfik_assemble_kernel   K1 + K2, ONE THREAD PER ROBOT (blfik_assemble.cuh).  Control flow depends on the model and the task
//                           list only, so it is uniform across a warp and its cost is shared by 32 robots; no shuffles, no
//                           idle lanes.  Leaves one "QP record" per robot in device memory (entry-major, so the stores of a
//                           warp coalesce): the dense task rows B (K x nv), their right-hand sides, the diagonal cost, its
//                           gradient and the joint box.
//   blfik_solve_kernel      K3 + K4, one warp per robot, task-agnostic linear algebra: G = E + B D^-1 B' (DMMA), Cholesky with
//                           the inverse factor riding along, U = D^-1 B' L^-T (DMMA), N = D^-1 - U U' (DMMA), the certified
//                           active-set / ADMM iterations of blfik_fast.cuh, ForwardEuler.  One-warp blocks, one robot each,
//                           no persistent loop: the hardware block scheduler is the dynamic hand-out.
//
// Reference computation replaced: the task update() calls + QPInverseKinematics.cpp:442-491 (kernel A), OsqpEigen update +
// solve, QPInverseKinematics.cpp:61-146, 494-528, and ForwardEuler.h:72-124 (kernel B).
#pragma once
#include "blfik_fast.cuh"

namespace blfik {

// ------------------------------------------------------------------------------------------------ the QP record
// One robot's QP, robot-major in device memory (kernel A writes it, kernel B loads it as is into shared memory), doubles:
//   Bt [NV][KP]   the dense task rows BY COLUMNS (column c = the KP entries of velocity variable c), rows beyond K are zero
//   cvec [KP]     right-hand sides of the rows
//   vec [32][4]   per velocity variable: diagonal cost, its gradient term (sum w b), lower and upper bound
template <int NV, int KP>
struct QpRec {
    static_assert(KP % 2 == 0, "rows come in pairs (16-byte alignment of the parts)");
    static constexpr int cvec = NV * KP;
    static constexpr int vec = cvec + KP;
    static constexpr int LEN = vec + 4 * 32;
};

struct AsmParams {
    TickParams T;
    double* qp;                 // [N][rec_len]
    int rec_len, off_c, off_v, K;
    const double* model_pack;   // pack_model(): link records [nl][20], parent / dof tables, ancestor masks, subtree masses
    int pack_len, sh_jump, sh_lod, sh_anc, sh_msub;
    const int* events;          // depth-first walk of the tree: (kind << 16) | (depth << 8) | link, kind 0 = visit, 1 = close
    int nevents;
    int tick;                   // index of this tick inside a multi-tick launch (0: first)
    double tnow;
    int* work_counter;          // kernel B's hand-out counter: zeroed here, one launch ahead
    int* slow_out_n;            // length of the slow-robot list the solve kernel of this tick writes (SolveParams): zeroed here; may be null
    int* bin_cnt;               // [64] robots per hardness bin (zeroed by the host before the launch), null: no binning
    int* bin_list;              // [64][bin_cap] robot indices per bin
    long long bin_cap;
    int jw_task;                // a JointTrackingTask whose weight comes with the launch (weight schedule), -1: none
    double jw[32];
};

constexpr int kAsmMaxDense = 10;   // dense tasks the assemble kernel keeps frame origins for
constexpr int kAsmMaxDepth = 12;   // levels of its link stack (tree depth + 1)

// ------------------------------------------------------------------------------------------------ kernel B
struct SolveParams {
    TickParams T;
    const double* qp;           // [N][QpRec::LEN]
    int* work_counter;          // next robot to hand out, relative to T.first
    int prefetch_ahead;         // a block asks the L2 for the record of the robot this many places further on (the blocks in flight)
    double evec[32];            // E_r of the dense rows: 0 (equality) or 1 / weight (cost); 1 on the padding rows
    int nv, K;
    int max_as_iters, max_polish_iters, max_total_iters, pdas_first;
    int tick;
    int admm_status;            // status of a robot whose ADMM iterates settled but whose active set could not be certified
    // Longest first: the robots that needed many iterations in the previous tick of the same batch (a closed loop: they are
    // mostly the same ones, e.g. those whose targets are out of reach) are handed out before the others, so that they run
    // beside the bulk instead of after it.  slow_in / slow_in_n: the list written by the previous tick; slow_out / slow_out_n:
    // the list this tick writes; slow_stamp_in[r] == stamp marks a robot of slow_in (the linear hand-out skips it);
    // slow_stamp_out[r] = stamp + 1 marks one of slow_out.  All null / zero when the launch does not cover the whole batch.
    const int* slow_in;
    const int* slow_in_n;
    int* slow_out;
    int* slow_out_n;
    const int* slow_stamp_in;
    int* slow_stamp_out;
    int stamp, slow_threshold;
    // hand-out order of the robots that are not in slow_in: the hardness bins the assemble kernel of this tick filled, from the
    // top bin down (null: by index)
    const int* bin_cnt;
    const int* bin_list;
    long long bin_cap;
    // Consecutive ticks a robot has ended without a solution inside the iteration budget (closed loops: mostly robots whose
    // constraints have no feasible point, tick after tick).  Such a robot gets a short budget, and the full one again every
    // sixteenth tick; a tick that solves it clears the count.  Null: every robot gets the full budget.
    int* unsolved_run;
    int short_budget;
};

// per-warp shared memory of kernel B, in doubles, all compile-time
template <int NV, int KP>
struct SolveLayout {
    using Rec = QpRec<NV, KP>;
    static constexpr int LDB = FastDims<NV, KP>::LDB;
    static constexpr int LDG = FastDims<NV, KP>::LDG;
    static constexpr int cmax(int a, int b) { return a > b ? a : b; }
    // region 1: the record image (Bm, cvec, vectors), later the rows of N
    static constexpr int R1 = 0;
    static constexpr int R1_LEN = (cmax(Rec::LEN, NV * LDB) + 1) & ~1;
    // region 2: G by columns -> L_G' / inverse factor (rows of KP + the rotating elimination's overshoot) -> U rows -> the
    // active-set matrix
    static constexpr int R2 = R1 + R1_LEN;
    static constexpr int NJ = NV - 6;
    static constexpr int MA_CAP = NJ < 32 ? NJ : 32;
    static constexpr int R2_LEN = (cmax(cmax(KP * LDG + KP, cmax(NV, KP) * LDG), MA_CAP * (MA_CAP | 1)) + 1) & ~1;
    static constexpr int dinv = R2 + R2_LEN;     // [LDB -> 34]
    static constexpr int rvec = dinv + 34;
    static constexpr int wv = rvec + 34;         // [LDG]
    static constexpr int dA = wv + ((LDG + 1) & ~1);
    static constexpr int muA = dA + 32;
    static constexpr int idxA = muA + 32;        // 32 ints
    static constexpr int TOTAL = idxA + 16;
};

// Factorisation of one robot's KKT system.  In: the record image (Bm, cvec) in shared memory, 1 / D.  Out: rows of N in Ns
// (over the record image), x_b.  Differences from fast_setup (blfik_fast.cuh):
//  * the Cholesky carries the inverse factor along: lane k, whose row of G is used up after step k, goes on as the forward
//    substitution of e_k, in the same instruction stream as the rows still being eliminated -- L^-1 costs nothing extra;
//  * U = (D^-1 B') L^-T is then a product on the FP64 tensor path (DMMA) instead of a second 29-lane substitution pass.
template <int NV, int KP>
__device__ __forceinline__ bool solve_setup(double* base, const double* evec, int nv, int lane, double& xb)
{
    using SL = SolveLayout<NV, KP>;
    constexpr int LDB = SL::LDB, LDG = SL::LDG;
    const double* Bt = base + SL::R1;   // Bt[c * KP + r]: entry (row r, velocity variable c) of the dense rows
    const double* cvec = Bt + QpRec<NV, KP>::cvec;
    double* Ns = base + SL::R1;
    double* Gs = base + SL::R2;
    double* LGt = base + SL::R2;
    double* Us = base + SL::R2;
    const double* dinv = base + SL::dinv;
    double* wv = base + SL::wv;
    const bool mine = lane < nv;
    const int gr = lane >> 2, tc = lane & 3;
    bool ok = true;
    constexpr int RT = (KP + 7) / 8;   // row tiles of B
    // ---- G = E + (B D^-1) B' as 8x8 DMMA tiles, lower tiles computed, off-diagonal ones mirrored (the riders below read
    // whole rows)
    {
        constexpr int KS = (NV + 3) / 4;
        double bf[RT][KS], dvv[KS];
#pragma unroll
        for (int s = 0; s < KS; s++) {
            const int col = 4 * s + tc;
            dvv[s] = (col < NV) ? dinv[col] : 0.0;
#pragma unroll
            for (int i = 0; i < RT; i++) {
                const int row = 8 * i + gr;
                bf[i][s] = (row < KP && col < NV) ? Bt[col * KP + row] : 0.0;
            }
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < RT; i++) {
            double ai[KS];
#pragma unroll
            for (int s = 0; s < KS; s++) ai[s] = bf[i][s] * dvv[s];
#pragma unroll
            for (int j = 0; j <= i; j++) {
                double c0 = 0.0, c1 = 0.0;
#pragma unroll
                for (int s = 0; s < KS; s++) dmma884(c0, c1, ai[s], bf[j][s]);
                const int row = 8 * i + gr, col = 8 * j + 2 * tc;
                if (row < KP) {
                    const double er = evec[row];
                    if (col < KP) {
                        const double v = c0 + (row == col ? er : 0.0);
                        Gs[col * LDG + row] = v;
                        if (j < i) Gs[row * LDG + col] = v;
                    }
                    if (col + 1 < KP) {
                        const double v = c1 + (row == col + 1 ? er : 0.0);
                        Gs[(col + 1) * LDG + row] = v;
                        if (j < i) Gs[row * LDG + col + 1] = v;
                    }
                }
            }
        }
    }
    __syncwarp();
    // ---- Cholesky G = L L' with riders.  Lane r < KP starts as row r of G.  At step c every lane scales its leading element
    // by 1 / l_cc: for a row r > c that is l_rc, for the right-hand side (lane KP, when there is a spare lane) it is
    // (L^-1 b)_c, and for a lane r < c, whose row was used up at step r and which has carried e_r's forward substitution ever
    // since, it is (L^-1)_{c r}.  Lane c itself turns into the rider of e_c: its row equals l_cc x (column c of L) at this
    // point, so eliminating with l_cc + 1 / l_cc leaves -(column c) / l_cc, the right-hand side e_c has after its own pivot.
    // Row c of the output (LGt[c][*]): (L^-1)_{c, 0..c} then L_{c+1.., c}.
    unsigned clamped = 0u;
    double floor_;
    {
        double gA[KP], gB[KP];
        const int myr = lane < KP ? lane : 0;
#pragma unroll
        for (int j = 0; j < KP; j++) gA[j] = Gs[j * LDG + myr];
        if (KP < 32 && lane == KP) {
#pragma unroll
            for (int j = 0; j < KP; j++) gA[j] = cvec[j];
        }
        // 1e-11: an exactly dependent row leaves a pivot of rounding noise, about 1e-14 of the largest diagonal entry
        floor_ = 1e-11 * warp_max(lane < KP ? Gs[myr * LDG + myr] : 0.0) + 1e-300;
        __syncwarp();
#pragma unroll 1
        for (int k = 0; k < KP; k += 2) {
            {
                double piv = __shfl_sync(kFull, gA[0], k);
                if (!(piv > floor_)) {
                    if (!(piv == piv)) ok = false;
                    clamped |= 1u << k;
                    piv = floor_;
                }
                const double rinv = rsqrt(piv);
                const double f = gA[0] * rinv;
                if (lane < KP) LGt[k * LDG + lane] = (lane == k) ? rinv : f;
                if (KP < 32 && lane == KP) wv[k] = f;
                __syncwarp();
                rot_update<KP, 0>(gA, gB, lane == k ? f + rinv : f, LGt + k * LDG + k);
            }
            {
                const int k1 = k + 1;
                double piv = __shfl_sync(kFull, gB[0], k1);
                if (!(piv > floor_)) {
                    if (!(piv == piv)) ok = false;
                    clamped |= 1u << k1;
                    piv = floor_;
                }
                const double rinv = rsqrt(piv);
                const double f = gB[0] * rinv;
                if (lane < KP) LGt[k1 * LDG + lane] = (lane == k1) ? rinv : f;
                if (KP < 32 && lane == KP) wv[k1] = f;
                __syncwarp();
                rot_update<KP, 1>(gB, gA, lane == k1 ? f + rinv : f, LGt + k1 * LDG + k1);
            }
        }
    }
    __syncwarp();
    if constexpr (KP == 32) {   // no spare lane carried b: w = L^-1 b from the inverse factor
        double a = 0.0;
        for (int r = 0; r < KP; r++) a = fma(r <= lane ? LGt[lane * LDG + r] : 0.0, cvec[r], a);
        wv[lane] = a;
        __syncwarp();
    }
    // A clamped pivot is an equality row that depends on the previous ones.  What is left of its right-hand side must vanish
    // -- consistent duplicates are harmless, the clamp acts like OSQP's regularised polish -- else the rows contradict each
    // other, there is no feasible point and the reference's advance() fails (QPInverseKinematics.cpp:518-523): status 2.
    if (clamped) {
        constexpr double kInconsistent = 1e-7;   // [m/s, rad/s]
        if (lane < KP && ((clamped >> lane) & 1u) && fabs(wv[lane]) * sqrt(floor_) > kInconsistent) ok = false;
    }
    // ---- U = (D^-1 B') L^-T on the tensor path: tile (i, j) = sum_s A(i, s) Linv'(s, j); Linv is lower triangular, so the
    // k-steps beyond the diagonal block are skipped.  All tiles are held in registers until the last read of LGt, whose
    // memory the rows of U then take.
    constexpr int MT = (NV + 7) / 8;
    {
        constexpr int KSU = (KP + 3) / 4;
        double uc[MT][RT][2];
        double lf[RT][KSU];
#pragma unroll
        for (int j = 0; j < RT; j++)
#pragma unroll
            for (int s = 0; s < KSU; s++) {
                const int c = 8 * j + gr, r = 4 * s + tc;   // (L^-1)_{c r}
                lf[j][s] = (c < KP && r <= c && 4 * s <= 8 * j + 7) ? LGt[c * LDG + r] : 0.0;
            }
#pragma unroll
        for (int i = 0; i < MT; i++) {
            double af[KSU];
            const int row = 8 * i + gr;
            const double di = row < NV ? dinv[row] : 0.0;
#pragma unroll
            for (int s = 0; s < KSU; s++) {
                const int r = 4 * s + tc;
                af[s] = (row < NV && r < KP) ? Bt[row * KP + r] * di : 0.0;
            }
#pragma unroll
            for (int j = 0; j < RT; j++) {
                double c0 = 0.0, c1 = 0.0;
#pragma unroll
                for (int s = 0; s < KSU; s++)
                    if (4 * s <= 8 * j + 7) dmma884(c0, c1, af[s], lf[j][s]);
                uc[i][j][0] = c0, uc[i][j][1] = c1;
            }
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
            for (int j = 0; j < RT; j++) {
                const int row = 8 * i + gr, col = 8 * j + 2 * tc;
                if (row < NV && col < LDG) *reinterpret_cast<double2*>(Us + row * LDG + col) = make_double2(uc[i][j][0], uc[i][j][1]);
            }
    }
    __syncwarp();
    // ---- N = D^-1 - U U' (DMMA, both fragments are the same registers), x_b = U w
    {
        constexpr int KS = (KP + 3) / 4;
        double uf[MT][KS];
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
            for (int s = 0; s < KS; s++) {
                const int row = 8 * i + gr, col = 4 * s + tc;
                uf[i][s] = (row < NV && col < KP) ? Us[row * LDG + col] : 0.0;
            }
#pragma unroll
        for (int i = 0; i < MT; i++) {
#pragma unroll
            for (int j = 0; j < MT; j++) {
                double c0 = 0.0, c1 = 0.0;
#pragma unroll
                for (int s = 0; s < KS; s++) dmma884(c0, c1, uf[i][s], uf[j][s]);
                const int row = 8 * i + gr, col = 8 * j + 2 * tc;
                if (row < NV && col < LDB) {
                    const double dr = dinv[row];
                    *reinterpret_cast<double2*>(Ns + row * LDB + col) = make_double2((row == col ? dr : 0.0) - c0, (row == col + 1 ? dr : 0.0) - c1);
                }
            }
        }
        const double2* myu = reinterpret_cast<const double2*>(Us + (lane < NV ? lane : 0) * LDG);
        const double2* w2 = reinterpret_cast<const double2*>(wv);
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int rr = 0; rr < KP / 2; rr++) {
            const double2 t2 = w2[rr], u2 = myu[rr];
            a0 = fma(u2.x, t2.x, a0);
            a1 = fma(u2.y, t2.y, a1);
        }
        xb = mine ? a0 + a1 : 0.0;
    }
    __syncwarp();
    return __all_sync(kFull, ok);
}

// Do the equality rows of the record hold at x (one variable per lane)?  Rare (active sets certified late), hence out of line:
// the solve kernel's instruction footprint is what its warps, each at a different place, share the instruction cache for.
static __device__ __noinline__ bool solve_rows_hold(const double* rec, int cvec_off, int nv, int kp, const double* evec, double x, int lane)
{
    const int row = lane < kp ? lane : 0;
    double acc = 0.0;
    for (int c = 0; c < nv; c++) acc = fma(__ldg(rec + c * kp + row), __shfl_sync(kFull, x, c), acc);
    const double cr = __ldg(rec + cvec_off + row);
    const bool eqrow = lane < kp && evec[row] == 0.0;
    return !__any_sync(kFull, eqrow && fabs(acc - cr) > 1e-8 * (1.0 + fabs(cr)));
}

template <int NV, int KP>
struct SolveDims {
    // Warps (robots) per block and resident blocks per SM.  The small instantiations: 16 warps at 128 registers, 12.5 KB of
    // shared memory each; with more dense rows a robot needs up to 19 KB and the 128-register build spills: 10 warps.
    // One-warp blocks: a robot that needs the ADMM fallback holds back nobody, and the slot of a finished robot is refilled
    // at once (with four-warp blocks 12.3 of 16 warps were resident on average).
    static constexpr bool SMALL = FastDims<NV, KP>::WARPS > 12;
#ifndef BLFIK_SOLVE_WPB
#define BLFIK_SOLVE_WPB 1
#endif
#ifndef BLFIK_SOLVE_BIG_BLOCKS
#define BLFIK_SOLVE_BIG_BLOCKS 11
#endif
    static constexpr int WPB = SMALL ? BLFIK_SOLVE_WPB : 1;
    static constexpr int BLOCKS = SMALL ? 16 / WPB : BLFIK_SOLVE_BIG_BLOCKS;
    static constexpr bool NROW_SMEM = FastDims<NV, KP>::NROW_SMEM;
};

// K3 + K4: one warp per robot, one robot per warp.
template <int NV, int KP>
__global__ void __launch_bounds__(SolveDims<NV, KP>::WPB * 32, SolveDims<NV, KP>::BLOCKS) blfik_solve_kernel(const SolveParams SP)
{
    using SL = SolveLayout<NV, KP>;
    using Rec = QpRec<NV, KP>;
    constexpr int LDB = SL::LDB;
    constexpr bool NROW_SMEM = SolveDims<NV, KP>::NROW_SMEM;
    extern __shared__ double smem[];
    __shared__ double s_evec[32];
    const TickParams& P = SP.T;
    const Settings& st = P.st;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    if (threadIdx.x < 32) s_evec[threadIdx.x] = SP.evec[threadIdx.x];
    __syncthreads();
    double* base = smem + (size_t)wib * SL::TOTAL;
    // Persistent warps: each takes its next robot from a device counter (zeroed by the assemble kernel of the same tick), so
    // a slot is never idle between two robots (with one block per robot 13.6 of 16 warps were resident on average).
  const int n_slow = SP.slow_in_n ? *SP.slow_in_n : 0;
  // s_bins[k]: how many robots the bins 63 ... 63 - k hold together (the top bin first), per warp of the block
  __shared__ int s_bins_all[SolveDims<NV, KP>::WPB][66];
  int* s_bins = s_bins_all[wib];
  if (lane == 0) s_bins[64] = 0;   // [64]: the bins say something (more than one of them is occupied)
  if (SP.bin_cnt) {
      int p_hi = SP.bin_cnt[63 - lane], p_lo = SP.bin_cnt[31 - lane];
      int cmax = p_hi > p_lo ? p_hi : p_lo;
      for (int o = 16; o > 0; o >>= 1) {
          const int other = __shfl_xor_sync(kFull, cmax, o);
          cmax = other > cmax ? other : cmax;
      }
      for (int o = 1; o < 32; o <<= 1) {
          const int a = __shfl_up_sync(kFull, p_hi, o), b = __shfl_up_sync(kFull, p_lo, o);
          if (lane >= o) p_hi += a, p_lo += b;
      }
      const int t_hi = __shfl_sync(kFull, p_hi, 31);
      s_bins[lane] = p_hi, s_bins[32 + lane] = p_lo + t_hi;
      if (lane == 31) s_bins[64] = cmax < p_lo + t_hi;   // all robots in one bin: the order by index is as good, and cheaper
  }
  __syncwarp();
  for (;;) {
    long long robot = 0;
    long long ahead = -1;   // the robot whose record the L2 is asked for: the one handed out prefetch_ahead items from now
    if (lane == 0) {
        const int wi = atomicAdd(SP.work_counter, 1);
        if (wi < n_slow) {
            robot = SP.slow_in[wi];
            if (SP.slow_stamp_in[robot] != SP.stamp) robot = -1;   // a stale entry: the ordinary hand-out takes this robot
        } else if (s_bins[64]) {
            const int v = wi - n_slow;
            if (v >= s_bins[63]) {
                robot = P.N;
            } else {
                int k = 0;   // the bin that holds item v: binary search over the running counts
                for (int step = 32; step > 0; step >>= 1)
                    if (v >= s_bins[k + step - 1]) k += step;
                robot = SP.bin_list[(size_t)(63 - k) * SP.bin_cap + (v - (k ? s_bins[k - 1] : 0))];
                if (n_slow > 0 && SP.slow_stamp_in[robot] == SP.stamp) robot = -1;   // handed out from the list
                const int va = v + SP.prefetch_ahead;
                if (va < s_bins[63]) {
                    int ka = k;   // not below the bin of v
                    while (va >= s_bins[ka]) ka++;
                    ahead = SP.bin_list[(size_t)(63 - ka) * SP.bin_cap + (va - (ka ? s_bins[ka - 1] : 0))];
                }
            }
        } else {
            robot = P.first + (wi - n_slow);
            if (robot + SP.prefetch_ahead < P.N) ahead = robot + SP.prefetch_ahead;
            if (robot < P.N && n_slow > 0 && SP.slow_stamp_in[robot] == SP.stamp) robot = -1;   // handed out from the list
        }
    }
    robot = __shfl_sync(kFull, robot, 0);
    ahead = __shfl_sync(kFull, ahead, 0);
    if (robot < 0) continue;
    if (robot >= P.N) break;
    int worst = 0, iters_prev = 0;
    int budget = SP.max_total_iters;
    int run = 0;
    if (SP.unsolved_run) {
        run = SP.unsolved_run[robot];
        if (run > 0 && (run & 15) != 0 && SP.short_budget < budget) budget = SP.short_budget;
    }
    if (SP.tick > 0) {
        worst = P.status[robot], iters_prev = P.iters[robot];
        if (worst == 2) continue;   // a failed tick ended this robot's rollout: nu stays zero, the state where the failure happened
    }
    double* Ns = base + SL::R1;
    double* MA = base + SL::R2;
    double* dinv = base + SL::dinv;
    double* rvec = base + SL::rvec;
    double* dA = base + SL::dA;
    double* muA = base + SL::muA;
    int* idxA = reinterpret_cast<int*>(base + SL::idxA);
    const int n = P.m.n, nv = SP.nv;
    const bool mine = lane < nv;
    const bool joint = mine && lane >= 6;
    const double* rec = SP.qp + (size_t)robot * Rec::LEN;
    // Ask the L2 for the record of the robot that some warp takes about one robot-time from now: its loads then hit the L2
    // instead of reading DRAM cold (which robot this very warp gets next is not known: the hand-out is dynamic).
    if (ahead >= 0) {
        const char* nxt = reinterpret_cast<const char*>(SP.qp + (size_t)ahead * Rec::LEN);
        for (int l = lane * 128; l < Rec::LEN * 8; l += 32 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(nxt + l));
    }

    const bool warm_flag = (P.flags & BLFIK_TICK_WARM_START) != 0;
    double x = 0.0, z = 0.0, y = 0.0;
    if (warm_flag && mine) {
        x = P.wx[(size_t)nv * robot + lane];
        if (joint) z = P.wz[(size_t)n * robot + lane - 6], y = P.wy[(size_t)n * robot + lane - 6];
    }
    // Two attempts, as in the fused kernel: attempt 0 factorises WITHOUT the ADMM penalty (rho = 0), so the proximal steps
    // contract by sigma / (sigma + lambda) ~ 1e-2 and each active-set guess is solved in 2-3 mat-vecs; the active set is refined
    // by primal-dual steps and accepted only with the KKT certificate.  Attempt 1 (the fallback): OSQP-style ADMM with rho on
    // the box rows predicts the active set, the same certified polish finishes.  The record is loaded again for it (the rows of
    // N have taken its place).
    bool done = false;
    int iters = 0, status = 1;
    const int first_attempt = SP.pdas_first ? 0 : 1;
    bool any_box = false;
    for (int attempt = first_attempt; attempt < 2 && !done && status != 2; attempt++) {
        const bool with_rho = attempt == 1;
        {   // the record image: contiguous 16-byte loads
            const double2* src = reinterpret_cast<const double2*>(rec);
            double2* dst = reinterpret_cast<double2*>(base + SL::R1);
#pragma unroll 6
            for (int e = lane; e < Rec::LEN / 2; e += 32) dst[e] = __ldg(src + e);
        }
        __syncwarp();
        const double2 v01 = *reinterpret_cast<const double2*>(base + SL::R1 + Rec::vec + 4 * lane);
        const double2 v23 = *reinterpret_cast<const double2*>(base + SL::R1 + Rec::vec + 4 * lane + 2);
        const double dh = v01.x, gd = v01.y, lo = v23.x, hi = v23.y;
        // per-variable ADMM step: 0 (no bound), rho (box), 1e3 rho (pinned: lo == hi)
        const bool bounded = joint && (lo > -kInfHalf || hi < kInfHalf);
        const bool pinned = bounded && lo == hi;
        const double rho_i = (bounded && with_rho) ? (pinned ? 1e3 * st.rho : st.rho) : 0.0;
        const double D_i = mine ? st.sigma + rho_i + dh : 1.0;
        const double Q_i = mine ? st.sigma + rho_i : 0.0;
        const double dinv_i = 1.0 / D_i;
        const double inv_rho = rho_i > 0.0 ? 1.0 / rho_i : 0.0;
        dinv[lane] = mine ? dinv_i : 0.0;
        if (lane < 2) dinv[32 + lane] = 0.0, rvec[32 + lane] = 0.0;
        any_box = __any_sync(kFull, bounded);
        const bool infeasible = __any_sync(kFull, bounded && lo > hi);
        const double gmax = warp_max(fabs(gd));
        __syncwarp();
        double nrow[NV];   // unused (and eliminated) when the row is streamed from shared memory
        const double* myrow = Ns + (lane < NV ? lane : 0) * LDB;
        double xb = 0.0;
        if (infeasible) {
            status = 2;
        } else {
            const bool fac_ok = solve_setup<NV, KP>(base, s_evec, nv, lane, xb);
            if (!fac_ok) status = 2;
            if constexpr (!NROW_SMEM) {
                const double2* myn = reinterpret_cast<const double2*>(myrow);
#pragma unroll
                for (int jj = 0; jj < NV / 2; jj++) {
                    const double2 t2 = myn[jj];
                    nrow[2 * jj] = t2.x, nrow[2 * jj + 1] = t2.y;
                }
                if (NV & 1) nrow[NV - 1] = myrow[NV - 1];
            }
        }
        if (status != 2) {
            // ============================================================ K3 iterations (blfik_fast.cuh)
            if (!warm_flag) x = 0.0, z = 0.0, y = 0.0;
            if (bounded) z = fmin(fmax(z, lo), hi);
            else z = 0.0, y = 0.0;
            const double tolx_rel = 1e-9;
            const double toly = 1e-8 * (1.0 + gmax);
            const double tol_fix = 1e-10 * (1.0 + gmax);
            double xp = x;
            double mu_mine = 0.0;
            int act = 0;   // 0 free, 1 lower, 2 upper, 3 pinned
            const int max_rounds = with_rho ? st.max_rounds : 1;
            for (int round = 0; round < max_rounds && !done && iters < budget; round++) {
                // ---- ADMM (skipped when there is no box row: the problem is a pure equality QP)
                double rp = 0.0, rd = 0.0;
                if (any_box && with_rho) {
                    for (int it = 0; it < st.admm_iters; it++) {
                        rvec[lane] = mine ? st.sigma * x + gd + (bounded ? rho_i * z - y : 0.0) : 0.0;
                        __syncwarp();
                        const double xt = NROW_SMEM ? fast_matvec_smem<NV>(myrow, rvec, xb) : fast_matvec<NV>(nrow, rvec, xb);
                        __syncwarp();
                        const double xn = st.alpha * xt + (1.0 - st.alpha) * x;
                        rd = st.sigma * fabs(xn - x);
                        rp = 0.0;
                        x = xn;
                        if (bounded) {
                            const double zr = st.alpha * xt + (1.0 - st.alpha) * z;
                            const double zn = fmin(fmax(zr + y * inv_rho, lo), hi);
                            y += rho_i * (zr - zn);
                            rd = fmax(rd, rho_i * fabs(zn - z));
                            z = zn;
                            rp = fabs(xn - zn);
                        }
                        iters++;
                    }
                    rp = warp_max(mine ? rp : 0.0), rd = warp_max(mine ? rd : 0.0);
                    if (!(rp == rp) || !(rd == rd)) {
                        status = 2;
                        break;
                    }
                }
                const bool admm_converged = any_box && with_rho && rp <= st.eps && rd <= st.eps * (1.0 + gmax);
                // ---- polish: OSQP's active-set rule, then primal-dual active-set refinement on the same N
                act = 0;
                if (bounded) {
                    if (pinned) act = 3;
                    else if (z - lo < -y) act = 1;
                    else if (hi - z < y) act = 2;
                }
                xp = x;
                for (int as_it = 0; as_it < SP.max_as_iters && !done && iters < budget; as_it++) {
                    const unsigned amask = __ballot_sync(kFull, act != 0);
                    const int a = __popc(amask);
                    const int rank = __popc(amask & ((1u << lane) - 1u));
                    const double bd = (act == 1) ? lo : hi;   // pinned: lo == hi
                    const int mld = a | 1;
                    bool small_ok = true;
                    if (a > SL::MA_CAP) break;               // cannot hold the active-set matrix: leave it to ADMM
                    if (a > 0) {
                        if (act) idxA[rank] = lane;
                        __syncwarp();
                        if (lane < a) {
                            const int ri = idxA[lane];
                            for (int uu = 0; uu < a; uu++) MA[lane * mld + uu] = Ns[ri * LDB + idxA[uu]];
                        }
                        __syncwarp();
                        // a joint whose N[j][j] is 1e-6 of its 1 / D_j is fixed by the equality rows and cannot be pinned to a
                        // bound: this active set has no solution, ADMM (and its verdict) takes over
                        small_ok = fast_small_inverse(MA, a, mld, lane, dinv, idxA);
                        small_ok = __all_sync(kFull, small_ok);
                    }
                    if (!small_ok) break;
                    if (act) xp = bd;
                    double dx = 0.0;
                    bool conv = false;
                    for (int it = 0; it < SP.max_polish_iters; it++) {
                        rvec[lane] = mine ? Q_i * xp + gd : 0.0;
                        __syncwarp();
                        double x1 = NROW_SMEM ? fast_matvec_smem<NV>(myrow, rvec, xb) : fast_matvec<NV>(nrow, rvec, xb);
                        __syncwarp();
                        if (a > 0) {
                            if (act) dA[rank] = x1 - bd;
                            __syncwarp();
                            if (lane < a) {
                                double mu = 0.0;
                                for (int uu = 0; uu < a; uu++) mu = fma(MA[lane * mld + uu], dA[uu], mu);
                                muA[lane] = mu;
                            }
                            __syncwarp();
                            if (mine)
                                for (int uu = 0; uu < a; uu++) x1 = fma(-Ns[lane * LDB + idxA[uu]], muA[uu], x1);
                        }
                        // measured before the pinned joints are set to their bounds (see blfik_fast.cuh)
                        dx = mine ? Q_i * fabs(x1 - xp) : 0.0;
                        if (a > 0 && act) x1 = bd, mu_mine = muA[rank];
                        xp = mine ? x1 : 0.0;
                        iters++;
                        if ((it & 1) == 1 || it + 1 == SP.max_polish_iters) {
                            const double dxm = warp_max(dx);
                            if (!(dxm == dxm)) break;
                            if (dxm <= tol_fix) {
                                conv = true;
                                break;
                            }
                        }
                    }
                    // KKT certificate: box feasibility of the free joints, multiplier signs of the pinned ones
                    bool viol_lo = false, viol_up = false, bad = false;
                    if (bounded && act == 0) {
                        const double tolx = tolx_rel * (1.0 + fabs(xp));
                        viol_up = xp > hi + tolx, viol_lo = xp < lo - tolx;
                    }
                    if (act == 1 && mu_mine > toly) bad = true;
                    if (act == 2 && mu_mine < -toly) bad = true;
                    if (!(xp == xp)) bad = true, conv = false;
                    const bool any_change = __any_sync(kFull, viol_lo || viol_up || bad);
                    // An active set found late pins joints that leave the equality rows nearly dependent: the factorisation
                    // then satisfies them to 1e-8 ... 1e-7 only and the point is off the optimum by 1e-5 and more.  Such a set
                    // gets the rows checked (the record is read again: its image has made room for N) and a point that misses
                    // them is handed out as inaccurate (status 1), not as certified.
                    bool rows_ok = true;
                    if (conv && !any_change && as_it >= 6) rows_ok = solve_rows_hold(rec, Rec::cvec, NV, KP, s_evec, xp, lane);
                    if (conv && !any_change && !rows_ok) {
                        done = true;
                        status = 1;
                        x = xp;
                        if (bounded) z = fmin(fmax(xp, lo), hi), y = act ? mu_mine : 0.0;
                    } else if (conv && !any_change) {
                        done = true;
                        status = 0;
                        x = xp;
                        if (bounded) {
                            z = fmin(fmax(xp, lo), hi);
                            y = act ? mu_mine : 0.0;
                        }
                    } else if (!conv && !any_change) {
                        break;   // right active set, fixed point not reached in the budget: continue with ADMM
                    } else {
                        if (bad) act = 0;
                        if (viol_lo) act = 1;
                        if (viol_up) act = 2;
                    }
                }
                if (!done && admm_converged) {
                    done = true;
                    status = SP.admm_status;
                }
                if (!any_box && !done) break;   // nothing ADMM could add
            }
            if (!done && status != 2) status = 1;
        }
        if (!any_box) break;   // a pure equality QP has no second attempt
    }
    if (status == 1 && (P.flags & BLFIK_TICK_STRICT)) status = 2;
    worst = status > worst ? status : worst;
    if (status == 2) x = 0.0;   // no velocity to report (the reference's advance() fails): hold still
    if (mine) {
        P.nu[(size_t)nv * robot + lane] = x;
        if (P.nu_host) P.nu_host[(size_t)nv * robot + lane] = x;
        P.wx[(size_t)nv * robot + lane] = x;
        if (joint) P.wz[(size_t)n * robot + lane - 6] = z, P.wy[(size_t)n * robot + lane - 6] = y;
    }
    if ((P.flags & BLFIK_TICK_INTEGRATE) && status != 2) {
        rvec[lane] = mine ? x : 0.0;
        __syncwarp();
        k4_integrate(n, P.base_pos + 3 * robot, P.base_quat + 4 * robot, P.s + (size_t)n * robot, rvec, P.dt, lane);
    }
    if (lane == 0) {
        P.status[robot] = worst;
        if (P.status_host) P.status_host[robot] = worst;
        P.iters[robot] = iters_prev + iters;
        if (SP.unsolved_run) SP.unsolved_run[robot] = status == 1 ? (run < (1 << 30) ? run + 1 : run) : 0;
        if (SP.slow_out && iters > SP.slow_threshold) {
            SP.slow_out[atomicAdd(SP.slow_out_n, 1)] = (int)robot;
            SP.slow_stamp_out[robot] = SP.stamp + 1;
        }
    }
    __syncwarp();
  }
}

}  // namespace blfik
