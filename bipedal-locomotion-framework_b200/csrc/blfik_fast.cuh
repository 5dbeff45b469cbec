// blfik_fast.cuh -- register-resident tick kernel for problems with nv <= 32 velocity variables and
// K <= 32 dense task rows (every humanoid configuration of BASELINE.json on ergoCub / iCub-26).
//
// Same algorithm family as blfik_device.cuh (K1 -> K2 -> K3 -> K4, one warp per robot, equality-projected ADMM
// + certified polish), re-derived so that the factorisation is small and every hot operand is a register:
//
//  * The Hessian of every problem on this path is "diagonal + a few dense rows": JointTrackingTask gives the
//    diagonal, SE3/SO3/CoM/R3 tasks give rows B_r (cost rows with weight w_r, or equality rows).  The x-update
//        x~ = argmin 1/2 x'Dx + sum_cost w_r/2 (B_r x - b_r)^2 - rhs'x   s.t.  B_eq x = b_eq,
//        D = diag(sigma + rho_i + joint weights)
//    is the KKT system [[D, B'], [B, -E]] (E_r = 1/w_r on cost rows, 0 on equality rows).  Eliminating x needs
//    only the K x K matrix G = E + B D^-1 B' = L_G L_G':
//        x~ = N rhs + x_b,   N = D^-1 - U U',   U = D^-1 B' L_G^-T,   x_b = U L_G^-1 b.
//    H is never formed.  ergoCub C3: K = 18 (15 equality + 3 chest rows) instead of an nv = 29 Cholesky, an nv x nv
//    triangular inverse and three nv x me products.
//  * Lane i keeps row i of N (and its x, z, y, lo, hi, rho) in registers; the only shared-memory operand of the
//    mat-vec is the broadcast right-hand side, read two doubles per LDS.128.
//  * The polish reuses the SAME factorisation: pinning the active joints is a rank-|A| correction of N
//    (x = x1 - N[:,A] (N[A,A])^-1 (x1[A] - bound)), and the multipliers of the pinned joints come out of that
//    correction, so the KKT certificate needs no second setup.  A rejected certificate updates the active set
//    (primal-dual active-set step); if that does not settle, ADMM continues.
#pragma once
#include "blfik_device.cuh"

namespace blfik {

__host__ __device__ constexpr int ld2odd(int n)
{
    int h = (n + 1) / 2;
    if ((h & 1) == 0) h++;
    return 2 * h;  // smallest 2*odd >= n: conflict-free for lane-strided 64/128-bit accesses, rows 16 B aligned
}

// FP64 tensor-core tile product C(8x8) += A(8x4) * B(4x8).  Fragments (g = lane >> 2, t = lane & 3):
// a = A[g][t], b = B[t][g], {c0, c1} = C[g][2t], C[g][2t + 1].  Measured on B200: 37.1 TFLOP/s, above the 33.8 TFLOP/s
// of a DFMA loop, at one issue slot per 256 FMAs (tcgen05 has no FP64 kind, so this legacy path is the FP64 tensor path).
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Phase alignment of the warps of a block (instruction-fetch sharing).  mode 1: the whole block; mode 2: only the
// warps that share a sub-partition (warp w lives on sub-partition w % 4), through a named barrier, so that a slow
// robot holds back two neighbours instead of the whole block; mode 0: none.
__device__ __forceinline__ void phase_barrier(int mode, int wib, int wpb)
{
    if (mode == 1) {
        __syncthreads();
    } else if (mode == 2) {
        const int grp = wib & 3;
        const int cnt = ((wpb - grp + 3) >> 2) * 32;   // warps grp, grp + 4, ... of this block
        asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(cnt) : "memory");
    } else if (mode == 3) {                            // pairs of warps of one sub-partition: grp, grp + 8
        const int grp = wib & 7;
        const int cnt = ((wpb - grp + 7) >> 3) * 32;
        asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(cnt) : "memory");
    }
}

// per-warp shared memory (offsets in doubles; the host fills this in)
struct FastLayout {
    int linkR, linkP, axisW, mcw;
    int LGt;    // KP x LDG : L_G stored by columns (column k contiguous), diagonal slot holds 1 / l_kk
    int Bm;     // KP x LDB : dense task rows
    int Us;     // NV x LDG : U rows
    int cvec;   // KP : right-hand sides b_r
    int evec;   // KP : E_r
    int dinv;   // LDB: 1 / D_i
    int rvec;   // LDB: broadcast vector of the mat-vec
    int wv;     // LDG: w = L_G^-1 b
    int dA, muA;  // 32 each
    int idxA;   // 32 ints (16 doubles)
    int Ns;     // alias: N rows for dynamic column access (over Bm/Us)
    int MA;     // alias: |A| x (|A| | 1) scratch for the active-set matrix (over the dead U region)
    int ma_cap; // largest |A| the alias can hold
    int total;
    // block-shared region (before the per-warp slices)
    int sh_model, sh_model_len;          // packed model constants: [nl][20] link records, then the tables below
    int sh_jump, sh_lod, sh_anc, sh_msub;  // offsets inside the model pack (doubles)
    int sh_tasks;                        // DevTask[ntasks]
    int sh_total;
};

struct FastParams {
    TickParams T;
    FastLayout F;
    int nv, K;          // actual sizes (<= NV, <= KP)
    int max_as_iters;   // active-set refinement steps per round
    int max_polish_iters;
    const double* model_pack;  // device copy of the packed model constants (layout in FastLayout)
    int njump;            // pointer-jumping steps = ceil(log2(tree depth + 1))
    int has_com;          // at least one CoMTask
    int pdas_first;       // 1: try the rho-free active-set attempt before ADMM (see the kernel)
    int sync_phases;      // 1: the warps of a block enter K1 and the setup together (block barriers)
    int max_total_iters;  // per-tick budget of mat-vecs (ADMM + polish); exceeding it ends the solve with status 1
    // Dynamic hand-out of robots: a team (the warps that meet at the phase barriers: a sub-partition group in sync
    // mode 2, the block in mode 1, a single warp in mode 0) takes the next `team size` robots from a device counter
    // instead of a fixed stride, so a team that drew expensive robots simply draws fewer.  The host zeroes the
    // counter before the launch.  nullptr: static striding.
    int* work_counter;
    // Streaming input (blfik_tick_packed, one persistent launch while the copy engine is still bringing the records in):
    // robots [first, first + *arrived) have their record on the device.  The copy stream raises *arrived after each piece
    // (a 4-byte copy ordered after the data copy); a team that drew robots beyond it polls the word -- a wait on a DMA
    // transfer, never on another kernel -- and gives up after arrive_timeout clocks (*abort_flag = 1, the host reports
    // the failure).  nullptr: everything is already resident.
    const int* arrived;
    int* abort_flag;        // device word: a team gave up, the others stop fetching
    int* abort_host;        // the same news in mapped host memory, read by the host after the launch
    long long arrive_timeout;
};

// Leader-side wait of the streaming-input mode: returns false on time-out.
__device__ __forceinline__ bool wait_arrived(const int* arrived, int need, long long timeout)
{
    int got;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(got) : "l"(arrived) : "memory");
    if (got >= need) return true;
    const long long t0 = clock64();
    while (true) {
        __nanosleep(200);
        asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(got) : "l"(arrived) : "memory");
        if (got >= need) return true;
        if (clock64() - t0 > timeout) return false;
    }
}

template <int NV, int KP>
struct FastDims {
    static constexpr int LDB = ld2odd(NV);
    static constexpr int LDG = ld2odd(KP);
    // Resident warps per SM (one block per SM).  The small instantiations run 16 warps at 128 registers with the row of N
    // re-read from shared memory in every mat-vec (measured on C3: 1.31 ms against 1.37 ms for 12 warps with the row in
    // registers); with more dense rows the 128-register build spills heavily (C4: 2.14 ms against 1.76 ms), so those
    // keep 12 warps at 168 registers and the row of N in registers.
    static constexpr int WARPS = (NV <= 29 && KP <= 18) ? 16 : 12;
    static constexpr bool NROW_SMEM = WARPS > 12;
};

// y_i = xb_i + sum_j N[i][j] * v[j]: own row streamed from shared memory (rows are 16-byte aligned and an odd number of
// 16-byte units apart, so the 128-bit loads of a warp are conflict-free), v broadcast
template <int NV>
__device__ __forceinline__ double fast_matvec_smem(const double* myrow, const double* v, double xb)
{
    double a0 = xb, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const double2* v2 = reinterpret_cast<const double2*>(v);
    const double2* n2 = reinterpret_cast<const double2*>(myrow);
#pragma unroll
    for (int jj = 0; jj < NV / 2; jj++) {
        const double2 t = v2[jj], m = n2[jj];
        if (jj & 1) {
            a2 = fma(m.x, t.x, a2);
            a3 = fma(m.y, t.y, a3);
        } else {
            a0 = fma(m.x, t.x, a0);
            a1 = fma(m.y, t.y, a1);
        }
    }
    if (NV & 1) a0 = fma(myrow[NV - 1], v[NV - 1], a0);
    return (a0 + a1) + (a2 + a3);
}

// y_i = xb_i + sum_j nrow[j] * v[j], v broadcast from shared memory two doubles per load
template <int NV>
__device__ __forceinline__ double fast_matvec(const double (&nrow)[NV], const double* v, double xb)
{
    double a0 = xb, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const double2* v2 = reinterpret_cast<const double2*>(v);
#pragma unroll
    for (int jj = 0; jj < NV / 2; jj++) {
        const double2 t = v2[jj];
        if (jj & 1) {
            a2 = fma(nrow[2 * jj], t.x, a2);
            a3 = fma(nrow[2 * jj + 1], t.y, a3);
        } else {
            a0 = fma(nrow[2 * jj], t.x, a0);
            a1 = fma(nrow[2 * jj + 1], t.y, a1);
        }
    }
    if (NV & 1) a0 = fma(nrow[NV - 1], v[NV - 1], a0);
    return (a0 + a1) + (a2 + a3);
}

// In-place inverse of the a x a SPD matrix M (row stride ld) in shared memory by Gauss-Jordan sweeps,
// row t by lane t.  a <= 32.  Returns false if a pivot is not positive.
// `floor`: pivots at or below it mean that the equality rows leave (a combination of) the selected joints no freedom:
// N[A,A] is then singular up to rounding noise and its "inverse" would scatter that noise over the whole solution.
// (N[j][j] <= 1 / D_j, so the floor of row k is 1e-6 / D of the joint it stands for.  N = D^-1 - U U' carries the rounding of a
// factorisation whose condition number reaches 1e7 .. 1e8 when the equality rows are nearly dependent -- an absolute error of
// about 1e-8 / D; with the former floor of 1e-10 a joint that those rows all but fix passed, and the correction through
// 1 / N[j][j] put the "certified" point 0.2 off the equality rows (fuzz seed 680: sigma_min(C_eq) = 5e-3).)
__device__ __forceinline__ bool fast_small_inverse(double* M, int a, int ld, int lane, const double* dinv, const int* idx)
{
    bool ok = true;
    for (int k = 0; k < a; k++) {
        const double p = M[k * ld + k];
        if (!(p > 1e-6 * dinv[idx[k]])) ok = false;
        const double ip = 1.0 / p;
        __syncwarp();
        if (lane == k) {
            for (int j = 0; j < a; j++) M[k * ld + j] = (j == k) ? ip : M[k * ld + j] * ip;
        }
        __syncwarp();
        if (lane < a && lane != k) {
            const double f = M[lane * ld + k];
            for (int j = 0; j < a; j++) M[lane * ld + j] = (j == k) ? -f * ip : M[lane * ld + j] - f * M[k * ld + j];
        }
        __syncwarp();
    }
    return ok;
}


// out[j-1] = in[j] - f * col[k + j] for j = 1 .. KP-1 (col = column k of L_G^T storage, 16 B aligned; PAR = k & 1).
// The arrays "rotate" by one slot per elimination step, so the pivot is always element 0 and the step loop
// can stay rolled with statically indexed registers (two steps per trip ping-pong between two arrays).
template <int KP, int PAR>
__device__ __forceinline__ void rot_update(const double (&in)[KP], double (&out)[KP], double f, const double* colk)
{
    if (PAR == 0) {
        out[0] = fma(-f, colk[1], in[1]);
#pragma unroll
        for (int j = 2; j + 1 < KP; j += 2) {
            const double2 c2 = *reinterpret_cast<const double2*>(colk + j);
            out[j - 1] = fma(-f, c2.x, in[j]);
            out[j] = fma(-f, c2.y, in[j + 1]);
        }
    } else {
#pragma unroll
        for (int j = 1; j + 1 < KP; j += 2) {
            const double2 c2 = *reinterpret_cast<const double2*>(colk + j);
            out[j - 1] = fma(-f, c2.x, in[j]);
            out[j] = fma(-f, c2.y, in[j + 1]);
        }
        out[KP - 2] = fma(-f, colk[KP - 1], in[KP - 1]);
    }
    out[KP - 1] = 0.0;
}

// One factorisation per tick.  In: task rows Bm (KP x LDB), 1/D (dinv), right-hand sides cvec, E (evec).
// Out: row `lane` of N in registers (nrow) and in shared memory (Ns), x_b.  Gs/Us share one region.
template <int NV, int KP>
__device__ __forceinline__ bool fast_setup(const double* Bm, double* Us, double* LGt, double* Ns, const double* dinv,
                                           const double* cvec, const double* evec, double* wv, double dinv_i, bool mine,
                                           int nv, int lane, double (&nrow)[NV], double& xb)
{
    constexpr int LDB = FastDims<NV, KP>::LDB;
    constexpr int LDG = FastDims<NV, KP>::LDG;
    double* Gs = Us;   // G by columns, consumed before U is written
    bool ok = true;
    [[maybe_unused]] bool clamped = false;   // KP == 32 only (no spare lane for the riding right-hand side)
    // ---- G = E + (B D^-1) B' on the FP64 tensor path (DMMA m8n8k4): ncu showed this contraction issue-bound
    // (34 % issue utilisation, 22 % FP64 pipe) as a row-per-lane DFMA loop; as 8x8 tiles it is 1/8 of the
    // instructions.  Only the tiles on and below the diagonal are formed (the Cholesky reads the lower triangle).
    {
        constexpr int RT = (KP + 7) / 8;   // row tiles of B
        constexpr int KS = (NV + 3) / 4;   // k steps over the velocity variables
        const int gr = lane >> 2, tc = lane & 3;
        double bf[RT][KS], dvv[KS];
#pragma unroll
        for (int s = 0; s < KS; s++) {
            const int col = 4 * s + tc;
            dvv[s] = (col < NV) ? dinv[col] : 0.0;
#pragma unroll
            for (int i = 0; i < RT; i++) {
                const int row = 8 * i + gr;
                bf[i][s] = (row < KP && col < NV) ? Bm[row * LDB + col] : 0.0;
            }
        }
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
                    if (col < KP) Gs[col * LDG + row] = c0 + (row == col ? er : 0.0);
                    if (col + 1 < KP) Gs[(col + 1) * LDG + row] = c1 + (row == col + 1 ? er : 0.0);
                }
            }
        }
    }
    __syncwarp();
    // ---- Cholesky G = L L' (rotating rows, rolled); column k of L goes to LGt[k][*], diagonal slot = 1 / l_kk
    {
        double gA[KP], gB[KP];
        const int myr = lane < KP ? lane : 0;
#pragma unroll
        for (int j = 0; j < KP; j++) gA[j] = Gs[j * LDG + myr];
        // With a spare lane (KP < 32) the right-hand side b rides along as one more row of the matrix: the same updates turn
        // it into L^-1 b step by step, so when a pivot is clamped (an equality row that depends on the previous ones) its
        // element 0 is what is left of that row's right-hand side.  It must vanish -- consistent duplicates are harmless, the
        // clamp acts like OSQP's regularised polish -- else the rows contradict each other, there is no feasible point and
        // the reference's advance() fails (QPInverseKinematics.cpp:518-523): status 2.
        if (KP < 32 && lane == KP) {
#pragma unroll
            for (int j = 0; j < KP; j++) gA[j] = cvec[j];
        }
        constexpr double kInconsistent = 1e-7;   // [m/s, rad/s]: exact duplicates leave ~1e-15, contradicting targets their difference
        // 1e-11: an exactly dependent row leaves a pivot of rounding noise, about 1e-14 of the largest diagonal entry after the
        // accumulations; a pivot this small in a full-rank set would mean singular values 3e-6 apart, beyond FP64 reach anyway
        const double floor_ = 1e-11 * warp_max(lane < KP ? Gs[myr * LDG + myr] : 0.0) + 1e-300;
        __syncwarp();
#pragma unroll 1
        for (int k = 0; k < KP; k += 2) {
            {
                double piv = __shfl_sync(kFull, gA[0], k);
                if (!(piv > floor_)) {
                    if (!(piv == piv) || (KP < 32 && lane == KP && fabs(gA[0]) > kInconsistent)) ok = false;
                    if (KP == 32) clamped = true;
                    piv = floor_;
                }
                const double rinv = rsqrt(piv);
                const double lrk = gA[0] * rinv;
                if (lane < KP) LGt[k * LDG + lane] = (lane == k) ? rinv : ((lane > k) ? lrk : 0.0);
                __syncwarp();
                rot_update<KP, 0>(gA, gB, lrk, LGt + k * LDG + k);
            }
            {
                const int k1 = k + 1;
                double piv = __shfl_sync(kFull, gB[0], k1);
                if (!(piv > floor_)) {
                    if (!(piv == piv) || (KP < 32 && lane == KP && fabs(gB[0]) > kInconsistent)) ok = false;
                    if (KP == 32) clamped = true;
                    piv = floor_;
                }
                const double rinv = rsqrt(piv);
                const double lrk = gB[0] * rinv;
                if (lane < KP) LGt[k1 * LDG + lane] = (lane == k1) ? rinv : ((lane > k1) ? lrk : 0.0);
                __syncwarp();
                rot_update<KP, 1>(gB, gA, lrk, LGt + k1 * LDG + k1);
            }
        }
    }
    // ---- U = D^-1 B' L^-T, row i by lane i (forward substitution, rotating); a spare lane solves L w = b
    {
        const bool spare = nv < 32;
        const bool wlane = spare && lane == nv;
        const int passes = spare ? 1 : 2;   // nv == 32: a second pass in which lane 0 solves for w
#pragma unroll 1
        for (int pass = 0; pass < passes; pass++) {
            const bool wl = pass == 1 ? (lane == 0) : wlane;
            double tA[KP], tB[KP];
#pragma unroll
            for (int j = 0; j < KP; j++) tA[j] = (pass == 0 && mine) ? Bm[j * LDB + lane] * dinv_i : (wl ? cvec[j] : 0.0);
#pragma unroll 1
            for (int c = 0; c < KP; c += 2) {
                {
                    const double uc = tA[0] * LGt[c * LDG + c];
                    if (pass == 0 && lane < NV) Us[lane * LDG + c] = mine ? uc : 0.0;
                    if (wl) wv[c] = uc;
                    rot_update<KP, 0>(tA, tB, uc, LGt + c * LDG + c);
                }
                {
                    const int c1 = c + 1;
                    const double uc = tB[0] * LGt[c1 * LDG + c1];
                    if (pass == 0 && lane < NV) Us[lane * LDG + c1] = mine ? uc : 0.0;
                    if (wl) wv[c1] = uc;
                    rot_update<KP, 1>(tB, tA, uc, LGt + c1 * LDG + c1);
                }
            }
        }
    }
    __syncwarp();
    if constexpr (KP == 32) {
        if (clamped) {
            // No spare lane carried b through the factorisation: check the same thing after the fact.  Every iterate is
            // x = N r + x_b with C_eq N = 0, so the equality residual of whatever the solve returns is C_eq x_b - b.
            double xbl = 0.0;
            if (mine)
                for (int r = 0; r < KP; r++) xbl = fma(Us[lane * LDG + r], wv[r], xbl);
            __syncwarp();
            LGt[lane] = xbl;   // L_G is not needed any more
            __syncwarp();
            bool bad = false;
            if (evec[lane] == 0.0) {
                double res = -cvec[lane];
                for (int j = 0; j < nv; j++) res = fma(Bm[lane * LDB + j], LGt[j], res);
                bad = !(fabs(res) <= 1e-6 * (1.0 + fabs(cvec[lane])));
            }
            if (__any_sync(kFull, bad)) ok = false;
            __syncwarp();
        }
    }
    // ---- N = D^-1 - U U' (row i by lane i, into Ns), x_b = U w
    {
        double u[KP];
        const double2* myu = reinterpret_cast<const double2*>(Us + (lane < NV ? lane : 0) * LDG);
#pragma unroll
        for (int rr = 0; rr < KP / 2; rr++) {
            const double2 t2 = myu[rr];
            u[2 * rr] = t2.x, u[2 * rr + 1] = t2.y;
        }
        {   // N = D^-1 - U U' as 8x8 DMMA tiles; the A and B fragments of U U' are the same registers
            constexpr int MT = (NV + 7) / 8;
            constexpr int KS = (KP + 3) / 4;
            const int gr = lane >> 2, tc = lane & 3;
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
        }
        const double2* w2 = reinterpret_cast<const double2*>(wv);
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int rr = 0; rr < KP / 2; rr++) {
            const double2 t2 = w2[rr];
            a0 = fma(u[2 * rr], t2.x, a0);
            a1 = fma(u[2 * rr + 1], t2.y, a1);
        }
        xb = mine ? a0 + a1 : 0.0;
    }
    __syncwarp();
    if constexpr (!FastDims<NV, KP>::NROW_SMEM) {
        const double2* myn = reinterpret_cast<const double2*>(Ns + (lane < NV ? lane : 0) * LDB);
#pragma unroll
        for (int jj = 0; jj < NV / 2; jj++) {
            const double2 t2 = myn[jj];
            nrow[2 * jj] = t2.x, nrow[2 * jj + 1] = t2.y;
        }
        if (NV & 1) nrow[NV - 1] = Ns[(lane < NV ? lane : 0) * LDB + NV - 1];
    }
    return __all_sync(kFull, ok);   // the consistency verdict is held by one lane
}

// Column `lane` of the angular rows of the centroidal momentum matrix (AngularMomentumTask.cpp:137-185; same formula as
// k1_angmom of the generic kernel): sum over the links distal to the DoF of I_k^W a + m_k (c_k - c) x (a x (c_k - o)).
// Rare task, long loop: kept out of line so that its registers do not weigh on the tick kernel.
static __device__ __noinline__ void fast_angmom_column(const double* s_link, const double* inertia, const unsigned long long* s_anc,
                                                const int* s_lod, const double* linkR, const double* linkP,
                                                const double* axisW, const double* mcw, const double* com, int nl, int lane,
                                                double* col)
{
    col[0] = col[1] = col[2] = 0.0;
    if (lane < 3) return;   // a translation of the whole robot carries no momentum about its centre of mass
    double a[3] = {0.0, 0.0, 0.0}, o[3];
    int j = -1;
    if (lane < 6) {
        a[lane - 3] = 1.0;
        o[0] = linkP[0], o[1] = linkP[1], o[2] = linkP[2];
    } else {
        j = lane - 6;
        const int kj = s_lod[j];
        for (int i = 0; i < 3; i++) a[i] = axisW[3 * kj + i], o[i] = linkP[3 * kj + i];
    }
    for (int k = 0; k < nl; k++) {
        if (j >= 0 && !((s_anc[k] >> j) & 1ull)) continue;
        const double mk = s_link[20 * k + 18];
        const double* R = linkR + 9 * k;
        if (inertia) {
            const double* I6 = inertia + 6 * k;
            double al[3], il[3], iw[3];
            mat3T_vec(R, a, al);
            il[0] = I6[0] * al[0] + I6[1] * al[1] + I6[2] * al[2];
            il[1] = I6[1] * al[0] + I6[3] * al[1] + I6[4] * al[2];
            il[2] = I6[2] * al[0] + I6[4] * al[1] + I6[5] * al[2];
            mat3_vec(R, il, iw);
            col[0] += iw[0], col[1] += iw[1], col[2] += iw[2];
        }
        if (mk > 0.0) {
            const double inv = 1.0 / mk;
            const double ck[3] = {mcw[3 * k] * inv, mcw[3 * k + 1] * inv, mcw[3 * k + 2] * inv};
            const double r[3] = {ck[0] - o[0], ck[1] - o[1], ck[2] - o[2]};
            const double d[3] = {ck[0] - com[0], ck[1] - com[1], ck[2] - com[2]};
            double v[3], t[3];
            cross3(a, r, v);
            cross3(d, v, t);
            col[0] += mk * t[0], col[1] += mk * t[1], col[2] += mk * t[2];
        }
    }
}

// Next `cnt` robots for a team (index relative to FP.T.first); in the streaming-input mode the caller also waits until
// their records are on the device.  A time-out ends the launch: every later fetch returns "past the end".
__device__ __forceinline__ int fetch_robots(const FastParams& FP, int cnt, int count_total)
{
    const int v = atomicAdd(FP.work_counter, cnt);
    if (FP.arrived && v < count_total) {
        const int need = v + cnt < count_total ? v + cnt : count_total;
        if (*reinterpret_cast<volatile int*>(FP.abort_flag) || !wait_arrived(FP.arrived, need, FP.arrive_timeout)) {
            *reinterpret_cast<volatile int*>(FP.abort_flag) = 1;
            *reinterpret_cast<volatile int*>(FP.abort_host) = 1;
            return 0x7fffffff - 64;
        }
    }
    return v;
}

// AM: the problem holds an AngularMomentumTask.  A separate instantiation, because even an out-of-line call in the row
// generation costs the common kernels 8 % (registers live across the call are spilled: 1.32 -> 1.43 ms on C3).
template <int NV, int KP, bool AM = false>
__global__ void __launch_bounds__(FastDims<NV, KP>::WARPS * 32, 1) blfik_tick_fast_kernel(const FastParams FP)
{
    constexpr int LDB = FastDims<NV, KP>::LDB;
    constexpr int LDG = FastDims<NV, KP>::LDG;
    extern __shared__ double smem[];
    const TickParams& P = FP.T;
    const FastLayout& F = FP.F;
    const DevModel& m = P.m;
    const Settings& st = P.st;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    // block-shared constants: the packed model (per-link records, pointer-jumping tables, ancestor masks, subtree
    // masses) and the task descriptors, staged once per block
    for (int e = threadIdx.x; e < F.sh_model_len; e += blockDim.x) smem[F.sh_model + e] = FP.model_pack[e];
    {
        const double* tsrc = reinterpret_cast<const double*>(P.tasks);
        const int tlen = P.ntasks * (int)(sizeof(DevTask) / sizeof(double));
        for (int e = threadIdx.x; e < tlen; e += blockDim.x) smem[F.sh_tasks + e] = tsrc[e];
    }
    __syncthreads();
    const double* s_link = smem + F.sh_model;                                      // [nl][20]
    const int* s_jump = reinterpret_cast<const int*>(smem + F.sh_model + F.sh_jump);   // [njump][32]
    const int* s_lod = reinterpret_cast<const int*>(smem + F.sh_model + F.sh_lod);     // link_of_dof[32]
    const int* s_dol = s_lod + 32;                                                     // dof_of_link[32]
    const unsigned long long* s_anc = reinterpret_cast<const unsigned long long*>(smem + F.sh_model + F.sh_anc);
    const double* s_msub = smem + F.sh_model + F.sh_msub;
    const DevTask* s_tasks = reinterpret_cast<const DevTask*>(smem + F.sh_tasks);
    double* base = smem + F.sh_total + (size_t)wib * F.total;
    double* linkR = base + F.linkR;
    double* linkP = base + F.linkP;
    double* axisW = base + F.axisW;
    double* mcw = base + F.mcw;
    double* LGt = base + F.LGt;
    double* Bm = base + F.Bm;
    double* Us = base + F.Us;
    double* cvec = base + F.cvec;
    double* evec = base + F.evec;
    double* dinv = base + F.dinv;
    double* rvec = base + F.rvec;
    double* wv = base + F.wv;
    double* dA = base + F.dA;
    double* muA = base + F.muA;
    int* idxA = reinterpret_cast<int*>(base + F.idxA);
    double* Ns = base + F.Ns;
    double* MA = base + F.MA;
    const int n = m.n, nv = FP.nv, K = FP.K;
    const bool mine = lane < nv;         // this lane owns velocity variable `lane`
    const bool joint = mine && lane >= 6;

    // Optional phase alignment (FP.sync_phases, env BLFIK_SYNC_PHASES): with the first, 120 KB version of this
    // kernel stall_no_instruction was the top stall, and warps that enter the long straight-line phases (K1/K2,
    // setup) together share the fetched lines.  After the code shrank, the barriers cost as much in load imbalance
    // (robots need different numbers of iterations) as they save in fetches (measured: 2.09 ms with, 2.00 ms without),
    // so they are off by default.  Warps past the end of the batch shadow the last robot without writing anything,
    // so that every warp reaches every barrier when they are on.
    __shared__ int s_next[8][2];
    const bool dynamic = FP.work_counter != nullptr;
    const int count_total = (int)(P.N - P.first);
    long long base0 = P.first + (long long)blockIdx.x * wpb;
    for (int trip = 0;; trip++) {
        int member = wib;
        if (dynamic) {
            int v;
            if (FP.sync_phases == 2 || FP.sync_phases == 3) {
                const int sh = FP.sync_phases == 2 ? 2 : 3, msk = (1 << sh) - 1;
                const int grp = wib & msk, cnt = (wpb - grp + msk) >> sh;
                if (wib == grp && lane == 0) s_next[grp][trip & 1] = fetch_robots(FP, cnt, count_total);
                asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(cnt * 32) : "memory");
                v = s_next[grp][trip & 1];
                member = wib >> sh;
            } else if (FP.sync_phases == 1) {
                if (threadIdx.x == 0) s_next[0][trip & 1] = fetch_robots(FP, wpb, count_total);
                __syncthreads();
                v = s_next[0][trip & 1];
            } else {
                v = lane == 0 ? fetch_robots(FP, 1, count_total) : 0;
                v = __shfl_sync(kFull, v, 0);
                member = 0;
            }
            base0 = P.first + v;
        } else if (trip > 0) {
            base0 += (long long)gridDim.x * wpb;
        }
        if (base0 >= P.N) break;
        const bool active = base0 + member < P.N;
        const long long robot = active ? base0 + member : P.N - 1;
        double* bp = P.base_pos + 3 * robot;
        double* bq = P.base_quat + 4 * robot;
        double* s = P.s + (size_t)n * robot;
        if (P.packed && active && !(P.flags & BLFIK_TICK_KEEP_STATE)) unpack_state(P, robot, bp, bq, s, n, lane);
        const bool warm_flag = (P.flags & BLFIK_TICK_WARM_START) != 0;
        double x = 0.0, z = 0.0, y = 0.0;
        if (warm_flag && mine) {
            x = P.wx[(size_t)nv * robot + lane];
            if (joint) z = P.wz[(size_t)n * robot + lane - 6], y = P.wy[(size_t)n * robot + lane - 6];
        }
        int status = 2, iters_total = 0, worst = 0;
        bool dead = false;
        for (int tick = 0; tick < P.nticks; tick++) {
          // Two attempts per tick.  Attempt 0 (FP.pdas_first): factorise WITHOUT the ADMM penalty (rho = 0), so the
          // proximal steps contract by sigma / (sigma + lambda) ~ 1e-2 and each active-set guess is solved in 2-3
          // mat-vecs; the active set is refined by primal-dual steps and accepted only with the KKT certificate.
          // Attempt 1 (the fallback, and the only attempt when pdas_first = 0): OSQP-style ADMM with rho on the box
          // rows predicts the active set, the same certified polish finishes.  The fallback redoes K1/K2 because the
          // factorisation reuses their shared memory.
          bool done = false;
          int iters = 0;
          status = 1;
          const int first_attempt = FP.pdas_first ? 0 : 1;
          for (int attempt = first_attempt; attempt < 2 && !done && status != 2; attempt++) {
            const bool with_rho = attempt == 1;
            if (attempt == first_attempt) phase_barrier(FP.sync_phases, wib, wpb);
            // ================================================================ K1
            // One link per lane: local joint transform, then composition along the ancestor chain by pointer jumping
            // (ceil(log2(depth + 1)) steps; the parent transform comes over warp shuffles, not shared memory).
            const int nl = m.nl;
            double com[3] = {0.0, 0.0, 0.0};
            {
                const double* lp = s_link + 20 * (lane < nl ? lane : 0);
                double R[9], p[3];
                if (lane == 0) {
                    const double q[4] = {bq[0], bq[1], bq[2], bq[3]};
                    quat_to_rot(q, R);
                    p[0] = bp[0], p[1] = bp[1], p[2] = bp[2];
                } else if (lane < nl) {
                    const double a0 = lp[12], a1 = lp[13], a2 = lp[14];
                    double sn, cs;
                    sincos(s[s_dol[lane]], &sn, &cs);
                    const double v = 1.0 - cs;
                    double Rj[9], X[9];
                    Rj[0] = cs + a0 * a0 * v, Rj[1] = a0 * a1 * v - a2 * sn, Rj[2] = a0 * a2 * v + a1 * sn;
                    Rj[3] = a1 * a0 * v + a2 * sn, Rj[4] = cs + a1 * a1 * v, Rj[5] = a1 * a2 * v - a0 * sn;
                    Rj[6] = a2 * a0 * v - a1 * sn, Rj[7] = a2 * a1 * v + a0 * sn, Rj[8] = cs + a2 * a2 * v;
#pragma unroll
                    for (int c = 0; c < 9; c++) X[c] = lp[c];
                    mat3_mul(X, Rj, R);
                    p[0] = lp[9], p[1] = lp[10], p[2] = lp[11];
                } else {
#pragma unroll
                    for (int c = 0; c < 9; c++) R[c] = (c % 4 == 0) ? 1.0 : 0.0;
                    p[0] = p[1] = p[2] = 0.0;
                }
                for (int stp = 0; stp < FP.njump; stp++) {
                    const int a = s_jump[stp * 32 + lane];
                    const int src = a < 0 ? lane : a;
                    double Ra[9], pa[3];
#pragma unroll
                    for (int c = 0; c < 9; c++) Ra[c] = __shfl_sync(kFull, R[c], src);
#pragma unroll
                    for (int c = 0; c < 3; c++) pa[c] = __shfl_sync(kFull, p[c], src);
                    if (a >= 0) {
                        double Rn[9], off[3];
                        mat3_mul(Ra, R, Rn);
                        mat3_vec(Ra, p, off);
#pragma unroll
                        for (int c = 0; c < 9; c++) R[c] = Rn[c];
#pragma unroll
                        for (int c = 0; c < 3; c++) p[c] = pa[c] + off[c];
                    }
                }
                const double ax[3] = {lp[12], lp[13], lp[14]}, cl[3] = {lp[15], lp[16], lp[17]};
                double aw[3], cw[3];
                mat3_vec(R, ax, aw);
                mat3_vec(R, cl, cw);
                const double mk = lane < nl ? lp[18] : 0.0;
                double mc[3];
#pragma unroll
                for (int c = 0; c < 3; c++) mc[c] = mk * (p[c] + cw[c]);
                if (lane < nl) {
#pragma unroll
                    for (int c = 0; c < 9; c++) linkR[9 * lane + c] = R[c];
#pragma unroll
                    for (int c = 0; c < 3; c++) linkP[3 * lane + c] = p[c], axisW[3 * lane + c] = aw[c], mcw[3 * lane + c] = mc[c];
                }
                if (FP.has_com) {
#pragma unroll
                    for (int c = 0; c < 3; c++) com[c] = warp_sum(mc[c]) * m.inv_total_mass;
                }
            }
            // ================================================================ K2
            // zero the row matrix (padding rows / columns must read as zero)
            for (int e = lane; e < KP * LDB / 2; e += 32) reinterpret_cast<double2*>(Bm)[e] = make_double2(0.0, 0.0);
            if (lane < KP) cvec[lane] = 0.0, evec[lane] = 1.0;
            __syncwarp();
            // data of this lane's own joint (column 6 + j of every Jacobian) and, for CoM tasks, its CoM column
            const int jj = joint ? lane - 6 : 0;
            const int kj = s_lod[jj];
            const double oj[3] = {linkP[3 * kj], linkP[3 * kj + 1], linkP[3 * kj + 2]};
            const double aj[3] = {axisW[3 * kj], axisW[3 * kj + 1], axisW[3 * kj + 2]};
            const double pB[3] = {linkP[0], linkP[1], linkP[2]};
            double comcol[3] = {0.0, 0.0, 0.0};
            if (FP.has_com && mine) {
                if (lane < 3) {
                    comcol[lane] = 1.0;
                } else if (lane < 6) {
                    const double r[3] = {com[0] - pB[0], com[1] - pB[1], com[2] - pB[2]};
                    double e[3] = {0.0, 0.0, 0.0};
                    e[lane - 3] = 1.0;
                    cross3(e, r, comcol);
                } else {
                    double mc[3] = {0.0, 0.0, 0.0};
                    for (int k = 1; k < nl; k++)
                        if ((s_anc[k] >> jj) & 1ull) mc[0] += mcw[3 * k], mc[1] += mcw[3 * k + 1], mc[2] += mcw[3 * k + 2];
                    const double ms = s_msub[jj];
                    const double d[3] = {(mc[0] - ms * oj[0]) * m.inv_total_mass, (mc[1] - ms * oj[1]) * m.inv_total_mass,
                                         (mc[2] - ms * oj[2]) * m.inv_total_mass};
                    cross3(aj, d, comcol);
                }
            }
            // control laws: lane t evaluates dense task t (frame pose, R3 / SO(3) P-controllers) once, for everybody
            double* tbuf = Us;   // [ntasks][12]: frame origin, linear control, angular control (U is not alive yet)
            if (lane < P.ntasks) {
                const DevTask& tk = s_tasks[lane];
                const bool dense = is_dense_type(tk.type);
                if (dense) {
                    double spl[13];
                    const double* rec = task_record(P, tk, robot);
                    const bool compact = task_compact(P, tk);
                    // the exogenous feedback sits behind the set-point: sp0[tk.fb_off + c] addresses it in both layouts
                    const double* sp0 = compact ? rec - (tk.fb_off - tk.base_c) : rec;
                    if (compact) {   // pose part only: the full record is rebuilt with zero feed-forward entries
                        expand_compact(tk, rec, spl);
                        rec = spl;
                    }
                    const double* sp = traj_eval(tk, robot, P.t0 + tick * P.dt, rec, spl);
                    const bool has_lin = tk.type == T_SE3 || tk.type == T_R3 || tk.type == T_COM, has_ang = tk.type == T_SE3 || tk.type == T_SO3;
                    double Rf[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, pf[3] = {com[0], com[1], com[2]};
                    if (tk.type != T_COM && tk.type != T_ANGMOM) {
                        const int Lk = tk.flink;
                        double Rl[9], off[3];
#pragma unroll
                        for (int c = 0; c < 9; c++) Rl[c] = linkR[9 * Lk + c];
                        mat3_mul(Rl, tk.Fr, Rf);
                        mat3_vec(Rl, tk.Fp, off);
#pragma unroll
                        for (int c = 0; c < 3; c++) pf[c] = linkP[3 * Lk + c] + off[c];
                    }
                    const int voff = tk.type == T_SE3 ? 7 : 3;
                    const bool lin_on = tk.type == T_COM || tk.type == T_R3 || tk.enabled != 0;   // only SE3Task has a controller mode
                    double bl[3] = {0.0, 0.0, 0.0}, ba[3];
                    if (has_lin) {
                        // exogenous feedback (SE3Task.cpp:266-269, CoMTask.cpp:192-195): the controller compares the set-point
                        // with the measured position stored behind it in the record; the rows still use the model's pf
                        const bool exo = tk.exo_pos && (tk.type == T_SE3 || tk.type == T_COM);
#pragma unroll
                        for (int c = 0; c < 3; c++) {
                            const double pc = exo ? sp0[tk.fb_off + c] : pf[c];
                            bl[c] = lin_on ? sp[voff + c] + tk.kp_lin[c] * (sp[c] - pc) : sp[voff + c];
                        }
                    }
                    double qd[4] = {0.0, 0.0, 0.0, 1.0}, wff[3] = {0.0, 0.0, 0.0};
                    if (has_ang) {
                        const int qo = tk.type == T_SE3 ? 3 : 0, wo = tk.type == T_SE3 ? 10 : 4;
#pragma unroll
                        for (int c = 0; c < 4; c++) qd[c] = sp[qo + c];
#pragma unroll
                        for (int c = 0; c < 3; c++) wff[c] = sp[wo + c];
                    }
                    if (tk.type == T_SE3 && tk.exo_ori) {   // SE3Task.cpp:261-265: measured orientation (Rf is not used again for an SE3 task)
                        const double qf[4] = {sp0[tk.fb_off + 3], sp0[tk.fb_off + 4], sp0[tk.fb_off + 5], sp0[tk.fb_off + 6]};
                        quat_to_rot(qf, Rf);
                    }
                    so3_control(Rf, qd, wff, tk.kp_ang, tk.type == T_SO3 ? true : (tk.enabled != 0), ba);
                    double aux[3] = {0.0, 0.0, 0.0};
                    if (tk.type == T_DISTANCE) {   // DistanceTask.cpp:178-229: u = d / max(thr, |d|), b = kp (d_des - |d|)
                        if (tk.flink2 >= 0) {
                            double off[3];
                            mat3_vec(linkR + 9 * tk.flink2, tk.Fp2, off);
#pragma unroll
                            for (int c = 0; c < 3; c++) aux[c] = linkP[3 * tk.flink2 + c] + off[c];   // origin of the reference frame
                        }
                        const double dv[3] = {pf[0] - aux[0], pf[1] - aux[1], pf[2] - aux[2]};
                        const double d = sqrt(dv[0] * dv[0] + dv[1] * dv[1] + dv[2] * dv[2]);
                        const double inv = 1.0 / fmax(d, tk.thr);
                        bl[0] = tk.kp_lin[0] * (sp[0] - d);
#pragma unroll
                        for (int c = 0; c < 3; c++) ba[c] = dv[c] * inv;   // the unit direction, reused by the row generation
                    } else if (tk.type == T_ANGMOM) {    // AngularMomentumTask.cpp:187-210: b is the desired momentum itself
                        bl[0] = sp[0], bl[1] = sp[1], bl[2] = sp[2];
                    } else if (tk.type == T_GRAVITY) {   // GravityTask.cpp:149-185
                        double zb[3] = {sp[0], sp[1], sp[2]}, fb[3] = {sp[3], sp[4], sp[5]}, z[3], ff[3];
                        const double zn = rsqrt(zb[0] * zb[0] + zb[1] * zb[1] + zb[2] * zb[2]);
                        zb[0] *= zn, zb[1] *= zn, zb[2] *= zn;
                        mat3_vec(Rf, zb, z);
                        mat3_vec(Rf, fb, ff);
                        bl[0] = ff[0] + tk.kp_lin[0] * z[1], bl[1] = ff[1] - tk.kp_lin[0] * z[0];
                    }
#pragma unroll
                    for (int c = 0; c < 3; c++)
                        tbuf[12 * lane + c] = pf[c], tbuf[12 * lane + 3 + c] = bl[c], tbuf[12 * lane + 6 + c] = ba[c], tbuf[12 * lane + 9 + c] = aux[c];
                }
            }
            __syncwarp();
            double dh = 0.0, gd = 0.0, lo = -kInf, hi = kInf;   // diagonal weight, sum w b, joint box
            const double sj = joint ? s[lane - 6] : 0.0;
            for (int t = 0; t < P.ntasks; t++) {
                const DevTask& tk = s_tasks[t];
                if (is_dense_type(tk.type)) {
                    // row map: masked linear rows first, then the three angular rows (SE3Task.cpp:289-316)
                    int map[6] = {-1, -1, -1, -1, -1, -1}, src = 0, rows = 0;
                    if (tk.type == T_DISTANCE) {          // one row: u'(J_lin,target - J_lin,reference), written through slot 0
                        map[0] = tk.brow, rows = 1;
                    } else if (tk.type == T_GRAVITY) {    // two rows: the first two angular rows
                        if (lane == 1) src = 1;
                        map[3] = tk.brow, map[4] = tk.brow + 1, rows = 2;
                    } else if (tk.type != T_SO3) {
#pragma unroll
                        for (int c = 0; c < 3; c++)
                            if (tk.mask[c]) {
                                if (lane == rows) src = c;
                                map[c] = tk.brow + rows++;
                            }
                    }
                    if (tk.type == T_SE3 || tk.type == T_SO3) {
#pragma unroll
                        for (int c = 0; c < 3; c++) {
                            if (lane == rows) src = 3 + c;
                            map[3 + c] = tk.brow + rows++;
                        }
                    }
                    if (mine) {
                        double lin[3] = {0.0, 0.0, 0.0}, ang[3] = {0.0, 0.0, 0.0};
                        if (tk.type == T_COM) {
                            lin[0] = comcol[0], lin[1] = comcol[1], lin[2] = comcol[2];
                        } else if (AM && tk.type == T_ANGMOM) {
                            if constexpr (AM) fast_angmom_column(s_link, m.inertia, s_anc, s_lod, linkR, linkP, axisW, mcw, com, nl, lane, lin);
                        } else {
                            const double pf[3] = {tbuf[12 * t], tbuf[12 * t + 1], tbuf[12 * t + 2]};
                            if (lane < 3) {
                                lin[lane] = 1.0;
                            } else if (lane < 6) {
                                const double r[3] = {pf[0] - pB[0], pf[1] - pB[1], pf[2] - pB[2]};
                                ang[lane - 3] = 1.0;
                                cross3(ang, r, lin);
                            } else if ((s_anc[tk.flink] >> jj) & 1ull) {
                                const double d[3] = {pf[0] - oj[0], pf[1] - oj[1], pf[2] - oj[2]};
                                ang[0] = aj[0], ang[1] = aj[1], ang[2] = aj[2];
                                cross3(aj, d, lin);
                            }
                            if (tk.type == T_DISTANCE) {
                                // subtract the reference frame's linear column, then project on the unit direction u
                                if (tk.flink2 >= 0) {
                                    const double pr[3] = {tbuf[12 * t + 9], tbuf[12 * t + 10], tbuf[12 * t + 11]};
                                    double l2[3] = {0.0, 0.0, 0.0};
                                    if (lane < 3) {
                                        l2[lane] = 1.0;
                                    } else if (lane < 6) {
                                        const double r[3] = {pr[0] - pB[0], pr[1] - pB[1], pr[2] - pB[2]};
                                        double e[3] = {0.0, 0.0, 0.0};
                                        e[lane - 3] = 1.0;
                                        cross3(e, r, l2);
                                    } else if ((s_anc[tk.flink2] >> jj) & 1ull) {
                                        const double d2[3] = {pr[0] - oj[0], pr[1] - oj[1], pr[2] - oj[2]};
                                        cross3(aj, d2, l2);
                                    }
                                    lin[0] -= l2[0], lin[1] -= l2[1], lin[2] -= l2[2];
                                }
                                lin[0] = tbuf[12 * t + 6] * lin[0] + tbuf[12 * t + 7] * lin[1] + tbuf[12 * t + 8] * lin[2];
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 3; i++) {
                            if (map[i] >= 0) Bm[map[i] * LDB + lane] = lin[i];
                            if (map[3 + i] >= 0) Bm[map[3 + i] * LDB + lane] = ang[i];
                        }
                    }
                    if (lane < rows) {
                        cvec[tk.brow + lane] = tbuf[12 * t + 3 + src];
                        evec[tk.brow + lane] = tk.priority == 0 ? 0.0 : 1.0 / tk.weight[lane];
                    }
                } else if (joint) {
                    const double* sp = task_record(P, tk, robot);
                    const int j = lane - 6;
                    if (tk.type == T_JOINT_TRACKING) {
                        double pj, vj;
                        joint_setpoint(tk, robot, P.t0 + tick * P.dt, sp, task_compact(P, tk), n, j, pj, vj);
                        const double bj = vj + tk.kp_joint[j] * (pj - sj);
                        if (tk.priority == 1) {
                            dh += tk.weight[j];
                            gd += tk.weight[j] * bj;
                        } else {
                            lo = fmax(lo, bj), hi = fmin(hi, bj);
                        }
                    } else if (tk.type == T_JOINT_LIMITS) {
                        hi = fmin(hi, tk.klim[j] * (tk.upper[j] - sj) * tk.inv_dt);
                        lo = fmax(lo, tk.klim[j] * (tk.lower[j] - sj) * tk.inv_dt);
                    } else {
                        hi = fmin(hi, tk.upper[j]);
                        lo = fmax(lo, tk.lower[j]);
                    }
                }
            }
            __syncwarp();   // the control buffer lives in the U region: everybody is done with it before the setup writes G there
            // per-variable ADMM step: 0 (no bound), rho (box), 1e3 rho (pinned: lo == hi)
            const bool bounded = joint && (lo > -kInfHalf || hi < kInfHalf);
            const bool pinned = bounded && lo == hi;
            const double rho_i = (bounded && with_rho) ? (pinned ? 1e3 * st.rho : st.rho) : 0.0;
            const double D_i = mine ? st.sigma + rho_i + dh : 1.0;
            const double Q_i = mine ? st.sigma + rho_i : 0.0;
            const double dinv_i = 1.0 / D_i;
            const double inv_rho = rho_i > 0.0 ? 1.0 / rho_i : 0.0;
            dinv[lane] = mine ? dinv_i : 0.0;
            if (LDB > 32 && lane < LDB - 32) dinv[32 + lane] = 0.0, rvec[32 + lane] = 0.0;
            const bool any_box = __any_sync(kFull, bounded);
            const bool infeasible = __any_sync(kFull, bounded && lo > hi);
            const double gmax = warp_max(fabs(gd));
            __syncwarp();

            if (P.dbg.H && tick == 0 && active && attempt == first_attempt) {   // parity introspection only: materialise H, g, C_eq, b_eq, box
                const int me = P.L.me;
                if (mine) {
                    for (int jx = 0; jx < nv; jx++) {
                        double h = (jx == lane) ? dh : 0.0;
                        for (int r = 0; r < K; r++)
                            if (evec[r] > 0.0) h += Bm[r * LDB + lane] * Bm[r * LDB + jx] / evec[r];
                        P.dbg.H[((size_t)robot * nv + lane) * nv + jx] = h;
                    }
                    double g = -gd;
                    int er = 0;
                    for (int r = 0; r < K; r++) {
                        if (evec[r] > 0.0) g -= Bm[r * LDB + lane] * cvec[r] / evec[r];
                        else {
                            P.dbg.Ceq[((size_t)robot * me + er) * nv + lane] = Bm[r * LDB + lane];
                            if (lane == 0) P.dbg.beq[(size_t)robot * me + er] = cvec[r];
                            er++;
                        }
                    }
                    P.dbg.g[(size_t)robot * nv + lane] = g;
                    if (joint) P.dbg.lo[(size_t)robot * n + lane - 6] = lo, P.dbg.hi[(size_t)robot * n + lane - 6] = hi;
                }
                __syncwarp();
            }

            if (attempt == first_attempt) phase_barrier(FP.sync_phases, wib, wpb);
            double nrow[NV];   // unused (and eliminated) when the row is streamed from shared memory
            const double* myrow = Ns + (lane < NV ? lane : 0) * LDB;
            double xb = 0.0;
            if (infeasible) {
                status = 2;
            } else {
                // ============================================================ K3 setup
                const bool fac_ok = fast_setup<NV, KP>(Bm, Us, LGt, Ns, dinv, cvec, evec, wv, dinv_i, mine, nv, lane, nrow, xb);
                if (!fac_ok) status = 2;
            }

            if (status != 2) {
                // ============================================================ K3 iterations
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
                for (int round = 0; round < max_rounds && !done && iters < FP.max_total_iters; round++) {
                    // ---- ADMM (skipped when there is no box row: the problem is a pure equality QP)
                    double rp = 0.0, rd = 0.0;
                    if (any_box && with_rho) {
                        for (int it = 0; it < st.admm_iters; it++) {
                            rvec[lane] = mine ? st.sigma * x + gd + (bounded ? rho_i * z - y : 0.0) : 0.0;
                            __syncwarp();
                            const double xt = FastDims<NV, KP>::NROW_SMEM ? fast_matvec_smem<NV>(myrow, rvec, xb) : fast_matvec<NV>(nrow, rvec, xb);
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
                    for (int as_it = 0; as_it < FP.max_as_iters && !done && iters < FP.max_total_iters; as_it++) {
                        const unsigned amask = __ballot_sync(kFull, act != 0);
                        const int a = __popc(amask);
                        const int rank = __popc(amask & ((1u << lane) - 1u));
                        const double bd = (act == 1) ? lo : hi;   // pinned: lo == hi
                        const int mld = a | 1;
                        bool small_ok = true;
                        if (a > FP.F.ma_cap) break;               // cannot hold the active-set matrix: leave it to ADMM
                        if (a > 0) {
                            if (act) idxA[rank] = lane;
                            __syncwarp();
                            if (lane < a) {
                                const int ri = idxA[lane];
                                for (int uu = 0; uu < a; uu++) MA[lane * mld + uu] = Ns[ri * LDB + idxA[uu]];
                            }
                            __syncwarp();
                            // a joint whose N[j][j] is 1e-6 of its 1 / D_j is fixed by the equality rows and cannot
                            // be pinned to a bound: this active set has no solution, ADMM (and its verdict) takes over
                            small_ok = fast_small_inverse(MA, a, mld, lane, dinv, idxA);
                            small_ok = __all_sync(kFull, small_ok);
                        }
                        if (!small_ok) break;
                        if (act) xp = bd;
                        double dx = 0.0;
                        bool conv = false;
                        for (int it = 0; it < FP.max_polish_iters; it++) {
                            rvec[lane] = mine ? Q_i * xp + gd : 0.0;
                            __syncwarp();
                            double x1 = FastDims<NV, KP>::NROW_SMEM ? fast_matvec_smem<NV>(myrow, rvec, xb) : fast_matvec<NV>(nrow, rvec, xb);
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
                            // measured before the pinned joints are set to their bounds: there x1 - xp is what the correction
                            // left of x1 - bound, which only vanishes if the equality rows allow this active set (a nearly
                            // singular N[A,A] otherwise passes the floor of fast_small_inverse and the certificate below would
                            // accept a point that is off the equality rows)
                            dx = mine ? Q_i * fabs(x1 - xp) : 0.0;
                            if (a > 0 && act) x1 = bd, mu_mine = muA[rank];
                            xp = mine ? x1 : 0.0;
                            iters++;
                            if ((it & 1) == 1 || it + 1 == FP.max_polish_iters) {
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
                        if (conv && !any_change) {
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
                        status = 0;
                    }
                    if (!any_box && !done) break;   // nothing ADMM could add
                }
                if (!done && status != 2) status = 1;
            }
            if (!any_box) break;   // a pure equality QP has no second attempt
          }
            iters_total += iters;
            // A failed tick ends this robot's rollout: no velocity to report (the reference's advance() fails), the state stays
            // where the failure happened.  The warp still walks through the remaining ticks -- its team meets at the phase
            // barriers of every tick -- but reports nothing from them.
            if (dead || (status == 1 && (P.flags & BLFIK_TICK_STRICT))) status = 2;
            worst = status > worst ? status : worst;
            if (status == 2) x = 0.0, dead = true;
            if (mine && active) {
                P.nu[(size_t)nv * robot + lane] = x;
                if (P.nu_host) P.nu_host[(size_t)nv * robot + lane] = x;
            }
            if ((P.flags & BLFIK_TICK_INTEGRATE) && active && !dead) {
                rvec[lane] = mine ? x : 0.0;
                __syncwarp();
                k4_integrate(n, bp, bq, s, rvec, P.dt, lane);
            }
            __syncwarp();
        }
        if (mine && active) {
            P.wx[(size_t)nv * robot + lane] = x;
            if (joint) P.wz[(size_t)n * robot + lane - 6] = z, P.wy[(size_t)n * robot + lane - 6] = y;
        }
        if (lane == 0 && active) {
            P.status[robot] = worst;   // the worst tick of the launch, not the last one
            if (P.status_host) P.status_host[robot] = worst;
            P.iters[robot] = iters_total;
        }
        __syncwarp();
    }
}

}  // namespace blfik
