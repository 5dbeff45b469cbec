// blfik_assemble.cuh -- K1 + K2 of the two-launch tick (see blfik_split.cuh): ONE THREAD PER ROBOT.
//
// Control flow depends on the model and the task list only, so it is uniform across a warp and its cost (type dispatch, row
// maps, tree bookkeeping) is shared by 32 robots; no shuffles, no idle lanes.  Replaces the task update() calls (SE3Task.cpp:231-321,
// CoMTask.cpp:171-235, SO3Task.cpp:153-192, R3Task.cpp:178-243, GravityTask.cpp:149-199, JointTrackingTask.cpp:127-155,
// JointLimitsTask.cpp:211-259, JointVelocityLimitsTask.cpp:134-145) and the stacking of QPInverseKinematics.cpp:442-491.
#pragma once
#include "blfik_split.cuh"
#include <type_traits>

namespace blfik {

// K1 + K2, one thread per robot, ONE depth-first pass over the tree.  What a thread has to remember is small: the poses of the
// links that are still open (a stack as deep as the tree), the mass-weighted CoM sums of their subtrees, and the origins of the
// task frames met so far.  When the walk leaves link k, every frame distal to k has been visited, so column 6 + dof(k) of every
// task's rows can be written there and then: [a_k x (p_F - o_k); a_k] if k is on the path to the frame, zeros otherwise, and for a
// CoMTask a_k x (sum_{sub(k)} m c - M_sub o_k) / M.  Nothing per link survives the walk.  The stack is a per-thread array in
// local memory (thread-interleaved by the hardware, served by the L1): measured on C3, 200 us against 425 us with the stack in
// shared memory, which holds only one block of 128 robots per SM (one warp per scheduler; every dependency stall exposed).
// The records are robot-major (kernel B loads one with contiguous 16-byte loads) and hold the rows by COLUMNS: a finished
// column is KP contiguous doubles, which the warp passes through a shared-memory tile so that its global stores are runs of
// KP doubles instead of single ones.
// Tasks this kernel does not assemble: DistanceTask (its direction needs both frames before any column) and
// AngularMomentumTask (needs the CoM before any column): problems holding one take the fused kernel.
template <int KP>
__global__ void __launch_bounds__(128, 4) blfik_assemble_kernel(const AsmParams A)
{
    extern __shared__ double smem[];
    const TickParams& P = A.T;
    const DevModel& m = P.m;
    const int n = m.n, nv = n + 6, tid = threadIdx.x, nthr = blockDim.x;
    if (blockIdx.x == 0 && tid == 0) {
        *A.work_counter = 0;   // the solve kernel of this tick hands its robots out from zero
        if (A.slow_out_n) *A.slow_out_n = 0;
    }
    // block-shared constants: the packed model, the task descriptors, the walk, the row maps
    for (int e = tid; e < A.pack_len; e += nthr) smem[e] = A.model_pack[e];
    const int tlen = P.ntasks * (int)(sizeof(DevTask) / sizeof(double));
    {
        const double* tsrc = reinterpret_cast<const double*>(P.tasks);
        for (int e = tid; e < tlen; e += nthr) smem[A.pack_len + e] = tsrc[e];
    }
    int* s_ev = reinterpret_cast<int*>(smem + A.pack_len + tlen);
    const int ev_doubles = (A.nevents + 1) / 2 + 1;
    for (int e = tid; e < A.nevents; e += nthr) s_ev[e] = A.events[e];
    int* s_map = reinterpret_cast<int*>(smem + A.pack_len + tlen + ev_doubles);   // [ntasks][8]: rows of the linear / angular components, dense ordinal, row count
    const DevTask* s_tasks = reinterpret_cast<const DevTask*>(smem + A.pack_len);
    __syncthreads();
    if (tid == 0) {
        int ord = 0;
        for (int t = 0; t < P.ntasks; t++) {
            const DevTask& tk = s_tasks[t];
            int* mp = s_map + 8 * t;
            for (int i = 0; i < 8; i++) mp[i] = -1;
            if (!is_dense_type(tk.type)) continue;
            int rows = 0;
            if (tk.type == T_GRAVITY) {           // two rows: the first two angular rows
                mp[3] = tk.brow, mp[4] = tk.brow + 1, rows = 2;
            } else if (tk.type != T_SO3) {        // masked linear rows first (SE3Task.cpp:289-316)
                for (int c = 0; c < 3; c++)
                    if (tk.mask[c]) mp[c] = tk.brow + rows++;
            }
            if (tk.type == T_SE3 || tk.type == T_SO3)
                for (int c = 0; c < 3; c++) mp[3 + c] = tk.brow + rows++;
            mp[6] = ord++, mp[7] = rows;
        }
    }
    __syncthreads();
    const double* s_link = smem;                                                   // [nl][20]: Xr Xp axis com mass
    const int* s_lod = reinterpret_cast<const int*>(smem + A.sh_lod);              // link_of_dof[32]
    const int* s_dol = s_lod + 32;                                                 // dof_of_link[32]
    const unsigned long long* s_anc = reinterpret_cast<const unsigned long long*>(smem + A.sh_anc);
    const double* s_msub = smem + A.sh_msub;
    constexpr int TS = KP | 1;                                                     // tile row stride (odd: conflict-free)
    const int lane = tid & 31;
    double* tile = smem + A.pack_len + tlen + ev_doubles + 4 * P.ntasks + (tid >> 5) * 32 * TS;   // this warp's [32][TS] staging tile
    double* trow = tile + lane * TS;
    // per level of the walk: world axis (3) and origin (3) of the open link, CoM sum of its subtree (3); the rotation of an open
    // link is only kept where the walk comes back to (links with two children or more), the others hand theirs to their only
    // child in registers
    double l_stk[kAsmMaxDepth * 6], l_acc[kAsmMaxDepth * 3], l_rot[kAsmMaxDepth * 9], l_pf[kAsmMaxDense * 3];
#define STK(d, c) l_stk[(d) * 6 + (c)]
#define ACC(d, c) l_acc[(d) * 3 + (c)]
#define ROT(d, c) l_rot[(d) * 9 + (c)]
#define PF(o, c) l_pf[(o) * 3 + (c)]

    const long long wfirst = P.first + (long long)blockIdx.x * nthr + (tid & ~31);   // first robot of this warp
    // Lanes past the end of the batch, or whose rollout ended at an earlier tick of the launch, shadow a valid robot (they are
    // needed in the warp-wide copies below) and write nothing.
    bool active = wfirst + lane < P.N;
    const long long robot = active ? wfirst + lane : P.N - 1;
    if (A.tick > 0 && P.status[robot] == 2) active = false;
    const unsigned valid = __ballot_sync(kFull, active);
    if (valid == 0u) return;
    double* wq = A.qp + (size_t)wfirst * A.rec_len;
    double* const myrhs = A.qp + (size_t)robot * A.rec_len + A.off_c;   // right-hand sides go out one by one as they are known
    // The 32 x LEN doubles the lanes have put into their tile rows go to doubles [off, off + LEN) of the 32 records: sixteen
    // lanes per robot write sixteen consecutive doubles (two robots per store instruction), and what is left of a row beyond
    // a multiple of sixteen goes out lane by lane.
    const int fl_e = lane & 15, fl_h = lane >> 4;
    auto flush = [&](int off, auto len_tag) {
        constexpr int LEN = decltype(len_tag)::value;
        __syncwarp();
#pragma unroll
        for (int pass = 0; pass < LEN / 16; pass++) {
            double* dst = wq + (size_t)fl_h * A.rec_len + off + 16 * pass + fl_e;
            const double* src = tile + fl_h * TS + 16 * pass + fl_e;
#pragma unroll 4
            for (int rr = 0; rr < 32; rr += 2) {
                if ((valid >> (rr + fl_h)) & 1u) *dst = *src;
                dst += 2 * (size_t)A.rec_len, src += 2 * TS;
            }
        }
        if (active) {
#pragma unroll
            for (int e = LEN / 16 * 16; e < LEN; e++) wq[(size_t)lane * A.rec_len + off + e] = trow[e];
        }
        __syncwarp();
    };
    using LenCol = std::integral_constant<int, KP>;
    using LenVec = std::integral_constant<int, 16>;
#pragma unroll
    for (int r = 0; r < KP; r++) trow[r] = 0.0;   // rows beyond K stay zero in every column
    if (active)
        for (int r = A.K; r < KP; r++) myrhs[r] = 0.0;

    // ---- state
    double* bp = P.base_pos + 3 * robot;
    double* bq = P.base_quat + 4 * robot;
    double* s = P.s + (size_t)n * robot;
    if (P.packed && active && !(P.flags & BLFIK_TICK_KEEP_STATE)) {   // packed input records carry the state: it also goes to the state arrays (K4, blfik_get_state)
        const double* pr = P.packed + (size_t)robot * P.packed_len;
        for (int i = 0; i < 3; i++) bp[i] = pr[i];
        for (int i = 0; i < 4; i++) bq[i] = pr[3 + i];
        for (int j = 0; j < n; j++) s[j] = pr[7 + j];
    }

    // frame pose of the dense tasks attached to link k (pose R, p): control law -> right-hand sides, origin kept for the columns
    auto frames_on_link = [&](int k, const double* R, const double* p) {
        for (int t = 0; t < P.ntasks; t++) {
            const DevTask& tk = s_tasks[t];
            const int* mp = s_map + 8 * t;
            if (mp[7] < 0 || tk.type == T_COM || tk.flink != k) continue;
            double Rf[9], off[3], pf[3];
            mat3_mul(R, tk.Fr, Rf);
            mat3_vec(R, tk.Fp, off);
#pragma unroll
            for (int c = 0; c < 3; c++) pf[c] = p[c] + off[c], PF(mp[6], c) = pf[c];
            double spl[13];
            const double* recsp = task_record(P, tk, robot);
            const bool compact = task_compact(P, tk);
            // the exogenous feedback sits behind the set-point: sp0[tk.fb_off + c] addresses it in both layouts
            const double* sp0 = compact ? recsp - (tk.fb_off - tk.base_c) : recsp;
            if (compact) {   // pose part only: the full record is rebuilt with zero feed-forward entries
                expand_compact(tk, recsp, spl);
                recsp = spl;
            }
            const double* sp = traj_eval(tk, robot, A.tnow, recsp, spl);
            if (tk.type == T_SE3 || tk.type == T_R3) {
                const int voff = tk.type == T_SE3 ? 7 : 3;
                const bool lin_on = tk.type == T_R3 || tk.enabled != 0;   // only SE3Task has a controller mode
                const bool exo = tk.exo_pos && tk.type == T_SE3;
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    const double pc = exo ? sp0[tk.fb_off + c] : pf[c];
                    const double bl = lin_on ? sp[voff + c] + tk.kp_lin[c] * (sp[c] - pc) : sp[voff + c];
                    if (mp[c] >= 0 && active) myrhs[mp[c]] = bl;
                }
            }
            if (tk.type == T_SE3 || tk.type == T_SO3) {
                const int qo = tk.type == T_SE3 ? 3 : 0, wo = tk.type == T_SE3 ? 10 : 4;
                const double qd[4] = {sp[qo], sp[qo + 1], sp[qo + 2], sp[qo + 3]}, wff[3] = {sp[wo], sp[wo + 1], sp[wo + 2]};
                double ba[3];
                if (tk.type == T_SE3 && tk.exo_ori) {   // SE3Task.cpp:261-265: measured orientation
                    const double qf[4] = {sp0[tk.fb_off + 3], sp0[tk.fb_off + 4], sp0[tk.fb_off + 5], sp0[tk.fb_off + 6]};
                    quat_to_rot(qf, Rf);
                }
                so3_control(Rf, qd, wff, tk.kp_ang, tk.type == T_SO3 ? true : (tk.enabled != 0), ba);
#pragma unroll
                for (int c = 0; c < 3; c++)
                    if (active) myrhs[mp[3 + c]] = ba[c];
            } else if (tk.type == T_GRAVITY) {   // GravityTask.cpp:149-185
                double zb[3] = {sp[0], sp[1], sp[2]}, fb[3] = {sp[3], sp[4], sp[5]}, z[3], ff[3];
                const double zn = rsqrt(zb[0] * zb[0] + zb[1] * zb[1] + zb[2] * zb[2]);
                zb[0] *= zn, zb[1] *= zn, zb[2] *= zn;
                mat3_vec(Rf, zb, z);
                mat3_vec(Rf, fb, ff);
                if (active) myrhs[mp[3]] = ff[0] + tk.kp_lin[0] * z[1], myrhs[mp[4]] = ff[1] - tk.kp_lin[0] * z[0];
            }
        }
    };

    // ---- K1 + the joint columns of K2: the walk
    double Rcur[9], pcur[3];   // pose of the link visited last
    int dcur = 0;
    {
        const double qb[4] = {bq[0], bq[1], bq[2], bq[3]};
        double R0[9];
        quat_to_rot(qb, R0);
        const double p0[3] = {bp[0], bp[1], bp[2]};
        const double* lp = s_link;
        const double cl[3] = {lp[15], lp[16], lp[17]};
        double cw[3];
        mat3_vec(R0, cl, cw);
#pragma unroll
        for (int c = 0; c < 9; c++) ROT(0, c) = R0[c], Rcur[c] = R0[c];
#pragma unroll
        for (int c = 0; c < 3; c++) STK(0, 3 + c) = p0[c], pcur[c] = p0[c], ACC(0, c) = lp[18] * (p0[c] + cw[c]);
        frames_on_link(0, R0, p0);
    }
    for (int ev = 0; ev < A.nevents; ev++) {
        const int code = s_ev[ev], k = code & 0xff, d = (code >> 8) & 0xff;
        const double* lp = s_link + 20 * k;
        if (((code >> 16) & 1) == 0) {
            // visit: pose of link k from its parent's (one level up)
            const double a0 = lp[12], a1 = lp[13], a2 = lp[14];
            double sn, cs;
            sincos(s[s_dol[k]], &sn, &cs);
            const double v = 1.0 - cs;
            double Rj[9], X[9], Rl[9], Rp[9], R[9], p[3];
            Rj[0] = cs + a0 * a0 * v, Rj[1] = a0 * a1 * v - a2 * sn, Rj[2] = a0 * a2 * v + a1 * sn;
            Rj[3] = a1 * a0 * v + a2 * sn, Rj[4] = cs + a1 * a1 * v, Rj[5] = a1 * a2 * v - a0 * sn;
            Rj[6] = a2 * a0 * v - a1 * sn, Rj[7] = a2 * a1 * v + a0 * sn, Rj[8] = cs + a2 * a2 * v;
            double pp[3];
            if (d == dcur + 1) {   // the parent is the link visited last: its pose is still in registers
#pragma unroll
                for (int c = 0; c < 9; c++) Rp[c] = Rcur[c];
#pragma unroll
                for (int c = 0; c < 3; c++) pp[c] = pcur[c];
            } else {               // the walk came back to a link with several children
#pragma unroll
                for (int c = 0; c < 9; c++) Rp[c] = ROT(d - 1, c);
#pragma unroll
                for (int c = 0; c < 3; c++) pp[c] = STK(d - 1, 3 + c);
            }
#pragma unroll
            for (int c = 0; c < 9; c++) X[c] = lp[c];
            mat3_mul(X, Rj, Rl);
            mat3_mul(Rp, Rl, R);
            const double xp[3] = {lp[9], lp[10], lp[11]}, cl[3] = {lp[15], lp[16], lp[17]}, ax[3] = {a0, a1, a2};
            double off[3], cw[3], aw[3];
            mat3_vec(Rp, xp, off);
            mat3_vec(R, cl, cw);
            mat3_vec(R, ax, aw);
            if (code & (1 << 17)) {   // two children or more: the rotation is needed again
#pragma unroll
                for (int c = 0; c < 9; c++) ROT(d, c) = R[c];
            }
#pragma unroll
            for (int c = 0; c < 9; c++) Rcur[c] = R[c];
#pragma unroll
            for (int c = 0; c < 3; c++) {
                p[c] = pp[c] + off[c];
                pcur[c] = p[c];
                STK(d, c) = aw[c], STK(d, 3 + c) = p[c];
                ACC(d, c) = lp[18] * (p[c] + cw[c]);
            }
            dcur = d;
            frames_on_link(k, R, p);
        } else {
            // close: column 6 + dof(k) of every dense task; the subtree's CoM sum goes up one level
            const int j = s_dol[k], col = 6 + j;
            double p[3], ms[3], aw[3];
#pragma unroll
            for (int c = 0; c < 3; c++) aw[c] = STK(d, c), p[c] = STK(d, 3 + c), ms[c] = ACC(d, c);
            for (int t = 0; t < P.ntasks; t++) {
                const DevTask& tk = s_tasks[t];
                const int* mp = s_map + 8 * t;
                if (mp[7] < 0) continue;
                double lin[3] = {0.0, 0.0, 0.0}, ang[3] = {0.0, 0.0, 0.0};
                if (tk.type == T_COM) {
                    const double msub = s_msub[j];
                    const double dd[3] = {(ms[0] - msub * p[0]) * m.inv_total_mass, (ms[1] - msub * p[1]) * m.inv_total_mass,
                                          (ms[2] - msub * p[2]) * m.inv_total_mass};
                    cross3(aw, dd, lin);
                } else if ((s_anc[tk.flink] >> j) & 1ull) {
                    const double dd[3] = {PF(mp[6], 0) - p[0], PF(mp[6], 1) - p[1], PF(mp[6], 2) - p[2]};
                    cross3(aw, dd, lin);
                    ang[0] = aw[0], ang[1] = aw[1], ang[2] = aw[2];
                }
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    if (mp[i] >= 0) trow[mp[i]] = lin[i];
                    if (mp[3 + i] >= 0) trow[mp[3 + i]] = ang[i];
                }
            }
            flush(col * KP, LenCol{});
#pragma unroll
            for (int c = 0; c < 3; c++) ACC(d - 1, c) += ms[c];
        }
    }
    // ---- the CoM task's right-hand side, then the base columns: MIXED base block [I, -S(p - p_B); 0, I]
    const double com[3] = {ACC(0, 0) * m.inv_total_mass, ACC(0, 1) * m.inv_total_mass, ACC(0, 2) * m.inv_total_mass};
    const double pB[3] = {STK(0, 3), STK(0, 4), STK(0, 5)};
    for (int t = 0; t < P.ntasks; t++) {
        const DevTask& tk = s_tasks[t];
        const int* mp = s_map + 8 * t;
        if (mp[7] < 0 || tk.type != T_COM) continue;
        double spl[13];
        const double* recsp = task_record(P, tk, robot);
        const bool compact = task_compact(P, tk);
        const double* sp0 = compact ? recsp - (tk.fb_off - tk.base_c) : recsp;
        if (compact) {
            expand_compact(tk, recsp, spl);
            recsp = spl;
        }
        const double* sp = traj_eval(tk, robot, A.tnow, recsp, spl);
#pragma unroll
        for (int c = 0; c < 3; c++) {   // CoMTask.cpp:192-203: b = v_ff + kp (p_d - p_com), measured CoM with exogenous feedback
            const double pc = tk.exo_pos ? sp0[tk.fb_off + c] : com[c];
            if (mp[c] >= 0 && active) myrhs[mp[c]] = sp[3 + c] + tk.kp_lin[c] * (sp[c] - pc);
        }
        PF(mp[6], 0) = com[0], PF(mp[6], 1) = com[1], PF(mp[6], 2) = com[2];
    }
#pragma unroll 1
    for (int c = 0; c < 6; c++) {
        for (int t = 0; t < P.ntasks; t++) {
            const DevTask& tk = s_tasks[t];
            const int* mp = s_map + 8 * t;
            if (mp[7] < 0) continue;
            double lin[3] = {0.0, 0.0, 0.0}, ang[3] = {0.0, 0.0, 0.0};
            if (c < 3) {
                lin[c] = 1.0;
            } else {
                const double r[3] = {PF(mp[6], 0) - pB[0], PF(mp[6], 1) - pB[1], PF(mp[6], 2) - pB[2]};
                double e[3] = {0.0, 0.0, 0.0};
                e[c - 3] = 1.0;
                cross3(e, r, lin);
                if (tk.type != T_COM) ang[c - 3] = 1.0;
            }
#pragma unroll
            for (int i = 0; i < 3; i++) {
                if (mp[i] >= 0) trow[mp[i]] = lin[i];
                if (mp[3 + i] >= 0) trow[mp[3 + i]] = ang[i];
            }
        }
        flush(c * KP, LenCol{});
    }

    // ---- K2: joint-space tasks -> diagonal cost, its gradient, the box; four variables per pass, [variable][dh gd lo hi]
    double inv_width = 0.0, max_width = 0.0;   // of the bounded joints: what the hand-out order of the solve kernel is made of
    int nbounded = 0;
    for (int g = 0; g < 8; g++) {
#pragma unroll 1
        for (int ii = 0; ii < 4; ii++) {
            const int i = 4 * g + ii;
            double dh = 0.0, gd = 0.0, lo = -kInf, hi = kInf;
            if (i >= 6 && i < nv) {
                const int j = i - 6;
                const double sj = s[j];
                for (int t = 0; t < P.ntasks; t++) {
                    const DevTask& tk = s_tasks[t];
                    if (s_map[8 * t + 7] >= 0) continue;
                    if (tk.type == T_JOINT_TRACKING) {
                        const double* sp = task_record(P, tk, robot);
                        double pj, vj;
                        joint_setpoint(tk, robot, A.tnow, sp, task_compact(P, tk), n, j, pj, vj);
                        const double bj = vj + tk.kp_joint[j] * (pj - sj);
                        if (tk.priority == 1) {
                            const double wj = t == A.jw_task ? A.jw[j] : tk.weight[j];
                            dh += wj;
                            gd += wj * bj;
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
            trow[4 * ii] = dh, trow[4 * ii + 1] = gd, trow[4 * ii + 2] = lo, trow[4 * ii + 3] = hi;
            if (lo > -kInfHalf && hi < kInfHalf) {
                const double wd = fmax(hi - lo, 1e-3);
                inv_width += 1.0 / wd, max_width = fmax(max_width, wd), nbounded++;
            }
        }
        flush(A.off_v + 16 * g, LenVec{});
    }
    // Hardest first: a robot whose joints sit near a position limit has a narrow box on those joints and needs more active-set
    // work (C3: ordering a cold batch by sum_j (widest box / box_j - 1) takes 13 % off the solve kernel, whose end is its slowest
    // robot).  The robots are binned by that score; the solve kernel hands the bins out from the top.
    if (A.bin_cnt) {
        const double score = max_width * inv_width - nbounded;
        const int b = active ? (score > 0.0 ? (score < 7.875 ? (int)(8.0 * score) : 63) : 0) : 64;   // 64: not in any bin
        // one atomic per bin and warp: the lanes of a bin elect a leader, who claims their places
        const unsigned peers = __match_any_sync(kFull, b);
        const int leader = __ffs(peers) - 1;
        int basepos = 0;
        if (lane == leader && b < 64) basepos = atomicAdd(A.bin_cnt + b, __popc(peers));
        basepos = __shfl_sync(kFull, basepos, leader);
        if (b < 64) A.bin_list[(size_t)b * A.bin_cap + basepos + __popc(peers & ((1u << lane) - 1u))] = (int)robot;
    }
#undef STK
#undef ACC
#undef ROT
#undef PF
}

}  // namespace blfik
