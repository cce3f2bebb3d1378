// Pair force / energy / virial evaluation over the Verlet list, the fused
// velocity-Verlet step, minimiser updates and the device reductions.
#pragma once
#include "b2_common.cuh"

// Block of the step / force kernels.  The blocks resident on an SM share its L1, where the
// neighbours' coordinates live; fewer, larger blocks cover a more compact piece of the (Morton-
// sorted) system and so need less of it: 0.241 / 0.224 / 0.216 / 0.213 ms per step (fp32, N = 1.2M)
// with 128 / 256 / 512 / 1024 threads at the same 48 warps per SM (1024: 32 warps).
#ifndef FORCE_THREADS
#define FORCE_THREADS 512
#endif
// neighbours in flight per lane (index -> position -> pair arithmetic, software-pipelined); fp32
// at 2 resident blocks of 512 threads has the registers for 6 (0.216 -> 0.205 ms per step against
// 4 at 3 blocks: fewer warps, but more loads in flight each and a smaller working set in L1)
#ifndef B2_UNROLL32
#define B2_UNROLL32 6
#endif
#ifndef B2_UNROLL64
#define B2_UNROLL64 4
#endif
#ifndef B2_PIPE
#define B2_PIPE 1
#endif
// resident blocks per SM asked of ptxas for the step kernel: fp32 fits 3 x 512 threads in 40
// registers; 2 x 512 with 64 registers and a deeper pipeline (B2_UNROLL32) is faster still; fp64 runs
// best at 2 x 512 in 64
#ifndef B2_MINBLK32
#define B2_MINBLK32 2
#endif
#ifndef B2_MINBLK64
#define B2_MINBLK64 2
#endif
#define RED_THREADS 256


template <typename R> struct PairTab {       // per-block copy of the pair table
    R c12x12[B2_MAXSP * B2_MAXSP], c6x6[B2_MAXSP * B2_MAXSP], rc2[B2_MAXSP * B2_MAXSP];
};

template <typename R>
__device__ __forceinline__ void load_pair_table(PairTab<R> &t)
{
    for (int a = threadIdx.x; a < B2_MAXSP * B2_MAXSP; a += blockDim.x) {
        t.c12x12[a] = (R)(12.0 * g_pair.c12[a]);
        t.c6x6[a] = (R)(6.0 * g_pair.c6[a]);
        t.rc2[a] = (R)g_pair.rc2[a];
    }
}

// Body of one pair evaluation, shared by the row walkers (pair_row here, pair_row_win in
// b2_win.cuh).  Uses the caller's pi, box, tab, ti, the register-held table row (a12_0 ...),
// the accumulators fx fy fz pe vir st.  #undef'd at the end of b2_win.cuh.
// tabulated potentials: value at r2 by linear interpolation between the two samples around it
__device__ __forceinline__ float tab_w(float, int a, float r2)
{
    const float t = r2 * (float)g_pair.tab_scale[a];
    const int k = min((int)t, g_pair.tab_np - 2);
    const float *w = g_pair.tab_wf + (size_t)a * g_pair.tab_np + k;
    const float w0 = __ldg(w), w1 = __ldg(w + 1);
    return fmaf(t - (float)k, w1 - w0, w0);
}
__device__ __forceinline__ double tab_w(double, int a, double r2)
{
    const double t = r2 * g_pair.tab_scale[a];
    const int k = min((int)t, g_pair.tab_np - 2);
    const double *w = g_pair.tab_w + (size_t)a * g_pair.tab_np + k;
    const double w0 = __ldg(w), w1 = __ldg(w + 1);
    return w0 + (t - (double)k) * (w1 - w0);
}
__device__ __forceinline__ void tab_uw(int a, double r2, double &u, double &w)
{
    const double t = r2 * g_pair.tab_scale[a];
    const int k = min((int)t, g_pair.tab_np - 2);
    const double f = t - (double)k;
    const double *pu = g_pair.tab_u + (size_t)a * g_pair.tab_np + k, *pw = g_pair.tab_w + (size_t)a * g_pair.tab_np + k;
    const double u0 = __ldg(pu), u1 = __ldg(pu + 1), w0 = __ldg(pw), w1 = __ldg(pw + 1);
    u = u0 + f * (u1 - u0);
    w = w0 + f * (w1 - w0);
}

#define B2_RCP(x) r_rcp(x)
#define B2_PAIR(PJ)                                                                   \
    {                                                                                 \
        R dx, dy, dz;                                                                 \
        delta(pi, PJ, box, dx, dy, dz);                                               \
        R r2 = dx * dx + dy * dy + dz * dz;                                           \
        const int tj = ptype(PJ);                                                     \
        R a12, a6, rc2;                                                               \
        if (FAST2) {                                                                  \
            a12 = tj ? a12_1 : a12_0; a6 = tj ? a6_1 : a6_0; rc2 = tj ? rc_1 : rc_0;  \
        } else {                                                                      \
            a12 = tab.c12x12[ti + tj]; a6 = tab.c6x6[ti + tj]; rc2 = tab.rc2[ti + tj]; \
        }                                                                             \
        R r2inv = B2_RCP(r2);                                                         \
        R fp;                                                                         \
        if (TAB) {                                                                    \
            /* tabulated: w = r.F interpolated linearly in r^2, F = (w / r^2) d */    \
            rc2 = (R)g_pair.rc2[ti + tj];                                             \
            fp = tab_w((R)0, ti + tj, r2) * r2inv;                                    \
        } else {                                                                      \
            R r6inv = r2inv * r2inv * r2inv;                                          \
            fp = r6inv * (a12 * r6inv - a6) * r2inv;                                  \
        }                                                                             \
        if (EV || STRESS) {                                                           \
            /* observables: fp64 pair arithmetic on the stored positions, in both */ \
            /* precisions (totals cancel heavily; fp32 would leave systematic    */ \
            /* errors that do not average out).  The cut-off decision of the     */ \
            /* REPORTED forces is taken on the same fp64 distance: the force is  */ \
            /* discontinuous at r_c (|F(r_c)| ~ 0.04), and among ~4e7 pairs a    */ \
            /* few sit within float round-off of it                              */ \
            const int a = ti + tj;                                                    \
            double ex, ey, ez;                                                        \
            delta_d(pi, PJ, box, ex, ey, ez);                                         \
            double q2 = ex * ex + ey * ey + ez * ez;                                  \
            const bool inside = q2 < g_pair.rc2[a];                                   \
            fp = inside ? fp : (R)0;                                                  \
            if (inside) {                                                             \
                double q2inv = 1.0 / q2;                                              \
                double wd, ud;                                                        \
                if (TAB) {                                                            \
                    tab_uw(a, q2, ud, wd);                                            \
                } else {                                                              \
                    double q6inv = q2inv * q2inv * q2inv;                             \
                    double d12 = g_pair.c12[a], d6 = g_pair.c6[a];                    \
                    wd = q6inv * (12.0 * d12 * q6inv - 6.0 * d6);                     \
                    ud = q6inv * (d12 * q6inv - d6) - g_pair.ushift[a];               \
                }                                                                     \
                if (EV) { pe += ud; vir += wd; }                                      \
                if (STRESS) {                                                         \
                    double gp = wd * q2inv;                                           \
                    double gx = gp * ex, gy = gp * ey, gz = gp * ez;                  \
                    st[0] += ex * gx; st[1] += ex * gy; st[2] += ex * gz;             \
                    st[3] += ey * gy; st[4] += ey * gz; st[5] += ez * gz;             \
                }                                                                     \
            }                                                                         \
        } else {                                                                      \
            fp = r2 < rc2 ? fp : (R)0;                                                \
        }                                                                             \
        fx += fp * dx; fy += fp * dy; fz += fp * dz;                                  \
    }

// One particle against its Verlet row.  Pair expressions are the oracle's:
//   r2inv = 1/r2 ; r6inv = r2inv^3
//   u = r6inv (c12 r6inv - c6) - ushift ;  w = r6inv (12 c12 r6inv - 6 c6) = r.F
//   F_i += (w r2inv) d ,  d = r_i - r_j  (minimum image, particle.py:129-130; see delta())
// The force part is branch-free (pairs beyond the cut-off contribute a zero
// through a select).  FAST2: at most two species — the particle's row of the
// table sits in registers and is picked with selects instead of shared loads.
template <typename R, bool EV, bool STRESS, bool FAST2, bool TAB>
__device__ __forceinline__ void pair_row(const typename Pos4<R>::type *__restrict__ pos,
                                         const int *__restrict__ row, int cnt,
                                         typename Pos4<R>::type pi, const Box<R> &box,
                                         const PairTab<R> &tab, int nsp,
                                         R &fx, R &fy, R &fz, double &pe, double &vir, double *st)
{
    typedef typename Pos4<R>::type P4;
    const int ti = ptype(pi) * nsp;
    R a12_0 = 0, a6_0 = 0, rc_0 = 0, a12_1 = 0, a6_1 = 0, rc_1 = 0;
    if (FAST2) {
        a12_0 = tab.c12x12[ti]; a6_0 = tab.c6x6[ti]; rc_0 = tab.rc2[ti];
        a12_1 = tab.c12x12[ti + nsp - 1]; a6_1 = tab.c6x6[ti + nsp - 1]; rc_1 = tab.rc2[ti + nsp - 1];
    }
    int k = 0;
    // the deeper pipeline only where the registers are there for it: the step kernel (no energies)
    constexpr int B2_UNROLL = (sizeof(R) == 4 && !EV) ? B2_UNROLL32 : B2_UNROLL64;
#if B2_PIPE
    // software pipeline: the index loads of group g+1 are in flight while the positions of
    // group g arrive and its pairs are evaluated (index -> position is a dependent chain)
    int jj[B2_UNROLL];
#pragma unroll
    for (int u = 0; u < B2_UNROLL; u++) jj[u] = ld_stream(row + (size_t)min(u, max(cnt - 1, 0)) * 32);
    for (; k + B2_UNROLL <= cnt; k += B2_UNROLL) {
        P4 pp[B2_UNROLL];
#pragma unroll
        for (int u = 0; u < B2_UNROLL; u++) pp[u] = pos[jj[u]];
#pragma unroll
        for (int u = 0; u < B2_UNROLL; u++) jj[u] = ld_stream(row + (size_t)min(k + B2_UNROLL + u, cnt - 1) * 32);
#pragma unroll
        for (int u = 0; u < B2_UNROLL; u++) B2_PAIR(pp[u])
    }
#else
    for (; k + B2_UNROLL <= cnt; k += B2_UNROLL) {
        int jj[B2_UNROLL];
        P4 pp[B2_UNROLL];
#pragma unroll
        for (int u = 0; u < B2_UNROLL; u++) jj[u] = ld_stream(row + (size_t)(k + u) * 32);
#pragma unroll
        for (int u = 0; u < B2_UNROLL; u++) pp[u] = pos[jj[u]];
#pragma unroll
        for (int u = 0; u < B2_UNROLL; u++) B2_PAIR(pp[u])
    }
#endif
    for (; k < cnt; k++) {
        int j = ld_stream(row + (size_t)k * 32);
        P4 p0 = pos[j];
        B2_PAIR(p0)
    }
}

template <typename R> struct ForceArgs {
    Ctl *ctl;
    const typename Pos4<R>::type *pos[2];
    typename Vec4<R>::type *frc;    // (fx, fy, fz, pe_i)
    R *vir;                          // vir_i
    const int *nbr, *nnb;
    double *partial;                 // [gridDim][9] stress partial sums (STRESS only)
    int n, cap, guarded;
    Box<R> box;
};

// Forces (+ per-particle energy / virial, + stress partials) of the current
// configuration, stored.  `guarded` launches obey the speculative-batch flags.
template <typename R, bool EV, bool STRESS, bool TAB>
__global__ void __launch_bounds__(FORCE_THREADS)
k_force(ForceArgs<R> a)
{
    typedef typename Pos4<R>::type P4;
    const Ctl *ctl = a.ctl;
    if (a.guarded && (ctl->halt_rebuild | ctl->overflow | ctl->done)) return;
    __shared__ PairTab<R> tab;
    load_pair_table(tab);
    __syncthreads();
    const P4 *__restrict__ pos = ctl->cur ? a.pos[1] : a.pos[0];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    double st[6] = {0, 0, 0, 0, 0, 0};
    if (i < a.n) {
        P4 pi = pos[i];
        R fx = 0, fy = 0, fz = 0;
        double pe = 0, vir = 0;
        const int *row = a.nbr + ((size_t)(i >> 5) * a.cap) * 32 + (i & 31);
        const int nsp = g_pair.nsp;
        // (a ghost has an empty row: nnb == 0)
        if (TAB) pair_row<R, EV, STRESS, false, true>(pos, row, a.nnb[i], pi, a.box, tab, nsp, fx, fy, fz, pe, vir, st);
        else if (nsp <= 2) pair_row<R, EV, STRESS, true, false>(pos, row, a.nnb[i], pi, a.box, tab, nsp, fx, fy, fz, pe, vir, st);
        else pair_row<R, EV, STRESS, false, false>(pos, row, a.nnb[i], pi, a.box, tab, nsp, fx, fy, fz, pe, vir, st);
        a.frc[i] = make4(fx, fy, fz, (R)(0.5 * pe));
        if (EV) a.vir[i] = (R)(0.5 * vir);
    }
    if (STRESS) {
        __shared__ double s_st[FORCE_THREADS / 32][6];
        int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
        for (int q = 0; q < 6; q++) {
            double v = warp_sum(st[q]);
            if (lane == 0) s_st[w][q] = v;
        }
        __syncthreads();
        if (threadIdx.x < 6) {
            double v = 0;
            for (int ww = 0; ww < FORCE_THREADS / 32; ww++) v += s_st[ww][threadIdx.x];
            a.partial[(size_t)blockIdx.x * 6 + threadIdx.x] = 0.5 * v;
        }
    }
}

template <typename R> struct StepArgs {
    Ctl *ctl;
    typename Pos4<R>::type *pos[2];
    typename Vec4<R>::type *vel;     // (vx, vy, vz, mass)
    const int *nbr, *nnb, *orig;
    int *img;                        // [3 n] box crossings, by ORIGINAL id
    int n, cap;
    R kick, drift;                   // v += kick F/m ; x += drift v
    Box<R> box;
    // domain decomposition: the step's bookkeeping waits for the GLOBAL max |v|^2
    // (k_dd_commit); this kernel only publishes the local one
    int dd;
    float *vmax_out;
    // peer-memory mode (b2_peer_*): the local max is also stored, tagged with the number of the
    // step (peer_state[0] + 1), straight into every rank's mailbox over NVLink
    unsigned long long *const *vmax_mail;   // [nranks] device pointers: rank r's slot for OUR value
    const unsigned *peer_state;             // see k_peer_commit
    int nranks;
    // the last block of the step kernel also WAITS for all ranks' values and settles the step (what k_peer_commit
    // does as a launch of its own): peer_commit_state != nullptr
    unsigned *peer_commit_state;            // LOCAL [steps taken, halo wanted]
    const unsigned long long *peer_mail_in; // LOCAL mailbox: slot r holds rank r's value
    float *peer_gv_out;
    double peer_drift;
    // ... and the step kernel PUSHES the halo itself: a particle on a send list stores its new position straight
    // into the neighbour's inbox over NVLink while the rest of the grid is still computing; the last block raises
    // the arrival flags (what k_peer_push does as a launch of its own).  halo_map[slot] = 2 q + side for entry q
    // of the left (0) / right (1) send list, -1 for everyone else; rebuilt with the list.
    const int *halo_map;
    double *halo_out_l[2], *halo_out_r[2];  // REMOTE inboxes by step parity
    unsigned long long *halo_flag_l, *halo_flag_r;
    double *ke_partial;              // [gridDim] sum m v^2 per block (thermostatted runs)
};

// mail[r] points at OUR slot in rank r's mailbox (parity 0); the parity-1 slot sits 8 * nranks
// words further (64-byte slots).  A rank that is not our neighbour may run one step ahead of
// us, hence the two parities.
__device__ __forceinline__ void peer_publish(unsigned long long *const *mail, int nranks, unsigned seq, float v)
{
    const unsigned long long packed = ((unsigned long long)seq << 32) | (unsigned long long)__float_as_uint(v);
    const size_t off = (size_t)(seq & 1u) * 8u * (size_t)nranks;
    __threadfence_system();
    for (int r = 0; r < nranks; r++) *(volatile unsigned long long *)(mail[r] + off) = packed;
}

#define PEER_TIMEOUT_CYCLES 20000000000LL      // ~10 s

// One thread: wait for every rank's max |v|^2 of step `seq` in the local mailbox, then the step's bookkeeping with
// the global maximum (all ranks decide alike).  Called by the last block of the step kernel; k_peer_commit is the
// same protocol as a kernel of one warp.
__device__ __forceinline__ void peer_commit_one(Ctl *ctl, unsigned *state, const unsigned long long *mail, int nranks,
                                                double drift, float *gv_out)
{
    const unsigned seq = state[0] + 1u;
    unsigned bits = 0u;
    bool ok = !ctl->peer_error;
    for (int r = 0; r < nranks && ok; r++) {
        const volatile unsigned long long *slot = mail + (size_t)(seq & 1u) * 8u * nranks + 8 * r;   // 64-byte slots
        const long long t0 = clock64();
        unsigned long long v;
        while (((v = *slot) >> 32) < (unsigned long long)seq) {
            if (clock64() - t0 > PEER_TIMEOUT_CYCLES) { ctl->peer_error = 1u; ok = false; break; }
            __nanosleep(40);
        }
        if ((v >> 32) == (unsigned long long)seq) bits = max(bits, (unsigned)(v & 0xffffffffu));
        else if (ok) { ctl->peer_error = 2u; ok = false; }            // a rank ran ahead of the protocol
    }
    const float gv = __uint_as_float(bits);                           // non-negative floats order like their bits
    *gv_out = gv;
    unsigned go = 0u;
    if (!ctl->peer_error) {
        double bound = ctl->disp_bound + fabs(drift) * sqrt((double)gv) * (1.0 + 1e-6);
        ctl->disp_bound = bound;
        ctl->cur ^= 1u;
        long long s = ctl->steps_done + 1;
        ctl->steps_done = s;
        state[0] = seq;
        if (s >= ctl->steps_target) ctl->done = 1u;
        // a refresh re-selects and re-sends the ghosts anyway: no halo after the step that asks for one
        if (bound > ctl->half_skin) ctl->halt_rebuild = 1u; else go = 1u;
    }
    state[1] = go;
    __threadfence();
}

// the new position of slot i, if it is on a send list, into the neighbour's inbox of this step's parity
template <typename R>
__device__ __forceinline__ bool peer_push_one(const StepArgs<R> &a, int i, const typename Pos4<R>::type &pnew)
{
    const int hm = a.halo_map[i];
    if (hm < 0) return false;
    const unsigned seq = a.peer_state[0] + 1u;
    double x, y, z;
    to_double(pnew, a.box, x, y, z);
    double *dst = ((hm & 1) ? a.halo_out_r[seq & 1u] : a.halo_out_l[seq & 1u]) + 3 * (size_t)(hm >> 1);
    // stores into the neighbour's HBM; ordered before this block's ticket by the block barrier and thread 0's
    // system-wide fence in step_epilogue (a fence of its own here stalls every warp that holds a boundary particle
    // until ALL its stores are visible system-wide: 0.297 ms per step against 0.234)
    dst[0] = x; dst[1] = y; dst[2] = z;
    return true;
}

// What every block of a step kernel does after its particles moved: block maxima / sums, the
// ticket, and - in the last block to finish - the step's bookkeeping (see k_step_nve).
template <typename R, bool NVT, int THREADS>
__device__ __forceinline__ void step_epilogue(const StepArgs<R> &a, Ctl *ctl, unsigned cur, float v2, double mv2,
                                              bool pushed = false)
{
    __shared__ float s_v2[THREADS / 32];
    __shared__ bool s_last;
    // (the windowed kernels run on a 3-D grid of super-blocks)
    const unsigned nblk = gridDim.x * gridDim.y * gridDim.z;
    const unsigned bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    __shared__ double s_ke[THREADS / 32];
    if (NVT) {
        mv2 = warp_sum(mv2);
        if ((threadIdx.x & 31) == 0) s_ke[threadIdx.x >> 5] = mv2;
    }
    v2 = warp_max(v2);
    if ((threadIdx.x & 31) == 0) s_v2[threadIdx.x >> 5] = v2;
    const bool block_pushed = __syncthreads_or(pushed) != 0;        // (only blocks at a face of the slab)
    if (threadIdx.x == 0) {
        float m = s_v2[0];
        for (int w = 1; w < THREADS / 32; w++) m = fmaxf(m, s_v2[w]);
        if (NVT) {
            double k = 0;
            for (int w = 0; w < THREADS / 32; w++) k += s_ke[w];
            a.ke_partial[bid] = k;
        }
        atomicMax(&ctl->vmax2_bits, __float_as_uint(m));
        if (block_pushed) __threadfence_system(); else __threadfence();  // (this block's halo stores, seen by thread 0 after the barrier)
        unsigned ticket = atomicAdd(&ctl->blocks_done, 1u);
        s_last = (ticket == nblk - 1);
        if (s_last && a.dd) {
            __threadfence();
            const float vm2 = __uint_as_float(atomicExch(&ctl->vmax2_bits, 0u));
            *a.vmax_out = vm2;
            ctl->blocks_done = 0u;
            if (a.vmax_mail) peer_publish(a.vmax_mail, a.nranks, a.peer_state[0] + 1u, vm2);
            if (a.halo_map) {                // every block's halo stores are out (ticket): tell both neighbours
                __threadfence_system();
                *(volatile unsigned long long *)a.halo_flag_l = (unsigned long long)(a.peer_state[0] + 1u);
                *(volatile unsigned long long *)a.halo_flag_r = (unsigned long long)(a.peer_state[0] + 1u);
            }
            if (a.vmax_mail && a.peer_commit_state)
                peer_commit_one(ctl, a.peer_commit_state, a.peer_mail_in, a.nranks, a.peer_drift, a.peer_gv_out);
        } else if (s_last) {
            __threadfence();
            float vm2 = __uint_as_float(atomicExch(&ctl->vmax2_bits, 0u));
            double bound = ctl->disp_bound + fabs((double)a.drift) * sqrt((double)vm2) * (1.0 + 1e-6);
            ctl->disp_bound = bound;
            ctl->cur = cur ^ 1u;
            long long s = ctl->steps_done + 1;
            ctl->steps_done = s;
            if (s >= ctl->steps_target) ctl->done = 1u;
            if (bound > ctl->half_skin) ctl->halt_rebuild = 1u;
            ctl->blocks_done = 0u;
        }
    }
    if (NVT) {
        // the last block (every partial is visible to it after the ticket) renews the thermostat
        __syncthreads();
        if (s_last) {
            double k = 0;
            for (int b = threadIdx.x; b < (int)nblk; b += blockDim.x) k += __ldcg(a.ke_partial + b);
            k = warp_sum(k);
            __syncthreads();
            if ((threadIdx.x & 31) == 0) s_ke[threadIdx.x >> 5] = k;
            __syncthreads();
            if (threadIdx.x == 0) {
                double tot = 0;
                for (int w = 0; w < THREADS / 32; w++) tot += s_ke[w];
                const double T = tot / ctl->th_ndof, dt = ctl->th_dt;
                ctl->th_T = T;
                if (ctl->th_kind == 0) {
                    const double l2 = 1.0 + dt / ctl->th_tau * (ctl->th_T0 / fmax(T, 1e-300) - 1.0);
                    ctl->th_a = ctl->th_b = sqrt(fmax(l2, 0.25));
                } else {
                    const double xi = ctl->th_xi + dt / (ctl->th_tau * ctl->th_tau) * (T / ctl->th_T0 - 1.0);
                    const double h = 0.5 * xi * dt;
                    ctl->th_xi = xi;
                    ctl->th_a = (1.0 - h) / (1.0 + h);
                    ctl->th_b = 1.0 / (1.0 + h);
                }
            }
        }
    }
}

// The fused velocity-Verlet step (kick-drift form): F(x) -> v += dt F/m ->
// x += dt v -> fold into the cell (particle.py:90-99) -> max |v|^2 for the
// displacement bound.  Reads the current position buffer, writes the other;
// the last block to finish flips the buffers and settles the step's
// bookkeeping on the device.
//
// NVT = true adds a thermostat at no extra pass over the data (the reference's
// backends honour system.thermostat, atooms/backends/lammps.py:229-241): the kick becomes
// v' = a v + b dt F/m with (a, b) read from the control block, each block leaves its sum of
// m v'^2, and the last block turns the total into the next step's (a, b):
//   Berendsen      lambda^2 = 1 + dt/tau (T0/T - 1),               a = b = lambda
//   Nose-Hoover    xi += dt/tau^2 (T/T0 - 1), h = xi dt/2,           a = (1-h)/(1+h), b = 1/(1+h)
// (the leap-frog form used by RUMD's IntegratorNVT); T is the kinetic temperature of the
// half-step velocities, (N-1) d degrees of freedom (atooms/system/system.py:164-179).
template <typename R, bool NVT, bool TAB>
__global__ void __launch_bounds__(FORCE_THREADS, (sizeof(R) == 4 ? B2_MINBLK32 : B2_MINBLK64))
k_step_nve(StepArgs<R> a)
{
    typedef typename Vec4<R>::type V4;
    typedef typename Pos4<R>::type P4;
    Ctl *ctl = a.ctl;
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) {
        // a no-op step moves nothing; with the peer-memory halo it also sends nothing: every rank
        // takes the same decision, and the mail sequence counts the steps actually taken
        if (a.dd && blockIdx.x == 0 && threadIdx.x == 0) {
            *a.vmax_out = 0.f;
            if (a.peer_commit_state) { a.peer_commit_state[1] = 0u; *a.peer_gv_out = 0.f; }     // (as k_peer_commit)
        }
        return;
    }
    __shared__ PairTab<R> tab;
    load_pair_table(tab);
    __syncthreads();
    const unsigned cur = ctl->cur;
    const P4 *__restrict__ pos = cur ? a.pos[1] : a.pos[0];
    P4 *__restrict__ posn = cur ? a.pos[0] : a.pos[1];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float v2 = 0.f;
    double mv2 = 0.0;
    bool pushed = false;
    const R th_a = NVT ? (R)ctl->th_a : (R)1, th_b = NVT ? (R)ctl->th_b : (R)1;
    if (i < a.n) {
        P4 pi = pos[i];
        if (is_ghost(pi)) {
            posn[i] = pi;            // overwritten by the halo exchange before the next step
        } else {
        R fx = 0, fy = 0, fz = 0;
        double pe = 0, vir = 0;
        const int *row = a.nbr + ((size_t)(i >> 5) * a.cap) * 32 + (i & 31);
        const int nsp = g_pair.nsp;
        if (TAB) pair_row<R, false, false, false, true>(pos, row, a.nnb[i], pi, a.box, tab, nsp, fx, fy, fz, pe, vir, nullptr);
        else if (nsp <= 2) pair_row<R, false, false, true, false>(pos, row, a.nnb[i], pi, a.box, tab, nsp, fx, fy, fz, pe, vir, nullptr);
        else pair_row<R, false, false, false, false>(pos, row, a.nnb[i], pi, a.box, tab, nsp, fx, fy, fz, pe, vir, nullptr);
        V4 v = a.vel[i];
        R s = a.kick / v.w;
        if (NVT) {
            s *= th_b;
            v.x = th_a * v.x + s * fx; v.y = th_a * v.y + s * fy; v.z = th_a * v.z + s * fz;
            mv2 = (double)v.w * ((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
        } else {
            v.x += s * fx; v.y += s * fy; v.z += s * fz;
        }
        int kx, ky, kz;
        const P4 pnew = advance(pi, a.drift * v.x, a.drift * v.y, a.drift * v.z, a.box, kx, ky, kz);
        posn[i] = pnew;
        if (a.halo_map) pushed = peer_push_one(a, i, pnew);
        if (kx | ky | kz) {
            int *im = a.img + 3 * (size_t)a.orig[i];
            im[0] += kx; im[1] += ky; im[2] += kz;
        }
        a.vel[i] = v;
        // rounded up: the bound must never be smaller than the true speed
        v2 = __double2float_ru((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
        }
    }
    step_epilogue<R, NVT, FORCE_THREADS>(a, ctl, cur, v2, mv2, pushed);
}

__global__ void k_thermostat_begin(Ctl *ctl, int kind, double T0, double tau, double dt, double ndof)
{
    ctl->th_kind = kind;
    ctl->th_T0 = T0;
    ctl->th_tau = tau;
    ctl->th_dt = dt;
    ctl->th_ndof = ndof;
    // xi carries over between calls (reset by b2_upload); the first step uses the coefficients it implies
    const double h = kind == 1 ? 0.5 * ctl->th_xi * dt : 0.0;
    ctl->th_a = (1.0 - h) / (1.0 + h);
    ctl->th_b = 1.0 / (1.0 + h);
}

// v += c F / m   (opening / closing half kick of a run)
template <typename R>
__global__ void k_kick(typename Vec4<R>::type *__restrict__ vel,
                       const typename Vec4<R>::type *__restrict__ frc, int n, R c)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Vec4<R>::type v = vel[i], f = frc[i];
    R s = c / v.w;
    v.x += s * f.x; v.y += s * f.y; v.z += s * f.z;
    vel[i] = v;
}

template <typename R>
__global__ void k_scale_vel(typename Vec4<R>::type *__restrict__ vel, int n, R c)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Vec4<R>::type v = vel[i];
    v.x *= c; v.y *= c; v.z *= c;
    vel[i] = v;
}

// ---- reductions: per-block partials in scratch, one block sums them in a
// fixed order (deterministic totals, SURVEY.md §7 "summation order") ----------
template <int NQ>
__device__ __forceinline__ void block_partials(double (&q)[NQ], double *__restrict__ partial)
{
    __shared__ double s_q[RED_THREADS / 32][NQ];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NQ; k++) {
        double v = warp_sum(q[k]);
        if (lane == 0) s_q[w][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < NQ) {
        double v = 0;
        for (int ww = 0; ww < RED_THREADS / 32; ww++) v += s_q[ww][threadIdx.x];
        partial[(size_t)blockIdx.x * NQ + threadIdx.x] = v;
    }
}

template <typename R>
__global__ void __launch_bounds__(RED_THREADS)
k_thermo_partial(const Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0,
                 const typename Pos4<R>::type *pos1,
                 const typename Vec4<R>::type *__restrict__ vel,
                 const typename Vec4<R>::type *__restrict__ frc, const R *__restrict__ vir,
                 const int *__restrict__ orig, const int *__restrict__ img,
                 const double *__restrict__ ref, int n, Box<R> box, double *__restrict__ partial)
{
    typedef typename Vec4<R>::type V4;
    typedef typename Pos4<R>::type P4;
    const P4 *__restrict__ pos = ctl->cur ? pos1 : pos0;
    double q[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        V4 v = vel[i], f = frc[i];
        P4 p = pos[i];
        if (is_ghost(p)) continue;
        double rx, ry, rz;
        to_double(p, box, rx, ry, rz);
        double m = (double)v.w;
        double vx = v.x, vy = v.y, vz = v.z, fx = f.x, fy = f.y, fz = f.z;
        q[Q_KIN] += 0.5 * m * ((vx * vx + vy * vy) + vz * vz);
        q[Q_POT] += (double)f.w;
        q[Q_VIR] += (double)vir[i];
        q[Q_FF] += (fx * fx + fy * fy) + fz * fz;
        q[Q_PX] += m * vx; q[Q_PY] += m * vy; q[Q_PZ] += m * vz;
        q[Q_MASS] += m;
        size_t o = 3 * (size_t)orig[i];
        double ux = (rx + (double)img[o + 0] * box.Ld[0]) - ref[o + 0];
        double uy = (ry + (double)img[o + 1] * box.Ld[1]) - ref[o + 1];
        double uz = (rz + (double)img[o + 2] * box.Ld[2]) - ref[o + 2];
        q[Q_MSD] += (ux * ux + uy * uy) + uz * uz;
    }
    block_partials<9>(q, partial);
}

// sum `nblk` rows of `nq` partials into out[0..nq)
__global__ void k_sum_partials(const double *__restrict__ partial, int nblk, int nq,
                               double *__restrict__ out)
{
    int q = threadIdx.x;
    if (q >= nq) return;
    double v = 0;
    for (int b = 0; b < nblk; b++) v += partial[(size_t)b * nq + q];
    out[q] = v;
}

// ---- minimisers -------------------------------------------------------------
// Shared by FIRE and steepest descent: P = F.v, |v|^2, |F|^2, max_i |F_i|^2.
template <typename R>
__global__ void __launch_bounds__(RED_THREADS)
k_min_partial(const Ctl *__restrict__ ctl, const typename Vec4<R>::type *__restrict__ vel,
              const typename Vec4<R>::type *__restrict__ frc, int n, double *__restrict__ partial)
{
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) return;
    typedef typename Vec4<R>::type V4;
    double q[4] = {0, 0, 0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        V4 v = vel[i], f = frc[i];
        double vx = v.x, vy = v.y, vz = v.z, fx = f.x, fy = f.y, fz = f.z;
        double f2 = (fx * fx + fy * fy) + fz * fz;
        q[0] += (fx * vx + fy * vy) + fz * vz;
        q[1] += (vx * vx + vy * vy) + vz * vz;
        q[2] += f2;
        q[3] = fmax(q[3], f2);
    }
    // sums for 0..2, max for 3
    __shared__ double s_q[RED_THREADS / 32][4];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 3; k++) q[k] = warp_sum(q[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q[3] = fmax(q[3], __shfl_xor_sync(0xffffffffu, q[3], o));
    if (lane == 0)
        for (int k = 0; k < 4; k++) s_q[w][k] = q[k];
    __syncthreads();
    if (threadIdx.x < 4) {
        double v = 0;
        for (int ww = 0; ww < RED_THREADS / 32; ww++)
            v = threadIdx.x < 3 ? v + s_q[ww][threadIdx.x] : fmax(v, s_q[ww][threadIdx.x]);
        partial[(size_t)blockIdx.x * 4 + threadIdx.x] = v;
    }
}

#define MIN_MODE_SD 0
#define MIN_MODE_FIRE 1

// One thread: finish the reduction and advance the minimiser's scalar state.
// FIRE: Bitzek et al., PRL 97, 170201 (2006): n_min 5, f_inc 1.1, f_dec 0.5,
// alpha0 0.1, f_alpha 0.99.  SD: step h = min(h_now, 0.1 / max|F_i|), h_now
// grows 1.1x while successive gradients agree (F.F_prev > 0), halves otherwise.
__global__ void k_min_logic(Ctl *ctl, const double *__restrict__ partial, int nblk, int n)
{
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) return;
    double P = 0, vv = 0, ff = 0, fmax2 = 0;
    for (int b = 0; b < nblk; b++) {
        P += partial[4 * (size_t)b + 0];
        vv += partial[4 * (size_t)b + 1];
        ff += partial[4 * (size_t)b + 2];
        fmax2 = fmax(fmax2, partial[4 * (size_t)b + 3]);
    }
    ctl->fn2 = ff / (double)n;
    if (ctl->fn2 < ctl->fn2_tol) {
        ctl->converged = 1u;
        ctl->done = 1u;
        return;
    }
    if (ctl->iter >= ctl->steps_target) {
        ctl->done = 1u;
        return;
    }
    if (ctl->min_mode == MIN_MODE_FIRE) {
        if (P > 0.0) {
            // mixing coefficients use alpha BEFORE this iteration's update
            ctl->red[0] = 1.0 - ctl->alpha;
            ctl->red[1] = ctl->alpha * sqrt(vv / ff);
            ctl->npos += 1;
            if (ctl->npos > 5) {
                ctl->dt_now = fmin(ctl->dt_now * 1.1, ctl->dt_max);
                ctl->alpha *= 0.99;
            }
        } else {
            ctl->red[0] = 0.0;
            ctl->red[1] = 0.0;
            ctl->npos = 0;
            ctl->dt_now *= 0.5;
            ctl->alpha = 0.1;
        }
        ctl->step_len = ctl->dt_now;
    } else {
        if (ctl->has_prev) ctl->dt_now *= (P > 0.0) ? 1.1 : 0.5;
        ctl->has_prev = 1;
        ctl->step_len = fmin(ctl->dt_now, 0.1 / sqrt(fmax2));
        ctl->red[2] = sqrt(fmax2);
    }
}

template <typename R> struct MinArgs {
    Ctl *ctl;
    typename Pos4<R>::type *pos[2];
    typename Vec4<R>::type *vel;
    const typename Vec4<R>::type *frc;
    const int *orig;
    int *img;
    int n;
    Box<R> box;
};

// FIRE: v = a v + b F ; v += dt F/m ; x += dt v.   SD: x += h F ; v = F.
// Positions are updated in place (forces come from the stored array).
template <typename R>
__global__ void __launch_bounds__(RED_THREADS)
k_min_step(MinArgs<R> a)
{
    typedef typename Vec4<R>::type V4;
    Ctl *ctl = a.ctl;
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) return;
    __shared__ float s_d2[RED_THREADS / 32];
    typedef typename Pos4<R>::type P4;
    P4 *__restrict__ pos = ctl->cur ? a.pos[1] : a.pos[0];
    const int mode = ctl->min_mode;
    const R h = (R)ctl->step_len, ca = (R)ctl->red[0], cb = (R)ctl->red[1];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float d2 = 0.f;
    if (i < a.n) {
        P4 p = pos[i];
        V4 v = a.vel[i], f = a.frc[i];
        R ux, uy, uz;
        if (mode == MIN_MODE_FIRE) {
            R s = h / v.w;
            v.x = ca * v.x + cb * f.x + s * f.x;
            v.y = ca * v.y + cb * f.y + s * f.y;
            v.z = ca * v.z + cb * f.z + s * f.z;
            ux = h * v.x; uy = h * v.y; uz = h * v.z;
        } else {
            ux = h * f.x; uy = h * f.y; uz = h * f.z;
            v.x = f.x; v.y = f.y; v.z = f.z;
        }
        int kx, ky, kz;
        pos[i] = advance(p, ux, uy, uz, a.box, kx, ky, kz);
        if (kx | ky | kz) {
            int *im = a.img + 3 * (size_t)a.orig[i];
            im[0] += kx; im[1] += ky; im[2] += kz;
        }
        a.vel[i] = v;
        d2 = __double2float_ru((double)ux * ux + (double)uy * uy + (double)uz * uz);
    }
    d2 = warp_max(d2);
    if ((threadIdx.x & 31) == 0) s_d2[threadIdx.x >> 5] = d2;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = s_d2[0];
        for (int w = 1; w < RED_THREADS / 32; w++) m = fmaxf(m, s_d2[w]);
        atomicMax(&ctl->vmax2_bits, __float_as_uint(m));
        __threadfence();
        unsigned ticket = atomicAdd(&ctl->blocks_done, 1u);
        if (ticket == gridDim.x - 1) {
            __threadfence();
            float dm2 = __uint_as_float(atomicExch(&ctl->vmax2_bits, 0u));
            double bound = ctl->disp_bound + sqrt((double)dm2) * (1.0 + 1e-6);
            ctl->disp_bound = bound;
            ctl->iter += 1;
            ctl->steps_done += 1;
            if (bound > ctl->half_skin) ctl->halt_rebuild = 1u;
            ctl->blocks_done = 0u;
        }
    }
}

// Domain decomposition: settle a step once the global max |v|^2 is known (the
// same number on every rank, so every rank takes the same decisions).
__global__ void k_dd_commit(Ctl *ctl, const float *__restrict__ gvmax2, double drift)
{
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) return;
    double bound = ctl->disp_bound + fabs(drift) * sqrt((double)*gvmax2) * (1.0 + 1e-6);
    ctl->disp_bound = bound;
    ctl->cur ^= 1u;
    long long s = ctl->steps_done + 1;
    ctl->steps_done = s;
    if (s >= ctl->steps_target) ctl->done = 1u;
    if (bound > ctl->half_skin) ctl->halt_rebuild = 1u;
}

// halo exchange: positions by LOCAL id (upload order), as 3 doubles relative to the cell centre
template <typename R>
__global__ void k_dd_pack(const Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0,
                          const typename Pos4<R>::type *pos1, const int *__restrict__ inv,
                          const int *__restrict__ ids, int n, Box<R> box, double *__restrict__ out)
{
    const typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    double x, y, z;
    to_double(pos[inv[ids[k]]], box, x, y, z);
    out[3 * k] = x; out[3 * k + 1] = y; out[3 * k + 2] = z;
}

template <typename R>
__global__ void k_dd_unpack(const Ctl *__restrict__ ctl, typename Pos4<R>::type *pos0,
                            typename Pos4<R>::type *pos1, const int *__restrict__ inv, int first, int n,
                            Box<R> box, const double *__restrict__ in)
{
    typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int s = inv[first + k];
    pos[s] = encode_pos(in[3 * k], in[3 * k + 1], in[3 * k + 2], praw(pos[s]), box);
}

__global__ void k_inverse(const Ctl *__restrict__ ctl, int guarded, const int *__restrict__ orig, int n,
                          int *__restrict__ inv)
{
    if (guarded && !ctl->halt_rebuild) return;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) inv[orig[s]] = s;
}

// ---------------------------------------------------------------------------
// Peer-memory halo (b2_peer_*): the per-step exchange without a communication
// library.  Every rank owns a mailbox block in its own HBM, mapped into its
// peers with CUDA IPC; senders store into it over NVLink, receivers spin on a
// sequence number.  Per step:   step kernel (publishes max|v|^2 to all ranks)
//   -> k_peer_commit  (waits for all ranks' values: the 4-byte all-reduce)
//   -> k_peer_push    (gathers boundary positions INTO the neighbours' inboxes)
//   -> k_peer_pull    (waits for both neighbours, scatters the inbox into the ghosts)
// Inboxes are double-buffered on the step parity: a neighbour can run at most
// one step ahead.  Every wait has a time-out that raises Ctl::peer_error.
// The sequence number is the count of steps TAKEN, kept on the device (state[0], which
// outlives re-uploads; state[1] tells push/pull whether the last commit wants a halo):
// a launch that finds the list invalid or the budget spent sends and awaits nothing
// (all ranks decide alike, from the same global max), and the step that asks for a
// refresh skips its halo, which the refresh replaces anyway.
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool peer_wait_ge(const volatile unsigned long long *flag, unsigned long long want, Ctl *ctl)
{
    const long long t0 = clock64();
    while (*flag < want) {
        if (clock64() - t0 > PEER_TIMEOUT_CYCLES) { ctl->peer_error = 1u; return false; }
        __nanosleep(100);
    }
    __threadfence_system();
    return true;
}

__global__ void k_peer_commit(Ctl *ctl, unsigned *state, const unsigned long long *mail, int nranks, double drift,
                              float *gv_out)
{
    // one warp; lane r waits for rank r's value of this step
    const int lane = threadIdx.x;
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) {              // the step kernel did nothing either
        if (lane == 0) { state[1] = 0u; *gv_out = 0.f; }
        return;
    }
    const unsigned seq = state[0] + 1u;
    unsigned bits = 0u;
    bool ok = true;
    if (ctl->peer_error) ok = false;
    for (int r = lane; r < nranks && ok; r += 32) {
        const volatile unsigned long long *slot = mail + (size_t)(seq & 1u) * 8u * nranks + 8 * r;   // 64-byte slots
        const long long t0 = clock64();
        unsigned long long v;
        while (((v = *slot) >> 32) < (unsigned long long)seq) {
            if (clock64() - t0 > PEER_TIMEOUT_CYCLES) { ctl->peer_error = 1u; ok = false; break; }
            __nanosleep(100);
        }
        if ((v >> 32) == (unsigned long long)seq) bits = max(bits, (unsigned)(v & 0xffffffffu));
        else if (ok) { ctl->peer_error = 2u; ok = false; }            // a neighbour ran ahead of the protocol
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bits = max(bits, __shfl_xor_sync(0xffffffffu, bits, o));
    if (lane == 0) {
        const float gv = __uint_as_float(bits);                       // non-negative floats order like their bits
        *gv_out = gv;
        unsigned go = 0u;
        if (!ctl->peer_error) {
            double bound = ctl->disp_bound + fabs(drift) * sqrt((double)gv) * (1.0 + 1e-6);
            ctl->disp_bound = bound;
            ctl->cur ^= 1u;
            long long s = ctl->steps_done + 1;
            ctl->steps_done = s;
            state[0] = seq;
            if (s >= ctl->steps_target) ctl->done = 1u;
            // a refresh re-selects and re-sends the ghosts anyway: no halo after the step that asks for one
            if (bound > ctl->half_skin) ctl->halt_rebuild = 1u; else go = 1u;
        }
        state[1] = go;
    }
}

template <typename R> struct PeerArgs {
    Ctl *ctl;
    typename Pos4<R>::type *pos[2];
    const int *inv;
    const int *send_l, *send_r;
    int n_send_l, n_send_r;
    int first_l, n_ghost_l, first_r, n_ghost_r;      // local ids of the ghosts
    Box<R> box;
    double *out_l[2], *out_r[2];                     // REMOTE: the neighbours' inboxes, by step parity
    unsigned long long *flag_l, *flag_r;             // REMOTE: the neighbours' arrival flags
    const double *in_l[2], *in_r[2];                 // LOCAL inboxes, by step parity
    const unsigned long long *my_flag_l, *my_flag_r; // LOCAL arrival flags (from the left / right neighbour)
    unsigned *ticket;                                // LOCAL counter
    const unsigned *state;                           // LOCAL [steps taken, halo wanted] (k_peer_commit)
};

// slot -> entry of a send list (StepArgs::halo_map), after a memset to -1
__global__ void k_halo_map(const int *__restrict__ inv, const int *__restrict__ send_l, int n_l,
                           const int *__restrict__ send_r, int n_r, int *__restrict__ map)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_l + n_r) return;
    const bool left = k < n_l;
    const int q = left ? k : k - n_l;
    map[inv[left ? send_l[q] : send_r[q]]] = 2 * q + (left ? 0 : 1);
}

template <typename R>
__global__ void __launch_bounds__(256)
k_peer_push(PeerArgs<R> a)
{
    if (!a.state[1]) return;
    const unsigned seq = a.state[0];
    const typename Pos4<R>::type *pos = a.ctl->cur ? a.pos[1] : a.pos[0];
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = a.n_send_l + a.n_send_r;
    if (k < total) {
        const bool left = k < a.n_send_l;
        const int q = left ? k : k - a.n_send_l;
        const int id = left ? a.send_l[q] : a.send_r[q];
        double x, y, z;
        to_double(pos[a.inv[id]], a.box, x, y, z);
        double *dst = (left ? (seq & 1u ? a.out_l[1] : a.out_l[0]) : (seq & 1u ? a.out_r[1] : a.out_r[0])) + 3 * (size_t)q;
        dst[0] = x; dst[1] = y; dst[2] = z;                          // stores into the neighbour's HBM
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(a.ticket, 1u);
        if (t == gridDim.x - 1) {                                     // everything of this rank is on its way
            *a.ticket = 0u;
            __threadfence_system();
            *(volatile unsigned long long *)a.flag_l = (unsigned long long)seq;
            *(volatile unsigned long long *)a.flag_r = (unsigned long long)seq;
        }
    }
}

template <typename R>
__global__ void __launch_bounds__(256)
k_peer_pull(PeerArgs<R> a)
{
    __shared__ bool s_ok;
    if (!a.state[1]) return;
    const unsigned seq = a.state[0];
    if (threadIdx.x == 0)
        s_ok = !a.ctl->peer_error &&
               peer_wait_ge((const volatile unsigned long long *)a.my_flag_l, seq, a.ctl) &&
               peer_wait_ge((const volatile unsigned long long *)a.my_flag_r, seq, a.ctl);
    __syncthreads();
    if (!s_ok) return;
    typename Pos4<R>::type *pos = a.ctl->cur ? a.pos[1] : a.pos[0];
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= a.n_ghost_l + a.n_ghost_r) return;
    const bool left = k < a.n_ghost_l;
    const int q = left ? k : k - a.n_ghost_l;
    const double *src = (left ? (seq & 1u ? a.in_l[1] : a.in_l[0]) : (seq & 1u ? a.in_r[1] : a.in_r[0])) + 3 * (size_t)q;
    const int s = a.inv[(left ? a.first_l : a.first_r) + q];
    // written by another GPU: read around L1
    pos[s] = encode_pos(__ldcg(src), __ldcg(src + 1), __ldcg(src + 2), praw(pos[s]), a.box);
}
