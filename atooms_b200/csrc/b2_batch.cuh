// Ensembles of small independent systems ("replicas"): one thread block per
// replica, the whole configuration in shared memory / registers, all pairs with
// the minimum image, and the complete minimisation loop inside one kernel
// launch — no host round trip, no neighbour list (a replica of ~150 particles
// in a box of two cut-offs has nothing to skip).
//
// Serves BASELINE.json configs[3] (4096 Kob-Andersen N = 150 replicas minimised
// with FIRE / steepest descent); the iteration is the one of k_min_logic /
// k_min_step (and of the oracle), evaluated redundantly by every thread from
// block-wide sums taken in a fixed order.
#pragma once
#include "b2_common.cuh"

#define BATCH_MAX_THREADS 1024

template <typename R> struct BatchArgs {
    typename Pos4<R>::type *pos;        // [nrep][n]
    const typename Vec4<R>::type *vel;  // [nrep][n]: .w is the mass
    int n, nrep;
    Box<R> box;
    int method;                         // MIN_MODE_FIRE / MIN_MODE_SD
    long long max_steps;
    double dt, fn2_tol;
    long long *steps_done;              // [nrep]
    int *converged;                     // [nrep]
    double *energy, *fn2, *virial;      // [nrep]
    typename Vec4<R>::type *frc;        // [nrep][n] forces of the final configuration (+ pe_i)
};

// block-wide sum / max over up to 32 warps, identical order in every replica
__device__ __forceinline__ double block_sum(double v, double *s_red)
{
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) s_red[w] = v;
    __syncthreads();
    double t = 0;
    for (int k = 0; k < nw; k++) t += s_red[k];
    return t;
}
__device__ __forceinline__ double block_max(double v, double *s_red)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) s_red[w] = v;
    __syncthreads();
    double t = 0;
    for (int k = 0; k < nw; k++) t = fmax(t, s_red[k]);
    return t;
}

template <typename R>
__global__ void k_batch_minimize(BatchArgs<R> a)
{
    typedef typename Pos4<R>::type P4;
    typedef typename Vec4<R>::type V4;
    extern __shared__ __align__(16) unsigned char s_raw[];
    P4 *s_pos = reinterpret_cast<P4 *>(s_raw);
    __shared__ double s_red[32];
    __shared__ R s_c12x12[B2_MAXSP * B2_MAXSP], s_c6x6[B2_MAXSP * B2_MAXSP], s_rc2[B2_MAXSP * B2_MAXSP];
    for (int k = threadIdx.x; k < B2_MAXSP * B2_MAXSP; k += blockDim.x) {
        s_c12x12[k] = (R)(12.0 * g_pair.c12[k]);
        s_c6x6[k] = (R)(6.0 * g_pair.c6[k]);
        s_rc2[k] = (R)g_pair.rc2[k];
    }
    const int rep = blockIdx.x, i = threadIdx.x, n = a.n, nsp = g_pair.nsp;
    const bool own = i < n;
    P4 *gpos = a.pos + (size_t)rep * n;
    P4 p = own ? gpos[i] : P4();
    const R m = own ? a.vel[(size_t)rep * n + i].w : (R)1;
    const int ti = own ? ptype(p) * nsp : 0;
    R vx = 0, vy = 0, vz = 0;             // FIRE velocity, or the previous gradient for SD
    R fx = 0, fy = 0, fz = 0;
    double dt_now = a.dt, alpha = 0.1;
    const double dt_max = 10.0 * a.dt;
    int npos = 0, has_prev = 0, conv = 0;
    long long it = 0;
    double fn2 = 0;
    for (;;) {
        // ---- forces of the current configuration (all pairs, minimum image) ----
        __syncthreads();
        if (own) s_pos[i] = p;
        __syncthreads();
        fx = fy = fz = 0;
        if (own) {
            for (int j = 0; j < n; j++) {
                if (j == i) continue;
                const P4 q = s_pos[j];
                R dx, dy, dz;
                delta(p, q, a.box, dx, dy, dz);
                const R r2 = dx * dx + dy * dy + dz * dz;
                const int t = ti + ptype(q);
                const R r2inv = r_rcp(r2);
                const R r6inv = r2inv * r2inv * r2inv;
                R fp = r6inv * (s_c12x12[t] * r6inv - s_c6x6[t]) * r2inv;
                fp = r2 < s_rc2[t] ? fp : (R)0;
                fx += fp * dx; fy += fp * dy; fz += fp * dz;
            }
        }
        // ---- block-wide sums ----
        const double dfx = fx, dfy = fy, dfz = fz, dvx = vx, dvy = vy, dvz = vz;
        const double f2 = (dfx * dfx + dfy * dfy) + dfz * dfz;
        const double ff = block_sum(f2, s_red);
        const double P = block_sum((dfx * dvx + dfy * dvy) + dfz * dvz, s_red);
        fn2 = ff / (double)n;
        if (fn2 < a.fn2_tol) { conv = 1; break; }
        if (it >= a.max_steps) break;
        R ux, uy, uz;
        if (a.method == 1) {                                  // FIRE (Bitzek et al. 2006)
            const double vv = block_sum((dvx * dvx + dvy * dvy) + dvz * dvz, s_red);
            R ca, cb;
            if (P > 0.0) {
                ca = (R)(1.0 - alpha);
                cb = (R)(alpha * sqrt(vv / ff));
                npos += 1;
                if (npos > 5) { dt_now = fmin(dt_now * 1.1, dt_max); alpha *= 0.99; }
            } else {
                ca = cb = 0;
                npos = 0; dt_now *= 0.5; alpha = 0.1;
            }
            const R h = (R)dt_now, s = h / m;
            vx = ca * vx + cb * fx + s * fx;
            vy = ca * vy + cb * fy + s * fy;
            vz = ca * vz + cb * fz + s * fz;
            ux = h * vx; uy = h * vy; uz = h * vz;
        } else {                                              // adaptive steepest descent
            const double fmax2 = block_max(f2, s_red);
            if (has_prev) dt_now *= (P > 0.0) ? 1.1 : 0.5;
            has_prev = 1;
            const R h = (R)fmin(dt_now, 0.1 / sqrt(fmax2));
            ux = h * fx; uy = h * fy; uz = h * fz;
            vx = fx; vy = fy; vz = fz;
        }
        if (own) {
            int kx, ky, kz;
            p = advance(p, ux, uy, uz, a.box, kx, ky, kz);
        }
        it++;
    }
    // ---- final configuration: positions, forces, fp64 energy and virial ----
    double pe = 0, vir = 0;
    if (own) {
        for (int j = 0; j < n; j++) {
            if (j == i) continue;
            double ex, ey, ez;
            delta_d(p, s_pos[j], a.box, ex, ey, ez);
            const double q2 = ex * ex + ey * ey + ez * ez;
            const int t = ti + ptype(s_pos[j]);
            if (q2 < g_pair.rc2[t]) {
                const double q2inv = 1.0 / q2, q6inv = q2inv * q2inv * q2inv;
                pe += q6inv * (g_pair.c12[t] * q6inv - g_pair.c6[t]) - g_pair.ushift[t];
                vir += q6inv * (12.0 * g_pair.c12[t] * q6inv - 6.0 * g_pair.c6[t]);
            }
        }
        gpos[i] = p;
        a.frc[(size_t)rep * n + i] = make4(fx, fy, fz, (R)(0.5 * pe));
    }
    const double e_tot = block_sum(0.5 * pe, s_red);
    const double w_tot = block_sum(0.5 * vir, s_red);
    if (threadIdx.x == 0) {
        a.steps_done[rep] = it;
        a.converged[rep] = conv;
        a.energy[rep] = e_tot;
        a.virial[rep] = w_tot;
        a.fn2[rep] = fn2;
    }
}

// all-pairs force on particle i of a replica whose positions sit in shared memory
template <typename R>
__device__ __forceinline__ void batch_force(const typename Pos4<R>::type *s_pos, int n, int i,
                                            const typename Pos4<R>::type &p, int ti, const Box<R> &box,
                                            const R *s_c12x12, const R *s_c6x6, const R *s_rc2,
                                            R &fx, R &fy, R &fz, int first = 0, int stride = 1)
{
    fx = fy = fz = 0;
    for (int j = first; j < n; j += stride) {
        if (j == i) continue;
        const typename Pos4<R>::type q = s_pos[j];
        R dx, dy, dz;
        delta(p, q, box, dx, dy, dz);
        const R r2 = dx * dx + dy * dy + dz * dz;
        const int t = ti + ptype(q);
        const R r2inv = r_rcp(r2);
        const R r6inv = r2inv * r2inv * r2inv;
        R fp = r6inv * (s_c12x12[t] * r6inv - s_c6x6[t]) * r2inv;
        fp = r2 < s_rc2[t] ? fp : (R)0;
        fx += fp * dx; fy += fp * dy; fz += fp * dz;
    }
}

template <typename R> struct BatchNveArgs {
    typename Pos4<R>::type *pos;        // [nrep][n]
    typename Vec4<R>::type *vel;        // [nrep][n] (vx, vy, vz, mass)
    typename Vec4<R>::type *frc;        // [nrep][n] forces of the final configuration (+ pe_i)
    int *img;                           // [nrep][n][3] box crossings, accumulated
    int n, nrep;
    Box<R> box;
    long long steps;
    double dt;
    double *kinetic, *energy, *virial;  // [nrep] of the final configuration
};

// `steps` velocity-Verlet steps of every replica in ONE launch: kick-drift form
// exactly as in k_step_nve (opening half kick, identical steps, closing half kick),
// two block barriers per step, nothing leaves the SM until the end.
// T threads share one particle (its neighbours dealt out round-robin, partial forces combined by
// a butterfly so that all T lanes hold the identical sum): a lone small system is bound by the
// latency of one thread's serial pair loop, which T divides.
template <typename R, int T>
__global__ void k_batch_nve(BatchNveArgs<R> a)
{
    typedef typename Pos4<R>::type P4;
    typedef typename Vec4<R>::type V4;
    extern __shared__ __align__(16) unsigned char s_raw[];
    P4 *s_pos = reinterpret_cast<P4 *>(s_raw);
    __shared__ double s_red[32];
    __shared__ R s_c12x12[B2_MAXSP * B2_MAXSP], s_c6x6[B2_MAXSP * B2_MAXSP], s_rc2[B2_MAXSP * B2_MAXSP];
    for (int k = threadIdx.x; k < B2_MAXSP * B2_MAXSP; k += blockDim.x) {
        s_c12x12[k] = (R)(12.0 * g_pair.c12[k]);
        s_c6x6[k] = (R)(6.0 * g_pair.c6[k]);
        s_rc2[k] = (R)g_pair.rc2[k];
    }
    const int rep = blockIdx.x, i = threadIdx.x / T, sub = threadIdx.x % T, n = a.n, nsp = g_pair.nsp;
    const bool own = i < n;
    const bool lead = own && sub == 0;
    const size_t g = (size_t)rep * n + (own ? i : 0);
    P4 p = own ? a.pos[g] : P4();
    V4 v = own ? a.vel[g] : make4((R)0, (R)0, (R)0, (R)1);
    const int ti = own ? ptype(p) * nsp : 0;
    const R dt = (R)a.dt, hdt = (R)(0.5 * a.dt);
    int ix = 0, iy = 0, iz = 0;
    R fx = 0, fy = 0, fz = 0;
#define B2_BATCH_FORCE()                                                                           \
    {                                                                                              \
        if (own) batch_force<R>(s_pos, n, i, p, ti, a.box, s_c12x12, s_c6x6, s_rc2, fx, fy, fz, sub, T); \
        _Pragma("unroll") for (int o = 1; o < T; o <<= 1) {                                        \
            fx += __shfl_xor_sync(0xffffffffu, fx, o);                                             \
            fy += __shfl_xor_sync(0xffffffffu, fy, o);                                             \
            fz += __shfl_xor_sync(0xffffffffu, fz, o);                                             \
        }                                                                                          \
    }
    if (lead) s_pos[i] = p;
    __syncthreads();
    B2_BATCH_FORCE()
    if (a.steps > 0) {
        R s = hdt / v.w;
        v.x -= s * fx; v.y -= s * fy; v.z -= s * fz;            // velocities to t - dt/2
        for (long long step = 0; step < a.steps; step++) {
            s = dt / v.w;
            v.x += s * fx; v.y += s * fy; v.z += s * fz;
            if (own) {
                int kx, ky, kz;
                p = advance(p, dt * v.x, dt * v.y, dt * v.z, a.box, kx, ky, kz);
                ix += kx; iy += ky; iz += kz;
            }
            __syncthreads();                                    // everybody has read the old positions
            if (lead) s_pos[i] = p;
            __syncthreads();
            B2_BATCH_FORCE()
        }
        s = hdt / v.w;
        v.x += s * fx; v.y += s * fy; v.z += s * fz;            // closing half kick
    }
    // final configuration: fp64 energy, virial, kinetic energy
#undef B2_BATCH_FORCE
    double pe = 0, vir = 0, kin = 0;
    if (lead) {
        for (int j = 0; j < n; j++) {
            if (j == i) continue;
            double ex, ey, ez;
            delta_d(p, s_pos[j], a.box, ex, ey, ez);
            const double q2 = ex * ex + ey * ey + ez * ez;
            const int t = ti + ptype(s_pos[j]);
            if (q2 < g_pair.rc2[t]) {
                const double q2inv = 1.0 / q2, q6inv = q2inv * q2inv * q2inv;
                pe += q6inv * (g_pair.c12[t] * q6inv - g_pair.c6[t]) - g_pair.ushift[t];
                vir += q6inv * (12.0 * g_pair.c12[t] * q6inv - 6.0 * g_pair.c6[t]);
            }
        }
        const double vx = v.x, vy = v.y, vz = v.z;
        kin = 0.5 * (double)v.w * ((vx * vx + vy * vy) + vz * vz);
        a.pos[g] = p;
        a.vel[g] = v;
        a.frc[g] = make4(fx, fy, fz, (R)(0.5 * pe));
        int *im = a.img + 3 * g;
        im[0] += ix; im[1] += iy; im[2] += iz;
    }
    const double e_tot = block_sum(0.5 * pe, s_red);
    const double w_tot = block_sum(0.5 * vir, s_red);
    const double k_tot = block_sum(kin, s_red);
    if (threadIdx.x == 0) {
        a.energy[rep] = e_tot;
        a.virial[rep] = w_tot;
        a.kinetic[rep] = k_tot;
    }
}

template <typename R>
__global__ void k_batch_ingest(const double *__restrict__ spos, const double *__restrict__ smass,
                               const int *__restrict__ sspe, long long tot, Box<R> box,
                               typename Pos4<R>::type *__restrict__ pos,
                               typename Vec4<R>::type *__restrict__ vel, int *__restrict__ img)
{
    long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= tot) return;
    double x = spos[3 * k], y = spos[3 * k + 1], z = spos[3 * k + 2];
    x -= rint(x / box.Ld[0]) * box.Ld[0];
    y -= rint(y / box.Ld[1]) * box.Ld[1];
    z -= rint(z / box.Ld[2]) * box.Ld[2];
    pos[k] = encode_pos(x, y, z, sspe[k], box);
    vel[k] = make4((R)0, (R)0, (R)0, (R)(smass ? smass[k] : 1.0));
    img[3 * k] = img[3 * k + 1] = img[3 * k + 2] = 0;
}

template <typename R>
__global__ void k_batch_export(const typename Pos4<R>::type *__restrict__ pos,
                               const typename Vec4<R>::type *__restrict__ frc, long long tot, Box<R> box,
                               double *__restrict__ opos, double *__restrict__ ofrc)
{
    long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= tot) return;
    if (opos) {
        double x, y, z;
        to_double(pos[k], box, x, y, z);
        opos[3 * k] = x; opos[3 * k + 1] = y; opos[3 * k + 2] = z;
    }
    if (ofrc) {
        typename Vec4<R>::type f = frc[k];
        ofrc[3 * k] = f.x; ofrc[3 * k + 1] = f.y; ofrc[3 * k + 2] = f.z;
    }
}

template <typename R>
__global__ void k_batch_set_vel(const double *__restrict__ svel, long long tot,
                                typename Vec4<R>::type *__restrict__ vel)
{
    long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= tot) return;
    typename Vec4<R>::type v = vel[k];
    vel[k] = make4((R)svel[3 * k], (R)svel[3 * k + 1], (R)svel[3 * k + 2], v.w);
}

template <typename R>
__global__ void k_batch_get_vel(const typename Vec4<R>::type *__restrict__ vel, long long tot,
                                double *__restrict__ ovel)
{
    long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= tot) return;
    typename Vec4<R>::type v = vel[k];
    ovel[3 * k] = v.x; ovel[3 * k + 1] = v.y; ovel[3 * k + 2] = v.z;
}
