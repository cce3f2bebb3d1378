// Slab refresh on the device (b2_peer_refresh): migration, ghost selection and
// re-ingestion without the host touching a particle.  The host-driven refresh of
// atooms_b200/dd.py (download -> ~100 torch operations -> exchange through
// torch.distributed -> re-upload) costs 1.7 ms for 1.2M particles per GPU, more than
// the list build that follows it; this path does the same work in five kernels and
// one host synchronisation, and exchanges through the peer-memory mailboxes.
//
//   k_ddr_classify   slab of every owned particle -> stay / go left / go right,
//                    near-left / near-right flags for those that stay; per-block counts
//   k_ddr_scan       exclusive scan of the five per-block count rows
//   k_ddr_scatter    ordered lists (slot order = Morton order: deterministic, and the
//                    re-ingested array is already nearly sorted for the next radix sort)
//   k_ddr_pack       migrants (position, velocity, mass, global id, images) and the ghost
//                    layer (positions) stored straight into the neighbours' inboxes
//   k_ddr_assemble   waits for both neighbours, writes the new particle arrays
//                    [stay | from left | from right | ghosts left | ghosts right]
//                    and the new send lists of the per-step halo
// Ghost order on both sides of a face: [the neighbour's staying near-face particles,
// then the particles that just crossed the face, in sending order] — the rank they
// left keeps their coordinates, so they are not sent back (as in dd.py).
#pragma once
#include "b2_common.cuh"

template <typename R> struct DdrMig {
    typename Pos4<R>::type p;
    typename Vec4<R>::type v;
    long long gid;
    int img[3];
    int pad_;
};

#define DDR_MIG_BYTES 96            // >= sizeof(DdrMig<double>), record stride in the inbox
#define DDR_GHOST_BYTES 32          // >= sizeof(Pos4<double>::type)
#define DDR_THREADS 256

// device counters (unsigned[32]) ----------------------------------------------
enum { DC_KEEP = 0, DC_GO_L, DC_GO_R, DC_NEAR_L, DC_NEAR_R, DC_ERR, DC_SEQ, DC_TICKET,
       DC_N_OWN = 8, DC_G_L, DC_G_R, DC_SEND_L, DC_SEND_R, DC_FROM_L, DC_FROM_R };
// error bits
enum { DE_HOP = 1, DE_MIG_CAP = 2, DE_GHOST_CAP = 4, DE_TIMEOUT = 8, DE_NCAP = 16 };

template <typename R> struct DdrArgs {
    typedef typename Pos4<R>::type P4;
    typedef typename Vec4<R>::type V4;
    const Ctl *ctl;
    const P4 *pos[2];
    const V4 *vel;
    const int *orig;              // slot -> local id
    const int *img;               // [3n] by id: box crossings since the last ingestion
    const int *cimg;              // [3n] by id: crossings before that
    const long long *gid;         // [n] by id
    int n, ncap;
    int rank, nranks;
    double lo, hi, margin;        // this slab and the ghost margin, as fractions of the box along x
    Box<R> box;
    // work arrays
    unsigned char *cls;           // [n]
    int *blk;                     // [5][nblk]
    int nblk;
    int *keep, *go_l, *go_r;      // slots
    int *near_l_slot, *near_l_new, *near_r_slot, *near_r_new;
    int mig_cap, ghost_cap;
    unsigned *cnt;                // device counters
    // mailboxes
    char *out_l, *out_r;          // REMOTE inbox blocks [parity 0]; parity 1 follows at `inbox_bytes`
    size_t inbox_bytes;
    unsigned long long *flag_l, *flag_r;           // REMOTE arrival flags
    const char *in_l, *in_r;      // LOCAL inbox blocks [parity 0]
    const unsigned long long *my_flag_l, *my_flag_r;
    // outputs (new particle arrays, by NEW local id)
    P4 *npos;
    V4 *nvel;
    long long *ngid;
    int *ncimg;
    int *nsend_l, *nsend_r;
};

__device__ __forceinline__ double ddr_frac(const int4 &p, const Box<float> &)
{
    return (double)((unsigned)p.x + 0x80000000u) * (1.0 / 4294967296.0);
}
__device__ __forceinline__ double ddr_frac(const double4 &p, const Box<double> &box)
{
    return p.x * box.invL[0] + 0.5;
}

// class byte: bits 0-1 = 0 nothing (ghost / out of range), 1 stay, 2 go left, 3 go right;
// bit 2 near the left face, bit 3 near the right face (only with "stay")
template <typename R>
__device__ __forceinline__ unsigned ddr_class(const DdrArgs<R> &a, const typename Pos4<R>::type &p, unsigned *cnt)
{
    if (is_ghost(p)) return 0u;
    const double u = ddr_frac(p, a.box);
    int slab = (int)floor(u * a.nranks);
    slab = min(max(slab, 0), a.nranks - 1);
    if (slab == a.rank) {
        unsigned c = 1u;
        if (u < a.lo + a.margin) c |= 4u;
        if (u >= a.hi - a.margin) c |= 8u;
        return c;
    }
    if (a.nranks == 2) return 3u;                       // two ranks: either way round the ring
    int hop = slab - a.rank;
    hop = (hop + a.nranks / 2 + a.nranks) % a.nranks - a.nranks / 2;
    if (hop == -1) return 2u;
    if (hop == 1) return 3u;
    atomicOr(cnt + DC_ERR, (unsigned)DE_HOP);           // moved more than one slab between two refreshes
    return 1u;
}

template <typename R>
__global__ void __launch_bounds__(DDR_THREADS)
k_ddr_classify(DdrArgs<R> a)
{
    const typename Pos4<R>::type *pos = a.ctl->cur ? a.pos[1] : a.pos[0];
    const int s = blockIdx.x * DDR_THREADS + threadIdx.x;
    unsigned c = 0u;
    if (s < a.n) {
        c = ddr_class(a, pos[s], a.cnt);
        a.cls[s] = (unsigned char)c;
    }
    const int k = __syncthreads_count((c & 3u) == 1u);
    const int gl = __syncthreads_count((c & 3u) == 2u);
    const int gr = __syncthreads_count((c & 3u) == 3u);
    const int nl = __syncthreads_count((c & 4u) != 0u);
    const int nr = __syncthreads_count((c & 8u) != 0u);
    if (threadIdx.x == 0) {
        a.blk[0 * a.nblk + blockIdx.x] = k;
        a.blk[1 * a.nblk + blockIdx.x] = gl;
        a.blk[2 * a.nblk + blockIdx.x] = gr;
        a.blk[3 * a.nblk + blockIdx.x] = nl;
        a.blk[4 * a.nblk + blockIdx.x] = nr;
    }
}

// one block of 1024 threads; row c of blk becomes its exclusive scan, the total goes to cnt[c]
__global__ void __launch_bounds__(1024)
k_ddr_scan(int *blk, int nblk, unsigned *cnt)
{
    __shared__ int s_part[1024];
    const int t = threadIdx.x;
    const int per = (nblk + 1023) / 1024;
    for (int c = 0; c < 5; c++) {
        int *row = blk + (size_t)c * nblk;
        const int b0 = min(t * per, nblk), b1 = min(b0 + per, nblk);
        int sum = 0;
        for (int b = b0; b < b1; b++) sum += row[b];
        s_part[t] = sum;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {            // Hillis-Steele, inclusive
            const int v = t >= o ? s_part[t - o] : 0;
            __syncthreads();
            s_part[t] += v;
            __syncthreads();
        }
        int run = s_part[t] - sum;
        if (t == 1023) cnt[c] = (unsigned)s_part[t];
        for (int b = b0; b < b1; b++) {
            const int v = row[b];
            row[b] = run;
            run += v;
        }
        __syncthreads();
    }
}

// rank of this thread among the threads of its block with `flag` set (block-wide, ordered by thread)
__device__ __forceinline__ int ddr_block_rank(bool flag, int *s_w /* [8] */)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    __syncthreads();                                   // s_w may still be read from the previous use
    if (lane == 0) s_w[w] = __popc(m);
    __syncthreads();
    int r = __popc(m & ((1u << lane) - 1u));
    for (int ww = 0; ww < w; ww++) r += s_w[ww];
    return r;
}

template <typename R>
__global__ void __launch_bounds__(DDR_THREADS)
k_ddr_scatter(DdrArgs<R> a)
{
    __shared__ int s_w[DDR_THREADS / 32];
    const int s = blockIdx.x * DDR_THREADS + threadIdx.x;
    const unsigned c = s < a.n ? a.cls[s] : 0u;
    const int b = blockIdx.x;
    const int ik = a.blk[0 * a.nblk + b] + ddr_block_rank((c & 3u) == 1u, s_w);
    const int il = a.blk[1 * a.nblk + b] + ddr_block_rank((c & 3u) == 2u, s_w);
    const int ir = a.blk[2 * a.nblk + b] + ddr_block_rank((c & 3u) == 3u, s_w);
    const int nl = a.blk[3 * a.nblk + b] + ddr_block_rank((c & 4u) != 0u, s_w);
    const int nr = a.blk[4 * a.nblk + b] + ddr_block_rank((c & 8u) != 0u, s_w);
    if ((c & 3u) == 1u) a.keep[ik] = s;
    if ((c & 3u) == 2u) { if (il < a.mig_cap) a.go_l[il] = s; else atomicOr(a.cnt + DC_ERR, (unsigned)DE_MIG_CAP); }
    if ((c & 3u) == 3u) { if (ir < a.mig_cap) a.go_r[ir] = s; else atomicOr(a.cnt + DC_ERR, (unsigned)DE_MIG_CAP); }
    if (c & 4u) {
        if (nl < a.ghost_cap) { a.near_l_slot[nl] = s; a.near_l_new[nl] = ik; }
        else atomicOr(a.cnt + DC_ERR, (unsigned)DE_GHOST_CAP);
    }
    if (c & 8u) {
        if (nr < a.ghost_cap) { a.near_r_slot[nr] = s; a.near_r_new[nr] = ik; }
        else atomicOr(a.cnt + DC_ERR, (unsigned)DE_GHOST_CAP);
    }
}

// inbox block of one (side, parity): [64-byte header: n_mig, n_ghost][mig_cap records][ghost_cap records]
template <typename R>
__device__ __forceinline__ DdrMig<R> *ddr_mig(char *inbox) { return (DdrMig<R> *)(inbox + 64); }
template <typename R>
__device__ __forceinline__ typename Pos4<R>::type *ddr_ghost(char *inbox, int mig_cap)
{
    return (typename Pos4<R>::type *)(inbox + 64 + (size_t)mig_cap * DDR_MIG_BYTES);
}

template <typename R>
__global__ void __launch_bounds__(DDR_THREADS)
k_ddr_pack(DdrArgs<R> a)
{
    typedef typename Pos4<R>::type P4;
    const P4 *pos = a.ctl->cur ? a.pos[1] : a.pos[0];
    const unsigned seq = a.cnt[DC_SEQ] + 1u;            // rewritten only after every block has passed here
    const size_t par = (size_t)(seq & 1u) * a.inbox_bytes;
    char *out_l = a.out_l + par, *out_r = a.out_r + par;
    const int ngl = min((int)a.cnt[DC_GO_L], a.mig_cap), ngr = min((int)a.cnt[DC_GO_R], a.mig_cap);
    const int nnl = min((int)a.cnt[DC_NEAR_L], a.ghost_cap), nnr = min((int)a.cnt[DC_NEAR_R], a.ghost_cap);
    int k = blockIdx.x * DDR_THREADS + threadIdx.x;
    if (k == 0) {
        ((int *)out_l)[0] = ngl; ((int *)out_l)[1] = nnl;
        ((int *)out_r)[0] = ngr; ((int *)out_r)[1] = nnr;
    }
    if (k < ngl + ngr) {                                // migrants
        const bool left = k < ngl;
        const int q = left ? k : k - ngl;
        const int s = left ? a.go_l[q] : a.go_r[q];
        const int o = a.orig[s];
        DdrMig<R> m;
        m.p = pos[s];
        m.v = a.vel[s];
        m.gid = a.gid[o];
        m.img[0] = a.cimg[3 * o] + a.img[3 * o];
        m.img[1] = a.cimg[3 * o + 1] + a.img[3 * o + 1];
        m.img[2] = a.cimg[3 * o + 2] + a.img[3 * o + 2];
        m.pad_ = 0;
        *(DdrMig<R> *)((char *)ddr_mig<R>(left ? out_l : out_r) + (size_t)q * DDR_MIG_BYTES) = m;
    } else {
        k -= ngl + ngr;
        if (k < nnl + nnr) {                            // ghost layer: particles that stay, near a face
            const bool left = k < nnl;
            const int q = left ? k : k - nnl;
            const int s = left ? a.near_l_slot[q] : a.near_r_slot[q];
            *(P4 *)((char *)ddr_ghost<R>(left ? out_l : out_r, a.mig_cap) + (size_t)q * DDR_GHOST_BYTES) = pos[s];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(a.cnt + DC_TICKET, 1u);
        if (t == gridDim.x - 1) {                       // everything of this rank is on its way
            a.cnt[DC_TICKET] = 0u;
            a.cnt[DC_SEQ] = seq;
            __threadfence_system();
            *(volatile unsigned long long *)a.flag_l = (unsigned long long)seq;
            *(volatile unsigned long long *)a.flag_r = (unsigned long long)seq;
        }
    }
}

__device__ __forceinline__ int4 ddr_as_ghost(int4 p) { p.w |= B2_GHOST_BIT; return p; }
__device__ __forceinline__ double4 ddr_as_ghost(double4 p)
{
    p.w = __longlong_as_double(__double_as_longlong(p.w) | (long long)B2_GHOST_BIT);
    return p;
}

template <typename R>
__global__ void __launch_bounds__(DDR_THREADS)
k_ddr_assemble(DdrArgs<R> a)
{
    typedef typename Pos4<R>::type P4;
    typedef typename Vec4<R>::type V4;
    __shared__ bool s_ok;
    const unsigned seq = a.cnt[DC_SEQ];
    if (threadIdx.x == 0) {
        bool ok = true;
        const long long t0 = clock64();
        while (*(const volatile unsigned long long *)a.my_flag_l < seq ||
               *(const volatile unsigned long long *)a.my_flag_r < seq) {
            if (clock64() - t0 > 20000000000LL) { ok = false; atomicOr(a.cnt + DC_ERR, (unsigned)DE_TIMEOUT); break; }
            __nanosleep(200);
        }
        __threadfence_system();
        s_ok = ok;
    }
    __syncthreads();
    if (!s_ok) return;
    const P4 *pos = a.ctl->cur ? a.pos[1] : a.pos[0];
    const size_t par = (size_t)(seq & 1u) * a.inbox_bytes;
    char *in_l = const_cast<char *>(a.in_l) + par, *in_r = const_cast<char *>(a.in_r) + par;
    // written by other GPUs: read around L1
    const int m_l = min(__ldcg((const int *)in_l), a.mig_cap), g_l = min(__ldcg((const int *)in_l + 1), a.ghost_cap);
    const int m_r = min(__ldcg((const int *)in_r), a.mig_cap), g_r = min(__ldcg((const int *)in_r + 1), a.ghost_cap);
    const int n_keep = (int)a.cnt[DC_KEEP];
    const int go_l = min((int)a.cnt[DC_GO_L], a.mig_cap), go_r = min((int)a.cnt[DC_GO_R], a.mig_cap);
    const int near_l = min((int)a.cnt[DC_NEAR_L], a.ghost_cap), near_r = min((int)a.cnt[DC_NEAR_R], a.ghost_cap);
    const int n_own = n_keep + m_l + m_r;
    const int G_l = g_l + go_l, G_r = g_r + go_r;
    const int total = n_own + G_l + G_r;
    const int t = blockIdx.x * DDR_THREADS + threadIdx.x;
    if (t == 0) {
        a.cnt[DC_N_OWN] = (unsigned)n_own;
        a.cnt[DC_G_L] = (unsigned)G_l;
        a.cnt[DC_G_R] = (unsigned)G_r;
        a.cnt[DC_SEND_L] = (unsigned)(near_l + m_l);
        a.cnt[DC_SEND_R] = (unsigned)(near_r + m_r);
        a.cnt[DC_FROM_L] = (unsigned)m_l;
        a.cnt[DC_FROM_R] = (unsigned)m_r;
        if (total > a.ncap) atomicOr(a.cnt + DC_ERR, (unsigned)DE_NCAP);
    }
    if (total > a.ncap) return;
    // the new send lists of the per-step halo: who stays near a face, then who just arrived through it
    if (t < near_l + m_l) a.nsend_l[t] = t < near_l ? a.near_l_new[t] : n_keep + (t - near_l);
    if (t < near_r + m_r) a.nsend_r[t] = t < near_r ? a.near_r_new[t] : n_keep + m_l + (t - near_r);
    if (t >= total) return;
    P4 p;
    V4 v = make4((R)0, (R)0, (R)0, (R)1);              // ghosts: no velocity, unit mass
    long long gid = -1;
    int i0 = 0, i1 = 0, i2 = 0;
    if (t < n_keep) {
        const int s = a.keep[t];
        const int o = a.orig[s];
        p = pos[s];
        v = a.vel[s];
        gid = a.gid[o];
        i0 = a.cimg[3 * o] + a.img[3 * o];
        i1 = a.cimg[3 * o + 1] + a.img[3 * o + 1];
        i2 = a.cimg[3 * o + 2] + a.img[3 * o + 2];
    } else if (t < n_own) {
        const bool left = t < n_keep + m_l;
        const int q = left ? t - n_keep : t - n_keep - m_l;
        const char *rec = (const char *)ddr_mig<R>(left ? in_l : in_r) + (size_t)q * DDR_MIG_BYTES;
        const int4 *w = (const int4 *)rec;
        DdrMig<R> m;
        int4 *dst = (int4 *)&m;
#pragma unroll
        for (int k = 0; k < (int)(sizeof(DdrMig<R>) / 16); k++) dst[k] = __ldcg(w + k);
        p = m.p; v = m.v; gid = m.gid; i0 = m.img[0]; i1 = m.img[1]; i2 = m.img[2];
    } else {
        const int q = t - n_own;
        if (q < G_l) {
            if (q < g_l) {
                const int4 *w = (const int4 *)((const char *)ddr_ghost<R>(in_l, a.mig_cap) + (size_t)q * DDR_GHOST_BYTES);
                int4 *dst = (int4 *)&p;
#pragma unroll
                for (int k = 0; k < (int)(sizeof(P4) / 16); k++) dst[k] = __ldcg(w + k);
            } else {
                p = pos[a.go_l[q - g_l]];
            }
        } else {
            const int qq = q - G_l;
            if (qq < g_r) {
                const int4 *w = (const int4 *)((const char *)ddr_ghost<R>(in_r, a.mig_cap) + (size_t)qq * DDR_GHOST_BYTES);
                int4 *dst = (int4 *)&p;
#pragma unroll
                for (int k = 0; k < (int)(sizeof(P4) / 16); k++) dst[k] = __ldcg(w + k);
            } else {
                p = pos[a.go_r[qq - g_r]];
            }
        }
        p = ddr_as_ghost(p);
    }
    a.npos[t] = p;
    a.nvel[t] = v;
    a.ngid[t] = gid;
    a.ncimg[3 * t] = i0; a.ncimg[3 * t + 1] = i1; a.ncimg[3 * t + 2] = i2;
}

// after the host has adopted the new arrays: identity order, crossing counters back to zero,
// rmsd origin at the current position (as k_ingest leaves them)
template <typename R>
__global__ void k_ddr_finish(const typename Pos4<R>::type *__restrict__ pos, int n, Box<R> box,
                             int *__restrict__ orig, int *__restrict__ inv, int *__restrict__ img,
                             double *__restrict__ ref)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orig[i] = i;
    inv[i] = i;
    img[3 * i] = img[3 * i + 1] = img[3 * i + 2] = 0;
    double x, y, z;
    to_double(pos[i], box, x, y, z);
    ref[3 * i] = x; ref[3 * i + 1] = y; ref[3 * i + 2] = z;
}
