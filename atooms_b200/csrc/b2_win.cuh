// Windowed Verlet list: neighbour positions come from shared memory, not from L1 gathers.
//
// The per-lane gathers of pair_row (b2_force.cuh) touch ~26 sectors in a dozen 128-byte lines per
// warp request and keep the L1 at 80 % of its throughput (profiles/r01_ncu_k_step_nve_fp32.txt).
// Here the grid is cut into SUPER-BLOCKS of 4x4x4 cells (cells are r_list,max / 2 wide, so a
// super-block is a cube of 2 r_list; its ~210 particles are 64 consecutive Morton codes, i.e. one
// run of slots).  Every neighbour of every particle of a super-block lies in the WINDOW of 8x8x8
// cells around it: 4x4x4 aligned 2x2x2 blocks of cells ("B8"), each of which is again one run of
// slots.  A thread block takes one super-block, copies the 64 runs of its window into shared
// memory with bulk asynchronous copies (cp.async.bulk, completion on an mbarrier) and walks
// neighbour rows whose entries are 16-BIT OFFSETS INTO THE WINDOW: the index stream is half as
// long, and the position reads are LDS.128 instead of global gathers.  The step kernel first turns
// the staged records into coordinates RELATIVE TO THE SUPER-BLOCK'S CENTRE in the working
// precision (one pass over the window, in place): a pair separation is then three subtractions,
// with no minimum-image arithmetic and no integer-to-float conversion per pair.
//
//   window order   run r = (rz * 4 + ry) * 4 + rx, B8 block (2 sx - 1 + rx, ...) wrapped
//                  periodically; offset of slot j of run r = P[r] + (j - start[r]), P = exclusive
//                  prefix of the run lengths.  The list build computes the same numbers.
//   rows           tiles of 32 lanes are LOCAL to a super-block (tile_base[] = prefix over the
//                  super-blocks of ceil(count / 32)); entry k of a lane sits at
//                  nbr16[((tile_base + t) * cap + k) * 32 + lane]: a warp reads one 64-byte row per k
//
// Requirements (else the context keeps the global-index list): half-width cells (m = 2), an EVEN
// number of cells per axis, at least 8 (setup_grid drops one cell where needed; the last super-
// block of an axis may then be two cells wide, its window wraps onto the first B8 blocks),
// windows of fewer than 65535 particles that fit the shared memory of an SM.
#pragma once
#include "b2_common.cuh"
#include "b2_force.cuh"

#define WIN_RUNS 64
#ifndef WIN_THREADS
#define WIN_THREADS 256
#endif
#ifndef WIN_UNROLL32
#define WIN_UNROLL32 10         // neighbours in flight per lane (index -> record -> pair arithmetic): 0.201 / 0.193 /
#endif                          // 0.190 ms per step with 6 / 8 / 10 (KA N = 1.2M, 64 registers, no spills)
#ifndef WIN_UNROLL64
#define WIN_UNROLL64 4
#endif
#ifndef WIN_MINBLK32
#define WIN_MINBLK32 4
#endif
#ifndef WIN_MINBLK64
#define WIN_MINBLK64 3
#endif

__device__ __forceinline__ int wrap_b8(int b, int nb) { return b < 0 ? b + nb : (b >= nb ? b - nb : b); }

// run r of the window of super-block (sx, sy, sz): where it starts and how many slots it has
__device__ __forceinline__ void win_run(const Grid &g, int sx, int sy, int sz, int r,
                                        const int *__restrict__ b8start, int &start, int &len)
{
    const int bx = wrap_b8(2 * sx - 1 + (r & 3), g.nc[0] >> 1);
    const int by = wrap_b8(2 * sy - 1 + ((r >> 2) & 3), g.nc[1] >> 1);
    const int bz = wrap_b8(2 * sz - 1 + (r >> 4), g.nc[2] >> 1);
    const uint32_t code = grid_key(g, 1, bx, by, bz);
    start = b8start[code];
    len = b8start[code + 1] - start;
}

// warp-level: the 64 runs (two per lane: r = lane and lane + 32) with their exclusive prefix, in
// registers; returns the window size W
__device__ __forceinline__ int win_layout_regs(const Grid &g, int sx, int sy, int sz, int lane,
                                               const int *__restrict__ b8start,
                                               int &p0, int &st0, int &ln0, int &p1, int &st1, int &ln1)
{
    win_run(g, sx, sy, sz, lane, b8start, st0, ln0);
    win_run(g, sx, sy, sz, lane + 32, b8start, st1, ln1);
    int i0 = ln0, i1 = ln1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v0 = __shfl_up_sync(0xffffffffu, i0, o), v1 = __shfl_up_sync(0xffffffffu, i1, o);
        if (lane >= o) { i0 += v0; i1 += v1; }
    }
    const int tot0 = __shfl_sync(0xffffffffu, i0, 31);
    p0 = i0 - ln0;
    p1 = tot0 + i1 - ln1;
    return tot0 + __shfl_sync(0xffffffffu, i1, 31);
}

// Per super-block table, written once per list refresh (k_sb_table) and read by every kernel that
// needs a window:  [0] first slot  [1] particles  [2] tile base  [3] window size W
// [4..9] the super-block's centre in the working precision (3 fixed-point words or 3 doubles),
// then from [SBT_HEAD] for each run r: its first slot, and (offset P of the run in the window) | (length << 16).
#define SBT_HEAD 12
#define SBT_STRIDE (SBT_HEAD + 2 * WIN_RUNS)

// the table's runs, left in shared memory for look-ups by offset (list build, export); one warp
__device__ __forceinline__ int win_layout_load(const int *__restrict__ T, int lane, int *sP, int *sS)
{
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int r = lane + 32 * h;
        sS[r] = T[SBT_HEAD + 2 * r];
        sP[r] = T[SBT_HEAD + 1 + 2 * r] & 0xffff;
    }
    const int W = T[3];
    if (lane == 0) sP[WIN_RUNS] = W;
    return W;
}

// slot of the window record at offset `off` (rare paths: band fix-up of the build, list export)
__device__ __forceinline__ int win_slot_of(const int *sP, const int *sS, int off)
{
    int r = 0;
#pragma unroll
    for (int step = 32; step; step >>= 1)
        if (r + step < WIN_RUNS && sP[r + step] <= off) r += step;
    return sS[r] + off - sP[r];
}

// first slot of each B8 block: lower bound of its first cell key in the sorted keys
__global__ void k_b8_bounds(const Ctl *__restrict__ ctl, const uint32_t *__restrict__ keys, int n, int ncodes,
                            int *__restrict__ b8start)
{
    if (!ctl->halt_rebuild) return;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > ncodes) return;
    const uint32_t target = (uint32_t)b << 3;
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (keys[mid] < target) lo = mid + 1; else hi = mid;
    }
    b8start[b] = lo;
}

// super-block of table row gi (= linear block index of the windowed kernels: x fastest over the launched
// sbn[0] x sbn[1] x sbn[2] super-blocks, which start at sb0 and wrap around the box)
__device__ __forceinline__ void sb_coords(const Grid &g, int gi, int &sx, int &sy, int &sz)
{
    sx = g.sb0[0] + gi % g.sbn[0];
    sy = g.sb0[1] + (gi / g.sbn[0]) % g.sbn[1];
    sz = g.sb0[2] + gi / (g.sbn[0] * g.sbn[1]);
    sx -= sx >= g.nsb[0] ? g.nsb[0] : 0;
    sy -= sy >= g.nsb[1] ? g.nsb[1] : 0;
    sz -= sz >= g.nsb[2] ? g.nsb[2] : 0;
}

// tile_base[gi] = tiles of 32 lanes before super-block gi (geometric index, x fastest)
__global__ void __launch_bounds__(1024)
k_sb_tiles(const Ctl *__restrict__ ctl, Grid g, const int *__restrict__ b8start, int *__restrict__ tile_base)
{
    if (!ctl->halt_rebuild) return;
    __shared__ int s_w[32];
    __shared__ int s_run;
    const int nsb = g.sbn[0] * g.sbn[1] * g.sbn[2];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_run = 0;
    __syncthreads();
    for (int base = 0; base < nsb; base += 1024) {
        const int gi = base + threadIdx.x;
        int t = 0;
        if (gi < nsb) {
            int sx, sy, sz;
            sb_coords(g, gi, sx, sy, sz);
            const uint32_t code = grid_key(g, 2, sx, sy, sz);
            t = (b8start[(code + 1) << 3] - b8start[code << 3] + 31) >> 5;
        }
        int incl = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_w[w] = incl;
        __syncthreads();
        if (w == 0) {
            int v = s_w[lane], i = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, i, o);
                if (lane >= o) i += u;
            }
            s_w[lane] = i - v;
        }
        __syncthreads();
        const int excl = s_run + s_w[w] + incl - t;
        if (gi < nsb) tile_base[gi] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_run = excl + t;
        __syncthreads();
    }
    if (threadIdx.x == 0) tile_base[nsb] = s_run;
}

// ---- bulk asynchronous copies -----------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@!p bra WAIT_%=;\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> this block's shared memory; bytes is a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct WinInfo { int first, cnt, tbase, W; };

// Block prologue (3-D grid of super-blocks).  No warp waits for another: after the barrier that
// publishes the initialised mbarrier every warp reads the block's row of the table and starts its
// share of the 64 bulk copies (run r goes to warp r mod #warps).  The caller requests whatever
// else it needs from global memory and then calls win_wait().  (With one warp laying the window
// out and issuing all 64 copies, the other seven spent a third of the kernel's warp-time at the
// hand-over barrier; with every warp computing the layout itself, a fifth of the instructions
// were prologue: profiles/r02_ncu_k_step_win_v1.txt, _v2.txt.)
template <typename R>
__device__ __forceinline__ WinInfo win_stage(const int *__restrict__ sbt,
                                             const typename Pos4<R>::type *__restrict__ pos,
                                             typename Pos4<R>::type *win, unsigned long long *bar)
{
    typedef typename Pos4<R>::type P4;
    const int gi = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    const int *__restrict__ T = sbt + (size_t)gi * SBT_STRIDE;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    if (threadIdx.x == 0) mbar_init(bar, 1);
    __syncthreads();
    WinInfo wi;
    wi.first = T[0];
    wi.cnt = T[1];
    wi.tbase = T[2];
    wi.W = T[3];
    if (threadIdx.x == 0) mbar_expect_tx(bar, (unsigned)wi.W * (unsigned)sizeof(P4));
    if (lane < WIN_RUNS / nwarp) {
        const int r = warp + nwarp * lane;
        const int st = T[SBT_HEAD + 2 * r], pl = T[SBT_HEAD + 1 + 2 * r];
        const int len = (int)((unsigned)pl >> 16);
        if (len > 0) bulk_g2s(win + (pl & 0xffff), pos + st, (unsigned)len * (unsigned)sizeof(P4), bar);
    }
    return wi;
}
__device__ __forceinline__ void win_wait(unsigned long long *bar) { mbar_wait(bar, 0u); }

__device__ __forceinline__ unsigned ld_stream_u16(const unsigned short *p)
{
    unsigned v;                 // zero-extended by the load itself
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// ---- coordinates relative to the super-block's centre -----------------------------------------
// Same size as the stored records (converted in place): (x, y, z, species bits).
struct __align__(16) relf4 { float x, y, z; int t; };
struct __align__(16) reld4 { double x, y, z; long long t; };
template <typename R> struct Rel4;
template <> struct Rel4<float>  { typedef relf4 type; };
template <> struct Rel4<double> { typedef reld4 type; };

__device__ __forceinline__ int ptype(const relf4 &p) { return p.t & 0xff; }
__device__ __forceinline__ int ptype(const reld4 &p) { return (int)p.t & 0xff; }
__device__ __forceinline__ void delta(const relf4 &a, const relf4 &b, const Box<float> &, float &dx, float &dy, float &dz)
{
    dx = a.x - b.x; dy = a.y - b.y; dz = a.z - b.z;
}
__device__ __forceinline__ void delta(const reld4 &a, const reld4 &b, const Box<double> &, double &dx, double &dy, double &dz)
{
    dx = a.x - b.x; dy = a.y - b.y; dz = a.z - b.z;
}
// (only to keep the observable branch of B2_PAIR well-formed: relative records are used by the step kernels, which
// evaluate no observables)
__device__ __forceinline__ void delta_d(const relf4 &a, const relf4 &b, const Box<float> &, double &dx, double &dy, double &dz)
{
    dx = (double)a.x - (double)b.x; dy = (double)a.y - (double)b.y; dz = (double)a.z - (double)b.z;
}
__device__ __forceinline__ void delta_d(const reld4 &a, const reld4 &b, const Box<double> &, double &dx, double &dy, double &dz)
{
    dx = a.x - b.x; dy = a.y - b.y; dz = a.z - b.z;
}

template <typename R> struct WinOrigin;
template <> struct WinOrigin<float>  { unsigned u[3]; };      // fixed point, as stored
template <> struct WinOrigin<double> { double d[3]; };

// centre of super-block (sx, sy, sz)
__device__ __forceinline__ WinOrigin<float> win_origin(const Grid &g, const Box<float> &, int sx, int sy, int sz)
{
    WinOrigin<float> o;
    // u = x + 2^31 is the fraction of the box from its lower face (cell_of)
    o.u[0] = (unsigned)((((unsigned long long)(4 * sx + 2)) << 32) / (unsigned)g.nc[0]) - 0x80000000u;
    o.u[1] = (unsigned)((((unsigned long long)(4 * sy + 2)) << 32) / (unsigned)g.nc[1]) - 0x80000000u;
    o.u[2] = (unsigned)((((unsigned long long)(4 * sz + 2)) << 32) / (unsigned)g.nc[2]) - 0x80000000u;
    return o;
}
__device__ __forceinline__ WinOrigin<double> win_origin(const Grid &g, const Box<double> &box, int sx, int sy, int sz)
{
    WinOrigin<double> o;
    o.d[0] = ((double)(4 * sx + 2) / (double)g.nc[0] - 0.5) * box.Ld[0];
    o.d[1] = ((double)(4 * sy + 2) / (double)g.nc[1] - 0.5) * box.Ld[1];
    o.d[2] = ((double)(4 * sz + 2) / (double)g.nc[2] - 0.5) * box.Ld[2];
    return o;
}
__device__ __forceinline__ void win_origin_store(int *T, const WinOrigin<float> &o)
{
    T[4] = (int)o.u[0]; T[5] = (int)o.u[1]; T[6] = (int)o.u[2];
}
__device__ __forceinline__ void win_origin_store(int *T, const WinOrigin<double> &o)
{
    double *D = reinterpret_cast<double *>(T + 4);        // 8-byte aligned: SBT_STRIDE is even
    D[0] = o.d[0]; D[1] = o.d[1]; D[2] = o.d[2];
}
__device__ __forceinline__ void win_origin_load(const int *__restrict__ T, WinOrigin<float> &o)
{
    o.u[0] = (unsigned)T[4]; o.u[1] = (unsigned)T[5]; o.u[2] = (unsigned)T[6];
}
__device__ __forceinline__ void win_origin_load(const int *__restrict__ T, WinOrigin<double> &o)
{
    const double *D = reinterpret_cast<const double *>(T + 4);
    o.d[0] = D[0]; o.d[1] = D[1]; o.d[2] = D[2];
}

// one warp per super-block: its row of the table
template <typename R>
__global__ void __launch_bounds__(256)
k_sb_table(const Ctl *__restrict__ ctl, Grid g, Box<R> box, const int *__restrict__ b8start,
           const int *__restrict__ tile_base, int *__restrict__ sbt)
{
    if (!ctl->halt_rebuild) return;
    const int lane = threadIdx.x & 31;
    const int gi = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (gi >= g.sbn[0] * g.sbn[1] * g.sbn[2]) return;
    int sx, sy, sz;
    sb_coords(g, gi, sx, sy, sz);
    int p0, st0, ln0, p1, st1, ln1;
    const int W = win_layout_regs(g, sx, sy, sz, lane, b8start, p0, st0, ln0, p1, st1, ln1);
    int *T = sbt + (size_t)gi * SBT_STRIDE;
    T[SBT_HEAD + 2 * lane] = st0;
    T[SBT_HEAD + 1 + 2 * lane] = (p0 & 0xffff) | (ln0 << 16);
    T[SBT_HEAD + 2 * (lane + 32)] = st1;
    T[SBT_HEAD + 1 + 2 * (lane + 32)] = (p1 & 0xffff) | (ln1 << 16);
    if (lane == 0) {
        const uint32_t code = grid_key(g, 2, sx, sy, sz);
        const int first = b8start[code << 3];
        T[0] = first;
        T[1] = b8start[(code + 1) << 3] - first;
        T[2] = tile_base[gi];
        T[3] = W;
        win_origin_store(T, win_origin(g, box, sx, sy, sz));
    }
}

// the wrap-around of the fixed-point difference IS the minimum image; every record of a window lies within
// 4 cells of the centre, far inside half a box
__device__ __forceinline__ relf4 to_rel(const int4 &p, const WinOrigin<float> &o, const Box<float> &box)
{
    relf4 r;
    r.x = (float)(int)((unsigned)p.x - o.u[0]) * box.s[0];
    r.y = (float)(int)((unsigned)p.y - o.u[1]) * box.s[1];
    r.z = (float)(int)((unsigned)p.z - o.u[2]) * box.s[2];
    r.t = p.w;
    return r;
}
__device__ __forceinline__ reld4 to_rel(const double4 &p, const WinOrigin<double> &o, const Box<double> &box)
{
    reld4 r;
    double dx = p.x - o.d[0], dy = p.y - o.d[1], dz = p.z - o.d[2];
    r.x = dx - box.L[0] * rint(dx * box.invL[0]);
    r.y = dy - box.L[1] * rint(dy * box.invL[1]);
    r.z = dz - box.L[2] * rint(dz * box.invL[2]);
    r.t = __double_as_longlong(p.w);
    return r;
}

// ---- list build on the staged window ------------------------------------------------------------
// One block per super-block, one warp per 2x2x2 block of cells ("B8") of it.  The window is staged as in
// the step kernel and turned, in place, into float coordinates relative to the super-block's centre with the
// species in the fourth lane.  The neighbours of a B8 block lie in the 3x3x3 B8 runs around it, and three
// runs adjacent along x are ONE stretch of window offsets: a warp walks nine stretches, every lane testing
// the same record (a broadcast LDS.128) against its own particle, and an entry of the list is the loop
// index itself.  No candidate is fetched from global memory by the loop (k_build_list_sc gathered its
// candidates per warp and spent 45 % of its time waiting for them: profiles/r02_ncu_k_build_list_sc_win_fp32.txt).
// Float test with a band around the threshold, band cases settled with the canonical predicate, as in
// k_build_list_sc: the list is bit-exact against the oracle's.
#ifndef BW_MINBLK
#define BW_MINBLK 2
#endif

// window records -> float4 (x, y, z relative to the centre; w = species: 0.f / 1.f when FAST2, else its integer bits)
template <bool FAST2>
__device__ __forceinline__ float4 bw_record(const relf4 &q)
{
    const int t = q.t & 0xff;
    return make_float4(q.x, q.y, q.z, FAST2 ? (float)t : __int_as_float(t));
}
template <bool FAST2>
__device__ __forceinline__ float4 bw_record(const reld4 &q)
{
    const int t = (int)q.t & 0xff;
    return make_float4((float)q.x, (float)q.y, (float)q.z, FAST2 ? (float)t : __int_as_float(t));
}
// Records are kept in PAIRS, transposed: records 2p and 2p + 1 are the two float4
//   (x0, x1, y0, y1)  (z0, z1, w0, w1)
// so that a 128-bit read yields two register pairs the packed fp32 instructions of sm_100 take as they are
// (add / mul / fma.f32x2: one issue slot for two candidates).
struct __align__(16) BwPair { float4 xy, zw; };
__device__ __forceinline__ BwPair bw_pair(const float4 &a, const float4 &b)
{
    BwPair p;
    p.xy = make_float4(a.x, b.x, a.y, b.y);
    p.zw = make_float4(a.z, b.z, a.w, b.w);
    return p;
}
#define BW_NOBODY make_float4(1e18f, 0.f, 0.f, 0.f)      // never within reach of anyone
template <bool FAST2>
__device__ __forceinline__ void bw_convert(unsigned char *raw, int W, const WinOrigin<float> &org, const Box<float> &box)
{
    // a thread takes both records of a pair: they occupy the 32 bytes the pair is written to
    const int4 *win = reinterpret_cast<const int4 *>(raw);
    BwPair *cand = reinterpret_cast<BwPair *>(raw);
    for (int p = threadIdx.x; 2 * p < W; p += WIN_THREADS) {
        const float4 a = bw_record<FAST2>(to_rel(win[2 * p], org, box));
        const float4 b = 2 * p + 1 < W ? bw_record<FAST2>(to_rel(win[2 * p + 1], org, box)) : BW_NOBODY;
        cand[p] = bw_pair(a, b);
    }
    __syncthreads();
}
template <bool FAST2>
__device__ __forceinline__ void bw_convert(unsigned char *raw, int W, const WinOrigin<double> &org, const Box<double> &box)
{
    // 64 bytes of records become a 32-byte pair at the front of the same buffer: a round of WIN_THREADS pairs is
    // read, then written behind a barrier; what a round writes lies below everything later rounds read
    const double4 *win = reinterpret_cast<const double4 *>(raw);
    BwPair *cand = reinterpret_cast<BwPair *>(raw);
    for (int base = 0; 2 * base < W; base += WIN_THREADS) {
        const int p = base + threadIdx.x;
        float4 a = BW_NOBODY, b = BW_NOBODY;
        if (2 * p < W) a = bw_record<FAST2>(to_rel(win[2 * p], org, box));
        if (2 * p + 1 < W) b = bw_record<FAST2>(to_rel(win[2 * p + 1], org, box));
        __syncthreads();
        if (2 * p < W) cand[p] = bw_pair(a, b);
    }
    __syncthreads();
}
// record k of the paired layout
__device__ __forceinline__ float4 bw_single(const unsigned char *raw, int k)
{
    const float *f = reinterpret_cast<const float *>(raw) + (size_t)(k >> 1) * 8 + (k & 1);
    return make_float4(f[0], f[2], f[4], f[6]);
}

// packed fp32 (sm_100): two lanes of a 64-bit register pair per instruction
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 x2_pack(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void x2_unpack(f32x2 v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 x2_add(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 x2_mul(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 x2_fma(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// entry `off` of a lane into the warp's row buffer (shared-memory address sptr), predicated like sc_append16
__device__ __forceinline__ void bw_append_s(bool hit, unsigned &sptr, int off)
{
    // the store's address in a register of its own: with `@q st [sptr]; @q add sptr, 64` the add waits until the
    // store unit has read sptr (short-scoreboard, 15 % of the loop's samples)
    const unsigned at = sptr;
    sptr = at + (hit ? 64u : 0u);
    if (hit) asm volatile("st.shared.b16 [%0], %1;" ::"r"(at), "h"((unsigned short)off) : "memory");
}
// the buffered rows of a warp -> the list, one row of 32 lanes per store.  Rows beyond a lane's own count carry
// leftovers: they land in slots of its row that are rewritten later or never read.
__device__ __forceinline__ void bw_flush(unsigned sbuf0, int lane, bool valid, int cnt, unsigned short *row16)
{
    int kmax = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
    if (valid) {
        const unsigned src = sbuf0 + 2u * (unsigned)lane;
#pragma unroll 4
        for (int k = 0; k < kmax; k++) {
            unsigned short v;
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(src + 64u * (unsigned)k));
            row16[(size_t)k * 32] = v;
        }
    }
    __syncwarp();
}

template <typename R, bool FAST2>
__global__ void __launch_bounds__(WIN_THREADS, BW_MINBLK)
k_build_list_win(Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0, const typename Pos4<R>::type *pos1,
                 Box<R> box, Grid g, const int *__restrict__ sbt, unsigned short *__restrict__ nbr16,
                 int *__restrict__ nnb, int cap, int kbuf)
{
    if (!ctl->halt_rebuild) return;
    typedef typename Pos4<R>::type P4;
    extern __shared__ __align__(128) unsigned char win_raw[];      // the window, then kbuf rows of 64 bytes per warp
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ int sP[WIN_RUNS + 1], sS[WIN_RUNS];
    __shared__ R s_rl2[B2_MAXSP * B2_MAXSP];
    __shared__ float s_hi[B2_MAXSP * B2_MAXSP], s_lo[B2_MAXSP * B2_MAXSP];
    const P4 *__restrict__ pos = ctl->cur ? pos1 : pos0;
    const int gi = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    const int *__restrict__ T = sbt + (size_t)gi * SBT_STRIDE;
    const int W = T[3];
    if (threadIdx.x == 0) atomicMax(&ctl->win_max, (unsigned)W);
    if (W > g.wcap || W > 65535) {                   // does not fit the shared memory: the host enlarges it
        if (threadIdx.x == 0) ctl->overflow = 1u;
        return;
    }
    if (T[1] == 0) return;
    const WinInfo wi = win_stage<R>(sbt, pos, reinterpret_cast<P4 *>(win_raw), &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp == 0) win_layout_load(T, lane, sP, sS);
    if (threadIdx.x < B2_MAXSP * B2_MAXSP) {
        const R rl2 = rl2_of(g_pair, threadIdx.x, (R)0);
        s_rl2[threadIdx.x] = rl2;
        // half-width of the band, relative to the pair's threshold: a coordinate (at most 4 cells from the
        // centre) is good to ~5e-7 of the cell width, r^2 to ~2e-6 r_list,max^2 in the worst case
        const float eps = SC_EPS * fmaxf(1.f, sqrtf(g.reach2 / fmaxf((float)rl2, 1e-30f)));
        s_hi[threadIdx.x] = (float)rl2 * (1.f + eps);
        s_lo[threadIdx.x] = (float)rl2 * (1.f - eps);
    }
    WinOrigin<R> org;
    win_origin_load(T, org);
    win_wait(&s_bar);
    bw_convert<FAST2>(win_raw, W, org, box);          // ends with a barrier: layout and thresholds are visible too
    const ulonglong2 *cand2 = reinterpret_cast<const ulonglong2 *>(win_raw);      // record pair p: entries 2p, 2p + 1
    const int nsp = g_pair.nsp;
    int sbx, sby, sbz;
    sb_coords(g, gi, sbx, sby, sbz);
    const unsigned sbuf0 = smem_u32(win_raw) + (unsigned)g.wcap * (unsigned)sizeof(P4) + 128u + (unsigned)(warp * kbuf) * 64u;
    for (int b = warp; b < 8; b += WIN_THREADS / 32) {
        const int ox = b & 1, oy = (b >> 1) & 1, oz = b >> 2;
        // the last super-block of an axis may be two cells wide
        if (4 * sbx + 2 * ox >= g.nc[0] || 4 * sby + 2 * oy >= g.nc[1] || 4 * sbz + 2 * oz >= g.nc[2]) continue;
        const int rx = 1 + ox, ry = 1 + oy, rz = 1 + oz;
        const int r_own = (rz * 4 + ry) * 4 + rx;
        const int own0 = sP[r_own], n_own = sP[r_own + 1] - own0, slot0 = sS[r_own];
        for (int b0 = 0; b0 < n_own; b0 += 32) {          // more than 32 particles in the block: again
            const bool valid = b0 + lane < n_own;
            const int self = own0 + (valid ? b0 + lane : 0);
            const int slot = slot0 + (self - own0);
            const P4 pi = pos[slot];
            const bool active = valid && !is_ghost(pi);   // ghosts act as neighbours but own no row
            const int ti = ptype(pi);
            const float4 me = bw_single(win_raw, self);
            const float px = me.x, py = me.y, pz = me.z;
            const float *hi_row = s_hi + ti * nsp, *lo_row = s_lo + ti * nsp;
            const float hi0 = active ? hi_row[0] : -1.f;
            const float dhi = active ? hi_row[nsp > 1 ? 1 : 0] - hi_row[0] : 0.f;
            const f32x2 npx = x2_pack(-px, -px), npy = x2_pack(-py, -py), npz = x2_pack(-pz, -pz);
            const f32x2 nhi0 = x2_pack(-hi0, -hi0), ndhi = x2_pack(-dhi, -dhi);
            float bandw = 0.f;                            // widest band of this particle's row of the pair table
            for (int tj = 0; tj < nsp; tj++) bandw = fmaxf(bandw, hi_row[tj] - lo_row[tj]);
            if (!active) bandw = -1.f;                    // (its thresholds are negative: t > 0 always)
            const int li = slot - wi.first;               // rows live in tiles local to the super-block
            unsigned short *row16 = nbr16 + ((size_t)(wi.tbase + (li >> 5)) * (size_t)cap) * 32 + (li & 31);
            // The listed pair closest to its threshold, with no select in the loop: t = r^2 - threshold is negative
            // for a listed pair, and among negative floats the one of smallest magnitude has the smallest bit
            // pattern read as a signed integer (pairs that are not listed have t >= 0, a non-negative integer).
            int nearbits = 0;
            // Entries go to the warp's buffer in shared memory first (rows of 32 lanes, as in the list) and from
            // there to the list one 64-byte row per store; a 2-byte store per entry straight into the list touches
            // a sector per lane, and the loop then waits for the store unit to release its operands (a quarter
            // of the loop's time, profiles/r02_ncu_k_build_list_win_v1.txt).  Once some lane has filled the
            // buffer, the rows are written out and every lane appends to the list directly (`direct`).
            const unsigned sptr0 = sbuf0 + 2u * (unsigned)lane;
            unsigned sptr = sptr0;                        // where this lane's next entry goes
            const unsigned sfull = sptr0 + 64u * (unsigned)(kbuf - 4);
            bool direct = kbuf < 8;                       // warp-uniform
            int cnt = 0;                                  // entries of this lane (direct mode)
#define BW_HIT(T, IDX, SELF)                                                          \
                const bool hit = SELF ? (T < 0.f && (IDX) != self) : T < 0.f;         \
                nearbits = min(nearbits, SELF && (IDX) == self ? 0 : __float_as_int(T));
#define BW_APPEND_S(IDX)   bw_append_s(hit, sptr, (IDX));
#define BW_APPEND_G(IDX)   sc_append16(hit, row16, cnt, cap, (IDX));
            // one record (the odd ends of a stretch)
#define BW_ONE(APPEND, K, SELF)                                                       \
            {                                                                         \
                const float4 q = bw_single(win_raw, (K));                             \
                const float dx = q.x - px, dy = q.y - py, dz = q.z - pz;              \
                const float hi = FAST2 ? fmaf(q.w, dhi, hi0)                          \
                                       : (active ? hi_row[__float_as_int(q.w)] : -1.f); \
                const float t = dx * dx + dy * dy + dz * dz - hi;                     \
                BW_HIT(t, (K), SELF) APPEND((K))                                      \
            }
            // the two records of a pair: (xy, zw) as they come from shared memory
#define BW_TWO(APPEND, XY, ZW, K, SELF)                                               \
            {                                                                         \
                const f32x2 dx = x2_add(XY.x, npx), dy = x2_add(XY.y, npy), dz = x2_add(ZW.x, npz); \
                const f32x2 r2 = x2_fma(dz, dz, x2_fma(dy, dy, x2_mul(dx, dx)));      \
                float ta, tb;                                                         \
                if (FAST2) {                                                          \
                    x2_unpack(x2_fma(ZW.y, ndhi, x2_add(r2, nhi0)), ta, tb);          \
                } else {                                                              \
                    float wa, wb;                                                     \
                    x2_unpack(r2, ta, tb);                                            \
                    x2_unpack(ZW.y, wa, wb);                                          \
                    ta -= active ? hi_row[__float_as_int(wa)] : -1.f;                 \
                    tb -= active ? hi_row[__float_as_int(wb)] : -1.f;                 \
                }                                                                     \
                { BW_HIT(ta, (K), SELF) APPEND((K)) }                                 \
                { BW_HIT(tb, (K) + 1, SELF) APPEND((K) + 1) }                         \
            }
            // two pairs per round, the next two already on their way (reads run up to two pairs past the
            // stretch: into the next run, or the slack behind the window); k is even here.  ROUNDS: how many rounds
            // may run before the buffer has to be looked at again (a round appends at most four entries per lane).
#define BW_STRETCH(APPEND, SELF, ROUNDS)                                              \
            {                                                                         \
                while (k + 4 <= end) {                                                \
                    int rounds = (end - k) >> 2;                                      \
                    ROUNDS                                                            \
                    const ulonglong2 *pp = cand2 + k;                                 \
                    ulonglong2 a0 = pp[0], a1 = pp[1], b0_ = pp[2], b1_ = pp[3];      \
                    k += 4 * rounds;                                                  \
                    int kk = k - 4 * rounds;                                          \
                    _Pragma("unroll 1")                                               \
                    for (; kk < k; kk += 4) {                                         \
                        pp += 4;                                                      \
                        const ulonglong2 n0 = pp[0], n1 = pp[1], n2 = pp[2], n3 = pp[3]; \
                        BW_TWO(APPEND, a0, a1, kk, SELF) BW_TWO(APPEND, b0_, b1_, kk + 2, SELF) \
                        a0 = n0; a1 = n1; b0_ = n2; b1_ = n3;                         \
                    }                                                                 \
                }                                                                     \
            }
#define BW_TAIL(APPEND, SELF)                                                         \
                _Pragma("unroll 1")                                                   \
                for (; k < end; k++) BW_ONE(APPEND, k, SELF)
            // rounds the fullest lane's buffer still takes: (rows left - 4) / 4 + 1, none once it is within four of the end
#define BW_ROOM   (sptr > sfull ? 0u : ((sfull - sptr) >> 8) + 1u)
#define BW_FULL   { const int room = (int)__reduce_min_sync(0xffffffffu, BW_ROOM); if (room == 0) break; rounds = min(rounds, room); }
#define BW_WALK(APPEND, SELF, GUARD)                                                  \
                {                                                                     \
                    if ((k & 1) && k < end) {                                         \
                        BW_ONE(APPEND, k, SELF)                                       \
                        k++;                                                          \
                    }                                                                 \
                    BW_STRETCH(APPEND, SELF, GUARD)                                   \
                }
#pragma unroll 1
            for (int s9 = 0; s9 < 9; s9++) {
                const int r0 = ((rz - 1 + s9 / 3) * 4 + (ry - 1 + s9 % 3)) * 4 + rx - 1;
                const int end = sP[r0 + 3];
                int k = sP[r0];
                if (!direct) {
                    // (a lane at most four entries short of the buffer's end stops the buffered walk: the checks
                    // come before the odd record at the front, before every round of four and before the last three)
                    if (!__any_sync(0xffffffffu, sptr > sfull)) {
                        if (s9 == 4) BW_WALK(BW_APPEND_S, true, BW_FULL)     // the stretch that holds the block itself
                        else BW_WALK(BW_APPEND_S, false, BW_FULL)
                    }
                    if (k + 4 > end && !__any_sync(0xffffffffu, sptr > sfull)) {
                        if (s9 == 4) BW_TAIL(BW_APPEND_S, true)
                        else BW_TAIL(BW_APPEND_S, false)
                    } else {
                        // rows out, and on in the list itself
                        cnt = (int)(sptr - sptr0) >> 6;
                        bw_flush(sbuf0, lane, valid, cnt, row16);
                        direct = true;
                    }
                }
                if (direct) {
                    if (s9 == 4) { BW_WALK(BW_APPEND_G, true, ) BW_TAIL(BW_APPEND_G, true) }
                    else { BW_WALK(BW_APPEND_G, false, ) BW_TAIL(BW_APPEND_G, false) }
                }
            }
#undef BW_WALK
#undef BW_FULL
#undef BW_ROOM
#undef BW_TAIL
#undef BW_STRETCH
#undef BW_TWO
#undef BW_ONE
#undef BW_APPEND_G
#undef BW_APPEND_S
#undef BW_HIT
            if (!direct) {
                cnt = (int)(sptr - sptr0) >> 6;
                bw_flush(sbuf0, lane, valid, cnt, row16);
            }
            if (nearbits < 0 && __int_as_float(nearbits) > -bandw) {
                // Some listed pair may sit inside the band around its threshold: settle those with the
                // canonical predicate and close the gaps (about one particle in a thousand).  The row is read
                // back eight entries at a time: the loads of a batch are in flight together.
                const int lim = min(cnt, cap);
                int kept = 0;
                for (int k0 = 0; k0 < lim; k0 += 8) {
                    int nm[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) nm[u] = k0 + u < lim ? (int)row16[(size_t)(k0 + u) * 32] : 0;
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        if (k0 + u >= lim) break;
                        const float4 q = bw_single(win_raw, nm[u]);
                        const int tj = FAST2 ? (int)q.w : __float_as_int(q.w);
                        const float dx = q.x - px, dy = q.y - py, dz = q.z - pz;
                        bool keep = true;
                        if (dx * dx + dy * dy + dz * dz > lo_row[tj])
                            keep = r2_canon(pi, pos[win_slot_of(sP, sS, nm[u])], box) < s_rl2[ti * nsp + tj];
                        if (keep) {
                            row16[(size_t)kept * 32] = (unsigned short)nm[u];     // kept <= k0 + u: behind the reads
                            kept++;
                        }
                    }
                }
                if (cnt <= cap) cnt = kept;
            }
            if (valid) {
                nnb[slot] = min(cnt, cap);
                if (cnt > cap) ctl->overflow = 1u;
            }
            int mx = valid ? cnt : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            if (lane == 0) atomicMax(&ctl->max_count, (unsigned)mx);
        }
    }
}

// reciprocal of the step kernels in fp32: the hardware approximation alone (1 ulp; the forces of a step are never
// reported, and the neighbour's coordinates carry a comparable rounding already).  fp64 keeps the full sequence.
template <bool FAST> __device__ __forceinline__ float win_rcp(float x)
{
    if (!FAST) return r_rcp(x);
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
template <bool FAST> __device__ __forceinline__ double win_rcp(double x) { return r_rcp(x); }
#undef B2_RCP
#define B2_RCP(x) win_rcp<FASTRCP>(x)

// coefficients of pair (a) as the force loop uses them, from the constant table
__device__ __forceinline__ void pair_coef(int a, float &a12, float &a6, float &rc2)
{
    a12 = g_pair.f12[a]; a6 = g_pair.f6[a]; rc2 = g_pair.frc2[a];
}
__device__ __forceinline__ void pair_coef(int a, double &a12, double &a6, double &rc2)
{
    a12 = 12.0 * g_pair.c12[a]; a6 = 6.0 * g_pair.c6[a]; rc2 = g_pair.rc2[a];
}

template <typename R, bool EV> struct WinUnroll {
    static constexpr int U = (sizeof(R) == 4 && !EV) ? WIN_UNROLL32 : WIN_UNROLL64;
};

// the first U entries of a row, requested ahead of use.  Reads may run up to U rows past the lane's
// count (into the next tile's rows, or the slack behind the list): such entries are never used.
template <int U>
__device__ __forceinline__ void win_row_begin(const unsigned short *__restrict__ row, unsigned (&jj)[U])
{
#pragma unroll
    for (int u = 0; u < U; u++) jj[u] = ld_stream_u16(row + (size_t)u * 32);
}

// One particle against its row of window offsets.  P is the record type of the window: the stored
// positions (force kernels: exact arithmetic, observables in fp64) or relative coordinates (step
// kernels).  `jj` holds the row's first U entries (win_row_begin).  Pair arithmetic: B2_PAIR.
template <typename R, typename P, bool EV, bool STRESS, bool FAST2, bool FASTRCP>
__device__ __forceinline__ void pair_row_win(const P *win, const unsigned short *__restrict__ row, int cnt,
                                             unsigned (&jj)[WinUnroll<R, EV>::U],
                                             P pi, const Box<R> &box, const PairTab<R> &tab, int nsp,
                                             R &fx, R &fy, R &fz, double &pe, double &vir, double *st)
{
    constexpr int U = WinUnroll<R, EV>::U;
    constexpr bool TAB = false;          // tabulated potentials run on the global-index list
    const int ti = ptype(pi) * nsp;
    R a12_0 = 0, a6_0 = 0, rc_0 = 0, a12_1 = 0, a6_1 = 0, rc_1 = 0;
    if (FAST2) {
        pair_coef(ti, a12_0, a6_0, rc_0);
        pair_coef(ti + nsp - 1, a12_1, a6_1, rc_1);
    }
    int k = 0;
    for (; k + U <= cnt; k += U) {
        P pp[U];
#pragma unroll
        for (int u = 0; u < U; u++) pp[u] = win[jj[u]];
#ifdef WIN_EXPERIMENT_DOUBLE_LDS       // timing experiment only (same physics): every record is read twice
#pragma unroll
        for (int u = 0; u < U; u++) {
            unsigned d0, d1, d2, d3;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(d0), "=r"(d1), "=r"(d2), "=r"(d3)
                         : "r"(smem_u32(win + jj[u])));
        }
#endif
#pragma unroll
        for (int u = 0; u < U; u++) jj[u] = ld_stream_u16(row + (size_t)(k + U + u) * 32);
#pragma unroll
        for (int u = 0; u < U; u++) B2_PAIR(pp[u])
    }
    // fewer than U entries left: their offsets are here already
#pragma unroll
    for (int u = 0; u < U - 1; u++)
        if (k + u < cnt) {
            const P p0 = win[jj[u]];
            B2_PAIR(p0)
        }
}
#undef B2_PAIR
#undef B2_RCP

template <typename R> struct WinArgs {
    Grid g;
    const int *sbt;
    const unsigned short *nbr16;
};

// The fused velocity-Verlet step of k_step_nve on the windowed list: one block per super-block
// (3-D grid of super-blocks).
template <typename R, bool NVT>
__global__ void __launch_bounds__(WIN_THREADS, (sizeof(R) == 4 ? WIN_MINBLK32 : WIN_MINBLK64))
k_step_win(StepArgs<R> a, WinArgs<R> wa)
{
    typedef typename Vec4<R>::type V4;
    typedef typename Pos4<R>::type P4;
    typedef typename Rel4<R>::type Q4;
    constexpr int U = WinUnroll<R, false>::U;
    static_assert(sizeof(P4) == sizeof(Q4), "relative records replace the stored ones in place");
    extern __shared__ __align__(128) unsigned char win_raw[];
    Ctl *ctl = a.ctl;
    if (ctl->halt_rebuild | ctl->overflow | ctl->done) {
        if (a.dd && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0) {
            *a.vmax_out = 0.f;
            if (a.peer_commit_state) { a.peer_commit_state[1] = 0u; *a.peer_gv_out = 0.f; }     // (as k_peer_commit)
        }
        return;
    }
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ PairTab<R> tab;                        // more than two species only
    const int nsp = g_pair.nsp;
    if (nsp > 2) load_pair_table(tab);
    const unsigned cur = ctl->cur;
    const P4 *__restrict__ pos = cur ? a.pos[1] : a.pos[0];
    P4 *__restrict__ posn = cur ? a.pos[0] : a.pos[1];
    P4 *win = reinterpret_cast<P4 *>(win_raw);
    Q4 *rel = reinterpret_cast<Q4 *>(win_raw);
    const WinInfo wi = win_stage<R>(wa.sbt, pos, win, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntile = (wi.cnt + 31) >> 5;
    // the first tile's own data, on its way while the window arrives
    int t = warp;
    bool valid = t < ntile && t * 32 + lane < wi.cnt;
    int i = wi.first + t * 32 + lane;
    const unsigned short *row = wa.nbr16 + ((size_t)(wi.tbase + t) * (size_t)a.cap) * 32 + lane;
    P4 pi;
    int cnt = 0;
    unsigned jj[U];
    if (valid) {
        pi = pos[i];
        cnt = a.nnb[i];
        win_row_begin<U>(row, jj);
    }
    WinOrigin<R> org;
    win_origin_load(wa.sbt + (size_t)((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * SBT_STRIDE, org);
    win_wait(&s_bar);
    for (int k = threadIdx.x; k < wi.W; k += WIN_THREADS) {
        const P4 p = win[k];
        rel[k] = to_rel(p, org, a.box);          // same address, same thread
    }
    __syncthreads();
    const R th_a = NVT ? (R)ctl->th_a : (R)1, th_b = NVT ? (R)ctl->th_b : (R)1;
    float v2 = 0.f;
    double mv2 = 0.0;
    bool pushed = false;
    for (;;) {
        if (valid) {
            if (is_ghost(pi)) {
                posn[i] = pi;
            } else {
                R fx = 0, fy = 0, fz = 0;
                double pe = 0, vir = 0;
                const Q4 qi = to_rel(pi, org, a.box);
                if (nsp <= 2) pair_row_win<R, Q4, false, false, true, true>(rel, row, cnt, jj, qi, a.box, tab, nsp, fx, fy, fz, pe, vir, nullptr);
                else pair_row_win<R, Q4, false, false, false, true>(rel, row, cnt, jj, qi, a.box, tab, nsp, fx, fy, fz, pe, vir, nullptr);
                V4 v = a.vel[i];
                R s = a.kick / v.w;
                if (NVT) {
                    s *= th_b;
                    v.x = th_a * v.x + s * fx; v.y = th_a * v.y + s * fy; v.z = th_a * v.z + s * fz;
                    mv2 += (double)v.w * ((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
                } else {
                    v.x += s * fx; v.y += s * fy; v.z += s * fz;
                }
                int kx, ky, kz;
                const P4 pnew = advance(pi, a.drift * v.x, a.drift * v.y, a.drift * v.z, a.box, kx, ky, kz);
                posn[i] = pnew;
                if (a.halo_map) pushed |= peer_push_one(a, i, pnew);
                if (kx | ky | kz) {
                    int *im = a.img + 3 * (size_t)a.orig[i];
                    im[0] += kx; im[1] += ky; im[2] += kz;
                }
                a.vel[i] = v;
                v2 = fmaxf(v2, __double2float_ru((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z));
            }
        }
        t += WIN_THREADS / 32;
        if (t >= ntile) break;
        valid = t * 32 + lane < wi.cnt;
        i = wi.first + t * 32 + lane;
        row = wa.nbr16 + ((size_t)(wi.tbase + t) * (size_t)a.cap) * 32 + lane;
        if (valid) {
            pi = pos[i];
            cnt = a.nnb[i];
            win_row_begin<U>(row, jj);
        }
    }
    step_epilogue<R, NVT, WIN_THREADS>(a, ctl, cur, v2, mv2, pushed);
}

// Forces (+ per-particle energy / virial, + stress partials) on the windowed list: k_force's twin.
// The window keeps the stored positions: reported forces and observables use the exact arithmetic.
template <typename R, bool EV, bool STRESS>
__global__ void __launch_bounds__(WIN_THREADS)
k_force_win(ForceArgs<R> a, WinArgs<R> wa)
{
    typedef typename Pos4<R>::type P4;
    constexpr int U = WinUnroll<R, EV>::U;
    extern __shared__ __align__(128) unsigned char win_raw[];
    const Ctl *ctl = a.ctl;
    if (a.guarded && (ctl->halt_rebuild | ctl->overflow | ctl->done)) return;
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ PairTab<R> tab;
    const int nsp = g_pair.nsp;
    if (nsp > 2) load_pair_table(tab);
    const P4 *__restrict__ pos = ctl->cur ? a.pos[1] : a.pos[0];
    P4 *win = reinterpret_cast<P4 *>(win_raw);
    const WinInfo wi = win_stage<R>(wa.sbt, pos, win, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntile = (wi.cnt + 31) >> 5;
    win_wait(&s_bar);
    __syncthreads();                 // the table of the many-species path
    double st[6] = {0, 0, 0, 0, 0, 0};
    for (int t = warp; t < ntile; t += WIN_THREADS / 32) {
        const int li = t * 32 + lane;
        if (li >= wi.cnt) continue;
        const int i = wi.first + li;
        const P4 pi = pos[i];
        R fx = 0, fy = 0, fz = 0;
        double pe = 0, vir = 0;
        const unsigned short *row = wa.nbr16 + ((size_t)(wi.tbase + t) * (size_t)a.cap) * 32 + lane;
        const int cnt = a.nnb[i];
        unsigned jj[U];
        win_row_begin<U>(row, jj);
        if (nsp <= 2) pair_row_win<R, P4, EV, STRESS, true, false>(win, row, cnt, jj, pi, a.box, tab, nsp, fx, fy, fz, pe, vir, st);
        else pair_row_win<R, P4, EV, STRESS, false, false>(win, row, cnt, jj, pi, a.box, tab, nsp, fx, fy, fz, pe, vir, st);
        a.frc[i] = make4(fx, fy, fz, (R)(0.5 * pe));
        if (EV) a.vir[i] = (R)(0.5 * vir);
    }
    if (STRESS) {
        __shared__ double s_st[WIN_THREADS / 32][6];
#pragma unroll
        for (int q = 0; q < 6; q++) {
            double v = warp_sum(st[q]);
            if (lane == 0) s_st[warp][q] = v;
        }
        __syncthreads();
        if (threadIdx.x < 6) {
            double v = 0;
            for (int ww = 0; ww < WIN_THREADS / 32; ww++) v += s_st[ww][threadIdx.x];
            const size_t bid = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
            a.partial[bid * 6 + threadIdx.x] = 0.5 * v;
        }
    }
}

// export of the windowed list in original numbering (b2_neighbors): one block per super-block
__global__ void __launch_bounds__(WIN_THREADS)
k_export_list_win(const int *__restrict__ sbt, const unsigned short *__restrict__ nbr16, const int *__restrict__ nnb,
                  const int *__restrict__ orig, int cap, int stride, int *__restrict__ ocount, int *__restrict__ onbr)
{
    __shared__ int sP[WIN_RUNS + 1], sS[WIN_RUNS];
    const int gi = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    const int *__restrict__ T = sbt + (size_t)gi * SBT_STRIDE;
    if (threadIdx.x < 32) win_layout_load(T, threadIdx.x, sP, sS);
    __syncthreads();
    const int first = T[0], cnt = T[1], tbase = T[2];
    for (int li = threadIdx.x; li < cnt; li += blockDim.x) {
        const int s = first + li, o = orig[s], c = nnb[s];
        ocount[o] = c;
        if (!onbr) continue;
        const unsigned short *row = nbr16 + ((size_t)(tbase + (li >> 5)) * (size_t)cap) * 32 + (li & 31);
        for (int k = 0; k < c; k++)
            onbr[(size_t)o * stride + k] = orig[win_slot_of(sP, sS, (int)row[(size_t)k * 32])];
    }
}
