// Neighbour search: Morton cell binning, LSD radix sort, cell table, Verlet list.
//
// All kernels of the refresh pipeline are guarded by Ctl::halt_rebuild: the
// host enqueues them speculatively between batches of MD steps and they do
// work only when the displacement bound has invalidated the list.
#pragma once
#include "b2_common.cuh"

#define SORT_THREADS 256
#define SORT_WARPS 8
#define SORT_ITEMS 16                                  // keys per thread
#define SORT_TILE (SORT_THREADS * SORT_ITEMS)          // keys per block
#define SORT_SEG (SORT_TILE / SORT_WARPS)              // contiguous keys per warp

// ---- keys -----------------------------------------------------------------
template <typename R>
__global__ void k_cell_keys(Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0,
                            const typename Pos4<R>::type *pos1, int n, Box<R> box, Grid g,
                            uint32_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    if (!ctl->halt_rebuild) return;
    const typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Pos4<R>::type p = pos[i];
    int cx, cy, cz;
    cell_of(box, g, p, cx, cy, cz);
    keys[i] = grid_key(g, 0, cx, cy, cz);
    vals[i] = (uint32_t)i;
    if (g.win) {
        // a particle outside the launched super-blocks would have no row: the host then launches them all
        int lx = (cx >> 2) - g.sb0[0], ly = (cy >> 2) - g.sb0[1], lz = (cz >> 2) - g.sb0[2];
        lx += lx < 0 ? g.nsb[0] : 0; ly += ly < 0 ? g.nsb[1] : 0; lz += lz < 0 ? g.nsb[2] : 0;
        if (lx >= g.sbn[0] || ly >= g.sbn[1] || lz >= g.sbn[2]) {
            ctl->region_miss = 1u;
            ctl->overflow = 1u;
        }
    }
}

// ---- radix sort, one 8-bit digit per pass -----------------------------------
__global__ void __launch_bounds__(SORT_THREADS)
k_sort_hist(const Ctl *__restrict__ ctl, const uint32_t *__restrict__ keys, int n, int shift,
            int nblk, int *__restrict__ hist)
{
    if (!ctl->halt_rebuild) return;
    __shared__ int sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * SORT_TILE;
#pragma unroll 4
    for (int it = 0; it < SORT_ITEMS; it++) {
        int idx = base + it * SORT_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&sh[(keys[idx] >> shift) & 255u], 1);
    }
    __syncthreads();
    hist[blockIdx.x * 256 + threadIdx.x] = sh[threadIdx.x];    // block-major: coalesced
}

// hist[b][d] -> start offset of digit d of block b in the sorted output
// (digit-major order: all blocks of digit 0, then digit 1, ...).  One block of
// 1024 threads: four groups of 256 digits, each group sweeps a quarter of the
// blocks with coalesced rows, eight independent loads in flight per thread.
#define SCAN_GROUPS 4
__global__ void __launch_bounds__(256 * SCAN_GROUPS)
k_sort_scan(const Ctl *__restrict__ ctl, int *hist, int nblk)
{
    if (!ctl->halt_rebuild) return;
    __shared__ int s_tot[SCAN_GROUPS][256];
    __shared__ int s_warp[8];
    const int d = threadIdx.x & 255, grp = threadIdx.x >> 8, lane = d & 31, w = d >> 5;
    const int per = (nblk + SCAN_GROUPS - 1) / SCAN_GROUPS;
    const int b0 = grp * per, b1 = min(nblk, b0 + per);
    int run = 0;
    for (int b = b0; b < b1; b += 8) {
        int t[8];
#pragma unroll
        for (int k = 0; k < 8; k++) t[k] = b + k < b1 ? hist[(b + k) * 256 + d] : 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (b + k < b1) hist[(b + k) * 256 + d] = run;
            run += t[k];
        }
    }
    s_tot[grp][d] = run;
    __syncthreads();
    // exclusive scan of the per-digit totals (first group), then each group's base
    int total = 0, before = 0;
#pragma unroll
    for (int g = 0; g < SCAN_GROUPS; g++) {
        const int t = s_tot[g][d];
        if (g < grp) before += t;
        total += t;
    }
    int incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (grp == 0 && lane == 31) s_warp[w] = incl;
    __syncthreads();
    int base = incl - total + before;
    for (int ww = 0; ww < w; ww++) base += s_warp[ww];
    for (int b = b0; b < b1; b += 8) {
        int t[8];
#pragma unroll
        for (int k = 0; k < 8; k++) t[k] = b + k < b1 ? hist[(b + k) * 256 + d] : 0;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (b + k < b1) hist[(b + k) * 256 + d] = t[k] + base;
    }
}

// stable scatter: each warp owns a contiguous segment of the block's tile
__global__ void __launch_bounds__(SORT_THREADS)
k_sort_scatter(const Ctl *__restrict__ ctl, const uint32_t *__restrict__ kin,
               const uint32_t *__restrict__ vin, uint32_t *__restrict__ kout,
               uint32_t *__restrict__ vout, int n, int shift, int nblk,
               const int *__restrict__ hist)
{
    if (!ctl->halt_rebuild) return;
    __shared__ int wh[SORT_WARPS][256];
    int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int k = tid; k < SORT_WARPS * 256; k += SORT_THREADS) ((int *)wh)[k] = 0;
    __syncthreads();
    int seg = blockIdx.x * SORT_TILE + w * SORT_SEG;
    uint32_t key[SORT_ITEMS];
    // pass 1: per-warp digit histogram
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; it++) {
        int idx = seg + it * 32 + lane;
        bool ok = idx < n;
        key[it] = ok ? kin[idx] : 0u;
        unsigned d = ok ? ((key[it] >> shift) & 255u) : 256u;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        if (ok && lane == __ffs(peers) - 1) wh[w][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    {   // exclusive prefix over the warps, on top of this block's global base
        int d = tid;
        int run = hist[blockIdx.x * 256 + d];
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ww++) {
            int t = wh[ww][d];
            wh[ww][d] = run;
            run += t;
        }
    }
    __syncthreads();
    // pass 2: ranks within the warp keep the order of equal digits
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; it++) {
        int idx = seg + it * 32 + lane;
        bool ok = idx < n;
        unsigned d = ok ? ((key[it] >> shift) & 255u) : 256u;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        int off = 0;
        if (ok && lane == leader) {
            off = wh[w][d];
            wh[w][d] = off + __popc(peers);
        }
        off = __shfl_sync(0xffffffffu, off, leader);
        int rank = __popc(peers & ((1u << lane) - 1u));
        if (ok) {
            kout[off + rank] = key[it];
            vout[off + rank] = vin[idx];
        }
        __syncwarp();
    }
}

// ---- permute the particle arrays into sorted order ---------------------------
template <typename R>
__global__ void k_gather(const Ctl *__restrict__ ctl, const uint32_t *__restrict__ perm, int n,
                         const typename Pos4<R>::type *pos0, const typename Pos4<R>::type *pos1,
                         const typename Vec4<R>::type *__restrict__ vel,
                         const int *__restrict__ orig,
                         typename Pos4<R>::type *__restrict__ tpos,
                         typename Vec4<R>::type *__restrict__ tvel, int *__restrict__ torig)
{
    if (!ctl->halt_rebuild) return;
    const typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    uint32_t src = perm[s];
    tpos[s] = pos[src];
    tvel[s] = vel[src];
    torig[s] = orig[src];
}

template <typename R>
__global__ void k_gather_commit(const Ctl *__restrict__ ctl, int n,
                                const typename Pos4<R>::type *__restrict__ tpos,
                                const typename Vec4<R>::type *__restrict__ tvel,
                                const int *__restrict__ torig,
                                typename Pos4<R>::type *pos0, typename Pos4<R>::type *pos1,
                                typename Vec4<R>::type *__restrict__ vel, int *__restrict__ orig)
{
    if (!ctl->halt_rebuild) return;
    typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    pos[s] = tpos[s];
    vel[s] = tvel[s];
    orig[s] = torig[s];
}

// ---- cell table ------------------------------------------------------------
__global__ void k_fill_int(const Ctl *__restrict__ ctl, int *__restrict__ a, int *__restrict__ b,
                           int count, int value)
{
    if (!ctl->halt_rebuild) return;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) {
        a[i] = value;
        b[i] = value;
    }
}

__global__ void k_cell_bounds(const Ctl *__restrict__ ctl, const uint32_t *__restrict__ keys, int n,
                              int *__restrict__ cell_start, int *__restrict__ cell_end)
{
    if (!ctl->halt_rebuild) return;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    uint32_t k = keys[s];
    if (s == 0 || keys[s - 1] != k) cell_start[k] = s;
    if (s == n - 1 || keys[s + 1] != k) cell_end[k] = s + 1;
}

// ---- Verlet list -----------------------------------------------------------
// Membership: r2 < (rcut_ab + skin)^2, strict.  r2 is the CANONICAL squared
// minimum-image distance (r2_canon): correctly rounded and uncontracted —
//   fp64: d - rint(d/L)*L (atooms/system/particle.py:129-130), (dx*dx + dy*dy) + dz*dz
//   fp32: d = float(wrapped int difference) * float(L / 2^32), same sum
// The cheap contracted estimate decides every pair outside a thin band around
// the threshold; the band is re-evaluated canonically, so the result is
// bit-identical to evaluating the canonical expression everywhere.
template <typename R> struct PredEps;
template <> struct PredEps<float>  { static constexpr float  v = 2e-5f; };
template <> struct PredEps<double> { static constexpr double v = 1e-11; };

template <typename R>
__device__ __forceinline__ bool in_list(const Box<R> &box, const typename Pos4<R>::type &pi,
                                        const typename Pos4<R>::type &pj, R rl2)
{
    R dx, dy, dz;
    delta(pi, pj, box, dx, dy, dz);
    R r2 = dx * dx + dy * dy + dz * dz;
    if (r2 > rl2 * ((R)1 + PredEps<R>::v)) return false;
    if (r2 < rl2 * ((R)1 - PredEps<R>::v)) return true;
    return r2_canon(pi, pj, box) < rl2;
}

__device__ __forceinline__ float  rl2_of(const PairConst &pc, int a, float)  { return pc.rl2_f[a]; }
__device__ __forceinline__ double rl2_of(const PairConst &pc, int a, double) { return pc.rl2_d[a]; }


// List layout: warp tile t = slot / 32 owns `cap` rows of 32 lanes:
//     nbr[(t * cap + k) * 32 + lane]   = k-th neighbour of slot t*32 + lane
// so a warp of the force kernel reads one coalesced 128-byte row per k.
//
// Cells are r_list,max / M wide and the stencil reaches M cells (M = 2 whenever
// the box allows it); cells whose nearest point is farther than r_list,max from
// the particle are skipped.  (A warp-cooperative walk over the union of the 32
// stencils — uniform candidate loads, no ragged cells — was measured 2x SLOWER:
// a Morton tile of 32 particles spans ~4 cells per axis, so the union holds
// ~1700 candidates against ~410 per lane here.)
__device__ __forceinline__ int wrap_cell(int c, int nc) { return c < 0 ? c + nc : (c >= nc ? c - nc : c); }

template <typename R, int M>
__global__ void __launch_bounds__(128)
k_build_list(Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0,
             const typename Pos4<R>::type *pos1, int n, Box<R> box, Grid g,
             const int *__restrict__ cell_start, const int *__restrict__ cell_end,
             int *__restrict__ nbr, int *__restrict__ nnb, int cap)
{
    if (!ctl->halt_rebuild) return;
    typedef typename Pos4<R>::type P4;
    const P4 *__restrict__ pos = ctl->cur ? pos1 : pos0;
    __shared__ R s_rl2[B2_MAXSP * B2_MAXSP];
    if (threadIdx.x < B2_MAXSP * B2_MAXSP) s_rl2[threadIdx.x] = rl2_of(g_pair, threadIdx.x, (R)0);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (((blockIdx.x * blockDim.x + threadIdx.x) & ~31) >= n) return;     // whole warp out of range
    const int nsp = g_pair.nsp;
    const P4 pi = pos[i < n ? i : n - 1];
    const bool active = i < n && !is_ghost(pi);      // ghosts act as neighbours but own no row
    const int ti = ptype(pi);
    const R *rl2_row = s_rl2 + ti * nsp;
    int *row = nbr + ((size_t)(i >> 5) * cap) * 32 + (i & 31);
    int cnt = 0;
    // the thin band around the threshold that must be decided canonically (see in_list)
    const R band_lo = (R)1 - PredEps<R>::v, band_hi = (R)1 + PredEps<R>::v;
#define B2_TEST(J)                                                        \
    {                                                                     \
        const P4 pj = pos[J];                                             \
        const R rl2 = rl2_row[ptype(pj)];                                 \
        R dx_, dy_, dz_;                                                  \
        delta(pi, pj, box, dx_, dy_, dz_);                                \
        const R r2_ = dx_ * dx_ + dy_ * dy_ + dz_ * dz_;                  \
        bool in_ = r2_ < rl2 * band_lo;                                   \
        if (!in_ && r2_ <= rl2 * band_hi) in_ = r2_canon(pi, pj, box) < rl2; \
        if (in_ && active && (J) != i) {                                  \
            if (cnt < cap) row[(size_t)cnt * 32] = (J);                   \
            cnt++;                                                        \
        }                                                                 \
    }
    if (M == 0) {       // all pairs
        for (int j = 0; j < n; j++) B2_TEST(j)
    } else {
        int cx, cy, cz;
        float fx, fy, fz;
        cell_frac(box, g, pi, cx, cy, cz, fx, fy, fz);
        // Morton bits of the 2M+1 x-columns of the stencil, once per particle
        uint32_t sx[2 * M + 1];
        float gx2[2 * M + 1];
#pragma unroll
        for (int dx = -M; dx <= M; dx++) {
            sx[dx + M] = grid_key_x(g, wrap_cell(cx + dx, g.nc[0]));
            const float gx = cell_gap(dx, fx) * g.w[0];
            gx2[dx + M] = gx * gx;
        }
        for (int dz = -M; dz <= M; dz++) {
            const float gz = cell_gap(dz, fz) * g.w[2];
            const float gz2 = gz * gz;
            if (gz2 > g.reach2) continue;
            const uint32_t sz = spread3((uint32_t)wrap_cell(cz + dz, g.nc[2])) << 2;
            for (int dy = -M; dy <= M; dy++) {
                const float gy = cell_gap(dy, fy) * g.w[1];
                const float gyz2 = gz2 + gy * gy;
                if (gyz2 > g.reach2) continue;
                const uint32_t syz = sz | (spread3((uint32_t)wrap_cell(cy + dy, g.nc[1])) << 1);
#pragma unroll
                for (int dx = 0; dx <= 2 * M; dx++) {
                    if (gyz2 + gx2[dx] > g.reach2) continue;
                    const uint32_t key = syz | sx[dx];
                    const int s = cell_start[key], e = cell_end[key];
                    for (int j = s; j < e; j++) B2_TEST(j)
                }
            }
        }
    }
#undef B2_TEST
    if (i < n) {
        nnb[i] = min(cnt, cap);
        if (cnt > cap) ctl->overflow = 1u;
        atomicMax(&ctl->max_count, (unsigned)cnt);
    }
}

// ---- Verlet list, one warp per 2x2x2 block of cells -------------------------------
// With M = 2 the eight cells of an aligned 2x2x2 block are consecutive Morton keys, so
// their particles are one contiguous run of slots (~26 for KA at rho = 1.2) and they
// all share the same candidate set: the 6x6x6 cells around the block.  One warp takes
// one block, one lane per particle; the candidates are staged in shared memory as
// float coordinates relative to the block's corner and every lane walks the SAME
// sequence with broadcast reads: no gathers and no ragged per-lane cell loops (the
// per-particle walk above keeps 14 of 32 lanes busy).  The float distance decides
// every pair outside a thin band around the threshold; the band is settled with the
// canonical predicate on the stored coordinates, exactly as in k_build_list.
#ifndef SC_WARPS
#define SC_WARPS 4
#endif
#ifndef SC_MINBLK
#define SC_MINBLK 5             // up to 102 registers per thread (7 blocks of 72 registers, with spills, timed the same)
#endif
#define SC_CHUNK 256
#define SC_CELLS 216
#define SC_EPS 4e-6f
#define SC_TAKE 248            // candidates staged per round: 31 lanes x 8, leaves a gap of 8 in the chunk

__device__ __forceinline__ uint32_t compact3(uint32_t x)
{
    x &= 0x09249249u;
    x = (x | (x >> 2)) & 0x030c30c3u;
    x = (x | (x >> 4)) & 0x0300f00fu;
    x = (x | (x >> 8)) & 0x030000ffu;
    x = (x | (x >> 16)) & 0x3ffu;
    return x;
}

struct ScOrigin { unsigned u[3]; double d[3]; };

__device__ __forceinline__ void rel_to(const int4 &p, const Box<float> &box, const ScOrigin &o,
                                       float &x, float &y, float &z)
{
    x = (float)(int)(((unsigned)p.x + 0x80000000u) - o.u[0]) * box.s[0];
    y = (float)(int)(((unsigned)p.y + 0x80000000u) - o.u[1]) * box.s[1];
    z = (float)(int)(((unsigned)p.z + 0x80000000u) - o.u[2]) * box.s[2];
}
__device__ __forceinline__ void rel_to(const double4 &p, const Box<double> &box, const ScOrigin &o,
                                       float &x, float &y, float &z)
{
    double dx = p.x - o.d[0], dy = p.y - o.d[1], dz = p.z - o.d[2];
    x = (float)(dx - box.L[0] * rint(dx * box.invL[0]));
    y = (float)(dy - box.L[1] * rint(dy * box.invL[1]));
    z = (float)(dz - box.L[2] * rint(dz * box.invL[2]));
}

// append j to the row when `hit`, as predicated instructions: the compiler's branch plus its
// reconvergence barrier cost more than the store they skip
__device__ __forceinline__ void sc_append(bool hit, int *row, int &cnt, int cap, int j)
{
    int *dst = row + (size_t)cnt * 32;
    asm volatile("{\n\t.reg .pred q, s;\n\tsetp.ne.s32 q, %1, 0;\n\tsetp.lt.and.s32 s, %0, %2, q;\n\t"
                 "@s st.global.b32 [%3], %4;\n\t@q add.s32 %0, %0, 1;\n\t}"
                 : "+r"(cnt) : "r"((int)hit), "r"(cap), "l"(dst), "r"(j) : "memory");
}

struct ScShared {
    int prefix[SC_CELLS + 8];
    int first[SC_CELLS + 8];
    float4 cand[SC_CHUNK + 4];       // + 4: the test loop reads one round ahead
};
// windowed list (b2_win.cuh): the window layout of the block's super-block, and where each candidate
// cell starts in it
struct ScWindow {
    int wP[64 + 1], wS[64];
    int woff[SC_CELLS + 8];
};

__device__ __forceinline__ void sc_append16(bool hit, unsigned short *row, int &cnt, int cap, int off)
{
    unsigned short *dst = row + (size_t)cnt * 32;
    asm volatile("{\n\t.reg .pred q, s;\n\tsetp.ne.s32 q, %1, 0;\n\tsetp.lt.and.s32 s, %0, %2, q;\n\t"
                 "@s st.global.b16 [%3], %4;\n\t@q add.s32 %0, %0, 1;\n\t}"
                 : "+r"(cnt) : "r"((int)hit), "r"(cap), "l"(dst), "h"((unsigned short)off) : "memory");
}

// forward declarations (b2_win.cuh)
__device__ __forceinline__ int win_layout_load(const int *__restrict__ T, int lane, int *sP, int *sS);
__device__ __forceinline__ int win_slot_of(const int *sP, const int *sS, int off);

// one block of cells (Morton code ms), one warp
template <typename R, bool FAST2, bool WIN>
__device__ __forceinline__ void sc_block(uint32_t ms, int lane, ScShared &sh, ScWindow &sw, const R *s_rl2, const float *s_hi, const float *s_lo,
                                         Ctl *__restrict__ ctl, const typename Pos4<R>::type *__restrict__ pos,
                                         const Box<R> &box, const Grid &g, const int *__restrict__ cell_start,
                                         const int *__restrict__ cell_end, int *__restrict__ nbr,
                                         int *__restrict__ nnb, int cap,
                                         const int *__restrict__ sbt)
{
    typedef typename Pos4<R>::type P4;
    // (block key -> block: the interleaved low bits, and the high bits of x above them)
    const int lowb = 3 * (g.cellbits - 1);
    const uint32_t mlo = ms & ((1u << lowb) - 1u);
    const int bx = 2 * (int)(compact3(mlo) | ((ms >> lowb) << (g.cellbits - 1))), by = 2 * (int)compact3(mlo >> 1), bz = 2 * (int)compact3(mlo >> 2);
    if (bx >= g.nc[0] || by >= g.nc[1] || bz >= g.nc[2]) return;
    // the block's own particles: one contiguous run of slots
    int first = 0x7fffffff, len = 0;
    if (lane < 8) {
        const uint32_t key = (ms << 3) | (uint32_t)lane;
        const int s0 = cell_start[key];
        len = cell_end[key] - s0;
        if (len > 0) first = s0;
    }
    int n_own = len;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
        n_own += __shfl_xor_sync(0xffffffffu, n_own, o);
    }
    first = __shfl_sync(0xffffffffu, first, 0);
    n_own = __shfl_sync(0xffffffffu, n_own, 0);
    if (n_own == 0) return;
    // windowed list: entries are offsets into the window of this block's super-block
    int W = 0, sb_first = 0, tbase = 0, self_base = 0;
    if (WIN) {
        const int sx = bx >> 2, sy = by >> 2, sz = bz >> 2;
        const int *__restrict__ T = sbt + (size_t)((sz * g.nsb[1] + sy) * g.nsb[0] + sx) * (12 + 2 * 64);   // SBT_STRIDE
        W = win_layout_load(T, lane, sw.wP, sw.wS);
        __syncwarp();
        sb_first = T[0];
        tbase = T[2];
        // the block's own run: B8 block ((bx & 2) + 2) / 2 ... of the window
        const int r_own = (((((bz & 2) + 2) >> 1) * 4) + (((by & 2) + 2) >> 1)) * 4 + (((bx & 2) + 2) >> 1);
        self_base = sw.wP[r_own] - sw.wS[r_own];      // offset of slot s = self_base + s
        if (lane == 0) {
            atomicMax(&ctl->win_max, (unsigned)W);
            if (W > g.wcap || W > 65535) ctl->overflow = 1u;
        }
    }
    // candidate cells: where each starts, and the running count of candidates before it
    int total = 0;
#pragma unroll
    for (int r = 0; r < (SC_CELLS + 31) / 32; r++) {
        const int q = r * 32 + lane;
        int s0 = 0, l = 0;
        if (q < SC_CELLS) {
            const int cx = wrap_cell(bx - 2 + q % 6, g.nc[0]);
            const int cy = wrap_cell(by - 2 + (q / 6) % 6, g.nc[1]);
            const int cz = wrap_cell(bz - 2 + q / 36, g.nc[2]);
            const uint32_t key = grid_key(g, 0, cx, cy, cz);
            s0 = cell_start[key];
            l = cell_end[key] - s0;
        }
        int incl = l;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        sh.prefix[q] = total + incl - l;
        sh.first[q] = s0;
        if (WIN && q < SC_CELLS) {
            // cells counted from the window's origin (4 sx - 2): the block starts (bx & 2) + 2 cells in
            const int ox = (bx & 2) + q % 6, oy = (by & 2) + (q / 6) % 6, oz = (bz & 2) + q / 36;
            const int r = ((oz >> 1) * 4 + (oy >> 1)) * 4 + (ox >> 1);
            sw.woff[q] = sw.wP[r] + (s0 - sw.wS[r]);          // used only when the cell holds particles
        }
        total += __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    ScOrigin org;
    // origin = centre of the block: every own and candidate coordinate is within 3 cells of
    // it, inside (-L/2, L/2) because the grid has at least 7 cells per axis
    org.u[0] = (unsigned)((((unsigned long long)(bx + 1)) << 32) / (unsigned)g.nc[0]);
    org.u[1] = (unsigned)((((unsigned long long)(by + 1)) << 32) / (unsigned)g.nc[1]);
    org.u[2] = (unsigned)((((unsigned long long)(bz + 1)) << 32) / (unsigned)g.nc[2]);
    org.d[0] = ((double)(bx + 1) / g.nc[0] - 0.5) * box.Ld[0];
    org.d[1] = ((double)(by + 1) / g.nc[1] - 0.5) * box.Ld[1];
    org.d[2] = ((double)(bz + 1) / g.nc[2] - 0.5) * box.Ld[2];
    const int nsp = g_pair.nsp;
    float4 *cand = sh.cand;
    const int *prefix = sh.prefix, *firsts = sh.first, *woffs = sw.woff;
    const float4 nobody = make_float4(1e18f, 0.f, 0.f, 0.f);      // never within reach of anyone

    for (int b0 = 0; b0 < n_own; b0 += 32) {          // more than 32 particles in the block: again
        const bool valid = b0 + lane < n_own;
        const int slot = first + (valid ? b0 + lane : 0);
        const P4 pi = pos[slot];
        const bool active = valid && !is_ghost(pi);   // ghosts act as neighbours but own no row
        const int ti = ptype(pi);
        float px, py, pz;
        rel_to(pi, box, org, px, py, pz);
        // thresholds of this particle's row of the pair table; a lane without a row never matches
        const float *hi_row = s_hi + ti * nsp;
        const float hi0 = active ? hi_row[0] : -1.f, hi1 = active ? hi_row[nsp > 1 ? 1 : 0] : -1.f;
        const float *lo_row = s_lo + ti * nsp;
        const float lo0 = lo_row[0], lo1 = lo_row[nsp > 1 ? 1 : 0];
        int *row = nbr + ((size_t)(slot >> 5) * cap) * 32 + (slot & 31);
        const int li = slot - sb_first;                   // windowed: rows live in tiles local to the super-block
        unsigned short *row16 = reinterpret_cast<unsigned short *>(nbr) +
                                ((size_t)(tbase + (li >> 5)) * (size_t)cap) * 32 + (li & 31);
        const int self = WIN ? self_base + slot : slot;   // the name candidates carry: window offset or slot
        int cnt = 0;
        float near = -1.f;        // > 0 once a listed pair sits inside the band around its threshold
        for (int c0 = 0; c0 < total; c0 += SC_TAKE) {
            const int m = min(SC_TAKE, total - c0);
            // ---- stage: lane takes 8 consecutive candidates; those farther than r_list,max from
            // the block itself (a quarter of the 6x6x6 cells' content) are dropped here
            float4 rec[8];
            unsigned ok = 0u, tyb = 0u;
            {
                const int f0 = c0 + lane * 8;
                int c = 0;
                if (f0 < c0 + m) {
                    int lo_c = 0, hi_c = SC_CELLS - 1;
                    while (lo_c < hi_c) {             // largest cell with prefix <= f0
                        const int mid = (lo_c + hi_c + 1) >> 1;
                        if (prefix[mid] <= f0) lo_c = mid; else hi_c = mid - 1;
                    }
                    c = lo_c;
                }
#pragma unroll
                for (int t = 0; t < 8; t++) {
                    const int f = f0 + t;
                    if (f < c0 + m) {
                        while (f >= prefix[c + 1]) c++;
                        const int j = firsts[c] + (f - prefix[c]);
                        const P4 pj = pos[j];
                        rel_to(pj, box, org, rec[t].x, rec[t].y, rec[t].z);
                        const int tj = ptype(pj);
                        const int name = WIN ? woffs[c] + (f - prefix[c]) : j;
                        rec[t].w = __int_as_float(FAST2 ? name : (name | (tj << 28)));
                        const float ex = fmaxf(fabsf(rec[t].x) - g.w[0], 0.f);
                        const float ey = fmaxf(fabsf(rec[t].y) - g.w[1], 0.f);
                        const float ez = fmaxf(fabsf(rec[t].z) - g.w[2], 0.f);
                        if (ex * ex + ey * ey + ez * ez <= g.reach2) {
                            ok |= 1u << t;
                            if (FAST2 && tj) tyb |= 1u << t;
                        }
                    }
                }
            }
            // species 0 packed from the front and (two species) species 1 from the back: each
            // stretch is then tested against ONE threshold held in a register
            const int mine = __popc(ok & ~tyb) | (__popc(ok & tyb) << 16);
            int incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            const int tot = __shfl_sync(0xffffffffu, incl, 31);
            int ia = (incl - mine) & 0xffff, ib = SC_CHUNK - 1 - ((incl - mine) >> 16);
#pragma unroll
            for (int t = 0; t < 8; t++)
                if (ok & (1u << t)) {
                    if (tyb & (1u << t)) cand[ib--] = rec[t];
                    else cand[ia++] = rec[t];
                }
            const int n0 = tot & 0xffff, n1 = tot >> 16;
            if (lane < 4) {                           // pad both stretches to a multiple of 4 (the gap is >= 8)
                cand[n0 + lane] = nobody;
                cand[SC_CHUNK - 1 - n1 - lane] = nobody;
            }
            __syncwarp();
            const int e0 = (n0 + 3) & ~3;
            if (FAST2) {
                float far0 = 0.f, far1 = 0.f;         // largest listed r^2 of either stretch
#define SC_TEST(Q, HI, FAR)                                                        \
                {                                                                  \
                    const float dx = Q.x - px, dy = Q.y - py, dz = Q.z - pz;       \
                    const float r2 = dx * dx + dy * dy + dz * dz;                  \
                    const int j = __float_as_int(Q.w);                             \
                    const bool hit = r2 < HI && j != self;                         \
                    if (WIN) sc_append16(hit, row16, cnt, cap, j);                 \
                    else sc_append(hit, row, cnt, cap, j);                         \
                    FAR = fmaxf(FAR, hit ? r2 : 0.f);                              \
                }
                // four candidates per round, the next four already on their way from shared memory
                // (the reads run up to 4 records past a stretch: the chunk is padded for that)
#define SC_STRETCH(FIRST, LAST, HI, FAR)                                           \
                if ((FIRST) < (LAST)) {                                            \
                    float4 q0 = cand[FIRST], q1 = cand[(FIRST) + 1], q2 = cand[(FIRST) + 2], q3 = cand[(FIRST) + 3]; \
                    _Pragma("unroll 1")                                            \
                    for (int r = (FIRST); r < (LAST); r += 4) {                    \
                        const float4 n0_ = cand[r + 4], n1_ = cand[r + 5], n2_ = cand[r + 6], n3_ = cand[r + 7]; \
                        SC_TEST(q0, HI, FAR) SC_TEST(q1, HI, FAR) SC_TEST(q2, HI, FAR) SC_TEST(q3, HI, FAR) \
                        q0 = n0_; q1 = n1_; q2 = n2_; q3 = n3_;                    \
                    }                                                              \
                }
                SC_STRETCH(0, e0, hi0, far0)
                SC_STRETCH(SC_CHUNK - ((n1 + 3) & ~3), SC_CHUNK, hi1, far1)
#undef SC_STRETCH
#undef SC_TEST
                near = fmaxf(near, fmaxf(far0 - lo0, far1 - lo1));
            } else {
#pragma unroll 4
                for (int r = 0; r < e0; r++) {
                    const float4 q = cand[r];
                    const float dx = q.x - px, dy = q.y - py, dz = q.z - pz;
                    const float r2 = dx * dx + dy * dy + dz * dz;
                    const int w = __float_as_int(q.w);
                    const float hi = active ? hi_row[(unsigned)w >> 28] : -1.f;
                    const int j = w & 0x0fffffff;
                    const bool hit = r2 < hi && j != self;
                    if (WIN) sc_append16(hit, row16, cnt, cap, j);
                    else sc_append(hit, row, cnt, cap, j);
                    near = fmaxf(near, hit ? r2 - lo_row[(unsigned)w >> 28] : -1.f);
                }
            }
            __syncwarp();
        }
        if (near > 0.f) {
            // Some listed pair is inside the band around its threshold: settle those with the
            // canonical predicate and close the gaps (about one particle in a thousand).  Pairs
            // beyond the band were rejected by the float test.
            const int lim = min(cnt, cap);
            int kept = 0;
            for (int k = 0; k < lim; k++) {
                const int name = WIN ? (int)row16[(size_t)k * 32] : row[(size_t)k * 32];
                const int e = WIN ? win_slot_of(sw.wP, sw.wS, name) : name;
                const P4 pj = pos[e];
                const int tj = ptype(pj);
                float qx, qy, qz;
                rel_to(pj, box, org, qx, qy, qz);
                const float dx = qx - px, dy = qy - py, dz = qz - pz;
                bool keep = true;
                if (dx * dx + dy * dy + dz * dz > lo_row[tj]) keep = r2_canon(pi, pj, box) < s_rl2[ti * nsp + tj];
                if (keep) {
                    if (WIN) row16[(size_t)kept * 32] = (unsigned short)name;
                    else row[(size_t)kept * 32] = e;
                    kept++;
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

// one warp per Morton code of a block of cells (the stride loop only matters for a smaller grid)
template <typename R, bool FAST2, bool WIN>
__global__ void __launch_bounds__(SC_WARPS * 32, SC_MINBLK)
k_build_list_sc(Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0,
                const typename Pos4<R>::type *pos1, int n, Box<R> box, Grid g,
                const int *__restrict__ cell_start, const int *__restrict__ cell_end,
                int *__restrict__ nbr, int *__restrict__ nnb, int cap,
                const int *__restrict__ sbt)
{
    if (!ctl->halt_rebuild) return;
    typedef typename Pos4<R>::type P4;
    const P4 *__restrict__ pos = ctl->cur ? pos1 : pos0;
    __shared__ R s_rl2[B2_MAXSP * B2_MAXSP];
    __shared__ float s_hi[B2_MAXSP * B2_MAXSP], s_lo[B2_MAXSP * B2_MAXSP];
    __shared__ ScShared s_warp[SC_WARPS];
    __shared__ ScWindow s_win[WIN ? SC_WARPS : 1];
    if (threadIdx.x < B2_MAXSP * B2_MAXSP) {
        const R rl2 = rl2_of(g_pair, threadIdx.x, (R)0);
        s_rl2[threadIdx.x] = rl2;
        // half-width of the band, relative to the pair's threshold: the float coordinates are good
        // to ~2.5e-7 of the cell width each, i.e. r^2 to ~1e-6 r_list,max^2 in the worst case
        const float eps = SC_EPS * fmaxf(1.f, sqrtf(g.reach2 / fmaxf((float)rl2, 1e-30f)));
        s_hi[threadIdx.x] = (float)rl2 * (1.f + eps);
        s_lo[threadIdx.x] = (float)rl2 * (1.f - eps);
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t codes = 1u << (g.keybits - 3);
    for (uint32_t ms = blockIdx.x * SC_WARPS + warp; ms < codes; ms += gridDim.x * SC_WARPS) {
        sc_block<R, FAST2, WIN>(ms, lane, s_warp[warp], s_win[WIN ? warp : 0], s_rl2, s_hi, s_lo, ctl, pos, box, g,
                                cell_start, cell_end, nbr, nnb, cap, sbt);
        __syncwarp();
    }
}

__global__ void k_rebuild_begin(Ctl *ctl)
{
    if (!ctl->halt_rebuild) return;
    ctl->max_count = 0u;
    ctl->win_max = 0u;
    ctl->overflow = 0u;
    ctl->region_miss = 0u;
}

__global__ void k_rebuild_finish(Ctl *ctl)
{
    if (!ctl->halt_rebuild) return;
    if (!ctl->overflow) {
        ctl->halt_rebuild = 0u;
        ctl->disp_bound = 0.0;
        ctl->rebuilds += 1;
    }
}
