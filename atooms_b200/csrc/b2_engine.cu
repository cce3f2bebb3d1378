// atooms_b200 engine: context, host orchestration and the C ABI
// (include/atooms_b200.h).  Pure CUDA runtime; no torch, no CPU fallback.
#include "../../include/atooms_b200.h"
#include "b2_common.cuh"
#include "b2_neigh.cuh"
#include "b2_force.cuh"
#include "b2_win.cuh"
#include "b2_batch.cuh"
#include "b2_ddr.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <mutex>
#include <thread>
#include <vector>

#define MAX_RED_BLOCKS 1024

struct b2_ctx {
    int device = 0, precision = 64, sms = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // model
    bool have_box = false, have_table = false;
    double side[3] = {0, 0, 0}, center[3] = {0, 0, 0};
    int nsp = 0, shifted = 1;
    double rcut[64], rcut_max = 0, skin = 0.3;
    PairConst pc;
    uint64_t const_serial = 0;
    bool tabulated = false;                  // b2_set_pair_tabulated: tables in tab_u / tab_w / tab_wf (device)
    double *tab_u = nullptr, *tab_w = nullptr;
    float *tab_wf = nullptr;
    // particles (sorted order unless noted)
    int64_t n = 0, ncap = 0;   // particles resident / allocated
    size_t nbr_alloc = 0;
    void *pos[2] = {nullptr, nullptr}, *vel = nullptr, *frc = nullptr, *vir = nullptr;
    void *tpos = nullptr, *tvel = nullptr;
    double *ke_partial = nullptr;   // per-block sum m v^2 of thermostatted steps
    void *mvel = nullptr;      // parked minimiser velocities, by original id
    bool min_resume = false;
    int min_method = -1;
    int *orig = nullptr, *torig = nullptr;
    int *inv = nullptr;        // slot of each local id (inverse of orig); used by the halo pack / unpack
    double dd_dt = 0;
    // NCCL fast path of the domain decomposition (b2_nccl_*)
    ncclComm_t comm = nullptr;
    int nranks = 1, myrank = 0, left = 0, right = 0;
    const int *send_l = nullptr, *send_r = nullptr;   // caller-owned device arrays of local ids
    int n_send_l = 0, n_send_r = 0, n_ghost_l = 0, n_ghost_r = 0;
    double *halo_buf = nullptr;                       // out_l | out_r | in_l | in_r
    size_t halo_cap = 0;
    float *dd_vmax = nullptr;
    // peer-memory halo (b2_peer_*)
    char *peer_buf = nullptr;                         // our mailbox block (IPC-exported)
    char *peer_map[64] = {nullptr};                   // the other ranks' blocks, mapped
    unsigned long long **peer_mail = nullptr;         // device array [nranks]: our slot in every rank's block
    unsigned *peer_ticket = nullptr;
    int64_t peer_cap = 0;                             // records per inbox
    bool peer_on = false;
    // slab refresh on the device (b2_peer_refresh, b2_ddr.cuh)
    unsigned char *ddr_cls = nullptr;
    int *ddr_blk = nullptr, *ddr_keep = nullptr, *ddr_go[2] = {nullptr, nullptr};
    int *ddr_near[4] = {nullptr, nullptr, nullptr, nullptr};      // near_l slot / new id, near_r slot / new id
    int *ddr_send[2] = {nullptr, nullptr};
    unsigned *ddr_cnt = nullptr, *ddr_hcnt = nullptr;             // device counters and their pinned copy
    long long *ddr_gid[2] = {nullptr, nullptr};                   // global ids by local id (current, next)
    int *ddr_cimg[2] = {nullptr, nullptr};                        // box crossings before the last ingestion
    int64_t ddr_ncap = 0;
    int ddr_mig_cap = 0;
    bool ddr_ids = false;                                         // gid / cimg describe the resident particles
    int *img = nullptr;        // [3n] by original id
    double *ref = nullptr;     // [3n] by original id, unfolded, relative to centre
    // neighbour search
    Grid grid;
    uint32_t *keys[2] = {nullptr, nullptr}, *vals[2] = {nullptr, nullptr};
    int *hist = nullptr, *cell_start = nullptr, *cell_end = nullptr;
    int ncelltab = 0, sort_blocks = 0, sort_passes = 0;
    int *nbr = nullptr, *nnb = nullptr;
    int cap = 0;
    // windowed list (b2_win.cuh)
    int *b8start = nullptr, *tile_base = nullptr, *sbt = nullptr;
    int nb8codes = 0, nsb_total = 0, win_tab_alloc = 0, win_sb_alloc = 0;
    bool win_off = false;            // windows do not fit: this context keeps the global-index list
    // the particles fill only the fractions lo <= u < hi (+ wrap) of the box along x: a slab of a decomposed system (b2_peer_refresh)
    bool peer_fused = false;         // (b2_peer_run -> dd_step: commit inside the step kernel)
    // ... and the halo push (StepArgs::halo_map): the map is rebuilt when the list (slots) or the send lists changed
    bool peer_push_fused = false;
    int *halo_map = nullptr;
    int64_t halo_map_cap = 0;
    long long list_epoch = 0, send_epoch = 0, halo_map_list = -1, halo_map_send = -1;
    double *halo_out_l[2] = {nullptr, nullptr}, *halo_out_r[2] = {nullptr, nullptr};
    unsigned long long *halo_flag_l = nullptr, *halo_flag_r = nullptr;
    bool region_on = false, region_off = false;
    double region_lo = 0, region_hi = 0, win_fill = 1.0;
    size_t smem_optin = 0, smem_sm = 0;      // shared memory a block may opt into, and of an SM
    // control / reductions / staging
    Ctl *ctl = nullptr, *hctl = nullptr;
    double *partial = nullptr;
    double *stage = nullptr;   // device staging, 6n doubles
    int *stage_i = nullptr;    // 3n ints
    // host-side knowledge
    bool list_dirty = true, forces_valid = false, ev_valid = false;
    // a run ended with the velocities still at t - dt/2: the closing half kick (and the force evaluation it
    // needs) is done by whoever next looks at velocities or forces — or never, if the next call is the
    // continuation of the same run (Simulation hands the backend one chunk per observer interval)
    bool half_pending = false;
    double half_dt = 0;
    int half_kind = -1;
    double half_T0 = 0, half_tau = 0;
    int batch = 8;
    // profiling (b2_set_profile): events around step batches / refresh chains
    bool profile = false;
    std::vector<cudaEvent_t> pev;
    double prof[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int64_t launches = 0, last_launches = 0, steps_total = 0;
    double last_ms = 0;
    char err[512] = "";
};

static char g_err[512] = "";
// The pair table sits in ONE __constant__ object per device, shared by every context (and replica
// batch) on it.  Ownership changes are serialised by a mutex and ORDERED ON THE DEVICE: every entry
// point that launches kernels leaves an event on its stream (TableGuard), and a new owner makes its
// stream wait for that event before it overwrites the table — kernels of the previous owner still
// queued on another stream keep reading their own coefficients.
static uint64_t g_const_owner[64];      // per device: serial of the ctx whose table is resident
static cudaEvent_t g_const_event[64];   // per device: after the owner's last launches
static bool g_const_event_set[64];
static uint64_t g_serial = 0;
static std::mutex g_const_mutex;

// make `serial` the owner of the device's table (the copy is enqueued on `stream`)
static cudaError_t table_acquire(int device, uint64_t serial, const PairConst *pc, cudaStream_t stream)
{
    std::lock_guard<std::mutex> lock(g_const_mutex);
    if (g_const_owner[device] == serial) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    if (g_const_event_set[device]) e = cudaStreamWaitEvent(stream, g_const_event[device], 0);
    if (e == cudaSuccess) e = cudaMemcpyToSymbolAsync(g_pair, pc, sizeof(PairConst), 0, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) g_const_owner[device] = serial;
    return e;
}

// mark the end of the owner's launches on `stream`
static void table_release(int device, uint64_t serial, cudaStream_t stream)
{
    std::lock_guard<std::mutex> lock(g_const_mutex);
    if (g_const_owner[device] != serial) return;
    if (!g_const_event[device] && cudaEventCreateWithFlags(&g_const_event[device], cudaEventDisableTiming) != cudaSuccess) return;
    g_const_event_set[device] = cudaEventRecord(g_const_event[device], stream) == cudaSuccess;
}

static uint64_t next_serial()
{
    std::lock_guard<std::mutex> lock(g_const_mutex);
    return ++g_serial;
}

static void table_forget(int device, uint64_t serial)
{
    std::lock_guard<std::mutex> lock(g_const_mutex);
    if (g_const_owner[device] == serial) g_const_owner[device] = 0;
}

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;


static int fail(b2_ctx *c, int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(c ? c->err : g_err, 512, fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(c, e_ == cudaErrorMemoryAllocation ? B2_ENOMEM : B2_ECUDA,        \
                        "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

#define LAUNCH_SMEM(kernel, grid, block, smem, ...)                                       \
    do {                                                                                  \
        kernel<<<(grid), (block), (smem), c->stream>>>(__VA_ARGS__);                      \
        c->launches++;                                                                    \
    } while (0)

#define LAUNCH(kernel, grid, block, ...)                                                  \
    do {                                                                                  \
        kernel<<<(grid), (block), 0, c->stream>>>(__VA_ARGS__);                           \
        c->launches++;                                                                    \
    } while (0)

static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

template <typename T> static int dev_alloc(b2_ctx *c, T **p, size_t count)
{
    if (*p) { cudaFree(*p); *p = nullptr; }
    CU(cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T)));
    return B2_OK;
}

template <typename R> static Box<R> make_box(const b2_ctx *c)
{
    Box<R> b;
    const bool fixed = sizeof(R) == 4;     // fp32 positions are 32-bit fixed point
    for (int d = 0; d < 3; d++) {
        b.L[d] = (R)c->side[d];
        b.invL[d] = (R)1 / b.L[d];
        b.Ld[d] = c->side[d];
        b.sd[d] = fixed ? c->side[d] / 4294967296.0 : 1.0;
        b.s[d] = (R)b.sd[d];
        b.is[d] = fixed ? (R)(4294967296.0 / c->side[d]) : (R)1;
    }
    return b;
}

static int bind_device(b2_ctx *c)
{
    CU(cudaSetDevice(c->device));
    return B2_OK;
}

// the pair table lives in __constant__ memory: (re)load it when another
// context used the device since our last launch
static int bind_table(b2_ctx *c)
{
    CU(table_acquire(c->device, c->const_serial, &c->pc, c->stream));
    return B2_OK;
}

// Every entry point that may launch kernels: the table is ours on entry, and the end of our launches is
// marked on leaving.
struct TableGuard {
    b2_ctx *c;
    explicit TableGuard(b2_ctx *ctx) : c(ctx) {}
    ~TableGuard() { if (c && c->have_table) table_release(c->device, c->const_serial, c->stream); }
};

static void rebuild_table(b2_ctx *c)
{
    for (int a = 0; a < c->nsp * c->nsp; a++) {
        double rl = c->rcut[a] + c->skin;
        c->pc.rl2_d[a] = rl * rl;
        float rlf = (float)rl;
        c->pc.rl2_f[a] = rlf * rlf;
        c->pc.f12[a] = (float)(12.0 * c->pc.c12[a]);      // = what load_pair_table computes per block
        c->pc.f6[a] = (float)(6.0 * c->pc.c6[a]);
        c->pc.frc2[a] = (float)c->pc.rc2[a];
    }
    c->const_serial = next_serial();
}

static int pull_ctl(b2_ctx *c)
{
    CU(cudaMemcpyAsync(c->hctl, c->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return B2_OK;
}

static int push_ctl(b2_ctx *c)
{
    CU(cudaMemcpyAsync(c->ctl, c->hctl, sizeof(Ctl), cudaMemcpyHostToDevice, c->stream));
    return B2_OK;
}

// ---------------------------------------------------------------------------
template <typename R> struct Impl {
    typedef typename Vec4<R>::type V4;
    typedef typename Pos4<R>::type P4;

    static int alloc_particles(b2_ctx *c, int64_t n)
    {
        int rc;
        c->ncap = n;
        if (c->ke_partial) { cudaFree(c->ke_partial); c->ke_partial = nullptr; }
        for (int k = 0; k < 2; k++)
            if ((rc = dev_alloc(c, (P4 **)&c->pos[k], n))) return rc;
        if ((rc = dev_alloc(c, (V4 **)&c->vel, n))) return rc;
        if ((rc = dev_alloc(c, (V4 **)&c->frc, n))) return rc;
        if ((rc = dev_alloc(c, (R **)&c->vir, n))) return rc;
        if ((rc = dev_alloc(c, (V4 **)&c->tpos, n))) return rc;   // >= n P4, also n doubles of scratch
        if ((rc = dev_alloc(c, (V4 **)&c->tvel, n))) return rc;
        if ((rc = dev_alloc(c, &c->orig, n))) return rc;
        if ((rc = dev_alloc(c, &c->torig, n))) return rc;
        if ((rc = dev_alloc(c, &c->inv, n))) return rc;
        if ((rc = dev_alloc(c, &c->img, 3 * n))) return rc;
        if ((rc = dev_alloc(c, &c->ref, 3 * n))) return rc;
        for (int k = 0; k < 2; k++) {
            if ((rc = dev_alloc(c, &c->keys[k], n))) return rc;
            if ((rc = dev_alloc(c, &c->vals[k], n))) return rc;
        }
        if ((rc = dev_alloc(c, &c->nnb, n))) return rc;
        if ((rc = dev_alloc(c, &c->stage, 6 * n))) return rc;
        if ((rc = dev_alloc(c, &c->stage_i, 3 * n))) return rc;
        c->sort_blocks = std::max(1, cdiv(n, SORT_TILE));
        if ((rc = dev_alloc(c, &c->hist, (size_t)256 * c->sort_blocks))) return rc;
        // stress partials: one row per block of k_force (n / FORCE_THREADS) or of k_force_win (a super-block of
        // 4^3 half-width cells each; never more than one per 8 particles in any sensible system, checked at launch)
        size_t np = std::max<size_t>((size_t)MAX_RED_BLOCKS * 9, (size_t)(cdiv(n, FORCE_THREADS) + n / 8 + 64) * 6);
        if ((rc = dev_alloc(c, &c->partial, np))) return rc;
        return B2_OK;
    }

    static int setup_grid(b2_ctx *c)
    {
        // cells of width >= r_list,max (1 + 1e-5) / m with a stencil of m cells: every listed
        // pair sits within the stencil.  m = 2 needs >= 5 cells per axis, m = 1 needs >= 3
        // (distinct periodic images in the stencil); below that, all pairs.
        const double rl = (c->rcut_max + c->skin) * (1.0 + 1e-5);
        Grid &g = c->grid;
        g.allpairs = 1;
        g.m = 0;
        for (int m = 2; m >= 1 && g.allpairs; m--) {
            bool ok = true;
            for (int d = 0; d < 3; d++) {
                // at most 256 cells along y and z; x may have up to 4096 (the long axis of a decomposed box: grid_key)
                int nc = std::min((int)std::floor(c->side[d] / (rl / m)), d == 0 ? 4096 : 256);
                if (nc < 2 * m + 1) ok = false;
                g.nc[d] = nc;
            }
            if (ok) { g.allpairs = 0; g.m = m; }
        }
        int mx = 1;
        if (g.allpairs) g.nc[0] = g.nc[1] = g.nc[2] = 1;
        for (int d = 0; d < 3; d++) {
            mx = std::max(mx, g.nc[d]);
            g.w[d] = (float)(c->side[d] / g.nc[d]);
        }
        // windowed list (b2_win.cuh): an even number (>= 8) of cells per axis, cut into super-blocks of 4 (the
        // last one of an axis may hold 2); the cells grow by at most 1 / nc to make it so.  B2_LIST_GLOBAL=1 keeps the global-index list (A/B timing).
        const bool force_global = getenv("B2_LIST_GLOBAL") && atoi(getenv("B2_LIST_GLOBAL")) != 0;
        g.win = 0;
        g.nsb[0] = g.nsb[1] = g.nsb[2] = 0;
        // Measured: fp32, one domain filling its box: step 0.190 ms against 0.207 with the global
        // list.  fp64 windows are twice as large (3 blocks per SM): 0.61 against 0.47 ms, so fp64 keeps the
        // global list unless B2_LIST_WINDOWED=1; so does a domain that fills only part of its box (a slab of a
        // decomposed system: half of the super-blocks are empty launches), see grow_after_overflow.
        const bool force_win = getenv("B2_LIST_WINDOWED") && atoi(getenv("B2_LIST_WINDOWED")) != 0;
        if (!g.allpairs && g.m == 2 && !c->win_off && !c->tabulated && !force_global && (c->precision == B2_FP32 || force_win) &&
            g.nc[0] >= 8 && g.nc[1] >= 8 && g.nc[2] >= 8) {
            double cellvol = 1.0;
            for (int d = 0; d < 3; d++) cellvol *= c->side[d] / (g.nc[d] & ~1);
            // super-blocks to launch along x when the particles are known to fill a slab only
            static const bool gathered = getenv("B2_BUILD_GATHERED") && atoi(getenv("B2_BUILD_GATHERED")) != 0;
            const int nsbx = ((g.nc[0] & ~1) + 3) / 4;
            int sb_first = 0, sb_count = nsbx;
            // B2_REGION_X="lo,hi": the same for a single context (timing experiments with a slab-shaped system)
            if (!c->region_on && getenv("B2_REGION_X") && sscanf(getenv("B2_REGION_X"), "%lf,%lf", &c->region_lo, &c->region_hi) == 2) {
                c->region_lo = (c->region_lo - c->center[0]) / c->side[0] + 0.5;
                c->region_hi = (c->region_hi - c->center[0]) / c->side[0] + 0.5;
                c->region_on = true;
            }
            if (c->region_on && !c->region_off && !gathered) {
                // region_lo / hi: fractions of the box along x from its lower face (may run past 0 or 1)
                const int ncx = g.nc[0] & ~1;
                const int c_lo = (int)floor(c->region_lo * ncx), c_hi = (int)floor(c->region_hi * ncx);
                const int s_lo = (int)floor(c_lo / 4.0) - 1, s_hi = (int)floor(c_hi / 4.0) + 1;     // + one super-block either side
                if (s_hi - s_lo + 1 < nsbx) {
                    sb_count = s_hi - s_lo + 1;
                    sb_first = ((s_lo % nsbx) + nsbx) % nsbx;
                }
            }
            c->win_fill = (double)sb_count / nsbx;
            cellvol /= c->win_fill;            // the density inside the launched part
            // records of a window (8^3 cells) at the mean density, with room for fluctuations; a slab of a
            // decomposed system is denser than its global mean: the first build reports what it needs
            const double expect = (double)c->n / (c->side[0] * c->side[1] * c->side[2]) * 512.0 * cellvol;
            int wcap = std::max(g.wcap, ((int)(1.25 * expect) + 128 + 7) / 8 * 8);
            const size_t rec = c->precision == B2_FP32 ? 16 : 32;
            const size_t lim = std::min<size_t>((c->smem_optin > 4096 ? c->smem_optin - 4096 : 0) / rec, 65528);
            const size_t nsbt = (size_t)sb_count * ((g.nc[1] + 3) / 4) * ((g.nc[2] + 3) / 4);
            if ((size_t)wcap <= lim && nsbt <= (size_t)c->ncap / 8 + 64) {
                g.win = 1;
                g.wcap = wcap;
                for (int d = 0; d < 3; d++) {
                    g.nc[d] &= ~1;
                    g.nsb[d] = (g.nc[d] + 3) / 4;
                    g.sb0[d] = 0;
                    g.sbn[d] = g.nsb[d];
                }
                g.sb0[0] = sb_first;
                g.sbn[0] = sb_count;
                mx = 1;
                for (int d = 0; d < 3; d++) {
                    mx = std::max(mx, g.nc[d]);
                    g.w[d] = (float)(c->side[d] / g.nc[d]);
                }
            }
        }
        g.reach2 = (float)(rl * rl * (1.0 + 1e-3));
        // Morton bits: what y and z need; further bits of x on top
        g.cellbits = 1;
        while ((1 << g.cellbits) < std::max(g.nc[1], g.nc[2])) g.cellbits++;
        int xbits = 1;
        while ((1 << xbits) < g.nc[0]) xbits++;
        g.keybits = 3 * g.cellbits + std::max(0, xbits - g.cellbits);
        int tab = 1 << g.keybits;
        c->sort_passes = (g.keybits + 7) / 8;
        if (tab != c->ncelltab || !c->cell_start) {
            int rc;
            if ((rc = dev_alloc(c, &c->cell_start, tab))) return rc;
            if ((rc = dev_alloc(c, &c->cell_end, tab))) return rc;
            c->ncelltab = tab;
        }
        if (g.win) {
            int rc;
            c->nb8codes = tab >> 3;
            c->nsb_total = g.sbn[0] * g.sbn[1] * g.sbn[2];
            if (c->nb8codes + 2 > c->win_tab_alloc) {
                if ((rc = dev_alloc(c, &c->b8start, (size_t)c->nb8codes + 2))) return rc;
                c->win_tab_alloc = c->nb8codes + 2;
            }
            if (c->nsb_total + 1 > c->win_sb_alloc) {
                if ((rc = dev_alloc(c, &c->tile_base, (size_t)c->nsb_total + 1))) return rc;
                if ((rc = dev_alloc(c, &c->sbt, (size_t)c->nsb_total * SBT_STRIDE))) return rc;
                c->win_sb_alloc = c->nsb_total + 1;
            }
            if ((rc = win_attributes(c))) return rc;
        }
        static const bool debug_grid = getenv("B2_DEBUG_GRID") && atoi(getenv("B2_DEBUG_GRID")) != 0;
        if (debug_grid)
            fprintf(stderr, "[b2 grid] rank %d n %lld cells %d %d %d m %d win %d super-blocks %d %d %d launched %d+%d wcap %d cap %d region %d (%g, %g) off %d\n",
                    c->myrank, (long long)c->n, g.nc[0], g.nc[1], g.nc[2], g.m, g.win, g.nsb[0], g.nsb[1], g.nsb[2], g.sb0[0], g.sbn[0],
                    g.wcap, c->cap, (int)c->region_on, c->region_lo, c->region_hi, (int)c->region_off);
        return B2_OK;
    }

    // dynamic shared memory of the windowed kernels (above the 48 KB default)
    static int win_attributes(b2_ctx *c)
    {
        const int bytes = (int)((size_t)c->grid.wcap * sizeof(P4));
        CU(cudaFuncSetAttribute(k_step_win<R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        CU(cudaFuncSetAttribute(k_step_win<R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        CU(cudaFuncSetAttribute(k_force_win<R, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        CU(cudaFuncSetAttribute(k_force_win<R, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        CU(cudaFuncSetAttribute(k_build_list_win<R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)build_smem(c)));
        CU(cudaFuncSetAttribute(k_build_list_win<R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)build_smem(c)));
        return B2_OK;
    }
#ifndef BW_MAXROWS
#define BW_MAXROWS 128
#endif
#ifndef BW_BLOCKS
#define BW_BLOCKS 2
#endif
    // Shared memory of the list build: the window (it reads up to four records past it), then per warp a buffer
    // of list rows (64 bytes each).  Up to 128 rows, as many as leave room for two blocks per SM.
    static int build_rows(const b2_ctx *c)
    {
        const size_t window = (size_t)c->grid.wcap * sizeof(P4) + 128, fixed = 4096;
        const size_t half = (c->smem_sm > 1024 * BW_BLOCKS ? c->smem_sm - 1024 * BW_BLOCKS : 0) / BW_BLOCKS;
        size_t room = half > window + fixed ? half - window - fixed : 0;
        if (room < 8 * 64 * (WIN_THREADS / 32) && c->smem_optin > window + fixed)          // one block per SM then
            room = c->smem_optin - window - fixed;
        return (int)std::min<size_t>(BW_MAXROWS, room / (64 * (WIN_THREADS / 32))) & ~3;
    }
    static size_t build_smem(const b2_ctx *c)
    {
        return (size_t)c->grid.wcap * sizeof(P4) + 128 + (size_t)build_rows(c) * 64 * (WIN_THREADS / 32);
    }

    // ints the list needs for `cap` slots per particle
    static size_t list_ints(const b2_ctx *c, int cap)
    {
        if (c->grid.win)      // 16-bit entries in tiles local to the super-blocks: at most one partial tile each
            return ((size_t)cdiv(c->ncap, 32) + (size_t)c->nsb_total + 1) * (size_t)cap * 16 + 1024;   // + slack: rows are read ahead
        return (size_t)cdiv(c->ncap, 32) * (size_t)cap * 32;
    }

    // the last build did not fit: more slots per row, a larger window, or no windows at all
    static int grow_after_overflow(b2_ctx *c)
    {
        int rc;
        Grid &g = c->grid;
        if (g.win && c->hctl->region_miss) {
            // a particle outside the slab the caller announced: launch every super-block from now on
            c->region_off = true;
            if ((rc = setup_grid(c))) return rc;
            return ensure_capacity(c, std::max(c->cap, 8));
        }
        if (g.win && ((int)c->hctl->win_max > g.wcap || c->hctl->win_max >= 65535u)) {
            const size_t rec = sizeof(P4);
            const size_t lim = std::min<size_t>((c->smem_optin > 4096 ? c->smem_optin - 4096 : 0) / rec, 65528);
            const size_t want = ((size_t)(c->hctl->win_max * 1.1) + 64 + 7) / 8 * 8;
            // windows far above the mean density: the particles fill only part of the box (a slab of a decomposed
            // system, a droplet) and many super-blocks are empty — the global list serves such systems better
            const bool force_win = getenv("B2_LIST_WINDOWED") && atoi(getenv("B2_LIST_WINDOWED")) != 0;
            const double expect = (double)c->n / (c->side[0] * c->side[1] * c->side[2]) * 512.0 *
                                  (c->side[0] / g.nc[0]) * (c->side[1] / g.nc[1]) * (c->side[2] / g.nc[2]);
            const bool sparse = (double)c->hctl->win_max > 1.6 * expect / c->win_fill + 256.0 && !force_win;
            if (want > lim || sparse) {
                c->win_off = true;
                c->cap = 0;
                if ((rc = setup_grid(c))) return rc;
                double rl = c->rcut_max + c->skin;
                double nbar = (double)c->n / (c->side[0] * c->side[1] * c->side[2]) * 4.18879 * rl * rl * rl;
                return ensure_capacity(c, std::max((int)(1.3 * nbar) + 24, (int)(c->hctl->max_count * 1.15) + 8));
            }
            g.wcap = (int)want;
            if ((rc = win_attributes(c))) return rc;
        }
        return ensure_capacity(c, (int)(c->hctl->max_count * 1.15) + 8);
    }

    static int ensure_capacity(b2_ctx *c, int want)
    {
        int64_t n = c->n;
        want = (int)std::min<int64_t>(want, std::max<int64_t>(n - 1, 1));
        want = (want + 7) / 8 * 8;
        if (want <= c->cap && c->nbr && list_ints(c, c->cap) <= c->nbr_alloc) return B2_OK;
        want = std::max(want, c->cap);
        int rc;
        size_t need = list_ints(c, want);
        if (need > c->nbr_alloc || !c->nbr) {
            if ((rc = dev_alloc(c, &c->nbr, need))) return rc;
            c->nbr_alloc = need;
        }
        c->cap = want;
        return B2_OK;
    }

    static int upload(b2_ctx *c, int64_t n, const double *pos, const double *vel,
                      const int32_t *species, const double *mass);
    static int update(b2_ctx *c, const double *pos, const double *vel);
    static int download(b2_ctx *c, double *pos, double *vel, int32_t *image);

    // -- neighbour refresh (guarded pipeline, enqueue only) -------------------
    static void enqueue_rebuild(b2_ctx *c)
    {
        int n = (int)c->n;
        Box<R> box = make_box<R>(c);
        const P4 *p0 = (const P4 *)c->pos[0], *p1 = (const P4 *)c->pos[1];
        LAUNCH(k_rebuild_begin, 1, 1, c->ctl);
        if (!c->grid.allpairs) {
            LAUNCH(k_cell_keys<R>, cdiv(n, 256), 256, c->ctl, p0, p1, n, box, c->grid, c->keys[0], c->vals[0]);
            int src = 0;
            for (int pass = 0; pass < c->sort_passes; pass++) {
                LAUNCH(k_sort_hist, c->sort_blocks, SORT_THREADS, c->ctl, c->keys[src], n, 8 * pass, c->sort_blocks, c->hist);
                LAUNCH(k_sort_scan, 1, 256 * SCAN_GROUPS, c->ctl, c->hist, c->sort_blocks);
                LAUNCH(k_sort_scatter, c->sort_blocks, SORT_THREADS, c->ctl, c->keys[src], c->vals[src],
                       c->keys[src ^ 1], c->vals[src ^ 1], n, 8 * pass, c->sort_blocks, c->hist);
                src ^= 1;
            }
            LAUNCH(k_gather<R>, cdiv(n, 256), 256, c->ctl, c->vals[src], n, p0, p1, (const V4 *)c->vel, c->orig,
                   (P4 *)c->tpos, (V4 *)c->tvel, c->torig);
            LAUNCH(k_gather_commit<R>, cdiv(n, 256), 256, c->ctl, n, (const P4 *)c->tpos, (const V4 *)c->tvel,
                   c->torig, (P4 *)c->pos[0], (P4 *)c->pos[1], (V4 *)c->vel, c->orig);
            LAUNCH(k_inverse, cdiv(n, 256), 256, c->ctl, 1, c->orig, n, c->inv);
            static const bool gathered_build = getenv("B2_BUILD_GATHERED") && atoi(getenv("B2_BUILD_GATHERED")) != 0;
            if (!c->grid.win || gathered_build) {       // (the windowed build goes by the runs of the 2x2x2 blocks alone)
                LAUNCH(k_fill_int, cdiv(c->ncelltab, 256), 256, c->ctl, c->cell_start, c->cell_end, c->ncelltab, 0);
                LAUNCH(k_cell_bounds, cdiv(n, 256), 256, c->ctl, c->keys[src], n, c->cell_start, c->cell_end);
            }
            if (c->grid.win) {
                LAUNCH(k_b8_bounds, cdiv(c->nb8codes + 1, 256), 256, c->ctl, c->keys[src], n, c->nb8codes, c->b8start);
                LAUNCH(k_sb_tiles, 1, 1024, c->ctl, c->grid, c->b8start, c->tile_base);
                LAUNCH(k_sb_table<R>, cdiv(c->nsb_total, 8), 256, c->ctl, c->grid, box, c->b8start, c->tile_base, c->sbt);
            }
        }
        // one warp per 2x2x2 block of cells when the grid allows it (at least 7 cells per axis: the
        // 6x6x6 candidate region then holds distinct cells within half a box of the block's centre); B2_LIST_WALK=1 forces the per-particle walk (A/B timing)
        static const bool force_walk = getenv("B2_LIST_WALK") && atoi(getenv("B2_LIST_WALK")) != 0;
        const Grid &g = c->grid;
        if (g.m == 2 && !g.allpairs && g.nc[0] >= 7 && g.nc[1] >= 7 && g.nc[2] >= 7 && n < (1 << 28) && (!force_walk || g.win)) {
            int codes = 1 << (g.keybits - 3);
            static const bool gathered = getenv("B2_BUILD_GATHERED") && atoi(getenv("B2_BUILD_GATHERED")) != 0;   // A/B timing
            if (g.win && !gathered) {
                auto kern = c->nsp <= 2 ? k_build_list_win<R, true> : k_build_list_win<R, false>;
                const int kbuf = std::min(build_rows(c), c->cap & ~3);
                LAUNCH_SMEM(kern, win_grid(c), WIN_THREADS, build_smem(c), c->ctl, p0, p1, box, c->grid, (const int *)c->sbt,
                            reinterpret_cast<unsigned short *>(c->nbr), c->nnb, c->cap, kbuf);
                LAUNCH(k_rebuild_finish, 1, 1, c->ctl);
                return;
            }
            auto build = g.win ? (c->nsp <= 2 ? k_build_list_sc<R, true, true> : k_build_list_sc<R, false, true>)
                               : (c->nsp <= 2 ? k_build_list_sc<R, true, false> : k_build_list_sc<R, false, false>);
            // one warp per code (a strided grid of a few blocks per SM was measured 10 % slower:
            // the valid codes are unevenly spread over the Morton range)
            LAUNCH(build, cdiv(codes, SC_WARPS), SC_WARPS * 32, c->ctl, p0, p1, n, box, c->grid,
                   c->cell_start, c->cell_end, c->nbr, c->nnb, c->cap, c->sbt);
        } else {
            auto build = g.m == 2 ? k_build_list<R, 2> : (g.m == 1 ? k_build_list<R, 1> : k_build_list<R, 0>);
            LAUNCH(build, cdiv(n, 128), 128, c->ctl, p0, p1, n, box, c->grid, c->cell_start, c->cell_end,
                   c->nbr, c->nnb, c->cap);
        }
        LAUNCH(k_rebuild_finish, 1, 1, c->ctl);
    }

    // make the list valid for the current configuration (synchronises)
    static int ensure_list(b2_ctx *c)
    {
        int rc;
        if ((rc = bind_table(c))) return rc;
        if (!c->list_dirty) return B2_OK;
        if (c->n == 0) { c->list_dirty = false; return B2_OK; }
        if ((rc = setup_grid(c))) return rc;
        {
            // first build: slots from the mean density; afterwards keep the slot count, but ALWAYS
            // re-check the allocation — a re-upload (domain decomposition) may have grown the arrays
            int want = c->cap;
            if (want == 0) {
                double rl = c->rcut_max + c->skin;
                double vol = c->side[0] * c->side[1] * c->side[2];
                double nbar = (double)c->n / vol * 4.18879 * rl * rl * rl;
                want = (int)(1.3 * nbar) + 24;
            }
            if ((rc = ensure_capacity(c, want))) return rc;
        }
        for (int attempt = 0; attempt < 8; attempt++) {
            if ((rc = pull_ctl(c))) return rc;
            c->hctl->halt_rebuild = 1u;
            c->hctl->overflow = 0u;
            c->hctl->half_skin = 0.5 * c->skin;
            if ((rc = push_ctl(c))) return rc;
            enqueue_rebuild(c);
            if ((rc = pull_ctl(c))) return rc;
            CU(cudaGetLastError());
            if (!c->hctl->overflow) {
                c->list_dirty = false;
                c->list_epoch++;
                return B2_OK;
            }
            if ((rc = grow_after_overflow(c))) return rc;
        }
        return fail(c, B2_ECUDA, "neighbour list did not fit after 8 attempts (max count %u, capacity %d)",
                    c->hctl->max_count, c->cap);
    }

    template <bool EV, bool STRESS>
    static void enqueue_force(b2_ctx *c, bool guarded)
    {
        ForceArgs<R> a;
        a.ctl = c->ctl;
        a.pos[0] = (const P4 *)c->pos[0];
        a.pos[1] = (const P4 *)c->pos[1];
        a.frc = (V4 *)c->frc;
        a.vir = (R *)c->vir;
        a.nbr = c->nbr;
        a.nnb = c->nnb;
        a.partial = c->partial;
        a.n = (int)c->n;
        a.cap = c->cap;
        a.guarded = guarded ? 1 : 0;
        a.box = make_box<R>(c);
        if (c->grid.win) {
            auto kern = k_force_win<R, EV, STRESS>;
            LAUNCH_SMEM(kern, win_grid(c), WIN_THREADS, (size_t)c->grid.wcap * sizeof(P4), a, win_args(c));
            return;
        }
        auto kern = c->tabulated ? k_force<R, EV, STRESS, true> : k_force<R, EV, STRESS, false>;
        LAUNCH(kern, cdiv(c->n, FORCE_THREADS), FORCE_THREADS, a);
    }

    static dim3 win_grid(const b2_ctx *c) { return dim3(c->grid.sbn[0], c->grid.sbn[1], c->grid.sbn[2]); }

    static WinArgs<R> win_args(const b2_ctx *c)
    {
        WinArgs<R> w;
        w.g = c->grid;
        w.sbt = c->sbt;
        w.nbr16 = (const unsigned short *)c->nbr;
        return w;
    }

    // blocks of the force kernels (rows of stress partials)
    static int force_blocks(const b2_ctx *c) { return c->grid.win ? c->nsb_total : cdiv(c->n, FORCE_THREADS); }

    template <bool NVT>
    static void enqueue_step(b2_ctx *c, const StepArgs<R> &a)
    {
        if (c->grid.win) {
            auto kern = k_step_win<R, NVT>;
            LAUNCH_SMEM(kern, win_grid(c), WIN_THREADS, (size_t)c->grid.wcap * sizeof(P4), a, win_args(c));
        } else {
            auto kern = c->tabulated ? k_step_nve<R, NVT, true> : k_step_nve<R, NVT, false>;
            LAUNCH(kern, cdiv(c->n, FORCE_THREADS), FORCE_THREADS, a);
        }
    }

    // forces + per-particle energy/virial of the current configuration
    static int ensure_forces(b2_ctx *c)
    {
        int rc;
        if ((rc = ensure_list(c))) return rc;
        if (c->forces_valid && c->ev_valid) return B2_OK;
        if (c->n > 0) enqueue_force<true, false>(c, false);
        c->forces_valid = c->ev_valid = true;
        return B2_OK;
    }

    static int settle(b2_ctx *c);
    static int thermo(b2_ctx *c, double *out, bool want_stress, double *stress9);
    static int run_nve(b2_ctx *c, int64_t steps, double dt, int th_kind = -1, double T0 = 0, double tau = 0);
    static int minimize(b2_ctx *c, int method, int64_t max_steps, double dt, double tol,
                        int64_t *steps_done, int *converged);
    static int neighbors(b2_ctx *c, int32_t *count, int32_t *nbr, int32_t stride, int32_t *max_count);
    static int mark_reference(b2_ctx *c);
    static int dd_begin(b2_ctx *c, int64_t steps, double dt, int opening_kick);
    static int dd_step(b2_ctx *c, float *vmax_out);
    static int dd_end(b2_ctx *c);
    static int dd_pack(b2_ctx *c, const int32_t *ids_dev, int64_t n, double *out_dev);
    static int dd_unpack(b2_ctx *c, int64_t first, int64_t n, const double *in_dev);
    static int compute(b2_ctx *c, int what, double *energy, double *virial, double *stress9,
                       double *forces, double *pe, double *vir);
};

// ---- staging kernels ---------------------------------------------------------
template <typename R>
__global__ void k_ingest(const double *__restrict__ spos, const double *__restrict__ svel,
                         const int *__restrict__ sspe, const double *__restrict__ smass, int n,
                         double cx, double cy, double cz, Box<R> box,
                         typename Pos4<R>::type *__restrict__ pos, typename Vec4<R>::type *__restrict__ vel,
                         int *__restrict__ orig, int *__restrict__ img, double *__restrict__ ref)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // fold in double (particle.py:90-99), then encode in the working precision
    double px = spos[3 * i] - cx, py = spos[3 * i + 1] - cy, pz = spos[3 * i + 2] - cz;
    px -= rint(px / box.Ld[0]) * box.Ld[0];
    py -= rint(py / box.Ld[1]) * box.Ld[1];
    pz -= rint(pz / box.Ld[2]) * box.Ld[2];
    typename Pos4<R>::type p = encode_pos(px, py, pz, sspe[i], box);
    pos[i] = p;
    vel[i] = make4((R)(svel ? svel[3 * i] : 0.0), (R)(svel ? svel[3 * i + 1] : 0.0),
                   (R)(svel ? svel[3 * i + 2] : 0.0), (R)(smass ? smass[i] : 1.0));
    orig[i] = i;
    img[3 * i] = img[3 * i + 1] = img[3 * i + 2] = 0;
    to_double(p, box, px, py, pz);
    ref[3 * i] = px; ref[3 * i + 1] = py; ref[3 * i + 2] = pz;
}

template <typename R>
__global__ void k_update(const Ctl *__restrict__ ctl, const double *__restrict__ spos,
                         const double *__restrict__ svel, int n, double cx, double cy, double cz,
                         Box<R> box, typename Pos4<R>::type *pos0, typename Pos4<R>::type *pos1,
                         typename Vec4<R>::type *__restrict__ vel, const int *__restrict__ orig,
                         int *__restrict__ img)
{
    typedef typename Vec4<R>::type V4;
    typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    size_t o = 3 * (size_t)orig[s];
    if (spos) {
        double px = spos[o] - cx, py = spos[o + 1] - cy, pz = spos[o + 2] - cz;
        double kx = rint(px / box.Ld[0]), ky = rint(py / box.Ld[1]), kz = rint(pz / box.Ld[2]);
        img[o] += (int)kx; img[o + 1] += (int)ky; img[o + 2] += (int)kz;
        pos[s] = encode_pos(px - kx * box.Ld[0], py - ky * box.Ld[1], pz - kz * box.Ld[2], praw(pos[s]), box);
    }
    if (svel) {
        V4 v = vel[s];
        vel[s] = make4((R)svel[o], (R)svel[o + 1], (R)svel[o + 2], v.w);
    }
}

template <typename R>
__global__ void k_export(const Ctl *__restrict__ ctl, int n, double cx, double cy, double cz, Box<R> box,
                         const typename Pos4<R>::type *pos0, const typename Pos4<R>::type *pos1,
                         const typename Vec4<R>::type *__restrict__ vel,
                         const typename Vec4<R>::type *__restrict__ frc, const R *__restrict__ vir,
                         const int *__restrict__ orig, double *__restrict__ opos,
                         double *__restrict__ ovel, double *__restrict__ ofrc,
                         double *__restrict__ ope, double *__restrict__ ovir)
{
    typedef typename Vec4<R>::type V4;
    const typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    size_t o = (size_t)orig[s];
    if (opos) {
        double x, y, z;
        to_double(pos[s], box, x, y, z);
        opos[3 * o] = x + cx; opos[3 * o + 1] = y + cy; opos[3 * o + 2] = z + cz;
    }
    if (ovel) {
        V4 v = vel[s];
        ovel[3 * o] = v.x; ovel[3 * o + 1] = v.y; ovel[3 * o + 2] = v.z;
    }
    if (ofrc || ope) {
        V4 f = frc[s];
        if (ofrc) { ofrc[3 * o] = f.x; ofrc[3 * o + 1] = f.y; ofrc[3 * o + 2] = f.z; }
        if (ope) ope[o] = f.w;
    }
    if (ovir) ovir[o] = vir[s];
}

template <typename R>
__global__ void k_mark_ref(const Ctl *__restrict__ ctl, int n, const typename Pos4<R>::type *pos0,
                           const typename Pos4<R>::type *pos1, const int *__restrict__ orig,
                           const int *__restrict__ img, Box<R> box, double *__restrict__ ref)
{
    const typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    size_t o = 3 * (size_t)orig[s];
    double x, y, z;
    to_double(pos[s], box, x, y, z);
    ref[o] = x + (double)img[o] * box.Ld[0];
    ref[o + 1] = y + (double)img[o + 1] * box.Ld[1];
    ref[o + 2] = z + (double)img[o + 2] * box.Ld[2];
}

// export the list in original numbering: out[orig[s]*stride + k] = orig[nbr]
__global__ void k_export_list(int n, const int *__restrict__ nbr, const int *__restrict__ nnb,
                              const int *__restrict__ orig, int cap, int stride,
                              int *__restrict__ ocount, int *__restrict__ onbr)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int o = orig[s], cnt = nnb[s];
    ocount[o] = cnt;
    if (!onbr) return;
    const int *row = nbr + ((size_t)(s >> 5) * cap) * 32 + (s & 31);
    for (int k = 0; k < cnt; k++) onbr[(size_t)o * stride + k] = orig[row[(size_t)k * 32]];
}

__global__ void k_ctl_begin(Ctl *ctl, long long target, int min_mode, double dt, double dt_max, double tol,
                            int resume)
{
    ctl->steps_done = 0;
    ctl->steps_target = target;
    ctl->done = target <= 0 ? 1u : 0u;
    ctl->converged = 0u;
    ctl->blocks_done = 0u;
    ctl->vmax2_bits = 0u;
    ctl->iter = 0;
    ctl->min_mode = min_mode;
    ctl->fn2_tol = tol;
    ctl->step_len = 0.0;
    if (!resume) {               // a resumed minimisation keeps dt, alpha, npos of the previous call
        ctl->dt_now = dt;
        ctl->dt_max = dt_max;
        ctl->alpha = 0.1;
        ctl->npos = 0;
        ctl->has_prev = 0;
    }
}

// minimiser velocities (FIRE) / previous gradient (SD) parked between two
// consecutive b2_minimize calls, by ORIGINAL id so that a re-sort in between is harmless
template <typename R>
__global__ void k_park_minvel(int n, int save, typename Vec4<R>::type *__restrict__ vel,
                              typename Vec4<R>::type *__restrict__ parked, const int *__restrict__ orig)
{
    typedef typename Vec4<R>::type V4;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int o = orig[s];
    if (save) {
        parked[o] = vel[s];
    } else {
        V4 p = parked[o], v = vel[s];
        vel[s] = make4(p.x, p.y, p.z, v.w);
    }
}

// ---------------------------------------------------------------------------
template <typename R>
int Impl<R>::upload(b2_ctx *c, int64_t n, const double *pos, const double *vel,
                    const int32_t *species, const double *mass)
{
    int rc;
    // a domain that refreshes itself on the device (b2_peer_refresh) needs room to grow in place
    const bool tight = c->peer_on && c->ncap < n + n / 16;
    if (n > c->ncap || !c->pos[0] || tight) {
        // grow with headroom: a decomposed domain re-uploads a slightly different n at every refresh
        if ((rc = alloc_particles(c, c->pos[0] ? n + n / 8 + 1024 : n))) return rc;
    }
    c->ddr_ids = false;
    c->n = n;
    c->list_dirty = true;
    c->forces_valid = c->ev_valid = false;
    c->half_pending = false;
    c->min_resume = false;
    if (c->mvel) { cudaFree(c->mvel); c->mvel = nullptr; }
    memset(c->hctl, 0, sizeof(Ctl));
    c->hctl->halt_rebuild = 1u;
    c->hctl->half_skin = 0.5 * c->skin;
    if ((rc = push_ctl(c))) return rc;
    if (n == 0) return B2_OK;
    double *spos = c->stage, *svel = c->stage + 3 * n;
    double *smass = (double *)c->tpos;     // scratch reuse: n doubles fit in n V4
    // cudaMemcpyDefault: the sources may be host (b2_upload) or device (b2_upload_device) memory
    CU(cudaMemcpyAsync(spos, pos, 3 * n * sizeof(double), cudaMemcpyDefault, c->stream));
    if (vel) CU(cudaMemcpyAsync(svel, vel, 3 * n * sizeof(double), cudaMemcpyDefault, c->stream));
    if (mass) CU(cudaMemcpyAsync(smass, mass, n * sizeof(double), cudaMemcpyDefault, c->stream));
    CU(cudaMemcpyAsync(c->stage_i, species, n * sizeof(int32_t), cudaMemcpyDefault, c->stream));
    LAUNCH(k_ingest<R>, cdiv(n, 256), 256, spos, vel ? svel : nullptr, c->stage_i, mass ? smass : nullptr,
           (int)n, c->center[0], c->center[1], c->center[2], make_box<R>(c),
           (P4 *)c->pos[0], (V4 *)c->vel, c->orig, c->img, c->ref);
    LAUNCH(k_inverse, cdiv(n, 256), 256, c->ctl, 0, c->orig, (int)n, c->inv);
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    return B2_OK;
}

template <typename R>
int Impl<R>::update(b2_ctx *c, const double *pos, const double *vel)
{
    int64_t n = c->n;
    if (n == 0 || (!pos && !vel)) return B2_OK;
    c->min_resume = false;
    double *spos = c->stage, *svel = c->stage + 3 * n;
    if (pos) CU(cudaMemcpyAsync(spos, pos, 3 * n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    if (vel) CU(cudaMemcpyAsync(svel, vel, 3 * n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    LAUNCH(k_update<R>, cdiv(n, 256), 256, c->ctl, pos ? spos : nullptr, vel ? svel : nullptr, (int)n,
           c->center[0], c->center[1], c->center[2], make_box<R>(c),
           (P4 *)c->pos[0], (P4 *)c->pos[1], (V4 *)c->vel, c->orig, c->img);
    CU(cudaStreamSynchronize(c->stream));
    if (pos) {
        c->list_dirty = true;
        c->forces_valid = c->ev_valid = false;
    }
    return B2_OK;
}

template <typename R>
int Impl<R>::download(b2_ctx *c, double *pos, double *vel, int32_t *image)
{
    int64_t n = c->n;
    if (n == 0) return B2_OK;
    double *spos = c->stage, *svel = c->stage + 3 * n;
    LAUNCH(k_export<R>, cdiv(n, 256), 256, c->ctl, (int)n, c->center[0], c->center[1], c->center[2], make_box<R>(c),
           (const P4 *)c->pos[0], (const P4 *)c->pos[1], (const V4 *)c->vel, (const V4 *)c->frc,
           (const R *)c->vir, c->orig, pos ? spos : nullptr, vel ? svel : nullptr, nullptr, nullptr, nullptr);
    if (pos) CU(cudaMemcpyAsync(pos, spos, 3 * n * sizeof(double), cudaMemcpyDefault, c->stream));
    if (vel) CU(cudaMemcpyAsync(vel, svel, 3 * n * sizeof(double), cudaMemcpyDefault, c->stream));
    if (image) CU(cudaMemcpyAsync(image, c->img, 3 * n * sizeof(int32_t), cudaMemcpyDefault, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return B2_OK;
}

template <typename R>
int Impl<R>::mark_reference(b2_ctx *c)
{
    if (c->n == 0) return B2_OK;
    LAUNCH(k_mark_ref<R>, cdiv(c->n, 256), 256, c->ctl, (int)c->n, (const P4 *)c->pos[0], (const P4 *)c->pos[1],
           c->orig, c->img, make_box<R>(c), c->ref);
    return B2_OK;
}

template <typename R>
int Impl<R>::thermo(b2_ctx *c, double *out, bool want_stress, double *stress9)
{
    int rc;
    for (int k = 0; k < B2_T_LEN; k++) out[k] = 0.0;
    if (c->n == 0) return B2_OK;
    if ((rc = ensure_forces(c))) return rc;
    int n = (int)c->n;
    int nblk = std::min(MAX_RED_BLOCKS, cdiv(n, RED_THREADS));
    LAUNCH(k_thermo_partial<R>, nblk, RED_THREADS, c->ctl, (const P4 *)c->pos[0], (const P4 *)c->pos[1],
           (const V4 *)c->vel, (const V4 *)c->frc, (const R *)c->vir, c->orig, c->img, c->ref, n,
           make_box<R>(c), c->partial);
    LAUNCH(k_sum_partials, 1, 32, c->partial, nblk, 9, c->ctl->red);
    if (want_stress) {
        enqueue_force<true, true>(c, false);
        LAUNCH(k_sum_partials, 1, 32, c->partial, force_blocks(c), 6, c->ctl->red + Q_S0);
    }
    if ((rc = pull_ctl(c))) return rc;
    CU(cudaGetLastError());
    const double *r = c->hctl->red;
    out[B2_T_KINETIC] = r[Q_KIN]; out[B2_T_POTENTIAL] = r[Q_POT]; out[B2_T_VIRIAL] = r[Q_VIR];
    out[B2_T_FNORM2] = r[Q_FF]; out[B2_T_PX] = r[Q_PX]; out[B2_T_PY] = r[Q_PY]; out[B2_T_PZ] = r[Q_PZ];
    out[B2_T_MASS] = r[Q_MASS]; out[B2_T_MSD] = r[Q_MSD];
    if (want_stress && stress9) {
        const double *s = r + Q_S0;   // xx xy xz yy yz zz
        stress9[0] = s[0]; stress9[1] = s[1]; stress9[2] = s[2];
        stress9[3] = s[1]; stress9[4] = s[3]; stress9[5] = s[4];
        stress9[6] = s[2]; stress9[7] = s[4]; stress9[8] = s[5];
    }
    return B2_OK;
}

template <typename R>
int Impl<R>::compute(b2_ctx *c, int what, double *energy, double *virial, double *stress9,
                     double *forces, double *pe, double *vir)
{
    int rc;
    double t[B2_T_LEN];
    bool want_stress = (what & B2_STRESS) && stress9;
    if ((rc = thermo(c, t, want_stress, stress9))) return rc;
    if (energy) *energy = t[B2_T_POTENTIAL];
    if (virial) *virial = t[B2_T_VIRIAL];
    if (c->n == 0) {
        if (want_stress) memset(stress9, 0, 9 * sizeof(double));
        return B2_OK;
    }
    if (forces || pe || vir) {
        int64_t n = c->n;
        double *sf = c->stage, *spe = c->stage + 3 * n, *sv = c->stage + 4 * n;
        LAUNCH(k_export<R>, cdiv(n, 256), 256, c->ctl, (int)n, 0.0, 0.0, 0.0, make_box<R>(c), (const P4 *)c->pos[0],
               (const P4 *)c->pos[1], (const V4 *)c->vel, (const V4 *)c->frc, (const R *)c->vir, c->orig,
               nullptr, nullptr, forces ? sf : nullptr, pe ? spe : nullptr, vir ? sv : nullptr);
        if (forces) CU(cudaMemcpyAsync(forces, sf, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        if (pe) CU(cudaMemcpyAsync(pe, spe, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        if (vir) CU(cudaMemcpyAsync(vir, sv, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    return B2_OK;
}

template <typename R>
int Impl<R>::neighbors(b2_ctx *c, int32_t *count, int32_t *nbr, int32_t stride, int32_t *max_count)
{
    int rc;
    int64_t n = c->n;
    if (max_count) *max_count = 0;
    if (n == 0) return B2_OK;
    // always the list of the CURRENT configuration
    c->list_dirty = true;
    if ((rc = ensure_list(c))) return rc;
    int mx = (int)c->hctl->max_count;
    if (max_count) *max_count = mx;
    if (nbr && stride < mx) return fail(c, B2_EINVAL, "b2_neighbors: stride %d smaller than max count %d", stride, mx);
    int *ocount = c->stage_i, *onbr = nullptr;
    if (nbr) CU(cudaMalloc((void **)&onbr, (size_t)n * stride * sizeof(int)));
    if (c->grid.win)
        LAUNCH(k_export_list_win, win_grid(c), WIN_THREADS, c->sbt, (const unsigned short *)c->nbr, c->nnb, c->orig, c->cap, stride, ocount, onbr);
    else
        LAUNCH(k_export_list, cdiv(n, 128), 128, (int)n, c->nbr, c->nnb, c->orig, c->cap, stride, ocount, onbr);
    std::vector<int32_t> hc((size_t)n);
    CU(cudaMemcpyAsync(hc.data(), ocount, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (nbr) CU(cudaMemcpyAsync(nbr, onbr, (size_t)n * stride * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (onbr) cudaFree(onbr);
    for (int64_t i = 0; i < n; i++) {
        if (count) count[i] = hc[i];
        if (nbr) std::sort(nbr + i * stride, nbr + i * stride + hc[i]);
    }
    return B2_OK;
}

template <typename R>
int Impl<R>::run_nve(b2_ctx *c, int64_t steps, double dt, int th_kind, double T0, double tau)
{
    int rc;
    if (steps == 0 || c->n == 0) return B2_OK;
    int n = (int)c->n;
    const bool nvt = th_kind >= 0;
    // continuation of the run the previous call left open: same integrator, same step — no closing and
    // no opening half kick, no force evaluation in between
    const bool cont = c->half_pending && c->half_dt == dt && c->half_kind == th_kind &&
                      (!nvt || (c->half_T0 == T0 && c->half_tau == tau));
    if (!cont) {
        if ((rc = settle(c))) return rc;
        if ((rc = ensure_forces(c))) return rc;
        // opening half kick (backwards): velocities move to t - dt/2
        LAUNCH(k_kick<R>, cdiv(n, 256), 256, (V4 *)c->vel, (const V4 *)c->frc, n, (R)(-0.5 * dt));
    }
    LAUNCH(k_ctl_begin, 1, 1, c->ctl, (long long)steps, 0, 0.0, 0.0, 0.0, 0);
    if (nvt) {
        if (!c->ke_partial && (rc = dev_alloc(c, &c->ke_partial, (size_t)cdiv(c->ncap, FORCE_THREADS) + c->ncap / 8 + 65))) return rc;
        // (a continued run keeps the coefficients its last step left in the control block)
        if (!cont) LAUNCH(k_thermostat_begin, 1, 1, c->ctl, th_kind, T0, tau, dt, 3.0 * (double)std::max<int64_t>(c->n - 1, 1));
    }
    c->min_resume = false;
    for (int k = 0; k < 8; k++) c->prof[k] = 0;
    long long rebuilds0 = c->hctl->rebuilds;
    StepArgs<R> a;
    a.ctl = c->ctl;
    a.pos[0] = (P4 *)c->pos[0];
    a.pos[1] = (P4 *)c->pos[1];
    a.vel = (V4 *)c->vel;
    a.nbr = c->nbr;
    a.nnb = c->nnb;
    a.orig = c->orig;
    a.img = c->img;
    a.n = n;
    a.cap = c->cap;
    a.kick = (R)dt;
    a.drift = (R)dt;
    a.box = make_box<R>(c);
    a.dd = 0;
    a.vmax_out = nullptr;
    a.vmax_mail = nullptr;
    a.peer_state = nullptr;
    a.peer_commit_state = nullptr;
    a.peer_mail_in = nullptr;
    a.peer_gv_out = nullptr;
    a.peer_drift = 0.0;
    a.halo_map = nullptr;
    a.nranks = 1;
    a.ke_partial = c->ke_partial;
    int64_t done = 0;
    while (done < steps) {
        // speculative batches: [guarded refresh][batch x step]; steps after the
        // displacement bound trips are no-ops until the next refresh
        int64_t remaining = steps - done;
        int64_t quota = std::min<int64_t>(remaining + c->batch, 4096);
        size_t nev = 0;
        auto mark = [&]() {
            if (!c->profile) return;
            if (nev == c->pev.size()) { cudaEvent_t e; cudaEventCreate(&e); c->pev.push_back(e); }
            cudaEventRecord(c->pev[nev++], c->stream);
        };
        for (int64_t launched = 0; launched < quota;) {
            mark();
            enqueue_rebuild(c);
            mark();
            int m = (int)std::min<int64_t>(c->batch, quota - launched);
            for (int k = 0; k < m; k++) {
                if (nvt) enqueue_step<true>(c, a); else enqueue_step<false>(c, a);
            }
            launched += m;
            c->prof[2] += m;
        }
        mark();
        long long rb0 = c->hctl->rebuilds;
        if ((rc = pull_ctl(c))) return rc;
        CU(cudaGetLastError());
        for (size_t k = 0; k + 2 < nev; k += 2) {   // [k] refresh [k+1] steps [k+2]
            float t_r = 0, t_s = 0;
            cudaEventElapsedTime(&t_r, c->pev[k], c->pev[k + 1]);
            cudaEventElapsedTime(&t_s, c->pev[k + 1], c->pev[k + 2]);
            c->prof[1] += t_r;
            c->prof[0] += t_s;
        }
        int64_t progressed = c->hctl->steps_done - done;
        done = c->hctl->steps_done;
        if (c->hctl->overflow) {
            if ((rc = grow_after_overflow(c))) return rc;
            a.nbr = c->nbr;
            a.cap = c->cap;
            c->hctl->overflow = 0u;
            if ((rc = push_ctl(c))) return rc;
        } else {
            long long nrb = c->hctl->rebuilds - rb0;
            if (nrb > 0 && progressed > 0)   // refresh period seen in this round -> batch length
                c->batch = (int)std::min<int64_t>(std::max<int64_t>(progressed / nrb, 1), 64);
        }
    }
    c->prof[3] = (double)steps;
    c->prof[4] = (double)(c->hctl->rebuilds - rebuilds0);
    // the closing half kick is left to settle()
    c->list_dirty = c->hctl->halt_rebuild != 0u;
    c->forces_valid = c->ev_valid = false;
    c->half_pending = true;
    c->half_dt = dt;
    c->half_kind = th_kind;
    c->half_T0 = T0;
    c->half_tau = tau;
    c->steps_total += steps;
    return B2_OK;
}

// velocities to the full step: forces of the final configuration, closing half kick
template <typename R>
int Impl<R>::settle(b2_ctx *c)
{
    int rc;
    if (!c->half_pending) return B2_OK;
    c->half_pending = false;
    if (c->n == 0) return B2_OK;
    const int n = (int)c->n;
    if ((rc = ensure_forces(c))) return rc;
    LAUNCH(k_kick<R>, cdiv(n, 256), 256, (V4 *)c->vel, (const V4 *)c->frc, n, (R)(0.5 * c->half_dt));
    return B2_OK;
}

template <typename R>
int Impl<R>::minimize(b2_ctx *c, int method, int64_t max_steps, double dt, double tol,
                      int64_t *steps_done, int *converged)
{
    int rc;
    *steps_done = 0;
    *converged = 0;
    if (c->n == 0) { *converged = 1; return B2_OK; }
    int n = (int)c->n;
    if ((rc = ensure_list(c))) return rc;
    const bool resume = c->min_resume && c->min_method == method && c->mvel;
    if (resume) {
        auto park = k_park_minvel<R>;
        LAUNCH(park, cdiv(n, 256), 256, n, 0, (V4 *)c->vel, (V4 *)c->mvel, c->orig);
    } else {
        LAUNCH(k_scale_vel<R>, cdiv(n, 256), 256, (V4 *)c->vel, n, (R)0);
    }
    LAUNCH(k_ctl_begin, 1, 1, c->ctl, (long long)max_steps, method == B2_MIN_FIRE ? MIN_MODE_FIRE : MIN_MODE_SD,
           dt, 10.0 * dt, tol, resume ? 1 : 0);
    // k_ctl_begin marks a zero budget as done; the first iteration must still test convergence
    MinArgs<R> a;
    a.ctl = c->ctl;
    a.pos[0] = (P4 *)c->pos[0];
    a.pos[1] = (P4 *)c->pos[1];
    a.vel = (V4 *)c->vel;
    a.frc = (const V4 *)c->frc;
    a.orig = c->orig;
    a.img = c->img;
    a.n = n;
    a.box = make_box<R>(c);
    int nblk = std::min(MAX_RED_BLOCKS, cdiv(n, RED_THREADS));
    if (max_steps == 0) {
        if ((rc = pull_ctl(c))) return rc;
        c->hctl->done = 0u;
        if ((rc = push_ctl(c))) return rc;
    }
    for (;;) {
        int64_t budget = std::min<int64_t>(max_steps + 1, 2048);
        for (int64_t launched = 0; launched < budget;) {
            enqueue_rebuild(c);
            int m = (int)std::min<int64_t>(c->batch, budget - launched);
            for (int k = 0; k < m; k++) {
                enqueue_force<true, false>(c, true);
                LAUNCH(k_min_partial<R>, nblk, RED_THREADS, c->ctl, (const V4 *)c->vel, (const V4 *)c->frc, n, c->partial);
                LAUNCH(k_min_logic, 1, 1, c->ctl, c->partial, nblk, n);
                LAUNCH(k_min_step<R>, cdiv(n, RED_THREADS), RED_THREADS, a);
            }
            launched += m;
        }
        if ((rc = pull_ctl(c))) return rc;
        CU(cudaGetLastError());
        if (c->hctl->overflow) {
            if ((rc = grow_after_overflow(c))) return rc;
            c->hctl->overflow = 0u;
            if ((rc = push_ctl(c))) return rc;
            continue;
        }
        if (c->hctl->done) break;
    }
    *steps_done = c->hctl->iter;
    *converged = (int)c->hctl->converged;
    c->steps_total += c->hctl->iter;
    // an unfinished minimisation can be continued by the next call (Optimization drives the
    // backend in chunks, optimization/core.py:56): park the minimiser's velocities
    c->min_resume = false;
    if (!c->hctl->converged) {
        if (!c->mvel && (rc = dev_alloc(c, (V4 **)&c->mvel, n))) return rc;
        auto park = k_park_minvel<R>;
        LAUNCH(park, cdiv(n, 256), 256, n, 1, (V4 *)c->vel, (V4 *)c->mvel, c->orig);
        c->min_resume = true;
        c->min_method = method;
    }
    // forces / energies of the final configuration are the ones just computed
    LAUNCH(k_scale_vel<R>, cdiv(n, 256), 256, (V4 *)c->vel, n, (R)0);
    c->list_dirty = c->hctl->halt_rebuild != 0u;
    c->forces_valid = c->ev_valid = true;
    return B2_OK;
}

// ---- domain decomposition primitives (no communication in here) ---------------
template <typename R>
int Impl<R>::dd_begin(b2_ctx *c, int64_t steps, double dt, int opening_kick)
{
    int rc;
    int n = (int)c->n;
    // after a mid-run refresh the velocities already live at half steps: only the list is needed
    if ((rc = opening_kick ? ensure_forces(c) : ensure_list(c))) return rc;
    c->dd_dt = dt;
    c->min_resume = false;
    if (n > 0 && opening_kick)
        LAUNCH(k_kick<R>, cdiv(n, 256), 256, (V4 *)c->vel, (const V4 *)c->frc, n, (R)(-0.5 * dt));
    LAUNCH(k_ctl_begin, 1, 1, c->ctl, (long long)steps, 0, 0.0, 0.0, 0.0, 0);
    return B2_OK;
}

template <typename R>
int Impl<R>::dd_step(b2_ctx *c, float *vmax_out)
{
    int n = (int)c->n;
    StepArgs<R> a;
    a.ctl = c->ctl;
    a.pos[0] = (P4 *)c->pos[0];
    a.pos[1] = (P4 *)c->pos[1];
    a.vel = (V4 *)c->vel;
    a.nbr = c->nbr;
    a.nnb = c->nnb;
    a.orig = c->orig;
    a.img = c->img;
    a.n = n;
    a.cap = c->cap;
    a.kick = (R)c->dd_dt;
    a.drift = (R)c->dd_dt;
    a.box = make_box<R>(c);
    a.dd = 1;
    a.vmax_out = vmax_out;
    a.vmax_mail = c->peer_on ? c->peer_mail : nullptr;
    a.peer_state = c->peer_on ? c->peer_ticket + 2 : nullptr;
    a.nranks = c->nranks;
    // b2_peer_run: the step kernel's last block settles the step itself (no k_peer_commit launch)
    const bool fused = c->peer_on && c->peer_fused && n > 0;
    a.peer_commit_state = fused ? c->peer_ticket + 2 : nullptr;
    a.peer_mail_in = fused ? (const unsigned long long *)c->peer_buf : nullptr;
    a.peer_gv_out = fused ? c->dd_vmax + 1 : nullptr;
    a.peer_drift = c->dd_dt;
    a.halo_map = fused && c->peer_push_fused ? c->halo_map : nullptr;
    for (int par = 0; par < 2; par++) {
        a.halo_out_l[par] = c->halo_out_l[par];
        a.halo_out_r[par] = c->halo_out_r[par];
    }
    a.halo_flag_l = c->halo_flag_l;
    a.halo_flag_r = c->halo_flag_r;
    a.ke_partial = nullptr;
    if (n > 0) enqueue_step<false>(c, a);
    return B2_OK;
}

template <typename R>
int Impl<R>::dd_end(b2_ctx *c)
{
    int rc;
    int n = (int)c->n;
    c->forces_valid = c->ev_valid = false;
    if ((rc = ensure_forces(c))) return rc;
    if (n > 0) LAUNCH(k_kick<R>, cdiv(n, 256), 256, (V4 *)c->vel, (const V4 *)c->frc, n, (R)(0.5 * c->dd_dt));
    return B2_OK;
}

template <typename R>
int Impl<R>::dd_pack(b2_ctx *c, const int32_t *ids_dev, int64_t n, double *out_dev)
{
    if (n > 0)
        LAUNCH(k_dd_pack<R>, cdiv(n, 256), 256, c->ctl, (const P4 *)c->pos[0], (const P4 *)c->pos[1], c->inv,
               ids_dev, (int)n, make_box<R>(c), out_dev);
    return B2_OK;
}

template <typename R>
int Impl<R>::dd_unpack(b2_ctx *c, int64_t first, int64_t n, const double *in_dev)
{
    if (n > 0)
        LAUNCH(k_dd_unpack<R>, cdiv(n, 256), 256, c->ctl, (P4 *)c->pos[0], (P4 *)c->pos[1], c->inv, (int)first,
               (int)n, make_box<R>(c), in_dev);
    return B2_OK;
}

// A new box for RESIDENT particles is an affine rescaling: fractional coordinates are kept, so distances
// scale with the sides (what a barostat step does; a re-upload follows anyway when the caller replaces the
// particles).  fp32 positions are stored as fractions of the cell and need no touch; fp64 positions and the
// rmsd origin are lengths relative to the centre and are multiplied by new / old.
template <typename R>
__global__ void k_rescale_box(int n, double rx, double ry, double rz, typename Pos4<R>::type *pos0,
                              typename Pos4<R>::type *pos1, double *__restrict__ ref)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (sizeof(R) == 8) {
        double4 *p0 = reinterpret_cast<double4 *>(pos0), *p1 = reinterpret_cast<double4 *>(pos1);
        double4 a = p0[i], b = p1[i];
        a.x *= rx; a.y *= ry; a.z *= rz;
        b.x *= rx; b.y *= ry; b.z *= rz;
        p0[i] = a; p1[i] = b;
    }
    ref[3 * i] *= rx; ref[3 * i + 1] *= ry; ref[3 * i + 2] *= rz;
}

// Hessian of the pair energy, dense (b2_hessian): block i takes particle with ORIGINAL id i, its threads
// the partners j; the diagonal block is minus the sum of the off-diagonal ones of the row.
template <typename R>
__global__ void __launch_bounds__(128)
k_hessian(const Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0, const typename Pos4<R>::type *pos1,
          const int *__restrict__ inv, int n, Box<R> box, double *__restrict__ out)
{
    typedef typename Pos4<R>::type P4;
    const P4 *__restrict__ pos = ctl->cur ? pos1 : pos0;
    const int i = blockIdx.x;
    const P4 pi = pos[inv[i]];
    const int ti = ptype(pi) * g_pair.nsp;
    double diag[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const size_t N = (size_t)n;
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        if (j != i) {
            const P4 pj = pos[inv[j]];
            const int a = ti + ptype(pj);
            double d[3];
            delta_d(pi, pj, box, d[0], d[1], d[2]);
            const double r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
            if (r2 < g_pair.rc2[a]) {
                const double i2 = 1.0 / r2, i6 = i2 * i2 * i2;
                const double c12 = g_pair.c12[a], c6 = g_pair.c6[a];
                const double up_r = (-12.0 * c12 * i6 + 6.0 * c6) * i6 * i2;        // u'(r) / r
                const double upp = (156.0 * c12 * i6 - 42.0 * c6) * i6 * i2;        // u''(r)
                const double k = (upp - up_r) * i2;
                for (int p = 0; p < 3; p++)
                    for (int q = 0; q < 3; q++) {
                        const double v = k * d[p] * d[q] + (p == q ? up_r : 0.0);   // d2u / dx_ip dx_iq
                        h[3 * p + q] = -v;                                          // ... / dx_ip dx_jq
                        diag[3 * p + q] += v;
                    }
            }
        }
        if (j != i)
            for (int p = 0; p < 3; p++)
                for (int q = 0; q < 3; q++) out[((p * N + i) * 3 + q) * N + j] = h[3 * p + q];
    }
    __shared__ double s_d[4][9];
    for (int k = 0; k < 9; k++) {
        const double v = warp_sum(diag[k]);
        if ((threadIdx.x & 31) == 0) s_d[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 9) {
        const int p = threadIdx.x / 3, q = threadIdx.x % 3;
        out[((p * N + i) * 3 + q) * N + i] = s_d[0][threadIdx.x] + s_d[1][threadIdx.x] + s_d[2][threadIdx.x] + s_d[3][threadIdx.x];
    }
}

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
#define DISPATCH(fn, ...) (c->precision == B2_FP32 ? Impl<float>::fn(__VA_ARGS__) : Impl<double>::fn(__VA_ARGS__))

struct Timed {
    b2_ctx *c;
    int64_t l0;
    explicit Timed(b2_ctx *ctx) : c(ctx), l0(ctx->launches) { cudaEventRecord(c->ev0, c->stream); }
    ~Timed()
    {
        cudaEventRecord(c->ev1, c->stream);
        cudaEventSynchronize(c->ev1);
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        c->last_ms = ms;
        c->last_launches = c->launches - l0;
    }
};

extern "C" {

int b2_abi_version(void) { return B2_ABI_VERSION; }
const char *b2_backend_name(void) { return "cuda-sm_100a"; }
const char *b2_last_error(const b2_ctx *c) { return c ? c->err : g_err; }

int b2_create(int device, int precision, b2_ctx **out)
{
    if (!out) return fail(nullptr, B2_EINVAL, "b2_create: out is NULL");
    *out = nullptr;
    if (precision != B2_FP32 && precision != B2_FP64)
        return fail(nullptr, B2_EINVAL, "b2_create: precision must be 32 or 64");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, B2_ENOCUDA, "b2_create: no usable CUDA device (%s); this library has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= ndev || device >= 64)
        return fail(nullptr, B2_EINVAL, "b2_create: device %d out of range (0..%d)", device, ndev - 1);
    b2_ctx *ctx = new b2_ctx();
    ctx->device = device;
    ctx->precision = precision;
    memset(&ctx->pc, 0, sizeof(PairConst));
    memset(&ctx->grid, 0, sizeof(Grid));
    int rc = B2_OK;
    do {
        if ((e = cudaSetDevice(device)) != cudaSuccess) break;
        if ((e = cudaDeviceGetAttribute(&ctx->sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) break;
        {
            int optin = 0;
            if ((e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)) != cudaSuccess) break;
            ctx->smem_optin = (size_t)std::max(optin, 0);
            int per_sm = 0;
            if ((e = cudaDeviceGetAttribute(&per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device)) != cudaSuccess) break;
            ctx->smem_sm = (size_t)std::max(per_sm, 0);
        }
        if ((e = cudaMalloc((void **)&ctx->ctl, sizeof(Ctl))) != cudaSuccess) break;
        if ((e = cudaMemset(ctx->ctl, 0, sizeof(Ctl))) != cudaSuccess) break;
        if ((e = cudaMallocHost((void **)&ctx->hctl, sizeof(Ctl))) != cudaSuccess) break;
        memset(ctx->hctl, 0, sizeof(Ctl));
        if ((e = cudaEventCreate(&ctx->ev0)) != cudaSuccess) break;
        if ((e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) break;
    } while (0);
    if (e != cudaSuccess) {
        rc = fail(nullptr, B2_ECUDA, "b2_create: %s", cudaGetErrorString(e));
        delete ctx;
        return rc;
    }
    *out = ctx;
    return B2_OK;
}

int b2_destroy(b2_ctx *c)
{
    if (!c) return B2_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    void *ptrs[] = {c->pos[0], c->pos[1], c->vel, c->frc, c->vir, c->tpos, c->tvel, c->orig, c->torig, c->inv, c->halo_buf, c->dd_vmax,
                    c->img, c->ref, c->mvel, c->ke_partial, c->keys[0], c->keys[1], c->vals[0], c->vals[1], c->hist,
                    c->cell_start, c->cell_end, c->nbr, c->nnb, c->ctl, c->partial, c->stage, c->stage_i, c->b8start,
                    c->tile_base, c->sbt, c->tab_u, c->tab_w, c->tab_wf, c->halo_map};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    for (int r = 0; r < 64; r++)
        if (c->peer_map[r] && c->peer_map[r] != c->peer_buf) cudaIpcCloseMemHandle(c->peer_map[r]);
    if (c->peer_buf) cudaFree(c->peer_buf);
    if (c->peer_mail) cudaFree(c->peer_mail);
    if (c->peer_ticket) cudaFree(c->peer_ticket);
    {
        void *ddr[] = {c->ddr_cls, c->ddr_blk, c->ddr_keep, c->ddr_go[0], c->ddr_go[1], c->ddr_near[0], c->ddr_near[1],
                       c->ddr_near[2], c->ddr_near[3], c->ddr_send[0], c->ddr_send[1], c->ddr_cnt, c->ddr_gid[0],
                       c->ddr_gid[1], c->ddr_cimg[0], c->ddr_cimg[1]};
        // the send lists may be the ones a context still points at; they die with it
        for (void *q : ddr)
            if (q) cudaFree(q);
        if (c->ddr_hcnt) cudaFreeHost(c->ddr_hcnt);
    }
    if (c->hctl) cudaFreeHost(c->hctl);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    for (cudaEvent_t e : c->pev) cudaEventDestroy(e);
    table_forget(c->device, c->const_serial);
    delete c;
    return B2_OK;
}

int b2_set_stream(b2_ctx *c, void *cuda_stream)
{
    if (!c) return fail(nullptr, B2_EINVAL, "NULL context");
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->stream = (cudaStream_t)cuda_stream;
    return B2_OK;
}

int b2_set_box(b2_ctx *c, const double side[3], const double center[3])
{
    if (!c || !side) return fail(c, B2_EINVAL, "b2_set_box: NULL argument");
    for (int d = 0; d < 3; d++)
        if (!(side[d] > 0.0)) return fail(c, B2_EINVAL, "b2_set_box: side must be > 0");
    if (c->have_box && c->pos[0] && c->n > 0 &&
        (side[0] != c->side[0] || side[1] != c->side[1] || side[2] != c->side[2])) {
        int rc_ = bind_device(c);
        if (!rc_ && c->have_table) rc_ = bind_table(c);
        if (!rc_) rc_ = DISPATCH(settle, c);
        if (rc_) return rc_;
        const int n = (int)c->n;
        const double rx = side[0] / c->side[0], ry = side[1] / c->side[1], rz = side[2] / c->side[2];
        if (c->precision == B2_FP32)
            LAUNCH(k_rescale_box<float>, cdiv(n, 256), 256, n, rx, ry, rz, (int4 *)c->pos[0], (int4 *)c->pos[1], c->ref);
        else
            LAUNCH(k_rescale_box<double>, cdiv(n, 256), 256, n, rx, ry, rz, (double4 *)c->pos[0], (double4 *)c->pos[1], c->ref);
    }
    for (int d = 0; d < 3; d++) {
        c->side[d] = side[d];
        c->center[d] = center ? center[d] : 0.0;
    }
    c->have_box = true;
    c->list_dirty = true;
    c->forces_valid = c->ev_valid = false;
    return B2_OK;
}

int b2_set_pair_table(b2_ctx *c, int nsp, const double *eps, const double *sigma, const double *rcut, int shifted)
{
    if (!c || !eps || !sigma || !rcut) return fail(c, B2_EINVAL, "b2_set_pair_table: NULL argument");
    if (nsp < 1 || nsp > B2_MAXSP) return fail(c, B2_EINVAL, "b2_set_pair_table: 1 <= nsp <= %d", B2_MAXSP);
    if (c->pos[0] && c->half_pending) {       // an open run is closed with the forces it was run with
        int rc_ = bind_device(c);
        if (!rc_) rc_ = DISPATCH(settle, c);
        if (rc_) return rc_;
    }
    for (int a = 0; a < nsp * nsp; a++)
        if (!(sigma[a] > 0.0) || !(rcut[a] > 0.0))
            return fail(c, B2_EINVAL, "b2_set_pair_table: sigma and rcut must be > 0");
    memset(&c->pc, 0, sizeof(PairConst));
    c->tabulated = false;
    c->nsp = c->pc.nsp = nsp;
    c->shifted = shifted;
    c->rcut_max = 0;
    for (int a = 0; a < nsp * nsp; a++) {
        // same expression order as the oracles (oracle_np.PairTable, oracle_c.c)
        double s2 = sigma[a] * sigma[a];
        double s6 = s2 * s2 * s2;
        c->pc.c6[a] = 4.0 * eps[a] * s6;
        c->pc.c12[a] = 4.0 * eps[a] * s6 * s6;
        c->rcut[a] = rcut[a];
        c->pc.rc2[a] = rcut[a] * rcut[a];
        double ir2 = 1.0 / c->pc.rc2[a];
        double ir6 = ir2 * ir2 * ir2;
        c->pc.ushift[a] = shifted ? ir6 * (c->pc.c12[a] * ir6 - c->pc.c6[a]) : 0.0;
        c->rcut_max = std::max(c->rcut_max, rcut[a]);
    }
    rebuild_table(c);
    c->have_table = true;
    c->list_dirty = true;
    c->forces_valid = c->ev_valid = false;
    c->cap = 0;
    return B2_OK;
}

int b2_set_pair_tabulated(b2_ctx *c, int nsp, int npoints, const double *rcut, const double *u, const double *w)
{
    if (!c || !rcut || !u || !w) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: NULL argument");
    if (nsp < 1 || nsp > B2_MAXSP) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: 1 <= nsp <= %d", B2_MAXSP);
    if (npoints < 2 || npoints > (1 << 24)) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: 2 <= npoints <= 2^24");
    for (int a = 0; a < nsp * nsp; a++)
        if (!(rcut[a] > 0.0)) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: rcut must be > 0");
    { int rc_ = bind_device(c); if (rc_) return rc_; }
    if (c->pos[0] && c->half_pending) { int rc_ = DISPATCH(settle, c); if (rc_) return rc_; }
    const size_t count = (size_t)nsp * nsp * npoints;
    if (c->tab_u) cudaFree(c->tab_u);
    if (c->tab_w) cudaFree(c->tab_w);
    if (c->tab_wf) cudaFree(c->tab_wf);
    c->tab_u = c->tab_w = nullptr;
    c->tab_wf = nullptr;
    CU(cudaMalloc((void **)&c->tab_u, count * sizeof(double)));
    CU(cudaMalloc((void **)&c->tab_w, count * sizeof(double)));
    CU(cudaMalloc((void **)&c->tab_wf, count * sizeof(float)));
    std::vector<float> wf(count);
    for (size_t k = 0; k < count; k++) wf[k] = (float)w[k];
    CU(cudaMemcpy(c->tab_u, u, count * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->tab_w, w, count * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->tab_wf, wf.data(), count * sizeof(float), cudaMemcpyHostToDevice));
    memset(&c->pc, 0, sizeof(PairConst));
    c->tabulated = true;
    c->nsp = c->pc.nsp = nsp;
    c->shifted = 0;
    c->rcut_max = 0;
    c->pc.tab_np = npoints;
    c->pc.tab_u = c->tab_u;
    c->pc.tab_w = c->tab_w;
    c->pc.tab_wf = c->tab_wf;
    for (int a = 0; a < nsp * nsp; a++) {
        c->rcut[a] = rcut[a];
        c->pc.rc2[a] = rcut[a] * rcut[a];
        c->pc.tab_scale[a] = (double)(npoints - 1) / c->pc.rc2[a];
        c->rcut_max = std::max(c->rcut_max, rcut[a]);
    }
    rebuild_table(c);
    c->have_table = true;
    c->list_dirty = true;
    c->forces_valid = c->ev_valid = false;
    c->cap = 0;
    return B2_OK;
}

int b2_set_skin(b2_ctx *c, double skin)
{
    if (!c || !(skin >= 0.0)) return fail(c, B2_EINVAL, "b2_set_skin: skin must be >= 0");
    c->skin = skin;
    if (c->have_table) rebuild_table(c);
    c->list_dirty = true;
    c->cap = 0;
    return B2_OK;
}

#define NEED_PARTICLES(who)                                                                   \
    if (!c) return fail(nullptr, B2_EINVAL, who ": NULL context");                            \
    if (!c->pos[0]) return fail(c, B2_ESTATE, who ": no particles resident (call b2_upload)"); \
    { int rc_ = bind_device(c); if (rc_) return rc_; }                                        \
    if (c->have_table) { int rc_ = bind_table(c); if (rc_) return rc_; }                      \
    TableGuard table_guard_(c);

// ... and that must see full-step velocities / the forces of the current configuration
#define NEED_SETTLED(who)                                                                     \
    NEED_PARTICLES(who)                                                                       \
    { int rc_ = DISPATCH(settle, c); if (rc_) return rc_; }

int b2_upload(b2_ctx *c, int64_t n, const double *pos, const double *vel, const int32_t *species,
              const double *mass)
{
    if (!c || n < 0 || (n > 0 && (!pos || !species))) return fail(c, B2_EINVAL, "b2_upload: bad argument");
    if (n > 0x7fffff00LL) return fail(c, B2_EINVAL, "b2_upload: at most 2^31 particles per context");
    if (!c->have_box) return fail(c, B2_ESTATE, "b2_upload: call b2_set_box first");
    if (!c->have_table) return fail(c, B2_ESTATE, "b2_upload: call b2_set_pair_table first");
    for (int64_t i = 0; i < n; i++) {
        if (species[i] < 0 || (species[i] & ~B2_GHOST) >= c->nsp)
            return fail(c, B2_EINVAL, "b2_upload: species id %d of particle %lld outside the pair table",
                        species[i], (long long)i);
        if (mass && !(mass[i] > 0.0)) return fail(c, B2_EINVAL, "b2_upload: mass must be > 0");
    }
    int rc = bind_device(c);
    if (rc) return rc;
    return DISPATCH(upload, c, n, pos, vel, species, mass);
}

int b2_update(b2_ctx *c, const double *pos, const double *vel)
{
    NEED_SETTLED("b2_update");
    return DISPATCH(update, c, pos, vel);
}

int b2_download(b2_ctx *c, double *pos, double *vel, int32_t *image)
{
    NEED_SETTLED("b2_download");
    return DISPATCH(download, c, pos, vel, image);
}

int b2_scale_velocities(b2_ctx *c, double factor)
{
    NEED_SETTLED("b2_scale_velocities");
    if (c->n == 0) return B2_OK;
    c->min_resume = false;
    if (c->precision == B2_FP32)
        LAUNCH(k_scale_vel<float>, cdiv(c->n, 256), 256, (float4 *)c->vel, (int)c->n, (float)factor);
    else
        LAUNCH(k_scale_vel<double>, cdiv(c->n, 256), 256, (double4 *)c->vel, (int)c->n, factor);
    return B2_OK;
}

int b2_mark_reference(b2_ctx *c)
{
    NEED_PARTICLES("b2_mark_reference");
    return DISPATCH(mark_reference, c);
}

int b2_compute(b2_ctx *c, int what, double *energy, double *virial, double *stress9, double *forces,
               double *pe, double *vir)
{
    NEED_SETTLED("b2_compute");
    if (what & ~(B2_ENERGY | B2_FORCES | B2_VIRIAL | B2_STRESS))
        return fail(c, B2_EUNSUPPORTED, "b2_compute: unsupported observable mask 0x%x", what);
    Timed t(c);
    return DISPATCH(compute, c, what, energy, virial, stress9, forces, pe, vir);
}

int b2_thermo(b2_ctx *c, double out[B2_T_LEN])
{
    NEED_SETTLED("b2_thermo");
    if (!out) return fail(c, B2_EINVAL, "b2_thermo: out is NULL");
    return DISPATCH(thermo, c, out, false, nullptr);
}

int b2_neighbors(b2_ctx *c, int32_t *count, int32_t *nbr, int32_t stride, int32_t *max_count)
{
    NEED_PARTICLES("b2_neighbors");
    return DISPATCH(neighbors, c, count, nbr, stride, max_count);
}

int b2_run_nve(b2_ctx *c, int64_t steps, double dt)
{
    NEED_PARTICLES("b2_run_nve");
    if (steps < 0 || !(dt > 0.0)) return fail(c, B2_EINVAL, "b2_run_nve: steps >= 0 and dt > 0 required");
    Timed t(c);
    return DISPATCH(run_nve, c, steps, dt);
}

int b2_run_nvt(b2_ctx *c, int64_t steps, double dt, double temperature, double relaxation_time, int kind)
{
    NEED_PARTICLES("b2_run_nvt");
    if (steps < 0 || !(dt > 0.0) || !(temperature > 0.0) || !(relaxation_time > 0.0) || (kind != 0 && kind != 1))
        return fail(c, B2_EINVAL, "b2_run_nvt: steps >= 0, dt, temperature, relaxation_time > 0, kind 0 | 1 required");
    Timed t(c);
    return DISPATCH(run_nve, c, steps, dt, kind, temperature, relaxation_time);
}

int b2_minimize(b2_ctx *c, int method, int64_t max_steps, double dt, double fn2_tol, int64_t *steps_done,
                int *converged)
{
    NEED_SETTLED("b2_minimize");
    if (max_steps < 0 || !(dt > 0.0) || !steps_done || !converged)
        return fail(c, B2_EINVAL, "b2_minimize: bad argument");
    if (method != B2_MIN_SD && method != B2_MIN_FIRE) return fail(c, B2_EINVAL, "b2_minimize: unknown method %d", method);
    Timed t(c);
    return DISPATCH(minimize, c, method, max_steps, dt, fn2_tol, steps_done, converged);
}

int b2_stats(b2_ctx *c, int64_t out[B2_S_LEN])
{
    if (!c || !out) return fail(c, B2_EINVAL, "b2_stats: NULL argument");
    for (int k = 0; k < B2_S_LEN; k++) out[k] = 0;
    out[B2_S_STEPS] = c->steps_total;
    out[B2_S_REBUILDS] = c->hctl ? c->hctl->rebuilds : 0;
    out[B2_S_LAUNCHES] = c->launches;
    out[B2_S_MAX_NEIGH] = c->hctl ? c->hctl->max_count : 0;
    out[B2_S_NEIGH_CAP] = c->cap;
    out[B2_S_LIST_MODE] = c->grid.allpairs ? 1 : (c->grid.win ? 2 : 0);
    out[B2_S_NCELLS] = (int64_t)c->grid.nc[0] * c->grid.nc[1] * c->grid.nc[2];
    out[B2_S_WIN_BLOCKS] = c->grid.win ? c->nsb_total : 0;
    out[B2_S_SMALL_BOX] = (c->have_box && c->have_table && 2.0 * c->rcut_max > std::min(c->side[0], std::min(c->side[1], c->side[2]))) ? 1 : 0;
    return B2_OK;
}

int b2_last_timing(b2_ctx *c, double *ms, int64_t *launches)
{
    if (!c) return fail(nullptr, B2_EINVAL, "NULL context");
    if (ms) *ms = c->last_ms;
    if (launches) *launches = c->last_launches;
    return B2_OK;
}

int b2_upload_device(b2_ctx *c, int64_t n, const double *pos_dev, const double *vel_dev,
                     const int32_t *species_dev, const double *mass_dev)
{
    if (!c || n < 0 || (n > 0 && (!pos_dev || !species_dev))) return fail(c, B2_EINVAL, "b2_upload_device: bad argument");
    if (n > 0x7fffff00LL) return fail(c, B2_EINVAL, "b2_upload_device: at most 2^31 particles per context");
    if (!c->have_box || !c->have_table) return fail(c, B2_ESTATE, "b2_upload_device: set the box and the pair table first");
    int rc = bind_device(c);
    if (rc) return rc;
    return DISPATCH(upload, c, n, pos_dev, vel_dev, species_dev, mass_dev);
}

int b2_download_device(b2_ctx *c, double *pos_dev, double *vel_dev, int32_t *image_dev)
{
    NEED_SETTLED("b2_download_device");
    return DISPATCH(download, c, pos_dev, vel_dev, image_dev);
}

int b2_dd_begin(b2_ctx *c, int64_t steps, double dt, int opening_kick)
{
    NEED_SETTLED("b2_dd_begin");
    if (steps < 0 || !(dt > 0.0)) return fail(c, B2_EINVAL, "b2_dd_begin: steps >= 0 and dt > 0 required");
    return DISPATCH(dd_begin, c, steps, dt, opening_kick);
}

int b2_dd_step(b2_ctx *c, float *vmax2_out_dev)
{
    NEED_PARTICLES("b2_dd_step");
    if (!vmax2_out_dev) return fail(c, B2_EINVAL, "b2_dd_step: NULL argument");
    if (!(c->dd_dt > 0.0)) return fail(c, B2_ESTATE, "b2_dd_step: call b2_dd_begin first");
    if (c->n == 0) CU(cudaMemsetAsync(vmax2_out_dev, 0, sizeof(float), c->stream));
    return DISPATCH(dd_step, c, vmax2_out_dev);
}

int b2_dd_commit(b2_ctx *c, const float *gvmax2_dev)
{
    NEED_PARTICLES("b2_dd_commit");
    if (!gvmax2_dev) return fail(c, B2_EINVAL, "b2_dd_commit: NULL argument");
    LAUNCH(k_dd_commit, 1, 1, c->ctl, gvmax2_dev, c->dd_dt);
    return B2_OK;
}

int b2_dd_poll(b2_ctx *c, int64_t *steps_done, int *need_refresh)
{
    NEED_PARTICLES("b2_dd_poll");
    int rc = pull_ctl(c);
    if (rc) return rc;
    CU(cudaGetLastError());
    if (steps_done) *steps_done = c->hctl->steps_done;
    if (need_refresh) *need_refresh = (c->hctl->halt_rebuild || c->hctl->overflow) ? 1 : 0;
    if (c->hctl->halt_rebuild) c->list_dirty = true;
    if (c->hctl->peer_error)
        return fail(c, B2_ECUDA, "peer-memory halo: %s", c->hctl->peer_error == 1u
                    ? "timed out waiting for a neighbouring rank" : "a rank ran ahead of the step protocol");
    return B2_OK;
}

int b2_dd_end(b2_ctx *c)
{
    NEED_PARTICLES("b2_dd_end");
    return DISPATCH(dd_end, c);
}

int b2_dd_pack(b2_ctx *c, const int32_t *ids_dev, int64_t n, double *out_dev)
{
    NEED_PARTICLES("b2_dd_pack");
    if (n < 0 || (n > 0 && (!ids_dev || !out_dev))) return fail(c, B2_EINVAL, "b2_dd_pack: bad argument");
    return DISPATCH(dd_pack, c, ids_dev, n, out_dev);
}

int b2_dd_unpack(b2_ctx *c, int64_t first_id, int64_t n, const double *in_dev)
{
    NEED_PARTICLES("b2_dd_unpack");
    if (n < 0 || first_id < 0 || first_id + n > c->n || (n > 0 && !in_dev))
        return fail(c, B2_EINVAL, "b2_dd_unpack: id range outside the resident particles");
    return DISPATCH(dd_unpack, c, first_id, n, in_dev);
}

int b2_set_images(b2_ctx *c, const int32_t *image)
{
    NEED_PARTICLES("b2_set_images");
    if (!image) return fail(c, B2_EINVAL, "b2_set_images: NULL argument");
    if (c->n > 0) CU(cudaMemcpyAsync(c->img, image, 3 * (size_t)c->n * sizeof(int32_t), cudaMemcpyDefault, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return B2_OK;
}

int b2_get_reference(b2_ctx *c, double *ref)
{
    NEED_PARTICLES("b2_get_reference");
    if (!ref) return fail(c, B2_EINVAL, "b2_get_reference: NULL argument");
    const size_t n3 = 3 * (size_t)c->n;
    if (n3) CU(cudaMemcpyAsync(ref, c->ref, n3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (size_t k = 0; k < n3; k++) ref[k] += c->center[k % 3];          // stored relative to the centre
    return B2_OK;
}

int b2_set_reference(b2_ctx *c, const double *ref)
{
    NEED_PARTICLES("b2_set_reference");
    if (!ref) return fail(c, B2_EINVAL, "b2_set_reference: NULL argument");
    const size_t n3 = 3 * (size_t)c->n;
    std::vector<double> rel(n3);
    for (size_t k = 0; k < n3; k++) rel[k] = ref[k] - c->center[k % 3];
    if (n3) CU(cudaMemcpy(c->ref, rel.data(), n3 * sizeof(double), cudaMemcpyHostToDevice));
    return B2_OK;
}

int b2_get_thermostat(b2_ctx *c, double *xi)
{
    NEED_PARTICLES("b2_get_thermostat");
    if (!xi) return fail(c, B2_EINVAL, "b2_get_thermostat: NULL argument");
    int rc = pull_ctl(c);
    if (rc) return rc;
    *xi = c->hctl->th_xi;
    return B2_OK;
}

int b2_set_thermostat(b2_ctx *c, double xi)
{
    NEED_SETTLED("b2_set_thermostat");
    int rc = pull_ctl(c);
    if (rc) return rc;
    c->hctl->th_xi = xi;
    return push_ctl(c);
}

int b2_hessian(b2_ctx *c, double *out)
{
    NEED_PARTICLES("b2_hessian");
    if (!out) return fail(c, B2_EINVAL, "b2_hessian: NULL argument");
    if (c->tabulated) return fail(c, B2_EUNSUPPORTED, "b2_hessian: analytic second derivatives exist for the 12-6 table only");
    const int64_t n = c->n;
    if (n > B2_HESSIAN_MAX_PARTICLES)
        return fail(c, B2_EINVAL, "b2_hessian: n = %lld exceeds B2_HESSIAN_MAX_PARTICLES = %d (the matrix is dense: 9 n^2 doubles)",
                    (long long)n, B2_HESSIAN_MAX_PARTICLES);
    if (n == 0) return B2_OK;
    double *dev = nullptr;
    const size_t count = 9 * (size_t)n * (size_t)n;
    CU(cudaMalloc((void **)&dev, count * sizeof(double)));
    if (c->precision == B2_FP32)
        LAUNCH(k_hessian<float>, (int)n, 128, c->ctl, (const int4 *)c->pos[0], (const int4 *)c->pos[1], c->inv, (int)n,
               make_box<float>(c), dev);
    else
        LAUNCH(k_hessian<double>, (int)n, 128, c->ctl, (const double4 *)c->pos[0], (const double4 *)c->pos[1], c->inv, (int)n,
               make_box<double>(c), dev);
    cudaError_t e = cudaMemcpyAsync(out, dev, count * sizeof(double), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(dev);
    if (e != cudaSuccess) return fail(c, B2_ECUDA, "b2_hessian: %s", cudaGetErrorString(e));
    return B2_OK;
}

int b2_format_xyz(char *out, int64_t cap, int64_t *written, int64_t n, const int32_t *ids, int nnames,
                  const char *const *names, const double *pos, const double *vel, int precision)
{
    if (!written || n < 0 || (n > 0 && (!ids || !names || !pos)) || precision < 0 || precision > 17)
        return fail(nullptr, B2_EINVAL, "b2_format_xyz: bad argument");
    *written = 0;
    const int per = vel ? 6 : 3;
    size_t longest = 1;
    for (int k = 0; k < nnames; k++) { size_t l = strlen(names[k]); if (l > longest) longest = l; }
    /* upper bound of one line from the largest magnitude: ' -' + integer digits + '.' + decimals per number */
    double big = 1.0;
    for (int64_t i = 0; i < 3 * n; i++) {
        double a = fabs(pos[i]), b = vel ? fabs(vel[i]) : 0.0;
        if (isfinite(a) && a > big) big = a;                             /* inf and nan print in four characters */
        if (isfinite(b) && b > big) big = b;
    }
    const size_t digits = (size_t)log10(big) + 2;
    const size_t line = longest + (size_t)per * (digits + (size_t)precision + 3) + 2;
    if (!out) { *written = (int64_t)(line * (size_t)n); return B2_OK; }       /* size query */
    if ((int64_t)(line * (size_t)n) > cap) return fail(nullptr, B2_EINVAL, "b2_format_xyz: output buffer too small");
    /* every chunk of rows is formatted at its worst-case offset, then the chunks are closed up in order */
    const int64_t chunk = 4096;
    const int64_t nchunk = (n + chunk - 1) / chunk;
    int64_t *len = (int64_t *)malloc((size_t)(nchunk > 0 ? nchunk : 1) * sizeof(int64_t));
    if (!len) return fail(nullptr, B2_ENOMEM, "b2_format_xyz: out of memory");
    auto work = [&](int64_t c) {
        char *p = out + (size_t)c * chunk * line, *p0 = p;
        const int64_t hi = (c + 1) * chunk < n ? (c + 1) * chunk : n;
        for (int64_t i = c * chunk; i < hi; i++) {
            const int id = ids[i];
            const char *name = (id >= 0 && id < nnames) ? names[id] : "?";
            size_t l = strlen(name);
            memcpy(p, name, l);
            p += l;
            for (int k = 0; k < 3; k++) p += snprintf(p, digits + (size_t)precision + 4, " %.*f", precision, pos[3 * i + k]);
            if (vel)
                for (int k = 0; k < 3; k++) p += snprintf(p, digits + (size_t)precision + 4, " %.*f", precision, vel[3 * i + k]);
            *p++ = '\n';
        }
        len[c] = (int64_t)(p - p0);
    };
    {
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        const int nthreads = (int)std::min<int64_t>((int64_t)std::min(hw, 64u), std::max<int64_t>(nchunk, 1));
        std::atomic<int64_t> next(0);
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; t++)
            pool.emplace_back([&]() { for (int64_t c = next++; c < nchunk; c = next++) work(c); });
        for (auto &th : pool) th.join();
    }
    int64_t at = 0;
    for (int64_t c = 0; c < nchunk; c++) {
        if (at != c * chunk * (int64_t)line) memmove(out + at, out + (size_t)c * chunk * line, (size_t)len[c]);
        at += len[c];
    }
    free(len);
    *written = at;
    return B2_OK;
}

int b2_host_alloc(void **out, int64_t bytes)
{
    if (!out || bytes < 0) return fail(nullptr, B2_EINVAL, "b2_host_alloc: bad argument");
    *out = nullptr;
    cudaError_t e = cudaMallocHost(out, (size_t)std::max<int64_t>(bytes, 1));
    if (e != cudaSuccess) return fail(nullptr, B2_ENOMEM, "b2_host_alloc: %s", cudaGetErrorString(e));
    return B2_OK;
}

int b2_host_free(void *ptr)
{
    if (ptr) cudaFreeHost(ptr);
    return B2_OK;
}

int b2_set_profile(b2_ctx *c, int on)
{
    if (!c) return fail(nullptr, B2_EINVAL, "NULL context");
    c->profile = on != 0;
    return B2_OK;
}

int b2_profile(b2_ctx *c, double out[8])
{
    if (!c || !out) return fail(c, B2_EINVAL, "b2_profile: NULL argument");
    for (int k = 0; k < 8; k++) out[k] = c->prof[k];
    return B2_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// b2_kinetic_static — the one entry point of this library that allocates
// nothing, neither at set-up nor per call.
//
// Who needs it: host programs that have given the whole GPU to an allocator of
// their own — PyTorch's caching allocator, a JAX-style up-front reservation, an
// MPI+CUDA code with a private arena — and call into this library from inside
// their step loop with raw device pointers.  For them a cudaMalloc issued by a
// library either fails (the pool already holds the memory) or stalls the whole
// device in the middle of a step, so every other entry point here, which sizes
// its reduction scratch with cudaMalloc when particles are uploaded, is off
// limits.  This variant of the thermo reduction (k_thermo_partial +
// k_sum_partials) keeps its per-block partial sums in the fixed array below
// instead.
//
// Limits and sharing rules
//   * at most B2_STATIC_MAX_PARTICLES particles per call; larger n is refused
//     with B2_EINVAL before anything is launched;
//   * the scratch is 256 KiB and is held by every GPU that loads this library
//     until the process exits;
//   * one array per GPU, shared by all calls: two calls that may overlap on the
//     same GPU (different streams, different host threads) race on it — order
//     them on one stream.  The array keeps whatever the previous call left;
//     nothing in it has to start at zero, because the first kernel overwrites
//     every row the second kernel reads;
//   * the first call on each GPU looks up the device address of the array
//     (cudaGetSymbolAddress).  Make that first call during set-up, before any
//     stream capture is begun on that GPU.
// ---------------------------------------------------------------------------
#define B2_STATIC_TILE 4096                                   // particles per block
#define B2_STATIC_ROWS 4096                                   // blocks at most
static_assert((int64_t)B2_STATIC_TILE * B2_STATIC_ROWS == B2_STATIC_MAX_PARTICLES,
              "scratch rows must cover B2_STATIC_MAX_PARTICLES (include/atooms_b200.h)");
#define B2_STATIC_NQ 8

static __device__ __align__(256) double g_static_scratch[B2_STATIC_ROWS * B2_STATIC_NQ];

// In host code `g_static_scratch` is only a handle for the runtime; the address
// a kernel can use comes from cudaGetSymbolAddress, once per device.
static double *g_static_ptr[64];

template <typename R>
__global__ void __launch_bounds__(RED_THREADS)
k_kinetic_static_partial(const typename Vec4<R>::type *__restrict__ vel, long long n,
                         double *__restrict__ scratch)
{
    typedef typename Vec4<R>::type V4;
    double q[5] = {0, 0, 0, 0, 0};
    long long base = (long long)blockIdx.x * B2_STATIC_TILE;
    for (int it = 0; it < B2_STATIC_TILE / RED_THREADS; it++) {
        long long i = base + (long long)it * RED_THREADS + threadIdx.x;
        if (i < n) {
            V4 v = vel[i];
            double m = (double)v.w, vx = v.x, vy = v.y, vz = v.z;
            q[0] += 0.5 * m * ((vx * vx + vy * vy) + vz * vz);
            q[1] += m * vx; q[2] += m * vy; q[3] += m * vz;
            q[4] += m;
        }
    }
    __shared__ double s_q[RED_THREADS / 32][5];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        double v = warp_sum(q[k]);
        if (lane == 0) s_q[w][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < B2_STATIC_NQ) {
        double v = 0;
        if (threadIdx.x < 5)
            for (int ww = 0; ww < RED_THREADS / 32; ww++) v += s_q[ww][threadIdx.x];
        scratch[(size_t)blockIdx.x * B2_STATIC_NQ + threadIdx.x] = v;    // all 8 slots of the row
    }
}

// rows are summed in block order: the result does not depend on scheduling
__global__ void k_kinetic_static_sum(const double *__restrict__ scratch, int rows, double *__restrict__ out)
{
    int q = threadIdx.x;
    if (q >= B2_STATIC_NQ) return;
    double v = 0;
    for (int b = 0; b < rows; b++) v += scratch[(size_t)b * B2_STATIC_NQ + q];
    out[q] = v;
}

extern "C" int b2_kinetic_static(const void *vel_mass_dev, int64_t n, int precision, void *cuda_stream,
                                 double *out_dev)
{
    if (precision != B2_FP32 && precision != B2_FP64)
        return fail(nullptr, B2_EINVAL, "b2_kinetic_static: precision must be 32 or 64");
    if (n < 0 || n > B2_STATIC_MAX_PARTICLES)
        return fail(nullptr, B2_EINVAL,
                    "b2_kinetic_static: n = %lld is outside 0..%lld (B2_STATIC_MAX_PARTICLES); the fixed scratch "
                    "array cannot hold more partial sums — split the call or use b2_thermo",
                    (long long)n, (long long)B2_STATIC_MAX_PARTICLES);
    if (!out_dev || (n > 0 && !vel_mass_dev))
        return fail(nullptr, B2_EINVAL, "b2_kinetic_static: NULL device pointer");
    // the records are read as float4 / double4: a pointer that is not 16-byte aligned (a tensor view at
    // an odd offset) would fault inside the kernel, and such a fault poisons the CALLER's CUDA context
    if (((uintptr_t)vel_mass_dev & 15u) != 0 || ((uintptr_t)out_dev & 7u) != 0)
        return fail(nullptr, B2_EINVAL, "b2_kinetic_static: vel_mass_dev must be 16-byte aligned and out_dev "
                                        "8-byte aligned");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64)
        return fail(nullptr, e == cudaSuccess ? B2_EINVAL : B2_ENOCUDA, "b2_kinetic_static: no current CUDA device (%s)",
                    cudaGetErrorString(e));
    if (!g_static_ptr[dev]) {
        // set-up, once per device, with that device current
        double *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(nullptr, B2_ECUDA, "b2_kinetic_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                        cudaGetErrorString(e));
        g_static_ptr[dev] = p;
    }
    double *scratch = g_static_ptr[dev];
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    int rows = (int)((n + B2_STATIC_TILE - 1) / B2_STATIC_TILE);
    if (rows > 0) {
        if (precision == B2_FP32)
            k_kinetic_static_partial<float><<<rows, RED_THREADS, 0, stream>>>((const float4 *)vel_mass_dev, n, scratch);
        else
            k_kinetic_static_partial<double><<<rows, RED_THREADS, 0, stream>>>((const double4 *)vel_mass_dev, n, scratch);
    }
    k_kinetic_static_sum<<<1, 32, 0, stream>>>(scratch, rows, out_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return fail(nullptr, B2_ECUDA, "b2_kinetic_static: launch failed: %s", cudaGetErrorString(e));
    return B2_OK;
}

// ---------------------------------------------------------------------------
// NCCL fast path of the slab decomposition: the whole per-step sequence
//   step kernel -> all-reduce(max) of one float -> commit -> pack -> ring
//   send/recv -> unpack
// is enqueued from C on the context's stream, `count` steps per call, so the
// host pays microseconds per step instead of a Python round trip per collective.
// NCCL is resolved at run time (dlopen): the library has no link-time
// dependency on it, and inside a torch process it picks up the NCCL torch loaded.
// ---------------------------------------------------------------------------
static int load_nccl(b2_ctx *c)
{
    if (g_nccl.handle) return B2_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(c, B2_EUNSUPPORTED, "NCCL not available: %s", dlerror());
#define B2_SYM(name)                                                                   \
    *(void **)(&g_nccl.name) = dlsym(h, "nccl" #name);                                 \
    if (!g_nccl.name) return fail(c, B2_EUNSUPPORTED, "NCCL symbol nccl" #name " not found");
    B2_SYM(GetUniqueId) B2_SYM(CommInitRank) B2_SYM(CommDestroy) B2_SYM(AllReduce) B2_SYM(Send) B2_SYM(Recv)
    B2_SYM(GroupStart) B2_SYM(GroupEnd) B2_SYM(GetErrorString)
#undef B2_SYM
    g_nccl.handle = h;
    return B2_OK;
}

#define NC(call)                                                                          \
    do {                                                                                  \
        ncclResult_t r_ = (call);                                                         \
        if (r_ != ncclSuccess)                                                            \
            return fail(c, B2_ECUDA, "%s failed: %s", #call, g_nccl.GetErrorString(r_));  \
    } while (0)

template <typename R> struct ImplDD {
    typedef typename Vec4<R>::type V4;
    typedef typename Pos4<R>::type P4;
    static int run(b2_ctx *c, int64_t count)
    {
        const int n = (int)c->n;
        // halo buffers: [out_l | out_r | in_l | in_r], each at least one record so that no message is empty
        const size_t sl = std::max(c->n_send_l, 1), sr = std::max(c->n_send_r, 1);
        const size_t gl = std::max(c->n_ghost_l, 1), gr = std::max(c->n_ghost_r, 1);
        const size_t need = 3 * (sl + sr + gl + gr);
        if (need > c->halo_cap) {
            if (c->halo_buf) cudaFree(c->halo_buf);
            c->halo_buf = nullptr;
            CU(cudaMalloc((void **)&c->halo_buf, (need + need / 4) * sizeof(double)));
            CU(cudaMemsetAsync(c->halo_buf, 0, (need + need / 4) * sizeof(double), c->stream));
            c->halo_cap = need + need / 4;
        }
        if (!c->dd_vmax) {
            CU(cudaMalloc((void **)&c->dd_vmax, 2 * sizeof(float)));
            CU(cudaMemsetAsync(c->dd_vmax, 0, 2 * sizeof(float), c->stream));
        }
        double *out_l = c->halo_buf, *out_r = out_l + 3 * sl, *in_l = out_r + 3 * sr, *in_r = in_l + 3 * gl;
        const int n_own = n - c->n_ghost_l - c->n_ghost_r;
        for (int64_t k = 0; k < count; k++) {
            int rc = Impl<R>::dd_step(c, c->dd_vmax);
            if (rc) return rc;
            NC(g_nccl.AllReduce(c->dd_vmax, c->dd_vmax, 1, ncclFloat32, ncclMax, c->comm, c->stream));
            LAUNCH(k_dd_commit, 1, 1, c->ctl, c->dd_vmax, c->dd_dt);
            if ((rc = Impl<R>::dd_pack(c, c->send_l, c->n_send_l, out_l))) return rc;
            if ((rc = Impl<R>::dd_pack(c, c->send_r, c->n_send_r, out_r))) return rc;
            // posting order [to-left, to-right] / [from-right, from-left]: unambiguous on a 2-rank ring
            NC(g_nccl.GroupStart());
            NC(g_nccl.Send(out_l, 3 * sl, ncclFloat64, c->left, c->comm, c->stream));
            NC(g_nccl.Send(out_r, 3 * sr, ncclFloat64, c->right, c->comm, c->stream));
            NC(g_nccl.Recv(in_r, 3 * gr, ncclFloat64, c->right, c->comm, c->stream));
            NC(g_nccl.Recv(in_l, 3 * gl, ncclFloat64, c->left, c->comm, c->stream));
            NC(g_nccl.GroupEnd());
            if ((rc = Impl<R>::dd_unpack(c, n_own, c->n_ghost_l, in_l))) return rc;
            if ((rc = Impl<R>::dd_unpack(c, n_own + c->n_ghost_l, c->n_ghost_r, in_r))) return rc;
        }
        return B2_OK;
    }
};

extern "C" {

int b2_nccl_unique_id(char out[128])
{
    b2_ctx *c = nullptr;
    int rc = load_nccl(nullptr);
    if (rc) return rc;
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    NC(g_nccl.GetUniqueId(&id));
    memcpy(out, &id, 128);
    return B2_OK;
}

int b2_nccl_init(b2_ctx *c, const char id[128], int rank, int nranks)
{
    if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks) return fail(c, B2_EINVAL, "b2_nccl_init: bad argument");
    int rc = load_nccl(c);
    if (rc) return rc;
    if ((rc = bind_device(c))) return rc;
    if (c->comm) { g_nccl.CommDestroy(c->comm); c->comm = nullptr; }
    ncclUniqueId uid;
    memcpy(&uid, id, 128);
    NC(g_nccl.CommInitRank(&c->comm, nranks, uid, rank));
    c->nranks = nranks;
    c->myrank = rank;
    c->left = (rank + nranks - 1) % nranks;
    c->right = (rank + 1) % nranks;
    return B2_OK;
}

int b2_nccl_set_halo(b2_ctx *c, const int32_t *send_l_dev, int64_t n_send_l, const int32_t *send_r_dev,
                     int64_t n_send_r, int64_t n_ghost_l, int64_t n_ghost_r)
{
    NEED_PARTICLES("b2_nccl_set_halo");
    if (n_send_l < 0 || n_send_r < 0 || n_ghost_l < 0 || n_ghost_r < 0 || n_ghost_l + n_ghost_r > c->n)
        return fail(c, B2_EINVAL, "b2_nccl_set_halo: bad counts");
    c->send_epoch++;
    c->send_l = send_l_dev; c->send_r = send_r_dev;
    c->n_send_l = (int)n_send_l; c->n_send_r = (int)n_send_r;
    c->n_ghost_l = (int)n_ghost_l; c->n_ghost_r = (int)n_ghost_r;
    return B2_OK;
}

int b2_nccl_run(b2_ctx *c, int64_t count)
{
    NEED_PARTICLES("b2_nccl_run");
    if (!c->comm) return fail(c, B2_ESTATE, "b2_nccl_run: call b2_nccl_init first");
    if (!(c->dd_dt > 0.0)) return fail(c, B2_ESTATE, "b2_nccl_run: call b2_dd_begin first");
    if (count < 0) return fail(c, B2_EINVAL, "b2_nccl_run: count >= 0 required");
    return c->precision == B2_FP32 ? ImplDD<float>::run(c, count) : ImplDD<double>::run(c, count);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Replica ensembles (b2_batch_*): one block per replica, see b2_batch.cuh
// ---------------------------------------------------------------------------
struct b2_batch {
    int device = 0, precision = 64;
    bool have_model = false;
    double side[3] = {0, 0, 0};
    PairConst pc;
    uint64_t const_serial = 0;
    int64_t nrep = 0;
    int n = 0;
    void *pos = nullptr, *frc = nullptr, *vel = nullptr;   // vel: (vx, vy, vz, mass)
    int *img = nullptr;
    double *stage = nullptr;
    int *stage_i = nullptr;
    long long *d_steps = nullptr;
    int *d_conv = nullptr;
    double *d_out = nullptr;     // energy | fn2 | virial
    int64_t launches = 0;
};

template <typename R> struct BatchImpl {
    typedef typename Vec4<R>::type V4;
    typedef typename Pos4<R>::type P4;

    static Box<R> box_of(const b2_batch *b)
    {
        b2_ctx tmp;
        for (int d = 0; d < 3; d++) tmp.side[d] = b->side[d];
        return make_box<R>(&tmp);
    }

    static int upload(b2_batch *b, int64_t nrep, int n, const double *pos, const int32_t *species, const double *mass)
    {
        b2_ctx *c = nullptr;
        const size_t tot = (size_t)nrep * n;
        void *old[] = {b->pos, b->frc, b->vel, b->img, b->stage, b->stage_i, b->d_steps, b->d_conv, b->d_out};
        for (void *p : old) if (p) cudaFree(p);
        b->pos = b->frc = b->vel = nullptr; b->img = nullptr; b->stage = nullptr; b->stage_i = nullptr;
        b->d_steps = nullptr; b->d_conv = nullptr; b->d_out = nullptr;
        CU(cudaMalloc(&b->pos, std::max<size_t>(tot, 1) * sizeof(P4)));
        CU(cudaMalloc(&b->frc, std::max<size_t>(tot, 1) * sizeof(V4)));
        CU(cudaMalloc(&b->vel, std::max<size_t>(tot, 1) * sizeof(V4)));
        CU(cudaMalloc((void **)&b->img, std::max<size_t>(tot, 1) * 3 * sizeof(int)));
        CU(cudaMalloc((void **)&b->stage, std::max<size_t>(tot, 1) * 4 * sizeof(double)));
        CU(cudaMalloc((void **)&b->stage_i, std::max<size_t>(tot, 1) * sizeof(int)));
        CU(cudaMalloc((void **)&b->d_steps, std::max<int64_t>(nrep, 1) * sizeof(long long)));
        CU(cudaMalloc((void **)&b->d_conv, std::max<int64_t>(nrep, 1) * sizeof(int)));
        CU(cudaMalloc((void **)&b->d_out, std::max<int64_t>(nrep, 1) * 3 * sizeof(double)));
        b->nrep = nrep;
        b->n = n;
        if (tot == 0) return B2_OK;
        CU(cudaMemcpy(b->stage, pos, tot * 3 * sizeof(double), cudaMemcpyHostToDevice));
        if (mass) CU(cudaMemcpy(b->stage + 3 * tot, mass, tot * sizeof(double), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(b->stage_i, species, tot * sizeof(int), cudaMemcpyHostToDevice));
        k_batch_ingest<R><<<cdiv(tot, 256), 256>>>(b->stage, mass ? b->stage + 3 * tot : nullptr, b->stage_i, (long long)tot,
                                                  box_of(b), (P4 *)b->pos, (V4 *)b->vel, b->img);
        b->launches++;
        CU(cudaDeviceSynchronize());
        return B2_OK;
    }

    static int minimize(b2_batch *b, int method, int64_t max_steps, double dt, double tol, int64_t *steps_done,
                        int *converged, double *energy, double *fn2, double *virial)
    {
        b2_ctx *c = nullptr;
        if (b->nrep == 0) return B2_OK;
        CU(table_acquire(b->device, b->const_serial, &b->pc, (cudaStream_t)0));
        struct Release {
            b2_batch *b;
            ~Release() { table_release(b->device, b->const_serial, (cudaStream_t)0); }
        } release_on_return{b};
        BatchArgs<R> a;
        a.pos = (P4 *)b->pos;
        a.vel = (const V4 *)b->vel;
        a.n = b->n;
        a.nrep = (int)b->nrep;
        a.box = box_of(b);
        a.method = method == B2_MIN_FIRE ? MIN_MODE_FIRE : MIN_MODE_SD;
        a.max_steps = max_steps;
        a.dt = dt;
        a.fn2_tol = tol;
        a.steps_done = b->d_steps;
        a.converged = b->d_conv;
        a.energy = b->d_out;
        a.fn2 = b->d_out + b->nrep;
        a.virial = b->d_out + 2 * b->nrep;
        a.frc = (V4 *)b->frc;
        const int threads = (b->n + 31) / 32 * 32;
        k_batch_minimize<R><<<(unsigned)b->nrep, threads, (size_t)b->n * sizeof(P4)>>>(a);
        b->launches++;
        CU(cudaGetLastError());
        std::vector<long long> hs((size_t)b->nrep);
        CU(cudaMemcpy(hs.data(), b->d_steps, b->nrep * sizeof(long long), cudaMemcpyDeviceToHost));
        for (int64_t r = 0; r < b->nrep; r++) if (steps_done) steps_done[r] = hs[r];
        if (converged) CU(cudaMemcpy(converged, b->d_conv, b->nrep * sizeof(int), cudaMemcpyDeviceToHost));
        if (energy) CU(cudaMemcpy(energy, b->d_out, b->nrep * sizeof(double), cudaMemcpyDeviceToHost));
        if (fn2) CU(cudaMemcpy(fn2, b->d_out + b->nrep, b->nrep * sizeof(double), cudaMemcpyDeviceToHost));
        if (virial) CU(cudaMemcpy(virial, b->d_out + 2 * b->nrep, b->nrep * sizeof(double), cudaMemcpyDeviceToHost));
        return B2_OK;
    }

    static int set_velocities(b2_batch *b, const double *vel)
    {
        b2_ctx *c = nullptr;
        const size_t tot = (size_t)b->nrep * b->n;
        if (tot == 0) return B2_OK;
        CU(cudaMemcpy(b->stage, vel, tot * 3 * sizeof(double), cudaMemcpyHostToDevice));
        k_batch_set_vel<R><<<cdiv(tot, 256), 256>>>(b->stage, (long long)tot, (V4 *)b->vel);
        b->launches++;
        CU(cudaDeviceSynchronize());
        return B2_OK;
    }

    static int run_nve(b2_batch *b, int64_t steps, double dt, double *kinetic, double *energy, double *virial)
    {
        b2_ctx *c = nullptr;
        if (b->nrep == 0) return B2_OK;
        CU(table_acquire(b->device, b->const_serial, &b->pc, (cudaStream_t)0));
        struct Release {
            b2_batch *b;
            ~Release() { table_release(b->device, b->const_serial, (cudaStream_t)0); }
        } release_on_return{b};
        BatchNveArgs<R> a;
        a.pos = (P4 *)b->pos;
        a.vel = (V4 *)b->vel;
        a.frc = (V4 *)b->frc;
        a.img = b->img;
        a.n = b->n;
        a.nrep = (int)b->nrep;
        a.box = box_of(b);
        a.steps = steps;
        a.dt = dt;
        a.kinetic = b->d_out;
        a.energy = b->d_out + b->nrep;
        a.virial = b->d_out + 2 * b->nrep;
        // few replicas: several threads per particle shorten each thread's serial pair loop
        int T = 1;
        while (T < 8 && b->n * 2 * T <= BATCH_MAX_THREADS && b->nrep * (int64_t)((b->n * T + 31) / 32) < 148 * 16) T *= 2;
        const int threads = (b->n * T + 31) / 32 * 32;
        const size_t smem = (size_t)b->n * sizeof(P4);
        if (T == 8) k_batch_nve<R, 8><<<(unsigned)b->nrep, threads, smem>>>(a);
        else if (T == 4) k_batch_nve<R, 4><<<(unsigned)b->nrep, threads, smem>>>(a);
        else if (T == 2) k_batch_nve<R, 2><<<(unsigned)b->nrep, threads, smem>>>(a);
        else k_batch_nve<R, 1><<<(unsigned)b->nrep, threads, smem>>>(a);
        b->launches++;
        CU(cudaGetLastError());
        if (kinetic) CU(cudaMemcpy(kinetic, b->d_out, b->nrep * sizeof(double), cudaMemcpyDeviceToHost));
        if (energy) CU(cudaMemcpy(energy, b->d_out + b->nrep, b->nrep * sizeof(double), cudaMemcpyDeviceToHost));
        if (virial) CU(cudaMemcpy(virial, b->d_out + 2 * b->nrep, b->nrep * sizeof(double), cudaMemcpyDeviceToHost));
        CU(cudaDeviceSynchronize());
        return B2_OK;
    }

    static int get_state(b2_batch *b, double *vel, int32_t *image)
    {
        b2_ctx *c = nullptr;
        const size_t tot = (size_t)b->nrep * b->n;
        if (tot == 0) return B2_OK;
        if (vel) {
            k_batch_get_vel<R><<<cdiv(tot, 256), 256>>>((const V4 *)b->vel, (long long)tot, b->stage);
            b->launches++;
            CU(cudaMemcpy(vel, b->stage, tot * 3 * sizeof(double), cudaMemcpyDeviceToHost));
        }
        if (image) CU(cudaMemcpy(image, b->img, tot * 3 * sizeof(int), cudaMemcpyDeviceToHost));
        return B2_OK;
    }

    static int download(b2_batch *b, double *pos, double *forces)
    {
        b2_ctx *c = nullptr;
        const size_t tot = (size_t)b->nrep * b->n;
        if (tot == 0) return B2_OK;
        k_batch_export<R><<<cdiv(tot, 256), 256>>>((const P4 *)b->pos, (const V4 *)b->frc, (long long)tot, box_of(b),
                                                  pos ? b->stage : nullptr, nullptr);
        b->launches++;
        // the staging buffer is reused: positions first, forces in a second trip
        if (pos) CU(cudaMemcpy(pos, b->stage, tot * 3 * sizeof(double), cudaMemcpyDeviceToHost));
        if (forces) {
            k_batch_export<R><<<cdiv(tot, 256), 256>>>((const P4 *)b->pos, (const V4 *)b->frc, (long long)tot, box_of(b),
                                                      nullptr, b->stage);
            b->launches++;
            CU(cudaMemcpy(forces, b->stage, tot * 3 * sizeof(double), cudaMemcpyDeviceToHost));
        }
        return B2_OK;
    }
};

extern "C" {

int b2_batch_create(int device, int precision, b2_batch **out)
{
    if (!out) return fail(nullptr, B2_EINVAL, "b2_batch_create: out is NULL");
    *out = nullptr;
    if (precision != B2_FP32 && precision != B2_FP64)
        return fail(nullptr, B2_EINVAL, "b2_batch_create: precision must be 32 or 64");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, B2_ENOCUDA, "b2_batch_create: no usable CUDA device (%s); this library has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= ndev || device >= 64) return fail(nullptr, B2_EINVAL, "b2_batch_create: bad device");
    b2_batch *b = new b2_batch();
    b->device = device;
    b->precision = precision;
    memset(&b->pc, 0, sizeof(PairConst));
    *out = b;
    return B2_OK;
}

int b2_batch_destroy(b2_batch *b)
{
    if (!b) return B2_OK;
    cudaSetDevice(b->device);
    void *ptrs[] = {b->pos, b->frc, b->vel, b->img, b->stage, b->stage_i, b->d_steps, b->d_conv, b->d_out};
    for (void *p : ptrs) if (p) cudaFree(p);
    table_forget(b->device, b->const_serial);
    delete b;
    return B2_OK;
}

int b2_batch_set_model(b2_batch *b, const double side[3], int nsp, const double *eps, const double *sigma,
                       const double *rcut, int shifted)
{
    if (!b || !side || !eps || !sigma || !rcut) return fail(nullptr, B2_EINVAL, "b2_batch_set_model: NULL argument");
    if (nsp < 1 || nsp > B2_MAXSP) return fail(nullptr, B2_EINVAL, "b2_batch_set_model: 1 <= nsp <= %d", B2_MAXSP);
    for (int d = 0; d < 3; d++) {
        if (!(side[d] > 0.0)) return fail(nullptr, B2_EINVAL, "b2_batch_set_model: side must be > 0");
        b->side[d] = side[d];
    }
    memset(&b->pc, 0, sizeof(PairConst));
    b->pc.nsp = nsp;
    for (int a = 0; a < nsp * nsp; a++) {
        if (!(sigma[a] > 0.0) || !(rcut[a] > 0.0)) return fail(nullptr, B2_EINVAL, "b2_batch_set_model: sigma, rcut > 0");
        double s2 = sigma[a] * sigma[a], s6 = s2 * s2 * s2;
        b->pc.c6[a] = 4.0 * eps[a] * s6;
        b->pc.c12[a] = 4.0 * eps[a] * s6 * s6;
        b->pc.rc2[a] = rcut[a] * rcut[a];
        double ir2 = 1.0 / b->pc.rc2[a], ir6 = ir2 * ir2 * ir2;
        b->pc.ushift[a] = shifted ? ir6 * (b->pc.c12[a] * ir6 - b->pc.c6[a]) : 0.0;
    }
    b->const_serial = next_serial();
    b->have_model = true;
    return B2_OK;
}

#define BDISPATCH(fn, ...) (b->precision == B2_FP32 ? BatchImpl<float>::fn(__VA_ARGS__) : BatchImpl<double>::fn(__VA_ARGS__))

int b2_batch_upload(b2_batch *b, int64_t nrep, int n, const double *pos, const int32_t *species, const double *mass)
{
    if (!b || nrep < 0 || n < 0 || (nrep * n > 0 && (!pos || !species))) return fail(nullptr, B2_EINVAL, "b2_batch_upload: bad argument");
    if (!b->have_model) return fail(nullptr, B2_ESTATE, "b2_batch_upload: call b2_batch_set_model first");
    if (n > BATCH_MAX_THREADS) return fail(nullptr, B2_EINVAL, "b2_batch_upload: at most %d particles per replica", BATCH_MAX_THREADS);
    for (int64_t k = 0; k < nrep * n; k++)
        if (species[k] < 0 || species[k] >= b->pc.nsp) return fail(nullptr, B2_EINVAL, "b2_batch_upload: species id outside the pair table");
    cudaSetDevice(b->device);
    return BDISPATCH(upload, b, nrep, n, pos, species, mass);
}

int b2_batch_minimize(b2_batch *b, int method, int64_t max_steps, double dt, double fn2_tol, int64_t *steps_done,
                      int *converged, double *energy, double *fn2, double *virial)
{
    if (!b || !b->pos) return fail(nullptr, B2_ESTATE, "b2_batch_minimize: no replicas resident");
    if (max_steps < 0 || !(dt > 0.0)) return fail(nullptr, B2_EINVAL, "b2_batch_minimize: bad argument");
    if (method != B2_MIN_SD && method != B2_MIN_FIRE) return fail(nullptr, B2_EINVAL, "b2_batch_minimize: unknown method");
    cudaSetDevice(b->device);
    return BDISPATCH(minimize, b, method, max_steps, dt, fn2_tol, steps_done, converged, energy, fn2, virial);
}

int b2_batch_download(b2_batch *b, double *pos, double *forces)
{
    if (!b || !b->pos) return fail(nullptr, B2_ESTATE, "b2_batch_download: no replicas resident");
    cudaSetDevice(b->device);
    return BDISPATCH(download, b, pos, forces);
}

int b2_batch_set_velocities(b2_batch *b, const double *vel)
{
    if (!b || !b->pos || !vel) return fail(nullptr, B2_ESTATE, "b2_batch_set_velocities: no replicas resident / NULL");
    cudaSetDevice(b->device);
    return BDISPATCH(set_velocities, b, vel);
}

int b2_batch_run_nve(b2_batch *b, int64_t steps, double dt, double *kinetic, double *energy, double *virial)
{
    if (!b || !b->pos) return fail(nullptr, B2_ESTATE, "b2_batch_run_nve: no replicas resident");
    if (steps < 0 || !(dt > 0.0)) return fail(nullptr, B2_EINVAL, "b2_batch_run_nve: steps >= 0 and dt > 0 required");
    cudaSetDevice(b->device);
    return BDISPATCH(run_nve, b, steps, dt, kinetic, energy, virial);
}

int b2_batch_get_state(b2_batch *b, double *vel, int32_t *image)
{
    if (!b || !b->pos) return fail(nullptr, B2_ESTATE, "b2_batch_get_state: no replicas resident");
    cudaSetDevice(b->device);
    return BDISPATCH(get_state, b, vel, image);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Peer-memory halo: host side.  Mailbox block layout (identical on every rank):
//   [0, 2*64*W)            max|v|^2 slots: [parity][source rank], 64 bytes each
//   [.., +2*64)            arrival flags: from the left, from the right neighbour
//   [.., +4*cap*24)        inboxes: [from-left | from-right][parity][cap records of 3 doubles]
// ---------------------------------------------------------------------------
static size_t peer_off_flags(int W) { return (size_t)2 * 64 * W; }
static size_t peer_off_inbox(int W) { return peer_off_flags(W) + 2 * 64; }
static size_t peer_inbox(int W, int64_t cap, int side, int parity)
{
    return peer_off_inbox(W) + ((size_t)side * 2 + parity) * (size_t)cap * 24;
}
static size_t peer_step_bytes(int W, int64_t cap) { return (peer_inbox(W, cap, 2, 0) + 255) & ~(size_t)255; }
// the refresh mailboxes (b2_peer_refresh) follow the per-step ones: two arrival flags, then one
// inbox per (side, parity) = [64-byte header][migrants][ghost layer]
static int64_t ddr_mig_cap_of(int64_t cap) { return cap / 4 + 1024; }
static size_t ddr_inbox_bytes(int64_t cap)
{
    return (64 + (size_t)ddr_mig_cap_of(cap) * DDR_MIG_BYTES + (size_t)cap * DDR_GHOST_BYTES + 255) & ~(size_t)255;
}
static size_t ddr_off_flags(int W, int64_t cap) { return peer_step_bytes(W, cap); }
static size_t ddr_inbox(int W, int64_t cap, int side, int parity)
{
    return ddr_off_flags(W, cap) + 256 + ((size_t)side * 2 + parity) * ddr_inbox_bytes(cap);
}
static size_t peer_bytes(int W, int64_t cap) { return ddr_inbox(W, cap, 2, 0); }

template <typename R> struct ImplPeer {
    typedef typename Pos4<R>::type P4;
    static int run(b2_ctx *c, int64_t count)
    {
        const int n = (int)c->n, W = c->nranks;
        const int n_own = n - c->n_ghost_l - c->n_ghost_r;
        if (std::max(c->n_send_l, c->n_send_r) > c->peer_cap || std::max(c->n_ghost_l, c->n_ghost_r) > c->peer_cap)
            return fail(c, B2_EINVAL, "b2_peer_run: halo of %d / %d records exceeds the mailbox capacity %lld",
                        std::max(c->n_send_l, c->n_send_r), std::max(c->n_ghost_l, c->n_ghost_r), (long long)c->peer_cap);
        if (!c->dd_vmax) {
            CU(cudaMalloc((void **)&c->dd_vmax, 2 * sizeof(float)));
            CU(cudaMemsetAsync(c->dd_vmax, 0, 2 * sizeof(float), c->stream));
        }
        char *L = c->peer_map[c->left], *Rt = c->peer_map[c->right];
        // The step kernel's last block settles the step (B2_PEER_COMMIT_KERNEL=1: k_peer_commit as a launch of its
        // own, 0.2364 against 0.2341 ms per step at 2 x 1.2M).  B2_PEER_PUSH_FUSED=1: the step kernel also stores
        // the boundary positions into the neighbours' inboxes (no k_peer_push) — measured SLOWER, 0.2374 ms: the
        // stores of a block at a face are scattered over its warps and have to be fenced system-wide in its tail,
        // where k_peer_push writes them as contiguous rows; kept as a switch, off by default.
        static const bool commit_kernel = getenv("B2_PEER_COMMIT_KERNEL") && atoi(getenv("B2_PEER_COMMIT_KERNEL")) != 0;
        static const bool push_fused = getenv("B2_PEER_PUSH_FUSED") && atoi(getenv("B2_PEER_PUSH_FUSED")) != 0;
        const bool fuse_commit = !commit_kernel && n > 0;
        const bool fuse_push = fuse_commit && push_fused && !c->list_dirty;
        if (fuse_push) {
            // the step kernel stores boundary positions into the neighbours' inboxes itself: slot -> send-list entry
            if (!c->halo_map || c->halo_map_cap < c->ncap) {
                if (c->halo_map) cudaFree(c->halo_map);
                c->halo_map = nullptr;
                CU(cudaMalloc((void **)&c->halo_map, (size_t)c->ncap * sizeof(int)));
                c->halo_map_cap = c->ncap;
                c->halo_map_list = -1;
            }
            if (c->halo_map_list != c->list_epoch || c->halo_map_send != c->send_epoch) {
                CU(cudaMemsetAsync(c->halo_map, 0xff, (size_t)n * sizeof(int), c->stream));
                if (c->n_send_l + c->n_send_r > 0)
                    LAUNCH(k_halo_map, cdiv(c->n_send_l + c->n_send_r, 256), 256, c->inv, c->send_l, c->n_send_l, c->send_r,
                           c->n_send_r, c->halo_map);
                c->halo_map_list = c->list_epoch;
                c->halo_map_send = c->send_epoch;
            }
            for (int par = 0; par < 2; par++) {
                c->halo_out_l[par] = (double *)(L + peer_inbox(W, c->peer_cap, 1, par));
                c->halo_out_r[par] = (double *)(Rt + peer_inbox(W, c->peer_cap, 0, par));
            }
            c->halo_flag_l = (unsigned long long *)(L + peer_off_flags(W) + 64);
            c->halo_flag_r = (unsigned long long *)(Rt + peer_off_flags(W));
        }
        for (int64_t k = 0; k < count; k++) {
            c->peer_fused = fuse_commit;
            c->peer_push_fused = fuse_push;
            int rc = Impl<R>::dd_step(c, c->dd_vmax);
            const bool fused = c->peer_fused;
            c->peer_fused = c->peer_push_fused = false;
            if (rc) return rc;
            if (!fused)
                LAUNCH(k_peer_commit, 1, 32, c->ctl, c->peer_ticket + 2, (const unsigned long long *)c->peer_buf, W, c->dd_dt,
                       c->dd_vmax + 1);
            PeerArgs<R> a;
            a.ctl = c->ctl;
            a.pos[0] = (P4 *)c->pos[0];
            a.pos[1] = (P4 *)c->pos[1];
            a.inv = c->inv;
            a.send_l = c->send_l; a.send_r = c->send_r;
            a.n_send_l = c->n_send_l; a.n_send_r = c->n_send_r;
            a.first_l = n_own; a.n_ghost_l = c->n_ghost_l;
            a.first_r = n_own + c->n_ghost_l; a.n_ghost_r = c->n_ghost_r;
            a.box = make_box<R>(c);
            // what we send to the LEFT neighbour arrives there "from the right" (side 1), and vice versa
            for (int par = 0; par < 2; par++) {
                a.out_l[par] = (double *)(L + peer_inbox(W, c->peer_cap, 1, par));
                a.out_r[par] = (double *)(Rt + peer_inbox(W, c->peer_cap, 0, par));
                a.in_l[par] = (const double *)(c->peer_buf + peer_inbox(W, c->peer_cap, 0, par));
                a.in_r[par] = (const double *)(c->peer_buf + peer_inbox(W, c->peer_cap, 1, par));
            }
            a.flag_l = (unsigned long long *)(L + peer_off_flags(W) + 64);
            a.flag_r = (unsigned long long *)(Rt + peer_off_flags(W));
            a.my_flag_l = (const unsigned long long *)(c->peer_buf + peer_off_flags(W));
            a.my_flag_r = (const unsigned long long *)(c->peer_buf + peer_off_flags(W) + 64);
            a.ticket = c->peer_ticket;
            a.state = c->peer_ticket + 2;
            if (!fuse_push) LAUNCH(k_peer_push<R>, std::max(1, cdiv(c->n_send_l + c->n_send_r, 256)), 256, a);
            LAUNCH(k_peer_pull<R>, std::max(1, cdiv(c->n_ghost_l + c->n_ghost_r, 256)), 256, a);
        }
        return B2_OK;
    }
};

extern "C" {

int b2_peer_export(b2_ctx *c, int64_t halo_capacity, int rank, int nranks, char handle_out[64])
{
    if (!c || !handle_out || nranks < 2 || nranks > 64 || rank < 0 || rank >= nranks || halo_capacity < 1)
        return fail(c, B2_EINVAL, "b2_peer_export: bad argument (2 <= nranks <= 64)");
    int rc = bind_device(c);
    if (rc) return rc;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (c->peer_buf) return fail(c, B2_ESTATE, "b2_peer_export: already exported");
    const size_t bytes = peer_bytes(nranks, halo_capacity);
    CU(cudaMalloc((void **)&c->peer_buf, bytes));
    CU(cudaMemset(c->peer_buf, 0, bytes));
    // [0] ticket of k_peer_push, [2] steps taken (the mail sequence), [3] halo wanted
    CU(cudaMalloc((void **)&c->peer_ticket, 4 * sizeof(unsigned)));
    CU(cudaMemset(c->peer_ticket, 0, 4 * sizeof(unsigned)));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, c->peer_buf));
    memcpy(handle_out, &h, 64);
    c->peer_cap = halo_capacity;
    c->nranks = nranks;
    c->myrank = rank;
    c->left = (rank + nranks - 1) % nranks;
    c->right = (rank + 1) % nranks;
    return B2_OK;
}

int b2_peer_connect(b2_ctx *c, const char *handles)
{
    if (!c || !handles || !c->peer_buf) return fail(c, B2_ESTATE, "b2_peer_connect: call b2_peer_export first");
    int rc = bind_device(c);
    if (rc) return rc;
    const int W = c->nranks;
    std::vector<unsigned long long *> mail((size_t)W);
    for (int r = 0; r < W; r++) {
        if (r == c->myrank) {
            c->peer_map[r] = c->peer_buf;
        } else {
            cudaIpcMemHandle_t h;
            memcpy(&h, handles + 64 * (size_t)r, 64);
            void *p = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess)
                return fail(c, B2_ECUDA, "b2_peer_connect: cannot map rank %d's mailbox (%s): peer access over "
                            "NVLink/PCIe is required, and two ranks must not share a GPU", r, cudaGetErrorString(e));
            c->peer_map[r] = (char *)p;
        }
        // our value lands in rank r's block, slot [parity 0][our rank]
        mail[r] = (unsigned long long *)(c->peer_map[r] + (size_t)64 * c->myrank);
    }
    CU(cudaMalloc((void **)&c->peer_mail, W * sizeof(unsigned long long *)));
    CU(cudaMemcpy(c->peer_mail, mail.data(), W * sizeof(unsigned long long *), cudaMemcpyHostToDevice));
    c->peer_on = true;
    return B2_OK;
}

}  // extern "C"

// ---- slab refresh on the device (kernels in b2_ddr.cuh) ---------------------------------
template <typename R>
__global__ void k_ddr_state(const Ctl *__restrict__ ctl, const typename Pos4<R>::type *pos0,
                            const typename Pos4<R>::type *pos1, const typename Vec4<R>::type *__restrict__ vel,
                            const int *__restrict__ orig, const int *__restrict__ img, const int *__restrict__ cimg,
                            int n, int *__restrict__ img_out, int *__restrict__ spe_out, double *__restrict__ mass_out)
{
    const typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int o = orig[s];
    if (spe_out) spe_out[o] = praw(pos[s]);
    if (mass_out) mass_out[o] = (double)vel[s].w;
    if (img_out)
        for (int d = 0; d < 3; d++) img_out[3 * o + d] = cimg[3 * o + d] + img[3 * o + d];
}

// k_update restricted to the local ids below n_first (the owned particles of a slab)
template <typename R>
__global__ void k_update_first(const Ctl *__restrict__ ctl, const double *__restrict__ spos,
                               const double *__restrict__ svel, int n, int n_first, double cx, double cy, double cz,
                               Box<R> box, typename Pos4<R>::type *pos0, typename Pos4<R>::type *pos1,
                               typename Vec4<R>::type *__restrict__ vel, const int *__restrict__ orig,
                               int *__restrict__ img)
{
    typedef typename Vec4<R>::type V4;
    typename Pos4<R>::type *pos = ctl->cur ? pos1 : pos0;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n || orig[s] >= n_first) return;
    const size_t o = 3 * (size_t)orig[s];
    if (spos) {
        const double px = spos[o] - cx, py = spos[o + 1] - cy, pz = spos[o + 2] - cz;
        const double kx = rint(px / box.Ld[0]), ky = rint(py / box.Ld[1]), kz = rint(pz / box.Ld[2]);
        img[o] += (int)kx; img[o + 1] += (int)ky; img[o + 2] += (int)kz;
        pos[s] = encode_pos(px - kx * box.Ld[0], py - ky * box.Ld[1], pz - kz * box.Ld[2], praw(pos[s]), box);
    }
    if (svel) {
        const V4 v = vel[s];
        vel[s] = make4((R)svel[o], (R)svel[o + 1], (R)svel[o + 2], v.w);
    }
}

static int ddr_alloc(b2_ctx *c)
{
    int rc;
    if (c->ddr_ncap == c->ncap && c->ddr_cls) return B2_OK;
    const int64_t n = c->ncap;
    const int64_t gc = c->peer_cap, mc = ddr_mig_cap_of(c->peer_cap);
    if ((rc = dev_alloc(c, &c->ddr_cls, n))) return rc;
    if ((rc = dev_alloc(c, &c->ddr_blk, 5 * (size_t)std::max<int64_t>(1, cdiv(n, DDR_THREADS))))) return rc;
    if ((rc = dev_alloc(c, &c->ddr_keep, n))) return rc;
    for (int k = 0; k < 2; k++) {
        if ((rc = dev_alloc(c, &c->ddr_go[k], mc))) return rc;
        if ((rc = dev_alloc(c, &c->ddr_send[k], gc + mc))) return rc;
        if ((rc = dev_alloc(c, &c->ddr_gid[k], n))) return rc;
        if ((rc = dev_alloc(c, &c->ddr_cimg[k], 3 * n))) return rc;
    }
    for (int k = 0; k < 4; k++)
        if ((rc = dev_alloc(c, &c->ddr_near[k], gc))) return rc;
    if (!c->ddr_cnt) {
        CU(cudaMalloc((void **)&c->ddr_cnt, 32 * sizeof(unsigned)));
        CU(cudaMemset(c->ddr_cnt, 0, 32 * sizeof(unsigned)));
        CU(cudaMallocHost((void **)&c->ddr_hcnt, 32 * sizeof(unsigned)));
    }
    c->ddr_ncap = n;
    c->ddr_mig_cap = (int)mc;
    c->ddr_ids = false;
    return B2_OK;
}

template <typename R> struct ImplDdr {
    typedef typename Pos4<R>::type P4;
    typedef typename Vec4<R>::type V4;
    static_assert(sizeof(DdrMig<R>) <= DDR_MIG_BYTES && sizeof(DdrMig<R>) % 16 == 0, "migrant record layout");
    static_assert(sizeof(P4) <= DDR_GHOST_BYTES, "ghost record layout");

    static int refresh(b2_ctx *c, double lo, double hi, double margin, int64_t *counts)
    {
        int rc;
        const int n = (int)c->n, W = c->nranks;
        DdrArgs<R> a;
        a.ctl = c->ctl;
        a.pos[0] = (const P4 *)c->pos[0];
        a.pos[1] = (const P4 *)c->pos[1];
        a.vel = (const V4 *)c->vel;
        a.orig = c->orig;
        a.img = c->img;
        a.cimg = c->ddr_cimg[0];
        a.gid = c->ddr_gid[0];
        a.n = n;
        a.ncap = (int)c->ncap;
        a.rank = c->myrank;
        a.nranks = W;
        a.lo = lo; a.hi = hi; a.margin = margin;
        a.box = make_box<R>(c);
        a.cls = c->ddr_cls;
        a.blk = c->ddr_blk;
        a.nblk = std::max(1, cdiv(n, DDR_THREADS));      // a rank may own nothing and must still take part
        a.keep = c->ddr_keep;
        a.go_l = c->ddr_go[0]; a.go_r = c->ddr_go[1];
        a.near_l_slot = c->ddr_near[0]; a.near_l_new = c->ddr_near[1];
        a.near_r_slot = c->ddr_near[2]; a.near_r_new = c->ddr_near[3];
        a.mig_cap = c->ddr_mig_cap;
        a.ghost_cap = (int)c->peer_cap;
        a.cnt = c->ddr_cnt;
        char *L = c->peer_map[c->left], *Rt = c->peer_map[c->right];
        // what we send to the LEFT neighbour arrives there "from the right" (side 1), and vice versa
        a.out_l = L + ddr_inbox(W, c->peer_cap, 1, 0);
        a.out_r = Rt + ddr_inbox(W, c->peer_cap, 0, 0);
        a.inbox_bytes = ddr_inbox_bytes(c->peer_cap);
        a.flag_l = (unsigned long long *)(L + ddr_off_flags(W, c->peer_cap) + 64);
        a.flag_r = (unsigned long long *)(Rt + ddr_off_flags(W, c->peer_cap));
        a.in_l = c->peer_buf + ddr_inbox(W, c->peer_cap, 0, 0);
        a.in_r = c->peer_buf + ddr_inbox(W, c->peer_cap, 1, 0);
        a.my_flag_l = (const unsigned long long *)(c->peer_buf + ddr_off_flags(W, c->peer_cap));
        a.my_flag_r = (const unsigned long long *)(c->peer_buf + ddr_off_flags(W, c->peer_cap) + 64);
        a.npos = (P4 *)c->tpos;
        a.nvel = (V4 *)c->tvel;
        a.ngid = c->ddr_gid[1];
        a.ncimg = c->ddr_cimg[1];
        a.nsend_l = c->ddr_send[0];
        a.nsend_r = c->ddr_send[1];

        CU(cudaMemsetAsync(c->ddr_cnt, 0, 6 * sizeof(unsigned), c->stream));        // counts and error bits
        LAUNCH(k_ddr_classify<R>, a.nblk, DDR_THREADS, a);
        LAUNCH(k_ddr_scan, 1, 1024, a.blk, a.nblk, a.cnt);
        LAUNCH(k_ddr_scatter<R>, a.nblk, DDR_THREADS, a);
        LAUNCH(k_ddr_pack<R>, cdiv(2 * a.mig_cap + 2 * a.ghost_cap, DDR_THREADS), DDR_THREADS, a);
        LAUNCH(k_ddr_assemble<R>, cdiv((int)c->ncap, DDR_THREADS), DDR_THREADS, a);
        CU(cudaMemcpyAsync(c->ddr_hcnt, c->ddr_cnt, 32 * sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaGetLastError());
        const unsigned *h = c->ddr_hcnt;
        if (h[DC_ERR]) {
            const unsigned e = h[DC_ERR];
            return fail(c, e & DE_TIMEOUT ? B2_ECUDA : B2_ESTATE,
                        "b2_peer_refresh failed:%s%s%s%s%s (stay %u, leave %u/%u, near-face %u/%u of capacity %d/%lld, "
                        "room for %lld particles)",
                        e & DE_HOP ? " a particle moved more than one slab;" : "",
                        e & DE_MIG_CAP ? " more migrants than the mailbox holds;" : "",
                        e & DE_GHOST_CAP ? " ghost layer larger than the mailbox;" : "",
                        e & DE_TIMEOUT ? " a neighbour did not answer within 10 s;" : "",
                        e & DE_NCAP ? " the new domain does not fit the particle arrays;" : "",
                        h[DC_KEEP], h[DC_GO_L], h[DC_GO_R], h[DC_NEAR_L], h[DC_NEAR_R], c->ddr_mig_cap,
                        (long long)c->peer_cap, (long long)c->ncap);
        }
        // adopt the new arrays: what b2_upload does, without the staging copies
        const int64_t nn = (int64_t)h[DC_N_OWN] + h[DC_G_L] + h[DC_G_R];
        std::swap(c->pos[0], c->tpos);
        std::swap(c->vel, c->tvel);
        std::swap(c->ddr_gid[0], c->ddr_gid[1]);
        std::swap(c->ddr_cimg[0], c->ddr_cimg[1]);
        c->n = nn;
        c->list_dirty = true;
        c->forces_valid = c->ev_valid = false;
        c->min_resume = false;
        // owned particles lie in [lo, hi), ghosts within `margin` of it: the windowed list launches that slab only
        if (!c->region_on && W > 1) {
            c->region_on = true;
            c->win_off = false;          // (the first, host-driven upload had to assume the whole box)
            c->cap = 0;
        }
        c->region_lo = lo - margin;
        c->region_hi = hi + margin;
        c->region_off = false;
        if (c->mvel) { cudaFree(c->mvel); c->mvel = nullptr; }
        memset(c->hctl, 0, sizeof(Ctl));
        c->hctl->halt_rebuild = 1u;
        c->hctl->half_skin = 0.5 * c->skin;
        if ((rc = push_ctl(c))) return rc;
        if (nn > 0)
            LAUNCH(k_ddr_finish<R>, cdiv((int)nn, 256), 256, (const P4 *)c->pos[0], (int)nn, make_box<R>(c),
                   c->orig, c->inv, c->img, c->ref);
        c->send_epoch++;
        c->send_l = c->ddr_send[0]; c->send_r = c->ddr_send[1];
        c->n_send_l = (int)h[DC_SEND_L]; c->n_send_r = (int)h[DC_SEND_R];
        c->n_ghost_l = (int)h[DC_G_L]; c->n_ghost_r = (int)h[DC_G_R];
        counts[0] = h[DC_N_OWN]; counts[1] = h[DC_G_L]; counts[2] = h[DC_G_R];
        counts[3] = h[DC_SEND_L]; counts[4] = h[DC_SEND_R];
        counts[5] = h[DC_GO_L] + h[DC_GO_R]; counts[6] = h[DC_FROM_L] + h[DC_FROM_R];
        return B2_OK;
    }

    static int update_first(b2_ctx *c, int64_t n_first, const double *pos, const double *vel)
    {
        const int n = (int)c->n;
        double *spos = c->stage, *svel = c->stage + 3 * (size_t)n;
        // cudaMemcpyDefault: pinned host memory or device memory
        if (pos) CU(cudaMemcpyAsync(spos, pos, 3 * n_first * sizeof(double), cudaMemcpyDefault, c->stream));
        if (vel) CU(cudaMemcpyAsync(svel, vel, 3 * n_first * sizeof(double), cudaMemcpyDefault, c->stream));
        LAUNCH(k_update_first<R>, cdiv(n, 256), 256, c->ctl, pos ? spos : nullptr, vel ? svel : nullptr, n,
               (int)n_first, c->center[0], c->center[1], c->center[2], make_box<R>(c), (P4 *)c->pos[0],
               (P4 *)c->pos[1], (V4 *)c->vel, c->orig, c->img);
        CU(cudaStreamSynchronize(c->stream));
        c->min_resume = false;
        if (pos) {
            c->list_dirty = true;
            c->forces_valid = c->ev_valid = false;
        }
        return B2_OK;
    }

    static int state(b2_ctx *c, int32_t *img, int32_t *spe, double *mass)
    {
        const int n = (int)c->n;
        LAUNCH(k_ddr_state<R>, cdiv(n, 256), 256, c->ctl, (const P4 *)c->pos[0], (const P4 *)c->pos[1],
               (const V4 *)c->vel, c->orig, c->img, c->ddr_cimg[0], n, img, spe, mass);
        return B2_OK;
    }
};

extern "C" {

int b2_peer_set_ids(b2_ctx *c, const int64_t *gid_dev, const int32_t *image_dev)
{
    NEED_SETTLED("b2_peer_set_ids");
    if (!c->peer_on) return fail(c, B2_ESTATE, "b2_peer_set_ids: call b2_peer_export / b2_peer_connect first");
    if (!gid_dev) return fail(c, B2_EINVAL, "b2_peer_set_ids: gid_dev is NULL");
    int rc = ddr_alloc(c);
    if (rc) return rc;
    static_assert(sizeof(long long) == sizeof(int64_t), "64-bit ids");
    CU(cudaMemcpyAsync(c->ddr_gid[0], gid_dev, c->n * sizeof(int64_t), cudaMemcpyDefault, c->stream));
    if (image_dev) CU(cudaMemcpyAsync(c->ddr_cimg[0], image_dev, 3 * c->n * sizeof(int32_t), cudaMemcpyDefault, c->stream));
    else CU(cudaMemsetAsync(c->ddr_cimg[0], 0, 3 * c->n * sizeof(int32_t), c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->ddr_ids = true;
    return B2_OK;
}

int b2_peer_state(b2_ctx *c, int64_t *gid_dev, int32_t *image_dev, int32_t *species_dev, double *mass_dev)
{
    NEED_SETTLED("b2_peer_state");
    if (!c->ddr_ids) return fail(c, B2_ESTATE, "b2_peer_state: no ids resident (b2_peer_set_ids, or a b2_upload since)");
    int rc = c->precision == B2_FP32 ? ImplDdr<float>::state(c, image_dev, species_dev, mass_dev)
                                     : ImplDdr<double>::state(c, image_dev, species_dev, mass_dev);
    if (rc) return rc;
    if (gid_dev) CU(cudaMemcpyAsync(gid_dev, c->ddr_gid[0], c->n * sizeof(int64_t), cudaMemcpyDefault, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return B2_OK;
}

int b2_peer_update(b2_ctx *c, int64_t n_first, const double *pos, const double *vel)
{
    NEED_SETTLED("b2_peer_update");
    if (n_first < 0 || n_first > c->n) return fail(c, B2_EINVAL, "b2_peer_update: n_first out of range");
    if (n_first == 0 || (!pos && !vel)) return B2_OK;
    return c->precision == B2_FP32 ? ImplDdr<float>::update_first(c, n_first, pos, vel)
                                   : ImplDdr<double>::update_first(c, n_first, pos, vel);
}

int b2_peer_refresh(b2_ctx *c, double x_lo, double x_hi, double margin, int64_t counts_out[7])
{
    NEED_PARTICLES("b2_peer_refresh");
    if (!c->peer_on) return fail(c, B2_ESTATE, "b2_peer_refresh: call b2_peer_export / b2_peer_connect first");
    if (!c->ddr_ids) return fail(c, B2_ESTATE, "b2_peer_refresh: call b2_peer_set_ids after the last b2_upload");
    if (!counts_out || !(x_hi > x_lo) || !(margin > 0.0)) return fail(c, B2_EINVAL, "b2_peer_refresh: bad argument");
    if (c->ddr_ncap != c->ncap) return fail(c, B2_ESTATE, "b2_peer_refresh: the particle arrays were re-allocated");
    return c->precision == B2_FP32 ? ImplDdr<float>::refresh(c, x_lo, x_hi, margin, counts_out)
                                   : ImplDdr<double>::refresh(c, x_lo, x_hi, margin, counts_out);
}

int b2_peer_run(b2_ctx *c, int64_t count)
{
    NEED_PARTICLES("b2_peer_run");
    if (!c->peer_on) return fail(c, B2_ESTATE, "b2_peer_run: call b2_peer_export / b2_peer_connect first");
    if (!(c->dd_dt > 0.0)) return fail(c, B2_ESTATE, "b2_peer_run: call b2_dd_begin first");
    if (count < 0) return fail(c, B2_EINVAL, "b2_peer_run: count >= 0 required");
    return c->precision == B2_FP32 ? ImplPeer<float>::run(c, count) : ImplPeer<double>::run(c, count);
}

}  // extern "C"
