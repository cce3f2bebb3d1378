// Shared device-side definitions for the atooms_b200 engine (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#define B2_MAXSP 8
#define B2_WARP 32

// ---------------------------------------------------------------------------
// precision traits: one kernel source, instantiated for float and double
// ---------------------------------------------------------------------------
template <typename R> struct Vec4;
template <> struct Vec4<float>  { typedef float4  type; };
template <> struct Vec4<double> { typedef double4 type; };

__device__ __forceinline__ float4  make4(float x, float y, float z, float w)     { return make_float4(x, y, z, w); }
__device__ __forceinline__ double4 make4(double x, double y, double z, double w) { return make_double4(x, y, z, w); }

__device__ __forceinline__ float  r_rint(float x)  { return rintf(x); }
__device__ __forceinline__ double r_rint(double x) { return rint(x); }
// reciprocal without the range-check / slow-path branch of the IEEE sequences:
// hardware approximation refined by Newton steps (x is a squared distance:
// finite, positive, far from the denormal range)
__device__ __forceinline__ float r_rcp(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return fmaf(r, fmaf(-x, r, 1.0f), r);                       // < 1 ulp
}
__device__ __forceinline__ double r_rcp(double x)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));     // ~20 bits
    r = fma(r, fma(-x, r, 1.0), r);                             // ~40 bits
    r = fma(r, fma(-x, r, 1.0), r);                             // ~80 bits -> rounding limited
    return fma(r, fma(-x, r, 1.0), r);
}

// round-to-nearest, never contracted into FMA: the neighbour predicate is
// compared bit-for-bit with the CPU oracle (particle.py:129-130 arithmetic)
__device__ __forceinline__ float  nc_sub(float a, float b)   { return __fsub_rn(a, b); }
__device__ __forceinline__ float  nc_mul(float a, float b)   { return __fmul_rn(a, b); }
__device__ __forceinline__ float  nc_add(float a, float b)   { return __fadd_rn(a, b); }
__device__ __forceinline__ float  nc_div(float a, float b)   { return __fdiv_rn(a, b); }
__device__ __forceinline__ double nc_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double nc_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double nc_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double nc_div(double a, double b) { return __ddiv_rn(a, b); }

// streaming (read-once) loads bypass L1 allocation so the neighbour-index
// stream does not evict the position working set
__device__ __forceinline__ int ld_stream(const int *p)
{
    int v;
    asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// ---------------------------------------------------------------------------
// model constants (per device): pair table and box
// ---------------------------------------------------------------------------
struct PairConst {
    int nsp;
    int pad_[3];
    // [ti * nsp + tj]
    double c12[B2_MAXSP * B2_MAXSP], c6[B2_MAXSP * B2_MAXSP];
    double rc2[B2_MAXSP * B2_MAXSP], ushift[B2_MAXSP * B2_MAXSP];
    double rl2_d[B2_MAXSP * B2_MAXSP];   // (rcut + skin)^2 evaluated in double
    float  rl2_f[B2_MAXSP * B2_MAXSP];   // ((float)(rcut + skin))^2 evaluated in float
    // force coefficients as the kernels use them: 12 c12, 6 c6, rc^2 in float (the fp64 kernels multiply themselves)
    float  f12[B2_MAXSP * B2_MAXSP], f6[B2_MAXSP * B2_MAXSP], frc2[B2_MAXSP * B2_MAXSP];
    // tabulated pair potentials (b2_set_pair_tabulated): per pair tab_np samples, uniform in r^2 over
    // [0, rc^2]; u = energy, w = -r du/dr = r.F; linear interpolation in r^2.  Tables live in global memory.
    int tab_np, tab_pad_;
    const double *tab_u, *tab_w;         // [nsp * nsp][tab_np]
    const float *tab_wf;                 // w in float, for the fp32 force loops
    double tab_scale[B2_MAXSP * B2_MAXSP];   // (tab_np - 1) / rc^2
};

// single translation unit: the table is defined here, once
__constant__ PairConst g_pair;

// ---------------------------------------------------------------------------
// Position storage.
//   fp64: double4 (x, y, z, species bits), coordinates relative to the cell
//         centre in [-L/2, L/2]; minimum image d - rint(d/L) L.
//   fp32: int4 (x, y, z, species): 32-bit FIXED-POINT fractions of the cell,
//         x_int = round(x / L * 2^32).  Two's-complement wrap-around IS the
//         periodic boundary: (x_i - x_j) mod 2^32 is the minimum-image
//         separation exactly, with a uniform resolution of L / 2^32 — no
//         loss of precision for pairs across the boundary or far from the
//         origin (float coordinates at |x| ~ 50 resolve only 4e-6).
// ---------------------------------------------------------------------------
template <typename R> struct Pos4;
template <> struct Pos4<float>  { typedef int4    type; };
template <> struct Pos4<double> { typedef double4 type; };

template <typename R> struct Box {
    R L[3], invL[3];
    R s[3], is[3];        // fp32: grid spacing L / 2^32 and its inverse; fp64: 1
    double Ld[3], sd[3];
};

// The species lane also carries the ghost flag (B2_GHOST in the C ABI): a ghost
// is a read-only copy of a particle owned by a neighbouring spatial domain — it
// acts on its neighbours but is neither integrated nor counted in any total.
#define B2_GHOST_BIT 256
__device__ __forceinline__ int praw(const int4 &p)     { return p.w; }
__device__ __forceinline__ int praw(const double4 &p)  { return (int)__double_as_longlong(p.w); }
__device__ __forceinline__ int ptype(const int4 &p)    { return p.w & 0xff; }
__device__ __forceinline__ int ptype(const double4 &p) { return (int)__double_as_longlong(p.w) & 0xff; }
__device__ __forceinline__ bool is_ghost(const int4 &p)    { return (p.w & B2_GHOST_BIT) != 0; }
__device__ __forceinline__ bool is_ghost(const double4 &p) { return ((int)__double_as_longlong(p.w) & B2_GHOST_BIT) != 0; }

__device__ __forceinline__ int wrap_sub(int a, int b) { return (int)((unsigned)a - (unsigned)b); }

// minimum-image separation a - b
__device__ __forceinline__ void delta(const int4 &a, const int4 &b, const Box<float> &box,
                                      float &dx, float &dy, float &dz)
{
    dx = (float)wrap_sub(a.x, b.x) * box.s[0];
    dy = (float)wrap_sub(a.y, b.y) * box.s[1];
    dz = (float)wrap_sub(a.z, b.z) * box.s[2];
}
__device__ __forceinline__ void delta(const double4 &a, const double4 &b, const Box<double> &box,
                                      double &dx, double &dy, double &dz)
{
    dx = a.x - b.x; dy = a.y - b.y; dz = a.z - b.z;
    dx -= box.L[0] * rint(dx * box.invL[0]);
    dy -= box.L[1] * rint(dy * box.invL[1]);
    dz -= box.L[2] * rint(dz * box.invL[2]);
}

// the same separation in double (observables are always summed in fp64)
__device__ __forceinline__ void delta_d(const int4 &a, const int4 &b, const Box<float> &box,
                                        double &dx, double &dy, double &dz)
{
    dx = (double)wrap_sub(a.x, b.x) * box.sd[0];
    dy = (double)wrap_sub(a.y, b.y) * box.sd[1];
    dz = (double)wrap_sub(a.z, b.z) * box.sd[2];
}
__device__ __forceinline__ void delta_d(const double4 &a, const double4 &b, const Box<double> &box,
                                        double &dx, double &dy, double &dz)
{
    delta(a, b, box, dx, dy, dz);
}

// canonical squared distance of the neighbour predicate: correctly rounded,
// never contracted, so that it can be compared bit-for-bit with the oracle
__device__ __forceinline__ float r2_canon(const int4 &a, const int4 &b, const Box<float> &box)
{
    float dx = __fmul_rn(__int2float_rn(wrap_sub(a.x, b.x)), box.s[0]);
    float dy = __fmul_rn(__int2float_rn(wrap_sub(a.y, b.y)), box.s[1]);
    float dz = __fmul_rn(__int2float_rn(wrap_sub(a.z, b.z)), box.s[2]);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
__device__ __forceinline__ double r2_canon(const double4 &a, const double4 &b, const Box<double> &box)
{
    // atooms/system/particle.py:129-130: d - rint(d / L) * L
    double dx = __dsub_rn(a.x, b.x), dy = __dsub_rn(a.y, b.y), dz = __dsub_rn(a.z, b.z);
    dx = __dsub_rn(dx, __dmul_rn(rint(__ddiv_rn(dx, box.L[0])), box.L[0]));
    dy = __dsub_rn(dy, __dmul_rn(rint(__ddiv_rn(dy, box.L[1])), box.L[1]));
    dz = __dsub_rn(dz, __dmul_rn(rint(__ddiv_rn(dz, box.L[2])), box.L[2]));
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// displace by u, fold into the cell (particle.py:90-99), report box crossings
__device__ __forceinline__ int4 advance(const int4 &p, float ux, float uy, float uz,
                                        const Box<float> &box, int &kx, int &ky, int &kz)
{
    int ix = __float2int_rn(ux * box.is[0]);
    int iy = __float2int_rn(uy * box.is[1]);
    int iz = __float2int_rn(uz * box.is[2]);
    int nx = (int)((unsigned)p.x + (unsigned)ix);
    int ny = (int)((unsigned)p.y + (unsigned)iy);
    int nz = (int)((unsigned)p.z + (unsigned)iz);
    kx = (ix > 0 && nx < p.x) ? 1 : ((ix < 0 && nx > p.x) ? -1 : 0);
    ky = (iy > 0 && ny < p.y) ? 1 : ((iy < 0 && ny > p.y) ? -1 : 0);
    kz = (iz > 0 && nz < p.z) ? 1 : ((iz < 0 && nz > p.z) ? -1 : 0);
    return make_int4(nx, ny, nz, p.w);
}
__device__ __forceinline__ double4 advance(const double4 &p, double ux, double uy, double uz,
                                           const Box<double> &box, int &kx, int &ky, int &kz)
{
    double x = p.x + ux, y = p.y + uy, z = p.z + uz;
    double fx = rint(x * box.invL[0]), fy = rint(y * box.invL[1]), fz = rint(z * box.invL[2]);
    kx = (int)fx; ky = (int)fy; kz = (int)fz;
    if (kx | ky | kz) { x -= fx * box.L[0]; y -= fy * box.L[1]; z -= fz * box.L[2]; }
    return make_double4(x, y, z, p.w);
}

// coordinates relative to the cell centre, in double
__device__ __forceinline__ void to_double(const int4 &p, const Box<float> &box, double &x, double &y, double &z)
{
    x = (double)p.x * box.sd[0]; y = (double)p.y * box.sd[1]; z = (double)p.z * box.sd[2];
}
__device__ __forceinline__ void to_double(const double4 &p, const Box<double> &, double &x, double &y, double &z)
{
    x = p.x; y = p.y; z = p.z;
}

// encode folded coordinates (relative to the centre, |x| <= L/2)
__device__ __forceinline__ int enc32(double x, double L)
{
    double f = x / L;
    f -= rint(f);
    long long k = llrint(f * 4294967296.0);          // in [-2^31, 2^31]
    return (int)(unsigned)(k & 0xffffffffLL);         // +2^31 wraps to -2^31: the same point
}
__device__ __forceinline__ int4 encode_pos(double x, double y, double z, int t, const Box<float> &box)
{
    return make_int4(enc32(x, box.Ld[0]), enc32(y, box.Ld[1]), enc32(z, box.Ld[2]), t);
}
__device__ __forceinline__ double4 encode_pos(double x, double y, double z, int t, const Box<double> &)
{
    return make_double4(x, y, z, __longlong_as_double((long long)t));
}

// Device-resident control block: every data-dependent decision of the inner
// loop (list refresh, step budget, minimiser state) lives here so that the
// host can enqueue kernels speculatively without synchronising.
struct Ctl {
    unsigned int halt_rebuild;   // list invalid: step kernels are no-ops until rebuilt
    unsigned int overflow;       // last list build ran out of slots
    unsigned int done;           // step budget reached / minimiser converged
    unsigned int cur;            // which of the two position buffers is current
    unsigned int vmax2_bits;     // max |v|^2 of the running step (float bits, atomicMax)
    unsigned int blocks_done;    // ticket for "last block finishes the step"
    unsigned int max_count;      // largest neighbour count seen by the last build
    unsigned int converged;
    unsigned int peer_error;     // a peer-memory wait timed out (1) or lost step (2)
    unsigned int win_max;        // windowed list: records of the largest candidate window seen by the last build (+1 dummy)
    long long steps_done, steps_target, rebuilds;
    double disp_bound, half_skin;
    // minimiser
    double dt_now, dt_max, alpha, fn2_tol, fn2, step_len;
    int npos, min_mode;
    int has_prev, pad1_;         // SD: the previous gradient is available (survives a resumed call)
    long long iter;
    // thermostat (b2_run_nvt): v' = th_a v + th_b dt F/m, coefficients renewed every step by the
    // last block from the kinetic energy of the half-step velocities just written
    int th_kind;                 // 0 Berendsen, 1 Nose-Hoover (leap-frog form)
    unsigned int region_miss;    // windowed list: a particle lies outside the launched super-blocks (Grid::sb0, sbn)
    double th_T0, th_tau, th_dt, th_xi, th_a, th_b, th_ndof, th_T;
    double red[32];              // reduction results
};

// quantities of the thermo reduction (index into Ctl::red)
enum { Q_KIN = 0, Q_POT, Q_VIR, Q_FF, Q_PX, Q_PY, Q_PZ, Q_MASS, Q_MSD, Q_S0, Q_THERMO = Q_S0 + 9 };

__device__ __forceinline__ uint32_t spread3(uint32_t x)
{
    x &= 0x3ffu;
    x = (x | (x << 16)) & 0x030000ffu;
    x = (x | (x << 8)) & 0x0300f00fu;
    x = (x | (x << 4)) & 0x030c30c3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}
__device__ __forceinline__ uint32_t morton3(int cx, int cy, int cz)
{
    return spread3((uint32_t)cx) | (spread3((uint32_t)cy) << 1) | (spread3((uint32_t)cz) << 2);
}

struct Grid {
    int nc[3];
    int cellbits;      // Morton bits per dimension (what y and z need; x may have more, see grid_key)
    int keybits;       // bits of a cell key; the cell table has 1 << keybits entries
    int allpairs;      // too few cells along some axis: O(N^2) list build
    int m;             // cells are (r_list,max / m) wide; the stencil reaches m cells (1 or 2)
    float w[3];        // cell widths
    float reach2;      // (r_list,max)^2, padded: cells farther than this cannot hold a neighbour
    // windowed list (b2_win.cuh): an even number of cells per axis, in super-blocks of 4x4x4 cells
    int win;           // 1: list entries are 16-bit offsets into the super-block's candidate window
    int nsb[3];        // super-blocks per axis (ceil(nc / 4), nc even)
    int sb0[3], sbn[3]; // the super-blocks that are launched and tabulated: sbn[d] of them from sb0[d] on, periodically
                       // (all of them, unless the particles are known to fill only a slab of the box)
    int wcap;          // records a window may hold, the dummy included (shared memory of the step kernel)
};

// Key of cell (x, y, z) at `level` (0: cells, 1: blocks of 2x2x2 cells, 2: super-blocks of 4x4x4): the low
// (cellbits - level) bits of the three coordinates interleaved, Morton order.  A box much longer along x than
// along y and z (the box of a slab decomposition) has more cells along x than 2^cellbits: the extra high bits of
// x go on top of the interleaved part, so the key space stays as small as the grid and blocks, super-blocks
// and their runs of slots are what they are in a cube.
__device__ __forceinline__ uint32_t grid_key(const Grid &g, int level, int x, int y, int z)
{
    const int b = g.cellbits - level;
    return (((uint32_t)x >> b) << (3 * b)) | morton3(x & ((1 << b) - 1), y, z);
}
// the x part of a cell key alone (y and z contribute spread3(y) << 1 | spread3(z) << 2)
__device__ __forceinline__ uint32_t grid_key_x(const Grid &g, int x)
{
    return (((uint32_t)x >> g.cellbits) << (3 * g.cellbits)) | spread3((uint32_t)x & ((1u << g.cellbits) - 1u));
}

__device__ __forceinline__ void cell_of(const Box<float> &, const Grid &g, const int4 &p, int &cx, int &cy, int &cz)
{
    // exact integer binning of the fixed-point coordinate
    cx = (int)__umulhi((unsigned)p.x + 0x80000000u, (unsigned)g.nc[0]);
    cy = (int)__umulhi((unsigned)p.y + 0x80000000u, (unsigned)g.nc[1]);
    cz = (int)__umulhi((unsigned)p.z + 0x80000000u, (unsigned)g.nc[2]);
}
__device__ __forceinline__ void cell_of(const Box<double> &box, const Grid &g, const double4 &p, int &cx, int &cy, int &cz)
{
    cx = (int)floor((p.x * box.invL[0] + 0.5) * (double)g.nc[0]);
    cy = (int)floor((p.y * box.invL[1] + 0.5) * (double)g.nc[1]);
    cz = (int)floor((p.z * box.invL[2] + 0.5) * (double)g.nc[2]);
    cx = min(max(cx, 0), g.nc[0] - 1);
    cy = min(max(cy, 0), g.nc[1] - 1);
    cz = min(max(cz, 0), g.nc[2] - 1);
}

// cell coordinates plus the position inside the cell as a fraction in [0, 1]
__device__ __forceinline__ void cell_frac(const Box<float> &box, const Grid &g, const int4 &p,
                                          int &cx, int &cy, int &cz, float &fx, float &fy, float &fz)
{
    cell_of(box, g, p, cx, cy, cz);
    const float k = 1.0f / 4294967296.0f;
    fx = __uint2float_rn((unsigned)p.x + 0x80000000u) * k * (float)g.nc[0] - (float)cx;
    fy = __uint2float_rn((unsigned)p.y + 0x80000000u) * k * (float)g.nc[1] - (float)cy;
    fz = __uint2float_rn((unsigned)p.z + 0x80000000u) * k * (float)g.nc[2] - (float)cz;
    fx = fminf(fmaxf(fx, 0.f), 1.f); fy = fminf(fmaxf(fy, 0.f), 1.f); fz = fminf(fmaxf(fz, 0.f), 1.f);
}
__device__ __forceinline__ void cell_frac(const Box<double> &box, const Grid &g, const double4 &p,
                                          int &cx, int &cy, int &cz, float &fx, float &fy, float &fz)
{
    cell_of(box, g, p, cx, cy, cz);
    fx = (float)((p.x * box.invL[0] + 0.5) * (double)g.nc[0] - (double)cx);
    fy = (float)((p.y * box.invL[1] + 0.5) * (double)g.nc[1] - (double)cy);
    fz = (float)((p.z * box.invL[2] + 0.5) * (double)g.nc[2] - (double)cz);
    fx = fminf(fmaxf(fx, 0.f), 1.f); fy = fminf(fmaxf(fy, 0.f), 1.f); fz = fminf(fmaxf(fz, 0.f), 1.f);
}

// distance (in cell widths, never over-estimated) from a point at fraction f of
// its cell to the cell d steps away along one axis
__device__ __forceinline__ float cell_gap(int d, float f)
{
    float g = d == 0 ? 0.f : (d > 0 ? (float)(d - 1) + (1.f - f) : (float)(-d - 1) + f);
    return fmaxf(g - 1e-3f, 0.f);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
