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

// species id travels in the .w lane of a position, bit-cast
__device__ __forceinline__ int   type_of(float w)  { return __float_as_int(w); }
__device__ __forceinline__ int   type_of(double w) { return (int)__double_as_longlong(w); }
__device__ __forceinline__ float  type_as(float,  int t) { return __int_as_float(t); }
__device__ __forceinline__ double type_as(double, int t) { return __longlong_as_double((long long)t); }

__device__ __forceinline__ float  r_rint(float x)  { return rintf(x); }
__device__ __forceinline__ double r_rint(double x) { return rint(x); }
__device__ __forceinline__ float  r_rcp(float x)   { return __frcp_rn(x); }
__device__ __forceinline__ double r_rcp(double x)  { return 1.0 / x; }

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
};

// single translation unit: the table is defined here, once
__constant__ PairConst g_pair;

template <typename R> struct Box {
    R L[3], invL[3];
};

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
    long long steps_done, steps_target, rebuilds;
    double disp_bound, half_skin;
    // minimiser
    double dt_now, dt_max, alpha, fn2_tol, fn2, step_len;
    int npos, min_mode;
    long long iter;
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
    int cellbits;      // Morton bits per dimension; table has 1 << (3 cellbits) entries
    int allpairs;      // fewer than 3 cells along some axis: O(N^2) list build
};

template <typename R>
__device__ __forceinline__ void cell_of(const Box<R> &box, const Grid &g, R x, R y, R z, int &cx, int &cy, int &cz)
{
    // positions are stored relative to the cell centre, in [-L/2, L/2]
    cx = (int)floor((x * box.invL[0] + (R)0.5) * (R)g.nc[0]);
    cy = (int)floor((y * box.invL[1] + (R)0.5) * (R)g.nc[1]);
    cz = (int)floor((z * box.invL[2] + (R)0.5) * (R)g.nc[2]);
    cx = min(max(cx, 0), g.nc[0] - 1);
    cy = min(max(cy, 0), g.nc[1] - 1);
    cz = min(max(cz, 0), g.nc[2] - 1);
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
