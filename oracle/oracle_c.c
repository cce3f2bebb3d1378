/*
 * Oracle-B: plain-C (OpenMP) CPU restatement of the hot path behind the SAME
 * C ABI as the product library (include/atooms_b200.h).
 *
 * TEST INFRASTRUCTURE ONLY — never loaded by the product package.  It is the
 * checker for the CUDA kernels at sizes the O(N^2) NumPy oracle cannot reach,
 * and the timed CPU baseline of bench.py (`cpu_baseline`, `--impl reference`).
 *
 * Parity status: the reference (atooms 3.14.1) has no force / neighbour /
 * integration code (CHANGELOG.md:151-152); this file restates the published
 * algorithms its call sites assume and is pinned, through
 * tests/test_oracle_pins.py, by the known-answer values of the reference's
 * own tests (tests/golden/pins.json).  Beyond those pins: "parity unpinned".
 *
 * Conventions (citations into /root/reference):
 *   outputs, zeroing ........ atooms/system/interaction.py:57-82
 *   virial / pressure ....... atooms/system/system.py:337-348, backends/lammps.py:128-133
 *   minimum image, fold ..... atooms/system/particle.py:90-99,129-130
 *   temperature dof ......... atooms/system/system.py:164-179
 *   NVE = velocity-Verlet ... atooms/backends/lammps.py:236-238
 *   KA / LJ tables .......... tests/test_rumd.py:26-31, tests/test_lammps.py:24-27
 *   stop rule ............... atooms/optimization/core.py:34-36
 *
 * Build: see oracle/Makefile (-O3 -march=x86-64-v3 -fopenmp).  The neighbour
 * predicate is compared bit-for-bit with the CUDA list, so the list build (and
 * nothing else) is compiled without FMA contraction: see the GCC optimize
 * pragma around in_list* / build_list.
 */
#include "../include/atooms_b200.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXSP 8
#define TYPE(c, i) ((c)->spe[i] & 0xff)
#define GHOST(c, i) (((c)->spe[i] & B2_GHOST) != 0)

struct b2_ctx {
    int precision;
    int64_t n;
    int have_box, have_table;
    double side[3], center[3];
    int nsp, shifted;
    double c6[MAXSP * MAXSP], c12[MAXSP * MAXSP], rc2[MAXSP * MAXSP], ushift[MAXSP * MAXSP];
    double rcut[MAXSP * MAXSP], rcut_max;
    /* tabulated potentials (b2_set_pair_tabulated): uniform in r^2 over [0, rc^2], linear interpolation */
    int tab_np;
    double *tab_u, *tab_w, tab_scale[MAXSP * MAXSP];
    double skin;
    double *pos, *vel, *mass, *frc, *pe, *vir, *ref;
    int32_t *spe, *img, *img_ref;
    /* Verlet list (CSR, original numbering, rows ascending) */
    int64_t *row;
    int32_t *nbr;
    int64_t nbr_cap;
    int list_valid, list_mode, forces_valid;
    int64_t ncells;
    int32_t max_count;
    double disp_bound;
    double energy, virial, stress[9];
    int64_t steps_total, rebuilds;
    double last_ms;
    /* resumable minimiser state (see b2_minimize in the header) */
    double *mvel;
    int min_resume, min_method, m_npos, m_has_prev;
    double m_dt, m_alpha, m_dt0;
    double th_xi;   /* Nose-Hoover friction, persists between b2_run_nvt calls */
    /* domain decomposition */
    double dd_dt;
    int64_t dd_steps, dd_target;
    int dd_halt;
    char err[512];
};

static char g_err[512] = "";

static int fail(b2_ctx *c, int code, const char *msg)
{
    snprintf(c ? c->err : g_err, 512, "%s", msg);
    return code;
}

int b2_abi_version(void) { return B2_ABI_VERSION; }
const char *b2_backend_name(void) { return "oracle-c"; }

/* Oracle-only (not part of the product's ABI): the OpenMP team size, for the timed CPU legs of
 * bench.py.  OMP_NUM_THREADS is read once per process, and torchrun pre-sets it to 1. */
int b2_oracle_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    (void)n;
    return 1;
#endif
}
int b2_oracle_threads(void)
{
#ifdef _OPENMP
    int t = 1;
#pragma omp parallel
    {
#pragma omp master
        t = omp_get_num_threads();
    }
    return t;
#else
    return 1;
#endif
}
const char *b2_last_error(const b2_ctx *c) { return c ? c->err : g_err; }

int b2_create(int device, int precision, b2_ctx **out)
{
    (void)device;
    if (!out) return fail(NULL, B2_EINVAL, "b2_create: out is NULL");
    if (precision != B2_FP32 && precision != B2_FP64)
        return fail(NULL, B2_EINVAL, "b2_create: precision must be 32 or 64");
    b2_ctx *c = (b2_ctx *)calloc(1, sizeof(b2_ctx));
    if (!c) return fail(NULL, B2_ENOMEM, "b2_create: out of memory");
    c->precision = precision;
    c->skin = 0.3;
    *out = c;
    return B2_OK;
}

static void free_particles(b2_ctx *c)
{
    free(c->pos); free(c->vel); free(c->mass); free(c->frc); free(c->pe); free(c->vir);
    free(c->ref); free(c->spe); free(c->img); free(c->img_ref);
    free(c->row); free(c->nbr); free(c->mvel);
    c->mvel = NULL; c->min_resume = 0;
    c->pos = c->vel = c->mass = c->frc = c->pe = c->vir = c->ref = NULL;
    c->spe = c->img = c->img_ref = NULL;
    c->row = NULL; c->nbr = NULL; c->nbr_cap = 0;
}

int b2_destroy(b2_ctx *c)
{
    if (!c) return B2_OK;
    free_particles(c);
    free(c->tab_u); free(c->tab_w);
    free(c);
    return B2_OK;
}

int b2_set_stream(b2_ctx *c, void *s) { (void)c; (void)s; return B2_OK; }

int b2_set_box(b2_ctx *c, const double side[3], const double center[3])
{
    if (!c || !side) return fail(c, B2_EINVAL, "b2_set_box: NULL argument");
    for (int d = 0; d < 3; d++)
        if (!(side[d] > 0.0)) return fail(c, B2_EINVAL, "b2_set_box: side must be > 0");
    if (c->have_box && c->pos && c->n > 0) {
        /* resident particles: affine rescaling about the centre (see include/atooms_b200.h) */
        for (int64_t i = 0; i < c->n; i++)
            for (int d = 0; d < 3; d++) {
                double r = side[d] / c->side[d], nc = center ? center[d] : 0.0;
                c->pos[3 * i + d] = nc + (c->pos[3 * i + d] - c->center[d]) * r;
                c->ref[3 * i + d] = nc + (c->ref[3 * i + d] - c->center[d]) * r;
            }
    }
    for (int d = 0; d < 3; d++) {
        c->side[d] = side[d];
        c->center[d] = center ? center[d] : 0.0;
    }
    c->have_box = 1;
    c->list_valid = c->forces_valid = 0;
    return B2_OK;
}

int b2_set_pair_table(b2_ctx *c, int nsp, const double *eps, const double *sigma,
                      const double *rcut, int shifted)
{
    if (!c || !eps || !sigma || !rcut) return fail(c, B2_EINVAL, "b2_set_pair_table: NULL argument");
    if (nsp < 1 || nsp > MAXSP) return fail(c, B2_EINVAL, "b2_set_pair_table: 1 <= nsp <= 8");
    c->tab_np = 0;
    c->nsp = nsp;
    c->shifted = shifted;
    c->rcut_max = 0.0;
    for (int a = 0; a < nsp * nsp; a++) {
        if (!(sigma[a] > 0.0) || !(rcut[a] > 0.0))
            return fail(c, B2_EINVAL, "b2_set_pair_table: sigma and rcut must be > 0");
        /* same expression order as oracle_np.PairTable */
        double s2 = sigma[a] * sigma[a];
        double s6 = s2 * s2 * s2;
        c->c6[a] = 4.0 * eps[a] * s6;
        c->c12[a] = 4.0 * eps[a] * s6 * s6;
        c->rcut[a] = rcut[a];
        c->rc2[a] = rcut[a] * rcut[a];
        double ir2 = 1.0 / c->rc2[a];
        double ir6 = ir2 * ir2 * ir2;
        c->ushift[a] = shifted ? ir6 * (c->c12[a] * ir6 - c->c6[a]) : 0.0;
        if (rcut[a] > c->rcut_max) c->rcut_max = rcut[a];
    }
    c->have_table = 1;
    c->list_valid = c->forces_valid = 0;
    return B2_OK;
}

int b2_set_pair_tabulated(b2_ctx *c, int nsp, int npoints, const double *rcut, const double *u, const double *w)
{
    if (!c || !rcut || !u || !w) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: NULL argument");
    if (nsp < 1 || nsp > MAXSP || npoints < 2) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: bad sizes");
    size_t count = (size_t)nsp * nsp * npoints;
    free(c->tab_u); free(c->tab_w);
    c->tab_u = (double *)malloc(count * sizeof(double));
    c->tab_w = (double *)malloc(count * sizeof(double));
    if (!c->tab_u || !c->tab_w) return fail(c, B2_ENOMEM, "b2_set_pair_tabulated: out of memory");
    memcpy(c->tab_u, u, count * sizeof(double));
    memcpy(c->tab_w, w, count * sizeof(double));
    c->tab_np = npoints;
    c->nsp = nsp;
    c->shifted = 0;
    c->rcut_max = 0.0;
    for (int a = 0; a < nsp * nsp; a++) {
        if (!(rcut[a] > 0.0)) return fail(c, B2_EINVAL, "b2_set_pair_tabulated: rcut must be > 0");
        c->rcut[a] = rcut[a];
        c->rc2[a] = rcut[a] * rcut[a];
        c->tab_scale[a] = (double)(npoints - 1) / c->rc2[a];
        c->c6[a] = c->c12[a] = c->ushift[a] = 0.0;
        if (rcut[a] > c->rcut_max) c->rcut_max = rcut[a];
    }
    c->have_table = 1;
    c->list_valid = c->forces_valid = 0;
    return B2_OK;
}

int b2_set_skin(b2_ctx *c, double skin)
{
    if (!c || !(skin >= 0.0)) return fail(c, B2_EINVAL, "b2_set_skin: skin must be >= 0");
    c->skin = skin;
    c->list_valid = 0;
    return B2_OK;
}

/* particle.py:90-99 */
static inline void fold3(const b2_ctx *c, double *x, int32_t *img)
{
    for (int d = 0; d < 3; d++) {
        double p = x[d] - c->center[d];
        double k = rint(p / c->side[d]);
        x[d] = p - k * c->side[d] + c->center[d];
        if (img) img[d] += (int32_t)k;
    }
}

/* fp32 engine: positions are 32-bit fixed-point fractions of the cell,
 * x_int = round(x / L * 2^32); the wrapped integer difference is the minimum
 * image, scaled by float(L / 2^32) (see b2_common.cuh, r2_canon). */
static inline int32_t fixed32(double x_rel, double L)
{
    double f = x_rel / L;
    f -= rint(f);
    long long k = llrint(f * 4294967296.0);
    return (int32_t)(uint32_t)(k & 0xffffffffLL);
}

int b2_upload(b2_ctx *c, int64_t n, const double *pos, const double *vel,
              const int32_t *species, const double *mass)
{
    if (!c || n < 0 || (n > 0 && (!pos || !species))) return fail(c, B2_EINVAL, "b2_upload: bad argument");
    if (!c->have_box) return fail(c, B2_ESTATE, "b2_upload: call b2_set_box first");
    if (!c->have_table) return fail(c, B2_ESTATE, "b2_upload: call b2_set_pair_table first");
    for (int64_t i = 0; i < n; i++) {
        if (species[i] < 0 || (species[i] & ~B2_GHOST) >= c->nsp)
            return fail(c, B2_EINVAL, "b2_upload: species id outside the pair table");
        if (mass && !(mass[i] > 0.0)) return fail(c, B2_EINVAL, "b2_upload: mass must be > 0");
    }
    free_particles(c);
    c->n = n;
    size_t n3 = (size_t)(n > 0 ? n : 1) * 3;
    c->pos = (double *)malloc(n3 * sizeof(double));
    c->vel = (double *)calloc(n3, sizeof(double));
    c->frc = (double *)calloc(n3, sizeof(double));
    c->ref = (double *)malloc(n3 * sizeof(double));
    c->mass = (double *)malloc(n3 / 3 * sizeof(double));
    c->pe = (double *)calloc(n3 / 3, sizeof(double));
    c->vir = (double *)calloc(n3 / 3, sizeof(double));
    c->spe = (int32_t *)malloc(n3 / 3 * sizeof(int32_t));
    c->img = (int32_t *)calloc(n3, sizeof(int32_t));
    c->img_ref = (int32_t *)calloc(n3, sizeof(int32_t));
    c->row = (int64_t *)calloc((size_t)n + 1, sizeof(int64_t));
    if (!c->pos || !c->vel || !c->frc || !c->ref || !c->mass || !c->pe || !c->vir || !c->spe ||
        !c->img || !c->img_ref || !c->row)
        return fail(c, B2_ENOMEM, "b2_upload: out of memory");
    for (int64_t i = 0; i < n; i++) {
        c->spe[i] = species[i];
        c->mass[i] = mass ? mass[i] : 1.0;
        for (int d = 0; d < 3; d++) {
            c->pos[3 * i + d] = pos[3 * i + d];
            if (vel) c->vel[3 * i + d] = vel[3 * i + d];
        }
        fold3(c, c->pos + 3 * i, NULL);
        if (c->precision == B2_FP32)   /* the fp32 engine holds fixed-point positions */
            for (int d = 0; d < 3; d++)
                c->pos[3 * i + d] = (double)fixed32(c->pos[3 * i + d] - c->center[d], c->side[d]) *
                                    (c->side[d] / 4294967296.0) + c->center[d];
    }
    memcpy(c->ref, c->pos, n3 * sizeof(double));
    c->list_valid = c->forces_valid = 0;
    c->disp_bound = 0.0;
    c->th_xi = 0.0;
    return B2_OK;
}

int b2_update(b2_ctx *c, const double *pos, const double *vel)
{
    if (!c || !c->pos) return fail(c, B2_ESTATE, "b2_update: no particles resident");
    c->min_resume = 0;
    if (pos) {
        for (int64_t i = 0; i < c->n; i++) {
            for (int d = 0; d < 3; d++) c->pos[3 * i + d] = pos[3 * i + d];
            fold3(c, c->pos + 3 * i, c->img + 3 * i);
        }
        c->list_valid = c->forces_valid = 0;
    }
    if (vel) memcpy(c->vel, vel, (size_t)c->n * 3 * sizeof(double));
    return B2_OK;
}

int b2_download(b2_ctx *c, double *pos, double *vel, int32_t *image)
{
    if (!c || !c->pos) return fail(c, B2_ESTATE, "b2_download: no particles resident");
    size_t n3 = (size_t)c->n * 3;
    if (pos) memcpy(pos, c->pos, n3 * sizeof(double));
    if (vel) memcpy(vel, c->vel, n3 * sizeof(double));
    if (image) memcpy(image, c->img, n3 * sizeof(int32_t));
    return B2_OK;
}

int b2_scale_velocities(b2_ctx *c, double factor)
{
    if (!c || !c->pos) return fail(c, B2_ESTATE, "b2_scale_velocities: no particles resident");
    c->min_resume = 0;
    for (int64_t i = 0; i < 3 * c->n; i++) c->vel[i] *= factor;
    return B2_OK;
}

int b2_mark_reference(b2_ctx *c)
{
    if (!c || !c->pos) return fail(c, B2_ESTATE, "b2_mark_reference: no particles resident");
    memcpy(c->ref, c->pos, (size_t)c->n * 3 * sizeof(double));
    memcpy(c->img_ref, c->img, (size_t)c->n * 3 * sizeof(int32_t));
    return B2_OK;
}

/* The list build is the one region that must not be FMA-contracted. */
#pragma GCC push_options
#pragma GCC optimize("fp-contract=off")
/* ------------------------------------------------------------------------
 * Neighbour predicate, evaluated in the context's precision.
 *   d = x_i - x_j ; d -= rint(d / L) * L ; r2 = (dx*dx + dy*dy) + dz*dz
 * ---------------------------------------------------------------------- */
static inline int in_list64(const double *xi, const double *xj, const double *L, double rl2)
{
    double dx = xi[0] - xj[0], dy = xi[1] - xj[1], dz = xi[2] - xj[2];
    dx = dx - rint(dx / L[0]) * L[0];
    dy = dy - rint(dy / L[1]) * L[1];
    dz = dz - rint(dz / L[2]) * L[2];
    double r2 = (dx * dx + dy * dy) + dz * dz;
    return r2 < rl2;
}

static inline int in_list32(const double *xi, const double *xj, const double *ctr,
                            const double *L, float rl2)
{
    float d[3];
    for (int k = 0; k < 3; k++) {
        int32_t a = fixed32(xi[k] - ctr[k], L[k]), b = fixed32(xj[k] - ctr[k], L[k]);
        int32_t w = (int32_t)((uint32_t)a - (uint32_t)b);
        d[k] = (float)w * (float)(L[k] / 4294967296.0);
    }
    float r2 = (d[0] * d[0] + d[1] * d[1]) + d[2] * d[2];
    return r2 < rl2;
}

static inline int in_list(const b2_ctx *c, int64_t i, int64_t j)
{
    int a = TYPE(c, i) * c->nsp + TYPE(c, j);
    if (c->precision == B2_FP64) {
        double rl = c->rcut[a] + c->skin;
        return in_list64(c->pos + 3 * i, c->pos + 3 * j, c->side, rl * rl);
    } else {
        float rl = (float)(c->rcut[a] + c->skin);
        return in_list32(c->pos + 3 * i, c->pos + 3 * j, c->center, c->side, rl * rl);
    }
}

static int cmp_i32(const void *a, const void *b)
{
    int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
    return (x > y) - (x < y);
}

static int build_list(b2_ctx *c)
{
    int64_t n = c->n;
    double rl = c->rcut_max + c->skin;
    int nc[3];
    int cells_ok = 1;
    for (int d = 0; d < 3; d++) {
        nc[d] = (int)floor(c->side[d] / rl);
        if (nc[d] < 3) cells_ok = 0;
        if (nc[d] > 1024) nc[d] = 1024;
    }
    /* Minimum-image semantics throughout (particle.py:129-130): exact when every
     * side >= 2 r_cut; for smaller boxes (data/lj_fcc_N108.xyz, L = 4.76 < 5) only
     * the nearest image of each pair interacts, as in the reference pin. */
    int64_t ncell = cells_ok ? (int64_t)nc[0] * nc[1] * nc[2] : 1;
    c->ncells = ncell;
    c->list_mode = cells_ok ? 0 : 1;
    int32_t *head = NULL, *next = NULL, *cell = NULL;
    if (cells_ok) {
        head = (int32_t *)malloc((size_t)ncell * sizeof(int32_t));
        next = (int32_t *)malloc((size_t)(n + 1) * sizeof(int32_t));
        cell = (int32_t *)malloc((size_t)(n + 1) * 3 * sizeof(int32_t));
        if (!head || !next || !cell) return fail(c, B2_ENOMEM, "build_list: out of memory");
        for (int64_t k = 0; k < ncell; k++) head[k] = -1;
        for (int64_t i = n - 1; i >= 0; i--) {
            int ic[3];
            for (int d = 0; d < 3; d++) {
                double u = (c->pos[3 * i + d] - c->center[d]) / c->side[d] + 0.5;
                int k = (int)floor(u * nc[d]);
                if (k < 0) k = 0;
                if (k >= nc[d]) k = nc[d] - 1;
                ic[d] = cell[3 * i + d] = k;
            }
            int64_t id = ((int64_t)ic[2] * nc[1] + ic[1]) * nc[0] + ic[0];
            next[i] = head[id];
            head[id] = (int32_t)i;
        }
    }
    for (int pass = 0; pass < 2; pass++) {
        if (pass == 1) {
            int64_t total = 0;
            int32_t mx = 0;
            for (int64_t i = 0; i < n; i++) {
                int64_t cnt = c->row[i + 1];
                if (cnt > mx) mx = (int32_t)cnt;
                c->row[i] = total;
                total += cnt;
            }
            c->row[n] = total;
            c->max_count = mx;
            if (total > c->nbr_cap) {
                free(c->nbr);
                c->nbr_cap = total + total / 8 + 16;
                c->nbr = (int32_t *)malloc((size_t)c->nbr_cap * sizeof(int32_t));
                if (!c->nbr) return fail(c, B2_ENOMEM, "build_list: out of memory");
            }
        }
#pragma omp parallel for schedule(dynamic, 64)
        for (int64_t i = 0; i < n; i++) {
            int64_t cnt = 0;
            int32_t *out = pass ? c->nbr + c->row[i] : NULL;
            if (GHOST(c, i)) {
                /* a ghost acts as a neighbour but owns no row */
            } else if (cells_ok) {
                for (int dz = -1; dz <= 1; dz++)
                    for (int dy = -1; dy <= 1; dy++)
                        for (int dx = -1; dx <= 1; dx++) {
                            int cx = (cell[3 * i] + dx + nc[0]) % nc[0];
                            int cy = (cell[3 * i + 1] + dy + nc[1]) % nc[1];
                            int cz = (cell[3 * i + 2] + dz + nc[2]) % nc[2];
                            int64_t id = ((int64_t)cz * nc[1] + cy) * nc[0] + cx;
                            for (int32_t j = head[id]; j >= 0; j = next[j])
                                if (j != i && in_list(c, i, j)) {
                                    if (out) out[cnt] = j;
                                    cnt++;
                                }
                        }
            } else {
                for (int64_t j = 0; j < n; j++)
                    if (j != i && in_list(c, i, j)) {
                        if (out) out[cnt] = (int32_t)j;
                        cnt++;
                    }
            }
            if (pass == 0) c->row[i + 1] = cnt;
            else qsort(out, (size_t)cnt, sizeof(int32_t), cmp_i32);
        }
    }
    free(head); free(next); free(cell);
    c->list_valid = 1;
    c->disp_bound = 0.0;
    c->rebuilds++;
    return B2_OK;
}

#pragma GCC pop_options

/* Pair expressions: identical to oracle_np.compute */
static int compute_forces(b2_ctx *c, const double *pos, double *frc, double *pe, double *vir,
                          double *energy, double *virial, double *stress)
{
    int64_t n = c->n;
    double e_tot = 0.0, w_tot = 0.0;
    double s[9] = {0};
#pragma omp parallel for schedule(static) reduction(+ : e_tot, w_tot, s[:9])
    for (int64_t i = 0; i < n; i++) {
        const double *xi = pos + 3 * i;
        int ti = TYPE(c, i);
        double fx = 0, fy = 0, fz = 0, ei = 0, wi = 0;
        double si[9] = {0};
        for (int64_t k = c->row[i]; k < c->row[i + 1]; k++) {
            int64_t j = c->nbr[k];
            int a = ti * c->nsp + TYPE(c, j);
            double dx = xi[0] - pos[3 * j], dy = xi[1] - pos[3 * j + 1], dz = xi[2] - pos[3 * j + 2];
            dx = dx - rint(dx / c->side[0]) * c->side[0];
            dy = dy - rint(dy / c->side[1]) * c->side[1];
            dz = dz - rint(dz / c->side[2]) * c->side[2];
            double r2 = (dx * dx + dy * dy) + dz * dz;
            if (r2 < c->rc2[a]) {
                double r2inv = 1.0 / r2;
                double u, w;
                if (c->tab_np) {
                    double t = r2 * c->tab_scale[a];
                    int kk = (int)t;
                    if (kk > c->tab_np - 2) kk = c->tab_np - 2;
                    double f = t - (double)kk;
                    const double *pu = c->tab_u + (size_t)a * c->tab_np + kk, *pw = c->tab_w + (size_t)a * c->tab_np + kk;
                    u = pu[0] + f * (pu[1] - pu[0]);
                    w = pw[0] + f * (pw[1] - pw[0]);
                } else {
                    double r6inv = r2inv * r2inv * r2inv;
                    u = r6inv * (c->c12[a] * r6inv - c->c6[a]) - c->ushift[a];
                    w = r6inv * (12.0 * c->c12[a] * r6inv - 6.0 * c->c6[a]);
                }
                double fp = w * r2inv;
                fx += fp * dx; fy += fp * dy; fz += fp * dz;
                ei += u; wi += w;
                if (stress) {
                    double d[3] = {dx, dy, dz};
                    for (int p = 0; p < 3; p++)
                        for (int q = 0; q < 3; q++) si[3 * p + q] += d[p] * (fp * d[q]);
                }
            }
        }
        frc[3 * i] = fx; frc[3 * i + 1] = fy; frc[3 * i + 2] = fz;
        if (pe) pe[i] = 0.5 * ei;
        if (vir) vir[i] = 0.5 * wi;
        e_tot += 0.5 * ei;
        w_tot += 0.5 * wi;
        for (int p = 0; p < 9; p++) s[p] += 0.5 * si[p];
    }
    if (energy) *energy = e_tot;
    if (virial) *virial = w_tot;
    if (stress) memcpy(stress, s, sizeof(s));
    return B2_OK;
}

static int ensure_ready(b2_ctx *c, const char *who)
{
    if (!c) return fail(NULL, B2_EINVAL, "NULL context");
    if (!c->pos) {
        snprintf(c->err, 512, "%s: no particles resident (call b2_upload)", who);
        return B2_ESTATE;
    }
    return B2_OK;
}

static int ensure_forces(b2_ctx *c)
{
    int rc;
    if (!c->list_valid && (rc = build_list(c)) != B2_OK) return rc;
    if (!c->forces_valid) {
        compute_forces(c, c->pos, c->frc, c->pe, c->vir, &c->energy, &c->virial, c->stress);
        c->forces_valid = 1;
    }
    return B2_OK;
}

static double now_ms(void)
{
#ifdef _OPENMP
    return omp_get_wtime() * 1e3;
#else
    return 0.0;
#endif
}

int b2_compute(b2_ctx *c, int what, double *energy, double *virial, double *stress9,
               double *forces, double *pe, double *vir)
{
    int rc = ensure_ready(c, "b2_compute");
    if (rc) return rc;
    if (what & ~(B2_ENERGY | B2_FORCES | B2_VIRIAL | B2_STRESS))
        return fail(c, B2_EUNSUPPORTED, "b2_compute: unsupported observable");
    double t0 = now_ms();
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    if (energy) *energy = c->energy;
    if (virial) *virial = c->virial;
    if (stress9) memcpy(stress9, c->stress, 9 * sizeof(double));
    if (forces) memcpy(forces, c->frc, (size_t)c->n * 3 * sizeof(double));
    if (pe) memcpy(pe, c->pe, (size_t)c->n * sizeof(double));
    if (vir) memcpy(vir, c->vir, (size_t)c->n * sizeof(double));
    c->last_ms = now_ms() - t0;
    return B2_OK;
}

int b2_thermo(b2_ctx *c, double out[B2_T_LEN])
{
    int rc = ensure_ready(c, "b2_thermo");
    if (rc) return rc;
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    for (int k = 0; k < B2_T_LEN; k++) out[k] = 0.0;
    double kin = 0, f2 = 0, px = 0, py = 0, pz = 0, mt = 0, msd = 0;
    for (int64_t i = 0; i < c->n; i++) {
        if (GHOST(c, i)) continue;
        const double *v = c->vel + 3 * i, *f = c->frc + 3 * i;
        double m = c->mass[i];
        kin += 0.5 * m * ((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
        f2 += (f[0] * f[0] + f[1] * f[1]) + f[2] * f[2];
        px += m * v[0]; py += m * v[1]; pz += m * v[2];
        mt += m;
        for (int d = 0; d < 3; d++) {
            double u = (c->pos[3 * i + d] + c->img[3 * i + d] * c->side[d]) -
                       (c->ref[3 * i + d] + c->img_ref[3 * i + d] * c->side[d]);
            msd += u * u;
        }
    }
    out[B2_T_KINETIC] = kin; out[B2_T_POTENTIAL] = c->energy; out[B2_T_VIRIAL] = c->virial;
    out[B2_T_FNORM2] = f2; out[B2_T_PX] = px; out[B2_T_PY] = py; out[B2_T_PZ] = pz;
    out[B2_T_MASS] = mt; out[B2_T_MSD] = msd;
    return B2_OK;
}

int b2_neighbors(b2_ctx *c, int32_t *count, int32_t *nbr, int32_t stride, int32_t *max_count)
{
    int rc = ensure_ready(c, "b2_neighbors");
    if (rc) return rc;
    if ((rc = build_list(c)) != B2_OK) return rc;   /* always of the CURRENT configuration */
    if (max_count) *max_count = c->max_count;
    for (int64_t i = 0; i < c->n; i++) {
        int64_t cnt = c->row[i + 1] - c->row[i];
        if (count) count[i] = (int32_t)cnt;
        if (nbr) {
            if (cnt > stride) return fail(c, B2_EINVAL, "b2_neighbors: stride smaller than max count");
            memcpy(nbr + i * (int64_t)stride, c->nbr + c->row[i], (size_t)cnt * sizeof(int32_t));
        }
    }
    return B2_OK;
}

/* After a drift with half-step velocities: bound += dt * max|v|; the list
 * stays valid while bound <= skin/2 (two particles approach by <= skin).   */
static void advance_bound(b2_ctx *c, double step_len)
{
    c->disp_bound += step_len;
    if (c->disp_bound > 0.5 * c->skin) c->list_valid = 0;
}

int b2_run_nve(b2_ctx *c, int64_t steps, double dt)
{
    int rc = ensure_ready(c, "b2_run_nve");
    if (rc) return rc;
    if (steps < 0 || !(dt > 0.0)) return fail(c, B2_EINVAL, "b2_run_nve: steps >= 0 and dt > 0 required");
    if (steps == 0) return B2_OK;
    c->min_resume = 0;
    double t0 = now_ms();
    int64_t n = c->n;
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    for (int64_t i = 0; i < n; i++)
        for (int d = 0; d < 3; d++) c->vel[3 * i + d] -= (0.5 * dt) * c->frc[3 * i + d] / c->mass[i];
    for (int64_t s = 0; s < steps; s++) {
        if (s > 0) {
            if (!c->list_valid && (rc = build_list(c)) != B2_OK) return rc;
            compute_forces(c, c->pos, c->frc, NULL, NULL, NULL, NULL, NULL);
        }
        double vmax2 = 0.0;
#pragma omp parallel for schedule(static) reduction(max : vmax2)
        for (int64_t i = 0; i < n; i++) {
            double *v = c->vel + 3 * i, *x = c->pos + 3 * i;
            if (GHOST(c, i)) continue;
            for (int d = 0; d < 3; d++) {
                v[d] += dt * c->frc[3 * i + d] / c->mass[i];
                x[d] += dt * v[d];
            }
            fold3(c, x, c->img + 3 * i);
            double v2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
            if (v2 > vmax2) vmax2 = v2;
        }
        advance_bound(c, dt * sqrt(vmax2));
    }
    c->forces_valid = 0;
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    for (int64_t i = 0; i < n; i++)
        for (int d = 0; d < 3; d++) c->vel[3 * i + d] += (0.5 * dt) * c->frc[3 * i + d] / c->mass[i];
    c->steps_total += steps;
    c->last_ms = now_ms() - t0;
    return B2_OK;
}

static double sum_f2(const b2_ctx *c, const double *f)
{
    double s = 0;
    for (int64_t i = 0; i < 3 * c->n; i++) s += f[i] * f[i];
    return s;
}

static void park(b2_ctx *c, int method, int converged)
{
    size_t n3 = (size_t)c->n * 3;
    c->min_resume = 0;
    if (!converged) {
        if (!c->mvel) c->mvel = (double *)malloc((n3 ? n3 : 1) * sizeof(double));
        memcpy(c->mvel, c->vel, n3 * sizeof(double));
        c->min_resume = 1;
        c->min_method = method;
    }
    memset(c->vel, 0, n3 * sizeof(double));
}

static int minimize_fire(b2_ctx *c, int64_t max_steps, double dt, double tol,
                         int64_t *steps_done, int *converged)
{
    /* Bitzek et al. 2006 defaults; see oracle_np.fire for the iteration */
    const int n_min = 5;
    const double f_inc = 1.1, f_dec = 0.5, alpha0 = 0.1, f_alpha = 0.99;
    int64_t n = c->n, it = 0;
    int rc, npos = 0;
    double dt_now = dt, alpha = alpha0, dt_max = 10.0 * dt;
    if (c->min_resume && c->min_method == B2_MIN_FIRE) {
        memcpy(c->vel, c->mvel, (size_t)n * 3 * sizeof(double));
        dt_now = c->m_dt; alpha = c->m_alpha; npos = c->m_npos;
    } else {
        memset(c->vel, 0, (size_t)n * 3 * sizeof(double));
        c->m_dt = dt;   /* dt_max refers to the dt of the fresh start */
    }
    dt_max = 10.0 * (c->min_resume && c->min_method == B2_MIN_FIRE ? c->m_dt0 : dt);
    *converged = 0;
    for (;;) {
        if ((rc = ensure_forces(c)) != B2_OK) return rc;
        double ff = sum_f2(c, c->frc);
        if (ff / (double)n < tol) { *converged = 1; break; }
        if (it >= max_steps) break;
        double power = 0, vv = 0;
        for (int64_t k = 0; k < 3 * n; k++) { power += c->frc[k] * c->vel[k]; vv += c->vel[k] * c->vel[k]; }
        if (power > 0.0) {
            double a = 1.0 - alpha, b = alpha * sqrt(vv / ff);
            for (int64_t k = 0; k < 3 * n; k++) c->vel[k] = a * c->vel[k] + b * c->frc[k];
            npos++;
            if (npos > n_min) { dt_now = fmin(dt_now * f_inc, dt_max); alpha *= f_alpha; }
        } else {
            npos = 0; dt_now *= f_dec; alpha = alpha0;
            memset(c->vel, 0, (size_t)n * 3 * sizeof(double));
        }
        double vmax2 = 0;
        for (int64_t i = 0; i < n; i++) {
            double *v = c->vel + 3 * i, *x = c->pos + 3 * i;
            for (int d = 0; d < 3; d++) {
                v[d] += dt_now * c->frc[3 * i + d] / c->mass[i];
                x[d] += dt_now * v[d];
            }
            fold3(c, x, c->img + 3 * i);
            double v2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
            if (v2 > vmax2) vmax2 = v2;
        }
        advance_bound(c, dt_now * sqrt(vmax2));
        c->forces_valid = 0;
        it++;
    }
    if (!(c->min_resume && c->min_method == B2_MIN_FIRE)) c->m_dt0 = dt;
    c->m_dt = dt_now; c->m_alpha = alpha; c->m_npos = npos;
    park(c, B2_MIN_FIRE, *converged);
    *steps_done = it;
    return B2_OK;
}

static int minimize_sd(b2_ctx *c, int64_t max_steps, double step, double tol,
                       int64_t *steps_done, int *converged)
{
    /* see oracle_np.steepest_descent: the previous force is kept in c->vel */
    const double max_disp = 0.1;
    int64_t n = c->n, it = 0;
    int rc, has_prev = 0;
    double h_now = step;
    if (c->min_resume && c->min_method == B2_MIN_SD) {
        memcpy(c->vel, c->mvel, (size_t)n * 3 * sizeof(double));
        h_now = c->m_dt; has_prev = c->m_has_prev;
    } else {
        memset(c->vel, 0, (size_t)n * 3 * sizeof(double));
    }
    *converged = 0;
    for (;;) {
        if ((rc = ensure_forces(c)) != B2_OK) return rc;
        double ff = sum_f2(c, c->frc);
        if (ff / (double)n < tol) { *converged = 1; break; }
        if (it >= max_steps) break;
        double dot = 0, fmax2 = 0;
        for (int64_t k = 0; k < 3 * n; k++) dot += c->frc[k] * c->vel[k];
        for (int64_t i = 0; i < n; i++) {
            const double *f = c->frc + 3 * i;
            double f2 = (f[0] * f[0] + f[1] * f[1]) + f[2] * f[2];
            if (f2 > fmax2) fmax2 = f2;
        }
        if (has_prev) h_now *= (dot > 0.0) ? 1.1 : 0.5;
        has_prev = 1;
        double h = fmin(h_now, max_disp / sqrt(fmax2));
        for (int64_t i = 0; i < n; i++) {
            for (int d = 0; d < 3; d++) {
                c->pos[3 * i + d] += h * c->frc[3 * i + d];
                c->vel[3 * i + d] = c->frc[3 * i + d];
            }
            fold3(c, c->pos + 3 * i, c->img + 3 * i);
        }
        advance_bound(c, h * sqrt(fmax2));
        c->forces_valid = 0;
        it++;
    }
    c->m_dt = h_now; c->m_has_prev = has_prev;
    park(c, B2_MIN_SD, *converged);
    *steps_done = it;
    return B2_OK;
}

int b2_minimize(b2_ctx *c, int method, int64_t max_steps, double dt, double fn2_tol,
                int64_t *steps_done, int *converged)
{
    int rc = ensure_ready(c, "b2_minimize");
    if (rc) return rc;
    if (max_steps < 0 || !(dt > 0.0) || !steps_done || !converged)
        return fail(c, B2_EINVAL, "b2_minimize: bad argument");
    double t0 = now_ms();
    if (c->n == 0) { *steps_done = 0; *converged = 1; return B2_OK; }
    if (method == B2_MIN_FIRE) rc = minimize_fire(c, max_steps, dt, fn2_tol, steps_done, converged);
    else if (method == B2_MIN_SD) rc = minimize_sd(c, max_steps, dt, fn2_tol, steps_done, converged);
    else return fail(c, B2_EINVAL, "b2_minimize: unknown method");
    if (rc == B2_OK) c->steps_total += *steps_done;
    c->last_ms = now_ms() - t0;
    return rc;
}

int b2_stats(b2_ctx *c, int64_t out[B2_S_LEN])
{
    if (!c) return fail(NULL, B2_EINVAL, "NULL context");
    for (int k = 0; k < B2_S_LEN; k++) out[k] = 0;
    out[B2_S_STEPS] = c->steps_total;
    out[B2_S_REBUILDS] = c->rebuilds;
    out[B2_S_LAUNCHES] = 0;
    out[B2_S_MAX_NEIGH] = c->max_count;
    out[B2_S_NEIGH_CAP] = c->max_count;
    out[B2_S_LIST_MODE] = c->list_mode;
    out[B2_S_NCELLS] = c->ncells;
    {
        double m = c->side[0] < c->side[1] ? c->side[0] : c->side[1];
        if (c->side[2] < m) m = c->side[2];
        out[B2_S_SMALL_BOX] = (c->have_box && c->have_table && 2.0 * c->rcut_max > m) ? 1 : 0;
    }
    return B2_OK;
}

int b2_last_timing(b2_ctx *c, double *ms, int64_t *launches)
{
    if (!c) return fail(NULL, B2_EINVAL, "NULL context");
    if (ms) *ms = c->last_ms;
    if (launches) *launches = 0;
    return B2_OK;
}

int b2_set_images(b2_ctx *c, const int32_t *image)
{
    if (!c || !c->pos || !image) return fail(c, B2_EINVAL, "b2_set_images: bad argument");
    memcpy(c->img, image, (size_t)c->n * 3 * sizeof(int32_t));
    return B2_OK;
}
int b2_get_reference(b2_ctx *c, double *ref)
{
    if (!c || !c->pos || !ref) return fail(c, B2_EINVAL, "b2_get_reference: bad argument");
    for (int64_t k = 0; k < 3 * c->n; k++) ref[k] = c->ref[k] + c->img_ref[k] * c->side[k % 3];
    return B2_OK;
}
int b2_set_reference(b2_ctx *c, const double *ref)
{
    if (!c || !c->pos || !ref) return fail(c, B2_EINVAL, "b2_set_reference: bad argument");
    memcpy(c->ref, ref, (size_t)c->n * 3 * sizeof(double));
    memset(c->img_ref, 0, (size_t)c->n * 3 * sizeof(int32_t));
    return B2_OK;
}
int b2_get_thermostat(b2_ctx *c, double *xi)
{
    if (!c || !xi) return fail(c, B2_EINVAL, "b2_get_thermostat: bad argument");
    *xi = c->th_xi;
    return B2_OK;
}
int b2_set_thermostat(b2_ctx *c, double xi)
{
    if (!c) return fail(c, B2_EINVAL, "b2_set_thermostat: bad argument");
    c->th_xi = xi;
    return B2_OK;
}

int b2_hessian(b2_ctx *c, double *out)
{
    if (!c || !c->pos || !out) return fail(c, B2_EINVAL, "b2_hessian: bad argument");
    if (c->tab_np) return fail(c, B2_EUNSUPPORTED, "b2_hessian: 12-6 table only");
    int64_t n = c->n;
    if (n > B2_HESSIAN_MAX_PARTICLES) return fail(c, B2_EINVAL, "b2_hessian: n exceeds B2_HESSIAN_MAX_PARTICLES");
    memset(out, 0, (size_t)(9 * n * n) * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++)
        for (int64_t j = 0; j < n; j++) {
            if (j == i) continue;
            int a = TYPE(c, i) * c->nsp + TYPE(c, j);
            double d[3];
            for (int k = 0; k < 3; k++) {
                d[k] = c->pos[3 * i + k] - c->pos[3 * j + k];
                d[k] -= rint(d[k] / c->side[k]) * c->side[k];
            }
            double r2 = (d[0] * d[0] + d[1] * d[1]) + d[2] * d[2];
            if (!(r2 < c->rc2[a])) continue;
            double i2 = 1.0 / r2, i6 = i2 * i2 * i2;
            double up_r = (-12.0 * c->c12[a] * i6 + 6.0 * c->c6[a]) * i6 * i2;
            double upp = (156.0 * c->c12[a] * i6 - 42.0 * c->c6[a]) * i6 * i2;
            double kk = (upp - up_r) * i2;
            for (int p = 0; p < 3; p++)
                for (int q = 0; q < 3; q++) {
                    double v = kk * d[p] * d[q] + (p == q ? up_r : 0.0);
                    out[((p * n + i) * 3 + q) * n + j] = -v;
                    out[((p * n + i) * 3 + q) * n + i] += v;
                }
        }
    return B2_OK;
}

int b2_format_xyz(char *out, int64_t cap, int64_t *written, int64_t n, const int32_t *ids, int nnames,
                  const char *const *names, const double *pos, const double *vel, int precision)
{
    if (!written || n < 0 || (n > 0 && (!ids || !names || !pos)) || precision < 0 || precision > 17)
        return fail(NULL, B2_EINVAL, "b2_format_xyz: bad argument");
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
    if ((int64_t)(line * (size_t)n) > cap) return fail(NULL, B2_EINVAL, "b2_format_xyz: output buffer too small");
    /* every chunk of rows is formatted at its worst-case offset, then the chunks are closed up in order */
    const int64_t chunk = 4096;
    const int64_t nchunk = (n + chunk - 1) / chunk;
    int64_t *len = (int64_t *)malloc((size_t)(nchunk > 0 ? nchunk : 1) * sizeof(int64_t));
    if (!len) return fail(NULL, B2_ENOMEM, "b2_format_xyz: out of memory");
    #pragma omp parallel for schedule(dynamic, 1)
    for (int64_t c = 0; c < nchunk; c++) {
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
    if (!out || bytes < 0) return fail(NULL, B2_EINVAL, "b2_host_alloc: bad argument");
    *out = malloc((size_t)(bytes > 0 ? bytes : 1));
    return *out ? B2_OK : fail(NULL, B2_ENOMEM, "b2_host_alloc: out of memory");
}
int b2_host_free(void *ptr) { free(ptr); return B2_OK; }

int b2_set_profile(b2_ctx *c, int on) { (void)c; (void)on; return B2_OK; }

int b2_profile(b2_ctx *c, double out[8])
{
    if (!c || !out) return fail(c, B2_EINVAL, "b2_profile: NULL argument");
    for (int k = 0; k < 8; k++) out[k] = 0.0;
    out[0] = c->last_ms;
    return B2_OK;
}

/* device-pointer entry point of the CUDA library: meaningless on the host */
int b2_kinetic_static(const void *vel_mass_dev, int64_t n, int precision, void *cuda_stream, double *out_dev)
{
    (void)vel_mass_dev; (void)n; (void)precision; (void)cuda_stream; (void)out_dev;
    return fail(NULL, B2_EUNSUPPORTED, "b2_kinetic_static takes device pointers; the CPU oracle has none");
}

/* ---- domain decomposition primitives: the "device" of the oracle is the host ---- */
int b2_upload_device(b2_ctx *c, int64_t n, const double *pos, const double *vel, const int32_t *species,
                     const double *mass)
{
    return b2_upload(c, n, pos, vel, species, mass);
}

int b2_download_device(b2_ctx *c, double *pos, double *vel, int32_t *image)
{
    return b2_download(c, pos, vel, image);
}

int b2_dd_begin(b2_ctx *c, int64_t steps, double dt, int opening_kick)
{
    int rc = ensure_ready(c, "b2_dd_begin");
    if (rc) return rc;
    c->dd_dt = dt;
    c->dd_steps = 0;
    c->dd_target = steps;
    c->min_resume = 0;
    if (opening_kick) {
        if ((rc = ensure_forces(c)) != B2_OK) return rc;
        for (int64_t i = 0; i < c->n; i++)
            for (int d = 0; d < 3; d++) c->vel[3 * i + d] -= (0.5 * dt) * c->frc[3 * i + d] / c->mass[i];
    } else if (!c->list_valid && (rc = build_list(c)) != B2_OK) {
        return rc;
    }
    c->dd_halt = 0;
    return B2_OK;
}

int b2_dd_step(b2_ctx *c, float *vmax2)
{
    int rc = ensure_ready(c, "b2_dd_step");
    if (rc) return rc;
    if (!vmax2) return fail(c, B2_EINVAL, "b2_dd_step: NULL argument");
    *vmax2 = 0.f;
    if (c->dd_halt || c->dd_steps >= c->dd_target) return B2_OK;
    double dt = c->dd_dt, vmax2v = 0.0;
    compute_forces(c, c->pos, c->frc, NULL, NULL, NULL, NULL, NULL);
    for (int64_t i = 0; i < c->n; i++) {
        double *v = c->vel + 3 * i, *x = c->pos + 3 * i;
        if (GHOST(c, i)) continue;
        for (int d = 0; d < 3; d++) {
            v[d] += dt * c->frc[3 * i + d] / c->mass[i];
            x[d] += dt * v[d];
        }
        fold3(c, x, c->img + 3 * i);
        double v2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
        if (v2 > vmax2v) vmax2v = v2;
    }
    *vmax2 = (float)vmax2v;
    if ((double)*vmax2 < vmax2v) *vmax2 = nextafterf(*vmax2, INFINITY);   /* round up */
    c->forces_valid = 0;
    return B2_OK;
}

int b2_dd_commit(b2_ctx *c, const float *gvmax2)
{
    int rc = ensure_ready(c, "b2_dd_commit");
    if (rc) return rc;
    if (c->dd_halt || c->dd_steps >= c->dd_target) return B2_OK;
    c->disp_bound += c->dd_dt * sqrt((double)*gvmax2) * (1.0 + 1e-6);
    c->dd_steps++;
    c->steps_total++;
    if (c->disp_bound > 0.5 * c->skin) { c->dd_halt = 1; c->list_valid = 0; }
    return B2_OK;
}

int b2_dd_poll(b2_ctx *c, int64_t *steps_done, int *need_refresh)
{
    int rc = ensure_ready(c, "b2_dd_poll");
    if (rc) return rc;
    if (steps_done) *steps_done = c->dd_steps;
    if (need_refresh) *need_refresh = c->dd_halt;
    return B2_OK;
}

int b2_dd_end(b2_ctx *c)
{
    int rc = ensure_ready(c, "b2_dd_end");
    if (rc) return rc;
    c->forces_valid = 0;
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    for (int64_t i = 0; i < c->n; i++)
        for (int d = 0; d < 3; d++) c->vel[3 * i + d] += (0.5 * c->dd_dt) * c->frc[3 * i + d] / c->mass[i];
    return B2_OK;
}

int b2_dd_pack(b2_ctx *c, const int32_t *ids, int64_t n, double *out)
{
    int rc = ensure_ready(c, "b2_dd_pack");
    if (rc) return rc;
    for (int64_t k = 0; k < n; k++)
        for (int d = 0; d < 3; d++) out[3 * k + d] = c->pos[3 * (int64_t)ids[k] + d] - c->center[d];
    return B2_OK;
}

int b2_dd_unpack(b2_ctx *c, int64_t first_id, int64_t n, const double *in)
{
    int rc = ensure_ready(c, "b2_dd_unpack");
    if (rc) return rc;
    if (first_id < 0 || first_id + n > c->n) return fail(c, B2_EINVAL, "b2_dd_unpack: id range outside the resident particles");
    for (int64_t k = 0; k < n; k++)
        for (int d = 0; d < 3; d++) c->pos[3 * (first_id + k) + d] = in[3 * k + d] + c->center[d];
    return B2_OK;
}

/* ---- replica ensembles: every replica is an ordinary context, minimised in turn ---- */
struct b2_batch {
    int precision, nsp, shifted, n, have_model;
    int64_t nrep;
    double side[3];
    double eps[MAXSP * MAXSP], sigma[MAXSP * MAXSP], rcut[MAXSP * MAXSP];
    b2_ctx **rep;
};

int b2_batch_create(int device, int precision, b2_batch **out)
{
    (void)device;
    if (!out || (precision != B2_FP32 && precision != B2_FP64)) return fail(NULL, B2_EINVAL, "b2_batch_create: bad argument");
    b2_batch *b = (b2_batch *)calloc(1, sizeof(b2_batch));
    if (!b) return fail(NULL, B2_ENOMEM, "b2_batch_create: out of memory");
    b->precision = precision;
    *out = b;
    return B2_OK;
}

static void batch_free(b2_batch *b)
{
    for (int64_t r = 0; r < b->nrep; r++) b2_destroy(b->rep[r]);
    free(b->rep);
    b->rep = NULL;
    b->nrep = 0;
}

int b2_batch_destroy(b2_batch *b)
{
    if (!b) return B2_OK;
    batch_free(b);
    free(b);
    return B2_OK;
}

int b2_batch_set_model(b2_batch *b, const double side[3], int nsp, const double *eps, const double *sigma,
                       const double *rcut, int shifted)
{
    if (!b || !side || !eps || !sigma || !rcut || nsp < 1 || nsp > MAXSP) return fail(NULL, B2_EINVAL, "b2_batch_set_model: bad argument");
    memcpy(b->side, side, 3 * sizeof(double));
    b->nsp = nsp; b->shifted = shifted;
    memcpy(b->eps, eps, (size_t)nsp * nsp * sizeof(double));
    memcpy(b->sigma, sigma, (size_t)nsp * nsp * sizeof(double));
    memcpy(b->rcut, rcut, (size_t)nsp * nsp * sizeof(double));
    b->have_model = 1;
    return B2_OK;
}

int b2_batch_upload(b2_batch *b, int64_t nrep, int n, const double *pos, const int32_t *species, const double *mass)
{
    if (!b || nrep < 0 || n < 0 || !b->have_model) return fail(NULL, B2_EINVAL, "b2_batch_upload: bad argument or no model");
    batch_free(b);
    b->rep = (b2_ctx **)calloc((size_t)(nrep > 0 ? nrep : 1), sizeof(b2_ctx *));
    b->n = n;
    for (int64_t r = 0; r < nrep; r++) {
        int rc = b2_create(0, b->precision, &b->rep[r]);
        if (rc) return rc;
        b->nrep = r + 1;
        b2_set_box(b->rep[r], b->side, NULL);
        b2_set_pair_table(b->rep[r], b->nsp, b->eps, b->sigma, b->rcut, b->shifted);
        rc = b2_upload(b->rep[r], n, pos + 3 * r * (int64_t)n, NULL, species + r * (int64_t)n, mass ? mass + r * (int64_t)n : NULL);
        if (rc) return fail(NULL, rc, b2_last_error(b->rep[r]));
    }
    return B2_OK;
}

int b2_batch_minimize(b2_batch *b, int method, int64_t max_steps, double dt, double fn2_tol, int64_t *steps_done,
                      int *converged, double *energy, double *fn2, double *virial)
{
    if (!b) return fail(NULL, B2_EINVAL, "b2_batch_minimize: NULL batch");
    int err = B2_OK;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t r = 0; r < b->nrep; r++) {
        int64_t it = 0;
        int conv = 0;
        double th[B2_T_LEN];
        int rc = b2_minimize(b->rep[r], method, max_steps, dt, fn2_tol, &it, &conv);
        if (rc == B2_OK) rc = b2_thermo(b->rep[r], th);
        if (rc != B2_OK) { err = rc; continue; }
        if (steps_done) steps_done[r] = it;
        if (converged) converged[r] = conv;
        if (energy) energy[r] = th[B2_T_POTENTIAL];
        if (fn2) fn2[r] = th[B2_T_FNORM2] / (double)(b->n > 0 ? b->n : 1);
        if (virial) virial[r] = th[B2_T_VIRIAL];
    }
    return err == B2_OK ? B2_OK : fail(NULL, err, "b2_batch_minimize: a replica failed");
}

int b2_batch_download(b2_batch *b, double *pos, double *forces)
{
    if (!b) return fail(NULL, B2_EINVAL, "b2_batch_download: NULL batch");
    for (int64_t r = 0; r < b->nrep; r++) {
        int rc = B2_OK;
        if (pos) rc = b2_download(b->rep[r], pos + 3 * r * (int64_t)b->n, NULL, NULL);
        if (rc == B2_OK && forces)
            rc = b2_compute(b->rep[r], B2_FORCES, NULL, NULL, NULL, forces + 3 * r * (int64_t)b->n, NULL, NULL);
        if (rc) return rc;
    }
    return B2_OK;
}

int b2_batch_set_velocities(b2_batch *b, const double *vel)
{
    if (!b || !vel) return fail(NULL, B2_EINVAL, "b2_batch_set_velocities: NULL argument");
    for (int64_t r = 0; r < b->nrep; r++) {
        int rc = b2_update(b->rep[r], NULL, vel + 3 * r * (int64_t)b->n);
        if (rc) return rc;
    }
    return B2_OK;
}

int b2_batch_run_nve(b2_batch *b, int64_t steps, double dt, double *kinetic, double *energy, double *virial)
{
    if (!b) return fail(NULL, B2_EINVAL, "b2_batch_run_nve: NULL batch");
    int err = B2_OK;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t r = 0; r < b->nrep; r++) {
        double th[B2_T_LEN];
        int rc = b2_run_nve(b->rep[r], steps, dt);
        if (rc == B2_OK) rc = b2_thermo(b->rep[r], th);
        if (rc != B2_OK) { err = rc; continue; }
        if (kinetic) kinetic[r] = th[B2_T_KINETIC];
        if (energy) energy[r] = th[B2_T_POTENTIAL];
        if (virial) virial[r] = th[B2_T_VIRIAL];
    }
    return err == B2_OK ? B2_OK : fail(NULL, err, "b2_batch_run_nve: a replica failed");
}

int b2_batch_get_state(b2_batch *b, double *vel, int32_t *image)
{
    if (!b) return fail(NULL, B2_EINVAL, "b2_batch_get_state: NULL batch");
    for (int64_t r = 0; r < b->nrep; r++) {
        int rc = b2_download(b->rep[r], NULL, vel ? vel + 3 * r * (int64_t)b->n : NULL,
                             image ? image + 3 * r * (int64_t)b->n : NULL);
        if (rc) return rc;
    }
    return B2_OK;
}

/* Thermostatted dynamics: same kick-drift loop as b2_run_nve with v' = a v + b dt F/m, the
 * coefficients renewed after every step from the kinetic temperature of the half-step
 * velocities just produced (one step late), exactly as the fused CUDA kernel does. */
int b2_run_nvt(b2_ctx *c, int64_t steps, double dt, double T0, double tau, int kind)
{
    int rc = ensure_ready(c, "b2_run_nvt");
    if (rc) return rc;
    if (steps < 0 || !(dt > 0.0) || !(T0 > 0.0) || !(tau > 0.0) || (kind != 0 && kind != 1))
        return fail(c, B2_EINVAL, "b2_run_nvt: bad argument");
    if (steps == 0) return B2_OK;
    c->min_resume = 0;
    int64_t n = c->n;
    const double ndof = 3.0 * (double)(n > 1 ? n - 1 : 1);
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    for (int64_t i = 0; i < n; i++)
        for (int d = 0; d < 3; d++) c->vel[3 * i + d] -= (0.5 * dt) * c->frc[3 * i + d] / c->mass[i];
    double h = kind == 1 ? 0.5 * c->th_xi * dt : 0.0;
    double a = (1.0 - h) / (1.0 + h), b = 1.0 / (1.0 + h);
    for (int64_t s = 0; s < steps; s++) {
        if (s > 0) {
            if (!c->list_valid && (rc = build_list(c)) != B2_OK) return rc;
            compute_forces(c, c->pos, c->frc, NULL, NULL, NULL, NULL, NULL);
        }
        double vmax2 = 0.0, mv2 = 0.0;
        for (int64_t i = 0; i < n; i++) {
            double *v = c->vel + 3 * i, *x = c->pos + 3 * i;
            if (GHOST(c, i)) continue;
            for (int d = 0; d < 3; d++) {
                v[d] = a * v[d] + b * (dt / c->mass[i]) * c->frc[3 * i + d];
                x[d] += dt * v[d];
            }
            fold3(c, x, c->img + 3 * i);
            double v2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
            mv2 += c->mass[i] * v2;
            if (v2 > vmax2) vmax2 = v2;
        }
        advance_bound(c, dt * sqrt(vmax2));
        const double T = mv2 / ndof;
        if (kind == 0) {
            double l2 = 1.0 + dt / tau * (T0 / (T > 1e-300 ? T : 1e-300) - 1.0);
            a = b = sqrt(l2 > 0.25 ? l2 : 0.25);
        } else {
            c->th_xi += dt / (tau * tau) * (T / T0 - 1.0);
            h = 0.5 * c->th_xi * dt;
            a = (1.0 - h) / (1.0 + h);
            b = 1.0 / (1.0 + h);
        }
    }
    c->forces_valid = 0;
    if ((rc = ensure_forces(c)) != B2_OK) return rc;
    for (int64_t i = 0; i < n; i++)
        for (int d = 0; d < 3; d++) c->vel[3 * i + d] += (0.5 * dt) * c->frc[3 * i + d] / c->mass[i];
    c->steps_total += steps;
    return B2_OK;
}
