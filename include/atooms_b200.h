/*
 * atooms_b200.h — C ABI of the B200-native pair-interaction / MD engine.
 *
 * This is the drop-in boundary for the ONE hot path of atooms-mirror/atooms
 * (SURVEY.md §8): pair force/energy/virial evaluation, neighbour search, the
 * NVE velocity-Verlet loop and SD/FIRE minimisation.  The reference reaches
 * that path through Python duck-typed plugins and external engines (LAMMPS
 * subprocess, RUMD SWIG module); it has no FFI of its own, so every entry
 * point below cites the reference *interface* it serves.  All citations are
 * `path:line` into the reference tree (atooms 3.14.1).
 *
 * Conventions
 *   - plain C: pointers + sizes, no C++/torch types.  All array arguments are
 *     HOST pointers unless a name ends in `_dev`.
 *   - positions / velocities / forces are N x 3, xyz interleaved
 *     ([x0,y0,z0,x1,...]) float64 — byte-identical to the (3,N) Fortran-order
 *     arrays `System.dump(..., order='F')` hands to `Interaction.compute`
 *     (atooms/system/system.py:245-259, :486-487).
 *   - every function returns 0 (B2_OK) or a negative B2_E* code, never throws
 *     or aborts; `b2_last_error` gives the message (Python side raises, like
 *     atooms/backends/lammps.py:68-73 re-raises engine failures).
 *   - a context is re-entrant per handle, not thread-safe on one handle
 *     (the reference host is single-threaded).
 *
 * Two libraries export this ABI: the product `libatooms_b200.so` (CUDA,
 * sm_100a) and the CPU oracle `oracle/_build/liboracle_b2.so` (test
 * infrastructure; never loaded by the product package).
 */
#ifndef ATOOMS_B200_H
#define ATOOMS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 1

/* error codes */
#define B2_OK            0
#define B2_EINVAL       -1   /* bad argument (-> ValueError)                         */
#define B2_ECUDA        -2   /* CUDA runtime failure                                 */
#define B2_ENOMEM       -3   /* allocation failed                                    */
#define B2_ESTATE       -4   /* call order: box / table / particles not set yet      */
#define B2_EUNSUPPORTED -5   /* e.g. the hessian of a tabulated potential               */
#define B2_ENOCUDA      -6   /* no CUDA device / driver: the product has no CPU path */

/* arithmetic the path computes in */
#define B2_FP32 32
#define B2_FP64 64

/* `what` bit mask for b2_compute; mirrors the observable names asserted at
 * atooms/system/interaction.py:65-66 ('energy','forces','virial','stress').
 * 'hessian' has an entry point of its own: b2_hessian.                      */
#define B2_ENERGY  1
#define B2_FORCES  2
#define B2_VIRIAL  4
#define B2_STRESS  8

/* minimiser methods: `method` of the EnergyMinimization surface
 * (atooms/backends/dryrun.py:129-144, atooms/backends/lammps.py:291-311).   */
#define B2_MIN_SD   0
#define B2_MIN_FIRE 1

/* b2_thermo slots */
#define B2_T_KINETIC     0   /* sum 1/2 m v^2          (system.py:232-243)            */
#define B2_T_POTENTIAL   1   /* U                      (system.py:261-274)            */
#define B2_T_VIRIAL      2   /* W = sum_{i<j} r.F      (system.py:321-335, :348)      */
#define B2_T_FNORM2      3   /* sum F^2                (system.py:305-319)            */
#define B2_T_PX          4   /* total momentum x,y,z   (particle.py:156-165)          */
#define B2_T_PY          5
#define B2_T_PZ          6
#define B2_T_MASS        7   /* total mass                                            */
#define B2_T_MSD         8   /* sum |unfolded x - x at b2_mark_reference|^2 (rumd.py:163-177) */
#define B2_T_LEN        16

/* b2_stats slots (int64) */
#define B2_S_STEPS       0   /* MD / minimiser steps since create                     */
#define B2_S_REBUILDS    1   /* neighbour-list rebuilds since create                  */
#define B2_S_LAUNCHES    2   /* kernels launched by this library since create         */
#define B2_S_MAX_NEIGH   3   /* largest neighbour count in the current list           */
#define B2_S_NEIGH_CAP   4   /* slots per particle in the current list                */
#define B2_S_LIST_MODE   5   /* 0 = cell list, 1 = all-pairs (box < 3 cells wide),
                                2 = cell list with windowed rows (CUDA engine, >= 8 cells per axis) */
#define B2_S_NCELLS      6
#define B2_S_SMALL_BOX   7   /* 1: some side < 2 r_cut,max (single minimum image only) */
#define B2_S_WIN_BLOCKS  8   /* list mode 2: super-blocks (4x4x4 cells) tabulated and launched: all of the grid,
                                or those of the slab a rank of the decomposition holds; else 0 */
#define B2_S_LEN        16

typedef struct b2_ctx b2_ctx;

/* ---- library ---------------------------------------------------------- */

/* ABI version of the loaded library (== B2_ABI_VERSION).
 * Serves: `backend.version` read by atooms/simulation/core.py:443-444.      */
int b2_abi_version(void);

/* "cuda-sm_100a" for the product, "oracle-c" for the CPU oracle.            */
const char *b2_backend_name(void);

/* Message of the last failure on `ctx` (or of the last failed b2_create when
 * ctx == NULL).  Replaces the engine-output echo of lammps.py:68-73.        */
const char *b2_last_error(const b2_ctx *ctx);

/* ---- context ----------------------------------------------------------- */

/* Create an engine on CUDA device `device` computing in `precision`
 * (B2_FP32 | B2_FP64).  Replaces engine construction in
 * atooms/backends/rumd.py:52-121 / lammps.py:151-202.  Fails with
 * B2_ENOCUDA when no GPU is usable — there is no CPU fallback.              */
int b2_create(int device, int precision, b2_ctx **out);
int b2_destroy(b2_ctx *ctx);

/* Run all work of this context on `cuda_stream` (a cudaStream_t; NULL = the
 * legacy default stream).  The caller keeps ownership of the stream.        */
int b2_set_stream(b2_ctx *ctx, void *cuda_stream);

/* ---- model -------------------------------------------------------------- */

/* Orthorhombic periodic cell: atooms/system/cell.py:13-33 (side, center).
 * All interactions follow the MINIMUM-IMAGE convention (particle.py:129-130):
 * exact while every side >= 2 r_cut,max; in a smaller box only the nearest
 * image of a pair interacts (the reference's own FCC N=108 fixture has
 * L = 4.76 < 5; none of its pairs has two images within the cut-off).
 * B2_S_SMALL_BOX of b2_stats is 1 in that case.
 * With particles RESIDENT a new box is an affine rescaling about the centre:
 * fractional coordinates are kept (positions and the rmsd origin scale with
 * the sides, velocities and box-crossing counters are untouched, the list
 * is rebuilt) — the volume move of a barostat.  To replace the particles as
 * well, call b2_upload afterwards, as before.                               */
int b2_set_box(b2_ctx *ctx, const double side[3], const double center[3]);

/* 12-6 Lennard-Jones table for `nsp` species, row-major nsp x nsp:
 * u = 4 eps [(sigma/r)^12 - (sigma/r)^6] (- u(rcut) if `shifted`), r < rcut.
 * `rcut` is absolute.  Replaces the potential definitions the reference
 * passes to its engines: tests/test_rumd.py:26-31 (Pot_LJ_12_6 SetParams),
 * tests/test_lammps.py:24-27 (pair_style lj/cut, pair_modify shift).        */
int b2_set_pair_table(b2_ctx *ctx, int nsp, const double *eps,
                      const double *sigma, const double *rcut, int shifted);

/* General pair potentials as lookup tables.  The reference describes an interaction as
 * func / params / species / cutoff with `npoints` = 20000 tabulation points
 * (atooms/trajectory/hdf5.py:21-43: _PairPotential, _CutOff; "potential tabulation",
 * CHANGELOG.md:151) and leaves the evaluation to its engines.  Here the caller samples
 *   u[a][k] = u_ab(r),  w[a][k] = -r du_ab/dr   at  r^2 = k rcut_ab^2 / (npoints - 1),
 * a = ti * nsp + tj, k = 0 .. npoints - 1 (whatever cut-off scheme it wants is already in u),
 * and the kernels interpolate linearly in r^2: F_i = (w / r^2) (r_i - r_j) for r < rcut_ab.
 * The first samples (r -> 0) may hold any finite value.  Replaces b2_set_pair_table for this
 * context; replica batches (b2_batch_*) stay 12-6 Lennard-Jones.                              */
int b2_set_pair_tabulated(b2_ctx *ctx, int nsp, int npoints, const double *rcut,
                          const double *u, const double *w);

/* Verlet skin (LAMMPS `neighbor 0.3 bin`, tests/test_lammps.py:27).         */
int b2_set_skin(b2_ctx *ctx, double skin);

/* ---- state -------------------------------------------------------------- */

/* Copy the particle state to the device (what LAMMPS.run does by writing a
 * data file, lammps.py:243-246, and RUMD by ReadConf, rumd.py:69-75).
 * vel == NULL -> zeros; mass == NULL -> 1.0; species are 0-based ids.
 * Positions are folded into the cell (particle.py:90-99); image counters and
 * the rmsd reference are reset.                                             */
int b2_upload(b2_ctx *ctx, int64_t n, const double *pos, const double *vel,
              const int32_t *species, const double *mass);

/* Replace only positions and/or velocities (either may be NULL) of the
 * resident particles; used when host views were modified between run()
 * calls (docs/simulations.org:302-308).                                     */
int b2_update(b2_ctx *ctx, const double *pos, const double *vel);

/* Copy state back in ORIGINAL particle order; any pointer may be NULL.
 * `image` is N x 3 int32 box-crossing counters (RUMD GetImages,
 * rumd.py:172-176).  Replaces the dump read-back of lammps.py:277-285.      */
int b2_download(b2_ctx *ctx, double *pos, double *vel, int32_t *image);

/* v *= factor   (System.scale_velocities, system.py:227-230)                */
int b2_scale_velocities(b2_ctx *ctx, double factor);

/* Remember the current unfolded positions as the rmsd origin
 * (rumd.py:117-121 snapshots `_initial_sample`).                            */
int b2_mark_reference(b2_ctx *ctx);

/* ---- single point ------------------------------------------------------- */

/* Evaluate the interaction on the resident state.  Serves
 * InteractionBase.compute (interaction.py:57-82) and the accessors of
 * system.py:261-335.  Outputs (each may be NULL): energy, virial (scalars),
 * stress9 (3x3 row-major, sum_{i<j} d_a F_b), forces (N x 3, original
 * order), pe / vir (N; 1/2 sum_j u_ij and 1/2 sum_j r.F — the RUMD per-
 * particle columns of data/ka_N256_rho1.185_rumd.xyz.gz).                   */
int b2_compute(b2_ctx *ctx, int what, double *energy, double *virial,
               double *stress9, double *forces, double *pe, double *vir);

/* K, U, W, sum F^2, momentum, mass, msd — device reductions behind
 * System.temperature / kinetic_energy / potential_energy / pressure
 * (observers.py:247-256).  Recomputes forces if the state changed.          */
int b2_thermo(b2_ctx *ctx, double out[B2_T_LEN]);

/* Full Verlet list of the current configuration in ORIGINAL numbering, each
 * row sorted ascending: count[i], nbr[i*stride + k].  Membership rule:
 * r2 < (rcut_ab + skin)^2; fp64: d - rint(d/L)*L minimum image
 * (particle.py:129-130); fp32: wrapped difference of the 32-bit fixed-point
 * coordinates times float(L/2^32); r2 = (dx*dx + dy*dy) + dz*dz, evaluated in the
 * context's precision.  nbr may be NULL (counts only).  Returns B2_EINVAL if
 * stride < max count (max count is written to *max_count first).            */
int b2_neighbors(b2_ctx *ctx, int32_t *count, int32_t *nbr, int32_t stride,
                 int32_t *max_count);

/* ---- dynamics ----------------------------------------------------------- */

/* `steps` velocity-Verlet NVE steps of size dt, entirely on the device.
 * Replaces LAMMPS.run (lammps.py:217-288, `fix nve` :236-238) and RUMD.run
 * (rumd.py:189-193) as called from Simulation.run_until
 * (simulation/core.py:308-315).  steps == 0 is a no-op (round-off pin,
 * tests/test_lammps.py:123-142).                                            */
int b2_run_nve(b2_ctx *ctx, int64_t steps, double dt);

/* Thermostatted dynamics: what the reference's backends do when
 * `system.thermostat` is set (`fix nvt temp T T relaxation_time`,
 * atooms/backends/lammps.py:229-241; IntegratorNVT, backends/rumd.py:105-118).
 * kind B2_NVT_BERENDSEN: velocities scaled by lambda, lambda^2 = 1 + dt/tau (T0/T - 1);
 * kind B2_NVT_NOSE_HOOVER: leap-frog Nose-Hoover, d xi/dt = (T/T0 - 1)/tau^2,
 * v' = [(1 - xi dt/2) v + dt F/m] / (1 + xi dt/2).  T is the kinetic temperature of the
 * half-step velocities with (N-1) d degrees of freedom (system.py:164-179), taken one
 * step late so that it costs no extra pass.  xi persists between calls; b2_upload resets it. */
#define B2_NVT_BERENDSEN   0
#define B2_NVT_NOSE_HOOVER 1
int b2_run_nvt(b2_ctx *ctx, int64_t steps, double dt, double temperature,
               double relaxation_time, int kind);

/* Minimise the potential energy with `method` for at most `max_steps`
 * iterations; stop early when sum F^2 / N < fn2_tol
 * (optimization/core.py:34-36).  *converged = 1 if the tolerance was met
 * (=> backend.reached_steps = False, lammps.py:350-354).  Velocities are
 * zero on return.  `dt` is the initial FIRE time step / SD step length.
 * A call that stopped on its budget can be continued: if nothing touched the
 * state in between, the next b2_minimize with the same method resumes with the
 * minimiser's internal velocities, step and mixing state (the reference's
 * Optimization drives backends in chunks, optimization/core.py:56, :69-74), so
 * the result does not depend on the chunking.                                */
int b2_minimize(b2_ctx *ctx, int method, int64_t max_steps, double dt,
                double fn2_tol, int64_t *steps_done, int *converged);

/* ---- introspection ------------------------------------------------------ */

int b2_stats(b2_ctx *ctx, int64_t out[B2_S_LEN]);

/* Device time (ms, CUDA events on the context's stream) and kernel-launch
 * count of the last b2_run_nve / b2_minimize / b2_compute call.             */
int b2_last_timing(b2_ctx *ctx, double *ms, int64_t *launches);

/* ---- spatial domain decomposition (one context = one slab of a larger system) --
 *
 * Nothing like it in the reference, whose only multi-process route is
 * `mpirun -n lammps_mpi_tasks` inside the LAMMPS subprocess
 * (atooms/backends/lammps.py:61-62).  These entry points do NO communication:
 * the caller moves the packed buffers between ranks (NCCL send/recv through
 * torch.distributed in atooms_b200/dd.py) and reduces the one float per step.
 * All `_dev` pointers are device pointers of the context's GPU.
 *
 * A species id with B2_GHOST set marks a ghost: a read-only copy of a particle
 * owned by a neighbouring domain.  Ghosts act on their neighbours but are not
 * integrated and are excluded from every total of b2_thermo / b2_compute.
 * Coordinates stay GLOBAL (same box on every rank), so packed positions are
 * meaningful on the receiving side as they are.                             */
#define B2_GHOST 256

/* b2_upload / b2_download with device-resident buffers (same layouts).      */
int b2_upload_device(b2_ctx *ctx, int64_t n, const double *pos_dev, const double *vel_dev,
                     const int32_t *species_dev, const double *mass_dev);
int b2_download_device(b2_ctx *ctx, double *pos_dev, double *vel_dev, int32_t *image_dev);

/* Open (opening_kick = 1: forces + opening half kick) or, after a mid-run
 * refresh re-uploaded the domain, continue (opening_kick = 0: list only) a run
 * of at most `steps` NVE steps.                                             */
int b2_dd_begin(b2_ctx *ctx, int64_t steps, double dt, int opening_kick);
/* Enqueue one fused step for the owned particles.  The device float
 * *vmax2_out_dev receives this rank's max |v|^2 of the step (0 if the step
 * was a no-op because the list is stale or the budget is spent).            */
int b2_dd_step(b2_ctx *ctx, float *vmax2_out_dev);
/* Settle the step with the GLOBAL max |v|^2 (device float): advances the
 * step count and the displacement bound, flips the position buffers, raises
 * the refresh flag — identically on every rank.                             */
int b2_dd_commit(b2_ctx *ctx, const float *gvmax2_dev);
/* Synchronise and report progress.                                          */
int b2_dd_poll(b2_ctx *ctx, int64_t *steps_done, int *need_refresh);
/* Close the run: forces of the final configuration, closing half kick.      */
int b2_dd_end(b2_ctx *ctx);
/* Halo: gather current positions of the particles with the given LOCAL ids
 * (upload order) as 3 doubles each / scatter received ones into the ghosts
 * with local ids first_id .. first_id + n - 1.                              */
int b2_dd_pack(b2_ctx *ctx, const int32_t *ids_dev, int64_t n, double *out_dev);
int b2_dd_unpack(b2_ctx *ctx, int64_t first_id, int64_t n, const double *in_dev);

/* ---- ensembles of small independent systems ("replicas") --------------------------
 *
 * BASELINE.json configs[3]: thousands of small configurations minimised at once
 * (the reference drives ONE system per Optimization, optimization/core.py:39-74,
 * one LAMMPS subprocess per chunk, backends/lammps.py:314-363).  All replicas
 * share the cell and the pair table; each is handled by one thread block with
 * all pairs under the minimum image, the whole minimisation inside one launch.
 * At most 1024 particles per replica.  Errors are reported through
 * b2_last_error(NULL).                                                        */
typedef struct b2_batch b2_batch;
int b2_batch_create(int device, int precision, b2_batch **out);
int b2_batch_destroy(b2_batch *batch);
/* cell sides (centre at the origin) and the pair table of b2_set_pair_table     */
int b2_batch_set_model(b2_batch *batch, const double side[3], int nsp, const double *eps,
                       const double *sigma, const double *rcut, int shifted);
/* pos: nrep x n x 3, species: nrep x n (0-based), mass: nrep x n or NULL (1.0)  */
int b2_batch_upload(b2_batch *batch, int64_t nrep, int n, const double *pos,
                    const int32_t *species, const double *mass);
/* Minimise every replica (method, max_steps, dt, fn2_tol as in b2_minimize).
 * Per-replica outputs, each nrep long or NULL: iterations, converged flag,
 * potential energy, sum F^2 / n, virial of the final configuration.           */
int b2_batch_minimize(b2_batch *batch, int method, int64_t max_steps, double dt, double fn2_tol,
                      int64_t *steps_done, int *converged, double *energy, double *fn2,
                      double *virial);
/* final positions (folded) and forces, nrep x n x 3 each, either may be NULL    */
int b2_batch_download(b2_batch *batch, double *pos, double *forces);
/* NVE for the whole ensemble: velocities (nrep x n x 3), then `steps` velocity-
 * Verlet steps of every replica inside ONE kernel launch (configs[0], LJ N=256
 * 1e4 steps, is launch-latency-bound on the list-based path: ~25 us per step;
 * here a step costs two block barriers).  Per-replica kinetic / potential energy
 * and virial of the final configuration, each nrep long or NULL.                */
int b2_batch_set_velocities(b2_batch *batch, const double *vel);
int b2_batch_run_nve(b2_batch *batch, int64_t steps, double dt, double *kinetic, double *energy,
                     double *virial);
/* velocities (nrep x n x 3) and accumulated box crossings (nrep x n x 3 int32)  */
int b2_batch_get_state(b2_batch *batch, double *vel, int32_t *image);

/* Measurement aid (nothing like it in the reference, whose only timer is the
 * wall-clock TSP of simulation/core.py:286-306): with profiling on,
 * b2_run_nve brackets the step kernels and the list-refresh pipeline with
 * CUDA events on the context's stream.  b2_profile returns, for the last
 * b2_run_nve: out[0] ms in step kernels, out[1] ms in list refreshes,
 * out[2] step-kernel launches, out[3] steps done, out[4] refreshes done.   */
int b2_set_profile(b2_ctx *ctx, int on);
int b2_profile(b2_ctx *ctx, double out[8]);

/* The 'hessian' observable of the reference's protocol (atooms/system/interaction.py:33,76-82: shape
 * (d, N, d, N)): second derivatives of the potential energy of the current configuration,
 *   out[((a * N + i) * 3 + b) * N + j] = d^2 U / dx_ia dx_jb ,
 * from the analytic pair terms, all pairs within the cut-off (minimum image), particles in the caller's
 * numbering.  Dense: 9 N^2 doubles, so n <= B2_HESSIAN_MAX_PARTICLES, else B2_EINVAL; the 12-6 table only
 * (B2_EUNSUPPORTED after b2_set_pair_tabulated).                                                      */
#define B2_HESSIAN_MAX_PARTICLES 4096
int b2_hessian(b2_ctx *ctx, double *out);

/* State that a restart must carry besides positions and velocities (the reference's fallback
 * checkpoint restores the whole System through a trajectory file, simulation/core.py:225-273; its RUMD
 * backend writes the engine's own restart file, backends/rumd.py:179-187):
 *   images     box-crossing counters, 3n int32 by particle (b2_download returns them)
 *   reference  the unfolded rmsd origin set by b2_mark_reference, 3n doubles, absolute coordinates
 *   thermostat the Nose-Hoover friction xi carried between b2_run_nvt calls
 * Set them after b2_upload: run(a); save; restore; run(b) then equals run(a + b), rmsd included.       */
int b2_set_images(b2_ctx *ctx, const int32_t *image);
int b2_get_reference(b2_ctx *ctx, double *ref);
int b2_set_reference(b2_ctx *ctx, const double *ref);
int b2_get_thermostat(b2_ctx *ctx, double *xi);
int b2_set_thermostat(b2_ctx *ctx, double xi);

/* Text formatting of a frame on the host cores, for the trajectory writers.  The reference formats
 * one particle at a time in Python (atooms/trajectory/xyz.py:437-474: seconds per frame at a million
 * particles); this writes the particle lines of its xyz layout, `NAME x y z [vx vy vz]\n` with
 * `precision` decimals, for all n particles into `out` (all host cores), and returns the number of
 * bytes in *written.  names[ids[i]] is the first column.  With out == NULL it only returns in
 * *written a capacity that is enough for these values; B2_EINVAL when `cap` is smaller.          */
int b2_format_xyz(char *out, int64_t cap, int64_t *written, int64_t n, const int32_t *ids, int nnames,
                  const char *const *names, const double *pos, const double *vel, int precision);

/* Host buffers the engine can copy to and from at full PCIe / C2C speed (page-locked memory).
 * The reference's System marshals through ordinary numpy arrays (atooms/system/system.py:398-530);
 * the array-backed System of this package keeps its host mirrors of position / velocity in such
 * buffers, so that the push before and the pull after every backend.run(n) with views handed out
 * (docs/simulations.org:302-308) are asynchronous DMA transfers.  b2_upload / b2_update /
 * b2_download accept any host pointer; these only make them faster.                          */
int b2_host_alloc(void **out, int64_t bytes);
int b2_host_free(void *ptr);

/* No counterpart in the reference.  Kinetic energy, total momentum and total
 * mass of n particles straight from a caller-owned DEVICE array, without a
 * context and without allocating anything: for hosts whose own allocator owns
 * the GPU (PyTorch caching allocator, up-front reservations, private arenas),
 * where a cudaMalloc inside a library fails or stalls the device mid-step.
 * The per-block partial sums that b2_thermo keeps in cudaMalloc'ed memory live
 * here in a fixed 256 KiB __device__ array, which every GPU that loads the
 * library holds until exit.
 *   vel_mass_dev : n x (vx, vy, vz, mass) as float4 (B2_FP32) or double4 (B2_FP64),
 *                  16-byte aligned (it is read with vector loads; a misaligned
 *                  pointer is refused with B2_EINVAL rather than left to fault
 *                  inside the caller's CUDA context)
 *   out_dev      : 8 doubles on the device (8-byte aligned): K, Px, Py, Pz, M, 0, 0, 0
 *   n            : at most B2_STATIC_MAX_PARTICLES, else B2_EINVAL
 * Runs asynchronously on `cuda_stream` of the CURRENT device.  Calls that can
 * overlap on one GPU share the scratch and race: keep them on one stream.  The
 * first call on a GPU resolves the array's device address; issue it at
 * set-up, before any stream capture on that GPU.                            */
#define B2_STATIC_MAX_PARTICLES 16777216
int b2_kinetic_static(const void *vel_mass_dev, int64_t n, int precision,
                      void *cuda_stream, double *out_dev);

#ifdef __cplusplus
}
#endif
#endif /* ATOOMS_B200_H */
