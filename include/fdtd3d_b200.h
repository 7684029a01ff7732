/*
 * fdtd3d_b200 -- C ABI of the B200-native Yee-grid leapfrog time step.
 *
 * This header is the drop-in boundary.  It replaces, for the GPU path, the C++
 * entry points of the reference's Source/Cuda/CudaInterface.{h,cu}; every declaration below cites
 * the reference interface it stands in for (paths relative to the reference repository root).
 * Plain C: POD structs, raw pointers and sizes only -- no C++ or torch types cross this line.
 *
 * Conventions
 *   - Every function returns 0 on success and a non-zero fdtd_status otherwise; the message of the
 *     last failure on the calling thread is available from fdtd_last_error().  The library never
 *     calls exit() (the reference does: Source/Cuda/CudaDefines.h:10-19).
 *   - "FPValue" is float or double (cfg.value_type = 4 or 8; Source/Kernels/FieldValue.h:7-17),
 *     "FieldValue" is FPValue or an interleaved (re,im) pair of FPValue when cfg.complex_values
 *     is set (Source/Kernels/FieldValue.h:22-27).
 *   - Host arrays are in the reference's linear order unless FDTD_ORDER_NATIVE is requested:
 *     3D  index = (i*Ny + j)*Nz + k   (Source/Grid/Grid.cpp:138, z fastest)
 *     2D  index =  i*Ny + j           (Source/Grid/Grid.cpp:82)
 *   - One value per cell per field crosses the boundary (the `cur` layer); after a run
 *     prev == cur in the reference, so that is the complete state.
 */
#ifndef FDTD3D_B200_H
#define FDTD3D_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDTD3D_B200_ABI_VERSION 1

/* Return codes.  FDTD_OK/FDTD_ERROR mirror CudaExitStatus {CUDA_OK, CUDA_ERROR}
 * (Source/Cuda/CudaDefines.h:4-8); the others refine CUDA_ERROR. */
typedef enum fdtd_status
{
  FDTD_OK = 0,
  FDTD_ERROR = 1,            /* CUDA runtime / launch failure (message has the CUDA error string) */
  FDTD_INVALID_ARGUMENT = 2, /* bad config, field id, size or pointer */
  FDTD_NO_DEVICE = 3,        /* no usable CUDA device: there is no CPU fallback */
  FDTD_COMM_ERROR = 4        /* NCCL failure or halo exchange requested without a communicator */
} fdtd_status;

/* Field identifiers.  Order of the E/H/D/B block follows Scheme3D's members
 * (Source/Scheme/Scheme3D.h:21-33). */
typedef enum fdtd_field
{
  FDTD_EX = 0, FDTD_EY = 1, FDTD_EZ = 2,
  FDTD_HX = 3, FDTD_HY = 4, FDTD_HZ = 5,
  FDTD_DX = 6, FDTD_DY = 7, FDTD_DZ = 8,   /* UPML auxiliary fields, only with use_pml */
  FDTD_BX = 9, FDTD_BY = 10, FDTD_BZ = 11,
  FDTD_EPS = 12, FDTD_MU = 13,             /* material grids, size N+1 per axis, FPValue (real part) */
  FDTD_EINC = 14, FDTD_HINC = 15,          /* 1-D incident-wave line, only with use_tfsf */
  FDTD_FIELD_COUNT = 16
} fdtd_field;

typedef enum fdtd_order
{
  FDTD_ORDER_REFERENCE = 0, /* z fastest (3D) / y fastest (2D): Source/Grid/Grid.cpp:82,138 */
  FDTD_ORDER_NATIVE = 1     /* x fastest, unpadded: ((k*Ny)+j)*Nx + i -- the device's own order */
} fdtd_order;

/*
 * Everything the time step needs that is not an array.  It carries what the reference passes to
 * cudaExecute3DSteps / cudaExecute2DTMzSteps (Source/Cuda/CudaInterface.cu:356-368, 20-29: layout,
 * dt, dx, totalStep, processId) plus the Scheme members the reference's CUDA path ignores because
 * it asserts on PML/TFSF (Source/Scheme/Scheme3D.cpp:3346-3349).
 * All FPValue-typed quantities are passed as double holding the exact FPValue value.
 */
typedef struct fdtd_config
{
  int32_t abi_version;     /* FDTD3D_B200_ABI_VERSION */
  int32_t dims;            /* 3: Scheme3D (GRID_3D); 2: SchemeTMz (GRID_2D), fields Ez,Hx,Hy,Dz,Bx,By */
  int32_t value_type;      /* 4: VALUE_TYPE f; 8: VALUE_TYPE d (sizeof FPValue) */
  int32_t complex_values;  /* COMPLEX_FIELD_VALUES */
  int32_t use_pml;         /* --use-pml  (Scheme3D::usePML)  */
  int32_t use_tfsf;        /* --use-tfsf (Scheme3D::useTFSF); when 0 the point source is active */
  int32_t exact_mode;      /* 1: reference evaluation order, IEEE division per cell, no FMA contraction
                              (bit-for-bit with the reference CPU solver in double and real float);
                              0: reciprocal tables + FMA (within the stated tolerance); with use_pml the
                                 cells where all sigma are zero are advanced without their D / B values
                                 (E += dt/dx / (eps eps0) curl), which the library reconstructs whenever D / B
                                 are read -- same tolerance, 18 instead of 30 field values of traffic per cell;
                              2: as 0 but D / B are carried through every cell as the reference does */
  int32_t device;          /* CUDA device ordinal; cudaInit(device) of Source/Cuda/CudaInterface.cu:934 */

  int64_t size[3];         /* --sizex/y/z; size[2] must be 1 for dims == 2 */
  int64_t pml_size[3];     /* --pml-sizex/y/z (YeeGridLayout::leftBorderPML) */
  int64_t tfsf_size[3];    /* --tfsf-sizex/y/z (YeeGridLayout::leftBorderTotalField) */
  int64_t source_pos[3];   /* point-source cell: (Nx/2,Ny/2,Nz/2) in Scheme3D.cpp:2011-2013,
                              (70,Ny/2) in SchemeTMz.cpp:1347 */

  double dx;               /* gridStep */
  double dt;               /* gridTimeStep */
  double source_frequency; /* sourceFrequency */
  double rel_phase_velocity; /* Scheme3D::relPhaseVelocity (Scheme3D.cpp:3401);
                                for dims == 2 the relPhi of SchemeTMz.cpp:35-49 */
  double angle_teta;       /* YeeGridLayout::incidentWaveAngle1, radians */
  double angle_phi;        /* incidentWaveAngle2 */
  double angle_psi;        /* incidentWaveAngle3 */

  /* z-slab decomposition over the GPUs of one box (replaces ParallelGrid's MPI exchange,
   * Source/Grid/ParallelGrid.cpp:1817-1823).  rank r owns planes [r*(Nz/n), ...), the last rank
   * takes the remainder (Source/Layout/ParallelYeeGridLayout.cpp:24-36).  nranks == 1: no exchange. */
  int32_t rank;
  int32_t nranks;

  int64_t time_step_start; /* value of t for the first step of the next fdtd_run (performNSteps' startStep) */
} fdtd_config;

typedef struct fdtd_solver *fdtd_handle;

/* Message of the last failure on this thread ("" if none). */
const char *fdtd_last_error (void);

/* cudaInfo() of Source/Cuda/CudaInterface.cu:939-959: writes a one-line description of `device`
 * (name, SM count, memory, theoretical bandwidth) into buf. */
int fdtd_device_info (int device, char *buf, int64_t buflen);

/* Host-only helpers (no device needed) that derive the time step's scalar inputs the way the reference
 * does.  fdtd_init_scheme: from cfg->dims, value_type, size, dx and the CLI-style wavelength / angles in
 * degrees (--wavelength, --angle-teta/phi/psi) fills dx (rounded to FPValue), dt, source_frequency,
 * rel_phase_velocity, the three angles in radians and the reference's point-source position --
 * Scheme3D::initScheme (Source/Scheme/Scheme3D.cpp:3387-3402), SchemeTMz::initScheme
 * (Source/Scheme/SchemeTMz.cpp:1948-1956, 35-49) and Source/main.cpp:79-85,148.
 * fdtd_pml_sigma_profile: the 1-D conductivity profile Scheme3D::initGrids writes along `axis`
 * (Source/Scheme/Scheme3D.cpp:3656-3818), size[axis]+1 FPValues, to be passed to fdtd_set_sigma. */
int fdtd_init_scheme (fdtd_config *cfg, double wavelength, double teta_deg, double phi_deg, double psi_deg);
int fdtd_pml_sigma_profile (const fdtd_config *cfg, int axis, void *out, int64_t count);

/* Allocates the device-resident state of one Scheme (all fields zero, eps = mu = 1, sigma = 0).
 * Stands in for the cudaMalloc block of cudaExecute3DSteps (Source/Cuda/CudaInterface.cu:420-470),
 * but the handle keeps the memory between calls so the timed loop has no allocation or copy. */
int fdtd_create (const fdtd_config *cfg, fdtd_handle *out);
int fdtd_destroy (fdtd_handle h);

/* 1-D PML conductivity profile along `axis` (0,1,2), length size[axis]+1, FPValue-typed: the
 * profile Scheme3D::initGrids writes into the SigmaX/Y/Z grids (Source/Scheme/Scheme3D.cpp:3656-3818;
 * sigma_x depends on i only, etc.).  Rebuilds the coefficient tables. */
int fdtd_set_sigma (fdtd_handle h, int axis, const void *profile, int64_t count);

/* Host -> device / device -> host copy of one field for global planes [z0, z0+nz) (dims == 2:
 * z0 = 0, nz = 1).  Planes outside this rank's slab are ignored.  Replaces the flatten + cudaMemcpy
 * blocks of Source/Cuda/CudaInterface.cu:492-561 and 816-873.  Host element type: FieldValue for
 * E/H/D/B/EINC/HINC, FPValue for EPS/MU.  For EINC/HINC the whole line is copied (z0, nz ignored). */
int fdtd_upload (fdtd_handle h, int field, const void *host, int order, int64_t z0, int64_t nz);
int fdtd_download (fdtd_handle h, int field, void *host, int order, int64_t z0, int64_t nz);

/* Sets every cell of a field to (re, im).  For FDTD_MU (and, with dims == 2, FDTD_EPS) this also tells the library that
 * the material is homogeneous: the kernels then take the value from here instead of reading the array (same results,
 * less memory traffic).  fdtd_upload of a material scans what it uploaded, so an uploaded array that holds one value
 * everywhere (the reference's own mu = 1) is recognised as well.  fdtd_create fills eps and mu with 1. */
int fdtd_fill (fdtd_handle h, int field, double re, double im);

/* eps := Approximation::approximateSphere (Source/Layout/Approximation.cpp:303-331) evaluated on
 * the device at every material cell, as Scheme3D::initGrids does (Source/Scheme/Scheme3D.cpp:3413-3442). */
int fdtd_init_eps_sphere (fdtd_handle h, double cx, double cy, double cz, double radius, double eps_inside);

/* Runs nsteps leapfrog steps: Scheme3D::performNSteps (Source/Scheme/Scheme3D.cpp:1966-2069) or
 * SchemeTMz::performNSteps (Source/Scheme/SchemeTMz.cpp:1311-1563).  This is what
 * cudaExecute3DSteps' `while (t < totalStep)` loop does (Source/Cuda/CudaInterface.cu:490-814),
 * here with PML, TFSF and the reference's point source.  fdtd_run blocks until the steps are done.
 * fdtd_run_async only enqueues them on the solver's stream and returns; several calls may be queued back to back
 * (chunks of steps with stream-ordered work of the caller's between them), fdtd_sync waits for all of them and
 * fdtd_busy tells without waiting whether queued steps are still running.  fdtd_run_async waits only where it has
 * to: when more source samples are queued than its pinned table holds (65536 complex-double steps between two
 * synchronisations), when a run is longer than any before (the device table grows), and -- like any CUDA launch --
 * when the driver's launch queue is full.  Every other entry point that touches the handle's arrays synchronises
 * first. */
int fdtd_run (fdtd_handle h, int64_t nsteps);
int fdtd_run_async (fdtd_handle h, int64_t nsteps);
int fdtd_sync (fdtd_handle h);
int fdtd_busy (fdtd_handle h, int *busy);

/* Turns the total field inside the TF/SF box into the scattered field, in place, for all six E/H components:
 * the post-processing the reference applies before it saves results with --save-res
 * (Source/Scheme/Scheme3D.cpp:2767-2921): cells whose real coordinate lies in
 * [leftBorderTotalField, rightBorderTotalField] get the projected incident field subtracted.  3D + use_tfsf only.
 * Like the reference's, it mutates the fields: call it after the last fdtd_run, before downloading for output. */
int fdtd_subtract_incident (fdtd_handle h);

/* Device time of the last completed fdtd_run / fdtd_run_async+fdtd_sync, from CUDA events on the
 * solver's stream (milliseconds); when several runs were queued, of the one queued last. */
int fdtd_last_run_ms (fdtd_handle h, float *ms);

/* Per-kernel timing for the roofline report: when enabled, the two half-step kernels of every step of a
 * single-rank run are bracketed by CUDA events on the solver's stream (capacity: the steps of one run).
 * fdtd_phase_times returns, for the last completed run, the summed device time of the E-phase and H-phase
 * kernels (milliseconds) and the number of steps they cover.  The reference's only timing is gettimeofday
 * around performSteps (Source/main.cpp:153-158). */
int fdtd_set_profiling (fdtd_handle h, int enable);
int fdtd_phase_times (fdtd_handle h, float *e_ms, float *h_ms, int64_t *steps);

/* Where the time of a multi-rank step goes (nranks > 1, two half-step kernels, profiling on): summed over the steps of
 * the last completed run, the device time between the nine events the time loop records per step on the compute
 * stream -- ms[0] incident line + face tables (E), [1] waiting for the H halo of the rank below, [2] E boundary plane,
 * [3] E interior, [4] incident line + face tables (H), [5] waiting for the E halo of the rank above, [6] H boundary
 * plane, [7] H interior.  *steps = 0 when no such run was profiled.  No reference counterpart (the reference has no
 * timing inside ParallelGrid::share, Source/Grid/ParallelGrid.cpp:1817-1823). */
int fdtd_step_breakdown (fdtd_handle h, float ms[8], int64_t *steps);

/* Number of kernels of this library launched by this handle since creation (bench.py's gpu_launches). */
int fdtd_launch_count (fdtd_handle h, int64_t *count);

/* Current value of t (number of steps done + time_step_start). */
int fdtd_get_time_step (fdtd_handle h, int64_t *t);

/*
 * The box of cells (global indices, inclusive) that the time loop currently advances without their D / B values
 * (exact_mode 0, use_pml: the region where all three sigma profiles are zero, shrunk to the kernels' tile
 * granularity); lo > hi on some axis when the box is empty or the optimisation is off.  No reference counterpart:
 * a diagnostic (bench.py reports the fraction of the grid it covers).
 */
int fdtd_get_core_box (fdtd_handle h, int64_t lo[3], int64_t hi[3]);

/*
 * One pass per time step in the core (by default only for complex double, where it was measured to pay): instead of an E and
 * an H half step, one kernel reads E and H once and writes them once inside a box that spans the whole x range, the
 * rows of the core box and its planes on this rank (12 instead of 18 field values per cell and step; exact_mode 0,
 * use_pml, homogeneous mu).  Results are bit-identical to the two half-step kernels.  On B200 it is 6 % faster than
 * they are for complex double but slower for real double and float -- the fused instruction stream is latency-bound at
 * the occupancy its registers allow (profiles/r2_one_pass.md) -- hence the default.  fdtd_get_fused_box: the box (global indices,
 * inclusive; lo > hi when the kernel is not in use); fdtd_fused_time: summed device time of that kernel over the last
 * completed run, with fdtd_set_profiling on (fdtd_phase_times then returns the E and H half steps of the remaining
 * shell).  No reference counterpart.
 */
int fdtd_get_fused_box (fdtd_handle h, int64_t lo[3], int64_t hi[3]);
int fdtd_fused_time (fdtd_handle h, float *ms);

/* Tuning aid, no reference counterpart: planes (2D: rows) marched per CTA by the half-step kernels and by the one-pass
 * kernel (0 = the library's rule), and whether the one-pass kernel is used where the geometry allows (0 = never,
 * 1 = always, -1 = the library's default: complex double only).  Results do not depend on any of it. */
int fdtd_set_tuning (fdtd_handle h, int march_planes, int one_pass_march_planes, int allow_one_pass);

/* The same box for a configuration and sigma profiles (FPValue arrays of size[a] + 1 values, NULL = all zero) without
 * creating a handle; host only, no GPU needed (planning, tests).  No reference counterpart. */
int fdtd_plan_core_box (const fdtd_config *cfg, const void *const sigma[3], int64_t lo[3], int64_t hi[3]);

/*
 * TF/SF face membership of every cell of component `field` (FDTD_EX .. FDTD_HZ) for cfg's grid and tfsf_size: the
 * integer boxes the half-step kernels use in place of YeeGridLayout::doNeedTFSFUpdate{Ex..Hz}Border
 * (Source/Layout/YeeGridLayout.cpp:327-809, floating-point predicates).  out[(i*Ny + j)*Nz + k] gets bit d set when the
 * reference's predicate holds for LayoutDirection d (LEFT 0, RIGHT 1, DOWN 2, UP 3, BACK 4, FRONT 5;
 * Source/Layout/GridLayout.h:9-16).  Host only, no GPU needed: this is what pins the index bookkeeping of the TF/SF
 * correction bit-exactly against the reference (tests/test_tfsf_faces_cpu.py).
 */
int fdtd_tfsf_face_masks (const fdtd_config *cfg, int field, uint8_t *out, int64_t count);

/* sum over this rank's cells of |cur|^2 of a field, accumulated in double (diagnostic used by the
 * parity tests at sizes where a full download is impractical; the known-answer vectors of
 * SURVEY.md §8c are quoted in this form). */
int fdtd_field_sumsq (fdtd_handle h, int field, double *out);

/*
 * The same sum of squares for callers that cannot allocate device memory or have no solver handle.
 * (No counterpart in the reference: it has no device-side reductions at all.)
 *
 * Computes sum(x[i]^2, i < count) over `count` float (value_type 4) or double (value_type 8) scalars at the
 * device address dev_in and writes the double result to the device address dev_out; a complex field is its
 * 2*N scalars.  Both pointers must belong to the device that is current in the calling thread.  The work is
 * queued on `cuda_stream` (a cudaStream_t, NULL for the default stream) and the function returns without
 * waiting: read *dev_out after that stream has been synchronised.
 *
 * It never calls cudaMalloc, cudaFree or any allocating routine.  Its scratch (one partial sum per thread
 * block) is a fixed 256 KiB array inside the library image, which every GPU that loads the library keeps
 * until the process ends.  Consequences a caller must know:
 *   - count may not exceed FDTD_SUMSQ_NOALLOC_MAX_COUNT; a larger count is refused with
 *     FDTD_INVALID_ARGUMENT and an explanation in errbuf (nothing is launched) -- split the array;
 *   - all calls on one GPU share that array, so two calls whose kernels may overlap on the GPU (different
 *     streams, or different host threads) race and return garbage: order them on one stream or with events.
 *     Nothing in the array needs clearing between calls: each call rewrites every slot it later reads;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress) and loads
 *     the kernels; make that first call outside, and before, any stream capture.  The address is remembered for
 *     the life of the process and belongs to the device's primary context of that moment: do not call this entry
 *     point again on a device after cudaDeviceReset (the array has moved and the remembered address is stale).
 * errbuf (may be NULL) receives the message of a failure; the return value is an fdtd_status.
 */
#define FDTD_SUMSQ_NOALLOC_MAX_COUNT 268435456LL
int fdtd_sumsq_noalloc (const void *dev_in, int value_type, int64_t count, double *dev_out, void *cuda_stream,
                        char *errbuf, int64_t errbuf_len);

/* Multi-GPU: NCCL communicator for the z-slab halo exchange (one rank per GPU).  Rank 0 calls
 * fdtd_comm_unique_id and distributes the 128 bytes out of band (torch.distributed, a file, ...);
 * then every rank calls fdtd_comm_init with the same id.  Replaces ParallelGridCore's MPI setup
 * (Source/Grid/ParallelGridCore.cpp) and ParallelGrid::share (Source/Grid/ParallelGrid.cpp:1817). */
int fdtd_comm_unique_id (void *id128);
int fdtd_comm_init (fdtd_handle h, const void *id128);

/* Host arrays for the one-shot call below; any pointer may be NULL (field starts at zero and is not
 * copied back; eps/mu default to 1; sigma to 0). */
typedef struct fdtd_host_arrays
{
  void *field[12];        /* indexed by fdtd_field FDTD_EX..FDTD_BZ, FieldValue, reference order, in/out */
  const void *eps;        /* FPValue, (Nx+1)(Ny+1)(Nz+1), reference order */
  const void *mu;
  const void *sigma[3];   /* FPValue 1-D profiles, size[axis]+1 */
  void *einc, *hinc;      /* FieldValue, 100*(Nx+Ny+Nz) cells (Source/Scheme/Scheme3D.h:278-279), in/out */
} fdtd_host_arrays;

/* One-shot drop-in for cudaExecute3DSteps / cudaExecute2DTMzSteps
 * (Source/Cuda/CudaInterface.cu:356-932, 20-339): allocate, upload host arrays, run total_steps,
 * copy everything back, free.  *status receives FDTD_OK / FDTD_ERROR like CudaExitStatus *retval. */
int fdtd_execute_steps (int *status, const fdtd_config *cfg, fdtd_host_arrays *arrays, int64_t total_steps);

#ifdef __cplusplus
}
#endif

#endif /* FDTD3D_B200_H */
