/*
 * Host-callable launch wrappers of the kernel translation units (kernels3d_tma.cu, kernels3d_fused.cu, kernels2d.cu, aux_kernels.cu).
 * Every wrapper launches exactly one kernel on `stream` and returns cudaGetLastError().
 */
#ifndef FDTD_LAUNCH_H
#define FDTD_LAUNCH_H

#include <cuda_runtime.h>
#include <string>
#include "fdtd_device.cuh"

namespace fdtd
{

/* TMA-staged 3D half steps (kernels3d_tma.cu).  TmaMaps3D holds the CUtensorMap descriptors (13 per phase,
 * 128 bytes each) of one solver's arrays; tma_prepare_3d encodes them once the arrays are allocated.
 * `core` selects the variants that skip D / B inside the box p.core_lo .. p.core_hi (fast mode with UPML only).
 * The CORE and EXACT variants use another tile height than the plain fast ones (kernels3d_tma.cu, TileRows): `alt`
 * = exact || core tells tma_prepare_3d / tma_tile_geometry which; the maps must be encoded for the variant launched.
 * `uniform_mu`: the mu array holds p.mu_const everywhere; the H kernels then do not load it. */
struct TmaMaps3D
{
  alignas (64) unsigned char e[13 * 128];
  alignas (64) unsigned char h[13 * 128];
  bool valid;
};
void tma_tile_geometry (int value_type, int nc, bool alt, int *tile_cells_x, int *tile_rows, int *quarter_cells_x,
                        int *ctas_per_sm);
bool tma_prepare_3d (TmaMaps3D *out, int value_type, int nc, bool pml, bool alt, void *const f[12], void *eps, void *mu,
                     int pitch, int ny, int planes, int pitchM, int planesM, std::string *err);
template <typename T>
cudaError_t launch_phase_3d_tma (bool e_phase, const StepParams<T> &p, const TmaMaps3D &maps, bool cplx, bool pml, bool tfsf,
                                 bool exact, bool core, bool uniform_mu, int kb, int ke, cudaStream_t stream);

/*
 * One-pass core kernel (kernels3d_fused.cu): E and H of one whole time step inside a box of the slab, reading the
 * `cur` buffers (p.f) and writing the `nxt` buffers (p.fo) of the E / H ping-pong pairs.  FusedBox: rows [fy0, fy1]
 * (a whole number of the kernel's tile rows, inside the y core), local planes [fz0, fz1] (inside the z core), kc planes
 * marched per CTA, nq = the largest number of quarter rows of one x tile that lie in the x absorber (sizes the D / B
 * part of a stage).  The box always spans the whole x range.  FusedMaps3D: the 13 tensor maps for one buffer parity.
 */
struct FusedBox
{
  int fy0, fy1, fz0, fz1, kc, nq;
  /* D of the absorber quarters.  A rim thread recomputes E' of a neighbour tile's cell from that cell's OLD D, which
   * the neighbour CTA may already have overwritten in place -- so, like E and H, the D values of the absorber quarters
   * inside the box live in a ping-pong pair while the kernel is in use: compact arrays that hold only those quarters
   * (x extent nqt quarter rows; quarter gq of the grid is compact quarter gq if gq < nlo, else nlo + gq - hq0), same
   * row / plane indexing as the field arrays.  The kernel reads the pair member its tensor maps describe and writes
   * dq_out; the solver gathers / scatters them from / to the D arrays when the kernel comes into / goes out of use. */
  int nlo, hq0, nqt;
  long long dq_pitch;   /* cells per row of a compact array = nqt * cells per quarter */
  void *dq_out[3];
};
/* copies the D values of the absorber quarters of the box between the field arrays and a compact array (to_dq = gather) */
cudaError_t launch_dq_copy (void *const main_d[3], void *const dq[3], int ncomp, bool to_dq, const FusedBox &fb, int cell_bytes,
                            int ny, int pitch, long long ps, cudaStream_t stream);
struct FusedMaps3D
{
  alignas (64) unsigned char set[13 * 128];
  bool valid;
};
void fused_tile_geometry (int value_type, int nc, int *tile_cells_x, int *tile_rows, int *ctas_per_sm, int *smem_per_quarter,
                          int *smem_base);
/* e, h: the E / H buffers being read; d: the compact D arrays being read (x extent dq_pitch cells); b: the B arrays */
bool fused_prepare_3d (FusedMaps3D *out, int value_type, int nc, void *const e[3], void *const h[3], void *const d[3],
                       long long dq_pitch, void *const b[3], void *eps, int pitch, int ny, int planes, int pitchM, int planesM,
                       std::string *err);
template <typename T>
cudaError_t launch_step_core_fused (const StepParams<T> &p, const FusedMaps3D &maps, const FusedBox &fb, bool cplx, bool tfsf,
                                    cudaStream_t stream);

/* The TMz counterpart (kernels2d_fused.cu): rows [fy0, fy1] of the 2D grid, kc tiles marched per CTA; fz0 / fz1 unused. */
void fused_tile_geometry_tmz (int value_type, int nc, int *tile_cells_x, int *tile_rows, int *smem_per_quarter, int *smem_base);
bool fused_prepare_tmz (FusedMaps3D *out, int value_type, int nc, void *ez, void *hx, void *hy, void *dz, long long dq_pitch,
                        void *bx, void *by, void *eps, int pitch, int ny, int pitchM, std::string *err);
template <typename T>
cudaError_t launch_step_core_tmz_fused (const StepParams<T> &p, const FusedMaps3D &maps, const FusedBox &fb, bool cplx, bool tfsf,
                                        bool uniform_eps, cudaStream_t stream);

/* D (B) -= / += eps eps0 E (mu mu0 H) over the core box, for the six components (aux_kernels.cu): `fold` before
 * running the CORE variants, unfold before anything else reads or replaces D / B / E / H / eps / mu */
template <typename T>
cudaError_t launch_fold_aux (const StepParams<T> &p, bool tmz, bool cplx, bool tfsf, bool fold, cudaStream_t stream);

/* 2D TMz half steps (kernels2d.cu) */
template <typename T>
cudaError_t launch_phase_tmz (bool e_phase, const StepParams<T> &p, bool cplx, bool pml, bool tfsf, bool exact, bool core,
                              bool uniform_material, cudaStream_t stream);

/* 1-D incident line: E update + source sample at cell 0, H update (aux_kernels.cu);
 * Scheme3D::performPlaneWaveESteps / HSteps, Scheme3D.cpp:24-82 */
template <typename T>
cudaError_t launch_incident (bool e_phase, const StepParams<T> &p, bool cplx, bool exact, cudaStream_t stream);

/* incident values of all TF/SF face cells of one half step into p.face_tab (24 tables of p.face_stride cells:
 * kernels3d_common.cuh, tfsf_curl_tab); run after launch_incident of the same half step, before the TMA half-step
 * kernel */
template <typename T>
cudaError_t launch_face_tables (bool e_phase, const StepParams<T> &p, bool cplx, cudaStream_t stream);

/* total -> scattered field inside the TF/SF box for one component (0..2 = Ex..Ez, 3..5 = Hx..Hz), in place:
 * the post-processing of Scheme3D.cpp:2767-2921 */
template <typename T>
cudaError_t launch_scattered (const StepParams<T> &p, bool cplx, int comp, cudaStream_t stream);

/*
 * Layout conversion between a staged host-order block and the device layout.
 * host-order block: cells (i in [ia, ib), j in [0, ny), k in [0, nzb)) stored as ((i-ia)*ny + j)*nzb + k
 * (reference order restricted to an i-range of a buffer that covers global planes [z0, z0+nzb));
 * device: ((k + kofs)*ny + j)*pitch + i, `cellBytes` bytes per cell.  to_device: scatter into the device
 * layout; otherwise gather from it.  Planes with k + kofs outside [kmin, kmax) are skipped.
 */
cudaError_t launch_reorder (bool to_device, void *dev, void *staged, int cellBytes, int ia, int ib, int ny, int nzb,
                            int pitch, long long ps, int kofs, int kmin, int kmax, cudaStream_t stream);

/* fill `count` cells of `cellBytes` (4, 8 or 16) with a 16-byte pattern */
cudaError_t launch_fill (void *dev, long long count, int cellBytes, const void *pattern16, cudaStream_t stream);

/* Approximation::approximateSphere over the material grid (Approximation.cpp:303-331) */
template <typename T>
cudaError_t launch_eps_sphere (T *eps, int mx, int my, int nzm, int kz0, int pitchM, long long psM, T cx, T cy, T cz,
                               T radius, T eps_in, cudaStream_t stream);

/*
 * sum of squares of `nscal` scalars per row over rows of a pitched array, accumulated in double.
 * Two kernels: per-CTA partial sums into `partials` (scratch, >= sumsq_partial_count() doubles), then a
 * single-CTA reduction into *out.
 */
int sumsq_partial_count ();
template <typename T>
cudaError_t launch_sumsq (const T *base, int nscal, long long nrows, long long pitch_scal, double *partials,
                          double *out, cudaStream_t stream);

/* out2[0] = 1 if the rows of a pitched array do not all hold the value base[0], else 0; out2[1] = base[0] (device doubles) */
template <typename T>
cudaError_t launch_uniform_check (const T *base, int nscal, long long nrows, long long pitch_scal, double *out2, cudaStream_t stream);

} /* namespace fdtd */

#endif /* FDTD_LAUNCH_H */
