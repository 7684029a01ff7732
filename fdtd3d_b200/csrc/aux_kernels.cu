/*
 * Small kernels around the two half-step kernels: the 1-D incident-wave line,
 * host-order <-> device-layout conversion, fills, the eps sphere and the |field|^2 reduction.
 * Compiled with -fmad=false like the rest (exact mode relies on it).
 */
#include "fdtd_device.cuh"
#include "launch.h"

namespace fdtd
{

namespace
{

/* ------------------------------------------------------------------ 1-D incident line */

/* Scheme3D::performPlaneWaveESteps, Scheme3D.cpp:24-59: E[i] = E[i] + ce * (H[i-1] - H[i]) for i >= 1,
 * E[0] = source sample of this step.  In place: the sweep reads only H and the cell's own E. */
template <typename T, int NC, bool EXACT>
__global__ void incident_e (const StepParams<T> p)
{
  const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.ninc) return;
  if (i == 0)
  {
    const long long s = current_step (p);
#pragma unroll
    for (int c = 0; c < NC; ++c) p.einc[c] = p.src[s * NC + c];
    return;
  }
#pragma unroll
  for (int c = 0; c < NC; ++c)
  {
    const T e = p.einc[i * NC + c];
    const T d = p.hinc[(i - 1) * NC + c] - p.hinc[i * NC + c];
    p.einc[i * NC + c] = EXACT ? e + p.inc_ce * d : fma_t<T> (p.inc_ce, d, e);
  }
}

/* Scheme3D::performPlaneWaveHSteps, Scheme3D.cpp:61-82: H[i] = H[i] + ch * (E[i] - E[i+1]), i < n-1 */
template <typename T, int NC, bool EXACT>
__global__ void incident_h (const StepParams<T> p)
{
  const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.ninc - 1) return;
#pragma unroll
  for (int c = 0; c < NC; ++c)
  {
    const T h = p.hinc[i * NC + c];
    const T d = p.einc[i * NC + c] - p.einc[(i + 1) * NC + c];
    p.hinc[i * NC + c] = EXACT ? h + p.inc_ch * d : fma_t<T> (p.inc_ch, d, h);
  }
}

/* ------------------------------------------------------------------ layout conversion */

/* 32x32 tiled transpose between the k-fastest staged block and the i-fastest device layout, one j per
 * blockIdx.y.  CellT = uint32_t / uint64_t / int4. */
template <typename CellT, bool TO_DEVICE>
__global__ void reorder_kernel (CellT *dev, CellT *staged, int ia, int ib, int ny, int nzb, int pitch, long long ps,
                                int kofs, int kmin, int kmax)
{
  __shared__ CellT tile[32][33];
  const int j = blockIdx.y;
  const int kt = blockIdx.x * 32, it = ia + blockIdx.z * 32;
  if (TO_DEVICE)
  {
    for (int r = threadIdx.y; r < 32; r += blockDim.y)
    {
      const int i = it + r, k = kt + threadIdx.x;
      if (i < ib && k < nzb) tile[r][threadIdx.x] = staged[((long long) (i - ia) * ny + j) * nzb + k];
    }
    __syncthreads ();
    for (int r = threadIdx.y; r < 32; r += blockDim.y)
    {
      const int k = kt + r, i = it + threadIdx.x;
      const int kd = k + kofs;
      if (i < ib && k < nzb && kd >= kmin && kd < kmax) dev[((long long) kd * ny + j) * pitch + i] = tile[threadIdx.x][r];
    }
  }
  else
  {
    for (int r = threadIdx.y; r < 32; r += blockDim.y)
    {
      const int k = kt + r, i = it + threadIdx.x;
      const int kd = k + kofs;
      if (i < ib && k < nzb && kd >= kmin && kd < kmax) tile[threadIdx.x][r] = dev[((long long) kd * ny + j) * pitch + i];
    }
    __syncthreads ();
    for (int r = threadIdx.y; r < 32; r += blockDim.y)
    {
      const int i = it + r, k = kt + threadIdx.x;
      const int kd = k + kofs;
      if (i < ib && k < nzb && kd >= kmin && kd < kmax) staged[((long long) (i - ia) * ny + j) * nzb + k] = tile[r][threadIdx.x];
    }
  }
}

template <typename CellT>
__global__ void fill_kernel (CellT *dev, long long count, CellT value)
{
  for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long) gridDim.x * blockDim.x)
    dev[i] = value;
}

/* Approximation::approximateSphere, Approximation.cpp:303-331, at material cell real coordinate m + 0.5
 * (YeeGridLayout.h:267-270); sqrt is the C double function there, rounded once to FPValue. */
template <typename T>
__global__ void eps_sphere_kernel (T *eps, int mx, int my, int nzm, int kz0, int pitchM, long long psM, T cx, T cy, T cz,
                                   T radius, T eps_in)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  if (i >= mx) return;
  for (int kl = blockIdx.z; kl < nzm; kl += gridDim.z)
  {
    const T px = (T) i + (T) 0.5, py = (T) j + (T) 0.5, pz = (T) (kz0 + kl) + (T) 0.5;
    const T dxv = px - cx, dyv = py - cy, dzv = pz - cz;
    const T sum = dxv * dxv + dyv * dyv + dzv * dzv;
    const T d = (T) sqrt ((double) sum);
    const T diff = d - radius;
    T v;
    if (diff < (T) -0.5) v = eps_in;
    else if (diff > (T) 0.5) v = (T) 1;
    else
    {
      const T prop = (T) 0.5 - diff;
      v = prop * eps_in + ((T) 1 - prop) * (T) 1;
    }
    eps[(long long) (kl + 1) * psM + (long long) j * pitchM + i] = v;
  }
}

/* ------------------------------------------------------------------ scattered field at save time */

/* Scheme3D.cpp:2767-2921 (the block run before results are saved with --save-res): every cell of an E or H
 * component whose real coordinate lies inside the TF/SF box [L, R] (inclusive) gets the projected incident field
 * subtracted in place.  One thread per cell of one component; comp = 0..2 (E) or 3..5 (H). */
template <typename T, int NC>
__global__ void scattered_kernel (const __grid_constant__ StepParams<T> p, int comp)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  if (i >= p.nx) return;
  const bool isE = comp < 3;
  const int a = comp % 3;
  const T off[3] = { isE ? (a == 0 ? (T) 1 : (T) 0.5) : (a == 0 ? (T) 0.5 : (T) 1),
                     isE ? (a == 1 ? (T) 1 : (T) 0.5) : (a == 1 ? (T) 0.5 : (T) 1),
                     isE ? (a == 2 ? (T) 1 : (T) 0.5) : (a == 2 ? (T) 0.5 : (T) 1) };
  const T rx = (T) i + off[0], ry = (T) j + off[1];
  if (rx < (T) p.L[0] || rx > (T) p.R[0] || ry < (T) p.L[1] || ry > (T) p.R[1]) return;
  T *F = p.f[comp];
  for (int kl = blockIdx.z; kl < p.nzl; kl += gridDim.z)
  {
    const int gk = p.kz0 + kl;
    const T rz = (T) gk + off[2];
    if (rz < (T) p.L[2] || rz > (T) p.R[2]) continue;
    T inc[NC];
    incident_sample<T, NC> (p, isE ? p.einc : p.hinc, isE ? (T) 0 : (T) 0.5, isE ? p.proj_e[a] : p.proj_h[a], i, j, gk,
                            off[0], off[1], off[2], inc);
    const long long o = (((long long) (kl + 1) * p.ny + j) * p.pitch + i) * NC;
#pragma unroll
    for (int c = 0; c < NC; ++c) F[o + c] = F[o + c] - inc[c];
  }
}

/* ------------------------------------------------------------------ TF/SF face tables */

/*
 * Incident values of every TF/SF face cell for one half step (layout: kernels3d_common.cuh, tfsf_curl_tab), so that
 * the half-step kernels apply a correction with one load.  One thread per in-face position (u, v) of one of the
 * twelve tables of the phase; the face boxes are those of tfsf_e_faces / tfsf_h_faces and the sample is the
 * incident_sample call of tfsf_curl (Scheme3D::calculate{Ex..Hz}TFSF), bit for bit.
 */
template <typename T, int NC, bool IS_E>
__global__ void face_tables_kernel (const __grid_constant__ StepParams<T> p)
{
  const int t = blockIdx.z, side = t & 1, pass = (t >> 1) & 1, a = t >> 2;
  const int b = (a + 1) % 3, c = (a + 2) % 3;
  const int n = pass == 0 ? b : c;  /* face normal */
  const int fc = pass == 0 ? c : b; /* corrected component of the other family = the remaining in-face axis */
  const int ua = n == 0 ? 1 : 0, va = n == 2 ? 1 : 2;
  const int ext[3] = { p.nx, p.ny, p.nzl };
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= ext[ua]) return;
  const bool low = side == 0;
  int pos[3];
  pos[n] = low ? p.L[n] - 1 : (IS_E ? p.R[n] : p.R[n] - 1);
  if (n == 2 && (pos[2] < p.kz0 || pos[2] >= p.kz0 + p.nzl)) return; /* that face plane belongs to another rank */
  const int kl_face = pos[2] - p.kz0;
  for (int v = blockIdx.y; v < ext[va]; v += gridDim.y)
  {
    pos[ua] = u;
    pos[va] = va == 2 ? p.kz0 + v : v;
    /* own axis and remaining axis of the face box */
    const int own_lo = IS_E ? p.L[a] - 1 : p.L[a], own_hi = p.R[a] - 1;
    const int rem_lo = IS_E ? p.L[fc] : p.L[fc] - 1, rem_hi = p.R[fc] - 1;
    if (pos[a] < own_lo || pos[a] > own_hi || pos[fc] < rem_lo || pos[fc] > rem_hi) continue;
    int sp[3] = { pos[0], pos[1], pos[2] };
    if (!low) sp[n] += IS_E ? -1 : 1;
    T off[3];
    for (int ax = 0; ax < 3; ++ax) off[ax] = IS_E ? (ax == fc ? (T) 0.5 : (T) 1) : (ax == fc ? (T) 1 : (T) 0.5);
    T inc[NC];
    incident_sample<T, NC> (p, IS_E ? p.hinc : p.einc, IS_E ? (T) 0.5 : (T) 0, IS_E ? p.proj_h[fc] : p.proj_e[fc],
                            sp[0], sp[1], sp[2], off[0], off[1], off[2], inc);
    const long long idx = (long long) v * ext[ua] + u;
    (void) kl_face;
#pragma unroll
    for (int cc = 0; cc < NC; ++cc) p.face_tab[(((IS_E ? 0 : 12) + t) * p.face_stride + idx) * NC + cc] = inc[cc];
  }
}

/* ------------------------------------------------------------------ sigma-free core: D/B <-> offsets */

/*
 * Where sigma = 0 the UPML recurrences (Kernels.h:13-29, 80-96 with Ca = 1) keep D - eps eps0 E (B - mu mu0 H)
 * constant in time, so the CORE variants of the half-step kernels never touch D/B there.  Between runs the core
 * cells of the D/B arrays hold that constant: fold (sign -1) stores D - m E before a run, unfold (sign +1) turns it
 * back into D = const + m E with the E of that moment.  m is the averaged material of YeeGridLayout.h:1005-1263 as
 * the half-step kernels form it.  The point-source cell of Dz is not folded (its Ez is overwritten every step).
 */
template <typename T, int NC>
__global__ void fold_aux_kernel (const __grid_constant__ StepParams<T> p, T sign, bool skip_source)
{
  const int i = p.core_lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  const int j = p.core_lo[1] + blockIdx.y;
  if (i > p.core_hi[0] || j > p.core_hi[1]) return;
  const T two = (T) 2;
  for (int gk = p.core_lo[2] + blockIdx.z; gk <= p.core_hi[2]; gk += gridDim.z)
  {
    const int kl = gk - p.kz0;
    if (kl < 0 || kl >= p.nzl) continue;
    const long long om = ((long long) (kl + 1) * (p.ny + 1) + j) * p.pitchM + i;
    const long long rowM = p.pitchM, plM = p.psM;
    const T e000 = p.eps[om], e100 = p.eps[om + 1], e010 = p.eps[om + rowM], e001 = p.eps[om + plM];
    const T m000 = p.mu[om], m100 = p.mu[om + 1], m010 = p.mu[om + rowM], m001 = p.mu[om + plM];
    const T m110 = p.mu[om + rowM + 1], m101 = p.mu[om + plM + 1], m011 = p.mu[om + plM + rowM];
    T mod[6];
    mod[0] = ((e000 + e100) / two) * p.eps0;
    mod[1] = ((e000 + e010) / two) * p.eps0;
    mod[2] = ((e000 + e001) / two) * p.eps0;
    mod[3] = (((m000 + m001) / two + (m010 + m011) / two) / two) * p.mu0;
    mod[4] = (((m000 + m001) / two + (m100 + m101) / two) / two) * p.mu0;
    mod[5] = (((m000 + m010) / two + (m100 + m110) / two) / two) * p.mu0;
    const long long o = (((long long) (kl + 1) * p.ny + j) * p.pitch + i) * NC;
    const bool src_cell = skip_source && i == p.srcpos[0] && j == p.srcpos[1] && gk == p.srcpos[2];
#pragma unroll
    for (int comp = 0; comp < 6; ++comp)
    {
      if (comp == 2 && src_cell) continue;
      const T m = sign * mod[comp];
#pragma unroll
      for (int c = 0; c < NC; ++c) p.f[6 + comp][o + c] = fma_t<T> (m, p.f[comp][o + c], p.f[6 + comp][o + c]);
    }
  }
}

/* TMz: Dz <-> eps[i,j] eps0 Ez (single cell, SchemeTMz.cpp:181,326,528-530); Bx <-> (mu[i,j+1] + mu[i,j]) / 2 mu0 Hx
 * (cpp:606-612); By <-> (mu[i,j] + mu[i+1,j]) / 2 mu0 Hy (cpp:975-981) -- as kernels2d.cu forms them */
template <typename T, int NC>
__global__ void fold_aux_tmz_kernel (const __grid_constant__ StepParams<T> p, T sign, bool skip_source)
{
  const int i = p.core_lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  if (i > p.core_hi[0]) return;
  const T two = (T) 2;
  for (int j = p.core_lo[1] + blockIdx.y; j <= p.core_hi[1]; j += gridDim.y)
  {
    const long long om = p.psM + (long long) j * p.pitchM + i;
    const T m00 = p.mu[om], m10 = p.mu[om + 1], m01 = p.mu[om + p.pitchM];
    const T mod_ez = p.eps[om] * p.eps0;
    const T mod_hx = ((m01 + m00) / two) * p.mu0;
    const T mod_hy = ((m00 + m10) / two) * p.mu0;
    const long long o = (p.ps + (long long) j * p.pitch + i) * NC;
    const bool src_cell = skip_source && i == p.srcpos[0] && j == p.srcpos[1];
#pragma unroll
    for (int c = 0; c < NC; ++c)
    {
      if (!src_cell) p.f[8][o + c] = fma_t<T> (sign * mod_ez, p.f[2][o + c], p.f[8][o + c]);
      p.f[9][o + c] = fma_t<T> (sign * mod_hx, p.f[3][o + c], p.f[9][o + c]);
      p.f[10][o + c] = fma_t<T> (sign * mod_hy, p.f[4][o + c], p.f[10][o + c]);
    }
  }
}

/* ------------------------------------------------------------------ reduction */

constexpr int SUMSQ_CTAS = 1184; /* 148 SMs x 8 */
constexpr int SUMSQ_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__ (SUMSQ_THREADS)
sumsq_partial (const T *base, int nscal, long long nrows, long long pitch_scal, double *partials)
{
  __shared__ double red[SUMSQ_THREADS / 32];
  double acc = 0.0;
  for (long long r = blockIdx.x; r < nrows; r += gridDim.x)
  {
    const T *rowp = base + r * pitch_scal;
    for (int s = threadIdx.x; s < nscal; s += SUMSQ_THREADS)
    {
      const double v = (double) rowp[s];
      acc += v * v;
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_down_sync (0xffffffffu, acc, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads ();
  if (threadIdx.x == 0)
  {
    double t = 0.0;
    for (int w = 0; w < SUMSQ_THREADS / 32; ++w) t += red[w];
    partials[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__ (SUMSQ_THREADS)
sumsq_final (const double *partials, int n, double *out)
{
  __shared__ double red[SUMSQ_THREADS];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += SUMSQ_THREADS) acc += partials[i];
  red[threadIdx.x] = acc;
  __syncthreads ();
  for (int s = SUMSQ_THREADS / 2; s > 0; s >>= 1)
  {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads ();
  }
  if (threadIdx.x == 0) *out = red[0];
}

/* does a pitched array hold one value everywhere?  out[0] = 1 if some value differs from the first one (else left at
 * the 0 the caller put there), out[1] = the first value */
template <typename T>
__global__ void __launch_bounds__ (SUMSQ_THREADS)
uniform_check (const T *base, int nscal, long long nrows, long long pitch_scal, double *out)
{
  const T v0 = base[0];
  bool differs = false;
  for (long long r = blockIdx.x; r < nrows; r += gridDim.x)
  {
    const T *rowp = base + r * pitch_scal;
    for (int s = threadIdx.x; s < nscal; s += SUMSQ_THREADS) differs |= rowp[s] != v0;
  }
  if (__syncthreads_or (differs) && threadIdx.x == 0) out[0] = 1.0;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[1] = (double) v0;
}

/* D of the absorber quarters inside the one-pass box <-> compact ping-pong array (launch.h, FusedBox): one thread per
 * 16-byte vector; a quarter row is 128 bytes = 8 vectors */
__global__ void dq_copy_kernel (int4 *main_d, int4 *dq, bool to_dq, FusedBox fb, int ny, long long pitch_v, long long ps_v)
{
  const int rows = fb.fy1 - fb.fy0 + 1;
  const long long per_plane = (long long) rows * fb.nqt * 8;
  const long long total = per_plane * (fb.fz1 - fb.fz0 + 1);
  for (long long t = (long long) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long) gridDim.x * blockDim.x)
  {
    const int k = fb.fz0 + (int) (t / per_plane);
    long long r = t % per_plane;
    const int j = fb.fy0 + (int) (r / (fb.nqt * 8));
    const int c = (int) (r % (fb.nqt * 8));
    const int cq = c >> 3, lane = c & 7;
    const int gq = cq < fb.nlo ? cq : cq - fb.nlo + fb.hq0;
    const long long om = ((long long) (k + 1) * ny + j) * pitch_v + (long long) gq * 8 + lane;
    const long long oq = ((long long) (k + 1) * ny + j) * (fb.nqt * 8LL) + c;
    if (to_dq) dq[oq] = main_d[om];
    else main_d[om] = dq[oq];
  }
  (void) ps_v;
}

template <typename CellT>
cudaError_t reorder_t (bool to_device, void *dev, void *staged, int ia, int ib, int ny, int nzb, int pitch, long long ps,
                       int kofs, int kmin, int kmax, cudaStream_t stream)
{
  dim3 block (32, 8);
  dim3 grid ((nzb + 31) / 32, ny, (ib - ia + 31) / 32);
  if (to_device)
    reorder_kernel<CellT, true><<<grid, block, 0, stream>>> ((CellT *) dev, (CellT *) staged, ia, ib, ny, nzb, pitch, ps, kofs, kmin, kmax);
  else
    reorder_kernel<CellT, false><<<grid, block, 0, stream>>> ((CellT *) dev, (CellT *) staged, ia, ib, ny, nzb, pitch, ps, kofs, kmin, kmax);
  return cudaGetLastError ();
}

} /* namespace */

template <typename T>
cudaError_t launch_incident (bool e_phase, const StepParams<T> &p, bool cplx, bool exact, cudaStream_t stream)
{
  const int threads = 256;
  const int blocks = (int) ((p.ninc + threads - 1) / threads);
  const int key = (e_phase ? 4 : 0) | (cplx ? 2 : 0) | (exact ? 1 : 0);
  switch (key)
  {
    case 0: incident_h<T, 1, false><<<blocks, threads, 0, stream>>> (p); break;
    case 1: incident_h<T, 1, true><<<blocks, threads, 0, stream>>> (p); break;
    case 2: incident_h<T, 2, false><<<blocks, threads, 0, stream>>> (p); break;
    case 3: incident_h<T, 2, true><<<blocks, threads, 0, stream>>> (p); break;
    case 4: incident_e<T, 1, false><<<blocks, threads, 0, stream>>> (p); break;
    case 5: incident_e<T, 1, true><<<blocks, threads, 0, stream>>> (p); break;
    case 6: incident_e<T, 2, false><<<blocks, threads, 0, stream>>> (p); break;
    default: incident_e<T, 2, true><<<blocks, threads, 0, stream>>> (p); break;
  }
  return cudaGetLastError ();
}
template cudaError_t launch_incident<float> (bool, const StepParams<float> &, bool, bool, cudaStream_t);
template cudaError_t launch_incident<double> (bool, const StepParams<double> &, bool, bool, cudaStream_t);

template <typename T>
cudaError_t launch_scattered (const StepParams<T> &p, bool cplx, int comp, cudaStream_t stream)
{
  dim3 block (128);
  dim3 grid ((p.nx + 127) / 128, p.ny, p.nzl < 64 ? p.nzl : 64);
  if (cplx) scattered_kernel<T, 2><<<grid, block, 0, stream>>> (p, comp);
  else scattered_kernel<T, 1><<<grid, block, 0, stream>>> (p, comp);
  return cudaGetLastError ();
}
template cudaError_t launch_scattered<float> (const StepParams<float> &, bool, int, cudaStream_t);
template cudaError_t launch_scattered<double> (const StepParams<double> &, bool, int, cudaStream_t);

template <typename T>
cudaError_t launch_face_tables (bool e_phase, const StepParams<T> &p, bool cplx, cudaStream_t stream)
{
  const int umax = p.nx > p.ny ? p.nx : p.ny, vmax = p.ny > p.nzl ? p.ny : p.nzl;
  dim3 block (128);
  dim3 grid ((umax + 127) / 128, vmax < 4096 ? vmax : 4096, 12);
  if (e_phase)
  {
    if (cplx) face_tables_kernel<T, 2, true><<<grid, block, 0, stream>>> (p);
    else face_tables_kernel<T, 1, true><<<grid, block, 0, stream>>> (p);
  }
  else
  {
    if (cplx) face_tables_kernel<T, 2, false><<<grid, block, 0, stream>>> (p);
    else face_tables_kernel<T, 1, false><<<grid, block, 0, stream>>> (p);
  }
  return cudaGetLastError ();
}
template cudaError_t launch_face_tables<float> (bool, const StepParams<float> &, bool, cudaStream_t);
template cudaError_t launch_face_tables<double> (bool, const StepParams<double> &, bool, cudaStream_t);

template <typename T>
cudaError_t launch_fold_aux (const StepParams<T> &p, bool tmz, bool cplx, bool tfsf, bool fold, cudaStream_t stream)
{
  if (p.core_lo[0] > p.core_hi[0] || p.core_lo[1] > p.core_hi[1] || p.core_lo[2] > p.core_hi[2]) return cudaSuccess;
  const int cx = p.core_hi[0] - p.core_lo[0] + 1, cy = p.core_hi[1] - p.core_lo[1] + 1, cz = p.core_hi[2] - p.core_lo[2] + 1;
  dim3 block (128);
  const T sign = fold ? (T) -1 : (T) 1;
  if (tmz)
  {
    dim3 grid ((cx + 127) / 128, cy < 16384 ? cy : 16384);
    if (cplx) fold_aux_tmz_kernel<T, 2><<<grid, block, 0, stream>>> (p, sign, !tfsf);
    else fold_aux_tmz_kernel<T, 1><<<grid, block, 0, stream>>> (p, sign, !tfsf);
    return cudaGetLastError ();
  }
  dim3 grid ((cx + 127) / 128, cy, cz < 64 ? cz : 64);
  if (cplx) fold_aux_kernel<T, 2><<<grid, block, 0, stream>>> (p, sign, !tfsf);
  else fold_aux_kernel<T, 1><<<grid, block, 0, stream>>> (p, sign, !tfsf);
  return cudaGetLastError ();
}
template cudaError_t launch_fold_aux<float> (const StepParams<float> &, bool, bool, bool, bool, cudaStream_t);
template cudaError_t launch_fold_aux<double> (const StepParams<double> &, bool, bool, bool, bool, cudaStream_t);

cudaError_t launch_reorder (bool to_device, void *dev, void *staged, int cellBytes, int ia, int ib, int ny, int nzb,
                            int pitch, long long ps, int kofs, int kmin, int kmax, cudaStream_t stream)
{
  switch (cellBytes)
  {
    case 4: return reorder_t<uint32_t> (to_device, dev, staged, ia, ib, ny, nzb, pitch, ps, kofs, kmin, kmax, stream);
    case 8: return reorder_t<unsigned long long> (to_device, dev, staged, ia, ib, ny, nzb, pitch, ps, kofs, kmin, kmax, stream);
    case 16: return reorder_t<int4> (to_device, dev, staged, ia, ib, ny, nzb, pitch, ps, kofs, kmin, kmax, stream);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_dq_copy (void *const main_d[3], void *const dq[3], int ncomp, bool to_dq, const FusedBox &fb, int cell_bytes,
                            int ny, int pitch, long long ps, cudaStream_t stream)
{
  if (fb.nqt <= 0) return cudaSuccess;
  for (int a = 0; a < ncomp; ++a)
  {
    dq_copy_kernel<<<592, 256, 0, stream>>> ((int4 *) main_d[a], (int4 *) dq[a], to_dq, fb, ny, (long long) pitch * cell_bytes / 16,
                                             ps * cell_bytes / 16);
    cudaError_t e = cudaGetLastError ();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_fill (void *dev, long long count, int cellBytes, const void *pattern16, cudaStream_t stream)
{
  const int threads = 256;
  long long want = (count + threads - 1) / threads;
  const int blocks = (int) (want < 148 * 16 ? (want > 0 ? want : 1) : 148 * 16);
  switch (cellBytes)
  {
    case 4: fill_kernel<uint32_t><<<blocks, threads, 0, stream>>> ((uint32_t *) dev, count, *(const uint32_t *) pattern16); break;
    case 8: fill_kernel<unsigned long long><<<blocks, threads, 0, stream>>> ((unsigned long long *) dev, count, *(const unsigned long long *) pattern16); break;
    case 16: fill_kernel<int4><<<blocks, threads, 0, stream>>> ((int4 *) dev, count, *(const int4 *) pattern16); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError ();
}

template <typename T>
cudaError_t launch_eps_sphere (T *eps, int mx, int my, int nzm, int kz0, int pitchM, long long psM, T cx, T cy, T cz,
                               T radius, T eps_in, cudaStream_t stream)
{
  dim3 block (128);
  dim3 grid ((mx + 127) / 128, my, nzm < 64 ? nzm : 64);
  eps_sphere_kernel<T><<<grid, block, 0, stream>>> (eps, mx, my, nzm, kz0, pitchM, psM, cx, cy, cz, radius, eps_in);
  return cudaGetLastError ();
}
template cudaError_t launch_eps_sphere<float> (float *, int, int, int, int, int, long long, float, float, float, float, float, cudaStream_t);
template cudaError_t launch_eps_sphere<double> (double *, int, int, int, int, int, long long, double, double, double, double, double, cudaStream_t);

int sumsq_partial_count () { return SUMSQ_CTAS; }

template <typename T>
cudaError_t launch_sumsq (const T *base, int nscal, long long nrows, long long pitch_scal, double *partials, double *out,
                          cudaStream_t stream)
{
  sumsq_partial<T><<<SUMSQ_CTAS, SUMSQ_THREADS, 0, stream>>> (base, nscal, nrows, pitch_scal, partials);
  cudaError_t e = cudaGetLastError ();
  if (e != cudaSuccess) return e;
  sumsq_final<<<1, SUMSQ_THREADS, 0, stream>>> (partials, SUMSQ_CTAS, out);
  return cudaGetLastError ();
}
template <typename T>
cudaError_t launch_uniform_check (const T *base, int nscal, long long nrows, long long pitch_scal, double *out2, cudaStream_t stream)
{
  cudaError_t e = cudaMemsetAsync (out2, 0, 2 * sizeof (double), stream);
  if (e != cudaSuccess) return e;
  uniform_check<T><<<SUMSQ_CTAS, SUMSQ_THREADS, 0, stream>>> (base, nscal, nrows, pitch_scal, out2);
  return cudaGetLastError ();
}
template cudaError_t launch_uniform_check<float> (const float *, int, long long, long long, double *, cudaStream_t);
template cudaError_t launch_uniform_check<double> (const double *, int, long long, long long, double *, cudaStream_t);

template cudaError_t launch_sumsq<float> (const float *, int, long long, long long, double *, double *, cudaStream_t);
template cudaError_t launch_sumsq<double> (const double *, int, long long, long long, double *, double *, cudaStream_t);

} /* namespace fdtd */
