// Consumers of the path's intermediates, evaluated tile by tile on the device so that nothing of size n_det x n_t in
// float64 is ever materialised:
//   matm_transmission  map step: pW per K_RJ of a frequency channel at (T0, zenith_scaled_pwv(d, t), el(d, t))
//                      (maria/sim/map.py:117-133 -> band.py:235-255), zenith_scaled_pwv sampled and linearly upsampled
//                      inside the kernel (sim/atmosphere.py:30-37)
//   matm_noise_scale   noise step: 1e12 (NEP + NEP_per_loading * loading) * unscaled_noise  (sim/noise.py:35-37,52-61)
//   matm_pointing      obs.coords on demand: az / el of a detector range over a sample range
//                      (sim/observation.py:55-58 -> coords/transforms.py:10-29)
#include <math.h>

#include "matm_synth.cuh"

#define K_B 1.380649e-23
#define CN_DT 32
#define CN_THREADS 256
#define CN_KT 64  // spline-free: intervals of the atmosphere-rate axis per tile

// H3[b][w][e] = 1e12 k_B * lerp_T0( trapz_{nu in [i0, i1)} passband_b(nu) exp(-opacity[T][w][e][nu]) dnu )
__global__ void fold_channel_kernel(const double* __restrict__ opacity, const double* __restrict__ passband,
                                    const double* __restrict__ nu, int n_bands, int npwv, int nel, int nnu, int iT, double wT,
                                    int i0, int i1, double* __restrict__ H3) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_bands * npwv * nel) return;
  const int b = e / (npwv * nel), w = (e / nel) % npwv, el = e % nel;
  const double* pb = passband + (size_t)b * nnu;
  double v[2];
  for (int dt = 0; dt < 2; dt++) {
    const double* tab = opacity + (((size_t)(iT + dt) * npwv + w) * nel + el) * nnu;
    double acc = 0.0;
    for (int i = i0 + 1; i < i1; i++)
      acc += (nu[i] - nu[i - 1]) * (pb[i] * exp(-tab[i]) + pb[i - 1] * exp(-tab[i - 1])) / 2.0;
    v[dt] = acc;
  }
  H3[e] = 1e12 * K_B * (v[0] * (1.0 - wT) + v[1] * wT);
}

// bilinear lookup on the (pwv, el) grid with jax RegularGridInterpolator semantics (SURVEY 8a notes): NaN outside
template <typename R>
__device__ __forceinline__ R lookup_pwv_el(const SynthParams& P, const double* __restrict__ tab, R zpwv, R el, int& ip, int& ie) {
  const R* gp = reinterpret_cast<const R*>(P.pwv_R);
  const R* ge = reinterpret_cast<const R*>(P.el_R);
  ip = find_interval<R>(gp, P.npwv, zpwv, ip);
  ie = find_interval<R>(ge, P.nel, el, ie);
  const R p0 = __ldg(gp + ip), p1 = __ldg(gp + ip + 1), e0 = __ldg(ge + ie), e1 = __ldg(ge + ie + 1);
  if (zpwv < __ldg(gp) || zpwv > __ldg(gp + P.npwv - 1) || el < __ldg(ge) || el > __ldg(ge + P.nel - 1) || !(zpwv == zpwv)) {
    if (zpwv == zpwv) atomicOr(&P.flags[0], 2);
    return r_nan<R>();
  }
  const R wp = r_div(zpwv - p0, p1 - p0), we = r_div(el - e0, e1 - e0);
  const double* g = tab + (size_t)ip * P.nel + ie;
  const R g00 = (R)__ldg(g), g01 = (R)__ldg(g + 1), g10 = (R)__ldg(g + P.nel), g11 = (R)__ldg(g + P.nel + 1);
  const R up = (R)1 - wp, ue = (R)1 - we;
  return g00 * (up * ue) + g01 * (up * we) + g10 * (wp * ue) + g11 * (wp * we);
}

// One CTA = CN_DT detectors x CN_KT intervals: zenith_scaled_pwv at the tile's knots (lane = detector) into shared memory,
// then every full-rate sample of the tile: linear interpolation in time, elevation of the detector, table lookup.
template <typename R>
__global__ void __launch_bounds__(CN_THREADS) transmission_kernel(const __grid_constant__ SynthParams P, const double* __restrict__ H3,
                                                                  int64_t det0, int64_t n_det, R* __restrict__ out) {
  constexpr int W = CN_KT + 1;
  __shared__ Knot<R> knots[W];
  __shared__ LayerS<R> ls[MATM_MAX_LAYERS];
  __shared__ R zp[W][CN_DT + 1];
  extern __shared__ __align__(16) unsigned char dyn[];
  LayerKnot<R>* lk = reinterpret_cast<LayerKnot<R>*>(dyn);
  const int n = P.n_t_atm, n_int = n - 1;
  const int n_tt = (n_int + CN_KT - 1) / CN_KT;
  const int64_t dtile = blockIdx.x / n_tt;
  const int tt = blockIdx.x % n_tt;
  const int k0 = tt * CN_KT, k1 = min(k0 + CN_KT, n_int);
  const int Wn = k1 - k0 + 1;
  load_window<R>(P, k0, Wn, W, knots, lk, ls);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t dbase = dtile * CN_DT;
  const int nd = (int)min((int64_t)CN_DT, n_det - dbase);
  if (lane < nd) {
    const int64_t d = det0 + dbase + lane;
    const double* da = P.det_atm + 4 * d;
    const R u = (R)da[0], Y = (R)da[1], wm1 = (R)da[2];
    for (int k = warp; k < Wn; k += CN_THREADS / 32) {
      R sin_el;
      zp[k][lane] = sample_zenith_pwv<R>(P, P.n_layers, ls, u, Y, wm1, knots[k], lk + k, W, d, k0 + k, &sin_el);
    }
  }
  __syncthreads();
  const int64_t j0 = P.jstart[k0], j1 = P.jstart[k1];
  int ip = 0, ie = 0;
  for (int64_t j = j0 + threadIdx.x; j < j1; j += CN_THREADS) {
    const float4 w = __ldg(P.samp_w + j);  // .x = (t - t_i) / (t_i+1 - t_i), .w = i: the linear weights too
    const int li = min(max(__float_as_int(w.w), k0), k1 - 1) - k0;
    const double ti = P.t_atm[k0 + li], s64 = (P.t_full[j] - ti) / (P.t_atm[k0 + li + 1] - ti);
    const R s = (R)s64;
    const double2 bf = *reinterpret_cast<const double2*>(P.bore_full + 2 * j);
    const R ce = (R)bf.x, se = (R)bf.y;
    for (int dd = 0; dd < nd; dd++) {
      const int64_t d = det0 + dbase + dd;
      const R z0 = zp[li][dd], z1 = zp[li + 1][dd];
      const R z = z0 + s * (z1 - z0);
      const R sin_el = (R)P.det_full[2 * d] * se + (R)P.det_full[2 * d + 1] * ce;
      const R el = r_asin01(sin_el);
      out[(dbase + dd) * P.n_t + j] = lookup_pwv_el<R>(P, H3 + (size_t)P.band_index[d] * P.npwv * P.nel, z, el, ip, ie);
    }
  }
}

extern "C" int matm_transmission(matm_handle* h, int64_t det0, int64_t n_det, double nu_min_Hz, double nu_max_Hz, void* out_dev,
                                 void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_scan || !h->have_layers || !h->have_dets || !h->have_bands)
    MATM_FAIL_ARG(h, "Atmosphere must be initialized with a simulation (scan, layers, detectors, bands not all set)");
  if (det0 < 0 || n_det < 1 || det0 + n_det > h->P.n_det || !out_dev) MATM_FAIL_ARG(h, "matm_transmission: bad arguments");
  const SynthParams& P = h->P;
  cudaStream_t st = (cudaStream_t)stream;
  // spectrum.side_nu >= nu_min and < nu_max (band.py:249-250): a contiguous range of the frequency axis
  int i0 = 0, i1 = h->nnu;
  while (i0 < h->nnu && !(h->nu_host[i0] >= nu_min_Hz)) i0++;
  while (i1 > i0 && !(h->nu_host[i1 - 1] < nu_max_Hz)) i1--;
  const size_t n_tab = (size_t)P.n_bands * P.npwv * P.nel;
  if (!h->H3) MATM_CUDA(h, cudaMalloc((void**)&h->H3, n_tab * sizeof(double)));
  fold_channel_kernel<<<(unsigned)((n_tab + 127) / 128), 128, 0, st>>>(h->opacity, h->passband, h->side_nu, P.n_bands, P.npwv, P.nel,
                                                                         h->nnu, h->iT, h->wT, i0, i1, h->H3);
  MATM_LAUNCHED(h);
  const int n_tt = (P.n_t_atm - 1 + CN_KT - 1) / CN_KT;
  const int64_t blocks = ((n_det + CN_DT - 1) / CN_DT) * n_tt;
  if (blocks > INT32_MAX) MATM_FAIL_ARG(h, "matm_transmission: too many tiles; split the detector range");
  if (h->dtype == MATM_F64) {
    const size_t smem = (size_t)P.n_layers * (CN_KT + 1) * sizeof(LayerKnot<double>);
    MATM_CUDA(h, cudaFuncSetAttribute(transmission_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    transmission_kernel<double><<<(unsigned)blocks, CN_THREADS, smem, st>>>(P, h->H3, det0, n_det, (double*)out_dev);
  } else {
    const size_t smem = (size_t)P.n_layers * (CN_KT + 1) * sizeof(LayerKnot<float>);
    MATM_CUDA(h, cudaFuncSetAttribute(transmission_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    transmission_kernel<float><<<(unsigned)blocks, CN_THREADS, smem, st>>>(P, h->H3, det0, n_det, (float*)out_dev);
  }
  MATM_LAUNCHED(h);
  return MATM_OK;
}

// ---------------------------------------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(256) noise_scale_kernel(const R* __restrict__ loading, const R* __restrict__ noise,
                                                          const int32_t* __restrict__ band_index, const double* __restrict__ nep,
                                                          int64_t det0, int64_t n_det, int64_t n_t, R* __restrict__ out) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_t) return;
  for (int64_t d = blockIdx.y; d < n_det; d += gridDim.y) {
    const int b = band_index[det0 + d];
    const double total = nep[2 * b] + nep[2 * b + 1] * (double)loading[d * n_t + j];  // W sqrt(s)  (noise.py:35-37)
    out[d * n_t + j] = (R)(1e12 * total * (double)noise[d * n_t + j]);               // pW           (noise.py:60)
  }
}

extern "C" int matm_noise_scale(matm_handle* h, int64_t det0, int64_t n_det, const void* loading_dev, const void* noise_dev,
                                const double* nep_host, const double* nep_per_loading_host, void* out_dev, void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_scan || !h->have_dets || !h->have_bands) MATM_FAIL_ARG(h, "matm_noise_scale: call matm_set_scan, matm_set_detectors and matm_set_bands first");
  if (det0 < 0 || n_det < 1 || det0 + n_det > h->P.n_det || !loading_dev || !noise_dev || !nep_host || !nep_per_loading_host || !out_dev)
    MATM_FAIL_ARG(h, "matm_noise_scale: bad arguments");
  std::vector<double> v(2 * (size_t)h->P.n_bands);
  for (int b = 0; b < h->P.n_bands; b++) {
    v[2 * b] = nep_host[b];
    v[2 * b + 1] = nep_per_loading_host[b];
  }
  double* nep_d = nullptr;
  cudaStream_t st = (cudaStream_t)stream;
  MATM_CUDA(h, cudaMallocAsync((void**)&nep_d, v.size() * sizeof(double), st));
  MATM_CUDA(h, cudaMemcpyAsync(nep_d, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, st));
  MATM_CUDA(h, cudaStreamSynchronize(st));  // `v` is a stack-lifetime staging buffer
  dim3 grid((unsigned)((h->P.n_t + 255) / 256), (unsigned)min((int64_t)2048, n_det));
  if (h->dtype == MATM_F64)
    noise_scale_kernel<double><<<grid, 256, 0, st>>>((const double*)loading_dev, (const double*)noise_dev, h->band_index, nep_d, det0,
                                                      n_det, h->P.n_t, (double*)out_dev);
  else
    noise_scale_kernel<float><<<grid, 256, 0, st>>>((const float*)loading_dev, (const float*)noise_dev, h->band_index, nep_d, det0,
                                                     n_det, h->P.n_t, (float*)out_dev);
  MATM_LAUNCHED(h);
  MATM_CUDA(h, cudaFreeAsync(nep_d, st));
  return MATM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// az / el of detectors [det0, det0 + n_det) at the full-rate samples [j0, j0 + n_j), float64 (transforms.py:10-29 with the
// roll-rotated offsets of observation.py:55-58): q = (u + i w) exp(i (el_b - pi/2)), az = atan2(v, Re q) + az_b, el = asin(Im q)
__global__ void __launch_bounds__(256) pointing_kernel(const __grid_constant__ SynthParams P, const double* __restrict__ az_full,
                                                       const double* __restrict__ det_v, int64_t det0, int64_t n_det, int64_t j0,
                                                       int64_t n_j, double* __restrict__ az, double* __restrict__ el) {
  const int64_t jj = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (jj >= n_j) return;
  const int64_t j = j0 + jj;
  const double ce = P.bore_full[2 * j], se = P.bore_full[2 * j + 1], azb = az_full[j];
  for (int64_t d = blockIdx.y; d < n_det; d += gridDim.y) {
    const int64_t dg = det0 + d;
    const double u = P.det_full[2 * dg], w = P.det_full[2 * dg + 1], v = det_v[dg];
    if (az) az[d * n_j + jj] = atan2(v, u * ce - w * se) + azb;
    if (el) el[d * n_j + jj] = asin(fmin(fmax(u * se + w * ce, -1.0), 1.0));
  }
}

extern "C" int matm_pointing(matm_handle* h, int64_t det0, int64_t n_det, int64_t j0, int64_t n_j, double* az_dev, double* el_dev,
                             void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_scan || !h->have_dets) MATM_FAIL_ARG(h, "matm_pointing: call matm_set_scan and matm_set_detectors first");
  if (det0 < 0 || n_det < 1 || det0 + n_det > h->P.n_det || j0 < 0 || n_j < 1 || j0 + n_j > h->P.n_t || (!az_dev && !el_dev))
    MATM_FAIL_ARG(h, "matm_pointing: bad arguments");
  if (az_dev && !h->az_full) MATM_FAIL_ARG(h, "matm_pointing: matm_set_scan was given no full-rate azimuth");
  dim3 grid((unsigned)((n_j + 255) / 256), (unsigned)min((int64_t)2048, n_det));
  pointing_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(h->P, h->az_full, h->det_v, det0, n_det, j0, n_j, az_dev, el_dev);
  MATM_LAUNCHED(h);
  return MATM_OK;
}
