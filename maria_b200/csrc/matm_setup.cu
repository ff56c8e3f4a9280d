// Set-up entry points: scan, layers, detectors, band tables.  Small host arrays are uploaded and the derived
// per-time / per-detector / per-band constants are evaluated on the device in float64.
#include <algorithm>
#include <math.h>

#include "matm_common.cuh"

#define K_B 1.380649e-23

// ---- scan ------------------------------------------------------------------------------------------------
// maria/coords/transforms.py:10-29: q = (sin r cos p + i cos r) * exp(i (el_b - pi/2)); az = atan2(sin r sin p, Re q) + az_b
__global__ void bore_atm_kernel(const double* __restrict__ az, const double* __restrict__ el, double* __restrict__ out, int n) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  double se, ce, sa, ca;
  sincos(el[k] - M_PI / 2, &se, &ce);
  sincos(az[k], &sa, &ca);
  out[4 * k + 0] = ce;
  out[4 * k + 1] = se;
  out[4 * k + 2] = ca;
  out[4 * k + 3] = sa;
}

__global__ void bore_full_kernel(const double* __restrict__ el, double* __restrict__ out, int64_t n) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  double se, ce;
  sincos(el[j] - M_PI / 2, &se, &ce);
  out[2 * j + 0] = ce;
  out[2 * j + 1] = se;
}

__host__ __device__ __forceinline__ int interval_of(double t, double t0, double ts, int n_atm) {
  int i = (int)floor((t - t0) / ts);
  int hi = n_atm - 2;
  return i < 0 ? 0 : (i > hi ? hi : i);
}

// jstart[k] = first full-rate sample whose interpolation interval (sim/atmosphere.py:30-37,73-80) is >= k
__global__ void jstart_kernel(const double* __restrict__ t, int64_t n_t, double t0, double ts, int n_atm,
                              int64_t* __restrict__ jstart) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_atm) return;
  if (k >= n_atm - 1) {
    jstart[k] = n_t;
    return;
  }
  int64_t lo = 0, hi = n_t;  // first j with interval_of(t[j]) >= k
  while (lo < hi) {
    int64_t mid = (lo + hi) / 2;
    if (interval_of(t[mid], t0, ts, n_atm) >= k)
      hi = mid;
    else
      lo = mid + 1;
  }
  jstart[k] = lo;
}

// el_rng[k] = (min, max) of the full-rate boresight elevation over the samples of spline interval k, i.e.
// [jstart[k], jstart[k + 1]): what the per-tile choice of the a14 evaluation mode has to bound (the elevation at the
// knots does not bound the full-rate samples at turnarounds or in the extrapolated tail).
__global__ void el_range_kernel(const double* __restrict__ el, const int64_t* __restrict__ jstart, int n_int,
                                double* __restrict__ rng) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_int) return;
  double lo = INFINITY, hi = -INFINITY;
  for (int64_t j = jstart[k]; j < jstart[k + 1]; j++) {
    lo = fmin(lo, el[j]);
    hi = fmax(hi, el[j]);
  }
  rng[2 * k] = lo;
  rng[2 * k + 1] = hi;
}

// Per full-rate sample: the interval i of the knot vector it falls in (the same rule as jstart) and the weights of the
// cubic spline in its second-derivative form, s = (t - x_i) / h_i:
//   S(t) = (1 - s) y_i + s y_i+1 + h_i^2/6 [((1-s)^3 - (1-s)) M_i + (s^3 - s) M_i+1]      (M in units of timestep^-2)
// evaluated in float64 and rounded once; they depend on the time axes only, so every tile of every detector reads them.
__global__ void sample_weights_kernel(const double* __restrict__ t, int64_t n_t, const double* __restrict__ x, double t0,
                                      double ts, int n_atm, float4* __restrict__ out) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_t) return;
  const int i = interval_of(t[j], t0, ts, n_atm);
  const double h = x[i + 1] - x[i], s = (t[j] - x[i]) / h, c = 1.0 - s, hk = h / ts, h26 = hk * hk / 6.0;
  out[j] = make_float4((float)s, (float)(h26 * (c * c * c - c)), (float)(h26 * (s * s * s - s)), __int_as_float(i));
}

__global__ void knot_f_kernel(const double* __restrict__ bore_atm, int n, float4* __restrict__ out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  out[k] = make_float4((float)bore_atm[4 * k], (float)bore_atm[4 * k + 1], (float)bore_atm[4 * k + 2], (float)bore_atm[4 * k + 3]);
}

// Not-a-knot cubic spline through the knots x_k (scipy interp1d kind="cubic" = make_interp_spline k=3, default
// boundary condition; sim/atmosphere.py:73-80), as the tridiagonal system for the second derivatives M_k in units of
// `timestep`: rows k = 1 .. n-2,  a_k M_{k-1} + b_k M_k + c_k M_{k+1} = 6 [(y_{k+1}-y_k)/h_k - (y_k-y_{k-1})/h_{k-1}],
// with M_0 and M_{n-1} eliminated through the continuity of the third derivative at x_1 and x_{n-2}.  The forward
// elimination factors depend on the knots only and are computed once, here (one thread, n steps).
__global__ void spline_coef_kernel(const double* __restrict__ x, int n, double timestep, double* __restrict__ spl) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  for (int k = 0; k < n - 1; k++) {
    double hk = (x[k + 1] - x[k]) / timestep;
    spl[5 * k + 0] = 1.0 / hk;
    spl[5 * k + 1] = hk * hk / 6.0;
  }
  spl[5 * (n - 1) + 0] = spl[5 * (n - 1) + 1] = 0.0;
  for (int j = 2; j < 5; j++) spl[j] = spl[5 * (n - 1) + j] = 0.0;
  double cp_prev = 0.0;
  for (int k = 1; k <= n - 2; k++) {
    double h0 = (x[k] - x[k - 1]) / timestep, h1 = (x[k + 1] - x[k]) / timestep;
    double a = h0, b = 2.0 * (h0 + h1), c = h1;
    if (k == 1) {  // M_0 = M_1 - (h_0 / h_1) (M_2 - M_1)
      a = 0.0;
      b = 3.0 * h0 + 2.0 * h1 + h0 * h0 / h1;
      c = h1 - h0 * h0 / h1;
    }
    if (k == n - 2) {  // M_{n-1} = M_{n-2} + (h_{n-2} / h_{n-3}) (M_{n-2} - M_{n-3});  n >= 4, so k > 1 here
      a = h0 - h1 * h1 / h0;
      b = 2.0 * h0 + 3.0 * h1 + h1 * h1 / h0;
      c = 0.0;
    }
    double w = 1.0 / (b - a * cp_prev);
    cp_prev = c * w;
    spl[5 * k + 2] = w;
    spl[5 * k + 3] = a * w;
    spl[5 * k + 4] = cp_prev;
  }
}

extern "C" int matm_set_scan(matm_handle* h, int n_t_atm, const double* t_atm_host, double timestep,
                             const double* az_atm_host, const double* el_atm_host, int64_t n_t, const double* t_host,
                             const double* az_host, const double* el_host) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (n_t_atm < 2 || n_t < 1 || !(timestep > 0)) MATM_FAIL_ARG(h, "matm_set_scan: need n_t_atm >= 2, n_t >= 1, timestep > 0");
  if (!t_atm_host || !az_atm_host || !el_atm_host || !t_host || !el_host) MATM_FAIL_ARG(h, "matm_set_scan: null pointer");
  for (int k = 1; k < n_t_atm; k++)
    if (!(t_atm_host[k] > t_atm_host[k - 1])) MATM_FAIL_ARG(h, "matm_set_scan: knot times must increase");
  for (int k = 0; k < n_t_atm; k++)
    if (!isfinite(az_atm_host[k]) || !isfinite(el_atm_host[k])) MATM_FAIL_ARG(h, "matm_set_scan: boresight az / el must be finite");
  for (int64_t j = 0; j < n_t; j++)
    if (!isfinite(el_host[j]) || !isfinite(t_host[j])) MATM_FAIL_ARG(h, "matm_set_scan: full-rate time / elevation must be finite");
  if (n_t > (int64_t)INT32_MAX) MATM_FAIL_ARG(h, "matm_set_scan: more than 2^31 - 1 samples per detector; split the scan in time");
  // the full-rate azimuth does not enter the synthesis (only the elevation does, tod.py:126-133); it is kept for
  // matm_pointing (obs.coords on demand)
  const double t_atm0 = t_atm_host[0];
  double *az_d = nullptr, *el_d = nullptr;
  int rc;
  if ((rc = matm_upload(h, &az_d, az_atm_host, (size_t)n_t_atm))) return rc;
  if ((rc = matm_upload(h, &el_d, el_atm_host, (size_t)n_t_atm))) return rc;
  if ((rc = matm_upload(h, &h->el_full, el_host, (size_t)n_t))) return rc;
  if (az_host) {
    if ((rc = matm_upload(h, &h->az_full, az_host, (size_t)n_t))) return rc;
  } else if (h->az_full) {
    cudaFree(h->az_full);
    h->az_full = nullptr;
  }
  if ((rc = matm_upload(h, &h->el_rng, (const double*)nullptr, (size_t)n_t_atm * 2))) return rc;
  if ((rc = matm_upload(h, &h->knot_f, (const float4*)nullptr, (size_t)n_t_atm))) return rc;
  if ((rc = matm_upload(h, &h->samp_w, (const float4*)nullptr, (size_t)n_t))) return rc;
  double* elf_d = h->el_full;
  if ((rc = matm_upload(h, &h->bore_atm, (const double*)nullptr, (size_t)n_t_atm * 4))) return rc;
  if ((rc = matm_upload(h, &h->bore_full, (const double*)nullptr, (size_t)n_t * 2))) return rc;
  if ((rc = matm_upload(h, &h->t_full, t_host, (size_t)n_t))) return rc;
  if ((rc = matm_upload(h, &h->jstart, (const int64_t*)nullptr, (size_t)n_t_atm + 1))) return rc;
  if ((rc = matm_upload(h, &h->t_atm, t_atm_host, (size_t)n_t_atm))) return rc;
  if ((rc = matm_upload(h, &h->spl, (const double*)nullptr, (size_t)n_t_atm * 5))) return rc;
  bore_atm_kernel<<<(n_t_atm + 255) / 256, 256>>>(az_d, el_d, h->bore_atm, n_t_atm);
  MATM_LAUNCHED(h);
  bore_full_kernel<<<(unsigned)((n_t + 255) / 256), 256>>>(elf_d, h->bore_full, n_t);
  MATM_LAUNCHED(h);
  jstart_kernel<<<(n_t_atm + 1 + 255) / 256, 256>>>(h->t_full, n_t, t_atm0, timestep, n_t_atm, h->jstart);
  MATM_LAUNCHED(h);
  if (n_t_atm >= 4) {
    spline_coef_kernel<<<1, 32>>>(h->t_atm, n_t_atm, timestep, h->spl);
    MATM_LAUNCHED(h);
  }
  el_range_kernel<<<(n_t_atm - 1 + 255) / 256, 256>>>(h->el_full, h->jstart, n_t_atm - 1, h->el_rng);
  MATM_LAUNCHED(h);
  knot_f_kernel<<<(n_t_atm + 255) / 256, 256>>>(h->bore_atm, n_t_atm, h->knot_f);
  MATM_LAUNCHED(h);
  sample_weights_kernel<<<(unsigned)((n_t + 255) / 256), 256>>>(h->t_full, n_t, h->t_atm, t_atm0, timestep, n_t_atm, h->samp_w);
  MATM_LAUNCHED(h);
  MATM_CUDA(h, cudaDeviceSynchronize());
  cudaFree(az_d);
  cudaFree(el_d);
  SynthParams& P = h->P;
  P.n_t_atm = n_t_atm;
  P.n_t = n_t;
  P.t_atm0 = t_atm0;
  P.timestep = timestep;
  P.bore_atm = h->bore_atm;
  P.bore_full = h->bore_full;
  P.t_full = h->t_full;
  P.jstart = h->jstart;
  {
    std::vector<int64_t> js((size_t)n_t_atm + 1);
    MATM_CUDA(h, cudaMemcpy(js.data(), h->jstart, js.size() * sizeof(int64_t), cudaMemcpyDeviceToHost));
    int64_t mn = INT64_MAX;
    for (int k = 1; k + 2 < n_t_atm; k++) mn = std::min(mn, js[k + 1] - js[k]);  // interior intervals only
    h->min_samples_per_interval = mn == INT64_MAX ? 0 : mn;
  }
  P.t_atm = h->t_atm;
  P.spl = h->spl;
  P.el_full = h->el_full;
  P.el_rng = h->el_rng;
  P.knot_f = h->knot_f;
  P.samp_w = h->samp_w;
  h->have_scan = true;
  h->have_layers = false;  // layer bases depend on the scan
  return MATM_OK;
}

// ---- layers ----------------------------------------------------------------------------------------------
// base[k] = ((h * pp_b[k] + cumsum(timestep * v)[k]) @ T - grid origin) / grid step   (atmosphere.py:318-319,346-347)
__global__ void layer_base_kernel(const double* __restrict__ bore_atm, const double* __restrict__ vx,
                                  const double* __restrict__ vy, int n, double timestep, double hgt, const double* T9,
                                  double ext0, double ext_step, double cs0, double cs_step, double* __restrict__ base) {
  // inclusive cumulative sum, sequential like np.cumsum, one thread per component
  if (threadIdx.x < 2) {
    const double* v = threadIdx.x == 0 ? vx : vy;
    double acc = 0.0;
    for (int k = 0; k < n; k++) {
      acc += timestep * v[k];
      base[2 * k + threadIdx.x] = acc;
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    double ce = bore_atm[4 * k + 0], se = bore_atm[4 * k + 1], ca = bore_atm[4 * k + 2], sa = bore_atm[4 * k + 3];
    // boresight direction (cos el cos az, cos el sin az, sin el) with cos el = -se, sin el = ce
    double px = hgt * ((-se * ca) / ce) + base[2 * k + 0];
    double py = hgt * ((-se * sa) / ce) + base[2 * k + 1];
    double pz = hgt;
    double qx = px * T9[0] + py * T9[3] + pz * T9[6];
    double qy = px * T9[1] + py * T9[4] + pz * T9[7];
    base[2 * k + 0] = (qx - ext0) / ext_step;
    base[2 * k + 1] = (qy - cs0) / cs_step;
  }
}

// the boresight's position in a layer grid split into integer cell + fraction, in float (the fp32 loading kernel reads
// one 16-byte record per layer and knot)
struct LayerKnotF {
  int ix, iy;
  float fx, fy;
};
__global__ void lk_tab_kernel(const double* __restrict__ base, int n, LayerKnotF* __restrict__ out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double cx = base[2 * k], cy = base[2 * k + 1];
  // clamp the integer part so that absurd (far out-of-grid) positions cannot overflow; they are flagged by the sampler
  const double flx = fmin(fmax(floor(cx), -1.0e9), 1.0e9), fly = fmin(fmax(floor(cy), -1.0e9), 1.0e9);
  LayerKnotF o;
  o.ix = (int)flx;
  o.iy = (int)fly;
  o.fx = (float)(cx - flx);
  o.fy = (float)(cy - fly);
  out[k] = o;
}

extern "C" int matm_set_layers(matm_handle* h, int n_layers, const matm_layer_desc* L, double pwv0) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_scan) MATM_FAIL_ARG(h, "matm_set_layers: call matm_set_scan first");
  if (n_layers < 0 || n_layers > MATM_MAX_LAYERS || (n_layers && !L)) MATM_FAIL_ARG(h, "matm_set_layers: bad layer count");
  SynthParams& P = h->P;
  const int n = P.n_t_atm;
  double *vx_d = nullptr, *vy_d = nullptr, *T_d = nullptr;
  int rc;
  if (h->lk_tab) MATM_CUDA(h, cudaFree(h->lk_tab));
  h->lk_tab = nullptr;
  if (n_layers > 0) MATM_CUDA(h, cudaMalloc(&h->lk_tab, (size_t)n_layers * n * sizeof(LayerKnotF)));
  P.lk_tab = h->lk_tab;
  for (int l = 0; l < n_layers; l++) {
    const matm_layer_desc& d = L[l];
    if (!d.data_dev || d.n_ext < 2 || d.n_cs < 2 || !d.vx_host || !d.vy_host || !(d.ext_step > 0) || !(d.cs_last > d.cs0))
      MATM_FAIL_ARG(h, "matm_set_layers: bad layer descriptor");
    if ((rc = matm_upload(h, &vx_d, d.vx_host, (size_t)n))) return rc;
    if ((rc = matm_upload(h, &vy_d, d.vy_host, (size_t)n))) return rc;
    if ((rc = matm_upload(h, &T_d, d.transform, (size_t)9))) return rc;
    if ((rc = matm_upload(h, &h->layer_base[l], (const double*)nullptr, (size_t)n * 2))) return rc;
    const double cs_step = (d.cs_last - d.cs0) / (d.n_cs - 1);
    layer_base_kernel<<<1, 256>>>(h->bore_atm, vx_d, vy_d, n, P.timestep, d.h, T_d, d.ext0, d.ext_step, d.cs0, cs_step,
                                  h->layer_base[l]);
    MATM_LAUNCHED(h);
    lk_tab_kernel<<<(n + 255) / 256, 256>>>(h->layer_base[l], n, (LayerKnotF*)h->lk_tab + (size_t)l * n);
    MATM_LAUNCHED(h);
    MATM_CUDA(h, cudaDeviceSynchronize());
    LayerDev& o = P.layers[l];
    if (h->dtype == MATM_F32)
      if ((rc = matm_upload(h, &h->layer_pairs[l], (const float2*)nullptr, (size_t)d.n_ext * d.n_cs))) return rc;
    o.data2 = h->layer_pairs[l];
    o.data = d.data_dev;
    o.base = h->layer_base[l];
    o.n_ext = d.n_ext;
    o.n_cs = d.n_cs;
    o.D[0] = d.h * d.transform[0] / d.ext_step;
    o.D[1] = d.h * d.transform[3] / d.ext_step;
    o.D[2] = d.h * d.transform[1] / cs_step;
    o.D[3] = d.h * d.transform[4] / cs_step;
    o.pwv_rms = d.pwv_rms;
  }
  cudaFree(vx_d);
  cudaFree(vy_d);
  cudaFree(T_d);
  P.n_layers = n_layers;
  P.pwv0 = pwv0;
  h->have_layers = true;
  return MATM_OK;
}

// ---- detectors -------------------------------------------------------------------------------------------
__global__ void det_kernel(const double* __restrict__ off_atm, const double* __restrict__ off_full, int64_t n,
                           double* __restrict__ det_atm, double* __restrict__ det_full, double* __restrict__ det_cal,
                           double* __restrict__ det_v) {
  int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n) return;
  {
    double dx = off_atm[2 * d], dy = off_atm[2 * d + 1];
    double r = sqrt(dx * dx + dy * dy), p = atan2(-dx, -dy);
    double sr, cr, sp, cp, sh = sin(0.5 * r);
    sincos(r, &sr, &cr);
    sincos(p, &sp, &cp);
    det_atm[4 * d + 0] = sr * cp;
    det_atm[4 * d + 1] = sr * sp;
    det_atm[4 * d + 2] = -2.0 * sh * sh;  // cos r - 1 without cancellation
    det_atm[4 * d + 3] = cr;
  }
  {
    double dx = off_full[2 * d], dy = off_full[2 * d + 1];
    double r = sqrt(dx * dx + dy * dy), p = atan2(-dx, -dy);
    double sr, cr, sp, cp;
    sincos(r, &sr, &cr);
    sincos(p, &sp, &cp);
    const double u = sr * cp, w = cr;
    det_full[2 * d + 0] = u;
    det_full[2 * d + 1] = w;
    // sin el_d = u sin(e) + w cos(e), e = el_b - pi/2, = rho sin(el_b - beta) with rho e^{i beta} = w + i u;
    // 1 - rho = (sin r sin p)^2 / (1 + rho) without cancellation
    const double rho = sqrt(u * u + w * w), q = sr * sp;
    det_cal[2 * d + 0] = atan2(u, w);
    det_cal[2 * d + 1] = q * q / (1.0 + rho);
    det_v[d] = q;  // sin r sin p: the azimuth of the line of sight is atan2(v, Re q) + az_b (transforms.py:10-29)
  }
}

extern "C" int matm_set_layer_data(matm_handle* h, int n_layers, const void* const* data_dev) {
  MATM_CHECK_HANDLE(h);
  if (!h->have_layers) MATM_FAIL_ARG(h, "matm_set_layer_data: call matm_set_layers first");
  if (n_layers != h->P.n_layers || !data_dev) MATM_FAIL_ARG(h, "matm_set_layer_data: layer count does not match matm_set_layers");
  for (int l = 0; l < n_layers; l++)
    if (!data_dev[l]) MATM_FAIL_ARG(h, "matm_set_layer_data: null layer pointer");
  for (int l = 0; l < n_layers; l++) h->P.layers[l].data = data_dev[l];
  return MATM_OK;
}

extern "C" int matm_set_detectors(matm_handle* h, int64_t n_det, const double* offsets_atm_host,
                                  const double* offsets_full_host, const int32_t* band_index_host,
                                  const double* mueller00_host, const double* gain_host) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (n_det < 1 || !offsets_atm_host || !offsets_full_host || !band_index_host || !mueller00_host)
    MATM_FAIL_ARG(h, "matm_set_detectors: bad arguments");
  for (int64_t i = 0; i < 2 * n_det; i++)
    if (!isfinite(offsets_atm_host[i]) || !isfinite(offsets_full_host[i])) MATM_FAIL_ARG(h, "matm_set_detectors: offsets must be finite");
  double *oa = nullptr, *of = nullptr;
  int rc;
  if ((rc = matm_upload(h, &oa, offsets_atm_host, (size_t)n_det * 2))) return rc;
  if ((rc = matm_upload(h, &of, offsets_full_host, (size_t)n_det * 2))) return rc;
  if ((rc = matm_upload(h, &h->det_atm, (const double*)nullptr, (size_t)n_det * 4))) return rc;
  if ((rc = matm_upload(h, &h->det_full, (const double*)nullptr, (size_t)n_det * 2))) return rc;
  if ((rc = matm_upload(h, &h->det_cal, (const double*)nullptr, (size_t)n_det * 2))) return rc;
  if ((rc = matm_upload(h, &h->det_v, (const double*)nullptr, (size_t)n_det))) return rc;
  if ((rc = matm_upload(h, &h->band_index, band_index_host, (size_t)n_det))) return rc;
  if ((rc = matm_upload(h, &h->mueller00, mueller00_host, (size_t)n_det))) return rc;
  std::vector<double> ones;
  if (!gain_host) {
    ones.assign((size_t)n_det, 1.0);
    gain_host = ones.data();
  }
  if ((rc = matm_upload(h, &h->gain, gain_host, (size_t)n_det))) return rc;
  det_kernel<<<(unsigned)((n_det + 255) / 256), 256>>>(oa, of, n_det, h->det_atm, h->det_full, h->det_cal, h->det_v);
  MATM_LAUNCHED(h);
  MATM_CUDA(h, cudaDeviceSynchronize());
  cudaFree(oa);
  cudaFree(of);
  h->det_xy_host.resize((size_t)n_det * 2);
  for (size_t i = 0; i < (size_t)n_det * 2; i++) h->det_xy_host[i] = offsets_atm_host[i];
  for (auto& kv : h->det_order) {
    cudaFree(kv.second.order);
    cudaFree(kv.second.tile_m);
  }
  h->det_order.clear();
  SynthParams& P = h->P;
  double rmax = 0.0;
  for (int64_t i = 0; i < n_det; i++) rmax = fmax(rmax, hypot(offsets_full_host[2 * i], offsets_full_host[2 * i + 1]));
  P.det_rmax_f = (float)(rmax * (1.0 + 1e-6));
  P.det_rmax = rmax;
  P.det_cal = h->det_cal;
  P.n_det = n_det;
  P.det_atm = h->det_atm;
  P.det_full = h->det_full;
  P.band_index = h->band_index;
  P.mueller00 = h->mueller00;
  P.gain = h->gain;
  h->have_dets = true;
  return MATM_OK;
}

extern "C" int matm_set_gain(matm_handle* h, const double* gain_host, void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_dets) MATM_FAIL_ARG(h, "matm_set_gain: call matm_set_detectors first");
  if (!gain_host) MATM_FAIL_ARG(h, "matm_set_gain: null pointer");
  // one copy into the existing buffer, ordered on the caller's stream: no reallocation, no device-wide synchronisation,
  // the processing order of the detectors stays cached
  MATM_CUDA(h, cudaMemcpyAsync(h->gain, gain_host, (size_t)h->P.n_det * sizeof(double), cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return MATM_OK;
}

static inline uint32_t spread_bits16(uint32_t v) {  // 16 bits -> every other bit of 32
  v = (v | (v << 8)) & 0x00FF00FFu;
  v = (v | (v << 4)) & 0x0F0F0F0Fu;
  v = (v | (v << 2)) & 0x33333333u;
  v = (v | (v << 1)) & 0x55555555u;
  return v;
}

int matm_detector_order(matm_handle* h, int64_t det0, int64_t n_det, const int32_t** order_dev, const uint8_t** tile_m_dev) {
  const auto key = std::make_pair(det0, n_det);
  auto it = h->det_order.find(key);
  if (it != h->det_order.end()) {
    *order_dev = it->second.order;
    *tile_m_dev = it->second.tile_m;
    return MATM_OK;
  }
  if (h->det_order.size() >= 256) {  // bounded cache (chunked delivery asks for a handful of ranges)
    for (auto& kv : h->det_order) {
      cudaFree(kv.second.order);
      cudaFree(kv.second.tile_m);
    }
    h->det_order.clear();
  }
  const double* xy = h->det_xy_host.data() + 2 * (size_t)det0;
  double lo[2] = {xy[0], xy[1]}, hi[2] = {xy[0], xy[1]};
  for (int64_t i = 0; i < n_det; i++)
    for (int a = 0; a < 2; a++) {
      lo[a] = fmin(lo[a], xy[2 * i + a]);
      hi[a] = fmax(hi[a], xy[2 * i + a]);
    }
  const double span = fmax(fmax(hi[0] - lo[0], hi[1] - lo[1]), 1e-300);
  std::vector<std::pair<uint32_t, int32_t>> keyed((size_t)n_det);
  for (int64_t i = 0; i < n_det; i++) {
    const uint32_t qx = (uint32_t)fmin(65535.0, (xy[2 * i] - lo[0]) / span * 65535.0);
    const uint32_t qy = (uint32_t)fmin(65535.0, (xy[2 * i + 1] - lo[1]) / span * 65535.0);
    keyed[(size_t)i] = std::make_pair(spread_bits16(qx) | (spread_bits16(qy) << 1), (int32_t)i);
  }
  // Morton order; inside one Morton cell by exact position, so that co-located detectors always end up adjacent
  std::sort(keyed.begin(), keyed.end(), [xy](const std::pair<uint32_t, int32_t>& a, const std::pair<uint32_t, int32_t>& b) {
    if (a.first != b.first) return a.first < b.first;
    const double ax = xy[2 * a.second], ay = xy[2 * a.second + 1], bx = xy[2 * b.second], by = xy[2 * b.second + 1];
    if (ax != bx) return ax < bx;
    if (ay != by) return ay < by;
    return a.second < b.second;
  });
  std::vector<int32_t> order((size_t)n_det);
  for (int64_t i = 0; i < n_det; i++) order[(size_t)i] = (int32_t)(det0 + keyed[(size_t)i].second);
  // per tile of 32 slots: the largest M in {4, 2, 1} such that every aligned group of M slots shares one position
  const int64_t n_tiles = (n_det + 31) / 32;
  std::vector<uint8_t> tile_m((size_t)n_tiles, 1);
  auto same = [&](int64_t s0, int64_t s1) {
    const int32_t a = keyed[(size_t)s0].second, b = keyed[(size_t)s1].second;
    return xy[2 * a] == xy[2 * b] && xy[2 * a + 1] == xy[2 * b + 1];
  };
  for (int64_t t = 0; t < n_tiles; t++) {
    if ((t + 1) * 32 > n_det) continue;  // partial tile
    for (int m = 4; m >= 2; m >>= 1) {
      bool ok = true;
      for (int g = 0; g < 32 && ok; g += m)
        for (int j = 1; j < m && ok; j++) ok = same(t * 32 + g, t * 32 + g + j);
      if (ok) {
        tile_m[(size_t)t] = (uint8_t)m;
        break;
      }
    }
  }
  matm_handle::DetOrder rec = {nullptr, nullptr};
  MATM_CUDA(h, cudaMalloc((void**)&rec.order, (size_t)n_det * sizeof(int32_t)));
  MATM_CUDA(h, cudaMemcpy(rec.order, order.data(), (size_t)n_det * sizeof(int32_t), cudaMemcpyHostToDevice));
  MATM_CUDA(h, cudaMalloc((void**)&rec.tile_m, (size_t)n_tiles));
  MATM_CUDA(h, cudaMemcpy(rec.tile_m, tile_m.data(), (size_t)n_tiles, cudaMemcpyHostToDevice));
  h->det_order[key] = rec;
  *order_dev = rec.order;
  *tile_m_dev = rec.tile_m;
  return MATM_OK;
}

// ---- band tables -----------------------------------------------------------------------------------------
// trapezoid over nu of f(table[..., nu]) * passband[nu]  (band.py:248-253,272-280)
__device__ double band_integral(const double* __restrict__ tab, const double* __restrict__ pb,
                                const double* __restrict__ nu, int nnu, bool transmission) {
  double acc = 0.0;
  double prev = transmission ? pb[0] * exp(-tab[0]) : tab[0] * pb[0];
  for (int i = 1; i < nnu; i++) {
    double cur = transmission ? pb[i] * exp(-tab[i]) : tab[i] * pb[i];
    acc += (nu[i] - nu[i - 1]) * (cur + prev) / 2.0;
    prev = cur;
  }
  return acc;
}

// jax RegularGridInterpolator index / weight on one axis (SURVEY 8a notes); returns false when out of bounds
__host__ __device__ inline bool axis_locate(const double* g, int n, double x, int* idx, double* w) {
  int lo = 0, hi = n;  // searchsorted(g, x), side="left"
  while (lo < hi) {
    int mid = (lo + hi) / 2;
    if (g[mid] < x)
      lo = mid + 1;
    else
      hi = mid;
  }
  int i = lo - 1;
  i = i < 0 ? 0 : (i > n - 2 ? n - 2 : i);
  *idx = i;
  *w = (x - g[i]) / (g[i + 1] - g[i]);
  return !(x < g[0] || x > g[n - 1]) && (x == x);
}

__global__ void fold_tables_kernel(const double* __restrict__ emission, const double* __restrict__ opacity,
                                   const double* __restrict__ passband, const double* __restrict__ nu, int n_bands,
                                   int nT, int npwv, int nel, int nnu, int iT, double wT, int iTc, double wTc, int iPc,
                                   double wPc, const double* __restrict__ cal_scale, double* __restrict__ G2,
                                   double* __restrict__ H1) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  int total_g = n_bands * npwv * nel;
  if (e < total_g) {
    int b = e / (npwv * nel), w = (e / nel) % npwv, el = e % nel;
    const double* pb = passband + (size_t)b * nnu;
    double g0 = band_integral(emission + (((size_t)iT * npwv + w) * nel + el) * nnu, pb, nu, nnu, false);
    double g1 = band_integral(emission + (((size_t)(iT + 1) * npwv + w) * nel + el) * nnu, pb, nu, nnu, false);
    G2[e] = 1e12 * K_B * (g0 * (1.0 - wT) + g1 * wT);
  } else if (e < total_g + n_bands * nel) {
    int f = e - total_g;
    int b = f / nel, el = f % nel;
    const double* pb = passband + (size_t)b * nnu;
    double acc = 0.0;
    for (int dt = 0; dt < 2; dt++)
      for (int dp = 0; dp < 2; dp++) {
        double v = band_integral(opacity + (((size_t)(iTc + dt) * npwv + (iPc + dp)) * nel + el) * nnu, pb, nu, nnu, true);
        acc += v * (dt ? wTc : 1.0 - wTc) * (dp ? wPc : 1.0 - wPc);
      }
    H1[f] = acc * cal_scale[b];
  }
}

// R-typed copies of the lookup tables, laid out back to back: pwv | el | sin el | G2 | Hc | Hs
template <typename R>
__global__ void typed_tables_kernel(const double* __restrict__ pwv, const double* __restrict__ el,
                                    const double* __restrict__ G2, const double* __restrict__ H1, int n_bands, int npwv,
                                    int nel, R* __restrict__ out) {
  R* o_pwv = out;
  R* o_el = o_pwv + npwv;
  R* o_sin = o_el + nel;
  R* o_G2 = o_sin + nel;
  R* o_Hc = o_G2 + (size_t)n_bands * npwv * nel;
  R* o_Hs = o_Hc + (size_t)n_bands * (nel - 1);
  int t = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
  for (int i = t; i < npwv; i += nt) o_pwv[i] = (R)pwv[i];
  for (int i = t; i < nel; i += nt) {
    o_el[i] = (R)el[i];
    o_sin[i] = el[i] >= M_PI / 2 ? (R)2 : (R)sin(el[i]);
  }
  for (int i = t; i < n_bands * npwv * nel; i += nt) o_G2[i] = (R)G2[i];
  for (int i = t; i < n_bands * (nel - 1); i += nt) {
    int b = i / (nel - 1), e = i % (nel - 1);
    double slope = (H1[b * nel + e + 1] - H1[b * nel + e]) / (el[e + 1] - el[e]);
    o_Hs[i] = (R)slope;
    o_Hc[i] = (R)(H1[b * nel + e] - slope * el[e]);
  }
}


extern "C" int matm_set_bands(matm_handle* h, int n_bands, int nT, int npwv, int nel, int nnu, const double* side_T_host,
                              const double* side_pwv_host, const double* side_el_host, const double* side_nu_host,
                              const double* emission_host, const double* opacity_host, const double* passband_host,
                              double T0, double cal_T0, double cal_pwv, const double* cal_factor_host) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (n_bands < 1 || n_bands > MATM_MAX_BANDS || nT < 2 || npwv < 2 || nel < 2 || nnu < 2)
    MATM_FAIL_ARG(h, "matm_set_bands: bad table shape");
  if (!side_T_host || !side_pwv_host || !side_el_host || !side_nu_host || !emission_host || !opacity_host ||
      !passband_host || !cal_factor_host)
    MATM_FAIL_ARG(h, "matm_set_bands: null pointer");
  const size_t ntab = (size_t)nT * npwv * nel * nnu;
  double *em = nullptr, *op = nullptr, *pb = nullptr, *nu = nullptr, *cs = nullptr;
  int rc;
  if ((rc = matm_upload(h, &em, emission_host, ntab))) return rc;
  if ((rc = matm_upload(h, &op, opacity_host, ntab))) return rc;
  if ((rc = matm_upload(h, &pb, passband_host, (size_t)n_bands * nnu))) return rc;
  if ((rc = matm_upload(h, &nu, side_nu_host, (size_t)nnu))) return rc;
  std::vector<double> scale((size_t)n_bands);
  for (int b = 0; b < n_bands; b++) scale[b] = cal_factor_host[b] * K_B * 1e12;  // K_RJ = pW / (factor k_B 1e12 H)
  if ((rc = matm_upload(h, &cs, scale.data(), (size_t)n_bands))) return rc;
  if ((rc = matm_upload(h, &h->side_pwv, side_pwv_host, (size_t)npwv))) return rc;
  if ((rc = matm_upload(h, &h->side_el, side_el_host, (size_t)nel))) return rc;
  if ((rc = matm_upload(h, &h->G2, (const double*)nullptr, (size_t)n_bands * npwv * nel))) return rc;
  if ((rc = matm_upload(h, &h->H1, (const double*)nullptr, (size_t)n_bands * nel))) return rc;
  int iT, iTc, iPc;
  double wT, wTc, wPc;
  bool ok = axis_locate(side_T_host, nT, T0, &iT, &wT);
  bool okc = axis_locate(side_T_host, nT, cal_T0, &iTc, &wTc);
  okc = axis_locate(side_pwv_host, npwv, cal_pwv, &iPc, &wPc) && okc;
  int total = n_bands * npwv * nel + n_bands * nel;
  fold_tables_kernel<<<(total + 127) / 128, 128>>>(em, op, pb, nu, n_bands, nT, npwv, nel, nnu, iT, wT, iTc, wTc, iPc, wPc,
                                                   cs, h->G2, h->H1);
  MATM_LAUNCHED(h);
  const size_t n_tab = (size_t)npwv + 2 * nel + (size_t)n_bands * npwv * nel + 2 * (size_t)n_bands * (nel - 1);
  const size_t esz = h->dtype == MATM_F64 ? 8 : 4;
  if (h->tabR) MATM_CUDA(h, cudaFree(h->tabR));
  h->tabR = nullptr;
  MATM_CUDA(h, cudaMalloc(&h->tabR, n_tab * esz));
  if (h->dtype == MATM_F64)
    typed_tables_kernel<double><<<8, 256>>>(h->side_pwv, h->side_el, h->G2, h->H1, n_bands, npwv, nel, (double*)h->tabR);
  else
    typed_tables_kernel<float><<<8, 256>>>(h->side_pwv, h->side_el, h->G2, h->H1, n_bands, npwv, nel, (float*)h->tabR);
  MATM_LAUNCHED(h);
  MATM_CUDA(h, cudaDeviceSynchronize());
  cudaFree(em);
  // the opacity table, the passbands and the frequency axis stay on the device: the map-step transmission of a frequency
  // channel (matm_transmission; sim/map.py:117-133) folds them again over [nu_min, nu_max)
  cudaFree(h->opacity);
  cudaFree(h->passband);
  cudaFree(h->side_nu);
  h->opacity = op;
  h->passband = pb;
  h->side_nu = nu;
  h->nT = nT;
  h->nnu = nnu;
  h->iT = iT;
  h->wT = wT;
  h->T_ok = ok;
  h->nu_host.assign(side_nu_host, side_nu_host + nnu);
  cudaFree(cs);
  SynthParams& P = h->P;
  P.n_bands = n_bands;
  P.npwv = npwv;
  P.nel = nel;
  P.side_pwv = h->side_pwv;
  P.side_el = h->side_el;
  P.G2 = h->G2;
  P.H1 = h->H1;
  {
    char* base = (char*)h->tabR;
    P.pwv_R = base;
    P.el_R = base + esz * (size_t)npwv;
    P.sinel_R = base + esz * ((size_t)npwv + nel);
    P.G2_R = base + esz * ((size_t)npwv + 2 * nel);
    P.Hc_R = base + esz * ((size_t)npwv + 2 * nel + (size_t)n_bands * npwv * nel);
    P.Hs_R = base + esz * ((size_t)npwv + 2 * nel + (size_t)n_bands * npwv * nel + (size_t)n_bands * (nel - 1));
  }
  P.table_ok = (ok ? 1 : 0) | (okc ? 2 : 0);
  P.pwv_lo_f = (float)side_pwv_host[0];
  P.pwv_hi_f = (float)side_pwv_host[npwv - 1];
  P.el_lo_f = (float)side_el_host[0];
  P.el_hi_f = (float)side_el_host[nel - 1];
  h->have_bands = true;
  return MATM_OK;
}
