// The detector x time stages: a6/a7/a9 (line-of-sight sampling), a10 (linear upsample), and the fused
// a6..a14 synthesis kernel (sampling -> emission lookup -> cubic not-a-knot upsample -> gain -> pW->K_RJ).
#include <mutex>

#include "matm_synth.cuh"
#include "matm_synth_f32.cuh"

#define DT 32          // detectors per tile (one lane per detector in the atmosphere-rate phase)
#define SY_THREADS 256

template <typename R>
struct Tile {
  static constexpr int KT = sizeof(R) == 4 ? 128 : 64;   // spline intervals per tile
  static constexpr int HALO = sizeof(R) == 4 ? 12 : 24;  // extra knots each side; the error of the artificial
                                                         // boundary decays as (2 - sqrt 3)^HALO
  static constexpr int W = KT + 1 + 2 * HALO;            // knots in the window
  static constexpr int WP = W + 1;                       // padded row length (bank-conflict free)
};

// ---------------------------------------------------------------------------------------------------------
// zenith_scaled_pwv at the atmosphere rate (atmosphere.py:293-380), float64 output
// ---------------------------------------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(SY_THREADS) zenith_pwv_kernel(const __grid_constant__ SynthParams P, int64_t det0,
                                                                int64_t n_det, double* __restrict__ out) {
  constexpr int W = 128;
  __shared__ Knot<R> knots[W];
  __shared__ LayerS<R> ls[MATM_MAX_LAYERS];
  extern __shared__ __align__(16) unsigned char dyn[];
  LayerKnot<R>* lk = reinterpret_cast<LayerKnot<R>*>(dyn);
  const int n_ttiles = (P.n_t_atm + W - 1) / W;
  const int64_t dtile = blockIdx.x / n_ttiles;
  const int ttile = blockIdx.x % n_ttiles;
  const int wa = ttile * W, Wn = min(W, P.n_t_atm - wa);
  load_window<R>(P, wa, Wn, W, knots, lk, ls);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t d = dtile * DT + lane;
  if (d >= n_det) return;
  const double* da = P.det_atm + 4 * (det0 + d);
  const R u = (R)da[0], Y = (R)da[1], wm1 = (R)da[2];
  for (int k = warp; k < Wn; k += SY_THREADS / 32) {
    R sin_el;
    R z = sample_zenith_pwv<R>(P, P.n_layers, ls, u, Y, wm1, knots[k], lk + k, W, det0 + d, wa + k, &sin_el);
    out[d * P.n_t_atm + wa + k] = (double)z;
  }
}

// ---------------------------------------------------------------------------------------------------------
// a10: scipy interp1d(kind="linear", fill_value="extrapolate") along time (sim/atmosphere.py:30-37)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) upsample_linear_kernel(const __grid_constant__ SynthParams P,
                                                              const double* __restrict__ in, int64_t n_det,
                                                              double* __restrict__ out) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P.n_t) return;
  const double t = P.t_full[j];
  int i = (int)floor((t - P.t_atm0) / P.timestep);
  i = min(max(i, 0), P.n_t_atm - 2);
  // the interval in the actual (rounded) knot vector, as np.searchsorted(x, t) - 1 clipped to [0, n - 2]
  while (i > 0 && !(P.t_atm[i] < t)) i--;
  while (i < P.n_t_atm - 2 && P.t_atm[i + 1] < t) i++;
  const double x0 = P.t_atm[i], x1 = P.t_atm[i + 1];
  const double s = (t - x0) / (x1 - x0);
  for (int64_t d = blockIdx.y; d < n_det; d += gridDim.y) {
    const double y0 = in[d * P.n_t_atm + i], y1 = in[d * P.n_t_atm + i + 1];
    out[d * P.n_t + j] = y0 + s * (y1 - y0);
  }
}

// ---------------------------------------------------------------------------------------------------------
// The fused kernel.  One CTA = DT detectors x KT spline intervals.
//   phase 1: loading [pW] at the window's knots (lane = detector, warp strides over knots)   -> smem y
//   phase 2: second derivatives of the interpolating cubic spline by a windowed Thomas solve -> smem m
//   phase 3: evaluate at the full-rate samples of the tile, gain, calibration, one store per sample
// ---------------------------------------------------------------------------------------------------------
template <typename R>
struct YM {  // knot value and spline second derivative (times the detector gain), interleaved: one 8/16-byte load
  R y, m;
};

template <typename R>
__global__ void __launch_bounds__(SY_THREADS) synth_kernel(const __grid_constant__ SynthParams P, int64_t det0,
                                                           int64_t n_det, R* __restrict__ out, int units,
                                                           int apply_gain) {
  using T = Tile<R>;
  extern __shared__ __align__(16) unsigned char dyn[];
  YM<R>* ym = reinterpret_cast<YM<R>*>(dyn);              // [DT][WP]
  R* spl = reinterpret_cast<R*>(ym + DT * T::WP);         // [W][5] spline system of the window (see SynthParams::spl)
  Knot<R>* knots = reinterpret_cast<Knot<R>*>(spl + 5 * T::W);
  LayerKnot<R>* lk = reinterpret_cast<LayerKnot<R>*>(knots + T::W);  // [n_layers][W]
  __shared__ LayerS<R> ls[MATM_MAX_LAYERS];
  __shared__ R dc_u[DT], dc_w[DT], dc_gain[DT];
  __shared__ int dc_hoff[DT];

  const int n = P.n_t_atm;
  const int n_int = n - 1;  // spline intervals
  const int n_layers = P.n_layers;
  const int n_ttiles = (n_int + T::KT - 1) / T::KT;
  const int64_t dtile = blockIdx.x / n_ttiles;
  const int ttile = blockIdx.x % n_ttiles;
  const int k0 = ttile * T::KT, k1 = min(k0 + T::KT, n_int);  // intervals [k0, k1), knots k0..k1
  const int wa = max(0, k0 - T::HALO), wb = min(n - 1, k1 + T::HALO);
  const int Wn = wb - wa + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t dbase = dtile * DT;  // relative to det0
  const int nd = (int)min((int64_t)DT, n_det - dbase);

  load_window<R>(P, wa, Wn, T::W, knots, lk, ls);
  if (threadIdx.x < DT) {
    int dd = threadIdx.x;
    if (dd < nd) {
      const int64_t d = det0 + dbase + dd;
      dc_u[dd] = (R)P.det_full[2 * d];
      dc_w[dd] = (R)P.det_full[2 * d + 1];
      dc_gain[dd] = apply_gain ? (R)P.gain[d] : (R)1;
      dc_hoff[dd] = P.band_index[d] * (P.nel - 1);
    }
  }
  for (int e = threadIdx.x; e < 5 * Wn; e += SY_THREADS) spl[e] = (R)P.spl[5 * (size_t)wa + e];
  __syncthreads();

  // ---- phase 1: loading [pW] at the window's knots; lane = detector, the warps stride over the knots ----
  if (lane < nd) {
    const int64_t d = det0 + dbase + lane;
    const double* da = P.det_atm + 4 * d;
    const R u = (R)da[0], Y = (R)da[1], wm1 = (R)da[2];
    const int band = P.band_index[d];
    const R mu = (R)P.mueller00[d];
    int ip = 0, ie = 0;
    for (int k = warp; k < Wn; k += SY_THREADS / 32) {
      R sin_el;
      R z = sample_zenith_pwv<R>(P, n_layers, ls, u, Y, wm1, knots[k], lk + k, T::W, d, wa + k, &sin_el);
      ym[lane * T::WP + k].y = loading_pW<R>(P, band, mu, z, sin_el, ip, ie);
    }
  }
  __syncthreads();

  // ---- phase 2: second derivatives M_k of the not-a-knot cubic spline through the window's knots ----
  // Rows of the global tridiagonal system, swept with the global elimination factors.  At a true end of the series
  // the sweep starts / ends exactly; at an artificial end it starts from the locally-constant guess
  // M = d / (a + b + c), whose error decays by ~(2 - sqrt 3) per knot across the halo.  The gain multiplies y and M
  // at the end (the spline is linear in the data), so phase 3 does not have to.
  if (warp == 0 && lane < nd) {
    YM<R>* row = ym + lane * T::WP;
    const int lo = 1, hi = Wn - 2;  // local rows; global row = wa + local
    auto yv = [&](int k) { return row[k].y; };
    auto invh = [&](int k) { return spl[5 * k + 0]; };
    auto rhs = [&](int k) { return (R)6 * ((yv(k + 1) - yv(k)) * invh(k) - (yv(k) - yv(k - 1)) * invh(k - 1)); };
    auto guess = [&](int k) { return rhs(k) / ((R)3 * ((R)1 / invh(k - 1) + (R)1 / invh(k))); };
    R dp = (wa == 0) ? rhs(lo) * spl[5 * lo + 2] : guess(lo) * ((R)1 + spl[5 * lo + 4]);
    row[lo].m = dp;
    for (int k = lo + 1; k <= hi; k++) {
      dp = rhs(k) * spl[5 * k + 2] - spl[5 * k + 3] * dp;
      row[k].m = dp;
    }
    if (wb != n - 1) row[hi].m -= spl[5 * hi + 4] * guess(hi);
    for (int k = hi - 1; k >= lo; k--) row[k].m -= spl[5 * k + 4] * row[k + 1].m;
    row[0].m = (wa == 0) ? row[1].m - (invh(1) / invh(0)) * (row[2].m - row[1].m) : row[1].m;
    row[Wn - 1].m = (wb == n - 1) ? row[Wn - 2].m + (invh(Wn - 3) / invh(Wn - 2)) * (row[Wn - 2].m - row[Wn - 3].m)
                                  : row[Wn - 2].m;
    const R g = dc_gain[lane];
    for (int k = 0; k < Wn; k++) {
      row[k].y *= g;
      row[k].m *= g;
    }
  }
  __syncthreads();

  // ---- phase 3: evaluate at the full-rate samples of the tile; thread = sample, loop over the tile's detectors ----
  const int64_t j0 = P.jstart[k0], j1 = P.jstart[k1];
  const double t_atm0 = P.t_atm0, inv_ts = 1.0 / P.timestep;
  const int64_t n_t = P.n_t;
  const bool krj = units == MATM_UNITS_KRJ;
  const R* sing = reinterpret_cast<const R*>(P.sinel_R);
  const R* hcs = reinterpret_cast<const R*>(P.Hc_R);
  const R* hss = reinterpret_cast<const R*>(P.Hs_R);
  int ie = 0;
  for (int64_t j = j0 + threadIdx.x; j < j1; j += SY_THREADS) {
    const double tj = P.t_full[j];
    int i = (int)floor((tj - t_atm0) * inv_ts);  // the tile's interval up to rounding; s is exact w.r.t. the knots used
    i = min(max(i, k0), k1 - 1);
    const double xi = P.t_atm[i];
    const R s = (R)((tj - xi) / (P.t_atm[i + 1] - xi));  // may leave [0, 1] by ~1e-6 (knot rounding) or at the ends
    const int li = i - wa;
    const R a0 = (R)1 - s, a1 = s;
    const R h26 = spl[5 * li + 1];
    const R a2 = h26 * (a0 * a0 * a0 - a0), a3 = h26 * (s * s * s - s);
    R cef = (R)0, sef = (R)0;
    if (krj) {
      cef = (R)P.bore_full[2 * j];
      sef = (R)P.bore_full[2 * j + 1];
    }
    const YM<R>* p = ym + li;
    R* o = out + dbase * n_t + j;
#pragma unroll 4
    for (int dd = 0; dd < nd; dd++) {
      const YM<R> q0 = p[dd * T::WP], q1 = p[dd * T::WP + 1];
      R val = a0 * q0.y + a1 * q1.y + a2 * q0.m + a3 * q1.m;
      if (krj) {
        const R sin_el = dc_u[dd] * sef + dc_w[dd] * cef;  // Im q of the rolled offsets at the full rate
        const int ho = dc_hoff[dd];
        val = r_div(val, cal_denominator<R>(P, sing, hcs + ho, hss + ho, sin_el, ie));
      }
      o[(int64_t)dd * n_t] = val;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// a14 alone (TOD.to): pW <-> K_RJ on an existing (n_det, n_t) array
// ---------------------------------------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(256) calibrate_kernel(const __grid_constant__ SynthParams P, int64_t det0,
                                                        int64_t n_det, const R* __restrict__ in, R* __restrict__ out,
                                                        int to_units) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P.n_t) return;
  const R cef = (R)P.bore_full[2 * j], sef = (R)P.bore_full[2 * j + 1];
  int ie = 0;
  for (int64_t d = blockIdx.y; d < n_det; d += gridDim.y) {
    const int64_t dg = det0 + d;
    const R sin_el = (R)P.det_full[2 * dg] * sef + (R)P.det_full[2 * dg + 1] * cef;
    const int ho = P.band_index[dg] * (P.nel - 1);
    const R den = cal_denominator<R>(P, reinterpret_cast<const R*>(P.sinel_R), reinterpret_cast<const R*>(P.Hc_R) + ho,
                                     reinterpret_cast<const R*>(P.Hs_R) + ho, sin_el, ie);
    const R x = in[d * P.n_t + j];
    out[d * P.n_t + j] = to_units == MATM_UNITS_KRJ ? x / den : x * den;
  }
}

// ---------------------------------------------------------------------------------------------------------
// entry points
// ---------------------------------------------------------------------------------------------------------
static int check_ready(matm_handle* h, bool need_bands, int64_t det0, int64_t n_det) {
  if (!h->have_scan || !h->have_layers || !h->have_dets || (need_bands && !h->have_bands))
    MATM_FAIL_ARG(h, "Atmosphere must be initialized with a simulation (scan, layers, detectors, bands not all set)");
  if (det0 < 0 || n_det < 1 || det0 + n_det > h->P.n_det) MATM_FAIL_ARG(h, "detector range out of bounds");
  return MATM_OK;
}

extern "C" int matm_zenith_pwv(matm_handle* h, int64_t det0, int64_t n_det, double* out_dev, void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  int rc = check_ready(h, false, det0, n_det);
  if (rc) return rc;
  if (!out_dev) MATM_FAIL_ARG(h, "matm_zenith_pwv: null output");
  const SynthParams& P = h->P;
  const int n_ttiles = (P.n_t_atm + 127) / 128;
  const int64_t blocks = ((n_det + DT - 1) / DT) * n_ttiles;
  if (blocks > INT32_MAX) MATM_FAIL_ARG(h, "matm_zenith_pwv: too many tiles; split the detector range");
  if (h->dtype == MATM_F64) {
    size_t smem = (size_t)P.n_layers * 128 * sizeof(LayerKnot<double>);
    MATM_CUDA(h, cudaFuncSetAttribute(zenith_pwv_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    zenith_pwv_kernel<double><<<(unsigned)blocks, SY_THREADS, smem, (cudaStream_t)stream>>>(P, det0, n_det, out_dev);
  } else {
    size_t smem = (size_t)P.n_layers * 128 * sizeof(LayerKnot<float>);
    MATM_CUDA(h, cudaFuncSetAttribute(zenith_pwv_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    zenith_pwv_kernel<float><<<(unsigned)blocks, SY_THREADS, smem, (cudaStream_t)stream>>>(P, det0, n_det, out_dev);
  }
  MATM_LAUNCHED(h);
  return MATM_OK;
}

extern "C" int matm_upsample_linear(matm_handle* h, const double* in_dev, int64_t n_det, double* out_dev, void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_scan) MATM_FAIL_ARG(h, "matm_upsample_linear: call matm_set_scan first");
  if (!in_dev || !out_dev || n_det < 1) MATM_FAIL_ARG(h, "matm_upsample_linear: bad arguments");
  dim3 grid((unsigned)((h->P.n_t + 255) / 256), (unsigned)min((int64_t)1024, n_det));
  upsample_linear_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(h->P, in_dev, n_det, out_dev);
  MATM_LAUNCHED(h);
  return MATM_OK;
}

template <typename R>
static size_t synth_smem(int n_layers) {
  using T = Tile<R>;
  return (size_t)DT * T::WP * sizeof(YM<R>) + 5 * T::W * sizeof(R) + T::W * sizeof(Knot<R>) +
         (size_t)n_layers * T::W * sizeof(LayerKnot<R>);
}

// The production float path: per chunk of detector tiles, loading_f32_kernel (atmosphere rate) writes the chunk's
// loading at the knots into an L2-sized scratch, synth_eval_f32_kernel turns it into the full-rate TOD rows.  Consecutive
// chunks alternate between two scratch buffers and two side streams, so that the latency-bound loading kernel of chunk
// c + 1 fills the SMs the HBM-bound evaluation kernel of chunk c leaves idle; the caller's stream forks into them and
// joins at the end (everything stays stream-ordered for the caller, nothing synchronises).
static int launch_synth_f32(matm_handle* h, int64_t det0, int64_t n_det, float* out, int units, int apply_gain,
                            cudaStream_t st) {
  const SynthParams& P = h->P;
  const int n = P.n_t_atm;
  const int32_t* order = nullptr;
  const uint8_t* tile_m = nullptr;
  int rc = matm_detector_order(h, det0, n_det, &order, &tile_m);
  if (rc) return rc;
  const int64_t n_tiles = (n_det + F32_DT - 1) / F32_DT;
  const size_t tile_bytes = (size_t)n * F32_DT * sizeof(float);
  int64_t chunk = (int64_t)(((size_t)h->chunk_mib << 20) / tile_bytes);
  if (chunk < 1) chunk = 1;
  if (chunk > n_tiles) chunk = n_tiles;
  const int n_kb = (n + LD_KA - 1) / LD_KA, n_tt = (n - 1 + F32_KT - 1) / F32_KT;
  if (chunk * (int64_t)max(n_kb, n_tt) > INT32_MAX) chunk = INT32_MAX / max(n_kb, n_tt);
  const size_t need = (size_t)chunk * tile_bytes;
  const int64_t n_chunks = (n_tiles + chunk - 1) / chunk;
  const bool two = h->overlap && n_chunks > 1;
  if (h->yk_bytes < need) {
    MATM_CUDA(h, cudaDeviceSynchronize());  // kernels of an earlier call may still read the old scratch
    for (int i = 0; i < 2; i++) {
      if (h->yk_buf[i]) MATM_CUDA(h, cudaFree(h->yk_buf[i]));
      h->yk_buf[i] = nullptr;
      MATM_CUDA(h, cudaMalloc((void**)&h->yk_buf[i], need));
    }
    h->yk_bytes = need;
  }
  for (int i = 0; i < 2; i++)
    if (!h->yk_stream[i]) MATM_CUDA(h, cudaStreamCreateWithFlags(&h->yk_stream[i], cudaStreamNonBlocking));
  for (int i = 0; i < 4; i++)
    if (!h->yk_ev[i]) MATM_CUDA(h, cudaEventCreateWithFlags(&h->yk_ev[i], cudaEventDisableTiming));
  const size_t smem = synth_f32_smem();
  // the 4-samples-per-thread kernel needs at most one knot inside any 4 consecutive samples
  const bool fast = h->min_samples_per_interval >= 4 && P.n_t < ((int64_t)1 << 30) && n_det < ((int64_t)1 << 31);
  auto eval_kernel = fast ? synth_eval_f32_kernel : synth_eval_general_f32_kernel;
  MATM_CUDA(h, cudaFuncSetAttribute(eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // the scratch buffers are shared by every call on this handle: order this call after the previous one
  if (h->yk_used) {
    MATM_CUDA(h, cudaStreamWaitEvent(st, h->yk_ev[3], 0));
  }
  {  // the sampler reads the smoothed layers through their pair-expanded copy
    int max_cells = 1;
    for (int l = 0; l < P.n_layers; l++) max_cells = max(max_cells, P.layers[l].n_ext * P.layers[l].n_cs);
    dim3 grid((unsigned)min(4 * h->sm_count, (max_cells + 255) / 256), (unsigned)P.n_layers);
    expand_layers_kernel<<<grid, 256, 0, st>>>(P);
    MATM_LAUNCHED(h);
  }
  const size_t smem_a = (size_t)P.n_layers * LD_KA * sizeof(int4);
  MATM_CUDA(h, cudaFuncSetAttribute(loading_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_a));
  if (two) {
    MATM_CUDA(h, cudaEventRecord(h->yk_ev[0], st));
    MATM_CUDA(h, cudaStreamWaitEvent(h->yk_stream[0], h->yk_ev[0], 0));
    MATM_CUDA(h, cudaStreamWaitEvent(h->yk_stream[1], h->yk_ev[0], 0));
  }
  for (int64_t c = 0; c < n_chunks; c++) {
    const int64_t t0 = c * chunk, nt = min(chunk, n_tiles - t0);
    cudaStream_t s = two ? h->yk_stream[c & 1] : st;
    float* yk = h->yk_buf[two ? (c & 1) : 0];
    cudaEvent_t* te = nullptr;
    if (h->timing) {
      while (h->tev.size() < h->tev_used + 3) {
        cudaEvent_t e;
        MATM_CUDA(h, cudaEventCreate(&e));
        h->tev.push_back(e);
      }
      te = &h->tev[h->tev_used];
      h->tev_used += 3;
      MATM_CUDA(h, cudaEventRecord(te[0], s));
    }
    loading_f32_kernel<<<(unsigned)(nt * n_kb), LD_THREADS, smem_a, s>>>(P, order, tile_m, t0, n_det, yk, apply_gain);
    MATM_LAUNCHED(h);
    if (te) MATM_CUDA(h, cudaEventRecord(te[1], s));
    eval_kernel<<<(unsigned)(nt * n_tt), F32_THREADS, smem, s>>>(P, order, t0, det0, n_det, yk, out, units);
    MATM_LAUNCHED(h);
    if (te) MATM_CUDA(h, cudaEventRecord(te[2], s));
  }
  if (two) {
    for (int i = 0; i < 2; i++) {
      MATM_CUDA(h, cudaEventRecord(h->yk_ev[1 + i], h->yk_stream[i]));
      MATM_CUDA(h, cudaStreamWaitEvent(st, h->yk_ev[1 + i], 0));
    }
  }
  MATM_CUDA(h, cudaEventRecord(h->yk_ev[3], st));
  h->yk_used = true;
  return MATM_OK;
}

static int launch_synth(matm_handle* h, int64_t det0, int64_t n_det, void* out_dev, int units, int apply_gain,
                        cudaStream_t st) {
  const SynthParams& P = h->P;
  if (P.n_t_atm < 4) MATM_FAIL_ARG(h, "cubic upsample needs at least 4 atmosphere-rate samples (scipy interp1d kind='cubic')");
  if (h->dtype == MATM_F64) {
    const int n_ttiles = (P.n_t_atm - 1 + Tile<double>::KT - 1) / Tile<double>::KT;
    const int64_t blocks = ((n_det + DT - 1) / DT) * n_ttiles;
    if (blocks > INT32_MAX) MATM_FAIL_ARG(h, "matm_synthesize: too many tiles; split the detector range");
    size_t smem = synth_smem<double>(P.n_layers);
    MATM_CUDA(h, cudaFuncSetAttribute(synth_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    synth_kernel<double><<<(unsigned)blocks, SY_THREADS, smem, st>>>(P, det0, n_det, (double*)out_dev, units, apply_gain);
  } else {
    return launch_synth_f32(h, det0, n_det, (float*)out_dev, units, apply_gain, st);
  }
  MATM_LAUNCHED(h);
  return MATM_OK;
}

extern "C" int matm_synthesize(matm_handle* h, int64_t det0, int64_t n_det, void* out_dev, int units, int apply_gain,
                               void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  int rc = check_ready(h, true, det0, n_det);
  if (rc) return rc;
  if (!out_dev || (units != MATM_UNITS_PW && units != MATM_UNITS_KRJ)) MATM_FAIL_ARG(h, "matm_synthesize: bad arguments");
  return launch_synth(h, det0, n_det, out_dev, units, apply_gain, (cudaStream_t)stream);
}

extern "C" int matm_calibrate(matm_handle* h, int64_t det0, int64_t n_det, const void* in_dev, void* out_dev, int to_units,
                              void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!h->have_scan || !h->have_dets || !h->have_bands) MATM_FAIL_ARG(h, "matm_calibrate: scan, detectors and bands must be set");
  if (det0 < 0 || n_det < 1 || det0 + n_det > h->P.n_det || !in_dev || !out_dev ||
      (to_units != MATM_UNITS_PW && to_units != MATM_UNITS_KRJ))
    MATM_FAIL_ARG(h, "matm_calibrate: bad arguments");
  dim3 grid((unsigned)((h->P.n_t + 255) / 256), (unsigned)min((int64_t)2048, n_det));
  if (h->dtype == MATM_F64)
    calibrate_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(h->P, det0, n_det, (const double*)in_dev, (double*)out_dev, to_units);
  else
    calibrate_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(h->P, det0, n_det, (const float*)in_dev, (float*)out_dev, to_units);
  MATM_LAUNCHED(h);
  return MATM_OK;
}

extern "C" int matm_set_tuning(matm_handle* h, int chunk_mib, int overlap) {
  MATM_CHECK_HANDLE(h);
  if (chunk_mib < 1 || chunk_mib > 65536) MATM_FAIL_ARG(h, "matm_set_tuning: chunk_mib must be in [1, 65536]");
  h->chunk_mib = chunk_mib;
  h->overlap = overlap != 0;
  return MATM_OK;
}

extern "C" int matm_set_timing(matm_handle* h, int enable) {
  MATM_CHECK_HANDLE(h);
  h->timing = enable != 0;
  h->tev_used = 0;
  return MATM_OK;
}

extern "C" int matm_get_timing(matm_handle* h, double* loading_ms, double* eval_ms, int64_t* n_chunks) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  MATM_CUDA(h, cudaDeviceSynchronize());
  double a = 0.0, b = 0.0;
  for (size_t i = 0; i + 2 < h->tev_used + 0 && i + 2 < h->tev.size() + 0; i += 3) {
    float ms = 0.f;
    MATM_CUDA(h, cudaEventElapsedTime(&ms, h->tev[i], h->tev[i + 1]));
    a += ms;
    MATM_CUDA(h, cudaEventElapsedTime(&ms, h->tev[i + 1], h->tev[i + 2]));
    b += ms;
  }
  if (loading_ms) *loading_ms = a;
  if (eval_ms) *eval_ms = b;
  if (n_chunks) *n_chunks = (int64_t)(h->tev_used / 3);
  h->tev_used = 0;
  return MATM_OK;
}

// End-to-end delivery into host memory: detector chunks, two staging buffers, copy of chunk c overlapped with
// the kernel of chunk c + 1.
extern "C" int matm_synthesize_host(matm_handle* h, int64_t det0, int64_t n_det, void* out_host, int units, int apply_gain) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  int rc = check_ready(h, true, det0, n_det);
  if (rc) return rc;
  if (!out_host) MATM_FAIL_ARG(h, "matm_synthesize_host: null output");
  const size_t esz = h->dtype == MATM_F64 ? 8 : 4;
  const size_t row = (size_t)h->P.n_t * esz;
  // chunk: a multiple of DT detectors, about 256 MiB
  int64_t chunk = (int64_t)(((size_t)256 << 20) / row);
  chunk = (chunk / DT) * DT;
  if (chunk < DT) chunk = DT;
  if (chunk > n_det) chunk = n_det;
  const size_t need = (size_t)chunk * row;
  if (h->stage_bytes < need) {
    for (int i = 0; i < 2; i++) {
      if (h->stage_dev[i]) MATM_CUDA(h, cudaFree(h->stage_dev[i]));
      h->stage_dev[i] = nullptr;
      MATM_CUDA(h, cudaMalloc(&h->stage_dev[i], need));
      if (!h->stage_stream[i]) MATM_CUDA(h, cudaStreamCreateWithFlags(&h->stage_stream[i], cudaStreamNonBlocking));
    }
    h->stage_bytes = need;
  }
  int c = 0;
  for (int64_t d = 0; d < n_det; d += chunk, c++) {
    const int b = c & 1;
    const int64_t nd = (n_det - d < chunk) ? (n_det - d) : chunk;
    cudaStream_t st = h->stage_stream[b];  // stream order protects the staging buffer from reuse before its copy ends
    rc = launch_synth(h, det0 + d, nd, h->stage_dev[b], units, apply_gain, st);
    if (rc) return rc;
    MATM_CUDA(h, cudaMemcpyAsync((char*)out_host + (size_t)d * row, h->stage_dev[b], (size_t)nd * row, cudaMemcpyDeviceToHost, st));
  }
  MATM_CUDA(h, cudaStreamSynchronize(h->stage_stream[0]));
  MATM_CUDA(h, cudaStreamSynchronize(h->stage_stream[1]));
  return MATM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant of the float synthesis (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that
// forbid it).  Same two kernels, same detector order, same results as matm_synthesize; the knot-value scratch
// yk[tile][n_t_atm][32] lives in ONE file-scope device array of MATM_NOALLOC_SCRATCH_BYTES that every GPU loading
// the library holds until exit.  Every call shares it: calls that overlap on the GPU race (the caller serialises
// them, e.g. by issuing them on one stream).  Nothing has to start at zero: loading_f32_kernel writes all 32 slots
// of every knot row of every tile of the launch (idle lanes of a partial tile shadow the last detector) before
// the evaluation kernel of the same stream reads them, so the previous call's contents are never observed.
// ---------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_noalloc_yk[MATM_NOALLOC_SCRATCH_BYTES / sizeof(float)];

#define MATM_NOALLOC_MAX_DEVICES 64
static float* g_noalloc_yk_addr[MATM_NOALLOC_MAX_DEVICES];  // device address of g_noalloc_yk, per device
static std::mutex g_noalloc_mutex;

// In host code the NAME g_noalloc_yk is a host-side stand-in, not a device address: the address a kernel can use is
// looked up once per device, with that device current (first call on the device = its set-up), and cached.
static int noalloc_scratch(matm_handle* h, float** out) {
  if (h->device < 0 || h->device >= MATM_NOALLOC_MAX_DEVICES) MATM_FAIL_ARG(h, "matm_synthesize_noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_yk_addr[h->device]) {
    void* p = nullptr;
    MATM_CUDA(h, cudaGetSymbolAddress(&p, g_noalloc_yk));
    if (!p) MATM_FAIL_ARG(h, "matm_synthesize_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_yk_addr[h->device] = static_cast<float*>(p);
  }
  *out = g_noalloc_yk_addr[h->device];
  return MATM_OK;
}

extern "C" int matm_synthesize_noalloc(matm_handle* h, int64_t det0, int64_t n_det, void* out_dev, int units, int apply_gain,
                                       void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (h->dtype != MATM_F32) MATM_FAIL_ARG(h, "matm_synthesize_noalloc: float32 handles only (the float64 kernel has no scratch)");
  int rc = check_ready(h, true, det0, n_det);
  if (rc) return rc;
  if (!out_dev || (units != MATM_UNITS_PW && units != MATM_UNITS_KRJ)) MATM_FAIL_ARG(h, "matm_synthesize_noalloc: bad arguments");
  const SynthParams& P = h->P;
  const int n = P.n_t_atm;
  if (n < 4) MATM_FAIL_ARG(h, "cubic upsample needs at least 4 atmosphere-rate samples (scipy interp1d kind='cubic')");
  const int64_t n_tiles = (n_det + F32_DT - 1) / F32_DT;
  const size_t need = (size_t)n_tiles * (size_t)n * F32_DT * sizeof(float);
  if (need > (size_t)MATM_NOALLOC_SCRATCH_BYTES) {
    h->err = "matm_synthesize_noalloc: scratch of " + std::to_string(need) + " B (ceil(n_det / 32) = " + std::to_string(n_tiles) +
             " tiles x n_t_atm = " + std::to_string(n) + " knots x 128 B) exceeds the static maximum of " +
             std::to_string((size_t)MATM_NOALLOC_SCRATCH_BYTES) + " B; split the detector range or use matm_synthesize";
    return MATM_ERR_ARG;
  }
  // the processing order of the range is built (and allocated) by matm_synthesize; here it is only looked up
  auto it = h->det_order.find(std::make_pair(det0, n_det));
  if (it == h->det_order.end())
    MATM_FAIL_ARG(h, "matm_synthesize_noalloc: no cached processing order for this detector range; call matm_synthesize once for "
                     "the same (det0, n_det) at set-up (it allocates the order), then this entry point allocates nothing");
  const int32_t* order = it->second.order;
  const uint8_t* tile_m = it->second.tile_m;
  float* yk = nullptr;
  rc = noalloc_scratch(h, &yk);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int n_kb = (n + LD_KA - 1) / LD_KA, n_tt = (n - 1 + F32_KT - 1) / F32_KT;
  if (n_tiles * (int64_t)max(n_kb, n_tt) > INT32_MAX) MATM_FAIL_ARG(h, "matm_synthesize_noalloc: too many tiles");
  const size_t smem = synth_f32_smem();
  const bool fast = h->min_samples_per_interval >= 4 && P.n_t < ((int64_t)1 << 30) && n_det < ((int64_t)1 << 31);
  auto eval_kernel = fast ? synth_eval_f32_kernel : synth_eval_general_f32_kernel;
  MATM_CUDA(h, cudaFuncSetAttribute(eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const size_t smem_a = (size_t)P.n_layers * LD_KA * sizeof(int4);
  MATM_CUDA(h, cudaFuncSetAttribute(loading_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_a));
  {  // pair-expanded copy of the smoothed layers (buffers allocated by matm_set_layers)
    int max_cells = 1;
    for (int l = 0; l < P.n_layers; l++) max_cells = max(max_cells, P.layers[l].n_ext * P.layers[l].n_cs);
    dim3 grid((unsigned)min(4 * h->sm_count, (max_cells + 255) / 256), (unsigned)P.n_layers);
    expand_layers_kernel<<<grid, 256, 0, st>>>(P);
    MATM_LAUNCHED(h);
  }
  loading_f32_kernel<<<(unsigned)(n_tiles * n_kb), LD_THREADS, smem_a, st>>>(P, order, tile_m, 0, n_det, yk, apply_gain);
  MATM_LAUNCHED(h);
  eval_kernel<<<(unsigned)(n_tiles * n_tt), F32_THREADS, smem, st>>>(P, order, 0, det0, n_det, yk, (float*)out_dev, units);
  MATM_LAUNCHED(h);
  return MATM_OK;
}
