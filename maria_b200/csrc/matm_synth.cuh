// Device-side building blocks of the sampler (a6, a7, a9, a11) shared by the zenith-PWV kernel and the fused
// synthesis kernel.  All arithmetic is templated on R (float: production, double: validation).
#pragma once
#include "matm_common.cuh"

template <typename R>
struct Vec2 {
  R x, y;
};

template <typename R>
struct Knot {  // per atmosphere-rate sample: boresight rotation, e = el_b - pi/2, a = az_b
  R ce, se, ca, sa;
};

template <typename R>
struct LayerKnot {  // boresight position in a layer grid, split into integer cell and fraction
  int ix, iy;
  R fx, fy;
};

__device__ __forceinline__ float r_floor(float x) { return floorf(x); }
__device__ __forceinline__ double r_floor(double x) { return floor(x); }
__device__ __forceinline__ float r_asin(float x) { return asinf(x); }
__device__ __forceinline__ double r_asin(double x) { return asin(x); }
template <typename R>
__device__ __forceinline__ R r_nan() {
  return (R)__longlong_as_double(0x7ff8000000000000LL);
}

__device__ __forceinline__ void record_out_of_grid(int* flags, int layer, int64_t det, int64_t sample) {
  atomicOr(&flags[0], 1);
  if (atomicCAS(&flags[6], 0, 1) == 0) {
    flags[1] = layer;
    flags[2] = (int)(uint32_t)((uint64_t)det & 0xffffffffu);
    flags[3] = (int)(uint32_t)((uint64_t)det >> 32);
    flags[4] = (int)(uint32_t)((uint64_t)sample & 0xffffffffu);
    flags[5] = (int)(uint32_t)((uint64_t)sample >> 32);
  }
}

// Stage the per-knot constants of the window [wa, wa + Wn) into shared memory.
//   knots : (Wn,)            boresight rotation
//   lk    : (n_layers, Wn)   split layer-grid coordinates of the boresight
template <typename R>
__device__ __forceinline__ void load_window(const SynthParams& P, int wa, int Wn, int Wstride, Knot<R>* knots,
                                            LayerKnot<R>* lk) {
  for (int k = threadIdx.x; k < Wn; k += blockDim.x) {
    const double* b = P.bore_atm + 4 * (size_t)(wa + k);
    knots[k].ce = (R)b[0];
    knots[k].se = (R)b[1];
    knots[k].ca = (R)b[2];
    knots[k].sa = (R)b[3];
  }
  for (int e = threadIdx.x; e < P.n_layers * Wn; e += blockDim.x) {
    int l = e / Wn, k = e % Wn;
    const double* b = P.layers[l].base + 2 * (size_t)(wa + k);
    double cx = b[0], cy = b[1];
    double flx = floor(cx), fly = floor(cy);
    // clamp the integer part so that absurd (far out-of-grid) positions cannot overflow; they are flagged later
    flx = fmin(fmax(flx, -1.0e9), 1.0e9);
    fly = fmin(fmax(fly, -1.0e9), 1.0e9);
    LayerKnot<R> o;
    o.ix = (int)flx;
    o.iy = (int)fly;
    o.fx = (R)(cx - flx);
    o.fy = (R)(cy - fly);
    lk[l * Wstride + k] = o;
  }
}

// jax RegularGridInterpolator axis lookup (linear): i = clip(searchsorted(g, x) - 1, 0, n - 2)
template <typename R>
__device__ __forceinline__ bool locate(const double* __restrict__ g, int n, R x, int* idx, R* w) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if ((R)__ldg(g + mid) < x)
      lo = mid + 1;
    else
      hi = mid;
  }
  int i = min(max(lo - 1, 0), n - 2);
  R g0 = (R)__ldg(g + i), g1 = (R)__ldg(g + i + 1);
  *idx = i;
  *w = (x - g0) / (g1 - g0);
  return !(x < (R)__ldg(g) || x > (R)__ldg(g + n - 1)) && (x == x);
}

// a6 + a7 + a9 for one (detector, atmosphere-rate sample): returns zenith_scaled_pwv and sin(el) of the detector.
//   (u, Y, wm1): detector direction in the boresight frame, (sin r cos p, sin r sin p, cos r - 1)
template <typename R>
__device__ __forceinline__ R sample_zenith_pwv(const SynthParams& P, R u, R Y, R wm1, const Knot<R>& kc,
                                               const LayerKnot<R>* lk, int Wstride, int64_t det, int64_t knot,
                                               R* sin_el) {
  // direction of the detector minus direction of the boresight, in the horizon frame (coords/transforms.py:10-29)
  const R dX = u * kc.ce - wm1 * kc.se;
  const R dZ = u * kc.se + wm1 * kc.ce;
  const R dNx = dX * kc.ca - Y * kc.sa;
  const R dNy = dX * kc.sa + Y * kc.ca;
  const R Nbx = -kc.se * kc.ca, Nby = -kc.se * kc.sa, Nbz = kc.ce;
  const R Nz = Nbz + dZ;
  // pp = (cos az / tan el, sin az / tan el) = (Nx / Nz, Ny / Nz)  (coordinates.py:333-349); difference to the boresight
  const R inv = (R)1 / (Nbz * Nz);
  const R dppx = (dNx * Nbz - Nbx * dZ) * inv;
  const R dppy = (dNy * Nbz - Nby * dZ) * inv;
  *sin_el = Nz;
  const bool f64 = sizeof(R) == 8;
  R acc = f64 ? (R)P.pwv0 : (R)0;
  for (int l = 0; l < P.n_layers; l++) {
    const LayerDev& L = P.layers[l];
    const LayerKnot<R> b = lk[l * Wstride];
    R fx = b.fx + ((R)L.D[0] * dppx + (R)L.D[1] * dppy);
    R fy = b.fy + ((R)L.D[2] * dppx + (R)L.D[3] * dppy);
    R flx = r_floor(fx), fly = r_floor(fy);
    int ix = b.ix + (int)flx, iy = b.iy + (int)fly;
    R wx = fx - flx, wy = fy - fly;
    if (ix == L.n_ext - 1 && wx == (R)0) {  // x == grid[-1] is inside (RGI clips the index to n - 2)
      ix--;
      wx = (R)1;
    }
    if (iy == L.n_cs - 1 && wy == (R)0) {
      iy--;
      wy = (R)1;
    }
    R y;
    if (ix < 0 || ix > L.n_ext - 2 || iy < 0 || iy > L.n_cs - 2 || !(fx == fx) || !(fy == fy)) {
      record_out_of_grid(P.flags, l, det, knot);  // atmosphere.py:368-369
      y = r_nan<R>();
    } else {
      const R* v = reinterpret_cast<const R*>(L.data) + (size_t)ix * L.n_cs + iy;
      const R v00 = __ldg(v), v01 = __ldg(v + 1), v10 = __ldg(v + L.n_cs), v11 = __ldg(v + L.n_cs + 1);
      const R ux = (R)1 - wx, uy = (R)1 - wy;
      y = v00 * (ux * uy) + v01 * (ux * wy) + v10 * (wx * uy) + v11 * (wx * wy);
    }
    acc += (R)L.pwv_rms * y;
  }
  return f64 ? acc : (R)P.pwv0 + acc;
}

// a11 + a12: emission lookup at (T0, pwv, el) on the folded (pwv, el) grid of the detector's band, times M[0,0]
template <typename R>
__device__ __forceinline__ R loading_pW(const SynthParams& P, int band, R mueller, R zpwv, R sin_el) {
  R el = r_asin(sin_el);  // <= pi/2 by construction, so the reference's clip(max=pi/2) is the identity
  int ip, ie;
  R wp, we;
  bool ok = locate<R>(P.side_pwv, P.npwv, zpwv, &ip, &wp);
  ok = locate<R>(P.side_el, P.nel, el, &ie, &we) && ok;
  if (!ok || !(P.table_ok & 1)) {
    if (zpwv == zpwv) atomicOr(&P.flags[0], 2);  // NaN from an out-of-grid layer is already flagged
    return r_nan<R>();
  }
  const double* g = P.G2 + ((size_t)band * P.npwv + ip) * P.nel + ie;
  const R g00 = (R)__ldg(g), g01 = (R)__ldg(g + 1), g10 = (R)__ldg(g + P.nel), g11 = (R)__ldg(g + P.nel + 1);
  const R up = (R)1 - wp, ue = (R)1 - we;
  return mueller * (g00 * (up * ue) + g01 * (up * we) + g10 * (wp * ue) + g11 * (wp * we));
}

// a14: 1-D lookup of the folded transmission integral at the full-rate elevation; returns factor k_B 1e12 H
template <typename R>
__device__ __forceinline__ R cal_denominator(const SynthParams& P, int band, R sin_el) {
  R el = r_asin(sin_el);
  int ie;
  R we;
  bool ok = locate<R>(P.side_el, P.nel, el, &ie, &we);
  if (!ok || !(P.table_ok & 2)) {
    atomicOr(&P.flags[0], 2);
    return r_nan<R>();
  }
  const double* hh = P.H1 + (size_t)band * P.nel + ie;
  return (R)__ldg(hh) * ((R)1 - we) + (R)__ldg(hh + 1) * we;
}
