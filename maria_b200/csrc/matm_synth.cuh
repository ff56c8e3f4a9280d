// Device-side building blocks of the sampler (a6, a7, a9, a11, a14) shared by the zenith-PWV kernel and the fused
// synthesis kernel.  All arithmetic is templated on R (float: production, double: validation).
#pragma once
#include "matm_common.cuh"

template <typename R>
struct Knot {  // per atmosphere-rate sample: boresight rotation, e = el_b - pi/2, a = az_b
  R ce, se, ca, sa;
};

template <typename R>
struct LayerKnot {  // boresight position in a layer grid, split into integer cell and fraction
  int ix, iy;
  R fx, fy;
};

template <typename R>
struct LayerS {  // per-layer constants in the kernel's arithmetic type (shared-memory copy of LayerDev)
  R D0, D1, D2, D3, rms;
  int n_ext, n_cs;
  const R* data;
};

__device__ __forceinline__ float r_floor(float x) { return floorf(x); }
__device__ __forceinline__ double r_floor(double x) { return floor(x); }
__device__ __forceinline__ double r_asin01(double x) { return asin(x); }
// asin on [-1, 1] for float: pi/2 - sqrt(1 - |x|) * P7(|x|)  (Abramowitz & Stegun 4.4.46, |error| <= 2e-8 rad, i.e. below
// one float ulp of an elevation); about half the instructions of asinf.
__device__ __forceinline__ float r_asin01(float x) {
  const float a = fabsf(x);
  float p = -0.0012624911f;
  p = fmaf(p, a, 0.0066700901f);
  p = fmaf(p, a, -0.0170881256f);
  p = fmaf(p, a, 0.0308918810f);
  p = fmaf(p, a, -0.0501743046f);
  p = fmaf(p, a, 0.0889789874f);
  p = fmaf(p, a, -0.2145988016f);
  p = fmaf(p, a, 1.5707963050f);
  const float r = 1.5707963268f - sqrtf(1.0f - a) * p;
  return copysignf(r, x);
}
__device__ __forceinline__ float r_div(float a, float b) { return __fdividef(a, b); }  // <= 2 ulp
__device__ __forceinline__ double r_div(double a, double b) { return a / b; }
template <typename R>
__device__ __forceinline__ R r_nan() {
  return (R)__longlong_as_double(0x7ff8000000000000LL);
}

static __device__ __noinline__ void record_out_of_grid(int* flags, int layer, int64_t det, int64_t sample) {
  atomicOr(&flags[0], 1);
  if (atomicCAS(&flags[6], 0, 1) == 0) {
    flags[1] = layer;
    flags[2] = (int)(uint32_t)((uint64_t)det & 0xffffffffu);
    flags[3] = (int)(uint32_t)((uint64_t)det >> 32);
    flags[4] = (int)(uint32_t)((uint64_t)sample & 0xffffffffu);
    flags[5] = (int)(uint32_t)((uint64_t)sample >> 32);
  }
}

// Stage the per-knot constants of the window [wa, wa + Wn) and the per-layer constants into shared memory.
template <typename R>
__device__ __forceinline__ void load_window(const SynthParams& P, int wa, int Wn, int Wstride, Knot<R>* knots,
                                            LayerKnot<R>* lk, LayerS<R>* ls) {
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
  if (threadIdx.x < P.n_layers) {
    const LayerDev& L = P.layers[threadIdx.x];
    LayerS<R> o;
    o.D0 = (R)L.D[0];
    o.D1 = (R)L.D[1];
    o.D2 = (R)L.D[2];
    o.D3 = (R)L.D[3];
    o.rms = (R)L.pwv_rms;
    o.n_ext = L.n_ext;
    o.n_cs = L.n_cs;
    o.data = reinterpret_cast<const R*>(L.data);
    ls[threadIdx.x] = o;
  }
}

// Interval of x in an increasing grid with np.searchsorted(g, x) - 1 semantics (g[i] < x <= g[i+1]), clipped to
// [0, n - 2] like jax's RegularGridInterpolator; `hint` is the previous answer (the inputs vary slowly).
template <typename R>
__device__ __forceinline__ int find_interval(const R* __restrict__ g, int n, R x, int hint) {
  int i = hint;
  while (i > 0 && !(__ldg(g + i) < x)) i--;
  while (i < n - 2 && __ldg(g + i + 1) < x) i++;
  return i;
}

// a6 + a7 + a9 for one (detector, atmosphere-rate sample): returns zenith_scaled_pwv and sin(el) of the detector.
//   (u, Y, wm1): detector direction in the boresight frame, (sin r cos p, sin r sin p, cos r - 1)
template <typename R>
__device__ __forceinline__ R sample_zenith_pwv(const SynthParams& P, int n_layers, const LayerS<R>* __restrict__ ls, R u,
                                               R Y, R wm1, const Knot<R>& kc, const LayerKnot<R>* __restrict__ lk,
                                               int Wstride, int64_t det, int64_t knot, R* sin_el) {
  // direction of the detector minus direction of the boresight, in the horizon frame (coords/transforms.py:10-29)
  const R dX = u * kc.ce - wm1 * kc.se;
  const R dZ = u * kc.se + wm1 * kc.ce;
  const R dNx = dX * kc.ca - Y * kc.sa;
  const R dNy = dX * kc.sa + Y * kc.ca;
  const R Nbx = -kc.se * kc.ca, Nby = -kc.se * kc.sa, Nbz = kc.ce;
  const R Nz = Nbz + dZ;
  // pp = (cos az / tan el, sin az / tan el) = (Nx / Nz, Ny / Nz)  (coordinates.py:333-349); difference to the boresight
  const R inv = r_div((R)1, Nbz * Nz);
  const R dppx = (dNx * Nbz - Nbx * dZ) * inv;
  const R dppy = (dNy * Nbz - Nby * dZ) * inv;
  *sin_el = Nz;
  const bool f64 = sizeof(R) == 8;
  R acc = f64 ? (R)P.pwv0 : (R)0;
  for (int l = 0; l < n_layers; l++) {
    const LayerS<R> L = ls[l];
    const LayerKnot<R> b = lk[l * Wstride];
    const R fx = b.fx + (L.D0 * dppx + L.D1 * dppy);
    const R fy = b.fy + (L.D2 * dppx + L.D3 * dppy);
    const R flx = r_floor(fx), fly = r_floor(fy);
    int ix = b.ix + (int)flx, iy = b.iy + (int)fly;
    R wx = fx - flx, wy = fy - fly;
    R y;
    if ((unsigned)ix >= (unsigned)(L.n_ext - 1) || (unsigned)iy >= (unsigned)(L.n_cs - 1)) {
      // rare path: exactly on the last grid line is inside (RGI clips the index to n - 2); anything else is outside
      if (ix == L.n_ext - 1 && wx == (R)0) {
        ix--;
        wx = (R)1;
      }
      if (iy == L.n_cs - 1 && wy == (R)0) {
        iy--;
        wy = (R)1;
      }
    }
    if ((unsigned)ix >= (unsigned)(L.n_ext - 1) || (unsigned)iy >= (unsigned)(L.n_cs - 1) || !(fx == fx) || !(fy == fy)) {
      record_out_of_grid(P.flags, l, det, knot);  // atmosphere.py:368-369
      y = r_nan<R>();
    } else {
      const R* v = L.data + (size_t)ix * L.n_cs + iy;
      const R v00 = __ldg(v), v01 = __ldg(v + 1), v10 = __ldg(v + L.n_cs), v11 = __ldg(v + L.n_cs + 1);
      const R ux = (R)1 - wx, uy = (R)1 - wy;
      y = v00 * (ux * uy) + v01 * (ux * wy) + v10 * (wx * uy) + v11 * (wx * wy);
    }
    acc += L.rms * y;
  }
  return f64 ? acc : (R)P.pwv0 + acc;
}

// a11 + a12: emission lookup at (T0, pwv, el) on the folded (pwv, el) grid of the detector's band, times M[0,0].
// ip / ie carry the previous intervals.
template <typename R>
__device__ __forceinline__ R loading_pW(const SynthParams& P, int band, R mueller, R zpwv, R sin_el, int& ip, int& ie) {
  const R* gp = reinterpret_cast<const R*>(P.pwv_R);
  const R* ge = reinterpret_cast<const R*>(P.el_R);
  const R el = r_asin01(sin_el);  // <= pi/2 by construction, so the reference's clip(max=pi/2) is the identity
  ip = find_interval<R>(gp, P.npwv, zpwv, ip);
  ie = find_interval<R>(ge, P.nel, el, ie);
  const R p0 = __ldg(gp + ip), p1 = __ldg(gp + ip + 1), e0 = __ldg(ge + ie), e1 = __ldg(ge + ie + 1);
  const bool ok = !(zpwv < __ldg(gp) || zpwv > __ldg(gp + P.npwv - 1) || el < __ldg(ge) || el > __ldg(ge + P.nel - 1)) &&
                  (zpwv == zpwv) && (P.table_ok & 1);
  if (!ok) {
    if (zpwv == zpwv) atomicOr(&P.flags[0], 2);  // NaN from an out-of-grid layer is already flagged
    return r_nan<R>();
  }
  const R wp = r_div(zpwv - p0, p1 - p0), we = r_div(el - e0, e1 - e0);
  const R* g = reinterpret_cast<const R*>(P.G2_R) + ((size_t)band * P.npwv + ip) * P.nel + ie;
  const R g00 = __ldg(g), g01 = __ldg(g + 1), g10 = __ldg(g + P.nel), g11 = __ldg(g + P.nel + 1);
  const R up = (R)1 - wp, ue = (R)1 - we;
  return mueller * (g00 * (up * ue) + g01 * (up * we) + g10 * (wp * ue) + g11 * (wp * we));
}

// a14: factor k_B 1e12 H at the full-rate elevation: the lerp of the folded transmission integral on the el grid,
// evaluated as Hc[i] + Hs[i] * el on the interval found by comparing sin(el) with sin(grid).
template <typename R>
__device__ __forceinline__ R cal_denominator(const SynthParams& P, const R* __restrict__ sing, const R* __restrict__ hc,
                                             const R* __restrict__ hs, R sin_el, int& ie) {
  ie = find_interval<R>(sing, P.nel, sin_el, ie);
  const R el = r_asin01(sin_el);
  R den = fma(__ldg(hs + ie), el, __ldg(hc + ie));
  if (sin_el < __ldg(sing) || !(P.table_ok & 2) || !(sin_el == sin_el)) {
    atomicOr(&P.flags[0], 2);
    den = r_nan<R>();
  }
  return den;
}
