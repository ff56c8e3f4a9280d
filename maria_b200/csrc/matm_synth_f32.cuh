// The production (float) instance of the fused synthesis kernel a6..a14, written for instruction count: the kernel is
// issue- and data-pipe-bound, not HBM-bound (ncu: ~140 thread-instructions per output sample in the generic template,
// ~37 here), so every per-output-sample operation is stripped to the minimum and done on two detectors at once with
// the packed fma.rn.f32x2 / mul / add instructions of sm_100.
//
//   phase 1  loading [pW] at the window's knots; lane = focal-plane position (co-located detectors share the line of
//            sight: tile multiplicity M), warps stride over knots, F32_G knots per thread                (per knot)
//   phase 2  not-a-knot cubic spline second derivatives by the windowed Thomas sweep                     (per knot)
//   phase 3  spline evaluation + pW -> K_RJ at the full-rate samples; thread = F32_S samples, loop over   (per sample)
//            detector PAIRS: 2 LDS.128 + 4 packed FMAs for the spline; the calibration factor is
//            rcp(Hc + Hs asin(sin el)) in packed arithmetic when the tile lies in one interval of the elevation
//            grid, else the t = sqrt(1 - sin el) table of matm_setup.cu (no search, no divide).
#pragma once
#include <type_traits>
#include "matm_synth.cuh"

#define F32_DT 32
#define F32_THREADS 256
#ifndef F32_KT
// spline intervals per tile.  Both per-tile loops are quantised: phase 1 works in passes of 8 M F32_G knots over the
// KT + 1 + 2 HALO knots of the window, phase 3 in iterations of 256 F32_S samples over the KT U samples of the tile
// (U = full-rate samples per interval).  135 fills both at the C4 cadence (U = 22.6: 156 knots = 2.5 passes of 64,
// 3056 samples = 3.98 iterations of 768) and keeps three CTAs per SM; any value is correct (-DF32_KT=... retunes).
#define F32_KT 135
#endif
#ifndef F32_HALO
#define F32_HALO 10  // extra knots each side; the error of the artificial boundary decays as (2 - sqrt 3)^HALO
#endif
#ifndef F32_MINB
#define F32_MINB 3   // CTAs per SM the register allocation is held to
#endif
#define F32_W (F32_KT + 1 + 2 * F32_HALO)
#define F32_NP (F32_DT / 2)  // detector pairs per tile
#ifndef F32_KS
#define F32_KS F32_NP  // float4 entries per knot row of ym4 (padding the rows by one entry was measured: no gain, 5.21 vs 5.09 ms)
#endif
#define F32_SPLN ((5 * F32_W + 3) & ~3)  // floats reserved for the spline system (keeps what follows 16-byte aligned)
#ifndef F32_S
#define F32_S 3  // full-rate samples per thread and iteration in phase 3 (4 spills at 80 registers)
#endif
#ifndef F32_G
#define F32_G 4  // knots per thread and iteration in phase 1
#endif

struct LayerF {  // per-layer constants, two 16-byte shared-memory loads + the pointer
  float D0, D1, D2, D3;
  float rms;
  int ne1, nc1, n_cs;  // n_ext - 1, n_cs - 1, n_cs
  const float* data;
  int pad[2];
};

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float fast_sqrt(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// *(sample_ptr + row_bytes) = v: the per-thread pointer already holds out + j, the row offset is the tile's constant
__device__ __forceinline__ void st_at(const char* sample_ptr, long long row_bytes, float v) {
  const char* a;
  asm("add.s64 %0, %1, %2;" : "=l"(a) : "l"(sample_ptr), "l"(row_bytes));
  asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(a), "f"(v) : "memory");
}
// asin on [-1, 1]: pi/2 - sqrt(1 - |x|) P7(|x|)  (Abramowitz & Stegun 4.4.46, |error| <= 2e-8 rad)
__device__ __forceinline__ float fast_asin(float x) {
  const float a = fabsf(x);
  float p = -0.0012624911f;
  p = fmaf(p, a, 0.0066700901f);
  p = fmaf(p, a, -0.0170881256f);
  p = fmaf(p, a, 0.0308918810f);
  p = fmaf(p, a, -0.0501743046f);
  p = fmaf(p, a, 0.0889789874f);
  p = fmaf(p, a, -0.2145988016f);
  p = fmaf(p, a, 1.5707963050f);
  const float r = fmaf(-fast_sqrt(fmaxf(1.0f - a, 0.0f)), p, 1.5707963268f);
  return copysignf(r, x);
}

// the same for two arguments in [0, 1] at once, given sq = sqrt(1 - x): pi/2 - sq P7(x), the polynomial in packed f32x2
__device__ __forceinline__ float2 asin01_f32x2(float2 x, float2 sq) {
  float2 p = f2(0.0012624911f, 0.0012624911f);  // -P7: the subtraction becomes the last FMA
  p = __ffma2_rn(p, x, f2(-0.0066700901f, -0.0066700901f));
  p = __ffma2_rn(p, x, f2(0.0170881256f, 0.0170881256f));
  p = __ffma2_rn(p, x, f2(-0.0308918810f, -0.0308918810f));
  p = __ffma2_rn(p, x, f2(0.0501743046f, 0.0501743046f));
  p = __ffma2_rn(p, x, f2(-0.0889789874f, -0.0889789874f));
  p = __ffma2_rn(p, x, f2(0.2145988016f, 0.2145988016f));
  p = __ffma2_rn(p, x, f2(-1.5707963050f, -1.5707963050f));
  return __ffma2_rn(sq, p, f2(1.5707963268f, 1.5707963268f));
}

// The calibration factor in a special bin of the table (calq_kind != 0): a bin holding one grid point of the elevation
// table has two lines, selected by the kink position; anything else goes through the exact searching code, which also
// raises the flag when the elevation leaves the table.
__device__ __noinline__ float cal_factor_special(const SynthParams& P, int idx, int band, float sin_el, float dt) {
  const int kind = P.calq_kind[idx];
  if (kind > 0) {
    const float4 a = __ldg(P.calq_aux + 2 * kind), k = __ldg(P.calq_aux + 2 * kind + 1);
    return dt > k.x ? fmaf(a.w, dt, a.z) : fmaf(a.y, dt, a.x);
  }
  int ie = 0;
  const int ho = band * (P.nel - 1);
  const float den = cal_denominator<float>(P, reinterpret_cast<const float*>(P.sinel_R),
                                           reinterpret_cast<const float*>(P.Hc_R) + ho,
                                           reinterpret_cast<const float*>(P.Hs_R) + ho, sin_el, ie);
  return 1.0f / den;
}

// a11 + a12 for this kernel, in two parts.  The cell of the folded (pwv, el) grid and the two weights depend on the line of
// sight only, so co-located detectors share them; the band's table is then one bilinear blend per detector.  The grid
// values met during the search are the cell's corners (no second fetch), the table ends come from the parameter block.
// Same interval as find_interval (np.searchsorted - 1, clipped), same NaN / flag behaviour as loading_pW.
__device__ __forceinline__ bool emission_cell_f32(const SynthParams& P, float zpwv, float sin_el, int& ip, int& ie, float& wp,
                                                  float& we) {
  const float* __restrict__ gp = reinterpret_cast<const float*>(P.pwv_R);
  const float* __restrict__ ge = reinterpret_cast<const float*>(P.el_R);
  const float el = fast_asin(sin_el);  // <= pi/2 by construction, so the reference's clip(max=pi/2) is the identity
  float p0 = __ldg(gp + ip), p1 = __ldg(gp + ip + 1);
  while (ip > 0 && !(p0 < zpwv)) {
    ip--;
    p1 = p0;
    p0 = __ldg(gp + ip);
  }
  while (ip < P.npwv - 2 && p1 < zpwv) {
    ip++;
    p0 = p1;
    p1 = __ldg(gp + ip + 1);
  }
  float e0 = __ldg(ge + ie), e1 = __ldg(ge + ie + 1);
  while (ie > 0 && !(e0 < el)) {
    ie--;
    e1 = e0;
    e0 = __ldg(ge + ie);
  }
  while (ie < P.nel - 2 && e1 < el) {
    ie++;
    e0 = e1;
    e1 = __ldg(ge + ie + 1);
  }
  const bool ok = !(zpwv < P.pwv_lo_f || zpwv > P.pwv_hi_f || el < P.el_lo_f || el > P.el_hi_f) && (zpwv == zpwv) &&
                  (P.table_ok & 1);
  if (!ok) {
    if (zpwv == zpwv) atomicOr(&P.flags[0], 2);  // NaN from an out-of-grid layer is already flagged
    return false;
  }
  wp = (zpwv - p0) * fast_rcp(p1 - p0);
  we = (el - e0) * fast_rcp(e1 - e0);
  return true;
}

__device__ __forceinline__ float emission_blend_f32(const SynthParams& P, int band, int ip, int ie, float wp, float we) {
  const float* __restrict__ g = reinterpret_cast<const float*>(P.G2_R) + ((band * P.npwv + ip) * P.nel + ie);
  const float g00 = __ldg(g), g01 = __ldg(g + 1), g10 = __ldg(g + P.nel), g11 = __ldg(g + P.nel + 1);
  const float a = fmaf(we, g01 - g00, g00), b = fmaf(we, g11 - g10, g10);
  return fmaf(wp, b - a, a);
}

// Phase 1 of the fused kernel: loading [pW] at the window's knots, for a tile whose detectors come in groups of M
// co-located ones (dichroic pixels, both polarisations: same offsets, hence the same line of sight through every layer).
// Each thread works on F32_G knots at once: F32_G independent gather chains per layer keep 4 F32_G loads in flight (the
// layers live in L2; a dependent chain per knot leaves the SM waiting on that latency), and the per-layer constants are
// fetched once per group.
template <int M>
__device__ __forceinline__ void phase1_f32(const SynthParams& P, const int32_t* __restrict__ order, int64_t dbase, int nd, int wa,
                                           int Wn, int n_layers, const Knot<float>* knots, const LayerKnot<float>* lk,
                                           const LayerF* ls, float* ymf, const float* dc_gain, const int* dc_band,
                                           const float* dc_mu, int lane, int warp) {
  {
    // M detectors share every line of sight: lanes [0, 32 / M) of a warp are the tile's positions, the other lanes take
    // further knots (M "virtual warps" per warp), and each lane finishes M detectors (band tables, Mueller, gain)
    constexpr int PL = 32 / M;
    const int slot0 = (lane % PL) * M, sub = lane / PL;
    const int64_t d = order[dbase + min(slot0, nd - 1)];  // idle lanes of a partial tile shadow the last detector
    const double* da = P.det_atm + 4 * d;
    const float u = (float)da[0], Y = (float)da[1], wm1 = (float)da[2];
    const float pwv0 = (float)P.pwv0;
    int ip = 0, ie = 0;
    constexpr int NW = (F32_THREADS / 32) * M;
    // one pass = G knots per thread, NW * G knots per CTA.  Full passes use F32_G; what is left at the end of the window
    // takes the smallest G that covers it, so a window of 148 knots costs 2.5 passes of 64 (M = 2) instead of 3.
    auto pass = [&](auto gtag, const int kb) {
      constexpr int G = decltype(gtag)::value;
      float dppx[G], dppy[G], sinel[G], acc[G];
      int kk[G];
      bool bad = false;
#pragma unroll
      for (int q = 0; q < G; q++) {
        kk[q] = min(kb + q * NW, Wn - 1);  // a short last group repeats the last knot (same value, same slot)
        const Knot<float> kc = knots[kk[q]];
        // direction of the detector minus direction of the boresight, horizon frame (coords/transforms.py:10-29)
        const float dX = u * kc.ce - wm1 * kc.se;
        const float dZ = u * kc.se + wm1 * kc.ce;
        const float dNx = dX * kc.ca - Y * kc.sa;
        const float dNy = dX * kc.sa + Y * kc.ca;
        const float Nbx = -kc.se * kc.ca, Nby = -kc.se * kc.sa, Nbz = kc.ce;
        const float Nz = Nbz + dZ;  // = sin(el) of the detector
        const float inv = fast_rcp(Nbz * Nz);
        dppx[q] = (dNx * Nbz - Nbx * dZ) * inv;  // pp_d - pp_b (coordinates.py:333-349)
        dppy[q] = (dNy * Nbz - Nby * dZ) * inv;
        sinel[q] = Nz;
        acc[q] = 0.f;
        bad = bad || !(dppx[q] == dppx[q]) || !(dppy[q] == dppy[q]);
      }
      unsigned badmask = 0;
      for (int l = 0; l < n_layers; l++) {
        const float4 Dm = *reinterpret_cast<const float4*>(&ls[l].D0);
        const int4 Ln = *reinterpret_cast<const int4*>(&ls[l].rms);
        const float* data = ls[l].data;
        float wx[G], wy[G];
        const float* v[G];
#pragma unroll
        for (int q = 0; q < G; q++) {
          const LayerKnot<float> b = lk[l * F32_W + kk[q]];
          const float fx = fmaf(Dm.x, dppx[q], fmaf(Dm.y, dppy[q], b.fx));
          const float fy = fmaf(Dm.z, dppx[q], fmaf(Dm.w, dppy[q], b.fy));
          const float flx = floorf(fx), fly = floorf(fy);
          int ix = b.ix + (int)flx, iy = b.iy + (int)fly;
          wx[q] = fx - flx;
          wy[q] = fy - fly;
          if ((unsigned)ix >= (unsigned)Ln.y || (unsigned)iy >= (unsigned)Ln.z) {
            // rare: exactly on the last grid line is inside (RGI clips the index to n - 2); anything else is outside
            if (ix == Ln.y && wx[q] == 0.f) {
              ix--;
              wx[q] = 1.f;
            }
            if (iy == Ln.z && wy[q] == 0.f) {
              iy--;
              wy[q] = 1.f;
            }
            if ((unsigned)ix >= (unsigned)Ln.y || (unsigned)iy >= (unsigned)Ln.z) {
              record_out_of_grid(P.flags, l, d, wa + kk[q]);  // atmosphere.py:368-369
              badmask |= 1u << q;
              ix = iy = 0;
            }
          }
          v[q] = data + (ix * Ln.w + iy);
        }
        float v00[G], v01[G], v10[G], v11[G];
#pragma unroll
        for (int q = 0; q < G; q++) {
          v00[q] = __ldg(v[q]);
          v01[q] = __ldg(v[q] + 1);
          v10[q] = __ldg(v[q] + Ln.w);
          v11[q] = __ldg(v[q] + Ln.w + 1);
        }
#pragma unroll
        for (int q = 0; q < G; q++) {
          const float a = fmaf(wy[q], v01[q] - v00[q], v00[q]), bb = fmaf(wy[q], v11[q] - v10[q], v10[q]);
          acc[q] = fmaf(__int_as_float(Ln.x), fmaf(wx[q], bb - a, a), acc[q]);
        }
      }
#pragma unroll
      for (int q = 0; q < G; q++) {
        if (bad || ((badmask >> q) & 1)) {
          if (!(dppx[q] == dppx[q]) || !(dppy[q] == dppy[q])) record_out_of_grid(P.flags, 0, d, wa + kk[q]);
          if (!(dppx[q] == dppx[q]) || !(dppy[q] == dppy[q]) || ((badmask >> q) & 1)) acc[q] = r_nan<float>();
        }
        float wp, we;
        const bool ok = emission_cell_f32(P, pwv0 + acc[q], sinel[q], ip, ie, wp, we);
#pragma unroll
        for (int j = 0; j < M; j++) {
          const int sl = slot0 + j;
          // times Mueller [0,0]; the spline being linear in the data, the detector gain goes in here too
          const float y = ok ? emission_blend_f32(P, dc_band[sl], ip, ie, wp, we) * dc_mu[sl] * dc_gain[sl] : r_nan<float>();
          if (kb + q * NW < Wn) ymf[(kk[q] * F32_KS + (sl >> 1)) * 4 + (sl & 1)] = y;
        }
      }
    };
    const int vw = warp * M + sub;
    for (int base = 0; base < Wn;) {
      const int rem = Wn - base;
      if (rem > (F32_G / 2) * NW) {
        pass(std::integral_constant<int, F32_G>{}, base + vw);
        base += F32_G * NW;
      } else if (rem > NW) {
        pass(std::integral_constant<int, F32_G / 2>{}, base + vw);
        base += (F32_G / 2) * NW;
      } else {
        pass(std::integral_constant<int, 1>{}, base + vw);
        base += NW;
      }
    }
  }
}

__global__ void __launch_bounds__(F32_THREADS, F32_MINB)
    synth_f32_kernel(const __grid_constant__ SynthParams P, const int32_t* __restrict__ order,
                     const uint8_t* __restrict__ tile_m, int64_t det0, int64_t n_det,
                     float* __restrict__ out, int units, int apply_gain) {
  extern __shared__ __align__(16) unsigned char dyn[];
  float4* ym4 = reinterpret_cast<float4*>(dyn);                      // [W][NP] {y_d0, y_d1, m_d0, m_d1}
  float* ymf = reinterpret_cast<float*>(dyn);
  float4* splq = ym4 + F32_W * F32_KS;                               // [W] {1/h_k, w_k, a_k w_k, c'_k} (SynthParams::spl)
  float* h26 = reinterpret_cast<float*>(splq + F32_W);               // [W] h_k^2 / 6
  Knot<float>* knots = reinterpret_cast<Knot<float>*>(reinterpret_cast<float*>(splq) + F32_SPLN);
  LayerKnot<float>* lk = reinterpret_cast<LayerKnot<float>*>(knots + F32_W);  // [n_layers][W]
  __shared__ __align__(16) LayerF ls[MATM_MAX_LAYERS];
  __shared__ float4 dc_uw[F32_NP];  // {u_d0, u_d1, w_d0, w_d1} of the rolled offsets
  __shared__ int2 dc_tab[F32_NP];   // calibration-table base of each detector's band, minus the rounding magic
  __shared__ float dc_gain[F32_DT];
  __shared__ int dc_band[F32_DT];
  __shared__ float dc_mu[F32_DT];
  __shared__ __align__(16) long long dc_row[F32_DT];  // byte offset of each detector's output row
  __shared__ float4 dc_cal[F32_NP];  // {Hc_d0, Hc_d1, Hs_d0, Hs_d1}: the line H(el) of the tile's elevation interval (s_cal_ie >= 0)
  __shared__ int s_cal_ie;

  const int n = P.n_t_atm;
  const int n_int = n - 1;
  const int n_layers = P.n_layers;
  const int n_ttiles = (n_int + F32_KT - 1) / F32_KT;
  const int64_t dtile = blockIdx.x / n_ttiles;
  const int ttile = blockIdx.x % n_ttiles;
  const int k0 = ttile * F32_KT, k1 = min(k0 + F32_KT, n_int);
  const int wa = max(0, k0 - F32_HALO), wb = min(n - 1, k1 + F32_HALO);
  const int Wn = wb - wa + 1;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t dbase = dtile * F32_DT;
  const int nd = (int)min((int64_t)F32_DT, n_det - dbase);
  constexpr int MAGIC_BITS = 0x4B400000;  // bits of 1.5 * 2^23

  // ---- stage the window: knots, per-layer boresight cells, per-layer constants, spline system, detector constants ----
  for (int k = tid; k < Wn; k += F32_THREADS) {
    const double* b = P.bore_atm + 4 * (size_t)(wa + k);
    Knot<float> o;
    o.ce = (float)b[0];
    o.se = (float)b[1];
    o.ca = (float)b[2];
    o.sa = (float)b[3];
    knots[k] = o;
  }
  for (int e = tid; e < n_layers * Wn; e += F32_THREADS) {
    const int l = e / Wn, k = e % Wn;
    const double* b = P.layers[l].base + 2 * (size_t)(wa + k);
    const double cx = b[0], cy = b[1];
    const double flx = fmin(fmax(floor(cx), -1.0e9), 1.0e9), fly = fmin(fmax(floor(cy), -1.0e9), 1.0e9);
    LayerKnot<float> o;
    o.ix = (int)flx;
    o.iy = (int)fly;
    o.fx = (float)(cx - flx);
    o.fy = (float)(cy - fly);
    lk[l * F32_W + k] = o;
  }
  if (tid < n_layers) {
    const LayerDev& L = P.layers[tid];
    LayerF o;
    o.D0 = (float)L.D[0];
    o.D1 = (float)L.D[1];
    o.D2 = (float)L.D[2];
    o.D3 = (float)L.D[3];
    o.rms = (float)L.pwv_rms;
    o.ne1 = L.n_ext - 1;
    o.nc1 = L.n_cs - 1;
    o.n_cs = L.n_cs;
    o.data = reinterpret_cast<const float*>(L.data);
    o.pad[0] = o.pad[1] = 0;
    ls[tid] = o;
  }
  if (tid < F32_DT) {
    const int dd = tid;
    float u = 0.f, w = 1.f, g = 0.f, mu = 0.f;
    int band = 0;
    long long row = 0;
    if (dd < nd) {
      const int64_t d = order[dbase + dd];  // absolute detector index of this slot
      row = (d - det0) * P.n_t * (long long)sizeof(float);
      u = (float)P.det_full[2 * d];
      w = (float)P.det_full[2 * d + 1];
      g = apply_gain ? (float)P.gain[d] : 1.f;
      band = P.band_index[d];
      mu = (float)P.mueller00[d];
    }
    dc_mu[dd] = mu;
    dc_gain[dd] = g;
    dc_band[dd] = band;
    dc_row[dd] = row;
    float* uw = reinterpret_cast<float*>(dc_uw);
    uw[(dd >> 1) * 4 + (dd & 1)] = u;
    uw[(dd >> 1) * 4 + 2 + (dd & 1)] = w;
    reinterpret_cast<int*>(dc_tab)[dd] = band * P.calq_n - MAGIC_BITS;
  }
  if (warp == F32_THREADS / 32 - 1) {
    // a14 without a table when the whole tile stays inside ONE interval of the elevation grid (the usual case: the grid
    // steps are degrees wide): sin(el) of any detector at any sample of the window lies within the boresight's range at
    // the knots, widened by the largest detector offset and by twice the largest knot-to-knot change (motion between knots)
    float lo = 2.f, hi = -2.f, dmax = 0.f;
    for (int k = lane; k < Wn; k += 32) {
      const float c = (float)P.bore_atm[4 * (size_t)(wa + k)];
      const float cn = (float)P.bore_atm[4 * (size_t)(wa + min(k + 1, Wn - 1))];
      lo = fminf(lo, c);
      hi = fmaxf(hi, c);
      dmax = fmaxf(dmax, fabsf(cn - c));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
      dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    }
    if (lane == 0) {
      const float m = P.det_rmax_f + 2.f * dmax + 1e-6f;
      lo -= m;
      hi += m;
      int ie = -1;
      if ((P.table_ok & 2) && units == MATM_UNITS_KRJ && lo == lo && hi == hi) {
        const float* __restrict__ sing = reinterpret_cast<const float*>(P.sinel_R);
        for (int e = 0; e + 1 < P.nel; e++)
          if (__ldg(sing + e) < lo && hi <= __ldg(sing + e + 1)) ie = e;
      }
      s_cal_ie = ie;
    }
  }
  for (int k = tid; k < Wn; k += F32_THREADS) {
    const double* sp = P.spl + 5 * (size_t)(wa + k);
    splq[k] = make_float4((float)sp[0], (float)sp[2], (float)sp[3], (float)sp[4]);
    h26[k] = (float)sp[1];
  }
  __syncthreads();
  if (tid < F32_DT && s_cal_ie >= 0) {
    const int ho = dc_band[tid] * (P.nel - 1) + s_cal_ie;
    float* cal = reinterpret_cast<float*>(dc_cal);
    cal[(tid >> 1) * 4 + (tid & 1)] = __ldg(reinterpret_cast<const float*>(P.Hc_R) + ho);
    cal[(tid >> 1) * 4 + 2 + (tid & 1)] = __ldg(reinterpret_cast<const float*>(P.Hs_R) + ho);
  }

  // ---- phase 1: loading [pW] at the window's knots (phase1_f32) ----
  {
    const int tm = nd == F32_DT ? tile_m[dtile] : 1;
    if (tm == 4)
      phase1_f32<4>(P, order, dbase, nd, wa, Wn, n_layers, knots, lk, ls, ymf, dc_gain, dc_band, dc_mu, lane, warp);
    else if (tm == 2)
      phase1_f32<2>(P, order, dbase, nd, wa, Wn, n_layers, knots, lk, ls, ymf, dc_gain, dc_band, dc_mu, lane, warp);
    else
      phase1_f32<1>(P, order, dbase, nd, wa, Wn, n_layers, knots, lk, ls, ymf, dc_gain, dc_band, dc_mu, lane, warp);
  }
  __syncthreads();

  // ---- phase 2: second derivatives M_k of the not-a-knot cubic spline through the window's knots; lane = detector.
  // Rows of the global tridiagonal system, swept with the global elimination factors.  At a true end of the series the
  // sweep starts / ends exactly; at an artificial end it starts from the locally-constant guess M = d / (a + b + c),
  // whose error decays by (2 - sqrt 3) per knot.  The same trick splits the window between the 8 warps: every warp
  // sweeps its own segment of rows plus F32_HALO rows of run-in on either side, so the serial chain is ~40 rows instead
  // of ~300 and no warp idles at the barrier for long.
  {
    float* row = ymf + (lane >> 1) * 4 + (lane & 1);
    constexpr int RS = F32_KS * 4;  // stride between knots, in floats
    constexpr int NW = F32_THREADS / 32;
    const int lo = 1, hi = Wn - 2, nrows = hi - lo + 1;
    const int ka = lo + (nrows * warp) / NW, kb = lo + (nrows * (warp + 1)) / NW - 1;  // this warp's rows [ka, kb]
    auto rhs3 = [](float ya, float yb, float yc, float iha, float ihb) { return 6.f * ((yc - yb) * ihb - (yb - ya) * iha); };
    auto guess_at = [&](int k) {  // locally-constant estimate of M_k
      const float iha = splq[k - 1].x, ihb = splq[k].x;
      return rhs3(row[(k - 1) * RS], row[k * RS], row[(k + 1) * RS], iha, ihb) / (3.f * (1.f / iha + 1.f / ihb));
    };
    float dh[F32_HALO];  // forward-eliminated right-hand sides of the rows just above this warp's segment
    if (ka <= kb) {
      // forward elimination from fs (run-in) to kb; only this warp's rows are stored
      const int fs = max(lo, ka - F32_HALO);
      float y0 = row[(fs - 1) * RS], y1 = row[fs * RS], y2 = row[(fs + 1) * RS];
      float4 sp0 = splq[fs - 1], sp1 = splq[fs];
      const float r0 = rhs3(y0, y1, y2, sp0.x, sp1.x);
      float dp = (fs == lo && wa == 0) ? r0 * sp1.y : r0 / (3.f * (1.f / sp0.x + 1.f / sp1.x)) * (1.f + sp1.w);
      if (fs >= ka) row[fs * RS + 2] = dp;
      for (int k = fs + 1; k <= kb; k++) {
        y0 = y1;
        y1 = y2;
        y2 = row[(k + 1) * RS];
        sp0 = sp1;
        sp1 = splq[k];
        dp = fmaf(rhs3(y0, y1, y2, sp0.x, sp1.x), sp1.y, -sp1.z * dp);
        if (k >= ka) row[k * RS + 2] = dp;
      }
    }
    __syncthreads();
    const int be = min(hi, kb + F32_HALO);  // back substitution starts here
    if (ka <= kb) {
#pragma unroll
      for (int i = 0; i < F32_HALO; i++) dh[i] = (kb + 1 + i <= be) ? row[(kb + 1 + i) * RS + 2] : 0.f;
    }
    __syncthreads();
    if (ka <= kb) {
      // M just above the segment: exact at the true end of the series, otherwise from the guess one row further up
      float mk;
      if (be == kb) {  // the segment ends at the last row of the window
        mk = row[kb * RS + 2];
        if (wb != n - 1) mk -= splq[kb].w * guess_at(kb);
      } else {
        mk = 0.f;
        const bool true_end = be == hi && wb == n - 1;
#pragma unroll
        for (int i = F32_HALO - 1; i >= 0; i--) {
          const int k = kb + 1 + i;
          if (k == be)
            mk = true_end ? dh[i] : dh[i] - splq[k].w * guess_at(k);
          else if (k < be)
            mk = fmaf(-splq[k].w, mk, dh[i]);
        }
        mk = fmaf(-splq[kb].w, mk, row[kb * RS + 2]);
      }
      row[kb * RS + 2] = mk;
      for (int k = kb - 1; k >= ka; k--) {
        mk = fmaf(-splq[k].w, mk, row[k * RS + 2]);
        row[k * RS + 2] = mk;
      }
    }
    __syncthreads();
    if (warp == 0) {
      const float m1 = row[RS + 2], m2 = row[2 * RS + 2];
      row[2] = (wa == 0) ? m1 - (splq[1].x / splq[0].x) * (m2 - m1) : m1;
    } else if (warp == 1) {
      const float me = row[(Wn - 2) * RS + 2], me2 = row[(Wn - 3) * RS + 2];
      row[(Wn - 1) * RS + 2] = (wb == n - 1) ? me + (splq[Wn - 3].x / splq[Wn - 2].x) * (me - me2) : me;
    }
  }
  __syncthreads();

  // ---- phase 3: evaluate at the full-rate samples of the tile ----
  // thread = F32_S samples (j, j + 256, ...), loop over detector PAIRS; the per-pair constants (offsets, table base,
  // output rows) are fetched once per F32_S samples: the kernel is bound by L1 wavefronts, not by arithmetic.
  const int64_t j0 = P.jstart[k0], j1 = P.jstart[k1];
  const double t_atm0 = P.t_atm0, inv_ts = 1.0 / P.timestep;
  const float inv_ts_f = (float)inv_ts;
  const bool krj = units == MATM_UNITS_KRJ;
  const int npairs = nd >> 1;
  const float2* __restrict__ calq = P.calq;
  const float2 scale2 = f2(MATM_CAL_SCALE, MATM_CAL_SCALE), magic2 = f2(12582912.f, 12582912.f);
  // a14 comes in three flavours, chosen per tile (the choice depends on the time window and on the whole focal plane, not
  // on the detector shard, so shards stay bit-identical): MODE 0 no calibration (pW); MODE 1 the table in t; MODE 2 the
  // whole tile lies in one interval of the elevation grid, H(el) = Hc + Hs el with el = asin(sin el) from a polynomial,
  // no load and no special bins.
  // warps start on multiples of 32 samples: a warp's store is then one 128-byte line per row whenever n_t % 32 == 0
  auto phase3 = [&](auto mode_tag) {
    constexpr int MODE = decltype(mode_tag)::value;
    for (int64_t jb = (j0 & ~(int64_t)31) + tid; jb < j1; jb += F32_THREADS * F32_S) {
      float a1[F32_S], a2[F32_S], a3[F32_S], se[F32_S], ce[F32_S];
      const float4* yk[F32_S];
      const char* oq[F32_S];  // out + j: a store is then one 64-bit add of the row's byte offset
#pragma unroll
      for (int q = 0; q < F32_S; q++) {
        // a sample index outside the tile repeats the tile's first / last sample: the same value is stored to the same address
        // by the thread that owns it, which is harmless and keeps the inner loop free of predicates
        const int64_t j = max(min(jb + q * F32_THREADS, j1 - 1), j0);
        const double tj = P.t_full[j];
        int i = (int)floor((tj - t_atm0) * inv_ts);
        i = min(max(i, k0), k1 - 1);
        const int li = i - wa;
        const float s = (float)(tj - P.t_atm[i]) * (splq[li].x * inv_ts_f);  // may leave [0, 1] by ~1e-6, or at the ends
        const float hh = h26[li], c = 1.f - s;
        a1[q] = s;
        a2[q] = hh * (c * c * c - c);
        a3[q] = hh * (s * s * s - s);
        se[q] = ce[q] = 0.f;
        if constexpr (MODE != 0) {
          const double2 bf = *reinterpret_cast<const double2*>(P.bore_full + 2 * j);
          ce[q] = (float)bf.x;
          se[q] = (float)bf.y;
        }
        yk[q] = ym4 + li * F32_KS;
        oq[q] = reinterpret_cast<const char*>(out + j);
      }
#pragma unroll 2
      for (int p = 0; p < npairs; p++) {
        const float4 uw = dc_uw[p];
        int2 tb = make_int2(0, 0);
        float4 cal = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (MODE == 1) tb = dc_tab[p];
        if constexpr (MODE == 2) cal = dc_cal[p];
        const longlong2 ro = *reinterpret_cast<const longlong2*>(&dc_row[2 * p]);
#pragma unroll
        for (int q = 0; q < F32_S; q++) {
          const float4 q0 = yk[q][p], q1 = yk[q][F32_KS + p];
          const float a0 = 1.f - a1[q];
          float2 v = __fmul2_rn(f2(a0, a0), f2(q0.x, q0.y));
          v = __ffma2_rn(f2(a1[q], a1[q]), f2(q1.x, q1.y), v);
          v = __ffma2_rn(f2(a2[q], a2[q]), f2(q0.z, q0.w), v);
          v = __ffma2_rn(f2(a3[q], a3[q]), f2(q1.z, q1.w), v);
          if constexpr (MODE == 1) {
            const float2 sn = __ffma2_rn(f2(uw.x, uw.y), f2(se[q], se[q]), __fmul2_rn(f2(uw.z, uw.w), f2(ce[q], ce[q])));
            // |1 - sin el|: sin el can exceed 1 by an ulp; inputs are finite (checked by matm_set_scan / matm_set_detectors)
            const float2 t = f2(fast_sqrt(fabsf(1.f - sn.x)), fast_sqrt(fabsf(1.f - sn.y)));
            const float2 x = __ffma2_rn(t, scale2, magic2);  // round(t * scale) in the low mantissa bits
            const float2 r = __fadd2_rn(x, f2(-12582912.f, -12582912.f));
            const float2 dt = __ffma2_rn(t, scale2, f2(-r.x, -r.y));
            const int i0 = __float_as_int(x.x) + tb.x, i1 = __float_as_int(x.y) + tb.y;
            const float2 c0 = __ldg(calq + i0), c1 = __ldg(calq + i1);
            float2 f = f2(fmaf(c0.y, dt.x, c0.x), fmaf(c1.y, dt.y, c1.x));
            if (!(f.x * f.y == f.x * f.y)) {  // a bin with a grid point of the elevation table, or outside the table
              if (!(f.x == f.x)) f.x = cal_factor_special(P, i0, dc_band[2 * p], sn.x, dt.x);
              if (!(f.y == f.y)) f.y = cal_factor_special(P, i1, dc_band[2 * p + 1], sn.y, dt.y);
            }
            v = __fmul2_rn(v, f);
          } else if constexpr (MODE == 2) {
            const float2 sn = __ffma2_rn(f2(uw.x, uw.y), f2(se[q], se[q]), __fmul2_rn(f2(uw.z, uw.w), f2(ce[q], ce[q])));
            const float2 sq = f2(fast_sqrt(fabsf(1.f - sn.x)), fast_sqrt(fabsf(1.f - sn.y)));  // |.|: sin el may exceed 1 by an ulp
            const float2 el = asin01_f32x2(sn, sq);
            const float2 den = __ffma2_rn(f2(cal.z, cal.w), el, f2(cal.x, cal.y));
            v = __fmul2_rn(v, f2(fast_rcp(den.x), fast_rcp(den.y)));
          }
          st_at(oq[q], ro.x, v.x);
          st_at(oq[q], ro.y, v.y);
        }
      }
      if (nd & 1) {  // odd tail detector of a partial tile; the same arithmetic as the paired path (bit-identical shards)
        const int p = npairs;
        const float uu = dc_uw[p].x, ww = dc_uw[p].z;
        const int tbx = dc_tab[p].x;
        const float hc = dc_cal[p].x, hs = dc_cal[p].z;
        const long long r0 = dc_row[2 * p];
#pragma unroll
        for (int q = 0; q < F32_S; q++) {
          const float4 q0 = yk[q][p], q1 = yk[q][F32_KS + p];
          float v = fmaf(a3[q], q1.z, fmaf(a2[q], q0.z, fmaf(a1[q], q1.x, (1.f - a1[q]) * q0.x)));
          if constexpr (MODE == 1) {
            const float sn = fmaf(uu, se[q], ww * ce[q]);
            const float t = fast_sqrt(fabsf(1.f - sn));
            const float x = fmaf(t, MATM_CAL_SCALE, 12582912.f);
            const float dt = fmaf(t, MATM_CAL_SCALE, -(x - 12582912.f));
            const int i0 = __float_as_int(x) + tbx;
            const float2 c0 = __ldg(calq + i0);
            float f = fmaf(c0.y, dt, c0.x);
            if (!(f == f)) f = cal_factor_special(P, i0, dc_band[2 * p], sn, dt);
            v *= f;
          } else if constexpr (MODE == 2) {
            const float sn = fmaf(uu, se[q], ww * ce[q]);
            const float sq = fast_sqrt(fabsf(1.f - sn));
            const float2 el = asin01_f32x2(f2(sn, sn), f2(sq, sq));
            v *= fast_rcp(fmaf(hs, el.x, hc));
          }
          st_at(oq[q], r0, v);
        }
      }
    }
  };
  if (!krj)
    phase3(std::integral_constant<int, 0>{});
  else if (s_cal_ie >= 0)
    phase3(std::integral_constant<int, 2>{});
  else
    phase3(std::integral_constant<int, 1>{});
}

static size_t synth_f32_smem(int n_layers) {
  return (size_t)F32_W * F32_KS * sizeof(float4) + F32_SPLN * sizeof(float) + F32_W * sizeof(Knot<float>) +
         (size_t)n_layers * F32_W * sizeof(LayerKnot<float>);
}
