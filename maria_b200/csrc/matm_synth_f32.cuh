// The production (float) path of the fused synthesis a6..a14, as TWO kernels per detector chunk, written for instruction
// count (ncu: the one-kernel version was issue- and latency-bound at 35 thread-instructions per output sample, not
// HBM-bound):
//
//   loading_f32_kernel   (per knot)    loading [pW] at the atmosphere-rate knots: lane = focal-plane position (co-located
//                                      detectors share the line of sight: tile multiplicity M), F32_G knots per thread,
//                                      8 bilinear gathers from the L2-resident layers, emission lookup, x Mueller x gain
//                                      -> yk[tile][knot][32 slots] (an L2-sized scratch of the chunk, no halo, own occupancy)
//   synth_eval_f32_kernel (per sample) window of yk -> shared memory; not-a-knot cubic spline second derivatives by the
//                                      windowed Thomas sweep; evaluation + pW -> K_RJ at the full-rate samples: thread =
//                                      4 consecutive samples, loop over the detectors, packed f32x2 arithmetic over sample
//                                      pairs, one 16-byte store per detector.  The calibration factor is rcp of a QUADRATIC
//                                      in the boresight elevation with per-(detector, tile) coefficients (every
//                                      transcendental depends on time only through el_b), else the exact search + asin.
#pragma once
#include <type_traits>
#include "matm_synth.cuh"

#define F32_DT 32
#define F32_THREADS 256
#ifndef F32_KT
// spline intervals per tile of the evaluation kernel.  Its loop over aligned groups of 4 samples is quantised in passes of
// 256 groups: 135 intervals are 3051 samples = 2.98 passes at the C4 cadence (22.6 samples per interval); the window of
// F32_KT + 1 + 2 F32_HALO knots x 32 detectors x 8 bytes is 42 KB.  Any value is correct (-DF32_KT=... retunes).
#define F32_KT 135
#endif
#ifndef F32_HALO
#define F32_HALO 10  // extra knots each side; the error of the artificial boundary decays as (2 - sqrt 3)^HALO
#endif
#ifndef F32_MINB
#define F32_MINB 3   // CTAs per SM the register allocation is held to
#endif
#define F32_W (F32_KT + 1 + 2 * F32_HALO)
// row stride of the {y, M} window in shared memory, in float2: the 32 lanes of a warp read 5-6 different knot rows at
// once (128 consecutive samples), and rows 256 bytes apart would all fall on the same banks (ncu: 93 % of the L1 data
// pipe, 1.2e9 bank conflicts); 34 shifts consecutive rows by 4 banks and keeps 16-byte alignment
#define F32_YS (F32_DT + 2)
#ifndef F32_UNROLL
#define F32_UNROLL 4  // detectors per unrolled body of the evaluation loop
#endif
#ifndef F32_G
#define F32_G 2  // knots per thread and pass of the loading kernel
#endif
#ifndef F32_LB
#define F32_LB 2  // layers whose gathers are in flight together (2 F32_G F32_LB loads per thread)
#endif

struct LayerF {  // per-layer constants, two 16-byte shared-memory loads + the pointer
  float D0, D2, D1, D3;  // (cell x, cell y) = D0, D2 per unit of pp_x and D1, D3 per unit of pp_y: packed pairs
  float rms;
  int ne1, nc1, n_cs;  // n_ext - 1, n_cs - 1, n_cs
  const float2* data2;  // pair-expanded layer: data2[i n_cs + j] = {v[i][j], v[i][j + 1]}
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


// ---------------------------------------------------------------------------------------------------------
// Kernel A: loading [pW] at the atmosphere-rate knots (a6 + a7 + a9 + a11 + a12 x Mueller, a13 gain folded in).
// One CTA = one tile of 32 processing-order slots x LD_KA knots.  Detectors that share a position (the bands of a
// dichroic pixel, both polarisations) share the line of sight through every layer: for a tile whose slots come in aligned
// groups of M co-located detectors, lanes [0, 32 / M) of a warp are the tile's positions, the other lanes take further
// knots (M "virtual warps" per warp), and each lane finishes M detectors (band table, Mueller, gain).  Each thread works
// on F32_G knots at once: F32_G independent gather chains per layer keep 4 F32_G loads in flight (the layers live in L2).
// ---------------------------------------------------------------------------------------------------------
#define LD_THREADS 256
#ifndef LD_KA
#define LD_KA 256  // knots per CTA (the per-knot cell table of all layers is staged once per CTA; 64: 8.9, 128: 8.1, 256: 7.7 ms)
#endif
#ifndef LD_MINB
// CTAs per SM the register allocation is held to.  Measured at C4 (ms per launch): G x LB x MINB = 2x2x4 8.1, 4x1x3 8.6,
// 4x2x3 9.4, 2x4x3 9.7, 1x8x3 9.9, 2x4x4 10.4, 4x4x2 12.0: more warps beat more loads in flight per warp.
#define LD_MINB 4
#endif

// One 8-byte load per grid row instead of two 4-byte loads: the sampler reads the layers through their pair-expanded copy
// (written from the smoothed layers at the start of every synthesis; 2 x 26 MB at C4, L2-resident).
__global__ void __launch_bounds__(256) expand_layers_kernel(const __grid_constant__ SynthParams P) {
  const LayerDev& L = P.layers[blockIdx.y];
  const float* __restrict__ v = reinterpret_cast<const float*>(L.data);
  float2* __restrict__ o = reinterpret_cast<float2*>(L.data2);
  const int n_cs = L.n_cs;
  const int64_t total = (int64_t)L.n_ext * n_cs;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(e % n_cs);
    const float a = v[e];
    o[e] = make_float2(a, j + 1 < n_cs ? v[e + 1] : a);
  }
}

template <int M>
__device__ __forceinline__ void loading_tile_f32(const SynthParams& P, const int32_t* __restrict__ order, int64_t dbase, int nd,
                                                 int ka, int Wn, int n_layers, const LayerF* ls, const int4* lks,
                                                 const float* dc_gain, const int* dc_band, const float* dc_mu,
                                                 float* __restrict__ yrow, int lane, int warp) {
  constexpr int PL = 32 / M;
  const int slot0 = (lane % PL) * M, sub = lane / PL;
  const int64_t d = order[dbase + min(slot0, nd - 1)];  // idle lanes of a partial tile shadow the last detector
  const double* da = P.det_atm + 4 * d;
  const float u = (float)da[0], Y = (float)da[1], wm1 = (float)da[2];
  const float pwv0 = (float)P.pwv0;
  int ip = 0, ie = 0;
  constexpr int NW = (LD_THREADS / 32) * M;
  // one pass = G knots per thread, NW * G knots per CTA.  Full passes use F32_G; what is left at the end of the block
  // takes the smallest G that covers it.
  auto pass = [&](auto gtag, const int kb) {
    constexpr int G = decltype(gtag)::value;
    float dppx[G], dppy[G], sinel[G], acc[G];
    int kk[G];  // knot within the block
    bool bad = false;
#pragma unroll
    for (int q = 0; q < G; q++) {
      kk[q] = min(kb + q * NW, Wn - 1);  // a short last group repeats the last knot (same value, same slot)
      const float4 kc = __ldg(P.knot_f + ka + kk[q]);  // ce, se, ca, sa
      // direction of the detector minus direction of the boresight, horizon frame (coords/transforms.py:10-29)
      const float dX = u * kc.x - wm1 * kc.y;
      const float dZ = u * kc.y + wm1 * kc.x;
      const float dNx = dX * kc.z - Y * kc.w;
      const float dNy = dX * kc.w + Y * kc.z;
      const float Nbx = -kc.y * kc.z, Nby = -kc.y * kc.w, Nbz = kc.x;
      const float Nz = Nbz + dZ;  // = sin(el) of the detector
      const float inv = fast_rcp(Nbz * Nz);
      dppx[q] = (dNx * Nbz - Nbx * dZ) * inv;  // pp_d - pp_b (coordinates.py:333-349)
      dppy[q] = (dNy * Nbz - Nby * dZ) * inv;
      sinel[q] = Nz;
      acc[q] = 0.f;
      bad = bad || !(dppx[q] == dppx[q]) || !(dppy[q] == dppy[q]);
    }
    unsigned badmask = 0;
    // The kernel is bound by the latency of the gathers (the layers live in L2): F32_LB layers are taken at a time, all
    // their 2 G F32_LB loads are issued before the first is used.
    for (int l0 = 0; l0 < n_layers; l0 += F32_LB) {
      float2 w[F32_LB][G], r0[F32_LB][G], r1[F32_LB][G];
      float rms[F32_LB];
#pragma unroll
      for (int i = 0; i < F32_LB; i++) {
        const int l = min(l0 + i, n_layers - 1);  // a short last block repeats the last layer with weight zero
        const float4 Dm = *reinterpret_cast<const float4*>(&ls[l].D0);  // D0, D2, D1, D3
        const int4 Ln = *reinterpret_cast<const int4*>(&ls[l].rms);
        const float2* __restrict__ data2 = ls[l].data2;
        const int4* __restrict__ lkl = lks + l * LD_KA;
        rms[i] = l0 + i < n_layers ? __int_as_float(Ln.x) : 0.f;
#pragma unroll
        for (int q = 0; q < G; q++) {
          const int4 b = lkl[kk[q]];  // {ix, iy, fx, fy} of the boresight in this layer's grid
          // (cell x, cell y) of the line of sight relative to the boresight's cell, both components at once
          const float2 f = __ffma2_rn(f2(dppx[q], dppx[q]), f2(Dm.x, Dm.y),
                                      __ffma2_rn(f2(dppy[q], dppy[q]), f2(Dm.z, Dm.w), f2(__int_as_float(b.z), __int_as_float(b.w))));
          const float2 fl = f2(floorf(f.x), floorf(f.y));
          int ix = b.x + (int)fl.x, iy = b.y + (int)fl.y;
          w[i][q] = __fadd2_rn(f, f2(-fl.x, -fl.y));
          if ((unsigned)ix >= (unsigned)Ln.y || (unsigned)iy >= (unsigned)Ln.z) {
            // rare: exactly on the last grid line is inside (RGI clips the index to n - 2); anything else is outside
            if (ix == Ln.y && w[i][q].x == 0.f) {
              ix--;
              w[i][q].x = 1.f;
            }
            if (iy == Ln.z && w[i][q].y == 0.f) {
              iy--;
              w[i][q].y = 1.f;
            }
            if ((unsigned)ix >= (unsigned)Ln.y || (unsigned)iy >= (unsigned)Ln.z) {
              if (l0 + i < n_layers) {
                record_out_of_grid(P.flags, l, d, ka + kk[q]);  // atmosphere.py:368-369
                badmask |= 1u << q;
              }
              ix = iy = 0;
            }
          }
          const unsigned idx = (unsigned)(ix * Ln.w + iy);
          r0[i][q] = __ldg(data2 + idx);                  // {v00, v01}
          r1[i][q] = __ldg(data2 + idx + (unsigned)Ln.w);  // {v10, v11}
        }
      }
#pragma unroll
      for (int i = 0; i < F32_LB; i++)
#pragma unroll
        for (int q = 0; q < G; q++) {
          // along the extrusion axis first, on both columns at once; then across
          const float2 c = __ffma2_rn(f2(w[i][q].x, w[i][q].x), __fadd2_rn(r1[i][q], f2(-r0[i][q].x, -r0[i][q].y)), r0[i][q]);
          acc[q] = fmaf(rms[i], fmaf(w[i][q].y, c.y - c.x, c.x), acc[q]);
        }
    }
#pragma unroll
    for (int q = 0; q < G; q++) {
      if (bad || ((badmask >> q) & 1)) {
        if (!(dppx[q] == dppx[q]) || !(dppy[q] == dppy[q])) record_out_of_grid(P.flags, 0, d, ka + kk[q]);
        if (!(dppx[q] == dppx[q]) || !(dppy[q] == dppy[q]) || ((badmask >> q) & 1)) acc[q] = r_nan<float>();
      }
      float wp, we;
      const bool ok = emission_cell_f32(P, pwv0 + acc[q], sinel[q], ip, ie, wp, we);
      float y[M];
#pragma unroll
      for (int j = 0; j < M; j++) {
        const int sl = slot0 + j;
        // times Mueller [0,0]; the spline being linear in the data, the detector gain goes in here too
        y[j] = ok ? emission_blend_f32(P, dc_band[sl], ip, ie, wp, we) * dc_mu[sl] * dc_gain[sl] : r_nan<float>();
        if (sl >= nd) y[j] = 0.f;  // unused slots of a partial tile
      }
      if (kb + q * NW < Wn) {
        float* o = yrow + (size_t)(ka + kk[q]) * F32_DT + slot0;
        if constexpr (M == 1) *o = y[0];
        if constexpr (M == 2) *reinterpret_cast<float2*>(o) = make_float2(y[0], y[1]);
        if constexpr (M == 4) *reinterpret_cast<float4*>(o) = make_float4(y[0], y[1], y[2], y[3]);
      }
    }
  };
  const int vw = warp * M + sub;
  for (int base = 0; base < Wn;) {
    const int rem = Wn - base;
    constexpr int GH = F32_G >= 2 ? F32_G / 2 : 1;
    if (rem > GH * NW || F32_G == 1) {
      pass(std::integral_constant<int, F32_G>{}, base + vw);
      base += F32_G * NW;
    } else if (rem > NW) {
      pass(std::integral_constant<int, GH>{}, base + vw);
      base += GH * NW;
    } else {
      pass(std::integral_constant<int, 1>{}, base + vw);
      base += NW;
    }
  }
}

// yk: [n_tiles of this launch][n_t_atm][32] float; tile0 = first tile (of the requested detector range) of this launch
__global__ void __launch_bounds__(LD_THREADS, LD_MINB)
    loading_f32_kernel(const __grid_constant__ SynthParams P, const int32_t* __restrict__ order,
                       const uint8_t* __restrict__ tile_m, int64_t tile0, int64_t n_det, float* __restrict__ yk,
                       int apply_gain) {
  extern __shared__ __align__(16) unsigned char dyn[];
  int4* lks = reinterpret_cast<int4*>(dyn);  // [n_layers][LD_KA] {ix, iy, bits of fx, fy}: the boresight in every layer grid
  __shared__ __align__(16) LayerF ls[MATM_MAX_LAYERS];
  __shared__ float dc_gain[F32_DT];
  __shared__ int dc_band[F32_DT];
  __shared__ float dc_mu[F32_DT];
  const int n = P.n_t_atm;
  const int n_kb = (n + LD_KA - 1) / LD_KA;
  const int64_t tl = blockIdx.x / n_kb;  // tile within this launch
  const int kb = blockIdx.x % n_kb;
  const int ka = kb * LD_KA, Wn = min(LD_KA, n - ka);
  const int64_t dtile = tile0 + tl;
  const int64_t dbase = dtile * F32_DT;
  const int nd = (int)min((int64_t)F32_DT, n_det - dbase);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_layers = P.n_layers;
  {
    const int4* __restrict__ src = reinterpret_cast<const int4*>(P.lk_tab);
    for (int e = tid; e < n_layers * Wn; e += LD_THREADS) {
      const int l = e / Wn, k = e % Wn;
      lks[l * LD_KA + k] = __ldg(src + (size_t)l * n + ka + k);
    }
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
    o.data2 = reinterpret_cast<const float2*>(L.data2);
    o.pad[0] = o.pad[1] = 0;
    ls[tid] = o;
  }
  if (tid >= 32 && tid < 32 + F32_DT) {
    const int dd = tid - 32;
    float g = 0.f, mu = 0.f;
    int band = 0;
    if (dd < nd) {
      const int64_t d = order[dbase + dd];
      g = apply_gain ? (float)P.gain[d] : 1.f;
      band = P.band_index[d];
      mu = (float)P.mueller00[d];
    }
    dc_mu[dd] = mu;
    dc_gain[dd] = g;
    dc_band[dd] = band;
  }
  __syncthreads();
  float* yrow = yk + (size_t)tl * n * F32_DT;
  const int tm = nd == F32_DT ? tile_m[dtile] : 1;
  if (tm == 4)
    loading_tile_f32<4>(P, order, dbase, nd, ka, Wn, n_layers, ls, lks, dc_gain, dc_band, dc_mu, yrow, lane, warp);
  else if (tm == 2)
    loading_tile_f32<2>(P, order, dbase, nd, ka, Wn, n_layers, ls, lks, dc_gain, dc_band, dc_mu, yrow, lane, warp);
  else
    loading_tile_f32<1>(P, order, dbase, nd, ka, Wn, n_layers, ls, lks, dc_gain, dc_band, dc_mu, yrow, lane, warp);
}

// ---------------------------------------------------------------------------------------------------------
// Kernel B: spline + evaluation + calibration.  One CTA = one tile of 32 slots x F32_KT spline intervals (+ F32_HALO knots
// of yk on either side for the windowed Thomas sweep).
//
// Evaluation (phase 3): a thread owns an aligned group of FOUR consecutive full-rate samples and loops over the tile's 32
// detectors; every arithmetic instruction is a packed f32x2 over two consecutive SAMPLES, the detector's numbers enter as
// scalar (broadcast) operands, and the four results leave as one 16-byte st.global (a warp writes 512 contiguous bytes of
// a row).  What depends on the sample only is computed once per group and kept in registers: the cubic weights (a float4
// table written once per scan by sample_weights_kernel), placed on a THREE-knot stencil so that a group that straddles a
// knot needs no branch, and x = el_b(t) - e_mid.  Per detector and group the kernel issues 3 LDS.64 (y, M at three knots),
// 2 loads of constants, 16 packed FMAs, 4 reciprocals and 1 store.
// The one-kernel version of round 1 spent 2 LDS.128 at ~4 shared-memory wavefronts each per detector pair and SAMPLE and
// was bound by the L1 data pipe (97 % of peak, ncu); here the shared-memory traffic per sample is ~5x lower.
//
// The three-knot stencil needs every interior spline interval to hold at least 4 full-rate samples (checked on the host);
// coarser scans go through synth_eval_general_f32_kernel (thread = sample, any spacing).
// ---------------------------------------------------------------------------------------------------------
struct EvalTile {
  int k0, k1, wa, wb, Wn, nd, mode;
  int64_t j0, j1, dbase;
  double emid;
};

// Shared prologue of the two evaluation kernels: stage the window of yk, choose the a14 mode of the tile, per-detector
// constants, spline second derivatives (phase 2).  Shared memory: ym [W][32] {y, M}, splq [W].
// a14 comes in three flavours, chosen per tile (the choice depends on the time window and on the whole focal plane, not
// on the detector shard, so shards stay bit-identical): MODE 0 no calibration (pW); MODE 2 the reciprocal of a quadratic
// in x = el_b(t) - e_mid; MODE 1 the exact evaluation per sample (tiles that cross a point of the elevation grid, or where
// the quadratic is not accurate enough: near the zenith, or a large change of elevation within the tile).
__device__ __forceinline__ bool eval_prologue(const SynthParams& P, const int32_t* __restrict__ order, int64_t tile0, int64_t det0,
                                              int64_t n_det, const float* __restrict__ yk, int units, float2* ym, float4* splq,
                                              float4* dc_cal, long long* dc_row, float2* dc_uw, int* dc_band, int* s_misc,
                                              double* s_emid, EvalTile& T) {
  const int n = P.n_t_atm;
  const int n_int = n - 1;
  const int n_ttiles = (n_int + F32_KT - 1) / F32_KT;
  const int64_t tl = blockIdx.x / n_ttiles;
  const int ttile = blockIdx.x % n_ttiles;
  T.k0 = ttile * F32_KT;
  T.k1 = min(T.k0 + F32_KT, n_int);
  T.wa = max(0, T.k0 - F32_HALO);
  T.wb = min(n - 1, T.k1 + F32_HALO);
  T.Wn = T.wb - T.wa + 1;
  T.dbase = (tile0 + tl) * F32_DT;
  T.nd = (int)min((int64_t)F32_DT, n_det - T.dbase);
  T.j0 = P.jstart[T.k0];
  T.j1 = P.jstart[T.k1];
  if (T.j0 >= T.j1) return false;  // a time tile that owns no full-rate sample (sample spacing > F32_KT knots)
  const int k0 = T.k0, k1 = T.k1, wa = T.wa, wb = T.wb, Wn = T.Wn, nd = T.nd;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool krj = units == MATM_UNITS_KRJ;
  float* ymf = reinterpret_cast<float*>(ym);

  // ---- stage the window: loading at the knots (kernel A's output), spline system, detector constants ----
  {
    const float* __restrict__ src = yk + ((size_t)tl * n + wa) * F32_DT;
    for (int e = tid; e < Wn * F32_DT; e += F32_THREADS)
      ymf[2 * ((e >> 5) * F32_YS + (e & 31))] = __ldcs(src + e);  // read once (plus the halo): streaming
  }
  if (tid < F32_DT) {
    const int dd = tid;
    float u = 0.f, w = 1.f;
    int band = 0;
    long long row = 0;
    if (dd < nd) {
      const int64_t d = order[T.dbase + dd];  // absolute detector index of this slot
      row = (d - det0) * P.n_t;
      u = (float)P.det_full[2 * d];
      w = (float)P.det_full[2 * d + 1];
      band = P.band_index[d];
    }
    dc_band[dd] = band;
    dc_row[dd] = row;
    dc_uw[dd] = f2(u, w);
  }
  if (warp == F32_THREADS / 32 - 1) {
    // MODE 2 needs (1) every detector inside ONE interval of the elevation grid over the tile's full-rate samples (the
    // usual case: the grid steps are degrees wide), so that H(el) = Hc + Hs el, and (2) the quadratic expansion of
    // el_d(el_b) about the tile's mid elevation exact to 5e-8 rad.  The range is taken over the FULL-RATE samples of the
    // tile (el_rng), not the knots: turnarounds between knots and the extrapolated tail are covered.
    double lo = INFINITY, hi = -INFINITY;
    for (int k = k0 + lane; k < k1; k += 32) {
      lo = fmin(lo, P.el_rng[2 * k]);
      hi = fmax(hi, P.el_rng[2 * k + 1]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) {
      int ie = -1;
      const double emid = 0.5 * (lo + hi), xmax = 0.5 * (hi - lo), r = P.det_rmax * (1.0 + 1e-6) + 1e-7;
      if ((P.table_ok & 2) && krj && lo == lo && hi == hi && hi + r < 1.5607963 /* pi/2 - 0.01 */ && lo - r > -1.5607963) {
        const float slo = (float)sin(lo - r) - 2e-7f, shi = (float)sin(hi + r) + 2e-7f;
        const float* __restrict__ sing = reinterpret_cast<const float*>(P.sinel_R);
        for (int e = 0; e + 1 < P.nel; e++)
          if (__ldg(sing + e) < slo && shi <= __ldg(sing + e + 1)) ie = e;
        // remainder of the quadratic: |d^3 delta / de^3| x^3 / 6 with delta = el_b - beta - el_d ~ eps tan(theta),
        // eps <= r^2 / 2: eps sec^2 (2 sec^2 + 4 tan^2) at the largest |elevation| a detector reaches
        const double em = fmax(fabs(lo - r), fabs(hi + r)), c = cos(em), t = tan(em), s2 = 1.0 / (c * c);
        const double rem = 0.5 * r * r * s2 * (2.0 * s2 + 4.0 * t * t) * xmax * xmax * xmax / 6.0;
        if (!(rem <= 5e-8)) ie = -1;
      }
      s_misc[0] = ie;
      *s_emid = emid;
    }
  }
  for (int k = tid; k < Wn; k += F32_THREADS) {
    const double* sp = P.spl + 5 * (size_t)(wa + k);
    splq[k] = make_float4((float)sp[0], (float)sp[2], (float)sp[3], (float)sp[4]);
  }
  __syncthreads();
  const int cal_ie = s_misc[0];
  T.mode = !krj ? 0 : (cal_ie >= 0 ? 2 : 1);
  T.emid = *s_emid;
  if (tid < F32_DT) {
    // el_d(e) = asin(rho sin(e - beta)) = theta - delta(e), theta = e - beta, rho = 1 - eps; with c = cos el_d:
    //   delta = asin(sin(theta) eps (2 - eps) / (c + rho cos theta))          (no cancellation; asin x = x to 1e-9)
    //   delta' = 1 - f',  f' = rho cos(theta) / c,  1 - f' = eps (2 - eps) / ((c + rho cos theta) c)
    //   delta'' = -f'' = tan(el_d) (1 - f') (1 + f')
    // H(el_d(e_mid + x)) = Hc + Hs (theta_mid - delta) + Hs (1 - delta') x - Hs delta''/2 x^2 + O(x^3)
    float p0 = 1.f, p1 = 0.f, p2 = 0.f;
    if (tid < nd && cal_ie >= 0) {
      const int64_t d = order[T.dbase + tid];
      const int ho = dc_band[tid] * (P.nel - 1) + cal_ie;
      const double hc = (double)__ldg(reinterpret_cast<const float*>(P.Hc_R) + ho), hs = (double)__ldg(reinterpret_cast<const float*>(P.Hs_R) + ho);
      const double beta = P.det_cal[2 * d], eps = P.det_cal[2 * d + 1];
      const double th = T.emid - beta, rho = 1.0 - eps;
      double st, ct;
      sincos(th, &st, &ct);
      const double sf = rho * st, c = sqrt(fmax(1.0 - sf * sf, 1e-30));
      const double g = eps * (2.0 - eps) / (c + rho * ct);
      const double delta = st * g, d1 = g / c, fp = 1.0 - d1, d2 = (sf / c) * d1 * (1.0 + fp);
      p0 = (float)(hc + hs * (th - delta));
      p1 = (float)(hs * (1.0 - d1));
      p2 = (float)(-0.5 * hs * d2);
    }
    dc_cal[tid] = make_float4(p0, p1, p2, __int_as_float((int)(uint32_t)(dc_row[tid] / P.n_t)));  // .w: output row index
  }
  // ---- phase 2: second derivatives M_k of the not-a-knot cubic spline through the window's knots; lane = detector.
  // Rows of the global tridiagonal system, swept with the global elimination factors.  At a true end of the series the
  // sweep starts / ends exactly; at an artificial end it starts from the locally-constant guess M = d / (a + b + c),
  // whose error decays by (2 - sqrt 3) per knot.  The same trick splits the window between the 8 warps: every warp
  // sweeps its own segment of rows plus F32_HALO rows of run-in on either side, so the serial chain is ~40 rows instead
  // of ~300 and no warp idles at the barrier for long.
  {
    float* row = ymf + lane * 2;
    constexpr int RS = F32_YS * 2;  // stride between knots, in floats
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
      if (fs >= ka) row[fs * RS + 1] = dp;
      for (int k = fs + 1; k <= kb; k++) {
        y0 = y1;
        y1 = y2;
        y2 = row[(k + 1) * RS];
        sp0 = sp1;
        sp1 = splq[k];
        dp = fmaf(rhs3(y0, y1, y2, sp0.x, sp1.x), sp1.y, -sp1.z * dp);
        if (k >= ka) row[k * RS + 1] = dp;
      }
    }
    __syncthreads();
    const int be = min(hi, kb + F32_HALO);  // back substitution starts here
    if (ka <= kb) {
#pragma unroll
      for (int i = 0; i < F32_HALO; i++) dh[i] = (kb + 1 + i <= be) ? row[(kb + 1 + i) * RS + 1] : 0.f;
    }
    __syncthreads();
    if (ka <= kb) {
      // M just above the segment: exact at the true end of the series, otherwise from the guess one row further up
      float mk;
      if (be == kb) {  // the segment ends at the last row of the window
        mk = row[kb * RS + 1];
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
        mk = fmaf(-splq[kb].w, mk, row[kb * RS + 1]);
      }
      row[kb * RS + 1] = mk;
      for (int k = kb - 1; k >= ka; k--) {
        mk = fmaf(-splq[k].w, mk, row[k * RS + 1]);
        row[k * RS + 1] = mk;
      }
    }
    __syncthreads();
    if (warp == 0) {
      const float m1 = row[RS + 1], m2 = row[2 * RS + 1];
      row[1] = (wa == 0) ? m1 - (splq[1].x / splq[0].x) * (m2 - m1) : m1;
    } else if (warp == 1) {
      const float me = row[(Wn - 2) * RS + 1], me2 = row[(Wn - 3) * RS + 1];
      row[(Wn - 1) * RS + 1] = (wb == n - 1) ? me + (splq[Wn - 3].x / splq[Wn - 2].x) * (me - me2) : me;
    }
  }
  __syncthreads();

  __syncthreads();
  return true;
}

// exact a14 factor of one detector sample (MODE 1): the generic searching code of matm_synth.cuh
__device__ __forceinline__ float cal_exact_f32(const SynthParams& P, int band, float u, float w, float ce, float se, int& ie) {
  const int ho = band * (P.nel - 1);
  const float den = cal_denominator<float>(P, reinterpret_cast<const float*>(P.sinel_R), reinterpret_cast<const float*>(P.Hc_R) + ho,
                                           reinterpret_cast<const float*>(P.Hs_R) + ho, fmaf(u, se, w * ce), ie);
  return 1.0f / den;
}

__global__ void __launch_bounds__(F32_THREADS, F32_MINB)
    synth_eval_f32_kernel(const __grid_constant__ SynthParams P, const int32_t* __restrict__ order, int64_t tile0,
                          int64_t det0, int64_t n_det, const float* __restrict__ yk, float* __restrict__ out, int units) {
  extern __shared__ __align__(16) unsigned char dyn[];
  float2* ym = reinterpret_cast<float2*>(dyn);                                // [W][32] {y, M}
  float4* splq = reinterpret_cast<float4*>(ym + (size_t)F32_W * F32_YS);      // [W] {1/h_k, w_k, a_k w_k, c'_k}
  __shared__ float4 dc_cal[F32_DT];                 // {P0, P1, P2, -}: 1 / factor = P0 + x (P1 + x P2)
  __shared__ __align__(16) long long dc_row[F32_DT];  // element offset of each detector's output row
  __shared__ float2 dc_uw[F32_DT];                  // rolled offsets: sin el_d = u sin(e) + w cos(e)
  __shared__ int dc_band[F32_DT];
  __shared__ int s_misc[2];
  __shared__ double s_emid;
  EvalTile T;
  if (!eval_prologue(P, order, tile0, det0, n_det, yk, units, ym, splq, dc_cal, dc_row, dc_uw, dc_band, s_misc, &s_emid, T)) return;

  const int tid = threadIdx.x;
  const int64_t j0 = T.j0, j1 = T.j1;
  const int nd = T.nd, wa = T.wa;
  const bool vec_ok = (P.n_t % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
  const uint32_t n_t32 = (uint32_t)P.n_t * 4u;  // bytes per output row
  auto phase3 = [&](auto mode_tag) {
    constexpr int MODE = decltype(mode_tag)::value;
    for (int64_t jb = (j0 & ~(int64_t)3) + 4 * (int64_t)tid; jb < j1; jb += 4 * F32_THREADS) {
      // the group's samples; those outside the tile (first / last group) repeat the nearest owned one and are not stored
      float2 W0[2], W1[2], W2[2], U0[2], U1[2], U2[2], X[2];
      int kb = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int64_t j = max(min(jb + q, j1 - 1), j0);
        const float4 w = __ldg(P.samp_w + j);  // {a1, a2, a3, interval}
        const int li = min(max(__float_as_int(w.w), T.k0), T.k1 - 1) - wa;
        if (q == 0) kb = li;
        const bool up = li != kb;  // the sample lies one interval further (at most one knot inside a group)
        const float a0 = 1.f - w.x;
        float* f;
        f = reinterpret_cast<float*>(W0) + q, *f = up ? 0.f : a0;
        f = reinterpret_cast<float*>(W1) + q, *f = up ? a0 : w.x;
        f = reinterpret_cast<float*>(W2) + q, *f = up ? w.x : 0.f;
        f = reinterpret_cast<float*>(U0) + q, *f = up ? 0.f : w.y;
        f = reinterpret_cast<float*>(U1) + q, *f = up ? w.y : w.z;
        f = reinterpret_cast<float*>(U2) + q, *f = up ? w.z : 0.f;
        if constexpr (MODE == 2) reinterpret_cast<float*>(X)[q] = (float)(P.el_full[j] - T.emid);
      }
      const float2* __restrict__ yb = ym + kb * F32_YS;
      // the third knot of the stencil: a group in the last interval of the window never looks further (its weights there
      // are zero, the address just has to be valid)
      const int o2 = (kb + 2 <= T.Wn - 1) ? 2 * F32_YS : F32_YS;
      const bool full = vec_ok && jb >= j0 && jb + 4 <= j1;
      const char* const obase = reinterpret_cast<const char*>(out + jb);
      int ie = 0;  // interval of the elevation grid found last (MODE 1)
      auto detectors = [&](auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
        constexpr int UNR = F32_UNROLL;
#pragma unroll UNR
        for (int d = 0; d < nd; d++) {
          const float2 q0 = yb[d], q1 = yb[F32_YS + d], q2 = yb[o2 + d];
          const float4 c = dc_cal[d];  // {P0, P1, P2, bits of the output row index}
          float2 v[2];
#pragma unroll
          for (int h = 0; h < 2; h++) {
            float2 t = __fmul2_rn(W0[h], f2(q0.x, q0.x));
            t = __ffma2_rn(W1[h], f2(q1.x, q1.x), t);
            t = __ffma2_rn(W2[h], f2(q2.x, q2.x), t);
            t = __ffma2_rn(U0[h], f2(q0.y, q0.y), t);
            t = __ffma2_rn(U1[h], f2(q1.y, q1.y), t);
            v[h] = __ffma2_rn(U2[h], f2(q2.y, q2.y), t);
          }
          if constexpr (MODE == 2) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              float2 den = __ffma2_rn(X[h], f2(c.z, c.z), f2(c.y, c.y));
              den = __ffma2_rn(den, X[h], f2(c.x, c.x));
              v[h] = __fmul2_rn(v[h], f2(fast_rcp(den.x), fast_rcp(den.y)));
            }
          } else if constexpr (MODE == 1) {
            const float2 uw = dc_uw[d];
            const int band = dc_band[d];
#pragma unroll
            for (int q = 0; q < 4; q++) {
              const int64_t j = max(min(jb + q, j1 - 1), j0);
              const double2 bf = *reinterpret_cast<const double2*>(P.bore_full + 2 * j);
              reinterpret_cast<float*>(v)[q] *= cal_exact_f32(P, band, uw.x, uw.y, (float)bf.x, (float)bf.y, ie);
            }
          }
          // row offset = row index x row bytes: one 32 x 32 + 64 -> 64-bit multiply-add (n_t < 2^30 on this path)
          float* o = reinterpret_cast<float*>(const_cast<char*>(obase) + (uint64_t)(uint32_t)__float_as_int(c.w) * (uint64_t)n_t32);
          if constexpr (FULL) {
            __stcs(reinterpret_cast<float4*>(o), make_float4(v[0].x, v[0].y, v[1].x, v[1].y));
          } else {
#pragma unroll
            for (int q = 0; q < 4; q++)
              if (jb + q >= j0 && jb + q < j1) __stcs(o + q, reinterpret_cast<const float*>(v)[q]);
          }
        }
      };
      if (full)
        detectors(std::true_type{});
      else
        detectors(std::false_type{});
    }
  };
  if (T.mode == 0)
    phase3(std::integral_constant<int, 0>{});
  else if (T.mode == 2)
    phase3(std::integral_constant<int, 2>{});
  else
    phase3(std::integral_constant<int, 1>{});
}

// Any sample spacing (fewer than four full-rate samples per spline interval): thread = sample, loop over the detectors.
__global__ void __launch_bounds__(F32_THREADS)
    synth_eval_general_f32_kernel(const __grid_constant__ SynthParams P, const int32_t* __restrict__ order, int64_t tile0,
                                  int64_t det0, int64_t n_det, const float* __restrict__ yk, float* __restrict__ out, int units) {
  extern __shared__ __align__(16) unsigned char dyn[];
  float2* ym = reinterpret_cast<float2*>(dyn);
  float4* splq = reinterpret_cast<float4*>(ym + (size_t)F32_W * F32_YS);
  __shared__ float4 dc_cal[F32_DT];
  __shared__ __align__(16) long long dc_row[F32_DT];
  __shared__ float2 dc_uw[F32_DT];
  __shared__ int dc_band[F32_DT];
  __shared__ int s_misc[2];
  __shared__ double s_emid;
  EvalTile T;
  if (!eval_prologue(P, order, tile0, det0, n_det, yk, units, ym, splq, dc_cal, dc_row, dc_uw, dc_band, s_misc, &s_emid, T)) return;
  for (int64_t j = T.j0 + threadIdx.x; j < T.j1; j += F32_THREADS) {
    const float4 w = __ldg(P.samp_w + j);
    const int li = min(max(__float_as_int(w.w), T.k0), T.k1 - 1) - T.wa;
    const float a0 = 1.f - w.x;
    const float x = T.mode == 2 ? (float)(P.el_full[j] - T.emid) : 0.f;
    float ce = 0.f, se = 0.f;
    if (T.mode == 1) {
      const double2 bf = *reinterpret_cast<const double2*>(P.bore_full + 2 * j);
      ce = (float)bf.x;
      se = (float)bf.y;
    }
    const float2* __restrict__ yb = ym + li * F32_YS;
    int ie = 0;
    for (int d = 0; d < T.nd; d++) {
      const float2 q0 = yb[d], q1 = yb[F32_YS + d];
      float v = fmaf(w.z, q1.y, fmaf(w.y, q0.y, fmaf(w.x, q1.x, a0 * q0.x)));
      if (T.mode == 2) {
        const float4 c = dc_cal[d];
        v *= fast_rcp(fmaf(fmaf(c.z, x, c.y), x, c.x));
      } else if (T.mode == 1) {
        const float2 uw = dc_uw[d];
        v *= cal_exact_f32(P, dc_band[d], uw.x, uw.y, ce, se, ie);
      }
      __stcs(out + dc_row[d] + j, v);
    }
  }
}

static size_t synth_f32_smem() { return (size_t)F32_W * F32_YS * sizeof(float2) + (size_t)F32_W * sizeof(float4); }
