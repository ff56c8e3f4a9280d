// Shared declarations of libmaria_atm.so (handle state, device-side descriptors, error helpers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <string>
#include <utility>
#include <vector>

#include "../../include/maria_atm.h"

#define MATM_ABI_VERSION 1

// Device-side layer descriptor: everything the sampler needs per layer, in cell units.
struct LayerDev {
  const void* data;   // (n_ext, n_cs) row-major, float or double
  void* data2;        // fp32 handles: (n_ext, n_cs) float2 = {v[i][j], v[i][j+1]}, refreshed from `data` by every synthesis
  const double* base; // (n_t_atm, 2): boresight position in the layer grid, in cells (f64)
  int n_ext, n_cs;
  double D[4];        // 2x2: cells per unit of (pp_d - pp_b): [x<-x, x<-y, y<-x, y<-y]
  double pwv_rms;
};

struct SynthParams {
  int n_layers;
  LayerDev layers[MATM_MAX_LAYERS];
  double pwv0;
  // scan
  int n_t_atm;
  int64_t n_t;
  double t_atm0, timestep;
  const double* bore_atm;  // (n_t_atm, 4): ce, se, ca, sa  (e = el_b - pi/2, a = az_b)
  const double* bore_full; // (n_t, 2): ce, se at the full rate
  const double* t_full;    // (n_t,)
  const int64_t* jstart;   // (n_t_atm + 1,): first full-rate sample with t >= t_atm[k]; jstart[n_t_atm] = n_t
  const double* t_atm;     // (n_t_atm,): the knot times exactly as np.arange produced them
  const double* spl;       // (n_t_atm, 5): not-a-knot cubic spline system in units of `timestep`:
                           //   [0] 1/h_k  [1] h_k^2/6 (interval k);  [2] w_k  [3] a_k w_k  [4] c'_k (row k, Thomas)
  const double* el_full;   // (n_t,): boresight elevation at the full rate
  const double* el_rng;    // (n_t_atm, 2): min / max of el_full over the samples of spline interval k (+inf / -inf if empty)
  const float4* knot_f;    // (n_t_atm,): bore_atm in float (fp32 handles)
  const float4* samp_w;    // (n_t,): {a1, a2, a3, bits of the interval index i}: cubic-spline weights of the full-rate sample,
                           //   value = y_i + a1 (y_i+1 - y_i) + a2 M_i + a3 M_i+1  (sample_weights_kernel)
  const void* lk_tab;      // (n_layers, n_t_atm) LayerKnot<float>: boresight cell + fraction in every layer grid (fp32 handles)
  // detectors
  int64_t n_det;
  const double* det_atm;   // (n_det, 4): u, Y, w-1, unused   with the atmosphere's offsets
  const double* det_full;  // (n_det, 2): u, w                with the rolled offsets
  const double* det_cal;   // (n_det, 2): beta, eps with u = (1 - eps) sin beta, w = (1 - eps) cos beta, i.e.
                           //   sin el_d = (1 - eps) sin(el_b - beta)
  const int32_t* band_index;
  const double* mueller00;
  const double* gain;
  // bands
  int n_bands, npwv, nel;
  const double* side_pwv;  // (npwv,)
  const double* side_el;   // (nel,)
  const double* G2;        // (n_bands, npwv, nel) emission grid folded at T0 [pW]
  const double* H1;        // (n_bands, nel) transmission integral folded at (cal_T0, cal_pwv), times factor*k_B*1e12
  // the same tables in the handle's arithmetic type R (float or double), read by the kernels
  const void* pwv_R;       // (npwv,)
  const void* el_R;        // (nel,)
  const void* sinel_R;     // (nel,): sin(side_el), 2.0 where side_el >= pi/2 (search key for sin(el))
  const void* G2_R;        // (n_bands, npwv, nel)
  const void* Hc_R;        // (n_bands, nel-1): H(el) = Hc + Hs * el on el-interval i  (same line as the lerp of H1)
  const void* Hs_R;        // (n_bands, nel-1)
  float pwv_lo_f, pwv_hi_f, el_lo_f, el_hi_f;  // ends of the fp32 (pwv, el) grids
  float det_rmax_f;        // largest angular distance of a detector (rolled offsets) from the boresight [rad]
  double det_rmax;
  int table_ok;            // 0 if T0 / cal_T0 / cal_pwv left the table (results are NaN, as the reference)
  int* flags;              // [0] = status bits, [1] = layer, [2..3] = det (lo, hi), [4..5] = sample (lo, hi)
};

struct matm_handle {
  int device = 0;
  int dtype = MATM_F32;
  std::string err;
  int64_t launches = 0;
  int sm_count = 148;
  bool have_scan = false, have_layers = false, have_dets = false, have_bands = false;
  SynthParams P;
  // owned device arrays
  double *bore_atm = nullptr, *bore_full = nullptr, *t_full = nullptr;
  int64_t* jstart = nullptr;
  double *t_atm = nullptr, *spl = nullptr, *el_full = nullptr, *el_rng = nullptr, *det_cal = nullptr;
  float4* knot_f = nullptr;
  float4* samp_w = nullptr;
  void* lk_tab = nullptr;
  // atmosphere-rate loading of the detector chunks in flight (fp32 production path): [tile][knot][32 slots]
  float* yk_buf[2] = {nullptr, nullptr};
  size_t yk_bytes = 0;
  bool yk_used = false;
  // optional per-kernel timing of the float synthesis (matm_set_timing / matm_get_timing): events around every launch
  bool timing = false;
  std::vector<cudaEvent_t> tev;  // triples (before loading, between, after evaluation) per chunk
  size_t tev_used = 0;
  int chunk_mib = 1024; // scratch per chunk.  Measured at C4 (ms per step): 16 MiB 28.6, 32 28.55, 128 28.2, 1024 28.0, one
                        // piece 28.0: keeping the scratch L2-resident buys nothing, fewer launch tails do
  bool overlap = true;  // alternate chunks between two side streams
  cudaStream_t yk_stream[2] = {nullptr, nullptr};
  cudaEvent_t yk_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  double* layer_base[MATM_MAX_LAYERS] = {};
  float2* layer_pairs[MATM_MAX_LAYERS] = {};  // pair-expanded copies of the smoothed layers (fp32 handles)
  double *det_atm = nullptr, *det_full = nullptr, *mueller00 = nullptr, *gain = nullptr;
  int32_t* band_index = nullptr;
  double *side_pwv = nullptr, *side_el = nullptr, *G2 = nullptr, *H1 = nullptr;
  // kept for the consumers of the path (matm_consumers.cu)
  double *opacity = nullptr, *passband = nullptr, *side_nu = nullptr, *az_full = nullptr, *det_v = nullptr, *H3 = nullptr;
  std::vector<double> nu_host;
  int nT = 0, nnu = 0, iT = 0;
  double wT = 0.0;
  bool T_ok = false;
  void* tabR = nullptr;  // one allocation holding the R-typed tables
  // processing order of the detectors of a requested range (spatially coherent tiles; matm_detector_order)
  std::vector<double> det_xy_host;  // (n_det, 2) the atmosphere's offsets, kept for the ordering
  struct DetOrder {
    int32_t* order;   // (n_det) absolute detector indices in processing order
    uint8_t* tile_m;  // (n_tiles) 1, 2 or 4: the 32 detectors of the tile come in aligned groups of that many co-located ones
  };
  std::map<std::pair<int64_t, int64_t>, DetOrder> det_order;
  // per stencil (device pointer of esi, n_sample, n_near): how many far samples the chain kernel serves from delay lines in
  // shared memory, their total length, and whether the rest may take the global path (matm_ar.cu)
  struct ArStencil {
    int n_ring, ring_doubles, fence_mask, ok;
  };
  std::map<std::pair<const void*, int64_t>, ArStencil> ar_stencil;
  int* flags = nullptr;
  std::vector<double> az_atm_host, el_atm_host;
  int64_t min_samples_per_interval = 0;  // over the interior spline intervals (decides the 4-samples-per-thread path)
  // e2e staging
  void* stage_dev[2] = {nullptr, nullptr};
  void* stage_pinned[2] = {nullptr, nullptr};
  size_t stage_bytes = 0;
  cudaStream_t stage_stream[2] = {nullptr, nullptr};
  cudaEvent_t stage_done[2] = {nullptr, nullptr};
};

#define MATM_CHECK_HANDLE(h) \
  if (!(h)) return MATM_ERR_ARG;

#define MATM_CUDA(h, call)                                                                      \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess) {                                                                   \
      (h)->err = std::string(#call) + ": " + cudaGetErrorString(e__) + " (" + __FILE__ + ":" +  \
                 std::to_string(__LINE__) + ")";                                                \
      return MATM_ERR_CUDA;                                                                     \
    }                                                                                           \
  } while (0)

#define MATM_FAIL_ARG(h, msg) \
  do {                        \
    (h)->err = (msg);         \
    return MATM_ERR_ARG;      \
  } while (0)

#define MATM_LAUNCHED(h)                         \
  do {                                           \
    (h)->launches++;                             \
    MATM_CUDA(h, cudaGetLastError());            \
  } while (0)

template <typename T>
static inline int matm_upload(matm_handle* h, T** dst, const T* src_host, size_t n) {
  if (*dst) {
    MATM_CUDA(h, cudaFree(*dst));
    *dst = nullptr;
  }
  if (n == 0) return MATM_OK;
  MATM_CUDA(h, cudaMalloc((void**)dst, n * sizeof(T)));
  if (src_host) MATM_CUDA(h, cudaMemcpy(*dst, src_host, n * sizeof(T), cudaMemcpyHostToDevice));
  return MATM_OK;
}

// Device array of n_det absolute detector indices: the order in which the detectors of [det0, det0 + n_det) are
// processed.  Consecutive entries are neighbours in the focal plane (Morton order of the offsets), so that the 32
// detectors of a warp sample the same few cells of every layer whatever order the caller's array is in.
int matm_detector_order(matm_handle* h, int64_t det0, int64_t n_det, const int32_t** order_dev, const uint8_t** tile_m_dev);

static inline int matm_set_device(matm_handle* h) {
  MATM_CUDA(h, cudaSetDevice(h->device));
  return MATM_OK;
}
