/* libmaria_atm.so -- C-ABI of the B200 atmospheric TOD synthesis path.
 *
 * The reference (thomaswmorris/maria) is pure Python and has no FFI; the seam this library sits behind is
 * the duck-typed Python surface of maria/atmosphere/atmosphere.py, maria/sim/atmosphere.py and
 * maria/tod/tod.py.  Every entry point cites the reference lines it replaces.  INTEGRATION.md shows the
 * ctypes binding a maintainer would add.
 *
 * Conventions
 *   - every function returns an int status (MATM_OK == 0); matm_last_error() gives the text;
 *   - pointers named *_dev are device pointers owned by the caller (torch tensors in the Python host);
 *     pointers named *_host are host pointers, copied during the call; the library never frees caller memory;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream); all work is enqueued on
 *     it and nothing synchronises unless stated;
 *   - one handle per GPU, no global state; calls on one handle are serialised by the caller.
 */
#ifndef MARIA_ATM_H
#define MARIA_ATM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct matm_handle matm_handle;

enum {
  MATM_OK = 0,
  MATM_ERR_CUDA = 1,        /* a CUDA runtime call failed */
  MATM_ERR_ARG = 2,         /* bad argument / call order (RuntimeError "must be initialized", atmosphere.py:294-295) */
  MATM_ERR_OUT_OF_GRID = 3, /* a line of sight left a layer grid (RuntimeError, atmosphere.py:368-369) */
  MATM_ERR_TABLE_RANGE = 4  /* (T0, pwv, el) left the spectrum table: NaN in the output (band.py:283-286) */
};

enum { MATM_F32 = 0, MATM_F64 = 1 };           /* arithmetic + storage type of layers / TOD */
enum { MATM_UNITS_PW = 0, MATM_UNITS_KRJ = 1 }; /* TOD units (tod.py:106-142) */

#define MATM_MAX_LAYERS 64  /* the "3d" model has one layer per resolution element of height (extrusion.py:69-77) */
#define MATM_MAX_BANDS 16

/* One turbulent layer as the sampler sees it (atmosphere.py:317-373). */
typedef struct {
  const void* data_dev; /* smoothed layer, (n_ext, n_cs) row-major, float or double per `dtype` of the handle */
  int32_t n_ext, n_cs;
  double h;            /* layer height above the telescope [m]            (atmosphere.py:346) */
  double pwv_rms;      /* weight of this layer in zenith_scaled_pwv [mm]  (atmosphere.py:373) */
  double transform[9]; /* row-major 3x3, q = p @ transform                (atmosphere.py:347) */
  double ext0, ext_step;   /* extrusion grid  = ext0 + i*ext_step, i < n_ext      (atmosphere.py:241-245) */
  double cs0, cs_last;     /* cross-section x grid = linspace(cs0, cs_last, n_cs) (atmosphere.py:210-216) */
  const double* vx_host;   /* (n_t_atm,) wind east  [m/s] of the layer's process  (atmosphere.py:141-151) */
  const double* vy_host;   /* (n_t_atm,) wind north [m/s] */
} matm_layer_desc;

/* ---- life cycle -------------------------------------------------------------------------------------- */
int matm_create(int device, int dtype, matm_handle** out);
int matm_destroy(matm_handle* h);
int matm_last_error(matm_handle* h, char* buf, int len);
int matm_abi_version(void);

/* ---- a5: AutoregressiveProcess.run (process.py:191-209) ----------------------------------------------- */
/* Standard-normal float64 noise, Philox4x32-10 counter RNG + Box-Muller (replaces np.random.standard_normal,
 * process.py:196,206, in production mode).  Element i depends only on (seed, subsequence, i). */
int matm_fill_normal(matm_handle* h, double* out_dev, size_t n, uint64_t seed, uint64_t subsequence, void* stream);

/* buffer_dev: (3*n_ext, n_cs) float64; on entry rows [2*n_ext, 3*n_ext) hold the initial white noise
 * (process.py:196-198; rows below are overwritten, so their initial content is never read);
 * z_dev: (2*n_ext, n_cs) float64, z_dev[k] = the vector drawn at the k-th executed step (buffer row
 * 2*n_ext-1-k, process.py:206).  On return (stream order) rows [0, n_ext) of buffer_dev are `values`.
 * A_dev (n_cs, n_sample), B_dev (n_cs, n_cs) float64 row-major; esi_dev/csi_dev (n_sample,) int32 are
 * extrusion_sample_index / cross_section_sample_index (process.py:44-67), esi non-decreasing as the reference builds
 * it.  n_near = number of leading stencil samples with esi + 1 <= 3 (they are served from an on-chip ring of the last
 * three rows); pass -1 to force the generic kernel.  The first call for a stencil (device pointer of esi + shape) copies
 * esi / csi back to the host once (a blocking cudaMemcpy) to size the on-chip delay lines of the far samples and caches the
 * result in the handle: the stencil arrays must not change behind the same pointer while the handle lives. */
int matm_ar_generate(matm_handle* h, int n_ext, int n_cs, int n_sample, int n_near, const double* A_dev, const double* B_dev,
                     const int32_t* esi_dev, const int32_t* csi_dev, double* buffer_dev, const double* z_dev,
                     void* stream);

/* ---- a8: beam smoothing (atmosphere.py:328-344; scipy.ndimage.gaussian_filter, reflect, truncate=4) ---- */
/* values_dev: (n_ext, ld) float64, the layer is columns [col0, col0+n_cols); out_dev: (n_ext, n_cols) in the
 * handle's dtype; scratch_dev: (n_ext, n_cols) float64. */
int matm_smooth_layer(matm_handle* h, const double* values_dev, int n_ext, int ld, int col0, int n_cols,
                      double sigma_ext, double sigma_x, void* out_dev, double* scratch_dev, void* stream);

/* ---- geometry / tables (set-up; small host arrays are copied) ----------------------------------------- */
/* Boresight at the atmosphere rate (Coordinates.downsample, coordinates.py:286-304) and at the full rate
 * (observation.py:46-52), and the time axes.  t_atm_host (n_t_atm,) are the knot times exactly as the host computed
 * them (np.arange(t.min(), t.max(), timestep): nominally uniform, but rounded at ulp(t) ~ 2e-7 s, which the
 * interpolating spline of sim/atmosphere.py:73-80 sees); any increasing knot vector is accepted.  Synchronises. */
int matm_set_scan(matm_handle* h, int n_t_atm, const double* t_atm_host, double timestep, const double* az_atm_host,
                  const double* el_atm_host, int64_t n_t, const double* t_host, const double* az_host,
                  const double* el_host);

/* Layers; must follow matm_set_scan (the wind translation cumsum, atmosphere.py:318-319, is evaluated here).
 * pwv0 = weather.pwv (atmosphere.py:309).  Synchronises. */
int matm_set_layers(matm_handle* h, int n_layers, const matm_layer_desc* layers_host, double pwv0);

/* Re-points the sampler at another set of smoothed-layer buffers of the same shapes (double buffering: the next
 * realisation is beam-smoothed into the spare set while a synthesis kernel still reads the current one).  Host-side
 * only: kernels launched afterwards see the new pointers; does not synchronise. */
int matm_set_layer_data(matm_handle* h, int n_layers, const void* const* data_dev);

/* Detectors: offsets_atm (n_det, 2) are the offsets the atmosphere uses (atmosphere.py:102-105), offsets_full
 * (n_det, 2) the roll-rotated offsets of obs.coords (observation.py:55-58) used by TOD.to for the elevation;
 * band_index (n_det,) int32; mueller00 (n_det,) = M[0,0] (array.py:204-218); gain (n_det,) = exp(gain_error*N(0,1))
 * (simulation.py:239-241), NULL = ones.  Synchronises. */
int matm_set_detectors(matm_handle* h, int64_t n_det, const double* offsets_atm_host, const double* offsets_full_host,
                       const int32_t* band_index_host, const double* mueller00_host, const double* gain_host);

/* Replaces the per-detector gain factors only (simulation.py:239-241 draws new ones for every observation): one
 * host -> device copy ordered on `stream`; does not synchronise the device. */
int matm_set_gain(matm_handle* h, const double* gain_host, void* stream);

/* Spectrum tables (band.py:235-255,264-300; spectrum/atmosphere.py:97-104).  emission/opacity: (nT, npwv, nel, nnu)
 * float64 row-major; passband: (n_bands, nnu) = Band.passband(side_nu); T0 = weather.temperature[0];
 * cal_T0 / cal_pwv = the rounded metadata values TOD.to uses (tod.py:90-104); cal_factor[b] = 0.5 if the band is
 * polarized else 1 (tod.py:94).  The band integrals (trapezoid over nu) are evaluated on the device.  Synchronises. */
int matm_set_bands(matm_handle* h, int n_bands, int nT, int npwv, int nel, int nnu, const double* side_T_host,
                   const double* side_pwv_host, const double* side_el_host, const double* side_nu_host,
                   const double* emission_host, const double* opacity_host, const double* passband_host, double T0,
                   double cal_T0, double cal_pwv, const double* cal_factor_host);

/* ---- a6 + a7 + a9: Atmosphere.simulate_pwv sampling (atmosphere.py:293-380) ---------------------------- */
/* out_dev: (n_det, n_t_atm) float64 zenith_scaled_pwv for detectors [det0, det0+n_det). */
int matm_zenith_pwv(matm_handle* h, int64_t det0, int64_t n_det, double* out_dev, void* stream);

/* ---- a10: linear upsample to the full rate (sim/atmosphere.py:30-37) ---------------------------------- */
int matm_upsample_linear(matm_handle* h, const double* in_dev, int64_t n_det, double* out_dev, void* stream);

/* ---- a6..a14 fused: loading -> cubic upsample -> gain -> (optionally) pW -> K_RJ ----------------------- */
/* out_dev: (n_det, n_t) in the handle's dtype.  apply_gain = 0 reproduces _compute_atmospheric_loading alone
 * (sim/atmosphere.py:39-84). */
int matm_synthesize(matm_handle* h, int64_t det0, int64_t n_det, void* out_dev, int units, int apply_gain,
                    void* stream);

/* Tuning of the float path of matm_synthesize: it runs per chunk of detector tiles as two kernels (loading at the
 * atmosphere-rate knots into a scratch of `chunk_mib` MiB, then spline + evaluation + calibration at the full rate);
 * overlap != 0 alternates consecutive chunks between two scratch buffers and two internal streams.  Defaults: 1024, 1. */
int matm_set_tuning(matm_handle* h, int chunk_mib, int overlap);

/* Per-kernel timing of the float path (measurement only): with enable != 0 every chunk's two launches are bracketed by
 * CUDA events on the stream they run on; matm_get_timing synchronises the device, returns the summed durations [ms] of the
 * loading and the evaluation kernels since the last call (they overlap in wall time when chunks alternate between the two
 * internal streams) and resets the counters. */
int matm_set_timing(matm_handle* h, int enable);
int matm_get_timing(matm_handle* h, double* loading_ms, double* eval_ms, int64_t* n_chunks);

/* Allocation-free variant of matm_synthesize for float32 handles.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same kernels, same detector order and bit-identical output to matm_synthesize, but the knot-value scratch
 * (ceil(n_det / 32) tiles x n_t_atm knots x 128 B) lives in one file-scope `static __device__` array of
 * MATM_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; the call makes no cudaMalloc / cudaFree /
 * stream or event creation.  Returns MATM_ERR_ARG (with the sizes in matm_last_error) when the scratch needed exceeds the
 * maximum -- split the detector range -- for float64 handles, and when the processing order of (det0, n_det) is not cached
 * yet: call matm_synthesize once for the same range at set-up (that call allocates the order), outside any capture; the first
 * call on a device also resolves the array's device address (cudaGetSymbolAddress).  Every call shares the array, which keeps
 * the last call's contents (nothing needs to start at zero: the loading kernel overwrites every element the evaluation kernel
 * reads); calls that overlap on the GPU RACE -- issue them on one stream or order them yourself. */
#define MATM_NOALLOC_SCRATCH_BYTES (32u << 20)
int matm_synthesize_noalloc(matm_handle* h, int64_t det0, int64_t n_det, void* out_dev, int units, int apply_gain,
                            void* stream);

/* Same, delivering into HOST memory (pinned or pageable): detector-chunked, device->host copies overlapped with
 * the kernels of the next chunk.  Synchronises before returning. */
int matm_synthesize_host(matm_handle* h, int64_t det0, int64_t n_det, void* out_host, int units, int apply_gain);

/* ---- a14 alone: TOD.to (tod.py:106-142 -> calibration/functions.py:73-90) ------------------------------ */
/* to_units = MATM_UNITS_KRJ: pW -> K_RJ; MATM_UNITS_PW: K_RJ -> pW (calibration/functions.py:48-70).  In place allowed. */
int matm_calibrate(matm_handle* h, int64_t det0, int64_t n_det, const void* in_dev, void* out_dev, int to_units,
                   void* stream);

/* ---- consumers of the path's intermediates, tile-wise on the device -------------- */
/* Map step (sim/map.py:117-133): out_dev (n_det, n_t) in the handle's dtype = 1e12 k_B * compute_transmission_integral
 * (band.py:235-255) of each detector's band over the channel [nu_min_Hz, nu_max_Hz), looked up at (T0, zenith_scaled_pwv
 * upsampled linearly to the full rate (sim/atmosphere.py:30-37), full-rate elevation of the detector): pW per K_RJ.  The
 * (n_det, n_t) float64 zenith_scaled_pwv of the reference is never materialised. */
int matm_transmission(matm_handle* h, int64_t det0, int64_t n_det, double nu_min_Hz, double nu_max_Hz, void* out_dev,
                      void* stream);

/* Noise step (sim/noise.py:35-37,52-61): out = 1e12 * (NEP[band] + NEP_per_loading[band] * loading) * unscaled_noise, all
 * (n_det, n_t) in the handle's dtype, loading in pW, NEP in W sqrt(s); nep_host / nep_per_loading_host: (n_bands,).
 * In place allowed (out_dev == noise_dev). */
int matm_noise_scale(matm_handle* h, int64_t det0, int64_t n_det, const void* loading_dev, const void* noise_dev,
                     const double* nep_host, const double* nep_per_loading_host, void* out_dev, void* stream);

/* obs.coords on demand (sim/observation.py:55-58 -> coords/transforms.py:10-29): float64 az / el (n_det, n_j) of the
 * detectors [det0, det0 + n_det) at the full-rate samples [j0, j0 + n_j); either output may be NULL. */
int matm_pointing(matm_handle* h, int64_t det0, int64_t n_det, int64_t j0, int64_t n_j, double* az_dev, double* el_dev,
                  void* stream);

/* ---- error flags --------------------------------------------------------------------------------------- */
/* Reads (and clears) the device flag word; synchronises the device.  status = MATM_OK, MATM_ERR_OUT_OF_GRID or
 * MATM_ERR_TABLE_RANGE; layer/det/sample describe the first recorded out-of-grid sample. */
int matm_check_flags(matm_handle* h, int* status, int* layer, int64_t* det, int64_t* sample);

/* Number of kernels this handle has launched since creation (bench.py's gpu_launches). */
int64_t matm_launch_count(matm_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* MARIA_ATM_H */
