// a8: beam smoothing of a generated layer (maria/atmosphere/atmosphere.py:328-344), i.e.
// scipy.ndimage.gaussian_filter(values[:, layer columns], sigma=(s_ext, s_x)) with its defaults
// (mode="reflect", truncate=4.0): separable, axis 0 first, weights exp(-x^2 / 2 s^2) / sum, radius int(4 s + 0.5),
// an axis with s <= 1e-15 is skipped.
#include "matm_common.cuh"

#define SM_MAX_RADIUS 2048

__device__ __forceinline__ int reflect_index(int i, int n) {
  // scipy "reflect": (d c b a | a b c d | d c b a), period 2n
  int p = 2 * n;
  i %= p;
  if (i < 0) i += p;
  return i < n ? i : p - 1 - i;
}

// one pass along `axis` (0: rows, 1: columns) of an (n_rows, n_cols) view with leading dimension ld_in
template <typename OUT>
__global__ void __launch_bounds__(256) gauss_pass_kernel(const double* __restrict__ in, int ld_in, OUT* __restrict__ out,
                                                         int n_rows, int n_cols, int axis, double sigma, int radius) {
  extern __shared__ double w[];  // 2 * radius + 1 normalised weights
  __shared__ double wsum;
  for (int k = threadIdx.x; k <= 2 * radius; k += blockDim.x) {
    double x = (double)(k - radius);
    w[k] = sigma > 1e-15 ? exp(-0.5 / (sigma * sigma) * x * x) : 1.0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int k = 0; k <= 2 * radius; k++) s += w[k];
    wsum = s;
  }
  __syncthreads();
  const double inv = 1.0 / wsum;
  const size_t total = (size_t)n_rows * n_cols;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int i = (int)(e / n_cols), j = (int)(e % n_cols);
    double acc = 0.0;
    if (axis == 0) {
      for (int k = -radius; k <= radius; k++) acc = fma(w[k + radius] * inv, in[(size_t)reflect_index(i + k, n_rows) * ld_in + j], acc);
    } else {
      const double* row = in + (size_t)i * ld_in;
      for (int k = -radius; k <= radius; k++) acc = fma(w[k + radius] * inv, row[reflect_index(j + k, n_cols)], acc);
    }
    out[e] = (OUT)acc;
  }
}

extern "C" int matm_smooth_layer(matm_handle* h, const double* values_dev, int n_ext, int ld, int col0, int n_cols,
                                 double sigma_ext, double sigma_x, void* out_dev, double* scratch_dev, void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (!values_dev || !out_dev || !scratch_dev) MATM_FAIL_ARG(h, "matm_smooth_layer: null pointer");
  if (n_ext < 1 || n_cols < 1 || col0 < 0 || col0 + n_cols > ld) MATM_FAIL_ARG(h, "matm_smooth_layer: bad shape");
  if (!(sigma_ext >= 0.0) || !(sigma_x >= 0.0)) MATM_FAIL_ARG(h, "matm_smooth_layer: bad sigma");
  cudaStream_t st = (cudaStream_t)stream;
  const int r0 = sigma_ext > 1e-15 ? (int)(4.0 * sigma_ext + 0.5) : 0;
  const int r1 = sigma_x > 1e-15 ? (int)(4.0 * sigma_x + 0.5) : 0;
  if (r0 > SM_MAX_RADIUS || r1 > SM_MAX_RADIUS) MATM_FAIL_ARG(h, "matm_smooth_layer: sigma too large");
  const size_t total = (size_t)n_ext * n_cols;
  int blocks = (int)((total + 255) / 256);
  if (blocks > h->sm_count * 16) blocks = h->sm_count * 16;
  gauss_pass_kernel<double><<<blocks, 256, (2 * r0 + 1) * sizeof(double), st>>>(values_dev + col0, ld, scratch_dev, n_ext,
                                                                                 n_cols, 0, sigma_ext, r0);
  MATM_LAUNCHED(h);
  if (h->dtype == MATM_F64)
    gauss_pass_kernel<double><<<blocks, 256, (2 * r1 + 1) * sizeof(double), st>>>(scratch_dev, n_cols, (double*)out_dev,
                                                                                   n_ext, n_cols, 1, sigma_x, r1);
  else
    gauss_pass_kernel<float><<<blocks, 256, (2 * r1 + 1) * sizeof(double), st>>>(scratch_dev, n_cols, (float*)out_dev,
                                                                                  n_ext, n_cols, 1, sigma_x, r1);
  MATM_LAUNCHED(h);
  return MATM_OK;
}
