// Handle life cycle, error reporting and the Philox normal generator of libmaria_atm.so.
#include "matm_common.cuh"

extern "C" int matm_abi_version(void) { return MATM_ABI_VERSION; }

extern "C" int matm_create(int device, int dtype, matm_handle** out) {
  if (!out || (dtype != MATM_F32 && dtype != MATM_F64)) return MATM_ERR_ARG;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return MATM_ERR_CUDA;
  matm_handle* h = new matm_handle();
  h->device = device;
  h->dtype = dtype;
  memset(&h->P, 0, sizeof(h->P));
  if (cudaSetDevice(device) != cudaSuccess) {
    delete h;
    return MATM_ERR_CUDA;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) h->sm_count = prop.multiProcessorCount;
  if (cudaMalloc((void**)&h->flags, 8 * sizeof(int)) != cudaSuccess ||
      cudaMemset(h->flags, 0, 8 * sizeof(int)) != cudaSuccess) {
    delete h;
    return MATM_ERR_CUDA;
  }
  h->P.flags = h->flags;
  *out = h;
  return MATM_OK;
}

extern "C" int matm_destroy(matm_handle* h) {
  MATM_CHECK_HANDLE(h);
  cudaSetDevice(h->device);
  cudaFree(h->bore_atm);
  cudaFree(h->bore_full);
  cudaFree(h->t_full);
  cudaFree(h->jstart);
  cudaFree(h->t_atm);
  cudaFree(h->spl);
  cudaFree(h->el_full);
  cudaFree(h->el_rng);
  cudaFree(h->knot_f);
  cudaFree(h->samp_w);
  cudaFree(h->lk_tab);
  cudaFree(h->det_cal);
  for (int i = 0; i < 2; i++) {
    cudaFree(h->yk_buf[i]);
    if (h->yk_stream[i]) cudaStreamDestroy(h->yk_stream[i]);
  }
  for (int i = 0; i < 4; i++)
    if (h->yk_ev[i]) cudaEventDestroy(h->yk_ev[i]);
  for (cudaEvent_t e : h->tev) cudaEventDestroy(e);
  for (int l = 0; l < MATM_MAX_LAYERS; l++) {
    cudaFree(h->layer_base[l]);
    cudaFree(h->layer_pairs[l]);
  }
  cudaFree(h->det_atm);
  cudaFree(h->det_full);
  cudaFree(h->mueller00);
  cudaFree(h->gain);
  cudaFree(h->band_index);
  cudaFree(h->side_pwv);
  cudaFree(h->side_el);
  cudaFree(h->G2);
  cudaFree(h->H1);
  cudaFree(h->tabR);
  cudaFree(h->opacity);
  cudaFree(h->passband);
  cudaFree(h->side_nu);
  cudaFree(h->az_full);
  cudaFree(h->det_v);
  cudaFree(h->H3);
  for (auto& kv : h->det_order) {
    cudaFree(kv.second.order);
    cudaFree(kv.second.tile_m);
  }
  cudaFree(h->flags);
  for (int i = 0; i < 2; i++) {
    cudaFree(h->stage_dev[i]);
    if (h->stage_pinned[i]) cudaFreeHost(h->stage_pinned[i]);
    if (h->stage_stream[i]) cudaStreamDestroy(h->stage_stream[i]);
    if (h->stage_done[i]) cudaEventDestroy(h->stage_done[i]);
  }
  delete h;
  return MATM_OK;
}

extern "C" int matm_last_error(matm_handle* h, char* buf, int len) {
  MATM_CHECK_HANDLE(h);
  if (!buf || len <= 0) return MATM_ERR_ARG;
  strncpy(buf, h->err.c_str(), (size_t)len - 1);
  buf[len - 1] = 0;
  return MATM_OK;
}

extern "C" int64_t matm_launch_count(matm_handle* h) { return h ? h->launches : -1; }

extern "C" int matm_check_flags(matm_handle* h, int* status, int* layer, int64_t* det, int64_t* sample) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  int f[8];
  MATM_CUDA(h, cudaDeviceSynchronize());
  MATM_CUDA(h, cudaMemcpy(f, h->flags, sizeof(f), cudaMemcpyDeviceToHost));
  MATM_CUDA(h, cudaMemset(h->flags, 0, sizeof(f)));
  int st = MATM_OK;
  if (f[0] & 1)
    st = MATM_ERR_OUT_OF_GRID;
  else if (f[0] & 2)
    st = MATM_ERR_TABLE_RANGE;
  if (status) *status = st;
  if (layer) *layer = f[1];
  if (det) *det = (int64_t)(((uint64_t)(uint32_t)f[3] << 32) | (uint32_t)f[2]);
  if (sample) *sample = (int64_t)(((uint64_t)(uint32_t)f[5] << 32) | (uint32_t)f[4]);
  return MATM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011), counter = element pair index, key = seed; Box-Muller in float64.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
  uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
  uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0;
  c[1] = n1;
  c[2] = n2;
  c[3] = n3;
}

__device__ __forceinline__ void philox4x32_10(uint64_t ctr, uint64_t subseq, uint64_t seed, uint32_t (&out)[4]) {
  uint32_t c[4] = {(uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)subseq, (uint32_t)(subseq >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; r++) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
#pragma unroll
  for (int i = 0; i < 4; i++) out[i] = c[i];
}

__global__ void __launch_bounds__(256) fill_normal_kernel(double* __restrict__ out, size_t n, uint64_t seed,
                                                          uint64_t subseq) {
  size_t pair = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t n_pairs = (n + 1) / 2;
  for (; pair < n_pairs; pair += stride) {
    uint32_t r[4];
    philox4x32_10(pair, subseq, seed, r);
    // two uniforms in (0, 1) with 53 significant bits
    double u1 = ((double)(((uint64_t)r[0] << 21) ^ (r[1] >> 11)) + 0.5) * (1.0 / 9007199254740992.0);
    double u2 = ((double)(((uint64_t)r[2] << 21) ^ (r[3] >> 11)) + 0.5) * (1.0 / 9007199254740992.0);
    double rad = sqrt(-2.0 * log(u1));
    double s, c;
    sincospi(2.0 * u2, &s, &c);
    out[2 * pair] = rad * c;
    if (2 * pair + 1 < n) out[2 * pair + 1] = rad * s;
  }
}

extern "C" int matm_fill_normal(matm_handle* h, double* out_dev, size_t n, uint64_t seed, uint64_t subsequence,
                                void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (n == 0) return MATM_OK;
  if (!out_dev) MATM_FAIL_ARG(h, "matm_fill_normal: null output");
  size_t n_pairs = (n + 1) / 2;
  int blocks = (int)((n_pairs + 255) / 256);
  int cap = h->sm_count * 8;
  if (blocks > cap) blocks = cap;
  fill_normal_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(out_dev, n, seed, subsequence);
  MATM_LAUNCHED(h);
  return MATM_OK;
}
