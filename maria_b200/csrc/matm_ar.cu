// a5: AutoregressiveProcess.run (maria/atmosphere/process.py:191-209) on the device.
//
//   BUFFER[i] = A @ BUFFER[i + 1 + esi, csi] + B @ z_i      for i = 2 n_ext - 1 ... 0
//
// Two kernels.  (1) bz_gemm_kernel: the only genuinely dense contraction, B @ z for ALL steps at once
// (M = 2 n_ext, N = K = n_cs), written straight into BUFFER rows [0, 2 n_ext).  (2) ar_chain_kernel: the
// sequential recurrence.  One thread-block cluster per process; every CTA keeps its slice of rows of A resident
// in shared memory (staged once with the TMA bulk-copy engine) for all 2 n_ext steps, the gather vector is
// rebuilt from the L2-resident BUFFER each step and the CTAs of the cluster meet at a hardware cluster barrier.
#include <cooperative_groups.h>

#include "matm_common.cuh"

namespace cg = cooperative_groups;

// ---------------------------------------------------------------------------------------------------------
// C[i][r] = sum_c Z[M-1-i][c] * B[r][c]     (both operands K-contiguous), 64x64x16 tiles, 4x4 per thread
// ---------------------------------------------------------------------------------------------------------
#define GT 64
#define GK 16
__global__ void __launch_bounds__(256) bz_gemm_kernel(const double* __restrict__ Z, const double* __restrict__ B,
                                                      double* __restrict__ C, int M, int N) {
  __shared__ double sZ[GK][GT + 1];
  __shared__ double sB[GK][GT + 1];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int i0 = blockIdx.x * GT, r0 = blockIdx.y * GT;
  double acc[4][4] = {};
  for (int k0 = 0; k0 < N; k0 += GK) {
    for (int e = threadIdx.x; e < GT * GK; e += 256) {
      int row = e / GK, k = e % GK;
      int i = i0 + row, r = r0 + row, c = k0 + k;
      sZ[k][row] = (i < M && c < N) ? Z[(size_t)(M - 1 - i) * N + c] : 0.0;
      // B is lower triangular (process.py:174): entries above the diagonal are exact zeros, read as stored
      sB[k][row] = (r < N && c < N) ? B[(size_t)r * N + c] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GK; k++) {
      double a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        a[u] = sZ[k][ty * 4 + u];
        b[u] = sB[k][tx * 4 + u];
      }
#pragma unroll
      for (int u = 0; u < 4; u++)
#pragma unroll
        for (int v = 0; v < 4; v++) acc[u][v] = fma(a[u], b[v], acc[u][v]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int u = 0; u < 4; u++) {
    int i = i0 + ty * 4 + u;
    if (i >= M) continue;
#pragma unroll
    for (int v = 0; v < 4; v++) {
      int r = r0 + tx * 4 + v;
      if (r < N) C[(size_t)i * N + r] = acc[u][v];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// TMA bulk copy (global -> shared, 1-D) completing on an mbarrier: stages the resident slice of A.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// The chain.  grid = CL CTAs forming ONE cluster; CTA c owns rows [c*rpc, min(n_cs, (c+1)*rpc)).
// Dynamic shared memory: [ A slice (rpc * n_sample doubles) if a_in_smem ] [ g (n_sample doubles) ]
//                        [ gidx (n_sample ints) ] [ mbarrier ]
// ---------------------------------------------------------------------------------------------------------
#define AR_THREADS 512
template <int CL>
__global__ void __launch_bounds__(AR_THREADS, 1)
    ar_chain_kernel(const double* __restrict__ A, const int* __restrict__ esi, const int* __restrict__ csi,
                    double* buffer, int n_ext, int n_cs, int n_sample, int rpc, int a_in_smem) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int cta = (CL > 1) ? (int)cg::this_cluster().block_rank() : 0;
  const int row0 = cta * rpc;
  const int rows = max(0, min(n_cs, row0 + rpc) - row0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = AR_THREADS / 32;

  double* sA = reinterpret_cast<double*>(smem_raw);
  double* g = sA + (a_in_smem ? (size_t)rpc * n_sample : 0);
  int* gidx = reinterpret_cast<int*>(g + n_sample);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + (((size_t)((unsigned char*)(gidx + n_sample) - smem_raw) + 15) & ~(size_t)15));

  const double* Arows = A + (size_t)row0 * n_sample;
  if (a_in_smem && rows > 0) {
    // rows are contiguous in A, so the slice is one contiguous range: bulk copies of <= 64 KiB, 16-byte granules
    const size_t total = (size_t)rows * n_sample * sizeof(double);
    const bool aligned = ((reinterpret_cast<uintptr_t>(Arows) & 15) == 0) && (total % 16 == 0);
    if (aligned) {
      if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      __syncthreads();
      if (tid == 0) {
        mbar_expect_tx(bar, (uint32_t)total);
        for (size_t off = 0; off < total; off += 65536) {
          uint32_t nb = (uint32_t)min((size_t)65536, total - off);
          tma_bulk_g2s(reinterpret_cast<unsigned char*>(sA) + off, reinterpret_cast<const unsigned char*>(Arows) + off,
                       nb, bar);
        }
      }
      mbar_wait(bar, 0);
    } else {
      for (size_t e = tid; e < (size_t)rows * n_sample; e += AR_THREADS) sA[e] = Arows[e];
    }
  }
  for (int s = tid; s < n_sample; s += AR_THREADS) gidx[s] = (esi[s] + 1) * n_cs + csi[s];
  __syncthreads();
  if (CL > 1) cg::this_cluster().sync();

  const double* Ause = a_in_smem ? sA : Arows;
  for (int i = 2 * n_ext - 1; i >= 0; i--) {
    double* brow = buffer + (size_t)i * n_cs;
    for (int s = tid; s < n_sample; s += AR_THREADS) g[s] = __ldcg(brow + gidx[s]);
    __syncthreads();
    for (int r = warp; r < rows; r += nwarps) {
      const double* a = Ause + (size_t)r * n_sample;
      double acc = 0.0;
      for (int s = lane; s < n_sample; s += 32) acc = fma(a[s], g[s], acc);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) brow[row0 + r] += acc;  // the row already holds B @ z_i
    }
    if (CL > 1) {
      cg::this_cluster().sync();  // release/acquire at cluster scope: the new row is visible to every CTA
    } else {
      __syncthreads();
    }
  }
}

template <int CL>
static int launch_chain(matm_handle* h, const double* A, const int* esi, const int* csi, double* buffer, int n_ext,
                        int n_cs, int n_sample, int rpc, int a_in_smem, size_t smem, cudaStream_t stream) {
  MATM_CUDA(h, cudaFuncSetAttribute(ar_chain_kernel<CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CL, 1, 1);
  cfg.blockDim = dim3(AR_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  MATM_CUDA(h, cudaLaunchKernelEx(&cfg, ar_chain_kernel<CL>, A, esi, csi, buffer, n_ext, n_cs, n_sample, rpc, a_in_smem));
  h->launches++;
  return MATM_OK;
}

extern "C" int matm_ar_generate(matm_handle* h, int n_ext, int n_cs, int n_sample, const double* A_dev,
                                const double* B_dev, const int32_t* esi_dev, const int32_t* csi_dev,
                                double* buffer_dev, const double* z_dev, void* stream) {
  MATM_CHECK_HANDLE(h);
  if (matm_set_device(h)) return MATM_ERR_CUDA;
  if (n_ext < 1 || n_cs < 1 || n_sample < 1) MATM_FAIL_ARG(h, "matm_ar_generate: empty process");
  if (!A_dev || !B_dev || !esi_dev || !csi_dev || !buffer_dev || !z_dev)
    MATM_FAIL_ARG(h, "matm_ar_generate: null pointer");
  if ((int64_t)3 * n_ext * n_cs > (int64_t)INT32_MAX) MATM_FAIL_ARG(h, "matm_ar_generate: process too large");
  cudaStream_t st = (cudaStream_t)stream;
  const int M = 2 * n_ext;
  dim3 grid((M + GT - 1) / GT, (n_cs + GT - 1) / GT);
  bz_gemm_kernel<<<grid, 256, 0, st>>>(z_dev, B_dev, buffer_dev, M, n_cs);
  MATM_LAUNCHED(h);

  // smallest cluster whose per-CTA slice of A fits in shared memory; above 8 CTAs A streams from L2
  const size_t budget = 200 * 1024;
  const size_t fixed = (size_t)n_sample * (sizeof(double) + sizeof(int)) + 64;
  int cl = 1, a_in_smem = 0;
  for (int c = 1; c <= 8; c *= 2) {
    int rpc = (n_cs + c - 1) / c;
    cl = c;
    if ((size_t)rpc * n_sample * sizeof(double) + fixed <= budget) {
      a_in_smem = 1;
      break;
    }
  }
  // more CTAs than needed only when the rows give every warp work: keeps the per-step dot product short
  while (cl < 8 && (n_cs + cl - 1) / cl > 64) cl *= 2;
  const int rpc = (n_cs + cl - 1) / cl;
  if (!a_in_smem && fixed > budget) MATM_FAIL_ARG(h, "matm_ar_generate: stencil too large for shared memory");
  const size_t smem = (a_in_smem ? (size_t)rpc * n_sample * sizeof(double) : 0) + fixed;
  switch (cl) {
    case 1: return launch_chain<1>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, rpc, a_in_smem, smem, st);
    case 2: return launch_chain<2>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, rpc, a_in_smem, smem, st);
    case 4: return launch_chain<4>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, rpc, a_in_smem, smem, st);
    default: return launch_chain<8>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, rpc, a_in_smem, smem, st);
  }
}
