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

// ---------------------------------------------------------------------------------------------------------
// The chain, latency-optimised (v2).  What bounds a step is not arithmetic but the round trips on its critical path,
// so everything that is not the previous row is taken off that path:
//   * A lives in REGISTERS: warp w of CTA c owns RW rows, lane l holds their entries for samples l, l+32, ...;
//   * the last three rows (stencil offsets 1..3) live in a 4-slot shared-memory ring (row R -> slot R & 3), written
//     by the owners of each element -- through distributed shared memory into every CTA of the cluster;
//   * stencil rows further back (offsets >= 4) and the row's own B z term are read from L2 two steps ahead of their
//     use (register pipeline of depth 2), so their latency never shows;
//   * one barrier per step: __syncthreads (1 CTA) or the hardware cluster barrier.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void st_dsmem_f64(double* local_ptr, int cta, double v) {
  uint32_t a = smem_u32(local_ptr), r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(cta));
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(r), "d"(v) : "memory");
}

template <int RW, int M>
__global__ void __launch_bounds__(AR_THREADS, 1)
    ar_chain_v2_kernel(const double* __restrict__ A, const int* __restrict__ esi, const int* __restrict__ csi,
                       double* buffer, int n_ext, int n_cs, int n_sample, int n_near, int cl) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* hist = reinterpret_cast<double*>(smem_raw);  // [4][n_cs]
  const int n_far = n_sample - n_near;
  double* far = hist + 4 * (size_t)n_cs;               // [2][n_far]
  int cta = 0;
  if (cl > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta));
  const int rpc = (n_cs + cl - 1) / cl;
  const int row0 = cta * rpc;
  const int rows = max(0, min(n_cs, row0 + rpc) - row0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- per-thread constants: A entries and the stencil descriptors of this lane's samples ----
  double a[RW][M];
  int desc[M];  // near sample: (offset << 24) | column ; far sample: index among the far samples ; -1: none
#pragma unroll
  for (int m = 0; m < M; m++) {
    const int s = lane + 32 * m;
    desc[m] = -1;
    if (s < n_sample) desc[m] = s < n_near ? (((esi[s] + 1) << 24) | csi[s]) : (s - n_near);
#pragma unroll
    for (int rr = 0; rr < RW; rr++) {
      const int r = warp * RW + rr;
      a[rr][m] = (s < n_sample && r < rows) ? A[(size_t)(row0 + r) * n_sample + s] : 0.0;
    }
  }
  // far sample f handled by thread f (f < n_far <= AR_THREADS * 2): global offset relative to the row being computed
  int foff[2];
#pragma unroll
  for (int q = 0; q < 2; q++) {
    const int f = tid + q * AR_THREADS;
    foff[q] = f < n_far ? (esi[n_near + f] + 1) * n_cs + csi[n_near + f] : -1;
  }

  const int i_top = 2 * n_ext - 1;
  // ---- prologue: ring <- rows i_top+1 .. i_top+3 (initial noise); far buffers and pipelines for the first steps ----
  for (int e = tid; e < 3 * n_cs; e += AR_THREADS) {
    const int R = i_top + 1 + e / n_cs, c = e % n_cs;
    hist[(R & 3) * n_cs + c] = buffer[(size_t)R * n_cs + c];
  }
#pragma unroll
  for (int q = 0; q < 2; q++)
    if (foff[q] >= 0) {
      const int f = tid + q * AR_THREADS;
      far[(i_top & 1) * n_far + f] = buffer[(size_t)i_top * n_cs + foff[q]];
      if (i_top >= 1) far[((i_top - 1) & 1) * n_far + f] = buffer[(size_t)(i_top - 1) * n_cs + foff[q]];
    }
  double far_p[2] = {0.0, 0.0};  // in flight: far values of step i - 2, loaded during step i
  double bz_p1[RW], bz_p2[RW];   // B z of the rows of step i - 1 / i - 2
#pragma unroll
  for (int rr = 0; rr < RW; rr++) {
    const int r = warp * RW + rr;
    bz_p1[rr] = bz_p2[rr] = 0.0;
    if (r < rows) {
      bz_p1[rr] = buffer[(size_t)i_top * n_cs + row0 + r];
      if (i_top >= 1) bz_p2[rr] = buffer[(size_t)(i_top - 1) * n_cs + row0 + r];
    }
  }
  __syncthreads();
  if (cl > 1) cluster_barrier();

  for (int i = i_top; i >= 0; i--) {
    // (1) issue the far / B z loads of step i - 2 (rows >= i + 2, complete since two barriers ago)
    double far_n[2] = {0.0, 0.0}, bz_n[RW];
    if (i >= 2) {
#pragma unroll
      for (int q = 0; q < 2; q++)
        if (foff[q] >= 0) far_n[q] = __ldcg(buffer + (size_t)(i - 2) * n_cs + foff[q]);
#pragma unroll
      for (int rr = 0; rr < RW; rr++) {
        const int r = warp * RW + rr;
        bz_n[rr] = r < rows ? __ldcg(buffer + (size_t)(i - 2) * n_cs + row0 + r) : 0.0;
      }
    } else {
#pragma unroll
      for (int rr = 0; rr < RW; rr++) bz_n[rr] = 0.0;
    }
    // (2) the dot products of this step, A from registers, x from shared memory
    double acc[RW];
#pragma unroll
    for (int rr = 0; rr < RW; rr++) acc[rr] = 0.0;
    const double* farp = far + (i & 1) * n_far;
#pragma unroll
    for (int m = 0; m < M; m++) {
      const int d = desc[m];
      double x = 0.0;
      if (d >= 0) {
        const int off = d >> 24;
        x = off ? hist[((i + off) & 3) * n_cs + (d & 0xffffff)] : farp[d];
      }
#pragma unroll
      for (int rr = 0; rr < RW; rr++) acc[rr] = fma(a[rr][m], x, acc[rr]);
    }
#pragma unroll
    for (int rr = 0; rr < RW; rr++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[rr] += __shfl_xor_sync(0xffffffffu, acc[rr], o);
    }
    // (3) publish: lane rr of the warp writes row warp*RW+rr to the ring of every CTA and to global memory
#pragma unroll
    for (int rr = 0; rr < RW; rr++) {
      const int r = warp * RW + rr;
      if (lane == rr && r < rows) {
        const double v = bz_p1[rr] + acc[rr];
        double* slot = hist + (i & 3) * n_cs + row0 + r;
        if (cl == 1) {
          *slot = v;
        } else {
          for (int c = 0; c < cl; c++) st_dsmem_f64(slot, c, v);
        }
        buffer[(size_t)i * n_cs + row0 + r] = v;
      }
    }
    // (4) land the far values loaded one step ago (they belong to step i - 1) and rotate the pipelines
    if (i >= 1) {
#pragma unroll
      for (int q = 0; q < 2; q++)
        if (foff[q] >= 0 && i <= i_top - 1) far[((i - 1) & 1) * n_far + tid + q * AR_THREADS] = far_p[q];
    }
#pragma unroll
    for (int q = 0; q < 2; q++) far_p[q] = far_n[q];
#pragma unroll
    for (int rr = 0; rr < RW; rr++) {
      bz_p1[rr] = bz_p2[rr];
      bz_p2[rr] = bz_n[rr];
    }
    if (cl == 1)
      __syncthreads();
    else
      cluster_barrier();
  }
}

template <int RW, int M>
static int launch_chain_v2(matm_handle* h, const double* A, const int* esi, const int* csi, double* buffer, int n_ext,
                           int n_cs, int n_sample, int n_near, int cl, cudaStream_t stream) {
  const size_t smem = (4 * (size_t)n_cs + 2 * (size_t)(n_sample - n_near)) * sizeof(double) + 16;
  MATM_CUDA(h, cudaFuncSetAttribute(ar_chain_v2_kernel<RW, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cl, 1, 1);
  cfg.blockDim = dim3(AR_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  MATM_CUDA(h, cudaLaunchKernelEx(&cfg, ar_chain_v2_kernel<RW, M>, A, esi, csi, buffer, n_ext, n_cs, n_sample, n_near, cl));
  h->launches++;
  return MATM_OK;
}

// picks (cluster size, rows per warp, samples per lane) for the register-resident kernel; false if it does not fit
static bool plan_v2(int n_cs, int n_sample, int* cl, int* rw, int* m) {
  const int warps = AR_THREADS / 32;
  int mm = (n_sample + 31) / 32;
  mm = mm <= 4 ? 4 : mm <= 8 ? 8 : mm <= 12 ? 12 : mm <= 16 ? 16 : 0;
  if (!mm) return false;
  const int max_rw = mm <= 8 ? 4 : 2;  // RW * M <= 32 doubles of A per thread (no spills at 512 threads)
  for (int c = 1; c <= 8; c *= 2) {
    const int rpc = (n_cs + c - 1) / c;
    int r = (rpc + warps - 1) / warps;
    r = r <= 1 ? 1 : r <= 2 ? 2 : r <= 4 ? 4 : 0;
    if (r && r <= max_rw) {
      *cl = c;
      *rw = r;
      *m = mm;
      return true;
    }
  }
  return false;
}

extern "C" int matm_ar_generate(matm_handle* h, int n_ext, int n_cs, int n_sample, int n_near, const double* A_dev,
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

  // register-resident kernel when the process fits it
  {
    int cl2, rw, mm;
    if (n_near >= 0 && n_near <= n_sample && n_ext >= 4 && (n_sample - n_near) <= 2 * AR_THREADS && plan_v2(n_cs, n_sample, &cl2, &rw, &mm)) {
#define V2_CASE(RWV, MV) \
  if (rw == RWV && mm == MV) \
    return launch_chain_v2<RWV, MV>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, n_near, cl2, st);
      V2_CASE(1, 4) V2_CASE(2, 4) V2_CASE(4, 4) V2_CASE(1, 8) V2_CASE(2, 8) V2_CASE(4, 8)
      V2_CASE(1, 12) V2_CASE(2, 12) V2_CASE(1, 16) V2_CASE(2, 16)
#undef V2_CASE
    }
  }
  // fallback (v1): smallest cluster whose per-CTA slice of A fits in shared memory; above 8 CTAs A streams from L2
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
