// a5: AutoregressiveProcess.run (maria/atmosphere/process.py:191-209) on the device.
//
//   BUFFER[i] = A @ BUFFER[i + 1 + esi, csi] + B @ z_i      for i = 2 n_ext - 1 ... 0
//
// Three kernels.  (1) bz_gemm_kernel: the only genuinely dense contraction, B @ z for ALL steps at once
// (M = 2 n_ext, N = K = n_cs), written straight into BUFFER rows [0, 2 n_ext).  (2) ar_chain_v4_kernel: the sequential
// recurrence of every process that fits it (all "2d" processes: n_sample <= 1280, n_cs <= 8 x 128): one thread-block
// cluster per process, A in REGISTERS (L lanes per row), the gather vectors in shared memory, every produced value sent
// with st.async into the gather vectors of the next three steps of every CTA of the cluster, completing bytes on mbarriers
// of the destination CTA -- no cluster barrier, no fence on the critical path.  (3) ar_chain_kernel: the general
// recurrence for larger processes (the coupled "3d" model: n_cs and n_sample in the thousands): every CTA of the cluster
// keeps its slice of rows of A in shared memory when it fits (staged once with the TMA bulk-copy engine), else streams
// it from L2; the gather vector is rebuilt from the L2-resident BUFFER each step and the CTAs meet at a hardware cluster
// barrier.  Both chain kernels are covered by parity tests (tests/test_gpu_parity.py, tests/test_gpu_configs.py).
#include <cooperative_groups.h>
#include <type_traits>

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

__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}

// ---------------------------------------------------------------------------------------------------------
// The chain, v4.  A step is bounded by the round trips on its critical path, not by arithmetic, so everything that is not
// "the previous row" is taken off that path and the rest is made short:
//   * A lives in REGISTERS.  L lanes share a row (32 / L rows per warp): lane `sub` holds the entries of samples
//     sub, sub + L, ...  ->  the cross-lane reduction is log2(L) shuffle stages instead of 5 per row;
//   * the gather vector of step i is a contiguous shared-memory array G[i & 3][n_sample]: whoever produces a value
//     writes it straight to where its consumers will read it; the stencil rows further back (offsets >= 4) plus the row's
//     own B z term are fetched from L2 ahead of their use through a register pipeline, so their latency never shows;
//   * no barrier: data-carrying signals.
// Every value travels with its own completion: the owner of (row i, column c) sends it with st.async into the gather
// vectors of steps i-1, i-2, i-3 of every CTA, each store completing bytes on an mbarrier of the DESTINATION CTA.  A
// gather vector has two mbarriers: EARLY (stencil offsets 2, 3 and the far samples: everything known one step ahead)
// and LATE (offset 1: the row just produced).  A step therefore waits only for the bytes it needs, there is no fence
// and no barrier round trip on the critical path, and the part of the dot product that uses the EARLY entries of the
// NEXT step (more than half of a row of A) is computed while the LATE entries are still in flight.
// Buffer reuse needs no extra synchronisation: a value of step i can only be sent after the LATE barrier of step i
// completed, i.e. after every publishing warp of every CTA has finished reading the vectors of steps > i.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mapa_u32(uint32_t a, int cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(cta));
  return r;
}
__device__ __forceinline__ void st_async_f64(uint32_t remote_addr, double v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(remote_addr), "d"(v),
               "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arm(uint32_t bar, uint32_t bytes) {
  if (bytes == 0)
    mbar_arrive(bar);
  else
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// wait for the phase of the given parity; data written by st.async of other CTAs is visible afterwards.  A wait that
// never ends is a bug in the byte accounting: trap instead of hanging the device.
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  int spins = 0;
  while (!done) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"  // transaction barrier: the bytes it counted are visible
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done && ++spins > (1 << 22)) __trap();
  }
}

// MC >= 0: the LATE part is registers [0, 4 MC), fixed at compile time; CL > 0: the cluster size, fixed at compile time
template <int L, int MP, int MC = -1, int CL = 0>
__global__ void __launch_bounds__(AR_THREADS, 1)
    ar_chain_v4_kernel(const double* __restrict__ A, const int* __restrict__ esi, const int* __restrict__ csi,
                       double* buffer, int n_ext, int n_cs, int n_sample, int n_near, int cl) {
  constexpr int RPW = 32 / L;  // rows per warp
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int gs = L * MP;                                     // padded length of one gather vector
  double* G = reinterpret_cast<double*>(smem_raw);           // [4][gs]
  uint64_t* bars = reinterpret_cast<uint64_t*>(G + 4 * (size_t)gs);  // [4] LATE, [4] EARLY
  int* cnt = reinterpret_cast<int*>(bars + 8);               // [4]: near samples of offset 1, 2, 3
  int* pos = cnt + 4;                                        // [3][n_cs]: position of column c among the samples of offset 1, 2, 3
  const int n_far = n_sample - n_near;
  int cta = 0;
  if (cl > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta));
  const int rpc = (n_cs + cl - 1) / cl;
  const int row0 = cta * rpc;
  const int rows = max(0, min(n_cs, row0 + rpc) - row0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rr = lane / L, sub = lane % L;
  const int r = warp * RPW + rr;  // row within this CTA
  const bool has_row = r < rows;
  const int i_top = 2 * n_ext - 1;
  const uint32_t G_s = smem_u32(G), bars_s = smem_u32(bars);

  // ---- set-up: A into registers, G zeroed, inverse stencil maps, sample counts per offset ----
  double a[MP];
#pragma unroll
  for (int m = 0; m < MP; m++) {
    const int s = sub + L * m;
    a[m] = (has_row && s < n_sample) ? A[(size_t)(row0 + r) * n_sample + s] : 0.0;
  }
  for (int e = tid; e < 4 * gs; e += AR_THREADS) G[e] = 0.0;
  for (int e = tid; e < 3 * n_cs; e += AR_THREADS) pos[e] = -1;
  if (tid < 4) cnt[tid] = 0;
  if (tid == 0) {
    for (int b = 0; b < 8; b++) mbar_init(bars + b, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  for (int s = tid; s < n_near; s += AR_THREADS) {
    pos[esi[s] * n_cs + csi[s]] = s;  // offset = esi + 1 in 1..3
    atomicAdd(&cnt[esi[s]], 1);
  }
  int foff[2];
#pragma unroll
  for (int q = 0; q < 2; q++) {
    const int f = tid + q * AR_THREADS;
    foff[q] = f < n_far ? (esi[n_near + f] + 1) * n_cs + csi[n_near + f] : -1;
  }
  __syncthreads();
  const int n1 = cnt[0], n2 = cnt[1], n3 = cnt[2];
  // registers [0, mcrit) hold every offset-1 sample (the samples are sorted by offset): the part of the dot product that
  // has to wait for the row just produced; the rest is computed one step ahead
  // (with MC given by the host from n_cs >= n1, the unrolled groups carry no branch and their loads are batched)
  const int mcrit = MC >= 0 ? min(MP, 4 * MC) : min(MP, 4 * ((n1 + 4 * L - 1) / (4 * L)));
  int p1 = -1, p2 = -1, p3 = -1;
  if (has_row && sub == 0) {
    p1 = pos[row0 + r];
    p2 = pos[n_cs + row0 + r];
    p3 = pos[2 * n_cs + row0 + r];
  }
  // With the cluster size known at compile time the 3 CL destinations of a value are spread over the L lanes of its row
  // (after the butterfly every lane holds the sum): lane `sub` sends to (CTA t / 3, gather vector of step i - 1 - t % 3)
  // for t = sub and t = sub + L.
  constexpr bool SPREAD = CL > 0 && 3 * CL <= 2 * L;
  uint32_t pubd[2] = {0, 0}, pubb[2] = {0, 0};
  int pubk[2] = {0, 0};
  if (SPREAD) {
#pragma unroll
    for (int j = 0; j < 2; j++) {
      const int t = sub + j * L, c = min(t / 3, cl - 1), k = t % 3 + 1;
      const int pk = (has_row && t < 3 * cl) ? pos[(k - 1) * n_cs + row0 + r] : -1;
      pubk[j] = pk >= 0 ? k : 0;
      pubd[j] = mapa_u32(G_s, c) + 8u * max(pk, 0);
      pubb[j] = mapa_u32(bars_s, c) + (k == 1 ? 0u : 32u);
    }
  }
  // ---- prologue: the initial-noise rows i_top+1 .. i_top+3 feed steps i_top, i_top-1, i_top-2 ----
  for (int e = tid; e < 3 * n_near; e += AR_THREADS) {
    const int k = e / n_near, s = e % n_near;  // step i_top - k, near sample s
    const int off = esi[s] + 1;
    if (off > k && i_top - k >= 0)
      G[((i_top - k) & 3) * gs + s] = buffer[(size_t)(i_top - k + off) * n_cs + csi[s]];
  }
#pragma unroll
  for (int q = 0; q < 2; q++)
    if (foff[q] >= 0) {
      const int f = n_near + tid + q * AR_THREADS;
      G[(i_top & 3) * gs + f] = buffer[(size_t)i_top * n_cs + foff[q]];
      if (i_top >= 1) G[((i_top - 1) & 3) * gs + f] = buffer[(size_t)(i_top - 1) * n_cs + foff[q]];
    }
  // Register pipelines, two sets each, used alternately by even / odd steps (the loop is unrolled by two):
  //   far_s[i & 1]: issued at step i + 4 for step i - 2, sent into G at step i, then re-issued for step i - 4
  //   bz_s[i & 1] : issued at step i + 2 for step i, consumed when step i publishes, then re-issued for step i - 2
  double far_s[2][2], bz_s[2];
  far_s[0][0] = far_s[0][1] = far_s[1][0] = far_s[1][1] = 0.0;
  bz_s[0] = bz_s[1] = 0.0;
  if (has_row) {  // i_top is odd: set 1 serves step i_top, set 0 step i_top - 1
    bz_s[1] = buffer[(size_t)i_top * n_cs + row0 + r];
    bz_s[0] = buffer[(size_t)(i_top - 1) * n_cs + row0 + r];
  }
#pragma unroll
  for (int q = 0; q < 2; q++)
    if (foff[q] >= 0) {
      if (i_top >= 2) far_s[1][q] = buffer[(size_t)(i_top - 2) * n_cs + foff[q]];
      if (i_top >= 3) far_s[0][q] = buffer[(size_t)(i_top - 3) * n_cs + foff[q]];
    }
  // bytes each barrier expects for step j = i_top - d: what the loop (not the prologue) delivers
  const uint32_t late_full = 8u * n1, early_full = 8u * (n2 + n3 + n_far);
  if (tid == 0) {
    for (int d = 0; d < 4 && d <= i_top; d++) {
      const int b = (i_top - d) & 3;
      mbar_arm(bars_s + 8 * b, d >= 1 ? late_full : 0u);
      mbar_arm(bars_s + 8 * (4 + b), (d >= 2 ? 8u * (n2 + n_far) : 0u) + (d >= 3 ? 8u * n3 : 0u));
    }
    // nobody waits for the EARLY barrier of step i_top (its part is computed right below): arm its next use here
    if (i_top >= 4) mbar_arm(bars_s + 8 * (4 + (i_top & 3)), early_full);
  }
  __syncthreads();
  cluster_barrier();

  // the EARLY part of the dot product of step j: registers [mcrit, MP)
  auto early_part = [&](const int j) -> double {
    const double* g = G + (j & 3) * gs + sub;
    double e0 = 0.0, e1 = 0.0, e2 = 0.0, e3 = 0.0;
#pragma unroll
    for (int m = 0; m < MP; m += 4)
      if (MC >= 0 ? (m >= 4 * MC) : (m >= mcrit)) {
        e0 = fma(a[m], g[L * m], e0);
        if (m + 1 < MP) e1 = fma(a[m + 1], g[L * (m + 1)], e1);
        if (m + 2 < MP) e2 = fma(a[m + 2], g[L * (m + 2)], e2);
        if (m + 3 < MP) e3 = fma(a[m + 3], g[L * (m + 3)], e3);
      }
    return (e0 + e1) + (e2 + e3);
  };
  // warps with neither a row nor a far sample only wait for the end
  const bool active = (warp * RPW < rows) || (warp * 32 < n_far);
  double pre = 0.0;
  if (active) {
    pre = early_part(i_top);

    // GD = false: the steady state (i >= 5), where every "does that step exist" test below is true
    auto step = [&](auto guard, const int i, double(&fs)[2], double& bz) {
      constexpr bool GD = decltype(guard)::value;
      const int ncl = CL > 0 ? CL : cl;
      const uint32_t par = ((i_top - i) >> 2) & 1;
      // (1) the row just produced has landed in G[i]
      mbar_wait_cluster(bars_s + 8 * (i & 3), par);
      if (tid == 0 && (!GD || i >= 4)) mbar_arm(bars_s + 8 * (i & 3), late_full);
      // (2) the LATE part of the dot product
      const double* g = G + (i & 3) * gs + sub;
      double acc0 = pre, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
      for (int m = 0; m < MP; m += 4)
        if (MC >= 0 ? (m < 4 * MC) : (m < mcrit)) {
          acc0 = fma(a[m], g[L * m], acc0);
          if (m + 1 < MP) acc1 = fma(a[m + 1], g[L * (m + 1)], acc1);
          if (m + 2 < MP) acc2 = fma(a[m + 2], g[L * (m + 2)], acc2);
          if (m + 3 < MP) acc3 = fma(a[m + 3], g[L * (m + 3)], acc3);
        }
      double acc = (acc0 + acc1) + (acc2 + acc3);
#pragma unroll
      for (int o = L / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      // (3) publish row r of step i into the gather vectors of steps i-1, i-2, i-3 of every CTA, and to global memory
      if (SPREAD) {
        const double v = bz + acc;
        if (has_row && sub == 0) buffer[(size_t)i * n_cs + row0 + r] = v;  // read again (far samples) three steps from now at the earliest
#pragma unroll
        for (int j = 0; j < 2; j++) {
          if (j == 1 && 3 * CL <= L) break;
          const int bsel = (i - pubk[j]) & 3;
          if (pubk[j] != 0 && (!GD || i >= pubk[j])) st_async_f64(pubd[j] + 8u * (bsel * gs), v, pubb[j] + 8u * bsel);
        }
      } else
      if (has_row && sub == 0) {
        const double v = bz + acc;
        const uint32_t d1 = G_s + 8u * (((i - 1) & 3) * gs + p1), b1 = bars_s + 8u * ((i - 1) & 3);
        const uint32_t d2 = G_s + 8u * (((i - 2) & 3) * gs + p2), b2 = bars_s + 8u * (4 + ((i - 2) & 3));
        const uint32_t d3 = G_s + 8u * (((i - 3) & 3) * gs + p3), b3 = bars_s + 8u * (4 + ((i - 3) & 3));
        buffer[(size_t)i * n_cs + row0 + r] = v;  // read again (far samples) no sooner than three steps from now
#pragma unroll
        for (int c = 0; c < (CL > 0 ? CL : 8); c++) {
          if (c >= ncl) break;
          if (p1 >= 0 && (!GD || i >= 1)) st_async_f64(mapa_u32(d1, c), v, mapa_u32(b1, c));
          if (p2 >= 0 && (!GD || i >= 2)) st_async_f64(mapa_u32(d2, c), v, mapa_u32(b2, c));
          if (p3 >= 0 && (!GD || i >= 3)) st_async_f64(mapa_u32(d3, c), v, mapa_u32(b3, c));
        }
      }
      // (4) off the critical path: send the far values of step i - 2 (in flight since step i + 4), re-issue the set for
      //     step i - 4 (rows >= i + 1, complete: their LATE barrier was passed above), re-issue B z for step i - 2
      if (!GD || i >= 2) {
        const uint32_t be = mapa_u32(bars_s + 8u * (4 + ((i - 2) & 3)), cta);
#pragma unroll
        for (int q = 0; q < 2; q++)
          if (foff[q] >= 0)
            st_async_f64(mapa_u32(G_s + 8u * (((i - 2) & 3) * gs + n_near + tid + q * AR_THREADS), cta), fs[q], be);
      }
      if (!GD || i >= 4) {
#pragma unroll
        for (int q = 0; q < 2; q++)
          if (foff[q] >= 0) fs[q] = __ldcg(buffer + (size_t)(i - 4) * n_cs + foff[q]);
      }
      if ((!GD || i >= 2) && has_row) bz = __ldcg(buffer + (size_t)(i - 2) * n_cs + row0 + r);
      // (5) the EARLY part of step i - 1, while the row of step i is on its way
      if (!GD || i >= 1) {
        const uint32_t pe = ((i_top - (i - 1)) >> 2) & 1;
        mbar_wait_cluster(bars_s + 8 * (4 + ((i - 1) & 3)), pe);
        if (tid == 0 && (!GD || i - 1 >= 4)) mbar_arm(bars_s + 8 * (4 + ((i - 1) & 3)), early_full);
        pre = early_part(i - 1);
      }
    };

    int i = i_top;  // 2 n_ext - 1: odd
    for (; i >= 7; i -= 2) {
      step(std::false_type{}, i, far_s[1], bz_s[1]);
      step(std::false_type{}, i - 1, far_s[0], bz_s[0]);
    }
    for (; i >= 1; i -= 2) {
      step(std::true_type{}, i, far_s[1], bz_s[1]);
      step(std::true_type{}, i - 1, far_s[0], bz_s[0]);
    }
  }
  // no CTA may leave while values are still on their way to it
  cluster_barrier();
}

template <int L, int MP>
static int launch_chain_v4(matm_handle* h, const double* A, const int* esi, const int* csi, double* buffer, int n_ext,
                           int n_cs, int n_sample, int n_near, int cl, cudaStream_t stream) {
  const size_t smem = 4 * (size_t)L * MP * sizeof(double) + 3 * (size_t)n_cs * sizeof(int) + 16 + 8 * sizeof(uint64_t) + 16;
  auto kernel = ar_chain_v4_kernel<L, MP>;
  if constexpr (L == 8 && MP >= 32) {
    // the long chains (n_cs ~ 100-160): the split between the LATE and EARLY registers as a compile-time constant.
    // n_cs bounds the number of offset-1 samples from above, so registers [0, 4 mc) always hold all of them.
    const int mc = (n_cs + 4 * L - 1) / (4 * L);
    if (mc == 3) kernel = ar_chain_v4_kernel<L, MP, 3>;
    if (mc == 4) kernel = cl == 2 ? ar_chain_v4_kernel<L, MP, 4, 2> : ar_chain_v4_kernel<L, MP, 4>;
    if (mc == 5) kernel = cl == 3 ? ar_chain_v4_kernel<L, MP, 5, 3> : ar_chain_v4_kernel<L, MP, 5>;
  }
  MATM_CUDA(h, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
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
  MATM_CUDA(h, cudaLaunchKernelEx(&cfg, kernel, A, esi, csi, buffer, n_ext, n_cs, n_sample, n_near, cl));
  h->launches++;
  return MATM_OK;
}

// picks (lanes per row, samples per lane, cluster size) for the register-resident kernel; false if it does not fit
static bool plan_v4(int n_cs, int n_sample, int* L, int* MP, int* cl) {
  static const int Ls[4] = {4, 8, 16, 32};
  static const int MPs[4] = {16, 24, 32, 40};
  for (int li = 0; li < 4; li++) {
    if (n_sample > Ls[li] * 40) continue;
    int mp = 0;
    for (int mi = 0; mi < 4; mi++)
      if (n_sample <= Ls[li] * MPs[mi]) {
        mp = MPs[mi];
        break;
      }
    const int rows_per_cta = (AR_THREADS / 32) * (32 / Ls[li]);
    const int c = (n_cs + rows_per_cta - 1) / rows_per_cta;
    if (c > 8) continue;  // a larger L only lowers rows_per_cta further
    *L = Ls[li];
    *MP = mp;
    *cl = c;
    return true;
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
    int L, mp, cl3;
    if (n_near >= 0 && n_near <= n_sample && n_ext >= 4 && (n_sample - n_near) <= 2 * AR_THREADS &&
        plan_v4(n_cs, n_sample, &L, &mp, &cl3)) {
#define V4_CASE(LV, MV) \
  if (L == LV && mp == MV) \
    return launch_chain_v4<LV, MV>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, n_near, cl3, st);
      V4_CASE(4, 16) V4_CASE(4, 24) V4_CASE(4, 32) V4_CASE(4, 40) V4_CASE(8, 16) V4_CASE(8, 24) V4_CASE(8, 32)
      V4_CASE(8, 40) V4_CASE(16, 24) V4_CASE(16, 32) V4_CASE(16, 40) V4_CASE(32, 24) V4_CASE(32, 32) V4_CASE(32, 40)
#undef V4_CASE
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
