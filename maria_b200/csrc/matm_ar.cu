// a5: AutoregressiveProcess.run (maria/atmosphere/process.py:191-209) on the device.
//
//   BUFFER[i] = A @ BUFFER[i + 1 + esi, csi] + B @ z_i      for i = 2 n_ext - 1 ... 0
//
// Three kernels.  (1) bz_gemm_kernel: the only genuinely dense contraction, B @ z for ALL steps at once
// (M = 2 n_ext, N = K = n_cs), written straight into BUFFER rows [0, 2 n_ext).  (2) ar_chain_v4_kernel: the sequential
// recurrence of every process that fits it (all "2d" processes: n_sample <= 1280, at most 128 far samples, rows that fit 8
// CTAs): one thread-block cluster per process, A in REGISTERS, the gather vectors in shared memory, every produced value
// sent with st.async into the gather vectors of the next three steps of every CTA of the cluster, completing bytes on an
// mbarrier of the destination CTA -- no cluster barrier, no fence on the critical path.  (3) ar_chain_kernel: the general
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
// The chain, v4.  A step is bounded by what lies between "the previous row has landed" and "this row has left", not by
// arithmetic, so everything that is not the previous row is taken off that path and the rest is made short:
//   * A lives in REGISTERS.  L lanes share a row (32 / L rows per warp).  At L = 8 (every process wider than 32 columns)
//     the layout is register-blocked: a lane holds all four rows of its warp for its own samples, so the warp reads a
//     gather vector once, with 16-byte loads, and the four partial rows are reduced by a reduce-scatter + butterfly;
//   * the gather vector of step i is a contiguous shared-memory array G[i & 3][n_sample]: whoever produces a value
//     writes it straight to where its consumers will read it;
//   * no cluster barrier, no fence on the critical path: data-carrying signals.  The owner of (row i, column c) sends the
//     value with st.async into the gather vectors of steps i-1, i-2, i-3 of every CTA; every store completes bytes on ONE
//     mbarrier of the destination CTA, the one step i-1 waits for.  The part of the dot product that uses the entries
//     known one step ahead (offsets 2, 3 and the far samples: more than half of a row of A) is computed while the row just
//     sent is still in flight;
//   * the far samples (offsets >= 4) are the business of the CTA's last warp, which holds no rows: per sample a delay line
//     in shared memory fed from the row that every CTA receives anyway (offsets < 600), BUFFER in global memory behind a
//     periodic fence for the rest.  Nothing is ever read back from global memory sooner than 566 steps after it was stored,
//     and what a step needs from global memory (those samples, the row's B z term) is prefetched 32 steps ahead with
//     cp.async by the same warp: the chain keeps its pace underneath kernels that saturate HBM;
//   * 256 threads per CTA (7 row warps + the far warp: two warps per scheduler, no register cap), the step loop unrolled by
//     four so that the phase (i & 3) of every shared-memory address is an immediate.
// Buffer reuse needs no extra synchronisation: a value of step i can only be sent after the barrier of step i completed,
// i.e. after every warp of every CTA has finished reading the vectors of steps > i.
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
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) {
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void cp_async_f64(uint32_t dst_smem, const double* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst_smem), "l"(src) : "memory");
}
// the same under a predicate inside the instruction stream: no branch, no reconvergence point around the store
__device__ __forceinline__ void st_async_f64_if(bool p, uint32_t remote_addr, double v, uint32_t remote_bar) {
  asm volatile(
      "{\n"
      ".reg .pred q;\n"
      "setp.ne.u32 q, %3, 0;\n"
      "@q st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];\n"
      "}\n" ::"r"(remote_addr),
      "d"(v), "r"(remote_bar), "r"((uint32_t)p)
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
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"  // transaction barrier: the bytes it counted are visible
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done;
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  // try_wait suspends the warp in hardware until the phase completes or a time limit passes: the first attempt almost always
  // succeeds; the counting loop is the rare path
  if (mbar_try_wait(bar, parity)) return;
  int spins = 0;
  while (!mbar_try_wait(bar, parity))
    if (++spins > (1 << 22)) __trap();
}

#ifdef MATM_AR_TRACE
// Diagnostic build only (-DMATM_AR_TRACE, scripts/chain_trace.py): SM clocks at the phase boundaries of 128 steps of the
// steady state, for warps 0 and 3 of every CTA.
static __device__ long long g_ar_trace[16][128][6];
__device__ __forceinline__ long long ar_clock_after(double x) {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) : "d"(x) : "memory");
  return t;
}
#define AR_TRACE(k, dep)                                                   \
  if (tr_on && i < tr_hi && i >= tr_hi - 128) g_ar_trace[tr_slot][i & 127][k] = ar_clock_after(dep)
extern "C" int matm_debug_ar_trace(long long* out_host) {
  return cudaMemcpyFromSymbol(out_host, g_ar_trace, sizeof(g_ar_trace)) == cudaSuccess ? 0 : 1;
}
#else
#define AR_TRACE(k, dep)
#endif

// the phase (step & 3) of a step: a compile-time constant in the unrolled steady state, a run-time value elsewhere
template <int V>
struct PhaseConst {};
template <int V>
__device__ __forceinline__ constexpr int phase_value(PhaseConst<V>) { return V; }
__device__ __forceinline__ constexpr int phase_value(int v) { return v; }

// MC >= 0: the LATE part is registers [0, 4 MC), fixed at compile time; CL > 0: the cluster size, fixed at compile time
// NT: threads per CTA.  512 = 128 registers per thread and up to 16 warps of rows per SM (the smallest cluster); 256 = 255
// registers, 8 warps (two per scheduler): fewer rows per SM, more CTAs per cluster, a shorter step.
// The LAST warp of the CTA owns the far samples and no rows, so that no warp on the critical path of a row carries them
// (profiles/r02t_chain_step_phases.txt: the warps that also served far samples were the ones everybody waited for); the
// other NT / 32 - 1 warps hold (NT / 32 - 1) * (32 / L) rows.  (A ninth warp next to eight row warps does not pay: registers
// are allocated as if the block had a multiple of four warps, 288 threads would be budgeted as 384.)
#define AR_FAR_Q 4  // far samples per lane of that warp: n_far <= 128
#define AR_PD 32       // look-ahead of the global prefetch [steps]
#define AR_PS 64       // slots of the prefetch rings (a slot is reused AR_PS steps later)
#define AR_PW 24       // cp.async groups allowed in flight: what was issued more than AR_PW steps ago has landed
#define AR_BZ_ROWS 64  // rows per CTA the B z ring holds (7 row warps x 8 rows at L = 4)
template <int L, int MP, int MC = -1, int CL = 0, int NT = AR_THREADS>
__global__ void __launch_bounds__(NT, 1)
    ar_chain_v4_kernel(const double* __restrict__ A, const int* __restrict__ esi, const int* __restrict__ csi,
                       double* buffer, int n_ext, int n_cs, int n_sample, int n_near, int cl, int n_ring, int fence_mask,
                       int ring_doubles) {
  constexpr int RPW = 32 / L;  // rows per warp
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int gs = L * MP;                                     // padded length of one gather vector
  double* G = reinterpret_cast<double*>(smem_raw);           // [4][gs]
  uint64_t* bars = reinterpret_cast<uint64_t*>(G + 4 * (size_t)gs);  // [4]: one transaction barrier per gather vector (slots 4..7 unused)
  int* cnt = reinterpret_cast<int*>(bars + 8);               // [4]: near samples of offset 1, 2, 3
  int* pos = cnt + 4;                                        // [3][n_cs]: position of column c among the samples of offset 1, 2, 3
  const int n_far = n_sample - n_near;
  int* roff = pos + 3 * n_cs;                                // [n_ring + 1]: start of the delay line of far sample f
  double* ring = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(roff + n_ring + 1) + 15) & ~(uintptr_t)15);
  // Everything a step needs from BUFFER in global memory -- the row's own B z term and the far samples beyond the delay lines
  // -- is fetched AR_PD steps ahead by the far warp with cp.async into two rings of AR_PS slots, so that no warp ever waits
  // for a global load: under a synthesis kernel that saturates HBM those loads take microseconds, and with a look-ahead of
  // two to four steps the chain ran at half speed (1.08 instead of 0.54 us/step, profiles/r02_chain_under_load.txt).
  double* bzr = ring + ring_doubles;                        // [AR_PS][AR_BZ_ROWS]: B z of step j for the rows of this CTA
  double* fstage = bzr + AR_PS * AR_BZ_ROWS;                // [AR_PS][n_glob]: far samples of the vector of step j
  const int n_glob = n_far - n_ring;
  int cta = 0;
  if (cl > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta));
  const int rpc = (n_cs + cl - 1) / cl;
  const int row0 = cta * rpc;
  const int rows = max(0, min(n_cs, row0 + rpc) - row0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NTT = NT;                            // all threads of the CTA (set-up loops)
  const bool far_warp = warp == NT / 32 - 1;         // the last warp: far samples only, no rows
  const int rr = lane / L, sub = lane % L;
  const int r = warp * RPW + rr;  // row within this CTA
  const bool has_row = r < rows && !far_warp;
  const int i_top = 2 * n_ext - 1;
  const uint32_t G_s = smem_u32(G), bars_s = smem_u32(bars);

  // ---- set-up: A into registers, G zeroed, inverse stencil maps, sample counts per offset ----
  // lane `sub` holds the sample PAIRS 2 (sub + L q), 2 (sub + L q) + 1 in registers 2 q, 2 q + 1: one 16-byte shared-memory
  // load feeds two FMAs, and the L lanes of a row read 16 L contiguous bytes (a full 128-byte wavefront at L = 8)
  static_assert(MP % 4 == 0, "registers are consumed in groups of four (two 16-byte loads)");
  // BLK (L = 8): register blocking.  The warp's RPW = 4 rows share every gather value, so a lane holds ALL four rows for its
  // own SPL = MP / 4 samples (pairs 2 (lane + 32 q), + 1): one 16-byte load feeds eight FMAs and the warp reads the gather
  // vector once instead of once per row -- the shared-memory pipe, not the FP64 pipe, was what the dot products waited for
  // (profiles/r02t_chain_step_phases.txt).  The four partial rows are then reduced over the 32 lanes by a reduce-scatter
  // (two exchanges that halve the rows a lane carries) followed by a three-stage butterfly, which leaves row lane / 8 in
  // lane, i.e. the same (rr, sub) roles as the unblocked layout.
  constexpr bool BLK = (L == 8);
  constexpr int SPL = MP / RPW, NQ = SPL / 2;
  static_assert(!BLK || SPL % 2 == 0, "blocked layout: whole sample pairs per lane");
  double a[MP];
#pragma unroll
  for (int m = 0; m < MP; m++) {
    if (BLK) {
      const int k = m / SPL, j = m % SPL, s = 2 * (lane + 32 * (j / 2)) + (j & 1), rk = warp * RPW + k;
      a[m] = (rk < rows && s < n_sample) ? A[(size_t)(row0 + rk) * n_sample + s] : 0.0;
    } else {
      const int s = 2 * (sub + L * (m / 2)) + (m & 1);
      a[m] = (has_row && s < n_sample) ? A[(size_t)(row0 + r) * n_sample + s] : 0.0;
    }
  }
  for (int e = tid; e < 4 * gs; e += NTT) G[e] = 0.0;
  for (int e = tid; e < 3 * n_cs; e += NTT) pos[e] = -1;
  if (tid < 4) cnt[tid] = 0;
  if (tid == 0) {
    for (int b = 0; b < 4; b++) mbar_init(bars + b, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  for (int s = tid; s < n_near; s += NTT) {
    pos[esi[s] * n_cs + csi[s]] = s;  // offset = esi + 1 in 1..3
    atomicAdd(&cnt[esi[s]], 1);
  }
  int foff[AR_FAR_Q];  // far sample lane + 32 q of the far warp: element offset of (row + d, column) relative to the row
#pragma unroll
  for (int q = 0; q < AR_FAR_Q; q++) {
    const int f = lane + 32 * q;
    foff[q] = (far_warp && f < n_far) ? (esi[n_near + f] + 1) * n_cs + csi[n_near + f] : -1;
  }
  __syncthreads();
  const int n1 = cnt[0], n2 = cnt[1], n3 = cnt[2];
  // ---- delay lines of the far samples ------------------------------------------------------------------------------
  // A far sample (offset d >= 4, column c) of step s is the value (s + d, c).  Reading it back from BUFFER in global memory
  // is only safe once the producer's store is known to be visible; a store followed by st.async carries no such guarantee
  // (measured: with 0.5 us steps and a synthesis kernel loading the memory system, the offset-5 samples, read one step after
  // they were produced, came back stale).  So the first n_ring far samples (the host checks that they include every offset
  // below 600) never touch global memory: every CTA receives the whole row i + 1 in the LATE part of its gather vector of
  // step i, the thread that owns far sample f copies (i + 1, c_f) from there into a private ring of d_f - 3 entries and
  // takes out (i - 2 + d_f, c_f), stored d_f - 3 steps earlier, which it sends into the vector of step i - 2.  The remaining
  // samples (offset >= 600, i.e. read >= 566 steps after they were produced) keep the global path, ordered by a
  // __threadfence() of every storing thread every (fence_mask + 1) = 512 steps: the store of row j is followed by a fence at
  // most 511 steps later, the fence by the thread's next st.async, and the reader has waited on a barrier those bytes completed.
  if (tid == 0) {
    int used = 0;
    for (int f = 0; f < n_ring; f++) {
      roff[f] = used;
      used += esi[n_near + f] + 1 - 3;
    }
    roff[n_ring] = used;
  }
  __syncthreads();
  for (int f = 0; f < n_ring; f++) {
    const int lr = esi[n_near + f] + 1 - 3, c = csi[n_near + f], ro = roff[f];
    for (int k = tid; k < lr; k += NTT) {  // rows i_top + 2 ... i_top - 2 + d: the initial noise, written by an earlier kernel
      const int row = i_top + 2 + k;
      ring[ro + row % lr] = buffer[(size_t)row * n_cs + c];
    }
  }
  uint32_t ring_p[AR_FAR_Q], ring_lo[AR_FAR_Q], ring_hi[AR_FAR_Q], ring_src[AR_FAR_Q];
#pragma unroll
  for (int q = 0; q < AR_FAR_Q; q++) {
    const int f = lane + 32 * q;
    ring_p[q] = ring_lo[q] = ring_hi[q] = ring_src[q] = 0u;
    if (far_warp && f < n_ring) {
      const int d = esi[n_near + f] + 1, lr = d - 3;
      ring_lo[q] = smem_u32(ring + roff[f]);
      ring_hi[q] = ring_lo[q] + 8u * (lr - 1);
      ring_p[q] = ring_lo[q] + 8u * ((i_top - 2 + d) % lr);
      ring_src[q] = 8u * pos[csi[n_near + f]];  // where the row just completed holds column c_f (offset-1 sample of c_f)
    }
  }
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
  // for t = sub, sub + L, ... < 3 CL.
  constexpr int NJ = CL > 0 ? (3 * CL + L - 1) / L : 1;  // destinations per lane
  constexpr bool SPREAD = CL > 0 && NJ <= 3;
  uint32_t pubd[NJ], pubb[NJ];
  int pubk[NJ];
#pragma unroll
  for (int j = 0; j < NJ; j++) pubd[j] = pubb[j] = 0u, pubk[j] = 0;
  if (SPREAD) {
#pragma unroll
    for (int j = 0; j < NJ; j++) {
      const int t = sub + j * L, c = min(t / 3, cl - 1), k = t % 3 + 1;
      const int pk = (has_row && t < 3 * cl) ? pos[(k - 1) * n_cs + row0 + r] : -1;
      pubk[j] = pk >= 0 ? k : 0;
      pubd[j] = mapa_u32(G_s, c) + 8u * max(pk, 0);
      pubb[j] = mapa_u32(bars_s, c);
    }
  }
  // ---- prologue: the initial-noise rows i_top+1 .. i_top+3 feed steps i_top, i_top-1, i_top-2 ----
  for (int e = tid; e < 3 * n_near; e += NTT) {
    const int k = e / n_near, s = e % n_near;  // step i_top - k, near sample s
    const int off = esi[s] + 1;
    if (off > k && i_top - k >= 0)
      G[((i_top - k) & 3) * gs + s] = buffer[(size_t)(i_top - k + off) * n_cs + csi[s]];
  }
#pragma unroll
  for (int q = 0; q < AR_FAR_Q; q++)
    if (foff[q] >= 0) {
      const int f = n_near + lane + 32 * q;
      G[(i_top & 3) * gs + f] = buffer[(size_t)i_top * n_cs + foff[q]];
      if (i_top >= 1) G[((i_top - 1) & 3) * gs + f] = buffer[(size_t)(i_top - 1) * n_cs + foff[q]];
    }
  // prologue of the prefetch rings: B z of steps i_top ... i_top - AR_PD + 1 and the far samples (beyond the delay lines) of the
  // vectors of steps i_top - 2 ... i_top - 1 - AR_PD, with plain loads (initial noise and B z, written by earlier kernels);
  // the far warp continues from there, one step per step
  for (int e = tid; e < AR_PD * rows; e += NTT) {
    const int j = i_top - e / rows, rl = e % rows;
    if (j >= 0) bzr[(j & (AR_PS - 1)) * AR_BZ_ROWS + rl] = buffer[(size_t)j * n_cs + row0 + rl];
  }
  for (int e = tid; e < AR_PD * n_glob; e += NTT) {
    const int j = i_top - 2 - e / n_glob, g = e % n_glob, sf = n_near + n_ring + g;
    if (j >= 0) fstage[(j & (AR_PS - 1)) * n_glob + g] = buffer[(size_t)j * n_cs + (size_t)(esi[sf] + 1) * n_cs + csi[sf]];
  }
  const uint32_t bzr_s = smem_u32(bzr), fstage_s = smem_u32(fstage);
  // bytes each barrier expects for step j = i_top - d: what the loop (not the prologue) delivers
  // ONE transaction barrier per step: everything sent during step j + 1 -- row j + 1 at its offsets 1, 2, 3 (vectors of
  // steps j, j - 1, j - 2) and the far samples of the vector of step j - 1 -- completes bytes on the barrier step j waits for.
  // Passing it therefore also proves the EARLY entries of the vector of step j - 1 complete (offset 2: this row; offset 3 and
  // far: a barrier passed one step earlier), so the EARLY part needs no wait of its own.
  auto step_bytes = [&](const int j) -> uint32_t {  // what the loop delivers for step j (vectors of negative steps are not sent)
    return 8u * (n1 + (j >= 1 ? n2 + n_far : 0) + (j >= 2 ? n3 : 0));
  };
  if (tid == 0) {
    for (int d = 0; d < 4 && d <= i_top; d++) {
      const int b = (i_top - d) & 3;
      mbar_arm(bars_s + 8 * b, d >= 1 ? step_bytes(i_top - d) : 0u);  // row i_top + 1 is initial noise: placed by the prologue
    }
  }
  __syncthreads();
  cluster_barrier();

  // the EARLY part of the dot product of step j: registers [mcrit, MP)
  // BLK: the LATE part is the first lq loads of a lane (64 samples each); every offset-1 sample must be inside
  constexpr int LQC = MC >= 0 ? (MC + 1) / 2 : -1;
  const int lq = LQC >= 0 ? LQC : min(NQ, (n1 + 63) / 64);
  auto early_part = [&](const int ph, double(&out)[RPW]) {  // ph = (step & 3): the gather vector of that step
    if (BLK) {
      const double2* g = reinterpret_cast<const double2*>(G + ph * gs) + lane;
      double ex[RPW], ey[RPW];
#pragma unroll
      for (int k = 0; k < RPW; k++) ex[k] = ey[k] = 0.0;
#pragma unroll
      for (int q = 0; q < NQ; q++)
        if (LQC >= 0 ? (q >= LQC) : (q >= lq)) {
          const double2 v = g[32 * q];
#pragma unroll
          for (int k = 0; k < RPW; k++) {
            ex[k] = fma(a[k * SPL + 2 * q], v.x, ex[k]);
            ey[k] = fma(a[k * SPL + 2 * q + 1], v.y, ey[k]);
          }
        }
#pragma unroll
      for (int k = 0; k < RPW; k++) out[k] = ex[k] + ey[k];
    } else {
      const double2* g = reinterpret_cast<const double2*>(G + ph * gs) + sub;
      double e0 = 0.0, e1 = 0.0, e2 = 0.0, e3 = 0.0;
#pragma unroll
      for (int m = 0; m < MP; m += 4)
        if (MC >= 0 ? (m >= 4 * MC) : (m >= mcrit)) {
          const double2 v0 = g[L * (m / 2)], v1 = g[L * (m / 2 + 1)];
          e0 = fma(a[m], v0.x, e0);
          e1 = fma(a[m + 1], v0.y, e1);
          e2 = fma(a[m + 2], v1.x, e2);
          e3 = fma(a[m + 3], v1.y, e3);
        }
      out[0] = (e0 + e1) + (e2 + e3);
    }
  };
  // warps with neither a row nor a far sample only wait for the end
  const bool active = warp * RPW < rows && !far_warp;
  double pre[RPW];
#pragma unroll
  for (int k = 0; k < RPW; k++) pre[k] = 0.0;
  if (active) {
    early_part(i_top & 3, pre);
    // this CTA's own window of the cluster's shared address space (the far samples and their barrier stay at home)

    // GD = false: the steady state (i >= 5), where every "does that step exist" test below is true
    // element offsets inside BUFFER fit 31 bits (3 n_ext n_cs <= INT32_MAX is checked by the host): 32-bit index arithmetic
    const int myrow = row0 + r;
#ifdef MATM_AR_TRACE
    const bool tr_on = lane == 0 && (warp == 0 || warp == 3) && cta < 8;
    const int tr_slot = cta * 2 + (warp == 3), tr_hi = i_top - 4096;
#endif
    // `phase` = i & 3, the gather vector of step i: a compile-time constant in the steady state (the loop below is unrolled
    // by four), so that every shared-memory address of the step is a register plus an immediate
    auto step = [&](auto guard, auto phase, const int i) {
      constexpr bool GD = decltype(guard)::value;
      const int PH = phase_value(phase), PH1 = (PH + 3) & 3, PH2 = (PH + 2) & 3, PH3 = (PH + 1) & 3;  // steps i, i - 1, i - 2, i - 3
      const int ncl = CL > 0 ? CL : cl;
      const uint32_t par = ((i_top - i) >> 2) & 1;
      // B z of this row and step: prefetched into shared memory by the far warp long ago
      const double bz = lds_f64(bzr_s + 8u * ((i & (AR_PS - 1)) * AR_BZ_ROWS + min(r, AR_BZ_ROWS - 1)));
      // (1) the row just produced has landed in G[i]
      AR_TRACE(0, pre[0]);
      mbar_wait_cluster(bars_s + 8 * PH, par);
      AR_TRACE(1, pre[0]);
      if (tid == 0 && (!GD || i >= 4)) mbar_arm(bars_s + 8 * PH, GD ? step_bytes(i - 4) : 8u * n_sample);
      // (2) the LATE part of the dot product
      double acc;
      if (BLK) {
        const double2* g = reinterpret_cast<const double2*>(G + PH * gs) + lane;
        double ax[RPW], ay[RPW];
#pragma unroll
        for (int k = 0; k < RPW; k++) ax[k] = pre[k], ay[k] = 0.0;
#pragma unroll
        for (int q = 0; q < NQ; q++)
          if (LQC >= 0 ? (q < LQC) : (q < lq)) {
            const double2 v = g[32 * q];
#pragma unroll
            for (int k = 0; k < RPW; k++) {
              ax[k] = fma(a[k * SPL + 2 * q], v.x, ax[k]);
              ay[k] = fma(a[k * SPL + 2 * q + 1], v.y, ay[k]);
            }
          }
#pragma unroll
        for (int k = 0; k < RPW; k++) ax[k] += ay[k];
        AR_TRACE(2, ax[0]);
        // reduce-scatter over lane bits 4 and 3 (rows 0, 1 stay in the lower half-warp, then one row per quarter), butterfly
        // over bits 2..0: lane ends up with the full sum of row lane / 8
        const bool hi = (lane & 16) != 0, h8 = (lane & 8) != 0;
        const double k0 = (hi ? ax[2] : ax[0]) + __shfl_xor_sync(0xffffffffu, hi ? ax[0] : ax[2], 16);
        const double k1 = (hi ? ax[3] : ax[1]) + __shfl_xor_sync(0xffffffffu, hi ? ax[1] : ax[3], 16);
        acc = (h8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, h8 ? k0 : k1, 8);
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
      } else {
        const double2* g = reinterpret_cast<const double2*>(G + PH * gs) + sub;
        double acc0 = pre[0], acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
        for (int m = 0; m < MP; m += 4)
          if (MC >= 0 ? (m < 4 * MC) : (m < mcrit)) {
            const double2 v0 = g[L * (m / 2)], v1 = g[L * (m / 2 + 1)];
            acc0 = fma(a[m], v0.x, acc0);
            acc1 = fma(a[m + 1], v0.y, acc1);
            acc2 = fma(a[m + 2], v1.x, acc2);
            acc3 = fma(a[m + 3], v1.y, acc3);
          }
        acc = (acc0 + acc1) + (acc2 + acc3);
        AR_TRACE(2, acc);
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      }
      AR_TRACE(3, acc);
      // (3) publish row r of step i into the gather vectors of steps i-1, i-2, i-3 of every CTA, and to global memory
      if (SPREAD) {
        const double v = bz + acc;
        if (has_row && sub == 0) buffer[(uint32_t)(i * n_cs + myrow)] = v;  // read again (far samples) three steps from now at the earliest
#pragma unroll
        for (int j = 0; j < NJ; j++) {
          const int bsel = (PH - pubk[j]) & 3;
          st_async_f64_if(pubk[j] != 0 && (!GD || i >= pubk[j]), pubd[j] + 8u * (bsel * gs), v, pubb[j] + 8u * PH1);
        }
      } else
      if (has_row && sub == 0) {
        const double v = bz + acc;
        const uint32_t d1 = G_s + 8u * (PH1 * gs + p1), b1 = bars_s + 8u * PH1;
        const uint32_t d2 = G_s + 8u * (PH2 * gs + p2), b2 = b1;
        const uint32_t d3 = G_s + 8u * (PH3 * gs + p3), b3 = b1;
        buffer[(uint32_t)(i * n_cs + myrow)] = v;  // read again (far samples) no sooner than three steps from now
#pragma unroll
        for (int c = 0; c < (CL > 0 ? CL : 8); c++) {
          if (c >= ncl) break;
          if (p1 >= 0 && (!GD || i >= 1)) st_async_f64(mapa_u32(d1, c), v, mapa_u32(b1, c));
          if (p2 >= 0 && (!GD || i >= 2)) st_async_f64(mapa_u32(d2, c), v, mapa_u32(b2, c));
          if (p3 >= 0 && (!GD || i >= 3)) st_async_f64(mapa_u32(d3, c), v, mapa_u32(b3, c));
        }
      }
      AR_TRACE(4, acc);
      // (4) off the critical path: the periodic fence that orders this thread's row stores for the far warp's reads of BUFFER,
      //     and B z for step i - 2 (the far samples themselves are the far warp's business)
      if (fence_mask >= 0 && (i & fence_mask) == 0) __threadfence();
      // (5) the EARLY part of step i - 1, while the row of step i is on its way
      if (!GD || i >= 1) {
        early_part(PH1, pre);
      }
      AR_TRACE(5, pre[0]);
    };

    // i is odd throughout (i_top = 2 n_ext - 1).  Steady state: four steps per trip, phases 3, 2, 1, 0, all four >= 5 (no
    // "does that step exist" test).  Two guarded steps with run-time phases bring i to 3 (mod 4) when i_top = 1 (mod 4),
    // and finish the chain (i = 7 ... 0) after the steady state; one copy of that code serves both.
    int i = i_top;
    while (true) {
      if ((i & 3) == 3)
        for (; i >= 8; i -= 4) {
          step(std::false_type{}, PhaseConst<3>{}, i);
          step(std::false_type{}, PhaseConst<2>{}, i - 1);
          step(std::false_type{}, PhaseConst<1>{}, i - 2);
          step(std::false_type{}, PhaseConst<0>{}, i - 3);
        }
      if (i < 1) break;
      step(std::true_type{}, i & 3, i);
      step(std::true_type{}, (i - 1) & 3, i - 1);
      i -= 2;
    }
  }
  if (far_warp) {
    // The far warp.  Step i: wait for the barrier of step i (row i + 1 is complete in the LATE part of this CTA's vector of
    // step i), move every delay line on by one, send the far values of the vector of step i - 2 (bytes complete on the
    // barrier of step i - 1, like everything else sent during step i).  Samples beyond the delay lines (offset >= 600) come
    // from BUFFER through the prefetch ring, like the B z terms of the CTA's rows.
    const uint32_t bars_self = mapa_u32(bars_s, cta);
    const uint32_t far_self = mapa_u32(G_s, cta) + 8u * (n_near + lane);
    int gidx[AR_FAR_Q];  // index among the samples beyond the delay lines
#pragma unroll
    for (int q = 0; q < AR_FAR_Q; q++) gidx[q] = lane + 32 * q - n_ring;
    auto far_step = [&](const int i) {
      const int PH = i & 3, PH1 = (PH + 3) & 3, PH2 = (PH + 2) & 3;
      mbar_wait_cluster(bars_s + 8 * PH, ((i_top - i) >> 2) & 1);
      // prefetch, AR_PD steps ahead: B z of step i - AR_PD for the rows of this CTA, far samples of the vector of step
      // i - 2 - AR_PD (rows produced >= 566 steps ago, behind the row warps' periodic fence)
      const int jb = i - AR_PD, jf = i - 2 - AR_PD;
      if (jb >= 0) {
#pragma unroll
        for (int hh = 0; hh < AR_BZ_ROWS / 32; hh++) {
          const int rl = lane + 32 * hh;
          if (rl < rows) cp_async_f64(bzr_s + 8u * ((jb & (AR_PS - 1)) * AR_BZ_ROWS + rl), buffer + (uint32_t)(jb * n_cs + row0 + rl));
        }
      }
#pragma unroll
      for (int q = 0; q < AR_FAR_Q; q++)
        if (foff[q] >= 0 && ring_hi[q] == 0u && jf >= 0)
          cp_async_f64(fstage_s + 8u * ((jf & (AR_PS - 1)) * n_glob + gidx[q]), buffer + (uint32_t)(jf * n_cs + foff[q]));
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group %0;" ::"n"(AR_PW) : "memory");  // everything issued more than AR_PW steps ago has landed
      const uint32_t be = bars_self + 8u * PH1;
#pragma unroll
      for (int q = 0; q < AR_FAR_Q; q++)
        if (foff[q] >= 0) {
          double v;
          if (ring_hi[q] != 0u) {  // delay line: take (i - 2 + d, c) out, put (i + 1, c) in its place
            const double vnew = lds_f64(G_s + 8u * (PH * gs) + ring_src[q]);
            v = lds_f64(ring_p[q]);
            sts_f64(ring_p[q], vnew);
            ring_p[q] = ring_p[q] == ring_lo[q] ? ring_hi[q] : ring_p[q] - 8u;
          } else {
            v = lds_f64(fstage_s + 8u * (((i - 2) & (AR_PS - 1)) * n_glob + gidx[q]));
          }
          if (i >= 2) st_async_f64(far_self + 8u * (PH2 * gs + 32 * q), v, be);
        }
    };
    for (int i = i_top; i >= 0; i--) far_step(i);
  }
  // no CTA may leave while values are still on their way to it
  cluster_barrier();
}

// dynamic shared memory of the register-resident kernel: gather vectors, barriers, stencil maps, delay lines, prefetch rings
static size_t chain_v4_smem(int L, int MP, int n_cs, int n_sample, int n_near, const matm_handle::ArStencil& sten) {
  return 4 * (size_t)L * MP * sizeof(double) + 3 * (size_t)n_cs * sizeof(int) + 16 + 8 * sizeof(uint64_t) + 16 +
         ((size_t)sten.n_ring + 1) * sizeof(int) + 16 + (size_t)sten.ring_doubles * sizeof(double) +
         (size_t)AR_PS * (AR_BZ_ROWS + (size_t)(n_sample - n_near - sten.n_ring)) * sizeof(double);
}

template <int L, int MP, int NT>
static int launch_chain_v4(matm_handle* h, const double* A, const int* esi, const int* csi, double* buffer, int n_ext,
                           int n_cs, int n_sample, int n_near, int cl, const matm_handle::ArStencil& sten,
                           cudaStream_t stream) {
  const size_t smem = chain_v4_smem(L, MP, n_cs, n_sample, n_near, sten);
  auto kernel = ar_chain_v4_kernel<L, MP, -1, 0, NT>;
  if constexpr (L == 8 && NT == 256) {
    // the production shape of every process wider than one CTA: 7 row warps x 4 rows = 28 rows per CTA.  mc = ceil(n_cs / 32)
    // bounds the LATE part (n_cs bounds the number of offset-1 samples from above: the first ceil(mc / 2) loads of a lane
    // always hold all of them); the cluster is mc or mc + 1 CTAs.  Both are compile-time constants of these instantiations.
    const int mc = (n_cs + 4 * L - 1) / (4 * L);
#define V4_PICK(MCV, CLV) \
  if (mc == MCV && cl == CLV) kernel = ar_chain_v4_kernel<L, MP, MCV, CLV, NT>;
    if constexpr (MP <= 32) { V4_PICK(2, 2) V4_PICK(2, 3) }
    if constexpr (MP >= 24 && MP <= 40) { V4_PICK(3, 3) V4_PICK(3, 4) }
    if constexpr (MP >= 32) { V4_PICK(4, 4) V4_PICK(4, 5) V4_PICK(5, 5) V4_PICK(5, 6) }
#undef V4_PICK
  }
  MATM_CUDA(h, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cl, 1, 1);
  cfg.blockDim = dim3(NT, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  MATM_CUDA(h, cudaLaunchKernelEx(&cfg, kernel, A, esi, csi, buffer, n_ext, n_cs, n_sample, n_near, cl, sten.n_ring, sten.fence_mask,
                                  sten.ring_doubles));
  h->launches++;
  return MATM_OK;
}

// Which far samples the register-resident kernel serves from delay lines in shared memory (see the kernel): the leading ones,
// in stencil order (offsets grow), as long as their column has an offset-1 sample and the lines fit AR_RING_DOUBLES.  What is
// left takes the global path, which needs its smallest offset to be >= AR_GLOBAL_MIN_OFFSET; otherwise the general kernel
// (cluster barrier per step) runs the process.  The stencil is read back from the device once and cached in the handle.
#define AR_RING_DOUBLES 12288
#define AR_FENCE_STEPS 512
#define AR_GLOBAL_MIN_OFFSET 600  // fence every 512 steps + AR_PD + 2 steps of look-ahead + margin
static int ar_stencil_info(matm_handle* h, const int32_t* esi_dev, const int32_t* csi_dev, int n_ext, int n_cs, int n_sample,
                           int n_near, matm_handle::ArStencil* out) {
  const int64_t shape = (((int64_t)n_sample * 1000003 + n_near) * 1000003 + n_ext) * 1000003 + n_cs;
  const auto key = std::make_pair((const void*)esi_dev, shape);
  auto it = h->ar_stencil.find(key);
  if (it != h->ar_stencil.end()) {
    *out = it->second;
    return MATM_OK;
  }
  std::vector<int32_t> esi((size_t)n_sample), csi((size_t)n_sample);
  MATM_CUDA(h, cudaMemcpy(esi.data(), esi_dev, (size_t)n_sample * sizeof(int32_t), cudaMemcpyDeviceToHost));
  MATM_CUDA(h, cudaMemcpy(csi.data(), csi_dev, (size_t)n_sample * sizeof(int32_t), cudaMemcpyDeviceToHost));
  std::vector<char> has1((size_t)n_cs, 0);
  bool sane = true;
  for (int s = 0; s < n_sample; s++) {
    if (csi[s] < 0 || csi[s] >= n_cs || esi[s] < 0) sane = false;
    else if (s < n_near && esi[s] == 0) has1[(size_t)csi[s]] = 1;
  }
  matm_handle::ArStencil st = {0, 0, -1, sane ? 1 : 0};
  if (sane) {
    int f = n_near;
    for (; f < n_sample; f++) {
      const int lr = esi[f] + 1 - 3;
      if (lr < 1 || !has1[(size_t)csi[f]] || st.ring_doubles + lr > AR_RING_DOUBLES) break;
      if (f > n_near && esi[f] < esi[f - 1]) break;
      st.ring_doubles += lr;
    }
    st.n_ring = f - n_near;
    if (f < n_sample) {
      int dmin = INT32_MAX;
      for (int g = f; g < n_sample; g++) dmin = min(dmin, esi[g] + 1);
      st.fence_mask = AR_FENCE_STEPS - 1;
      if (dmin < AR_GLOBAL_MIN_OFFSET) st.ok = 0;
    }
  }
  h->ar_stencil[key] = st;
  *out = st;
  return MATM_OK;
}

// Picks (lanes per row, samples per lane, cluster size, threads per CTA) for the register-resident kernel; false if the
// process does not fit it.  A step of the chain is bound by what the warps of one SM do between two barrier phases, so the
// rows are spread thin: 256 threads per CTA (two warps per scheduler, 255 registers per thread) and at least 8 lanes per row
// once the process is wider than one warp's worth of rows -- measured on the longest C4 layer (138 x 314, 37 368 steps):
// 512 threads, L = 8, 3 CTAs 0.92 us/step; 512, L = 16, 5 CTAs 0.68; 256, L = 8, 5 CTAs 0.60 (profiles/r02q_chain_latency.txt).
// 512 threads remain for the shapes whose rows would not fit 8 CTAs of 256.
static bool plan_v4(int n_cs, int n_sample, int n_far, int* L, int* MP, int* cl, int* nt) {
  static const int Ls[4] = {4, 8, 16, 32};
  static const int MPs[5] = {16, 24, 32, 40, 48};
  static const int NTs[2] = {256, AR_THREADS};
  for (int ti = 0; ti < 2; ti++) {
    const int threads = NTs[ti];
    if (n_far > 32 * AR_FAR_Q) continue;
    for (int li = 0; li < 4; li++) {
      // 48 samples per lane (96 registers of A) only where the registers are: L = 8 with 256 threads.  It keeps the six-hour
      // layers of C5 (n_sample = 323, 327) on the 8-lane shape instead of 16 lanes x 512 threads (0.60 vs 1.35 us/step).
      const int max_mp = (Ls[li] == 8 && threads == 256) ? 48 : 40;
      if ((n_cs > 32 && Ls[li] < 8) || n_sample > Ls[li] * max_mp) continue;
      if (threads == AR_THREADS && Ls[li] < 16) continue;  // instantiated for the wide stencils only
      int mp = 0;
      for (int mi = 0; mi < 5; mi++)
        if (MPs[mi] <= max_mp && n_sample <= Ls[li] * MPs[mi]) {
          mp = MPs[mi];
          break;
        }
      if (Ls[li] >= 16 && mp < 24) mp = 24;
      const int rows_per_cta = (threads / 32 - 1) * (32 / Ls[li]);  // the last warp serves the far samples
      const int c = (n_cs + rows_per_cta - 1) / rows_per_cta;
      if (c > 8) continue;  // a larger L only lowers rows_per_cta further
      *L = Ls[li];
      *MP = mp;
      *cl = c;
      *nt = threads;
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
    int L, mp, cl3, nt;
    matm_handle::ArStencil sten = {0, 0, -1, 0};
    if (n_near >= 0 && n_near <= n_sample && n_ext >= 4 && plan_v4(n_cs, n_sample, n_sample - n_near, &L, &mp, &cl3, &nt)) {
      int rc = ar_stencil_info(h, esi_dev, csi_dev, n_ext, n_cs, n_sample, n_near, &sten);
      if (rc) return rc;
      // 227 KB of shared memory per CTA on sm_100: a stencil whose rings do not fit runs on the general kernel
      if (chain_v4_smem(L, mp, n_cs, n_sample, n_near, sten) > (size_t)225 * 1024) sten.ok = 0;
    }
    if (sten.ok) {
#define V4_CASE(LV, MV, NV) \
  if (L == LV && mp == MV && nt == NV) \
    return launch_chain_v4<LV, MV, NV>(h, A_dev, esi_dev, csi_dev, buffer_dev, n_ext, n_cs, n_sample, n_near, cl3, sten, st);
      V4_CASE(4, 16, 256) V4_CASE(4, 24, 256) V4_CASE(4, 32, 256) V4_CASE(4, 40, 256)
      V4_CASE(8, 16, 256) V4_CASE(8, 24, 256) V4_CASE(8, 32, 256) V4_CASE(8, 40, 256) V4_CASE(8, 48, 256)
      V4_CASE(16, 24, 256) V4_CASE(16, 32, 256) V4_CASE(16, 40, 256)
      V4_CASE(32, 24, 256) V4_CASE(32, 32, 256) V4_CASE(32, 40, 256)
      V4_CASE(16, 24, 512) V4_CASE(16, 32, 512) V4_CASE(16, 40, 512)
      V4_CASE(32, 24, 512) V4_CASE(32, 32, 512) V4_CASE(32, 40, 512)
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
