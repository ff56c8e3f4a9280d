// Which store pattern writes a (rows x n_t) float matrix at HBM speed?  The synthesis kernel owns tiles of 32 rows x ~3000
// consecutive samples; this microbenchmark writes the same tiles with different thread -> element mappings.
//   0  lane = sample (128 B per warp and row), loop over the 32 rows              (what phase 3 does)
//   1  lane = 4 consecutive samples, st.v4 (512 B per warp and row), loop over rows
//   2  row-major inside the tile: a warp walks one row segment of 1 KB at a time
//   3  shared-memory staging + cp.async.bulk (TMA) row segments of SEG bytes
//   4  linear fill (upper bound)
// usage: store_pattern <rows> <n_t> <tile_samples>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ void st_cs(float* p, float v) { asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }
__device__ __forceinline__ void st_cs4(float* p, float4 v) {
  asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(256, 3) k_tile(float* out, int64_t n_t, int ts, int n_tt) {
  const int64_t dt = blockIdx.x / n_tt;
  const int tt = blockIdx.x % n_tt;
  const int64_t j0 = (int64_t)tt * ts, j1 = min(j0 + ts, n_t);
  float* base = out + dt * 32 * n_t;
  const int tid = threadIdx.x;
  if (MODE == 0) {
    for (int64_t j = j0 + tid; j < j1; j += 256) {
      float v = (float)j * 1e-6f;
#pragma unroll 4
      for (int r = 0; r < 32; r++) st_cs(base + r * n_t + j, v + r);
    }
  } else if (MODE == 1) {
    for (int64_t j = j0 + 4 * tid; j < j1; j += 1024) {
      float v = (float)j * 1e-6f;
#pragma unroll 4
      for (int r = 0; r < 32; r++) st_cs4(base + r * n_t + j, make_float4(v + r, v, v, v));
    }
  } else if (MODE == 2) {
    const int warp = tid >> 5, lane = tid & 31;
    for (int r = warp; r < 32; r += 8) {
      for (int64_t j = j0 + 4 * lane; j < j1; j += 128) st_cs4(base + r * n_t + j, make_float4((float)j, 1.f, 2.f, 3.f));
    }
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// staging: [32 rows][SEG floats]; after a chunk is computed one thread per row issues a bulk store
template <int SEGF>
__global__ void __launch_bounds__(256, 3) k_tma(float* out, int64_t n_t, int ts, int n_tt) {
  extern __shared__ __align__(128) float stage[];  // 2 x 32 x SEGF
  const int64_t dt = blockIdx.x / n_tt;
  const int tt = blockIdx.x % n_tt;
  const int64_t j0 = (int64_t)tt * ts, j1 = min(j0 + ts, n_t);
  float* base = out + dt * 32 * n_t;
  const int tid = threadIdx.x;
  int buf = 0;
  for (int64_t jc = j0; jc < j1; jc += SEGF, buf ^= 1) {
    float* st = stage + buf * 32 * SEGF;
    // the buffer was handed to the TMA two chunks ago: wait until it has been read
    if (tid < 32) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    __syncthreads();
    const int nseg = (int)min((int64_t)SEGF, j1 - jc);
    for (int e = tid; e < 32 * SEGF / 4; e += 256) {
      const int r = e / (SEGF / 4), c = (e % (SEGF / 4)) * 4;
      *reinterpret_cast<float4*>(st + r * SEGF + c) = make_float4((float)(jc + c), (float)r, 2.f, 3.f);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid < 32) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(base + tid * n_t + jc),
                   "r"(smem_u32(st + tid * SEGF)), "r"(nseg * 4)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (tid < 32) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__global__ void k_linear(float4* out, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = make_float4(1.f, 2.f, 3.f, 4.f);
}

int main(int argc, char** argv) {
  const int64_t rows = argc > 1 ? atoll(argv[1]) : 32768, n_t = argc > 2 ? atoll(argv[2]) : 180000;
  const int ts = argc > 3 ? atoi(argv[3]) : 3072;
  float* out;
  const size_t bytes = (size_t)rows * n_t * 4;
  CK(cudaMalloc(&out, bytes));
  const int n_tt = (int)((n_t + ts - 1) / ts);
  const unsigned grid = (unsigned)(rows / 32 * n_tt);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](const char* name, auto launch) {
    for (int w = 0; w < 2; w++) launch();
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    for (int w = 0; w < 5; w++) launch();
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms / 5, bytes / (ms / 5 * 1e-3) / 1e9);
  };
  printf("rows %ld n_t %ld tile %d samples, %.1f GB\n", (long)rows, (long)n_t, ts, bytes / 1e9);
  run("0 lane=sample st.cs.f32, loop rows", [&] { k_tile<0><<<grid, 256>>>(out, n_t, ts, n_tt); });
  run("1 lane=4 samples st.cs.v4, loop rows", [&] { k_tile<1><<<grid, 256>>>(out, n_t, ts, n_tt); });
  run("2 row-major in tile st.cs.v4", [&] { k_tile<2><<<grid, 256>>>(out, n_t, ts, n_tt); });
  CK(cudaFuncSetAttribute(k_tma<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 32 * 256 * 4));
  CK(cudaFuncSetAttribute(k_tma<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 32 * 512 * 4));
  CK(cudaFuncSetAttribute(k_tma<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 32 * 128 * 4));
  run("3 smem + TMA bulk, 512 B segments", [&] { k_tma<128><<<grid, 256, 2 * 32 * 128 * 4>>>(out, n_t, ts, n_tt); });
  run("3 smem + TMA bulk, 1 KB segments", [&] { k_tma<256><<<grid, 256, 2 * 32 * 256 * 4>>>(out, n_t, ts, n_tt); });
  run("3 smem + TMA bulk, 2 KB segments", [&] { k_tma<512><<<grid, 256, 2 * 32 * 512 * 4>>>(out, n_t, ts, n_tt); });
  run("4 linear fill", [&] { k_linear<<<148 * 8, 256>>>((float4*)out, (int64_t)(bytes / 16)); });
  CK(cudaGetLastError());
  return 0;
}
