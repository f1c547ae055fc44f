// tests/microbench/tmem_bw.cu -- development microbenchmark (not part of the product library).
// Tensor memory (TMEM, 256 KB/SM) used as per-thread private scratch: tcgen05.st once, tcgen05.ld in a loop.
// Question: latency and bytes/clk/SM of tcgen05.ld.32x32b.xN compared with 128-bit shared loads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bw tmem_bw.cu && ./tmem_bw
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

template <int N> struct Ld;
template <> struct Ld<4> {
  __device__ static void ld(uint32_t a, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(a));
  }
  __device__ static void st(uint32_t a, const float* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]));
  }
};
template <> struct Ld<8> {
  __device__ static void ld(uint32_t a, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "r"(a));
  }
  __device__ static void st(uint32_t a, const float* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(a), "f"(v[0]), "f"(v[1]),
                 "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]));
  }
};

// each thread owns COLS words: warps w and w+4 share a lane quadrant and split the 512 columns
template <int N, int WAITEVERY>
__global__ void tmem_kernel(float* out, long long* cyc, int iters, int cols_per_thread) {
  __shared__ uint32_t tbase_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbase_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tb = tbase_s;
  const uint32_t mine = tb + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * cols_per_thread);
  float v[N];
  for (int c = 0; c < cols_per_thread; c += N) {
    for (int i = 0; i < N; ++i) v[i] = (float)(threadIdx.x * 1000 + c + i);
    Ld<N>::st(mine + c, v);
  }
  asm volatile("tcgen05.wait::st.sync.aligned;");
  float acc = 0.0f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll 1
    for (int c = 0; c < cols_per_thread; c += N * WAITEVERY) {
      float u[WAITEVERY][N];
#pragma unroll
      for (int k = 0; k < WAITEVERY; ++k) Ld<N>::ld(mine + c + k * N, u[k]);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int k = 0; k < WAITEVERY; ++k)
#pragma unroll
        for (int i = 0; i < N; ++i) acc += u[k][i];
    }
  }
  long long t1 = clock64();
  // check: sum over my columns of (tid*1000 + col) * iters
  double expect = 0;
  for (int c = 0; c < cols_per_thread; ++c) expect += (double)(threadIdx.x * 1000 + c);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc - (float)(expect * iters);
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tb));
}

// same traffic from shared memory with 128-bit loads, [group][thread][4] layout
__global__ void smem_kernel(float* out, long long* cyc, int iters, int words) {
  extern __shared__ float4 sm4[];
  const int groups = words / 4;
  for (int g = 0; g < groups; ++g) sm4[g * blockDim.x + threadIdx.x] = make_float4(1.f, 2.f, 3.f, (float)g);
  __syncthreads();
  float acc = 0.0f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll 4
    for (int g = 0; g < groups; ++g) {
      float4 v;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                   : "r"((uint32_t)__cvta_generic_to_shared(&sm4[g * blockDim.x + threadIdx.x])));
      acc += v.x + v.y + v.z + v.w;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int N, int W>
static int run_tmem(const char* name, int threads, int cols, int iters, float* d_out, long long* d_cyc, int grid) {
  tmem_kernel<N, W><<<grid, threads>>>(d_out, d_cyc, iters, cols);
  CK(cudaDeviceSynchronize());
  long long cyc[1024];
  float out[256];
  CK(cudaMemcpy(cyc, d_cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(out, d_out, sizeof(out), cudaMemcpyDeviceToHost));
  double bytes = (double)threads * cols * 4.0 * iters;
  float maxdiff = 0;
  for (int i = 0; i < threads && i < 256; ++i) maxdiff = fmaxf(maxdiff, fabsf(out[i]));
  printf("%-34s threads %3d cols/thr %3d : %8.1f cyc/iter, %6.1f B/clk/SM, %.1f cyc per ld   (check diff %.1f)\n", name, threads, cols,
         (double)cyc[0] / iters, bytes / cyc[0], (double)cyc[0] / iters / (cols / N), maxdiff);
  return 0;
}

int main() {
  float* d_out; long long* d_cyc;
  CK(cudaMalloc(&d_out, 1024 * 1024 * 4)); CK(cudaMalloc(&d_cyc, 1024 * 8));
  const int grid = 148, iters = 200;
  run_tmem<4, 1>("tmem x4 wait every ld", 32, 64, iters, d_out, d_cyc, grid);
  run_tmem<8, 1>("tmem x8 wait every ld", 32, 64, iters, d_out, d_cyc, grid);
  run_tmem<4, 1>("tmem x4 wait every ld", 128, 64, iters, d_out, d_cyc, grid);
  run_tmem<4, 4>("tmem x4 wait every 4 lds", 128, 64, iters, d_out, d_cyc, grid);
  run_tmem<8, 2>("tmem x8 wait every 2 lds", 128, 64, iters, d_out, d_cyc, grid);
  run_tmem<8, 4>("tmem x8 wait every 4 lds", 128, 128, iters, d_out, d_cyc, grid);
  run_tmem<4, 4>("tmem x4 wait every 4 lds", 256, 64, iters, d_out, d_cyc, grid);
  run_tmem<8, 4>("tmem x8 wait every 4 lds", 256, 128, iters, d_out, d_cyc, grid);
  run_tmem<8, 4>("tmem x8 wait every 4 lds", 256, 256, iters, d_out, d_cyc, grid);
  for (int threads : {32, 128, 256}) {
    int words = 128;
    CK(cudaFuncSetAttribute(smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    smem_kernel<<<grid, threads, threads * words * 4>>>(d_out, d_cyc, iters, words);
    CK(cudaDeviceSynchronize());
    long long cyc[1024];
    CK(cudaMemcpy(cyc, d_cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost));
    printf("smem LDS.128 threads %3d words/thr %d : %8.1f cyc/iter, %6.1f B/clk/SM\n", threads, words, (double)cyc[0] / iters,
           (double)threads * words * 4.0 * iters / cyc[0]);
  }
  return 0;
}
