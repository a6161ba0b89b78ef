// Microbenchmark behind a design decision of point_stream_kernel:
// can the TMA engine take over the camera-side gather (x rows, 80 bytes each) and
// scatter (y rows, 80 bytes each, reduce-add) of one 32-row tile, instead of the
// LSU path (coalesced-by-piece LDG.128 + smem park, smem transposition + RED.ADD.F64)?
// Every variant processes the same (tile, lane) -> camera stream.
//
// Build:  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o build/tma_microbench tools/tma_microbench.cu
// Run  :  build/tma_microbench [cameras] [tiles_per_warp]
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); std::exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  uint32_t ok;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(phase) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void bulk_reduce_add_f64(void* dst_global, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(dst_global),
               "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int kWarps = 4;

// ---- gather, LSU path: as point_stream_kernel does it today ----
__global__ void __launch_bounds__(128, 3) gather_lsu(const int* cams, int tiles_per_warp, const double* x, double* out) {
  __shared__ __align__(16) double sx[kWarps][320];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gw = blockIdx.x * kWarps + warp;
  const int* my = cams + (size_t)gw * tiles_per_warp * 32;
  double acc = 0.0;
  double2 xq[5];
  auto gather = [&](int t) {
    const int c_own = my[t * 32 + lane];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int idx = j * 32 + lane, o = idx / 5, h = idx - 5 * o;
      const int c = __shfl_sync(0xffffffffu, c_own, o);
      xq[j] = __ldg(reinterpret_cast<const double2*>(x + c) + h);
    }
  };
  gather(0);
  for (int t = 0; t < tiles_per_warp; ++t) {
#pragma unroll
    for (int j = 0; j < 5; ++j) reinterpret_cast<double2*>(sx[warp])[j * 32 + lane] = xq[j];
    __syncwarp();
    if (t + 1 < tiles_per_warp) gather(t + 1);
#pragma unroll
    for (int k = 0; k < 5; ++k) { const double2 v = reinterpret_cast<const double2*>(sx[warp])[lane * 5 + k]; acc += v.x + v.y; }
    __syncwarp();
  }
  if (acc == 12345.678) out[0] = acc;
}

// ---- gather, TMA path: every lane issues one 80-byte bulk copy of its row's camera vector ----
__global__ void __launch_bounds__(128, 3) gather_tma(const int* cams, int tiles_per_warp, const double* x, double* out) {
  __shared__ __align__(128) double sx[kWarps][2][320];
  __shared__ uint64_t bars[kWarps][2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gw = blockIdx.x * kWarps + warp;
  const int* my = cams + (size_t)gw * tiles_per_warp * 32;
  if (lane == 0) { mbar_init(&bars[warp][0], 1); mbar_init(&bars[warp][1], 1); }
  __syncwarp();
  double acc = 0.0;
  auto issue = [&](int t) {
    const int b = t & 1;
    const int c = my[t * 32 + lane];
    if (lane == 0) mbar_expect_tx(&bars[warp][b], 2560);
    __syncwarp();
    bulk_g2s(&sx[warp][b][lane * 10], x + c, 80, &bars[warp][b]);
  };
  issue(0);
  for (int t = 0; t < tiles_per_warp; ++t) {
    if (t + 1 < tiles_per_warp) issue(t + 1);
    mbar_wait(&bars[warp][t & 1], (t >> 1) & 1);
#pragma unroll
    for (int k = 0; k < 5; ++k) { const double2 v = reinterpret_cast<const double2*>(sx[warp][t & 1])[lane * 5 + k]; acc += v.x + v.y; }
    __syncwarp();
  }
  if (acc == 12345.678) out[0] = acc;
}

// ---- scatter, LSU path: smem transposition + 9 coalesced RED.ADD.F64 per tile ----
__global__ void __launch_bounds__(128, 3) scatter_lsu(const int* cams, int tiles_per_warp, double* yrep, int nrep, int rep_stride) {
  __shared__ __align__(16) double sx[kWarps][320];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gw = blockIdx.x * kWarps + warp;
  const int* my = cams + (size_t)gw * tiles_per_warp * 32;
  double* yacc = yrep + (size_t)(gw % nrep) * rep_stride;
  for (int t = 0; t < tiles_per_warp; ++t) {
    const int cam = my[t * 32 + lane];
#pragma unroll
    for (int k = 0; k < 9; ++k) sx[warp][lane * 9 + k] = 1.0 + k;
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 9; ++j) {
      const int idx = j * 32 + lane, o = idx / 9, k = idx - 9 * o;
      const int c = __shfl_sync(0xffffffffu, cam, o);
      atomicAdd(yacc + c + k, sx[warp][idx]);
    }
    __syncwarp();
  }
}

// ---- scatter, TMA path: every lane issues one 80-byte bulk reduce-add of its row ----
__global__ void __launch_bounds__(128, 3) scatter_tma(const int* cams, int tiles_per_warp, double* yrep, int nrep, int rep_stride) {
  __shared__ __align__(128) double sx[kWarps][2][320];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gw = blockIdx.x * kWarps + warp;
  const int* my = cams + (size_t)gw * tiles_per_warp * 32;
  double* yacc = yrep + (size_t)(gw % nrep) * rep_stride;
  for (int t = 0; t < tiles_per_warp; ++t) {
    const int cam = my[t * 32 + lane];
    double2* row = reinterpret_cast<double2*>(&sx[warp][t & 1][lane * 10]);
    bulk_wait_read<1>();  // the reduce that read this buffer two tiles ago is done with it
#pragma unroll
    for (int k = 0; k < 4; ++k) row[k] = make_double2(1.0 + 2 * k, 2.0 + 2 * k);
    row[4] = make_double2(9.0, 0.0);
    fence_async_smem();
    bulk_reduce_add_f64(yacc + cam, row, 80);
    bulk_commit();
  }
  bulk_wait_read<0>();
}

int main(int argc, char** argv) {
  const int C = argc > 1 ? std::atoi(argv[1]) : 1778;
  const int tiles_per_warp = argc > 2 ? std::atoi(argv[2]) : 96;
  int dev_sms = 0;
  CK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0));
  const int grid = dev_sms * 3;
  const int warps = grid * kWarps;
  const size_t n = (size_t)warps * tiles_per_warp * 32;
  // camera of each (tile, lane): a random walk so that neighbouring rows see nearby cameras, like BAL tracks
  std::vector<int> cams(n);
  uint64_t s = 88172645463325252ull;
  auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
  for (size_t i = 0; i < n; ++i) cams[i] = 10 * (int)(rnd() % (uint64_t)C);
  int* d_cams; double *d_x, *d_y, *d_out;
  const int nrep = 8, rep_stride = 10 * C + 6;
  CK(cudaMalloc(&d_cams, n * sizeof(int)));
  CK(cudaMalloc(&d_x, (size_t)10 * C * 8));
  CK(cudaMalloc(&d_y, (size_t)nrep * rep_stride * 8));
  CK(cudaMalloc(&d_out, 8));
  CK(cudaMemcpy(d_cams, cams.data(), n * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemset(d_x, 0, (size_t)10 * C * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  auto time = [&](const char* name, auto launch) {
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    const int reps = 20;
    for (int i = 0; i < reps; ++i) launch();
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    const double us = 1e3 * ms / reps;
    std::printf("%-12s %8.1f us per launch  %6.2f ns per tile per SM  (%d tiles per SM)\n", name, us,
                1e3 * us / ((double)warps * tiles_per_warp / dev_sms), warps * tiles_per_warp / dev_sms);
  };
  time("gather_lsu", [&] { gather_lsu<<<grid, 128>>>(d_cams, tiles_per_warp, d_x, d_out); });
  time("gather_tma", [&] { gather_tma<<<grid, 128>>>(d_cams, tiles_per_warp, d_x, d_out); });
  CK(cudaMemset(d_y, 0, (size_t)nrep * rep_stride * 8));
  time("scatter_lsu", [&] { scatter_lsu<<<grid, 128>>>(d_cams, tiles_per_warp, d_y, nrep, rep_stride); });
  std::vector<double> y1((size_t)nrep * rep_stride), y2((size_t)nrep * rep_stride);
  CK(cudaMemset(d_y, 0, y1.size() * 8));
  scatter_lsu<<<grid, 128>>>(d_cams, tiles_per_warp, d_y, nrep, rep_stride);
  CK(cudaMemcpy(y1.data(), d_y, y1.size() * 8, cudaMemcpyDeviceToHost));
  time("scatter_tma", [&] { scatter_tma<<<grid, 128>>>(d_cams, tiles_per_warp, d_y, nrep, rep_stride); });
  CK(cudaMemset(d_y, 0, y1.size() * 8));
  scatter_tma<<<grid, 128>>>(d_cams, tiles_per_warp, d_y, nrep, rep_stride);
  CK(cudaMemcpy(y2.data(), d_y, y2.size() * 8, cudaMemcpyDeviceToHost));
  size_t bad = 0;
  for (size_t i = 0; i < y1.size(); ++i) bad += (y1[i] != y2[i]);
  std::printf("scatter results differ at %zu of %zu entries\n", bad, y1.size());
  CK(cudaGetLastError());
  return 0;
}
