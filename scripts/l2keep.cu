// Microbenchmark (dev tool): does a set S read with one L2 eviction priority survive a stream of Y bytes read (or
// written) with another?  Prints the time of the first and of the second pass over S.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2keep l2keep.cu && ./l2keep
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

enum { kNone = 0, kNormal = 1, kLast = 2, kFirst = 3 };

__device__ __forceinline__ uint64_t policy(int kind) {
  uint64_t p;
  if (kind == kLast) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  else if (kind == kFirst) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ float4 ld4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ld4h(const float* p, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ float4 ldk(const float* p, int kind, uint64_t pol) {
  return kind == kNone ? ld4(p) : ld4h(p, pol);
}
__device__ __forceinline__ void st4h(float* p, float4 v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}

// every CTA reads a contiguous chunk; `rev` hands the chunks out in reverse so that another SM reads each one
__global__ void k_read(const float* __restrict__ a, size_t n4, int kind, int rev, float* out) {
  const uint64_t pol = policy(kind);
  const int b = rev ? (gridDim.x - 1 - blockIdx.x) : blockIdx.x;
  const size_t per = (n4 + gridDim.x - 1) / gridDim.x;
  const size_t lo = (size_t)b * per, hi = (lo + per < n4) ? lo + per : n4;
  float s = 0.f;
  constexpr int U = 8;
  size_t i = lo + threadIdx.x;
  for (; i + (U - 1) * blockDim.x < hi; i += U * blockDim.x) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ldk(a + 4 * (i + u * blockDim.x), kind, pol);
#pragma unroll
    for (int u = 0; u < U; ++u) s += v[u].x + v[u].y + v[u].z + v[u].w;
  }
  for (; i < hi; i += blockDim.x) {
    const float4 v = ldk(a + 4 * i, kind, pol);
    s += v.x + v.y + v.z + v.w;
  }
  if (s == 123.456f) out[0] = s;
}
__global__ void k_write(float* __restrict__ a, size_t n4, int kind) {
  const uint64_t pol = policy(kind);
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    if (kind == kNone) *reinterpret_cast<float4*>(a + 4 * i) = make_float4(1.f, 2.f, 3.f, 4.f);
    else st4h(a + 4 * i, make_float4(1.f, 2.f, 3.f, 4.f), pol);
  }
}

static const char* nm(int k) {
  return k == kNone ? "none" : k == kNormal ? "normal" : k == kLast ? "last" : "first";
}

int main(int argc, char** argv) {
  const size_t MB = 1 << 20;
  const int setaside_mb = argc > 1 ? atoi(argv[1]) : 0;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  printf("L2 %d MB, persisting max %d MB, set-aside requested %d MB\n", (int)(prop.l2CacheSize / MB),
         (int)(prop.persistingL2CacheMaxSize / MB), setaside_mb);
  if (setaside_mb > 0) {
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)setaside_mb * MB);
    size_t got = 0;
    cudaDeviceGetLimit(&got, cudaLimitPersistingL2CacheSize);
    printf("set-aside now %d MB\n", (int)(got / MB));
  }
  float *S, *Y, *F, *out;
  cudaMalloc(&S, 128 * MB);
  cudaMalloc(&Y, 1024 * MB);
  cudaMalloc(&F, 512 * MB);
  cudaMalloc(&out, 4096);
  cudaMemset(S, 0, 128 * MB);
  cudaMemset(Y, 0, 1024 * MB);
  cudaMemset(F, 0, 512 * MB);
  const int grid = 148 * 8, thr = 256;
  cudaEvent_t e[4];
  for (auto& x : e) cudaEventCreate(&x);
  printf("%6s %7s %7s %5s %6s %4s | %9s %9s %9s\n", "S_MB", "S_pol", "Y_pol", "Y_MB", "Y_op", "rev", "pass1_us", "pass2_us", "pass2GB/s");
  const int sizes[] = {32, 48};
  struct Cfg { int ps, py, ymb, wr; };
  const Cfg cfgs[] = {{kNone, kNone, 256, 0},  {kNormal, kFirst, 256, 0}, {kLast, kFirst, 256, 0},  {kLast, kNone, 256, 0},
                      {kLast, kFirst, 512, 1},  {kLast, kNone, 512, 1}};
  for (const Cfg& c : cfgs)
    for (int smb : sizes)
      for (int rev = 0; rev < 1; ++rev) {
        const size_t n4 = (size_t)smb * MB / 16;
        k_read<<<grid, thr>>>(F, 512 * MB / 16, kNone, 0, out);  // previous kernels' leftovers: normal lines
        cudaEventRecord(e[0]);
        k_read<<<grid, thr>>>(S, n4, c.ps, 0, out);
        cudaEventRecord(e[1]);
        if (c.wr) k_write<<<grid, thr>>>(Y, (size_t)c.ymb * MB / 16, c.py);
        else k_read<<<grid, thr>>>(Y, (size_t)c.ymb * MB / 16, c.py, 0, out);
        cudaEventRecord(e[2]);
        k_read<<<grid, thr>>>(S, n4, c.ps, rev, out);
        cudaEventRecord(e[3]);
        cudaDeviceSynchronize();
        float t1, t2;
        cudaEventElapsedTime(&t1, e[0], e[1]);
        cudaEventElapsedTime(&t2, e[2], e[3]);
        printf("%6d %7s %7s %5d %6s %4d | %9.1f %9.1f %9.0f\n", smb, nm(c.ps), nm(c.py), c.ymb, c.wr ? "write" : "read", rev,
               t1 * 1e3, t2 * 1e3, smb * 1.048576 / t2);
      }
  // reset evict_last lines so that nothing stays pinned for whoever runs next on this GPU
  cudaCtxResetPersistingL2Cache();
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
