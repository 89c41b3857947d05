// Microbenchmark (dev tool): read-only HBM bandwidth ceilings on this GPU.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o readbw readbw.cu && ./readbw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ld4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

template <int U>
__global__ void k_read(const float* __restrict__ a, size_t n4, float* out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  float s = 0.f;
  for (; i + (U - 1) * stride < n4; i += U * stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld4(a + 4 * (i + u * stride));
#pragma unroll
    for (int u = 0; u < U; ++u) s += v[u].x + v[u].y + v[u].z + v[u].w;
  }
  if (s == 123.456f) out[0] = s;
}

__global__ void k_copy(const float4* __restrict__ a, float4* __restrict__ b, size_t n4) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) b[i] = a[i];
}

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// bulk-copy ring reader: 1 producer thread, consumers just touch one word per slot
template <int SLOTS>
__global__ void __launch_bounds__(128, 1) k_bulk(const float* __restrict__ a, size_t bytes_per_cta, int slot_bytes, float* out) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* full = (uint64_t*)(sm + (size_t)SLOTS * slot_bytes);
  uint64_t* empty = full + SLOTS;
  if (threadIdx.x == 0) {
    for (int s = 0; s < SLOTS; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&full[s])), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&empty[s])), "r"(1));
    }
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  const unsigned char* src = (const unsigned char*)a + (size_t)blockIdx.x * bytes_per_cta;
  const int n = (int)(bytes_per_cta / slot_bytes);
  auto wait = [](uint64_t* bar, uint32_t par) {
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(s32(bar)), "r"(par) : "memory");
  };
  if (threadIdx.x == 0) {
    for (int it = 0; it < n; ++it) {
      const int s = it % SLOTS;
      wait(&empty[s], ((it / SLOTS) & 1) ^ 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&full[s])), "r"(slot_bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(sm + (size_t)s * slot_bytes)), "l"(src + (size_t)it * slot_bytes), "r"(slot_bytes), "r"(s32(&full[s])) : "memory");
    }
  } else if (threadIdx.x == 32) {
    float acc = 0.f;
    for (int it = 0; it < n; ++it) {
      const int s = it % SLOTS;
      wait(&full[s], (it / SLOTS) & 1);
      acc += *(volatile float*)(sm + (size_t)s * slot_bytes);
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&empty[s])) : "memory");
    }
    if (acc == 123.456f) out[0] = acc;
  }
}

int main() {
  const size_t bytes = (size_t)1 << 30;  // 1 GiB
  float *a, *b, *out;
  cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMalloc(&out, 4);
  cudaMemset(a, 1, bytes); cudaMemset(b, 0, bytes);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto timeit = [&](const char* name, auto launch, double moved) {
    for (int i = 0; i < 3; ++i) launch();
    cudaDeviceSynchronize();
    float best = 1e9;
    for (int rep = 0; rep < 10; ++rep) {
      cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    printf("%-40s %8.3f ms  %8.1f GB/s  (%s)\n", name, best, moved / best / 1e6, cudaGetErrorString(cudaGetLastError()));
  };
  const size_t n4 = bytes / 16;
  for (int ctas : {148 * 2, 148 * 4, 148 * 8, 148 * 16}) {
    char nm[64];
    snprintf(nm, 64, "read LDG.128 U=4 256thr x %d", ctas);
    timeit(nm, [&] { k_read<4><<<ctas, 256>>>(a, n4, out); }, (double)bytes);
    snprintf(nm, 64, "read LDG.128 U=8 256thr x %d", ctas);
    timeit(nm, [&] { k_read<8><<<ctas, 256>>>(a, n4, out); }, (double)bytes);
  }
  timeit("read LDG.128 U=8 512thr x 592", [&] { k_read<8><<<592, 512>>>(a, n4, out); }, (double)bytes);
  timeit("read LDG.128 U=16 256thr x 1184", [&] { k_read<16><<<1184, 256>>>(a, n4, out); }, (double)bytes);
  timeit("copy float4 (read+write bytes)", [&] { k_copy<<<148 * 16, 256>>>((const float4*)a, (float4*)b, n4); }, 2.0 * bytes);
  timeit("cudaMemcpy D2D (read+write bytes)", [&] { cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice); }, 2.0 * bytes);
  for (int slot : {16384, 32768, 49152}) {
    const size_t per = (bytes / 148 / slot) * slot;
    char nm[64];
    {
      constexpr int S = 3;
      const size_t smem = (size_t)S * slot + 2 * S * 8;
      cudaFuncSetAttribute(k_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      snprintf(nm, 64, "bulk ring 3 x %d B, 148 CTAs", slot);
      timeit(nm, [&] { k_bulk<S><<<148, 128, smem>>>(a, per, slot, out); }, (double)per * 148);
    }
    if (slot <= 32768) {
      constexpr int S = 6;
      const size_t smem = (size_t)S * slot + 2 * S * 8;
      cudaFuncSetAttribute(k_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      snprintf(nm, 64, "bulk ring 6 x %d B, 148 CTAs", slot);
      timeit(nm, [&] { k_bulk<S><<<148, 128, smem>>>(a, per, slot, out); }, (double)per * 148);
    }
  }
  {
    constexpr int S = 12;
    const int slot = 16384;
    const size_t per = (bytes / 148 / slot) * slot;
    const size_t smem = (size_t)S * slot + 2 * S * 8;
    cudaFuncSetAttribute(k_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    timeit("bulk ring 12 x 16384 B, 148 CTAs", [&] { k_bulk<S><<<148, 128, smem>>>(a, per, slot, out); }, (double)per * 148);
  }
  {
    constexpr int S = 3;
    const int slot = 32768;
    const size_t per = (bytes / 296 / slot) * slot;
    const size_t smem = (size_t)S * slot + 2 * S * 8;
    cudaFuncSetAttribute(k_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    timeit("bulk ring 3 x 32768 B, 296 CTAs (2/SM)", [&] { k_bulk<S><<<296, 128, smem>>>(a, per, slot, out); }, (double)per * 296);
  }
  return 0;
}
