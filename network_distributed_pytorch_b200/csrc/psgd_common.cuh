// Shared device-side types and small helpers for the PowerSGD kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <stdint.h>

namespace psgd {

constexpr int kThreads = 256;      // CTA size of the streaming kernels (8 warps)
constexpr int kWarps = kThreads / 32;
constexpr int kConsumerWarps = 8;   // consumer warps of the streamed kernels (psgd_stream.cuh); warp 8 is the producer
constexpr int kOrthThreads = 512;  // CTA size of the generic (global-memory) Gram-Schmidt kernel
constexpr int kOrthSmallThreads = 1024;  // CTA size of the register-resident Gram-Schmidt kernels (R <= 8)
// At R = 16 the same kernels run 512 threads x 4 rows (same rows per CTA): 128 registers per thread leave room for the
// rolled step loop with rotating columns; the unrolled 1024 x 2 form spent its time fetching instructions (K2 118 us).
__host__ __device__ constexpr int orth_threads(int R) { return R >= 16 ? 512 : kOrthSmallThreads; }
constexpr int kOrthCluster = 8;    // portable cluster size: the generic Gram-Schmidt kernel, and K2's fallback hardware cluster
constexpr int kOrthClusterMax = 16;  // non-portable cluster size: K2's hardware clusters (groups of 1-16 CTAs per matrix inside)

// per-matrix constants, uploaded once per plan
struct MatDesc {
  long long p_off;   // element offset of P_i in p_mem   (reducer.py:95)
  long long q_off;   // element offset of Q_i in q_mem   (reducer.py:96)
  long long ws_off;  // element offset of the K3 partials of this matrix in the workspace
  int n, m, r;       // rows, columns, r = min(n, m, rank) (reducer.py:76-78)
  int cgw;           // K3: column groups (threads) per tile row, power of two <= kThreads
  int cpt;           // K3: columns per thread (4 when m % 4 == 0, else 1)
  int t3_rows;       // K3: rows per tile
  int nblk3;         // K3: row blocks = ceil(n / t3_rows)
  int pad_;
};
static_assert(sizeof(MatDesc) == 56, "MatDesc layout");

// K3 tile: matrix, first row, row count, first column, row-block index
struct ColTile {
  int mat, row0, nrows, col0;
  int blk, pad0, pad1, pad2;
};

// flat-buffer tile: `count` elements of tensor `tensor` starting at `start`, living at `flat_off`
struct VecTile {
  long long flat_off;
  int tensor, start, count, pad_;
};
constexpr int kVecTileElems = 2048;

// K3 stage-2 tile: `count` consecutive Q elements of matrix `mat` starting at local element `e0`
struct QTile {
  int mat, e0, count, pad_;
};
constexpr int kQTileElems = 1024;

// device views of the per-step pointer arrays (one ring slot)
struct PtrSlot {
  const float* const* in;   // M_i   (grad_in,  reducer.py:43)
  float* const* out;        // grad_out
  float* const* mem;        // memory_out
  const float* const* g;    // raw gradient  (fuse_add)
  const float* const* e;    // error memory  (fuse_add)
  const unsigned* flags;    // alignment flags per matrix
  const float* const* vin;  // <=1-D inputs
  float* const* vout;       // <=1-D outputs
  const float* const* vg;   // <=1-D raw gradients (fuse_add)
  const float* const* ve;   // <=1-D error memories (fuse_add)
};

enum : unsigned {
  kFlagInVec = 1u,   // M_i 16-byte aligned and m % 4 == 0
  kFlagOutVec = 2u,  // out_i and mem_i 16-byte aligned
  kFlagGeVec = 4u,   // grad_i and err_i 16-byte aligned
};

struct DevPlan {
  const MatDesc* mats;
  const int4* tiles1;      // K1 row tiles {mat, row0, nrows, -}
  const int4* tiles4;      // K4 row tiles
  const ColTile* tiles3;   // K3 tiles
  const VecTile* vtiles;   // tiles of the flat <=1-D buffer
  const QTile* qtiles;     // tiles of q_mem for the K3 second stage
  int n_mat, n_vec;
  int wc_cap;              // max columns of Q staged in shared memory at once
  int pad_;
};

// ---- programmatic dependent launch ------------------------------------------------------
// The five kernels of a step are launched with programmatic stream serialisation: each lets its successor's CTAs
// take SMs as they free up (launch latency, barrier set-up, descriptor and first-task fetches happen in the
// shadow of the predecessor's tail) and every CTA of the successor waits here, before it touches anything a
// previous kernel of the stream wrote, until the predecessor has completed and flushed.  Both are no-ops for a
// launch without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// named barrier over the consumer warps of a streamed kernel only (threads 0 .. 255)
__device__ __forceinline__ void consumer_sync() {
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
}

// ---- mbarrier / shared-memory PTX wrappers --------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_n(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Spins until the phase with the given parity has completed.  A protocol error (or a peer GPU that never answers)
// would otherwise hang the kernel for good: after ~4 s of waiting the thread reports and traps instead.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  long long t0 = 0;
  for (unsigned spins = 1;; ++spins) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
    if ((spins & 0x3fffu) == 0) {
      if (t0 == 0) {
        t0 = clock64();
      } else if (clock64() - t0 > 8000000000LL) {
        printf("psgd: mbarrier wait timed out (CTA %d, thread %d, smem barrier 0x%x, parity %u)\n", (int)blockIdx.x,
               (int)threadIdx.x, addr, parity);
        __trap();
      }
    }
  }
}
// the same for waits inside register-tight loops: no report, just the trap after ~2^27 polls (seconds)
__device__ __forceinline__ void mbar_wait_lite(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  for (unsigned spins = 1;; ++spins) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
    if (spins == (1u << 27)) __trap();
  }
}
// ---- distributed shared memory: deliver a value into another CTA of the cluster and signal its mbarrier ----------
// (one one-way message instead of a store followed by a cluster-wide barrier round trip)
__device__ __forceinline__ uint32_t cluster_map_u32(uint32_t smem_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_async_f32(uint32_t remote_addr, float v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(remote_addr),
               "r"(__float_as_uint(v)), "r"(remote_bar)
               : "memory");
}

// ---- peer (NVLink) memory: flags, loads, in-switch reductions --------------------------------------------------
constexpr int kMaxPeers = 8;
struct PeerArgs {
  const float* in[kMaxPeers];   // input buffer of every rank, mapped into this process
  unsigned* sig[kMaxPeers];     // signal pad of every rank (uint32 slots), mapped into this process
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_sys4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_sys1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ float4 mm_ld_reduce4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w) : "memory");
}

// all threads of the CTA return once every rank has posted `epoch` in this rank's flags[0 .. world)
__device__ __forceinline__ void peer_flags_wait(const unsigned* flags, int world, unsigned epoch, int rank, int which) {
  if ((int)threadIdx.x < world) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0) {
      if (clock64() - t0 > 20000000000LL) {  // ~10 s: a crashed peer must not hang the GPU
        printf("psgd k_peer_allreduce_mc: rank %d timed out waiting for rank %d (barrier %d, epoch %u)\n", rank,
               (int)threadIdx.x, which, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void mm_st1(float* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
// A result element goes either to ordinary memory or, through a multicast address, into the same place of every
// rank's receive buffer at once ("push": the all-reduce's transfer happens while the producing kernel still runs).
__device__ __forceinline__ void st_out(float* p, float v, bool mc) {
  if (mc) mm_st1(p, v); else *p = v;
}
__device__ __forceinline__ float mm_ld_reduce1(const float* mc) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
  return v;
}

// What a kernel needs to read a buffer "summed over all ranks" (SUM all-reduce on load) from symmetric memory: every
// rank's copy as mapped into this process, the NVSwitch multicast address of the same buffer (or null), this rank's
// signal pad and everybody else's.  mode 1: one multimem.ld_reduce per element (the switch adds the copies); mode 2:
// one load per rank, added in rank order (bit-identical on every rank by construction, needs no multicast mapping).
// Peer data is only ever touched with plain / multimem loads and stores.
struct PeerLoad {
  const float* in[kMaxPeers];  // modes 1, 2: rank r's copy of the buffer; mode 3: in[0] = this rank's receive buffer
  unsigned* sig[kMaxPeers];
  const float* mc;
  long long stride;            // mode 3: elements between the slots of the receive buffer
  int rank, world, slot0, mode;
  unsigned epoch;
  int light;  // flag protocol: 0 = st.release.sys / ld.acquire.sys, 1 = device-scope fences around relaxed system-scope accesses
};
// mode 3 ("push"): every rank has ALREADY written its contribution into slot [rank] of every rank's receive buffer
// (multimem.st from the producing kernel), so after the flag barrier the sum is W local loads in slot order.
enum { kPeerOff = 0, kPeerMultimem = 1, kPeerPull = 2, kPeerSlots = 3 };

// "This rank's input is complete" to every peer (one CTA does it, after its own dependency wait), then every CTA
// waits until all ranks have said so.  All threads of the CTA must call it (it ends in __syncthreads()).
__device__ __forceinline__ void peer_barrier(const PeerLoad& pm, bool announce) {
  if (announce && (int)threadIdx.x < pm.world) {
    // everything this thread has observed (the finished grids before its dependency wait) is ordered before the flag
    if (pm.light) {
      // peers read this GPU's memory through this GPU's L2, so a device-scope fence orders the data before the flag
      __threadfence();
      asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(pm.sig[threadIdx.x] + pm.slot0 + pm.rank), "r"(pm.epoch) : "memory");
    } else {
      st_release_sys(pm.sig[threadIdx.x] + pm.slot0 + pm.rank, pm.epoch);
    }
  }
  if (pm.light) {
    if ((int)threadIdx.x < pm.world) {
      const unsigned* f = pm.sig[pm.rank] + pm.slot0 + threadIdx.x;
      const long long t0 = clock64();
      for (;;) {
        unsigned v;
        asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if ((int)(v - pm.epoch) >= 0) break;
        if (clock64() - t0 > 20000000000LL) {
          printf("psgd peer barrier: rank %d timed out waiting for rank %d (epoch %u)\n", pm.rank, (int)threadIdx.x, pm.epoch);
          __trap();
        }
      }
      __threadfence();
    }
    __syncthreads();
    return;
  }
  peer_flags_wait(pm.sig[pm.rank] + pm.slot0, pm.world, pm.epoch, pm.rank, 0);
}
__device__ __forceinline__ float4 peer_sum4(const PeerLoad& pm, long long e) {  // e % 4 == 0, 16-byte aligned buffers
  if (pm.mode == kPeerMultimem) return mm_ld_reduce4(pm.mc + e);
  if (pm.mode == kPeerSlots) {
    float4 s = __ldcg(reinterpret_cast<const float4*>(pm.in[0] + e));
    for (int r = 1; r < pm.world; ++r) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(pm.in[0] + (long long)r * pm.stride + e));
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    return s;
  }
  float4 s = ld_sys4(pm.in[0] + e);
  for (int r = 1; r < pm.world; ++r) {
    const float4 v = ld_sys4(pm.in[r] + e);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  return s;
}
__device__ __forceinline__ float peer_sum1(const PeerLoad& pm, long long e) {
  if (pm.mode == kPeerMultimem) return mm_ld_reduce1(pm.mc + e);
  if (pm.mode == kPeerSlots) {
    float s = __ldcg(pm.in[0] + e);
    for (int r = 1; r < pm.world; ++r) s += __ldcg(pm.in[0] + (long long)r * pm.stride + e);
    return s;
  }
  float s = ld_sys1(pm.in[0] + e);
  for (int r = 1; r < pm.world; ++r) s += ld_sys1(pm.in[r] + e);
  return s;
}

// ---- memory helpers -------------------------------------------------------------------

// streaming 128-bit load: read-only path, do not allocate in L1 (M is touched once per pass)
__device__ __forceinline__ float4 ld_stream4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_stream1(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream4(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_stream1(float* p, float v) {
  asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

// V consecutive floats; V = 4 uses 128-bit accesses, V = 1 is the unaligned fallback
template <int V>
struct Pack {
  float v[V];
};
template <int V>
__device__ __forceinline__ Pack<V> pk_zero() {
  Pack<V> r;
#pragma unroll
  for (int i = 0; i < V; ++i) r.v[i] = 0.f;
  return r;
}
template <int V>
__device__ __forceinline__ Pack<V> pk_ld_stream(const float* p);
template <>
__device__ __forceinline__ Pack<4> pk_ld_stream<4>(const float* p) {
  float4 t = ld_stream4(p);
  Pack<4> r;
  r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
  return r;
}
template <>
__device__ __forceinline__ Pack<1> pk_ld_stream<1>(const float* p) {
  Pack<1> r;
  r.v[0] = ld_stream1(p);
  return r;
}
template <int V>
__device__ __forceinline__ void pk_st_stream(float* p, const Pack<V>& a);
template <>
__device__ __forceinline__ void pk_st_stream<4>(float* p, const Pack<4>& a) {
  st_stream4(p, make_float4(a.v[0], a.v[1], a.v[2], a.v[3]));
}
template <>
__device__ __forceinline__ void pk_st_stream<1>(float* p, const Pack<1>& a) {
  st_stream1(p, a.v[0]);
}
template <int V>
__device__ __forceinline__ Pack<V> pk_ld_shared(const float* p);
template <>
__device__ __forceinline__ Pack<4> pk_ld_shared<4>(const float* p) {
  float4 t = *reinterpret_cast<const float4*>(p);
  Pack<4> r;
  r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
  return r;
}
template <>
__device__ __forceinline__ Pack<1> pk_ld_shared<1>(const float* p) {
  Pack<1> r;
  r.v[0] = *p;
  return r;
}

// Shared-memory row pitch (floats) of the transposed Q stage: multiple of 4 (128-bit reads)
// and skewed so the transposing stores of `R` interleaved columns spread over the banks.
__host__ __device__ inline int q_pitch(int wc, int R) {
  int skew = 32 / R;
  if (skew < 4) skew = 4;
  if (R == 1) skew = 0;
  return ((wc + 31) / 32) * 32 + skew;
}

// Stage columns [c0, c0+wc) of the row-major m x r matrix Q into shared memory transposed:
// Qs[c * pitch + j] = Q[(c0 + j) * r + c] / div ; rows c in [r, R) are zero filled so the
// compute loops can run with the compile-time rank R.
template <int R>
__device__ __forceinline__ void stage_q(float* Qs, const float* __restrict__ Q, int c0, int wc,
                                        int r, int pitch, float div) {
  const float* src = Q + (size_t)c0 * r;
  const int total = wc * r;
  for (int e = threadIdx.x; e < total; e += blockDim.x) {
    const int j = e / r;
    const int c = e - j * r;
    float v = __ldg(src + e);
    if (div != 1.0f) v = v / div;
    Qs[c * pitch + j] = v;
  }
  if (r < R) {
    const int zt = (R - r) * wc;
    for (int e = threadIdx.x; e < zt; e += blockDim.x) {
      const int c = r + e / wc;
      const int j = e % wc;
      Qs[c * pitch + j] = 0.f;
    }
  }
}

// Butterfly "transpose-reduce": every lane enters with NV partial sums; on return lane L
// holds in v[0] the sum over all 32 lanes of partial number (L / (32 / NV)).  NV-1 shuffles
// for the folding steps plus log2(32/NV) plain steps instead of 5 * NV.
template <int NV>
__device__ __forceinline__ float warp_fold(float (&v)[NV], int lane) {
  static_assert(NV == 1 || NV == 2 || NV == 4 || NV == 8 || NV == 16 || NV == 32, "NV");
  int bit = 16;
#pragma unroll
  for (int half = NV / 2; half >= 1; half /= 2) {
    const bool upper = (lane & bit) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = upper ? v[i] : v[i + half];
      const float keep = upper ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
    bit >>= 1;
  }
  float x = v[0];
#pragma unroll
  for (; bit >= 1; bit >>= 1) x += __shfl_xor_sync(0xffffffffu, x, bit);
  return x;
}

__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

}  // namespace psgd
