// Persistent, warp-specialised streaming kernels for the two row-streaming phases (K1: P = M Q and
// K4+K5: out = P Q^T, mem = M - out).  One CTA per SM; one producer thread keeps a ring of shared-
// memory slots full with bulk asynchronous copies (cp.async.bulk, the 1-D TMA path) signalled through
// mbarriers; eight consumer warps take sub-spans of ROWS rows each, read M and the transposed Q copy
// from shared memory and run the rank-r FMAs.  Bytes in flight per SM are set by the ring (~150 KB)
// instead of by registers, which is what lets a read stream approach the HBM limit.
//
// Everything the consumers need arrives through the ring, in order: when the producer moves to a new
// matrix it first sends that matrix' Q (raw, row-major) as ring units, the consumers transpose it from
// the slot into the Q stage (shared -> shared), then the M units follow.  No consumer ever waits on a
// global load in steady state.  Work is handed out dynamically (guided self-scheduling over a global
// task counter): with equal static shares the SMs finish up to 40 % apart.
//
// Matrices that do not qualify (m % 4 != 0, unaligned base, a row that does not fit a slot, Q too large
// for the stage) are left to the direct-load kernels in psgd_kernels.cuh.
#pragma once
#include "psgd_common.cuh"

namespace psgd {

constexpr int kStreamThreads = (kConsumerWarps + 1) * 32;  // + 1 producer warp
constexpr int kMaxSlots = 8;
constexpr int kMatCache = 256;  // matrices whose {base, q, m, r} are cached in shared memory

// A task is a run of rows of ONE matrix, cut by the producer into units of at most one slot.
struct StreamTask {
  int mat, row_begin, row_end, rot0;
};

struct StreamPlan {
  const StreamTask* tasks;
  int* counters;         // [0] next task, [1] CTAs finished (the last one re-arms both for the next launch)
  int n_tasks;
  int slot_bytes;        // bytes per ring slot (multiple of 128)
  int n_slots;
  int q_bytes;           // bytes reserved for the transposed Q stage
  unsigned long long* dbg;  // optional per-CTA cycle breakdown (PSGD_STREAM_DEBUG), else nullptr
  // side job folded into the launch (saves a kernel and its gap): the flat <=1-D buffer
  const VecTile* vtiles;
  float* flat;
  int n_vtiles;
  int vec_mode;          // 0 none, 1 pack, 2 pack with grad + err, 3 unpack (/ world)
  float vec_world;
  // second side job: tiny matrices that do not qualify for the ring (e.g. a 64 x 147 first convolution) are done
  // straight from global memory by the consumer threads of one CTA each, instead of costing an extra launch
  const int* direct_ids;
  const int* direct_row_pre;  // n_direct + 1 prefix sums of their row counts
  int n_direct;
  int split_max_sub;     // K1: units with at most this many row groups are also cut by columns (0: never)
  int out_mc;            // K1: p_mem and flat are multicast addresses (results are pushed to every rank, see st_out)
};

// The rows of all such matrices form one list of items; warp w of CTA b takes items (k * gridDim + b) * 8 + w, so a
// 64-row matrix is spread over eight CTAs and costs each of them about a microsecond.
__device__ __forceinline__ bool direct_item(const StreamPlan& sp, int k, int& mat, int& row) {
  const int item = (k * (int)gridDim.x + (int)blockIdx.x) * kConsumerWarps + (int)(threadIdx.x >> 5);
  if (item >= sp.direct_row_pre[sp.n_direct]) return false;
  int d = 0;
  while (item >= sp.direct_row_pre[d + 1]) ++d;
  mat = sp.direct_ids[d];
  row = item - sp.direct_row_pre[d];
  return true;
}

// P = M Q, one row per warp, lanes over the columns, straight from global memory.
template <int R, bool FUSE>
__device__ __forceinline__ void direct_p_from_m(const DevPlan& pl, const PtrSlot& ps, const StreamPlan& sp,
                                                const float* __restrict__ q_mem, float* __restrict__ p_mem, float q_div) {
  const int lane = threadIdx.x & 31;
  const int rounds = (sp.direct_row_pre[sp.n_direct] + (int)gridDim.x * kConsumerWarps - 1) / ((int)gridDim.x * kConsumerWarps);
  for (int k = 0; k < rounds; ++k) {
    int mat, row;
    if (!direct_item(sp, k, mat, row)) continue;
    const MatDesc md = pl.mats[mat];
    const int m = md.m, r = md.r;
    const float* Q = q_mem + md.q_off;
    const float* M = ps.in[mat] + (size_t)row * m;
    float acc[R];
#pragma unroll
    for (int c = 0; c < R; ++c) acc[c] = 0.f;
    for (int j = lane; j < m; j += 32) {
      float x;
      if (FUSE) {
        x = ps.g[mat][(size_t)row * m + j] + ps.e[mat][(size_t)row * m + j];
        const_cast<float*>(M)[j] = x;
      } else {
        x = M[j];
      }
#pragma unroll
      for (int c = 0; c < R; ++c) {
        if (c < r) {
          float q = __ldg(Q + (size_t)j * r + c);
          if (q_div != 1.0f) q = q / q_div;
          acc[c] = fmaf(x, q, acc[c]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < R; ++c) {
      const float t = warp_sum(acc[c]);
      if (lane == 0 && c < r) st_out(p_mem + md.p_off + (size_t)row * r + c, t, sp.out_mc != 0);
    }
  }
}

// out = P (Q / world)^T, mem = M - out, one row per warp.
template <int R>
__device__ __forceinline__ void direct_finalize(const DevPlan& pl, const PtrSlot& ps, const StreamPlan& sp,
                                                const float* __restrict__ p_mem, const float* __restrict__ q_mem, float world) {
  const int lane = threadIdx.x & 31;
  const int rounds = (sp.direct_row_pre[sp.n_direct] + (int)gridDim.x * kConsumerWarps - 1) / ((int)gridDim.x * kConsumerWarps);
  for (int k = 0; k < rounds; ++k) {
    int mat, row;
    if (!direct_item(sp, k, mat, row)) continue;
    const MatDesc md = pl.mats[mat];
    const int m = md.m, r = md.r;
    const float* Q = q_mem + md.q_off;
    float p[R];
#pragma unroll
    for (int c = 0; c < R; ++c) p[c] = (c < r) ? __ldg(p_mem + md.p_off + (size_t)row * r + c) : 0.f;
    const float* M = ps.in[mat] + (size_t)row * m;
    float* Out = ps.out[mat] + (size_t)row * m;
    float* Mem = ps.mem[mat] + (size_t)row * m;
    for (int j = lane; j < m; j += 32) {
      float o = 0.f;
#pragma unroll
      for (int c = 0; c < R; ++c) {
        if (c < r) {
          float q = __ldg(Q + (size_t)j * r + c);
          if (world != 1.0f) q = q / world;
          o = fmaf(p[c], q, o);
        }
      }
      const float x = M[j];
      Out[j] = o;
      Mem[j] = x - o;
    }
  }
}

// The flat-buffer gather / scatter of the <=1-D tensors as a side job of a streamed kernel: the consumer threads
// of CTA b take tiles b, b + gridDim.x, ... (a few KB each) while the producer's first copies are in flight.
__device__ __forceinline__ void stream_vec_side_job(const StreamPlan& sp, const PtrSlot& ps) {
  constexpr int NT = kConsumerWarps * 32;
  for (int t = blockIdx.x; t < sp.n_vtiles; t += gridDim.x) {
    const VecTile vt = sp.vtiles[t];
    float* f = sp.flat + vt.flat_off;
    if (sp.vec_mode == 1) {
      const float* s = ps.vin[vt.tensor] + vt.start;
      for (int i = threadIdx.x; i < vt.count; i += NT) st_out(f + i, s[i], sp.out_mc != 0);
    } else if (sp.vec_mode == 2) {
      const float* g = ps.vg[vt.tensor] + vt.start;
      const float* e = ps.ve[vt.tensor] + vt.start;
      float* send = const_cast<float*>(ps.vin[vt.tensor]) + vt.start;
      for (int i = threadIdx.x; i < vt.count; i += NT) {
        const float v = g[i] + e[i];
        send[i] = v;
        st_out(f + i, v, sp.out_mc != 0);
      }
    } else if (sp.vec_mode == 3) {
      float* d = ps.vout[vt.tensor] + vt.start;
      if (sp.vec_world != 1.0f) {
        for (int i = threadIdx.x; i < vt.count; i += NT) d[i] = f[i] / sp.vec_world;
      } else {
        for (int i = threadIdx.x; i < vt.count; i += NT) d[i] = f[i];
      }
    }
  }
}
constexpr int kDbgWords = 8;  // per CTA: t0, t1 (ns), cycles: desc wait, data wait, Q units, total, units, switches

// what travels with a slot: an M unit {mat, row0, nrows > 0, rot}, a Q unit {mat, j0, -count, flags},
// or the end marker {mat < 0}
constexpr int kQFlagBulk = 1 << 16;  // Q unit: data was bulk-copied into the slot (else: fetch from global)
constexpr int kQFlagChunk1 = 1 << 17;  // Q unit: second column half of a column-chunked matrix

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// ---- PTX wrappers (mbarrier helpers: psgd_common.cuh) ---------------------------------------------
// global -> shared bulk copy (bytes % 16 == 0, both addresses 16-byte aligned), completion on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

#ifndef PSGD_K1_UNROLL_R4
#define PSGD_K1_UNROLL_R4 2
#endif
#ifndef PSGD_K1_UNROLL_R8
#define PSGD_K1_UNROLL_R8 2
#endif
#ifndef PSGD_ROWS1_R8
#define PSGD_ROWS1_R8 4
#endif
#ifndef PSGD_ROWS1_R16
#define PSGD_ROWS1_R16 2
#endif
#ifndef PSGD_ROWS4_R16
#define PSGD_ROWS4_R16 2
#endif
// rows of M one consumer warp works on at a time (every Q value read from shared memory is used ROWS times)
__host__ __device__ constexpr int rows_s1(int R) { return (R <= 2) ? 8 : (R <= 4) ? 4 : (R <= 8) ? PSGD_ROWS1_R8 : (R == 16) ? PSGD_ROWS1_R16 : 1; }
__host__ __device__ constexpr int rows_s4(int R) { return (R <= 8) ? 4 : (R == 16) ? PSGD_ROWS4_R16 : 1; }
template <int R> struct RowsS1 { static constexpr int v = rows_s1(R); };
template <int R> struct RowsS4 { static constexpr int v = rows_s4(R); };

// shared memory map: [Q stage | ring slots | full, empty, desc, partial barriers | descriptor per slot | matrix cache |
// partial sums of column-split units]
struct StreamSmem {
  float* qs;
  unsigned char* ring;
  uint64_t* full;   // data of the slot has landed (tx bytes)
  uint64_t* empty;  // all consumer warps are done with the slot
  uint64_t* desc;   // the slot's descriptor is published (well before its data lands)
  uint64_t* pbar;   // K1, column-split units: every consumer warp has delivered its partial sums of the slot's unit
  volatile int4* slot_desc;
  struct MatInfo { const float* base; const float* q; const float* g; const float* e; int m; int r; int nchunk; int pad; };
  MatInfo* matinfo;
  float* part;      // K1, column-split units: [slot][item][32] partial sums
};
constexpr int kSplitItems = 12;                  // most (row group, column slice) items of a column-split unit
constexpr int kPartBytes = kSplitItems * 32 * 4;  // per ring slot
constexpr int kStreamSmemExtra = 4 * kMaxSlots * 8 + kMaxSlots * 16 + kMatCache * 48;  // + n_slots * kPartBytes
static_assert(sizeof(StreamSmem::MatInfo) == 48, "MatInfo layout");

__device__ __forceinline__ StreamSmem carve(unsigned char* base, const StreamPlan& sp) {
  StreamSmem s;
  s.qs = reinterpret_cast<float*>(base);
  s.ring = base + sp.q_bytes;
  s.full = reinterpret_cast<uint64_t*>(base + sp.q_bytes + (size_t)sp.n_slots * sp.slot_bytes);
  s.empty = s.full + kMaxSlots;
  s.desc = s.empty + kMaxSlots;
  s.pbar = s.desc + kMaxSlots;
  s.slot_desc = reinterpret_cast<volatile int4*>(s.pbar + kMaxSlots);
  s.matinfo = reinterpret_cast<StreamSmem::MatInfo*>(const_cast<int4*>(s.slot_desc) + kMaxSlots);
  s.part = reinterpret_cast<float*>(s.matinfo + kMatCache);
  return s;
}

__device__ __forceinline__ StreamSmem::MatInfo mat_info(const DevPlan& pl, const PtrSlot& ps, const StreamSmem& sm,
                                                        const float* q_mem, int mat) {
  if (mat < kMatCache) return sm.matinfo[mat];
  StreamSmem::MatInfo mi;
  const MatDesc md = pl.mats[mat];
  mi.base = ps.in[mat];
  mi.q = q_mem + md.q_off;
  mi.g = ps.g[mat];
  mi.e = ps.e[mat];
  mi.m = md.m;
  mi.r = md.r;
  mi.nchunk = md.pad_ > 1 ? md.pad_ : 1;
  mi.pad = 0;
  return mi;
}

// ---- Q all-reduce as the prologue of a streamed kernel (reducer.py:145-147 without a launch of its own) -----------
// After the flag barrier ("every rank's Q = M^T P is written") CTA b sums the b-th slice of Q over the ranks, divides
// by n_workers and stores the mean to the local q_mem; a grid-wide arrival count (the grid is one CTA per SM, all
// resident) then tells every CTA that the whole of q_mem is in place, and the kernel goes on to stage Q from there as
// it would at one worker.  One barrier across ranks, one across the grid, no copy-out, no separate divide.
struct PeerQ {
  PeerLoad pm;       // pm.mode == kPeerOff: no prologue
  float* q_out;      // local q_mem (receives the mean)
  long long n;       // elements of Q
  float div;         // n_workers
  unsigned gtarget;  // value *gsync reaches once every CTA of this launch has arrived (monotonic counter)
  unsigned* gsync;
};

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async;" ::: "memory");
}

__device__ __forceinline__ void peer_q_prologue(const PeerQ& pq) {  // all threads of the CTA
  pdl_wait();  // this rank's Q partials are complete
  peer_barrier(pq.pm, blockIdx.x == 0);
  const long long n4 = (pq.n + 3) >> 2;
  const long long per = (n4 + gridDim.x - 1) / gridDim.x;
  const long long lo = (long long)blockIdx.x * per;
  const long long hi = (lo + per < n4) ? lo + per : n4;
  for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    float4 v = peer_sum4(pq.pm, 4 * i);  // the symmetric buffers are padded to whole float4s
    v.x = v.x / pq.div; v.y = v.y / pq.div; v.z = v.z / pq.div; v.w = v.w / pq.div;
    if (4 * i + 4 <= pq.n) {
      *reinterpret_cast<float4*>(pq.q_out + 4 * i) = v;
    } else {
      const float t[4] = {v.x, v.y, v.z, v.w};
      for (int k = 0; 4 * i + k < pq.n; ++k) pq.q_out[4 * i + k] = t[k];
    }
  }
  __threadfence();
  fence_proxy_async();  // q_mem is read next by bulk copies (async proxy), possibly of another SM
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(pq.gsync, 1u);
    const long long t0 = clock64();
    while ((int)(ld_acquire_gpu(pq.gsync) - pq.gtarget) < 0) {
      if (clock64() - t0 > 20000000000LL) {
        printf("psgd: grid-wide Q arrival timed out in CTA %d (rank %d)\n", (int)blockIdx.x, pq.pm.rank);
        __trap();
      }
    }
    __threadfence();
  }
  __syncthreads();
  fence_proxy_async();
}

// common prologue: barriers + matrix cache
__device__ __forceinline__ void stream_setup(const DevPlan& pl, const PtrSlot& ps, const StreamPlan& sp,
                                             const StreamSmem& sm, const float* q_mem) {
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    for (int s = 0; s < sp.n_slots; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], kConsumerWarps);
      mbar_init(&sm.desc[s], 1);
      mbar_init(&sm.pbar[s], kConsumerWarps);
    }
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < pl.n_mat && i < kMatCache; i += blockDim.x) {
    const MatDesc md = pl.mats[i];
    sm.matinfo[i].base = ps.in[i];
    sm.matinfo[i].q = q_mem + md.q_off;
    sm.matinfo[i].g = ps.g[i];
    sm.matinfo[i].e = ps.e[i];
    sm.matinfo[i].m = md.m;
    sm.matinfo[i].r = md.r;
    sm.matinfo[i].nchunk = md.pad_ > 1 ? md.pad_ : 1;
  }
  __syncthreads();
}


// The producer: one thread takes tasks from the global counter, cuts them into units and refills slots
// as the consumers release them.  Nothing on its per-unit path touches global memory: the task after
// the current one is requested while the current one streams.
// The producer: one thread takes tasks from the global counter, cuts them into units and refills slots
// as the consumers release them.  Nothing on its per-unit path touches global memory: the task after
// the current one is requested while the current one streams.
template <int ROWS, bool FUSE = false>
__device__ __forceinline__ void stream_producer(const DevPlan& pl, const PtrSlot& ps, const StreamPlan& sp,
                                                const StreamSmem& sm, const float* q_mem) {
  int it = 0;
  int prev_mat = -1;
  int t = atomicAdd(&sp.counters[0], 1);
  StreamTask task = sp.tasks[t < sp.n_tasks ? t : 0];
  pdl_wait();  // nothing above depends on an earlier kernel of the stream; the copies below do
  int prev_chunk = 0;
  while (t < sp.n_tasks) {
    const int t_next = atomicAdd(&sp.counters[0], 1);  // in flight while this task streams
    const StreamSmem::MatInfo mi = mat_info(pl, ps, sm, q_mem, task.mat);
    const int ld = mi.m;                 // row length in global memory
    const int nchunk = mi.nchunk;        // 1, or 2: the Q stage holds half of the columns at a time (wide matrix, high rank)
    const int m = ld / nchunk;           // columns per pass = row length inside a slot
    // Column-chunked matrix: the task's rows are streamed once per column half -- [Q_a][rows, left half][Q_b][rows,
    // right half] -- each row segment with its own bulk copy.  Both passes cut the rows into the same units with the
    // same warp rotation, so in K1 the lane that stored a row's partial sum in pass a adds pass b to its own store.
    for (int chunk = 0; chunk < nchunk; ++chunk) {
      const int col0 = chunk * m;
      if (task.mat != prev_mat || chunk != prev_chunk) {  // new matrix (or column half): its Q goes first
        const float* qsrc = mi.q + (size_t)col0 * mi.r;
        const bool bulk = (reinterpret_cast<uintptr_t>(qsrc) & 15) == 0;
        const int qrows = max(4, (sp.slot_bytes / (mi.r * 4)) & ~3);
        for (int j0 = 0; j0 < m; j0 += qrows, ++it) {
          const int cnt = min(qrows, m - j0);
          const int slot = it % sp.n_slots;
          const uint32_t phase = (uint32_t)(it / sp.n_slots) & 1u;
          mbar_wait(&sm.empty[slot], phase ^ 1u);
          sm.slot_desc[slot].x = task.mat; sm.slot_desc[slot].y = j0;
          sm.slot_desc[slot].z = -cnt; sm.slot_desc[slot].w = (bulk ? kQFlagBulk : 0) | (chunk ? kQFlagChunk1 : 0);
          mbar_arrive(&sm.desc[slot]);
          if (bulk) {
            const uint32_t bytes = (uint32_t)cnt * (uint32_t)mi.r * 4u;
            mbar_arrive_expect_tx(&sm.full[slot], bytes);
            bulk_g2s(sm.ring + (size_t)slot * sp.slot_bytes, qsrc + (size_t)j0 * mi.r, bytes, &sm.full[slot]);
          } else {
            mbar_arrive(&sm.full[slot]);  // every use of a slot completes one phase of each of its barriers
          }
        }
        prev_mat = task.mat;
        prev_chunk = chunk;
      }
      // FUSE: a unit carries the rows of grad and, right behind them, the same rows of err
      const int rpu = max(1, sp.slot_bytes / (m * 4 * (FUSE ? 2 : 1)));
      const int sub_per_unit = (rpu + ROWS - 1) / ROWS;
      int rot = task.rot0;
      for (int row0 = task.row_begin; row0 < task.row_end; row0 += rpu, ++it) {
        const int nrows = min(rpu, task.row_end - row0);
        const int slot = it % sp.n_slots;
        const uint32_t phase = (uint32_t)(it / sp.n_slots) & 1u;
        mbar_wait(&sm.empty[slot], phase ^ 1u);  // passes at once on the first lap
        const uint32_t bytes = (uint32_t)nrows * (uint32_t)m * 4u;
        sm.slot_desc[slot].x = task.mat; sm.slot_desc[slot].y = row0;
        sm.slot_desc[slot].z = nrows; sm.slot_desc[slot].w = rot | (chunk << 8);
        mbar_arrive(&sm.desc[slot]);
        unsigned char* dst = sm.ring + (size_t)slot * sp.slot_bytes;
        mbar_arrive_expect_tx(&sm.full[slot], FUSE ? 2u * bytes : bytes);
        if (nchunk == 1) {
          if (FUSE) {
            bulk_g2s(dst, mi.g + (size_t)row0 * m, bytes, &sm.full[slot]);
            bulk_g2s(dst + bytes, mi.e + (size_t)row0 * m, bytes, &sm.full[slot]);
          } else {
            bulk_g2s(dst, mi.base + (size_t)row0 * m, bytes, &sm.full[slot]);
          }
        } else {
          const uint32_t rb = (uint32_t)m * 4u;  // one bulk copy per row segment
          for (int i = 0; i < nrows; ++i) {
            const size_t goff = (size_t)(row0 + i) * ld + col0;
            if (FUSE) {
              bulk_g2s(dst + (size_t)i * rb, mi.g + goff, rb, &sm.full[slot]);
              bulk_g2s(dst + bytes + (size_t)i * rb, mi.e + goff, rb, &sm.full[slot]);
            } else {
              bulk_g2s(dst + (size_t)i * rb, mi.base + goff, rb, &sm.full[slot]);
            }
          }
        }
        rot = (rot + sub_per_unit) % kConsumerWarps;
      }
    }
    t = t_next;
    if (t < sp.n_tasks) task = sp.tasks[t];  // one round trip, covered by the slots already in flight
  }
  {  // end marker
    const int slot = it % sp.n_slots;
    const uint32_t phase = (uint32_t)(it / sp.n_slots) & 1u;
    mbar_wait(&sm.empty[slot], phase ^ 1u);
    sm.slot_desc[slot].x = -1;
    mbar_arrive(&sm.desc[slot]);
  }
  // the last CTA to get here re-arms the counters, so the next launch needs no memset
  __threadfence();
  if (atomicAdd(&sp.counters[1], 1) == (int)gridDim.x - 1) {
    sp.counters[0] = 0;
    sp.counters[1] = 0;
    __threadfence();
  }
}

// Consumers, Q unit: rows [j0, j0 + cnt) of the row-major m x r matrix Q -> transposed stage
// Qs[c * pitch + j] = Q[j * r + c] / div, rows c in [r, R) zeroed.  Called by all consumer threads.
template <int R>
__device__ __forceinline__ void consume_q_unit(float* Qs, const float* slot, const float* __restrict__ qglobal,
                                               bool bulk, int j0, int cnt, int m, int r, int pitch, float div) {
  constexpr int NT = kConsumerWarps * 32;
  const float* src = bulk ? slot : qglobal + (size_t)j0 * r;
  if (bulk && r == R && (R % 4) == 0) {  // one row of Q per thread and step: 128-bit reads, conflict-free writes
    for (int j = threadIdx.x; j < cnt; j += NT) {
#pragma unroll
      for (int c4 = 0; c4 < R; c4 += 4) {
        float4 v = *reinterpret_cast<const float4*>(src + (size_t)j * R + c4);
        if (div != 1.0f) { v.x = v.x / div; v.y = v.y / div; v.z = v.z / div; v.w = v.w / div; }
        Qs[(c4 + 0) * pitch + j0 + j] = v.x;
        Qs[(c4 + 1) % R * pitch + j0 + j] = v.y;
        Qs[(c4 + 2) % R * pitch + j0 + j] = v.z;
        Qs[(c4 + 3) % R * pitch + j0 + j] = v.w;
      }
    }
  } else {
    for (int j = threadIdx.x; j < cnt; j += NT) {
      for (int c = 0; c < r; ++c) {
        float v = bulk ? src[(size_t)j * r + c] : __ldg(src + (size_t)j * r + c);
        if (div != 1.0f) v = v / div;
        Qs[c * pitch + j0 + j] = v;
      }
    }
  }
  if (r < R && j0 == 0) {
    const int zt = (R - r) * m;
    for (int e = threadIdx.x; e < zt; e += NT) Qs[(r + e / m) * pitch + (e % m)] = 0.f;
  }
}

// =======================================================================================
// K1 (streamed):  P = M Q ; with FUSE the caller's M = grad + err (ddp_init.py:155-157) is formed on the way:
// both operands arrive in the slot, the sum is stored to the send buffer and fed to the FMAs
// =======================================================================================
template <int R, bool FUSE>
__global__ void __launch_bounds__(kStreamThreads, 1)
k_p_from_m_stream(DevPlan pl, PtrSlot ps, StreamPlan sp, const float* __restrict__ q_mem, float* __restrict__ p_mem,
                  float q_div, PeerQ pq) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int ROWS = RowsS1<R>::v;
  constexpr int RG = (ROWS * R > 32) ? 32 / R : ROWS;  // rows per butterfly fold (at most 32 values)
  constexpr int UNR1 = (R <= 4) ? PSGD_K1_UNROLL_R4 : PSGD_K1_UNROLL_R8;  // column steps in flight per warp
  constexpr int NV = RG * R;
  const StreamSmem sm = carve(smem_raw, sp);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  stream_setup(pl, ps, sp, sm, q_mem);
  if (pq.pm.mode != kPeerOff) peer_q_prologue(pq);  // extra power iterations at N > 1: Q still has to be summed

  if (warp == kConsumerWarps) {
    if (lane == 0) stream_producer<ROWS, FUSE>(pl, ps, sp, sm, q_mem);
    return;
  }

  pdl_wait();
  if (sp.vec_mode) stream_vec_side_job(sp, ps);
  if (sp.n_direct) direct_p_from_m<R, FUSE>(pl, ps, sp, q_mem, p_mem, q_div);
  int cur_mat = -1, cur_chunk = 0, m = 0, ld = 0, col0 = 0, r = 0, pitch = 0;  // m: columns per pass (row length in a slot)
  unsigned pphase = 0;  // bit s: parity of the partial barrier of slot s
  float* P = nullptr;
  float* Mw = nullptr;  // FUSE: the send buffer M is written to
  const bool dbg = sp.dbg != nullptr && threadIdx.x == 0;
  unsigned long long d_t0 = 0, d_desc = 0, d_data = 0, d_q = 0, d_units = 0, d_sw = 0, d_work = 0;
  long long c0 = 0, c_begin = 0;
  if (dbg) { d_t0 = gtime_ns(); c_begin = clock64(); }
  for (int it = 0;; ++it) {
    const int slot = it % sp.n_slots;
    const uint32_t phase = (uint32_t)(it / sp.n_slots) & 1u;
    if (dbg) c0 = clock64();
    mbar_wait(&sm.desc[slot], phase);
    if (dbg) d_desc += clock64() - c0;
    const int u_mat = sm.slot_desc[slot].x, u_a = sm.slot_desc[slot].y;
    const int u_n = sm.slot_desc[slot].z, u_w = sm.slot_desc[slot].w;
    if (u_mat < 0) break;
    const float* base = reinterpret_cast<const float*>(sm.ring + (size_t)slot * sp.slot_bytes);
    if (u_n < 0) {  // Q unit (uniform over the consumer warps: every warp visits every unit)
      if (dbg) c0 = clock64();
      const StreamSmem::MatInfo mi = mat_info(pl, ps, sm, q_mem, u_mat);
      const int chunk = (u_w & kQFlagChunk1) ? 1 : 0;
      if (u_mat != cur_mat || chunk != cur_chunk) {
        P = p_mem + pl.mats[u_mat].p_off;  // in flight while the Q data lands
        Mw = const_cast<float*>(mi.base);
        ld = mi.m; m = ld / mi.nchunk; col0 = chunk * m; r = mi.r; pitch = q_pitch(m, R);
        cur_mat = u_mat;
        cur_chunk = chunk;
        consumer_sync();  // everyone is done with the previous Q
      }
      mbar_wait(&sm.full[slot], phase);
      consume_q_unit<R>(sm.qs, base, mi.q + (size_t)col0 * r, (u_w & kQFlagBulk) != 0, u_a, -u_n, m, r, pitch, q_div);
      consumer_sync();
      if (lane == 0) mbar_arrive(&sm.empty[slot]);
      if (dbg) { d_q += clock64() - c0; ++d_sw; }
      continue;
    }
    const int nsub = (u_n + ROWS - 1) / ROWS;
    const int kmax = m >> 2;
    // Work items of a unit.  Normally one item = ROWS rows x all columns (sub-span j belongs to warp (rot + j) %
    // kConsumerWarps).  A wide matrix puts only a few rows into a unit and would keep one or two warps busy, so there
    // the columns are cut into `cs` slices as well: the warps leave their partial sums in shared memory, arrive on the
    // slot's partial barrier WITHOUT waiting and move on; only the warp that owns the last slice of a row group waits
    // for the barrier, adds the slices up in slice order and writes P (deterministic, no full-CTA barrier per unit).
    int cs = 1;
    if (nsub <= sp.split_max_sub)
      while (nsub * cs < kConsumerWarps && (kmax >> 6) >= cs) cs <<= 1;  // a slice keeps >= 32 float4 columns
    const int nitems = nsub * cs;
    // Every warp arrives on every unit, with or without an item: that keeps any slot's barriers within one lap of
    // every warp (a warp that skipped arrivals could find a slot two laps ahead and mis-read the parity).
    const int j0 = (warp - (u_w & 0xff) % kConsumerWarps + kConsumerWarps) % kConsumerWarps;
    const bool accumulate = (u_w >> 8) != 0;  // second column half: add to what this lane stored in the first pass
    if (j0 >= nitems && cs == 1) {
      if (lane == 0) mbar_arrive(&sm.empty[slot]);
      continue;
    }
    if (dbg) c0 = clock64();
    if (j0 < nitems) mbar_wait(&sm.full[slot], phase);
    if (dbg) { d_data += clock64() - c0; ++d_units; c0 = clock64(); }
    const int kslice = (kmax + cs - 1) / cs;
    for (int j = j0; j < nitems; j += kConsumerWarps) {
      const int sub = j / cs;
      const int k_lo = (j - sub * cs) * kslice;
      const int k_hi = (k_lo + kslice < kmax) ? k_lo + kslice : kmax;
      const int rbase = sub * ROWS;
      const int nr = (u_n - rbase < ROWS) ? (u_n - rbase) : ROWS;
      const float* rowp[ROWS];
#pragma unroll
      for (int a = 0; a < ROWS; ++a) rowp[a] = base + (size_t)(rbase + (a < nr ? a : nr - 1)) * m;
      float acc[ROWS][R];
#pragma unroll
      for (int a = 0; a < ROWS; ++a)
#pragma unroll
        for (int c = 0; c < R; ++c) acc[a][c] = 0.f;
#pragma unroll(UNR1)
      for (int k = k_lo + lane; k < k_hi; k += 32) {
        float4 x[ROWS];
#pragma unroll
        for (int a = 0; a < ROWS; ++a) x[a] = *reinterpret_cast<const float4*>(rowp[a] + 4 * k);
        if (FUSE) {
#pragma unroll
          for (int a = 0; a < ROWS; ++a) {
            const float4 e = *reinterpret_cast<const float4*>(rowp[a] + (size_t)u_n * m + 4 * k);
            x[a].x += e.x; x[a].y += e.y; x[a].z += e.z; x[a].w += e.w;
            if (a < nr) st_stream4(Mw + (size_t)(u_a + rbase + a) * ld + col0 + 4 * k, x[a]);
          }
        }
#pragma unroll
        for (int c = 0; c < R; ++c) {
          const float4 q = *reinterpret_cast<const float4*>(sm.qs + c * pitch + 4 * k);
#pragma unroll
          for (int a = 0; a < ROWS; ++a) {
            acc[a][c] = fmaf(x[a].x, q.x, acc[a][c]);
            acc[a][c] = fmaf(x[a].y, q.y, acc[a][c]);
            acc[a][c] = fmaf(x[a].z, q.z, acc[a][c]);
            acc[a][c] = fmaf(x[a].w, q.w, acc[a][c]);
          }
        }
      }
#pragma unroll
      for (int g = 0; g < ROWS; g += RG) {
        float flat[NV];
#pragma unroll
        for (int a = 0; a < RG; ++a)
#pragma unroll
          for (int c = 0; c < R; ++c) flat[a * R + c] = acc[g + a][c];
        const float tot = warp_fold<NV>(flat, lane);
        constexpr int LPV = 32 / NV;
        if ((lane % LPV) == 0) {
          const int idx = lane / LPV;
          const int a = g + idx / R, c = idx % R;
          if (cs == 1) {
            if (a < nr && c < r) {
              float* dst = P + (size_t)(u_a + rbase + a) * r + c;
              if (accumulate) *dst = *dst + tot; else st_out(dst, tot, sp.out_mc != 0);
            }
          } else {
            sm.part[(slot * kSplitItems + j) * 32 + g * R + idx] = tot;
          }
        }
      }
    }
    if (cs > 1) {
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.pbar[slot]);
      for (int sub = 0; sub < nsub; ++sub) {
        if (warp != ((u_w & 0xff) + sub * cs + cs - 1) % kConsumerWarps) continue;
        mbar_wait(&sm.pbar[slot], (pphase >> slot) & 1u);
        if (lane < ROWS * R) {
          const float* pp = sm.part + (slot * kSplitItems + sub * cs) * 32 + lane;
          float t = pp[0];
          for (int sl = 1; sl < cs; ++sl) t += pp[sl * 32];
          const int a = lane / R, c = lane % R;
          if (sub * ROWS + a < u_n && c < r) {
            float* dst = P + (size_t)(u_a + sub * ROWS + a) * r + c;
            if (accumulate) *dst = *dst + t; else st_out(dst, t, sp.out_mc != 0);
          }
        }
      }
      pphase ^= 1u << slot;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.empty[slot]);
    if (dbg) d_work += clock64() - c0;
  }
  if (dbg) {
    unsigned long long* o = sp.dbg + (size_t)blockIdx.x * kDbgWords;
    o[0] = d_t0; o[1] = gtime_ns(); o[2] = d_desc; o[3] = d_data; o[4] = d_q;
    o[5] = (unsigned long long)(clock64() - c_begin); o[6] = d_units; o[7] = d_work;
  }
}

// =======================================================================================
// K4 + K5 (streamed):  out = P (Q / n_workers)^T ; mem = M - out
// =======================================================================================
template <int R>
__global__ void __launch_bounds__(kStreamThreads, 1)
k_finalize_stream(DevPlan pl, PtrSlot ps, StreamPlan sp, const float* __restrict__ p_mem,
                  const float* __restrict__ q_mem, float world, PeerQ pq) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int ROWS = RowsS4<R>::v;
  const StreamSmem sm = carve(smem_raw, sp);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  stream_setup(pl, ps, sp, sm, q_mem);
  if (pq.pm.mode != kPeerOff) peer_q_prologue(pq);

  if (warp == kConsumerWarps) {
    if (lane == 0) stream_producer<ROWS>(pl, ps, sp, sm, q_mem);
    return;
  }

  pdl_wait();
  if (sp.vec_mode) stream_vec_side_job(sp, ps);
  if (sp.n_direct) direct_finalize<R>(pl, ps, sp, p_mem, q_mem, world);
  int cur_mat = -1, cur_chunk = 0, m = 0, ld = 0, col0 = 0, r = 0, pitch = 0;
  const float* P = nullptr;
  float* Out = nullptr;
  float* Mem = nullptr;
  for (int it = 0;; ++it) {
    const int slot = it % sp.n_slots;
    const uint32_t phase = (uint32_t)(it / sp.n_slots) & 1u;
    mbar_wait(&sm.desc[slot], phase);
    const int u_mat = sm.slot_desc[slot].x, u_a = sm.slot_desc[slot].y;
    const int u_n = sm.slot_desc[slot].z, u_w = sm.slot_desc[slot].w;
    if (u_mat < 0) break;
    const float* base = reinterpret_cast<const float*>(sm.ring + (size_t)slot * sp.slot_bytes);
    if (u_n < 0) {  // Q unit
      const StreamSmem::MatInfo mi = mat_info(pl, ps, sm, q_mem, u_mat);
      const int chunk = (u_w & kQFlagChunk1) ? 1 : 0;
      if (u_mat != cur_mat || chunk != cur_chunk) {
        P = p_mem + pl.mats[u_mat].p_off;  // these three loads are in flight while the Q data lands
        Out = ps.out[u_mat];
        Mem = ps.mem[u_mat];
        ld = mi.m; m = ld / mi.nchunk; col0 = chunk * m; r = mi.r; pitch = q_pitch(m, R);
        cur_mat = u_mat;
        cur_chunk = chunk;
        consumer_sync();
      }
      mbar_wait(&sm.full[slot], phase);
      consume_q_unit<R>(sm.qs, base, mi.q + (size_t)col0 * r, (u_w & kQFlagBulk) != 0, u_a, -u_n, m, r, pitch, world);
      consumer_sync();
      if (lane == 0) mbar_arrive(&sm.empty[slot]);
      continue;
    }
    // Work items of a unit = (group of ROWS rows) x (column slice).  Wide matrices give few rows per unit, so
    // the rows alone would keep one or two warps busy; this kernel has no reduction, so the columns are simply
    // cut into `cs` slices until all eight warps have an item.
    const int nsub = (u_n + ROWS - 1) / ROWS;
    int cs = 1;
    while (nsub * cs < kConsumerWarps && cs < 8) cs <<= 1;
    const int nitems = nsub * cs;
    const int j0 = (warp - (u_w & 0xff) % kConsumerWarps + kConsumerWarps) % kConsumerWarps;
    if (j0 >= nitems) {  // see k_p_from_m_stream
      if (lane == 0) mbar_arrive(&sm.empty[slot]);
      continue;
    }
    const int kmax = m >> 2;
    const int kslice = (kmax + cs - 1) / cs;
    bool landed = false;
    for (int j = j0; j < nitems; j += kConsumerWarps) {
      const int rbase = (j / cs) * ROWS;
      const int k_lo = (j % cs) * kslice;
      const int k_hi = (k_lo + kslice < kmax) ? k_lo + kslice : kmax;
      const int nr = (u_n - rbase < ROWS) ? (u_n - rbase) : ROWS;
      float p[ROWS][R];  // the P rows do not live in the slot: request them before waiting for M
#pragma unroll
      for (int a = 0; a < ROWS; ++a) {
        const int rr = u_a + rbase + (a < nr ? a : nr - 1);
#pragma unroll
        for (int c = 0; c < R; ++c) p[a][c] = (c < r) ? __ldg(P + (size_t)rr * r + c) : 0.f;
      }
      if (!landed) {
        mbar_wait(&sm.full[slot], phase);
        landed = true;
      }
#pragma unroll 2
      for (int k = k_lo + lane; k < k_hi; k += 32) {
        float4 q[R];
#pragma unroll
        for (int c = 0; c < R; ++c) q[c] = *reinterpret_cast<const float4*>(sm.qs + c * pitch + 4 * k);
#pragma unroll
        for (int a = 0; a < ROWS; ++a) {
          if (a < nr) {
            const float4 x = *reinterpret_cast<const float4*>(base + (size_t)(rbase + a) * m + 4 * k);
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int c = 0; c < R; ++c) {
              o.x = fmaf(p[a][c], q[c].x, o.x);
              o.y = fmaf(p[a][c], q[c].y, o.y);
              o.z = fmaf(p[a][c], q[c].z, o.z);
              o.w = fmaf(p[a][c], q[c].w, o.w);
            }
            const size_t g = (size_t)(u_a + rbase + a) * ld + col0 + 4 * k;
            st_stream4(Out + g, o);
            st_stream4(Mem + g, make_float4(x.x - o.x, x.y - o.y, x.z - o.z, x.w - o.w));
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.empty[slot]);
  }
}

}  // namespace psgd
