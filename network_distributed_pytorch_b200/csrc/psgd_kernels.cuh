// Hand-written sm_100a kernels of the PowerSGD reducer hot path.
//
// One launch per phase over a device-side table of all gradient matrices ("grouped
// launch"): blockIdx.x selects a tile, the tile names its matrix.  All contractions have
// rank r <= 32 and are HBM-bound, so the kernels are plain FFMA streaming kernels:
// 128-bit coalesced loads of M, Q staged (transposed) in shared memory, warp-shuffle
// reductions.  Citations are to ddp_powersgd_guide_cifar10/reducer.py upstream.
#pragma once
#include <cooperative_groups.h>

#include "psgd_common.cuh"
#include "psgd_tc.cuh"

namespace psgd {
namespace cg = cooperative_groups;

// rows a warp carries at once in K1 / K4 for compile-time rank R
#ifndef PSGD_K3_THREADS
#define PSGD_K3_THREADS 256
#endif
#ifndef PSGD_K3_MINBLK
#define PSGD_K3_MINBLK 3
#endif
constexpr int kThreads3 = PSGD_K3_THREADS;   // CTA size of K3
constexpr int kMinBlocks3 = PSGD_K3_MINBLK;  // K3 CTAs per SM the register budget is set for (R <= 8)
template <int R> struct RowsK1 { static constexpr int v = (R <= 2) ? 8 : (R <= 8) ? 4 : (R == 16) ? 2 : 1; };
template <int R> struct RowsK4 { static constexpr int v = (R <= 8) ? 4 : (R == 16) ? 2 : 1; };

// =======================================================================================
// K1:  P = M Q   (reducer.py:121-123), optionally M = grad + err first (ddp_init.py:155-157)
// Tile = `nrows` full rows of one matrix (a contiguous span of M).  Each warp carries ROWS
// rows at a time; lanes stride over the columns in 128-bit steps, so a warp reads 512
// contiguous bytes per row per step.  Q values come from the transposed shared copy and
// are reused for the ROWS rows.  Partial sums are folded across the warp with shuffles.
// =======================================================================================
template <int R, int ROWS, int V, bool FUSE>
__device__ __forceinline__ void p_tile(const MatDesc& md, const float* __restrict__ M,
                                       const float* __restrict__ G, const float* __restrict__ E,
                                       float* __restrict__ Mw, const float* __restrict__ Q,
                                       float* __restrict__ P, int row0, int nrows, int wc_cap,
                                       float q_div, float* Qs) {
  constexpr int NV = R * ROWS;
  const int m = md.m, r = md.r;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nsets = (nrows + ROWS - 1) / ROWS;
  const int nchunks = (m + wc_cap - 1) / wc_cap;
  const int pitch = q_pitch(m < wc_cap ? m : wc_cap, R);

  // Column chunks outermost: a chunk of Q is staged once per tile and serves all of its row sets.  With more than
  // one chunk a row's partial sums are added up in P itself, chunk after chunk, by the lane that owns the element
  // (same warp, same lane every time: program order, no race, fixed order).
  for (int ch = 0; ch < nchunks; ++ch) {
    const int c0 = ch * wc_cap;
    const int wc = (m - c0 < wc_cap) ? (m - c0) : wc_cap;
    if (ch) __syncthreads();
    stage_q<R>(Qs, Q, c0, wc, r, pitch, q_div);
    __syncthreads();
    const int kmax = wc / V;
    for (int set = warp; set < nsets; set += kWarps) {
      const int rbase = set * ROWS;  // tile-relative
      float acc[ROWS][R];
#pragma unroll
      for (int a = 0; a < ROWS; ++a)
#pragma unroll
        for (int c = 0; c < R; ++c) acc[a][c] = 0.f;
      size_t roff[ROWS];
#pragma unroll
      for (int a = 0; a < ROWS; ++a) {
        int rr = rbase + a;
        if (rr > nrows - 1) rr = nrows - 1;  // clamp: loads stay in bounds, result discarded
        roff[a] = (size_t)(row0 + rr) * m + c0;
      }
      for (int k = lane; k < kmax; k += 64) {
        const bool two = (k + 32) < kmax;
        const int k1 = two ? k + 32 : k;
        Pack<V> a0[ROWS], a1[ROWS];
#pragma unroll
        for (int a = 0; a < ROWS; ++a) {
          if (FUSE) {
            Pack<V> g = pk_ld_stream<V>(G + roff[a] + (size_t)k * V);
            Pack<V> e = pk_ld_stream<V>(E + roff[a] + (size_t)k * V);
#pragma unroll
            for (int x = 0; x < V; ++x) a0[a].v[x] = g.v[x] + e.v[x];
          } else {
            a0[a] = pk_ld_stream<V>(M + roff[a] + (size_t)k * V);
          }
        }
#pragma unroll
        for (int a = 0; a < ROWS; ++a) {
          if (two) {
            if (FUSE) {
              Pack<V> g = pk_ld_stream<V>(G + roff[a] + (size_t)k1 * V);
              Pack<V> e = pk_ld_stream<V>(E + roff[a] + (size_t)k1 * V);
#pragma unroll
              for (int x = 0; x < V; ++x) a1[a].v[x] = g.v[x] + e.v[x];
            } else {
              a1[a] = pk_ld_stream<V>(M + roff[a] + (size_t)k1 * V);
            }
          } else {
            a1[a] = pk_zero<V>();
          }
        }
        if (FUSE) {
#pragma unroll
          for (int a = 0; a < ROWS; ++a) {
            if (rbase + a < nrows) {
              pk_st_stream<V>(Mw + roff[a] + (size_t)k * V, a0[a]);
              if (two) pk_st_stream<V>(Mw + roff[a] + (size_t)k1 * V, a1[a]);
            }
          }
        }
#pragma unroll
        for (int c = 0; c < R; ++c) {
          const Pack<V> q0 = pk_ld_shared<V>(Qs + c * pitch + k * V);
          const Pack<V> q1 = pk_ld_shared<V>(Qs + c * pitch + k1 * V);
#pragma unroll
          for (int a = 0; a < ROWS; ++a) {
#pragma unroll
            for (int x = 0; x < V; ++x) {
              acc[a][c] = fmaf(a0[a].v[x], q0.v[x], acc[a][c]);
              acc[a][c] = fmaf(a1[a].v[x], q1.v[x], acc[a][c]);
            }
          }
        }
      }
      float flat[NV];
#pragma unroll
      for (int a = 0; a < ROWS; ++a)
#pragma unroll
        for (int c = 0; c < R; ++c) flat[a * R + c] = acc[a][c];
      const float tot = warp_fold<NV>(flat, lane);
      constexpr int LPV = 32 / NV;  // lanes per value
      if ((lane % LPV) == 0) {
        const int idx = lane / LPV;
        const int a = idx / R, c = idx % R;
        if (rbase + a < nrows && c < r) {
          float* dst = P + (size_t)(row0 + rbase + a) * r + c;
          *dst = ch ? *dst + tot : tot;
        }
      }
    }
  }
}

template <int R, bool FUSE>
__global__ void __launch_bounds__(kThreads, 2)
k_p_from_m(DevPlan pl, PtrSlot ps, const float* __restrict__ q_mem, float* __restrict__ p_mem, float q_div) {
  extern __shared__ __align__(16) float smem[];
  const int4 t = pl.tiles1[blockIdx.x];
  const MatDesc md = pl.mats[t.x];
  const unsigned fl = ps.flags[t.x];
  const float* M = ps.in[t.x];
  const float* G = FUSE ? ps.g[t.x] : nullptr;
  const float* E = FUSE ? ps.e[t.x] : nullptr;
  float* Mw = FUSE ? const_cast<float*>(M) : nullptr;
  const float* Q = q_mem + md.q_off;
  float* P = p_mem + md.p_off;
  const bool vec = (fl & kFlagInVec) && (!FUSE || (fl & kFlagGeVec));
  if (vec)
    p_tile<R, RowsK1<R>::v, 4, FUSE>(md, M, G, E, Mw, Q, P, t.y, t.z, pl.wc_cap, q_div, smem);
  else
    p_tile<R, RowsK1<R>::v, 1, FUSE>(md, M, G, E, Mw, Q, P, t.y, t.z, pl.wc_cap, q_div, smem);
}

// =======================================================================================
// K2:  p_mem /= n_workers ; modified Gram-Schmidt of every P_i  (reducer.py:128,136-137,180-191)
// One thread-block cluster per matrix; each CTA owns a contiguous slice of the rows and
// every thread always touches the same rows, so P stays coherent without fences.  Step i
// needs ONE cluster-wide reduction: ||col_i||^2 and the dots col_i . col_j (j > i) are
// summed together, the projection coefficient is then dot / (norm + eps).  Column order and
// normalise-then-project order are the reference's.  Partials travel through distributed
// shared memory and are summed in a fixed order, so the result is deterministic.
// =======================================================================================
template <int R>
__device__ __forceinline__ void orth_body(float* __restrict__ P, int n, int r, float world,
                                          int crank, int csize, bool use_cluster) {
  __shared__ float wsum[kOrthThreads / 32][R];
  __shared__ float part[2][kOrthCluster][R];
  __shared__ float tot_s[R];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nthr = blockDim.x, nwarp = nthr >> 5;
  const int per = (n + csize - 1) / csize;
  const int lo = crank * per;
  const int hi = (lo + per < n) ? lo + per : n;
  const bool divide = world != 1.0f;
  const float inv_world = 1.0f / world;

  for (int i = 0; i < r; ++i) {
    float s[R];
#pragma unroll
    for (int j = 0; j < R; ++j) s[j] = 0.f;
    for (int row = lo + tid; row < hi; row += nthr) {
      float* pr = P + (size_t)row * r;
      float v[R];
#pragma unroll
      for (int j = 0; j < R; ++j) {
        v[j] = 0.f;
        if (j >= i && j < r) {
          v[j] = pr[j];
          if (i == 0 && divide) {
            v[j] = v[j] * inv_world;
            pr[j] = v[j];
          }
        }
      }
      float vi = 0.f;
#pragma unroll
      for (int j = 0; j < R; ++j)
        if (j == i) vi = v[j];
#pragma unroll
      for (int j = 0; j < R; ++j) s[j] = fmaf(vi, v[j], s[j]);
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (j >= i && j < r) {
        const float w = warp_sum(s[j]);
        if (lane == 0) wsum[warp][j] = w;
      }
    }
    __syncthreads();
    const int par = i & 1;
    if (use_cluster) {
      if (tid < csize * R) {
        const int dst = tid / R, j = tid % R;
        if (j >= i && j < r) {
          float t = 0.f;
          for (int w = 0; w < nwarp; ++w) t += wsum[w][j];
          cg::cluster_group cluster = cg::this_cluster();
          float* remote = cluster.map_shared_rank(&part[0][0][0], dst);
          remote[(par * kOrthCluster + crank) * R + j] = t;
        }
      }
      cg::this_cluster().sync();
    } else {
      if (tid < R) {
        const int j = tid;
        if (j >= i && j < r) {
          float t = 0.f;
          for (int w = 0; w < nwarp; ++w) t += wsum[w][j];
          part[par][0][j] = t;
        }
      }
      __syncthreads();
    }
    float tot[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      tot[j] = 0.f;
      if (j >= i && j < r)
        for (int c = 0; c < csize; ++c) tot[j] += part[par][c][j];
    }
    float nrm2 = 0.f;
#pragma unroll
    for (int j = 0; j < R; ++j)
      if (j == i) nrm2 = tot[j];
    const float inv = 1.0f / (sqrtf(nrm2) + 1e-8f);  // eps outside the root, reducer.py:186
    float coef[R];
#pragma unroll
    for (int j = 0; j < R; ++j) coef[j] = (j > i && j < r) ? tot[j] * inv : 0.f;

    for (int row = lo + tid; row < hi; row += nthr) {
      float* pr = P + (size_t)row * r;
      const float vi = pr[i] * inv;
      pr[i] = vi;
#pragma unroll
      for (int j = 0; j < R; ++j)
        if (j > i && j < r) pr[j] = fmaf(-coef[j], vi, pr[j]);
    }
    (void)tot_s;
  }
}

#ifdef PSGD_ORTH_TS
__device__ unsigned long long g_orth_ts[64][12];
__device__ __forceinline__ void orth_ts(int slot) {
  if (threadIdx.x == 0 && blockIdx.x < 64 && slot < 12) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_orth_ts[blockIdx.x][slot] = t;
  }
}
#define ORTH_TS(s) orth_ts(s)
#else
#define ORTH_TS(s)
#endif
// The step's R sums over the csize partials in `part` ([c][R]), for every thread.  One CTA: read them.  A cluster: the
// warp shares the work -- lane l adds the partials c = l / R, l / R + 32 / R, ... of value l % R, xor-shuffles join the
// 32 / R lanes of a value, R more hand every total to every lane -- instead of csize * R additions in every thread
// (at R = 16 and 16 CTAs those were half of a step's instructions).  Same order in every warp of every CTA.
template <int R>
__device__ __forceinline__ void orth_totals(const float* part, int csize, int lane, float (&tot)[R]) {
  if (csize == 1) {
    if (R % 4 == 0) {
#pragma unroll
      for (int j = 0; j < R; j += 4) {
        const float4 q = *reinterpret_cast<const float4*>(part + j);
        tot[j] = q.x; tot[(j + 1) % R] = q.y; tot[(j + 2) % R] = q.z; tot[(j + 3) % R] = q.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < R; ++j) tot[j] = part[j];
    }
    return;
  }
  constexpr int LPV = 32 / R;
  const int j = lane % R;
  float x = 0.f;
  for (int c = lane / R; c < csize; c += LPV) x += part[c * R + j];
#pragma unroll
  for (int o = R; o < 32; o <<= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
#pragma unroll
  for (int k = 0; k < R; ++k) tot[k] = __shfl_sync(0xffffffffu, x, k);
}

// Register-resident variant: every thread keeps its (at most RPT) rows in registers for all r steps,
// so P is read once and written once and a step costs one block (+ cluster) reduction.  csize == 1
// runs a whole matrix in one CTA with block barriers only; csize > 1 spreads one tall matrix over a
// thread-block cluster and exchanges the partials through distributed shared memory.  All sums are
// formed in a fixed order (xor-shuffle trees, then CTA rank order), so results are deterministic.
//   wsum: 32 * R floats, part: 2 * kOrthClusterMax * R floats, both 16-byte aligned shared memory
//   cbar (csize > 1): two mbarriers (count 1), initialised and made known to the cluster by the caller (orth_cluster_init).
// The cluster exchange of a step is one-way: warp 0 of every CTA sends its partial sums into every CTA's `part`
// with st.async, which also counts the bytes on the receiver's mbarrier; a CTA goes on as soon as its own barrier has
// seen all csize partials.  (A cluster-wide barrier after plain remote stores cost 1.05 us per step, twice the step
// itself.)  `part` and the barriers alternate between two sets: a CTA that is one step ahead writes the other set, and
// it cannot be two ahead because it needs this CTA's next partial first.
// The csize cooperating CTAs are the cluster ranks cbase .. cbase + csize - 1 of the hardware cluster (crank counts
// from cbase): groups smaller than the hardware cluster share one, each exchanging only among its own CTAs.
// pm != nullptr: the rows are read "summed over all ranks" from symmetric memory (element offset sym_off of the
// buffer pm describes: the SUM all-reduce of reducer.py:126 happens on load), the result goes to the local P.
template <int R, int RPT>
__device__ __forceinline__ void orth_cached(float* __restrict__ P, int n, int r, float world,
                                            int crank, int csize, float* wsum, float* part, uint64_t* cbar,
                                            const PeerLoad* pm = nullptr, long long sym_off = 0, int cbase = 0) {
  constexpr int LPV = 32 / R;  // lanes per value after a warp_fold<R>
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int per = (n + csize - 1) / csize;
  // Only as many warps as the rows need stay: a step is bound by the instructions the CTA issues (fold, barrier, scale
  // factors: the same for every warp, with or without rows), not by its arithmetic.  The others leave before the first
  // barrier; block and cluster barriers count the threads that have not exited.
  const int want = (((per + RPT - 1) / RPT) + 31) & ~31;
  const int nthr = want < (int)blockDim.x ? want : (int)blockDim.x, nwarp = nthr >> 5;
  if (tid >= nthr) return;
  const int lo = crank * per;
  const int hi = (lo + per < n) ? lo + per : n;
  auto cta_sync = [] { __syncthreads(); };

  // IEEE fp32 division costs ~40 instructions; one reciprocal per scale factor and multiplies instead
  // (exact for power-of-two worker counts, otherwise within 1 ulp of the reference's true division)
  const bool vec4 = (R % 4 == 0) && (r == R) && ((reinterpret_cast<uintptr_t>(P) & 15) == 0) && ((sym_off & 3) == 0);
  float v[RPT][R];
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int row = lo + tid + k * nthr;
    if (vec4) {
#pragma unroll
      for (int c = 0; c < R; c += 4) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < hi) {
          if (pm) t = peer_sum4(*pm, sym_off + (long long)row * R + c);
          else t = __ldcg(reinterpret_cast<const float4*>(P + (size_t)row * R + c));
        }
        v[k][c] = t.x; v[k][(c + 1) % R] = t.y; v[k][(c + 2) % R] = t.z; v[k][(c + 3) % R] = t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < R; ++j) {
        v[k][j] = 0.f;
        if (row < hi && j < r) v[k][j] = pm ? peer_sum1(*pm, sym_off + (long long)row * r + j) : __ldcg(P + (size_t)row * r + j);
      }
    }
  }
  if (world != 1.0f) {
    const float inv_world = 1.0f / world;
#pragma unroll
    for (int k = 0; k < RPT; ++k)
#pragma unroll
      for (int j = 0; j < R; ++j) v[k][j] *= inv_world;
  }
  ORTH_TS(2);

  // For R <= 8 the step loop stays ROLLED: unrolled, the R steps are R copies of straight-line code that every warp runs exactly
  // once, and the kernel spent most of its time waiting for instruction fetches (ncu: stall_no_inst 42 % at R = 8,
  // 160 KB of SASS).  To keep the rows in registers under a runtime step index the columns rotate instead: the column
  // being normalised is always register 0, the unfinished ones follow, finished ones queue up at the back, and after R
  // rotations everything is back in place.
  if constexpr (R <= 16) {
#pragma unroll 1
  for (int i = 0; i < R; ++i) {
    if (i < r) {  // uniform over the CTA / cluster
      // partial sums of col_i . col_j for all register positions j; one butterfly fold reduces all R of them
      float s[R];
#pragma unroll
      for (int j = 0; j < R; ++j) s[j] = 0.f;
#pragma unroll
      for (int k = 0; k < RPT; ++k)
#pragma unroll
        for (int j = 0; j < R; ++j) s[j] = fmaf(v[k][0], v[k][j], s[j]);
      {
        const float w = warp_fold<R>(s, lane);
        if ((lane % LPV) == 0) wsum[warp * R + lane / LPV] = w;
      }
      cta_sync();
      const int par = i & 1;
      if (warp == 0) {
        float t[R];
#pragma unroll
        for (int j = 0; j < R; ++j) t[j] = (lane < nwarp) ? wsum[lane * R + j] : 0.f;
        const float w = warp_fold<R>(t, lane);
        if (csize > 1) {
          // one message per (value, destination CTA): it lands in the receiver's `part` and counts on its mbarrier.
          // All LPV lanes of a group hold the group's value after the fold, so they share out the destinations.
          const int j = lane / LPV;
          const uint32_t slot = smem_u32(part + (par * kOrthClusterMax + crank) * R + j), bar = smem_u32(&cbar[par]);
          for (int dst = lane % LPV; dst < csize; dst += LPV)
            st_async_f32(cluster_map_u32(slot, cbase + dst), w, cluster_map_u32(bar, cbase + dst));
          if (lane == 0) mbar_arrive_expect_tx(&cbar[par], (uint32_t)csize * R * 4u);
        } else if ((lane % LPV) == 0) {
          part[(par * kOrthClusterMax) * R + lane / LPV] = w;
        }
      }
      if (csize > 1) mbar_wait_lite(&cbar[par], (uint32_t)(i >> 1) & 1u); else cta_sync();
      float tot[R];
      orth_totals<R>(part + par * kOrthClusterMax * R, csize, lane, tot);
      const float inv = 1.0f / (sqrtf(tot[0]) + 1e-8f);  // eps outside the root, reducer.py:186
      const int live = R - i;  // positions 1 .. live-1 hold the columns still to be projected
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        const float vi = v[k][0] * inv;
        v[k][0] = vi;
#pragma unroll
        for (int j = 1; j < R; ++j) {
          const float upd = fmaf(-(tot[j] * inv), vi, v[k][j]);
          v[k][j] = (j < live) ? upd : v[k][j];
        }
      }
    }
    ORTH_TS(3 + i);
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const float head = v[k][0];
#pragma unroll
      for (int j = 0; j + 1 < R; ++j) v[k][j] = v[k][j + 1];
      v[k][R - 1] = head;
    }
  }
  } else {
  // larger R: unrolled steps
#pragma unroll
  for (int i = 0; i < R; ++i) {
    if (i < r) {  // uniform over the CTA / cluster
      // partial sums of col_i . col_j for j >= i (j < i stays 0); one butterfly fold reduces all R of them
      float s[R];
#pragma unroll
      for (int j = 0; j < R; ++j) s[j] = 0.f;
#pragma unroll
      for (int k = 0; k < RPT; ++k)
#pragma unroll
        for (int j = i; j < R; ++j) s[j] = fmaf(v[k][i], v[k][j], s[j]);
      {
        const float w = warp_fold<R>(s, lane);
        if ((lane % LPV) == 0) wsum[warp * R + lane / LPV] = w;
      }
      cta_sync();
      const int par = i & 1;
      if (warp == 0) {
        float t[R];
#pragma unroll
        for (int j = 0; j < R; ++j) t[j] = (lane < nwarp) ? wsum[lane * R + j] : 0.f;
        const float w = warp_fold<R>(t, lane);
        if (csize > 1) {
          // one message per (value, destination CTA): it lands in the receiver's `part` and counts on its mbarrier.
          // All LPV lanes of a group hold the group's value after the fold, so they share out the destinations.
          const int j = lane / LPV;
          const uint32_t slot = smem_u32(part + (par * kOrthClusterMax + crank) * R + j), bar = smem_u32(&cbar[par]);
          for (int dst = lane % LPV; dst < csize; dst += LPV)
            st_async_f32(cluster_map_u32(slot, cbase + dst), w, cluster_map_u32(bar, cbase + dst));
          if (lane == 0) mbar_arrive_expect_tx(&cbar[par], (uint32_t)csize * R * 4u);
        } else if ((lane % LPV) == 0) {
          part[(par * kOrthClusterMax) * R + lane / LPV] = w;
        }
      }
      if (csize > 1) mbar_wait_lite(&cbar[par], (uint32_t)(i >> 1) & 1u); else cta_sync();
      float tot[R];
      orth_totals<R>(part + par * kOrthClusterMax * R, csize, lane, tot);
      const float inv = 1.0f / (sqrtf(tot[i]) + 1e-8f);  // eps outside the root, reducer.py:186
      float coef[R];
#pragma unroll
      for (int j = i + 1; j < R; ++j) coef[j] = tot[j] * inv;
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        const float vi = v[k][i] * inv;
        v[k][i] = vi;
#pragma unroll
        for (int j = i + 1; j < R; ++j) v[k][j] = fmaf(-coef[j], vi, v[k][j]);
      }
    }
  }
  }
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int row = lo + tid + k * nthr;
    if (row < hi) {
      if (vec4) {
#pragma unroll
        for (int c = 0; c < R; c += 4)
          *reinterpret_cast<float4*>(P + (size_t)row * R + c) =
              make_float4(v[k][c], v[k][(c + 1) % R], v[k][(c + 2) % R], v[k][(c + 3) % R]);
      } else {
#pragma unroll
        for (int j = 0; j < R; ++j)
          if (j < r) P[(size_t)row * r + j] = v[k][j];
      }
    }
  }
  ORTH_TS(11);
}

// the two exchange barriers of a cooperating cluster: initialised, then made known to all its CTAs before any of them
// can send (all threads of all CTAs of the cluster call this; it does not depend on an earlier kernel)
__device__ __forceinline__ void orth_cluster_init(uint64_t* cbar) {
  if (threadIdx.x == 0) {
    mbar_init(&cbar[0], 1);
    mbar_init(&cbar[1], 1);
    mbar_fence_init();
  }
  cg::this_cluster().sync();
}

// one CTA per "small" matrix (n <= blockDim * RPT)
template <int R, int RPT>
__global__ void __launch_bounds__(orth_threads(R))
k_orth_small(DevPlan pl, const int* __restrict__ ids, float* __restrict__ p_mem, float world) {
  __shared__ __align__(16) float wsum[32 * R];
  __shared__ __align__(16) float part[2 * kOrthClusterMax * R];
  pdl_launch_dependents();
  const int mat = ids[blockIdx.x];
  const MatDesc md = pl.mats[mat];
  pdl_wait();
  orth_cached<R, RPT>(p_mem + md.p_off, md.n, md.r, world, 0, 1, wsum, part, nullptr);
}

// The flat <=1-D buffer travels behind P in the same symmetric buffer (reducer.py:131-133): CTAs past the matrices sum
// it over the ranks, form the mean (reducer.py:167) and leave it in the local flat buffer.
struct PeerFlat {
  float* out;          // local flat buffer (mean over ranks); nullptr: nothing to do
  long long sym_off;   // element offset of the flat part in the symmetric buffer (= sum n*r)
  long long n;         // elements
  float div;           // n_workers
};
__host__ __device__ constexpr int peer_flat_per_cta(int R) { return orth_threads(R) * 4; }

__device__ __forceinline__ void peer_flat_job(const PeerLoad& pm, const PeerFlat& pf, int idx) {
  const long long e = ((long long)idx * blockDim.x + threadIdx.x) * 4;
  if (e >= pf.n) return;
  if (e + 4 <= pf.n && ((pf.sym_off & 3) == 0) && ((reinterpret_cast<uintptr_t>(pf.out) & 15) == 0)) {
    float4 v = peer_sum4(pm, pf.sym_off + e);
    v.x = v.x / pf.div; v.y = v.y / pf.div; v.z = v.z / pf.div; v.w = v.w / pf.div;
    *reinterpret_cast<float4*>(pf.out + e) = v;
  } else {
    for (long long k = e; k < e + 4 && k < pf.n; ++k) pf.out[k] = peer_sum1(pm, pf.sym_off + k) / pf.div;
  }
}

// K2 with the P all-reduce on load (reducer.py:126-128 + 136-137 in one kernel): after its dependency wait the
// grid tells every peer "my P is written" and waits for theirs (one flag barrier, no extra launch, no copy), then
// every CTA reads its rows summed over the ranks.  One CTA per small matrix.
template <int R, int RPT>
__global__ void __launch_bounds__(orth_threads(R))
k_orth_small_peer(DevPlan pl, const int* __restrict__ ids, int n_small, float* __restrict__ p_mem, float world,
                  PeerLoad pm, PeerFlat pf) {
  __shared__ __align__(16) float wsum[32 * R];
  __shared__ __align__(16) float part[2 * kOrthClusterMax * R];
  pdl_launch_dependents();
  pdl_wait();
  peer_barrier(pm, blockIdx.x == 0);
  if ((int)blockIdx.x < n_small) {
    const int mat = ids[blockIdx.x];
    const MatDesc md = pl.mats[mat];
    orth_cached<R, RPT>(p_mem + md.p_off, md.n, md.r, world, 0, 1, wsum, part, nullptr, &pm, md.p_off);
  } else if (pf.out) {
    peer_flat_job(pm, pf, (int)blockIdx.x - n_small);
  }
}

// Matrices too tall for one CTA, each on a group of 2, 4, 8 or 16 CTAs (n <= group * blockDim * RPT) of a 16-CTA
// hardware cluster, per the plan's job table {matrix, first cluster rank of the group, CTAs in the group, rank in the
// group}.  When the plan also has small matrices and everything fits one wave, they ride on the same launch as groups
// of one.  CTAs past the table have nothing to do.
template <int R, int RPT>
__global__ void __launch_bounds__(orth_threads(R))
k_orth_tall(DevPlan pl, const int4* __restrict__ jobs, int n_jobs, float* __restrict__ p_mem, float world) {
  __shared__ __align__(16) float wsum[32 * R];
  __shared__ __align__(16) float part[2 * kOrthClusterMax * R];
  __shared__ uint64_t cbar[2];
  ORTH_TS(0);
  pdl_launch_dependents();
  orth_cluster_init(cbar);  // every CTA of every cluster; under the predecessor's tail
  if ((int)blockIdx.x >= n_jobs) {
    pdl_wait();  // every CTA waits: the grid must not be able to complete before its predecessor has
    return;
  }
  const int4 job = jobs[blockIdx.x];
  const MatDesc md = pl.mats[job.x];
  pdl_wait();
  ORTH_TS(1);
  orth_cached<R, RPT>(p_mem + md.p_off, md.n, md.r, world, job.w, job.z, wsum, part, cbar, nullptr, 0, job.y);
}

// the same with the P all-reduce on load; CTAs past the job table sum the flat buffer (see k_orth_small_peer)
template <int R, int RPT>
__global__ void __launch_bounds__(orth_threads(R))
k_orth_tall_peer(DevPlan pl, const int4* __restrict__ jobs, int n_jobs, float* __restrict__ p_mem, float world,
                 PeerLoad pm, PeerFlat pf) {
  __shared__ __align__(16) float wsum[32 * R];
  __shared__ __align__(16) float part[2 * kOrthClusterMax * R];
  __shared__ uint64_t cbar[2];
  pdl_launch_dependents();
  orth_cluster_init(cbar);  // every CTA of every cluster; under the predecessor's tail
  const bool has_job = (int)blockIdx.x < n_jobs;
  const int4 job = has_job ? jobs[blockIdx.x] : make_int4(0, 0, 1, 0);
  const MatDesc md = pl.mats[job.x];  // plan constants: not behind the dependency wait
  pdl_wait();
  peer_barrier(pm, blockIdx.x == 0);
  if (has_job) {
    orth_cached<R, RPT>(p_mem + md.p_off, md.n, md.r, world, job.w, job.z, wsum, part, cbar, &pm, md.p_off, job.y);
  } else if (pf.out) {
    peer_flat_job(pm, pf, (int)blockIdx.x - n_jobs);
  }
}

template <int R>
__global__ void __launch_bounds__(kOrthThreads)
k_orthogonalize(DevPlan pl, const int* __restrict__ ids, float* __restrict__ p_mem, float world) {
  cg::cluster_group cluster = cg::this_cluster();
  const int csize = (int)cluster.num_blocks();
  const int mat = ids[blockIdx.x / csize];
  const MatDesc md = pl.mats[mat];
  orth_body<R>(p_mem + md.p_off, md.n, md.r, world, (int)cluster.block_rank(), csize, true);
}

// single matrix, single CTA: the module-level orthogonalize(matrix) of reducer.py:180-191
template <int R>
__global__ void __launch_bounds__(kOrthThreads)
k_orthogonalize_one(float* __restrict__ mat, int n, int r) {
  orth_body<R>(mat, n, r, 1.0f, 0, 1, false);
}

// =======================================================================================
// K3:  Q = M^T P   (reducer.py:140-142) as a split-n reduction.
// Tile = t3_rows rows x (cgw * cpt) columns.  Thread (rl, cg) owns `cpt` adjacent columns and
// walks rows rl, rl + RL, ... accumulating cpt x R partial dot products in registers; the
// P rows of the tile sit in shared memory and are read as broadcasts.  The RL row lanes are
// combined through shared memory in fixed order and the tile's partial Q goes to the
// workspace; k_q_reduce sums the row blocks in order (deterministic, no atomics).
// =======================================================================================
template <int R, int CPT, bool VECLD, int NT = kThreads>
__device__ __forceinline__ void q_tile(const MatDesc& md, const ColTile& t, const float* __restrict__ M,
                                       const float* __restrict__ P, float* __restrict__ ws, float* smem) {
#ifndef PSGD_K3_U_R8
#define PSGD_K3_U_R8 6
#endif
#ifndef PSGD_K3_U_R4
#define PSGD_K3_U_R4 8
#endif
  // rows in flight per thread: 3 CTAs per SM x 256 threads x 8 x 16 B = 96 KB per SM (two CTAs with 16 rows each
  // measured 1.5 % slower over the step: a third resident CTA hides more of a tile's fixed prologue / epilogue latency)
  constexpr int U = (R <= 4 && CPT == 4) ? PSGD_K3_U_R4 : (R == 8 && CPT == 4) ? PSGD_K3_U_R8 : 8;
  constexpr int NV = CPT * R;
  const int m = md.m, r = md.r, cgw = md.cgw;
  const int RL = NT / cgw;
  const int tid = threadIdx.x;
  const int cgi = tid % cgw, rl = tid / cgw;
  const int nrows = t.nrows;
  float* Ps = smem;                       // [nrows][R]
  float* red = smem + (size_t)md.t3_rows * R;  // [RL][NV][cgw + 1]

  // stage P rows (contiguous nrows * r floats), widening r -> R with zeros
  {
    const float* src = P + (size_t)t.row0 * r;
    const int total = nrows * R;
    for (int e = tid; e < total; e += NT) {
      const int i = e / R, c = e % R;
      Ps[e] = (c < r) ? __ldcg(src + (size_t)i * r + c) : 0.f;
    }
  }
  __syncthreads();

  const int col = t.col0 + cgi * CPT;
  const bool active = col < m;  // CPT == 4 implies m % 4 == 0, so a group is all-in or all-out
  float acc[CPT][R];
#pragma unroll
  for (int x = 0; x < CPT; ++x)
#pragma unroll
    for (int c = 0; c < R; ++c) acc[x][c] = 0.f;

  if (active) {
    const float* base = M + (size_t)t.row0 * m + col;
    for (int i0 = rl; i0 < nrows; i0 += RL * U) {
      float a[U][CPT];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * RL;
        if (i < nrows) {
          const float* p = base + (size_t)i * m;
          if (CPT == 4 && VECLD) {
            const float4 v = ld_stream4(p);
            a[u][0] = v.x; a[u][1 % CPT] = v.y; a[u][2 % CPT] = v.z; a[u][3 % CPT] = v.w;
          } else {
#pragma unroll
            for (int x = 0; x < CPT; ++x) a[u][x] = ld_stream1(p + x);
          }
        } else {
#pragma unroll
          for (int x = 0; x < CPT; ++x) a[u][x] = 0.f;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        int i = i0 + u * RL;
        if (i > nrows - 1) i = nrows - 1;  // a == 0 there
        const float* pp = Ps + (size_t)i * R;
        float p[R];
        if (R % 4 == 0) {
#pragma unroll
          for (int c = 0; c < R; c += 4) {
            const float4 v = *reinterpret_cast<const float4*>(pp + c);
            p[c] = v.x; p[(c + 1) % R] = v.y; p[(c + 2) % R] = v.z; p[(c + 3) % R] = v.w;
          }
        } else {
#pragma unroll
          for (int c = 0; c < R; ++c) p[c] = pp[c];
        }
#pragma unroll
        for (int x = 0; x < CPT; ++x)
#pragma unroll
          for (int c = 0; c < R; ++c) acc[x][c] = fmaf(a[u][x], p[c], acc[x][c]);
      }
    }
  }

  // combine the RL row lanes: red[rl][v][cg], v = x * R + c
  const int rp = cgw + 1;
#pragma unroll
  for (int x = 0; x < CPT; ++x)
#pragma unroll
    for (int c = 0; c < R; ++c) red[((size_t)rl * NV + x * R + c) * rp + cgi] = acc[x][c];
  __syncthreads();

  // partial layout in the workspace: [blk][m][r]
  float* dst = ws + md.ws_off + (size_t)t.blk * m * r;
  const int wcols = cgw * CPT;
  const int total = wcols * R;
  for (int e = tid; e < total; e += NT) {
    const int j = e / R, c = e % R;  // tile-relative column, rank index
    const int g = j / CPT, x = j % CPT;
    const int cc = t.col0 + j;
    if (cc < m && c < r) {
      float s = 0.f;
      for (int q = 0; q < RL; ++q) s += red[((size_t)q * NV + x * R + c) * rp + g];
      dst[(size_t)cc * r + c] = s;
    }
  }
}

// `done` (optional): one counter per (matrix, column tile).  The CTA that completes the last row block of a
// column tile sums that tile's partials over the row blocks, in block order, straight into q_mem -- the second
// stage without a second launch.  Which CTA does the summing varies from run to run, the order of the sum does
// not, so the result stays deterministic.  The counter is left at zero for the next launch.
template <int R>
__global__ void __launch_bounds__(kThreads3, (R <= 8) ? kMinBlocks3 : (R == 16) ? 2 : 1)
k_q_from_mt(DevPlan pl, PtrSlot ps, const float* __restrict__ p_mem, float* __restrict__ ws, float* __restrict__ q_mem,
            int* __restrict__ done, int use_tc, int q_mc) {
  extern __shared__ __align__(16) float smem[];
  pdl_launch_dependents();
  const ColTile t = pl.tiles3[blockIdx.x];
  const MatDesc md = pl.mats[t.mat];
  const float* M = ps.in[t.mat];
  const float* P = p_mem + md.p_off;
  const unsigned in_flags = ps.flags[t.mat];
  pdl_wait();
  bool done_tc = false;
  if constexpr (R >= 16 && kThreads3 == kTcThreads) {
    if (use_tc && md.cpt == 4 && (in_flags & kFlagInVec)) {  // tensor pipe (3 x TF32), psgd_tc.cuh
      q_tile_tc<R>(md, t, M, P, ws, smem);
      done_tc = true;
    }
  }
  if (done_tc) {
  } else if (md.cpt == 4) {
    if (in_flags & kFlagInVec)
      q_tile<R, 4, true, kThreads3>(md, t, M, P, ws, smem);
    else
      q_tile<R, 4, false, kThreads3>(md, t, M, P, ws, smem);
  } else {
    q_tile<R, 1, false, kThreads3>(md, t, M, P, ws, smem);
  }
  if (done == nullptr) return;

  __shared__ int s_last;
  __threadfence();  // this CTA's partial is visible device-wide before it is counted
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(&done[t.pad0], 1) == md.nblk3 - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int r = md.r, m = md.m;
  const int wcols = min(md.cgw * md.cpt, m - t.col0);
  const size_t stride = (size_t)m * r;
  const size_t e0 = (size_t)t.col0 * r;
  const int count = wcols * r;
  const float* src = ws + md.ws_off + e0;
  float* dst = q_mem + md.q_off + e0;
  const int nb = md.nblk3;
  if ((stride & 3) == 0 && (count & 3) == 0) {
    for (int f = threadIdx.x; f < (count >> 2); f += kThreads3) {
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
      int b = 0;
      for (; b + 4 <= nb; b += 4) {
        const float4 v0 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)b * stride) + f);
        const float4 v1 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(b + 1) * stride) + f);
        const float4 v2 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(b + 2) * stride) + f);
        const float4 v3 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(b + 3) * stride) + f);
        sum.x = (((sum.x + v0.x) + v1.x) + v2.x) + v3.x;
        sum.y = (((sum.y + v0.y) + v1.y) + v2.y) + v3.y;
        sum.z = (((sum.z + v0.z) + v1.z) + v2.z) + v3.z;
        sum.w = (((sum.w + v0.w) + v1.w) + v2.w) + v3.w;
      }
      for (; b < nb; ++b) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(src + (size_t)b * stride) + f);
        sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
      }
      if (q_mc) {  // pushed into every rank's receive buffer (q_mem is a multicast address)
        if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
          mm_st4(dst + 4 * f, sum);
        } else {
          mm_st1(dst + 4 * f, sum.x); mm_st1(dst + 4 * f + 1, sum.y); mm_st1(dst + 4 * f + 2, sum.z); mm_st1(dst + 4 * f + 3, sum.w);
        }
      } else {
        dst[4 * f] = sum.x; dst[4 * f + 1] = sum.y; dst[4 * f + 2] = sum.z; dst[4 * f + 3] = sum.w;
      }
    }
  } else {
    for (int e = threadIdx.x; e < count; e += kThreads3) {
      float sum = 0.f;
      for (int b = 0; b < nb; ++b) sum += __ldcg(src + (size_t)b * stride + e);
      st_out(dst + e, sum, q_mc != 0);
    }
  }
  if (threadIdx.x == 0) done[t.pad0] = 0;
}

// stage 2: q_mem[e] = sum over row blocks, in block order.  One tile = kQTileElems consecutive elements
// of one matrix; a thread owns 4 of them and keeps 4 row blocks in flight.
__global__ void __launch_bounds__(kThreads)
k_q_reduce(DevPlan pl, const float* __restrict__ ws, float* __restrict__ q_mem) {
  const QTile t = pl.qtiles[blockIdx.x];
  const MatDesc md = pl.mats[t.mat];
  const size_t stride = (size_t)md.m * md.r;
  const int e = t.e0 + threadIdx.x * 4;
  if (e >= t.e0 + t.count) return;
  const float* src = ws + md.ws_off + e;
  float* dst = q_mem + md.q_off + e;
  const int nb = md.nblk3;
  if ((stride & 3) == 0 && e + 4 <= t.e0 + t.count) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    int b = 0;
    for (; b + 4 <= nb; b += 4) {
      const float4 v0 = *reinterpret_cast<const float4*>(src + (size_t)b * stride);
      const float4 v1 = *reinterpret_cast<const float4*>(src + (size_t)(b + 1) * stride);
      const float4 v2 = *reinterpret_cast<const float4*>(src + (size_t)(b + 2) * stride);
      const float4 v3 = *reinterpret_cast<const float4*>(src + (size_t)(b + 3) * stride);
      s.x = (((s.x + v0.x) + v1.x) + v2.x) + v3.x;
      s.y = (((s.y + v0.y) + v1.y) + v2.y) + v3.y;
      s.z = (((s.z + v0.z) + v1.z) + v2.z) + v3.z;
      s.w = (((s.w + v0.w) + v1.w) + v2.w) + v3.w;
    }
    for (; b < nb; ++b) {
      const float4 v = *reinterpret_cast<const float4*>(src + (size_t)b * stride);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    dst[0] = s.x; dst[1] = s.y; dst[2] = s.z; dst[3] = s.w;
  } else {
    const int lim = min(4, t.e0 + t.count - e);
    for (int x = 0; x < lim; ++x) {
      float s = 0.f;
      for (int b = 0; b < nb; ++b) s += src[(size_t)b * stride + x];
      dst[x] = s;
    }
  }
}

// =======================================================================================
// K4 + K5:  out = P (Q / n_workers)^T ; mem = M - out   (reducer.py:147,158-163), one sweep:
// M is read once, out and mem are written once.  Same tiling as K1.
// =======================================================================================
template <int R, int ROWS, int V>
__device__ __forceinline__ void fin_tile(const MatDesc& md, const float* __restrict__ M,
                                         float* __restrict__ Out, float* __restrict__ Mem,
                                         const float* __restrict__ Q, const float* __restrict__ P,
                                         float world, int row0, int nrows, int wc_cap, float* Qs) {
  const int m = md.m, r = md.r;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nsets = (nrows + ROWS - 1) / ROWS;
  const int nchunks = (m + wc_cap - 1) / wc_cap;
  const int pitch = q_pitch(m < wc_cap ? m : wc_cap, R);

  for (int ch = 0; ch < nchunks; ++ch) {
    const int c0 = ch * wc_cap;
    const int wc = (m - c0 < wc_cap) ? (m - c0) : wc_cap;
    if (ch) __syncthreads();
    stage_q<R>(Qs, Q, c0, wc, r, pitch, world);
    __syncthreads();
    const int kmax = wc / V;
    for (int set = warp; set < nsets; set += kWarps) {
      const int rbase = set * ROWS;
      float p[ROWS][R];
      size_t roff[ROWS];
#pragma unroll
      for (int a = 0; a < ROWS; ++a) {
        int rr = rbase + a;
        if (rr > nrows - 1) rr = nrows - 1;
        roff[a] = (size_t)(row0 + rr) * m + c0;
#pragma unroll
        for (int c = 0; c < R; ++c) p[a][c] = (c < r) ? __ldg(P + (size_t)(row0 + rr) * r + c) : 0.f;
      }
      for (int k = lane; k < kmax; k += 64) {
        const bool two = (k + 32) < kmax;
        const int k1 = two ? k + 32 : k;
        Pack<V> a0[ROWS], a1[ROWS];
#pragma unroll
        for (int a = 0; a < ROWS; ++a) a0[a] = pk_ld_stream<V>(M + roff[a] + (size_t)k * V);
#pragma unroll
        for (int a = 0; a < ROWS; ++a) a1[a] = two ? pk_ld_stream<V>(M + roff[a] + (size_t)k1 * V) : pk_zero<V>();
        Pack<V> o0[ROWS], o1[ROWS];
#pragma unroll
        for (int a = 0; a < ROWS; ++a) { o0[a] = pk_zero<V>(); o1[a] = pk_zero<V>(); }
#pragma unroll
        for (int c = 0; c < R; ++c) {
          const Pack<V> q0 = pk_ld_shared<V>(Qs + c * pitch + k * V);
          const Pack<V> q1 = pk_ld_shared<V>(Qs + c * pitch + k1 * V);
#pragma unroll
          for (int a = 0; a < ROWS; ++a)
#pragma unroll
            for (int x = 0; x < V; ++x) {
              o0[a].v[x] = fmaf(p[a][c], q0.v[x], o0[a].v[x]);
              o1[a].v[x] = fmaf(p[a][c], q1.v[x], o1[a].v[x]);
            }
        }
#pragma unroll
        for (int a = 0; a < ROWS; ++a) {
          if (rbase + a < nrows) {
            Pack<V> e0, e1;
#pragma unroll
            for (int x = 0; x < V; ++x) { e0.v[x] = a0[a].v[x] - o0[a].v[x]; e1.v[x] = a1[a].v[x] - o1[a].v[x]; }
            pk_st_stream<V>(Out + roff[a] + (size_t)k * V, o0[a]);
            pk_st_stream<V>(Mem + roff[a] + (size_t)k * V, e0);
            if (two) {
              pk_st_stream<V>(Out + roff[a] + (size_t)k1 * V, o1[a]);
              pk_st_stream<V>(Mem + roff[a] + (size_t)k1 * V, e1);
            }
          }
        }
      }
    }
  }
}

template <int R>
__global__ void __launch_bounds__(kThreads, 2)
k_finalize(DevPlan pl, PtrSlot ps, const float* __restrict__ p_mem, const float* __restrict__ q_mem, float world) {
  extern __shared__ __align__(16) float smem[];
  const int4 t = pl.tiles4[blockIdx.x];
  const MatDesc md = pl.mats[t.x];
  const unsigned fl = ps.flags[t.x];
  const float* M = ps.in[t.x];
  float* Out = ps.out[t.x];
  float* Mem = ps.mem[t.x];
  const float* Q = q_mem + md.q_off;
  const float* P = p_mem + md.p_off;
  if ((fl & kFlagInVec) && (fl & kFlagOutVec))
    fin_tile<R, RowsK4<R>::v, 4>(md, M, Out, Mem, Q, P, world, t.y, t.z, pl.wc_cap, smem);
  else
    fin_tile<R, RowsK4<R>::v, 1>(md, M, Out, Mem, Q, P, world, t.y, t.z, pl.wc_cap, smem);
}

// =======================================================================================
// K6: flat-buffer gather / scatter of the <=1-D tensors (tensor_buffer.py:19,34-36) as one kernel over
// a pointer table and a tile table (one CTA = up to kVecTileElems elements of one tensor, no search);
// T is the element word (1/2/4/8 bytes) so TensorBuffer keeps the upstream "dtype follows the inputs"
// behaviour bit-exactly.
// =======================================================================================
template <typename T>
__global__ void __launch_bounds__(kThreads)
k_gather(const VecTile* __restrict__ tiles, const void* const* __restrict__ src, T* __restrict__ flat) {
  const VecTile t = tiles[blockIdx.x];
  const T* s = static_cast<const T*>(src[t.tensor]) + t.start;
  T* d = flat + t.flat_off;
  for (int i = threadIdx.x; i < t.count; i += kThreads) d[i] = s[i];
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
k_scatter(const VecTile* __restrict__ tiles, const T* __restrict__ flat, void* const* __restrict__ dst) {
  const VecTile t = tiles[blockIdx.x];
  T* d = static_cast<T*>(dst[t.tensor]) + t.start;
  const T* s = flat + t.flat_off;
  for (int i = threadIdx.x; i < t.count; i += kThreads) d[i] = s[i];
}

// K6a with the caller's error-feedback add fused (ddp_init.py:155-157 for the <=1-D tensors):
// send = grad + err is stored to the send buffer AND to the flat buffer in one pass
__global__ void __launch_bounds__(kThreads)
k_gather_sum(const VecTile* __restrict__ tiles, PtrSlot ps, float* __restrict__ flat) {
  const VecTile t = tiles[blockIdx.x];
  const float* g = ps.vg[t.tensor] + t.start;
  const float* e = ps.ve[t.tensor] + t.start;
  float* send = const_cast<float*>(ps.vin[t.tensor]) + t.start;
  float* d = flat + t.flat_off;
  for (int i = threadIdx.x; i < t.count; i += kThreads) {
    const float v = g[i] + e[i];
    send[i] = v;
    d[i] = v;
  }
}

// K6b: vec_out = flat / n_workers (reducer.py:167-168); blocks past the vector tiles scale q_mem in place
// (q_mem /= n_workers, reducer.py:147)
__global__ void __launch_bounds__(kThreads)
k_finish(DevPlan pl, PtrSlot ps, const float* __restrict__ flat, int n_vtiles, float* __restrict__ q_mem,
         long long q_total, float world) {
  if ((int)blockIdx.x < n_vtiles) {
    const VecTile t = pl.vtiles[blockIdx.x];
    float* d = ps.vout[t.tensor] + t.start;
    const float* s = flat + t.flat_off;
    if (world != 1.0f) {
      for (int i = threadIdx.x; i < t.count; i += kThreads) d[i] = s[i] / world;
    } else {
      for (int i = threadIdx.x; i < t.count; i += kThreads) d[i] = s[i];
    }
  } else {
    const long long base = (long long)(blockIdx.x - n_vtiles) * kVecTileElems;
    for (int i = threadIdx.x; i < kVecTileElems; i += kThreads) {
      const long long e = base + i;
      if (e < q_total) q_mem[e] = q_mem[e] / world;
    }
  }
}

// =======================================================================================
// Caller-side update that follows reduce() upstream (ddp_init.py:166-178), one sweep instead of four
// launches per parameter:   m = first ? g : lambda * m + g ;  g += m ;  p -= lr * g
// Tile = up to kEwTileElems elements of one tensor; 128-bit accesses when the three tensors allow it.
// =======================================================================================
constexpr int kEwTileElems = 4096;

__global__ void __launch_bounds__(kThreads)
k_momentum_step(const VecTile* __restrict__ tiles, float* const* __restrict__ params, float* const* __restrict__ grads,
                float* const* __restrict__ momenta, float lr, float lambda, int first) {
  const VecTile t = tiles[blockIdx.x];
  float* p = params[t.tensor] + t.start;
  float* g = grads[t.tensor] + t.start;
  float* m = momenta[t.tensor] + t.start;
  const bool vec = (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m) & 15) == 0;
  if (vec) {
    const int n4 = t.count >> 2;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
      float4 gv = reinterpret_cast<float4*>(g)[i];
      float4 pv = reinterpret_cast<float4*>(p)[i];
      float4 mv;
      if (first) {
        mv = gv;
      } else {
        mv = reinterpret_cast<float4*>(m)[i];
        mv.x = fmaf(lambda, mv.x, gv.x); mv.y = fmaf(lambda, mv.y, gv.y);
        mv.z = fmaf(lambda, mv.z, gv.z); mv.w = fmaf(lambda, mv.w, gv.w);
      }
      gv.x += mv.x; gv.y += mv.y; gv.z += mv.z; gv.w += mv.w;
      pv.x = fmaf(-lr, gv.x, pv.x); pv.y = fmaf(-lr, gv.y, pv.y);
      pv.z = fmaf(-lr, gv.z, pv.z); pv.w = fmaf(-lr, gv.w, pv.w);
      reinterpret_cast<float4*>(m)[i] = mv;
      reinterpret_cast<float4*>(g)[i] = gv;
      reinterpret_cast<float4*>(p)[i] = pv;
    }
    for (int i = (n4 << 2) + threadIdx.x; i < t.count; i += kThreads) {
      const float mv = first ? g[i] : fmaf(lambda, m[i], g[i]);
      const float gv = g[i] + mv;
      m[i] = mv; g[i] = gv; p[i] = fmaf(-lr, gv, p[i]);
    }
  } else {
    for (int i = threadIdx.x; i < t.count; i += kThreads) {
      const float mv = first ? g[i] : fmaf(lambda, m[i], g[i]);
      const float gv = g[i] + mv;
      m[i] = mv; g[i] = gv; p[i] = fmaf(-lr, gv, p[i]);
    }
  }
}

// =======================================================================================
// One-shot SUM all-reduce over peer-mapped (symmetric) memory for the two small packed buffers of a step
// (P || flat: reducer.py:126,132 and Q: reducer.py:145).  Every rank reads all ranks' input buffers over
// NVLink and adds them in rank order, so all ranks produce bit-identical sums (the property the upstream loop
// relies on after dist.all_reduce).  The messages are 0.7-8 MB: latency-bound, which is why one kernel with a
// flag barrier beats a general collective here.
//   barrier: block 0 publishes "my input is complete" (epoch) into every peer's signal pad, every block waits
//   until its own pad shows the epoch from every rank.  No trailing barrier is needed: the P and Q phases
//   alternate, and a rank can only overwrite its P input after it has passed the Q barrier (and vice versa),
//   which every peer reaches only after it finished reading.
// =======================================================================================
__global__ void __launch_bounds__(kThreads)
k_peer_allreduce(PeerArgs a, float* __restrict__ out, long long n, int rank, int world, int slot0, unsigned epoch) {
  if (blockIdx.x == 0 && (int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(a.sig[threadIdx.x] + slot0 + rank, epoch);  // tell peer threadIdx.x: rank's input is ready
  }
  if ((int)threadIdx.x < world) {
    const unsigned* mine = a.sig[rank] + slot0 + threadIdx.x;
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
      // a peer that never arrives (crashed rank) must not hang the GPU: give up after ~10 s and fault loudly
      if (clock64() - t0 > 20000000000LL) {
        printf("psgd k_peer_allreduce: rank %d timed out waiting for rank %d (epoch %u)\n", rank, (int)threadIdx.x, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n4; i += stride) {
    float4 s = ld_sys4(a.in[0] + 4 * i);
    for (int r = 1; r < world; ++r) {
      const float4 v = ld_sys4(a.in[r] + 4 * i);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    reinterpret_cast<float4*>(out)[i] = s;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    float s = ld_sys1(a.in[0] + i);
    for (int r = 1; r < world; ++r) s += ld_sys1(a.in[r] + i);
    out[i] = s;
  }
}


// "Push" of a locally produced buffer: this rank's slot of every rank's receive buffer (one multimem.st per element).
// Used when the producing kernels could not push their results themselves (direct-load fallbacks, separate second
// stage of K3): same values, one small extra launch.
__global__ void __launch_bounds__(kThreads)
k_push_to_all(const float* __restrict__ local, float* __restrict__ mc, long long n) {
  const long long stride = (long long)gridDim.x * kThreads;
  const long long n4 = n >> 2;
  const bool vec = ((reinterpret_cast<uintptr_t>(local) | reinterpret_cast<uintptr_t>(mc)) & 15) == 0;
  if (vec) {
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n4; i += stride)
      mm_st4(mc + 4 * i, __ldcg(reinterpret_cast<const float4*>(local) + i));
    for (long long i = (n4 << 2) + (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) mm_st1(mc + i, __ldcg(local + i));
  } else {
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) mm_st1(mc + i, __ldcg(local + i));
  }
}

// Stand-alone form of the all-reduce-on-load (fallback for steps whose consumer kernel cannot take it in: direct-load
// kernels on unaligned tensors, matrices too tall for the register-resident Gram-Schmidt): one flag barrier, then
// out[i] = sum over ranks of in_r[i], divided by `div` from element div_from on.
__global__ void __launch_bounds__(kThreads)
k_peer_reduce_to_local(PeerLoad pm, float* __restrict__ out, long long n, float div, long long div_from) {
  peer_barrier(pm, blockIdx.x == 0);
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    const float v = peer_sum1(pm, i);
    out[i] = (i >= div_from && div != 1.0f) ? v / div : v;
  }
}

// ---- the same sum through the NVSwitch multicast address of the symmetric buffer ----------------------------
// Two-shot, in place: after a flag barrier ("every rank's input is written") rank k owns the k-th slice of the
// buffer: one multimem.ld_reduce returns the switch-side sum of all ranks' copies of an element, one multimem.st
// writes it back into every rank's copy.  A second flag barrier ("every rank's slice is stored everywhere") and the
// result is copied from this rank's copy to `out`.  Each element is summed exactly once, so all ranks hold the
// same bits; a rank moves n/W elements over NVLink instead of (W-1) n.  `blocks_done` is a zero-initialised word owned
// by the call site (one per exchange buffer): the kernel leaves it at zero.

__global__ void __launch_bounds__(kThreads)
k_peer_allreduce_mc(PeerArgs a, float* __restrict__ mc, float* __restrict__ out, long long n, int rank, int world,
                    int slot0, unsigned epoch, float out_div, long long div_from, unsigned* __restrict__ blocks_done) {
  __shared__ int s_last;
  const long long n4 = (n + 3) >> 2;  // the symmetric buffers are padded to whole float4s, `out` is not
  if (blockIdx.x == 0 && (int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(a.sig[threadIdx.x] + slot0 + rank, epoch);  // this rank's input is complete (stream order)
  }
  peer_flags_wait(a.sig[rank] + slot0, world, epoch, rank, 0);

  const long long per = (n4 + world - 1) / world;
  const long long lo = (long long)rank * per;
  const long long hi = (lo + per < n4) ? lo + per : n4;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = lo + (long long)blockIdx.x * kThreads + threadIdx.x; i < hi; i += stride) {
    const float4 v = mm_ld_reduce4(mc + 4 * i);
    mm_st4(mc + 4 * i, v);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(blocks_done, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last) {  // the whole slice of this rank is stored in every copy
    if (threadIdx.x == 0) *blocks_done = 0;
    if ((int)threadIdx.x < world) {
      __threadfence_system();
      st_release_sys(a.sig[threadIdx.x] + slot0 + kMaxPeers + rank, epoch);
    }
  }
  peer_flags_wait(a.sig[rank] + slot0 + kMaxPeers, world, epoch, rank, 1);

  const float* mine = a.in[rank];
  const long long full4 = n >> 2;
  // the mean of reducer.py:128,147,167 can be formed on the way out: elements from div_from on leave as sum / out_div
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < full4; i += stride) {
    float4 v = ld_sys4(mine + 4 * i);
    if (out_div != 1.0f) {
      const long long e = 4 * i;
      if (e >= div_from) v.x = v.x / out_div;
      if (e + 1 >= div_from) v.y = v.y / out_div;
      if (e + 2 >= div_from) v.z = v.z / out_div;
      if (e + 3 >= div_from) v.w = v.w / out_div;
    }
    reinterpret_cast<float4*>(out)[i] = v;
  }
  for (long long i = (full4 << 2) + (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    float v = ld_sys1(mine + i);
    if (out_div != 1.0f && i >= div_from) v = v / out_div;
    out[i] = v;
  }
}

}  // namespace psgd
