// Tensor-pipe variants of the three contractions for compression ranks 16 and 32 (BASELINE.json north_star: "tensor
// cores are used only if a rank >= 16 config shows tensor-pipe benefit").  At rank 16 the FFMA kernels issue 16 FMAs
// per gradient element and are issue-bound at a third of the HBM roofline; here the products run on the tensor pipe
// as 3 x TF32 (a = a_hi + a_lo with both halves in TF32; a b ~ a_lo b_hi + a_hi b_lo + a_hi b_hi, fp32 accumulate),
// which keeps the error at the fp32 level the 1e-5 parity gate needs (plain TF32 would not).
//
// mma.sync.m16n8k8 (tf32) fragments, g = lane / 4, t = lane % 4:
//   A (16 x 8, row):  a0 = A[g][t]      a1 = A[g + 8][t]   a2 = A[g][t + 4]   a3 = A[g + 8][t + 4]
//   B (8 x 8, col):   b0 = B[t][g]      b1 = B[t + 4][g]
//   C (16 x 8):       c0 = C[g][2t]     c1 = C[g][2t + 1]  c2 = C[g + 8][2t]  c3 = C[g + 8][2t + 1]
// Gradient elements go from global memory straight into fragment registers with 128-bit loads: a contraction index
// (k) or a free column index (n) may be permuted at will as long as both operands agree, so "lane t owns four
// consecutive columns" is mapped onto the fragment positions and every load / store stays fully coalesced.
#pragma once
#include "psgd_common.cuh"

namespace psgd {

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
// x = hi + lo (+ O(2^-22 |x|)), both representable in TF32
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = to_tf32(x);
  lo = to_tf32(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// c += a b with both operands split: small terms first
__device__ __forceinline__ void mma_3xtf32(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], uint32_t bh0,
                                           uint32_t bh1, uint32_t bl0, uint32_t bl1) {
  mma_tf32(c, al[0], al[1], al[2], al[3], bh0, bh1);
  mma_tf32(c, ah[0], ah[1], ah[2], ah[3], bl0, bl1);
  mma_tf32(c, ah[0], ah[1], ah[2], ah[3], bh0, bh1);
}

constexpr int kTcThreads = 256;
constexpr int kTcWarps = kTcThreads / 32;

// shared-memory row pitch (words) of the transposed Q stage of K1: = 16 (mod 32), so that the 128-bit reads of a
// quarter warp (two rows x four lanes) fall into eight distinct bank groups
__host__ __device__ inline int tc_q_pitch(int wc) { return ((wc + 31) / 32) * 32 + 16; }

// =======================================================================================================
// K1 (tensor pipe):  P = M Q.   Tile = nrows full rows of one matrix (the row tiles of the direct kernel); a warp takes
// 16 rows at a time, lane (g, t) loads rows g and g + 8, four consecutive columns of every 16.  The staged Q is the
// transposed copy Qs[c][j] = Q[j][c] / q_div, split once into its TF32 halves.  Columns are processed in chunks of at
// most wc_cap; with several chunks a row's partial sums are added up in P by the owning lane, in chunk order.
// =======================================================================================================
template <int R, bool FUSE>
__global__ void __launch_bounds__(kTcThreads, 2)
k_p_from_m_tc(DevPlan pl, PtrSlot ps, const float* __restrict__ q_mem, float* __restrict__ p_mem, float q_div, int wc_cap) {
  static_assert(R == 16 || R == 32, "tensor-pipe kernels are for ranks 16 and 32");
  constexpr int NT = R / 8;  // n8 tiles
  extern __shared__ __align__(16) float smem[];
  const int4 tile = pl.tiles1[blockIdx.x];
  const MatDesc md = pl.mats[tile.x];
  const int m = md.m, r = md.r, row0 = tile.y, nrows = tile.z;
  const float* M = ps.in[tile.x];
  const float* G = FUSE ? ps.g[tile.x] : nullptr;
  const float* E = FUSE ? ps.e[tile.x] : nullptr;
  float* Mw = FUSE ? const_cast<float*>(M) : nullptr;
  const float* Q = q_mem + md.q_off;
  float* P = p_mem + md.p_off;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int nchunks = (m + wc_cap - 1) / wc_cap;
  const int pitch = tc_q_pitch(m < wc_cap ? m : wc_cap);
  uint32_t* Qh = reinterpret_cast<uint32_t*>(smem);
  uint32_t* Ql = Qh + (size_t)R * pitch;
  const int nblk = (nrows + 15) / 16;

  for (int ch = 0; ch < nchunks; ++ch) {
    const int c0 = ch * wc_cap;
    const int wc = (m - c0 < wc_cap) ? (m - c0) : wc_cap;  // multiple of 4; padded to 16 with zeros below
    const int wc16 = (wc + 15) & ~15;
    if (ch) __syncthreads();
    for (int e = threadIdx.x; e < wc16 * R; e += kTcThreads) {
      const int j = e / R, c = e - j * R;
      float v = 0.f;
      if (j < wc && c < r) {
        v = __ldg(Q + (size_t)(c0 + j) * r + c);
        if (q_div != 1.0f) v = v / q_div;
      }
      uint32_t h, l;
      split_tf32(v, h, l);
      Qh[c * pitch + j] = h;
      Ql[c * pitch + j] = l;
    }
    __syncthreads();
    for (int blk = warp; blk < nblk; blk += kTcWarps) {
      const int ra = min(blk * 16 + g, nrows - 1), rb = min(blk * 16 + g + 8, nrows - 1);  // clamped: results discarded
      const size_t offa = (size_t)(row0 + ra) * m + c0 + 4 * t, offb = (size_t)(row0 + rb) * m + c0 + 4 * t;
      float acc[NT][4];
#pragma unroll
      for (int i = 0; i < NT; ++i)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[i][k] = 0.f;
#pragma unroll 4
      for (int kb = 0; kb < wc16; kb += 16) {
        float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
        if (kb + 4 * t < wc) {
          if (FUSE) {
            const float4 ga = ld_stream4(G + offa + kb), ea = ld_stream4(E + offa + kb);
            const float4 gb = ld_stream4(G + offb + kb), eb = ld_stream4(E + offb + kb);
            va = make_float4(ga.x + ea.x, ga.y + ea.y, ga.z + ea.z, ga.w + ea.w);
            vb = make_float4(gb.x + eb.x, gb.y + eb.y, gb.z + eb.z, gb.w + eb.w);
            if (blk * 16 + g < nrows) st_stream4(Mw + offa + kb, va);
            if (blk * 16 + g + 8 < nrows) st_stream4(Mw + offb + kb, vb);
          } else {
            va = ld_stream4(M + offa + kb);
            vb = ld_stream4(M + offb + kb);
          }
        }
        // two k8 steps: logical k = t <-> column 4t (+2), k = t + 4 <-> column 4t + 1 (+2)
        uint32_t ah0[4], al0[4], ah1[4], al1[4];
        split_tf32(va.x, ah0[0], al0[0]); split_tf32(vb.x, ah0[1], al0[1]);
        split_tf32(va.y, ah0[2], al0[2]); split_tf32(vb.y, ah0[3], al0[3]);
        split_tf32(va.z, ah1[0], al1[0]); split_tf32(vb.z, ah1[1], al1[1]);
        split_tf32(va.w, ah1[2], al1[2]); split_tf32(vb.w, ah1[3], al1[3]);
#pragma unroll
        for (int i = 0; i < NT; ++i) {
          const uint4 qh = *reinterpret_cast<const uint4*>(Qh + (size_t)(8 * i + g) * pitch + kb + 4 * t);
          const uint4 ql = *reinterpret_cast<const uint4*>(Ql + (size_t)(8 * i + g) * pitch + kb + 4 * t);
          mma_3xtf32(acc[i], ah0, al0, qh.x, qh.y, ql.x, ql.y);
          mma_3xtf32(acc[i], ah1, al1, qh.z, qh.w, ql.z, ql.w);
        }
      }
      // C fragment: rows g / g + 8, rank columns 8 i + 2 t, + 1
#pragma unroll
      for (int i = 0; i < NT; ++i) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int row = blk * 16 + g + ((k & 2) ? 8 : 0);
          const int c = 8 * i + 2 * t + (k & 1);
          if (row < nrows && c < r) {
            float* dst = P + (size_t)(row0 + row) * r + c;
            *dst = ch ? *dst + acc[i][k] : acc[i][k];
          }
        }
      }
    }
  }
}

// =======================================================================================================
// K3 (tensor pipe):  Q = M^T P_hat, computed as Q^T [R x cols] = P_hat^T [R x rows] . M [rows x cols].   Tile = the
// column tile of the FFMA kernel (t3_rows rows x 256 columns): each of the eight warps owns a strip of 32 columns and
// walks all rows of the tile, eight at a time (lane (g, t): rows t and t + 4, columns 4g .. 4g + 3), so no reduction
// across warps is needed.  P_hat rows of the tile sit in shared memory in fp32 and are split on the fly.  The tile's
// partial Q goes to the workspace [row block][m][r]; the completion-counter second stage of k_q_from_mt follows.
// =======================================================================================================
constexpr int kTcPsPitch = 24;  // words per staged P_hat row for R = 16 (banks 24 t + g distinct); R = 32 uses 40

template <int R>
__device__ __forceinline__ void q_tile_tc(const MatDesc& md, const ColTile& tl, const float* __restrict__ M,
                                          const float* __restrict__ P, float* __restrict__ ws, float* smem) {
  constexpr int MT = R / 16;            // m16 tiles over the rank
  constexpr int PP = R == 16 ? 24 : 40;  // pitch of the staged P rows
  const int m = md.m, r = md.r, nrows = tl.nrows;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  float* Ps = smem;  // [nrows8][PP], rows beyond nrows and columns beyond r zero
  const int nrows8 = (nrows + 7) & ~7;
  for (int e = threadIdx.x; e < nrows8 * R; e += kTcThreads) {
    const int i = e / R, c = e - i * R;
    Ps[i * PP + c] = (i < nrows && c < r) ? __ldcg(P + (size_t)(tl.row0 + i) * r + c) : 0.f;
  }
  __syncthreads();
  const int tile_w = md.cgw * md.cpt;           // columns of a tile (<= 256 = 8 warps x 32)
  const int col = tl.col0 + 32 * warp + 4 * g;  // this lane's four columns
  const bool active = col < m && 32 * warp + 4 * g < tile_w;  // m % 4 == 0: all four in or all out
  float acc[MT][4][4];
#pragma unroll
  for (int a = 0; a < MT; ++a)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[a][j][k] = 0.f;
  if (tl.col0 + 32 * warp < m && 32 * warp < tile_w) {  // warp-uniform: the strip has at least one column
    const float* base = M + (size_t)tl.row0 * m + (active ? col : 0);
#pragma unroll 2
    for (int k0 = 0; k0 < nrows8; k0 += 8) {
      float4 u = make_float4(0.f, 0.f, 0.f, 0.f), v = u;
      if (active && k0 + t < nrows) u = ld_stream4(base + (size_t)(k0 + t) * m);
      if (active && k0 + t + 4 < nrows) v = ld_stream4(base + (size_t)(k0 + t + 4) * m);
      uint32_t bh0[4], bl0[4], bh1[4], bl1[4];
      split_tf32(u.x, bh0[0], bl0[0]); split_tf32(u.y, bh0[1], bl0[1]);
      split_tf32(u.z, bh0[2], bl0[2]); split_tf32(u.w, bh0[3], bl0[3]);
      split_tf32(v.x, bh1[0], bl1[0]); split_tf32(v.y, bh1[1], bl1[1]);
      split_tf32(v.z, bh1[2], bl1[2]); split_tf32(v.w, bh1[3], bl1[3]);
#pragma unroll
      for (int a = 0; a < MT; ++a) {
        // A = P_hat^T: a0 = P[k0 + t][16a + g], a1 = P[k0 + t][16a + g + 8], a2 / a3 the same of row k0 + t + 4
        uint32_t ah[4], al[4];
        split_tf32(Ps[(k0 + t) * PP + 16 * a + g], ah[0], al[0]);
        split_tf32(Ps[(k0 + t) * PP + 16 * a + g + 8], ah[1], al[1]);
        split_tf32(Ps[(k0 + t + 4) * PP + 16 * a + g], ah[2], al[2]);
        split_tf32(Ps[(k0 + t + 4) * PP + 16 * a + g + 8], ah[3], al[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j) mma_3xtf32(acc[a][j], ah, al, bh0[j], bh1[j], bl0[j], bl1[j]);
      }
    }
  }
  // C of MMA j: rank rows 16a + g / + 8, n = 2t, 2t + 1 <-> columns strip + 4 n + j
  float* dst = ws + md.ws_off + (size_t)tl.blk * m * r;
#pragma unroll
  for (int a = 0; a < MT; ++a)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int c = 16 * a + g + ((k & 2) ? 8 : 0);
        const int cj = 32 * warp + 4 * (2 * t + (k & 1)) + j;  // tile-relative column
        const int cc = tl.col0 + cj;
        if (cc < m && cj < tile_w && c < r) dst[(size_t)cc * r + c] = acc[a][j][k];
      }
}

// =======================================================================================================
// K4 + K5 (tensor pipe):  out = P_hat (Q / W)^T, mem = M - out.   Tile = nrows full rows; the CTA walks the columns in
// groups of 256, each warp owning a strip of 32: the strip's Q fragments (both TF32 halves) are loaded once from the
// staged copy (Qn[col][k], pitch = 1 mod 8: banks 4g + t distinct) and reused for every 16-row block of the tile.
// The n <-> column permutation (n = 2t, 2t + 1 of MMA j <-> columns 8t + j, 8t + 4 + j) gives every lane eight
// consecutive columns of rows g and g + 8: M is read and out / mem are written with 128-bit accesses.
// =======================================================================================================
template <int R>
__global__ void __launch_bounds__(kTcThreads, (R == 16) ? 2 : 1)
k_finalize_tc(DevPlan pl, PtrSlot ps, const float* __restrict__ p_mem, const float* __restrict__ q_mem, float world,
              int wc_cap) {
  static_assert(R == 16 || R == 32, "tensor-pipe kernels are for ranks 16 and 32");
  constexpr int KS = R / 8;   // k8 steps
  constexpr int QP = R + 1;   // pitch of the staged Q rows (R + 1 = 1 mod 8)
  extern __shared__ __align__(16) float smem[];
  const int4 tile = pl.tiles4[blockIdx.x];
  const MatDesc md = pl.mats[tile.x];
  const int m = md.m, r = md.r, row0 = tile.y, nrows = tile.z;
  const float* M = ps.in[tile.x];
  float* Out = ps.out[tile.x];
  float* Mem = ps.mem[tile.x];
  const float* Q = q_mem + md.q_off;
  const float* P = p_mem + md.p_off;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  uint32_t* Qh = reinterpret_cast<uint32_t*>(smem);  // [wc][QP]
  uint32_t* Ql = Qh + (size_t)wc_cap * QP;
  const int nblk = (nrows + 15) / 16;
  const int nchunks = (m + wc_cap - 1) / wc_cap;

  for (int ch = 0; ch < nchunks; ++ch) {
    const int c0 = ch * wc_cap;
    const int wc = (m - c0 < wc_cap) ? (m - c0) : wc_cap;
    if (ch) __syncthreads();
    for (int e = threadIdx.x; e < wc * R; e += kTcThreads) {
      const int j = e / R, c = e - j * R;
      float v = 0.f;
      if (c < r) {
        v = __ldg(Q + (size_t)(c0 + j) * r + c);
        if (world != 1.0f) v = v / world;
      }
      uint32_t h, l;
      split_tf32(v, h, l);
      Qh[j * QP + c] = h;
      Ql[j * QP + c] = l;
    }
    __syncthreads();
    for (int cg0 = 0; cg0 < wc; cg0 += 32 * kTcWarps) {
      const int s0 = cg0 + 32 * warp;  // chunk-relative first column of this warp's strip
      if (s0 >= wc) continue;           // warp-uniform
      // B fragments of the strip: MMA j, k8 step s: b0 = Qn[s0 + 4g + j][8s + t], b1 = Qn[..][8s + t + 4]
      uint32_t bh[4][KS][2], bl[4][KS][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int qc = min(s0 + 4 * g + j, wc - 1);
#pragma unroll
        for (int s = 0; s < KS; ++s) {
          bh[j][s][0] = Qh[qc * QP + 8 * s + t]; bh[j][s][1] = Qh[qc * QP + 8 * s + t + 4];
          bl[j][s][0] = Ql[qc * QP + 8 * s + t]; bl[j][s][1] = Ql[qc * QP + 8 * s + t + 4];
        }
      }
      const int col = c0 + s0 + 8 * t;          // this lane's eight columns: col .. col + 7
      const bool in0 = s0 + 8 * t < wc, in1 = s0 + 8 * t + 4 < wc;
      for (int blk = 0; blk < nblk; ++blk) {
        const int ra = blk * 16 + g, rb = ra + 8;
        const int rac = min(ra, nrows - 1), rbc = min(rb, nrows - 1);
        // issue the M loads first, then build the A fragments from P while they are in flight
        const size_t oa = (size_t)(row0 + rac) * m + col, ob = (size_t)(row0 + rbc) * m + col;
        float4 ma0 = make_float4(0.f, 0.f, 0.f, 0.f), ma1 = ma0, mb0 = ma0, mb1 = ma0;
        if (in0) { ma0 = ld_stream4(M + oa); mb0 = ld_stream4(M + ob); }
        if (in1) { ma1 = ld_stream4(M + oa + 4); mb1 = ld_stream4(M + ob + 4); }
        uint32_t ah[KS][4], al[KS][4];
        const float* pa = P + (size_t)(row0 + rac) * r;
        const float* pb = P + (size_t)(row0 + rbc) * r;
#pragma unroll
        for (int s = 0; s < KS; ++s) {
          const int k1 = 8 * s + t, k2 = 8 * s + t + 4;
          split_tf32(k1 < r ? __ldg(pa + k1) : 0.f, ah[s][0], al[s][0]);
          split_tf32(k1 < r ? __ldg(pb + k1) : 0.f, ah[s][1], al[s][1]);
          split_tf32(k2 < r ? __ldg(pa + k2) : 0.f, ah[s][2], al[s][2]);
          split_tf32(k2 < r ? __ldg(pb + k2) : 0.f, ah[s][3], al[s][3]);
        }
        float acc[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#pragma unroll
          for (int k = 0; k < 4; ++k) acc[j][k] = 0.f;
#pragma unroll
          for (int s = 0; s < KS; ++s) mma_3xtf32(acc[j], ah[s], al[s], bh[j][s][0], bh[j][s][1], bl[j][s][0], bl[j][s][1]);
        }
        // acc[j][0] = out[ra][col + j], acc[j][1] = out[ra][col + 4 + j], acc[j][2] / [3] the same of row rb
        const float4 oa0 = make_float4(acc[0][0], acc[1][0], acc[2][0], acc[3][0]);
        const float4 oa1 = make_float4(acc[0][1], acc[1][1], acc[2][1], acc[3][1]);
        const float4 ob0 = make_float4(acc[0][2], acc[1][2], acc[2][2], acc[3][2]);
        const float4 ob1 = make_float4(acc[0][3], acc[1][3], acc[2][3], acc[3][3]);
        if (ra < nrows) {
          if (in0) { st_stream4(Out + oa, oa0); st_stream4(Mem + oa, make_float4(ma0.x - oa0.x, ma0.y - oa0.y, ma0.z - oa0.z, ma0.w - oa0.w)); }
          if (in1) { st_stream4(Out + oa + 4, oa1); st_stream4(Mem + oa + 4, make_float4(ma1.x - oa1.x, ma1.y - oa1.y, ma1.z - oa1.z, ma1.w - oa1.w)); }
        }
        if (rb < nrows) {
          if (in0) { st_stream4(Out + ob, ob0); st_stream4(Mem + ob, make_float4(mb0.x - ob0.x, mb0.y - ob0.y, mb0.z - ob0.z, mb0.w - ob0.w)); }
          if (in1) { st_stream4(Out + ob + 4, ob1); st_stream4(Mem + ob + 4, make_float4(mb1.x - ob1.x, mb1.y - ob1.y, mb1.z - ob1.z, mb1.w - ob1.w)); }
        }
      }
    }
  }
}

}  // namespace psgd
