/* psgd_b200.h -- C ABI of the B200-native PowerSGD reducer hot path.
 *
 * The upstream project (jaeyong-song/network_distributed_pytorch) has no FFI of its own:
 * its boundary is the Python class PowerSGDReducer in
 * ddp_powersgd_guide_cifar10/reducer.py (identical copy under
 * ddp_powersgd_distillBERT_IMDb/).  Each entry point below replaces the Python loop cited
 * next to it; the host mirror in network_distributed_pytorch_b200/reducer.py calls them
 * through ctypes with raw device pointers and the current CUDA stream.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; psgd_last_error() then
 *     holds a human readable message (thread local);
 *   - all gradient data is fp32, row-major, contiguous; "matrix i" is the i-th tensor with
 *     more than one dimension viewed as n[i] x m[i] (reducer.py:72-76), "vector j" the j-th
 *     tensor with at most one dimension (reducer.py:53-57);
 *   - p_mem / q_mem use the upstream packed layout: matrix i's P is n[i] x r[i] row-major
 *     at element offset sum_{k<i} n[k]*r[k], Q is m[i] x r[i] at sum_{k<i} m[k]*r[k],
 *     r[i] = min(n[i], m[i], rank), no padding (reducer.py:78-98);
 *   - a plan is not thread safe; different plans may be used concurrently; all calls on
 *     one plan must be issued to one stream at a time;
 *   - `psgd_stream` is a cudaStream_t passed as void*.
 */
#ifndef PSGD_B200_H
#define PSGD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct psgd_plan psgd_plan;
typedef struct psgd_tb psgd_tb;
typedef struct psgd_sgd psgd_sgd;
typedef void* psgd_stream;

#define PSGD_ABI_VERSION 2
#define PSGD_MAX_RANK 32

/* phases, used by the tile and timing queries */
enum {
  PSGD_PHASE_PACK = 0,      /* K6a */
  PSGD_PHASE_P_FROM_M = 1,  /* K1  */
  PSGD_PHASE_ORTH = 2,      /* K2  */
  PSGD_PHASE_Q_FROM_MT = 3, /* K3 stage 1 */
  PSGD_PHASE_Q_REDUCE = 4,  /* K3 stage 2 */
  PSGD_PHASE_FINALIZE = 5,  /* K4+K5 */
  PSGD_PHASE_FINISH = 6,    /* K6b (+ in-place Q mean) */
  PSGD_N_PHASES = 7
};

int psgd_abi_version(void);
const char* psgd_last_error(void);

/* Replaces the size pass and view construction of reducer.py:69-98 (done once per shape
 * set instead of every call).  n/m: the 2-D views of the matrices; vec_len: element
 * counts of the <=1-D tensors (tensor_buffer.py:10-16 prefix sums).  device >= 0 selects
 * the CUDA device and uploads the tables; device < 0 builds the host-side plan only
 * (no CUDA call is made; compute entry points then fail) so the layout can be inspected
 * on a machine without a GPU. */
int psgd_plan_create(int n_mat, const int64_t* n, const int64_t* m, int rank,
                     int n_vec, const int64_t* vec_len, int device, psgd_plan** out);
int psgd_plan_destroy(psgd_plan* plan);

/* layout queries (host only) */
int64_t psgd_p_elems(const psgd_plan* plan);         /* sum n*r  (reducer.py:79) */
int64_t psgd_q_elems(const psgd_plan* plan);         /* sum m*r  (reducer.py:80) */
int64_t psgd_vec_elems(const psgd_plan* plan);       /* tensor_buffer.py:10-13    */
int64_t psgd_workspace_bytes(const psgd_plan* plan); /* scratch for psgd_q_from_mt */
/* out[0..7] = n, m, r, p_off, q_off, ws_off, k3 row blocks, k3 columns per tile */
int psgd_matrix_info(const psgd_plan* plan, int i, int64_t* out8);
int psgd_tile_count(const psgd_plan* plan, int phase);
/* out[0..5] = matrix, row0, nrows, col0, ncols, row-block index.
 * PSGD_PHASE_ORTH: the Gram-Schmidt job of CTA idx -- matrix, first cluster rank of the group of CTAs that shares the
 * matrix, CTAs in the group, this CTA's rank in the group, CTAs per hardware cluster, 1 for the generic kernel. */
int psgd_tile_get(const psgd_plan* plan, int phase, int idx, int32_t* out6);

/* Per-step pointer refresh: parameter.grad is re-allocated by zero_grad()
 * (ddp_init.py:144), so addresses may move every step.  Host arrays of device pointers,
 * n_mat / n_vec entries each.  grad / err may be NULL unless fuse_add is used. */
int psgd_set_ptrs(psgd_plan* plan, const void* const* mat_in, void* const* mat_out,
                  void* const* mat_mem, const void* const* vec_in, void* const* vec_out,
                  const void* const* grad, const void* const* err, psgd_stream stream);

/* Same as psgd_set_ptrs for a step whose error-feedback add is fused (ddp_init.py:155-157,
 * send = grad + err for every parameter): mat_send / vec_send are the caller's send buffers (written by
 * psgd_p_from_m(fuse_add=1) and psgd_pack_vec_fused), grad / err the two operands for matrices and vectors. */
int psgd_set_ptrs_fused(psgd_plan* plan, const void* const* mat_send, void* const* mat_out,
                        void* const* mat_mem, const void* const* vec_send, void* const* vec_out,
                        const void* const* mat_grad, const void* const* mat_err,
                        const void* const* vec_grad, const void* const* vec_err, psgd_stream stream);

/* K6a: flat = cat(vec_in)             (tensor_buffer.py:19, reducer.py:131) */
int psgd_pack_vec(psgd_plan* plan, float* flat, psgd_stream stream);
/* K6a with ddp_init.py:155-157 fused for the <=1-D tensors: vec_send = vec_grad + vec_err, flat = cat(vec_send) */
int psgd_pack_vec_fused(psgd_plan* plan, float* flat, psgd_stream stream);
/* K1: P_i = M_i Q_i for every matrix  (reducer.py:121-123).  fuse_add != 0 first forms
 * M = grad + err (ddp_init.py:155-157) and stores it to mat_in.  q_div divides Q while it
 * is staged (1 for the normal path; the worker count when Q still holds an all-reduce
 * SUM, used by extra power iterations). */
int psgd_p_from_m(psgd_plan* plan, float* p_mem, const float* q_mem, int fuse_add,
                  float q_div, psgd_stream stream);
/* psgd_p_from_m with psgd_pack_vec (or psgd_pack_vec_fused when fuse_add != 0) folded into the same launch:
 * one kernel and one launch gap less per step.  Falls back to two launches when the streamed kernel is not in use. */
int psgd_p_from_m_pack(psgd_plan* plan, float* p_mem, const float* q_mem, int fuse_add, float q_div,
                       float* flat, psgd_stream stream);
/* Single-worker convenience: psgd_p_from_m_pack (flat may be NULL: no gather) followed by
 * psgd_orthogonalize(n_workers = 1) as ONE call, for the case where no all-reduce sits between reducer.py:121-123 and
 * :136-137 (one worker).  Results are identical to the two separate calls. */
int psgd_p_from_m_orth(psgd_plan* plan, float* p_mem, const float* q_mem, int fuse_add, float q_div,
                       float* flat, psgd_stream stream);
/* K2: p_mem /= n_workers then modified Gram-Schmidt of every P_i in place
 * (reducer.py:128, 136-137, 180-191). */
int psgd_orthogonalize(psgd_plan* plan, float* p_mem, float n_workers, psgd_stream stream);
/* K3: Q_i = M_i^T P_i                 (reducer.py:140-142): split-n partials into
 * `workspace` (psgd_workspace_bytes, 16-byte aligned) then a fixed-order sum. */
int psgd_q_from_mt(psgd_plan* plan, const float* p_mem, float* q_mem, void* workspace,
                   psgd_stream stream);
/* K4+K5: out_i = P_i (Q_i / n_workers)^T ; mem_i = M_i - out_i   (reducer.py:147,158-163).
 * q_mem itself is left untouched (still the SUM); psgd_finish scales it. */
int psgd_finalize(psgd_plan* plan, const float* p_mem, const float* q_mem, float n_workers,
                  psgd_stream stream);
/* psgd_finalize with the scatter half of psgd_finish (vec_out_j = flat_j / n_workers) folded into the same launch.
 * q_mem /= n_workers still needs psgd_finish(plan, NULL-able flat, q_mem, ...) afterwards when n_workers > 1. */
int psgd_finalize_unpack(psgd_plan* plan, const float* p_mem, const float* q_mem, float n_workers,
                         const float* flat, psgd_stream stream);
/* K6b: if flat != NULL, vec_out_j = flat_j / n_workers (reducer.py:167-168, tensor_buffer.py:34-36); if
 * q_mem != NULL, q_mem /= n_workers in place (reducer.py:147). */
int psgd_finish(psgd_plan* plan, const float* flat, float* q_mem, float n_workers,
                psgd_stream stream);

/* per-phase device timing with CUDA events on the launch stream (bench.py roofline): `mask` bit k
 * brackets phase k (PSGD_PHASE_*) with a pair of events, 0 switches timing off.  Each bracketed phase costs
 * a few microseconds of stream time, so the benchmark brackets only the dominant kernel in its timed region. */
int psgd_timing_enable(psgd_plan* plan, int mask);
/* synchronises the recorded events; ms_sum/count have PSGD_N_PHASES entries; resets */
int psgd_timing_read(psgd_plan* plan, double* ms_sum, int64_t* count);
/* development aid: with PSGD_STREAM_DEBUG=1 in the environment at plan creation the streamed P = M Q
 * kernel records a per-CTA cycle breakdown; this copies it out (8 words per CTA), returns the word count */
int psgd_debug_read(psgd_plan* plan, uint64_t* out, int max_words);
/* kernels launched through this library since load (all plans) */
int64_t psgd_launch_count(void);

/* Stand-alone flat packer behind tensor_buffer.TensorBuffer (tensor_buffer.py:4-36):
 * element-size generic (1, 2, 4 or 8 bytes), layout = torch.cat of the flattened inputs. */
int psgd_tb_create(int n, const int64_t* nelem, int elem_size, int device, psgd_tb** out);
int psgd_tb_destroy(psgd_tb* tb);
int psgd_tb_pack(psgd_tb* tb, const void* const* src, void* flat, psgd_stream stream);
int psgd_tb_unpack(psgd_tb* tb, const void* flat, void* const* dst, psgd_stream stream);

/* The caller-side update that follows reduce() upstream (ddp_init.py:166-178), over all parameters in one
 * launch instead of four per parameter:  m = first_step ? g : momentum * m + g ;  g += m ;  p -= lr * g.
 * nelem: element counts of the parameters; params / grads / momenta: host arrays of fp32 device pointers. */
int psgd_sgd_create(int n, const int64_t* nelem, int device, psgd_sgd** out);
int psgd_sgd_destroy(psgd_sgd* h);
int psgd_sgd_step(psgd_sgd* h, void* const* params, void* const* grads, void* const* momenta, float lr,
                  float momentum, int first_step, psgd_stream stream);

/* One-shot SUM all-reduce of a small fp32 buffer over peer-mapped memory, replacing dist.all_reduce on the packed
 * P || flat and Q buffers (reducer.py:126,132,145) inside one NVLink domain.  peer_in[r] / peer_signal[r]: rank r's
 * input buffer and uint32 signal pad as mapped into THIS process (e.g. torch symmetric memory), world <= 8 entries;
 * out: local result (must not alias any input); slot0: first of `world` pad slots reserved for this call site;
 * epoch: a counter the caller increments on every call at this site (same value on all ranks).  Sums are formed in
 * rank order, so every rank obtains the same bits.  Call sites must alternate between two (buffer, slot0) pairs,
 * as the reducer's P and Q phases do: that alternation is what guarantees a peer has finished reading an input
 * before its owner overwrites it. */
int psgd_peer_allreduce(const void* const* peer_in, void* const* peer_signal, float* out, int64_t n, int rank,
                        int world, int slot0, uint32_t epoch, int device, psgd_stream stream);

/* The same all-reduce through the NVSwitch multicast mapping of the symmetric buffer (`multicast` = the multicast
 * address of the buffer whose per-rank addresses are peer_in[]; torch: symmetric-memory handle .multicast_ptr):
 * rank k sums the k-th slice inside the switch (multimem.ld_reduce) and stores it into every rank's copy
 * (multimem.st), in place; `out[0 .. n)` then receives this rank's copy.  The symmetric buffers must be padded to a
 * multiple of 4 floats (the padding takes part in the sum).  out[i] = sum / out_div for i >= div_from, the plain sum
 * below (the `/= n_workers` of reducer.py:147,167 on the way out; out_div = 1 leaves everything a sum).  Uses signal
 * slots slot0 .. slot0 + 15.  Same alternation rule as above.  blocks_done: one zero-initialised device word owned by
 * this call site (the kernel counts its finished CTAs there and leaves it at zero); two call sites that may overlap
 * on the GPU (different streams, different reducers) must not share one. */
int psgd_peer_allreduce_mc(const void* const* peer_in, void* const* peer_signal, void* multicast, float* out, int64_t n,
                           int rank, int world, int slot0, uint32_t epoch, float out_div, int64_t div_from,
                           uint32_t* blocks_done, int device, psgd_stream stream);

/* ---- all-reduces taken into the kernels -------------------------------------------------------------
 * Upstream runs dist.all_reduce on p_memory, on the flat <=1-D buffer and on q_memory (reducer.py:126, 132, 145), each
 * followed by `/= n_workers`.  Inside one NVLink domain those three collectives need no launch of their own: the
 * kernels that consume the sums read them "summed over all ranks" straight from symmetric memory.
 *
 * psgd_peer describes the two exchange buffers of a reducer as mapped into THIS process (e.g. torch symmetric memory):
 *   pf_bufs[r] / q_bufs[r]  rank r's  P || flat  buffer (sum n*r + vec elements) and Q buffer (sum m*r elements), each
 *                           padded to a multiple of 4 floats and 16-byte aligned; every rank writes its own copy
 *                           (psgd_p_from_m* into pf_bufs[rank], psgd_q_from_mt into q_bufs[rank]);
 *   pf_signals[r] / q_signals[r]  rank r's uint32 signal pad of that buffer; slots slot0 .. slot0 + world - 1 are used;
 *   pf_multicast / q_multicast    NVSwitch multicast address of the buffer (mode 1), else NULL;
 *   mode 1: one multimem.ld_reduce per element (the switch adds the ranks' copies); mode 2: one load per rank, added in
 *           rank order (bit-identical on all ranks by construction, no multicast mapping needed).
 * Every rank must issue the same sequence of *_peer calls (the handle counts barrier epochs).  Peer memory is read with
 * plain and multimem loads only.  A crashed peer makes the waiting kernel trap after ~10 s instead of hanging. */
typedef struct psgd_peer psgd_peer;
int psgd_peer_create(int world, int rank, int device, const void* const* pf_bufs, void* const* pf_signals,
                     const void* pf_multicast, const void* const* q_bufs, void* const* q_signals,
                     const void* q_multicast, int slot0, int mode, psgd_peer** out);
int psgd_peer_destroy(psgd_peer* peer);
/* reducer.py:126-128 + 136-137 (+ 131-133, 167 for the flat buffer) as ONE launch: a flag barrier ("every rank's P is
 * written"), then each Gram-Schmidt CTA loads its rows summed over the ranks, scales by 1 / n_workers, orthogonalises
 * and stores P_hat to the local p_mem; spare CTAs leave mean(flat) in flat_out (NULL: the flat part is not touched,
 * e.g. later power iterations).  Matrices too tall for the register-resident kernels fall back to a stand-alone
 * reduction into p_mem followed by psgd_orthogonalize. */
int psgd_orthogonalize_peer(psgd_plan* plan, psgd_peer* peer, float* p_mem, float* flat_out, psgd_stream stream);
/* reducer.py:145-147 + 158-163 (+ 168) as ONE launch: the streamed finalize kernel first takes the flag barrier ("every
 * rank's Q is written"), its CTAs sum one slice of Q each over the ranks, divide by n_workers and store the mean to the
 * local q_mem, a grid-wide arrival count follows, then the kernel runs as at one worker.  flat: the averaged flat
 * buffer psgd_orthogonalize_peer produced (scattered to vec_out as is). */
int psgd_finalize_peer(psgd_plan* plan, psgd_peer* peer, const float* p_mem, float* q_mem, const float* flat,
                       psgd_stream stream);
/* K1 of a later power iteration at N > 1: the same Q prologue, then P = M mean(Q) into p_mem (this rank's part of
 * pf_bufs).  q_mem: local buffer receiving mean(Q). */
int psgd_p_from_m_peer(psgd_plan* plan, psgd_peer* peer, float* p_mem, float* q_mem, int fuse_add, psgd_stream stream);

/* Push form of the same idea (needs the NVSwitch multicast mapping): every rank owns a RECEIVE buffer with one slot per
 * source rank (slot stride in floats, a multiple of 4; rx_*_local = this rank's buffer, rx_*_multicast = the multicast
 * address of the same symmetric allocation).  The producing kernels store their results with multimem.st into slot
 * [rank] of every rank's buffer while they run (psgd_p_from_m_push, psgd_q_from_mt_push), so that after the flag barrier
 * psgd_orthogonalize_peer / psgd_finalize_peer / the Q prologue only add up W LOCAL slots, in slot order (bit-identical
 * on all ranks by construction): no load crosses NVLink on the critical path. */
int psgd_peer_create_push(int world, int rank, int device, void* rx_pf_local, void* rx_pf_multicast, int64_t pf_stride,
                          void* const* pf_signals, void* rx_q_local, void* rx_q_multicast, int64_t q_stride,
                          void* const* q_signals, int slot0, psgd_peer** out);
/* K1 (reducer.py:121-123; with_flat: + the flat gather of :131) with its output pushed to every rank.  reduce_q_first:
 * a later power iteration, Q is summed over the ranks in the kernel's prologue first (as psgd_p_from_m_peer). */
int psgd_p_from_m_push(psgd_plan* plan, psgd_peer* peer, float* q_mem, int fuse_add, int with_flat, int reduce_q_first,
                       psgd_stream stream);
/* K3 (reducer.py:140-142) with its output pushed to every rank. */
int psgd_q_from_mt_push(psgd_plan* plan, psgd_peer* peer, const float* p_mem, void* workspace, psgd_stream stream);

/* In-place modified Gram-Schmidt of one n x r row-major matrix: the module-level
 * orthogonalize(matrix) of reducer.py:180-191. */
int psgd_orthogonalize_one(float* mat, int64_t n, int r, int device, psgd_stream stream);

/* No counterpart in the reference (which allocates freely through PyTorch).  Q = M^T P for ONE n x m
 * row-major fp32 matrix (p_hat: n x r, q_out: m x r, both row-major) without a plan and without any
 * allocation, for callers that cannot obtain device memory at this point: a process whose framework
 * allocator already owns the GPU, a hot loop that must not take the device-wide synchronisation of
 * cudaMalloc / cudaFree, a host that forbids driver allocations after start-up.  The split-n partial sums
 * live in an array built into the library (PSGD_NOALLOC_SCRATCH_BYTES per GPU, held until exit).
 * Limit: ceil(n / T) * m * r * 4 bytes must fit that array, T = clamp(ceil(n / 32), 128, 512); larger
 * requests fail with a message naming both numbers.  The first call on a device looks the array's address
 * up and should be made before any stream capture starts.  All calls share the array: nothing in it needs
 * clearing between calls (every partial that is read is overwritten first), but calls that overlap on the
 * GPU race -- keep them on one stream or order them. */
#define PSGD_NOALLOC_SCRATCH_BYTES (4 * 1024 * 1024)
int psgd_q_from_mt_noalloc(const float* mat, const float* p_hat, float* q_out, int64_t n, int64_t m,
                           int r, int device, psgd_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* PSGD_B200_H */
