/*
 * conceptattention_b200.h — C ABI of the B200-native ConceptAttention hot path.
 *
 * Drop-in boundary.  The reference (manuragkhullar/ConceptAttention @ 401dc36) is
 * pure Python/PyTorch and has no FFI; the entry points below are what a ctypes
 * binding on the reference side would call in place of the PyTorch op sequences
 * cited on each function (paths under concept_attention/).  INTEGRATION.md shows
 * that binding.
 *
 * Conventions
 *  - every pointer is a raw DEVICE pointer unless the name ends in _host;
 *  - activations are bf16 (uint16 storage), accumulators/heatmaps fp32, masks
 *    int64 / uint8; all tensors dense row-major with the shapes given;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *  - functions never allocate device memory (except *_host wrappers), never
 *    synchronise (except *_host wrappers) and never throw; they return
 *    CA_OK or a negative ca_status code.  Thread-safe for disjoint buffers;
 *  - sm_100a only: on any other device ca_device_check / the kernels' launch
 *    return CA_ERR_ARCH / CA_ERR_CUDA.  There is no fallback path.
 *  - token order inside one (batch, head) sequence is [img | txt | concept]:
 *    N_tot = P + T + C.  Attention is permutation-equivariant, so this differs
 *    from the reference's cat(txt,img) / cat(concept,img) only in layout.
 */
#ifndef CONCEPTATTENTION_B200_H_
#define CONCEPTATTENTION_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int ca_status;
enum {
  CA_OK = 0,
  CA_ERR_INVALID_ARG = -1,   /* null pointer, negative size, bad enum            */
  CA_ERR_UNSUPPORTED = -2,   /* shape outside what the kernels are built for    */
  CA_ERR_ALIGNMENT = -3,     /* pointer / stride not 16-byte aligned            */
  CA_ERR_ARCH = -4,          /* device is not sm_100                            */
  CA_ERR_CUDA = -5,          /* CUDA runtime/driver error; see ca_last_cuda_error */
  CA_ERR_SCRATCH_LIMIT = -6  /* the call needs more scratch than the static no-alloc array holds */
};

#define CA_HEAD_DIM 128      /* Flux head dim (hidden 3072 / 24 heads)           */
#define CA_MAX_CONCEPTS 32

/* concept_attention_pipeline.py:64-71 */
enum { CA_NORM_SOFTMAX = 0, CA_NORM_SPARSEMAX = 1, CA_NORM_ENTMAX15 = 2, CA_NORM_NONE = 3 };
/* modified_double_stream_block.py:126-168 (joint_attention_kwargs ablations) */
enum { CA_ATTN_BOTH = 0, CA_ATTN_CROSS_ONLY = 1, CA_ATTN_SELF_ONLY = 2, CA_ATTN_NEITHER = 3 };

const char* ca_version(void);
const char* ca_status_string(ca_status s);
int ca_last_cuda_error(void);                 /* cudaError_t / CUresult of the last failure */
ca_status ca_device_check(int device);        /* CA_OK iff compute capability 10.x */

/* ---- K1a: LayerNorm(eps 1e-6, no affine) + adaLN modulate -------------------
 * out[b,n,:] = (1 + scale) * LN(x[b,n,:]) + shift, fp32 maths, bf16 out.
 * Rows n <  split take (shift,scale) from mod0[b], rows n >= split from mod1[b]
 * (concept rows use txt weights but their own modulation vector).
 * Replaces modified_double_stream_block.py:88-89,94-95,100-101 and
 * modified_single_stream_block.py:47.  mod rows are bf16 [*, mod_stride];
 * shift at +shift_off, scale at +scale_off (elements). */
ca_status ca_ln_modulate(const void* x, void* out, int B, int N, int D,
                         const void* mod0, const void* mod1, int split,
                         int64_t mod_stride, int shift_off, int scale_off, void* stream);

/* ---- K1c: gated residual, optionally followed by LN + modulate --------------
 * x[b,n,:] += gate * (y[b,n,:] (+ 0));  if out != NULL:
 * out[b,n,:] = (1 + scale2) * LN(x) + shift2.   y has row stride y_stride.
 * Replaces modified_double_stream_block.py:194-202 (the elementwise parts) and
 * modified_single_stream_block.py:56. */
ca_status ca_gate_residual_ln_modulate(void* x, const void* y, int64_t y_stride, void* out,
                                       int B, int N, int D,
                                       const void* mod0, const void* mod1, int split,
                                       int64_t mod_stride, int gate_off, int shift_off,
                                       int scale_off, void* stream);

/* ---- K1b: qkv split + QK-RMSNorm + RoPE + head-major layout -----------------
 * qkv_a: bf16 [B, Na, row_stride_a] rows of the image stream (first 3*H*128
 * columns are "(K H D)"), normalised with (q_scale_a, k_scale_a);
 * qkv_b: bf16 [B, Nb, row_stride_b] rows of [txt | concept] with (q_scale_b,
 * k_scale_b); may be NULL with Nb = 0 (single-stream blocks).
 * rope: fp32 [Na+Nb, 64, 2] = (cos, sin) of the per-token angles
 * (flux/src/flux/math.py:15-22), token order [a | b].
 * q,k,v: bf16 [B,H,Na+Nb,128]; q_pre (nullable): post-RMSNorm PRE-RoPE q, the
 * reference's cross_attention_*_vectors (modified_double_stream_block.py:189-190).
 * Replaces :91-92,97-98,103-104,110,125 and flux/src/flux/modules/layers.py:63-84,
 * flux/src/flux/math.py:25-30. */
ca_status ca_qkv_prep(const void* qkv_a, int64_t row_stride_a, int Na,
                      const void* qkv_b, int64_t row_stride_b, int Nb,
                      const void* q_scale_a, const void* k_scale_a,
                      const void* q_scale_b, const void* k_scale_b,
                      const float* rope, int B, int H,
                      void* q, void* k, void* v, void* q_pre, void* stream);
/* Grouped form for the sequence-parallel exchange (B = 1): heads [g*G, (g+1)*G) of q/k/v are
 * written at  base + g*group_stride + ((h % G)*(Na+Nb) + n)*128  (elements), so that with
 * q = buf, k = buf + G*N*128, v = buf + 2*G*N*128 and group_stride = 3*G*N*128 the buffer is
 * [H/G][3][G][N][128]: one contiguous all-to-all chunk per destination rank.  q_pre keeps the
 * plain [H][N][128] layout.  group_heads = 0 is ca_qkv_prep. */
ca_status ca_qkv_prep_grouped(const void* qkv_a, int64_t row_stride_a, int Na,
                              const void* qkv_b, int64_t row_stride_b, int Nb,
                              const void* q_scale_a, const void* k_scale_a,
                              const void* q_scale_b, const void* k_scale_b,
                              const float* rope, int B, int H,
                              void* q, void* k, void* v, void* q_pre,
                              int group_heads, int64_t group_stride, void* stream);

/* ---- K2: one-way joint attention (tcgen05 / TMEM / TMA) ---------------------
 * q,k,v bf16 [B,H,P+T+C,128], order [img|txt|concept].
 *   rows [0,P+T)      attend keys [0,P+T)                     (txt (+) img, :106-119)
 *   rows [P+T,P+T+C)  attend keys [0,P) U [P+T,P+T+C)         (concept (+) img, :121-168)
 * mode selects the ablation for the concept rows (CA_ATTN_*).  Image rows of the
 * reference's second attention are never computed.
 * out_img bf16 [B,P,H*128]; out_txtc bf16 [B,T+C,H*128] ("B L (H D)", :170-176).
 * softmax scale = 1/sqrt(128).  C may be 0 (single-stream blocks; then T rows
 * still go to out_txtc). */
ca_status ca_joint_attention(const void* q, const void* k, const void* v,
                             int B, int H, int P, int T, int C, int mode,
                             void* out_img, void* out_txtc, void* stream);

/* Same, with a caller-owned scratch buffer that lets the launch balance its last wave:
 * when the query-tile units do not fill a whole number of waves of the GPU's SMs, the key
 * range of the left-over units is cut into one equal share per SM (stream-K) and the pieces
 * of a unit are merged in fp32 by whichever piece finishes last (fixed merge order: results
 * do not depend on timing).  `workspace` needs ca_joint_attention_workspace_bytes(...) bytes
 * (0 = this shape is never split), 16-byte aligned, contents irrelevant; with a null or too
 * small workspace the call behaves exactly like ca_joint_attention. */
size_t ca_joint_attention_workspace_bytes(int B, int H, int P, int T, int C, int mode);
ca_status ca_joint_attention_ws(const void* q, const void* k, const void* v,
                                int B, int H, int P, int T, int C, int mode,
                                void* out_img, void* out_txtc,
                                void* workspace, size_t workspace_bytes, void* stream);

/* Same, with the kernel and the tail plan chosen PER CALL (no process-wide switches):
 *   impl     CA_K2_AUTO  the default of the two calls above (measured inside the pipeline on B200):
 *                        CA_K2_V11, except CA_K2_V11P for many short units (<= 16 key steps in >= 4
 *                        waves of the SMs) and for long units (> 64 key steps) in fewer than eight
 *                        waves whose left-over wave is worth balancing with its key-step ranges
 *            CA_K2_V3    round-1 kernel: P overlays S in TMEM, two softmax warpgroups share a
 *                        tile; the only one with the stream-K tail (uses `workspace`)
 *            CA_K2_V7    P through shared memory, a tile's score buffer is free once loaded
 *            CA_K2_V11   one score buffer shared by the two tiles, P in TMEM, five K/V stages
 *            CA_K2_V11P  the same kernel persistent: one CTA per SM walks the unit list, the mbarrier
 *                        phases run on from unit to unit (no TMEM / barrier set-up or CTA launch per unit);
 *                        with `workspace`: whole waves of units in lockstep, then the left-over units
 *                        cut into equal key-step ranges per CTA (fp32 merge in fixed order)
 *   plan_sms 0 = plan the stream-K tail for the real SM count with the measured thresholds,
 *            < 0 = never split, n > 0 = plan for n SMs (parity tests of the tail on any GPU).
 *            CA_K2_V11P balances with key-step ranges over its persistent CTAs instead (same
 *            workspace mechanism): 0 = when the measured rule says they pay, < 0 = never,
 *            > 0 = whenever structurally possible (tests and A/B runs)
 * All kernels compute the same function to bf16 round-off; CA_ERR_INVALID_ARG for an unknown impl. */
enum { CA_K2_AUTO = 0, CA_K2_V3 = 1, CA_K2_V7 = 2, CA_K2_V11 = 3, CA_K2_V11P = 4 };
size_t ca_joint_attention_workspace_bytes_ex(int B, int H, int P, int T, int C, int mode,
                                             int impl, int plan_sms);
ca_status ca_joint_attention_ex(const void* q, const void* k, const void* v,
                                int B, int H, int P, int T, int C, int mode,
                                void* out_img, void* out_txtc,
                                void* workspace, size_t workspace_bytes,
                                int impl, int plan_sms, void* stream);

/* ---- K2 without any allocation by anybody -------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  ca_joint_attention_ex with the scratch of its tail balancing (the stream-K partials of
 * CA_K2_V3 and the key-step ranges of CA_K2_V11P) taken from ONE file-scope static __device__ array of
 * CA_K2_NOALLOC_MAX_BYTES (40 MiB, held by every GPU that loads the library until exit) instead of
 * a caller-owned workspace.  The entry point calls no cudaMalloc / cudaFree / allocating library
 * routine; inputs and outputs are still the caller's.  A call whose plan needs more scratch than
 * the array holds is rejected with CA_ERR_SCRATCH_LIMIT before anything is launched
 * (ca_joint_attention_workspace_bytes_ex tells the need in advance; plan_sms < 0 never needs any).
 * Every call on a device shares the array: the arrival counters are cleared on the stream at the
 * start of each call, the partials are overwritten before they are read, but two calls that
 * overlap on the GPU (different streams, or graphs replayed concurrently) RACE on it — serialise
 * them.  Results are identical to ca_joint_attention_ex with a caller workspace. */
#define CA_K2_NOALLOC_MAX_BYTES (40u << 20)
ca_status ca_joint_attention_noalloc(const void* q, const void* k, const void* v,
                                     int B, int H, int P, int T, int C, int mode,
                                     void* out_img, void* out_txtc,
                                     int impl, int plan_sms, void* stream);

/* ---- K3+K4: heatmap logits + normalise over concepts + accumulate -----------
 * logits[b,c,p] = sum_{s<nseg} sum_{j<seg_len} img[b,s,p,j] * con[b,s,c,j]
 * where element (b,s,p,j) sits at img + b*img_stride_b + s*img_stride_s +
 * p*img_stride_p + j (bf16 elements); same for con with con_stride_{b,s,c}.
 * Output space: nseg=1, seg_len=H*128; cross-attention space: nseg=H, seg_len=128
 * (concept_attention_pipeline.py:43-61).
 * acc[b,c,p] (fp32) += norm_c(logits[b,:,p])  with norm in CA_NORM_* (:64-71);
 * CA_NORM_NONE accumulates raw logits (multi_class_segmentation.py:59-63).
 * round_bf16 != 0 reproduces the reference's bf16 rounding points (logits and
 * probabilities rounded to bf16 before use). */
ca_status ca_heatmap_accumulate(const void* img, int64_t img_stride_b, int64_t img_stride_s,
                                int64_t img_stride_p,
                                const void* con, int64_t con_stride_b, int64_t con_stride_s,
                                int64_t con_stride_c,
                                int nseg, int seg_len, int B, int P, int C,
                                int norm, int round_bf16, float* acc, void* stream);

/* Same pass, additionally keeping THIS (timestep, layer) pair's normalised map: the
 * reference's heatmaps[t][:, l] before the mean (concept_attention_pipeline.py:76-77;
 * experiments/per_layer_segmentation), BASELINE config 4's per-layer / per-timestep mode.
 *   pair[slot][b,c,p] = norm_c(logits[b,:,p]),  slot = (*pair_slot_dev or 0) + pair_slot_off,
 * slots `pair_stride` floats apart (>= B*C*P).  pair_slot_dev is a DEVICE int read by the
 * kernel, so one captured CUDA graph of a forward serves every timestep; pair == NULL is
 * ca_heatmap_accumulate. */
ca_status ca_heatmap_accumulate_pair(const void* img, int64_t img_stride_b, int64_t img_stride_s,
                                     int64_t img_stride_p,
                                     const void* con, int64_t con_stride_b, int64_t con_stride_s,
                                     int64_t con_stride_c,
                                     int nseg, int seg_len, int B, int P, int C,
                                     int norm, int round_bf16, float* acc, float* pair,
                                     int64_t pair_stride, const int* pair_slot_dev,
                                     int pair_slot_off, void* stream);

/* Both spaces of one layer in ONE launch: problem 0 (output space: img0 / con0, nseg0 = 1) and
 * problem 1 (cross-attention space: the pre-RoPE queries, nseg1 = H) share B, P, C, the total vector
 * length nseg * seg_len, the normaliser and the per-pair buffer; each has its own accumulator and
 * its own slot offset.  Same results as two ca_heatmap_accumulate_pair calls (which is what this
 * entry does when a problem does not fit the TMA/tcgen05 kernel). */
ca_status ca_heatmap_accumulate_both(
    const void* img0, int64_t img0_stride_b, int64_t img0_stride_s, int64_t img0_stride_p,
    const void* con0, int64_t con0_stride_b, int64_t con0_stride_s, int64_t con0_stride_c,
    int nseg0, int seg_len0, float* acc0, int pair_slot_off0,
    const void* img1, int64_t img1_stride_b, int64_t img1_stride_s, int64_t img1_stride_p,
    const void* con1, int64_t con1_stride_b, int64_t con1_stride_s, int64_t con1_stride_c,
    int nseg1, int seg_len1, float* acc1, int pair_slot_off1,
    int B, int P, int C, int norm, int round_bf16, float* pair, int64_t pair_stride,
    const int* pair_slot_dev, void* stream);

/* K3 only: partial logits of a head/feature slice, written (not accumulated) as
 * fp32 [B,C,P] — the buffer a head-sharded run all-reduces before K4. */
ca_status ca_heatmap_logits(const void* img, int64_t img_stride_b, int64_t img_stride_s,
                            int64_t img_stride_p,
                            const void* con, int64_t con_stride_b, int64_t con_stride_s,
                            int64_t con_stride_c,
                            int nseg, int seg_len, int B, int P, int C,
                            float* logits, void* stream);

/* K4 only: acc += norm_c(logits) on fp32 [B,C,P]. */
ca_status ca_norm_accumulate(const float* logits, int B, int P, int C, int norm,
                             int round_bf16, float* acc, void* stream);

/* K4 with the per-pair output of ca_heatmap_accumulate_pair (pair fp32 [B,C,P], nullable). */
ca_status ca_norm_accumulate_pair(const float* logits, int B, int P, int C, int norm,
                                  int round_bf16, float* acc, float* pair, void* stream);

/* utils.py:35-44 applied over the concept axis of a [C,K] bf16 matrix
 * (normalize_concepts=True, concept_attention_pipeline.py:53-54). */
ca_status ca_concept_linear_normalize(const void* con, void* out, int C, int64_t K,
                                      void* stream);

/* ---- K5: finalise and heatmap -> mask ----------------------------------------
 * heat = acc * inv_count (mean over selected (t,l), :78-82); optional bf16 rounding. */
ca_status ca_finalize_heatmaps(const float* acc, float* heat, int64_t n, float inv_count,
                               int round_bf16, void* stream);
/* 3x3 gaussian blur sigma=1, reflect padding (segmentation.py:57-58). maps [M,h,w] fp32. */
ca_status ca_gaussian_blur3(const float* in, float* out, int M, int h, int w, void* stream);
/* mask[b,p] = argmax_c heat[b,c,p] (first index on ties), indices < num_background -> 0
 * (multi_class_segmentation.py:77; run_multi_class_seg_experiment.py:25-31). */
ca_status ca_argmax_mask(const float* heat, int64_t* mask, int B, int C, int P,
                         int num_background, void* stream);
/* mask[m,p] = heat[m,p] > mean_p(heat[m,:])   (segmentation.py:60-61). M = B*C maps. */
ca_status ca_binary_masks(const float* heat, uint8_t* mask, int M, int P, void* stream);

/* ---- K6: segmentation-metric counters ------------------------
 * One pass over a predicted and a ground-truth label map (integer-valued floats, ignore label -1)
 * replaces the reference's host-side numpy counting (concept_attention/utils.py:71-85
 * batch_pix_accuracy and :88-108 batch_intersection_union, which shift labels by +1 and histogram
 * over (1, nclass)).  counts is device memory, uint64 [2 + 3*nclass], overwritten:
 *   [0] pixel_correct  [1] pixel_labeled  [2..) area_inter[nclass], area_pred[nclass], area_lab[nclass]
 * (area_union = area_pred + area_lab - area_inter is left to the caller, as in the reference). */
ca_status ca_segmentation_counts(const float* predict, const float* target, int64_t n, int nclass,
                                 uint64_t* counts, void* stream);

/* ---- K7: layout kernels of the sequence-parallel exchange (multi-GPU, B = 1) ----
 * W ranks each own Pl image rows, Tl text rows (and a replica of the C concept rows) for every
 * token-wise op; attention runs head-split (Hl = H/W heads per rank over all rows).  An all-to-all
 * (NCCL, caller side) moves contiguous per-rank chunks; these are the permutations around it.
 * The reference has no multi-GPU path; K2 row order is [img | txt | concept].
 *   ca_seq_unpack_qkv  recv [W][3][Hl][Pl+Tl+C][128] -> q,k,v [Hl][W*Pl + W*Tl + C][128]
 *                      (concept rows from chunk `self`)
 *   ca_seq_pack_out    K2 outputs out_img [W*Pl][Hl*128], out_txtc [W*Tl+C][Hl*128]
 *                      -> send [W][Pl+Tl+C][Hl*128]
 *   ca_seq_unpack_out  recv [W][Pl+Tl+C][Hl*128] -> attn_img [Pl][W*Hl*128], attn_tc [Tl+C][W*Hl*128] */
ca_status ca_seq_unpack_qkv(const void* recv, int W, int self, int Hl, int Pl, int Tl, int C,
                            void* q, void* k, void* v, void* stream);
ca_status ca_seq_pack_out(const void* out_img, const void* out_txtc, int W, int Hl, int Pl, int Tl,
                          int C, void* send, void* stream);
ca_status ca_seq_unpack_out(const void* recv, int W, int Hl, int Pl, int Tl, int C,
                            void* attn_img, void* attn_tc, void* stream);

/* Fused form of the two exchanges: the kernels PULL their input straight out of the peers' buffers
 * over NVLink instead of unpacking what a collective delivered.  `peers_dev` is a DEVICE array of W
 * pointers, entry w = base of rank w's symmetric allocation as mapped into this process (CUDA VMM /
 * IPC handles exchanged once by the host side; torch.distributed._symmetric_memory provides both
 * the mapping and the cross-GPU barrier); all ranks use the same byte offsets inside it.
 *   ca_seq_gather_qkv  == ca_seq_unpack_qkv with recv[w] := (rank w's send [W][3][Hl][Nl][128] at
 *                      send_off)[self]: replaces all-to-all + unpack
 *   ca_seq_gather_out  attn_img / attn_tc of this rank's rows from every peer's K2 outputs out_img
 *                      [W*Pl][Hl*128] (img_off) and out_txtc [W*Tl+C][Hl*128] (tc_off): replaces
 *                      ca_seq_pack_out + all-to-all + ca_seq_unpack_out
 * The caller orders the ranks: a barrier between the writes of the buffers and these reads; the next
 * write of a buffer happens after the NEXT barrier, which every peer reaches only after its reads. */
ca_status ca_seq_gather_qkv(const void* const* peers_dev, int64_t send_off, int W, int self, int Hl,
                            int Pl, int Tl, int C, void* q, void* k, void* v, void* stream);
ca_status ca_seq_gather_out(const void* const* peers_dev, int64_t img_off, int64_t tc_off, int W,
                            int self, int Hl, int Pl, int Tl, int C, void* attn_img, void* attn_tc,
                            void* stream);

/* ---- host-buffer form of compute_heatmaps_from_vectors (:29-91) --------------
 * image_host  bf16 [Nt,L,B,P,K]   (or per-head [Nt,L,B,H,P,d] when per_head != 0)
 * concept_host bf16 [Nt,L,B,C,K]  (or [Nt,L,B,H,C,d])
 * Selected (timestep, layer) pairs are copied H2D, reduced on the device with
 * K3+K4+K5 and the [B,C,P] fp32 mean heatmap is copied back.  Synchronises. */
ca_status ca_heatmaps_from_vectors_host(const void* image_host, const void* concept_host,
                                        int Nt, int L, int B, int H, int P, int C, int d,
                                        int per_head, const int* timesteps, int n_timesteps,
                                        const int* layers, int n_layers, int norm,
                                        int round_bf16, float* heat_host);

#ifdef __cplusplus
}
#endif
#endif /* CONCEPTATTENTION_B200_H_ */
