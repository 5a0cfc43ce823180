/*
 * tss_b200.h -- C ABI of the B200-native self-training label path.
 *
 * The reference (MuliangDu-sudo/thermal-semantic-segmentation @ 8a37af1) is pure
 * Python and has no FFI of its own; the boundary it offers is a set of Python
 * callables on torch tensors / numpy arrays.  Each entry
 * point below replaces the arithmetic of the reference lines cited beside it and is
 * what a ctypes binding on the reference side binds.
 *
 * Conventions (all entry points):
 *   - every pointer is a raw DEVICE pointer unless its name ends in _host;
 *   - tensors are dense, row-major ("NCHW contiguous" as the reference keeps them);
 *   - the caller owns all memory: the library never allocates, frees or synchronises;
 *     scratch space is passed in, its size comes from the matching *_workspace_bytes;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*) and returns at once;
 *   - return value: 0 on success, a TSS_E_* code for a rejected argument, or
 *     1000 + cudaError_t for a CUDA failure; tss_last_error_string() explains it
 *     (thread-local).  No C++ exceptions cross this boundary.  The library keeps no mutable state between
 *     calls other than that message and an atomic launch counter (tss_launch_count): kernel variants are
 *     selected per call (`variant` / `flags` / `form` arguments, 0 = automatic), never through globals;
 *   - num_classes is limited to TSS_MAX_CLASSES.
 */
#ifndef TSS_B200_H_
#define TSS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TSS_VERSION 100
#define TSS_MAX_CLASSES 32

/* status codes */
#define TSS_OK 0
#define TSS_E_ARG 1        /* null pointer / non-positive extent / bad enum */
#define TSS_E_CLASSES 2    /* num_classes outside [1, TSS_MAX_CLASSES] or unsupported by this kernel */
#define TSS_E_ALIGN 3      /* pointer alignment requirement not met */
#define TSS_E_WORKSPACE 4  /* workspace too small */
#define TSS_E_CUDA 1000    /* 1000 + cudaError_t */

/* integer element types of label maps */
#define TSS_U8 0
#define TSS_I32 1
#define TSS_I64 2

int tss_version(void);
const char* tss_last_error_string(void);
/* number of kernels this library has launched in the calling process (bench "gpu_launches"; atomic) */
uint64_t tss_launch_count(void);
/* sha256 of the sources this library was compiled from (the host mirror refuses a stale library) */
const char* tss_source_digest(void);

/* ---------------------------------------------------------------------------------------------
 * A14  confusion matrix -- replaces utils/eval_tools.py:4-9 (_fast_hist) and the per-image
 * accumulation of evaluate() :49-51.  hist[C*C] (int64, rows = ground truth, cols = prediction) is
 * ACCUMULATED (+=).  A pixel counts iff 0 <= gt < C.  The reference does not range-check pred: a
 * counted pixel with pred outside [0, C) makes np.bincount / reshape raise (or silently alias another
 * bin); here such pixels are left out of hist and counted in *bad_keys (uint64, accumulated; may be
 * NULL), and the host mirror raises ValueError when it is non-zero.
 * pred/gt: n elements of `dtype` (TSS_U8 / TSS_I32 / TSS_I64).
 * variant: TSS_HIST_AUTO, or force the warp-private shared-memory atomics kernel (TSS_HIST_ATOMIC) / the packed
 * per-thread 8-bit counters kernel where it applies (TSS_HIST_PACKED); all give the same matrix. */
#define TSS_HIST_AUTO 0
#define TSS_HIST_ATOMIC 1
#define TSS_HIST_PACKED 2
int tss_confusion(const void* pred, const void* gt, int dtype, int64_t n, int num_classes,
                  int64_t* hist, uint64_t* bad_keys, int variant, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A2   low-resolution softmax -- generate_pseudo_label.py:75 (soft branch) and the first softmax
 * of :82-83.  logits/probs: [N, C, P] fp32 (P = h*w). */
int tss_softmax_nchw(const float* logits, float* probs, int64_t n_img, int num_classes, int64_t plane,
                     void* stream);

/* A1-A4 fused pseudo-label kernel.  One pass: bilinear upsample (align_corners=True) of the C planes
 * of `src` [N,C,h,w] to [H,W] -> per-pixel reduction over classes -> storage casts -> threshold.
 *   mode TSS_PL_UP_SOFTMAX       src = logits;  softmax AFTER upsampling, then max/argmax
 *                                (self_training.py:53-55 order + generate_pseudo_label.py:85; with
 *                                 H==h, W==w this is exactly the no-flip hard branch :85)
 *   mode TSS_PL_UP_ARGMAX        src = logits;  argmax only, conf = max logit
 *                                (segmentation_evaluate.py:38-42)
 *   mode TSS_PL_PROB_UP          src = probabilities; conf = max, label = argmax
 *   mode TSS_PL_PROB_UP_SOFTMAX  src = probabilities; a second softmax over the upsampled
 *                                probabilities, then max (generate_pseudo_label.py:82-85 as committed)
 *   flip TSS_FLIP_NONE | TSS_FLIP_CLASS (torch.fliplr on NCHW = reversed class axis, what
 *        generate_pseudo_label.py:10,84 really does) | TSS_FLIP_WIDTH (the intended mirror):
 *        out = (up(src) + flip(up(src_flip))) / 2 before the reduction (:84).  src_flip may be NULL
 *        iff flip == TSS_FLIP_NONE.
 * Outputs (each may be NULL): label_out [N,H,W] uint8 (:91), conf16_out [N,H,W] IEEE fp16 bits,
 * round-to-nearest-even (:94), conf32_out fp32, label_thr_out [N,H,W] uint8 = the threshold block
 * data/freiburg_dataset.py:115-123 applied to (label, fp16 conf): ignore_index where
 * fp16(conf) <= fp16(thr_host[label]) or keep_host[label] == 0.  thr_host / keep_host: HOST arrays of
 * num_classes entries (NULL = no threshold / keep all).
 * flags: 0, or TSS_PL_FLAG_FORCE_TILE (use the shared-memory tile kernel even where a faster form applies: A/B
 * of the kernels) | TSS_PL_FLAG_SM_WARPS(w) (size the one-wave grid of the register-resident kernel for w warps
 * per SM instead of the whole register file, so that the HBM-bound accumulate kernel of the prototype stage can
 * run on the same SMs at the same time: tss_proto_accumulate(consumer_warps = 4)). */
#define TSS_PL_FLAG_FORCE_TILE 1
/* two output rows that share a bracket of source rows are processed in one instruction stream (instruction-level
 * parallelism across their dependency chains; 170 registers per thread, 12 warps per SM) */
#define TSS_PL_FLAG_ROWS2 2
/* the 13-class register-resident kernel compiled under a 96-register cap (20 warps per SM alone): more of its warps fit
 * beside a co-resident kernel's CTA; same results bit for bit */
#define TSS_PL_FLAG_REGS96 4
#define TSS_PL_FLAG_SM_WARPS(w) (((w) & 0xff) << 8)
/* co-resident scheduling: fixed CTA size in warps (3..8), and the number of waves the rows are cut into (1..15): many
 * small CTAs let the block scheduler refill an SM with pseudo-label CTAs as the accumulate kernel's CTAs retire */
#define TSS_PL_FLAG_CTA_WARPS(c) (((c) & 0xf) << 16)
#define TSS_PL_FLAG_WAVES(n) (((n) & 0xf) << 20)
#define TSS_PL_UP_SOFTMAX 0
#define TSS_PL_UP_ARGMAX 1
#define TSS_PL_PROB_UP 2
#define TSS_PL_PROB_UP_SOFTMAX 3
#define TSS_FLIP_NONE 0
#define TSS_FLIP_CLASS 1
#define TSS_FLIP_WIDTH 2
int tss_pseudo_label(const float* src, const float* src_flip, int64_t n_img, int num_classes, int h, int w,
                     int H, int W, int mode, int flip, uint8_t* label_out, uint16_t* conf16_out,
                     float* conf32_out, uint8_t* label_thr_out, const float* thr_host,
                     const uint8_t* keep_host, int ignore_index, int flags, void* stream);

/* A1-A4 + A14 in one pass: tss_pseudo_label plus the confusion matrix of (gt, label) -- utils/eval_tools.py:4-9
 * (_fast_hist) applied to the labels of generate_pseudo_label.py:85 / segmentation_evaluate.py:42 (upsample ->
 * argmax -> evaluate is the loop body of seg_validate, segmentation_evaluate.py:38-44,77-79).
 * gt [N,H,W] uint8; hist [C*C] int64 ACCUMULATED, rows = ground truth; a pixel counts iff gt < C (labels are
 * always in range here, so there is no bad-key counter).  Where the register-resident kernel applies the histogram is
 * accumulated inside it (one 16-bit load per pixel pair, shared-memory atomics, one global atomic per non-zero bin and
 * CTA); for every other shape the histogram kernel of tss_confusion runs on label_out (which must then be non-NULL). */
int tss_pseudo_label_hist(const float* src, const float* src_flip, int64_t n_img, int num_classes, int h, int w,
                          int H, int W, int mode, int flip, uint8_t* label_out, uint16_t* conf16_out,
                          float* conf32_out, uint8_t* label_thr_out, const float* thr_host,
                          const uint8_t* keep_host, int ignore_index, const uint8_t* gt, int64_t* hist, int flags,
                          void* stream);

/* A1   plain bilinear upsample, align_corners=True: self_training.py:62,69-77 (soft labels, feat, out).
 * src [planes, h, w] -> dst [planes, H, W] fp32. */
int tss_upsample_bilinear(const float* src, float* dst, int64_t planes, int h, int w, int H, int W,
                          void* stream);

/* adjoint of tss_upsample_bilinear (the backward of nn.Upsample at self_training.py:53-57,75):
 * grad_in [planes,h,w] is zeroed by the call, then receives taps^T grad_out [planes,H,W]. */
int tss_upsample_bilinear_adjoint(const float* grad_out, float* grad_in, int64_t planes, int h, int w,
                                  int H, int W, void* stream);

/* A4   threshold block on STORED labels: data/freiburg_dataset.py:115-123.  label_in uint8 [n],
 * conf16 fp16 bits [n] -> label_out uint8 [n]. */
int tss_threshold_labels(const uint8_t* label_in, const uint16_t* conf16, int64_t n, int num_classes,
                         const float* thr_host, const uint8_t* keep_host, int ignore_index,
                         uint8_t* label_out, void* stream);

/* A4   nearest-neighbour gather used for PIL's Image.resize(NEAREST) (data/freiburg_dataset.py:124)
 * and F.interpolate(mode='nearest') (self_training.py:103): dst[p, y, x] = src[p, ytab[y], xtab[x]].
 * ytab/xtab are DEVICE int32 tables the host mirror fills with the library's exact index rule. */
int tss_gather_nearest_u8(const uint8_t* src, uint8_t* dst, int64_t planes, int h, int w, int H, int W,
                          const int32_t* ytab, const int32_t* xtab, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A5   per-(image, class) segmented sum of features by label -- replaces the (n, t) double loop of
 * Class_Features.calculate_mean_vector (cal_prototype.py:118-144) and
 * SelfTrain.calculate_mean_vector (self_training.py:176-202).
 *
 * Step 1, tss_proto_labels: logits [N,C,P] -> per pixel argmax (ties -> lowest class, as torch),
 *   optionally intersected with labels_val [N,P] fp32 (values >= C go to `overflow_channel`,
 *   process_label :146-151 / :204-209) and gated by conf >= conf_thresh (use_conf_gate != 0,
 *   self_training.py:180-181).  Writes seg_label [N,P] uint8 (class whose sum the pixel joins, 255 =
 *   none) and, in `records` (tss_proto_records_bytes(N, P) bytes, 16-byte aligned), per tile of
 *   TSS_PROTO_TILE pixels: the pixel offsets sorted by class in groups of eight (class segments padded with the
 *   offset of an always-zero slot), the gated per-class counts (the divisor :128,:140 / :189,:198) and the
 *   ungated one-hot counts (the "< 10" rule :135 / :196).  No atomics, no zero-initialised buffers.
 *   sums_to_zero (may be NULL): sums_elems floats
 *   zeroed by the same launch, so that tss_proto_accumulate(zero_sums = 0) can follow without a memset.
 * Step 2, tss_proto_accumulate: feat [N,K,P] fp32 (K % 32 == 0) streamed once through shared memory
 *   (bulk async copies + mbarrier ring); sums [N,C,K] fp32 are zeroed first when zero_sums != 0, then
 *   accumulated.  consumer_warps: 0 (= 16, the kernel owns the SM) | 4 | 8 | 16 -- with 4 the CTA is 160 threads
 *   x 96 registers and shares every SM with the pseudo-label kernel (TSS_PL_FLAG_SM_WARPS(12)).
 * Step 3, tss_proto_finalize: count_all / count_gated [N,C] int32 (OUT, summed over the tile records);
 *   vectors [N,C,K] = sums / count_gated; valid [N,C] uint8 = count_gated > 0 && count_all >= min_pixels
 *   (the two `continue`s); vector_sum [N,C] fp32 (for the `vector.sum() == 0` skip of
 *   update_objective_SingleVector :154 / :212).
 * Step 3+A6 in one launch, tss_proto_finalize_update: the same outputs, then the sequential prototype
 *   update of tss_proto_update below with them (cal_prototype.py:72-75: calculate_mean_vector followed by
 *   the update loop). */
#define TSS_PROTO_TILE 512
#define TSS_PROTO_RECORD_BYTES 1776 /* per tile: 768 uint16 offsets, 96 group classes, 2 x 32 counts, group count */
size_t tss_proto_records_bytes(int64_t n_img, int64_t plane);
int tss_proto_labels(const float* logits, const float* labels_val, int64_t n_img, int num_classes,
                     int64_t plane, int overflow_channel, int use_conf_gate, float conf_thresh,
                     uint8_t* seg_label, void* records, float* sums_to_zero, int64_t sums_elems,
                     void* stream);
int tss_proto_accumulate(const float* feat, const void* records, int64_t n_img, int64_t channels,
                         int64_t plane, int num_classes, float* sums, int zero_sums, int consumer_warps,
                         void* stream);
int tss_proto_finalize(const float* sums, const void* records, int64_t n_img, int num_classes,
                       int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                       int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum, void* stream);
/* Step 3 on a sharded batch: the same outputs, and `vectors` is ALSO stored through
 * peer_vectors[0 .. n_peers) (n_peers <= 7) -- device pointers into other GPUs' memory that this process has
 * mapped (CUDA IPC / symmetric memory), each the [N,C,K] slot of this rank in that GPU's gather buffer;
 * peer_vector_sums (or NULL) likewise receives vector_sum [N,C].  The
 * exchange of the reference's per-image mean vectors is then the kernel's own NVLink store stream; a barrier
 * and tss_proto_update over the gathered [ranks*N,C,K] buffer follow on every rank. */
int tss_proto_finalize_scatter(const float* sums, const void* records, int64_t n_img, int num_classes,
                               int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                               int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum,
                               float* const* peer_vectors, float* const* peer_vector_sums, int n_peers,
                               void* stream);
int tss_proto_finalize_update(const float* sums, const void* records, int64_t n_img, int num_classes,
                              int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                              int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum,
                              float* protos, float* num, int mode, int start_mean, double momentum,
                              double cap, void* stream);

/* A5 without allocations or caller workspace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Steps 1-3 above in one call -- tss_proto_labels (no labels_val, no confidence gate, overflow channel =
 * num_classes: the Class_Features form, cal_prototype.py:118-144), tss_proto_accumulate, tss_proto_finalize -- with the
 * scratch they pass to each other (tile records, segment labels, the [N,C,K] sums) in a file-scope
 * `static __device__` array of TSS_NOALLOC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it
 * until the process exits).  Outputs are those of tss_proto_finalize, all caller-owned.  The call allocates and frees
 * nothing.  Needed scratch = up256(tss_proto_records_bytes(N, P)) + up256(N*P) + up256(4*N*C*K) bytes (up256 = rounded
 * up to 256); a batch that needs more is rejected with TSS_E_ARG and a message naming both numbers (cfg2, 16 x 13 x 256
 * x 81 x 241, needs 1.63 MB).  Every call on a device shares the one array, and it keeps the last call's contents (the
 * label kernel rewrites all it will read and zeroes the sums in its own launch): calls that can overlap on the GPU --
 * different streams, or graph branches without an edge between them -- RACE; issue them on one stream. */
#define TSS_NOALLOC_SCRATCH_BYTES (4 << 20)
int tss_class_means_noalloc(const float* feat, const float* logits, int64_t n_img, int num_classes,
                            int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                            int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum, void* stream);

/* A6   sequential prototype update -- update_objective_SingleVector (cal_prototype.py:153-169,
 * self_training.py:211-227) applied to the vectors of one batch in the reference's list order
 * (image-major, then class).  Classes are independent, so one thread block walks each class's images
 * in order with the reference's own fp32 operation sequence.  protos [C,K] and num [C] fp32 are
 * updated in place.  mode: TSS_UPDATE_MEAN | TSS_UPDATE_MOVING_AVERAGE; start_mean != 0 forces 'mean'
 * while num < 100; cap = 3000 in the reference. */
#define TSS_UPDATE_MEAN 0
#define TSS_UPDATE_MOVING_AVERAGE 1
/* valid == NULL: every vector counts; vector_sum == NULL: the kernel sums each vector itself (the
 * single-vector form update_objective_SingleVector(id, vector) passes protos + id*K, num + id, C = 1). */
int tss_proto_update(float* protos, float* num, const float* vectors, const uint8_t* valid,
                     const float* vector_sum, int64_t n_img, int num_classes, int64_t channels, int mode,
                     int start_mean, double momentum, double cap, void* stream);
/* Sharded form: the same update expressed through its closed-form weights so
 * that ranks can combine with one all-reduce.  For every class, with the contributing vectors of THIS
 * rank occupying positions [first_index[c], first_index[c] + m_local) of the global order of
 * total[c] contributions: partial[c,:] = sum_j w_j * v_j (fp32 out, fp64 weights), and scale0[c] = the
 * factor the old prototype keeps.  The caller all-reduces partial, then
 * protos = scale0 * protos + partial, num = num_after.  contrib_count [C] int32 receives m_local
 * when first_index == NULL (counting pass). */
int tss_proto_update_partial(const float* vectors, const uint8_t* valid, const float* vector_sum,
                             int64_t n_img, int num_classes, int64_t channels, int mode, int start_mean,
                             double momentum, double cap, const float* num_before,
                             const int32_t* first_index, const int32_t* total, int32_t* contrib_count,
                             float* partial, float* scale0, float* num_after, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A7/A8/A10  feature-to-prototype distance, soft weights, label rectification -- replaces
 * SelfTrain.feat_prototype_distance (self_training.py:160-165), get_prototype_weight :139-142 and
 * the rectify block :106-112.  feat [N,K,P] fp32 (K % 32 == 0), protos [C,K] fp32.
 *   dist_out   [N,C,P] fp32 or NULL : ||proto_c - feat||_2 in fp32 (see `form`)
 *   weight_out [N,C,P] fp32 or NULL : softmax_c(-(d - min_c d) * temperature)
 *   if soft != NULL: rect = weight * soft [N,C,P]; label = argmax_c rect; label = ignore_index where
 *   max_c(rect / sum_c rect) < train_thred; label = hard[px] where hard != NULL and
 *   hard[px] != ignore_index.  hard: [N,P] of hard_dtype; label_out [N,P] int64.
 * form: 0 = automatic, else TSS_DIST_FORM(variant, prefetch, split, algo) --
 *   variant  0 auto | 1 eight-warp register-pipeline kernel | 2 / 4 thread-level-parallel kernel, 2 / 4 px per lane |
 *            8 streaming kernel (TMA bulk copies of feat tiles + prototype slices into a shared-memory ring; expanded form)
 *   prefetch 0 auto | 1 + depth (0, 1, 2 four-channel groups) of the thread-level-parallel kernel; with variant 8: the
 *            tile shape of the streaming kernel (1: 512 px x 32 ch, 8 warps x 2 px; 2: 512 x 32, 4 warps x 4 px;
 *            3: 1024 x 16, 8 warps x 4 px; 4: 1024 x 16, 16 warps x 2 px)
 *   split    0 auto | K split 1, 2, 4 (8: register-pipeline kernel only)
 *   algo     0 auto | TSS_DIST_DIRECT: sum (p - f)^2 as torch.norm does | TSS_DIST_EXPANDED: |f|^2 + |p|^2 - 2 f.p
 *            (half the FP32 work) with every (pixel, class) whose d^2 < 0.5 (|f|^2 + |p|^2) -- the cancellation
 *            region -- recomputed in the direct form; automatic = the streaming kernel (expanded) for K >= 512, the
 *            thread-level-parallel kernel in its expanded form for 128 <= K < 512, direct below. */
#define TSS_DIST_DIRECT 1
#define TSS_DIST_EXPANDED 2
#define TSS_DIST_FORM(variant, prefetch_plus1, split, algo) \
    (((variant) & 0xf) | (((prefetch_plus1) & 0xf) << 4) | (((split) & 0xf) << 8) | (((algo) & 0xf) << 12))
int tss_proto_distance(const float* feat, const float* protos, int64_t n_img, int64_t channels,
                       int64_t plane, int num_classes, float temperature, float* dist_out,
                       float* weight_out, const float* soft, const void* hard, int hard_dtype,
                       float train_thred, int ignore_index, int64_t* label_out, int form, void* stream);

/* A7-A10 from the network's own 1/8-stride features: the chain self_training.py:69 (bilinear
 * align_corners=True interpolation of feat8 [N,K,h8,w8] to the quarter map [Hq,Wq]) -> :145-158 (full2weak: per image
 * resize to (rh, rw), crop at (y0, x0), optional mirror; `params` as in tss_full2weak) -> :160-165, :139-142, :106-112
 * evaluated per OUTPUT pixel [N, TH*TW] with the two ATen operation sequences reproduced on the fly: neither the
 * [N,K,Hq,Wq] nor the [N,K,TH,TW] tensor is written.  Outputs / soft / hard / label as in tss_proto_distance
 * (P = TH*TW); direct-difference form; channels % 4 == 0.  flags: 0, or 1 = four channels' taps in flight per thread
 * (more registers, half the resident warps: A/B of the two forms). */
int tss_proto_distance_sampled(const float* feat8, const float* protos, int64_t n_img, int64_t channels, int h8,
                               int w8, int Hq, int Wq, const int32_t* params, int TH, int TW, int num_classes,
                               float temperature, float* dist_out, float* weight_out, const float* soft,
                               const void* hard, int hard_dtype, float train_thred, int ignore_index,
                               int64_t* label_out, int flags, void* stream);

/* A9   full2weak -- SelfTrain.full2weak (self_training.py:145-158): per image bilinear
 * (align_corners=True) resize of feat [N,K,h,w] to (rh, rw), crop [y0:y0+TH, x0:x0+TW], optional mirror
 * along W, concatenated -> out [N,K,TH,TW]; one gather pass, no resized intermediate.
 * params: DEVICE int32 [N,6] rows {rh, rw, y0, x0, flip, 0} (the weak_params layout of
 * utils/self_train_augmentation.py after the reference's int(./4) arithmetic, done by the host mirror). */
int tss_full2weak(const float* feat, int64_t n_img, int64_t channels, int h, int w, const int32_t* params,
                  int TH, int TW, float* out, void* stream);

/* process_label -- cal_prototype.py:146-151 / self_training.py:204-209: one-hot [N, C+1, P] fp32 of a
 * float label map [N,P]; labels >= C go to `overflow_channel` (C in Class_Features, args.ignore_index in
 * SelfTrain; the reference's scatter_ raises when that exceeds C, and so does this call). */
int tss_one_hot_overflow(const float* label, int64_t n_img, int64_t plane, int num_classes,
                         int overflow_channel, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A11/A12  cross-entropy with ignore_index and reverse cross-entropy -- nn.CrossEntropyLoss
 * (self_training.py:27,57,114; utils/loss.py:145) and SelfTrain.rce (self_training.py:167-174).
 * logits [N,C,P] fp32, target [N,P] (target_dtype).  Results land in `out` (DEVICE, 8 doubles, 5 used):
 * {sum_px ce, sum_px rce, pixels with a label in [0, C) and != ignore, pixels whose label the reference raises
 * on (outside [0, C], or == C without rce), pixels labelled exactly C with rce: one_hot(label, C+1)[:, :C]
 * accepts them, their clamped one-hot row is all 1e-4, so they add -log(1e-4) sum_c clamp(p_c) to the rce sum and
 * count in its mask, while CrossEntropyLoss has no term for them}.  The host mirror forms
 * ce = out[0]/out[2], rce = out[1]/(out[2] + out[4] + 1e-6).  workspace: tss_loss_workspace_bytes() bytes; its trailing
 * 16 bytes hold a ticket counter that MUST BE ZERO when the call is issued and is left zero by every call (allocate once
 * with zeros, reuse; one workspace per stream): the CTA that draws the last ticket does the fixed-order final reduction,
 * so forward = ONE launch. */
size_t tss_loss_workspace_bytes(void);
int tss_ce_rce_fwd(const float* logits, const void* target, int target_dtype, int64_t n_img,
                   int num_classes, int64_t plane, int ignore_index, int with_rce, double* out,
                   void* workspace, void* stream);
/* The same, and losses (DEVICE, 3 floats) = {alpha ce + beta rce, ce, rce} formed in the reduction kernel with the
 * reference's arithmetic (fp64 division, fp32 rounding, fp32 mix; ce_sum != 0: reduction='sum' for ce), so the op
 * is one launch with no scalar follow-up (self_training.py:114-118). */
int tss_ce_rce_fwd_mix(const float* logits, const void* target, int target_dtype, int64_t n_img,
                       int num_classes, int64_t plane, int ignore_index, int with_rce, double alpha,
                       double beta, int ce_sum, double* out, float* losses, void* workspace, void* stream);
/* grad[N,C,P] = ce_scale * d(sum ce)/dx + rce_scale * d(sum rce)/dx, scales read from the DEVICE
 * array scales[2] (so that no host sync is needed between forward and backward). */
int tss_ce_rce_bwd(const float* logits, const void* target, int target_dtype, int64_t n_img,
                   int num_classes, int64_t plane, int ignore_index, const float* scales, float* grad,
                   void* stream);

/* The same with the scales formed in the kernel from the upstream gradients upstream3 = {d/d total, d/d ce, d/d rce}
 * (DEVICE, 3 floats) and the forward's `out` sums: ce_scale = (g_total alpha + g_ce) / valid, rce_scale =
 * (g_total beta + g_rce) / (valid + 1e-6) -- the autograd backward is one launch. */
int tss_ce_rce_bwd_mix(const float* logits, const void* target, int target_dtype, int64_t n_img,
                       int num_classes, int64_t plane, int ignore_index, int with_rce, double alpha,
                       double beta, int ce_sum, const float* upstream3, const double* fwd_out, float* grad,
                       void* stream);

/* A13  focal loss -- FocalLoss.forward (utils/loss.py:164-184): loss_px = -(1-pt)^gamma * alpha[y] *
 * log pt with pt detached; no ignore handling.  out (DEVICE, 8 doubles) = {sum loss, pixel count, -, labels
 * outside [0, C) (the reference's gather raises), -}.
 * alpha: DEVICE [C] fp32 or NULL.  bwd: grad = scale[0] * d(sum loss)/dx. */
int tss_focal_fwd(const float* logits, const void* target, int target_dtype, int64_t n_img,
                  int num_classes, int64_t plane, float gamma, const float* alpha, double* out,
                  void* workspace, void* stream);
int tss_focal_bwd(const float* logits, const void* target, int target_dtype, int64_t n_img,
                  int num_classes, int64_t plane, float gamma, const float* alpha, const float* scale,
                  float* grad, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TSS_B200_H_ */
