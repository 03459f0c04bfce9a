/* yv3_head.h -- C ABI of the B200-native YOLOv3 detection-head path.
 *
 * One shared library (libyv3head.so), plain C linkage, plain pointers and sizes.  The reference
 * (zheng-yuwei/YOLOv3-tensorflow) has no FFI for this path: its boundary is a Python class surface
 * executed by TensorFlow.  Each entry point below names the reference interface it replaces
 * (paths relative to the reference repo root).
 *
 * Conventions
 *  - every tensor pointer is a DEVICE pointer unless the name ends in `_host`;
 *  - the caller owns every buffer, including the workspace; the library allocates nothing on the
 *    device, keeps no global state and never synchronises the stream (except yv3_read_status);
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *  - functions return 0 (YV3_OK) or a negative yv3_error; nothing throws across the ABI;
 *  - all arithmetic is fp32, one rounding per operator in the reference's operator order
 *    (no FMA contraction), except NMS IoU which is fp64 like the reference's NumPy code.
 *
 * Packed head layout (yolov3/yolov3_detector.py:80-85, yolov3/yolov3_decoder.py:99-117):
 *   predicts[n][r][c][k], r < H/32, c < W/32, k < Ctot; every /32 row holds 16*B8 records of the /8
 *   head (16 raster-consecutive cells), then 4*B16 records of /16, then B32 records of /32; a record is
 *   L = 5 + class_num floats [t_x, t_y, t_w, t_h, t_conf, z_0 .. z_{C-1}].  The whole tensor is thus a
 *   flat array of N * (H/32*W/32) * (16*B8 + 4*B16 + B32) records.  Gradients use the same layout.
 */
#ifndef YV3_HEAD_H_
#define YV3_HEAD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YV3_ABI_VERSION 6
#define YV3_MAX_BOX 8      /* anchors per head (reference ships 3 / 2 / 3, configs.py:37-41) */
#define YV3_MAX_GT 256     /* label slots per image (M) */

typedef enum yv3_error {
    YV3_OK = 0,
    YV3_ERR_NULL = -1,          /* required pointer is NULL */
    YV3_ERR_SHAPE = -2,         /* grids are not /8,/16,/32 of one input; N, M, box_num out of range */
    YV3_ERR_CONFIG = -3,        /* weights / thresholds / eps not finite or out of range */
    YV3_ERR_ALIGN = -4,         /* device pointer not aligned as documented */
    YV3_ERR_WORKSPACE = -5,     /* workspace NULL or smaller than yv3_query_workspace reports */
    YV3_ERR_CUDA = -6,          /* a CUDA runtime call failed (launch error, wrong device, ...) */
    YV3_ERR_CAPACITY = -7,      /* candidate capacity exceeded / problem too large for the kernel */
    YV3_ERR_GT_RANGE = -8,      /* (status) a valid label's centre falls outside the grid; TF's
                                   gather_nd raises here on CPU (yolov3/yolov3_loss.py:271) */
    YV3_ERR_NONFINITE = -9,     /* (status) inf/NaN met in the loss (reported, not "fixed") */
    YV3_ERR_STATIC_SCRATCH = -10 /* yv3_nms_static: n x capacity needs more than YV3_NMS_STATIC_SCRATCH_BYTES of scratch */
} yv3_error;

/* The switch surface the path consumes: YOLOv3Loss.__init__ (yolov3/yolov3_loss.py:16-67),
 * YOLOv3Decoder.__init__ (yolov3/yolov3_decoder.py:14-45), configs.py:36-72, and
 * keras.backend.epsilon() as set by run.py:26.  Heads are ordered /8, /16, /32. */
typedef struct yv3_config {
    int32_t grid_h[3];
    int32_t grid_w[3];
    int32_t box_num[3];
    int32_t class_num;
    float anchors[3][YV3_MAX_BOX][2];   /* normalised (w, h) priors, fp32-rounded like tf.constant */
    float iou_thresh;
    float loss_weights[3][5];           /* per head: xy, wh, noobj, obj, cls  (configs.py:52) */
    float rectified_loss_weight[3];     /* yolov3_loss.py:63-67 */
    int32_t is_focal_loss;
    float focal_alpha;
    float focal_gamma;
    int32_t is_tiou_recall;
    double eps;                         /* keras.backend.epsilon(): clip_by_value(x, fp32(eps),
                                           fp32(1 - eps)), the Python doubles rounded once */
} yv3_config;

int yv3_abi_version(void);
const char* yv3_strerror(int code);

/* Validates a config; replaces the ValueError checks of YOLOv3Loss.__init__ (yolov3_loss.py:63-66)
 * and the implicit "input is a multiple of 32" contract of configs.py:46-48. */
int yv3_config_check(const yv3_config* cfg);

/* Floats per image in the packed layout ((H/32)*(W/32)*Ctot) and records (anchors) per image. */
int64_t yv3_packed_floats_per_image(const yv3_config* cfg);
int64_t yv3_records_per_image(const yv3_config* cfg);

/* Workspace needed by yv3_assign / yv3_loss_fwd_bwd for a local batch of n images, m label slots. */
int yv3_query_workspace(const yv3_config* cfg, int n, int m, size_t* bytes);

/* YOLOv3Decoder.decode (yolov3/yolov3_decoder.py:62-87): out[3*h + 0] = raw[...,0:4] (N,H,W,B,4),
 * out[3*h + 1] = decoded (N,H,W,B,L) in grid units, out[3*h + 2] = corner boxes (N,H,W,B,4). */
int yv3_decode(const yv3_config* cfg, const float* predicts, int n, float* const out[9], void* stream);

/* LabelDecoder.decode (yolov3/label_decoder.py:26-60): targets (N, M*5) ->
 * out[2*h] = (N,M,5) labels in head-h grid units, out[2*h + 1] = (N,M,4) corner boxes. */
int yv3_label_decode(const yv3_config* cfg, const float* targets, int n, int m, float* const out[6],
                     void* stream);

/* Label ingestion, the step that produces `targets`: the letter-box label transform of
 * FileUtil._parse_image_with_label (dataset/file_util.py:47-52) and the -1 padded batch of
 * FileUtil.get_dataset (dataset/file_util.py:97), for n images at once.
 *   labels       device, the images' label rows back to back, 5 floats each [x, y, w, h, cls] normalised to
 *                the SOURCE image (the numbers of a label.txt line, file_util.py:22-32)
 *   row_offsets  device int32[n + 1]: image i owns rows row_offsets[i] .. row_offsets[i + 1]
 *   image_hw     device int32[n][2]: source image height, width in pixels (tf.shape(image)[0:2])
 *   target_h/w   the network input size (configs.py:36)
 *   targets      device (n, m * 5) fp32: transformed rows, then -1 up to m slots.  An image with more than m
 *                rows is an error (YV3_ERR_SHAPE is returned by the host check of the caller-supplied maximum).
 * fp32 arithmetic in the reference's order: r = (w_src / W, h_src / H) reversed and divided by its max;
 * xy' = xy * r + (1 - r) / 2;  wh' = wh * r. */
int yv3_label_batch(const float* labels, const int32_t* row_offsets, const int32_t* image_hw, int n, int m,
                    int target_h, int target_w, float* targets, void* stream);

/* Ground-truth -> cell / anchor / head assignment: YOLOv3Loss._get_valid_target, _calc_iou (response
 * part) and the cross-head ownership test (yolov3/yolov3_loss.py:225-252, :269-302, :203-208).
 * table[n][j][h] = {row, col, best_anchor, owned} (int32; all -1/-1/-1/0 for a padding slot),
 * iou[n][j][h] = best IoU of the centre cell's anchors (0 for padding).  Either output may be NULL.
 * Tie rules: lowest anchor index among equal IoUs; `>=` both ways across heads (ties -> every tied
 * head owns the label). */
int yv3_assign(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
               int32_t* table, float* iou, void* workspace, size_t workspace_bytes, void* stream);

/* YOLOv3Loss.loss forward (yolov3/yolov3_loss.py:81-138, :140-164, :166-222, :305-369) fused with
 * the backward TensorFlow autodiff derives from it.
 *   n_global           images in the whole (all-rank) batch: the reduce_mean denominator (:112)
 *   rectified_active   value of `current_num <= rectified_coord_num` for this step (:125)
 *   upstream           d(final objective)/d(loss) (1.0 under Keras)
 *   loss_out20         device, 20 floats.  [0..17] = the (6,3) loss detail: rows [xy, wh, noobj,
 *                      obj, cls, rectified] x heads [/8,/16,/32] (the loss_detail variables of
 *                      :71-79, :115-134); THIS RANK's share of the batch mean (sum over local
 *                      images / n_global).  [18] = their sum = this rank's share of the scalar
 *                      loss (:137).  [19] = 0.  Summing (all-reducing) the 20 floats over the
 *                      ranks of a batch-sharded job gives the global values.
 *   phases             bit mask, 7 = the whole loss.  The path is three launches -- 1: the per-label
 *                      assignment kernel (centre-cell IoU, ownership, owner-side terms; latency-bound,
 *                      tiny), 2: the fused dense sweep (HBM-bound; leaves one row of partial sums per
 *                      CTA in the workspace), 4: the fold of those rows into loss_out20 (one CTA) --
 *                      and a caller that wants to time them separately may issue them as separate
 *                      calls with the same arguments, in that order.
 *   grad_predicts      device, packed layout, n images; must be 16-byte aligned; may be NULL
 *                      (forward only).  Every element is written. */
int yv3_loss_fwd_bwd(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
                     int n_global, int rectified_active, float upstream, float* loss_out20,
                     float* grad_predicts, void* workspace, size_t workspace_bytes, int phases, void* stream);

/* The same loss on the three head tensors BEFORE they are packed: head_predicts[h] is the output of head h's
 * last convolution, (n, H_h, W_h, B_h * (5 + C)) fp32 contiguous, h = /8, /16, /32.  Replaces the Reshape + concatenate
 * of yolov3/yolov3_detector.py:80-85 on the way in (a full read + write of the tensor the packed form needs first)
 * and the matching split of the gradient on the way out: head_grads[h] has the shape of head_predicts[h] and every
 * element is written (may be NULL as a whole: forward only).  A head tensor alone is the packed layout of a geometry
 * that has only that head (the Reshape is a reinterpretation), so the same kernels run once per tensor; results
 * per record are bit-identical to the packed call, the loss sums are folded in a different (still fixed) order.
 * All other arguments as in yv3_loss_fwd_bwd.  SURVEY 8(f) rank 2; yv3_filter_heads is the inference-side twin. */
int yv3_loss_fwd_bwd_heads(const yv3_config* cfg, const float* const head_predicts[3], const float* targets, int n,
                           int m, int n_global, int rectified_active, float upstream, float* loss_out20,
                           float* const head_grads[3], void* workspace, size_t workspace_bytes, int phases,
                           void* stream);

/* Deferred status of the last yv3_assign / yv3_loss_fwd_bwd that used `workspace`: synchronises the
 * stream and returns YV3_OK, YV3_ERR_GT_RANGE or YV3_ERR_NONFINITE (the errors TF would raise /
 * the NaNs it would silently propagate). */
int yv3_read_status(void* workspace, void* stream);

/* ---- post-process ---------------------------------------------------------------------------- */

/* One candidate row = the 8 columns of YOLOv3PostProcessor._filter_single_head_boxes
 * (yolov3/yolov3_post_process.py:65-75): x0/W, y0/H, x1/W, y1/H, conf, prob, class, score -- fp32.
 * The reference's rows are float64 only because NumPy promotes fp32 columns next to the int64
 * arg-max column (:58, :74); every value is exactly an fp32, so fp32 rows lose nothing. */
#define YV3_CAND_COLS 8
#define YV3_MAX_CANDIDATES (1 << 18)   /* rows per candidate list (NMS sort key keeps the row in 18 bits) */

/* Workspace for the three calls below for n lists of `capacity` rows; cfg may be NULL when only
 * yv3_nms is used. */
int yv3_query_postprocess_workspace(const yv3_config* cfg, int n, int32_t capacity, size_t* bytes);

/* YOLOv3PostProcessor.filter_boxes on already-decoded heads of ONE image
 * (yolov3/yolov3_post_process.py:21-76).  decoded[h] is (H,W,B,L), boxes[h] is (H,W,B,4), 16-B aligned.
 * cand (capacity x 8 fp32, 16-B aligned) receives the survivors in the reference's order (head
 * /8,/16,/32, raster (y,x,b) within a head); counts (device int32[3]) their number per head (the TRUE
 * number: if their sum exceeds capacity the surplus rows are dropped and yv3_read_status reports
 * YV3_ERR_CAPACITY).  Keep iff score > fp32(thresh); class = lowest index among equal maxima. */
int yv3_filter_decoded(const yv3_config* cfg, const float* const decoded[3], const float* const boxes[3],
                       float score_thresh, float* cand, int32_t capacity, int32_t* counts,
                       void* workspace, size_t workspace_bytes, void* stream);

/* Fused decode + score + threshold straight from the packed logits of n images (the inference path
 * run.py:59-69 takes, without materialising the decoded tensors).  cand is (n, capacity, 8) fp32,
 * counts (n, 3) int32.  Per image the order is the reference's; results are identical to
 * yv3_decode followed by yv3_filter_decoded.  phases: bit mask, 3 = the whole call.  It is two launches -- 1: the
 * score sweep over the logits (HBM-bound; stages the survivors per 32-record tile in the workspace), 2: the gather
 * that turns the staged survivors into cand / counts (tiny) -- and a scheduler may issue them as two calls with the
 * same arguments, e.g. to start another HBM-bound kernel between them. */
int yv3_filter_packed(const yv3_config* cfg, const float* predicts, int n, float score_thresh,
                      float* cand, int32_t capacity, int32_t* counts, void* workspace, size_t workspace_bytes,
                      int phases, void* stream);

/* yv3_filter_packed on the three head tensors before they are packed (head_predicts[h]: (n, H_h, W_h, B_h * (5 + C))
 * fp32 contiguous, h = /8, /16, /32; see yv3_loss_fwd_bwd_heads): same rows, same order, same counts.  Every head must
 * have at least 4 records per image (YV3_ERR_SHAPE otherwise: a 32x32 input with fewer than 4 anchors at /32). */
int yv3_filter_heads(const yv3_config* cfg, const float* const head_predicts[3], int n, float score_thresh, float* cand,
                     int32_t capacity, int32_t* counts, void* workspace, size_t workspace_bytes, int phases, void* stream);

/* Evaluation while training: yv3_loss_fwd_bwd (phases 7) and yv3_filter_packed (phases 3) on the SAME packed predicts
 * in one call -- what a trainer.py step followed by run.py's filter on the same head output does (yolov3_loss.py:81-138,
 * yolov3_post_process.py:21-76).  The two HBM-bound sweeps become one: a warp lands a tile of 32 records in shared
 * memory once, scores it, runs the loss record phase, turns the tile in place into the gradient and stores it back
 * whole (4*A*L bytes read + 4*A*L bytes written per image instead of two reads of the tensor and a gradient written
 * in 32-byte pieces).  Arguments as in the two calls (loss_workspace from yv3_query_workspace, post_workspace from
 * yv3_query_postprocess_workspace); grad_predicts is required.  cand / counts / grad_predicts and every integer
 * decision are bit-identical to the separate calls; the loss sums are folded per warp, so a detail value may differ
 * in its last bit.  Launches: assignment, the fused sweep, the fold, the gather.
 * phases: bit mask, 15 = the whole call; 1 = the assignment, 2 = the sweep, 4 = the fold, 8 = the gather, for a caller that
 * times them or puts them on streams of its choice (same arguments; 2 after 1, 4 and 8 after 2).  Falls back to the two separate sweeps
 * (same results) when the fused tile shape does not apply (predicts not 16-byte aligned, fewer than 32 records per
 * image, more than 262144 tiles of 32 records per call); *fused_out (host, may be NULL) says which ran.  yv3_nms follows as usual. */
int yv3_loss_filter_fused(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
                          int n_global, int rectified_active, float upstream, float* loss_out20, float* grad_predicts,
                          void* loss_workspace, size_t loss_workspace_bytes, float score_thresh, float* cand,
                          int32_t capacity, int32_t* counts, void* post_workspace, size_t post_workspace_bytes,
                          int32_t* fused_out, int phases, void* stream);

/* yv3_loss_filter_fused on the three head tensors BEFORE yolov3/yolov3_detector.py:80-85 packs them (head_predicts /
 * head_grads as for yv3_loss_fwd_bwd_heads): the one-sweep kernel runs once per head tensor, so an evaluation-while-
 * training step neither concatenates the heads nor reads any of them twice.  Gradients, candidate rows and counts are
 * bit-identical to yv3_loss_fwd_bwd_heads + yv3_filter_heads, to which the call falls back (same results) when a head
 * tensor is not 16-byte aligned or has fewer than 32 records per image; *fused_out says which ran. */
int yv3_loss_filter_fused_heads(const yv3_config* cfg, const float* const head_predicts[3], const float* targets, int n,
                                int m, int n_global, int rectified_active, float upstream, float* loss_out20,
                                float* const head_grads[3], void* loss_workspace, size_t loss_workspace_bytes,
                                float score_thresh, float* cand, int32_t capacity, int32_t* counts, void* post_workspace,
                                size_t post_workspace_bytes, int32_t* fused_out, int phases, void* stream);

/* Class-aware greedy NMS, YOLOv3PostProcessor._apply_nms (yolov3/yolov3_post_process.py:108-159), for
 * n independent candidate lists.  cand is (n, capacity, 8) fp32; list i holds
 * sum(counts[i*count_cols .. +count_cols)) rows (pass yv3_filter_packed's (n,3) counts with
 * count_cols = 3).  Class ids must be integers in [0, num_classes).  keep (n, capacity) uint8 receives
 * 1 for the TRUE greedy survivors: walking candidates by score descending (ties: lower row first, i.e.
 * the reference's stable sort), a box is dropped iff an already kept box of the same class has
 * IoU > nms_thresh, IoU evaluated in float64 in the reference's operation order.  The id-collision
 * quirk of apply_nms (:81-104) is layered on the host. */
int yv3_nms(const yv3_config* cfg, const float* cand, const int32_t* counts, int count_cols, int n,
            int32_t capacity, int num_classes, double nms_thresh, uint8_t* keep, void* workspace,
            size_t workspace_bytes, void* stream);

/* yv3_nms for callers that cannot allocate device memory and have no workspace to hand over (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels and the same keep flags as yv3_nms(cfg = NULL, ...), but the
 * scratch (sort keys, sorted boxes, row indices, spill list, segment queues) is a file-scope device array of
 * YV3_NMS_STATIC_SCRATCH_BYTES inside the library (csrc/static_scratch.cu), which every GPU that loads the library holds
 * until the process exits.  The call allocates nothing: no cudaMalloc / cudaFree, no allocating library routine.
 *   limit    the scratch for n lists of `capacity` rows is, each term rounded up to 256 bytes,
 *              8 * n * pow2ceil(capacity) + 16 * r + 4 * r + 16 * r + 8 * (r + r / 256 + 16) + 256,   r = n * capacity
 *            (about 60 bytes per row: 64 lists of 1024 rows, 16 of 4096, 1 of 65536).  A larger request returns
 *            YV3_ERR_STATIC_SCRATCH before anything is launched; use yv3_nms with a workspace for those.
 *   sharing  all calls on one device use the same array and it keeps the last call's contents (the call clears the few
 *            words that must start at zero).  Calls that OVERLAP on the GPU -- two streams with no ordering between them --
 *            race on it: keep the calls of one device on one stream, or order them with events.
 *   set-up   the first call on a device asks the runtime for the array's device address (cudaGetSymbolAddress, cached per
 *            device); make that call outside stream capture. */
#define YV3_NMS_STATIC_SCRATCH_BYTES (4u << 20)
int yv3_nms_static(const float* cand, const int32_t* counts, int count_cols, int n, int32_t capacity, int num_classes,
                   double nms_thresh, uint8_t* keep, void* stream);

/* The rows run.py gets from apply_nms + resize_boxes (yolov3_post_process.py:94-104, :162-172), on the device: per
 * image and per head (/8, /16, /32) the returned rows in local-id order,
 *   [x0 * target_w, y0 * target_h, x1 * target_w, y1 * target_h, conf, prob, class, score, id]   float64
 * (the reference's rows are float64 and an fp32 coordinate times an integer size is exact there).  cand / counts as
 * written by yv3_filter_*, keep as written by yv3_nms.  bug_compat != 0 reproduces the shipped id collision of :81-104
 * (a row is returned iff its local id equals the local id of a greedy survivor of ANY head); 0 returns the survivors.
 *   rows_out    device, n * capacity * 9 doubles; image i's rows start at i * capacity * 9, heads back to back
 *   counts_out  device, n * 3 int32: rows returned per head
 * SURVEY 8(f) rank 3 (the mAP accumulation named there is an open TODO in the reference, run.py:78-79: nothing to match). */
int yv3_collect_kept(const float* cand, const int32_t* counts, const uint8_t* keep, int n, int32_t capacity,
                     double target_w, double target_h, int bug_compat, double* rows_out, int32_t* counts_out,
                     void* stream);

/* ---- evaluation after NMS: AP / mAP accumulation on the device ---------------------------------------
 * The reference leaves this as a TODO that names the third-party tool Cartucho/mAP (run.py:78-79, README.md:137; not
 * vendored, no version pinned).  These entry points run that tool's published algorithm (main.py: per class, detections
 * by confidence descending, each matched to the ground-truth box of its image and class with the largest IoU, true
 * positive iff IoU >= min_overlap and the box is not yet claimed; VOC2012 all-point AP; mAP over the classes that have
 * ground truth) on the rows yv3_collect_kept leaves on the device and the label batch the loss takes.
 *   yv3_map_state_bytes   bytes of the accumulator for up to max_detections kept rows over the whole evaluation
 *   yv3_map_reset         start an evaluation (zeroes the counters)
 *   yv3_map_accumulate    one batch: kept_rows (n, capacity, 9) float64 and kept_counts (n, 3) as written by
 *                         yv3_collect_kept (boxes in pixels of image_w x image_h); targets (n, m * 5) fp32 as the loss
 *                         takes them ([x, y, w, h, class] normalised, -1 padded; boxes become pixels as (x -+ w/2) * image_w);
 *                         image_base = index of the batch's first image in the evaluation (ties keep image order);
 *                         pixel_inclusive != 0: the tool's "+1" box widths; score_col 7 (conf * prob) or 4 (conf);
 *                         class_agnostic != 0: every box is class 0 (class_num 0 heads)
 *   yv3_map_finalize      ap_out[num_classes] (AP, -1 for a class without ground truth), gt_count_out[num_classes] (may be
 *                         NULL), map_out2[2] = {mAP, 1.0 if more than max_detections rows arrived}.  Ends the evaluation. */
int yv3_map_state_bytes(int max_detections, int num_classes, size_t* bytes);
int yv3_map_reset(void* state, size_t state_bytes, int max_detections, int num_classes, void* stream);
int yv3_map_accumulate(const double* kept_rows, const int32_t* kept_counts, int n, int32_t capacity, const float* targets,
                       int m, double image_w, double image_h, int image_base, double min_overlap, int pixel_inclusive,
                       int score_col, int class_agnostic, void* state, void* stream);
int yv3_map_finalize(void* state, int num_classes, double* ap_out, int32_t* gt_count_out, double* map_out2, void* stream);

/* ---- the one exchange step of the sharded path: the batch mean of yolov3/yolov3_loss.py:112 over ALL shards ----------
 * Sum of every rank's loss rows (20 floats per step: 18 detail values, their sum, a pad) over the GPUs of one NVLink
 * domain, as ONE kernel over peer-mapped memory (csrc/peer_reduce.cu): ranks push their rows into each other's
 * mailboxes, raise a flag, poll their own flags and add the rows in rank order (bit-identical sums on every rank).
 * An ordinary stream-ordered launch: it can be captured in the CUDA graph that holds the step, which the library
 * collective could not.  Replaces multi_gpu_model's gather-to-CPU of yolov3/trainer.py:40-41 for this path.
 *   yv3_peer_mailbox_bytes  size of the mailbox every rank provides: device memory mapped into every peer (e.g. a
 *                           symmetric-memory allocation), zero-initialised once, 128-byte aligned
 *   yv3_peer_allreduce      mailboxes[r] = rank r's mailbox as mapped into THIS process; out[0..count) = sum over ranks
 *                           of local[0..count), count <= cap; local may equal out.  Every rank makes the same calls.
 *   yv3_peer_status         1 when a peer failed to arrive within the poll bound (~2 s) in some call, else 0 */
int yv3_peer_mailbox_bytes(int world, int cap, size_t* bytes);
int yv3_peer_allreduce(void* const* mailboxes, int world, int rank, int cap, const float* local, int count, float* out,
                       void* stream);
int yv3_peer_status(const void* mailbox, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* YV3_HEAD_H_ */
