// Launcher declarations shared between the kernel translation units and the C-ABI (yv3_api.cu).
#pragma once
#include "yv3_common.cuh"

namespace yv3 {

// One staged survivor of the score sweep: the finished candidate row (yolov3_post_process.py:65-75), written while the
// record's logits are in shared memory -- the gather then only compacts rows and never goes back to the tensor.
struct __align__(16) StageRow {
    float4 box;                 // x0 / W, y0 / H, x1 / W, y1 / H
    float4 tail;                // conf, prob, class, score
};

// ---- assign.cu ---------------------------------------------------------------------------------
// Where the records of head h live: the packed tensor (one base, the head's segment inside every /32 row), or one
// tensor per head (N, H, W, B*L) -- which is the packed layout of a geometry that has only that head (a /32 row of
// it holds fold*B records starting at 0).  record = row32 * rpr + seg + cell * B + b inside an image of rpi records.
struct HeadLayout {
    const float* pred[3];
    int rpi[3], rpr[3], seg[3];
    uint32_t* objbits[3];      // object mask bits by record index in that tensor (zeroed by the caller)
};

struct AssignArgs {
    const float* predicts;    // packed form (lay is filled from it by launch_assign when lay.pred[0] is NULL)
    HeadLayout lay;
    const float* targets;
    int n, m;
    float kappa;          // upstream / n_global
    OwnEntry* own;        // [n][3][m]
    float* img_terms;     // [n][3][3]  weighted xy, wh, obj sums per head
    int32_t* status;
    uint32_t* objbits;    // object mask bits by global record index (zeroed by the caller)
    float* gtab;          // [n][3][m][8] valid labels per head (x0, y0, x1, y1, area), label order
    int32_t* vcount;      // [n]
    int32_t* table;       // optional public output [n][m][3][4]
    float* iou_out;       // optional public output [n][m][3]
};
cudaError_t launch_assign(const Geo& g, const AssignArgs& a, cudaStream_t stream);

// ---- loss_dense.cu -----------------------------------------------------------------------------
struct DenseArgs {
    const float* predicts;
    float* grad;                      // may be NULL: forward only
    const OwnEntry* own;
    const float* img_terms;
    const uint32_t* objbits;
    const float* gtab;
    const int32_t* vcount;
    float* partials;
    double* fold_slices;              // [64][9] first-level sums of the two-level fold (may be NULL: one level)
    int32_t* status;
    float* detail;                    // 20 floats: 18 detail values, their sum, pad
    int n, m, n_global;
    int num_chunks;                   // 32-record chunks (one per warp iteration)
    int cpw;                          // consecutive chunks per warp
    int bulk;                         // zero fill by bulk (TMA) stores from a shared-memory block of zeros
    int zero_bytes;                   // size of that block (multiple of 16)
    unsigned zlo[32], zhi[32];        // per lane: zero-fill pattern of a chunk (set by the launcher)
    long long total_rec;
    float kappa;                      // upstream / n_global
    float upstream;
    float rect_c[3];                  // (upstream * rw_h) / n_global: what reaches every squared logit of head h (:157-162)
    int rect_active;
    int vec_ok;                       // predicts and grad are 16-byte aligned
    int row_base;                     // first partial row of this launch (un-packed form: one launch per head tensor)
    int max_grid;                     // CTAs this launch may use (rows of the partial table it owns)
    int debug;                        // developer switches (YV3_DEBUG): 1 scalar path, 4 skip class terms, 8 skip the gradient stores, 16 labels from global memory, 64 zero fill by the lanes instead of bulk stores, 128 zero-fill warp per CTA, 256 skip value stores, 512 skip the zero warp's stores
};
cudaError_t launch_loss_dense(const Geo& g, DenseArgs a, int sm_count, bool sweep, bool fold, cudaStream_t stream, int* grid_out = nullptr);
cudaError_t launch_loss_fold(const Geo& g, const DenseArgs& a, int nrows, cudaStream_t stream);

// ---- fused_sweep.cu ----------------------------------------------------------------------------
// score sweep + dense loss sweep on one read of the packed tensor (a: as for launch_loss_dense, grad required)
int fused_row_count(const Geo& g, const float* predicts, int n, float thresh);       // spans (= partial rows) of the sweep
int packed_tile_count(const Geo& g, const float* predicts, int n, float thresh);     // tiles of 32 records of the sweep
bool fused_sweep_supported(const Geo& g, const DenseArgs& a, int n, float thresh);
cudaError_t launch_fused_sweep(const Geo& g, DenseArgs a, int n, float thresh, StageRow* stage, int32_t* tcount,
                               int32_t* post_status, int* rows_out, cudaStream_t stream);

// ---- decode.cu ---------------------------------------------------------------------------------
struct DecodeArgs {
    const float* predicts;
    int n;
    float* raw4[3];       // (N,H,W,B,4) or NULL
    float* decoded[3];    // (N,H,W,B,L) or NULL
    float* boxes[3];      // (N,H,W,B,4) or NULL
};
cudaError_t launch_decode(const Geo& g, const DecodeArgs& a, int sm_count, cudaStream_t stream);
cudaError_t launch_label_decode(const Geo& g, const float* targets, int n, int m, float* const out[6],
                                cudaStream_t stream);

cudaError_t launch_label_batch(const float* labels, const int32_t* row_offsets, const int32_t* image_hw, int n, int m,
                               int target_h, int target_w, float* targets, cudaStream_t stream);

// ---- postprocess.cu ----------------------------------------------------------------------------
struct PostWs {
    int32_t* status;            // [kStatusInts]
    float* score;               // [n][rpi]  head-raster order (filter only)
    float* prob;                // [n][rpi]
    int32_t* cls;               // [n][rpi]
    StageRow* stage;            // [num_tiles][tr] staged survivors of the filter, in ballot-rank order per (image slot, head)
    int32_t* tcount;            // [num_tiles][8] survivors per tile and (image slot, head)
    unsigned long long* keys;   // [n][cap2] sorted keys (class | ~score | row)
    float4* sbox;               // [n][capacity] boxes in sorted order
    int32_t* sidx;              // [n][capacity] original row of a sorted position
    float4* kspill;             // [n][capacity] kept boxes beyond the shared-memory list
    int2* wl;                   // [n][capacity] queue of (image, first sorted position) of segments for nms_kernel
    int32_t* wl_count;
    size_t bytes;
};
int nms_cap2(int capacity);
int score_num_tiles(const Geo& g, int n);
cudaError_t launch_collect(const float* cand, const int32_t* counts, const unsigned char* keep, int n, int capacity,
                           double tw, double th, int bug_compat, double* rows_out, int32_t* counts_out, cudaStream_t stream);
int heads_tr(const Geo& g, int n);      // records per tile of the per-head sweeps; < 4: a head has fewer than 4 records per image
cudaError_t launch_filter_packed(const Geo& g, const float* const pred[3], bool unpacked, int n, float thresh, float* cand,
                                 int capacity, int32_t* counts, int32_t* status, StageRow* stage, int32_t* tcount,
                                 bool sweep, bool gather, cudaStream_t stream);
cudaError_t launch_filter_decoded(const Geo& g, const float* const decoded[3], const float* const boxes[3],
                                  float thresh, float* cand, int capacity, int32_t* counts, int32_t* status,
                                  float* score, float* prob, int32_t* cls, int sm_count, cudaStream_t stream);
cudaError_t launch_nms(const float* cand, const int32_t* count, int count_cols, int n, int capacity,
                       int num_classes, double thresh, unsigned char* keep, const PostWs& ws,
                       cudaStream_t stream);

}  // namespace yv3
