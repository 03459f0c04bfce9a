// C ABI of the head path (include/yv3_head.h): argument checks, geometry, workspace carving, launches.
// No device allocation, no global state, no stream synchronisation except yv3_read_status.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "yv3_kernels.cuh"

using namespace yv3;

namespace {

int sm_count_cached() {
    // one attribute query per process and device; not "state" of the path (a hardware constant)
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cached[dev] = v;
    }
    return cached[dev];
}

bool finite_f(float v) { return isfinite(v); }

// YV3_TRACE=1 in the environment: say which CUDA call failed (stderr) before returning YV3_ERR_CUDA.
bool cuda_ok(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    if (getenv("YV3_TRACE")) fprintf(stderr, "[yv3] %s failed: %s\n", what, cudaGetErrorString(e));
    return false;
}

int check_shape(const yv3_config* c) {
    for (int h = 0; h < 3; ++h) {
        if (c->grid_h[h] <= 0 || c->grid_w[h] <= 0) return YV3_ERR_SHAPE;
        if (c->box_num[h] < 1 || c->box_num[h] > YV3_MAX_BOX) return YV3_ERR_SHAPE;
    }
    // /8, /16, /32 of one input that is a multiple of 32 (configs.py:46-48; the Reshape of
    // yolov3_detector.py:80-85 is impossible otherwise)
    if (c->grid_h[0] != 4 * c->grid_h[2] || c->grid_w[0] != 4 * c->grid_w[2]) return YV3_ERR_SHAPE;
    if (c->grid_h[1] != 2 * c->grid_h[2] || c->grid_w[1] != 2 * c->grid_w[2]) return YV3_ERR_SHAPE;
    if (c->class_num < 0 || c->class_num > 16383) return YV3_ERR_SHAPE;
    long long rpi = (long long)c->grid_h[2] * c->grid_w[2] *
                    (16LL * c->box_num[0] + 4LL * c->box_num[1] + c->box_num[2]);
    if (rpi >= (1LL << 22)) return YV3_ERR_SHAPE;      // record ids are packed into 22-bit fields
    if (c->grid_w[0] > 1024) return YV3_ERR_SHAPE;     // exact reciprocal division in locate()
    return YV3_OK;
}

int check_values(const yv3_config* c) {
    if (!finite_f(c->iou_thresh)) return YV3_ERR_CONFIG;
    for (int h = 0; h < 3; ++h) {
        for (int k = 0; k < 5; ++k)
            if (!finite_f(c->loss_weights[h][k])) return YV3_ERR_CONFIG;
        if (!finite_f(c->rectified_loss_weight[h])) return YV3_ERR_CONFIG;
        for (int b = 0; b < c->box_num[h]; ++b) {
            float w = c->anchors[h][b][0], hh = c->anchors[h][b][1];
            if (!finite_f(w) || !finite_f(hh) || w <= 0.f || hh <= 0.f) return YV3_ERR_CONFIG;
        }
    }
    if (!finite_f(c->focal_alpha) || !finite_f(c->focal_gamma)) return YV3_ERR_CONFIG;
    if (!(c->eps >= 0.0) || !(c->eps < 0.5)) return YV3_ERR_CONFIG;
    return YV3_OK;
}

Geo make_geo(const yv3_config* c) {
    Geo g;
    memset(&g, 0, sizeof(g));
    const int fold[3] = {16, 4, 1};
    g.C = c->class_num;
    g.L = 5 + c->class_num;
    g.rows32 = c->grid_h[2] * c->grid_w[2];
    int base = 0;
    for (int h = 0; h < 3; ++h) {
        g.H[h] = c->grid_h[h];
        g.W[h] = c->grid_w[h];
        g.B[h] = c->box_num[h];
        g.fold[h] = fold[h];
        g.seg_base[h] = base;
        base += fold[h] * g.B[h];
        g.head_rec[h] = g.H[h] * g.W[h] * g.B[h];
        g.fW[h] = (float)g.W[h];
        g.fH[h] = (float)g.H[h];
        g.fHW[h] = (float)(g.H[h] * g.W[h]);             // python int height*width -> fp32 (:350)
        for (int b = 0; b < g.B[h]; ++b) {
            // tf.constant(anchors, fp32) * np.array([W, H]) -> fp32 product (yolov3_decoder.py:35-40)
            g.aw[h][b] = c->anchors[h][b][0] * g.fW[h];
            g.ah[h][b] = c->anchors[h][b][1] * g.fH[h];
        }
        for (int k = 0; k < 5; ++k) g.lw[h][k] = c->loss_weights[h][k];
        g.rw[h] = c->rectified_loss_weight[h];
    }
    g.rpr = base;
    g.rpi = g.rows32 * g.rpr;
    auto magic = [](int d) { return d == 1 ? 0u : (unsigned)((1ULL << 32) / (unsigned long long)d + 1ULL); };
    g.mg_rpr = magic(g.rpr);
    for (int h = 0; h < 3; ++h) { g.mg_B[h] = magic(g.B[h]); g.mg_W[h] = magic(g.W[h]); }
    g.lo = (float)c->eps;                                 // python double -> fp32 constant
    g.hi = (float)(1.0 - c->eps);
    g.iou_thresh = c->iou_thresh;
    g.focal = c->is_focal_loss ? 1 : 0;
    g.alpha = c->focal_alpha;
    g.gamma = c->focal_gamma;
    g.tiou = c->is_tiou_recall ? 1 : 0;
    return g;
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15u) == 0; }
bool aligned4(const void* p) { return ((uintptr_t)p & 3u) == 0; }

}  // namespace

extern "C" {

int yv3_abi_version(void) { return YV3_ABI_VERSION; }

const char* yv3_strerror(int code) {
    switch (code) {
        case YV3_OK: return "ok";
        case YV3_ERR_NULL: return "required pointer is NULL";
        case YV3_ERR_SHAPE: return "bad shape: grids must be /8,/16,/32 of an input that is a multiple of 32; "
                                   "box_num in [1,8]; n, m in range";
        case YV3_ERR_CONFIG: return "bad config value (non-finite weight/threshold, anchor <= 0, eps outside [0,0.5))";
        case YV3_ERR_ALIGN: return "device pointer not aligned as documented";
        case YV3_ERR_WORKSPACE: return "workspace NULL or smaller than the query function reports";
        case YV3_ERR_CUDA: return "CUDA runtime error";
        case YV3_ERR_CAPACITY: return "candidate capacity exceeded / problem too large";
        case YV3_ERR_GT_RANGE: return "a valid label's centre lies outside the grid (tf.gather_nd would raise)";
        case YV3_ERR_NONFINITE: return "inf/NaN in the loss";
        case YV3_ERR_STATIC_SCRATCH: return "yv3_nms_static: n x capacity needs more than the library's static scratch array "
                                            "(YV3_NMS_STATIC_SCRATCH_BYTES = 4 MiB); use yv3_nms with a workspace";
        default: return "unknown yv3 error";
    }
}

int yv3_config_check(const yv3_config* cfg) {
    if (!cfg) return YV3_ERR_NULL;
    int rc = check_shape(cfg);
    if (rc != YV3_OK) return rc;
    return check_values(cfg);
}

int64_t yv3_records_per_image(const yv3_config* cfg) {
    if (!cfg || check_shape(cfg) != YV3_OK) return -1;
    return (int64_t)make_geo(cfg).rpi;
}

int64_t yv3_packed_floats_per_image(const yv3_config* cfg) {
    if (!cfg || check_shape(cfg) != YV3_OK) return -1;
    Geo g = make_geo(cfg);
    return (int64_t)g.rpi * g.L;
}

int yv3_query_workspace(const yv3_config* cfg, int n, int m, size_t* bytes) {
    if (!cfg || !bytes) return YV3_ERR_NULL;
    int rc = yv3_config_check(cfg);
    if (rc != YV3_OK) return rc;
    if (n < 1 || m < 1 || m > YV3_MAX_GT) return YV3_ERR_SHAPE;
    *bytes = carve_workspace(nullptr, n, m, make_geo(cfg).rpi).bytes;
    return YV3_OK;
}

}  // extern "C"

// ---- assignment and loss ---------------------------------------------------------------------
namespace {

int loss_common_checks(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
                       void* workspace, size_t workspace_bytes) {
    if (!cfg || !predicts || !targets) return YV3_ERR_NULL;
    int rc = yv3_config_check(cfg);
    if (rc != YV3_OK) return rc;
    if (n < 1 || m < 1 || m > YV3_MAX_GT) return YV3_ERR_SHAPE;
    if (!workspace) return YV3_ERR_WORKSPACE;
    if (!aligned16(workspace)) return YV3_ERR_ALIGN;
    if (workspace_bytes < carve_workspace(nullptr, n, m, make_geo(cfg).rpi).bytes) return YV3_ERR_WORKSPACE;
    if (!aligned4(predicts) || !aligned4(targets)) return YV3_ERR_ALIGN;
    return YV3_OK;
}

}  // namespace

extern "C" {

int yv3_assign(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
               int32_t* table, float* iou, void* workspace, size_t workspace_bytes, void* stream) {
    int rc = loss_common_checks(cfg, predicts, targets, n, m, workspace, workspace_bytes);
    if (rc != YV3_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const Geo g = make_geo(cfg);
    Workspace ws = carve_workspace(workspace, n, m, g.rpi);
    if (!cuda_ok(cudaMemsetAsync(ws.status, 0, ws.zero_bytes, st), "cudaMemsetAsync")) return YV3_ERR_CUDA;
    AssignArgs a;
    memset(&a, 0, sizeof(a));
    a.predicts = predicts; a.targets = targets; a.n = n; a.m = m;
    a.kappa = 1.0f / (float)n;
    a.own = ws.own; a.img_terms = ws.img_terms; a.status = ws.status;
    a.objbits = ws.objbits; a.gtab = ws.gtab; a.vcount = ws.vcount;
    a.table = table; a.iou_out = iou;
    if (!cuda_ok(launch_assign(g, a, st), "launch_assign")) return YV3_ERR_CUDA;
    return YV3_OK;
}

static int loss_impl(const yv3_config* cfg, const float* const pred[3], bool unpacked, const float* targets, int n,
                     int m, int n_global, int rectified_active, float upstream, float* loss_out20,
                     float* const grad[3], void* workspace, size_t workspace_bytes, int phases, void* stream) {
    int rc = loss_common_checks(cfg, pred[0], targets, n, m, workspace, workspace_bytes);
    if (rc != YV3_OK) return rc;
    if (!loss_out20) return YV3_ERR_NULL;
    if (n_global < n) return YV3_ERR_SHAPE;
    if ((long long)n * yv3_records_per_image(cfg) >= (1LL << 31)) return YV3_ERR_CAPACITY;
    const int tensors = unpacked ? 3 : 1;
    const bool want_grad = grad[0] != nullptr;
    for (int h = 0; h < tensors; ++h) {
        if (!pred[h]) return YV3_ERR_NULL;
        if (!aligned4(pred[h])) return YV3_ERR_ALIGN;
        if (want_grad && (!grad[h] || !aligned16(grad[h]))) return grad[h] ? YV3_ERR_ALIGN : YV3_ERR_NULL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const Geo g = make_geo(cfg);
    Workspace ws = carve_workspace(workspace, n, m, g.rpi);
    if ((phases & 7) == 0) return YV3_ERR_SHAPE;
    const float kappa = upstream / (float)n_global;
    // one object-bit region per tensor
    Geo gh[3];
    uint32_t* bits[3];
    {
        size_t word = 0;
        for (int h = 0; h < 3; ++h) {
            gh[h] = unpacked ? head_geo(g, h) : g;
            bits[h] = ws.objbits + (unpacked ? word : 0);
            word += (size_t)(((long long)n * gh[h].rpi + 31) / 32 + 1);
        }
    }
    if (phases & 1) {
        if (!cuda_ok(cudaMemsetAsync(ws.status, 0, ws.zero_bytes, st), "cudaMemsetAsync")) return YV3_ERR_CUDA;
        AssignArgs a;
        memset(&a, 0, sizeof(a));
        a.predicts = pred[0]; a.targets = targets; a.n = n; a.m = m;
        a.kappa = kappa;
        a.own = ws.own; a.img_terms = ws.img_terms; a.status = ws.status;
        a.objbits = ws.objbits; a.gtab = ws.gtab; a.vcount = ws.vcount;
        if (unpacked) {
            for (int h = 0; h < 3; ++h) {
                a.lay.pred[h] = pred[h];
                a.lay.rpi[h] = gh[h].rpi; a.lay.rpr[h] = gh[h].rpr; a.lay.seg[h] = 0;
                a.lay.objbits[h] = bits[h];
            }
        }
        if (!cuda_ok(launch_assign(g, a, st), "launch_assign")) return YV3_ERR_CUDA;
    }
    if (!(phases & 6)) return YV3_OK;

    DenseArgs d;
    memset(&d, 0, sizeof(d));
    d.own = ws.own; d.img_terms = ws.img_terms; d.partials = ws.partials; d.status = ws.status;
    d.gtab = ws.gtab; d.vcount = ws.vcount;
    d.detail = loss_out20;
    d.n = n; d.m = m; d.n_global = n_global;
    d.kappa = kappa; d.upstream = upstream;
    for (int h = 0; h < 3; ++h) d.rect_c[h] = (upstream * g.rw[h]) / (float)n_global;
    d.rect_active = rectified_active ? 1 : 0;
    const char* dbg = getenv("YV3_DEBUG");
    d.debug = dbg ? atoi(dbg) : 0;
    int rows = 0;
    for (int h = 0; h < tensors; ++h) {
        d.predicts = pred[h]; d.grad = grad[h];
        d.objbits = bits[h];
        d.total_rec = (long long)n * gh[h].rpi;
        d.row_base = rows;
        d.max_grid = unpacked ? kMaxGrid / 4 : kMaxGrid;
        // both tensors 16-byte aligned: the box logits of a record are read (and its gradient written) as two
        // aligned 128-bit words; otherwise the scalar path
        d.vec_ok = (aligned16(pred[h]) && (!grad[h] || aligned16(grad[h])) && !(d.debug & 1)) ? 1 : 0;
        int grid = 0;
        if (!cuda_ok(launch_loss_dense(gh[h], d, sm_count_cached(), (phases & 2) != 0, false, st, &grid), "launch_loss_dense"))
            return YV3_ERR_CUDA;
        rows += grid;
    }
    if (phases & 4) {
        if (!cuda_ok(launch_loss_fold(g, d, rows, st), "launch_loss_fold")) return YV3_ERR_CUDA;
    }
    return YV3_OK;
}

int yv3_loss_fwd_bwd(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
                     int n_global, int rectified_active, float upstream, float* loss_out20,
                     float* grad_predicts, void* workspace, size_t workspace_bytes, int phases, void* stream) {
    const float* pred[3] = {predicts, nullptr, nullptr};
    float* grad[3] = {grad_predicts, nullptr, nullptr};
    return loss_impl(cfg, pred, false, targets, n, m, n_global, rectified_active, upstream, loss_out20, grad, workspace,
                     workspace_bytes, phases, stream);
}

int yv3_loss_fwd_bwd_heads(const yv3_config* cfg, const float* const head_predicts[3], const float* targets, int n,
                           int m, int n_global, int rectified_active, float upstream, float* loss_out20,
                           float* const head_grads[3], void* workspace, size_t workspace_bytes, int phases,
                           void* stream) {
    if (!head_predicts) return YV3_ERR_NULL;
    float* none[3] = {nullptr, nullptr, nullptr};
    return loss_impl(cfg, head_predicts, true, targets, n, m, n_global, rectified_active, upstream, loss_out20,
                     head_grads ? head_grads : none, workspace, workspace_bytes, phases, stream);
}

int yv3_read_status(void* workspace, void* stream) {
    if (!workspace) return YV3_ERR_WORKSPACE;
    int32_t host[kStatusInts];
    cudaStream_t st = (cudaStream_t)stream;
    if (!cuda_ok(cudaMemcpyAsync(host, workspace, sizeof(host), cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync")) return YV3_ERR_CUDA;
    if (!cuda_ok(cudaStreamSynchronize(st), "cudaStreamSynchronize")) return YV3_ERR_CUDA;
    unsigned bits = (unsigned)host[0];
    if (bits & kStatusGtRange) return YV3_ERR_GT_RANGE;
    if (bits & kStatusNonFinite) return YV3_ERR_NONFINITE;
    if (bits & kStatusCapacity) return YV3_ERR_CAPACITY;
    return YV3_OK;
}

}  // extern "C"

// ---- decode, labels, post-process ------------------------------------------------------------------
namespace {

PostWs carve_post(void* base, long long rpi, int n, int capacity, int num_tiles = 0) {
    PostWs w;
    size_t off = 0;
    char* p = (char*)base;
    auto take = [&](size_t bytes) {
        char* r = p + off;
        off = align_up(off + bytes, 256);
        return r;
    };
    const size_t recs = (size_t)n * (size_t)rpi;
    const size_t rows = (size_t)n * (size_t)capacity;
    w.status = (int32_t*)take(sizeof(int32_t) * kStatusInts);
    w.score = (float*)take(recs * 4);
    w.prob = (float*)take(recs * 4);
    w.cls = (int32_t*)take(recs * 4);
    w.stage = (StageRow*)take((recs + 256) * sizeof(StageRow));
    w.tcount = (int32_t*)take(((size_t)num_tiles + 1) * 8 * 4);
    w.keys = (unsigned long long*)take((size_t)n * nms_cap2(capacity) * 8);
    w.sbox = (float4*)take(rows * 16);
    w.sidx = (int32_t*)take(rows * 4);
    w.kspill = (float4*)take(rows * 16);
    w.wl = (int2*)take((rows + rows / 256 + 16) * 8);      // medium-segment queue [rows], then the large-segment queue
    w.wl_count = (int32_t*)take(256);
    w.bytes = off;
    return w;
}

}  // namespace

extern "C" {

int yv3_decode(const yv3_config* cfg, const float* predicts, int n, float* const out[9], void* stream) {
    if (!cfg || !predicts || !out) return YV3_ERR_NULL;
    int rc = yv3_config_check(cfg);
    if (rc != YV3_OK) return rc;
    if (n < 1) return YV3_ERR_SHAPE;
    if (!aligned4(predicts)) return YV3_ERR_ALIGN;
    const Geo g = make_geo(cfg);
    DecodeArgs a;
    a.predicts = predicts;
    a.n = n;
    for (int h = 0; h < 3; ++h) {
        a.raw4[h] = out[3 * h + 0];
        a.decoded[h] = out[3 * h + 1];
        a.boxes[h] = out[3 * h + 2];
        if ((a.raw4[h] && !aligned16(a.raw4[h])) || (a.boxes[h] && !aligned16(a.boxes[h]))) return YV3_ERR_ALIGN;
        if (a.decoded[h] && !aligned4(a.decoded[h])) return YV3_ERR_ALIGN;
    }
    if (!cuda_ok(launch_decode(g, a, sm_count_cached(), (cudaStream_t)stream), "launch_decode")) return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_label_decode(const yv3_config* cfg, const float* targets, int n, int m, float* const out[6],
                     void* stream) {
    if (!cfg || !targets || !out) return YV3_ERR_NULL;
    int rc = yv3_config_check(cfg);
    if (rc != YV3_OK) return rc;
    if (n < 1 || m < 1) return YV3_ERR_SHAPE;
    for (int h = 0; h < 3; ++h)
        if (out[2 * h + 1] && !aligned16(out[2 * h + 1])) return YV3_ERR_ALIGN;
    const Geo g = make_geo(cfg);
    if (!cuda_ok(launch_label_decode(g, targets, n, m, out, (cudaStream_t)stream), "launch_label_decode")) return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_label_batch(const float* labels, const int32_t* row_offsets, const int32_t* image_hw, int n, int m,
                    int target_h, int target_w, float* targets, void* stream) {
    if (!row_offsets || !image_hw || !targets) return YV3_ERR_NULL;       // labels may be NULL when no image has rows
    if (n < 1 || m < 1 || m > YV3_MAX_GT || target_h < 1 || target_w < 1) return YV3_ERR_SHAPE;
    if (!cuda_ok(launch_label_batch(labels, row_offsets, image_hw, n, m, target_h, target_w, targets,
                                    (cudaStream_t)stream), "launch_label_batch")) return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_query_postprocess_workspace(const yv3_config* cfg, int n, int32_t capacity, size_t* bytes) {
    if (!bytes) return YV3_ERR_NULL;
    if (n < 1 || capacity < 1 || capacity > (1 << 18)) return YV3_ERR_CAPACITY;
    long long rpi = 0;
    if (cfg) {
        int rc = yv3_config_check(cfg);
        if (rc != YV3_OK) return rc;
        rpi = make_geo(cfg).rpi;
    }
    *bytes = carve_post(nullptr, rpi, n, capacity, cfg ? score_num_tiles(make_geo(cfg), n) : 0).bytes;
    return YV3_OK;
}

static int post_checks(const yv3_config* cfg, int n, int32_t capacity, void* workspace, size_t workspace_bytes,
                       PostWs* ws) {
    if (n < 1) return YV3_ERR_SHAPE;
    if (capacity < 1 || capacity > (1 << 18)) return YV3_ERR_CAPACITY;
    if (!workspace) return YV3_ERR_WORKSPACE;
    if (!aligned16(workspace)) return YV3_ERR_ALIGN;
    long long rpi = 0;
    if (cfg) {
        int rc = yv3_config_check(cfg);
        if (rc != YV3_OK) return rc;
        rpi = make_geo(cfg).rpi;
    }
    *ws = carve_post(workspace, rpi, n, capacity, cfg ? score_num_tiles(make_geo(cfg), n) : 0);
    if (workspace_bytes < ws->bytes) return YV3_ERR_WORKSPACE;
    return YV3_OK;
}

static int filter_impl(const yv3_config* cfg, const float* const pred[3], bool unpacked, int n, float score_thresh,
                       float* cand, int32_t capacity, int32_t* counts, void* workspace, size_t workspace_bytes,
                       int phases, void* stream) {
    if ((phases & 3) == 0) return YV3_ERR_SHAPE;
    if (!cfg || !cand || !counts) return YV3_ERR_NULL;
    for (int h = 0; h < (unpacked ? 3 : 1); ++h) {
        if (!pred[h]) return YV3_ERR_NULL;
        if (!aligned4(pred[h])) return YV3_ERR_ALIGN;
    }
    PostWs ws;
    int rc = post_checks(cfg, n, capacity, workspace, workspace_bytes, &ws);
    if (rc != YV3_OK) return rc;
    if (!aligned16(cand)) return YV3_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if ((phases & 1) && !cuda_ok(cudaMemsetAsync(ws.status, 0, sizeof(int32_t) * kStatusInts, st), "cudaMemsetAsync"))
        return YV3_ERR_CUDA;
    const Geo g = make_geo(cfg);
    if (unpacked && heads_tr(g, n) < 4) return YV3_ERR_SHAPE;      // a head with fewer than 4 records per image
    if (!cuda_ok(launch_filter_packed(g, pred, unpacked, n, score_thresh, cand, capacity, counts, ws.status, ws.stage,
                                      ws.tcount, (phases & 1) != 0, (phases & 2) != 0, st), "launch_filter_packed"))
        return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_filter_packed(const yv3_config* cfg, const float* predicts, int n, float score_thresh, float* cand,
                      int32_t capacity, int32_t* counts, void* workspace, size_t workspace_bytes, int phases,
                      void* stream) {
    const float* pred[3] = {predicts, nullptr, nullptr};
    return filter_impl(cfg, pred, false, n, score_thresh, cand, capacity, counts, workspace, workspace_bytes, phases, stream);
}

int yv3_filter_heads(const yv3_config* cfg, const float* const head_predicts[3], int n, float score_thresh, float* cand,
                     int32_t capacity, int32_t* counts, void* workspace, size_t workspace_bytes, int phases,
                     void* stream) {
    if (!head_predicts) return YV3_ERR_NULL;
    return filter_impl(cfg, head_predicts, true, n, score_thresh, cand, capacity, counts, workspace, workspace_bytes, phases,
                       stream);
}

int yv3_loss_filter_fused(const yv3_config* cfg, const float* predicts, const float* targets, int n, int m,
                          int n_global, int rectified_active, float upstream, float* loss_out20, float* grad_predicts,
                          void* loss_workspace, size_t loss_workspace_bytes, float score_thresh, float* cand,
                          int32_t capacity, int32_t* counts, void* post_workspace, size_t post_workspace_bytes,
                          int32_t* fused_out, int phases, void* stream) {
    if ((phases & 15) == 0) return YV3_ERR_SHAPE;
    int rc = loss_common_checks(cfg, predicts, targets, n, m, loss_workspace, loss_workspace_bytes);
    if (rc != YV3_OK) return rc;
    if (!loss_out20 || !grad_predicts || !cand || !counts) return YV3_ERR_NULL;
    if (!aligned16(grad_predicts) || !aligned16(cand)) return YV3_ERR_ALIGN;
    if (n_global < n) return YV3_ERR_SHAPE;
    if ((long long)n * yv3_records_per_image(cfg) >= (1LL << 31)) return YV3_ERR_CAPACITY;
    PostWs pws;
    rc = post_checks(cfg, n, capacity, post_workspace, post_workspace_bytes, &pws);
    if (rc != YV3_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const Geo g = make_geo(cfg);
    Workspace ws = carve_workspace(loss_workspace, n, m, g.rpi);
    DenseArgs d;
    memset(&d, 0, sizeof(d));
    d.predicts = predicts; d.grad = grad_predicts;
    d.own = ws.own; d.img_terms = ws.img_terms; d.partials = ws.partials; d.status = ws.status;
    d.objbits = ws.objbits; d.gtab = ws.gtab; d.vcount = ws.vcount;
    d.fold_slices = ws.fold_slices;
    d.detail = loss_out20;
    d.n = n; d.m = m; d.n_global = n_global;
    d.kappa = upstream / (float)n_global; d.upstream = upstream;
    for (int h = 0; h < 3; ++h) d.rect_c[h] = (upstream * g.rw[h]) / (float)n_global;
    d.rect_active = rectified_active ? 1 : 0;
    const char* dbg = getenv("YV3_DEBUG");
    d.debug = dbg ? atoi(dbg) : 0;
    d.vec_ok = (aligned16(predicts) && !(d.debug & 1)) ? 1 : 0;
    const bool fused = fused_sweep_supported(g, d, n, score_thresh);
    if (fused_out) *fused_out = fused ? 1 : 0;
    if (!fused) {
        // the two sweeps one after the other (unaligned tensors, records too long for a tile, more tiles than partial rows)
        rc = yv3_loss_fwd_bwd(cfg, predicts, targets, n, m, n_global, rectified_active, upstream, loss_out20, grad_predicts,
                              loss_workspace, loss_workspace_bytes, phases & 7, stream);
        if ((phases & 7) && rc != YV3_OK) return rc;
        const int fp = ((phases & 2) ? 1 : 0) | ((phases & 8) ? 2 : 0);
        if (!fp) return YV3_OK;
        return yv3_filter_packed(cfg, predicts, n, score_thresh, cand, capacity, counts, post_workspace, post_workspace_bytes,
                                 fp, stream);
    }
    // 1: assignment (labels, object bits, owner terms); 2: the one sweep; 4, 8: the two small tails (fold, gather)
    if (phases & 1) {
        rc = yv3_loss_fwd_bwd(cfg, predicts, targets, n, m, n_global, rectified_active, upstream, loss_out20, grad_predicts,
                              loss_workspace, loss_workspace_bytes, 1, stream);
        if (rc != YV3_OK) return rc;
    }
    int rows = fused_row_count(g, predicts, n, score_thresh);
    if (phases & 2) {
        if (!cuda_ok(launch_fused_sweep(g, d, n, score_thresh, pws.stage, pws.tcount, pws.status, &rows, st), "launch_fused_sweep"))
            return YV3_ERR_CUDA;
    }
    if ((phases & 4) && !cuda_ok(launch_loss_fold(g, d, rows, st), "launch_loss_fold")) return YV3_ERR_CUDA;
    if (phases & 8) {
        const float* pred[3] = {predicts, nullptr, nullptr};
        if (!cuda_ok(launch_filter_packed(g, pred, false, n, score_thresh, cand, capacity, counts, pws.status, pws.stage,
                                          pws.tcount, false, true, st), "launch_filter_packed"))
            return YV3_ERR_CUDA;
    }
    return YV3_OK;
}

int yv3_filter_decoded(const yv3_config* cfg, const float* const decoded[3], const float* const boxes[3],
                       float score_thresh, float* cand, int32_t capacity, int32_t* counts, void* workspace,
                       size_t workspace_bytes, void* stream) {
    if (!cfg || !decoded || !boxes || !cand || !counts) return YV3_ERR_NULL;
    for (int h = 0; h < 3; ++h) {
        if (!decoded[h] || !boxes[h]) return YV3_ERR_NULL;
        if (!aligned16(boxes[h]) || !aligned4(decoded[h])) return YV3_ERR_ALIGN;
    }
    PostWs ws;
    int rc = post_checks(cfg, 1, capacity, workspace, workspace_bytes, &ws);
    if (rc != YV3_OK) return rc;
    if (!aligned16(cand)) return YV3_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (!cuda_ok(cudaMemsetAsync(ws.status, 0, sizeof(int32_t) * kStatusInts, st), "cudaMemsetAsync")) return YV3_ERR_CUDA;
    const Geo g = make_geo(cfg);
    if (!cuda_ok(launch_filter_decoded(g, decoded, boxes, score_thresh, cand, capacity, counts, ws.status, ws.score,
                              ws.prob, ws.cls, sm_count_cached(), st), "launch_filter_decoded")) return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_collect_kept(const float* cand, const int32_t* counts, const uint8_t* keep, int n, int32_t capacity,
                     double target_w, double target_h, int bug_compat, double* rows_out, int32_t* counts_out,
                     void* stream) {
    if (!cand || !counts || !keep || !rows_out || !counts_out) return YV3_ERR_NULL;
    if (n < 1 || capacity < 1 || capacity > YV3_MAX_CANDIDATES) return YV3_ERR_SHAPE;
    if (!aligned16(cand) || ((uintptr_t)rows_out & 7u)) return YV3_ERR_ALIGN;
    if (!cuda_ok(launch_collect(cand, counts, keep, n, capacity, target_w, target_h, bug_compat ? 1 : 0, rows_out, counts_out,
                                (cudaStream_t)stream), "launch_collect")) return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_loss_filter_fused_heads(const yv3_config* cfg, const float* const head_predicts[3], const float* targets, int n,
                                int m, int n_global, int rectified_active, float upstream, float* loss_out20,
                                float* const head_grads[3], void* loss_workspace, size_t loss_workspace_bytes,
                                float score_thresh, float* cand, int32_t capacity, int32_t* counts, void* post_workspace,
                                size_t post_workspace_bytes, int32_t* fused_out, int phases, void* stream) {
    if ((phases & 15) == 0) return YV3_ERR_SHAPE;
    if (!head_predicts || !head_grads) return YV3_ERR_NULL;
    int rc = loss_common_checks(cfg, head_predicts[0], targets, n, m, loss_workspace, loss_workspace_bytes);
    if (rc != YV3_OK) return rc;
    if (!loss_out20 || !cand || !counts) return YV3_ERR_NULL;
    if (!aligned16(cand)) return YV3_ERR_ALIGN;
    if (n_global < n) return YV3_ERR_SHAPE;
    if ((long long)n * yv3_records_per_image(cfg) >= (1LL << 31)) return YV3_ERR_CAPACITY;
    for (int h = 0; h < 3; ++h) {
        if (!head_predicts[h] || !head_grads[h]) return YV3_ERR_NULL;
        if (!aligned4(head_predicts[h]) || !aligned16(head_grads[h])) return YV3_ERR_ALIGN;
    }
    PostWs pws;
    rc = post_checks(cfg, n, capacity, post_workspace, post_workspace_bytes, &pws);
    if (rc != YV3_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const Geo g = make_geo(cfg);
    Workspace ws = carve_workspace(loss_workspace, n, m, g.rpi);
    const char* dbg = getenv("YV3_DEBUG");
    const int debug = dbg ? atoi(dbg) : 0;
    // every head tensor is a packed tensor of a geometry that has just that head (yolov3_detector.py:80-85 without the
    // concatenate): the one-sweep kernel runs once per tensor, on that tensor's object bits, tile range and partial rows
    Geo gh[3];
    uint32_t* bits[3];
    DenseArgs d[3];
    bool fused = heads_tr(g, n) == 32;
    {
        size_t word = 0;
        for (int h = 0; h < 3; ++h) {
            gh[h] = head_geo(g, h);
            bits[h] = ws.objbits + word;
            word += (size_t)(((long long)n * gh[h].rpi + 31) / 32 + 1);
            memset(&d[h], 0, sizeof(DenseArgs));
            d[h].predicts = head_predicts[h]; d[h].grad = head_grads[h];
            d[h].own = ws.own; d[h].img_terms = ws.img_terms; d[h].partials = ws.partials; d[h].status = ws.status;
            d[h].objbits = bits[h]; d[h].gtab = ws.gtab; d[h].vcount = ws.vcount; d[h].fold_slices = ws.fold_slices;
            d[h].detail = loss_out20;
            d[h].n = n; d[h].m = m; d[h].n_global = n_global;
            d[h].kappa = upstream / (float)n_global; d[h].upstream = upstream;
            for (int k = 0; k < 3; ++k) d[h].rect_c[k] = (upstream * g.rw[k]) / (float)n_global;
            d[h].rect_active = rectified_active ? 1 : 0;
            d[h].debug = debug;
            d[h].vec_ok = (aligned16(head_predicts[h]) && !(debug & 1)) ? 1 : 0;
            fused = fused && fused_sweep_supported(gh[h], d[h], n, score_thresh);
        }
    }
    int rows_total = 0;
    for (int h = 0; h < 3 && fused; ++h) rows_total += fused_row_count(gh[h], head_predicts[h], n, score_thresh);
    if (rows_total > kMaxGrid) fused = false;
    if (fused_out) *fused_out = fused ? 1 : 0;
    if (!fused) {
        // the two sweeps per head tensor, same results (unaligned tensors, a head with fewer than 32 records per image)
        rc = yv3_loss_fwd_bwd_heads(cfg, head_predicts, targets, n, m, n_global, rectified_active, upstream, loss_out20,
                                    head_grads, loss_workspace, loss_workspace_bytes, phases & 7, stream);
        if ((phases & 7) && rc != YV3_OK) return rc;
        const int fp = ((phases & 2) ? 1 : 0) | ((phases & 8) ? 2 : 0);
        if (!fp) return YV3_OK;
        return yv3_filter_heads(cfg, head_predicts, n, score_thresh, cand, capacity, counts, post_workspace,
                                post_workspace_bytes, fp, stream);
    }
    if (phases & 1) {
        rc = yv3_loss_fwd_bwd_heads(cfg, head_predicts, targets, n, m, n_global, rectified_active, upstream, loss_out20,
                                    head_grads, loss_workspace, loss_workspace_bytes, 1, stream);
        if (rc != YV3_OK) return rc;
    }
    if (phases & 2) {
        int rows = 0, tbase = 0;
        for (int h = 0; h < 3; ++h) {
            d[h].row_base = rows;
            int r = 0;
            if (!cuda_ok(launch_fused_sweep(gh[h], d[h], n, score_thresh, pws.stage + (size_t)tbase * 32, pws.tcount + (size_t)tbase * 8,
                                            pws.status, &r, st), "launch_fused_sweep"))
                return YV3_ERR_CUDA;
            rows += r;
            tbase += packed_tile_count(gh[h], head_predicts[h], n, score_thresh);
        }
    }
    if ((phases & 4) && !cuda_ok(launch_loss_fold(g, d[0], rows_total, st), "launch_loss_fold")) return YV3_ERR_CUDA;
    if (phases & 8) {
        if (!cuda_ok(launch_filter_packed(g, head_predicts, true, n, score_thresh, cand, capacity, counts, pws.status, pws.stage,
                                          pws.tcount, false, true, st), "launch_filter_packed"))
            return YV3_ERR_CUDA;
    }
    return YV3_OK;
}

int yv3_nms(const yv3_config* cfg, const float* cand, const int32_t* counts, int count_cols, int n,
            int32_t capacity, int num_classes, double nms_thresh, uint8_t* keep, void* workspace,
            size_t workspace_bytes, void* stream) {
    if (!cand || !counts || !keep) return YV3_ERR_NULL;
    if (count_cols < 1 || count_cols > 8 || num_classes < 1 || num_classes > 16384) return YV3_ERR_SHAPE;
    PostWs ws;
    int rc = post_checks(cfg, n, capacity, workspace, workspace_bytes, &ws);
    if (rc != YV3_OK) return rc;
    if (!aligned16(cand)) return YV3_ERR_ALIGN;
    if (!cuda_ok(launch_nms(cand, counts, count_cols, n, capacity, num_classes, nms_thresh, keep, ws,
                   (cudaStream_t)stream), "launch_nms")) return YV3_ERR_CUDA;
    return YV3_OK;
}

}  // extern "C"
