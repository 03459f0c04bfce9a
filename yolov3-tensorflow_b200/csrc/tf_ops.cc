// TensorFlow custom-op shim over the C ABI of include/yv3_head.h  (tf.load_op_library).
//
// SOURCE ONLY on this image: TensorFlow (headers, libtensorflow_framework.so) is not installed and cannot be
// installed, so this file is NOT part of build(), has never been compiled here and is untested.  INTEGRATION.md §4
// holds the build line and the Python side (gradient registration, YOLOv3Loss wrapper).  It exists because the
// reference drives the head path from a TF graph (yolov3/trainer.py:76-84 compiles the model with
// YOLOv3Loss(...).loss; run.py:62 calls YOLOv3Decoder.decode eagerly): the ops below let that code keep its shape.
//
//   Yolov3HeadLoss(predicts, targets, rectified_active) -> (loss, loss_detail[6,3], grad_predicts)
//       yolov3/yolov3_loss.py:81-138 + the backward TF autodiff derives; gradient op: dy * grad_predicts
//   Yolov3Decode(predicts) -> 9 tensors (raw4, decoded, boxes per head)      yolov3/yolov3_decoder.py:62-87
//   Yolov3FilterNms(predicts) -> (cand[N,cap,8], counts[N,3], keep[N,cap])    run.py:63-74 batched
//
// Attributes mirror struct yv3_config field for field.
#if defined(YV3_WITH_TENSORFLOW)

#define EIGEN_USE_GPU
#include "tensorflow/core/framework/op.h"
#include "tensorflow/core/framework/op_kernel.h"
#include "tensorflow/core/framework/shape_inference.h"
#include <stdlib.h>
#include <string.h>

#include <string>

#include "yv3_head.h"

namespace tf = tensorflow;

namespace {

#define YV3_CONFIG_ATTRS                                                                     \
    .Attr("grid_h: list(int)")                                                              \
    .Attr("grid_w: list(int)")                                                              \
    .Attr("box_num: list(int)")                                                             \
    .Attr("class_num: int")                                                                 \
    .Attr("anchors: list(float)") /* flattened per head: w0,h0,w1,h1,... in head order */   \
    .Attr("iou_thresh: float = 0.5")                                                        \
    .Attr("loss_weights: list(float) = [1,1,1,1,1, 1,1,1,1,1, 1,1,1,1,1]")                  \
    .Attr("rectified_loss_weight: list(float) = [0.01, 0.01, 0.01]")                        \
    .Attr("is_focal_loss: bool = false")                                                    \
    .Attr("focal_alpha: float = 0.25")                                                      \
    .Attr("focal_gamma: float = 2.0")                                                       \
    .Attr("is_tiou_recall: bool = false")                                                   \
    .Attr("eps: string = '1e-7'") /* decimal text of the Python double (K.epsilon()): a float attr would round it to fp32 */

tf::Status ReadConfig(tf::OpKernelConstruction* c, yv3_config* cfg) {
    std::vector<int> gh, gw, bn;
    std::vector<float> anchors, lw, rw;
    bool focal, tiou;
    std::string eps;
    memset(cfg, 0, sizeof(*cfg));               // anchors past box_num and the padding bytes are defined
    TF_RETURN_IF_ERROR(c->GetAttr("grid_h", &gh));
    TF_RETURN_IF_ERROR(c->GetAttr("grid_w", &gw));
    TF_RETURN_IF_ERROR(c->GetAttr("box_num", &bn));
    TF_RETURN_IF_ERROR(c->GetAttr("class_num", &cfg->class_num));
    TF_RETURN_IF_ERROR(c->GetAttr("anchors", &anchors));
    TF_RETURN_IF_ERROR(c->GetAttr("iou_thresh", &cfg->iou_thresh));
    TF_RETURN_IF_ERROR(c->GetAttr("loss_weights", &lw));
    TF_RETURN_IF_ERROR(c->GetAttr("rectified_loss_weight", &rw));
    TF_RETURN_IF_ERROR(c->GetAttr("is_focal_loss", &focal));
    TF_RETURN_IF_ERROR(c->GetAttr("focal_alpha", &cfg->focal_alpha));
    TF_RETURN_IF_ERROR(c->GetAttr("focal_gamma", &cfg->focal_gamma));
    TF_RETURN_IF_ERROR(c->GetAttr("is_tiou_recall", &tiou));
    TF_RETURN_IF_ERROR(c->GetAttr("eps", &eps));
    if (gh.size() != 3 || gw.size() != 3 || bn.size() != 3 || lw.size() != 15 || rw.size() != 3)
        return tf::errors::InvalidArgument("grid_h/grid_w/box_num need 3 entries, loss_weights 15, rectified 3");
    size_t need = 0;
    for (int h = 0; h < 3; ++h) need += 2 * (size_t)bn[h];
    if (anchors.size() != need) return tf::errors::InvalidArgument("anchors must hold 2 floats per box");
    size_t k = 0;
    for (int h = 0; h < 3; ++h) {
        cfg->grid_h[h] = gh[h]; cfg->grid_w[h] = gw[h]; cfg->box_num[h] = bn[h];
        if (bn[h] < 1 || bn[h] > YV3_MAX_BOX) return tf::errors::InvalidArgument("box_num out of range");
        for (int b = 0; b < bn[h]; ++b) { cfg->anchors[h][b][0] = anchors[k++]; cfg->anchors[h][b][1] = anchors[k++]; }
        for (int j = 0; j < 5; ++j) cfg->loss_weights[h][j] = lw[h * 5 + j];
        cfg->rectified_loss_weight[h] = rw[h];
    }
    cfg->is_focal_loss = focal ? 1 : 0;
    cfg->is_tiou_recall = tiou ? 1 : 0;
    {
        char* end = nullptr;
        cfg->eps = strtod(eps.c_str(), &end);   // the double the Python side holds, not its fp32 rounding
        if (end == eps.c_str() || *end != '\0') return tf::errors::InvalidArgument("eps must be a decimal number");
    }
    const int rc = yv3_config_check(cfg);
    if (rc != YV3_OK) return tf::errors::InvalidArgument(yv3_strerror(rc));
    return tf::Status::OK();
}

void* StreamOf(tf::OpKernelContext* ctx) { return (void*)ctx->eigen_device<Eigen::GpuDevice>().stream(); }

}  // namespace

REGISTER_OP("Yolov3HeadLoss")
    .Input("predicts: float")            // (N, H/32, W/32, Ctot) packed layout
    .Input("targets: float")             // (N, M*5), -1 padded
    .Input("rectified_active: int32")    // scalar, HOST memory: value of current_num <= rectified_coord_num (:125)
    .Output("loss: float")
    .Output("loss_detail: float")        // (6, 3): xy, wh, noobj, obj, cls, rectified x /8,/16,/32
    .Output("grad_predicts: float")
    YV3_CONFIG_ATTRS
    .SetShapeFn([](tf::shape_inference::InferenceContext* c) {
        c->set_output(0, c->Scalar());
        c->set_output(1, c->Matrix(6, 3));
        c->set_output(2, c->input(0));
        return tf::Status::OK();
    });

class Yolov3HeadLossOp : public tf::OpKernel {
 public:
    explicit Yolov3HeadLossOp(tf::OpKernelConstruction* c) : tf::OpKernel(c) { OP_REQUIRES_OK(c, ReadConfig(c, &cfg_)); }
    void Compute(tf::OpKernelContext* ctx) override {
        const tf::Tensor& p = ctx->input(0);
        const tf::Tensor& t = ctx->input(1);
        const int active = ctx->input(2).scalar<int>()();
        OP_REQUIRES(ctx, p.dims() == 4 && t.dims() == 2 && t.dim_size(0) == p.dim_size(0) && t.dim_size(1) % 5 == 0,
                    tf::errors::InvalidArgument("predicts (N,H32,W32,Ctot), targets (N,M*5)"));
        const int n = (int)p.dim_size(0), m = (int)(t.dim_size(1) / 5);
        OP_REQUIRES(ctx, p.NumElements() == (tf::int64)n * yv3_packed_floats_per_image(&cfg_),
                    tf::errors::InvalidArgument("predicts does not have the packed head layout of this configuration"));
        tf::Tensor *loss, *detail, *grad, out20, ws;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, tf::TensorShape({}), &loss));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(1, tf::TensorShape({6, 3}), &detail));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(2, p.shape(), &grad));
        size_t bytes = 0;
        OP_REQUIRES(ctx, yv3_query_workspace(&cfg_, n, m, &bytes) == YV3_OK, tf::errors::Internal("workspace query"));
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(tf::DT_UINT8, tf::TensorShape({(tf::int64)bytes}), &ws));
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(tf::DT_FLOAT, tf::TensorShape({20}), &out20));
        void* st = StreamOf(ctx);
        const int rc = yv3_loss_fwd_bwd(&cfg_, p.flat<float>().data(), t.flat<float>().data(), n, m, /*n_global=*/n,
                                        active, 1.0f, out20.flat<float>().data(), grad->flat<float>().data(),
                                        ws.flat<tf::uint8>().data(), bytes, 7, st);
        OP_REQUIRES(ctx, rc == YV3_OK, tf::errors::Internal(yv3_strerror(rc)));
        cudaError_t e = cudaMemcpyAsync(detail->flat<float>().data(), out20.flat<float>().data(), 18 * sizeof(float),
                                        cudaMemcpyDeviceToDevice, (cudaStream_t)st);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(loss->flat<float>().data(), out20.flat<float>().data() + 18, sizeof(float),
                                cudaMemcpyDeviceToDevice, (cudaStream_t)st);
        OP_REQUIRES(ctx, e == cudaSuccess, tf::errors::Internal(cudaGetErrorString(e)));
    }

 private:
    yv3_config cfg_;
};
REGISTER_KERNEL_BUILDER(Name("Yolov3HeadLoss").Device(tf::DEVICE_GPU).HostMemory("rectified_active"), Yolov3HeadLossOp);

REGISTER_OP("Yolov3Decode")
    .Input("predicts: float")
    .Output("raw8: float").Output("decoded8: float").Output("boxes8: float")
    .Output("raw16: float").Output("decoded16: float").Output("boxes16: float")
    .Output("raw32: float").Output("decoded32: float").Output("boxes32: float")
    YV3_CONFIG_ATTRS;

class Yolov3DecodeOp : public tf::OpKernel {
 public:
    explicit Yolov3DecodeOp(tf::OpKernelConstruction* c) : tf::OpKernel(c) { OP_REQUIRES_OK(c, ReadConfig(c, &cfg_)); }
    void Compute(tf::OpKernelContext* ctx) override {
        const tf::Tensor& p = ctx->input(0);
        OP_REQUIRES(ctx, p.dims() == 4 && p.dim_size(0) >= 1 &&
                             p.NumElements() == p.dim_size(0) * yv3_packed_floats_per_image(&cfg_),
                    tf::errors::InvalidArgument("predicts does not have the packed head layout of this configuration"));
        const int n = (int)p.dim_size(0), L = 5 + cfg_.class_num;
        float* out[9];
        for (int h = 0; h < 3; ++h) {
            const tf::int64 H = cfg_.grid_h[h], W = cfg_.grid_w[h], B = cfg_.box_num[h];
            const tf::TensorShape shapes[3] = {{n, H, W, B, 4}, {n, H, W, B, L}, {n, H, W, B, 4}};
            for (int k = 0; k < 3; ++k) {
                tf::Tensor* o;
                OP_REQUIRES_OK(ctx, ctx->allocate_output(3 * h + k, shapes[k], &o));
                out[3 * h + k] = o->flat<float>().data();
            }
        }
        const int rc = yv3_decode(&cfg_, p.flat<float>().data(), n, out, StreamOf(ctx));
        OP_REQUIRES(ctx, rc == YV3_OK, tf::errors::Internal(yv3_strerror(rc)));
    }

 private:
    yv3_config cfg_;
};
REGISTER_KERNEL_BUILDER(Name("Yolov3Decode").Device(tf::DEVICE_GPU), Yolov3DecodeOp);

REGISTER_OP("Yolov3FilterNms")
    .Input("predicts: float")
    .Output("candidates: float")         // (N, capacity, 8)  rows of yolov3_post_process.py:65-75
    .Output("counts: int32")             // (N, 3) per head
    .Output("keep: uint8")               // (N, capacity) true greedy survivors (the F9 quirk is layered in Python)
    .Attr("score_thresh: float")         // compared in fp32 (:62): the float attr IS the reference's fp32(thresh)
    .Attr("nms_thresh: string")          // decimal text of the Python double: IoU > thresh is decided in fp64 (:124), and
                                         // 0.45f widened would be 0.44999998807907104, another threshold
    .Attr("capacity: int")
    YV3_CONFIG_ATTRS;

class Yolov3FilterNmsOp : public tf::OpKernel {
 public:
    explicit Yolov3FilterNmsOp(tf::OpKernelConstruction* c) : tf::OpKernel(c) {
        OP_REQUIRES_OK(c, ReadConfig(c, &cfg_));
        OP_REQUIRES_OK(c, c->GetAttr("score_thresh", &score_));
        std::string nms;
        OP_REQUIRES_OK(c, c->GetAttr("nms_thresh", &nms));
        char* end = nullptr;
        nms_ = strtod(nms.c_str(), &end);
        OP_REQUIRES(c, end != nms.c_str() && *end == '\0', tf::errors::InvalidArgument("nms_thresh must be a decimal number"));
        OP_REQUIRES_OK(c, c->GetAttr("capacity", &cap_));
    }
    void Compute(tf::OpKernelContext* ctx) override {
        const tf::Tensor& p = ctx->input(0);
        OP_REQUIRES(ctx, p.dims() == 4 && p.dim_size(0) >= 1 &&
                             p.NumElements() == p.dim_size(0) * yv3_packed_floats_per_image(&cfg_),
                    tf::errors::InvalidArgument("predicts does not have the packed head layout of this configuration"));
        const int n = (int)p.dim_size(0);
        tf::Tensor *cand, *counts, *keep, ws;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, tf::TensorShape({n, cap_, YV3_CAND_COLS}), &cand));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(1, tf::TensorShape({n, 3}), &counts));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(2, tf::TensorShape({n, cap_}), &keep));
        size_t bytes = 0;
        OP_REQUIRES(ctx, yv3_query_postprocess_workspace(&cfg_, n, cap_, &bytes) == YV3_OK, tf::errors::Internal("workspace"));
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(tf::DT_UINT8, tf::TensorShape({(tf::int64)bytes}), &ws));
        void* st = StreamOf(ctx);
        int rc = yv3_filter_packed(&cfg_, p.flat<float>().data(), n, score_, cand->flat<float>().data(), cap_,
                                   counts->flat<tf::int32>().data(), ws.flat<tf::uint8>().data(), bytes, 3, st);
        OP_REQUIRES(ctx, rc == YV3_OK, tf::errors::Internal(yv3_strerror(rc)));
        rc = yv3_nms(&cfg_, cand->flat<float>().data(), counts->flat<tf::int32>().data(), 3, n, cap_,
                     cfg_.class_num > 0 ? cfg_.class_num : 1, nms_, keep->flat<tf::uint8>().data(),
                     ws.flat<tf::uint8>().data(), bytes, st);
        OP_REQUIRES(ctx, rc == YV3_OK, tf::errors::Internal(yv3_strerror(rc)));
    }

 private:
    yv3_config cfg_;
    float score_;
    double nms_;
    int cap_;
};
REGISTER_KERNEL_BUILDER(Name("Yolov3FilterNms").Device(tf::DEVICE_GPU), Yolov3FilterNmsOp);

#endif  // YV3_WITH_TENSORFLOW
