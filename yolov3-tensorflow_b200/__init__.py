"""B200-native detection-head path of zheng-yuwei/YOLOv3-tensorflow.

Host-side mirror of the reference's interface for the path (same class names, arguments, return
structures) over the C-ABI library ``lib/libyv3head.so`` (include/yv3_head.h):

    YOLOv3Decoder        yolov3/yolov3_decoder.py
    LabelDecoder         yolov3/label_decoder.py
    LabelBatcher         dataset/file_util.py:47-52,97 (letter-box label transform + padded batch)
    YOLOv3Loss           yolov3/yolov3_loss.py   (forward + analytic backward, fused)
    YOLOv3PostProcessor  yolov3/yolov3_post_process.py
    HeadPostProcessor    batched device-resident filter + NMS (the run.py:59-74 loop as one call)
    FusedPipeline        loss fwd+bwd + filter + NMS of consecutive steps on three streams (what bench.py times)
    backend              keras.backend.epsilon() / set_epsilon()
    configs              the configs.py switches the path consumes
"""
from . import backend  # noqa: F401
from ._lib import to_device_f32 as to_device  # noqa: F401
from .label_decoder import LabelDecoder  # noqa: F401
from .label_ingest import LabelBatcher, parse_label_line, read_label_file  # noqa: F401
from .loss_detail_log import detail_log_values, detail_scalars, format_detail  # noqa: F401
from .yolov3_decoder import YOLOv3Decoder  # noqa: F401
from .yolov3_loss import YOLOv3Loss  # noqa: F401
from .yolov3_post_process import HeadPostProcessor, YOLOv3PostProcessor  # noqa: F401
from .pipeline import FusedPipeline  # noqa: F401
from .peer_reduce import PeerAllReduce  # noqa: F401
from .map_eval import MeanAPAccumulator  # noqa: F401

__all__ = ["YOLOv3Decoder", "LabelDecoder", "LabelBatcher", "parse_label_line", "read_label_file", "YOLOv3Loss", "YOLOv3PostProcessor", "HeadPostProcessor", "FusedPipeline", "PeerAllReduce", "MeanAPAccumulator", "backend", "to_device",
           "detail_log_values", "detail_scalars", "format_detail"]
