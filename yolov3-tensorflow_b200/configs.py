"""The switches of the reference's configs.py that the head path consumes (configs.py:36-72, :99-100),
same names and shipped values, as a plain attribute bag that imports without easydict / np.int.

    from yolov3_tensorflow_b200.configs import FLAGS
    loss = YOLOv3Loss(FLAGS.head_grid_sizes, FLAGS.class_num, FLAGS.anchor_boxes, FLAGS.iou_thresh,
                      FLAGS.loss_weights, rectified_coord_num=FLAGS.rectified_coord_num, ...)   # trainer.py:76-83
"""


class _Flags(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def derive(flags):
    """configs.py:44-48: box_len, head_channel_nums, head_grid_sizes from input_image_size."""
    h, w = flags.input_image_size[0], flags.input_image_size[1]
    if h % 32 or w % 32:
        raise ValueError("input_image_size must be a multiple of 32 (configs.py:46-48)")
    flags.box_len = 4 + 1 + flags.class_num
    flags.head_channel_nums = [len(a) * flags.box_len for a in flags.anchor_boxes]
    flags.head_grid_sizes = [(h // 8, w // 8), (h // 16, w // 16), (h // 32, w // 32)]
    return flags


FLAGS = _Flags()
FLAGS.input_image_size = (384, 480, 3)                      # configs.py:36
FLAGS.anchor_boxes = [                                       # configs.py:37-41 (normalised w, h; 3 / 2 / 3)
    [(0.06618181818181816, 0.1025177510694752), (0.18544278606965178, 0.13160367921287464),
     (0.13, 0.32733333333333337)],
    [(0.13, 0.32733333333333337), (0.303806787732042, 0.34370030784316496)],
    [(0.303806787732042, 0.34370030784316496), (0.4667050847457627, 0.5281262429095761),
     (0.7906945888923907, 0.7888860433597275)],
]
FLAGS.class_num = 0                                          # configs.py:42
FLAGS.iou_thresh = 0.8                                       # configs.py:50
FLAGS.loss_weights = [(5, 5, 0.05, 3, 1), (8, 8, 0.05, 2, 1), (10, 10, 0.05, 2, 1)]   # configs.py:52
FLAGS.batch_size = 3                                         # configs.py:56
FLAGS.rectified_coord_num = 1464                             # configs.py:58
FLAGS.rectified_loss_weight = [1, 1, 1]                      # configs.py:59
FLAGS.is_label_smoothing = False                             # configs.py:67 (never read by the path)
FLAGS.is_focal_loss = False                                  # configs.py:68
FLAGS.focal_alpha = 1.0                                      # configs.py:69
FLAGS.focal_gamma = 2.0                                      # configs.py:70
FLAGS.is_gradient_harmonized = False                         # configs.py:71 (never read by the path)
FLAGS.is_tiou_recall = False                                 # configs.py:72
FLAGS.confidence_thresh = 0.8                                # configs.py:99
FLAGS.nms_thresh = 0.4                                       # configs.py:100
FLAGS.epsilon = 1e-8                                         # run.py:26
derive(FLAGS)
