"""TEST INFRASTRUCTURE ONLY (oracle): the four ``keras.backend`` symbols the reference's head path uses.

  epsilon / set_epsilon   yolov3_decoder.py:154,179,191 read it; run.py:26 sets 1e-8 (Keras default 1e-7)
  variable                yolov3_loss.py:69 (``current_num``, int32)
  update_add              yolov3_loss.py:152; queued until ``tensorflow.standin.end_step()`` (assumption A1)
"""
_EPSILON = 1e-7


def epsilon():
    return _EPSILON


def set_epsilon(value):
    global _EPSILON
    _EPSILON = float(value)


def variable(value, dtype=None, name=None):
    import tensorflow as tf
    return tf.Variable(value, tf._as_dtype(dtype), name)


def update_add(x, increment):
    import tensorflow as tf
    tf._STATE["pending"].append((x, int(increment)))
    return x
