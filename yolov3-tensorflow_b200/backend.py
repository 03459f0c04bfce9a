"""The one piece of Keras backend state the head path reads: keras.backend.epsilon().

The reference clips sigmoid/softmax outputs to [eps, 1 - eps] (yolov3_decoder.py:154,179,191) with
``keras.backend.epsilon()``, which run.py:26 sets to 1e-8 before anything is built; Keras' own
default is 1e-7.  The value is read when a decoder / loss object is constructed, as in the
reference (where it is baked into the graph at build time).
"""
_EPSILON = 1e-7


def epsilon():
    return _EPSILON


def set_epsilon(value):
    global _EPSILON
    _EPSILON = float(value)
