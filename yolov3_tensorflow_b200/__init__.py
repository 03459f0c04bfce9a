"""Import alias: the product package lives in ``yolov3-tensorflow_b200/`` (a directory name
Python cannot import directly); this shim makes ``import yolov3_tensorflow_b200`` resolve to it."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "yolov3-tensorflow_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
