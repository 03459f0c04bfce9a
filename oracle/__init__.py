"""TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference head path (the oracle).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import anything from this package; the product (yolov3-tensorflow_b200/) never does.
"""
