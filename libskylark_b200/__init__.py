"""libskylark_b200 -- B200-native sketch-apply path behind libSkylark's Python API.

    from libskylark_b200 import sketch
    sketch.initialize(123834)
    S = sketch.CWT(n, s)
    SA = S.apply(A, None, "columnwise")     # numpy / scipy.sparse / torch CUDA tensors

The package mirrors python-skylark/skylark/{lib,sketch,errors}.py; all arithmetic runs in
libcskylark_b200.so (hand-written sm_100a CUDA behind a C ABI).  There is no CPU path.
"""
from . import errors  # noqa: F401
from . import lib  # noqa: F401
from . import sketch  # noqa: F401
