"""goofit_b200 — B200-native (sm_100a) implementation of GooFit's likelihood-evaluation hot path.

`from goofit_b200 import *` mirrors `from goofit import *` for the classes on that path: the pybind11 module
`_goofit` wraps the host C++ API (GooFit class names), which calls hand-written CUDA kernels through the C ABI
in include/goofit_b200.h. There is no CPU fallback: evaluating a PDF without a CUDA device raises GeneralError.
"""
import os as _os

_here = _os.path.dirname(_os.path.abspath(__file__))
LIB_DIR = _os.path.join(_here, "lib")
CAPI_PATH = _os.path.join(LIB_DIR, "libgoofit_b200.so")

try:
    from ._goofit import *  # noqa: F401,F403
    from . import _goofit
    from ._goofit import Resonances, Lineshapes, minuit2  # noqa: F401
    from ._goofit import __version__, __main__  # noqa: F401  (the star import skips underscored names)
    import sys as _sys
    _sys.modules[__name__ + ".minuit2"] = minuit2  # `import goofit_b200.minuit2` / `from goofit_b200.minuit2 import MnMigrad` (upstream: goofit.minuit2)
except ImportError as e:  # pragma: no cover - build missing
    raise ImportError(
        "goofit_b200: the native extension is not built (run `make` or `python -c 'import __graft_entry__ as g; g.build()'`): %s" % e
    )



def _help(cls):
    """Upstream's PDFs carry a static help() that prints their documentation; here it prints the class's docstring and the
    signatures of its constructors."""
    text = (cls.__doc__ or cls.__name__).strip() + "\n" + (cls.__init__.__doc__ or "")
    print(text)
    return text


for _name in dir(_goofit):
    _cls = getattr(_goofit, _name)
    if isinstance(_cls, type) and issubclass(_cls, _goofit.PdfBase) and "help" not in _cls.__dict__:
        _cls.help = classmethod(_help)
for _sub in (Resonances, Lineshapes):
    for _name in dir(_sub):
        _cls = getattr(_sub, _name)
        if isinstance(_cls, type) and issubclass(_cls, _goofit.PdfBase) and "help" not in _cls.__dict__:
            _cls.help = classmethod(_help)
del _name, _cls, _sub

from . import capi  # noqa: E402,F401
from . import landau  # noqa: E402,F401  (upstream: goofit.landau, python/extras/landau.cpp)
