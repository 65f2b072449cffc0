"""`import goofit` — the reference's Python package name, served by goofit_b200: a script written for the reference
(`from goofit import *`, `import goofit.minuit2`) runs against the B200 implementation by putting this repository on sys.path."""
import sys as _sys

import goofit_b200 as _impl
from goofit_b200 import *  # noqa: F401,F403
from goofit_b200 import Lineshapes, Resonances, landau, minuit2  # noqa: F401
from goofit_b200 import __main__, __version__  # noqa: F401

__all__ = [n for n in dir(_impl) if not n.startswith("_")]
_sys.modules[__name__ + ".minuit2"] = minuit2
_sys.modules[__name__ + ".landau"] = landau  # `from goofit.landau import landau_quantile` (python/examples/simpleFit.py)
