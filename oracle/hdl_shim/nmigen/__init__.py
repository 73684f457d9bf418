"""TEST INFRASTRUCTURE -- a small interpreter of the nMigen constructs the reference
design uses, so that /root/reference/mfcc/core/*.py and mfcc/misc/*.py can be executed
here without the nmigen package (tests/golden/make_rtl_golden.py).  It is NOT nMigen
and implements one clock domain, exact integer semantics and a cycle simulator only.
The product never imports it."""
from ._ast import *            # noqa: F401,F403
from ._ast import __all__ as _a
from ._dsl import *            # noqa: F401,F403
from ._dsl import __all__ as _d
from .hdl.rec import Record    # noqa: F401

__all__ = list(_a) + list(_d) + ["Record"]
__version__ = "0.3.shim"
