"""Alias of ``b200flow.ops``: importing this package yields that very module object."""
import sys as _sys

import b200flow.ops as _impl

_sys.modules[__name__] = _impl
