"""Alias of ``b200flow.utils``: importing this package yields that very module object."""
import sys as _sys

import b200flow.utils as _impl

_sys.modules[__name__] = _impl
