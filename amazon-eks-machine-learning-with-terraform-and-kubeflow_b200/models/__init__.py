"""Alias of ``b200flow.models``: importing this package yields that very module object."""
import sys as _sys

import b200flow.models as _impl

_sys.modules[__name__] = _impl
