"""Import alias: the package sources live in ``vllm-metal_b200/`` (a directory
name Python cannot import); this shim makes them importable as
``vllm_metal_b200`` without copying anything."""

import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "vllm-metal_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _os, _f, _real
