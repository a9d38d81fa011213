"""Importable alias of the package directory `sglang-diffusion_b200/` (a hyphen cannot be imported).

`import sglang_diffusion_b200` executes `sglang-diffusion_b200/__init__.py` as THIS module
and points `__path__` at that directory, so `sglang_diffusion_b200.ops`, `.capi`, `.linear`,
`.build` are the files that live there.  No code lives here.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "sglang-diffusion_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
