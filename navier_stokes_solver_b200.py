"""Import shim: the package directory is ``navier-stokes-solver_b200/`` (the name the project
layout prescribes), which is not a valid Python identifier.  ``import navier_stokes_solver_b200``
resolves to that directory."""
import os as _os

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "navier-stokes-solver_b200")
__path__ = [_dir]
__package__ = __name__
if __spec__ is not None:
    __spec__.submodule_search_locations = __path__
_init = _os.path.join(_dir, "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"))
