"""Import alias for the package whose sources live in ``../mast3r-slam_b200/`` (a hyphen is not a
legal Python identifier, so ``import mast3r_slam_b200`` lands here and is redirected)."""
import os as _os

_impl = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "mast3r-slam_b200")
__path__[:] = [_impl]
_init = _os.path.join(_impl, "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"), globals())
del _os, _f
