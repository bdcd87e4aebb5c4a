"""insb200 -- B200-native time-stepping core behind the operator API of IncompressibleNavierStokes.jl.

Package directory: `incompressiblenavierstokes.jl_b200/` (import it through the `insb200` alias
module at the repo root).  Contents: `csrc/` (CUDA kernels + C ABI), `lib/` (built
libinsb200.so), `_lib.py` (ctypes prototypes), `api.py` / `processors.py` / `initializers.py`
(host mirror of the reference's Julia API for the hot path).
"""
from . import _lib  # noqa: F401
from .api import *  # noqa: F401,F403
from .api import Observable, ExplicitRungeKuttaMethod, decomposition, get_state, slab_bounds  # noqa: F401
from .initializers import create_spectrum, random_field, velocityfield  # noqa: F401
from .processors import checkpointer, load_checkpoint, save_checkpoint  # noqa: F401
from .processors import energy_history, fieldsaver, observefield, observespectrum, processor, spectral_stuff, timelogger  # noqa: F401
