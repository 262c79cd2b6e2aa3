"""tamcmc_b200 -- B200-native hot path of TAMCMC behind the C-ABI of include/tamcmc_gpu.h.

The directory name (``tamcmc-c-_b200``) is not an importable identifier; load it with::

    import importlib.util, sys
    spec = importlib.util.spec_from_file_location("tamcmc_b200", "<repo>/tamcmc-c-_b200/__init__.py",
                                                  submodule_search_locations=["<repo>/tamcmc-c-_b200"])
    mod = importlib.util.module_from_spec(spec); sys.modules["tamcmc_b200"] = mod; spec.loader.exec_module(mod)

(``tests/conftest.py``, ``bench.py`` and ``__graft_entry__.py`` do exactly that.)

Python here is harness only: ctypes marshalling over ``libtamcmc_gpu.so`` and the host-side
mirror of the reference's operator interface (``Model_def`` / ``MALA``).  Nothing in this package
computes a model, a likelihood or a prior on the CPU; if the CUDA library is missing or no GPU is
present the calls raise.
"""
from . import workloads  # noqa: F401
from . import sharding  # noqa: F401
from .build import build_library, LIB_PATH  # noqa: F401
from .gpu import (GpuLibrary, GpuModelDef, GpuMALA, GpuBatch, MalaConfig, TgError, load_library,  # noqa: F401
                  ST_EMPTY_WINDOW, ST_NAN_WINDOW, ST_NAN_LOGL, ST_BAD_PRIOR, ST_CHOL_FAIL, ST_MODE_OVERFLOW)

from . import outputs  # noqa: F401,E402
from . import frontend  # noqa: F401,E402
from . import config_tree  # noqa: F401,E402

__all__ = ["workloads", "build_library", "LIB_PATH", "GpuLibrary", "GpuModelDef", "GpuMALA", "GpuBatch", "MalaConfig",
           "TgError", "load_library"]
