"""ecolab_b200 -- B200-native implementation of the EcoLab model's per-timestep
population update (generate / mutate / migrate / condense of
highperformancecoder/ecolab, models/ecolab_model.cc) behind the reference's
scripted model API.

    from ecolab_b200 import PanmicticModel, SpatialModel      # explicit objects
    import ecolab_b200.compat                                 # or: make `import ecolab_model`,
                                                              # `import ecolab`, `GUI`, `plot` resolve
The compute path is libecolab_b200.so (hand-written sm_100a CUDA + C++ host model)
behind the C ABI in include/ecolab_b200.h.  There is no CPU fallback.
"""
from ._cabi import EcolabError, LIB_PATH, lib
from .model import EcoList, PanmicticModel
from .spatial import Comm, SpatialModel, slab_bounds

__all__ = ["EcolabError", "LIB_PATH", "lib", "EcoList", "PanmicticModel", "SpatialModel", "Comm", "slab_bounds"]
