"""`ecolab_model`: the two global model objects the reference registers with
CLASSDESC_ADD_GLOBAL (models/ecolab_model.cc:32-41), created lazily on first use
so that importing the module does not touch the GPU."""
import os as _os


class _Lazy:
    def __init__(self, factory):
        object.__setattr__(self, "_f", factory)
        object.__setattr__(self, "_o", None)

    def _obj(self):
        if object.__getattribute__(self, "_o") is None:
            object.__setattr__(self, "_o", object.__getattribute__(self, "_f")())
        return object.__getattribute__(self, "_o")

    def __getattr__(self, n):
        return getattr(self._obj(), n)

    def __setattr__(self, n, v):
        setattr(self._obj(), n, v)


def _device():
    return int(_os.environ.get("LOCAL_RANK", _os.environ.get("ECOLAB_DEVICE", "0")))


def _make_panmictic():
    from ecolab_b200 import PanmicticModel
    return PanmicticModel(_device())


def _make_spatial():
    from ecolab_b200 import Comm, SpatialModel
    comm = None
    if int(_os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            torch.cuda.set_device(_device())
            dist.init_process_group("nccl")
        comm = Comm()
    return SpatialModel(_device(), comm)


panmictic_ecolab = _Lazy(_make_panmictic)
spatial_ecolab = _Lazy(_make_spatial)
