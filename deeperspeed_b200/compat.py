"""`deepspeed` import alias. Maps the reference module paths onto this package so that user scripts doing
`import deepspeed` / `from deepspeed.runtime.pipe.schedule import TrainSchedule` and pickled checkpoint objects
(`deepspeed.runtime.fp16.loss_scaler.DynamicLossScaler`) resolve to the B200-native implementation."""
import importlib
import importlib.abc
import importlib.util
import sys

_PKG = "deeperspeed_b200"
# reference path prefix -> our path prefix (longest prefix wins)
_REMAP = {
    "deepspeed.pt": _PKG + ".pt",
    # attributes of the reference package that people also import as modules
    "deepspeed.checkpointing": _PKG + ".runtime.activation_checkpointing.checkpointing",
    "deepspeed.zero": _PKG + ".runtime.zero",
    "deepspeed.ops.op_builder": _PKG + ".ops.op_builder",
    "deepspeed": _PKG,
}


def _translate(name):
    best = None
    for src in _REMAP:
        if name == src or name.startswith(src + "."):
            if best is None or len(src) > len(best):
                best = src
    if best is None:
        return None
    return _REMAP[best] + name[len(best):]


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, target):
        self.target = target

    def create_module(self, spec):
        return importlib.import_module(self.target)

    def exec_module(self, module):
        pass


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        if fullname != "deepspeed" and not fullname.startswith("deepspeed."):
            return None
        real = _translate(fullname)
        if real is None:
            return None
        try:
            real_spec = importlib.util.find_spec(real)
        except (ImportError, ValueError):
            return None
        if real_spec is None:
            return None
        spec = importlib.util.spec_from_loader(fullname, _AliasLoader(real), is_package=real_spec.submodule_search_locations is not None)
        return spec


_installed = False


def install(force=False):
    global _installed
    if _installed:
        return
    if "deepspeed" in sys.modules and not force:
        mod = sys.modules["deepspeed"]
        if getattr(mod, "__name__", "") != _PKG:
            raise ImportError("a different `deepspeed` package is already imported; pass force=True to override")
    sys.meta_path.insert(0, _AliasFinder())
    for k in [k for k in sys.modules if k == "deepspeed" or k.startswith("deepspeed.")]:
        del sys.modules[k]
    _installed = True
