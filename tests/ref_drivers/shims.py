"""Import shims for third-party packages the reference's drivers import but this image does not have (plotting,
augmentation, CNN zoos): permissive stand-ins, installed only for modules that are really missing.  None of them
is on the hot path — the drivers use them for figures, TensorBoard scalars and model classes the tests do not
instantiate.  Test infrastructure only."""
import importlib
import sys
import types


class _Stub:
    """Accepts any attribute access, call, indexing or 2-way unpacking (fig, ax = plt.subplots(...))."""
    def __init__(self, *a, **k):
        pass

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Stub()

    def __call__(self, *a, **k):
        return _Stub()

    def __getitem__(self, key):
        return _Stub()

    def __iter__(self):
        return iter((_Stub(), _Stub()))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Stub


MISSING_OK = [
    "matplotlib", "matplotlib.pyplot", "seaborn", "albumentations", "albumentations.pytorch",
    "albumentations.pytorch.transforms", "trimesh", "efficientnet_pytorch", "perceiver_pytorch", "timm",
]


def install():
    installed = []
    for name in MISSING_OK:
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
        except Exception:
            mod = _StubModule(name)
            mod.__path__ = []          # looks like a package, so that submodule imports resolve through sys.modules
            sys.modules[name] = mod
            parent, _, child = name.rpartition(".")
            if parent and parent in sys.modules:
                setattr(sys.modules[parent], child, mod)
            installed.append(name)
    # pyplot: module-level functions must be callables returning stubs, not the _Stub class itself
    plt = sys.modules.get("matplotlib.pyplot")
    if isinstance(plt, _StubModule):
        for fn in ("subplots", "figure", "savefig", "close", "show", "plot", "imshow", "legend", "title"):
            setattr(plt, fn, _Stub())
    sb = sys.modules.get("seaborn")
    if isinstance(sb, _StubModule):
        sb.barplot = _Stub()
    return installed
