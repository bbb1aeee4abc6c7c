"""Import shim for the live reference (TEST INFRASTRUCTURE ONLY — never imported by the product).

Makes the unmodified reference package importable by stubbing the five non-arithmetic modules it imports at
module scope (h5py, matplotlib, mpl_toolkits, mplhep, torchsde) when they are absent.  The reference is looked
up in `oracle/_ref/` first (the verbatim copy staged by `oracle/build_ref.sh`, which travels to the GPU box with
the gpurun snapshot) and in `/root/reference` otherwise (build container).  Users: `oracle/make_golden*.py`,
`tests/` (drop-in checks) and `bench.py`'s reference legs.  Never imported by `calodiffusion_b200/`.
"""
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(_HERE, "_ref") if os.path.isdir(os.path.join(_HERE, "_ref", "calodiffusion")) else "/root/reference"


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "calodiffusion"))


def _make_stub(name):
    mod = types.ModuleType(name)
    mod.__path__ = []

    def __getattr__(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        if attr[:1].isupper():
            return type(attr, (), {"__init__": lambda self, *a, **k: None})
        return lambda *a, **k: None

    mod.__getattr__ = __getattr__
    return mod


def install():
    import torch, numpy, scipy, einops  # noqa: F401  (real modules first)

    names = [
        "h5py", "matplotlib", "matplotlib.pyplot", "matplotlib.gridspec", "matplotlib.ticker",
        "matplotlib.colors", "mpl_toolkits", "mpl_toolkits.axes_grid1",
        "mpl_toolkits.axes_grid1.inset_locator", "mplhep", "torchsde",
    ]
    for n in names:
        top = n.split(".")[0]
        try:
            if top not in sys.modules:
                __import__(top)
                continue
        except ImportError:
            pass
        if n not in sys.modules:
            sys.modules[n] = _make_stub(n)
            if "." in n:
                parent, child = n.rsplit(".", 1)
                setattr(sys.modules[parent], child, sys.modules[n])
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)


def load_reference_model(config_name, sampler="DDPM", overrides=None):
    """Build the reference CaloDiffusion (CPU) for one of its shipped configs."""
    install()
    from calodiffusion.utils import utils as ref_utils
    from calodiffusion.models.calodiffusion import CaloDiffusion

    cfg = ref_utils.LoadJson(f"{REF_ROOT}/calodiffusion/configs/{config_name}")
    cfg["SAMPLER"] = sampler
    if overrides:
        cfg.update(overrides)
    model = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"])
    model.eval()
    return model, cfg
