"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Imports the *unmodified* reference LightGlue (gluefactory/models/matchers/lightglue.py)
straight from the read-only tree at /root/reference, in THIS container only (the tree
does not exist on the GPU box).  Used by oracle/make_golden.py and by the CPU tests that
pin oracle/lightglue_oracle.py against the real reference (they skip when the tree is
absent).

Recipe:
  * `omegaconf` is not installed -> register a ~30-line stub that offers exactly what
    lightglue.py:369 and utils/losses.py:36 use (merge / create / to_container, attribute access);
  * register empty parent packages with __path__ pointing into the tree, so that
    gluefactory/__init__.py (which drags in h5py, kornia ...) is never executed;
  * adaptive configs need two shims because the reference crashes without them
    (SURVEY.md 8(a) row A12): a `confidence_thresholds` buffer, and a `torch.stack([])`
    guard for the early-exit return at lightglue.py:572.
"""
import importlib
import os
import sys
import types
import warnings

import torch

REF_ROOT = os.environ.get("LG_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "gluefactory/models/matchers/lightglue.py"))


class _Cfg(dict):
    """dict with attribute access (stands in for omegaconf.DictConfig)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:  # pragma: no cover
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def _wrap(d):
    if isinstance(d, dict):
        return _Cfg({k: _wrap(v) for k, v in d.items()})
    return d


def _merge(*cfgs):
    out = _Cfg()
    for c in cfgs:
        for k, v in dict(c).items():
            if isinstance(v, dict) and isinstance(out.get(k), dict):
                out[k] = _merge(out[k], v)
            else:
                out[k] = _wrap(v)
    return out


def _install_omegaconf_stub():
    try:
        import omegaconf  # noqa: F401

        return
    except ImportError:
        pass
    m = types.ModuleType("omegaconf")

    import contextlib

    class OmegaConf:
        merge = staticmethod(_merge)
        create = staticmethod(lambda d=None: _wrap({} if d is None else d))
        # base_model.py:72,83-84 (used by the reference's NearestNeighborMatcher, a BaseModel): no-ops here
        set_struct = staticmethod(lambda conf, flag: None)
        set_readonly = staticmethod(lambda conf, flag: None)

        @staticmethod
        def to_container(c, resolve=False):
            return {k: (OmegaConf.to_container(v) if isinstance(v, dict) else v) for k, v in c.items()}

    m.OmegaConf = OmegaConf
    m.DictConfig = _Cfg
    m.read_write = lambda conf: contextlib.nullcontext()
    m.open_dict = lambda conf: contextlib.nullcontext()
    sys.modules["omegaconf"] = m


_REF_MOD = None


def load_reference_module():
    """Return the reference module object `gluefactory.models.matchers.lightglue`."""
    global _REF_MOD
    if _REF_MOD is not None:
        return _REF_MOD
    if not reference_available():
        raise RuntimeError(f"reference tree not found under {REF_ROOT}")
    _install_omegaconf_stub()
    for name, rel in [
        ("gluefactory", "gluefactory"),
        ("gluefactory.models", "gluefactory/models"),
        ("gluefactory.models.matchers", "gluefactory/models/matchers"),
        ("gluefactory.models.utils", "gluefactory/models/utils"),
    ]:
        if name not in sys.modules:
            pkg = types.ModuleType(name)
            pkg.__path__ = [os.path.join(REF_ROOT, rel)]
            sys.modules[name] = pkg
    flash_was = torch.backends.cuda.flash_sdp_enabled()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _REF_MOD = importlib.import_module("gluefactory.models.matchers.lightglue")
    torch.backends.cuda.enable_flash_sdp(flash_was)
    return _REF_MOD


class _TorchProxy:
    """Module-global `torch` as seen by the reference, with stack([]) -> empty tensor."""

    def __init__(self, real):
        self._real = real

    def __getattr__(self, k):
        return getattr(self._real, k)

    def stack(self, tensors, *a, **kw):
        if len(tensors) == 0:
            return self._real.empty(0)
        return self._real.stack(tensors, *a, **kw)


def build_reference(conf: dict, state_dict=None, dtype=torch.float32):
    """Instantiate the reference LightGlue in eval mode (CPU) and load `state_dict`."""
    import contextlib
    import io

    lg = load_reference_module()
    with contextlib.redirect_stdout(io.StringIO()):  # swallow the fork's debug prints
        model = lg.LightGlue(dict(conf))
    if state_dict is not None:
        missing, unexpected = model.load_state_dict(state_dict, strict=False)
        missing = [k for k in missing if not k.startswith("loss_fn")]
        assert not missing and not unexpected, (missing, unexpected)
    model = model.eval().to(dtype)
    # A12(i): the attribute the adaptive path reads but the fork never defines
    model.confidence_thresholds = [
        float(model.confidence_threshold(i)) for i in range(model.conf.n_layers)
    ]
    # A12(ii): early exit returns instead of raising at torch.stack([])
    if not isinstance(lg.torch, _TorchProxy):
        lg.torch = _TorchProxy(torch)
    return model


@torch.no_grad()
def run_reference(model, data: dict):
    d = dict(data)
    d.setdefault("view0", {})
    d.setdefault("view1", {})
    return model(d)


def load_reference_nn_module():
    """The reference module `gluefactory.models.matchers.nearest_neighbor_matcher` (find_nn, mutual_check,
    NearestNeighborMatcher), imported from the read-only tree like the LightGlue module."""
    load_reference_module()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return importlib.import_module("gluefactory.models.matchers.nearest_neighbor_matcher")
