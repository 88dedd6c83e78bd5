"""Load the reference's OWN hot-path sources, by path and unmodified, on top of the keras stand-in (keras_torch.py).

TEST INFRASTRUCTURE — never imported by bayesflow_b200/ or by bench.py; works only where /root/reference exists (this
container; NOT the GPU box — the `-m gpu` tests use the golden vectors tools/make_golden.py writes from here).

How: a meta-path finder claims every `bayesflow[.x.y]` import.
  * a plain module (`bayesflow/utils/integrate.py`, `.../time_mlp.py`, `.../flow_matching.py`, ...) is executed from
    the reference file exactly as it lies on disk;
  * a PACKAGE (`bayesflow`, `bayesflow.utils`, `bayesflow.networks`, ...) is synthesised empty — the reference's
    `__init__.py` files import the whole library (matplotlib, every network family) and are not on the path — and
    resolves `from package import name` lazily: to what the package's real `__init__.py` binds that name to (the file
    is PARSED for its relative imports, never executed), else a sub-module of that name, else the module inside the
    package's directory whose source defines `name` at top level (that file is then executed, again unmodified);
  * `bayesflow.adapters.Adapter` (a type annotation in utils/integrate.py:10) is a placeholder class.
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.util
import os
import re
import sys
import types

REF_ROOT = os.environ.get("BAYESFLOW_REFERENCE", "/root/reference")
PKG_DIR = os.path.join(REF_ROOT, "bayesflow")

_OVERRIDES = {
    ("bayesflow.adapters", "Adapter"): type("Adapter", (), {}),
}


class ReferenceUnavailable(RuntimeError):
    pass


def available() -> bool:
    return os.path.isfile(os.path.join(PKG_DIR, "utils", "integrate.py"))


def _path_of(fullname):
    rel = fullname.split(".")[1:]
    base = os.path.join(PKG_DIR, *rel)
    if os.path.isdir(base):
        return base, True
    if os.path.isfile(base + ".py"):
        return base + ".py", False
    return None, False


def _defines(path, name):
    try:
        src = open(path, encoding="utf-8").read()
    except OSError:
        return False
    return re.search(rf"^(?:def|class)\s+{re.escape(name)}\b|^{re.escape(name)}\s*(?::[^=\n]+)?=", src, re.M) is not None


def _find_definer(pkg_fullname, pkg_dir, name):
    """Depth-first search of the package directory for the module that defines `name` at top level."""
    for entry in sorted(os.listdir(pkg_dir)):
        p = os.path.join(pkg_dir, entry)
        if entry.endswith(".py") and entry != "__init__.py" and _defines(p, name):
            return f"{pkg_fullname}.{entry[:-3]}"
    for entry in sorted(os.listdir(pkg_dir)):
        p = os.path.join(pkg_dir, entry)
        if os.path.isdir(p) and not entry.startswith("_") and not entry.startswith("."):
            hit = _find_definer(f"{pkg_fullname}.{entry}", p, name)
            if hit:
                return hit
    return None


def _init_exports(directory):
    """name -> (level, module, attr) for every relative `from .x import a as b` of the package's real __init__.py"""
    import ast
    path = os.path.join(directory, "__init__.py")
    out = {}
    try:
        tree = ast.parse(open(path, encoding="utf-8").read())
    except (OSError, SyntaxError):
        return out
    for node in tree.body:
        if isinstance(node, ast.ImportFrom) and node.level >= 1:
            for a in node.names:
                if a.name != "*":
                    out[a.asname or a.name] = (node.level, node.module, a.name)
    return out


class _SynthPackage(types.ModuleType):
    """When the import system binds a freshly imported sub-module `x` onto its parent, keep what the reference's
    __init__.py binds `x` to instead (e.g. `from .integrate import integrate` rebinds the name to the function)."""

    def __setattr__(self, name, value):
        exports = self.__dict__.get("_exports", {})
        if isinstance(value, types.ModuleType) and name in exports and exports[name][1] is not None:
            level, mod, attr = exports[name]
            if level == 1 and mod == name and hasattr(value, attr):
                value = getattr(value, attr)
        super().__setattr__(name, value)


class _SynthPackageLoader(importlib.abc.Loader):
    def __init__(self, directory):
        self.directory = directory

    def create_module(self, spec):
        mod = _SynthPackage(spec.name)
        mod.__path__ = [self.directory]
        mod.__package__ = spec.name
        return mod

    def exec_module(self, module):
        fullname, directory = module.__name__, self.directory

        exports = _init_exports(directory)
        module.__dict__["_exports"] = exports

        def __getattr__(name):
            if name.startswith("__"):
                raise AttributeError(name)
            if (fullname, name) in _OVERRIDES:
                return _OVERRIDES[(fullname, name)]
            value = None
            if name in exports:  # what the reference's own __init__.py binds this name to (parsed, not executed)
                level, mod, attr = exports[name]
                base = fullname if level == 1 else ".".join(fullname.split(".")[: 1 - level])
                if mod is None:
                    value = importlib.import_module(f"{base}.{attr}")
                else:
                    value = getattr(importlib.import_module(f"{base}.{mod}"), attr)
            else:
                sub, _ = _path_of(f"{fullname}.{name}")
                if sub is not None:
                    return importlib.import_module(f"{fullname}.{name}")
                definer = _find_definer(fullname, directory, name) if os.path.isdir(directory) else None
                if definer is None:
                    raise AttributeError(f"reference package {fullname!r} has no attribute {name!r}")
                value = getattr(importlib.import_module(definer), name)
            setattr(module, name, value)
            return value

        module.__getattr__ = __getattr__


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname != "bayesflow" and not fullname.startswith("bayesflow."):
            return None
        if fullname == "bayesflow":
            return importlib.util.spec_from_loader(fullname, _SynthPackageLoader(PKG_DIR), is_package=True)
        p, is_dir = _path_of(fullname)
        if p is None:
            if fullname == "bayesflow.adapters":
                return importlib.util.spec_from_loader(fullname, _SynthPackageLoader(os.path.join(PKG_DIR, "adapters")), is_package=True)
            return None
        if is_dir:
            return importlib.util.spec_from_loader(fullname, _SynthPackageLoader(p), is_package=True)
        return importlib.util.spec_from_file_location(fullname, p)  # the reference file itself, unmodified


_INSTALLED = False


def install():
    """keras stand-in + the finder; returns the (synthesised) top-level `bayesflow` package."""
    global _INSTALLED
    if not available():
        raise ReferenceUnavailable(f"{PKG_DIR} not found (the reference exists only in the build container)")
    here = os.path.dirname(os.path.abspath(__file__))
    if here not in sys.path:
        sys.path.insert(0, here)
    import keras_torch
    keras_torch.install()
    if not _INSTALLED:
        if "bayesflow" in sys.modules and not isinstance(getattr(sys.modules["bayesflow"], "__spec__", None) and
                                                          sys.modules["bayesflow"].__spec__.loader, _SynthPackageLoader):
            raise RuntimeError("a real bayesflow is already imported")
        sys.meta_path.insert(0, _Finder())
        _INSTALLED = True
    return importlib.import_module("bayesflow")


def reference():
    """-> namespace with the reference classes / functions the pin tests use (each executed from its own file)."""
    install()
    ns = types.SimpleNamespace()
    ns.integrate_mod = importlib.import_module("bayesflow.utils.integrate")
    ns.integrate = ns.integrate_mod.integrate
    ns.TimeMLP = importlib.import_module("bayesflow.networks.subnets.mlp.time_mlp").TimeMLP
    ns.ConditionalDenseBlock = importlib.import_module("bayesflow.networks.helpers.blocks.conditional_dense_block").ConditionalDenseBlock
    ns.FourierEmbedding = importlib.import_module("bayesflow.networks.helpers.embeddings.fourier_embedding").FourierEmbedding
    ns.FiLM = importlib.import_module("bayesflow.networks.helpers.embeddings.featurewise_linear_modulation").FiLM
    ns.FlowMatching = importlib.import_module("bayesflow.networks.inference.flow_matching.flow_matching").FlowMatching
    ns.DiffusionModel = importlib.import_module("bayesflow.networks.inference.diffusion.diffusion_model").DiffusionModel
    ns.DiagonalNormal = importlib.import_module("bayesflow.distributions.diagonal_normal").DiagonalNormal
    sched = "bayesflow.networks.inference.diffusion.schedules."
    ns.EDMNoiseSchedule = importlib.import_module(sched + "edm_noise_schedule").EDMNoiseSchedule
    ns.CosineNoiseSchedule = importlib.import_module(sched + "cosine_noise_schedule").CosineNoiseSchedule
    return ns


def load_params(ref_net, params):
    """Copy an oracle parameter dict ('subnet.blocks.0.dense.kernel', ... — the reference's variable paths, kernels
    (in, out)) into a BUILT reference FlowMatching / DiffusionModel.  Raises on any name that has no home."""
    import torch
    alias = {"norm": "norm_layer", "frequencies": "w"}
    for name, value in params.items():
        obj = ref_net
        parts = name.split(".")
        for part in parts[:-1]:
            part = alias.get(part, part)
            obj = obj[int(part)] if part.isdigit() else getattr(obj, part)
        leaf = alias.get(parts[-1], parts[-1])
        target = getattr(obj, leaf)
        if target is None or tuple(target.shape) != tuple(value.shape):
            raise ValueError(f"{name}: reference variable {None if target is None else tuple(target.shape)} vs {tuple(value.shape)}")
        with torch.no_grad():
            target.copy_(value)
    n_ref = sum(w.numel() for w in ref_net.subnet.weights) + sum(w.numel() for w in ref_net.output_projector.weights)
    n_par = sum(v.numel() for v in params.values())
    if n_ref != n_par:
        raise ValueError(f"parameter count mismatch: reference holds {n_ref}, dict holds {n_par}")
