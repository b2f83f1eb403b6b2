"""Loads the UNMODIFIED reference modules by explicit file path — TEST / BASELINE INFRASTRUCTURE, NOT PRODUCT.

The reference's `utils/` is a namespace package and this repository ships a drop-in package of the same name, so a
plain `import utils.fusion_model` says nothing about whose code ran.  Everything that needs the reference itself (the
golden generator, the `--impl reference` arm of bench.py, the cpu_baseline leg, the tests that run the reference's
loops over the drop-ins) goes through this loader instead: each module is executed from its file under a private name
(`tdssf_ref.<name>`), its `__file__` is asserted to lie under the reference root, and the imports it makes itself
(`from LovaszSoftmax.pytorch.lovasz_losses import ...`, `from utils.evaluation import evaluate`, `import wandb`,
`from utils.plot import ...`) are satisfied from an explicit table that is installed in `sys.modules` only for the
duration of the load.

Roots, in order: $TDSSF_REFERENCE or /root/reference (build container), else oracle/_ref/ (the verbatim copy made by
oracle/make_ref.py, hashes checked against its MANIFEST.json; this is what exists on the GPU box).
"""
from __future__ import annotations

import contextlib
import importlib.util
import json
import os
import sys
import types
import warnings

from . import make_ref

HERE = os.path.dirname(os.path.abspath(__file__))
_REF_COPY = os.path.join(HERE, "_ref")


def ref_root(allow_copy: bool = True) -> str | None:
    """Path of the reference tree to load from, or None."""
    root = make_ref.reference_root()
    if root is not None:
        return root
    if allow_copy and os.path.isfile(os.path.join(_REF_COPY, "MANIFEST.json")):
        with open(os.path.join(_REF_COPY, "MANIFEST.json")) as f:
            manifest = json.load(f)
        for rel, digest in manifest["files"].items():
            path = os.path.join(_REF_COPY, rel)
            if not os.path.isfile(path) or make_ref.sha256(path) != digest:
                raise RuntimeError(f"oracle/_ref/{rel} does not match its MANIFEST.json hash: re-run oracle/make_ref.py")
        return _REF_COPY
    return None


def kind(root: str | None) -> str:
    """How bench.py labels a run from this root: the mounted tree and its verbatim copy are both 'reference'."""
    return "reference" if root else "port"


@contextlib.contextmanager
def _patched_modules(table: dict):
    saved = {k: sys.modules.get(k) for k in table}
    sys.modules.update(table)
    try:
        yield
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_file(root: str, rel: str, name: str, imports: dict | None = None):
    """Executes `root/rel` as module `tdssf_ref.<name>` with `imports` visible in sys.modules while it loads."""
    path = os.path.join(root, rel)
    spec = importlib.util.spec_from_file_location("tdssf_ref." + name, path)
    mod = importlib.util.module_from_spec(spec)
    with _patched_modules(imports or {}), warnings.catch_warnings():
        warnings.simplefilter("ignore", SyntaxWarning)  # lovasz_losses.py:79,95,139,186 ("is" with a literal)
        spec.loader.exec_module(mod)
    assert os.path.abspath(mod.__file__).startswith(os.path.abspath(root) + os.sep), (mod.__file__, root)
    return mod


def _pkg(name: str, **attrs):
    m = types.ModuleType(name)
    m.__path__ = []  # a package, so that `from name.sub import x` consults sys.modules['name.sub']
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


def load_hot_path(root: str | None = None):
    """The reference's hot-path modules: fusion_model, losses, evaluation, lovasz_losses."""
    root = root or ref_root()
    if root is None:
        raise FileNotFoundError("no reference tree: neither $TDSSF_REFERENCE / /root/reference nor oracle/_ref/ exists")
    lovasz = load_file(root, "LovaszSoftmax/pytorch/lovasz_losses.py", "lovasz_losses")
    fusion_model = load_file(root, "utils/fusion_model.py", "fusion_model")
    evaluation = load_file(root, "utils/evaluation.py", "evaluation")
    losses = load_file(root, "utils/losses.py", "losses", {
        "LovaszSoftmax": _pkg("LovaszSoftmax"),
        "LovaszSoftmax.pytorch": _pkg("LovaszSoftmax.pytorch", lovasz_losses=lovasz),
        "LovaszSoftmax.pytorch.lovasz_losses": lovasz,
    })
    return types.SimpleNamespace(root=root, fusion_model=fusion_model, losses=losses, evaluation=evaluation,
                                 lovasz_losses=lovasz)


class _Wandb(types.ModuleType):
    """Stand-in for wandb (absent, and a network side effect: utils/train.py:17-24): records what the loop logs."""

    def __init__(self):
        super().__init__("wandb")
        self.logged = []

    def init(self, **kwargs):
        return self

    def log(self, payload, **kwargs):
        self.logged.append(dict(payload))


def load_loops(evaluation_module, root: str | None = None):
    """The reference's loops, unmodified: utils/train.py (train_model) and utils/test.py (test_model, test_sample).
    `evaluation_module` is what their `from utils.evaluation import evaluate` resolves to — the drop-in module in the
    "runs unchanged over tdssf" tests, the reference's own in a pure-reference run.  wandb and utils.plot (matplotlib /
    plotly / cv2, absent here) are replaced by recording stand-ins."""
    root = root or ref_root()
    if root is None:
        raise FileNotFoundError("no reference tree to load the loops from")
    wandb = _Wandb()
    plot_calls = []
    plot = types.ModuleType("utils.plot")
    plot.plot_images_with_point_cloud = lambda **kw: plot_calls.append(kw)
    imports = {"wandb": wandb, "utils": _pkg("utils", evaluation=evaluation_module, plot=plot),
               "utils.evaluation": evaluation_module, "utils.plot": plot}
    train = load_file(root, "utils/train.py", "train", imports)
    test = load_file(root, "utils/test.py", "test", imports)
    return types.SimpleNamespace(root=root, train=train, test=test, wandb=wandb, plot_calls=plot_calls)
