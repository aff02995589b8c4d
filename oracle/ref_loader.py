"""TEST INFRASTRUCTURE ONLY -- loads the UNMODIFIED reference module
/root/reference/code/models/TransMIL.py by path, with `oracle/` on sys.path so
that `from nystrom_attention import NystromAttention` (TransMIL.py:5) resolves
to the restated package.  Nothing is copied; the file is executed where it lies.

Usable only in the build container (the GPU box has no /root/reference): it is
the tool that PINS oracle/transmil_oracle.py and generates tests/golden/*.
"""
import importlib.util
import os
import sys

import torch

REF_FILE = "/root/reference/code/models/TransMIL.py"
_HERE = os.path.dirname(os.path.abspath(__file__))


def available():
    return os.path.isfile(REF_FILE)


def load_reference_module():
    if _HERE not in sys.path:
        sys.path.insert(0, _HERE)
    spec = importlib.util.spec_from_file_location("_reference_TransMIL", REF_FILE)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _NoCuda:
    """TransMIL.py:184 hard-codes `.cuda()` on the cls token; on a CPU-only
    host that raises.  Neutralise that one call for the duration of a forward."""

    def __enter__(self):
        self._orig = torch.Tensor.cuda
        if not torch.cuda.is_available():
            torch.Tensor.cuda = lambda self, *a, **k: self
        return self

    def __exit__(self, *exc):
        torch.Tensor.cuda = self._orig
        return False


def build_reference(n_classes, in_features, seed=2021, fc1_1024_else_form=True, quiet=True):
    """Construct the reference TransMIL under torch.manual_seed(seed).

    in_features == 1024: the reference's own branch (TransMIL.py:117-121) raises
    at the first forward (LayerNorm(512) on a 1024-wide tensor).  SURVEY.md 8a/A1
    decision: replace `_fc1` post-construction by Sequential(Linear(1024,512),
    GELU()) -- the form of code/models/MDMIL.py:66 and of TransMIL's own else
    branch (TransMIL.py:130-133).  The replacement Linear is drawn from a
    generator seeded `seed + 1` so the rest of the init stream is untouched.
    """
    mod = load_reference_module()
    torch.manual_seed(seed)
    if quiet:
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            model = mod.TransMIL(n_classes=n_classes, in_features=in_features)
    else:
        model = mod.TransMIL(n_classes=n_classes, in_features=in_features)
    if in_features == 1024 and fc1_1024_else_form:
        model._fc1 = make_fc1_1024(seed)
    model.eval()
    return model


def make_fc1_1024(seed):
    import torch.nn as nn
    state = torch.random.get_rng_state()
    torch.manual_seed(seed + 1)
    fc1 = nn.Sequential(nn.Linear(1024, 512), nn.GELU())
    torch.random.set_rng_state(state)
    return fc1


def reference_forward(model, x, return_attn=False):
    with _NoCuda(), torch.no_grad():
        return model(x, return_attn=return_attn)


def load_reference_sibling(name):
    """Executes the unmodified /root/reference/code/models/<name>.py (CTMIL, TransformerMIL) where it lies.  Those files import
    `from ._transformer import ...`, so they are loaded as members of a synthetic package whose __path__ is the reference's
    models directory (its own __init__.py, which pulls in Lightning, is not executed)."""
    import types
    if _HERE not in sys.path:
        sys.path.insert(0, _HERE)
    pkg = "_reference_models"
    if pkg not in sys.modules:
        m = types.ModuleType(pkg)
        m.__path__ = [os.path.dirname(REF_FILE)]
        sys.modules[pkg] = m
    return importlib.import_module(f"{pkg}.{name}")
