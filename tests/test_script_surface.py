"""Script-level acceptance, static half: every attribute the reference's own driver scripts touch
on the model object (`from ecolab_model import spatial_ecolab as ecolab`, models/*.py) and every name they import from
the `ecolab`, `GUI` and `plot` modules exists on the host-side mirror.  The scripts are only PARSED (ast), where they
lie under /root/reference -- that tree exists in the build container, not on the GPU box, so the test skips there.
The dynamic half (the same call sequences executed on the GPU) is tests/test_gpu_scripts.py.
"""
import ast
import glob
import os
from unittest import mock

import pytest

REF = "/root/reference/models"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present on this machine")

MODEL_MODULE = "ecolab_model"
SHIM_MODULES = ("ecolab", "GUI", "plot")


def _chains(tree, roots):
    """Attribute chains rooted at one of `roots`: ecolab.interaction.val -> ('ecolab', ['interaction', 'val']).
    A call in the middle of a chain is kept as the marker '()' so that cell(i, j).density resolves on the cell object."""
    out = []

    def walk(node):
        parts = []
        while True:
            if isinstance(node, ast.Attribute):
                parts.append(node.attr)
                node = node.value
            elif isinstance(node, ast.Call):
                parts.append("()")
                node = node.func
            elif isinstance(node, ast.Name):
                return (node.id, parts[::-1]) if node.id in roots else None
            else:
                return None

    for n in ast.walk(tree):
        if isinstance(n, ast.Attribute):
            c = walk(n)
            if c:
                out.append(c)
    return out


def _scripts():
    return sorted(glob.glob(os.path.join(REF, "*.py")))


def _mock_models():
    """Mirror objects built without a GPU: the C ABI is replaced by a mock whose calls all succeed (status 0)."""
    import ecolab_b200
    from ecolab_b200 import _cabi
    lib = mock.MagicMock()
    lib.eco_last_error.return_value = b""
    for name in dir(lib):
        pass
    lib.configure_mock(**{})
    with mock.patch.object(_cabi, "lib", return_value=lib):
        lib.eco_create.return_value = 0
        for f in ("eco_set_grid", "eco_set_grid_slab", "eco_get_scalar", "eco_set_scalar", "eco_destroy"):
            getattr(lib, f).return_value = 0
        pan = ecolab_b200.PanmicticModel(0)
        spa = ecolab_b200.SpatialModel(0, None)
    return pan, spa


def _resolve(obj, parts):
    for p in parts:
        if p == "()":
            # only object-returning calls matter for the surface: cell(i, j) and nsp()
            name = getattr(obj, "__name__", "")
            if name == "cell":
                from ecolab_b200.model import _Cell
                obj = _Cell(obj.__self__, 0)
                continue
            return True   # value-returning call: the chain ends here for the purposes of this test
        if not hasattr(obj, p):
            return False
        obj = getattr(obj, p)
    return True


def test_reference_scripts_are_found():
    assert len(_scripts()) >= 6


@pytest.mark.parametrize("path", _scripts() if os.path.isdir(REF) else [])
def test_model_surface_of_script(path):
    tree = ast.parse(open(path).read())
    alias = {}
    for n in ast.walk(tree):
        if isinstance(n, ast.ImportFrom) and n.module == MODEL_MODULE:
            for a in n.names:
                alias[a.asname or a.name] = a.name
    if not alias:
        pytest.skip("script does not drive the ecolab model")
    pan, spa = _mock_models()
    missing = []
    for root, parts in _chains(tree, set(alias)):
        obj = spa if alias[root] == "spatial_ecolab" else pan
        if not _resolve(obj, parts):
            missing.append(root + "." + ".".join(parts))
    assert not missing, f"{os.path.basename(path)} uses members the mirror lacks: {sorted(set(missing))}"


@pytest.mark.parametrize("path", _scripts() if os.path.isdir(REF) else [])
def test_shim_module_names_of_script(path):
    import ecolab_b200.compat  # noqa: F401  (puts the shim modules on sys.path)
    import importlib
    tree = ast.parse(open(path).read())
    missing = []
    for n in ast.walk(tree):
        if isinstance(n, ast.ImportFrom) and n.module in SHIM_MODULES:
            mod = importlib.import_module(n.module)
            missing += [f"{n.module}.{a.name}" for a in n.names if a.name != "*" and not hasattr(mod, a.name)]
    assert not missing, f"{os.path.basename(path)} imports names the shims lack: {missing}"
