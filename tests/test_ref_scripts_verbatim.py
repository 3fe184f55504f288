"""The driver scripts under tests/ref_scripts/ are the reference's own files, byte for byte (checked where the reference
checkout exists: the build container; the GPU box has none and skips)."""
import os
import re

import pytest

REF = "/root/reference"
HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_scripts")
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present on this machine")


def test_model_scripts_are_verbatim():
    for f in ("spatial_ecolab", "steady_migration", "fragmentation_pump", "SAR", "SAR1", "panmictic_ecolab", "pred-prey", "migration_test"):
        assert open(os.path.join(HERE, f + ".py"), "rb").read() == open(os.path.join(REF, "models", f + ".py"), "rb").read(), f


def test_shell_test_bodies_are_verbatim():
    for name in ("predator-prey", "condense", "spatial_ecolab"):
        src = open(os.path.join(REF, "test", "00", name + ".sh")).read()
        body = re.search(r"cat >input\.py <<EOF\n(.*?)\nEOF\n", src, re.S).group(1) + "\n"
        assert open(os.path.join(HERE, "test00_" + name.replace("-", "_") + ".py")).read() == body, name
