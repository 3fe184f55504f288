"""GPU: the bounds-assert build.  compute-sanitizer cannot run on the GPU pool; libecolab_b200_bounds.so is the same
sources compiled with -DECO_BOUNDS, in which every global access of the kernels that move densities (generate's bulk copies
and stores, migrate's loads and stores incl. the cells across tile boundaries, compaction, re-layout, transposes, births,
mutate's records) is checked against the registered buffer ranges.  A worker process runs the stages over a spread of
shapes (all tile widths, ragged tiles, forced unit lengths, capacity growth, host streaming in place) through that
library: not one access may fall outside, and the negative control (one probe 4 bytes past the end) must be caught."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(os.path.dirname(HERE), "ecolab_b200", "libecolab_b200_bounds.so")


def test_no_access_outside_the_registered_buffers():
    assert os.path.exists(LIB), "build() makes libecolab_b200_bounds.so"
    env = dict(os.environ, ECOLAB_B200_LIB=LIB)
    r = subprocess.run([sys.executable, os.path.join(HERE, "bounds_worker.py")], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("BOUNDS_RESULT ")][-1]
    res = json.loads(line[len("BOUNDS_RESULT "):])
    assert res["selftest"] == 1, "the checker must catch the deliberate out-of-range probe (and only that one)"
    assert res["cases"] == 12 and res["bad"] == 0, res
