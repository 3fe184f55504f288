"""Drop-in module names of the reference.

Importing this package puts its directory on sys.path so that the reference's
driver scripts (models/*.py) run unchanged:

    import ecolab_b200.compat
    from ecolab_model import spatial_ecolab as ecolab     # models/spatial_ecolab.py:1
    from ecolab import array_urand, myid, nprocs          # src/ecolab.cc:225-266
    from GUI import gui, statusBar                        # lib/GUI.py   (headless)
    from plot import plot                                 # lib/plot.py  (headless)
"""
import os
import sys

import importlib.util

_here = os.path.dirname(os.path.abspath(__file__))
if _here not in sys.path:
    sys.path.insert(0, _here)
# models/panmictic_ecolab.py imports tkinter unconditionally; interpreters built without it get an inert stand-in
if importlib.util.find_spec("tkinter") is None:
    sys.path.append(os.path.join(_here, "_stubs"))
