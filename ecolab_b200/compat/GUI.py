"""Headless stand-in for lib/GUI.py (Tk): there is no graphics stack on this box.
`gui(step, model)` runs `step` ECOLAB_STEPS times (default 0) instead of opening a window."""
import os as _os

windows = []


class _Label:
    text = ""

    def configure(self, **kw):
        self.text = kw.get("text", self.text)


statusBar = _Label()
performance = _Label()


def gui(step, model=None):
    for _ in range(int(_os.environ.get("ECOLAB_STEPS", "0"))):
        step()
