"""Inert stand-in for tkinter, used ONLY when the interpreter has none (this image): models/panmictic_ecolab.py does
`from tkinter import Tk, ttk` at import time although its window code is commented out.  There is no graphics stack on
this box; creating a window raises."""


class Tk:
    def __init__(self, *a, **k):
        raise RuntimeError("no graphics stack on this box: tkinter windows are unavailable")


class _Ttk:
    def __getattr__(self, n):
        raise RuntimeError("no graphics stack on this box: tkinter widgets are unavailable")


ttk = _Ttk()
