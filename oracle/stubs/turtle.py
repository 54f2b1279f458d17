"""Import stub (see README.md): ffmlp/ffmlp.py:2 does `from turtle import backward, forward` (an editor auto-import; turtle needs
tkinter, absent here).  The two names are never used."""


def backward(*a, **k):
    raise RuntimeError("stub turtle.backward")


def forward(*a, **k):
    raise RuntimeError("stub turtle.forward")
