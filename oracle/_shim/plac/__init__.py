"""Stub of the `plac` CLI helper (imported, never used on the forward path, by
/root/reference/ReSIDE/models/lasinger2019.py:1).  TEST INFRASTRUCTURE."""


class Annotation(tuple):
    def __new__(cls, *a, **k):
        return super().__new__(cls, a)


def annotations(**_kw):
    def deco(f):
        return f
    return deco


def call(f, *a, **k):
    raise RuntimeError("plac stub: CLIs are out of scope")
