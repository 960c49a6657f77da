"""Loader of the committed golden fixtures (made by make_golden.py from oracle/_ref)."""
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def path(name):
    return os.path.join(_HERE, name + ".npz")


def load(name):
    with np.load(path(name)) as z:
        return {k: z[k] for k in z.files}
