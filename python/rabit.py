"""Drop-in location of the reference's ``python/rabit.py``: re-exports :mod:`rabit_b200.rabit`."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from rabit_b200.rabit import *  # noqa: F401,F403,E402
from rabit_b200.rabit import __version__, _loadlib, _unloadlib  # noqa: F401,E402
