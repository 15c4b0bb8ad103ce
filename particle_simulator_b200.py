"""Import alias: the package directory is ``particle-simulator_b200/`` (hyphenated, as the
project layout names it), which Python cannot import by name.  This module loads it and
installs it in sys.modules under the importable name ``particle_simulator_b200``."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "particle-simulator_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
