"""Import alias: the product package lives in the directory ``wrf-partmc_b200/``
(the name the project layout prescribes; a hyphen is not importable), so
``import wrf_partmc_b200`` loads that directory as a package."""
import importlib.util
import os
import sys

_d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "wrf-partmc_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_d, "__init__.py"),
                                               submodule_search_locations=[_d])
_m = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _m
_spec.loader.exec_module(_m)
