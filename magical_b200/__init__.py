"""Import alias: the product package lives in ``magical-1_b200/`` (a directory name Python cannot
import directly); ``import magical_b200`` loads it from there."""
import importlib.util
import os
import sys

_real = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'magical-1_b200')
_spec = importlib.util.spec_from_file_location('magical_b200', os.path.join(_real, '__init__.py'),
                                               submodule_search_locations=[_real])
_mod = importlib.util.module_from_spec(_spec)
sys.modules['magical_b200'] = _mod
_spec.loader.exec_module(_mod)
