"""Import shim: the framework lives in ``distributed-llama-multiusers_b200/`` (a directory
name Python cannot import directly because of the hyphens).  ``import dllama_b200`` resolves
every submodule from that directory.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "distributed-llama-multiusers_b200")
__path__.insert(0, _real)
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
