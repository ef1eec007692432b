"""Import shim: the product package lives in ``int-act_b200/`` (the name the task fixes, which is not a
valid Python identifier); this package makes it importable as ``int_act_b200``."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "int-act_b200")
__path__ = [_real]
with open(_os.path.join(_real, "_init.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "_init.py"), "exec"))
