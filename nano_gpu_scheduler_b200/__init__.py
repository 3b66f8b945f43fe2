"""Import alias: the product package lives in ``nano-gpu-scheduler_b200/`` (the name the build
brief fixes); Python cannot import a hyphenated directory, so this package points its search
path there and re-exports it."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "nano-gpu-scheduler_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
