"""Import alias: the product package lives in `instancesegmentation-_b200/` (a name Python
cannot import directly); `import isb` exposes it under an importable name."""
import os as _os

_pkg = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "instancesegmentation-_b200")
__path__.insert(0, _pkg)
with open(_os.path.join(_pkg, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_pkg, "__init__.py"), "exec"))
del _os, _f
