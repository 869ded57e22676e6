"""Import shim: the package lives in `radio-sim_b200/` (the name the layout asks for), which
Python cannot import directly because of the hyphen.  This module makes that directory the
package path of `radio_sim_b200` and runs its __init__."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "radio-sim_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
