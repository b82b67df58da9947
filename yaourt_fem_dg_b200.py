"""Import shim: the package directory is `yaourt-fem-dg_b200/` (not a Python identifier), so this
module turns itself into that package.  `import yaourt_fem_dg_b200 as yfd`."""
import os as _os

__package__ = __name__
__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "yaourt-fem-dg_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
