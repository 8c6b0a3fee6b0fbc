"""Import alias: the package directory is named ``gaussian-process_b200`` (not a valid Python
identifier), so ``import gaussian_process_b200`` resolves here and this module turns itself into
that package (its ``__path__`` points at the hyphenated directory)."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "gaussian-process_b200")]
__file__ = _os.path.join(__path__[0], "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
