"""Import shim: the package directory is `dart-ray_b200/` (not a valid Python identifier)."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "dart-ray_b200")]
__file__ = _os.path.join(__path__[0], "__init__.py")
exec(compile(open(__file__).read(), __file__, "exec"))
