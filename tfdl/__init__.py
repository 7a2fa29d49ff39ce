"""Import alias: ``import tfdl`` loads the package that lives in
``tensorflow-distributed-learning_b200/`` (a directory name Python cannot import directly
because of the hyphens). Sub-modules resolve through ``__path__`` so ``tfdl.keras`` etc. work."""
import os as _os

_root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
_real = _os.path.join(_root, "tensorflow-distributed-learning_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
