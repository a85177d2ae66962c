"""Import alias: ``import ppb200`` loads the package that lives in the (non-importable, hyphenated) directory
``distributed-training-with-pipeline-parallelism_b200/``.  Sub-modules resolve through ``__path__``, so
``ppb200.ops``, ``ppb200.parallel.schedule`` ... are the files of that directory."""
import os as _os

_PKG_DIR = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                         "distributed-training-with-pipeline-parallelism_b200")
__path__ = [_PKG_DIR]
__file__ = _os.path.join(_PKG_DIR, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
