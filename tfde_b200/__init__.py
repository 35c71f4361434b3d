"""Import alias for the framework package.

The product package lives in ``tensorflow-distributed-example_b200/`` (the directory name the
project mandates).  A hyphen cannot appear in a Python ``import`` statement, so this tiny
package re-points its ``__path__`` at that directory and executes its ``__init__``: after
``import tfde_b200`` every submodule (``tfde_b200.ops``, ``tfde_b200.keras`` ...) is loaded from
the real tree.  Nothing else lives here.
"""
import os as _os

_REAL = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "tensorflow-distributed-example_b200")
__path__[:] = [_REAL]
__file__ = _os.path.join(_REAL, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
