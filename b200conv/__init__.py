"""Importable alias for the package that lives in
``parallel-image-convolution-using-mpi-openmp-and-cuda_b200/`` (a directory name Python
cannot import directly).  ``import b200conv`` executes that directory's ``__init__.py``
as this module and makes its sub-packages (``b200conv.ops``, ``b200conv.parallel``,
``b200conv.models``, ``b200conv.utils``) resolve there.
"""
import os as _os

_here = _os.path.dirname(_os.path.abspath(__file__))
_pkg = _os.path.join(_os.path.dirname(_here), "parallel-image-convolution-using-mpi-openmp-and-cuda_b200")
__path__.insert(0, _pkg)
with open(_os.path.join(_pkg, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_pkg, "__init__.py"), "exec"))
del _f
