"""Drop-in for `simple_knn._C` (submodules/simple-knn/ext.cpp:15-17): exports distCUDA2."""
from .._C import distCUDA2  # noqa: F401
