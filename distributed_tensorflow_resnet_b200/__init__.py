"""Importable alias of the package directory `distributed-tensorflow-resnet_b200/` (a hyphen is
not a legal Python identifier).  All code lives there; this shim only redirects the import."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                          "distributed-tensorflow-resnet_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
