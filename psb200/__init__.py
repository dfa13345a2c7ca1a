"""Importable alias of the ``parameter-server-architecture-on-apache-kafka_b200`` source tree.

The product directory carries the project's full (hyphenated) name, which the
``import`` statement cannot spell.  This shim is a regular package whose
``__path__`` points at that directory, so ``import psb200.engine`` resolves to
``parameter-server-architecture-on-apache-kafka_b200/engine``.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
PACKAGE_DIR = _os.path.join(_ROOT, "parameter-server-architecture-on-apache-kafka_b200")
REPO_ROOT = _ROOT
__path__ = [PACKAGE_DIR]

with open(_os.path.join(PACKAGE_DIR, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(PACKAGE_DIR, "__init__.py"), "exec"))
