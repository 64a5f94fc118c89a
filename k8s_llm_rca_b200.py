"""Import alias for the package directory ``k8s-llm-rca_b200/``.

The directory name carries a hyphen (it is the name the project brief fixes), which Python
cannot import directly.  This module makes ``import k8s_llm_rca_b200`` resolve to that
directory: it sets ``__path__`` so that sub-modules are found there and runs the package's
own ``__init__.py`` in this namespace.
"""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "k8s-llm-rca_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
del _f
