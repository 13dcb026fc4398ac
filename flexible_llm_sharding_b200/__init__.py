"""Import alias for the package directory `flexible-llm-sharding_b200/`.

The product lives in `flexible-llm-sharding_b200/` (the name the project layout prescribes); a
hyphen cannot appear in a Python import, so this stub points `__path__` at that directory and
runs its `__init__.py`.  `import flexible_llm_sharding_b200.utils` therefore loads
`flexible-llm-sharding_b200/utils.py`.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "flexible-llm-sharding_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
