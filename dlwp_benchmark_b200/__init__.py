"""Import alias: the package directory is ``dlwp-benchmark_b200/`` (not a valid Python
identifier), so this shim makes ``import dlwp_benchmark_b200`` resolve to it."""
import os as _os

_impl = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "dlwp-benchmark_b200")
__path__ = [_impl]
with open(_os.path.join(_impl, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_impl, "__init__.py"), "exec"))
