"""Import shim: the package directory is ``llm-mcp_b200/`` (not a legal Python identifier),
so ``import llm_mcp_b200`` resolves here and points the package path at that directory."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "llm-mcp_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _f.name, "exec"))
del _f
