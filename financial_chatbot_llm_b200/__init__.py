"""Importable alias for the package directory ``financial-chatbot-llm_b200/`` (a hyphen
cannot appear in a Python module name). ``import financial_chatbot_llm_b200 as fcb``
exposes exactly the contents of that directory."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "financial-chatbot-llm_b200")
__path__.insert(0, _real)
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
