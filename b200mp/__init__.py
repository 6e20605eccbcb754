"""Import alias: ``b200mp`` -> ``models-for-relational-multimodal-data_b200/`` (whose name is not
a valid Python identifier).  The package body lives there."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "models-for-relational-multimodal-data_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
