"""Importable alias of the hyphen-named package directory
`distributed-continuous-training-with-airflow-pytorch-distributed-ddp-_b200/` (see its __init__)."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "distributed-continuous-training-with-airflow-pytorch-distributed-ddp-_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _fh:
    exec(compile(_fh.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _fh
