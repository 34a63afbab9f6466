"""Import shim: ``import fps_b200`` loads the package that lives in
``flink-parameter-server-1_b200/`` (a directory name Python cannot import directly).

Sub-modules resolve through ``__path__``, so ``fps_b200.parallel.engine`` is the file
``flink-parameter-server-1_b200/parallel/engine.py``.
"""
import pathlib as _pathlib

_real = _pathlib.Path(__file__).resolve().parent.parent / "flink-parameter-server-1_b200"
__path__.insert(0, str(_real))  # type: ignore[name-defined]
_init = _real / "__init__.py"
exec(compile(_init.read_text(), str(_init), "exec"))
