"""Importable alias for the package directory `audio-analysis_b200/` (a hyphen is not a
legal Python identifier, so `import audio_analysis_b200` resolves here and every
submodule is served from `audio-analysis_b200/`)."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "audio-analysis_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
