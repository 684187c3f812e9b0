"""Execute the reference's own src/*.py UNCHANGED under the stand-in modules in
oracle/stubs (librosa / tensorflow / audioread).  Only works where /root/reference is
mounted (the build container); the GPU box uses the committed fixtures instead."""
import importlib
import os
import sys

REFERENCE_SRC = os.environ.get("CACOPHONY_REFERENCE_SRC", "/root/reference/src")
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")


def available():
    return os.path.isfile(os.path.join(REFERENCE_SRC, "identify_tracks.py"))


def load(name):
    """Import reference module `name` (custommel, identify_tracks, identify_morepork,
    cacophony_index, common, magtransform, magtransformv2) under a private prefix-free
    namespace; returns the module object."""
    if not available():
        raise RuntimeError("reference sources not present at %s" % REFERENCE_SRC)
    saved_path = list(sys.path)
    saved_mods = {k: sys.modules.get(k) for k in
                  ("librosa", "librosa.feature", "tensorflow", "audioread", "audioread.ffdec",
                   "custommel", "common", name)}
    try:
        for k in saved_mods:
            sys.modules.pop(k, None)
        sys.path.insert(0, REFERENCE_SRC)
        sys.path.insert(0, _STUBS)
        mod = importlib.import_module(name)
        return mod
    finally:
        sys.path[:] = saved_path
        for k, v in saved_mods.items():
            if v is not None:
                sys.modules[k] = v
            else:
                sys.modules.pop(k, None)
