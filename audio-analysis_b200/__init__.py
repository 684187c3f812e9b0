"""cacophony-b200: B200-native feature hot path of TheCacophonyProject/audio-analysis.

Submodules mirror the reference's module names for this path (same call signatures):
  custommel, identify_tracks, identify_morepork, cacophony_index, magtransform
plus `engine` (ctypes binding of the C-ABI library built from csrc/) and `synth`
(seeded synthetic recordings).  All arithmetic runs in hand-written sm_100a kernels;
there is no CPU fallback - importing `engine` without the built library raises.
"""
__version__ = "0.1.0"
