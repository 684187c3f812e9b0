"""librosa.feature stand-in: the htk=False / use_mfcc branches of get_spect
(src/identify_tracks.py:229-238, 269-280) are off by default and not restated."""


def melspectrogram(*a, **k):
    raise NotImplementedError("librosa.feature.melspectrogram not restated (htk=False branch)")


def mfcc(*a, **k):
    raise NotImplementedError("librosa.feature.mfcc not restated (use_mfcc branch)")
