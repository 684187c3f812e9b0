class FFmpegAudioFile:  # pragma: no cover - decode is out of scope
    def __init__(self, *a, **k):
        raise NotImplementedError("ffmpeg decode is out of scope")
