def __getattr__(name):
    def _unavailable(*args, **kwargs):
        raise RuntimeError("matplotlib is not installed in this image (import shim)")
    return _unavailable
