def __getattr__(name):
    from .pyplot import _Anything
    return _Anything()
